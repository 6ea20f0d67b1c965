"""train.py of the reference's documented command line (/root/reference/README.md:59; the file itself is absent from
the snapshot):

    python train.py --dataroot=./datasets/cheetah --env_type=cheetah --netG=s2p --batchSize=16 --gpu_ids=0

GAN training of the S2P generator (L1 + hinge GAN + feature matching + VGG19 perceptual, Adam) on
(image_observations[i], next_observations[i]) -> image_observations_tp1[i] rows.  `--dataroot` is a directory of
`<key>.npy` files (s2p_b200.synthesis.save_dataset_dir: the flat twin of the reference's `cheetah.hdf5`; h5py is not
in this image).  Checkpoints: `<checkpoints_dir>/<env_type>_<epoch>.pth` holds the generator's state_dict (the layout of
the reference's released `cheetah_30.pth`, README.md:19-25) and `<env_type>_latest_trainer.pth` the full resume state.
Multi-GPU: launch with torchrun; one process per GPU, gradients all-reduced with NCCL.
"""
import argparse
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def extra_args(argv):
    p = argparse.ArgumentParser(add_help=False)
    p.add_argument("--niter", type=int, default=30, help="epochs (the released weights are *_30.pth)")
    p.add_argument("--max_iters", type=int, default=0, help="stop after this many iterations (0 = run all epochs)")
    p.add_argument("--checkpoints_dir", type=str, default="./checkpoints")
    p.add_argument("--save_epoch_freq", type=int, default=10)
    p.add_argument("--print_freq", type=int, default=100)
    p.add_argument("--continue_train", action="store_true")
    p.add_argument("--no_vgg_loss", action="store_true")
    p.add_argument("--seed", type=int, default=0)
    return p.parse_known_args(argv)[0]


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    import s2p_b200
    from s2p_b200.synthesis import load_dataset_dir, shard_range
    from s2p_b200.training import S2PTrainer
    ex = extra_args(argv)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    opt = s2p_b200.make_options(argv)
    if not opt.gpu_ids:
        raise RuntimeError("--gpu_ids=-1: s2p_b200 has no CPU path (run the oracle for a CPU reference)")
    dev_index = opt.gpu_ids[local % len(opt.gpu_ids)] if world > 1 else opt.gpu_ids[0]
    torch.cuda.set_device(dev_index)
    dev = torch.device("cuda", dev_index)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    opt.gpu_ids = [dev_index]
    data = load_dataset_dir(opt.dataroot, keys=["image_observations", "next_observations", "image_observations_tp1"])
    n = data["image_observations"].shape[0]
    if data["next_observations"].shape[1] != opt.state_dim:
        raise RuntimeError(f"dataset states are {data['next_observations'].shape[1]}-d, --env_type={opt.env_type} expects {opt.state_dim}")
    torch.manual_seed(ex.seed)                    # same initial weights on every rank
    netG, netD = s2p_b200.define_G(opt), s2p_b200.define_D(opt)
    vgg = None if ex.no_vgg_loss else s2p_b200.VGG19FeaturesB200().to(dev)
    trainer = S2PTrainer(netG, netD, world_size=world, vgg=vgg)
    os.makedirs(ex.checkpoints_dir, exist_ok=True)
    latest = os.path.join(ex.checkpoints_dir, f"{opt.env_type}_latest_trainer.pth")
    start_epoch, it = 1, 0
    if ex.continue_train and os.path.exists(latest):
        ck = torch.load(latest, map_location=dev)
        trainer.load_state_dict(ck["trainer"])
        start_epoch, it = ck["epoch"] + 1, ck["iters"]
        if rank == 0:
            print(f"resumed from {latest}: epoch {ck['epoch']}, {it} iterations")
    B = opt.batchSize
    lo, hi = shard_range(n, rank, world)          # each rank draws batches from its own contiguous rows
    rs = np.random.RandomState(ex.seed + 1000 * rank)
    to_f = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev).permute(0, 3, 1, 2).float().div_(127.5).sub_(1.0)
    graphed = False
    t0 = time.time()
    # every rank runs the SAME number of steps per epoch (each step is a collective when world > 1): shard_range gives
    # the first n % world ranks one extra row, which can cross a batch boundary (n=1001, world=2, B=167: 3 vs 2 steps)
    steps_per_epoch = (n // world) // B
    for epoch in range(start_epoch, ex.niter + 1):
        perm = lo + rs.permutation(hi - lo)
        for b0 in range(0, steps_per_epoch * B, B):  # drop the ragged tail (fixed batch: the step is a CUDA graph)
            idx = np.sort(perm[b0:b0 + B])
            prev, real = to_f(data["image_observations"][idx]), to_f(data["image_observations_tp1"][idx])
            state = torch.from_numpy(np.ascontiguousarray(data["next_observations"][idx], dtype=np.float32)).to(dev)
            if world == 1 and B <= 32:
                if not graphed:
                    trainer.capture(prev, state, real)
                    graphed = True
                g_losses, d_loss = trainer.train_step_graphed(prev, state, real)
            else:
                g_losses, d_loss = trainer.train_step(prev, state, real)
            it += 1
            if rank == 0 and it % ex.print_freq == 0:
                msg = " ".join(f"{k}={float(v):.4f}" for k, v in g_losses.items())
                print(f"epoch {epoch} iter {it} {msg} D={float(d_loss):.4f} ({(time.time() - t0) / it * 1e3:.1f} ms/iter)", flush=True)
            if ex.max_iters and it >= ex.max_iters:
                break
        if rank == 0 and (epoch % ex.save_epoch_freq == 0 or epoch == ex.niter or (ex.max_iters and it >= ex.max_iters)):
            torch.save(netG.state_dict(), os.path.join(ex.checkpoints_dir, f"{opt.env_type}_{epoch}.pth"))
            torch.save({"trainer": trainer.state_dict(), "epoch": epoch, "iters": it}, latest)
        if ex.max_iters and it >= ex.max_iters:
            break
    if world > 1:
        dist.destroy_process_group()
    return it


if __name__ == "__main__":
    main()

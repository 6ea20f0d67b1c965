"""simple_test.py of the reference's documented command line (/root/reference/README.md:33; the file itself is absent):

    python simple_test.py --env_type=cheetah --dataroot=./datasets/cheetah --netG=s2p --start_idx=0 --seq_len=5 --gpu_ids=0

Loads `<checkpoints_dir>/<env_type>_<which_epoch>.pth` (README.md:19-25: `cheetah_30.pth`, `walker_30.pth`; random-init
weights when the file is missing), takes frame `start_idx` of the dataset and synthesises `seq_len` frames
autoregressively from the dataset's next states (img_{k+1} = netG(img_k, s_{k+1}), README.md:30-33).  Results go to
`<results_dir>/<env_type>_<start_idx>/` as PNG frames and one `frames.npy` [seq_len, H, W, 3] uint8.
`--gpu_ids=-1` is refused: this package has no CPU path (oracle/ holds the CPU restatement for tests).
"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    import s2p_b200
    from s2p_b200.synthesis import load_dataset_dir
    p = argparse.ArgumentParser(add_help=False)
    p.add_argument("--checkpoints_dir", type=str, default="./checkpoints")
    p.add_argument("--which_epoch", type=str, default="30")
    p.add_argument("--results_dir", type=str, default="./results")
    ex = p.parse_known_args(argv)[0]
    opt = s2p_b200.make_options(argv)
    if not opt.gpu_ids:
        raise RuntimeError("--gpu_ids=-1: s2p_b200 has no CPU path")
    dev = torch.device("cuda", opt.gpu_ids[0])
    torch.cuda.set_device(dev)
    netG = s2p_b200.define_G(opt).eval()
    ck = os.path.join(ex.checkpoints_dir, f"{opt.env_type}_{ex.which_epoch}.pth")
    if os.path.exists(ck):
        netG.load_state_dict(torch.load(ck, map_location=dev))
        print(f"loaded {ck}")
    else:
        print(f"{ck} not found: random-init weights")
    data = load_dataset_dir(opt.dataroot, keys=["image_observations", "next_observations"])
    i0, n = opt.start_idx, opt.seq_len
    if i0 < 0 or i0 + n > data["image_observations"].shape[0]:
        raise RuntimeError(f"--start_idx {i0} + --seq_len {n} runs past the dataset ({data['image_observations'].shape[0]} rows)")
    frame = torch.from_numpy(np.ascontiguousarray(data["image_observations"][i0:i0 + 1])).to(dev)
    img = frame.permute(0, 3, 1, 2).float().div_(127.5).sub_(1.0)
    states = torch.from_numpy(np.ascontiguousarray(data["next_observations"][i0:i0 + n], dtype=np.float32)).to(dev)
    with torch.no_grad():
        seq = netG.rollout(img, states.unsqueeze(0))[0]                       # [n, 3, H, W] in [-1, 1], fed back unquantised
    out = ((seq.permute(0, 2, 3, 1) + 1.0) * 127.5 + 0.5).floor().clamp_(0, 255).to(torch.uint8).cpu().numpy()
    dst = os.path.join(ex.results_dir, f"{opt.env_type}_{i0}")
    os.makedirs(dst, exist_ok=True)
    np.save(os.path.join(dst, "frames.npy"), out)
    try:
        from PIL import Image
        for k in range(n):
            Image.fromarray(out[k]).save(os.path.join(dst, f"frame_{k + 1:02d}.png"))
    except ImportError:
        pass
    print(f"wrote {n} frames to {dst}")
    return out


if __name__ == "__main__":
    main()

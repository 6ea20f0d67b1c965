"""CPU: host logic of the multi-GPU bulk path (world_size 2, gloo).  Generation shards
by contiguous index range with no data-path collective (SURVEY.md section 8 e); what
has to be right on the host is the partition and the max-over-ranks timing plumbing."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from s2p_b200.synthesis import BulkSynthesizer, shard_range


@pytest.mark.parametrize("n,world", [(1_000_000, 8), (10, 3), (7, 8), (0, 2), (1024, 1), (1001, 4)])
def test_shard_range_partitions_exactly(n, world):
    ranges = [shard_range(n, r, world) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == n
    for (a0, a1), (b0, b1) in zip(ranges, ranges[1:]):
        assert a1 == b0 and a0 <= a1
    sizes = [hi - lo for lo, hi in ranges]
    assert max(sizes) - min(sizes) <= 1


def test_shard_range_rejects_bad_rank():
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)
    with pytest.raises(ValueError):
        shard_range(10, 0, 0)


def test_bulk_synthesizer_needs_cuda():
    import s2p_b200
    net = s2p_b200.S2PGeneratorB200(s2p_b200.make_options())
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        BulkSynthesizer(net)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_rows, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_range(n_rows, rank, world)
    # each rank "synthesises" its slice: mark rows with rank+1 (stand-in for the GPU step)
    mine = torch.zeros(n_rows, dtype=torch.int64)
    mine[lo:hi] = rank + 1
    dist.all_reduce(mine, op=dist.ReduceOp.SUM)          # test-only gather of the coverage map
    t = torch.tensor([10.0 * (rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)             # bench.py's max-over-ranks timing
    if rank == 0:
        q.put((mine.tolist(), float(t.item())))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_gloo():
    world, n_rows = 2, 11
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_rows, q)) for r in range(world)]
    for p in procs:
        p.start()
    cover, tmax = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert cover == [1] * 6 + [2] * 5       # every row owned by exactly one rank, contiguous
    assert tmax == 20.0


def _grad_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from s2p_b200.training import allreduce_gradients
    torch.manual_seed(0)
    params = [torch.nn.Parameter(torch.zeros(s)) for s in ((3, 5), (7,), (2, 2, 2), (11,))]
    for i, p in enumerate(params):
        p.grad = torch.full_like(p, float((rank + 1) * (i + 1)))
    params[1].grad = None                                   # a parameter without gradient is skipped
    allreduce_gradients(params, world, bucket_elems=16)     # tiny buckets: several all-reduces
    if rank == 0:
        q.put([None if p.grad is None else p.grad.flatten()[0].item() for p in params])
    dist.barrier()
    dist.destroy_process_group()


def test_gradient_allreduce_two_ranks_gloo():
    """data-parallel training (BASELINE configs[4]): gradients are averaged over ranks in flat buckets."""
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_grad_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got == [1.5, None, 4.5, 6.0]       # mean over ranks of (rank+1)*(i+1)


def _bucket_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from s2p_b200.training import GradBuckets
    torch.manual_seed(0)                                     # same weights on both ranks
    net = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 4), torch.nn.Tanh(), torch.nn.Linear(4, 1))
    gb = GradBuckets(net.parameters(), world, bucket_elems=30)   # several buckets
    views_ok = all(p.grad.data_ptr() >= gb.buckets[gb.where[p]]["flat"].data_ptr() for p in net.parameters())
    outs = []
    for step in range(2):                                    # two steps: zero() must re-arm the buckets
        gb.zero()
        x = torch.full((3, 6), float(rank + 1 + step))
        net(x).sum().backward()
        launched_by_hooks = gb.launched
        gb.finish()
        outs.append([p.grad.clone() for p in net.parameters()])
    # reference: mean over ranks of the local gradients
    refs = []
    for step in range(2):
        acc = None
        for r in range(world):
            ref = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 4), torch.nn.Tanh(), torch.nn.Linear(4, 1))
            ref.load_state_dict(net.state_dict())
            ref(torch.full((3, 6), float(r + 1 + step))).sum().backward()
            g = [p.grad for p in ref.parameters()]
            acc = g if acc is None else [a + b for a, b in zip(acc, g)]
        refs.append([a / world for a in acc])
    err = max((a - b).abs().max().item() for o, r in zip(outs, refs) for a, b in zip(o, r))
    if rank == 0:
        q.put((len(gb.buckets), launched_by_hooks, views_ok, err))
    dist.barrier()
    dist.destroy_process_group()


def test_grad_buckets_overlapped_allreduce_two_ranks_gloo():
    """GradBuckets: p.grad are views into flat buckets, every bucket's all-reduce is launched from the backward
    hooks (not after backward), the result is the mean over ranks, and a second step works after zero()."""
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_bucket_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    nb, launched, views_ok, err = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert nb >= 3 and launched == nb and views_ok
    assert err <= 1e-6


def test_every_rank_runs_the_same_number_of_training_steps():
    """train.py: steps per epoch = (n // world) // B on EVERY rank; the per-rank row counts of shard_range can differ
    by one, which used to give ranks different step counts (a hung all-reduce)."""
    from s2p_b200.synthesis import shard_range
    for n, world, B in ((1001, 2, 167), (1000, 8, 16), (1023, 4, 51), (50, 3, 16)):
        steps = (n // world) // B
        old = {(shard_range(n, r, world)[1] - shard_range(n, r, world)[0]) // B for r in range(world)}
        assert all(shard_range(n, r, world)[1] - shard_range(n, r, world)[0] >= steps * B for r in range(world))
        if (n, world, B) == (1001, 2, 167):
            assert old == {2, 3} and steps == 2
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "train.py")).read()
    assert "steps_per_epoch = (n // world) // B" in src and "range(0, steps_per_epoch * B, B)" in src


def test_flat_dataset_dir_roundtrip(tmp_path):
    """the .npy-per-key twin of the HDF5 schema: exact round trip, row limit, memory-mapped frames"""
    import numpy as np
    from s2p_b200.synthesis import load_dataset_dir, save_dataset_dir, shard_range
    rs = np.random.RandomState(0)
    d = {"image_observations": rs.randint(0, 256, size=(10, 100, 100, 3)).astype(np.uint8),
         "next_observations": rs.randn(10, 17).astype(np.float32), "timeouts": np.zeros(10, np.float32),
         "slac_observation_indices": rs.randint(0, 10, size=(10, 9)).astype(np.int64)}
    save_dataset_dir(str(tmp_path / "ds"), d)
    back = load_dataset_dir(str(tmp_path / "ds"))
    assert set(back) == set(d)
    for k in d:
        assert back[k].dtype == d[k].dtype and np.array_equal(np.asarray(back[k]), d[k])
    assert isinstance(back["image_observations"], np.memmap)
    part = load_dataset_dir(str(tmp_path / "ds"), keys=["image_observations", "next_observations"], data_num=7, mmap=False)
    assert part["image_observations"].shape[0] == 7 and set(part) == {"image_observations", "next_observations"}
    lo, hi = shard_range(10, 1, 3)
    assert np.array_equal(np.asarray(back["image_observations"][lo:hi]), d["image_observations"][lo:hi])

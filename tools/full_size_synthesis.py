"""BASELINE configs[3] at its full size: bulk synthesis of N (default 1 000 000) dataset rows through the public
BulkSynthesizer (host uint8 frames + fp32 states in, host uint8 frames out), wall-clock timed end to end, with
size-independent checks on the way:
  * determinism   - the first chunk synthesised again at the end is byte-identical;
  * batch / shard invariance - 1024 rows of the first chunk synthesised on their own equal the same rows of the chunk;
  * checksum of checksums over every output chunk (printed, so two runs / two shardings can be compared).
Host memory stays bounded: one pool of `--chunk` distinct synthetic frames is reused as the input of every chunk
(each chunk gets its own states, so every output frame differs); one pinned output chunk is reused.
usage: python tools/full_size_synthesis.py [--rows 1000000] [--chunk 65536] [--batch 1024]
under torchrun each rank takes shard_range(rows, rank, world) and rank 0 reports the max wall-clock over ranks."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import s2p_b200
from s2p_b200.synthesis import BulkSynthesizer, shard_range

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=1_000_000)
ap.add_argument("--chunk", type=int, default=65536)
ap.add_argument("--batch", type=int, default=1024)
a = ap.parse_args()

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)

torch.manual_seed(1234)
opt = s2p_b200.make_options(env_type="cheetah", gpu_ids=str(local), init_type="scaled")
net = s2p_b200.define_G(opt).eval()
net.pack()
synth = BulkSynthesizer(net, batch=a.batch)

lo, hi = shard_range(a.rows, rank, world)
n = hi - lo
chunk = min(a.chunk, n)
rng = np.random.default_rng(2024)            # the SAME pool on every rank; states depend on the global row index
pool = torch.from_numpy(rng.integers(0, 256, size=(chunk, 100, 100, 3), dtype=np.uint8)).pin_memory()
outs = [torch.empty_like(pool).pin_memory() for _ in range(2)]      # chunk i is checksummed while chunk i+1 runs
out = outs[0]


def states_for(row0, m):
    g = torch.Generator().manual_seed(1_000_003 + row0)
    return torch.randn(m, opt.state_dim, generator=g).pin_memory()


# warm-up (kernel / workspace set-up) outside the timed region
synth.synthesize(pool[: a.batch], states_for(lo, a.batch), out[: a.batch])
torch.cuda.synchronize()
if world > 1:
    dist.barrier()

def checksum(t):
    v = t.numpy().reshape(-1)
    return int(v.view(np.uint64).sum(dtype=np.uint64)) if v.size % 8 == 0 else int(v.sum(dtype=np.uint64))


sums, first, pending = [], None, None
t0 = time.perf_counter()
for i, c0 in enumerate(range(0, n, chunk)):
    m = min(chunk, n - c0)
    st = states_for(lo + c0, m)
    buf = outs[i % 2]
    synth.synthesize(pool[:m], st, buf[:m])              # queues the whole chunk and returns
    done = torch.cuda.Event()
    done.record()
    if pending is not None:                              # host work on the previous chunk overlaps this one
        ev, pbuf, pm, pst = pending
        ev.synchronize()
        sums.append(checksum(pbuf[:pm]))
        if first is None:
            first = (pst, pbuf[:pm].clone())
    pending = (done, buf, m, st)
ev, pbuf, pm, pst = pending
ev.synchronize()
sums.append(checksum(pbuf[:pm]))
if first is None:
    first = (pst, pbuf[:pm].clone())
t1 = time.perf_counter()
wall = t1 - t0
if world > 1:
    t = torch.tensor([wall], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    wall = float(t.item())

# ---- size-independent checks on this rank's first chunk
st0, ref0 = first
m0 = ref0.shape[0]
again = synth.synthesize(pool[:m0], st0, torch.empty_like(ref0).pin_memory())
torch.cuda.synchronize()
deterministic = bool(torch.equal(again, ref0))
k = min(1024, m0)
alone = BulkSynthesizer(net, batch=128).synthesize(pool[:k].clone().pin_memory(), st0[:k].clone().pin_memory())
torch.cuda.synchronize()
batch_invariant = bool(torch.equal(alone, ref0[:k]))
cc = int(np.bitwise_xor.reduce(np.array(sums, dtype=np.uint64)))
ok = torch.tensor([int(deterministic and batch_invariant)], device=dev)
if world > 1:
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
if rank == 0:
    print(json.dumps({"workload": "BASELINE configs[3] bulk synthesis, full size", "rows": a.rows, "n_gpus": world,
                      "rows_per_rank": n, "chunk_rows": chunk, "wall_s": wall, "frames_per_sec": a.rows / wall,
                      "h2d_bytes": synth.h2d_bytes, "d2h_bytes": synth.d2h_bytes, "deterministic": deterministic,
                      "batch_invariant": batch_invariant, "all_ranks_ok": bool(ok.item()),
                      "rank0_checksum_of_checksums": f"{cc:016x}", "chunks": len(sums),
                      "includes": "pinned-host H2D of every row, generator, D2H of every frame, per-chunk host checksum"}))
if world > 1:
    dist.destroy_process_group()
sys.exit(0 if bool(ok.item()) else 1)

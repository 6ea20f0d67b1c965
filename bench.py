"""Benchmark of the S2P generator hot path (BASELINE.json metric: generated frames/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl reference]

One "step" = one pass of the generator over one batch of B synthetic dataset rows
per rank (uint8 HWC frame + fp32 state -> uint8 HWC next frame): the unit of work of
BASELINE.json configs[3] (bulk synthesis, batch-sharded over 1/2/4/8 GPUs, weak
scaling: B rows per rank per step) and, per frame, the same forward as configs[2]
(the 5-step rollout at batch 1024, also timed as a side measurement `rollout`).  configs[1]
(GAN training step, batch 16) is timed beside it as `training_step`; `--workload train` makes
the training step the headline line (configs[4]: batch 256 per GPU, NCCL gradient all-reduce).
  value : frames/s, inputs resident in HBM, CUDA-event timed, max over ranks.
  e2e   : frames/s through s2p_b200.synthesis.BulkSynthesizer with pinned HOST
          buffers: H2D of every row and D2H of every result inside the timed region.
  roofline : the tcgen05 implicit-GEMM conv (dominant kernel class), measured live with
          CUDA events around every conv launch of one instrumented step.
  cpu_baseline : the fp32 oracle on the host cores (bounded sample), rank 0, N=1 only.
`--impl reference` times the oracle (the reference's generator source is absent, see
DESIGN.md) on the host cores with all threads.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# rank 0 prints exactly ONE JSON line on stdout: NCCL's own lines (banner at WARN, rank/ring lines at INFO) go to
# stderr.  NCCL_DEBUG itself is only DEFAULTED: a driver that asks for INFO keeps its rank evidence.
os.environ.setdefault("NCCL_DEBUG", "WARN")
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
# ... and because a library can still write to file descriptor 1 behind Python's back (NCCL's version banner does when
# the launcher exported its own NCCL_DEBUG settings), the real stdout is kept aside for the JSON line and descriptor 1
# is pointed at stderr for everything else
_JSON_OUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(line: dict):
    _JSON_OUT.write(json.dumps(line) + "\n")
    _JSON_OUT.flush()


# diagnostic switches of the library / tools change what a step does or how fast it runs: a bench line measured with
# one of them set is not the product's number
_DIAGNOSTIC_ENV = ("S2P_CONV_CLUSTER", "S2P_CONV_DEBUG", "S2P_NORM_PIX_PER_THREAD", "S2P_WGRAD_ROWS", "S2P_WGRAD_TAP_SPLIT",
                   "S2P_WGRAD_PAD16", "S2P_WGRAD_PAIRS")


def refuse_diagnostic_env():
    bad = [k for k in _DIAGNOSTIC_ENV if os.environ.get(k) not in (None, "")]
    if bad:
        raise SystemExit(f"bench.py refuses to run with diagnostic switches set: {bad} (unset them)")

METRIC = "generated_frames_per_sec"
UNIT = "frames/s"
ENV = "cheetah"
GMAC_PER_FRAME_IGEMM = None  # filled from the layer list at run time


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--batch", type=int, default=1024)
    p.add_argument("--impl", type=str, default="b200", choices=["b200", "reference"])
    p.add_argument("--workload", type=str, default="synthesis", choices=["synthesis", "train"],
                   help="synthesis = BASELINE configs[2]/[3] (the metric's workload, default); "
                        "train = configs[1]/[4], one G + D step per step (use --batch 16 / 256)")
    p.add_argument("--cpu-sample", type=int, default=16, help="frames per CPU-baseline forward")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-train", action="store_true", help="skip the configs[1] training-step side measurement")
    p.add_argument("--precision", type=str, default="high", choices=["high", "standard"],
                   help="generator precision schedule (s2p_b200/options.py); 'standard' is the round-1 single-bf16 "
                        "schedule that misses the 50 dB tolerance, for A/B timing only")
    return p.parse_args()


def workload_config(args, world):
    return {"workload": "bulk synthesis (BASELINE configs[3]): uint8 HWC frame + fp32 state -> uint8 HWC next frame, "
                        f"{ENV}, {args.batch} rows per rank per step; per-frame forward identical to configs[2]",
            "env_type": ENV, "rows_per_rank_per_step": args.batch, "global_rows_per_step": args.batch * world,
            "image": "100x100x3 uint8", "state_dim": 17, "parallelism": f"shard{world}",
            "weights": "random init (ORACLE_SPEC v1 'scaled' init, seed 1234)",
            "rows_per_launch": "512 (options.max_microbatch: rows are independent, outputs bit-identical for every value)",
            "l2_policy": "inputs+activations per step (>=3 GB) exceed the 126 MB L2; no explicit flush"}


# ---------------------------------------------------------------------------- CPU legs
def cpu_oracle_fps(sample, repeats, threads=None):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch
    import s2p_oracle as O
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    torch.manual_seed(1234)
    net = O.define_G(O.make_options(env_type=ENV, init_type="scaled")).eval()
    g = torch.Generator().manual_seed(0)
    frames = torch.randint(0, 256, (sample, 100, 100, 3), dtype=torch.uint8, generator=g)
    states = torch.randn(sample, 17, generator=g)
    times = []
    with torch.no_grad():
        for _ in range(repeats):
            t0 = time.perf_counter()
            out = O.to_uint8_hwc(net(O.from_uint8_hwc(frames), states))
            times.append(time.perf_counter() - t0)
    assert out.shape == frames.shape
    return sample / min(times), sample / (sum(times) / len(times)), threads, times


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample = args.cpu_sample
    # warmup + steps forwards of `sample` frames each; bounded: ~0.6 s/frame/8 cores
    fps_best, fps_mean, threads, times = cpu_oracle_fps(sample, args.warmup + args.steps)
    timed = times[args.warmup:] or times
    ms = 1e3 * sum(timed) / len(timed)
    value = sample / (ms / 1e3)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, max(1, args.gpus)),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"{sample} frames per step (the oracle: reference generator source is absent)"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ---------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw,power.limit")

    def __init__(self, index, enabled=True):
        # only rank 0 samples: eight ranks each forking nvidia-smi five times a second compete with the ranks' own
        # launch threads for the host cores
        self.index, self.samples, self._stop, self._t, self.enabled = index, [], threading.Event(), None, enabled

    def _loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [t.strip() for t in out.strip().split(",")]
                if len(parts) >= 6:
                    self.samples.append(parts)
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        if self.enabled:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t is not None:
            self._t.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = sorted(float(s[0]) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        out = {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons,
               "samples": len(sm)}
        try:   # board power under load next to its limit: says whether the clock is where the power cap puts it
            pw = sorted(float(s[6]) for s in self.samples if len(s) >= 8)
            out["power_w"], out["power_limit_w"] = pw[len(pw) // 2], float(self.samples[0][7])
        except (ValueError, IndexError):
            pass
        return out


# ---------------------------------------------------------------------------- GPU arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    import s2p_b200
    from s2p_b200 import _lib, ops
    from s2p_b200.synthesis import BulkSynthesizer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch N>1 with torch.distributed.run (see module docstring)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    refuse_diagnostic_env()
    opt = s2p_b200.make_options(env_type=ENV, gpu_ids=str(local), init_type="scaled", precision=args.precision)
    torch.manual_seed(1234)
    net = s2p_b200.define_G(opt).eval()
    net.pack()
    B = args.batch
    g = torch.Generator().manual_seed(100 + rank)
    frames_h = torch.randint(0, 256, (B, 100, 100, 3), dtype=torch.uint8, generator=g).pin_memory()
    states_h = torch.randn(B, 17, generator=g).pin_memory()
    out_h = torch.empty_like(frames_h).pin_memory()
    frames_d, states_d = frames_h.to(dev), states_h.to(dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident timing
    for _ in range(args.warmup):
        net.forward_u8(frames_d, states_d)
    barrier()
    with ClockSampler(local, enabled=rank == 0) as clk:
        l0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            res = net.forward_u8(frames_d, states_d)
        e1.record()
        launches = _lib.launch_count() - l0
        # ---- instrumented steps, back to back with the timed ones (no barrier, no pause: same clocks, same power
        # state): CUDA events around every conv / norm launch give each kernel class's SHARE of a step; the roofline
        # is then algorithmic flops / (share x the timed ms_per_step)
        inst = instrumented_steps(net, frames_d, states_d, ops, torch, n_steps=3) if rank == 0 else None
        barrier()
        ms_dev = max_over_ranks(e0.elapsed_time(e1))
        # ---- end to end through the public bulk API with host buffers
        synth = BulkSynthesizer(net, batch=min(B, 256))
        for _ in range(max(1, args.warmup // 2)):
            synth.synthesize(frames_h, states_h, out_h)
        barrier()
        synth.h2d_bytes = synth.d2h_bytes = 0
        e0.record()
        for _ in range(args.steps):
            synth.synthesize(frames_h, states_h, out_h)
        e1.record()
        barrier()
        ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    assert torch.equal(out_h.to(dev), res), "e2e result differs from the device-resident result"
    clocks = clk.summary()

    roof, roof_norm, shares = None, None, None
    if rank == 0:
        roof, roof_norm, shares = rooflines(inst, ms_dev / args.steps)

    # ---- N > 1: BASELINE configs[4] beside the headline -- data-parallel GAN training, 256 frames per GPU, gradient
    # all-reduce over NVLink overlapped with backward (every rank takes part; rank 0 reports)
    train_dp = None
    if world > 1 and not args.no_train:
        try:
            train_dp = train_dp_side_measurement(torch, dist, s2p_b200, dev, world, rank)
        except Exception as e:  # noqa: BLE001
            train_dp = {"error": repr(e)[:200]}
    if world > 1:
        dist.barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    frames_total = B * world * args.steps
    value = frames_total / (ms_dev / 1e3)
    e2e = frames_total / (ms_e2e / 1e3)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload_config(args, world),
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": synth.h2d_bytes // args.steps,
                    "d2h_bytes_per_step": synth.d2h_bytes // args.steps, "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "roofline_norm": roof_norm,
            "kernel_time_share": shares}
    # ---- parity of THIS run's output: 64 of the rows just synthesised against the fp32 oracle (the checker; on the
    # GPU, TF32 off), as PSNR on [0,1]-mapped frames.  north_star: >= 50 dB.
    try:
        line["parity_check"] = parity_spot_check(torch, net, frames_d, states_d, res, dev)
    except Exception as e:  # noqa: BLE001
        line["parity_check"] = {"error": repr(e)[:200]}
    if train_dp is not None:
        line["training_step"] = train_dp
    if world == 1 and not args.no_train:
        try:
            line["latency_b16"] = latency_b16_side_measurement(torch, s2p_b200, dev)
        except Exception as e:  # noqa: BLE001
            line["latency_b16"] = {"error": repr(e)[:200]}
        try:
            line["eager_gpu_baseline"] = eager_gpu_baseline(torch, dev)
        except Exception as e:  # noqa: BLE001
            line["eager_gpu_baseline"] = {"error": repr(e)[:200]}
    if world == 1 and not args.no_train:
        try:
            line["training_step"] = train_side_measurement(torch, s2p_b200, dev)
        except Exception as e:  # noqa: BLE001  (a side measurement must never cost the headline line)
            line["training_step"] = {"error": repr(e)[:200]}
    if world == 1 and not args.no_train:
        try:
            line["rollout"] = rollout_side_measurement(torch, net, dev)
        except Exception as e:  # noqa: BLE001
            line["rollout"] = {"error": repr(e)[:200]}
    if world == 1 and not args.no_train:
        try:
            line["adjacent_components"] = adjacent_side_measurements(torch, dev)
        except Exception as e:  # noqa: BLE001
            line["adjacent_components"] = {"error": repr(e)[:200]}
    if world == 1 and not args.no_cpu_baseline:
        cpu_n = 4 * args.cpu_sample            # 64 frames x 2 forwards: ~12 s of CPU work on 16 cores
        best, mean, threads, times = cpu_oracle_fps(cpu_n, 2)
        line["cpu_baseline"] = {"value": best, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"{cpu_n} frames x 2 forwards of the fp32 oracle (best of 2), "
                                          "same uint8->uint8 row workload"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def _oracle_on(torch, dev, env=ENV):
    """the fp32 oracle (test infrastructure, used here only as the checker / the eager baseline) with the weights of
    the benchmarked generator: same seeded 'scaled' init"""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import s2p_oracle as O
    torch.manual_seed(1234)
    return O, O.define_G(O.make_options(env_type=env, init_type="scaled")).eval().to(dev)


def parity_spot_check(torch, net, frames_d, states_d, res, dev, rows=64):
    O, ora = _oracle_on(torch, dev)
    # bench.py builds the generator from its own seeded init: load the oracle's weights' twin check instead of
    # trusting two inits to agree -- compare the oracle run ON THE GENERATOR'S OWN WEIGHTS
    ora.load_state_dict(net.state_dict())
    tf32 = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        with torch.no_grad():
            ref = ora(O.from_uint8_hwc(frames_d[:rows]), states_d[:rows])
        got = net(O.from_uint8_hwc(frames_d[:rows]), states_d[:rows])
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32
    mse = ((got - ref).double() * 0.5).pow(2).mean().item()
    import math
    psnr = 10.0 * math.log10(1.0 / max(mse, 1e-30))
    u8_ref = O.to_uint8_hwc(ref)
    d = (res[:rows].int() - u8_ref.int()).abs()
    return {"rows": rows, "psnr_db": psnr, "max_abs": float((got - ref).abs().max()), "tolerance_db": 50.0,
            "ok": bool(psnr >= 50.0), "u8_max_lsb": int(d.max()), "u8_mean_lsb": float(d.float().mean()),
            "against": "fp32 oracle on the same weights and rows (cuDNN, TF32 off), fp32 [-1,1] outputs; the u8 line "
                       "compares the benchmarked uint8 result rows with the oracle's quantised output"}


def latency_b16_side_measurement(torch, s2p_b200, dev, batch=16, reps=20):
    """BASELINE configs[0]/[1] size: one-step forward at batch 16 (README.md:59's batch size), cheetah, fp32 NCHW in /
    out through netG(image, state).  ms_graph = the default path (CUDA-graph replay of the launch sequence, state path
    = one fused launch); ms_eager = the same kernels launched one by one from Python."""
    from s2p_b200 import _lib
    torch.manual_seed(1234)
    nets = {}
    for name, graph in (("graph", 1), ("eager", 0)):
        n = s2p_b200.define_G(s2p_b200.make_options(env_type=ENV, gpu_ids=str(dev.index), init_type="scaled", cuda_graph=graph)).eval()
        if nets:
            n.load_state_dict(nets["graph"].state_dict())
        nets[name] = n
    g = torch.Generator().manual_seed(16)
    image = (torch.rand(batch, 3, 100, 100, generator=g) * 2 - 1).to(dev)
    state = torch.randn(batch, 17, generator=g).to(dev)
    out = {"config": f"one-step forward, cheetah, batch {batch}, fp32 NCHW in/out, 1 GPU (BASELINE configs[0] on the GPU)"}
    results = {}
    for name, n in nets.items():
        for _ in range(3):
            results[name] = n(image, state)
        torch.cuda.synchronize()
        l0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t_host = time.perf_counter()
        e0.record()
        for _ in range(reps):
            n(image, state)
        e1.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t_host) / reps * 1e3
        out[f"ms_{name}"] = e0.elapsed_time(e1) / reps
        out[f"ms_{name}_wall"] = wall
        out[f"launches_{name}"] = (_lib.launch_count() - l0) // reps
    out["bit_identical"] = bool(torch.equal(results["graph"], results["eager"]))
    out["frames_per_sec_graph"] = batch / out["ms_graph"] * 1e3
    return out


def eager_gpu_baseline(torch, dev, batch=128, reps=3):
    """the practical bar of BASELINE.md section 3: the oracle module (plain torch.nn -> cuDNN) on the same GPU under
    bf16 autocast, 128-frame batches"""
    O, ora = _oracle_on(torch, dev)
    g = torch.Generator().manual_seed(5)
    image = (torch.rand(batch, 3, 100, 100, generator=g) * 2 - 1).to(dev)
    state = torch.randn(batch, 17, generator=g).to(dev)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        for _ in range(2):
            ora(image, state)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            ora(image, state)
        e1.record()
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    return {"what": "oracle torch.nn module on cuda under bf16 autocast (cuDNN / cuBLAS library kernels)", "batch": batch,
            "reps": reps, "ms_per_batch": ms, "frames_per_sec": batch / ms * 1e3}


def train_dp_side_measurement(torch, dist, s2p_b200, dev, world, rank, batch=256, steps=3, warmup=2):
    """BASELINE configs[4]: data-parallel G + D training step, walker, 256 frames per GPU, NCCL gradient all-reduce
    (GradBuckets: issued from the backward hooks on a side stream).  Max over ranks, CUDA events."""
    from s2p_b200.training import S2PTrainer
    torch.manual_seed(1234)                      # same initial weights on every rank
    opt = s2p_b200.make_options(env_type="walker", gpu_ids=str(dev.index), init_type="scaled")
    tr = S2PTrainer(s2p_b200.define_G(opt), s2p_b200.define_D(opt), world_size=world, vgg=s2p_b200.VGG19FeaturesB200().to(dev))
    g = torch.Generator().manual_seed(300 + rank)
    prev = (torch.rand(batch, 3, 100, 100, generator=g) * 2 - 1).to(dev)
    real = (torch.rand(batch, 3, 100, 100, generator=g) * 2 - 1).to(dev)
    state = torch.randn(batch, opt.state_dim, generator=g).to(dev)
    for _ in range(warmup):
        tr.train_step(prev, state, real)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        gl, dl = tr.train_step(prev, state, real)
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    # same weights on every rank after the steps <=> the all-reduce really averaged the gradients
    w = torch.cat([p.detach().flatten()[:64].double() for p in tr.netG.parameters()]).sum().reshape(1)
    lo, hi = w.clone(), w.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    flops_frame = 3 * 112.9e9 * 1.03 + 3 * 7.006e9
    del tr
    torch.cuda.empty_cache()
    return {"config": f"BASELINE configs[4]: data-parallel generator + multiscale-discriminator training step, walker, batch {batch} per GPU, "
                      f"{world} GPUs, NCCL gradient all-reduce overlapped with backward",
            "n_gpus": world, "batch_per_gpu": batch, "ms_per_step": ms, "frames_per_sec": batch * world / ms * 1e3, "steps": steps,
            "warmup": warmup, "tflops_per_gpu": flops_frame * batch / (ms / 1e3) / 1e12, "weights_in_sync": bool((hi - lo).abs().item() == 0.0),
            "losses": {k: float(v) for k, v in gl.items()} | {"D": float(dl)}}


def adjacent_side_measurements(torch, dev):
    """SURVEY.md section 8 f rows either side of the generator, timed with CUDA events beside the headline:
    state-transition producer (1M rows, one launch), SLAC window gather (HBM roofline) and SLAC encoder straight
    from the resident uint8 dataset."""
    import numpy as np
    from s2p_b200.networks.slac_encoder import SlacEncoderB200
    from s2p_b200.slac_buffer import SlacReplayBufferB200
    from s2p_b200.transition import EnsembleTransitionB200
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)

    def timed(fn, reps=5, warm=2):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    out = {}
    T = 1_000_000
    torch.manual_seed(0)
    tm = EnsembleTransitionB200(17, 6, 256, 3).to(dev)
    norm = dict(obs_mean=np.zeros(17, np.float32), obs_std=np.ones(17, np.float32), next_obs_mean=np.zeros(17, np.float32),
                next_obs_std=np.ones(17, np.float32), reward_mean=0.0, reward_std=1.0)
    obs, act = torch.randn(T, 17, device=dev), torch.rand(T, 6, device=dev) * 2 - 1
    idx = torch.randint(0, 7, (T,), device=dev, dtype=torch.int32)
    ms = timed(lambda: tm.rollout_rows(obs, act, idx, norm))
    flops = 2.0 * T * 7 * (32 * 256 + 2 * 256 * 256 + 256 * 64)
    out["state_transition_producer"] = {
        "kernel": "ensemble_transition_kernel (7 x MLP 23-256-256-256-36 + selection + uncertainties, fp32 CUDA cores)",
        "rows": T, "ms": ms, "rows_per_sec": T / ms * 1e3, "fp32_tflops": flops / ms / 1e9, "launches_per_call": 1}
    del obs, act, idx
    # multi-step generated dataset (random_state_5step_random_action): 5 chained ensemble steps + 5-step generator
    # rollout + row assembly, host arrays in / out (the whole producer call)
    try:
        import s2p_b200
        from s2p_b200.synthesis import synthesize_multistep_dataset
        netg = s2p_b200.define_G(s2p_b200.make_options(env_type=ENV, gpu_ids=str(dev.index), init_type="scaled")).eval()
        rs = np.random.RandomState(3)
        n_real, J = 4000, 1024
        real = {"observations": rs.randn(n_real, 17).astype(np.float32), "actions": rs.uniform(-1, 1, (n_real, 6)).astype(np.float32),
                "rewards": rs.randn(n_real).astype(np.float32), "timeouts": np.zeros(n_real, np.float32),
                "image_observations": rs.randint(0, 256, (n_real, 100, 100, 3)).astype(np.uint8)}
        real["timeouts"][249::250] = 1
        lo_a, hi_a = -np.ones(6, np.float32), np.ones(6, np.float32)
        synthesize_multistep_dataset(netg, tm, real, norm, lo_a, hi_a, n_starts=64, rng=np.random.RandomState(0))   # warm-up
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ms_d = synthesize_multistep_dataset(netg, tm, real, norm, lo_a, hi_a, n_starts=J, rng=np.random.RandomState(1))
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        out["multistep_dataset_producer"] = {
            "what": "synthesize_multistep_dataset: random_state_5step_random_action (5 chained ensemble steps + 5-step generator "
                    "rollout + row assembly), host arrays in and out, wall clock",
            "starts": J, "steps": 5, "rows_out": int(ms_d["actions"].shape[0]), "generated_frames": J * 5, "seconds": dt,
            "generated_frames_per_sec": J * 5 / dt, "rows_per_sec": int(ms_d["actions"].shape[0]) / dt}
        del netg, real, ms_d
    except Exception as e:  # noqa: BLE001
        out["multistep_dataset_producer"] = {"error": repr(e)[:200]}
    N, Bw = 40_000, 1024
    buf = SlacReplayBufferB200(N, 8, (3, 100, 100), (6,), dev)
    buf.frames = torch.randint(0, 256, (N, 100, 100, 3), dtype=torch.uint8, device=dev)
    buf.frames_tp1 = torch.randint(0, 256, (N, 100, 100, 3), dtype=torch.uint8, device=dev)
    rows = torch.arange(8, N, device=dev)
    buf.codes = torch.cat([2 * (rows.unsqueeze(1) - 8 + torch.arange(8, device=dev)), (2 * (rows - 1) + 1).unsqueeze(1)], 1)
    buf._n = buf.codes.shape[0]
    pick = np.random.RandomState(0).randint(0, buf._n, size=Bw)
    ms = timed(lambda: buf.gather_states(pick), reps=10)
    byts = Bw * 9 * 30000 * 5
    out["slac_window_gather"] = {
        "kernel": "window_gather_quad_kernel (gather + HWC->CHW + uint8->float/255; whole sample call incl. index upload)",
        "batch": Bw, "frames": Bw * 9, "ms": ms, "bound": "hbm", "achieved": byts / ms / 1e6, "peak": hbm_peak, "unit": "GB/s",
        "frac": byts / ms / 1e6 / hbm_peak, "bytes_per_frame": 150000}
    enc = SlacEncoderB200(3, 256, 100).to(dev)
    pick = pick[:256]
    ms = timed(lambda: enc.forward_from_buffer(buf, pick), reps=10)
    out["slac_encoder"] = {"kernel": "slac_conv1_kernel (uint8 gather fused) + 5 x conv_igemm_kernel", "batch": 256,
                           "frames": 256 * 9, "ms": ms, "frames_per_sec": 256 * 9 / ms * 1e3}
    del buf, enc
    # RAD augmentations through the reference-named calls (draws on the host + upload + one launch)
    from s2p_b200 import data_augs as rad
    np.random.seed(0)
    stacks = torch.randint(0, 256, (Bw, 9, 100, 100), dtype=torch.uint8, device=dev)
    ms_crop = timed(lambda: rad.random_crop(stacks, out=84), reps=10)
    stacks_f = torch.rand(Bw, 9, 100, 100, device=dev)
    ms_gray = timed(lambda: rad.random_grayscale(stacks_f, p=.3), reps=10)
    b_crop, b_gray = Bw * 9 * (84 * 84 * 2), Bw * 9 * 10000 * 8
    out["rad_augmentations"] = {
        "kernel": "augment_u8_staged_kernel<crop> (uint8, rows bulk-copied to shared memory) / grayscale_kernel<4> (float32); whole call incl. host draws + upload",
        "stacks": Bw, "shape": "9x100x100", "bound": "hbm", "peak": hbm_peak, "unit": "GB/s",
        "random_crop_u8": {"ms": ms_crop, "achieved": b_crop / ms_crop / 1e6, "frac": b_crop / ms_crop / 1e6 / hbm_peak},
        "random_grayscale_f32": {"ms": ms_gray, "achieved": b_gray / ms_gray / 1e6, "frac": b_gray / ms_gray / 1e6 / hbm_peak}}
    return out


def rollout_side_measurement(torch, net, dev, batch=1024, seq_len=5, steps=3, warmup=1):
    """BASELINE configs[2]: N-step autoregressive rollout, seq_len 5, 1024 start frames, through the public
    `rollout(image, states)` call (fp32 NCHW frames fed back unquantised, all 5 frames of every row returned)."""
    g = torch.Generator().manual_seed(11)
    image = (torch.rand(batch, 3, 100, 100, generator=g) * 2 - 1).to(dev)
    states = torch.randn(batch, seq_len, net.opt.state_dim, generator=g).to(dev)
    for _ in range(warmup):
        out = net.rollout(image, states)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        out = net.rollout(image, states)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return {"config": f"BASELINE configs[2]: {seq_len}-step autoregressive rollout of {batch} start frames, cheetah, 1 GPU",
            "ms_per_rollout": ms, "frames_per_sec": batch * seq_len / ms * 1e3, "steps": steps, "warmup": warmup,
            "output": list(out.shape), "finite": bool(torch.isfinite(out).all())}


def train_side_measurement(torch, s2p_b200, dev, batch=16, steps=5, warmup=3):
    """BASELINE configs[1] (G + D training step, walker, batch 16, 1 GPU) timed beside the headline
    generation workload: CUDA-graph replay of the whole step, CUDA events."""
    from s2p_b200.training import S2PTrainer
    torch.manual_seed(1234)
    opt = s2p_b200.make_options(env_type="walker", gpu_ids=str(dev.index), init_type="scaled")
    tr = S2PTrainer(s2p_b200.define_G(opt), s2p_b200.define_D(opt), vgg=s2p_b200.VGG19FeaturesB200().to(dev))
    g = torch.Generator().manual_seed(7)
    prev = (torch.rand(batch, 3, 100, 100, generator=g) * 2 - 1).to(dev)
    real = (torch.rand(batch, 3, 100, 100, generator=g) * 2 - 1).to(dev)
    state = torch.randn(batch, opt.state_dim, generator=g).to(dev)
    tr.capture(prev, state, real)
    for _ in range(warmup):
        tr.train_step_graphed(prev, state, real)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        gl, dl = tr.train_step_graphed(prev, state, real)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return {"config": "BASELINE configs[1]: generator + multiscale-discriminator training step, walker, batch 16, 1 GPU "
                      "(L1 + hinge GAN + feature matching + VGG19 perceptual with a random-init extractor, Adam; CUDA-graph replay)",
            "ms_per_step": ms, "frames_per_sec": batch / ms * 1e3, "steps": steps, "warmup": warmup,
            "losses": {k: float(v) for k, v in gl.items()} | {"D": float(dl)}}


def instrumented_steps(net, frames_d, states_d, ops, torch, n_steps=3):
    """n_steps extra steps with a CUDA-event pair around every conv / norm launch, enqueued right behind the timed
    steps.  Returns the raw records; rooflines() turns them into the JSON objects once the events have completed."""
    rec = {"conv": [], "norm": [], "other": []}
    orig = {n: getattr(ops, n) for n in ("conv_igemm", "conv_mat_fused", "mat_norm", "stem_conv", "torgb_tanh", "torgb_igemm",
                                         "linear", "pos_encode", "in_stats_finalize", "mat_prepare", "state_encoder_fused")}

    def wrap(name, kind, work_fn):
        fn = orig[name]

        def inner(*a, **k):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            r = fn(*a, **k)
            e.record()
            rec[kind].append((s, e, work_fn(*a, **k) if work_fn else 0.0))
            return r
        return inner

    def conv_flops(x, w_packed, bias, cout, ksize, stride=1, *a, **k):
        # ALGORITHMIC flops: the convolution the oracle defines.  A split-precision launch walks 3x the channels of
        # the conv it computes (three products of one weight x activation pair): counted once.
        b, h, w, cin = x.shape
        if k.get("split"):
            cin //= 2
        ho = (h + 2 * (ksize // 2) - ksize) // stride + 1
        wo = (w + 2 * (ksize // 2) - ksize) // stride + 1
        cin2 = k["x2"].shape[3] if k.get("x2") is not None else 0     # the 1x1 shortcut riding along as a second operand
        return 2.0 * b * ho * wo * cout * (cin * ksize * ksize + cin2)

    def norm_bytes(x, mean_rstd, gs=None, bs=None, gb_spatial=None, up=1, act=2, out=None, x_lo=None, out_lo=None):
        # bytes the launch has to move: a wide (hi + lo) tensor is 4 B per element
        b, hx, wx, c = x.shape
        px = b * hx * up * wx * up
        rd = (4.0 if x_lo is not None else 2.0) * c * (b * hx * wx)
        wr = (4.0 if out_lo is not None else 2.0) * c * px
        return rd + wr + (4.0 * c * px if gb_spatial is not None else 0.0)

    def fused_flops(h, w_packed_gb, x, mod_table, up=1, act=2, ksize=3, out=None, x_lo=None):
        b, hh, ww, cin = h.shape
        return 2.0 * b * hh * ww * (2 * x.shape[3]) * cin * ksize * ksize

    ops.conv_igemm = wrap("conv_igemm", "conv", conv_flops)
    ops.conv_mat_fused = wrap("conv_mat_fused", "conv", fused_flops)
    ops.mat_norm = wrap("mat_norm", "norm", norm_bytes)
    for n in ("stem_conv", "torgb_tanh", "torgb_igemm", "linear", "pos_encode", "in_stats_finalize", "mat_prepare",
              "state_encoder_fused"):
        setattr(ops, n, wrap(n, "other", None))
    try:
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(n_steps):
            net.forward_u8(frames_d, states_d)
        t1.record()
    finally:
        for n, fn in orig.items():
            setattr(ops, n, fn)
    return {"rec": rec, "t0": t0, "t1": t1, "n_steps": n_steps,
            "rows_per_launch": min(int(net.max_microbatch), int(frames_d.shape[0]))}


def rooflines(inst, ms_per_step_timed):
    """roofline objects from the instrumented steps: share of each kernel class inside a step (events), applied to the
    TIMED ms_per_step.  achieved = algorithmic work per step / (share x ms_per_step)."""
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    tf_peak = peaks.get("bf16_tflops_sustained", 1400.0)
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    src = "measured (MEASURED_PEAKS.json, sustained)" if peaks else "fallback (B200_PROFILING.md)"
    rec, n_steps = inst["rec"], inst["n_steps"]
    total = inst["t0"].elapsed_time(inst["t1"])                       # all instrumented steps
    tms = {k: sum(s.elapsed_time(e) for s, e, _ in v) for k, v in rec.items()}
    work = {k: sum(w for _, _, w in v) / n_steps for k, v in rec.items()}     # per step
    nconv, nnorm = len(rec["conv"]) // n_steps, len(rec["norm"]) // n_steps
    shares = {k: tms[k] / total for k in tms}
    traffic_conv = traffic_norm = None
    traffic_file = None
    for cand in ("r02_kernel_metrics.csv", "r01_kernel_metrics.csv"):
        if os.path.exists(os.path.join(ROOT, "profiles", cand)):
            traffic_file = cand
            break
    try:   # DRAM bytes per launch from the committed ncu pass of the same forward (128-frame micro-batch)
        import csv
        with open(os.path.join(ROOT, "profiles", traffic_file)) as f:
            rows = list(csv.DictReader(l for l in f if not l.startswith("==")))
        by = {}
        for r in rows:
            if r["Metric Name"] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                by.setdefault((r["ID"], r["Kernel Name"]), 0.0)
                by[(r["ID"], r["Kernel Name"])] += float(r["Metric Value"].replace(",", "")) * \
                    {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}[r["Metric Unit"]]
        c = [v for (i, k), v in by.items() if "conv_igemm" in k]
        n = [v for (i, k), v in by.items() if "mat_norm" in k]
        # the ncu pass profiles ONE forward of 128 frames; a launch of the benchmarked step covers rows_per_launch rows,
        # and these kernels' DRAM traffic is activations: scaled linearly
        scale = inst.get("rows_per_launch", 128) / 128.0
        traffic_conv = sum(c) / len(c) * scale if c else None
        traffic_norm = sum(n) / len(n) * scale if n else None
    except Exception:
        pass
    conv_ms = shares["conv"] * ms_per_step_timed
    norm_ms = shares["norm"] * ms_per_step_timed
    ach_tf = work["conv"] / (conv_ms / 1e3) / 1e12
    ach_gb = work["norm"] / (norm_ms / 1e3) / 1e9
    roof = {"kernel": "conv_igemm_kernel (tcgen05 implicit GEMM, plain and MAT-fused epilogues; all launches of one step)",
            "bound": "tensor", "achieved": ach_tf, "peak": tf_peak, "unit": "TFLOP/s", "frac": ach_tf / tf_peak,
            "traffic": traffic_conv,
            "traffic_note": f"avg dram__bytes_read+write per conv launch, profiles/{traffic_file} (ncu of a 128-frame forward) "
                            f"x {inst.get('rows_per_launch', 128)}/128 rows per launch of this step",
            "peak_source": src, "launches": nconv, "avg_launch_ms": conv_ms / nconv,
            "flops_per_launch_avg": work["conv"] / nconv, "flops_per_step": work["conv"],
            "how": f"algorithmic flops of a step / (conv share {shares['conv']:.4f} x timed ms_per_step {ms_per_step_timed:.3f}); "
                   f"share = CUDA-event time of the conv launches / total, over {n_steps} instrumented steps enqueued "
                   "directly behind the timed steps"}
    roof_norm = {"kernel": "mat_norm_kernel (fused IN + LeakyReLU of the image encoder; all launches of one step)",
                 "bound": "hbm", "achieved": ach_gb, "peak": hbm_peak, "unit": "GB/s", "frac": ach_gb / hbm_peak,
                 "traffic": traffic_norm, "peak_source": src, "launches": nnorm, "avg_launch_ms": norm_ms / nnorm,
                 "bytes_per_launch_avg": work["norm"] / nnorm}
    shares["step_ms_instrumented"] = total / n_steps
    return roof, roof_norm, shares


def run_train(args):
    """BASELINE configs[1] (walker, batch 16, 1 GPU) / configs[4] (batch 256 per GPU, NCCL gradient
    all-reduce): one step = one generator step + one discriminator step incl. Adam."""
    import torch
    import torch.distributed as dist
    import s2p_b200
    from s2p_b200 import _lib
    from s2p_b200.training import S2PTrainer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.batch
    env = "walker"
    torch.manual_seed(1234)                      # same initial weights on every rank
    opt = s2p_b200.make_options(env_type=env, gpu_ids=str(local), init_type="scaled")
    tr = S2PTrainer(s2p_b200.define_G(opt), s2p_b200.define_D(opt), world_size=world,
                    vgg=s2p_b200.VGG19FeaturesB200().to(dev))
    g = torch.Generator().manual_seed(100 + rank)
    host = [(torch.rand(B, 3, 100, 100, generator=g) * 2 - 1).pin_memory(),
            torch.randn(B, opt.state_dim, generator=g).pin_memory(),
            (torch.rand(B, 3, 100, 100, generator=g) * 2 - 1).pin_memory()]
    devt = [t.to(dev) for t in host]
    use_graph = world == 1 and B <= 32           # launch-bound regime: replay a CUDA graph of the step
    if use_graph:
        tr.capture(*devt)
        step = lambda: tr.train_step_graphed(*devt)   # noqa: E731
    else:
        step = lambda: tr.train_step(*devt)           # noqa: E731

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(args.warmup):
        step()
    barrier()
    with ClockSampler(local, enabled=rank == 0) as clk:
        l0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            step()
        e1.record()
        barrier()
        launches = _lib.launch_count() - l0
        ms_dev = max_over_ranks(e0.elapsed_time(e1))
        # end to end: batch from pinned host memory in, loss scalars back out, every step
        e0.record()
        for _ in range(args.steps):
            for d, h in zip(devt, host):
                d.copy_(h, non_blocking=True)
            gl, dl = step()
            losses_host = [float(v.detach()) for v in gl.values()] + [float(dl.detach())]
        e1.record()
        barrier()
        ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    clocks = clk.summary()
    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        tf_peak = peaks.get("bf16_tflops_sustained", 1400.0)
        frames = B * world * args.steps
        # algorithmic flops per frame: generator forward 112.9 GFLOP, its
        # backward = dgrad + wgrad = 2x; discriminator (fake+real forward, backward) ~ 3 % on top
        # G fwd+dgrad+wgrad (x1.03 for D) + VGG19 to relu5_1 (3.503 GMAC per image): fwd(fake) + fwd(real) + dgrad(fake)
        flops_frame = 3 * 112.9e9 * 1.03 + 3 * 7.006e9
        ach = flops_frame * B / (ms_dev / args.steps / 1e3) / 1e12
        h2d = sum(t.numel() * t.element_size() for t in host)
        line = {"metric": "training_frames_per_sec", "value": frames / (ms_dev / 1e3), "unit": "frames/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": "GAN training step (BASELINE configs[1]/[4]): generator step (L1 + hinge GAN + "
                                       f"feature matching + VGG19 perceptual) + discriminator step + Adam, {env}, batch {B} per GPU",
                           "env_type": env, "batch_per_gpu": B, "global_batch": B * world, "parallelism": f"dp{world}",
                           "cuda_graph": use_graph, "weights": "random init ('scaled'); VGG19 extractor seeded random init (ORACLE_SPEC v1.2)"},
                "e2e": {"value": frames / (ms_e2e / 1e3), "unit": "frames/s", "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": 4 * len(losses_host), "ms_per_step": ms_e2e / args.steps},
                "gpu_launches": int(launches) if not use_graph else None,
                "gpu_launches_note": "CUDA-graph replay: the library's launches are inside the graph" if use_graph else None,
                "clocks": clocks,
                "roofline": {"kernel": "whole training step (conv fwd/dgrad/wgrad dominate)", "bound": "tensor",
                             "achieved": ach, "peak": tf_peak, "unit": "TFLOP/s", "frac": ach / tf_peak, "traffic": None}}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.workload == "train":
        if a.batch == 1024:
            a.batch = 16
        run_train(a)
    else:
        run_b200(a)

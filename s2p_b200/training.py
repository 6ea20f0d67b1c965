"""GAN training step of train.py --netG=s2p (README.md:59): one generator step (L1 + hinge GAN +
feature matching + VGG19 perceptual when a `vgg` extractor is given; rebuttal.md:71,131-135) then one discriminator step, SPADE Pix2PixTrainer order.
Mirrors oracle/s2p_oracle.py `training_step` (ORACLE_SPEC v1.1) term by term.

Multi-GPU (BASELINE configs[4]): one process per GPU, data parallel.  Gradients live in flat fp32 buckets
(`GradBuckets`: every p.grad is a view into its bucket); a bucket's NCCL all-reduce is issued on a side stream the
moment the backward pass has produced its last gradient, so communication overlaps the rest of the backward and
nothing is staged through torch.cat copies.  Parameters are updated by the repo's fused Adam kernel (fp32 master
weights, torch.optim.Adam arithmetic)."""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import autograd as A
from . import ops

LAMBDA_FEAT = 10.0
LAMBDA_L1 = 10.0


class FusedAdam:
    """Adam(lr, betas, eps) with torch.optim.Adam's update rule, one fused launch per 64 tensors.  The
    step count lives in device memory so that a captured CUDA graph of the step stays valid."""

    def __init__(self, params, lr=2e-4, betas=(0.0, 0.9), eps=1e-8):
        self.params = [p for p in params if p.requires_grad]
        self.lr, self.betas, self.eps = lr, betas, eps
        dev = self.params[0].device
        self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        self.m = [torch.zeros_like(p, memory_format=torch.contiguous_format) for p in self.params]
        self.v = [torch.zeros_like(p, memory_format=torch.contiguous_format) for p in self.params]

    @property
    def step_count(self):
        return int(self.step_dev.item())

    def zero_grad(self, set_to_none=True):
        if getattr(self, "buckets", None) is not None:
            self.buckets.zero()           # gradients are views into flat buckets: clear those, keep the views
            return
        for p in self.params:
            p.grad = None

    def state_dict(self):
        return {"step": self.step_dev.clone(), "exp_avg": [t.clone() for t in self.m], "exp_avg_sq": [t.clone() for t in self.v]}

    def load_state_dict(self, sd):
        self.step_dev.copy_(sd["step"])
        for dst, src in zip(self.m, sd["exp_avg"]):
            dst.copy_(src)
        for dst, src in zip(self.v, sd["exp_avg_sq"]):
            dst.copy_(src)

    @torch.no_grad()
    def step(self):
        self.step_dev += 1
        live = [(p, p.grad if p.grad.is_contiguous() else p.grad.contiguous(), m, v)
                for p, m, v in zip(self.params, self.m, self.v) if p.grad is not None]
        if live:
            ps, gs, ms, vs = zip(*live)
            ops.adam_step_multi(ps, gs, ms, vs, self.lr, self.betas[0], self.betas[1], self.eps, self.step_dev)
            ops.bump_param_epoch()      # packed-weight caches keyed on the parameters are stale now


def allreduce_gradients(params, world_size, bucket_elems=32 * 1024 * 1024):
    """average .grad over ranks after the fact (flat fp32 buckets through staging copies, one all-reduce each).
    The trainer uses GradBuckets (overlapped, in place); this is the simple form for callers with ordinary .grad
    tensors."""
    import torch.distributed as dist
    grads = [p.grad for p in params if p.grad is not None]
    i = 0
    while i < len(grads):
        j, n = i, 0
        while j < len(grads) and (n == 0 or n + grads[j].numel() <= bucket_elems):
            n += grads[j].numel()
            j += 1
        flat = torch.cat([g.reshape(-1) for g in grads[i:j]])
        dist.all_reduce(flat)
        flat.mul_(1.0 / world_size)
        off = 0
        for g in grads[i:j]:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()
        i = j


class GradBuckets:
    """Data-parallel gradient exchange overlapped with backward (SURVEY.md section 8 e, training).

    The parameters' gradients are views into a few flat fp32 buckets, filled in reverse parameter order (roughly the
    order backward produces them).  A post-accumulate hook counts a bucket's gradients as they arrive; the last one
    issues that bucket's all-reduce on a communication stream (which first waits for the producing stream), so the
    reduce runs under the remaining backward kernels and reads / writes the gradient memory in place.  `finish()`
    joins the streams and scales by 1 / world_size before the optimiser step.  On CPU tensors (gloo tests) the same
    logic runs with async work handles instead of streams."""

    def __init__(self, params, world_size, bucket_elems=16 * 1024 * 1024):
        import torch.distributed as dist
        self.dist = dist
        self.world = world_size
        self.params = [p for p in params if p.requires_grad]
        self.cuda = self.params[0].is_cuda
        dev = self.params[0].device
        order = list(reversed(self.params))
        self.buckets, self.where = [], {}
        i = 0
        while i < len(order):
            j, n = i, 0
            while j < len(order) and (n == 0 or n + order[j].numel() <= bucket_elems):
                n += order[j].numel()
                j += 1
            flat = torch.zeros(n, dtype=torch.float32, device=dev)
            off = 0
            for p in order[i:j]:
                p.grad = flat[off:off + p.numel()].view_as(p)
                self.where[p] = len(self.buckets)
                off += p.numel()
            self.buckets.append(dict(flat=flat, size=j - i, ready=0, work=None))
            i = j
        self.comm = torch.cuda.Stream(dev) if self.cuda else None
        self.handles = [p.register_post_accumulate_grad_hook(self._hook) for p in self.params]
        self.launched = 0

    def zero(self):
        for b in self.buckets:
            b["flat"].zero_()
            b["ready"], b["work"] = 0, None
        self.launched = 0

    def _launch(self, b):
        if self.cuda:
            self.comm.wait_stream(torch.cuda.current_stream())      # the bucket's gradients are complete on the compute stream
            with torch.cuda.stream(self.comm):
                self.dist.all_reduce(b["flat"])
                b["flat"].mul_(1.0 / self.world)
        else:
            b["work"] = self.dist.all_reduce(b["flat"], async_op=True)
        self.launched += 1

    def _hook(self, p):
        b = self.buckets[self.where[p]]
        b["ready"] += 1
        if b["ready"] == b["size"]:
            self._launch(b)

    def finish(self):
        """every bucket reduced and averaged before the optimiser reads the gradients.  A bucket whose parameters did
        not all receive a gradient this step (none do in this model) is reduced here, so ranks never disagree on the
        set of collectives."""
        for b in self.buckets:
            if b["ready"] != b["size"]:
                self._launch(b)
        if self.cuda:
            torch.cuda.current_stream().wait_stream(self.comm)
        else:
            for b in self.buckets:
                b["work"].wait()
                b["flat"].mul_(1.0 / self.world)


def _split(pred):
    fake = [[t[: t.shape[0] // 2] for t in p] for p in pred]
    real = [[t[t.shape[0] // 2:] for t in p] for p in pred]
    return fake, real


def hinge_d_loss(pred_real, pred_fake):
    loss = 0.0
    for r, f in zip(pred_real, pred_fake):
        loss = loss + F.relu(1.0 - r[-1].float()).mean() + F.relu(1.0 + f[-1].float()).mean()
    return loss / len(pred_real)


def generator_losses(pred_fake, pred_real, fake, real, vgg=None):
    num_d = len(pred_fake)
    gan = sum(-f[-1].float().mean() for f in pred_fake) / num_d
    feat = 0.0
    for f, r in zip(pred_fake, pred_real):
        for a, b in zip(f[:-1], r[:-1]):
            feat = feat + A.l1_mean(a, b) * LAMBDA_FEAT / num_d
    losses = {"GAN": gan, "GAN_Feat": feat, "L1": F.l1_loss(fake, real) * LAMBDA_L1}
    if vgg is not None:
        from .networks.vgg import LAMBDA_VGG, vgg_loss
        losses["VGG"] = vgg_loss(vgg, fake, real) * LAMBDA_VGG
    return losses


class S2PTrainer:
    """run_generator_one_step / run_discriminator_one_step, as SPADE's Pix2PixTrainer."""

    def __init__(self, netG, netD, lr=2e-4, betas=(0.0, 0.9), world_size=1, vgg=None):
        self.netG, self.netD, self.vgg = netG, netD, vgg
        self.optG = FusedAdam(netG.parameters(), lr, betas)
        self.optD = FusedAdam(netD.parameters(), lr, betas)
        self.world_size = world_size
        self.fake = None
        if world_size > 1:      # gradients of G and of D live in flat buckets whose all-reduce overlaps backward
            self.optG.buckets = GradBuckets(self.optG.params, world_size)
            self.optD.buckets = GradBuckets(self.optD.params, world_size)

    def _d(self, prev, fake, real):
        return _split(self.netD(torch.cat([torch.cat([prev, fake], 1), torch.cat([prev, real], 1)], 0)))

    def run_generator_one_step(self, prev_image, state, real_image):
        self.netG.train()
        fake = self.netG(prev_image, state)
        for p in self.optD.params:            # the G step needs D's data gradients only
            p.requires_grad_(False)
        pred_fake, pred_real = self._d(prev_image, fake, real_image)
        for p in self.optD.params:
            p.requires_grad_(True)
        losses = generator_losses(pred_fake, pred_real, fake, real_image, self.vgg)
        self.optG.zero_grad()
        sum(losses.values()).backward()
        if self.world_size > 1:
            self.optG.buckets.finish()
        self.optG.step()
        self.fake = fake.detach()
        return losses

    def run_discriminator_one_step(self, prev_image, state, real_image):
        if self.fake is None:
            with torch.no_grad():
                self.netG.eval()
                self.fake = self.netG(prev_image, state)
        pred_fake, pred_real = self._d(prev_image, self.fake, real_image)
        loss = hinge_d_loss(pred_real, pred_fake)
        self.optD.zero_grad()
        loss.backward()
        if self.world_size > 1:
            self.optD.buckets.finish()
        self.optD.step()
        self.fake = None
        return loss

    def state_dict(self):
        """everything a resumed run needs (SPADE's `--continue_train`): G, D, both optimiser states"""
        return {"netG": self.netG.state_dict(), "netD": self.netD.state_dict(), "optG": self.optG.state_dict(),
                "optD": self.optD.state_dict()}

    def load_state_dict(self, sd):
        self.netG.load_state_dict(sd["netG"])
        self.netD.load_state_dict(sd["netD"])
        self.optG.load_state_dict(sd["optG"])
        self.optD.load_state_dict(sd["optD"])

    def train_step(self, prev_image, state, real_image):
        g = self.run_generator_one_step(prev_image, state, real_image)
        d = self.run_discriminator_one_step(prev_image, state, real_image)
        return g, d

    # ------------------------------------------------------------------ CUDA graph of the whole step
    def capture(self, prev_image, state, real_image, warmup=3):
        """Capture one full G + D step (forward, backward, Adam) into a CUDA graph with static
        input buffers.  At batch 16 the step is launch bound (~800 kernels of this library plus
        the tape's glue); replaying the graph removes the host from the loop."""
        if self.world_size > 1:
            raise RuntimeError("capture(): single-process only (the NCCL all-reduce is issued eagerly)")
        self._static = [t.clone() for t in (prev_image, state, real_image)]
        # the warm-up steps below are real optimiser steps: snapshot and restore, so capture() leaves weights, Adam
        # moments and the step count exactly as it found them
        snap = {"netG": {k: v.clone() for k, v in self.netG.state_dict().items()},
                "netD": {k: v.clone() for k, v in self.netD.state_dict().items()},
                "optG": self.optG.state_dict(), "optD": self.optD.state_dict()}
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self.train_step(*self._static)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        with torch.no_grad():
            self.load_state_dict(snap)
        self.fake = None
        self.optG.zero_grad()
        self.optD.zero_grad()
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph):
            g, d = self.train_step(*self._static)
            self._static_out = ({k: v.detach() for k, v in g.items()}, d.detach())
        return self

    def train_step_graphed(self, prev_image, state, real_image):
        for dst, src in zip(self._static, (prev_image, state, real_image)):
            dst.copy_(src, non_blocking=True)
        self._graph.replay()
        ops.bump_param_epoch()          # the replayed graph moved the weights without touching their version counters
        return self._static_out

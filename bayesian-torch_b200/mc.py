"""Monte-Carlo prediction over S weight samples, sharded across the GPUs of one box.

Replaces the sequential driver loop of the reference
(examples/main_bayesian_cifar_dnn2bnn.py:536-557: S forwards, torch.stack, softmax,
mean) and the off-line numpy uncertainty metrics (utils/util.py:41-60):

  * rank r of G evaluates samples [r*S/G, (r+1)*S/G) on the FULL batch; the Philox
    sample id is the global sample index, so the prediction does not depend on G;
  * each rank accumulates sum p, sum p^2 and sum_c p log p on device (bt_mc_accumulate);
  * ONE all-reduce (NCCL over NVLink) of 2*B*C + B floats joins the ranks -- the path
    has no other exchange step;
  * bt_mc_finalize turns the sums into predictive mean, variance, entropy and mutual
    information.
"""
from __future__ import annotations

from typing import NamedTuple, Optional, Tuple

import torch

from . import _lib, ops


class MCResult(NamedTuple):
    mean: torch.Tensor                  # [B, C] predictive mean (of softmax, or of raw outputs)
    var: torch.Tensor                   # [B, C] predictive variance  E[p^2] - E[p]^2
    entropy: Optional[torch.Tensor]     # [B]   H[mean p]
    mutual_info: Optional[torch.Tensor]  # [B]  H[mean p] - mean H[p]


def shard_samples(num_mc: int, rank: int, world: int) -> Tuple[int, int]:
    """[first, last) global sample indices of `rank`: contiguous, balanced to within one."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(num_mc, world)
    first = rank * base + min(rank, extra)
    return first, first + base + (1 if rank < extra else 0)


def allreduce_moments(buf: torch.Tensor, group=None) -> torch.Tensor:
    """Sum the packed accumulator across ranks (no-op without an initialised process group)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return buf


def allreduce_gradients(model, group=None, average: bool = True, bucket_mb: float = 256.0) -> int:
    """Data-parallel training step of the path: every rank ran forward/backward on its
    shard of the batch (and of the S weight samples); the gradients of the variational parameters
    (mu_*, rho_*, and whatever else the model trains) are summed across ranks in flat fp32 buckets --
    ONE NCCL all-reduce per bucket over NVLink instead of one per tensor (ResNet-18: 93 MB, ResNet-50:
    204 MB of mu/rho gradients fit one bucket) -- and divided by the world size when `average`.

    The KL term and its gradient are replica-identical: scale the KL by 1/world in the loss (or leave
    `average` on and keep kl/batch_size as the reference does, main_bayesian_cifar_dnn2bnn.py:411-415,
    with batch_size the GLOBAL batch).  Replaces the DistributedDataParallel wrapper of
    utils/utils.py:482-495.  Returns the number of collectives issued (0 without a process group)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return 0
    world = dist.get_world_size(group)
    grads = [p.grad for p in model.parameters() if p.grad is not None]
    limit = int(bucket_mb * (1 << 20))
    buckets, cur, cur_bytes = [], [], 0
    for g in grads:
        nbytes = g.numel() * 4
        if cur and (cur_bytes + nbytes > limit or g.device != cur[0].device):
            buckets.append(cur); cur, cur_bytes = [], 0
        cur.append(g); cur_bytes += nbytes
    if cur:
        buckets.append(cur)
    for b in buckets:
        flat = torch.cat([g.detach().reshape(-1).to(torch.float32) for g in b])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        if average:
            flat.div_(world)
        off = 0
        for g in b:
            n = g.numel()
            g.copy_(flat[off:off + n].view_as(g))
            off += n
    return len(buckets)


class MCAccumulator:
    """Packed fp32 device buffer [sum_p | sum_p2 | sum_plogp] fed by bt_mc_accumulate."""

    def __init__(self, batch: int, classes: int, device, softmax: bool = True):
        self.B, self.C, self.softmax = int(batch), int(classes), bool(softmax)
        self.buf = torch.zeros(2 * self.B * self.C + self.B, dtype=torch.float32, device=device)
        self.count = 0

    def _parts(self):
        bc = self.B * self.C
        return self.buf[:bc], self.buf[bc:2 * bc], self.buf[2 * bc:]

    def add(self, logits: torch.Tensor, n_samples: int = 1) -> None:
        """logits: [n_samples*B, C] (sample-major) fp32 or bf16."""
        ops._require_cuda(logits)
        if logits.dim() != 2 or logits.shape[0] != n_samples * self.B or logits.shape[1] != self.C:
            raise RuntimeError(f"expected logits [{n_samples}*{self.B}, {self.C}], got {tuple(logits.shape)}")
        lg = logits.detach().contiguous()
        p, p2, pl = self._parts()
        with torch.cuda.device(lg.device):
            rc = _lib.lib().bt_mc_accumulate(lg.data_ptr(), ops._DT[lg.dtype], n_samples, self.B, self.C,
                                             int(self.softmax), p.data_ptr(), p2.data_ptr(), pl.data_ptr(),
                                             ops._stream())
        _lib.check(rc, "bt_mc_accumulate")
        self.count += n_samples

    def finalize(self, total_samples: int, uncertainty: bool = True) -> MCResult:
        p, p2, pl = self._parts()
        dev = self.buf.device
        mean = torch.empty(self.B, self.C, dtype=torch.float32, device=dev)
        var = torch.empty_like(mean)
        ent = torch.empty(self.B, dtype=torch.float32, device=dev) if uncertainty else None
        mi = torch.empty(self.B, dtype=torch.float32, device=dev) if uncertainty else None
        with torch.cuda.device(dev):
            rc = _lib.lib().bt_mc_finalize(p.data_ptr(), p2.data_ptr(), pl.data_ptr(), int(total_samples), self.B,
                                           self.C, mean.data_ptr(), var.data_ptr(), ops._ptr(ent), ops._ptr(mi),
                                           ops._stream())
        _lib.check(rc, "bt_mc_finalize")
        return MCResult(mean, var, ent, mi)


@torch.no_grad()
def mc_predict(model, x: torch.Tensor, num_mc: int, *, softmax: bool = True, group=None, first_sample: int = 0,
               batch_samples: int = 1, uncertainty: bool = True) -> MCResult:
    """Predictive mean / variance / entropy / MI of `model` on `x` from num_mc weight samples.

    batch_samples > 1 folds that many samples into the batch axis of one model pass
    (ops.mc_batched) so late, small layers still fill the 148 SMs."""
    import torch.distributed as dist
    rank, world = 0, 1
    if dist.is_available() and dist.is_initialized():
        rank, world = dist.get_rank(group), dist.get_world_size(group)
    first, last = shard_samples(num_mc, rank, world)
    acc = None
    s = first
    while s < last:
        n = min(batch_samples, last - s)
        if n == 1:
            with ops.mc_sample(first_sample + s):
                logits = model(x)
        else:
            with ops.mc_batched(n, first_sample + s, shared_batch=x.shape[0]):
                logits = model(x)
        if isinstance(logits, tuple):   # native Bayesian models return (out, kl)
            logits = logits[0]
        logits = logits.reshape(logits.shape[0], -1)
        if acc is None:
            acc = MCAccumulator(x.shape[0], logits.shape[1], x.device, softmax)
        acc.add(logits, n)
        s += n
    if acc is None:  # more ranks than samples: contribute zeros
        with ops.mc_sample(first_sample):
            probe = model(x)
        probe = probe[0] if isinstance(probe, tuple) else probe
        acc = MCAccumulator(x.shape[0], probe.reshape(probe.shape[0], -1).shape[1], x.device, softmax)
    allreduce_moments(acc.buf, group)
    return acc.finalize(num_mc, uncertainty)


class HostBatchPrefetcher:
    """Double-buffered host -> device input pipeline for MC inference over a stream of batches.

    The reference's evaluate() loops copy every batch synchronously before its S forwards
    (examples/main_bayesian_cifar_dnn2bnn.py:537-541: `data.cuda()`).  Here the copy of batch i+1 (pinned host
    memory, a side stream, converted to the model's dtype / memory format on that stream) runs while batch i
    computes, so the PCIe time (77 MB per 128 fp32 ImageNet-size images) leaves the step's critical path.

        pf = HostBatchPrefetcher(device, dtype=torch.bfloat16, memory_format=torch.channels_last)
        pf.submit(first_batch)
        for nxt in rest:            # pinned host tensors
            x = pf.take()           # device tensor of the batch submitted earlier
            pf.submit(nxt)          # its copy overlaps the forwards of x
            res = mc_predict(model, x, S)
    """

    def __init__(self, device, dtype=None, memory_format=torch.contiguous_format):
        self.device = torch.device(device)
        self.dtype, self.memory_format = dtype, memory_format
        self.stream = torch.cuda.Stream(self.device)
        self._pending = None

    def submit(self, host_batch: torch.Tensor) -> None:
        if self._pending is not None:
            raise RuntimeError("HostBatchPrefetcher: take() the pending batch before submitting another")
        if host_batch.is_cuda:
            raise RuntimeError("HostBatchPrefetcher.submit expects a host tensor (pin it for an asynchronous copy)")
        with torch.cuda.stream(self.stream):
            x = host_batch.to(self.device, non_blocking=True)
            if self.dtype is not None:
                x = x.to(self.dtype)
            x = x.contiguous(memory_format=self.memory_format)
            done = torch.cuda.Event()
            done.record(self.stream)
        self._pending = (x, done)

    def take(self) -> torch.Tensor:
        if self._pending is None:
            raise RuntimeError("HostBatchPrefetcher: nothing submitted")
        x, done = self._pending
        self._pending = None
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(done)
        x.record_stream(cur)     # allocated on the side stream, consumed on the compute stream
        return x

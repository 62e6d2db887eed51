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


class MCPredictor:
    """mc_predict() for a frozen model and a fixed input shape, with the step's launches captured ONCE into a CUDA
    graph and replayed: the S/G folded forwards (54 fused layer kernels + pooling each for ResNet-50) and the
    accumulate kernels cost one graph launch of host time per step instead of ~80 us of Python / ctypes per
    kernel -- at 8 GPUs a step is ~9 ms of device time and its enqueue took 4.5 ms.  The C ABI is capturable by
    construction (no allocation, no synchronisation, everything on the caller's stream; include/bt_b200.h).

        predict = MCPredictor(model, 32, batch_samples=32)
        res = predict(x)          # first call per input shape: warm-up + capture; later calls: copy-in + replay

    The all-reduce of the moments and the finalize kernel stay outside the graph.  Sample ids are part of the
    captured launches: like mc_predict(first_sample=k), every call evaluates the same S draws.  The model's
    parameters must not change between calls (a changed parameter version or storage re-captures)."""

    def __init__(self, model, num_mc: int, *, softmax: bool = True, group=None, first_sample: int = 0, batch_samples: int = 1,
                 uncertainty: bool = True, use_graph: bool = True):
        self.model, self.num_mc, self.softmax, self.group = model, int(num_mc), softmax, group
        self.first_sample, self.batch_samples, self.uncertainty = int(first_sample), int(batch_samples), uncertainty
        self.use_graph = bool(use_graph)
        self._cap = None          # (key, graph, static_x, acc, keep-alive list)
        self.launches_per_step = 0
        self.graphed = False

    def _shard(self):
        import torch.distributed as dist
        rank, world = 0, 1
        if dist.is_available() and dist.is_initialized():
            rank, world = dist.get_rank(self.group), dist.get_world_size(self.group)
        return shard_samples(self.num_mc, rank, world)

    def _forwards(self, x, acc):
        """This rank's share of the samples: folded model passes + accumulation (what the graph holds).
        Returns the accumulator (created from the first logits when `acc` is None)."""
        first, last = self._shard()
        s = first
        if first >= last and acc is None:   # more ranks than samples: this rank contributes zeros
            with ops.mc_sample(self.first_sample):
                probe = self.model(x)
            probe = probe[0] if isinstance(probe, tuple) else probe
            acc = MCAccumulator(x.shape[0], probe.reshape(probe.shape[0], -1).shape[1], x.device, self.softmax)
        while s < last:
            n = min(self.batch_samples, last - s)
            if n == 1:
                with ops.mc_sample(self.first_sample + s):
                    logits = self.model(x)
            else:
                with ops.mc_batched(n, self.first_sample + s, shared_batch=x.shape[0]):
                    logits = self.model(x)
            if isinstance(logits, tuple):
                logits = logits[0]
            logits = logits.reshape(logits.shape[0], -1)
            if acc is None:
                acc = MCAccumulator(x.shape[0], logits.shape[1], x.device, self.softmax)
            acc.add(logits, n)
            s += n
        return acc

    def _key(self, x):
        params = tuple((p.data_ptr(), p._version) for p in self.model.parameters())
        return (tuple(x.shape), x.dtype, tuple(x.stride()), x.device, ops.get_seed(), self._shard(), params)

    def invalidate(self):
        self._cap = None

    @torch.no_grad()
    def __call__(self, x: torch.Tensor) -> MCResult:
        ops._require_cuda(x)
        if not self.use_graph:
            return mc_predict(self.model, x, self.num_mc, softmax=self.softmax, group=self.group, first_sample=self.first_sample,
                              batch_samples=self.batch_samples, uncertainty=self.uncertainty)
        key = self._key(x)
        if self._cap is None or self._cap[0] != key:
            self._capture(x, key)
        _, graph, static_x, acc, _ = self._cap
        if x.data_ptr() != static_x.data_ptr():
            static_x.copy_(x)
        graph.replay()
        allreduce_moments(acc.buf, self.group)
        return acc.finalize(self.num_mc, self.uncertainty)

    def _capture(self, x, key):
        dev = x.device
        static_x = torch.empty_strided(x.shape, x.stride(), dtype=x.dtype, device=dev)
        static_x.copy_(x)
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):                   # warm-up ON the capture stream: frozen layers pack their
            acc = self._forwards(static_x, None)        # parameters now, the captured launches find them valid
        side.synchronize()
        graph = torch.cuda.CUDAGraph()
        before = ops.PROFILE.launches
        with torch.cuda.graph(graph, stream=side, capture_error_mode="thread_local"):
            acc.buf.zero_()
            acc.count = 0
            self._forwards(static_x, acc)
        self.launches_per_step = ops.PROFILE.launches - before + 1      # + the finalize kernel outside the graph
        torch.cuda.current_stream(dev).wait_stream(side)
        # the captured launches hold raw pointers into the layers' packed-parameter workspaces: keep those alive
        # even if a later eager call on another stream replaces the cache entries
        keep = [m._pack_cache["ws"][1] for m in self.model.modules()
                if isinstance(getattr(m, "_pack_cache", None), dict) and "ws" in m._pack_cache]
        self._cap = (key, graph, static_x, acc, keep)
        self.graphed = True


class GradientBuckets:
    """Data-parallel gradient exchange of the training path: fixed, pre-allocated flat buckets
    filled from inside backward; a bucket's NCCL all-reduce (average) is issued the moment its last gradient has
    been produced, so it overlaps the dgrad / wgrad kernels of the layers still to come.  Replaces the
    DistributedDataParallel wrapper of the reference's training drivers (utils/utils.py:482-495) around the loop
    of examples/main_bayesian_cifar_dnn2bnn.py:404-420.

        buckets = GradientBuckets(model)             # once, after the model is on its device
        loss.backward()                              # hooks copy each gradient into its bucket, all-reduce per bucket
        buckets.finish()                             # waits for the collectives; p.grad are views of the averaged buckets
        opt.step()

    Buckets follow the REVERSE parameter order (the order backward produces gradients in).  p.grad is re-pointed at
    its slice of the bucket, so no copy back is needed.  The KL term is replica-identical: scale it by the GLOBAL
    batch in the loss (kl / (B * world)) and the averaged gradient equals the single-process one."""

    def __init__(self, model, group=None, bucket_mb: float = 25.0):
        import torch.distributed as dist
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self._avg = self.world > 1 and dist.get_backend(group) == "nccl"      # gloo has no AVG: sum, then divide in finish()
        params = [p for p in model.parameters() if p.requires_grad]
        limit = int(bucket_mb * (1 << 20)) // 4
        self.buckets = []        # dicts: flat, params [(p, off, n)], pending, work
        cur, cur_n = [], 0
        for p in reversed(params):
            # a bucket is one flat buffer of ONE dtype on one device: fp32 for the variational parameters (the
            # kernels' masters), the model's own dtype for what .bfloat16() converted (BatchNorm affine terms)
            if cur and (cur_n + p.numel() > limit or p.device != cur[0].device or p.dtype != cur[0].dtype):
                self._close(cur, cur_n); cur, cur_n = [], 0
            cur.append(p); cur_n += p.numel()
        if cur:
            self._close(cur, cur_n)
        self._where = {}
        self._handles = []
        for bi, b in enumerate(self.buckets):
            for p, off, n in b["params"]:
                self._where[id(p)] = (bi, off, n)
                self._handles.append(p.register_post_accumulate_grad_hook(self._hook))
        self.collectives = 0

    def _close(self, ps, n):
        flat = torch.zeros(n, dtype=ps[0].dtype, device=ps[0].device)
        offs, off = [], 0
        for p in ps:
            offs.append((p, off, p.numel())); off += p.numel()
        self.buckets.append({"flat": flat, "params": offs, "pending": len(ps), "work": None, "fired": False})

    def _hook(self, p):
        bi, off, n = self._where[id(p)]
        b = self.buckets[bi]
        view = b["flat"][off:off + n].view_as(p)
        view.copy_(p.grad)
        p.grad = view                       # the optimizer reads the averaged gradient straight from the bucket
        b["pending"] -= 1
        if b["pending"] == 0:
            self._fire(b)

    def _fire(self, b):
        import torch.distributed as dist
        if self.world > 1 and not b["fired"]:
            op = dist.ReduceOp.AVG if self._avg else dist.ReduceOp.SUM
            b["work"] = dist.all_reduce(b["flat"], op=op, group=self.group, async_op=True)
            self.collectives += 1
        b["fired"] = True

    def finish(self) -> int:
        """Flush buckets some of whose parameters produced no gradient this step, wait for every collective (the
        current stream waits, not the host) and re-arm the buckets.  Returns the number of collectives of the step."""
        n = 0
        for b in self.buckets:
            if not b["fired"] and b["pending"] < len(b["params"]):
                self._fire(b)
        for b in self.buckets:
            if b["work"] is not None:
                b["work"].wait()
                if not self._avg:
                    b["flat"].div_(self.world)
                n += 1
            b["work"], b["pending"], b["fired"] = None, len(b["params"]), False
        return n

    def remove(self):
        for h in self._handles:
            h.remove()
        self._handles = []


class HostBatchPrefetcher:
    """Double-buffered host -> device input pipeline for MC inference over a stream of batches.

    The reference's evaluate() loops copy every batch synchronously before its S forwards
    (examples/main_bayesian_cifar_dnn2bnn.py:537-541: `data.cuda()`).  Here the copy of batch i+1 (pinned host
    memory, a side stream, converted to the model's dtype / memory format on that stream) runs while batch i
    computes, so the PCIe time (77 MB per 128 fp32 ImageNet-size images) leaves the step's critical path.

    shard=True (sample-sharded multi-GPU inference: every rank needs the WHOLE batch): rank r copies only rows
    [r*B/G, (r+1)*B/G) over PCIe (overlapped with the previous step), converts them, and one NCCL all-gather over
    NVLink (in take(), between two steps) assembles the batch on every GPU -- G ranks pulling the same 77 MB through
    one host's PCIe complex was what held the 8-GPU end-to-end number 15 % under the device-resident one.

        pf = HostBatchPrefetcher(device, dtype=torch.bfloat16, memory_format=torch.channels_last)
        pf.submit(first_batch)
        for nxt in rest:            # pinned host tensors
            x = pf.take()           # device tensor of the batch submitted earlier
            pf.submit(nxt)          # its copy overlaps the forwards of x
            res = mc_predict(model, x, S)
    """

    def __init__(self, device, dtype=None, memory_format=torch.contiguous_format, shard: bool = False, group=None):
        self.device = torch.device(device)
        self.dtype, self.memory_format = dtype, memory_format
        self.stream = torch.cuda.Stream(self.device)
        self.shard, self.group = bool(shard), group
        self._pending = None

    def _world(self):
        import torch.distributed as dist
        if self.shard and dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            return dist.get_rank(self.group), dist.get_world_size(self.group)
        return 0, 1

    @staticmethod
    def _flat(t):
        """1-d view of a dense tensor in MEMORY order (channels-last 4-d: [B, H, W, C]); a batch shard is one chunk."""
        if t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last) and not t.is_contiguous():
            return t.permute(0, 2, 3, 1).reshape(-1)
        return t.reshape(-1)

    def submit(self, host_batch: torch.Tensor) -> None:
        if self._pending is not None:
            raise RuntimeError("HostBatchPrefetcher: take() the pending batch before submitting another")
        if host_batch.is_cuda:
            raise RuntimeError("HostBatchPrefetcher.submit expects a host tensor (pin it for an asynchronous copy)")
        rank, world = self._world()
        B = host_batch.shape[0]
        work = None
        with torch.cuda.stream(self.stream):
            if world > 1 and B % world == 0:
                import torch.distributed as dist
                lo = rank * (B // world)
                part = host_batch[lo:lo + B // world].to(self.device, non_blocking=True)
                if self.dtype is not None:
                    part = part.to(self.dtype)
                part = part.contiguous(memory_format=self.memory_format)
                x = torch.empty((B, *part.shape[1:]), dtype=part.dtype, device=self.device, memory_format=self.memory_format)
                work = part        # gathered in take(), on the compute stream
            else:
                x = host_batch.to(self.device, non_blocking=True)
                if self.dtype is not None:
                    x = x.to(self.dtype)
                x = x.contiguous(memory_format=self.memory_format)
            done = torch.cuda.Event()
            done.record(self.stream)
        self._pending = (x, done, work)

    def take(self) -> torch.Tensor:
        if self._pending is None:
            raise RuntimeError("HostBatchPrefetcher: nothing submitted")
        x, done, work = self._pending
        self._pending = None
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(done)
        if work is not None:
            # The all-gather runs ON the compute stream, between two steps: ~0.1 ms over NVLink.  Issued from the side
            # stream it overlapped the step in principle, but its CTAs then compete with the persistent layer kernels
            # (one CTA per SM, all of shared memory): a rank whose peers were late kept SMs busy-waiting and every layer
            # kernel behind it started short of SMs -- 17 % of the step at 4 GPUs.
            import torch.distributed as dist
            work.record_stream(cur)
            dist.all_gather_into_tensor(self._flat(x), self._flat(work), group=self.group)
        x.record_stream(cur)     # allocated on the side stream, consumed on the compute stream
        return x

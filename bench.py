#!/usr/bin/env python
"""Headline benchmark: Bayesian ResNet-50 Monte-Carlo inference, S=32 weight samples,
batch 128 x 3 x 224 x 224 synthetic images (BASELINE.json configs[3]; the metric is quoted
on it and it fits one B200).  torchvision.resnet50 + dnn_to_bnn(Reparameterization), eval.

    python bench.py --gpus N --steps K --warmup W            # this repo (CUDA, sm_100a)
    python bench.py --impl reference --gpus N ...            # the reference's own eager code on the host cores
    python bench.py --workload train --gpus N ...            # configs[2] extended to N GPUs (ResNet-18 Flipout training step)

One step = one MC prediction of one batch: S weight samples sharded over the N ranks (S/N
full-batch forwards each), predictive mean/variance/entropy/MI joined by one NCCL
all-reduce.  images/s counts every image ONCE however many samples it was evaluated with
(the reference's definition, examples/main_bayesian_cifar_dnn2bnn.py:549).
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

# stdout carries ONE JSON line.  Libraries write there too (NCCL's version banner under NCCL_DEBUG=VERSION / WARN / INFO comes
# from C code, past sys.stdout): file descriptor 1 is pointed at stderr for the whole run and the line goes to the saved one.
_STDOUT_FD = None


def _claim_stdout():
    global _STDOUT_FD
    if _STDOUT_FD is None:
        sys.stdout.flush()
        _STDOUT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    os.write(_STDOUT_FD if _STDOUT_FD is not None else 1, data)


ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "bayesian_resnet50_mc_images_per_s_S32"
UNIT = "images/s"
GFLOP_PER_IMAGE_SAMPLE = 8.178368512  # 2 * 4 089 184 256 MACs, Bayesian layers only
BNN_PARAMS = {"prior_mu": 0.0, "prior_sigma": 1.0, "posterior_mu_init": 0.0, "posterior_rho_init": -3.0,
              "type": "Reparameterization", "moped_enable": True, "moped_delta": 0.5}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="mc_inference", choices=["mc_inference", "train"])
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--samples", type=int, default=32)
    ap.add_argument("--mc-batch", type=int, default=int(os.environ.get("BT_MC_BATCH", "32")),
                    help="weight samples folded into one model pass (ops.mc_batched)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-eager", action="store_true", help="skip the reference-eager-on-this-GPU comparison")
    ap.add_argument("--no-extras", action="store_true", help="skip the configs[2] / configs[4] / KL side measurements")
    ap.add_argument("--no-fold-bn", action="store_true", help="keep BatchNorm / ReLU as separate torch ops")
    ap.add_argument("--no-graph", action="store_true", help="enqueue every step from Python instead of replaying a CUDA graph")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"tflops": float(p.get("bf16_tflops_sustained", 1381.2)), "gbs": float(p.get("hbm_gbs", 6541.5)),
                "source": "MEASURED_PEAKS.json (sustained bf16: kernel timed inside a long step)"}
    return {"tflops": 1400.0, "gbs": 6650.0, "source": "fallback of B200_PROFILING.md (sustained ~1.4 PFLOP/s)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def seed_resnet50():
    """The deterministic network both arms convert: torchvision ResNet-50, random init (no checkpoints offline).
    gamma of the last BatchNorm of every block is 0.5: that keeps the random-init residual stack (weight noise
    sigma = 0.5 |w| on top) in a regime where the logits are O(1) and the softmax is not saturated, so the
    predictive mean / variance a step returns (and the checksum printed with it) mean something.  No effect on
    the work done."""
    import torch
    import torchvision
    torch.manual_seed(0)
    model = torchvision.models.resnet50()
    for name, mod in model.named_modules():
        if name.endswith("bn3"):
            mod.weight.data.fill_(0.5)
    return model


# ------------------------------------------------------------------------------------------
# reference arm / cpu baseline / eager-on-this-GPU: the reference's OWN code (oracle/_ref, staged by
# oracle/build_ref.py), never this repo's kernels
# ------------------------------------------------------------------------------------------
def reference_model(device):
    """torchvision ResNet-50 converted by the reference's dnn_to_bnn (models/dnn_to_bnn.py:127-154, MOPED on),
    eval mode -- the model examples/main_bayesian_imagenet_dnn2bnn.py builds."""
    from oracle.build_ref import import_reference
    L, D, U, where = import_reference()
    model = seed_resnet50()
    D.dnn_to_bnn(model, dict(BNN_PARAMS))
    return model.to(device).eval(), where


def cpu_reference_images_per_s(batch: int, samples: int, forwards: int, warm: int = 1, budget_s: float = 120.0):
    """Time the reference's eager CPU path on this host's cores: `forwards` Monte-Carlo forwards of the FULL batch
    (fp32, all threads).  One step of the metric is `samples` such forwards (the evaluate() loop,
    examples/main_bayesian_cifar_dnn2bnn.py:542-556), so images/s = batch / (t_forward * samples)."""
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    try:
        model, where = reference_model("cpu")
        kind = "reference"
    except ImportError:   # oracle/_ref not staged (a clone that never saw /root/reference): the oracle port
        from oracle.layers import oracle_dnn_to_bnn
        model, where, kind = oracle_dnn_to_bnn(seed_resnet50()).eval(), "oracle/layers.py", "port"
    g = torch.Generator().manual_seed(0)
    x = torch.randn(batch, 3, 224, 224, generator=g)
    times = []
    t_all = time.perf_counter()
    with torch.no_grad():
        for i in range(warm + forwards):
            t0 = time.perf_counter()
            model(x)
            dt = time.perf_counter() - t0
            if i >= warm:
                times.append(dt)
            if time.perf_counter() - t_all > budget_s and times:
                break
    t = sum(times) / len(times)
    return {"value": batch / (t * samples), "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"{len(times)} Monte-Carlo forward(s) of the full batch ({batch} images, fp32, torch CPU, {cores} threads, "
                      f"{warm} warm-up) of the {samples} a step needs: {t:.2f} s each; the step time is {samples} x that",
            "source": where, "seconds_per_forward": t, "forwards_timed": len(times)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1 to every rank; the reference arm is ONE process that should use every host
    # core, so the thread pools are sized before torch is imported (nothing above this line imports it)
    for var in ("OMP_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[var] = str(os.cpu_count() or 1)
    warm = max(1, min(args.warmup, 2))
    base = cpu_reference_images_per_s(args.batch, args.samples, forwards=args.steps, warm=warm, budget_s=200.0)
    t = base.pop("seconds_per_forward")
    n = base.pop("forwards_timed")
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": n,
            "warmup": warm, "ms_per_step": 1e3 * t * args.samples, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "resnet50_dnn_to_bnn_reparam_mc_inference", "batch": args.batch, "samples": args.samples,
                       "image": "3x224x224", "note": "the reference's own eager code on the host cores; each timed step is ONE of the "
                                                     f"{args.samples} full-batch Monte-Carlo forwards of a step (bounded sample), "
                                                     f"ms_per_step = {args.samples} x the measured forward"},
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def eager_on_gpu(dev, batch: int, samples: int, timed_forwards: int = 2):
    """SURVEY.md section 2's bar: the reference's eager op sequence ON THIS GPU (`.cuda()`, what a user of the
    reference gets today): softplus, normal_(), mul, add, F.conv2d per layer (conv_variational.py:362-385) x 54
    layers x S sequential forwards (main_bayesian_cifar_dnn2bnn.py:542-556), CUDA-event timed."""
    import torch
    out = {}
    g = torch.Generator().manual_seed(0)
    x32 = torch.randn(batch, 3, 224, 224, generator=g).to(dev)
    for tag in ("fp32_nchw", "bf16_channels_last"):
        try:
            model, where = reference_model(dev)
            x = x32
            if tag == "bf16_channels_last":   # the reference's --half analogue on this hardware
                model = model.to(torch.bfloat16).to(memory_format=torch.channels_last)
                x = x32.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
            with torch.no_grad():
                for _ in range(2):
                    model(x)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                outs = [model(x) for _ in range(timed_forwards)]
                torch.softmax(torch.stack(outs).float(), dim=-1).mean(0)
                e1.record()
                torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / timed_forwards
            out[tag] = {"value": batch * 1e3 / (ms * samples), "unit": UNIT, "ms_per_forward": ms,
                        "sample": f"{timed_forwards} of the {samples} full-batch forwards of a step, scaled",
                        "tf32_conv": bool(torch.backends.cudnn.allow_tf32)}
            del model
        except Exception as e:   # the reference's own limits (e.g. dtype handling) are reported, not hidden
            out[tag] = {"error": f"{type(e).__name__}: {e}"[:300]}
        torch.cuda.empty_cache()
    out["source"] = "oracle/_ref (the unmodified reference), torch eager on this GPU"
    return out


# ------------------------------------------------------------------------------------------
# this repo's arm
# ------------------------------------------------------------------------------------------
def measured_traffic(n_gpus, launches_per_step, samples_per_gpu):
    """DRAM bytes per launch of the dominant kernel from an ncu pass over one timed step of THIS command at THIS
    GPU count (profiles/r2_tc2_traffic.json, tools/summarize_launches.py); None when no capture matches."""
    path = os.path.join(ROOT, "profiles", "r2_tc2_traffic.json")
    try:
        with open(path) as f:
            for t in json.load(f).get("captures", []):
                if t.get("n_gpus") == n_gpus and t.get("launches") == launches_per_step and t.get("samples_per_gpu") == samples_per_gpu:
                    return t["dram_bytes_per_launch"]
    except (OSError, ValueError, KeyError):
        pass
    return None


def build_model(device, fold_bn=True, prepare=None):
    """The model every arm of this benchmark (and tests/test_config3_gpu.py) runs.  `prepare(model)` is called on
    the fp32 eval-mode model on `device` before the bf16 / channels-last conversion and the BatchNorm fold (the
    parity test builds its CPU oracle twin there)."""
    import torch
    import bayesian_torch_b200 as bt
    model = seed_resnet50()
    bt.dnn_to_bnn(model, dict(BNN_PARAMS))         # MOPED: mu <- DNN init, rho <- get_rho (ImageNet examples force it)
    model = model.to(device).eval()
    if prepare is not None:
        prepare(model)
    # bf16 activations, channels-last; variational parameters stay fp32 masters (layers/_apply)
    model = model.to(memory_format=torch.channels_last).to(torch.bfloat16)
    if fold_bn:   # eval-mode BatchNorm / ReLU / residual join folded into the layer kernels (exact up to rounding)
        model = bt.fold_batchnorm(model, inplace=True)
    return model


def roofline_of(records, ms_step, pk, tc_path, n_gpus, samples_per_gpu):
    tc = [(f, b, e0.elapsed_time(e1)) for path, f, b, e0, e1 in records if path == tc_path]
    other_ms = sum(e0.elapsed_time(e1) for path, f, b, e0, e1 in records if path != tc_path)
    tc_ms, tc_flops = sum(t for _, _, t in tc), sum(f for f, _, _ in tc)
    achieved = tc_flops / (tc_ms * 1e-3) / 1e12 if tc_ms > 0 else 0.0
    # every layer against ITS OWN binding bound: tensor pipe or HBM, whichever takes longer
    t_tc = [f / (pk["tflops"] * 1e12) * 1e3 for f, _, _ in tc]
    t_hbm = [b / (pk["gbs"] * 1e9) * 1e3 for _, b, _ in tc]
    comp = [i for i in range(len(tc)) if t_tc[i] >= t_hbm[i]]
    mem = [i for i in range(len(tc)) if t_tc[i] < t_hbm[i]]
    mixed = sum(max(a, b) for a, b in zip(t_tc, t_hbm))

    def frac(idx, bound):
        return (sum(bound[i] for i in idx) / sum(tc[i][2] for i in idx)) if idx else None
    return {"bound": "tensor", "kernel": "tc2_fwd_kernel (all launches of one step)", "achieved": achieved,
            "peak": pk["tflops"], "unit": "TFLOP/s", "frac": achieved / pk["tflops"],
            "traffic": measured_traffic(n_gpus, len(tc), samples_per_gpu),
            "algorithmic_bytes_per_launch": (sum(b for _, b, _ in tc) / len(tc)) if tc else None,
            "peak_source": pk["source"], "launches_per_step": len(tc), "ms_per_step_in_kernel": tc_ms,
            "ms_per_step_other_bt_kernels": other_ms, "share_of_step": tc_ms / ms_step if ms_step > 0 else None,
            "mixed_bound_ms": mixed, "mixed_bound_frac": mixed / tc_ms if tc_ms > 0 else None,
            "compute_bound_layers": {"count": len(comp), "ms": sum(tc[i][2] for i in comp), "frac_of_tensor_peak": frac(comp, t_tc)},
            "hbm_bound_layers": {"count": len(mem), "ms": sum(tc[i][2] for i in mem), "frac_of_hbm_peak": frac(mem, t_hbm),
                                 "hbm_peak_gbs": pk["gbs"]}}


def cuda_time(fn, iters, warm=2):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def extras(dev, model, pk):
    """Side measurements on the same box, N = 1: BASELINE.json configs[4] (LinearReparameterization sweep, one weight
    sample per call, bf16 tensor-core path and fp32 activations, fused KL on), the KL kernel over ResNet-50's table,
    configs[2] (ResNet-18 Flipout training step).  Reported under `extra`; none of it is in the timed region."""
    import torch
    import bayesian_torch_b200 as bt
    from bayesian_torch_b200 import layers as L
    out = {}
    sweep = []
    for n in (1024, 2048, 4096, 8192, 16384):
        torch.manual_seed(0)
        layer = L.LinearReparameterization(n, n, bias=True).to(dev)
        row = {"n": n, "batch": 4096}
        xb = torch.randn(4096, n, device=dev).bfloat16()
        with torch.no_grad():
            for tag, x, iters in (("bf16", xb, 10), ("fp32", xb.float(), 10 if n <= 4096 else 2)):
                ms = cuda_time(lambda: layer(x, return_kl=True), iters)
                tf = 2.0 * 4096 * n * n / ms / 1e9
                row[tag] = {"ms_fwd_plus_kl": ms, "tflops": tf, "frac_of_tensor_peak": tf / pk["tflops"]}
            kms = cuda_time(lambda: layer.kl_loss(), 10, warm=3)
            row["kl"] = {"ms": kms, "gbs": 8.0 * (n * n + n) / kms / 1e6, "frac_of_hbm_peak": 8.0 * (n * n + n) / kms / 1e6 / pk["gbs"]}
        sweep.append(row)
        del layer, xb, x
        torch.cuda.empty_cache()
    out["cfg4_linear_sweep_S1"] = sweep
    with torch.no_grad():
        nbytes = 8 * sum(p.numel() for n_, p in model.named_parameters() if n_.rsplit(".", 1)[-1].startswith("mu_"))
        kms = cuda_time(lambda: bt.get_kl_loss(model), 20, warm=3)
    out["kl_resnet50_table"] = {"ms": kms, "gbs": nbytes / kms / 1e6, "frac_of_hbm_peak": nbytes / kms / 1e6 / pk["gbs"], "bytes": nbytes}
    out["cfg2_resnet18_flipout_train_step"] = train_step_time(dev, 128, "Flipout", iters=5)
    # the reference's README usage, untouched: dnn_to_bnn(model); model.cuda(); S sequential forwards on fp32 NCHW input
    # (no bf16 cast, no BatchNorm fold, no sample folding) -- fp32 activations run the tf32 tensor-core instantiation
    m32 = seed_resnet50()
    bt.dnn_to_bnn(m32, dict(BNN_PARAMS))
    m32 = m32.to(dev).eval()
    x32 = torch.randn(128, 3, 224, 224, device=dev)

    def fwd32():
        with torch.no_grad(), bt.mc_sample(0):
            m32(x32)
    ms = cuda_time(fwd32, 3)
    out["fp32_nchw_default_path"] = {"value": 128 * 1e3 / (ms * 32), "unit": UNIT, "ms_per_forward": ms,
                                     "note": "drop-in default: fp32 NCHW input, path auto (tf32 tensor cores), torch BatchNorm / ReLU, "
                                             "32 sequential forwards per image scaled from 3; compare eager_b200.fp32_nchw"}
    return out


def make_train_step(dev, batch, typ="Flipout", arch="resnet18"):
    """BASELINE.json configs[2]: torchvision ResNet-18 + dnn_to_bnn(type=Flipout), synthetic 3x224x224, one step =
    forward, get_kl_loss, CE + kl / batch, backward, SGD (examples/main_bayesian_cifar_dnn2bnn.py:374-449, num_mc 1)."""
    import torch
    import torchvision
    import bayesian_torch_b200 as bt
    torch.manual_seed(0)
    model = getattr(torchvision.models, arch)()
    bt.dnn_to_bnn(model, dict(BNN_PARAMS, type=typ, moped_enable=False))
    model = model.to(dev).bfloat16().to(memory_format=torch.channels_last).train()
    opt = torch.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9)
    return model, opt


def train_step_time(dev, batch, typ, iters):
    import torch
    import bayesian_torch_b200 as bt
    model, opt = make_train_step(dev, batch, typ)
    x = torch.randn(batch, 3, 224, 224, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 1000, (batch,), device=dev)

    def step():
        opt.zero_grad(set_to_none=True)
        out = model(x)
        loss = torch.nn.functional.cross_entropy(out.float(), y) + bt.get_kl_loss(model) / batch
        loss.backward()
        opt.step()
    ms = cuda_time(step, iters)
    return {"ms": ms, "images_per_s": batch * 1e3 / ms, "batch": batch, "dtype": "bf16 activations, fp32 parameters"}


def run_b200(args):
    import torch
    import torch.distributed as dist
    import bayesian_torch_b200 as bt
    from bayesian_torch_b200 import _lib, ops

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus or world == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    if args.workload == "train":
        return run_train(args, dev, rank, world)

    model = build_model(dev, fold_bn=not args.no_fold_bn)
    B, S = args.batch, args.samples
    g = torch.Generator().manual_seed(0)
    x_host = torch.randn(B, 3, 224, 224, generator=g).pin_memory()          # synthetic images, host resident
    x_dev = x_host.to(dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    bt.manual_seed(1234)

    # One step through the public API.  bt.MCPredictor captures the step's launches (the S/N folded forwards, the
    # accumulate / finalize kernels) into a CUDA graph on first use and replays it: at 8 GPUs a step is ~9 ms of
    # device time and enqueueing its 58 launches from Python took 4.5 ms of host time.
    predictor = bt.MCPredictor(model, S, batch_samples=args.mc_batch, use_graph=not args.no_graph)

    def step_resident():
        return predictor(x_dev)

    # end to end through the public API: every step brings its batch in from pinned host memory -- each rank copies
    # its 1/N of the images over PCIe and an all-gather over NVLink assembles the batch on every GPU
    # (bt.HostBatchPrefetcher: the copy of the NEXT step's batch overlaps this step's forwards) -- and reads the
    # predictive mean / variance back to the host
    prefetch = bt.HostBatchPrefetcher(dev, dtype=torch.bfloat16, memory_format=torch.channels_last, shard=world > 1)

    # results come back the way the batches go in: asynchronously, into two pairs of pinned host buffers, on a side stream
    # that waits for the step's last kernel -- the host never blocks the next step's launches on a read-back (a blocking
    # .cpu() per step left the GPUs idle for the host round trip: 8 % of the step at 8 GPUs).  A pair is reused two steps
    # later, after its event; the timed region ends with a device synchronize, so every read-back has landed inside it.
    d2h_stream = torch.cuda.Stream(dev)
    host_out = []
    turn = [0]

    def step_e2e():
        if prefetch._pending is None:          # first step of a timed run: nothing to overlap with, the copy is exposed
            prefetch.submit(x_host)
        xd = prefetch.take()
        res = predictor(xd)
        prefetch.submit(x_host)                # next step's batch (same synthetic images): H2D on the side stream
        if not host_out:
            for _ in range(2):
                host_out.append((torch.empty(res.mean.shape, dtype=res.mean.dtype).pin_memory(),
                                 torch.empty(res.var.shape, dtype=res.var.dtype).pin_memory(), torch.cuda.Event()))
        m, v, ev = host_out[turn[0] & 1]
        turn[0] += 1
        ev.synchronize()
        d2h_stream.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(d2h_stream):
            m.copy_(res.mean, non_blocking=True)
            v.copy_(res.var, non_blocking=True)
            ev.record(d2h_stream)
        res.mean.record_stream(d2h_stream)
        res.var.record_stream(d2h_stream)
        return m, v

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, tail=None):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            out = fn()
        if tail is not None:
            tail()                 # e.g. the last read-back on the side stream joins the timed stream before the closing event
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms) / steps, out

    for _ in range(max(args.warmup, 3)):
        step_resident()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ops.PROFILE.reset()
    ncu_range = os.environ.get("BT_NCU_RANGE") == "1"   # `ncu --profile-from-start off`: capture the timed steps only
    if ncu_range:
        torch.cuda.cudart().cudaProfilerStart()
    ms_step, res = timed(step_resident, args.steps)
    if ncu_range:
        torch.cuda.cudart().cudaProfilerStop()
    launches = predictor.launches_per_step * args.steps if predictor.graphed else ops.PROFILE.launches
    clocks = sampler.stop() if rank == 0 else None
    # what the step returned, as numbers the N = 1 / 2 / 4 / 8 lines can be compared by (sample s is the same draw on
    # any rank: the sums differ only by the order of the fp32 additions)
    checksum = {"sum_mean_sq": float((res.mean.double() ** 2).sum()), "sum_var": float(res.var.double().sum()),
                "sum_entropy": float(res.entropy.double().sum()), "sum_mutual_info": float(res.mutual_info.double().sum()),
                "max_mean": float(res.mean.max()), "argmax_sum": int(res.mean.argmax(-1).sum())}

    # roofline of the dominant kernel (tc2_fwd_kernel): CUDA events around every layer call of one more step
    # enqueued eagerly on the launching stream (the same launches the graph replays)
    ops.PROFILE.reset()
    ops.PROFILE.timing = True
    barrier()
    with torch.no_grad():
        bt.mc_predict(model, x_dev, S, batch_samples=args.mc_batch)
    torch.cuda.synchronize()
    ops.PROFILE.timing = False
    records = ops.PROFILE.records

    # the end-to-end arm warms up too (the first copies allocate the side stream's device buffers: a one-off 40-200 ms that
    # a 5-step timed region would otherwise carry); the pending batch is drained so that the first timed copy is exposed
    for _ in range(max(args.warmup, 3)):
        step_e2e()
    prefetch.take()
    ms_e2e, _ = timed(step_e2e, args.steps, tail=lambda: torch.cuda.current_stream(dev).wait_stream(d2h_stream))
    prefetch.take()                            # the batch submitted by the last step (one copy more than steps timed)

    if rank == 0:
        pk = peaks()
        value = B * 1e3 / ms_step
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": "resnet50_dnn_to_bnn_reparam_mc_inference", "batch": B, "samples": S, "image": "3x224x224",
                       "samples_per_gpu": S / world, "mc_batch": args.mc_batch, "bn_folded": not args.no_fold_bn, "parallelism": f"mc_samples_sharded_x{world}",
                       "cuda_graph": predictor.graphed,
                       "l2": "activations per step (GBs) far exceed the 126 MB L2; no flush needed",
                       "image_samples_per_s": value * S,
                       "tflops_bayesian_layers": value * S * GFLOP_PER_IMAGE_SAMPLE / 1e3},
            "clocks": clocks,
            "e2e": {"value": B * 1e3 / ms_e2e, "unit": UNIT, "h2d_bytes_per_step": x_host.numel() * 4,
                    "d2h_bytes_per_step": 2 * res.mean.numel() * 4 * world,
                    "note": "whole-job bytes: each rank copies batch/N images H2D (an NVLink all-gather assembles the batch) "
                            "and reads mean + variance back"},
            "gpu_launches": launches,
            "checksum": checksum,
            "roofline": roofline_of(records, ms_step, pk, _lib.PATH_TC_BF16, world, S / world),
        }
        if world == 1:   # rank-0, N = 1 side figures (torchrun pins OMP threads to 1, and they need a free GPU)
            torch.cuda.empty_cache()
            if not args.no_cpu_baseline:
                base = cpu_reference_images_per_s(B, S, forwards=2, warm=1, budget_s=60.0)
                line["cpu_baseline"] = {k: v for k, v in base.items() if k not in ("seconds_per_forward", "forwards_timed")}
            if not args.no_eager:
                try:
                    line["eager_b200"] = eager_on_gpu(dev, B, S)
                except ImportError as e:
                    line["eager_b200"] = {"error": str(e)}
            if not args.no_extras:
                line["extra"] = extras(dev, model, pk)
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_train(args, dev, rank, world):
    """configs[2] extended to N GPUs (training): data-parallel ResNet-18 Flipout, batch 128 PER GPU
    (weak scaling), the mu / rho gradients summed by NCCL in fixed fp32 buckets issued from inside backward
    (bt.GradientBuckets) so the all-reduce overlaps the remaining dgrad / wgrad kernels."""
    import torch
    import torch.distributed as dist
    import bayesian_torch_b200 as bt
    from bayesian_torch_b200 import ops
    B = args.batch
    model, opt = make_train_step(dev, B)
    bt.manual_seed(1234 + rank)        # each rank draws its own weight noise, like the reference's DDP ranks
    buckets = bt.GradientBuckets(model, bucket_mb=25.0) if world > 1 else None
    g = torch.Generator().manual_seed(rank)
    x_host = torch.randn(B, 3, 224, 224, generator=g).pin_memory()
    y = torch.randint(0, 1000, (B,), generator=g).to(dev)
    x_dev = x_host.to(dev).bfloat16().contiguous(memory_format=torch.channels_last)
    prefetch = bt.HostBatchPrefetcher(dev, dtype=torch.bfloat16, memory_format=torch.channels_last)

    def step(x):
        opt.zero_grad(set_to_none=True)
        out = model(x)
        loss = torch.nn.functional.cross_entropy(out.float(), y) + bt.get_kl_loss(model) / (B * world)
        loss.backward()
        if buckets is not None:
            buckets.finish()           # the last buckets' all-reduces + the copy-back of the averaged gradients
        opt.step()
        return loss

    def step_e2e():
        if prefetch._pending is None:
            prefetch.submit(x_host)
        xd = prefetch.take()
        loss = step(xd)
        prefetch.submit(x_host)
        return loss.cpu()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            out = fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms) / steps, out

    for _ in range(max(args.warmup, 3)):
        step(x_dev)
    sampler = ClockSampler(dev.index)
    if rank == 0:
        sampler.start()
    ops.PROFILE.reset()
    ms_step, loss = timed(lambda: step(x_dev), args.steps)
    launches = ops.PROFILE.launches
    clocks = sampler.stop() if rank == 0 else None
    for _ in range(max(args.warmup, 3)):       # the end-to-end arm warms up too (first copies allocate the side stream's buffers)
        step_e2e()
    prefetch.take()
    ms_e2e, _ = timed(step_e2e, args.steps)
    prefetch.take()
    if rank == 0:
        value = B * world * 1e3 / ms_step
        emit({
            "metric": "bayesian_resnet18_flipout_train_images_per_s", "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "resnet18_dnn_to_bnn_flipout_training_step", "batch_per_gpu": B, "global_batch": B * world,
                       "image": "3x224x224", "parallelism": f"dp{world}", "grad_allreduce": "bucketed fp32, issued inside backward" if world > 1 else "none",
                       "l2": "activations per step far exceed the 126 MB L2; no flush needed"},
            "clocks": clocks, "loss": float(loss),
            "e2e": {"value": B * world * 1e3 / ms_e2e, "unit": UNIT, "h2d_bytes_per_step": x_host.numel() * 4 * world, "d2h_bytes_per_step": 4 * world},
            "gpu_launches": launches})
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    _claim_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)

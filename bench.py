#!/usr/bin/env python
"""Headline benchmark: Bayesian ResNet-50 Monte-Carlo inference, S=32 weight samples,
batch 128 x 3 x 224 x 224 synthetic images (BASELINE.json configs[3]; the metric is quoted
on it and it fits one B200).  torchvision.resnet50 + dnn_to_bnn(Reparameterization), eval.

    python bench.py --gpus N --steps K --warmup W            # this repo (CUDA, sm_100a)
    python bench.py --impl reference --gpus N ...            # reference CPU path (oracle port)

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

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "bayesian_resnet50_mc_images_per_s_S32"
UNIT = "images/s"
GFLOP_PER_IMAGE_SAMPLE = 8.178368512  # 2 * 4 089 184 256 MACs, Bayesian layers only


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--samples", type=int, default=32)
    ap.add_argument("--mc-batch", type=int, default=int(os.environ.get("BT_MC_BATCH", "32")),
                    help="weight samples folded into one model pass (ops.mc_batched)")
    ap.add_argument("--cpu-batch", type=int, default=8, help="images in the bounded CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-fold-bn", action="store_true", help="keep BatchNorm / ReLU as separate torch ops")
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


# ------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the oracle port of the reference's CPU path on host cores
# ------------------------------------------------------------------------------------------
def cpu_reference_images_per_s(batch: int, samples: int, cpu_batch: int, repeats: int = 1):
    """Time the oracle port (oracle/layers.py) of the reference's eager path on this host's
    cores: torchvision.resnet50 with every Conv2d/Linear swapped for the CPU restatement of
    the Reparameterization layers, fp32, all threads.  Bounded sample: `cpu_batch` images x 1
    weight sample per timed forward; one image needs `samples` forwards, so
    images/s = cpu_batch / (t_forward * samples)."""
    import torch
    import torchvision
    from oracle.layers import oracle_dnn_to_bnn
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model = oracle_dnn_to_bnn(torchvision.models.resnet50()).eval()
    x = torch.randn(cpu_batch, 3, 224, 224)
    with torch.no_grad():
        model(x)  # warm-up
        best = float("inf")
        for _ in range(max(1, repeats)):
            t0 = time.perf_counter()
            model(x)
            best = min(best, time.perf_counter() - t0)
    ips = cpu_batch / (best * samples)
    return {"value": ips, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{cpu_batch} images x 1 weight sample per timed forward (fp32, torch CPU, {cores} threads), "
                      f"best of {max(1, repeats)}; scaled to S={samples} forwards per image",
            "seconds_per_forward": best}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1 to every rank; the reference arm is ONE process that should use every host
    # core, so the thread pools are sized before torch is imported (nothing above this line imports it)
    for var in ("OMP_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[var] = str(os.cpu_count() or 1)
    t_all = time.perf_counter()
    per_step = []
    base = None
    for i in range(args.warmup + args.steps):
        base = cpu_reference_images_per_s(args.batch, args.samples, args.cpu_batch)
        if i >= args.warmup:
            per_step.append(base["seconds_per_forward"])
        if time.perf_counter() - t_all > 240:   # keep the whole run within a few minutes
            break
    if not per_step:
        per_step = [base["seconds_per_forward"]]
    t = sum(per_step) / len(per_step)
    value = args.cpu_batch / (t * args.samples)
    ms_full = 1e3 * t * args.samples * args.batch / args.cpu_batch
    base["value"] = value
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": len(per_step),
            "warmup": args.warmup, "ms_per_step": ms_full, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "resnet50_dnn_to_bnn_reparam_mc_inference", "batch": args.batch, "samples": args.samples,
                       "image": "3x224x224", "note": "reference CPU path (oracle port); each step is a bounded sample, "
                                                     "ms_per_step extrapolated to the full batch x S"},
            "cpu_baseline": {k: v for k, v in base.items() if k != "seconds_per_forward"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------
# this repo's arm
# ------------------------------------------------------------------------------------------
def tc_traffic(launches_per_step):
    """DRAM bytes per launch of the dominant kernel from the committed ncu pass over one timed step
    of this same command (profiles/r1_tc2_traffic.json, tools/summarize_launches.py); None when the
    capture does not match the launch count of this run."""
    path = os.path.join(ROOT, "profiles", "r1_tc2_traffic.json")
    try:
        with open(path) as f:
            t = json.load(f)
        if t.get("launches") == launches_per_step:
            return t["dram_bytes_per_launch"]
    except (OSError, ValueError, KeyError):
        pass
    return None


def build_model(device, fold_bn=True, prepare=None):
    """The model every arm of this benchmark (and tests/test_config3_gpu.py) runs.  `prepare(model)` is called on
    the fp32 eval-mode model on `device` before the bf16 / channels-last conversion and the BatchNorm fold (the
    parity test builds its CPU oracle twin there)."""
    import torch
    import torchvision
    import bayesian_torch_b200 as bt
    torch.manual_seed(0)
    model = torchvision.models.resnet50()          # random init: no checkpoints offline
    # gamma of the last BatchNorm of every block: 0.5 keeps the random-init residual stack (weight noise sigma =
    # 0.5 |w| on top) in a regime where the logits are O(1) and the softmax is not saturated, so the predictive
    # mean / variance the step returns (and the checksum printed with it) mean something.  No effect on the work done.
    for name, mod in model.named_modules():
        if name.endswith("bn3"):
            mod.weight.data.fill_(0.5)
    params = {"prior_mu": 0.0, "prior_sigma": 1.0, "posterior_mu_init": 0.0, "posterior_rho_init": -3.0,
              "type": "Reparameterization", "moped_enable": True, "moped_delta": 0.5}
    bt.dnn_to_bnn(model, params)                   # MOPED: mu <- DNN init, rho <- get_rho (ImageNet examples force it)
    model = model.to(device).eval()
    if prepare is not None:
        prepare(model)
    # bf16 activations, channels-last; variational parameters stay fp32 masters (layers/_apply)
    model = model.to(memory_format=torch.channels_last).to(torch.bfloat16)
    if fold_bn:   # eval-mode BatchNorm / ReLU / residual join folded into the layer kernels (exact up to rounding)
        model = bt.fold_batchnorm(model, inplace=True)
    return model


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

    model = build_model(dev, fold_bn=not args.no_fold_bn)
    B, S = args.batch, args.samples
    g = torch.Generator().manual_seed(0)
    x_host = torch.randn(B, 3, 224, 224, generator=g).pin_memory()          # synthetic images, host resident
    x_dev = x_host.to(dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    bt.manual_seed(1234)

    def step_resident():
        return bt.mc_predict(model, x_dev, S, batch_samples=args.mc_batch)

    # end to end through the public API: every step copies its own batch from pinned host memory (bt.HostBatchPrefetcher:
    # the copy of the NEXT step's batch is issued right after this step's forwards are enqueued and overlaps them) and
    # reads the predictive mean / variance back to the host
    prefetch = bt.HostBatchPrefetcher(dev, dtype=torch.bfloat16, memory_format=torch.channels_last)

    def step_e2e():
        if prefetch._pending is None:          # first step of a timed run: nothing to overlap with, the copy is exposed
            prefetch.submit(x_host)
        xd = prefetch.take()
        res = bt.mc_predict(model, xd, S, batch_samples=args.mc_batch)
        prefetch.submit(x_host)                # next step's batch (same synthetic images): H2D on the side stream
        return res.mean.cpu(), res.var.cpu()

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
    launches = ops.PROFILE.launches
    clocks = sampler.stop() if rank == 0 else None

    # roofline of the dominant kernel (tc_fwd_kernel): CUDA events around every layer call of
    # one more timed step, on the launching stream
    ops.PROFILE.reset()
    ops.PROFILE.timing = True
    barrier()
    step_resident()
    torch.cuda.synchronize()
    ops.PROFILE.timing = False
    tc = [(f, b, e0.elapsed_time(e1)) for path, f, b, e0, e1 in ops.PROFILE.records if path == _lib.PATH_TC_BF16]
    other = [(f, b, e0.elapsed_time(e1)) for path, f, b, e0, e1 in ops.PROFILE.records if path != _lib.PATH_TC_BF16]
    tc_ms, tc_flops = sum(t for _, _, t in tc), sum(f for f, _, _ in tc)
    other_ms = sum(t for _, _, t in other)

    ms_e2e, _ = timed(step_e2e, args.steps)
    prefetch.take()                            # the batch submitted by the last step (one copy more than steps timed)

    if rank == 0:
        pk = peaks()
        achieved = tc_flops / (tc_ms * 1e-3) / 1e12 if tc_ms > 0 else 0.0
        value = B * 1e3 / ms_step
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": "resnet50_dnn_to_bnn_reparam_mc_inference", "batch": B, "samples": S, "image": "3x224x224",
                       "samples_per_gpu": S / world, "mc_batch": args.mc_batch, "bn_folded": not args.no_fold_bn, "parallelism": f"mc_samples_sharded_x{world}",
                       "l2": "activations per step (GBs) far exceed the 126 MB L2; no flush needed",
                       "image_samples_per_s": value * S,
                       "tflops_bayesian_layers": value * S * GFLOP_PER_IMAGE_SAMPLE / 1e3},
            "clocks": clocks,
            "e2e": {"value": B * 1e3 / ms_e2e, "unit": UNIT, "h2d_bytes_per_step": x_host.numel() * 4,
                    "d2h_bytes_per_step": 2 * res.mean.numel() * 4},
            "gpu_launches": launches,
            "roofline": {"bound": "tensor", "kernel": "tc2_fwd_kernel (all launches of one step)", "achieved": achieved,
                         "peak": pk["tflops"], "unit": "TFLOP/s", "frac": achieved / pk["tflops"], "traffic": tc_traffic(len(tc)), "algorithmic_bytes_per_launch": (sum(b for _, b, _ in tc) / len(tc)) if tc else None,
                         "peak_source": pk["source"], "launches_per_step": len(tc), "ms_per_step_in_kernel": tc_ms,
                         "ms_per_step_other_bt_kernels": other_ms,
                         "share_of_step": tc_ms / ms_step if ms_step > 0 else None},
        }
        if not args.no_cpu_baseline and world == 1:   # the CPU baseline is a rank-0, N = 1 figure (torchrun pins OMP threads to 1)
            line["cpu_baseline"] = {k: v for k, v in cpu_reference_images_per_s(B, S, 16, repeats=5).items()
                                    if k != "seconds_per_forward"}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)

"""bt_kl_normal_multi_static: the allocation-free variant of the multi-tensor KL kernel (scratch in a file-scope
device array instead of a caller workspace; include/bt_b200.h).  No counterpart in the reference: the test pins it
to the existing entry point, bt_kl_normal_multi, through the C ABI, on two calls in a row (the array keeps the
previous call's sums, so the second call checks the clearing too).  Collected last on purpose."""
import ctypes as C

import pytest
import torch

from bayesian_torch_b200 import _lib

MAX_TENSORS = 192   # BT_KL_STATIC_MAX_TENSORS


def _arr(tensors):
    a = (C.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        a[i] = None if t is None else t.data_ptr()
    return a


def _tables(sizes, seed, dev, tensor_prior=False, grads=False):
    g = torch.Generator(device="cpu").manual_seed(seed)
    mk = lambda n, scale, shift: (torch.randn(n, generator=g) * scale + shift).to(dev)
    t = {"mu": [mk(n, 0.3, 0.0) for n in sizes], "rho": [mk(n, 0.5, -3.0) for n in sizes], "sizes": sizes}
    t["pmu"] = [mk(n, 0.1, 0.0) if tensor_prior and i % 2 == 0 else None for i, n in enumerate(sizes)]
    t["psig"] = [mk(n, 0.2, 0.0).abs_().add_(0.5 + 0.1 * i) if tensor_prior and i % 3 == 0 else None for i, n in enumerate(sizes)]
    t["grads"] = grads
    return t


def _outputs(t, dev):
    n = len(t["sizes"])
    o = {"each": torch.full((n,), float("nan"), device=dev), "total": torch.full((), float("nan"), device=dev)}
    o["dmu"] = [torch.full((k,), float("nan"), device=dev) for k in t["sizes"]] if t["grads"] else None
    o["drho"] = [torch.full((k,), float("nan"), device=dev) for k in t["sizes"]] if t["grads"] else None
    o["scale"] = torch.full((), 0.25, device=dev) if t["grads"] else None
    return o


def _call(t, o, static, ws=None):
    n = len(t["sizes"])
    numel = (C.c_int64 * n)(*t["sizes"])
    any_pt = any(p is not None for p in t["pmu"] + t["psig"])
    head = (n, _arr(t["mu"]), _arr(t["rho"]), numel, 0, 0.1, 0.7, _arr(t["pmu"]) if any_pt else None,
            _arr(t["psig"]) if any_pt else None, o["each"].data_ptr(), o["total"].data_ptr(),
            _arr(o["dmu"]) if o["dmu"] else None, _arr(o["drho"]) if o["drho"] else None,
            o["scale"].data_ptr() if o["scale"] is not None else None)
    st = torch.cuda.current_stream().cuda_stream
    if static:
        return _lib.lib().bt_kl_normal_multi_static(*head, st)
    return _lib.lib().bt_kl_normal_multi(*head, ws.data_ptr(), st)


def _same(a, b):
    torch.testing.assert_close(a["each"], b["each"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(a["total"], b["total"], rtol=1e-5, atol=1e-6)
    for key in ("dmu", "drho"):
        if a[key]:
            for x, y in zip(a[key], b[key]):
                torch.testing.assert_close(x, y, rtol=1e-5, atol=1e-7)


@pytest.mark.gpu
def test_static_scratch_kl_matches_the_workspace_entry_point_twice_in_a_row():
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        # two different tables: ragged sizes (tail chunks, an unaligned-length tensor, one past a chunk boundary)
        first = _tables([5, 8192, 8193, 70001, 1 << 18, 33], 11, dev)
        second = _tables([300000, 7, 12345, 8192 * 3], 12, dev, tensor_prior=True, grads=True)
        want, got = [], []
        for t in (first, second):
            ws = torch.empty(int(_lib.lib().bt_workspace_bytes(b"kl", None, len(t["sizes"]))), dtype=torch.uint8, device=dev)
            o = _outputs(t, dev)
            assert _call(t, o, static=False, ws=ws) == 0, _lib.lib().bt_last_error()
            want.append(o)
            got.append(_outputs(t, dev))
        torch.cuda.synchronize()
        # a warm-up call loads the module and looks the array's address up; everything the two checked calls touch
        # exists by now, so free device memory must not move across them
        warm = _outputs(first, dev)
        assert _call(first, warm, static=True) == 0, _lib.lib().bt_last_error()
        torch.cuda.synchronize()
        free0, _ = torch.cuda.mem_get_info(dev)
        launches0 = _lib.lib().bt_launch_count()
        assert _call(first, got[0], static=True) == 0, _lib.lib().bt_last_error()
        assert _call(second, got[1], static=True) == 0, _lib.lib().bt_last_error()   # back to back: the array still holds call 1's sums
        torch.cuda.synchronize()
        free1, _ = torch.cuda.mem_get_info(dev)
        assert _lib.lib().bt_launch_count() - launches0 == 2
        assert free1 >= free0, (free0, free1)
        _same(got[0], want[0])
        _same(got[1], want[1])
        _same(warm, want[0])
        assert torch.isfinite(got[1]["total"]) and float(got[1]["total"]) > 0.0


@pytest.mark.gpu
def test_static_scratch_kl_refuses_more_tensors_than_its_array_holds():
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        n = MAX_TENSORS + 1
        t = _tables([16] * n, 13, dev)
        o = _outputs(t, dev)
        rc = _call(t, o, static=True)
        assert rc == -1
        msg = _lib.lib().bt_last_error()
        assert b"bt_kl_normal_multi_static" in msg and str(MAX_TENSORS).encode() in msg
        torch.cuda.synchronize()
        assert bool(torch.isnan(o["total"]))          # nothing ran
        # the maximum itself is served
        t = _tables([16] * MAX_TENSORS, 14, dev)
        o, w = _outputs(t, dev), _outputs(t, dev)
        ws = torch.empty(int(_lib.lib().bt_workspace_bytes(b"kl", None, MAX_TENSORS)), dtype=torch.uint8, device=dev)
        assert _call(t, w, static=False, ws=ws) == 0
        assert _call(t, o, static=True) == 0, _lib.lib().bt_last_error()
        torch.cuda.synchronize()
        _same(o, w)

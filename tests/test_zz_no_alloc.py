"""smsgpu_small_cycles_scan_noalloc (include/smsgpu.h): the K6 scan on the caller's device buffers with its scratch in a
file-scope __device__ array -- no allocation inside the call.  Checked through the C ABI against the existing entry
point smsgpu_small_cycles_scan (which the other tests pin to the reference), bit for bit like every other test of the
scan, on two calls in a row that share the scratch array (the second with a shorter list and another x*, so leftovers
of the first would show).  This file is collected last on purpose."""
import numpy as np
import pytest

from sms_b200 import instances as I

pytestmark = pytest.mark.gpu

FIELDS = ("off", "node", "eid", "is_stay", "n_cross", "src", "violation")


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from sms_b200 import build
    build.build()
    return torch


def _eids_of(inst, cycles):
    """edge id of (cycle[i], cycle[(i + 1) % k]) from the instance's CSR"""
    src = np.repeat(np.arange(inst.n), np.diff(inst.rowptr))
    key = src.astype(np.int64) * inst.n + inst.col
    order = np.argsort(key)
    k = cycles.shape[1]
    q = cycles.astype(np.int64) * inst.n + np.roll(cycles, -1, axis=1)
    pos = np.searchsorted(key[order], q.ravel())
    assert np.array_equal(key[order][pos], q.ravel())
    return inst.eid[order][pos].reshape(-1, k).astype(np.int32)


class DeviceOut:
    def __init__(self, torch, k, cap):
        dev = "cuda:0"
        self.k, self.cap = k, cap
        self.node = torch.full((max(k * cap, 1),), -7, dtype=torch.int32, device=dev)
        self.eid = torch.full((max(k * cap, 1),), -7, dtype=torch.int32, device=dev)
        self.is_stay = torch.full((max(k * cap, 1),), 9, dtype=torch.uint8, device=dev)
        self.n_cross = torch.full((max(cap, 1),), -7, dtype=torch.int32, device=dev)
        self.src = torch.full((max(cap, 1),), -7, dtype=torch.int32, device=dev)
        self.violation = torch.full((max(cap, 1),), -7.0, dtype=torch.float64, device=dev)
        self.off = torch.full((cap + 1,), -7, dtype=torch.int64, device=dev)
        self.count = torch.full((1,), -7, dtype=torch.int64, device=dev)

    def ptrs(self):
        return [t.data_ptr() for t in (self.node, self.eid, self.is_stay, self.n_cross, self.src, self.violation, self.off, self.count)]

    def host(self, u):
        k = self.k
        return dict(off=self.off[:u + 1].cpu().numpy(), node=self.node[:k * u].cpu().numpy(), eid=self.eid[:k * u].cpu().numpy(),
                    is_stay=self.is_stay[:k * u].cpu().numpy(), n_cross=self.n_cross[:u].cpu().numpy(), src=self.src[:u].cpu().numpy(),
                    violation=self.violation[:u].cpu().numpy())


def _compare(got, ref, u, what):
    assert ref.ncuts == u, what
    for f in FIELDS:
        a, b = got[f], np.asarray(getattr(ref, f))[:len(got[f])]
        assert a.dtype == b.dtype or f == "is_stay", f"{what}: dtype of {f}"
        assert np.array_equal(a, b), f"{what}: {f} differs"  # bit-exact, fp64 violation included


@pytest.mark.parametrize("k", [3, 4])
def test_noalloc_scan_equals_the_existing_entry_point_twice_in_a_row(torch_cuda, k):
    torch = torch_cuda
    from sms_b200 import capi
    inst = I.random_graph(300, 9000, seed=2)
    ctx = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid)
    try:
        ctx.small_cycles_enumerate(True)
        cycles = ctx.small_cycles_get(k).copy()
        if k == 4:
            cycles = cycles[:200000].copy()
        assert len(cycles) > 5000
        eids = _eids_of(inst, cycles)
        d_cycles = torch.from_numpy(cycles.ravel().copy()).cuda()
        d_eids = torch.from_numpy(eids.ravel().copy()).cuda()
        bl = np.zeros(inst.n, np.uint8)
        # call 1: the whole list, LP-like x*; call 2: a shorter list (not a multiple of the tile) and iid x*, so that stale
        # codes / tile counts of call 1 sit behind the part of the scratch call 2 uses
        calls = [(len(cycles), I.xstar_lp_like(inst.m, seed=3)), (len(cycles) // 3 + 17, I.xstar_uniform(inst.m, seed=4)),
                 (len(cycles), I.xstar_half_integral(inst.m, seed=5))]
        free_before = None
        for idx, (T, x) in enumerate(calls):
            ctx.small_cycles_set(cycles[:T])
            ctx.set_round_input(x, bl)
            ref = ctx.small_cycles_scan(k)
            d_x = torch.from_numpy(np.ascontiguousarray(x, np.float64)).cuda()
            out = DeviceOut(torch, k, T)
            torch.cuda.synchronize()
            if idx > 0:
                free_before = torch.cuda.mem_get_info()[0]
            u = capi.small_cycles_scan_noalloc(k, T, d_cycles.data_ptr(), d_eids.data_ptr(), d_x.data_ptr(), T, *out.ptrs())
            if idx > 0:  # (the first call on a device may load the module lazily)
                assert torch.cuda.mem_get_info()[0] == free_before, "device memory was allocated inside the call"
            assert u == int(out.count.item()) and u > 0
            _compare(out.host(u), ref, u, f"k={k} call {idx}")
    finally:
        ctx.close()


def test_noalloc_scan_limits_and_truncation(torch_cuda):
    torch = torch_cuda
    from sms_b200 import capi
    inst = I.random_graph(80, 700, seed=4)
    ctx = capi.Context(inst.n, inst.rowptr, inst.col, inst.eid)
    try:
        ctx.small_cycles_enumerate(False)
        cycles = ctx.small_cycles_get(3).copy()
        T = len(cycles)
        eids = _eids_of(inst, cycles)
        x = I.xstar_lp_like(inst.m, seed=1)
        ctx.set_round_input(x, np.zeros(inst.n, np.uint8))
        ref = ctx.small_cycles_scan(3)
        assert ref.ncuts >= 4
        d_cycles, d_eids = torch.from_numpy(cycles.ravel().copy()).cuda(), torch.from_numpy(eids.ravel().copy()).cuda()
        d_x = torch.from_numpy(x).cuda()
        # beyond the static array: refused with a message, nothing launched
        out = DeviceOut(torch, 3, 4)
        with pytest.raises(capi.SmsGpuError, match="SMSGPU_NOALLOC_MAX_CYCLES"):
            capi.small_cycles_scan_noalloc(3, capi.NOALLOC_MAX_CYCLES + 1, d_cycles.data_ptr(), d_eids.data_ptr(), d_x.data_ptr(), 4, *out.ptrs())
        with pytest.raises(capi.SmsGpuError, match="k must be 3 or 4"):
            capi.small_cycles_scan_noalloc(5, T, d_cycles.data_ptr(), d_eids.data_ptr(), d_x.data_ptr(), 4, *out.ptrs())
        assert int(out.count.item()) == -7
        # fewer output slots than violated cycles: the first `cap` in list order, the count still complete
        cap = ref.ncuts // 2
        out = DeviceOut(torch, 3, cap + 3)  # 3 guard slots behind the capacity
        u = capi.small_cycles_scan_noalloc(3, T, d_cycles.data_ptr(), d_eids.data_ptr(), d_x.data_ptr(), cap, *out.ptrs())
        assert u == ref.ncuts
        got = out.host(cap)
        for f in ("node", "eid", "is_stay", "n_cross", "src", "violation"):
            assert np.array_equal(got[f], np.asarray(getattr(ref, f))[:len(got[f])]), f
        assert np.array_equal(got["off"], np.asarray(ref.off)[:cap + 1])
        assert (out.src[cap:] == -7).all().item() and (out.node[3 * cap:] == -7).all().item() and (out.off[cap + 1:] == -7).all().item()
        # empty list, and the asynchronous form (count read from the device afterwards)
        out = DeviceOut(torch, 3, 2)
        assert capi.small_cycles_scan_noalloc(3, 0, 0, 0, 0, 2, *out.ptrs()) == 0
        assert int(out.off[0].item()) == 0
        out = DeviceOut(torch, 3, T)
        s = torch.cuda.Stream()
        assert capi.small_cycles_scan_noalloc(3, T, d_cycles.data_ptr(), d_eids.data_ptr(), d_x.data_ptr(), T, *out.ptrs(),
                                              stream=s.cuda_stream, sync=False) is None
        s.synchronize()
        u = int(out.count.item())
        _compare(out.host(u), ref, u, "async")
    finally:
        ctx.close()

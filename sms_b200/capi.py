"""ctypes binding of libsmsgpu.so (include/smsgpu.h) -- the same stub a maintainer would write for any FFI.

This module never imports anything under oracle/ and has no CPU fallback: if the CUDA library is missing or
no device is present, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libsmsgpu.so")

# every symbol include/smsgpu.h declares (tests/test_abi.py checks the header against this list and the .so)
SYMBOLS = [
    "smsgpu_create", "smsgpu_destroy", "smsgpu_last_error", "smsgpu_abi_version", "smsgpu_default_params",
    "smsgpu_set_round_input", "smsgpu_input_buffers", "smsgpu_set_round_input_device", "smsgpu_set_sources", "smsgpu_separate",
    "smsgpu_separate_async", "smsgpu_fetch_cuts", "smsgpu_get_cuts", "smsgpu_get_stats", "smsgpu_get_labels",
    "smsgpu_get_src_stats", "smsgpu_small_cycles_set", "smsgpu_small_cycles_enumerate", "smsgpu_small_cycles_get", "smsgpu_small_cycles_scan", "smsgpu_small_cycles_scan_device", "smsgpu_get_small_cycle_cuts", "smsgpu_small_cycles_timing",
    "smsgpu_small_cycles_scan_noalloc",
    "smsgpu_nccl_unique_id", "smsgpu_comm_init", "smsgpu_comm_destroy",
]


class SmsGpuError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libsmsgpu error {code}: {msg}")
        self.code = code


class Params(C.Structure):
    _fields_ = [
        ("max_intermediates", C.c_int32),
        ("best_dist_init", C.c_double),
        ("dedup", C.c_int32),
        ("export_labels", C.c_int32),
        ("export_src_stats", C.c_int32),
        ("start_tier", C.c_int32),
        ("mode", C.c_int32),
        ("debug_flags", C.c_int32),
        ("min_efficacy", C.c_double),
        ("reserved", C.c_int32 * 6),
    ]


class CutsStruct(C.Structure):
    _fields_ = [
        ("ncuts", C.c_int64),
        ("nnodes", C.c_int64),
        ("off", C.POINTER(C.c_int64)),
        ("node", C.POINTER(C.c_int32)),
        ("eid", C.POINTER(C.c_int32)),
        ("is_stay", C.POINTER(C.c_uint8)),
        ("n_cross", C.POINTER(C.c_int32)),
        ("src", C.POINTER(C.c_int32)),
        ("violation", C.POINTER(C.c_double)),
    ]


class StatsStruct(C.Structure):
    _fields_ = [
        ("sources_run", C.c_int64),
        ("arcs_scanned", C.c_int64),
        ("settled", C.c_int64),
        ("raw_walks", C.c_int64),
        ("unique_cuts", C.c_int64),
        ("tier_sources", C.c_int64 * 4),
        ("ms_weights", C.c_double),
        ("ms_search", C.c_double),
        ("ms_dedup", C.c_double),
        ("ms_total_device", C.c_double),
        ("ms_h2d", C.c_double),
        ("ms_d2h", C.c_double),
        ("search_launches", C.c_int64),
        ("total_launches", C.c_int64),
        ("mirrored_sources", C.c_int64),
        ("mirrored_arcs", C.c_int64),
        ("ms_tier", C.c_double * 4),
        ("verbatim_sweeps", C.c_int64),
    ]


@dataclass
class Cuts:
    """Host copy of one round's result (layout documented at smsgpu_cuts in include/smsgpu.h)."""
    off: np.ndarray
    node: np.ndarray
    eid: np.ndarray
    is_stay: np.ndarray
    n_cross: np.ndarray
    src: np.ndarray
    violation: np.ndarray

    @property
    def ncuts(self) -> int:
        return int(self.off.shape[0] - 1)

    def walks(self):
        """[(src, nodes, stay)] in list order"""
        o = self.off
        return [(int(self.src[k]), tuple(self.node[o[k]:o[k + 1]].tolist()), tuple(self.is_stay[o[k]:o[k + 1]].tolist()))
                for k in range(self.ncuts)]


_lib = None


def load_library(path: str | None = None) -> C.CDLL:
    """dlopen libsmsgpu.so.  Works without a GPU (cudart is linked statically); calls need a device."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("SMSGPU_LIB") or LIB_PATH  # SMSGPU_LIB: e.g. the device-assert build libsmsgpu_dbg.so
    if not os.path.exists(p):
        raise FileNotFoundError(f"{p} is missing: build it with `python -m sms_b200.build` (there is no CPU fallback)")
    lib = C.CDLL(p)
    vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int
    lib.smsgpu_create.argtypes = [C.POINTER(vp), i64, i64, vp, vp, vp, i32]
    lib.smsgpu_destroy.argtypes = [vp]
    lib.smsgpu_destroy.restype = None
    lib.smsgpu_last_error.argtypes = [vp]
    lib.smsgpu_last_error.restype = C.c_char_p
    lib.smsgpu_default_params.argtypes = [C.POINTER(Params)]
    lib.smsgpu_default_params.restype = None
    lib.smsgpu_set_round_input.argtypes = [vp, vp, vp]
    lib.smsgpu_set_round_input_device.argtypes = [vp, vp, vp]
    lib.smsgpu_input_buffers.argtypes = [vp, C.POINTER(vp), C.POINTER(vp)]
    lib.smsgpu_set_sources.argtypes = [vp, vp, i64, i32, i32]
    lib.smsgpu_separate.argtypes = [vp, C.POINTER(Params)]
    lib.smsgpu_separate_async.argtypes = [vp, C.POINTER(Params)]
    lib.smsgpu_fetch_cuts.argtypes = [vp]
    lib.smsgpu_get_cuts.argtypes = [vp, C.POINTER(CutsStruct)]
    lib.smsgpu_get_stats.argtypes = [vp, C.POINTER(StatsStruct)]
    lib.smsgpu_get_labels.argtypes = [vp, C.POINTER(i64), C.POINTER(C.POINTER(i64)), C.POINTER(C.POINTER(C.c_int32)),
                                      C.POINTER(C.POINTER(C.c_double)), C.POINTER(C.POINTER(C.c_int32))]
    lib.smsgpu_get_src_stats.argtypes = [vp, C.POINTER(i64), C.POINTER(C.POINTER(i64)), C.POINTER(C.POINTER(i64)),
                                         C.POINTER(C.POINTER(C.c_int32))]
    lib.smsgpu_small_cycles_set.argtypes = [vp, i32, i64, vp]
    lib.smsgpu_small_cycles_enumerate.argtypes = [vp, i32, C.POINTER(i64), C.POINTER(i64)]
    lib.smsgpu_small_cycles_get.argtypes = [vp, i32, C.POINTER(i64), C.POINTER(C.POINTER(C.c_int32))]
    lib.smsgpu_small_cycles_scan.argtypes = [vp, i32]
    lib.smsgpu_small_cycles_scan_device.argtypes = [vp, i32]
    lib.smsgpu_get_small_cycle_cuts.argtypes = [vp, i32, C.POINTER(CutsStruct)]
    lib.smsgpu_small_cycles_timing.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.smsgpu_small_cycles_scan_noalloc.argtypes = [i32, vp, i32, i64, vp, vp, vp, i64, vp, vp, vp, vp, vp, vp, vp, vp,
                                                     C.POINTER(i64), C.c_char_p, C.c_size_t]
    lib.smsgpu_nccl_unique_id.argtypes = [vp]
    lib.smsgpu_comm_init.argtypes = [vp, vp, i32, i32]
    lib.smsgpu_comm_destroy.argtypes = [vp]
    if path is None:
        _lib = lib
    return lib


def _arr(ptr, n, dtype):
    if n == 0:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True)


def _cuts_from_struct(cs: CutsStruct) -> Cuts:
    U, N = int(cs.ncuts), int(cs.nnodes)
    return Cuts(off=_arr(cs.off, U + 1, np.int64), node=_arr(cs.node, N, np.int32), eid=_arr(cs.eid, N, np.int32),
                is_stay=_arr(cs.is_stay, N, np.uint8), n_cross=_arr(cs.n_cross, U, np.int32), src=_arr(cs.src, U, np.int32),
                violation=_arr(cs.violation, U, np.float64))


NOALLOC_MAX_CYCLES = 4194304  # SMSGPU_NOALLOC_MAX_CYCLES


def small_cycles_scan_noalloc(k: int, T: int, d_cycles: int, d_cycle_eids: int, d_x: int, out_capacity: int, d_node: int, d_eid: int,
                              d_is_stay: int, d_n_cross: int, d_src: int, d_violation: int, d_off: int, d_count: int,
                              device: int = 0, stream: int = 0, sync: bool = True, lib=None):
    """smsgpu_small_cycles_scan_noalloc: the K6 scan on the caller's DEVICE buffers (raw addresses), no allocation inside.
    Returns the number of violated cycles when `sync`, else None (the count is in *d_count once the stream gets there)."""
    lib = lib or load_library()
    err = C.create_string_buffer(512)
    cnt = C.c_int64(-1)
    rc = lib.smsgpu_small_cycles_scan_noalloc(device, stream or None, k, T, d_cycles or None, d_cycle_eids or None, d_x or None,
                                              out_capacity, d_node or None, d_eid or None, d_is_stay or None, d_n_cross or None,
                                              d_src or None, d_violation or None, d_off or None, d_count or None,
                                              C.byref(cnt) if sync else None, err, len(err))
    if rc != 0:
        raise SmsGpuError(rc, err.value.decode(errors="replace"))
    return int(cnt.value) if sync else None


class Context:
    """One smsgpu_ctx: a graph replica on one GPU."""

    def __init__(self, n: int, rowptr, col, eid, device: int = 0):
        self.lib = load_library()
        self.n = int(n)
        self._rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
        self._col = np.ascontiguousarray(col, dtype=np.int32)
        self._eid = np.ascontiguousarray(eid, dtype=np.int32)
        self.m = int(self._col.shape[0] // 2)
        h = C.c_void_p()
        rc = self.lib.smsgpu_create(C.byref(h), self.n, self.m, self._rowptr.ctypes.data, self._col.ctypes.data,
                                    self._eid.ctypes.data, int(device))
        if rc != 0:
            raise SmsGpuError(rc, self.lib.smsgpu_last_error(None).decode())
        self.h = h
        self.local_source_count = self.n

    def close(self):
        if getattr(self, "h", None):
            self.lib.smsgpu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int):
        if rc != 0:
            raise SmsGpuError(rc, self.lib.smsgpu_last_error(self.h).decode())

    # ---- round input ----
    def set_round_input(self, x_by_eid, blacklist):
        x = np.ascontiguousarray(x_by_eid, dtype=np.float64)
        bl = np.ascontiguousarray(blacklist, dtype=np.uint8)
        if x.shape[0] != self.m or bl.shape[0] != self.n:
            raise ValueError("x_by_eid must have m entries and blacklist n")
        self._check(self.lib.smsgpu_set_round_input(self.h, x.ctypes.data, bl.ctypes.data))

    def input_buffers(self):
        """(x[m] float64, blacklist[n] uint8) numpy views of the context's pinned staging buffers: fill them in place and
        pass them to set_round_input (no pageable -> pinned copy)."""
        px, pb = C.c_void_p(), C.c_void_p()
        self._check(self.lib.smsgpu_input_buffers(self.h, C.byref(px), C.byref(pb)))
        x = np.ctypeslib.as_array(C.cast(px, C.POINTER(C.c_double)), shape=(max(self.m, 1),))[:self.m]
        bl = np.ctypeslib.as_array(C.cast(pb, C.POINTER(C.c_uint8)), shape=(self.n,))
        return x, bl

    def set_round_input_device(self, x_ptr: int, bl_ptr: int):
        self._check(self.lib.smsgpu_set_round_input_device(self.h, C.c_void_p(x_ptr), C.c_void_p(bl_ptr)))

    def set_sources(self, sources=None, rank: int = 0, nranks: int = 1):
        if sources is None:
            self._check(self.lib.smsgpu_set_sources(self.h, None, 0, rank, nranks))
            g = self.n
        else:
            s = np.ascontiguousarray(sources, dtype=np.int32)
            self._check(self.lib.smsgpu_set_sources(self.h, s.ctypes.data, s.shape[0], rank, nranks))
            g = int(s.shape[0])
        from .sharding import local_count
        self.local_source_count = local_count(g, rank, nranks)

    # ---- a round ----
    def params(self, **kw) -> Params:
        p = Params()
        self.lib.smsgpu_default_params(C.byref(p))
        for k, v in kw.items():
            if not hasattr(p, k):
                raise TypeError(f"unknown parameter {k}")
            setattr(p, k, v)
        return p

    def separate(self, **kw) -> Cuts:
        p = self.params(**kw)
        self._check(self.lib.smsgpu_separate(self.h, C.byref(p)))
        return self.get_cuts()

    def separate_async(self, **kw):
        p = self.params(**kw)
        self._check(self.lib.smsgpu_separate_async(self.h, C.byref(p)))

    def separate_nocopy(self, p: Params):
        """One blocking round through the C ABI without building numpy copies (bench e2e leg)."""
        self._check(self.lib.smsgpu_separate(self.h, C.byref(p)))

    def fetch(self):
        self._check(self.lib.smsgpu_fetch_cuts(self.h))

    def get_cuts(self) -> Cuts:
        cs = CutsStruct()
        self._check(self.lib.smsgpu_get_cuts(self.h, C.byref(cs)))
        return _cuts_from_struct(cs)

    def cut_counts(self):
        cs = CutsStruct()
        self._check(self.lib.smsgpu_get_cuts(self.h, C.byref(cs)))
        return int(cs.ncuts), int(cs.nnodes)

    def stats(self) -> dict:
        s = StatsStruct()
        self._check(self.lib.smsgpu_get_stats(self.h, C.byref(s)))
        d = {k: getattr(s, k) for k, _ in StatsStruct._fields_ if k not in ("tier_sources", "ms_tier")}
        d["tier_sources"] = list(s.tier_sources)
        d["ms_tier"] = [round(v, 4) for v in s.ms_tier]
        return d

    def labels(self):
        """(label_off[nsrc+1], node, dist, hops) of the last round run with export_labels=1"""
        n = C.c_int64()
        po = C.POINTER(C.c_int64)()
        pn = C.POINTER(C.c_int32)()
        pd = C.POINTER(C.c_double)()
        ph = C.POINTER(C.c_int32)()
        self._check(self.lib.smsgpu_get_labels(self.h, C.byref(n), C.byref(po), C.byref(pn), C.byref(pd), C.byref(ph)))
        L = int(n.value)
        nsrc = self.local_source_count
        return (_arr(po, nsrc + 1, np.int64), _arr(pn, L, np.int32), _arr(pd, L, np.float64), _arr(ph, L, np.int32))

    def src_stats(self):
        n = C.c_int64()
        pa = C.POINTER(C.c_int64)()
        ps = C.POINTER(C.c_int64)()
        pt = C.POINTER(C.c_int32)()
        self._check(self.lib.smsgpu_get_src_stats(self.h, C.byref(n), C.byref(pa), C.byref(ps), C.byref(pt)))
        k = int(n.value)
        return _arr(pa, k, np.int64), _arr(ps, k, np.int64), _arr(pt, k, np.int32)

    # ---- small cycles ----
    def small_cycles_set(self, cycles):
        c = np.ascontiguousarray(cycles, dtype=np.int32)
        k = c.shape[1] if c.ndim == 2 else 3
        self._check(self.lib.smsgpu_small_cycles_set(self.h, k, c.shape[0] if c.ndim == 2 else 0, c.ctypes.data))
        return k

    def small_cycles_enumerate(self, want_holes: bool = True):
        """computeSmallCCs on the device; returns (number of triangles, number of four-holes); both lists are registered"""
        nt, nh = C.c_int64(), C.c_int64()
        self._check(self.lib.smsgpu_small_cycles_enumerate(self.h, 1 if want_holes else 0, C.byref(nt), C.byref(nh)))
        return int(nt.value), int(nh.value)

    def small_cycles_get(self, k: int) -> np.ndarray:
        n = C.c_int64()
        p = C.POINTER(C.c_int32)()
        self._check(self.lib.smsgpu_small_cycles_get(self.h, k, C.byref(n), C.byref(p)))
        return _arr(p, int(n.value) * k, np.int32).reshape(-1, k)

    def small_cycles_scan(self, k: int) -> Cuts:
        self._check(self.lib.smsgpu_small_cycles_scan(self.h, k))
        cs = CutsStruct()
        self._check(self.lib.smsgpu_get_small_cycle_cuts(self.h, k, C.byref(cs)))
        return _cuts_from_struct(cs)

    def small_cycles_scan_nocopy(self, k: int):
        """the scan through the C ABI without building numpy copies (bench); returns (violated cycles, walk nodes)"""
        self._check(self.lib.smsgpu_small_cycles_scan(self.h, k))
        cs = CutsStruct()
        self._check(self.lib.smsgpu_get_small_cycle_cuts(self.h, k, C.byref(cs)))
        return int(cs.ncuts), int(cs.nnodes)

    def small_cycles_timing(self):
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        self._check(self.lib.smsgpu_small_cycles_timing(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return {"ms_scan": a.value, "ms_gather": b.value, "ms_d2h": c.value}

    # ---- multi GPU ----
    def nccl_unique_id(self) -> bytes:
        buf = C.create_string_buffer(128)
        rc = self.lib.smsgpu_nccl_unique_id(buf)
        if rc != 0:
            raise SmsGpuError(rc, self.lib.smsgpu_last_error(None).decode())
        return buf.raw

    def comm_init(self, uid: bytes, rank: int, nranks: int):
        buf = C.create_string_buffer(uid, 128)
        self._check(self.lib.smsgpu_comm_init(self.h, buf, rank, nranks))

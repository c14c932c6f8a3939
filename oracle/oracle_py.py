"""TEST INFRASTRUCTURE ONLY -- ctypes bindings of the two CPU checkers.

  * ``orc``: oracle/liborc.so, the restatement (oracle/ocw_oracle.hpp)
  * ``ref``: oracle/_ref/libsmsref.so, the UNMODIFIED reference compiled from /root/reference

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module; the product package sms_b200 never does (tests/test_no_oracle_in_product.py enforces it).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORC_PATH = os.path.join(_HERE, "liborc.so")
REF_PATH = os.path.join(_HERE, "_ref", "libsmsref.so")

_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


def build(ref: bool = True) -> None:
    """Compile the checkers (oracle/Makefile).  The reference part needs /root/reference (CPU container only)."""
    subprocess.run(["make", "-C", _HERE, "orc"], check=True, capture_output=True)
    if ref and os.path.isdir("/root/reference/src"):
        subprocess.run(["make", "-C", _HERE, "-j8", "ref"], check=True, capture_output=True)


@dataclass
class Sweep:
    cut_off: np.ndarray
    cut_src: np.ndarray
    node: np.ndarray
    stay: np.ndarray
    label_off: np.ndarray | None
    label_node: np.ndarray | None
    label_dist: np.ndarray | None
    label_hops: np.ndarray | None
    stats: np.ndarray  # [nsrc, 6]: arcs, settled, labelled, heapMax, intermediates, maxWalkLen
    seconds: float
    n_walks: int
    total_len: int
    timing: tuple = ()  # (set-up s, sum of per-source s, sum of squares, max per-source s, sources timed)

    @property
    def ncuts(self) -> int:
        return int(self.cut_src.shape[0])

    def walks(self):
        """list of (src_index, nodes tuple, stay tuple) in emission order"""
        out = []
        for k in range(self.ncuts):
            a, b = int(self.cut_off[k]), int(self.cut_off[k + 1])
            out.append((int(self.cut_src[k]), tuple(self.node[a:b].tolist()), tuple(self.stay[a:b].tolist())))
        return out


class _Lib:
    def __init__(self, path: str, prefix: str):
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing -- run `make -C oracle` (the reference part only builds where /root/reference exists)")
        self.lib = C.CDLL(path)
        self.p = prefix
        f = self._f
        f("csr").argtypes = [C.c_int64, C.c_int64, _i32p, _i32p, _i64p, _i32p, _i32p]
        f("csr").restype = None
        f("sweep").argtypes = [C.c_int64, C.c_int64, _i32p, _i32p, _f64p, _u8p, _i32p, C.c_int64, C.c_int, C.c_int, C.c_int]
        f("sweep").restype = C.c_void_p
        f("sizes").argtypes = [C.c_void_p, _i64p]
        f("seconds").argtypes = [C.c_void_p]
        f("seconds").restype = C.c_double
        f("timing").argtypes = [C.c_void_p, _f64p]
        f("copy_cuts").argtypes = [C.c_void_p, _i64p, _i32p, _i32p, _u8p]
        f("copy_labels").argtypes = [C.c_void_p, _i64p, _i32p, _f64p, _i32p]
        f("copy_stats").argtypes = [C.c_void_p, _i64p]
        f("free").argtypes = [C.c_void_p]
        if prefix == "orc":
            f("sweep_fixpoint").argtypes = [C.c_int64, C.c_int64, _i32p, _i32p, _f64p, _u8p, _i32p, C.c_int64, C.c_int, C.c_int, _i64p]
            f("sweep_fixpoint").restype = C.c_void_p
        for name in ("simple_odd_cycles", "extract_cycles"):
            f(name).argtypes = [C.c_int64, C.c_int64, _i32p, _i32p, _f64p, _i32p, _u8p, C.c_int64]
            f(name).restype = C.c_void_p
        f("most_violated").argtypes = [C.c_int64, C.c_int64, _i32p, _i32p, _f64p, C.c_int64]
        f("most_violated").restype = C.c_void_p
        f("odd_selection").argtypes = [_i32p, _f64p, C.c_int64, _u8p]
        f("odd_selection").restype = C.c_int
        f("small_ccs").argtypes = [C.c_int64, C.c_int64, _i32p, _i32p, _i64p, C.c_void_p, C.c_void_p]
        f("small_cycle_scan").argtypes = [C.c_int64, C.c_int64, _i32p, _i32p, _f64p, C.c_int, _i32p, C.c_int64]
        f("small_cycle_scan").restype = C.c_void_p

    def _f(self, name):
        return getattr(self.lib, f"{self.p}_{name}")

    # ------------------------------------------------------------------------------------------
    def csr(self, n, eu, ev):
        m = len(eu)
        rowptr = np.zeros(n + 1, np.int64)
        col = np.zeros(2 * m, np.int32)
        eid = np.zeros(2 * m, np.int32)
        self._f("csr")(n, m, _c(eu, np.int32), _c(ev, np.int32), rowptr, col, eid)
        return rowptr, col, eid

    def _unpack(self, h, nsrc, want_labels):
        sz = np.zeros(8, np.int64)
        self._f("sizes")(h, sz)
        ncuts, total, nlab, nstat = (int(v) for v in sz[:4])
        cut_off = np.zeros(ncuts + 1, np.int64)
        cut_src = np.zeros(ncuts, np.int32)
        node = np.zeros(total, np.int32)
        stay = np.zeros(total, np.uint8)
        self._f("copy_cuts")(h, cut_off, cut_src, node, stay)
        lo = ln = ld = lh = None
        if want_labels:
            lo = np.zeros(nsrc + 1, np.int64)
            ln = np.zeros(nlab, np.int32)
            ld = np.zeros(nlab, np.float64)
            lh = np.zeros(nlab, np.int32)
            self._f("copy_labels")(h, lo, ln, ld, lh)
        stats = np.zeros(nstat * 6, np.int64)
        if nstat:
            self._f("copy_stats")(h, stats)
        secs = float(self._f("seconds")(h))
        tm = np.zeros(5, np.float64)
        self._f("timing")(h, tm)
        self._f("free")(h)
        return Sweep(cut_off, cut_src, node, stay, lo, ln, ld, lh, stats.reshape(-1, 6), secs, int(sz[4]), int(sz[5]),
                     tuple(float(v) for v in tm))

    def sweep(self, inst, x, blacklist=None, sources=None, mode="I", want_labels=False, want_cuts=True,
              canonical=False) -> Sweep:
        """mode "I": blacklist re-initialised before every source; "V": verbatim sequential sweep.
        canonical=True (restatement only): canonical tie-breaking, the contract of the CUDA path."""
        if canonical and self.p != "orc":
            raise ValueError("the unmodified reference has no canonical tie mode")
        bl = inst.root_blacklist() if blacklist is None else _c(blacklist, np.uint8)
        src = inst.handler_sources() if sources is None else _c(sources, np.int32)
        h = self._f("sweep")(inst.n, inst.m, inst.eu, inst.ev, _c(x, np.float64), bl, src, len(src),
                             {"I": 0, "V": 1}[mode] | (2 if canonical else 0), int(want_labels), int(want_cuts))
        return self._unpack(h, len(src), want_labels)

    def sweep_fixpoint(self, inst, x, blacklist=None, sources=None, canonical=True, max_sweeps=10000):
        """Mode V as the fixed point of independent sweeps over a time-stamped blacklist (restatement only): the CPU
        statement of SMSGPU_MODE_VERBATIM.  Returns (Sweep, info) with info = dict(sweeps, changed, ran_last, blacklisted)."""
        bl = inst.root_blacklist() if blacklist is None else _c(blacklist, np.uint8)
        src = inst.handler_sources() if sources is None else _c(sources, np.int32)
        info = np.zeros(4, np.int64)
        h = self._f("sweep_fixpoint")(inst.n, inst.m, inst.eu, inst.ev, _c(x, np.float64), bl, src, len(src),
                                      int(canonical), int(max_sweeps), info)
        return self._unpack(h, len(src), False), dict(sweeps=int(info[0]), changed=int(info[1]), ran_last=int(info[2]),
                                                       blacklisted=int(info[3]))

    def _walk_call(self, name, inst, x, nodes, stay):
        h = self._f(name)(inst.n, inst.m, inst.eu, inst.ev, _c(x, np.float64), _c(nodes, np.int32),
                          _c(stay, np.uint8), len(nodes))
        return self._unpack(h, 0, False).walks()

    def simple_odd_cycles(self, inst, x, nodes, stay):
        return self._walk_call("simple_odd_cycles", inst, x, nodes, stay)

    def extract_cycles(self, inst, x, nodes, stay):
        return self._walk_call("extract_cycles", inst, x, nodes, stay)

    def most_violated(self, inst, x, source):
        h = self._f("most_violated")(inst.n, inst.m, inst.eu, inst.ev, _c(x, np.float64), source)
        return self._unpack(h, 0, False).walks()

    def odd_selection(self, nodes, weights):
        out = np.zeros(len(nodes), np.uint8)
        ok = self._f("odd_selection")(_c(nodes, np.int32), _c(weights, np.float64), len(nodes), out)
        return out if ok else None

    def small_ccs(self, inst):
        cnt = np.zeros(2, np.int64)
        self._f("small_ccs")(inst.n, inst.m, inst.eu, inst.ev, cnt, None, None)
        tri = np.zeros(3 * int(cnt[0]), np.int32)
        holes = np.zeros(4 * int(cnt[1]), np.int32)
        self._f("small_ccs")(inst.n, inst.m, inst.eu, inst.ev, cnt, tri.ctypes.data_as(C.c_void_p),
                             holes.ctypes.data_as(C.c_void_p))
        return tri.reshape(-1, 3), holes.reshape(-1, 4)

    def small_cycle_scan(self, inst, x, cycles):
        cycles = _c(cycles, np.int32)
        K = cycles.shape[1]
        h = self._f("small_cycle_scan")(inst.n, inst.m, inst.eu, inst.ev, _c(x, np.float64), K, cycles.reshape(-1),
                                        cycles.shape[0])
        return self._unpack(h, 0, False).walks()

    def small_cycle_scan_timed(self, inst, x, cycles):
        """(number of violated cycles, set-up seconds = the m updateWeights of a round, scan seconds)"""
        cycles = _c(cycles, np.int32)
        K = cycles.shape[1]
        h = self._f("small_cycle_scan")(inst.n, inst.m, inst.eu, inst.ev, _c(x, np.float64), K, cycles.reshape(-1), cycles.shape[0])
        s = self._unpack(h, 0, False)
        return s.ncuts, s.timing[0], s.timing[1]


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


_cache = {}


def orc() -> _Lib:
    if "orc" not in _cache:
        if not os.path.exists(ORC_PATH):
            build(ref=False)
        _cache["orc"] = _Lib(ORC_PATH, "orc")
    return _cache["orc"]


def ref() -> _Lib:
    if "ref" not in _cache:
        _cache["ref"] = _Lib(REF_PATH, "ref")
    return _cache["ref"]


def have_ref() -> bool:
    return os.path.exists(REF_PATH)

"""ctypes binding of the CPU oracle (oracle/libcurvix_oracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package (3d_curvix_b200/) never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


class OrcConfig(C.Structure):
    # same field order as orc_config (oracle/curvix_oracle.hpp) and cvx_config (include/curvix_b200.h)
    _fields_ = [
        ("delta", C.c_double), ("delta_theta_deg", C.c_double), ("parallel_angle_deg", C.c_double),
        ("lowes_ratio", C.c_double), ("epip_tangency_displ_thresh", C.c_double),
        ("location_perturbation", C.c_double), ("orient_perturbation", C.c_double),
        ("cluster_dist_thresh", C.c_double), ("cluster_orient_thresh_deg", C.c_double),
        ("cluster_gauss_sigma", C.c_double), ("pi_truncated", C.c_double),
        ("n_views", C.c_int32), ("img_rows", C.c_int32), ("img_cols", C.c_int32),
        ("max_support_views", C.c_int32), ("max_cluster_size", C.c_int32), ("reserved", C.c_int32),
    ]


class OrcCounts(C.Structure):
    _fields_ = [(n, C.c_int64) for n in (
        "n_h1_total", "n_h1_processed", "n_candidates", "n_hypotheses", "n_validations", "n_valid_rows",
        "n_pairs", "n_degenerate_wedges", "n_supported", "n_parallel", "n_inliers", "n_unique_hypotheses")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


def build(fast=False, quiet=True):
    """Compile the oracle with oracle/Makefile (g++, no reference sources involved)."""
    target = "libcurvix_oracle_fast.so" if fast else "libcurvix_oracle.so"
    subprocess.run(["make", "-C", _HERE, target], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)
    return os.path.join(_HERE, target)


def _ptr(a, t):
    if a is None:
        return None
    return a.ctypes.data_as(C.POINTER(t))


_d = lambda a: _ptr(a, C.c_double)
_i = lambda a: _ptr(a, C.c_int32)
_l = lambda a: _ptr(a, C.c_int64)


def load(fast=False):
    path = os.path.join(_HERE, "libcurvix_oracle_fast.so" if fast else "libcurvix_oracle.so")
    if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(os.path.join(_HERE, "curvix_oracle.cpp")):
        build(fast)
    lib = C.CDLL(path)
    lib.orc_create.restype = C.c_void_p
    lib.orc_create.argtypes = [C.POINTER(OrcConfig)]
    for name in ("orc_destroy", "orc_get_counts"):
        getattr(lib, name).restype = None
    return lib


def default_config(lib=None, **kw):
    lib = lib or load()
    cfg = OrcConfig()
    lib.orc_default_config(C.byref(cfg))
    for k, v in kw.items():
        setattr(cfg, k, v)
    return cfg


class Oracle:
    """Thin object wrapper; mirrors the C-ABI call sequence of the product library."""

    def __init__(self, cfg, fast=False):
        self.lib = load(fast)
        self.cfg = cfg
        self.h = C.c_void_p(self.lib.orc_create(C.byref(cfg)))
        self.V = cfg.n_views - 2
        self.n_edges = [0] * cfg.n_views

    def close(self):
        if self.h:
            self.lib.orc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_cameras(self, R, T, K):
        R = np.ascontiguousarray(R, np.float64).reshape(-1, 9)
        T = np.ascontiguousarray(T, np.float64).reshape(-1, 3)
        K = np.ascontiguousarray(K, np.float64)
        per_view = 1 if K.size == 9 * R.shape[0] and R.shape[0] > 1 else 0
        rc = self.lib.orc_set_cameras(self.h, R.shape[0], _d(R), _d(T), _d(K), per_view)
        assert rc == 0
        self._keep = (R, T, K)

    def set_view_edges(self, v, edges):
        e = np.asarray(edges, np.float64)
        x, y, th = (np.ascontiguousarray(e[:, k]) for k in range(3))
        s = np.ascontiguousarray(e[:, 3]) if e.shape[1] > 3 else np.ones(len(e))
        rc = self.lib.orc_set_view_edges(self.h, v, len(e), _d(x), _d(y), _d(th), _d(s))
        assert rc == 0
        self.n_edges[v] = len(e)

    def set_pair(self, h1, h2):
        rc = self.lib.orc_set_hypothesis_pair(self.h, h1, h2)
        assert rc == 0
        self.h1, self.h2 = h1, h2

    def run(self, h1_begin=0, h1_end=-1, mode=1, threads=0, keep=False):
        if threads <= 0:
            threads = os.cpu_count() or 1
        rc = self.lib.orc_run_edge_sketch(self.h, h1_begin, h1_end, mode, threads, 1 if keep else 0)
        assert rc == 0
        self.keep = keep
        return self.counts()

    def counts(self):
        c = OrcCounts()
        self.lib.orc_get_counts(self.h, C.byref(c))
        return c.as_dict()

    def h2_angles(self):
        out = np.empty(self.n_edges[self.h2])
        self.lib.orc_get_h2_angles(self.h, _d(out))
        return out

    def h1_wedges(self):
        E = self.n_edges[self.h1]
        w = np.empty((E, 2)); f = np.empty(E, np.int32)
        self.lib.orc_get_h1_wedges(self.h, _d(w), _i(f))
        return w, f

    def h1_table(self):
        E = self.n_edges[self.h1]
        m = np.empty(E, np.int32); n = np.empty(E, np.int32); off = np.empty(E, np.int64); fm = np.empty(E)
        self.lib.orc_get_h1_table(self.h, _i(m), _i(n), _l(off), _d(fm))
        return dict(m=m, n=n, off=off, form_margin=fm)

    def rows(self):
        H = self.counts()["n_hypotheses"]
        ints = {k: np.empty(H, np.int32) for k in ("h1", "h2_orig", "local_idx", "raw_at_local", "label", "h2_rep", "count")}
        dbl = {k: np.empty(H) for k in ("xc", "yc", "theta_orig", "xbar", "ybar", "thbar")}
        self.lib.orc_get_rows(self.h, _i(ints["h1"]), _i(ints["h2_orig"]), _i(ints["local_idx"]), _i(ints["raw_at_local"]),
                              _d(dbl["xc"]), _d(dbl["yc"]), _d(dbl["theta_orig"]), _d(dbl["xbar"]), _d(dbl["ybar"]),
                              _d(dbl["thbar"]), _i(ints["label"]), _i(ints["h2_rep"]), _i(ints["count"]))
        ints.update(dbl)
        return ints

    def support_matrix(self):
        H = self.counts()["n_hypotheses"]
        out = np.empty((H, self.V), np.int32)
        self.lib.orc_get_support_matrix(self.h, _i(out))
        return out

    def intermediates(self):
        c = self.counts()
        H, G = c["n_hypotheses"], c["n_h1_processed"]
        w31 = np.empty((G, self.V, 2)); w32 = np.empty((H, self.V, 2)); tgt = np.empty((H, self.V, 2))
        mg = np.empty((H, self.V)); wf = np.empty((H, self.V), np.int32)
        rc = self.lib.orc_get_intermediates(self.h, _d(w31), _d(w32), _d(tgt), _d(mg), _i(wf))
        assert rc == 0, "run(keep=True) first"
        return dict(wedge31=w31, wedge32=w32, tgt=tgt, margin=mg, wflag=wf)

    def valid_rows(self):
        n = self.counts()["n_valid_rows"]
        a = {k: np.empty(n, np.int32) for k in ("h1", "row", "count", "selected")}
        self.lib.orc_get_valid_rows(self.h, _i(a["h1"]), _i(a["row"]), _i(a["count"]), _i(a["selected"]))
        return a

    def pairs(self):
        P = self.counts()["n_pairs"]
        a = {k: np.empty(P, np.int32) for k in ("h1", "h2rep", "row", "count")}
        sup = np.empty((P, self.V), np.int32)
        self.lib.orc_get_pairs(self.h, _i(a["h1"]), _i(a["h2rep"]), _i(a["row"]), _i(a["count"]), _i(sup))
        a["supports"] = sup
        return a

    def finalize(self):
        P = self.counts()["n_pairs"]
        gc = np.empty((P, 3)); gw = np.empty((P, 3)); tw = np.empty((P, 3)); h2c = np.empty((P, 2)); nf = np.empty(P, np.int32)
        self.lib.orc_finalize(self.h, _d(gc), _d(gw), _d(tw), _d(h2c), _i(nf))
        return dict(gamma_cam=gc, gamma_world=gw, tangent_world=tw, h2_corrected=h2c, nan_flag=nf)

    def accumulate_claims(self, masks):
        """masks: list of n_views uint8 arrays (edge count of each view); OR-ed in place.  Returns #pairs that inserted."""
        arr = (C.c_void_p * len(masks))(*[m.ctypes.data for m in masks])
        return self.lib.orc_accumulate_claims(self.h, arr)

    def pair_cluster(self, p, cap=4096):
        buf = np.empty(cap, np.int32)
        n = self.lib.orc_get_pair_cluster(self.h, p, _i(buf), cap)
        return buf[:min(n, cap)].copy()


def select_next_views(claimed, total, history, stop_ratio, fast=False):
    """EdgeSketch_Core.cpp:799-868 -> (best1, best2, avg_ratio, abort)."""
    lib = load(fast)
    claimed = np.ascontiguousarray(claimed, np.int64); total = np.ascontiguousarray(total, np.int64)
    hist = np.ascontiguousarray(history, np.int32)
    b1, b2, avg = C.c_int32(), C.c_int32(), C.c_double()
    lib.orc_select_next_views.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p]
    ab = lib.orc_select_next_views(len(claimed), claimed.ctypes.data, total.ctypes.data, hist.ctypes.data, len(hist), float(stop_ratio),
                                   C.addressof(b1), C.addressof(b2), C.addressof(avg))
    return b1.value, b2.value, avg.value, bool(ab)


def support_test(group_off, view_off, ang31, ang32, cs, sn, wedge31, wedge32, tgt, thr, parallel_deg, threads=0, fast=False):
    lib = load(fast)
    group_off = np.ascontiguousarray(group_off, np.int64); view_off = np.ascontiguousarray(view_off, np.int64)
    G = len(group_off) - 1; V = len(view_off) - 1; H = int(group_off[-1])
    out = np.empty((V, H), np.int32)
    lib.orc_support_test.argtypes = [C.c_int, C.POINTER(C.c_int64), C.c_int, C.POINTER(C.c_int64)] + [C.POINTER(C.c_double)] * 7 + \
        [C.c_double, C.c_double, C.POINTER(C.c_int32), C.c_int]
    lib.orc_support_test(G, _l(group_off), V, _l(view_off), _d(ang31), _d(ang32), _d(cs), _d(sn), _d(wedge31), _d(wedge32),
                         _d(tgt), thr, parallel_deg, _i(out), threads or (os.cpu_count() or 1))
    return out

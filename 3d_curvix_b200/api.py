"""ctypes binding of libcurvix_b200.so (the C-ABI in include/curvix_b200.h).

There is no CPU path: if the CUDA library is missing or no device is present this module raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CVX_LIB_PATH") or os.path.join(_HERE, "libcurvix_b200.so")   # env override: kernel-variant experiments

# every symbol include/curvix_b200.h declares
SYMBOLS = [
    "cvx_default_config", "cvx_create", "cvx_destroy", "cvx_last_error", "cvx_set_cameras", "cvx_set_view_edges", "cvx_set_all_view_edges",
    "cvx_set_hypothesis_pair", "cvx_run_edge_sketch", "cvx_get_counts", "cvx_get_h2_angles", "cvx_get_h1_wedges",
    "cvx_get_h1_table", "cvx_get_rows", "cvx_get_support_matrix", "cvx_set_keep_intermediates", "cvx_get_intermediates", "cvx_get_valid_rows",
    "cvx_get_pairs", "cvx_get_reference_order", "cvx_finalize", "cvx_get_device_results", "cvx_get_timings", "cvx_synchronize", "cvx_get_stream",
    "cvx_claims_reset", "cvx_claims_accumulate", "cvx_claims_get_counts", "cvx_claims_get_mask", "cvx_claims_allreduce",
    "cvx_split_h1_by_rows", "cvx_comm_unique_id", "cvx_comm_init", "cvx_comm_attach", "cvx_comm_destroy", "cvx_gather_reset",
    "cvx_gather_post", "cvx_gather_wait", "cvx_gather_get", "cvx_gather_fetch", "cvx_gather_fetch_async", "cvx_gather_stats",
    "cvx_support_problem_create", "cvx_support_problem_run", "cvx_support_problem_stats", "cvx_support_problem_result",
    "cvx_support_problem_destroy",
    "cvx_graph_create", "cvx_graph_build_links", "cvx_graph_prune", "cvx_graph_get_links", "cvx_graph_get_link_details", "cvx_graph_get_neighbors",
    "cvx_graph_set_neighbors", "cvx_graph_align", "cvx_graph_destroy", "cvx_graph_align_no_alloc",
]


class CvxConfig(C.Structure):
    _fields_ = [
        ("delta", C.c_double), ("delta_theta_deg", C.c_double), ("parallel_angle_deg", C.c_double),
        ("lowes_ratio", C.c_double), ("epip_tangency_displ_thresh", C.c_double),
        ("location_perturbation", C.c_double), ("orient_perturbation", C.c_double),
        ("cluster_dist_thresh", C.c_double), ("cluster_orient_thresh_deg", C.c_double),
        ("cluster_gauss_sigma", C.c_double), ("pi_truncated", C.c_double),
        ("n_views", C.c_int32), ("img_rows", C.c_int32), ("img_cols", C.c_int32),
        ("max_support_views", C.c_int32), ("max_cluster_size", C.c_int32), ("strict_libm_order", C.c_int32),
    ]


class CvxCounts(C.Structure):
    _fields_ = [(n, C.c_int64) for n in (
        "n_h1_total", "n_h1_processed", "n_candidates", "n_hypotheses", "n_validations", "n_valid_rows",
        "n_pairs", "n_degenerate_wedges", "n_supported", "n_parallel", "n_inliers", "n_unique_hypotheses",
        "n_entries_visited", "n_heavy_queries")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class CvxTimings(C.Structure):
    _fields_ = [("set_pair_ms", C.c_float), ("run_ms", C.c_float), ("finalize_ms", C.c_float), ("form_ms", C.c_float),
                ("prep_ms", C.c_float), ("validate_ms", C.c_float), ("select_ms", C.c_float),
                ("kernel_launches", C.c_int32), ("pad", C.c_int32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "pad"}


class CvxDeviceResults(C.Structure):
    _fields_ = [("n_pairs", C.c_int64), ("n_vali_views", C.c_int32), ("pad", C.c_int32),
                ("pair_h1", C.c_void_p), ("pair_h2rep", C.c_void_p), ("pair_count", C.c_void_p), ("pair_supports", C.c_void_p),
                ("gamma_cam", C.c_void_p), ("gamma_world", C.c_void_p), ("tangent_world", C.c_void_p)]


class CvxGatheredUnit(C.Structure):
    _fields_ = [("meta", C.c_int64 * 8), ("pair_h1", C.c_void_p), ("pair_h2rep", C.c_void_p), ("pair_count", C.c_void_p), ("nan_flag", C.c_void_p),
                ("support_off", C.c_void_p), ("support_entries", C.c_void_p), ("gamma_world", C.c_void_p), ("tangent_world", C.c_void_p),
                ("gamma_cam", C.c_void_p), ("h2_corrected_xy", C.c_void_p)]


_lib = None


def load():
    """Load the CUDA library.  Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python 3d_curvix_b200/build.py` (there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    lib.cvx_last_error.restype = C.c_char_p
    lib.cvx_last_error.argtypes = [C.c_void_p]
    lib.cvx_create.argtypes = [C.POINTER(CvxConfig), C.c_int, C.POINTER(C.c_void_p)]
    lib.cvx_support_problem_create.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p] + [C.c_void_p] * 7 + \
        [C.c_double, C.c_double, C.POINTER(C.c_void_p)]
    for f in ("cvx_support_problem_run", "cvx_support_problem_stats", "cvx_support_problem_result", "cvx_support_problem_destroy"):
        getattr(lib, f).argtypes = None
    lib.cvx_support_problem_run.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
    lib.cvx_support_problem_stats.argtypes = [C.c_void_p] + [C.POINTER(C.c_int64)] * 4
    lib.cvx_support_problem_result.argtypes = [C.c_void_p, C.c_void_p]
    lib.cvx_support_problem_destroy.argtypes = [C.c_void_p]
    lib.cvx_graph_create.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]
    lib.cvx_graph_build_links.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64)]
    lib.cvx_graph_prune.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    lib.cvx_graph_get_links.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64)]
    lib.cvx_graph_get_link_details.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.cvx_graph_get_neighbors.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.cvx_graph_set_neighbors.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.cvx_graph_align.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.POINTER(C.c_float)]
    lib.cvx_graph_destroy.argtypes = [C.c_void_p]
    lib.cvx_graph_align_no_alloc.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_double,
                                             C.c_double, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
    lib.cvx_graph_align_no_alloc.restype = C.c_int
    lib.cvx_gather_post.argtypes = [C.c_void_p, C.c_int64, C.c_int]
    lib.cvx_gather_wait.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.cvx_gather_get.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(CvxGatheredUnit)]
    lib.cvx_gather_fetch.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 10
    lib.cvx_gather_fetch_async.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 10
    lib.cvx_comm_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    lib.cvx_comm_attach.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    lib.cvx_split_h1_by_rows.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    _lib = lib
    return lib


def default_config(**kw):
    cfg = CvxConfig()
    load().cvx_default_config(C.byref(cfg))
    for k, v in kw.items():
        setattr(cfg, k, v)
    return cfg


def _vp(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


class CurvixError(RuntimeError):
    pass


class EdgeSketch:
    """One context on one GPU.  Call order mirrors EdgeSketch_Core as driven by cmd/main.cpp:114-155:
    set_cameras -> set_view_edges (all views) -> per pass: set_pair -> run -> getters / finalize."""

    def __init__(self, cfg, device=0):
        self.lib = load()
        self.cfg = cfg
        self.h = C.c_void_p()
        rc = self.lib.cvx_create(C.byref(cfg), device, C.byref(self.h))
        if rc != 0:
            raise CurvixError(f"cvx_create failed with status {rc} (no CUDA device? there is no CPU fallback)")
        self.V = cfg.n_views - 2
        self.n_edges = [0] * cfg.n_views
        self._keep = []

    def _ck(self, rc):
        if rc != 0:
            raise CurvixError(f"status {rc}: {self.lib.cvx_last_error(self.h).decode()}")

    def close(self):
        if self.h:
            self.lib.cvx_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_cameras(self, R, T, K):
        R = np.ascontiguousarray(R, np.float64).reshape(-1, 9)
        T = np.ascontiguousarray(T, np.float64).reshape(-1, 3)
        K = np.ascontiguousarray(K, np.float64)
        per_view = 1 if (K.size == 9 * R.shape[0] and R.shape[0] > 1) else 0
        self._ck(self.lib.cvx_set_cameras(self.h, R.shape[0], _vp(R), _vp(T), _vp(K), per_view))

    def set_view_edges(self, v, edges):
        e = np.asarray(edges, np.float64)
        cols = [np.ascontiguousarray(e[:, k]) for k in range(3)]
        self.set_view_edges_soa(v, *cols)

    def set_view_edges_soa(self, v, x, y, th):
        """x, y, th: contiguous float64 arrays (numpy) or raw host pointers with explicit n via (ptr, n) tuples."""
        n = len(x)
        self._ck(self.lib.cvx_set_view_edges(self.h, v, n, _vp(x), _vp(y), _vp(th), None))
        self.n_edges[v] = n

    def set_view_edges_ptr(self, v, n, px, py, pth):
        self._ck(self.lib.cvx_set_view_edges(self.h, v, n, C.c_void_p(px), C.c_void_p(py), C.c_void_p(pth), None))
        self.n_edges[v] = n

    def set_all_view_edges(self, edges):
        """edges: list (one per view) of E x >=3 arrays [x y theta ...]; one packed upload for all views."""
        n = np.array([len(e) for e in edges], np.int32)
        cat = np.concatenate([np.asarray(e, np.float64)[:, :3] for e in edges], axis=0) if n.sum() else np.zeros((0, 3))
        cols = [np.ascontiguousarray(cat[:, k]) for k in range(3)]
        self.set_all_view_edges_ptr(n, cols[0].ctypes.data, cols[1].ctypes.data, cols[2].ctypes.data)

    def set_all_view_edges_ptr(self, n, px, py, pth):
        """n: int32 array of per-view edge counts; px/py/pth: raw host pointers to the packed columns (may be pinned)."""
        n = np.ascontiguousarray(n, np.int32)
        self._ck(self.lib.cvx_set_all_view_edges(self.h, len(n), _vp(n), C.c_void_p(px), C.c_void_p(py), C.c_void_p(pth)))
        self.n_edges = [int(k) for k in n]

    def set_pair(self, h1, h2):
        self._ck(self.lib.cvx_set_hypothesis_pair(self.h, h1, h2))
        self.h1, self.h2 = h1, h2

    def run(self, h1_begin=0, h1_end=-1):
        self._ck(self.lib.cvx_run_edge_sketch(self.h, h1_begin, h1_end))
        return self

    def counts(self):
        c = CvxCounts()
        self._ck(self.lib.cvx_get_counts(self.h, C.byref(c)))
        return c.as_dict()

    def timings(self):
        t = CvxTimings()
        self._ck(self.lib.cvx_get_timings(self.h, C.byref(t)))
        return t.as_dict()

    def synchronize(self):
        self._ck(self.lib.cvx_synchronize(self.h))

    def stream_ptr(self):
        """cudaStream_t of the context as an integer (for torch.cuda.ExternalStream)."""
        s = C.c_void_p()
        self._ck(self.lib.cvx_get_stream(self.h, C.byref(s)))
        return int(s.value or 0)

    def h2_angles(self):
        out = np.empty(self.n_edges[self.h2])
        self._ck(self.lib.cvx_get_h2_angles(self.h, _vp(out)))
        return out

    def h1_wedges(self):
        E = self.n_edges[self.h1]
        w = np.empty((E, 2)); f = np.empty(E, np.int32)
        self._ck(self.lib.cvx_get_h1_wedges(self.h, _vp(w), _vp(f)))
        return w, f

    def h1_table(self):
        E = self.n_edges[self.h1]
        m = np.empty(E, np.int32); n = np.empty(E, np.int32); off = np.empty(E, np.int64)
        self._ck(self.lib.cvx_get_h1_table(self.h, _vp(m), _vp(n), _vp(off)))
        return dict(m=m, n=n, off=off)

    def rows(self):
        H = self.counts()["n_hypotheses"]
        ints = {k: np.empty(H, np.int32) for k in ("h1", "h2_orig", "local_idx", "raw_at_local", "label", "h2_rep", "count")}
        dbl = {k: np.empty(H) for k in ("xc", "yc", "theta_orig", "xbar", "ybar", "thbar")}
        self._ck(self.lib.cvx_get_rows(self.h, _vp(ints["h1"]), _vp(ints["h2_orig"]), _vp(ints["local_idx"]), _vp(ints["raw_at_local"]),
                                       _vp(dbl["xc"]), _vp(dbl["yc"]), _vp(dbl["theta_orig"]), _vp(dbl["xbar"]), _vp(dbl["ybar"]),
                                       _vp(dbl["thbar"]), _vp(ints["label"]), _vp(ints["h2_rep"]), _vp(ints["count"])))
        ints.update(dbl)
        return ints

    def support_matrix(self):
        H = self.counts()["n_hypotheses"]
        out = np.empty((H, self.V), np.int32)
        self._ck(self.lib.cvx_get_support_matrix(self.h, _vp(out)))
        return out

    def keep_intermediates(self, on=True):
        self._ck(self.lib.cvx_set_keep_intermediates(self.h, 1 if on else 0))

    def intermediates(self):
        H = self.counts()["n_hypotheses"]
        E1 = self.n_edges[self.h1]
        w31 = np.empty((E1, self.V, 2)); w32 = np.empty((H, self.V, 2)); tgt = np.empty((H, self.V, 2))
        self._ck(self.lib.cvx_get_intermediates(self.h, _vp(w31), _vp(w32), _vp(tgt)))
        return dict(wedge31=w31, wedge32=w32, tgt=tgt)

    def valid_rows(self):
        n = self.counts()["n_valid_rows"]
        a = {k: np.empty(n, np.int32) for k in ("h1", "row", "count", "selected")}
        self._ck(self.lib.cvx_get_valid_rows(self.h, _vp(a["h1"]), _vp(a["row"]), _vp(a["count"]), _vp(a["selected"])))
        return a

    def results_into(self, ptrs):
        """Accepted pairs + supports + 3-D edges + tangents into caller-owned host buffers (raw pointers, e.g. pinned memory sized for
        n_pairs): ptrs = dict with any of h1, h2rep, row, count, supports, gamma_cam, gamma_world, tangent_world, h2_corrected, nan_flag."""
        q = lambda k: C.c_void_p(ptrs[k]) if ptrs.get(k) else None
        self._ck(self.lib.cvx_get_pairs(self.h, q("h1"), q("h2rep"), q("row"), q("count"), q("supports")))
        self._ck(self.lib.cvx_finalize(self.h, q("gamma_cam"), q("gamma_world"), q("tangent_world"), q("h2_corrected"), q("nan_flag")))

    def pairs(self, with_supports=True):
        P = self.counts()["n_pairs"]
        a = {k: np.empty(P, np.int32) for k in ("h1", "h2rep", "row", "count")}
        sup = np.empty((P, self.V), np.int32) if with_supports else None
        self._ck(self.lib.cvx_get_pairs(self.h, _vp(a["h1"]), _vp(a["h2rep"]), _vp(a["row"]), _vp(a["count"]), _vp(sup)))
        a["supports"] = sup
        return a

    def reference_order(self):
        """Indices into pairs()/finalize() arrays in the reference's paired_edge order (host std::sort step + capacity)."""
        P = self.counts()["n_pairs"]
        idx = np.empty(max(P, 1), np.int64); n = C.c_int64()
        self._ck(self.lib.cvx_get_reference_order(self.h, _vp(idx), C.byref(n)))
        return idx[:n.value].copy()

    def finalize(self):
        P = self.counts()["n_pairs"]
        gc = np.empty((P, 3)); gw = np.empty((P, 3)); tw = np.empty((P, 3)); h2c = np.empty((P, 2)); nf = np.empty(P, np.int32)
        self._ck(self.lib.cvx_finalize(self.h, _vp(gc), _vp(gw), _vp(tw), _vp(h2c), _vp(nf)))
        return dict(gamma_cam=gc, gamma_world=gw, tangent_world=tw, h2_corrected=h2c, nan_flag=nf)

    def claims_reset(self):
        self._ck(self.lib.cvx_claims_reset(self.h))

    def claims_accumulate(self):
        """Mark the edges claimed by the accepted pairs of the last run (EdgeSketch_Core.cpp:682-726, 785-797)."""
        self._ck(self.lib.cvx_claims_accumulate(self.h))

    def claims_counts(self):
        n = self.cfg.n_views
        claimed = np.empty(n, np.int64); total = np.empty(n, np.int64)
        self._ck(self.lib.cvx_claims_get_counts(self.h, _vp(claimed), _vp(total)))
        return claimed, total

    def claims_mask(self, view):
        out = np.zeros(self.n_edges[view], np.uint8)
        self._ck(self.lib.cvx_claims_get_mask(self.h, view, _vp(out)))
        return out.astype(bool)

    def split_h1_by_rows(self, parts):
        """parts + 1 bounds of contiguous view-1 ranges with about equal hypothesis-row counts (current pair)."""
        b = np.empty(parts + 1, np.int32)
        self._ck(self.lib.cvx_split_h1_by_rows(self.h, parts, _vp(b)))
        return b

    # ---- multi-GPU gather (cvx_comm.cu): NCCL behind the C-ABI
    @staticmethod
    def comm_unique_id():
        buf = C.create_string_buffer(128)
        rc = load().cvx_comm_unique_id(buf)
        if rc != 0:
            raise CurvixError("cvx_comm_unique_id failed (libnccl.so.2 not loadable?)")
        return buf.raw

    def comm_init(self, world, rank, unique_id):
        self._ck(self.lib.cvx_comm_init(self.h, world, rank, C.c_char_p(unique_id)))
        self.world, self.rank = world, rank

    def gather_reset(self):
        self._ck(self.lib.cvx_gather_reset(self.h))

    def gather_post(self, unit_id=0, empty=False):
        self._ck(self.lib.cvx_gather_post(self.h, int(unit_id), 1 if empty else 0))

    def gather_wait(self, join_stream=True, host_sync=False):
        self._ck(self.lib.cvx_gather_wait(self.h, 1 if join_stream else 0, 1 if host_sync else 0))

    def gather_get(self, rnd, rank):
        u = CvxGatheredUnit()
        self._ck(self.lib.cvx_gather_get(self.h, rnd, rank, C.byref(u)))
        return u

    GATHER_FIELDS = ("h1", "h2rep", "count", "nan_flag", "support_off", "support_entries", "gamma_world", "tangent_world", "gamma_cam", "h2_corrected")

    @staticmethod
    def gather_widths(P, S):
        """bytes of every field of a gathered unit with P rows and S supports"""
        return dict(h1=4 * P, h2rep=4 * P, count=4 * P, nan_flag=4 * P, support_off=8 * (P + 1), support_entries=8 * S, gamma_world=24 * P,
                    tangent_world=24 * P, gamma_cam=24 * P, h2_corrected=16 * P)

    def gather_fetch(self, rnd, rank):
        """Host copies of what `rank` contributed in round `rnd` (rank 0 only).  `supports` is the CSR list expanded to the reference's
        dense P x V layout (-2 = no support)."""
        u = self.gather_get(rnd, rank)
        P, V, S = int(u.meta[0]), int(u.meta[6]), int(u.meta[7])
        out = dict(meta=[int(x) for x in u.meta], h1=np.empty(P, np.int32), h2rep=np.empty(P, np.int32), count=np.empty(P, np.int32),
                   nan_flag=np.empty(P, np.int32), support_off=np.zeros(P + 1, np.int64), support_entries=np.empty((S, 2), np.int32),
                   gamma_world=np.empty((P, 3)), tangent_world=np.empty((P, 3)), gamma_cam=np.empty((P, 3)), h2_corrected=np.empty((P, 2)))
        if P:
            self._ck(self.lib.cvx_gather_fetch(self.h, rnd, rank, *[_vp(out[k]) for k in self.GATHER_FIELDS]))
        dense = np.full((P, V), -2, np.int32)
        if S:
            rows = np.repeat(np.arange(P), np.diff(out["support_off"]))
            dense[rows, out["support_entries"][:, 0]] = out["support_entries"][:, 1]
        out["supports"] = dense
        return out

    def gather_fetch_into(self, rnd, rank, ptrs, wait=True):
        """Same fields (supports as the CSR list), into caller-owned host buffers (raw pointers, e.g. pinned memory): ptrs = dict with any
        of GATHER_FIELDS.  wait=False: enqueued behind the round's transfers, completed by gather_wait(host_sync=True).  Returns the
        number of bytes copied."""
        u = self.gather_get(rnd, rank)
        P, S = int(u.meta[0]), int(u.meta[7])
        if P == 0:
            return 0
        width = self.gather_widths(P, S)
        fn = self.lib.cvx_gather_fetch if wait else self.lib.cvx_gather_fetch_async
        self._ck(fn(self.h, rnd, rank, *[C.c_void_p(ptrs[k]) if ptrs.get(k) else None for k in self.GATHER_FIELDS]))
        return sum(width[k] for k in self.GATHER_FIELDS if ptrs.get(k))

    def gather_stats(self):
        a, b, ms = C.c_int64(), C.c_int64(), C.c_float()
        self._ck(self.lib.cvx_gather_stats(self.h, C.byref(a), C.byref(b), C.byref(ms)))
        return dict(bytes_sent=a.value, bytes_received=b.value, comm_ms=ms.value)

    def claims_allreduce(self):
        self._ck(self.lib.cvx_claims_allreduce(self.h))

    def device_results(self):
        r = CvxDeviceResults()
        self._ck(self.lib.cvx_get_device_results(self.h, C.byref(r)))
        return r


ALIGN_NO_ALLOC_MAX_NODES = 65536


def graph_align_no_alloc(device, stream_ptr, n_nodes, d_offsets, d_neighbors, d_location, d_orientation, iterations, step_force, step_torque,
                         d_location_out, d_orientation_out):
    """cvx_graph_align_no_alloc: every d_* argument is a raw DEVICE address (int).  Enqueues only; returns (rc, message)."""
    lib = load()
    err = C.create_string_buffer(512)
    rc = lib.cvx_graph_align_no_alloc(int(device), C.c_void_p(stream_ptr or 0), int(n_nodes), C.c_void_p(d_offsets), C.c_void_p(d_neighbors),
                                      C.c_void_p(d_location), C.c_void_p(d_orientation), int(iterations), float(step_force), float(step_torque),
                                      C.c_void_p(d_location_out), C.c_void_p(d_orientation_out), err, len(err))
    return rc, err.value.decode()


class EdgeGraph:
    """The 3-D edge graph of EdgeMapping::consolidate_3D_edges on the device: links from curvelets, the two prunings, neighbour
    lists, the alignment iterations.  Nodes are the caller's numbering of the distinct 3-D edges."""

    def __init__(self, sketch, location, first_tangent):
        self.sk, self.lib = sketch, sketch.lib
        self.loc = np.ascontiguousarray(location, np.float64).reshape(-1, 3)
        self.t0 = np.ascontiguousarray(first_tangent, np.float64).reshape(-1, 3)
        self.n = len(self.loc)
        self.p = C.c_void_p()
        sketch._ck(self.lib.cvx_graph_create(sketch.h, self.n, _vp(self.loc), _vp(self.t0), C.byref(self.p)))

    def build_links(self, key_off, key_nodes, cv_self, cv_neighbor):
        ko = np.ascontiguousarray(key_off, np.int64); kn = np.ascontiguousarray(key_nodes, np.int32)
        a = np.ascontiguousarray(cv_self, np.int32); b = np.ascontiguousarray(cv_neighbor, np.int32)
        n = C.c_int64(0)
        self.sk._ck(self.lib.cvx_graph_build_links(self.p, len(ko) - 1, _vp(ko), _vp(kn), len(a), _vp(a), _vp(b), C.byref(n)))
        return int(n.value)

    def prune(self, R, T, K, lambda1=1.0, lambda2=1.0, lambda3=1.0, proj_thresh=6.0):
        R = np.ascontiguousarray(R, np.float64).reshape(-1, 9); T = np.ascontiguousarray(T, np.float64).reshape(-1, 3)
        K = np.ascontiguousarray(np.asarray(K, np.float64).reshape(-1)[:9])
        st = np.zeros(6); n1 = C.c_int64(0); n2 = C.c_int64(0)
        self.sk._ck(self.lib.cvx_graph_prune(self.p, lambda1, lambda2, lambda3, proj_thresh, len(R), _vp(R), _vp(T), _vp(K), _vp(st), C.byref(n1), C.byref(n2)))
        return dict(mu_t=st[0], sigma_t=st[1], mu_n=st[2], sigma_n=st[3], mu_theta=st[4], sigma_theta=st[5], n_after_3d=int(n1.value), n_after_projection=int(n2.value))

    def links(self, stage):
        n = C.c_int64(0)
        self.sk._ck(self.lib.cvx_graph_get_links(self.p, stage, None, None, C.byref(n)))
        a = np.zeros(max(n.value, 1), np.int32); b = np.zeros(max(n.value, 1), np.int32)
        self.sk._ck(self.lib.cvx_graph_get_links(self.p, stage, _vp(a), _vp(b), C.byref(n)))
        return a[:n.value], b[:n.value]

    def link_details(self):
        n = len(self.links(0)[0])
        pos = np.zeros(max(n, 1), np.int64); k3 = np.zeros(max(n, 1), np.uint8); kp = np.zeros(max(n, 1), np.uint8)
        self.sk._ck(self.lib.cvx_graph_get_link_details(self.p, _vp(pos), _vp(k3), _vp(kp)))
        return pos[:n], k3[:n].astype(bool), kp[:n].astype(bool)

    def neighbors(self, n_links2):
        off = np.zeros(self.n + 1, np.int64); nb = np.zeros(max(2 * n_links2, 1), np.int32)
        self.sk._ck(self.lib.cvx_graph_get_neighbors(self.p, _vp(off), _vp(nb)))
        return off, nb[:off[-1]]

    def set_neighbors(self, off, nb):
        off = np.ascontiguousarray(off, np.int64); nb = np.ascontiguousarray(nb, np.int32)
        self.sk._ck(self.lib.cvx_graph_set_neighbors(self.p, _vp(off), _vp(nb)))

    def align(self, orientation, iterations=1000, step_force=0.01, step_torque=0.01):
        o = np.ascontiguousarray(orientation, np.float64).reshape(-1, 3)
        lo = np.zeros((self.n, 3)); oo = np.zeros((self.n, 3)); ms = C.c_float(0)
        self.sk._ck(self.lib.cvx_graph_align(self.p, _vp(o), iterations, step_force, step_torque, _vp(lo), _vp(oo), C.byref(ms)))
        return lo, oo, float(ms.value)

    def close(self):
        if self.p:
            self.lib.cvx_graph_destroy(self.p)
            self.p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class SupportProblem:
    """Stand-alone support test on pre-formed hypotheses (BASELINE.json config 5)."""

    def __init__(self, sketch, group_off, view_off, ang31, ang32, cs, sn, wedge31, wedge32, tgt, cos_thresh, parallel_deg):
        self.sk = sketch
        self.lib = sketch.lib
        a = [np.ascontiguousarray(group_off, np.int64), np.ascontiguousarray(view_off, np.int64)] + \
            [np.ascontiguousarray(t, np.float64) for t in (ang31, ang32, cs, sn, wedge31, wedge32, tgt)]
        self.G = len(a[0]) - 1; self.V = len(a[1]) - 1; self.H = int(a[0][-1])
        self.p = C.c_void_p()
        rc = self.lib.cvx_support_problem_create(sketch.h, self.G, _vp(a[0]), self.V, _vp(a[1]), *[_vp(t) for t in a[2:]],
                                                 float(cos_thresh), float(parallel_deg), C.byref(self.p))
        sketch._ck(rc)

    def run(self):
        ms = C.c_float()
        self.sk._ck(self.lib.cvx_support_problem_run(self.p, C.byref(ms)))
        return ms.value

    def stats(self):
        v = [C.c_int64() for _ in range(4)]
        self.sk._ck(self.lib.cvx_support_problem_stats(self.p, *[C.byref(x) for x in v]))
        return dict(zip(("n_inliers", "n_visited", "n_supported", "n_parallel"), (int(x.value) for x in v)))

    def result(self):
        out = np.empty((self.V, self.H), np.int32)
        self.sk._ck(self.lib.cvx_support_problem_result(self.p, _vp(out)))
        return out

    def close(self):
        if self.p:
            self.lib.cvx_support_problem_destroy(self.p)
            self.p = C.c_void_p()

"""ctypes binding of oracle/_ref/libcurvix_ref*.so: the UNMODIFIED 3D Curvix reference sources compiled against the stand-in
Eigen / yaml-cpp headers of oracle/ref_shim/ (recipe: oracle/ref_shim/Makefile).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
The product package (3d_curvix_b200/) never imports this module.  /root/reference exists only in the build container; on the
GPU box the prebuilt libraries under oracle/_ref/ (they travel with gpurun) are used as they are.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("CURVIX_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(_HERE, "_ref")
_VARIANTS = {"strict": "libcurvix_ref.so", "fast": "libcurvix_ref_fast.so", "tree": "libcurvix_ref_tree.so", "chk": "libcurvix_ref_chk.so"}


class RefConfig(C.Structure):
    _fields_ = [
        ("delta", C.c_double), ("delta_theta_deg", C.c_double), ("parallel_angle_deg", C.c_double),
        ("reproj_dist_thresh", C.c_double), ("stop_ratio", C.c_double),
        ("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
        ("n_views", C.c_int32), ("img_rows", C.c_int32), ("img_cols", C.c_int32), ("max_support_views", C.c_int32),
        ("num_threads", C.c_int32), ("toed_thresh", C.c_int32), ("max_passes", C.c_int32), ("reserved", C.c_int32),
        ("dataset_path", C.c_char * 512), ("dataset_name", C.c_char * 64), ("scene_name", C.c_char * 64),
    ]


def have_sources():
    return os.path.exists(os.path.join(REF_ROOT, "Edge_Reconst", "EdgeSketch_Core.cpp"))


def available(variant="strict"):
    return os.path.exists(os.path.join(OUT, _VARIANTS[variant])) or have_sources()


def build(variants=("strict",), quiet=True, main=False):
    """Compile the reference where it lies (never copies its sources).  No-op without /root/reference (GPU box)."""
    if not have_sources():
        return False
    targets = [os.path.join(OUT, _VARIANTS[v]) for v in variants]
    if main:
        targets.append(os.path.join(OUT, "curvix_main_ref"))
    subprocess.run(["make", "-C", os.path.join(_HERE, "ref_shim"), "-j8", "REF=" + REF_ROOT] + targets, check=True,
                   stdout=subprocess.DEVNULL if quiet else None)
    return True


def _ptr(a, t):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


_d = lambda a: _ptr(a, C.c_double)
_i = lambda a: _ptr(a, C.c_int32)
_l = lambda a: _ptr(a, C.c_int64)
_libs = {}


def load(variant="strict"):
    if variant in _libs:
        return _libs[variant]
    build((variant,))
    path = os.path.join(OUT, _VARIANTS[variant])
    if not os.path.exists(path):
        raise FileNotFoundError(path + " (needs /root/reference to build, or the prebuilt library)")
    lib = C.CDLL(path)
    lib.ref_create.restype = C.c_void_p
    lib.ref_create.argtypes = [C.POINTER(RefConfig)]
    lib.ref_build_info.restype = C.c_char_p
    for n in ("ref_n_pairs", "ref_n_support_blocks", "ref_n_cluster_keys", "ref_n_3d_keys"):
        getattr(lib, n).restype = C.c_int64
    for n in ("ref_destroy", "ref_get_times", "ref_relative_pose", "ref_fundamental", "ref_inverse3", "ref_ore_list",
              "ref_ore_list_bar", "ref_gamma3", "ref_triangulate", "ref_tangent3d"):
        getattr(lib, n).restype = None
    lib.ref_hypo2_idx_ore.argtypes = [C.c_int, C.POINTER(C.c_double), C.c_double, C.c_double, C.POINTER(C.c_int32)]
    lib.ref_support_idx.argtypes = [C.c_double, C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_double)]
    lib.ref_ore_list.argtypes = [C.c_int, C.c_int] + [C.POINTER(C.c_double)] * 4 + [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double),
                                 C.c_double, C.c_int, C.c_int, C.POINTER(C.c_double)]
    lib.ref_ore_list_bar.argtypes = [C.c_int] + [C.POINTER(C.c_double)] * 4 + [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double),
                                     C.c_double, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    _libs[variant] = lib
    return lib


def build_info(variant="strict"):
    return load(variant).ref_build_info().decode()


def config_for(scene, *, delta=0.4, delta_theta_deg=15.0, parallel_angle_deg=15.0, max_support_views=4, num_threads=0,
               stop_ratio=0.7, max_passes=25, toed_thresh=1, dataset_path="", dataset_name="ABC-NEF", scene_name="00000006"):
    K = np.asarray(scene.K, np.float64).reshape(-1)[:9].reshape(3, 3)
    c = RefConfig()
    c.delta, c.delta_theta_deg, c.parallel_angle_deg = delta, delta_theta_deg, parallel_angle_deg
    c.reproj_dist_thresh, c.stop_ratio = 2.0, stop_ratio
    c.fx, c.fy, c.cx, c.cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    c.n_views, c.img_rows, c.img_cols = scene.n_views, scene.rows, scene.cols
    c.max_support_views, c.num_threads = max_support_views, num_threads or (os.cpu_count() or 1)
    c.toed_thresh, c.max_passes = toed_thresh, max_passes
    c.dataset_path, c.dataset_name, c.scene_name = dataset_path.encode(), dataset_name.encode(), scene_name.encode()
    return c


class _run_dir:
    """The reference writes to the relative path ../../outputs (definitions.h:4): run those calls from oracle/_ref/run/bin so the
    files land in oracle/_ref/outputs (git-ignored) and nowhere outside the repository."""

    def __enter__(self):
        self.prev = os.getcwd()
        d = os.path.join(OUT, "run", "bin")
        os.makedirs(d, exist_ok=True)
        os.makedirs(os.path.join(OUT, "outputs"), exist_ok=True)
        os.chdir(d)

    def __exit__(self, *a):
        os.chdir(self.prev)


class Reference:
    """The reference's EdgeSketch_Core driven in the call order of cmd/main.cpp:114-155."""

    def __init__(self, cfg, variant="strict"):
        self.lib = load(variant)
        self.cfg = cfg
        self.V = cfg.n_views - 2
        self.h = C.c_void_p(self.lib.ref_create(C.byref(cfg)))
        self.n_edges = [0] * cfg.n_views

    @classmethod
    def for_scene(cls, scene, variant="strict", **kw):
        r = cls(config_for(scene, **kw), variant)
        r.set_cameras(scene.R, scene.T)
        for v, e in enumerate(scene.edges):
            r.set_view_edges(v, e)
        return r

    def close(self):
        if self.h:
            self.lib.ref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_cameras(self, R, T):
        R = np.ascontiguousarray(R, np.float64).reshape(-1, 9)
        T = np.ascontiguousarray(T, np.float64).reshape(-1, 3)
        assert self.lib.ref_set_cameras(self.h, R.shape[0], _d(R), _d(T)) == 0

    def set_view_edges(self, v, edges):
        e = np.asarray(edges, np.float64)
        x, y, th = (np.ascontiguousarray(e[:, k]) for k in range(3))
        s = np.ascontiguousarray(e[:, 3]) if e.shape[1] > 3 else np.ones(len(e))
        assert self.lib.ref_set_view_edges(self.h, v, len(e), _d(x), _d(y), _d(th), _d(s)) == 0
        self.n_edges[v] = len(e)

    def limit_h1(self, n):
        self.lib.ref_limit_h1(self.h, int(n))

    def set_pair(self, h1, h2):
        assert self.lib.ref_set_pair(self.h, h1, h2) == 0
        self.h1, self.h2 = h1, h2

    def run(self):
        assert self.lib.ref_run(self.h) == 0

    def finalize(self):
        with _run_dir():
            assert self.lib.ref_finalize(self.h) == 0

    def clear(self):
        self.lib.ref_clear(self.h)

    def select_next_views(self):
        b1, b2, ab, avg = C.c_int32(), C.c_int32(), C.c_int32(), C.c_double()
        self.lib.ref_select_next_views(self.h, C.byref(b1), C.byref(b2), C.byref(avg), C.byref(ab))
        return b1.value, b2.value, avg.value, bool(ab.value)

    def mapping_add_many(self, rec, image, edge_idx):
        """rec: n x 12 doubles {X Y Z, tx ty tz, x y, ux uy utheta, 0}: EdgeMapping::add3DToSupportingEdgesMapping for every record."""
        rec = np.ascontiguousarray(rec, np.float64).reshape(-1, 12)
        image = np.ascontiguousarray(image, np.int32); edge_idx = np.ascontiguousarray(edge_idx, np.int32)
        assert self.lib.ref_mapping_add_many(self.h, C.c_int64(len(rec)), _d(rec), _i(image), _i(edge_idx)) == 0

    def consolidate(self):
        with _run_dir():
            self.lib.ref_consolidate(self.h)
        return os.path.join(OUT, "outputs")

    def times(self):
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        self.lib.ref_get_times(self.h, C.byref(a), C.byref(b), C.byref(c))
        return dict(pair_edges=a.value, finalize=b.value, next_views=c.value)

    def pairs(self):
        P = int(self.lib.ref_n_pairs(self.h))
        h1 = np.empty(P, np.int32); h2 = np.empty(P, np.int32); sup = np.empty((P, self.V), np.int32)
        self.lib.ref_get_pairs(self.h, _i(h1), _i(h2), _i(sup))
        return dict(h1=h1, h2rep=h2, supports=sup)

    def points(self):
        P = int(self.lib.ref_n_pairs(self.h))
        gc = np.empty((P, 3)); tw = np.empty((P, 3)); gw = np.empty((P, 3))
        self.lib.ref_get_3d(self.h, _d(gc), _d(tw), _d(gw))
        return dict(gamma_cam=gc, tangent_world=tw, gamma_world=gw)

    def support_blocks(self):
        """all_supported_indices: list of (m x V) arrays, one per processed view-1 edge (NaN = never written)."""
        nb = int(self.lib.ref_n_support_blocks(self.h))
        rows = np.empty(nb, np.int64)
        self.lib.ref_get_support_block_rows(self.h, _l(rows))
        flat = np.empty((int(rows.sum()), self.V))
        self.lib.ref_get_support_blocks(self.h, _d(flat))
        return np.split(flat, np.cumsum(rows)[:-1]) if nb else []

    def cluster_map(self):
        tot = C.c_int64()
        nk = int(self.lib.ref_n_cluster_keys(self.h, C.byref(tot)))
        k1 = np.empty(nk, np.int32); k2 = np.empty(nk, np.int32); off = np.empty(nk + 1, np.int64); mem = np.empty(tot.value, np.int32)
        self.lib.ref_get_cluster_map(self.h, _i(k1), _i(k2), _l(off), _i(mem))
        return {(int(k1[i]), int(k2[i])): mem[off[i]:off[i + 1]].copy() for i in range(nk)}

    def accumulate_claims(self, masks):
        arr = (C.c_void_p * len(masks))(*[m.ctypes.data for m in masks])
        self.lib.ref_accumulate_claims(self.h, arr)

    def n_3d_keys(self):
        return int(self.lib.ref_n_3d_keys(self.h))


# ------------------------------------------------------------------------------------- helper classes, stand-alone
def _cam(scene_or_RT):
    R, T = scene_or_RT
    R = np.ascontiguousarray(R, np.float64).reshape(-1, 9)
    T = np.ascontiguousarray(T, np.float64).reshape(-1, 3)
    return R, T


def _k9(K):
    return np.ascontiguousarray(np.asarray(K, np.float64).reshape(-1)[:9])


def relative_pose(R1, T1, R2, T2, variant="strict"):
    lib = load(variant)
    a = [np.ascontiguousarray(x, np.float64).reshape(-1) for x in (R1, T1, R2, T2)]
    R21 = np.empty(9); T21 = np.empty(3); R12 = np.empty(9); T12 = np.empty(3)
    lib.ref_relative_pose(_d(a[0]), _d(a[1]), _d(a[2]), _d(a[3]), _d(R21), _d(T21), _d(R12), _d(T12))
    return R21.reshape(3, 3), T21, R12.reshape(3, 3), T12


def fundamental(K1, K2, R21, T21, variant="strict"):
    lib = load(variant)
    F = np.empty(9)
    lib.ref_fundamental(_d(_k9(K1)), _d(_k9(K2)), _d(np.ascontiguousarray(R21, np.float64).reshape(-1)), _d(np.ascontiguousarray(T21, np.float64)), _d(F))
    return F.reshape(3, 3)


def inverse3(A, variant="strict"):
    out = np.empty(9)
    load(variant).ref_inverse3(_d(_k9(A)), _d(out))
    return out.reshape(3, 3)


def ore_list(which, R, T, K1, K2, valid_idx, ref_idx, edges, delta, rows, cols, variant="strict"):
    """which 0: getOreList(hyp01=ref_idx, hyp02=valid_idx, ...); 1: getOreListVali(..., VALID_INDX=valid_idx, REFIDX=ref_idx)."""
    R, T = _cam((R, T))
    e = np.ascontiguousarray(np.asarray(edges, np.float64)[:, :4])
    if e.shape[1] < 4:
        e = np.ascontiguousarray(np.hstack([e, np.ones((len(e), 4 - e.shape[1]))]))
    out = np.empty(len(e))
    load(variant).ref_ore_list(which, R.shape[0], _d(R), _d(T), _d(_k9(K1)), _d(_k9(K2)), valid_idx, ref_idx, len(e), _d(e), float(delta), rows, cols, _d(out))
    return out


def ore_list_bar(R, T, K1, K2, valid_idx, ref_idx, edges, delta, rows, cols, variant="strict"):
    R, T = _cam((R, T))
    e = np.ascontiguousarray(np.asarray(edges, np.float64)[:, :4])
    out = np.empty((len(e), 2)); ep = np.empty(3)
    load(variant).ref_ore_list_bar(R.shape[0], _d(R), _d(T), _d(_k9(K1)), _d(_k9(K2)), valid_idx, ref_idx, len(e), _d(e), float(delta), rows, cols, _d(out), _d(ep))
    return out, ep


def hypo2_idx_ore(angles, lo, hi, variant="strict"):
    a = np.ascontiguousarray(angles, np.float64)
    out = np.empty(len(a), np.int32)
    n = load(variant).ref_hypo2_idx_ore(len(a), _d(a), float(lo), float(hi), _i(out))
    return out[:n].copy()


def epipolar_correction(edgels_h2, edgel_h1, F21, F12, idx_raw, variant="strict"):
    e2 = np.ascontiguousarray(edgels_h2, np.float64).reshape(-1, 4)
    e1 = np.ascontiguousarray(edgel_h1, np.float64).reshape(4)
    raw = np.ascontiguousarray(idx_raw, np.float64)
    out = np.empty((len(e2), 10))
    n = load(variant).ref_epipolar_correction(len(e2), _d(e2), _d(e1), _d(np.ascontiguousarray(F21, np.float64).reshape(-1)),
                                              _d(np.ascontiguousarray(F12, np.float64).reshape(-1)), _d(raw), _d(out))
    return out[:n].copy()


def cluster(h2_final, h1_idx, idx_raw, e2_rows, corrected, variant="strict"):
    h2 = np.ascontiguousarray(h2_final, np.float64).reshape(-1, 4)
    n = len(h2)
    raw = np.ascontiguousarray(idx_raw, np.float64)
    cor = np.ascontiguousarray(corrected, np.float64).reshape(n, 10)
    labels = np.empty(n, np.int32); hidx = np.empty(n); avg = np.empty((n, 4)); nc = C.c_int32(); nk = C.c_int32()
    key = np.empty(max(n, 1), np.int32); off = np.empty(n + 1, np.int64); mem = np.empty(max(n * n, 1), np.int32)
    load(variant).ref_cluster(n, _d(h2), int(h1_idx), len(raw), _d(raw), int(e2_rows), _d(cor), _i(labels), _d(hidx), _d(avg),
                              C.byref(nc), _i(key), _l(off), _i(mem), C.byref(nk))
    keys = {int(key[i]): mem[off[i]:off[i + 1]].copy() for i in range(nk.value)}
    return dict(labels=labels, hypo2_idx=hidx, averaged=avg, n_clusters=nc.value, keys=keys)


def gamma3(R, T, K1, K2, h1, h2, valid_idx, edges_h1, edges_h2, variant="strict"):
    R, T = _cam((R, T))
    a = np.ascontiguousarray(edges_h1, np.float64).reshape(-1, 4); b = np.ascontiguousarray(edges_h2, np.float64).reshape(-1, 4)
    out = np.empty((len(b), 4))
    load(variant).ref_gamma3(R.shape[0], _d(R), _d(T), _d(_k9(K1)), _d(_k9(K2)), h1, h2, valid_idx, len(b), _d(a), _d(b), _d(out))
    return out


def support_idx(delta_theta_deg, tgt, tangents, inliers, variant="strict"):
    t = np.ascontiguousarray(tgt, np.float64); tv = np.ascontiguousarray(tangents, np.float64).reshape(-1, 2)
    inl = np.ascontiguousarray(inliers, np.float64)
    return load(variant).ref_support_idx(float(delta_theta_deg), _d(t), len(tv), _d(tv), len(inl), _d(inl))


def triangulate(p1, p2, R21, T21, K, variant="strict"):
    out = np.empty(3)
    load(variant).ref_triangulate(_d(np.ascontiguousarray(p1, np.float64)), _d(np.ascontiguousarray(p2, np.float64)),
                                  _d(np.ascontiguousarray(R21, np.float64).reshape(-1)), _d(np.ascontiguousarray(T21, np.float64)), _d(_k9(K)), _d(out))
    return out


def closest_vector_to_line(P1, d1, P2, variant="strict"):
    out = np.empty(3)
    f = lambda x: _d(np.ascontiguousarray(x, np.float64).reshape(-1))
    load(variant).ref_closest_vector_to_line(f(P1), f(d1), f(P2), _d(out))
    return out


def shortest_align_euler_deg(v1, v2, variant="strict"):
    out = np.empty(3)
    f = lambda x: _d(np.ascontiguousarray(x, np.float64).reshape(-1))
    load(variant).ref_shortest_align_euler_deg(f(v1), f(v2), _d(out))
    return out


def euler_to_rotation(roll, pitch, yaw, variant="strict"):
    out = np.empty(9)
    lib = load(variant)
    lib.ref_euler_to_rotation.argtypes = [C.c_double, C.c_double, C.c_double, C.c_void_p]
    lib.ref_euler_to_rotation(float(roll), float(pitch), float(yaw), _d(out))
    return out.reshape(3, 3)


def tangent3d(e1, e2, K1, K2, R1, T1, R2, T2, variant="strict"):
    out = np.empty(3)
    f = lambda x: _d(np.ascontiguousarray(x, np.float64).reshape(-1))
    load(variant).ref_tangent3d(f(e1), f(e2), _d(_k9(K1)), _d(_k9(K2)), f(R1), f(T1), f(R2), f(T2), _d(out))
    return out

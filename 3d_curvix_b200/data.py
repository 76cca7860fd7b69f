"""Input side of the edge-sketch path: reference dataset layout readers and the seeded synthetic
multiview edge generator used by the tests and bench.py.

File formats follow the reference readers:
  Edges/Edge_<i>_t<thr>.txt   4 doubles per edge  x y theta strength   (Edge_Reconst/file_reader.cpp:50-78)
  RnT/R_matrix.txt            3 doubles per line, 3 lines per view      (file_reader.cpp:113-145)
  RnT/T_matrix.txt            1 double per line, 3 per view             (file_reader.cpp:148-174)
"""
import os

import numpy as np


class Scene:
    """A multiview edge scene: cameras (world->camera R, T; intrinsics K) and per-view edge tables."""

    def __init__(self, R, T, K, edges, rows, cols, name="scene"):
        self.R = np.ascontiguousarray(R, np.float64).reshape(-1, 3, 3)
        self.T = np.ascontiguousarray(T, np.float64).reshape(-1, 3)
        self.K = np.ascontiguousarray(K, np.float64).reshape(3, 3)
        self.edges = [np.ascontiguousarray(e, np.float64) for e in edges]   # each E x 4 [x y theta strength]
        self.rows, self.cols = int(rows), int(cols)
        self.name = name

    @property
    def n_views(self):
        return len(self.edges)

    def save_npz(self, path):
        d = {"R": self.R, "T": self.T, "K": self.K, "rows": self.rows, "cols": self.cols,
             "counts": np.array([len(e) for e in self.edges], np.int64),
             "edges": np.concatenate(self.edges, 0)}
        np.savez_compressed(path, **d)

    @staticmethod
    def load_npz(path, name=None):
        z = np.load(path)
        off = np.concatenate([[0], np.cumsum(z["counts"])])
        edges = [z["edges"][off[i]:off[i + 1]] for i in range(len(off) - 1)]
        return Scene(z["R"], z["T"], z["K"], edges, int(z["rows"]), int(z["cols"]), name or os.path.basename(path))

    def write_reference_layout(self, root, dataset="SYNTH", scene="s0", thresh=1):
        """Write the scene in the reference's on-disk layout (for the C++ curvix_main drop-in)."""
        base = os.path.join(root, dataset, scene)
        os.makedirs(os.path.join(base, "Edges"), exist_ok=True)
        os.makedirs(os.path.join(base, "RnT"), exist_ok=True)
        for i, e in enumerate(self.edges):
            np.savetxt(os.path.join(base, "Edges", f"Edge_{i}_t{thresh}.txt"), e, fmt="%.17g", delimiter="\t")
        np.savetxt(os.path.join(base, "RnT", "R_matrix.txt"), self.R.reshape(-1, 3), fmt="%.17g", delimiter="\t")
        np.savetxt(os.path.join(base, "RnT", "T_matrix.txt"), self.T.reshape(-1, 1), fmt="%.17g")
        cv = getattr(self, "curvelets", None)
        if cv is not None:   # curvelets/Curvelets_<i>_t<thr>.txt: 4 ints per line (file_reader.cpp:81-110 reads them as a stream of triples)
            os.makedirs(os.path.join(base, "curvelets"), exist_ok=True)
            for i, c in enumerate(cv):
                np.savetxt(os.path.join(base, "curvelets", f"Curvelets_{i}_t{thresh}.txt"), np.asarray(c, np.int64).reshape(-1, 4), fmt="%d")
        return base

    def load_curvelets(self, path):
        """Attach the curvelet files of the scene (npz made by tests/golden/make_abc_fixture.py): per view a flat int stream."""
        z = np.load(path)
        off = np.concatenate([[0], np.cumsum(z["counts"])])
        self.curvelets = [z["flat"][off[i]:off[i + 1]] for i in range(len(off) - 1)]
        return self

    def curvelet_triples(self, view):
        """(self edge, neighbour edge) per record as the reference's reader forms them: the int stream in triples, third value unused."""
        f = np.asarray(self.curvelets[view], np.int64)
        n = len(f) // 3
        t = f[:3 * n].reshape(n, 3)
        return t[:, 0].astype(np.int32), t[:, 1].astype(np.int32)


def read_reference_scene(scene_dir, n_views, fx, fy, cx, cy, rows, cols, thresh=1, refined=False):
    """Read <Dataset_Path>/<Dataset_Name>/<Scene_Name> as the reference's file_reader does."""
    rn = "refined_R.txt" if refined else "R_matrix.txt"
    tn = "refined_T.txt" if refined else "T_matrix.txt"
    R = np.loadtxt(os.path.join(scene_dir, "RnT", rn)).reshape(-1, 3, 3)[:n_views]
    T = np.loadtxt(os.path.join(scene_dir, "RnT", tn)).reshape(-1, 3)[:n_views]
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1.0]])
    edges = []
    for i in range(n_views):
        e = np.loadtxt(os.path.join(scene_dir, "Edges", f"Edge_{i}_t{thresh}.txt")).reshape(-1, 4)
        edges.append(e)
    return Scene(R, T, K, edges, rows, cols, os.path.basename(scene_dir.rstrip("/")))


# ----------------------------------------------------------------------------------------------
# Synthetic generator
# ----------------------------------------------------------------------------------------------
def _look_at(c):
    """World->camera rotation for a camera at c looking at the origin (x right, y down, z forward)."""
    z = -c / np.linalg.norm(c)
    up = np.array([0.0, 0.0, 1.0])
    if abs(z @ up) > 0.95:
        up = np.array([0.0, 1.0, 0.0])
    x = np.cross(z, up); x /= np.linalg.norm(x)
    y = np.cross(z, x)
    return np.stack([x, y, z], 0)


def _curves(rng, n_curves):
    """Random 3-D curves inside the unit ball: circles, segments, helices.  Returns callables t->(P, dP)."""
    out = []
    for k in range(n_curves):
        kind = k % 3
        c = rng.normal(size=3); c *= rng.uniform(0, 0.55) / np.linalg.norm(c)
        a = rng.normal(size=3); a /= np.linalg.norm(a)
        b = np.cross(a, rng.normal(size=3)); b /= np.linalg.norm(b)
        n = np.cross(a, b)
        if kind == 0:      # circle
            r = rng.uniform(0.15, 0.42)
            f = lambda t, c=c, a=a, b=b, r=r: (c + r * (np.cos(2 * np.pi * t)[:, None] * a + np.sin(2 * np.pi * t)[:, None] * b),
                                                r * 2 * np.pi * (-np.sin(2 * np.pi * t)[:, None] * a + np.cos(2 * np.pi * t)[:, None] * b))
            length = 2 * np.pi * r
        elif kind == 1:    # segment
            L = rng.uniform(0.4, 0.85)
            f = lambda t, c=c, a=a, L=L: (c + (t[:, None] - 0.5) * L * a, np.repeat((L * a)[None], len(t), 0))
            length = L
        else:              # helix
            r = rng.uniform(0.1, 0.3); h = rng.uniform(0.3, 0.6); turns = rng.uniform(1.0, 2.5)
            w = 2 * np.pi * turns
            f = lambda t, c=c, a=a, b=b, n=n, r=r, h=h, w=w: (
                c + r * (np.cos(w * t)[:, None] * a + np.sin(w * t)[:, None] * b) + ((t[:, None] - 0.5) * h) * n,
                r * w * (-np.sin(w * t)[:, None] * a + np.cos(w * t)[:, None] * b) + h * n)
            length = np.hypot(r * w, h)
        out.append((f, length))
    return out


def make_synthetic_scene(n_views, edges_per_view, rows=800, cols=800, seed=20251106, n_curves=24, pair_baseline_deg=25.0,
                         outlier_frac=0.10, loc_sigma=0.25, ori_sigma_deg=2.0, cam_radius=4.0, name=None, fy_scale=1.0, c_shift=(0.0, 0.0)):
    """Seeded synthetic multiview edge scene.

    Views 0 and 1 form a hypothesis pair with `pair_baseline_deg` of baseline; the remaining cameras are
    spread over the sphere.  Every view gets exactly `edges_per_view` edgels: (1-outlier_frac) sampled along
    the projected curves with Gaussian location/orientation noise, the rest uniformly random.  No edgel lies
    within 10 px of the border.
    """
    rng = np.random.default_rng(seed)
    f = 1111.11136542426 * (cols / 800.0)
    # fy_scale / c_shift: anisotropic focal length and off-centre principal point (the DTU intrinsics have both)
    fy = f * fy_scale
    K = np.array([[f, 0, (cols - 1) / 2.0 + c_shift[0]], [0, fy, (rows - 1) / 2.0 + c_shift[1]], [0, 0, 1.0]])
    curves = _curves(rng, n_curves)
    tot_len = sum(L for _, L in curves)
    # camera centres
    C = np.zeros((n_views, 3))
    d0 = rng.normal(size=3); d0 /= np.linalg.norm(d0)
    ax = np.cross(d0, rng.normal(size=3)); ax /= np.linalg.norm(ax)
    ang = np.deg2rad(pair_baseline_deg)
    d1 = d0 * np.cos(ang) + np.cross(ax, d0) * np.sin(ang)
    C[0], C[1] = cam_radius * d0, cam_radius * d1
    for v in range(2, n_views):
        if v % 2 == 0:
            d = rng.normal(size=3); d /= np.linalg.norm(d)
        else:   # (2k, 2k+1) are usable as hypothesis pairs: same baseline as (0, 1)
            p = C[v - 1] / cam_radius
            ax = np.cross(p, rng.normal(size=3)); ax /= np.linalg.norm(ax)
            d = p * np.cos(ang) + np.cross(ax, p) * np.sin(ang)
        C[v] = cam_radius * d
    R = np.stack([_look_at(c) for c in C], 0)
    T = np.stack([-R[v] @ C[v] for v in range(n_views)], 0)
    n_in_target = int(round(edges_per_view * (1.0 - outlier_frac)))
    edges = []
    for v in range(n_views):
        xs, ys, ths = [], [], []
        over = 1.6   # oversample, then trim to the exact target
        for fcurve, L in curves:
            nc = max(4, int(n_in_target * over * L / tot_len))
            t = (np.arange(nc) + rng.uniform()) / nc
            P, dP = fcurve(t)
            Pc = P @ R[v].T + T[v]
            dPc = dP @ R[v].T
            z = Pc[:, 2]
            u = f * Pc[:, 0] / z + K[0, 2]
            w = fy * Pc[:, 1] / z + K[1, 2]
            # image tangent = d/dt of the projection
            du = f * (dPc[:, 0] * z - Pc[:, 0] * dPc[:, 2]) / (z * z)
            dw = fy * (dPc[:, 1] * z - Pc[:, 1] * dPc[:, 2]) / (z * z)
            xs.append(u); ys.append(w); ths.append(np.arctan2(dw, du))
        x = np.concatenate(xs); y = np.concatenate(ys); th = np.concatenate(ths)
        x = x + rng.normal(0, loc_sigma, x.shape); y = y + rng.normal(0, loc_sigma, y.shape)
        th = th + rng.normal(0, np.deg2rad(ori_sigma_deg), th.shape)
        # third-order edge orientations live in (-pi/2, pi/2] like the reference's Edge_*.txt
        th = np.arctan(np.tan(th))
        ok = (x > 10.5) & (x < cols - 10.5) & (y > 10.5) & (y < rows - 10.5)
        x, y, th = x[ok], y[ok], th[ok]
        if len(x) > n_in_target:
            sel = np.sort(rng.choice(len(x), n_in_target, replace=False))
            x, y, th = x[sel], y[sel], th[sel]
        n_out = edges_per_view - len(x)
        xo = rng.uniform(10.5, cols - 10.5, n_out); yo = rng.uniform(10.5, rows - 10.5, n_out)
        tho = rng.uniform(-np.pi / 2, np.pi / 2, n_out)
        e = np.stack([np.concatenate([x, xo]), np.concatenate([y, yo]), np.concatenate([th, tho]),
                      np.ones(edges_per_view)], 1)
        e = e[rng.permutation(edges_per_view)]   # detection order carries no structure
        edges.append(e)
    return Scene(R, T, K, edges, rows, cols, name or f"synth_{n_views}x{edges_per_view}_s{seed}")


# Named workloads of BASELINE.json / BASELINE.md section 3
def workload(name):
    if name == "cfg2":   # 50 views x 10k edges, single pair (0,1)
        return dict(scene=dict(n_views=50, edges_per_view=10000, rows=800, cols=800, seed=20251106), pairs=[(0, 1)],
                    max_support_views=4)
    if name == "cfg3":   # DTU-shaped 49 x 100k, 1600x1200, N=6, 24 disjoint pairs
        return dict(scene=dict(n_views=49, edges_per_view=100000, rows=1200, cols=1600, seed=20251107, n_curves=96),
                    pairs=[(2 * k, 2 * k + 1) for k in range(24)], max_support_views=6)
    if name == "cfg4":   # 100 views x 50k, fixed 16-pair schedule
        return dict(scene=dict(n_views=100, edges_per_view=50000, rows=800, cols=800, seed=20251108, n_curves=64),
                    pairs=[(2 * k, 2 * k + 1) for k in range(16)], max_support_views=4)
    if name == "tiny":   # CPU-test sized
        return dict(scene=dict(n_views=8, edges_per_view=400, rows=800, cols=800, seed=7, n_curves=6), pairs=[(0, 1)],
                    max_support_views=2)
    raise KeyError(name)

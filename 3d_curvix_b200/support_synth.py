"""Seeded generator of pre-formed support-test problems (BASELINE.json config 5: "support-test stress").

A problem is what the inner body of the validation loop consumes (EdgeSketch_Core.cpp:348-384): per validation view the
two epipolar angles and the unit tangent (cos, sin) of every edge; per group (= one view-1 edge) and view the wedge31
angle range; per hypothesis row and view the wedge32 range and the reprojected unit tangent.  Geometry is synthetic but
shaped like the real thing: edges are points in an image, the two epipoles are far outside it, wedge31 of a group is a
thin strip through the image and the rows of a group (view-2 candidates at different depths) sit at different positions
along that strip.
"""
import numpy as np


def _ang(x, y, ex, ey):
    a = np.degrees(np.arctan((y - ey) / (x - ex)))
    return np.where(a < 0, a + 180.0, a)


def make_support_problem(n_views=200, edges_per_view=200000, n_groups=50000, rows_per_group=20, seed=20251109, rows=800, cols=800,
                         wedge_deg=0.026, delta_theta_deg=2.0, parallel_deg=15.0, hit_frac=0.5, pi_truncated=3.1415926):
    """Returns a dict with the arguments of oracle_py.support_test / api.SupportProblem (view-major layouts)."""
    rng = np.random.default_rng(seed)
    V, E, G = n_views, edges_per_view, n_groups
    n_rows = rng.integers(max(1, rows_per_group // 2), rows_per_group * 3 // 2 + 1, G)
    group_off = np.concatenate([[0], np.cumsum(n_rows)]).astype(np.int64)
    H = int(group_off[-1])
    view_off = (np.arange(V + 1, dtype=np.int64) * E)
    ang31 = np.empty(V * E); ang32 = np.empty(V * E); cs = np.empty(V * E); sn = np.empty(V * E)
    wedge31 = np.empty((V, G, 2)); wedge32 = np.empty((V, H, 2)); tgt = np.empty((V, H, 2))
    grp_of_row = np.repeat(np.arange(G), n_rows)
    for v in range(V):
        x = rng.uniform(10, cols - 10, E); y = rng.uniform(10, rows - 10, E); th = rng.uniform(-np.pi / 2, np.pi / 2, E)
        # two far epipoles in roughly orthogonal directions (per view), as for cameras on a sphere around the object
        r1, r2 = rng.uniform(4000, 9000, 2)
        p1 = rng.uniform(0, 2 * np.pi); p2 = p1 + np.pi / 2 + rng.uniform(-0.5, 0.5)
        e1 = (cols / 2 + r1 * np.cos(p1), rows / 2 + r1 * np.sin(p1)); e2 = (cols / 2 + r2 * np.cos(p2), rows / 2 + r2 * np.sin(p2))
        a31 = _ang(x, y, *e1); a32 = _ang(x, y, *e2)
        s = slice(v * E, (v + 1) * E)
        ang31[s] = a31; ang32[s] = a32; cs[s] = np.cos(th); sn[s] = np.sin(th)
        # each group's strip passes through a random point of the image
        gx = rng.uniform(10, cols - 10, G); gy = rng.uniform(10, rows - 10, G)
        c31 = _ang(gx, gy, *e1)
        wedge31[v, :, 0] = c31 - wedge_deg / 2; wedge31[v, :, 1] = c31 + wedge_deg / 2
        # rows: a fraction is centred on a real edge of the strip (so that supports exist), the rest at random positions
        tx = rng.uniform(10, cols - 10, H); ty = rng.uniform(10, rows - 10, H)
        c32 = _ang(tx, ty, *e2)
        tth = rng.uniform(-np.pi / 2, np.pi / 2, H)
        hit = rng.random(H) < hit_frac
        # pick, for "hit" rows, an edge inside the group's strip if there is one
        order = np.argsort(a31, kind="stable")
        a31s = a31[order]
        lo = np.searchsorted(a31s, wedge31[v, grp_of_row, 0], "right"); hi = np.searchsorted(a31s, wedge31[v, grp_of_row, 1], "left")
        has = hit & (hi > lo)
        pick = order[np.minimum(lo + (rng.random(H) * np.maximum(hi - lo, 1)).astype(np.int64), E - 1)]
        c32 = np.where(has, a32[pick] + rng.normal(0, wedge_deg / 6, H), c32)
        tth = np.where(has, th[pick] + rng.normal(0, np.radians(delta_theta_deg) * 0.7, H), tth)
        wedge32[v, :, 0] = c32 - wedge_deg / 2; wedge32[v, :, 1] = c32 + wedge_deg / 2
        tgt[v, :, 0] = np.cos(tth); tgt[v, :, 1] = np.sin(tth)
    thr = float(np.cos(delta_theta_deg / 180 * pi_truncated))
    return dict(group_off=group_off, view_off=view_off, ang31=ang31, ang32=ang32, cs=cs, sn=sn, wedge31=wedge31, wedge32=wedge32,
                tgt=tgt, thr=thr, parallel_deg=float(parallel_deg))

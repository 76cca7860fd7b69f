"""Independent numpy restatement of the angle / wedge / support functions (written from the reference sources, not from the
oracle) against the oracle's stand-alone entry points: guards oracle transcription errors.
getOrientationList.cpp:33-353,391-461; getSupportedEdgels.cpp:32-47; PairEdgeHypo.cpp:88-102."""
import ctypes as C

import numpy as np

D = C.c_double
PD = C.POINTER(C.c_double)


def _p(a):
    return np.ascontiguousarray(a, np.float64).ctypes.data_as(PD)


def np_pair_geometry(R1, T1, R2, T2, K1, K2):
    R21 = R2 @ R1.T                                       # util.cpp:50-56
    C1 = -R1.T @ T1; C2 = -R2.T @ T2
    T21 = R2 @ (C1 - C2)                                  # util.cpp:58-66
    Tx = np.array([[0, -T21[2], T21[1]], [T21[2], 0, -T21[0]], [-T21[1], T21[0], 0]])
    F = np.linalg.inv(K2).T @ (Tx @ R21) @ np.linalg.inv(K1)    # util.cpp:44-48
    e1 = R21.T @ T21; e1 = K1 @ (e1 / e1[2])              # getOrientationList.cpp:47,49
    e2 = K2 @ (T21 / T21[2])                              # :48,50
    return R21, T21, F, e1, e2


def np_ore(x, y, e):                                      # getOrientationList.cpp:391-425
    a = np.degrees(np.arctan((y - e[1]) / (x - e[0])))
    return a + 180 if a < 0 else a


def np_wedge(x, y, F, e1, e2, delta, rows, cols):         # getOrientationList.cpp:33-353, one row
    dy, dx = y - e1[1], x - e1[0]
    d12 = np.sqrt(dy * dy + dx * dx - delta * delta)
    th = np.arcsin(delta / d12); ang = np.arctan(dy / dx)
    pts = []
    for a in (ang + th, ang - th):
        ddx, ddy = abs(np.cos(a) * d12), abs(np.sin(a) * d12)
        if dx < 0: ddx = -ddx
        if dy < 0: ddy = -ddy
        pts.append((ddx + e1[0], ddy + e1[1]))
    lines = [F @ np.array([p[0], p[1], 1.0]) for p in pts]
    ores = []
    for l in lines:
        o = np.degrees(np.arctan(l[1] / l[0]))
        ores.append(o + 180 if o < 0 else o)
    ore_lo, ore_hi = min(ores), max(ores)
    out = []
    for l, ore in zip(lines, (ore_lo, ore_hi)):
        A, B, Cc = l
        cand = [(0.0, -Cc / B), (float(cols), -(Cc + A * cols) / B), (-Cc / A, 0.0), (-(Cc + B * rows) / A, float(rows))]
        hits = [c for k, c in enumerate(cand) if (0 <= (c[1] if k < 2 else c[0]) <= (rows if k < 2 else cols))]
        if len(hits) != 2:
            return None
        d = [(h[0] - e2[0]) ** 2 + (h[1] - e2[1]) ** 2 for h in hits]
        px, py = hits[0] if d[0] < d[1] else hits[1]
        dl = abs(np.cos(np.radians(ore)) * delta)
        if px == 0 or px == cols:
            s = [(py - e2[1] - dl) / (px - e2[0]), (py - e2[1] + dl) / (px - e2[0])]
        else:
            s = [(py - e2[1]) / (px - e2[0] - dl), (py - e2[1]) / (px - e2[0] + dl)]
        for q in s:
            a = np.degrees(np.arctan(q))
            out.append(a + 180 if a < 0 else a)
    return min(out), max(out)


def test_geometry_angles_and_wedges(oracle_mod, tiny_scene):
    scene, _ = tiny_scene
    lib = oracle_mod.load()
    lib.orc_wedge_row.argtypes = [D, D, PD, PD, PD, D, C.c_int, C.c_int, PD, PD, C.POINTER(C.c_int32)]
    lib.orc_ore_angle.restype = D; lib.orc_ore_angle.argtypes = [D, D, D, D]
    n_ok = n_flag = 0
    for (a, b) in [(0, 1), (3, 2), (5, 6), (7, 0)]:
        R21 = np.empty(9); T21 = np.empty(3); F = np.empty(9); e1 = np.empty(3); e2 = np.empty(3)
        lib.orc_pair_geometry(_p(scene.R[a]), _p(scene.T[a]), _p(scene.R[b]), _p(scene.T[b]), _p(scene.K), _p(scene.K), _p(R21), _p(T21),
                              _p(F), _p(e1), _p(e2))
        nR21, nT21, nF, ne1, ne2 = np_pair_geometry(scene.R[a], scene.T[a], scene.R[b], scene.T[b], scene.K, scene.K)
        assert np.allclose(R21.reshape(3, 3), nR21, rtol=0, atol=1e-12) and np.allclose(T21, nT21, rtol=1e-12)
        assert np.allclose(F.reshape(3, 3), nF, rtol=1e-9, atol=1e-18) and np.allclose(e1, ne1, rtol=1e-10) and np.allclose(e2, ne2, rtol=1e-10)
        for x, y, _, _ in scene.edges[b][:100]:
            assert abs(lib.orc_ore_angle(x, y, e2[0], e2[1]) - np_ore(x, y, ne2)) < 1e-9
        for x, y, _, _ in scene.edges[a][:200]:
            lo = D(); hi = D(); fl = C.c_int32()
            lib.orc_wedge_row(x, y, _p(F), _p(e1), _p(e2), 0.4, scene.rows, scene.cols, C.byref(lo), C.byref(hi), C.byref(fl))
            w = np_wedge(x, y, nF, ne1, ne2, 0.4, scene.rows, scene.cols)
            if w is None:
                n_flag += 1
                assert fl.value != 0
            else:
                n_ok += 1
                assert fl.value == 0 and abs(lo.value - w[0]) < 1e-7 and abs(hi.value - w[1]) < 1e-7
    assert n_ok > 600


def test_support_test_against_numpy(oracle_mod):
    """orc_support_test vs a direct numpy statement of EdgeSketch_Core.cpp:348-384 + getSupportedEdgels.cpp:32-47."""
    rng = np.random.default_rng(11)
    V, E, G = 3, 400, 25
    n_rows = rng.integers(1, 6, G); group_off = np.concatenate([[0], np.cumsum(n_rows)]).astype(np.int64); H = int(group_off[-1])
    view_off = np.arange(V + 1, dtype=np.int64) * E
    a31 = rng.uniform(0, 180, V * E); a32 = rng.uniform(0, 180, V * E); th = rng.uniform(-np.pi / 2, np.pi / 2, V * E)
    cs, sn = np.cos(th), np.sin(th)
    c31 = rng.uniform(5, 175, (V, G)); w31 = np.stack([c31 - 4, c31 + 4], -1)
    c32 = rng.uniform(5, 175, (V, H)); w32 = np.stack([c32 - 30, c32 + 30], -1)
    tt = rng.uniform(-np.pi / 2, np.pi / 2, (V, H)); tgt = np.stack([np.cos(tt), np.sin(tt)], -1)
    thr = float(np.cos(15 / 180 * 3.1415926))
    got = oracle_mod.support_test(group_off, view_off, a31, a32, cs, sn, w31, w32, tgt, thr, 15.0, threads=2)
    grp = np.repeat(np.arange(G), n_rows)
    exp = np.full((V, H), -2, np.int32)
    for v in range(V):
        s = slice(v * E, (v + 1) * E)
        for r in range(H):
            l1, h1 = w31[v, grp[r]]; l2, h2 = w32[v, r]
            if max(abs(l1 - l2), abs(l1 - h2), abs(h1 - l2), abs(h1 - h2)) <= 15.0:
                continue
            inl = np.nonzero((a31[s] > l1) & (a31[s] < h1) & (a32[s] > l2) & (a32[s] < h2))[0]
            prev = 0.0
            for i in inl:
                d = abs(tgt[v, r, 0] * cs[s][i] + tgt[v, r, 1] * sn[s][i])
                if d >= thr and d > prev:
                    prev = d; exp[v, r] = i
    assert np.array_equal(got, exp)
    assert (exp != -2).sum() > 20

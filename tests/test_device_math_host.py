"""CPU-side transcription check: the product's per-element FP64 math (cvx_math.cuh, compiled for the host by
tests/host_harness.cpp) must reproduce the oracle bit for bit -- on the host both use glibc, so any difference is a
transcription error, not a libm difference."""
import ctypes as C

import numpy as np

from conftest import make_oracle

D = C.c_double
PD = C.POINTER(C.c_double)


def _p(a):
    return np.ascontiguousarray(a, np.float64).ctypes.data_as(PD)


def _geom(oracle_mod, scene, a, b):
    lib = oracle_mod.load()
    R21 = np.empty(9); T21 = np.empty(3); F = np.empty(9); e1 = np.empty(3); e2 = np.empty(3)
    lib.orc_pair_geometry(_p(scene.R[a]), _p(scene.T[a]), _p(scene.R[b]), _p(scene.T[b]), _p(scene.K), _p(scene.K),
                          _p(R21), _p(T21), _p(F), _p(e1), _p(e2))
    return R21, T21, F, e1, e2


def test_wedge_and_angle_bit_exact(oracle_mod, host_harness, tiny_scene):
    scene, _ = tiny_scene
    olib = oracle_mod.load()
    olib.orc_wedge_row.argtypes = [D, D, PD, PD, PD, D, C.c_int, C.c_int, PD, PD, C.POINTER(C.c_int32)]
    host_harness.hh_wedge_row.argtypes = [D, D, PD, PD, PD, D, C.c_int, C.c_int, PD, PD]
    olib.orc_ore_angle.restype = D; olib.orc_ore_angle.argtypes = [D, D, D, D]
    host_harness.hh_ore_angle.restype = D; host_harness.hh_ore_angle.argtypes = [D, D, D, D]
    n_flag = 0
    for (a, b) in [(0, 1), (0, 3), (1, 5), (2, 7), (6, 0)]:
        _, _, F, e1, e2 = _geom(oracle_mod, scene, a, b)
        for x, y, _, _ in scene.edges[a][:150]:
            lo = D(); hi = D(); fl = C.c_int32(); lo2 = D(); hi2 = D()
            olib.orc_wedge_row(x, y, _p(F), _p(e1), _p(e2), 0.4, scene.rows, scene.cols, C.byref(lo), C.byref(hi), C.byref(fl))
            f2 = host_harness.hh_wedge_row(x, y, _p(F), _p(e1), _p(e2), 0.4, scene.rows, scene.cols, C.byref(lo2), C.byref(hi2))
            assert fl.value == f2
            n_flag += bool(f2)
            if not f2:
                assert lo.value == lo2.value and hi.value == hi2.value
            else:
                assert np.isnan(lo2.value) and np.isnan(hi2.value)
        for x, y, _, _ in scene.edges[b][:150]:
            assert olib.orc_ore_angle(x, y, e2[0], e2[1]) == host_harness.hh_ore_angle(x, y, e2[0], e2[1])
    assert n_flag < 750   # not everything degenerate


def test_fast_wedge_agrees_with_reference_order(oracle_mod, host_harness, tiny_scene, abc_scene):
    """wedge_row_fast (libm round trips removed; the validation kernel's default) against the reference-order wedge_row:
    same degenerate set, bounds equal to libm-noise level (well inside parity.ABS_ANGLE = 1e-10 degrees)."""
    host_harness.hh_wedge_row.argtypes = [D, D, PD, PD, PD, D, C.c_int, C.c_int, PD, PD]
    host_harness.hh_wedge_row_fast.argtypes = [D, D, PD, PD, PD, D, C.c_int, C.c_int, PD, PD]
    worst, n_ok, n_flag, worst_at, worst_far = 0.0, 0, 0, None, 0.0
    rng = np.random.default_rng(3)
    for scene, pairs in ((tiny_scene[0], [(0, 1), (0, 3), (1, 5), (2, 7), (6, 0)]), (abc_scene, [(47, 42), (42, 14), (28, 48), (0, 40), (9, 39)])):
        for (a, b) in pairs:
            _, _, F, e1, e2 = _geom(oracle_mod, scene, a, b)
            pts = [tuple(e[:2]) for e in scene.edges[a][:400]]
            pts += [(e1[0] + dx, e1[1] + dy) for dx, dy in rng.normal(0, 1.0, (40, 2))]        # around / inside the disc about the epipole
            pts += [(float(x), float(y)) for x, y in rng.uniform(0, scene.cols, (200, 2))]
            for x, y in pts:
                lo = D(); hi = D(); lo2 = D(); hi2 = D()
                f1 = host_harness.hh_wedge_row(x, y, _p(F), _p(e1), _p(e2), 0.4, scene.rows, scene.cols, C.byref(lo), C.byref(hi))
                f2 = host_harness.hh_wedge_row_fast(x, y, _p(F), _p(e1), _p(e2), 0.4, scene.rows, scene.cols, C.byref(lo2), C.byref(hi2))
                assert f1 == f2, (x, y, f1, f2)
                n_flag += bool(f1)
                if not f1:
                    d = max(abs(lo.value - lo2.value), abs(hi.value - hi2.value))
                    if d > worst:
                        worst, worst_at = d, (x, y, float(np.hypot(x - e1[0], y - e1[1])), lo.value, hi.value)
                    if np.hypot(x - e1[0], y - e1[1]) > 5.0:
                        worst_far = max(worst_far, d)
                    n_ok += 1
    # within a few pixels of the epipole the wedge is tens of degrees wide and both formulations are ill-conditioned
    assert worst_far < 2e-12, worst_far
    assert n_ok > 2000 and worst < 1e-10, (n_ok, n_flag, worst, worst_at)


def test_correction_reprojection_bit_exact(oracle_mod, host_harness, tiny_scene):
    scene, _ = tiny_scene
    olib = oracle_mod.load()
    cfg = oracle_mod.default_config(n_views=scene.n_views)
    olib.orc_correct_one.argtypes = [C.c_void_p, PD, D, D, D, D, D, PD, PD]
    host_harness.hh_correct_one.argtypes = [PD, D, D, D, D, D, D, D, D, PD, PD]
    olib.orc_reproject_tangent.argtypes = [D] * 6 + [PD] * 7
    host_harness.hh_reproject.argtypes = [D] * 6 + [PD] * 7
    R21, T21, F21, _, _ = _geom(oracle_mod, scene, 0, 1)
    R31, T31, _, _, _ = _geom(oracle_mod, scene, 0, 4)
    Kinv = np.empty(9); olib.orc_inverse3(_p(scene.K), _p(Kinv))
    rng = np.random.default_rng(0)
    n_ok = 0
    for i in range(300):
        x1, y1, th1, _ = scene.edges[0][i]
        j = rng.integers(len(scene.edges[1]))
        x2, y2, th2, _ = scene.edges[1][j]
        # put the view-2 edge near the epipolar line half of the time so every branch is exercised
        if i % 2 == 0:
            l = F21.reshape(3, 3) @ np.array([x1, y1, 1.0])
            y2 = -(l[0] * x2 + l[2]) / l[1] + rng.normal(0, 1.5)
        cx = D(); cy = D(); cx2 = D(); cy2 = D()
        ok1 = olib.orc_correct_one(C.byref(cfg), _p(F21), x1, y1, x2, y2, th2, C.byref(cx), C.byref(cy))
        ok2 = host_harness.hh_correct_one(_p(F21), x1, y1, x2, y2, th2, cfg.location_perturbation, cfg.epip_tangency_displ_thresh,
                                          cfg.orient_perturbation, C.byref(cx2), C.byref(cy2))
        assert ok1 == ok2
        if ok1:
            n_ok += 1
            assert cx.value == cx2.value and cy.value == cy2.value
            t3a = np.empty(2); t3b = np.empty(2)
            olib.orc_reproject_tangent(x1, y1, th1, cx.value, cy.value, th2, _p(scene.K), _p(scene.K), _p(R21), _p(T21), _p(R31), _p(T31), _p(t3a))
            host_harness.hh_reproject(x1, y1, th1, cx.value, cy.value, th2, _p(Kinv), _p(Kinv), _p(R21), _p(T21), _p(R31), _p(T31), _p(t3b))
            assert np.array_equal(t3a, t3b, equal_nan=True)
    assert n_ok > 50


def test_triangulation_tangent_bit_exact_and_kat(oracle_mod, host_harness):
    """Fixed numeric case of the reference (test/test_functions.cpp:117-145: R1,T1,R2,T2,K, 10 point pairs).  The reference
    only prints the results; here the case pins oracle == product math and the reprojection error of the triangulated
    points against an independent numpy SVD of the same DLT system."""
    olib = oracle_mod.load()
    R1 = np.array([0.229714, 0.973258, -1.29728e-07, -0.220415, 0.0520239, 0.974018, -0.947971, 0.223746, -0.226472])
    T1 = np.array([-0.601486, -0.402814, -3.93395])
    R2 = np.array([0.159322, -0.987227, -3.64782e-08, -0.0530916, -0.00856814, 0.998553, 0.985798, 0.159092, 0.0537785])
    T2 = np.array([0.413953, -0.468447, -3.95085])
    K = np.array([2584.93250981950, 0, 249.771375872214, 0, 2584.79186060577, 278.312679379194, 0, 0, 1.0])
    p1 = np.array([[520.590800, 428.498217], [520.581695, 427.998471], [520.573102, 427.498616], [520.564490, 426.998731],
                   [520.555647, 426.498837], [520.645733, 430.996517], [520.634568, 430.496660], [520.623192, 429.996899],
                   [520.611786, 429.497305], [520.600795, 428.997801]])
    p2 = np.array([[240.855540, 429.000920], [240.849805, 428.001160], [240.839928, 427.001388], [240.839928, 427.001388],
                   [240.834057, 426.501440], [240.885916, 432.001500], [240.872839, 431.001501], [240.866802, 430.501425],
                   [240.861729, 430.001200], [240.855540, 429.000920]])
    R21 = (R2.reshape(3, 3) @ R1.reshape(3, 3).T)
    T21 = T2 - R21 @ T1          # test_functions.cpp:143-144
    Kinv = np.empty(9); olib.orc_inverse3(_p(K), _p(Kinv))
    Km = K.reshape(3, 3)
    for a, b in zip(p1, p2):
        g1 = np.empty(3); g2 = np.empty(3)
        olib.orc_triangulate(_p(a), _p(b), _p(R21), _p(T21), _p(K), _p(g1))
        host_harness.hh_triangulate(_p(a), _p(b), _p(R21), _p(T21), _p(Kinv), _p(g2))
        assert np.array_equal(g1, g2)
        # independent check of the solver: numpy SVD of the same A^T A (util.cpp:283-308) gives the same point (1e-6 relative)
        Kin = np.linalg.inv(Km); m1 = Kin @ np.array([a[0], a[1], 1.0]); m2 = Kin @ np.array([b[0], b[1], 1.0])
        A = np.array([[0, -1, m1[1], 0], [1, 0, -m1[0], 0],
                      [m2[1] * R21[2, 0] - R21[1, 0], m2[1] * R21[2, 1] - R21[1, 1], m2[1] * R21[2, 2] - R21[1, 2], m2[1] * T21[2] - T21[1]],
                      [R21[0, 0] - m2[0] * R21[2, 0], R21[0, 1] - m2[0] * R21[2, 1], R21[0, 2] - m2[0] * R21[2, 2], T21[0] - m2[0] * T21[2]]])
        _, _, Vt = np.linalg.svd(A.T @ A)
        ref = Vt[-1, :3] / Vt[-1, 3]
        assert np.linalg.norm(ref - g1) <= 1e-6 * np.linalg.norm(ref)
        assert 3.0 < abs(g1[2]) < 5.0   # cameras sit ~4 units from the object (this dataset has negative depths)
        e1 = np.array([a[0], a[1], 0.3]); e2 = np.array([b[0], b[1], -0.2])
        t1 = np.empty(3); t2 = np.empty(3)
        olib.orc_tangent3d(_p(e1), _p(e2), _p(K), _p(K), _p(R1), _p(T1), _p(R2), _p(T2), _p(t1))
        R21o = np.empty(9); dummy = [np.empty(3), np.empty(9), np.empty(3), np.empty(3)]
        olib.orc_pair_geometry(_p(R1), _p(T1), _p(R2), _p(T2), _p(K), _p(K), _p(R21o), _p(dummy[0]), _p(dummy[1]), _p(dummy[2]), _p(dummy[3]))
        host_harness.hh_tangent3d(_p(e1), _p(e2), _p(Kinv), _p(Kinv), _p(R1), _p(R21o), _p(t2))
        assert np.array_equal(t1, t2)
        assert abs(np.linalg.norm(t1) - 1.0) < 1e-6   # R1 is given to 6 digits only


def test_graph_alignment_arithmetic_equals_the_reference(host_harness):
    """cvx_graph_math.cuh (what k_align_step / k_link_metrics run) against the REFERENCE's own util.cpp functions in oracle/_ref, both on the host
    with glibc: findClosestVectorFromPointToLine, getShortestAlignEulerAnglesDegrees (Rodrigues rotation, Eigen's eulerAngles(0,1,2), the
    shortest-path rule incl. nearly opposite directions) and euler_to_rotation_matrix (AngleAxis -> quaternion product -> matrix): bit for bit."""
    import ctypes as C
    from oracle import ref_py
    if not ref_py.available():
        pytest.skip("oracle/_ref not built and /root/reference absent")
    rng = np.random.default_rng(3)
    P = lambda a: np.ascontiguousarray(a, np.float64).ctypes.data_as(C.c_void_p)
    host_harness.hh_euler_to_rotation.argtypes = [C.c_double, C.c_double, C.c_double, C.c_void_p]
    n_large = 0
    for k in range(4000):
        a = rng.normal(size=3); b = rng.normal(size=3)
        if k % 4 == 1:
            b = a + 0.05 * rng.normal(size=3)                  # neighbouring orientations (the common case of the alignment)
        if k % 4 == 2:
            b = -a + 0.05 * rng.normal(size=3)                 # nearly opposite: the reversed-direction branch (util.cpp:193-203)
        if k % 2 == 0:
            a /= np.linalg.norm(a); b /= np.linalg.norm(b)
        got = np.empty(3)
        host_harness.hh_shortest_align_euler_deg(P(a), P(b), P(got))
        want = ref_py.shortest_align_euler_deg(a, b)
        assert np.array_equal(got, want), (k, a, b, got, want)
        n_large += int(np.abs(want).max() > 45)
        p1, p2 = rng.uniform(-1, 1, 3), rng.uniform(-1, 1, 3)
        host_harness.hh_closest_vector_to_line(P(p1), P(a), P(p2), P(got))
        assert np.array_equal(got, ref_py.closest_vector_to_line(p1, a, p2))
        e = rng.uniform(-0.2, 0.2, 3) if k % 3 else rng.uniform(-3.1, 3.1, 3)
        R = np.empty(9)
        host_harness.hh_euler_to_rotation(float(e[0]), float(e[1]), float(e[2]), P(R))
        assert np.array_equal(R.reshape(3, 3), ref_py.euler_to_rotation(*e)), (k, e)
    assert n_large > 100                                       # the large-angle branches were exercised

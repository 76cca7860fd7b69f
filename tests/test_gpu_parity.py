"""GPU parity tests proper: the CUDA path, called through the C-ABI, against the CPU oracle on the same inputs."""
import json
import os

import numpy as np
import pytest

from conftest import ROOT, make_oracle, pkg
from parity import compare

pytestmark = pytest.mark.gpu


def _gpu(scene, **cfg_kw):
    api = pkg("api")
    cfg = api.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols, **cfg_kw)
    g = api.EdgeSketch(cfg, 0)
    g.keep_intermediates(True)    # parity runs compare wedge32 / reprojected tangents too (production keeps them in registers)
    g.set_cameras(scene.R, scene.T, scene.K)
    for v, e in enumerate(scene.edges):
        g.set_view_edges(v, e)
    return g


def _report(name, rep):
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"parity_{name}.json"), "w") as f:
        json.dump(rep, f, indent=1, default=str)


def test_tiny_scene_full_parity(oracle_mod, tiny_scene):
    scene, w = tiny_scene
    o = make_oracle(oracle_mod, scene, max_support_views=w["max_support_views"])
    g = _gpu(scene, max_support_views=w["max_support_views"])
    for (a, b) in [(0, 1), (2, 3), (5, 1)]:
        o.set_pair(a, b); o.run(mode=1, keep=True)
        g.set_pair(a, b); g.run()
        rep = compare(o, g, check_intermediates=True)
        _report(f"tiny_{a}_{b}", rep)
        assert rep["counts_gpu"]["n_hypotheses"] > 0


def test_large_hypothesis_groups_parity(oracle_mod, data_mod):
    """Dense views: hypothesis groups of 129..256 and 257..512 rows are clustered by 4 / 8 warps per edge inside
    k_cluster_classes, 513..1024 rows by k_cluster_coop (block-cooperative merge, named barriers); all must reproduce the
    oracle's greedy merge order exactly.  Also exercises the packed all-view upload."""
    scene = data_mod.make_synthetic_scene(n_views=6, edges_per_view=40000, rows=800, cols=800, seed=11, n_curves=12)
    o = make_oracle(oracle_mod, scene, max_support_views=2)
    api = pkg("api")
    g = api.EdgeSketch(api.default_config(n_views=6, img_rows=800, img_cols=800, max_support_views=2), 0)
    g.keep_intermediates(True)
    g.set_cameras(scene.R, scene.T, scene.K)
    g.set_all_view_edges(scene.edges)
    o.set_pair(0, 1); g.set_pair(0, 1)
    n = g.run(0, 1500).h1_table()["n"]
    n = n[:1500]
    assert ((n > 128) & (n <= 256)).sum() > 50 and ((n > 256) & (n <= 512)).sum() > 50 and ((n > 512) & (n <= 1024)).sum() > 5, \
        "scene no longer populates every clustering size class"
    o.run(0, 1500, mode=1, keep=True)
    rep = compare(o, g, check_intermediates=True)
    _report("large_groups_0_1", rep)
    assert rep["counts_gpu"]["n_hypotheses"] > 100000


def test_abc_nef_first_pass_parity(oracle_mod, abc_scene):
    """BASELINE.json config 1: ABC-NEF 00000006, configs/3D_Curvix_ABC_NEF.yaml parameters, initial pair (47, 42)."""
    o = make_oracle(oracle_mod, abc_scene)
    g = _gpu(abc_scene)
    o.set_pair(47, 42); o.run(mode=1, keep=True)
    g.set_pair(47, 42); g.run()
    rep = compare(o, g, check_intermediates=True)
    _report("abc_47_42", rep)
    c = rep["counts_gpu"]
    assert c["n_hypotheses"] == 54254 and c["n_validations"] == 2604192   # oracle-generated golden counts for this pass
    # Disagreements are allowed only within 1e-9 of a threshold (compare() asserts that) and are counted: on this pass they
    # are the representatives of 2-member clusters, an exact tie that the last bit of tan() decides (oracle margin ~1e-14).
    assert rep["threshold_disagreements_form"] <= 0.02 * c["n_h1_processed"] and rep["threshold_disagreements_support"] <= 5
    assert rep["n_pairs_compared"] >= 0.97 * rep["n_pairs_oracle"]


@pytest.mark.parametrize("pair", [(1, 4), (31, 8), (9, 6), (19, 17)])
def test_abc_later_passes_of_the_reference_schedule_parity(oracle_mod, abc_scene, pair):
    """The other four pairs the reference's own run of ABC-NEF 00000006 visits (tests/golden/ref_abc_full_run.npz): every stage against the port
    oracle, disagreements only within 1e-9 of a threshold (asserted by compare) and counted."""
    o = make_oracle(oracle_mod, abc_scene)
    g = _gpu(abc_scene)
    o.set_pair(*pair); o.run(mode=1, keep=True)
    g.set_pair(*pair); g.run()
    rep = compare(o, g, check_intermediates=True)
    _report(f"abc_{pair[0]}_{pair[1]}", rep)
    c = rep["counts_gpu"]
    assert rep["threshold_disagreements_form"] <= 0.03 * c["n_h1_processed"] and rep["threshold_disagreements_support"] <= 5
    assert rep["n_pairs_compared"] >= 0.96 * rep["n_pairs_oracle"]


@pytest.mark.parametrize("pair", [(47, 42), (28, 48)])
def test_abc_against_reference_goldens(abc_scene, pair):
    """The CUDA path against vectors written by the REFERENCE ITSELF (tests/golden/ref_abc_*.npz: its unmodified sources built against
    oracle/ref_shim): accepted view-1 edges, the support index of every accepted row in every validation view and the support counts
    bit-exact; the view-2 representative exact except for tied 2-member clusters (the other member of the same cluster, counted);
    3-D points / tangents to 1e-6 on the rows with the same representative."""
    gold = np.load(os.path.join(ROOT, "tests", "golden", f"ref_abc_{pair[0]}_{pair[1]}.npz"))
    g = _gpu(abc_scene)
    g.set_pair(*pair); g.run()
    order = g.reference_order()
    p, f, r, t = g.pairs(), g.finalize(), g.rows(), g.h1_table()
    assert len(order) == len(gold["pair_h1"])
    assert np.array_equal(p["h1"][order], gold["pair_h1"])
    assert np.array_equal(p["supports"][order], gold["pair_supports"])
    same = p["h2rep"][order] == gold["pair_h2rep"]
    n_tied = int((~same).sum())
    assert n_tied <= 0.03 * len(order)
    for k in np.nonzero(~same)[0]:                       # the reference's representative is another member of the SAME cluster
        row = int(p["row"][order[k]]); h1 = int(p["h1"][order[k]])
        s0, n = int(t["off"][h1]), int(t["n"][h1])
        members = r["h2_orig"][s0:s0 + n][r["label"][s0:s0 + n] == r["label"][row]]
        assert int(gold["pair_h2rep"][k]) in members.tolist() and len(members) == 2
    ok = same & (f["nan_flag"][order] == 0)
    for key in ("gamma_world", "tangent_world"):
        a, b = f[key][order][ok], gold[key][ok]
        assert np.max(np.abs(a - b) / np.maximum(np.linalg.norm(b, axis=1, keepdims=True), 1e-3)) <= 1e-6, key
    _report(f"abc_{pair[0]}_{pair[1]}_vs_reference", {"pairs": int(len(order)), "tied_representatives": n_tied})


def test_strict_libm_order_switch(oracle_mod, abc_scene):
    """cvx_config.strict_libm_order = 1: the validation kernel evaluates the view-2 wedge in the reference's libm call order
    (wedge_row) instead of wedge_row_fast; both must satisfy the parity rule, and they select the same pairs on this pass."""
    o = make_oracle(oracle_mod, abc_scene)
    o.set_pair(47, 42); o.run(mode=1, keep=True)
    sel = []
    for strict in (1, 0):
        g = _gpu(abc_scene, strict_libm_order=strict)
        g.set_pair(47, 42); g.run()
        rep = compare(o, g, check_intermediates=True)
        _report(f"abc_47_42_strict{strict}", rep)
        p = g.pairs()
        sel.append(np.stack([p["h1"], p["h2rep"]], 1))
        g.close()
    assert np.array_equal(sel[0], sel[1])


def test_abc_nef_other_pairs(oracle_mod, abc_scene):
    o = make_oracle(oracle_mod, abc_scene)
    g = _gpu(abc_scene)
    for (a, b) in [(28, 48), (0, 40), (9, 39)]:   # pairs the authors' run visited (doc/00000006.png legend)
        o.set_pair(a, b); o.run(mode=1, keep=True)
        g.set_pair(a, b); g.run()
        rep = compare(o, g)
        _report(f"abc_{a}_{b}", rep)


def test_h1_range_sharding_is_exact(oracle_mod, abc_scene):
    """Sharding by view-1 edge range (the multi-GPU decomposition) reproduces the unsharded pass exactly."""
    g = _gpu(abc_scene)
    g.set_pair(47, 42); g.run()
    full = g.pairs(); fin = g.finalize(); E1 = g.n_edges[47]
    parts = []
    for (a, b) in [(0, E1 // 3), (E1 // 3, 2 * E1 // 3), (2 * E1 // 3, E1)]:
        g.run(a, b)
        p = g.pairs(); f = g.finalize()
        parts.append((p, f))
    for k in ("h1", "h2rep", "count"):
        assert np.array_equal(np.concatenate([p[k] for p, _ in parts]), full[k])
    assert np.array_equal(np.concatenate([p["supports"] for p, _ in parts]), full["supports"])
    for k in ("gamma_world", "tangent_world"):
        assert np.array_equal(np.concatenate([f[k] for _, f in parts]), fin[k])


def test_intermediates_off_gives_identical_results(oracle_mod, abc_scene):
    """The production path (wedge32 / tangent kept in registers) and the parity path (materialised) agree bit for bit."""
    g = _gpu(abc_scene)
    g.set_pair(47, 42); g.run()
    s_full = g.support_matrix(); p_full = g.pairs()
    g.keep_intermediates(False)
    g.run()
    assert np.array_equal(g.support_matrix(), s_full)
    assert np.array_equal(g.pairs()["row"], p_full["row"]) and np.array_equal(g.pairs()["supports"], p_full["supports"])
    with pytest.raises(Exception):
        g.intermediates()


def test_edge_cases(oracle_mod, data_mod):
    """Empty and ragged inputs: a view with zero edges, a hypothesis view with no candidates, an empty H1 range."""
    scene = data_mod.make_synthetic_scene(n_views=6, edges_per_view=200, seed=3, n_curves=4)
    scene.edges[3] = scene.edges[3][:0]          # a validation view without edges
    scene.edges[4] = scene.edges[4][:17]         # ragged
    o = make_oracle(oracle_mod, scene, max_support_views=1)
    g = _gpu(scene, max_support_views=1)
    o.set_pair(0, 1); o.run(mode=1, keep=True)
    g.set_pair(0, 1); g.run()
    compare(o, g, check_intermediates=True)
    g.run(5, 5)                                   # empty range
    assert g.counts()["n_hypotheses"] == 0 and g.counts()["n_pairs"] == 0
    scene.edges[1] = scene.edges[1][:0]           # hypothesis view 2 empty -> no candidates at all
    o = make_oracle(oracle_mod, scene, max_support_views=1)
    g = _gpu(scene, max_support_views=1)
    o.set_pair(0, 1); o.run(mode=1, keep=True)
    g.set_pair(0, 1); g.run()
    assert g.counts()["n_hypotheses"] == 0 == o.counts()["n_hypotheses"]


def test_support_problem_against_oracle(oracle_mod):
    """Stand-alone support test (BASELINE.json config 5 shape, scaled down) against orc_support_test."""
    api = pkg("api"); synth = pkg("support_synth")
    prob = synth.make_support_problem(n_views=6, edges_per_view=5000, n_groups=400, rows_per_group=12, seed=5, delta_theta_deg=2.0)
    exp = oracle_mod.support_test(**prob)
    cfg = api.default_config(n_views=8)
    g = api.EdgeSketch(cfg, 0)
    sp = api.SupportProblem(g, prob["group_off"], prob["view_off"], prob["ang31"], prob["ang32"], prob["cs"], prob["sn"],
                            prob["wedge31"], prob["wedge32"], prob["tgt"], prob["thr"], prob["parallel_deg"])
    sp.run()
    got = sp.result()
    assert np.array_equal(got, exp)
    assert (exp != -2).sum() > 0
    sp.close()


def test_support_problem_cfg5_width(oracle_mod):
    """BASELINE.json config 5 at full width per view (200 000 edges, ~1 M pre-formed hypothesis rows in 50 000 groups, delta_theta 2 deg) over 20
    validation views: (1) the oracle's Theta(E3) scans on a bounded sample of groups in two views, exact; (2) size-independent properties of EVERY
    answer of two whole views, checked in numpy: a returned edge lies strictly inside both wedges and reaches the orientation threshold, and no
    inlier of the row beats it; (3) a second run returns the same matrix."""
    api = pkg("api"); synth = pkg("support_synth")
    prob = synth.make_support_problem(n_views=20, edges_per_view=200000, n_groups=50000, rows_per_group=20, seed=20251109, delta_theta_deg=2.0)
    g = api.EdgeSketch(api.default_config(n_views=22), 0)
    sp = api.SupportProblem(g, prob["group_off"], prob["view_off"], prob["ang31"], prob["ang32"], prob["cs"], prob["sn"],
                            prob["wedge31"], prob["wedge32"], prob["tgt"], prob["thr"], prob["parallel_deg"])
    sp.run()
    got = sp.result()
    H, E = int(prob["group_off"][-1]), 200000
    assert got.shape == (20, H) and H > 900000
    sp.run()
    assert np.array_equal(sp.result(), got)
    grp = np.repeat(np.arange(50000), np.diff(prob["group_off"]))
    for v in (0, 13):
        gs = 300                                                        # (1) oracle on the first groups of this view
        hs = int(prob["group_off"][gs])
        sl = slice(v * E, (v + 1) * E)
        sub = dict(group_off=prob["group_off"][:gs + 1], view_off=np.array([0, E], np.int64), ang31=prob["ang31"][sl], ang32=prob["ang32"][sl],
                   cs=prob["cs"][sl], sn=prob["sn"][sl], wedge31=prob["wedge31"][v:v + 1, :gs], wedge32=prob["wedge32"][v:v + 1, :hs],
                   tgt=prob["tgt"][v:v + 1, :hs], thr=prob["thr"], parallel_deg=prob["parallel_deg"])
        exp = oracle_mod.support_test(**sub)
        assert np.array_equal(exp[0], got[v, :hs])
        a31, a32, cs, sn = prob["ang31"][sl], prob["ang32"][sl], prob["cs"][sl], prob["sn"][sl]   # (2) properties of every row of the view
        w1, w2, tg, r = prob["wedge31"][v][grp], prob["wedge32"][v], prob["tgt"][v], got[v]
        hit = r != -2
        assert 0.05 * H < hit.sum() < 0.95 * H
        e = r[hit]
        assert np.all((a31[e] > w1[hit, 0]) & (a31[e] < w1[hit, 1]) & (a32[e] > w2[hit, 0]) & (a32[e] < w2[hit, 1]))
        d = np.abs(tg[hit, 0] * cs[e] + tg[hit, 1] * sn[e])
        assert np.all(d >= prob["thr"])
        # rows without support: brute force over a random subset, no edge of the view qualifies
        miss = np.nonzero(~hit)[0]
        for h in np.random.default_rng(v).choice(miss, 200, replace=False):
            inl = (a31 > w1[h, 0]) & (a31 < w1[h, 1]) & (a32 > w2[h, 0]) & (a32 < w2[h, 1])
            par = max(abs(w1[h, 0] - w2[h, 0]), abs(w1[h, 0] - w2[h, 1]), abs(w1[h, 1] - w2[h, 0]), abs(w1[h, 1] - w2[h, 1])) <= prob["parallel_deg"]
            assert par or not np.any(np.abs(tg[h, 0] * cs[inl] + tg[h, 1] * sn[inl]) >= prob["thr"])
        for h in np.random.default_rng(v + 1).choice(np.nonzero(hit)[0], 200, replace=False):   # supported rows: the returned edge is the first maximum
            inl = np.nonzero((a31 > w1[h, 0]) & (a31 < w1[h, 1]) & (a32 > w2[h, 0]) & (a32 < w2[h, 1]))[0]
            dd = np.abs(tg[h, 0] * cs[inl] + tg[h, 1] * sn[inl])
            assert r[h] == inl[np.argmax(dd)]
    sp.close(); g.close()


def test_abc_against_committed_goldens(abc_scene):
    """CUDA path vs tests/golden/oracle_abc_47_42.npz (no oracle run involved).  Index work must be identical; the only
    tolerated difference is the representative of tied 2-member clusters (see test_abc_nef_first_pass_parity)."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "oracle_abc_47_42.npz"))
    gold = dict(zip(g["count_names"].tolist(), g["counts"].tolist()))
    s = _gpu(abc_scene)
    s.set_pair(47, 42); s.run()
    c = s.counts()
    for k in ("n_h1_total", "n_h1_processed", "n_candidates", "n_hypotheses", "n_validations", "n_unique_hypotheses", "n_degenerate_wedges"):
        assert c[k] == gold[k], k
    t, r = s.h1_table(), s.rows()
    assert np.array_equal(t["m"], g["m"]) and np.array_equal(t["n"], g["n"])
    assert np.array_equal(r["label"], g["label"]) and np.array_equal(r["h2_orig"], g["h2_orig"])
    assert (r["h2_rep"] != g["h2_rep"]).mean() < 0.01
    same_rep = r["h2_rep"] == g["h2_rep"]
    assert (r["count"][same_rep] != g["row_count"][same_rep]).mean() < 1e-3
    p = s.pairs(); f = s.finalize()
    key_g = {(int(a), int(b)) for a, b in zip(g["pair_h1"], g["pair_h2rep"])}
    key_s = {(int(a), int(b)) for a, b in zip(p["h1"], p["h2rep"])}
    assert len(key_g & key_s) >= 0.98 * len(key_g)
    # geometry of the common pairs (first occurrence) within 1e-6 relative
    first_g = {}
    for i, k in enumerate(zip(g["pair_h1"].tolist(), g["pair_h2rep"].tolist())):
        first_g.setdefault(k, i)
    idx_s, idx_g = [], []
    seen = set()
    for i, k in enumerate(zip(p["h1"].tolist(), p["h2rep"].tolist())):
        if k in first_g and k not in seen:
            seen.add(k); idx_s.append(i); idx_g.append(first_g[k])
    a, b = f["gamma_world"][idx_s], g["gamma_world"][idx_g]
    assert np.max(np.abs(a - b) / np.maximum(np.linalg.norm(b, axis=1, keepdims=True), 1e-12)) <= 1e-6
    a, b = f["tangent_world"][idx_s], g["tangent_world"][idx_g]
    assert np.max(np.abs(a - b)) <= 1e-6


@pytest.mark.parametrize("name", ["cfg4", "cfg3"])
def test_full_size_properties(oracle_mod, data_mod, name):
    """BASELINE.json configs 4 and 3 at full size (100 views x 50k edges/view; DTU-shaped 49 views x 100k edges/view 1600x1200 with the DTU
    parameters; one pair each): the oracle cannot finish this in seconds, so
    the size-independent properties are checked instead -- determinism (two runs identical), exact view-1-range sharding,
    vote consistency of every accepted pair, finite unit tangents -- and a bounded view-1 range is compared with the oracle."""
    w = data_mod.workload(name)
    scene = data_mod.make_synthetic_scene(**w["scene"])
    api = pkg("api")
    cfg = api.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols, max_support_views=w["max_support_views"])
    g = api.EdgeSketch(cfg, 0)
    g.set_cameras(scene.R, scene.T, scene.K)
    g.set_all_view_edges(scene.edges)
    g.set_pair(0, 1)
    E1 = g.n_edges[0]
    g.run(); c = g.counts(); full = g.pairs(); fin = g.finalize()
    assert c["n_validations"] == c["n_hypotheses"] * (scene.n_views - 2) and c["n_hypotheses"] > 10_000_000
    g.run(); again = g.pairs(); fin2 = g.finalize()
    for k in ("h1", "h2rep", "count", "supports"):
        assert np.array_equal(full[k], again[k]), k
    assert np.array_equal(fin["gamma_world"], fin2["gamma_world"], equal_nan=True)
    parts = []
    for (a, b) in [(0, E1 // 2), (E1 // 2, E1)]:
        g.run(a, b); parts.append((g.pairs(), g.finalize()))
    for k in ("h1", "h2rep", "count", "supports"):
        assert np.array_equal(np.concatenate([p[0][k] for p in parts]), full[k]), k
    assert np.array_equal(np.concatenate([p[1]["tangent_world"] for p in parts]), fin["tangent_world"], equal_nan=True)
    # vote consistency: count = number of supporting views >= N_support; support indices address real edges of their view
    sup = full["supports"]
    assert np.array_equal((sup != -2).sum(1), full["count"]) and (full["count"] >= w["max_support_views"]).all()
    vali = [v for v in range(scene.n_views) if v not in (0, 1)]
    for j, v in enumerate(vali[:8]):
        s = sup[:, j]
        assert ((s == -2) | ((s >= 0) & (s < len(scene.edges[v])))).all()
    ok = fin["nan_flag"] == 0
    tn = np.linalg.norm(fin["tangent_world"][ok], axis=1)
    assert ok.mean() > 0.99 and np.allclose(tn, 1.0, atol=1e-9)
    # bounded oracle comparison on the same full-size scene
    o = make_oracle(oracle_mod, scene, max_support_views=w["max_support_views"])
    o.set_pair(0, 1); o.run(0, 24, mode=1, keep=True)
    g.keep_intermediates(True)
    g.run(0, 24)
    rep = compare(o, g, check_intermediates=True)
    _report(f"{name}_0_1_h1_0_24", rep)
    g.close()


def test_global_memory_clustering_fallback(oracle_mod, tiny_scene):
    """MAX_CLUSTER_SIZE > 32 (member lists no longer fit a warp) routes every hypothesis group to the global-memory clustering
    fallback -- the same path groups of more than 1024 rows take; it must reproduce the oracle's merge order as well."""
    scene, w = tiny_scene
    o = make_oracle(oracle_mod, scene, max_support_views=w["max_support_views"], max_cluster_size=40)
    g = _gpu(scene, max_support_views=w["max_support_views"], max_cluster_size=40)
    o.set_pair(0, 1); o.run(mode=1, keep=True)
    g.set_pair(0, 1); g.run()
    rep = compare(o, g, check_intermediates=True)
    _report("tiny_0_1_maxcluster40", rep)
    assert rep["counts_gpu"]["n_hypotheses"] > 0


def test_dtu_shaped_scene_parity(oracle_mod, data_mod):
    """DTU-shaped inputs (configs/3D_Curvix_DTU.yaml): non-square image, fx != fy, off-centre principal point, N_support = 6;
    full parity against the oracle on two hypothesis pairs, with the packed all-view upload."""
    scene = data_mod.make_synthetic_scene(n_views=14, edges_per_view=2500, rows=1200, cols=1600, seed=31, n_curves=16,
                                          fy_scale=922.616455078125 / 925.546142578125, c_shift=(7.4, -5.9))
    o = make_oracle(oracle_mod, scene, max_support_views=6)
    api = pkg("api")
    g = api.EdgeSketch(api.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols, max_support_views=6), 0)
    g.keep_intermediates(True)
    g.set_cameras(scene.R, scene.T, scene.K)
    g.set_all_view_edges(scene.edges)
    for (a, b) in [(0, 1), (6, 7)]:
        o.set_pair(a, b); o.run(mode=1, keep=True)
        g.set_pair(a, b); g.run()
        rep = compare(o, g, check_intermediates=True)
        _report(f"dtu_shaped_{a}_{b}", rep)
        assert rep["counts_gpu"]["n_hypotheses"] > 1000
    g.close()

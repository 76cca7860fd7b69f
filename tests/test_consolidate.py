"""EdgeMapping::consolidate_3D_edges of the drop-in (links from curvelets, prunings and alignment on the device through
cvx_graph_*, connectivity graph / curve tracing / merging / writers on the host) against the REFERENCE's own consolidate_3D_edges
(its unmodified edge_mapping.cpp in oracle/_ref) on the SAME 3-D edge -> 2-D edge links.  The drop-in keeps the reference's container orders, so
the comparison is file against file: nodes, the three link graphs, aligned locations / orientations after 1000 iterations, the connectivity
graph, the traced curves and final_curves.txt."""
import gzip
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, pkg

sys.path.insert(0, os.path.join(ROOT, "scripts"))

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dataset(abc_scene, tmp_path_factory):
    from test_host_io import ABC_YAML
    if getattr(abc_scene, "curvelets", None) is None:
        pytest.skip("curvelet fixture missing")
    root = tmp_path_factory.mktemp("consolidate")
    abc_scene.write_reference_layout(str(root), dataset="ABC-NEF", scene="00000006", thresh=1)
    cfg = root / "abc.yaml"
    cfg.write_text(ABC_YAML.replace("/some/where/Datasets/", str(root) + "/"))
    return root, cfg


def test_consolidation_equals_the_reference_on_the_same_links(abc_scene, dataset):
    """One sketch pass of the drop-in (pair (47,42)); its links go to the reference's EdgeMapping; every output file of the two consolidations
    is compared (3-D edge graphs, smoothing, connectivity graph, curves, final_curves.txt)."""
    from oracle import ref_py
    import consolidate_check as cc
    if not ref_py.available():
        pytest.skip("oracle/_ref not built and /root/reference absent")
    root, cfg = dataset
    exe = pkg("build").build_host(force=False)
    out, links = root / "out_one_pass", root / "links.txt"
    r = subprocess.run([exe, "-c", str(cfg), "--outputs", str(out), "--max-passes", "1", "--support-links", str(links), "--quiet"],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "final curves" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    L = np.loadtxt(links).reshape(-1, 13)            # X Y Z image idx x y tx ty tz ux uy utheta
    assert len(L) > 100000
    rec = np.concatenate([L[:, 0:3], L[:, 7:10], L[:, 5:7], L[:, 10:13], np.zeros((len(L), 1))], axis=1)
    ref = ref_py.Reference.for_scene(abc_scene, dataset_path=str(root) + "/")
    ref.mapping_add_many(rec, L[:, 3].astype(np.int32), L[:, 4].astype(np.int32))
    ref_out = ref.consolidate()
    rep, ok = cc.compare_dirs(str(out), ref_out)
    print(rep)
    assert ok, rep
    assert rep["final_curves"]["same_curve_and_node_ids"] and rep["final_curves"]["equal_to_print_precision"]
    assert rep["3D_edges_after_smoothing"]["equal_to_print_precision"] and rep["3D_edge_pruned_graph_by_proj"]["links"] > 1000
    ref.close()


def test_device_graph_stages_against_numpy(abc_scene):
    """The device stages on a small synthetic graph against a plain numpy restatement: links from curvelet records, link statistics and the 3-D
    pruning, the projection pruning, and neighbour lists; one alignment step leaves isolated nodes untouched."""
    api = pkg("api")
    rng = np.random.default_rng(11)
    n, n_keys, n_cv = 400, 300, 2000
    loc = rng.uniform(-0.5, 0.5, (n, 3)); tan = rng.normal(size=(n, 3)); tan /= np.linalg.norm(tan, axis=1, keepdims=True)
    cnt = rng.integers(0, 6, n_keys)
    key_off = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int64)
    key_nodes = rng.integers(0, n, key_off[-1]).astype(np.int32)
    cv_a = rng.integers(-1, n_keys, n_cv).astype(np.int32); cv_b = rng.integers(0, n_keys, n_cv).astype(np.int32)
    c = api.default_config(n_views=abc_scene.n_views, img_rows=abc_scene.rows, img_cols=abc_scene.cols)
    sk = api.EdgeSketch(c, 0)
    g = api.EdgeGraph(sk, loc, tan)
    L0 = g.build_links(key_off, key_nodes, cv_a, cv_b)
    want = set()
    first = {}
    pos = 0
    for a, b in zip(cv_a, cv_b):
        recs = ([] if a < 0 else list(key_nodes[key_off[a]:key_off[a + 1]])) + list(key_nodes[key_off[b]:key_off[b + 1]])
        for i in range(len(recs) - 1):
            for j in range(i + 1, len(recs)):
                x, y = int(recs[i]), int(recs[j])
                if x != y:                                   # random doubles: distinct points never share a coordinate
                    k = (min(x, y), max(x, y)); want.add(k); first.setdefault(k, pos)
                pos += 1
    a0, b0 = g.links(0)
    assert L0 == len(want) and set(zip(a0.tolist(), b0.tolist())) == want
    fp, k3, kp = g.link_details()
    assert [first[k] for k in zip(a0.tolist(), b0.tolist())] == fp.tolist()
    K = np.asarray(abc_scene.K, np.float64).reshape(-1)[:9].reshape(3, 3)
    st = g.prune(abc_scene.R, abc_scene.T, K)
    _, k3, kp = g.link_details()
    # numpy restatement of the metrics (edge_mapping.cpp:247-277) and of the thresholds (:319)
    sw = (loc[a0] > loc[b0])
    lex = sw[:, 0] | ((loc[a0, 0] == loc[b0, 0]) & sw[:, 1])
    f, s = np.where(lex, b0, a0), np.where(lex, a0, b0)
    p1, p2, t1, t2 = loc[f], loc[s], tan[f], tan[s]
    ang = np.degrees(np.arccos(np.clip(np.abs((t1 * t2).sum(1)), -1, 1))) * np.pi / 3.1415926
    al = ((p1 - p2) * t2).sum(1)
    nd = np.linalg.norm(p2 + al[:, None] * t2 - p1, axis=1)
    td = np.linalg.norm(p1 - p2, axis=1)
    assert np.isclose(st["mu_n"], nd.mean(), rtol=1e-12) and np.isclose(st["sigma_t"], td.std(), rtol=1e-10) and np.isclose(st["mu_theta"], ang.mean(), rtol=1e-10)
    keep = (nd < nd.mean() + nd.std()) & (td < td.mean() + td.std()) & (ang < ang.mean() + ang.std())
    near = np.abs(nd - nd.mean() - nd.std()) + np.abs(td - td.mean() - td.std()) < 1e-12
    assert np.array_equal(k3[~near], keep[~near]) and st["n_after_3d"] == int(k3.sum())
    proj = np.ones(len(a0), bool)
    for v in range(abc_scene.n_views):
        q1 = (K @ (abc_scene.R[v] @ p1.T + abc_scene.T[v][:, None])).T; q2 = (K @ (abc_scene.R[v] @ p2.T + abc_scene.T[v][:, None])).T
        d = np.linalg.norm(q1[:, :2] / q1[:, 2:3] - q2[:, :2] / q2[:, 2:3], axis=1)
        proj &= ~(d > 6)
    assert np.array_equal(kp, proj)
    a2, b2 = g.links(2)
    assert st["n_after_projection"] == len(a2) == int((k3 & kp).sum())
    off, nb = g.neighbors(len(a2))
    deg = np.bincount(np.concatenate([a2, b2]), minlength=n)
    assert np.array_equal(np.diff(off), deg) and all(np.all(np.diff(nb[off[i]:off[i + 1]]) > 0) for i in range(n))
    lo, oo, _ = g.align(tan, iterations=1)
    iso = deg == 0
    assert np.array_equal(lo[iso], loc[iso]) and np.array_equal(oo[iso], tan[iso])
    g.close(); sk.close()

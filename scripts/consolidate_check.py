#!/usr/bin/env python
"""Compares the drop-in's EdgeMapping::consolidate_3D_edges (device graph + host curve tracing) with a directory of the REFERENCE's output
files for the same 3-D edge -> 2-D edge links.

    python scripts/consolidate_check.py LINKS_FILE REFERENCE_OUTPUT_DIR [--keep DIR]

LINKS_FILE is curvix_main's --support-links format; REFERENCE_OUTPUT_DIR holds what the reference's consolidate_3D_edges wrote for the same
links (oracle/ref_py.Reference.mapping_add_many + consolidate).  Prints one JSON line; exit code 1 when a comparison fails."""
import importlib
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def numeric_rows(path, skip=0, sep=None):
    rows = []
    for l in open(path).read().splitlines()[skip:]:
        l = l.replace(";", " ")
        if l.strip():
            rows.append([float(x) for x in l.split(sep)])
    return rows


def compare_dirs(mine, ref):
    """Returns (report, ok).  Files carry 6 significant digits; numeric comparisons allow one unit of the last printed digit."""
    rep, ok = {}, True

    def close(a, b):
        a, b = np.asarray(a, float), np.asarray(b, float)
        return a.shape == b.shape and bool(np.all(np.abs(a - b) <= 2e-6 * np.maximum(1.0, np.abs(b)) + 1e-9))

    for name in ("3D_edges_before_smoothing", "3D_edges_after_smoothing"):
        a, b = numeric_rows(os.path.join(mine, name + ".txt")), numeric_rows(os.path.join(ref, name + ".txt"))
        same = len(a) == len(b) and close(a, b)
        rep[name] = {"rows": len(a), "rows_reference": len(b), "equal_to_print_precision": same,
                     "max_abs_diff": float(np.max(np.abs(np.asarray(a) - np.asarray(b)))) if len(a) == len(b) and len(a) else None}
        ok &= same
    for name in ("3D_edge_graph", "3D_edge_pruned_graph", "3D_edge_pruned_graph_by_proj"):
        pa, pb = os.path.join(mine, name + ".txt"), os.path.join(ref, name + ".txt")
        if not (os.path.exists(pa) and os.path.exists(pb)):
            continue
        ta, tb = open(pa).read(), open(pb).read()
        na, nb = ta.count("\n"), tb.count("\n")
        rep[name] = {"links": na, "links_reference": nb, "identical_text": ta == tb}
        if ta != tb:   # same links in the same order up to the last printed digit?
            a, b = numeric_rows(pa), numeric_rows(pb)
            rep[name]["equal_to_print_precision"] = len(a) == len(b) and close(a, b)
            ok &= rep[name]["equal_to_print_precision"]
    a, b = numeric_rows(os.path.join(mine, "connectivity_graph.txt"), 2), numeric_rows(os.path.join(ref, "connectivity_graph.txt"), 2)
    A, B = np.asarray(a), np.asarray(b)
    topo = A.shape == B.shape and bool(np.all(A[:, [0, 7, 11]] == B[:, [0, 7, 11]]))          # node, left neighbour, right neighbour
    rep["connectivity_graph"] = {"nodes": len(a), "nodes_reference": len(b), "same_order_and_neighbours": topo,
                                 "equal_to_print_precision": A.shape == B.shape and close(A, B)}
    ok &= topo
    for name in ("curves_from_connectivity_graph", "final_curves"):
        ta, tb = open(os.path.join(mine, name + ".txt")).read(), open(os.path.join(ref, name + ".txt")).read()
        a, b = numeric_rows(os.path.join(mine, name + ".txt"), 2), numeric_rows(os.path.join(ref, name + ".txt"), 2)
        A, B = np.asarray(a), np.asarray(b)
        ids = A.shape == B.shape and bool(np.all(A[:, :2] == B[:, :2]))
        rep[name] = {"curves": int(A[:, 0].max()) + 1 if len(A) else 0, "curves_reference": int(B[:, 0].max()) + 1 if len(B) else 0,
                     "rows": len(a), "rows_reference": len(b), "same_curve_and_node_ids": ids, "identical_text": ta == tb,
                     "equal_to_print_precision": A.shape == B.shape and close(A, B)}
        ok &= ids
    ta, tb = open(os.path.join(mine, "curve_indices_for_merging.txt")).read(), open(os.path.join(ref, "curve_indices_for_merging.txt")).read()
    rep["curve_indices_for_merging"] = {"identical_text": ta == tb}
    ok &= ta == tb
    return rep, ok


def write_abc_dataset(root):
    data = importlib.import_module("3d_curvix_b200.data")
    sc = data.Scene.load_npz(os.path.join(ROOT, "tests", "golden", "abc_nef_00000006_t1.npz"), "ABC-NEF/00000006")
    sc.load_curvelets(os.path.join(ROOT, "tests", "golden", "abc_nef_00000006_curvelets_t1.npz"))
    sc.write_reference_layout(root, dataset="ABC-NEF", scene="00000006", thresh=1)
    from test_host_io import ABC_YAML
    cfg = os.path.join(root, "abc.yaml")
    open(cfg, "w").write(ABC_YAML.replace("/some/where/Datasets/", root + "/"))
    return sc, cfg


if __name__ == "__main__":
    links, ref_dir = sys.argv[1], sys.argv[2]
    keep = sys.argv[sys.argv.index("--keep") + 1] if "--keep" in sys.argv else None
    exe = importlib.import_module("3d_curvix_b200.build").build_host(force=False)
    with tempfile.TemporaryDirectory() as tmp:
        _, cfg = write_abc_dataset(tmp)
        out = keep or os.path.join(tmp, "outputs")
        r = subprocess.run([exe, "-c", cfg, "--outputs", out, "--consolidate-links", links], capture_output=True, text=True, timeout=900)
        print(r.stdout[-1500:], file=sys.stderr)
        if r.returncode != 0:
            print(json.dumps({"error": (r.stderr or r.stdout)[-800:]}))
            sys.exit(1)
        rep, ok = compare_dirs(out, ref_dir)
        rep["ok"] = ok
        print(json.dumps(rep))
        sys.exit(0 if ok else 1)

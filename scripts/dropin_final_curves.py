"""The drop-in boundary, end to end, on the reference's OWN host consumer.

A producer of the edge-sketch path -- the CUDA path through the C-ABI (--gpu) or the port oracle (default, CPU) -- runs ABC-NEF 00000006
over the reference's pass schedule; every accepted row is handed to the reference's EdgeMapping::add3DToSupportingEdgesMapping exactly
as Finalize_Edge_Pairs_and_Reconstruct_3D_Edges does (EdgeSketch_Core.cpp:678-726: the two hypothesis edges, the other members of the
representative's cluster, the supporting edge of every validation view); then the reference's consolidate_3D_edges (its unmodified
edge_mapping.cpp built against oracle/ref_shim) groups the 3-D edges into curves and writes outputs/final_curves.txt.  The file is compared
with the one the reference wrote on its own (tests/golden/ref_final_curves_abc.txt.gz).  Needs /root/reference (curvelet files).
    python scripts/dropin_final_curves.py [--gpu]        (~5 minutes: the reference's 1000-iteration alignment dominates)
"""
import gzip
import importlib
import json
import os
import shutil
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle_py, ref_py  # noqa: E402

data = importlib.import_module("3d_curvix_b200.data")
use_gpu = "--gpu" in sys.argv
scene = data.Scene.load_npz(os.path.join(ROOT, "tests", "golden", "abc_nef_00000006_t1.npz"))
gold = np.load(os.path.join(ROOT, "tests", "golden", "ref_abc_full_run.npz"))
work = tempfile.mkdtemp(prefix="dropin_curves_")
# the dataset folder EdgeMapping re-reads (curvelets, edges); poses under the file names the reference build expects (definitions.h:11)
sd = os.path.join(work, "ABC-NEF", "00000006")
os.makedirs(os.path.join(sd, "RnT"))
for d in ("Edges", "curvelets"):
    os.symlink(os.path.join(ref_py.REF_ROOT, "datasets", "ABC-NEF", "00000006", d), os.path.join(sd, d))
for a, b in (("R_matrix.txt", "refined_R.txt"), ("T_matrix.txt", "refined_T.txt")):
    shutil.copy(os.path.join(ref_py.REF_ROOT, "datasets", "ABC-NEF", "00000006", "RnT", a), os.path.join(sd, "RnT", b))
ref = ref_py.Reference.for_scene(scene, dataset_path=work + "/")

if use_gpu:
    api = importlib.import_module("3d_curvix_b200.api")
    prod = api.EdgeSketch(api.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols), 0)
    prod.set_cameras(scene.R, scene.T, scene.K)
    prod.set_all_view_edges(scene.edges)
else:
    cfg = oracle_py.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols)
    prod = oracle_py.Oracle(cfg)
    prod.set_cameras(scene.R, scene.T, scene.K)
    for v, e in enumerate(scene.edges):
        prod.set_view_edges(v, e)

n_links = 0
t0 = time.time()
n_passes = int(os.environ.get("DROPIN_PASSES", "0")) or len(gold["schedule"])
for (h1, h2) in [tuple(int(x) for x in p) for p in gold["schedule"]][:n_passes]:
    prod.set_pair(h1, h2)
    prod.run() if use_gpu else prod.run(mode=1)
    p, f = prod.pairs(), prod.finalize()
    order = prod.reference_order() if use_gpu else np.arange(len(p["h1"]))
    vviews = [v for v in range(scene.n_views) if v not in (h1, h2)]
    E1, E2 = scene.edges[h1], scene.edges[h2]
    rec, img, idx = [], [], []
    for k in order:
        if f["nan_flag"][k]:
            continue
        a, b = int(p["h1"][k]), int(p["h2rep"][k])
        W, T = f["gamma_world"][k], f["tangent_world"][k]
        def add(pt, un, image, e):
            rec.append(np.concatenate([W, T, pt, un, [0.0]])); img.append(image); idx.append(e)
        add(E1[a, :2], E1[a, :3], h1, a)                                           # :678-682
        add(f["h2_corrected"][k], E2[b, :3], h2, b)                                # :683
        members = prod.pair_cluster(int(k)) if not use_gpu else None
        if members is None:                                                        # CUDA path: cluster of the representative from the row table
            members = prod.cluster_members_of_pair(int(k)) if hasattr(prod, "cluster_members_of_pair") else np.array([b])
        for m in members:                                                          # :686-692
            m = int(m)
            if m < 0 or m >= len(E2) or m == b:
                continue
            add(f["h2_corrected"][k], E2[m, :3], h2, m)
        for vi, v in enumerate(vviews):                                            # :698-724
            s = int(p["supports"][k, vi])
            if s != -2:
                add(scene.edges[v][s, :2], scene.edges[v][s, :3], v, s)
    ref.mapping_add_many(np.array(rec), img, idx)
    if os.environ.get("DROPIN_LINKS"):     # the same records in curvix_main's --support-links / --consolidate-links format
        with open(os.environ["DROPIN_LINKS"], "a") as lf:
            for r, im, ix in zip(rec, img, idx):
                lf.write(" ".join(repr(float(x)) for x in (*r[0:3], im, ix, *r[6:8], *r[3:6], *r[8:11])) + "\n")
    n_links += len(rec)
    print(f"pass ({h1},{h2}): {len(order)} accepted rows, {len(rec)} links", flush=True)
print(f"{ref.n_3d_keys()} distinct 3-D edges, {n_links} links, producer {'CUDA path' if use_gpu else 'port oracle'}, {time.time() - t0:.1f} s; "
      "running the reference's consolidate_3D_edges ...", flush=True)
t0 = time.time()
out_dir = ref.consolidate()
t_cons = time.time() - t0
mine = open(os.path.join(out_dir, "final_curves.txt")).read().splitlines()
theirs = gzip.open(os.path.join(ROOT, "tests", "golden", "ref_final_curves_abc.txt.gz"), "rt").read().splitlines()


def parse(lines):
    rows = np.array([[float(x) for x in l.split()] for l in lines[2:] if l.strip()])
    return rows, len({int(r[0]) for r in rows})


A, ca = parse(mine)
B, cb = parse(theirs)
key = lambda M: {tuple(np.round(r[2:5], 3)) for r in M}
ka, kb = key(A), key(B)
rep = {"producer": "cuda" if use_gpu else "port oracle", "curves": ca, "curves_reference": cb, "nodes": len(A), "nodes_reference": len(B),
       "node_positions_in_common_3dp": len(ka & kb), "node_positions": len(ka), "node_positions_reference": len(kb),
       "headers_equal": mine[:2] == theirs[:2], "consolidate_s": round(t_cons, 1)}
print(json.dumps(rep))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(rep, open(os.path.join(ROOT, "gpurun_out", "dropin_final_curves.json"), "w"), indent=1)
shutil.rmtree(work, ignore_errors=True)

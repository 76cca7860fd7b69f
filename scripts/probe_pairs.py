"""Per-pair pass times and counts of a workload's fixed hypothesis-pair schedule on one GPU (load-balance planning for --gpus N)."""
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
api = importlib.import_module("3d_curvix_b200.api")
data = importlib.import_module("3d_curvix_b200.data")

name = sys.argv[1] if len(sys.argv) > 1 else "cfg4"
w = data.workload(name)
if len(sys.argv) > 2:
    w["pairs"] = [tuple(int(t) for t in sys.argv[2].split(","))]
t0 = time.time()
scene = data.make_synthetic_scene(**w["scene"])
print("scene built in %.1f s" % (time.time() - t0), flush=True)
cfg = api.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols, max_support_views=w["max_support_views"])
g = api.EdgeSketch(cfg, 0)
g.set_cameras(scene.R, scene.T, scene.K)
g.set_all_view_edges(scene.edges)
out = []
for rep in range(2):
    for (a, b) in w["pairs"]:
        g.set_pair(a, b); g.run(); g.device_results(); g.synchronize()
        tm = g.timings(); c = g.counts()
        if rep == 1:
            out.append(dict(pair=[a, b], ms=tm["set_pair_ms"] + tm["run_ms"] + tm["finalize_ms"], phases={k: round(v, 3) for k, v in tm.items() if k.endswith("_ms")},
                            H=c["n_hypotheses"], U=c["n_unique_hypotheses"], P=c["n_pairs"], vals=c["n_validations"]))
            print(out[-1], flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"probe_pairs_{name}.json"), "w"), indent=1)
ms = [o["ms"] for o in out]
print("sum %.1f ms  max %.1f  min %.1f  mean %.1f" % (sum(ms), max(ms), min(ms), sum(ms) / len(ms)))

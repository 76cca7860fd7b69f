#!/usr/bin/env python
"""A/B of kernel variants on one GPU: the cfg2 single-pair pass (bench.secondary_cfg2) and the cfg5 stand-alone support test
(bench.support_problem_bench) for the library named by CVX_LIB_PATH (default: the in-tree build).  One JSON line per workload.

    CVX_LIB_OUT=/tmp/v.so CVX_NVCC_EXTRA="-DGRID_DUP=4" python 3d_curvix_b200/build.py --force   # build a variant here
    CVX_LIB_PATH=.../v.so python scripts/ab_kernels.py --tag dup4 [--cfg4-pairs 2]                 # time it on the GPU box
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tag", default="")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--cfg5-views", type=int, default=40)
    ap.add_argument("--pair", default="")
    ap.add_argument("--only", default="", help="cfg2 | cfg5 | cfg4")
    ap.add_argument("--cfg4-pairs", type=int, default=0, help="also time the first n pairs of the cfg4 schedule")
    a = ap.parse_args()
    args = argparse.Namespace(steps=a.steps, warmup=3, pair=a.pair, cfg5_views=a.cfg5_views, cpu_seconds=0.0, no_cpu_baseline=True)
    if a.only in ("", "cfg2"):
      s = bench.secondary_cfg2(args, 0)
      print(json.dumps({"tag": a.tag, "what": "cfg2", "ms": s["ms_per_step"], "e2e_ms": s.get("e2e", {}).get("ms_per_step"), "phases": s.get("phases_ms")}))
    if a.only in ("", "cfg5"):
      r = bench.support_problem_bench(args, 0, 1, 0, a.cfg5_views, a.steps, quiet=True)
      print(json.dumps({"tag": a.tag, "what": "cfg5", "ms": r.get("kernel_ms"), "frac": r.get("frac"), "visited": r.get("entries_visited_per_validation")}))
    if a.cfg4_pairs and a.only in ("", "cfg4"):
        import torch
        api = bench.pkg("api")
        scene, w = bench.build_scene("cfg4")
        cfg = api.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols, max_support_views=w["max_support_views"])
        g = api.EdgeSketch(cfg, 0)
        g.set_cameras(scene.R, scene.T, scene.K)
        g.set_all_view_edges(scene.edges)
        tot = {}
        for rep in range(2):
            tot = {}
            for p in w["pairs"][:a.cfg4_pairs]:
                g.set_pair(*p); g.run(); g.device_results()
                torch.cuda.synchronize()
                tm = g.timings()
                for k in ("form_ms", "prep_ms", "validate_ms", "select_ms", "run_ms"):
                    tot[k] = tot.get(k, 0.0) + tm[k]
        print(json.dumps({"tag": a.tag, "what": f"cfg4 first {a.cfg4_pairs} pairs", "phases": {k: round(v, 2) for k, v in tot.items()}}))


if __name__ == "__main__":
    main()

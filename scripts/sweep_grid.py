"""Experiment: validation-kernel time and grid statistics versus grid resolution (cfg2, pair (0,1))."""
import importlib, os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
api = importlib.import_module("3d_curvix_b200.api"); data = importlib.import_module("3d_curvix_b200.data")
w = data.workload(sys.argv[1] if len(sys.argv) > 1 else "cfg2")
scene = data.make_synthetic_scene(**w["scene"])
for f in [int(x) for x in os.environ.get("SWEEP", "16").split(",")]:
    os.environ["CVX_GRID_CELLS_PER_EDGE"] = str(f)
    cfg = api.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols, max_support_views=w["max_support_views"])
    g = api.EdgeSketch(cfg, 0)
    g.set_cameras(scene.R, scene.T, scene.K)
    for v, e in enumerate(scene.edges):
        g.set_view_edges(v, e)
    best = None
    for _ in range(5):
        g.set_pair(*w["pairs"][0]); g.run()
        t = g.timings()
        if best is None or t["validate_ms"] < best["validate_ms"]:
            best = t
    c = g.counts()
    ev = c["n_unique_hypotheses"] * (scene.n_views - 2)
    print(json.dumps({"cells_per_edge": f, "validate_ms": round(best["validate_ms"], 4), "set_pair_ms": round(best["set_pair_ms"], 4),
                      "visited_per_eval": c["n_entries_visited"] / ev, "heavy_frac": c["n_heavy_queries"] / ev, "evals": ev,
                      "inliers_weighted": c["n_inliers"], "n_parallel": c["n_parallel"], "n_validations": c["n_validations"]}))
    g.close()

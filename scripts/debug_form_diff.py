import sys, os, json, importlib, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import oracle_py as op
api = importlib.import_module("3d_curvix_b200.api"); data = importlib.import_module("3d_curvix_b200.data")
sc = data.Scene.load_npz(os.path.join(ROOT, "tests/golden/abc_nef_00000006_t1.npz"))
o = op.Oracle(op.default_config(n_views=50)); g = api.EdgeSketch(api.default_config(n_views=50), 0)
o.set_cameras(sc.R, sc.T, sc.K); g.set_cameras(sc.R, sc.T, sc.K)
for v, e in enumerate(sc.edges): o.set_view_edges(v, e); g.set_view_edges(v, e)
o.set_pair(47, 42); g.set_pair(47, 42); o.run(mode=1, keep=True); g.run()
to, tg = o.h1_table(), g.h1_table(); ro, rg = o.rows(), g.rows()
out = []
for h in range(len(to["m"])):
    n = to["n"][h]
    if n == 0: continue
    so, sg = to["off"][h], tg["off"][h]
    diff = [k for k in ("h2_orig", "local_idx", "raw_at_local", "label", "h2_rep") if not np.array_equal(ro[k][so:so+n], rg[k][sg:sg+n])]
    if to["m"][h] != tg["m"][h] or to["n"][h] != tg["n"][h]: diff.append("mn")
    if diff:
        d = dict(h1=h, n=int(n), margin=float(to["form_margin"][h]), diff=diff)
        for k in ("label", "h2_rep", "h2_orig"):
            d[k + "_o"] = ro[k][so:so+n].tolist(); d[k + "_g"] = rg[k][sg:sg+n].tolist()
        for k in ("xc", "yc", "theta_orig", "xbar", "ybar"):
            d[k + "_o"] = ro[k][so:so+n].tolist(); d[k + "_g"] = rg[k][sg:sg+n].tolist()
        out.append(d)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "form_diff.json"), "w"), indent=0)
print(len(out), "groups differ"); print(json.dumps(out[0])[:1500] if out else "")

"""Generates tests/golden/oracle_abc_47_42.npz: outputs of the CPU oracle (oracle/) on the bundled ABC-NEF scene, pass (47,42).
These goldens are ORACLE-generated (the reference ships none and cannot be built here: DESIGN.md section 2); they pin the
oracle and the CUDA path against regressions, not against the reference binary.
    python tests/golden/make_oracle_goldens.py
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle_py  # noqa: E402

data = importlib.import_module("3d_curvix_b200.data")
scene = data.Scene.load_npz(os.path.join(ROOT, "tests", "golden", "abc_nef_00000006_t1.npz"))
cfg = oracle_py.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols)
o = oracle_py.Oracle(cfg)
o.set_cameras(scene.R, scene.T, scene.K)
for v, e in enumerate(scene.edges):
    o.set_view_edges(v, e)
o.set_pair(47, 42)
c = o.run(mode=1)
p = o.pairs(); f = o.finalize(); t = o.h1_table(); r = o.rows()
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "oracle_abc_47_42.npz"),
                    counts=np.array([c[k] for k in sorted(c)], np.int64), count_names=np.array(sorted(c)),
                    m=t["m"], n=t["n"], label=r["label"].astype(np.int16), h2_orig=r["h2_orig"], h2_rep=r["h2_rep"], row_count=r["count"].astype(np.int8),
                    pair_h1=p["h1"], pair_h2rep=p["h2rep"], pair_count=p["count"].astype(np.int8), pair_supports=p["supports"].astype(np.int16),
                    gamma_world=f["gamma_world"], tangent_world=f["tangent_world"], nan_flag=f["nan_flag"].astype(np.int8))
print(c)

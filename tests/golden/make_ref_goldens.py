"""Generates tests/golden/ref_abc_47_42.npz and ref_abc_28_48.npz: outputs of the REFERENCE ITSELF -- /root/reference's unmodified
Edge_Reconst sources compiled against the stand-in headers of oracle/ref_shim (oracle/_ref/libcurvix_ref.so) -- on the bundled
ABC-NEF scene 00000006 (t1 edges, un-refined poses), one edge-sketch pass each, driven in the call order of cmd/main.cpp:131-144.
These vectors pin the port oracle and the CUDA path on machines where /root/reference is absent (the GPU box).
    python tests/golden/make_ref_goldens.py            (needs /root/reference; ~1 min per pass on 8 cores)
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_py  # noqa: E402

data = importlib.import_module("3d_curvix_b200.data")
scene = data.Scene.load_npz(os.path.join(ROOT, "tests", "golden", "abc_nef_00000006_t1.npz"))


def one_pass(h1, h2):
    r = ref_py.Reference.for_scene(scene, num_threads=1)   # one thread: all_supported_indices in ascending view-1 order
    r.set_pair(h1, h2)
    r.run()
    r.finalize()
    p, q = r.pairs(), r.points()
    blocks = r.support_blocks()
    rows = np.array([len(b) for b in blocks], np.int32)
    sup = np.concatenate(blocks) if blocks else np.zeros((0, scene.n_views - 2))
    cm = r.cluster_map()
    keys = sorted(cm)
    np.savez_compressed(
        os.path.join(ROOT, "tests", "golden", f"ref_abc_{h1}_{h2}.npz"),
        build_info=np.array(ref_py.build_info()), pair=np.array([h1, h2]),
        pair_h1=p["h1"], pair_h2rep=p["h2rep"], pair_supports=p["supports"].astype(np.int16),
        gamma_cam=q["gamma_cam"], gamma_world=q["gamma_world"], tangent_world=q["tangent_world"],
        # all_supported_indices (EdgeSketch_Core.cpp:402,555): m rows per processed view-1 edge; NaN = cells the reference never wrote
        block_rows=rows, support_blocks=np.where(np.isnan(sup), -32768, sup).astype(np.int16),
        cluster_key_h1=np.array([k[0] for k in keys], np.int32), cluster_key_h2=np.array([k[1] for k in keys], np.int32),
        cluster_off=np.cumsum([0] + [len(cm[k]) for k in keys]).astype(np.int64),
        cluster_members=np.concatenate([cm[k] for k in keys]).astype(np.int32) if keys else np.zeros(0, np.int32))
    print((h1, h2), "pairs", len(p["h1"]), "blocks", len(blocks), "cluster keys", len(keys), r.times())


if __name__ == "__main__":
    for pr in ((47, 42), (28, 48)):
        one_pass(*pr)

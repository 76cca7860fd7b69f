"""Generates tests/golden/abc_nef_00000006_t1.npz from the reference's bundled dataset (data only, no code):
/root/reference/datasets/ABC-NEF/00000006 (Edges/Edge_<i>_t1.txt, RnT/R_matrix.txt, RnT/T_matrix.txt) with the intrinsics of
configs/3D_Curvix_ABC_NEF.yaml:18-25.  Run in the build container (the GPU box has no /root/reference)."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
data = importlib.import_module("3d_curvix_b200.data")

if __name__ == "__main__":
    sc = data.read_reference_scene("/root/reference/datasets/ABC-NEF/00000006", 50, 1111.11136542426, 1111.11136542426,
                                   399.5, 399.5, 800, 800, thresh=1, refined=False)
    out = os.path.join(ROOT, "tests", "golden", "abc_nef_00000006_t1.npz")
    sc.save_npz(out)
    print(out, os.path.getsize(out) / 1e6, "MB", sum(len(e) for e in sc.edges), "edges")
    # curvelets/Curvelets_<i>_t1.txt (consumed by EdgeMapping::build3DEdgeWeightedGraph): the int stream of every file
    import numpy as np
    flat, counts = [], []
    for i in range(50):
        f = np.loadtxt(f"/root/reference/datasets/ABC-NEF/00000006/curvelets/Curvelets_{i}_t1.txt", dtype=np.int64).reshape(-1)
        flat.append(f.astype(np.int32)); counts.append(len(f))
    out = os.path.join(ROOT, "tests", "golden", "abc_nef_00000006_curvelets_t1.npz")
    np.savez_compressed(out, flat=np.concatenate(flat), counts=np.asarray(counts, np.int64))
    print(out, os.path.getsize(out) / 1e6, "MB", sum(counts) // 4, "lines")

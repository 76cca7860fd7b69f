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


def full_run(workdir="/tmp/refrun"):
    """The reference's own main (oracle/_ref/curvix_main_ref = cmd/main.cpp + all of Edge_Reconst, unmodified, built against
    oracle/ref_shim) on the bundled scene with configs/3D_Curvix_ABC_NEF.yaml (only Dataset_Path and the thread count changed; the
    un-refined poses are presented under the refined_{R,T}.txt names the build expects, definitions.h:11).  ~11 min on 8 cores.
    -> tests/golden/ref_abc_full_run.npz (schedule, claimed ratios, 3-D edges / tangents of every pass as printed, 6 digits) and
       tests/golden/ref_final_curves_abc.txt.gz (outputs/final_curves.txt)."""
    import gzip
    import re
    import shutil
    import subprocess
    ref = ref_py.REF_ROOT
    out = os.path.join(workdir, "outputs")
    log = os.path.join(workdir, "full_run.log")
    if not os.path.exists(os.path.join(out, "final_curves.txt")):
        ref_py.build(("strict",), main=True)
        scene_dir = os.path.join(workdir, "data", "ABC-NEF", "00000006")
        os.makedirs(os.path.join(scene_dir, "RnT"), exist_ok=True)
        os.makedirs(os.path.join(workdir, "run", "bin"), exist_ok=True)
        os.makedirs(out, exist_ok=True)
        for d in ("Edges", "curvelets"):
            if not os.path.exists(os.path.join(scene_dir, d)):
                os.symlink(os.path.join(ref, "datasets", "ABC-NEF", "00000006", d), os.path.join(scene_dir, d))
        shutil.copy(os.path.join(ref, "datasets/ABC-NEF/00000006/RnT/R_matrix.txt"), os.path.join(scene_dir, "RnT", "refined_R.txt"))
        shutil.copy(os.path.join(ref, "datasets/ABC-NEF/00000006/RnT/T_matrix.txt"), os.path.join(scene_dir, "RnT", "refined_T.txt"))
        cfg = open(os.path.join(ref, "configs", "3D_Curvix_ABC_NEF.yaml")).read()
        cfg = re.sub(r"^Dataset_Path:.*$", "Dataset_Path: " + os.path.join(workdir, "data") + "/", cfg, flags=re.M)
        cfg = re.sub(r"^Num_Of_OMP_Threads:.*$", "Num_Of_OMP_Threads: %d" % (os.cpu_count() or 1), cfg, flags=re.M)
        open(os.path.join(workdir, "cfg.yaml"), "w").write(cfg)
        with open(log, "w") as f:
            subprocess.run([os.path.join(ref_py.OUT, "curvix_main_ref"), "-c", "../../cfg.yaml"], cwd=os.path.join(workdir, "run", "bin"), stdout=f,
                           stderr=subprocess.STDOUT, check=True)
    text = open(log).read()
    sched = [(int(a), int(b)) for a, b in re.findall(r"Selected views for hypotheses are (\d+) and (\d+)", text)]
    ratios = [float(x) for x in re.findall(r"Average ratio of claimed edges: ([0-9.eE+-]+)", text)]
    npairs = [int(x) for x in re.findall(r"Number of edge pairs: (\d+)", text)]
    times = {k: float(v) for k, v in re.findall(r"Time for (Pairing up Edges|Finalizing Edge Pairs|Finding Next Hypothesis Views):\s+([0-9.eE+-]+)", text)}
    d = dict(schedule=np.array(sched), avg_ratio=np.array(ratios), n_pairs=np.array(npairs), build_info=np.array(ref_py.build_info()),
             pair_edges_s=times.get("Pairing up Edges", np.nan), finalize_s=times.get("Finalizing Edge Pairs", np.nan),
             next_views_s=times.get("Finding Next Hypothesis Views", np.nan), threads=os.cpu_count() or 1)
    for i, (a, b) in enumerate(sched):
        name = f"ABC-NEF_00000006_hypo1_{a}_hypo2_{b}.txt"
        d[f"edges_{i}"] = np.loadtxt(os.path.join(out, "3D_edges_" + name)).reshape(-1, 3)
        d[f"tangents_{i}"] = np.loadtxt(os.path.join(out, "3D_tangents_" + name)).reshape(-1, 3)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "ref_abc_full_run.npz"), **d)
    with open(os.path.join(out, "final_curves.txt"), "rb") as f, gzip.GzipFile(os.path.join(ROOT, "tests", "golden", "ref_final_curves_abc.txt.gz"), "wb", 9, mtime=0) as g:
        g.write(f.read())
    print("full run:", sched, ratios, npairs, times)


if __name__ == "__main__":
    if "--full" in sys.argv:
        full_run()
    else:
        for pr in ((47, 42), (28, 48)):
            one_pass(*pr)

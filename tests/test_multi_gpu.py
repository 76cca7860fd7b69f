"""The multi-GPU exchange behind the C-ABI (cvx_comm_* / cvx_gather_* / cvx_claims_allreduce / cvx_split_h1_by_rows) on real GPUs:
scripts/dist_check.py under torchrun, one process per GPU, compared bit for bit with single-GPU passes.  Needs >= 2 GPUs."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def test_gather_and_sharding_match_single_gpu():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs (run with gpurun --gpus 2)")
    world = min(n, 4)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", os.path.join(ROOT, "scripts", "dist_check.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "[dist_check] OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_curvix_main_devices_equals_single_device(abc_scene, tmp_path):
    """curvix_main --devices 0,1[,2,3]: one context and one host thread per device, every pass split by view-1 range
    (cvx_split_h1_by_rows), rows gathered through cvx_gather_* and claims OR-ed by cvx_claims_allreduce, then the same
    add3DToSupportingEdgesMapping loop on rank 0.  Every output file and the pass schedule equal the single-device run's byte for byte."""
    import torch
    from conftest import pkg
    from test_host_dropin import ABC_YAML
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs (run with gpurun --gpus 2)")
    exe = pkg("build").build_host(force=False)
    root = tmp_path
    abc_scene.write_reference_layout(str(root), dataset="ABC-NEF", scene="00000006", thresh=1)
    cfg = root / "abc.yaml"
    cfg.write_text(ABC_YAML.replace("/some/where/Datasets/", str(root) + "/"))
    runs = {}
    for tag, dev in (("one", ["--device", "0"]), ("many", ["--devices", ",".join(str(i) for i in range(min(n, 4)))])):
        out, sched = root / ("out_" + tag), root / (tag + ".txt")
        r = subprocess.run([exe, "-c", str(cfg), "--outputs", str(out), "--schedule", str(sched), "--max-passes", "3", "--quiet"] + dev,
                           capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
        runs[tag] = ({f: (out / f).read_bytes() for f in sorted(os.listdir(out))}, [l.split()[:8] for l in sched.read_text().splitlines()])
    assert runs["one"][1] == runs["many"][1], (runs["one"][1], runs["many"][1])
    assert list(runs["one"][0]) == list(runs["many"][0]) and len(runs["one"][0]) >= 6
    for f in runs["one"][0]:
        assert runs["one"][0][f] == runs["many"][0][f], f

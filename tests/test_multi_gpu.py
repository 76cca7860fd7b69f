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

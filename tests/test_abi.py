"""The C-ABI library loads on a CPU-only box and exports exactly the symbols include/curvix_b200.h declares
(no compute calls here: cvx_create must fail cleanly without a GPU -- there is no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT, pkg


def _declared():
    src = open(os.path.join(ROOT, "include", "curvix_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(cvx_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    pkg("build").build_cuda(force=False)
    api = pkg("api")
    lib = C.CDLL(api.LIB_PATH)
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in curvix_b200.h but not exported"
    assert sorted(api.SYMBOLS) == names, "api.SYMBOLS out of sync with the header"


def test_struct_layouts_match_header():
    api = pkg("api")
    assert C.sizeof(api.CvxConfig) == 11 * 8 + 6 * 4
    assert C.sizeof(api.CvxCounts) == 14 * 8
    assert C.sizeof(api.CvxTimings) == 7 * 4 + 2 * 4
    cfg = api.default_config()
    assert (cfg.delta, cfg.delta_theta_deg, cfg.lowes_ratio, cfg.pi_truncated) == (0.4, 15.0, 0.6, 3.1415926)
    assert (cfg.max_cluster_size, cfg.max_support_views, cfg.n_views) == (10, 4, 50)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    api = pkg("api")
    with pytest.raises(api.CurvixError):
        api.EdgeSketch(api.default_config(), 0)

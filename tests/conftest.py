import importlib
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pkg(mod=None):
    """The package directory name starts with a digit, so it is imported through importlib."""
    return importlib.import_module("3d_curvix_b200" + ("." + mod if mod else ""))


@pytest.fixture(scope="session")
def oracle_mod():
    from oracle import oracle_py
    oracle_py.build(fast=False)
    return oracle_py


@pytest.fixture(scope="session")
def data_mod():
    return pkg("data")


@pytest.fixture(scope="session")
def tiny_scene(data_mod):
    w = data_mod.workload("tiny")
    return data_mod.make_synthetic_scene(**w["scene"]), w


@pytest.fixture(scope="session")
def abc_scene(data_mod):
    """The bundled ABC-NEF object 00000006 (t1 edges, un-refined poses) as a committed binary fixture."""
    p = os.path.join(ROOT, "tests", "golden", "abc_nef_00000006_t1.npz")
    if not os.path.exists(p):
        pytest.skip("ABC-NEF fixture missing")
    sc = data_mod.Scene.load_npz(p, "ABC-NEF/00000006")
    cv = os.path.join(ROOT, "tests", "golden", "abc_nef_00000006_curvelets_t1.npz")   # curvelets/Curvelets_<i>_t1.txt of the same object
    if os.path.exists(cv):
        sc.load_curvelets(cv)
    return sc


def make_oracle(oracle_mod, scene, **cfg_kw):
    cfg = oracle_mod.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols, **cfg_kw)
    o = oracle_mod.Oracle(cfg)
    o.set_cameras(scene.R, scene.T, scene.K)
    for v, e in enumerate(scene.edges):
        o.set_view_edges(v, e)
    return o


@pytest.fixture(scope="session")
def host_harness():
    """Host build of the product's per-element math header, for CPU-side transcription checks."""
    import ctypes as C
    so = os.path.join(ROOT, "tests", "libhost_harness.so")
    src = os.path.join(ROOT, "tests", "host_harness.cpp")
    hdrs = [os.path.join(ROOT, "3d_curvix_b200", "csrc", h) for h in ("cvx_math.cuh", "cvx_graph_math.cuh")]
    if not os.path.exists(so) or os.path.getmtime(so) < max([os.path.getmtime(src)] + [os.path.getmtime(h) for h in hdrs]):
        cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        subprocess.run([cxx, "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-std=c++17", "-o", so, src], check=True)
    return C.CDLL(so)

"""CPU tests of the C++ host drop-in's I/O: settings reader, dataset readers, edge cache, Eigen-format writer.
(The class itself needs a GPU: tests/test_host_dropin.py, -m gpu.)"""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

ABC_YAML = """# %YAML:1.0

#> 3D Edge Sketch Settings (ABC-NEF)
Num_Of_OMP_Threads: 32                #> Number of CPU cores
Init_Hypo1_View_Index: 47 #47            #> Initial hypothesis view 1 index
Init_Hypo2_View_Index: 42 #42
delta: 0.4                            #> edge location perturbation
delta_theta: 15
Max_Num_Of_Support_Views: 4
TOED_Thresh: 1
Parallel_Epipolar_Line_Angle: 15
Reproj_Dist_Thresh: 2
Ratio_Of_Claimed_Edges_to_Stop: 0.7   #> stopping ratio
Max_Num_Of_3D_Edge_Sketch_Passes: 25

#> Dataset Settings
Dataset_Path: /some/where/Datasets/
Dataset_Name: ABC-NEF
Scene_Name: 00000006
Total_Num_Of_Images: 50
fx: 1111.11136542426
fy: 1111.11136542426
cx: 399.500000000000
cy: 399.500000000000
Image_Rows: 800
Image_Cols: 800
Use_Multiple_K: false
"""


@pytest.fixture(scope="module")
def io_check(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("hostio") / "host_io_check")
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.run([cxx, "-O1", "-std=c++17", "-Wall", "-o", exe, os.path.join(ROOT, "tests", "host_io_check.cpp")], check=True)
    return exe


def test_flat_yaml_reads_the_reference_settings_layout(io_check, tmp_path):
    p = tmp_path / "cfg.yaml"
    p.write_text(ABC_YAML)
    out = subprocess.run([io_check, "yaml", str(p)], check=True, capture_output=True, text=True).stdout.splitlines()
    a = out[0].split()
    assert [int(a[0]), int(a[1]), int(a[2])] == [32, 47, 42]
    assert float(a[3]) == 0.4 and float(a[4]) == 15 and int(a[5]) == 4 and int(a[6]) == 1 and float(a[7]) == 15
    assert float(a[8]) == 0.7 and float(a[9]) == 1111.11136542426 and int(a[10]) == 25
    assert out[1] == "/some/where/Datasets/|ABC-NEF|00000006|50|800|800|0"


def test_edge_reader_and_cache_round_trip(io_check, tmp_path):
    rng = np.random.default_rng(3)
    e = np.column_stack([rng.uniform(0, 800, 257), rng.uniform(0, 800, 257), rng.uniform(-np.pi, np.pi, 257), rng.uniform(0, 9, 257)])
    p = tmp_path / "Edge_0_t1.txt"
    np.savetxt(p, e, fmt="%.17g", delimiter="\t")
    with open(p, "a") as f:
        f.write("1.5 2.5\n")          # trailing incomplete record: dropped (file_reader.cpp:64-73)
    out = subprocess.run([io_check, "edges", str(p)], check=True, capture_output=True, text=True).stdout.splitlines()
    assert int(out[0]) == 257
    got = np.array([[float(v) for v in l.split()] for l in out[1:]])
    assert np.array_equal(got, e)
    # empty file -> zero edges
    q = tmp_path / "Edge_1_t1.txt"
    q.write_text("")
    out = subprocess.run([io_check, "edges", str(q)], check=True, capture_output=True, text=True).stdout.splitlines()
    assert int(out[0]) == 0


def test_rt_reader(io_check, tmp_path):
    rng = np.random.default_rng(4)
    R = rng.normal(size=(5, 3, 3)); T = rng.normal(size=(5, 3))
    np.savetxt(tmp_path / "R_matrix.txt", R.reshape(-1, 3), fmt="%.17g", delimiter="\t")
    np.savetxt(tmp_path / "T_matrix.txt", T.reshape(-1, 1), fmt="%.17g")
    out = subprocess.run([io_check, "rt", str(tmp_path)], check=True, capture_output=True, text=True).stdout.splitlines()
    assert out[0] == "5 5"
    v = np.array([float(x) for x in out[1:]])
    assert np.array_equal(v[:45], R.ravel()) and np.array_equal(v[45:], T.ravel())


def test_eigen_default_format_writer(io_check, tmp_path):
    """`ofstream << MatrixXd`: precision 6, single-space separated, every coefficient right-aligned to the widest one of the
    whole matrix, newline between rows, none at the end (EdgeSketch_Core.cpp:750-763)."""
    p = tmp_path / "m.txt"
    m = [[1, -2.5, 3.14159265], [100000, 1e-5, 1234567.0], [-0.000123456789, 0.5, -7]]
    txt = "3 3 " + " ".join(repr(float(v)) for r in m for v in r)
    subprocess.run([io_check, "eigen", str(p)], input=txt, text=True, check=True)
    exp = ("           1         -2.5      3.14159\n"
           "      100000        1e-05  1.23457e+06\n"
           "-0.000123457          0.5           -7")
    assert p.read_text() == exp
    # readable by the consumers (visualization/plot_3D_edge_sketch.m:39 reads %f %f %f)
    assert np.allclose(np.loadtxt(p), np.array(m), rtol=1e-5)
    # empty matrix -> empty file
    subprocess.run([io_check, "eigen", str(p)], input="0 3", text=True, check=True)
    assert p.read_text() == ""


def test_edge_mapping_insertion_api(io_check, tmp_path):
    p = tmp_path / "links.txt"
    out = subprocess.run([io_check, "mapping", str(p)], check=True, capture_output=True, text=True).stdout.split()
    assert out == ["2", "3"]          # keyed on the exact doubles of the 3-D point, like the reference's unordered_map
    assert len(p.read_text().splitlines()) == 3


def test_final_curves_file_layout_round_trips_the_reference_file(io_check, tmp_path):
    """tests/golden/ref_final_curves_abc.txt.gz is outputs/final_curves.txt as the REFERENCE wrote it (its own cmd/main.cpp built against
    oracle/ref_shim, ABC-NEF 00000006, 5 passes; tests/golden/make_ref_goldens.py --full).  The host reader (the loop of
    visualization/plot_curves.m:17-38) parses it, and the host writer (edge_mapping.cpp:1126-1164 layout) reproduces it byte for byte."""
    import gzip
    gold = os.path.join(ROOT, "tests", "golden", "ref_final_curves_abc.txt.gz")
    if not os.path.exists(gold):
        pytest.skip("reference final_curves golden missing")
    text = gzip.open(gold, "rb").read()
    p = tmp_path / "final_curves.txt"
    p.write_bytes(text)
    out = subprocess.run([io_check, "curves", str(p)], check=True, capture_output=True, text=True).stdout.split()
    lines = text.decode().splitlines()
    assert lines[0].startswith("# Curves") and lines[1].startswith("# Format: CurveID")
    body = [l for l in lines[2:] if l.strip()]
    n_curves = len({l.split()[0] for l in body})
    assert int(out[0]) == n_curves == 780 and int(out[1]) == len(body)
    assert (tmp_path / "final_curves.txt.out").read_bytes() == text

#!/usr/bin/env python
"""Run bench.py against every kernel-variant library under build_variants/ (CVX_LIB_PATH) and print the phase times.
   python scripts/sweep_variants.py [cfg5-views]      (GPU box; variants are built with CVX_LIB_OUT / CVX_NVCC_EXTRA, see build.py)"""
import glob, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
views = sys.argv[1] if len(sys.argv) > 1 else "20"
libs = [os.path.join(ROOT, "3d_curvix_b200", "libcurvix_b200.so")] + sorted(glob.glob(os.path.join(ROOT, "build_variants", "lib_*.so")))
for lib in libs:
    env = dict(os.environ, CVX_LIB_PATH=lib)
    out = {}
    for name, extra in (("cfg2", []), ("cfg5", ["--workload", "cfg5", "--cfg5-views", views])):
        r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "5", "--no-cpu-baseline"] + extra, env=env,
                           capture_output=True, text=True)
        try:
            d = json.loads(r.stdout.strip().splitlines()[-1])
            out[name] = d
        except Exception:
            print(os.path.basename(lib), name, "FAILED", r.stderr[-300:])
    if "cfg2" in out and "cfg5" in out:
        print("%-22s cfg2 step %.3f ms validate %.3f form %.3f | cfg5(%sv) %.3f ms frac %.3f" % (
            os.path.basename(lib), out["cfg2"]["ms_per_step"], out["cfg2"]["phases_ms"]["validate_ms"], out["cfg2"]["phases_ms"]["form_ms"],
            views, out["cfg5"]["ms_per_step"], out["cfg5"]["roofline"]["frac"]), flush=True)

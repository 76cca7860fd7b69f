#!/usr/bin/env python
"""bench.py -- throughput of the edge-sketch hot path (BASELINE.json metric: hypothesis x view validations / s).

    python bench.py --gpus N --steps K --warmup W            # the CUDA path (one process per GPU; torchrun for N > 1)
    python bench.py --impl reference --gpus N --steps K ...  # the CPU arm: the oracle in faithful mode on the host cores

A "step" is one full pass of the hot path over one hypothesis view pair of the named synthetic scene:
Set_Hypothesis_Views_Edgels (per-pass staging) + Run_3D_Edge_Sketch (formation, validation, vote/ratio selection) +
Finalize (epipolar correction of the representative, 2-view triangulation, 3-D tangent, world transform).
At N = 1 the workload is BASELINE.json configs[1] (50 views x 10k edges/view, single pair (0, 1)).  At N > 1 every rank
runs one pair of the same scene (rank r -> pair (2r, 2r+1): "sharded by hypothesis view pair", weak scaling, no
data-path collective) and the accepted pairs + 3-D edges + tangents are gathered to rank 0 with NCCL inside the timed
region (the path's only exchange step).

value  = validations of all ranks / device time of the step (CUDA events on the context's stream, max over ranks);
         inputs resident in HBM, results left on the device.
e2e    = the same metric through the C-ABI with HOST buffers: every step re-uploads all views' edges from pinned host
         memory (cvx_set_view_edges), runs the pass and reads the accepted pairs, supports, 3-D edges and tangents back.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "hypothesis_x_view_validations_per_sec"
UNIT = "validations/s"
L2_FLUSH_BYTES = 512 << 20     # > 126 MB of L2


def pkg(mod=None):
    return importlib.import_module("3d_curvix_b200" + ("." + mod if mod else ""))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--pair", default="", help="N=1 only: hypothesis view pair a,b instead of the workload's first pair")
    ap.add_argument("--cfg5-views", type=int, default=0, help="cfg5 only: number of validation views (default 200)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.005)

    def start(self):
        if self.nv:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------------ CPU arm
def build_scene(name):
    data = pkg("data")
    w = data.workload(name)
    return data.make_synthetic_scene(**w["scene"]), w


def make_oracle(scene, w):
    """The CPU oracle (oracle/, test infrastructure): used here ONLY as the timed CPU baseline / reference arm."""
    from oracle import oracle_py
    oracle_py.build(fast=False)
    cfg = oracle_py.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols,
                                   max_support_views=w["max_support_views"])
    o = oracle_py.Oracle(cfg)
    o.set_cameras(scene.R, scene.T, scene.K)
    for v, e in enumerate(scene.edges):
        o.set_view_edges(v, e)
    return o


def cpu_sample_rate(o, pair, h1_begin, h1_end, mode, threads):
    o.set_pair(*pair)
    t0 = time.perf_counter()
    c = o.run(h1_begin, h1_end, mode=mode, threads=threads)
    dt = time.perf_counter() - t0
    return c["n_validations"], dt


def cpu_baseline(scene, w, pair, target_s, threads):
    """Faithful-mode oracle (the reference's loop nest: Theta(E3) scans per validation, per-(H1,view) recomputation,
    OpenMP schedule(static,T)) on a bounded range of view-1 edges of the same workload."""
    o = make_oracle(scene, w)
    E1 = len(scene.edges[pair[0]])
    n0 = min(E1, 4 * threads)
    v0, t0 = cpu_sample_rate(o, pair, 0, n0, 0, threads)          # calibration
    per_edge = max(t0 / max(n0, 1), 1e-6)
    n1 = int(min(E1, max(n0, target_s / per_edge)))
    n1 = max(threads, (n1 // threads) * threads)
    v1, t1 = cpu_sample_rate(o, pair, 0, n1, 0, threads)
    vh, th = cpu_sample_rate(o, pair, 0, min(E1, 8 * n1), 1, threads)   # hoisted ("fair optimised CPU") for context
    o.close()
    return {"value": v1 / t1, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"view-1 edges [0,{n1}) of pair {pair} ({v1} validations, {t1:.1f} s), oracle faithful mode, "
                      f"-O2 -ffp-contract=off, OpenMP {threads} threads",
            "hoisted_value": vh / th, "host_cores": os.cpu_count()}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU algorithm for the path.  The reference cannot be compiled here (Eigen +
    yaml-cpp absent) so this is the oracle port in faithful mode, all host threads; each step is a bounded sample."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    if args.workload == "cfg5":
        return run_reference_support(args, threads)
    scene, w = build_scene(args.workload)
    pair = tuple(w["pairs"][0])
    o = make_oracle(scene, w)
    E1 = len(scene.edges[pair[0]])
    n0 = min(E1, 4 * threads)
    v0, t0 = cpu_sample_rate(o, pair, 0, n0, 0, threads)
    # whole run (steps + warmup) within ~2 minutes
    budget = 120.0 / max(args.steps + args.warmup, 1)
    n1 = int(min(E1, max(threads, budget / max(t0 / n0, 1e-6))))
    n1 = max(threads, (n1 // threads) * threads)
    for _ in range(args.warmup):
        cpu_sample_rate(o, pair, 0, n1, 0, threads)
    tot_v, tot_t = 0, 0.0
    for _ in range(args.steps):
        v, t = cpu_sample_rate(o, pair, 0, n1, 0, threads)
        tot_v += v; tot_t += t
    val = tot_v / tot_t
    sample = f"view-1 edges [0,{n1}) of pair {pair} per step ({tot_v // max(args.steps, 1)} validations), oracle faithful mode"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(args.steps, 1), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.workload, scene, 1), "threads": threads},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_reference_support(args, threads):
    """--impl reference --workload cfg5: orc_support_test (the reference's two Theta(E3) scans + set_intersection + getSupportIdx
    per validation) on one cfg5-shaped view and a bounded range of groups per step."""
    from oracle import oracle_py
    oracle_py.build(fast=False)
    synth = pkg("support_synth")
    shape = dict(CFG5, n_views=1)
    prob = synth.make_support_problem(seed=20251109, **shape)
    gs, budget = 64, 120.0 / max(args.steps + args.warmup, 1)

    def sample(gs):
        hs = int(prob["group_off"][gs])
        sub = dict(prob, group_off=prob["group_off"][:gs + 1], wedge31=prob["wedge31"][:, :gs], wedge32=prob["wedge32"][:, :hs], tgt=prob["tgt"][:, :hs])
        t0 = time.perf_counter()
        oracle_py.support_test(threads=threads, **sub)
        return hs, time.perf_counter() - t0

    hs, dt = sample(gs)
    gs = int(max(threads, min(shape["n_groups"], gs * budget / max(dt, 1e-3))))
    for _ in range(args.warmup):
        sample(gs)
    tot_v, tot_t = 0, 0.0
    for _ in range(args.steps):
        v, t = sample(gs)
        tot_v += v; tot_t += t
    val = tot_v / tot_t
    smp = f"one cfg5-shaped view ({shape['edges_per_view']} edges), groups [0,{gs}) per step ({tot_v // max(args.steps, 1)} validations), orc_support_test"
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": 1e3 * tot_t / max(args.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                      "data": "synthetic", "config": {"workload": "cfg5: support-test stress (bounded sample)", "threads": threads},
                      "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": smp},
                      "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}), flush=True)


def workload_name(name, scene, world):
    e = len(scene.edges[0])
    base = f"{name}: synthetic {scene.n_views} views x {e} edges/view {scene.cols}x{scene.rows}"
    if world == 1:
        return base + ", single hypothesis view pair (0,1)"
    return base + f", {world} hypothesis view pairs (2r,2r+1), one per GPU"


# ------------------------------------------------------------------------------------------------ CUDA arm
def algorithmic_bytes(kernel, c, V, edges_vali):
    """Algorithmic HBM bytes of one launch of `kernel` (DESIGN.md "Kernels and rooflines").
    c = counts of the pass, edges_vali = sum over validation views of E3."""
    HV, H = c["n_validations"], c["n_hypotheses"]
    if kernel == "validate":      # fused reprojection + wedge + support test, one evaluation per (cluster, view):
        U = c["n_unique_hypotheses"]   # support index out (4 B) + wedge31 in (16 B) per evaluation, hypothesis (56 B) once,
        return (4 + 16) * U * V + 56 * U + 16 * c["n_inliers"] + 72 * edges_vali   # (cos,sin) per inlier, indexes once
    if kernel == "form":          # view-2 angle list (8 B / edge) per warp is L2-resident: compulsory = once + rows out
        return 8 * c["n_h1_total"] * 0 + 100 * H + 40 * c["n_candidates"]
    raise KeyError(kernel)


def run_b200(args, rank, world, local_rank):
    import torch
    api = pkg("api")
    pkg("build").build_cuda(force=False) if rank == 0 and not os.path.exists(api.LIB_PATH) else None
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    use_dist = world > 1
    if use_dist:
        import torch.distributed as tdist
        tdist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        distm = pkg("dist")
    scene, w = build_scene(args.workload)
    V = scene.n_views - 2
    pair = (2 * rank, 2 * rank + 1) if use_dist else tuple(w["pairs"][0])
    if args.pair and not use_dist:
        pair = tuple(int(t) for t in args.pair.split(","))
    if pair[1] >= scene.n_views:
        raise SystemExit("scene has too few views for this many ranks")
    cfg = api.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols, max_support_views=w["max_support_views"])
    g = api.EdgeSketch(cfg, local_rank)
    g.set_cameras(scene.R, scene.T, scene.K)
    # pinned host copy of every view's columns, packed in view order (the e2e leg uploads from it every step)
    n_per_view = np.array([len(e) for e in scene.edges], np.int32)
    total = int(n_per_view.sum())
    pinned = torch.empty((3, total), dtype=torch.float64).pin_memory()
    pinned.copy_(torch.from_numpy(np.ascontiguousarray(np.concatenate([e[:, :3] for e in scene.edges], axis=0).T)))
    h2d_bytes = pinned.numel() * 8

    def upload():
        g.set_all_view_edges_ptr(n_per_view, pinned[0].data_ptr(), pinned[1].data_ptr(), pinned[2].data_ptr())

    out_spec = {"h1": (1, 4), "h2rep": (1, 4), "row": (1, 4), "count": (1, 4), "supports": (V, 4), "gamma_cam": (3, 8), "gamma_world": (3, 8),
                "tangent_world": (3, 8), "h2_corrected": (2, 8), "nan_flag": (1, 4)}
    out_buf = {"cap": 0}

    def out_ptrs(P):
        """Pinned host result buffers (caller-owned, as the C-ABI expects), grown geometrically."""
        if P > out_buf["cap"]:
            out_buf["cap"] = max(2 * P, 1024)
            for k, (w, b) in out_spec.items():
                out_buf[k] = torch.empty(out_buf["cap"] * w * b, dtype=torch.uint8).pin_memory()
        return {k: out_buf[k].data_ptr() for k in out_spec}, sum(P * w * b for w, b in out_spec.values())

    upload()
    stream = torch.cuda.ExternalStream(g.stream_ptr(), device=torch.device("cuda", local_rank))
    flush = torch.empty(L2_FLUSH_BYTES // 4, dtype=torch.int32, device="cuda")

    gather_plan = {}

    def gather(res):
        """The path's only exchange: accepted pairs + 3-D edges + tangents to rank 0 (NCCL), rank order = pair order.  The first
        call gathers with exact counts (two collectives) and agrees a row capacity; the following ones use ONE collective."""
        P, Vv = int(res.n_pairs), int(res.n_vali_views)
        loc = {"h1": distm.device_tensor(res.pair_h1, (P,), torch.int32), "h2": distm.device_tensor(res.pair_h2rep, (P,), torch.int32),
               "sup": distm.device_tensor(res.pair_supports, (P, Vv), torch.int32),
               "gw": distm.device_tensor(res.gamma_world, (P, 3), torch.float64), "tw": distm.device_tensor(res.tangent_world, (P, 3), torch.float64)}
        if "pg" not in gather_plan:
            out = distm.gather_results(loc)
            cap = torch.tensor([P], dtype=torch.int64, device="cuda")
            tdist.all_reduce(cap, op=tdist.ReduceOp.MAX)
            fields = {k: (tuple(t.shape[1:]), t.dtype) for k, t in loc.items()}
            gather_plan["pg"] = distm.PackedGather(fields, int(cap.item()) * 5 // 4 + 64, torch.device("cuda", local_rank))
            return out
        return gather_plan["pg"].run(loc)

    gather_ev = []

    def step_device():
        g.set_pair(*pair)
        g.run()
        res = g.device_results()        # runs the finalize kernel; results stay on the device
        if use_dist:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            gather(res)
            e1.record(stream)
            gather_ev.append((e0, e1))
        return res

    def step_e2e():
        upload()
        g.set_pair(*pair)
        g.run()
        P = int(g.device_results().n_pairs)      # runs the finalize kernel
        ptrs, nbytes = out_ptrs(P)
        g.results_into(ptrs)                     # accepted pairs, supports, 3-D edges, tangents -> pinned host buffers
        return nbytes

    def barrier():
        if use_dist:
            tdist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        for _ in range(max(args.warmup, 3)):
            flush.zero_()
            step_device()
        barrier()
        counts = g.counts()
        launches0 = g.timings()["kernel_launches"]
        sampler = ClockSampler(local_rank)
        sampler.start()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        parts = {k: 0.0 for k in ("set_pair_ms", "form_ms", "prep_ms", "validate_ms", "select_ms", "finalize_ms", "run_ms")}
        barrier()
        gather_ev.clear()
        wall0 = time.perf_counter()
        for i in range(args.steps):
            flush.zero_()                     # L2 flush between timed iterations (outside the events)
            ev[i][0].record(stream)
            step_device()
            ev[i][1].record(stream)
            tm = g.timings()
            for k in parts:
                parts[k] += tm[k]
        barrier()
        wall = time.perf_counter() - wall0
        clocks = sampler.stop()
        dev_ms = sum(a.elapsed_time(b) for a, b in ev)
        launches = g.timings()["kernel_launches"] - launches0
        # ---- e2e leg: host buffers in, host buffers out, copies inside the timed region (wall clock, max over ranks)
        for _ in range(2):
            step_e2e()
        barrier()
        t0 = time.perf_counter()
        d2h = 0
        for _ in range(args.steps):
            flush.zero_()
            d2h = step_e2e()
        barrier()
        e2e_s = time.perf_counter() - t0

    vals = counts["n_validations"]
    per_rank = None
    if use_dist:
        mine = {"rank": rank, "pair": list(pair), "ms_per_step": dev_ms / args.steps, "validations": vals,
                "gather_ms": sum(a.elapsed_time(b) for a, b in gather_ev) / max(len(gather_ev), 1),
                "phases_ms": {k: round(v / args.steps, 4) for k, v in parts.items()}}
        per_rank = [None] * world
        tdist.all_gather_object(per_rank, mine)
        t = torch.tensor([dev_ms, e2e_s, wall], dtype=torch.float64, device="cuda")
        tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
        dev_ms, e2e_s, wall = (float(x) for x in t.tolist())
        s = torch.tensor([vals, launches, h2d_bytes, d2h], dtype=torch.int64, device="cuda")
        tdist.all_reduce(s, op=tdist.ReduceOp.SUM)
        vals_all, launches_all, h2d_all, d2h_all = (int(x) for x in s.tolist())
    else:
        vals_all, launches_all, h2d_all, d2h_all = vals, launches, h2d_bytes, d2h
    if rank != 0:
        if use_dist:
            tdist.destroy_process_group()
        return
    ms_per_step = dev_ms / args.steps
    value = vals_all / (ms_per_step * 1e-3)
    # ---- roofline of the dominant kernel (rank 0's pass), event-timed inside the library on the launching stream
    ktime = {k: parts[k + "_ms"] / args.steps for k in ("form", "prep", "validate", "select")}
    ktime = {k: v for k, v in ktime.items() if v > 0}
    dom = max(ktime, key=ktime.get)
    kname = {"validate": "k_validate<false>", "form": "k_form<true>", "select": "k_select", "prep": "k_wedge31"}.get(dom, dom)
    edges_vali = sum(len(scene.edges[v]) for v in range(scene.n_views) if v not in pair)
    peak, peak_src = peaks()
    try:
        abytes = algorithmic_bytes(dom, counts, V, edges_vali)
    except KeyError:
        abytes = 0
    ach = abytes / (ktime[dom] * 1e-3) / 1e9
    roof = {"bound": "hbm", "kernel": kname, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic(kname),
            "peak_source": peak_src, "kernel_ms": ktime[dom], "algorithmic_bytes": abytes, "ncu": ncu_notes(kname),
            "kernel_ms_all": {k: round(v, 4) for k, v in ktime.items()},
            "note": "FP64-transcendental bound, not HBM bound (DESIGN.md): ~1.3k FP64 flop-equivalents per validation vs a few bytes"}
    val_ms = parts["validate_ms"] / args.steps
    if val_ms > 0 and dom != "validate":
        sb = algorithmic_bytes("validate", counts, V, edges_vali)
        roof["validate_kernel"] = {"kernel": "k_validate<false>", "kernel_ms": val_ms, "algorithmic_bytes": sb,
                                   "achieved": sb / (val_ms * 1e-3) / 1e9, "frac": sb / (val_ms * 1e-3) / 1e9 / peak}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": workload_name(args.workload, scene, world), "l2": f"flushed between timed iterations ({L2_FLUSH_BYTES >> 20} MiB write)",
                       "validations_per_step": vals_all, "hypotheses_rank0": counts["n_hypotheses"], "unique_hypotheses_rank0": counts["n_unique_hypotheses"], "pairs_rank0": counts["n_pairs"],
                       "parallelism": "1 GPU" if world == 1 else f"{world} GPUs, one hypothesis view pair per GPU, NCCL gather of results"},
            "clocks": clocks,
            "e2e": {"value": vals_all * args.steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all,
                    "ms_per_step": 1e3 * e2e_s / args.steps, "timer": "host wall clock around the C-ABI calls, max over ranks"},
            "gpu_launches": launches_all, "wall_s_timed_region": wall, "roofline": roof,
            "phases_ms": {k: round(v / args.steps, 4) for k, v in parts.items()}, "counts_rank0": counts}
    if per_rank:
        line["per_rank"] = per_rank
    if not args.no_cpu_baseline and world == 1:
        g.close()
        line["cpu_baseline"] = cpu_baseline(scene, w, pair, args.cpu_seconds, os.cpu_count() or 1)
    print(json.dumps(line), flush=True)
    if use_dist:
        tdist.destroy_process_group()


# ------------------------------------------------------------------------------------------------ cfg5: support-test stress
CFG5 = dict(n_views=200, edges_per_view=200000, n_groups=50000, rows_per_group=20, delta_theta_deg=2.0)


def run_support_stress(args, rank, world, local_rank):
    """BASELINE.json configs[4]: the isolated support test (getHYPO2_idx_Ore x2 + set_intersection + getSupportIdx,
    EdgeSketch_Core.cpp:348-384) on pre-formed hypotheses, 200 validation views x 200k edges/view, ~1M hypothesis rows,
    delta_theta = 2 deg.  HBM bound: SURVEY.md 8d bytes = 52 + 16 n per validation + 32 B per staged edge.  Every rank runs its
    own problem (seed + rank): weak scaling, no collective on the data path."""
    import torch
    api = pkg("api"); synth = pkg("support_synth")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    use_dist = world > 1
    if use_dist:
        import torch.distributed as tdist
        tdist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    shape = dict(CFG5)
    if args.cfg5_views:
        shape["n_views"] = args.cfg5_views
    prob = synth.make_support_problem(seed=20251109 + rank, **shape)
    V, H = shape["n_views"], int(prob["group_off"][-1])
    g = api.EdgeSketch(api.default_config(n_views=8), local_rank)
    order = ("group_off", "view_off", "ang31", "ang32", "cs", "sn", "wedge31", "wedge32", "tgt")
    t0 = time.perf_counter()
    sp = api.SupportProblem(g, *[prob[k] for k in order], prob["thr"], prob["parallel_deg"])
    create_s = time.perf_counter() - t0
    h2d = sum(np.asarray(prob[k]).nbytes for k in order)
    stats = sp.stats()

    def barrier():
        if use_dist:
            tdist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        sp.run()
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    wall0 = time.perf_counter()
    ms = [sp.run() for _ in range(args.steps)]      # inputs (7 GB) >> L2 (126 MB): no flush needed between launches
    barrier()
    wall = time.perf_counter() - wall0
    clocks = sampler.stop()
    # e2e: host buffers in (problem upload + staging), kernel, support matrix back to the host
    t0 = time.perf_counter()
    sp2 = api.SupportProblem(g, *[prob[k] for k in order], prob["thr"], prob["parallel_deg"])
    sp2.run()
    res = sp2.result()
    e2e_s = time.perf_counter() - t0
    sp2.close()
    dev_ms = float(sum(ms))
    vals = V * H
    if use_dist:
        t = torch.tensor([dev_ms, e2e_s], dtype=torch.float64, device="cuda")
        tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
        dev_ms, e2e_s = (float(x) for x in t.tolist())
        sv = torch.tensor([vals, h2d, res.nbytes], dtype=torch.int64, device="cuda")
        tdist.all_reduce(sv, op=tdist.ReduceOp.SUM)
        vals_all, h2d_all, d2h_all = (int(x) for x in sv.tolist())
    else:
        vals_all, h2d_all, d2h_all = vals, h2d, res.nbytes
    if rank != 0:
        if use_dist:
            tdist.destroy_process_group()
        return
    ms_per_step = dev_ms / args.steps
    peak, peak_src = peaks()
    n_bar = stats["n_inliers"] / vals
    abytes = vals * 52 + 16 * stats["n_inliers"] + 32 * int(prob["view_off"][-1])
    ach = abytes / (ms_per_step * 1e-3) / 1e9
    line = {"metric": METRIC, "value": vals_all / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"cfg5: support-test stress, {V} validation views x {shape['edges_per_view']} edges/view, {H} pre-formed "
                                   f"hypothesis rows in {shape['n_groups']} groups, delta_theta {shape['delta_theta_deg']} deg",
                       "l2": "inputs per launch (7 GB at full size) exceed L2; no flush", "validations_per_step": vals_all,
                       "parallelism": "1 GPU" if world == 1 else f"{world} GPUs, one problem per GPU, no collective"},
            "clocks": clocks,
            "e2e": {"value": vals_all / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all,
                    "ms_per_step": 1e3 * e2e_s, "timer": "host wall clock around create (pageable host arrays -> device, staging) + run + result, one step"},
            "gpu_launches": args.steps, "wall_s_timed_region": wall,
            "roofline": {"bound": "hbm", "kernel": "k_validate<true>", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                         "traffic": traffic("k_validate<true>", V), "peak_source": peak_src, "kernel_ms": ms_per_step, "algorithmic_bytes": abytes,
                         "bytes_formula": "validations x 52 + 16 x wedge-intersection members + 32 x staged edges (SURVEY.md 8d)",
                         "mean_intersection": n_bar, "entries_visited_per_validation": stats["n_visited"] / vals, "ncu": ncu_notes("k_validate<true>")},
            "counts_rank0": dict(stats, n_validations=vals, n_rows=H, n_edges=int(prob["view_off"][-1])), "create_s": create_s}
    if not args.no_cpu_baseline and world == 1:
        # bounded CPU sample: one view, the first groups; same arrays
        from oracle import oracle_py
        oracle_py.build(fast=False)
        threads = os.cpu_count() or 1
        E = shape["edges_per_view"]
        gs = 256
        while True:
            hs = int(prob["group_off"][gs])
            sub = dict(group_off=prob["group_off"][:gs + 1], view_off=prob["view_off"][:2], ang31=prob["ang31"][:E], ang32=prob["ang32"][:E],
                       cs=prob["cs"][:E], sn=prob["sn"][:E], wedge31=prob["wedge31"][:1, :gs], wedge32=prob["wedge32"][:1, :hs],
                       tgt=prob["tgt"][:1, :hs], thr=prob["thr"], parallel_deg=prob["parallel_deg"])
            t0 = time.perf_counter()
            exp = oracle_py.support_test(threads=threads, **sub)
            dt = time.perf_counter() - t0
            if dt > args.cpu_seconds / 3 or gs >= shape["n_groups"]:
                break
            gs = min(shape["n_groups"], int(gs * min(8.0, max(2.0, args.cpu_seconds / max(dt, 1e-3)))))
        line["cpu_baseline"] = {"value": hs / dt, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"view 0, groups [0,{gs}) ({hs} validations, {dt:.1f} s), orc_support_test (Theta(E3) scans per validation as the reference), OpenMP {threads} threads",
                                "gpu_matches_on_sample": bool(np.array_equal(exp[0], res[0, :hs]))}
    print(json.dumps(line), flush=True)
    sp.close(); g.close()
    if use_dist:
        tdist.destroy_process_group()


def ncu_notes(kernel):
    """Pipe / issue utilisation of `kernel` from the committed ncu capture (profiles/traffic.json): what actually bounds it."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    return json.load(open(p)).get(kernel, {}).get("ncu") if os.path.exists(p) else None


def traffic(kernel, views=0):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel`, from the committed ncu --set full capture
    (profiles/traffic.json, written by scripts/ncu_summary.py from the .ncu-rep of the same bench command), else None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        e = json.load(open(p)).get(kernel, {})
        if "bytes_per_view" in e and views:
            return e["bytes_per_view"] * views
        return e.get("bytes")
    return None


def main():
    # NCCL prints its version banner on fd 1: keep the real stdout for the ONE JSON line, send everything else to stderr
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w")
    args = parse()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank, world)
    elif args.workload == "cfg5":
        run_support_stress(args, rank, world, local_rank)
    else:
        run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()

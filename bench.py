#!/usr/bin/env python
"""bench.py -- throughput of the edge-sketch hot path (BASELINE.json metric: hypothesis x view validations / s).

    python bench.py --gpus N --steps K --warmup W            # the CUDA path (one process per GPU; torchrun for N > 1)
    python bench.py --impl reference --gpus N --steps K ...  # the CPU arm: the reference's own Run_3D_Edge_Sketch on the host cores

Workload (every N, so that the driver's 1 -> 8 curve is strong scaling on north_star's configuration): cfg4 = synthetic 100 views x
50k edges/view, the FIXED 16-pair schedule (2k, 2k+1).  A "step" is one pass of the hot path over the whole schedule: per pair
Set_Hypothesis_Views_Edgels (staging) + Run_3D_Edge_Sketch (formation, validation, vote / ratio selection) + Finalize (correction of
the representative, 2-view triangulation, 3-D tangent, world transform).  At N > 1 the schedule is cut into units (a pair, or a
row-balanced view-1 range of a pair), dealt longest-first from times measured in the warm-up, and every unit's accepted rows + 3-D
edges + tangents are gathered to rank 0 through the C-ABI (cvx_gather_*: NCCL send/recv on a side stream, overlapped with the next
unit) inside the timed region.  torch.distributed (gloo) only carries the rendezvous and the bookkeeping.

value  = validations of the whole schedule / device time of the step (CUDA events on the context stream, max over ranks);
         inputs resident in HBM, results left on the device (rank 0's at N > 1).
e2e    = the same metric through the C-ABI with HOST buffers: every step re-uploads all views' edges from pinned host memory, runs
         the schedule and reads every accepted pair, its supports, 3-D edge and tangent back into pinned host memory (wall clock).
Extra objects of the N = 1 line: `secondary` (cfg2 single pair: BASELINE.json configs[1], round-1 continuity), `roofline_support_test`
(cfg5 stand-alone support test: the >= 60 % HBM target), `sec_per_object` (cfg1: curvix_main on ABC-NEF 00000006), `cpu_baseline`.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import tempfile
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
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg4", help="cfg4 (default, 16-pair schedule), cfg2, cfg3, cfg5 (stand-alone support test)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip secondary / roofline_support_test / sec_per_object")
    ap.add_argument("--parts", type=int, default=0, help="view-1 ranges per pair (0 = whole pairs while there are two per rank)")
    ap.add_argument("--shard", default="pairs", choices=["pairs", "h1"], help="h1: every pair sharded by view-1 range over all ranks")
    ap.add_argument("--pair", default="", help="cfg2 secondary / single-pair runs: hypothesis view pair a,b")
    ap.add_argument("--cfg5-views", type=int, default=0, help="cfg5: number of validation views (default 200; 40 inside the default line)")
    ap.add_argument("--ref-variant", default="strict", choices=["strict", "fast"], help="--impl reference: build flags of oracle/_ref")
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


# ------------------------------------------------------------------------------------------------ workloads
def build_scene(name):
    data = pkg("data")
    w = data.workload(name)
    return data.make_synthetic_scene(**w["scene"]), w


def workload_name(name, scene, w):
    e = len(scene.edges[0])
    base = f"{name}: synthetic {scene.n_views} views x {e} edges/view {scene.cols}x{scene.rows}"
    n = len(w["pairs"])
    if n == 1:
        return base + f", single hypothesis view pair {tuple(w['pairs'][0])}"
    return base + f", fixed schedule of {n} hypothesis view pairs (2k,2k+1)"


# ------------------------------------------------------------------------------------------------ CPU arm
class SubScene:
    """Views [h1, h2] + the first n_vali validation views of a scene, renumbered 0 .. n_vali + 1.  One validation costs Theta(E3)
    whatever the number of views, so validations / s of a sub-scene is the rate of the full one; only the sample shrinks."""

    def __init__(self, scene, pair, n_vali):
        others = [v for v in range(scene.n_views) if v not in pair][:n_vali]
        self.views = list(pair) + others
        self.R = scene.R[self.views]; self.T = scene.T[self.views]; self.K = scene.K
        self.edges = [scene.edges[v] for v in self.views]
        self.rows, self.cols = scene.rows, scene.cols
        self.n_views = len(self.views)


def cpu_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))      # the cores this process may use (a container can own fewer than os.cpu_count())
    except (AttributeError, OSError):
        return os.cpu_count() or 1


class CpuArm:
    """The timed CPU implementation of the path: oracle/_ref (the reference's own sources, kind "reference") when its library is there,
    else the port oracle in faithful mode (kind "port").  Test infrastructure used here ONLY as the measured baseline."""

    def __init__(self, variant="strict"):
        self.variant = variant
        self.kind = "port"
        try:
            from oracle import ref_py
            if ref_py.available(variant):
                ref_py.load(variant)
                self.ref_py, self.kind = ref_py, "reference"
        except Exception as e:      # missing library on a box without /root/reference
            sys.stderr.write(f"[bench] oracle/_ref unavailable ({e}); falling back to the port oracle\n")
        if self.kind == "port":
            from oracle import oracle_py
            oracle_py.build(fast=(variant == "fast"))
            self.oracle_py = oracle_py

    def describe(self):
        if self.kind == "reference":
            return self.ref_py.build_info(self.variant)
        return "port oracle (oracle/curvix_oracle.cpp), faithful mode, " + ("-O3 -march=native -ffast-math" if self.variant == "fast" else "-O2 -ffp-contract=off")

    def sample(self, scene, w, pair, n_edges, n_vali, threads):
        """Run_3D_Edge_Sketch on view-1 edges [0, n_edges) of `pair` against n_vali validation views.  -> (validations, seconds)."""
        sub = SubScene(scene, pair, n_vali)
        if self.kind == "reference":
            r = self.ref_py.Reference.for_scene(sub, variant=self.variant, num_threads=threads, max_support_views=w["max_support_views"])
            r.limit_h1(n_edges)
            r.set_pair(0, 1)
            t0 = r.times()["pair_edges"]
            r.run()
            dt = r.times()["pair_edges"] - t0         # the reference's own timer (EdgeSketch_Core.cpp:170,572)
            vals = sum(int((~np.isnan(b[:, 0])).sum()) for b in r.support_blocks()) * n_vali   # rows >= n are never written
            r.close()
            return vals, dt
        o = self.oracle_py
        cfg = o.default_config(n_views=sub.n_views, img_rows=sub.rows, img_cols=sub.cols, max_support_views=w["max_support_views"])
        orc = o.Oracle(cfg, fast=(self.variant == "fast"))
        orc.set_cameras(sub.R, sub.T, sub.K)
        for v, e in enumerate(sub.edges):
            orc.set_view_edges(v, e)
        orc.set_pair(0, 1)
        t0 = time.perf_counter()
        c = orc.run(0, n_edges, mode=0, threads=threads)
        dt = time.perf_counter() - t0
        orc.close()
        return c["n_validations"], dt

    def plan(self, scene, w, pair, threads, step_seconds):
        """Sample size that keeps EVERY OpenMP thread busy: the reference's loop is schedule(static, T) with chunk = T
        (EdgeSketch_Core.cpp:198), so T x T view-1 edges give each thread exactly one chunk; the time is fitted with the number of
        validation views.  -> (n_edges, n_vali, calibration)."""
        V = scene.n_views - 2
        n_edges = min(threads * threads, len(scene.edges[pair[0]]))
        # pilot at the FULL thread count on the smallest sample that keeps every thread busy (one validation view): a single-thread
        # calibration overestimates the rate of 32 contending threads by 2-3x (measured on the 8-GPU box), and a step cannot be interrupted
        vals, dt = self.sample(scene, w, pair, n_edges, 1, threads)
        n_vali = int(max(1, min(V, step_seconds / max(2.0 * dt, 1e-6))))   # factor 2: shared hosts run the long sample slower than the pilot
        return n_edges, n_vali, {"edges": n_edges, "views": 1, "threads": threads, "seconds": dt}


def cpu_baseline(scene, w, pair, target_s, variant="strict"):
    arm = CpuArm(variant)
    T = cpu_threads()
    n_edges, n_vali, cal = arm.plan(scene, w, pair, T, target_s)
    vals, dt = arm.sample(scene, w, pair, n_edges, n_vali, T)
    return {"value": vals / dt, "unit": UNIT, "cores": T, "kind": arm.kind, "threads_busy": min(T, -(-n_edges // T)),
            "sample": f"view-1 edges [0,{n_edges}) of pair {tuple(pair)} against {n_vali} of {scene.n_views - 2} validation views "
                      f"({vals} validations, {dt:.1f} s): Run_3D_Edge_Sketch, schedule(static,{T}) -> one {T}-edge chunk per thread",
            "build": arm.describe(), "host_cores": os.cpu_count()}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on this box's host cores, all threads busy, on the
    b200 arm's config / metric / unit; each step a bounded sample (see CpuArm.plan).  Rank 0 alone runs it."""
    if rank != 0:
        return
    if args.workload == "cfg5":
        return run_reference_support(args, cpu_threads())
    scene, w = build_scene(args.workload)
    pair = tuple(w["pairs"][0])
    arm = CpuArm(args.ref_variant)
    T = cpu_threads()
    budget = 150.0 / max(args.steps + args.warmup, 1)           # the whole run within a few minutes
    n_edges, n_vali, cal = arm.plan(scene, w, pair, T, budget)
    t_start, guard_s = time.perf_counter(), 240.0                # a step cannot be interrupted: stop starting new ones after 4 minutes
    for _ in range(args.warmup):
        if time.perf_counter() - t_start < guard_s / 3:
            arm.sample(scene, w, pair, n_edges, n_vali, T)
    tot_v, tot_t, steps_run = 0, 0.0, 0
    for _ in range(args.steps):
        if steps_run and time.perf_counter() - t_start > guard_s:
            break
        v, t = arm.sample(scene, w, pair, n_edges, n_vali, T)
        tot_v += v; tot_t += t; steps_run += 1
    val = tot_v / tot_t
    sample = (f"per step: view-1 edges [0,{n_edges}) of pair {pair} against {n_vali} of {scene.n_views - 2} validation views "
              f"({tot_v // max(steps_run, 1)} validations); schedule(static,{T}) -> one {T}-edge chunk per thread; {steps_run} of {args.steps} steps run "
              f"(pilot {cal['seconds']:.1f} s on {cal['threads']} threads)")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(steps_run, 1), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.workload, scene, w)},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": T, "kind": arm.kind, "threads_busy": min(T, -(-n_edges // T)), "sample": sample,
                             "build": arm.describe(), "host_cores": os.cpu_count()},
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


# ------------------------------------------------------------------------------------------------ CUDA arm: helpers
def algorithmic_bytes_fused(c, edges_vali):
    """SURVEY.md 8(d), fused kernel: 4 B support index out per validation (the reference materialises the full matrix,
    EdgeSketch_Core.cpp:402,555) + 48 B per hypothesis (read once, reused over the V views) + 32 B per staged validation edge."""
    return 4 * c["n_validations"] + 48 * c["n_hypotheses"] + 32 * edges_vali


def lpt_assign(costs, world):
    return pkg("dist").lpt_assign(costs, world)


def ncu_notes(kernel):
    """Pipe / issue utilisation of `kernel` from the committed ncu capture (profiles/traffic.json): what actually bounds it."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    return json.load(open(p)).get(kernel, {}).get("ncu") if os.path.exists(p) else None


def traffic(kernel, views=0, workload="", launches=0):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel`, from the committed ncu --set full capture
    (profiles/traffic.json, written by scripts/ncu_summary.py from the .ncu-rep of the same bench command), else None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        e = json.load(open(p)).get(kernel, {})
        if "bytes_per_view" in e and views:
            return e["bytes_per_view"] * views
        if workload == "cfg4":          # a step = `launches` launches of the kernel (one per unit of this rank): per-launch capture x launches
            return e["cfg4_bytes_per_launch"] * launches if "cfg4_bytes_per_launch" in e and launches else None
        return e.get("bytes") if workload in ("", "cfg2") else None
    return None


class HostBuffers:
    """Pinned host memory on both sides of the C-ABI for the e2e leg: the packed edge columns going in, result buffers coming out."""

    def __init__(self, torch, scene, V):
        self.torch = torch
        self.n_per_view = np.array([len(e) for e in scene.edges], np.int32)
        total = int(self.n_per_view.sum())
        self.pinned = torch.empty((3, total), dtype=torch.float64).pin_memory()
        self.pinned.copy_(torch.from_numpy(np.ascontiguousarray(np.concatenate([e[:, :3] for e in scene.edges], axis=0).T)))
        self.h2d_bytes = self.pinned.numel() * 8
        self.spec = {"h1": (1, 4), "h2rep": (1, 4), "count": (1, 4), "supports": (V, 4), "gamma_cam": (3, 8), "gamma_world": (3, 8),
                     "tangent_world": (3, 8), "nan_flag": (1, 4)}
        self.cap, self.buf = 0, {}
        self.arenas, self.chunk, self.arena_off, self.arena_bytes = [], 0, 0, 1 << 20

    def upload(self, g):
        g.set_all_view_edges_ptr(self.n_per_view, self.pinned[0].data_ptr(), self.pinned[1].data_ptr(), self.pinned[2].data_ptr())

    def arena_reset(self):
        self.chunk, self.arena_off = 0, 0

    def arena_ptrs(self, widths):
        """Room for one more unit (field -> bytes) in a growing list of pinned chunks: the asynchronous fetches of a whole step stay in
        flight together, so nothing is reused before arena_reset."""
        need = sum(((b + 255) // 256) * 256 for b in widths.values())
        while True:
            if self.chunk == len(self.arenas):
                self.arenas.append(self.torch.empty(max(need, self.arena_bytes if not self.arenas else self.arena_bytes // 4), dtype=self.torch.uint8).pin_memory())
            if self.arena_off + need <= self.arenas[self.chunk].numel():
                break
            self.chunk += 1
            self.arena_off = 0
        base, ptrs = self.arenas[self.chunk].data_ptr() + self.arena_off, {}
        for k, b in widths.items():
            ptrs[k] = base
            base += ((b + 255) // 256) * 256
        self.arena_off += need
        return ptrs

    def out_ptrs(self, P):
        if P > self.cap:
            self.cap = max(P + P // 4, 1024)
            for k, (wd, b) in self.spec.items():
                self.buf[k] = self.torch.empty(self.cap * wd * b, dtype=self.torch.uint8).pin_memory()
        return {k: self.buf[k].data_ptr() for k in self.spec}, sum(P * wd * b for wd, b in self.spec.values())


# ------------------------------------------------------------------------------------------------ CUDA arm: the schedule
def run_schedule(args, rank, world, local_rank):
    import torch
    use_dist = world > 1
    if use_dist:
        import torch.distributed as tdist
        tdist.init_process_group("gloo")          # rendezvous + bookkeeping only; the data plane is cvx_gather_* (NCCL in the library)
    # one rank builds, the others wait for it (a cold tree must not be dlopen-ed half written)
    if rank == 0:
        pkg("build").build_cuda(force=False)
    if use_dist:
        tdist.barrier()
    api = pkg("api")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    scene, w = build_scene(args.workload)
    V = scene.n_views - 2
    pairs = [tuple(p) for p in w["pairs"]]
    if args.pair:
        pairs = [tuple(int(t) for t in args.pair.split(","))]
    cfg = api.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols, max_support_views=w["max_support_views"])
    g = api.EdgeSketch(cfg, local_rank)
    g.set_cameras(scene.R, scene.T, scene.K)
    host = HostBuffers(torch, scene, V)
    host.upload(g)
    if use_dist:
        ids = [api.EdgeSketch.comm_unique_id() if rank == 0 else None]
        tdist.broadcast_object_list(ids, src=0)
        g.comm_init(world, rank, ids[0])
    stream = torch.cuda.ExternalStream(g.stream_ptr(), device=torch.device("cuda", local_rank))
    flush = torch.empty(L2_FLUSH_BYTES // 4, dtype=torch.int32, device="cuda")

    # ---- units: (pair, part k of S).  Bounds of the parts are balanced by hypothesis rows (cvx_split_h1_by_rows), deterministic on every rank.
    if args.shard == "h1":
        S = world
    else:
        # whole pairs while there are at least two per rank (splitting a pair costs ~8 % in launch tails: profiles/r02g); finer otherwise
        S = args.parts if args.parts > 0 else max(1, -(-2 * world // len(pairs)))
    units = [dict(id=i * S + k, pair=p, part=k) for i, p in enumerate(pairs) for k in range(S)]
    bounds_of = {}
    if args.shard == "h1":
        assign = [[u["id"] for u in units if u["part"] == r] for r in range(world)]     # rank r: its range of every pair, in schedule order
    else:
        assign = [[u["id"] for j, u in enumerate(units) if j % world == r] for r in range(world)]
    state = {"pair": None}

    def run_unit(u):
        pair = u["pair"]
        if S > 1 and pair not in bounds_of:           # once per pair and rank (warm-up): the count pass + scan of cvx_split_h1_by_rows
            g.set_pair(*pair); state["pair"] = pair
            bounds_of[pair] = [int(b) for b in g.split_h1_by_rows(S)]
        if state["pair"] != pair:
            g.set_pair(*pair); state["pair"] = pair
        if S > 1:
            g.run(bounds_of[pair][u["part"]], bounds_of[pair][u["part"] + 1])
        else:
            g.run()
        return g.device_results()                    # runs the finalize kernel; results stay on the device

    def fetch_rounds(r0, r1, to_host):
        """Rank 0 hands the gathered units of rounds [r0, r1) to the host consumer: asynchronous copies into the pinned arena."""
        n = 0
        for r in range(r0, r1):
            for q in range(world):
                u = g.gather_get(r, q)
                if int(u.meta[0]):
                    n += g.gather_fetch_into(r, q, to_host.arena_ptrs(g.gather_widths(int(u.meta[0]), int(u.meta[7]))), wait=False)
        return n

    def step(mine, rounds, measure=None, to_host=None):
        """One pass over this rank's units (+ the gather rounds).  measure: dict unit id -> (ms, counts, phases); to_host: HostBuffers."""
        d2h = 0
        fetched = [0]
        if use_dist:
            g.gather_reset()
            if to_host is not None:
                to_host.arena_reset()
        for r in range(rounds):
            if r < len(mine):
                u = units[mine[r]]
                res = run_unit(u)
                if measure is not None:
                    tm = g.timings()
                    measure[u["id"]] = (tm["set_pair_ms"] + tm["run_ms"] + tm["finalize_ms"], g.counts(), tm)
                if use_dist:
                    g.gather_post(u["id"])            # also hands round r - 1 to NCCL
                elif to_host is not None:
                    ptrs, nbytes = to_host.out_ptrs(int(res.n_pairs))
                    g.results_into(ptrs)              # accepted pairs, supports, 3-D edges, tangents -> pinned host buffers
                    d2h += nbytes
            elif use_dist:
                g.gather_post(empty=True)
            if use_dist and to_host is not None and rank == 0 and r >= 1:
                d2h += fetch_rounds(fetched[0], r, to_host)   # rounds < r: device -> pinned host behind their transfers, not waited for
                fetched[0] = r
        if use_dist:
            if to_host is not None and rank == 0:     # the last round(s) still on the device
                g.gather_wait(join_stream=True, host_sync=False)
                d2h += fetch_rounds(fetched[0], rounds, to_host)
            g.gather_wait(join_stream=True, host_sync=to_host is not None)
        return d2h

    def barrier():
        if use_dist:
            tdist.barrier()
        torch.cuda.synchronize()

    def rounds_of(a):
        return max(len(x) for x in a)

    with torch.cuda.stream(stream):
        # ---- warm-up: one cold pass (allocations), then every unit measured on the rank it was dealt to; longest-first assignment
        # from the measured times; the remaining warm-up passes run the final assignment
        state["pair"] = None
        step(assign[rank], rounds_of(assign))
        barrier()
        meas = {}
        step(assign[rank], rounds_of(assign), measure=meas)
        barrier()
        mine_meas = {k: (v[0], v[1]["n_validations"], v[1]["n_pairs"]) for k, v in meas.items()}
        if use_dist:
            allm = [None] * world
            tdist.all_gather_object(allm, mine_meas)
        else:
            allm = [mine_meas]
        cost = {k: v[0] for m in allm for k, v in m.items()}
        vals_of = {k: v[1] for m in allm for k, v in m.items()}
        # pinned arena of the e2e leg at N > 1: 100 B per accepted row + its supports as a CSR list (at most the dense V x 8 B)
        host.arena_bytes = int(sum(v[2] * (100 + 8 * min(V, 24)) + 4096 for m in allm for v in m.values())) + (1 << 20)
        if use_dist and args.shard == "pairs":
            assign, load = pkg("dist").plan_units([cost[u["id"]] for u in units], [u["pair"] for u in units], world)
        else:
            load = [sum(cost[i] for i in a) for a in assign]
        mine, rounds = assign[rank], rounds_of(assign)
        vals_all = int(sum(vals_of.values()))
        for _ in range(max(args.warmup, 3) - 2):
            flush.zero_()
            step(mine, rounds)
        barrier()
        launches0 = g.timings()["kernel_launches"]
        sampler = ClockSampler(local_rank)
        sampler.start()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        barrier()
        wall0 = time.perf_counter()
        for i in range(args.steps):
            flush.zero_()                     # L2 flush between timed iterations (outside the events)
            ev[i][0].record(stream)
            step(mine, rounds)
            ev[i][1].record(stream)
        barrier()
        wall = time.perf_counter() - wall0
        clocks = sampler.stop()
        dev_ms = sum(a.elapsed_time(b) for a, b in ev)
        launches = g.timings()["kernel_launches"] - launches0
        gstats = g.gather_stats() if use_dist else None
        # ---- one profiling step: per-unit phase times of this rank (event-timed inside the library on the launching stream)
        prof = {}
        step(mine, rounds, measure=prof)
        barrier()
        # ---- e2e leg: host buffers in, host buffers out, copies inside the timed region (wall clock, max over ranks)
        for _ in range(2):
            host.upload(g); state["pair"] = None
            step(mine, rounds, to_host=host)
        barrier()
        t0 = time.perf_counter()
        d2h = 0
        for _ in range(args.steps):
            flush.zero_()
            host.upload(g); state["pair"] = None     # a new upload invalidates the staged pair
            d2h = step(mine, rounds, to_host=host)
        barrier()
        e2e_s = time.perf_counter() - t0

    busy_ms = sum(v[0] for v in prof.values())
    per_rank = None
    if use_dist:
        me = {"rank": rank, "units": [[units[i]["pair"], units[i]["part"]] for i in mine], "ms_per_step": dev_ms / args.steps, "busy_ms": busy_ms,
              "planned_ms": load[rank], "validations": int(sum(vals_of[i] for i in mine)), "gather": gstats,
              "phases_ms": {k: round(sum(v[2][k] for v in prof.values()), 3) for k in ("set_pair_ms", "form_ms", "prep_ms", "validate_ms", "select_ms", "finalize_ms")}}
        per_rank = [None] * world
        tdist.all_gather_object(per_rank, me)
        red = [None] * world
        tdist.all_gather_object(red, (dev_ms, e2e_s, wall, launches, host.h2d_bytes, d2h))
        dev_ms, e2e_s, wall = (max(x[k] for x in red) for k in range(3))
        launches_all, h2d_all, d2h_all = (sum(x[k] for x in red) for k in (3, 4, 5))
    else:
        launches_all, h2d_all, d2h_all = launches, host.h2d_bytes, d2h
    if rank != 0:
        g.close()
        if use_dist:
            tdist.destroy_process_group()
        return
    ms_per_step = dev_ms / args.steps
    value = vals_all / (ms_per_step * 1e-3)
    # ---- roofline of the dominant kernel of rank 0's units
    ph = {k: sum(v[2][k + "_ms"] for v in prof.values()) for k in ("form", "prep", "validate", "select")}
    csum = {k: int(sum(v[1][k] for v in prof.values())) for k in ("n_validations", "n_hypotheses", "n_unique_hypotheses", "n_candidates", "n_pairs", "n_inliers")}
    edges_vali = sum(sum(len(scene.edges[v]) for v in range(scene.n_views) if v not in units[i]["pair"]) for i in mine) // max(S, 1)
    dom = max(ph, key=ph.get)
    peak, peak_src = peaks()
    kname = {"validate": "k_validate<false,false>", "form": "k_form + k_cluster_classes", "select": "k_select", "prep": "k_wedge31"}[dom]
    fused = algorithmic_bytes_fused(csum, edges_vali)
    form_bytes = 100 * csum["n_hypotheses"] + 40 * csum["n_candidates"]
    abytes = fused if dom == "validate" else form_bytes
    ach = abytes / (ph[dom] * 1e-3) / 1e9
    roof = {"bound": "hbm", "kernel": kname, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic(kname, workload=args.workload, launches=len(mine)),
            "peak_source": peak_src, "kernel_ms": ph[dom], "algorithmic_bytes": abytes,
            "bytes_formula": ("4 x validations + 48 x hypotheses + 32 x staged validation edges (SURVEY.md 8d, fused kernel)" if dom == "validate"
                              else "100 B per hypothesis row written + 40 B per raw candidate read (DESIGN.md section 4)"),
            "kernel_ms_all": {k: round(v, 3) for k, v in ph.items()}, "ncu": ncu_notes(kname),
            "validate_kernel": {"kernel": "k_validate<false,false>", "kernel_ms": ph["validate"], "algorithmic_bytes": fused,
                                "achieved": fused / (ph["validate"] * 1e-3) / 1e9, "frac": fused / (ph["validate"] * 1e-3) / 1e9 / peak,
                                "ncu": ncu_notes("k_validate<false,false>")},
            "note": "the fused path is FP64-issue bound, not HBM bound (SURVEY.md 8d: ~1.3k FP64 flop-equivalents per validation against a few "
                    "bytes); the HBM-roofline target of north_star is the stand-alone support test: see roofline_support_test"}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": workload_name(args.workload, scene, w)},
            "run": {"l2": f"flushed between timed iterations ({L2_FLUSH_BYTES >> 20} MiB write)", "validations_per_step": vals_all,
                    "evaluated_queries_per_step": int(sum(v[1]["n_unique_hypotheses"] for v in prof.values())) * V if not use_dist else None,
                    "units": len(units), "parts_per_pair": S, "shard": args.shard,
                    "parallelism": "1 GPU" if world == 1 else f"{world} GPUs, units dealt longest-first from measured times, cvx_gather (NCCL) of every unit to rank 0",
                    "unit_ms": {str(k): round(v, 3) for k, v in sorted(cost.items())}},
            "clocks": clocks,
            "e2e": {"value": vals_all * args.steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all,
                    "ms_per_step": 1e3 * e2e_s / args.steps, "timer": "host wall clock around the C-ABI calls, max over ranks"},
            "gpu_launches": launches_all, "wall_s_timed_region": wall, "roofline": roof,
            "phases_ms": {k: round(v, 3) for k, v in ph.items()}, "counts_rank0": csum}
    if per_rank:
        line["per_rank"] = per_rank
        busy = [p["busy_ms"] for p in per_rank]
        line["run"]["balance"] = {"mean_over_max_busy": sum(busy) / len(busy) / max(busy), "idle_frac_worst_rank": 1 - min(busy) / max(busy)}
    g.close()
    if world == 1 and not args.no_extras:
        extras(args, line, local_rank)
    if not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_baseline(scene, w, pairs[0], args.cpu_seconds)
        if not args.no_extras:
            try:
                line["cpu_baseline"]["fast_math"] = cpu_baseline(scene, w, pairs[0], args.cpu_seconds / 2, variant="fast")
            except Exception as e:
                line["cpu_baseline"]["fast_math"] = {"error": str(e)}
    print(json.dumps(line), flush=True)
    if use_dist:
        tdist.destroy_process_group()


# ------------------------------------------------------------------------------------------------ N = 1 extras
def extras(args, line, local_rank):
    """secondary (cfg2 single pair), roofline_support_test (cfg5), sec_per_object (cfg1) -- each guarded: a failure is recorded in the
    line instead of losing it."""
    for key, fn in (("secondary", lambda: secondary_cfg2(args, local_rank)),
                    ("roofline_support_test", lambda: support_problem_bench(args, 0, 1, local_rank, args.cfg5_views or 40, args.steps, quiet=True)),
                    ("sec_per_object", lambda: sec_per_object(local_rank))):
        try:
            t0 = time.perf_counter()
            line[key] = fn()
            line[key]["bench_wall_s"] = round(time.perf_counter() - t0, 2)
        except Exception as e:      # noqa: BLE001
            line[key] = {"error": f"{type(e).__name__}: {e}"}


def secondary_cfg2(args, local_rank):
    """BASELINE.json configs[1]: synthetic 50 views x 10k edges/view, single hypothesis view pair (0,1) -- the round-1 headline workload."""
    import torch
    api = pkg("api")
    scene, w = build_scene("cfg2")
    V = scene.n_views - 2
    pair = tuple(int(t) for t in args.pair.split(",")) if args.pair else tuple(w["pairs"][0])
    cfg = api.default_config(n_views=scene.n_views, img_rows=scene.rows, img_cols=scene.cols, max_support_views=w["max_support_views"])
    g = api.EdgeSketch(cfg, local_rank)
    g.set_cameras(scene.R, scene.T, scene.K)
    host = HostBuffers(torch, scene, V)
    host.upload(g)
    stream = torch.cuda.ExternalStream(g.stream_ptr(), device=torch.device("cuda", local_rank))
    flush = torch.empty(L2_FLUSH_BYTES // 4, dtype=torch.int32, device="cuda")
    steps = max(args.steps, 5)
    with torch.cuda.stream(stream):
        for _ in range(3):
            flush.zero_(); g.set_pair(*pair); g.run(); g.device_results()
        torch.cuda.synchronize()
        counts = g.counts()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        parts = {k: 0.0 for k in ("set_pair_ms", "form_ms", "prep_ms", "validate_ms", "select_ms", "finalize_ms", "run_ms")}
        l0 = g.timings()["kernel_launches"]
        for i in range(steps):
            flush.zero_()
            ev[i][0].record(stream)
            g.set_pair(*pair); g.run(); g.device_results()
            ev[i][1].record(stream)
            tm = g.timings()
            for k in parts:
                parts[k] += tm[k]
        torch.cuda.synchronize()
        launches = (g.timings()["kernel_launches"] - l0) // steps
        dev_ms = sum(a.elapsed_time(b) for a, b in ev) / steps
        for _ in range(2):
            host.upload(g); g.set_pair(*pair); g.run(); P = int(g.device_results().n_pairs); g.results_into(host.out_ptrs(P)[0])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            flush.zero_()
            host.upload(g); g.set_pair(*pair); g.run()
            P = int(g.device_results().n_pairs)
            ptrs, d2h = host.out_ptrs(P)
            g.results_into(ptrs)
        torch.cuda.synchronize()
        e2e_ms = 1e3 * (time.perf_counter() - t0) / steps
    g.close()
    vals = counts["n_validations"]
    edges_vali = sum(len(scene.edges[v]) for v in range(scene.n_views) if v not in pair)
    peak, _ = peaks()
    val_ms = parts["validate_ms"] / steps
    fused = algorithmic_bytes_fused(counts, edges_vali)
    return {"workload": workload_name("cfg2", scene, w), "value": vals / (dev_ms * 1e-3), "unit": UNIT, "ms_per_step": dev_ms,
            "evaluated_queries_per_sec": counts["n_unique_hypotheses"] * V / (dev_ms * 1e-3),
            "e2e": {"value": vals / (e2e_ms * 1e-3), "ms_per_step": e2e_ms, "h2d_bytes_per_step": host.h2d_bytes, "d2h_bytes_per_step": d2h},
            "gpu_launches_per_step": launches, "phases_ms": {k: round(v / steps, 4) for k, v in parts.items()},
            "validate_kernel": {"kernel": "k_validate<false,false>", "kernel_ms": val_ms, "algorithmic_bytes_fused": fused,
                                "achieved_gbs": fused / (val_ms * 1e-3) / 1e9, "frac": fused / (val_ms * 1e-3) / 1e9 / peak,
                                "ncu": ncu_notes("k_validate<false,false>"), "traffic": traffic("k_validate<false,false>")},
            "counts": counts}


def sec_per_object(local_rank):
    """BASELINE.json metric (ii) on configs[0]: the C++ drop-in (curvix_main) on ABC-NEF 00000006 with configs/3D_Curvix_ABC_NEF.yaml's
    parameters to the 0.7 stop ratio / 25 passes, timers as cmd/main.cpp:165-170 prints them; beside it the reference's own
    EdgeSketch_Core (oracle/_ref) on the FIRST pass of the same schedule (a full reference run is ~25 x that: BASELINE.md)."""
    data = pkg("data")
    exe = pkg("build").build_host(force=False)
    fixture = os.path.join(ROOT, "tests", "golden", "abc_nef_00000006_t1.npz")
    if not exe or not os.path.exists(fixture):
        return {"error": "curvix_main or the ABC-NEF fixture is missing"}
    scene = data.Scene.load_npz(fixture, "ABC-NEF/00000006")
    cvf = os.path.join(ROOT, "tests", "golden", "abc_nef_00000006_curvelets_t1.npz")
    if os.path.exists(cvf):
        scene.load_curvelets(cvf)
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        scene.write_reference_layout(tmp, dataset="ABC-NEF", scene="00000006", thresh=1)
        yaml = os.path.join(tmp, "abc.yaml")
        ref_yaml = os.path.join(ROOT, "tests", "golden", "3D_Curvix_ABC_NEF.yaml")
        text = open(ref_yaml).read() if os.path.exists(ref_yaml) else ABC_YAML_FALLBACK
        lines = [("Dataset_Path: " + tmp + "/") if l.startswith("Dataset_Path:") else l for l in text.splitlines()]
        open(yaml, "w").write("\n".join(lines) + "\n")
        sched = os.path.join(tmp, "schedule.txt")
        cmd = [exe, "-c", yaml, "--outputs", os.path.join(tmp, "outputs"), "--schedule", sched, "--device", str(local_rank), "--refined", "0", "--quiet"]
        subprocess.run(cmd, capture_output=True, text=True, timeout=600)           # warm run (edge parsing cache, CUDA context)
        t0 = time.perf_counter()
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        wall = time.perf_counter() - t0
        if r.returncode != 0:
            return {"error": "curvix_main failed: " + (r.stderr or r.stdout)[-400:]}
        tm = {}
        for l in r.stdout.splitlines():
            for key, name in (("Pairing up Edges", "pair_edges_s"), ("Finalizing Edge Pairs", "finalize_s"), ("Finding Next Hypothesis Views", "next_views_s"),
                              ("Stacking 3D edges", "stack_s"), ("Total computation time", "total_s")):
                if key in l:
                    try:
                        tm[name] = float(l.replace("(s)", " ").split(":")[-1].split()[0])
                    except ValueError:
                        pass
            if "consolidation:" in l:      # EdgeMapping::consolidate_3D_edges of the drop-in (links / prunings / alignment on the device, curves on the host)
                try:
                    tm["consolidation_s"] = float(l.split("consolidation:")[1].split()[0])
                    tm["consolidation"] = l.split("(s):", 1)[1].strip()
                except (ValueError, IndexError):
                    pass
        rows = [l.split() for l in open(sched).read().splitlines()] if os.path.exists(sched) else []
        out["b200"] = dict(tm, passes=len(rows), process_wall_s=round(wall, 3),
                           sec_per_object=tm.get("pair_edges_s", 0.0) + tm.get("finalize_s", 0.0) + tm.get("stack_s", 0.0),
                           definition="pair_edges_time + finalize_edge_pair_time + Stack_3D_Edges over all passes (SURVEY.md 8d (ii))",
                           schedule=[[int(x[1]), int(x[2])] for x in rows], validations=int(sum(int(x[4]) for x in rows)) if rows and len(rows[0]) > 5 else None)
    try:
        from oracle import ref_py
        if ref_py.available():
            T = cpu_threads()
            ref = ref_py.Reference.for_scene(scene, num_threads=T)
            ref.set_pair(47, 42)
            ref.run(); ref.finalize()
            t = ref.times()
            out["reference_first_pass"] = dict(pair=[47, 42], pair_edges_s=t["pair_edges"], finalize_s=t["finalize"], cores=T,
                                               n_pairs=int(len(ref.pairs()["h1"])), build=ref_py.build_info())
            ref.close()
    except Exception as e:      # noqa: BLE001
        out["reference_first_pass"] = {"error": str(e)}
    return out


ABC_YAML_FALLBACK = """Num_Of_OMP_Threads: 32
Init_Hypo1_View_Index: 47
Init_Hypo2_View_Index: 42
delta: 0.4
delta_theta: 15
Max_Num_Of_Support_Views: 4
TOED_Thresh: 1
Parallel_Epipolar_Line_Angle: 15
Reproj_Dist_Thresh: 2
Ratio_Of_Claimed_Edges_to_Stop: 0.7
Max_Num_Of_3D_Edge_Sketch_Passes: 25
Dataset_Path: /tmp/
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


# ------------------------------------------------------------------------------------------------ cfg5: support-test stress
CFG5 = dict(n_views=200, edges_per_view=200000, n_groups=50000, rows_per_group=20, delta_theta_deg=2.0)


def support_problem_bench(args, rank, world, local_rank, n_views, steps, quiet=False):
    """BASELINE.json configs[4]: the isolated support test (getHYPO2_idx_Ore x2 + set_intersection + getSupportIdx,
    EdgeSketch_Core.cpp:348-384) on pre-formed hypotheses, n_views validation views x 200k edges/view, ~1M hypothesis rows,
    delta_theta = 2 deg.  HBM bound: SURVEY.md 8d bytes = 52 + 16 n per validation + 32 B per staged edge.  -> dict."""
    import torch
    api = pkg("api"); synth = pkg("support_synth")
    shape = dict(CFG5, n_views=n_views)
    prob = synth.make_support_problem(seed=20251109 + rank, **shape)
    V, H = shape["n_views"], int(prob["group_off"][-1])
    g = api.EdgeSketch(api.default_config(n_views=8), local_rank)
    order = ("group_off", "view_off", "ang31", "ang32", "cs", "sn", "wedge31", "wedge32", "tgt")
    t0 = time.perf_counter()
    sp = api.SupportProblem(g, *[prob[k] for k in order], prob["thr"], prob["parallel_deg"])
    create_s = time.perf_counter() - t0
    h2d = sum(np.asarray(prob[k]).nbytes for k in order)
    stats = sp.stats()
    for _ in range(3):
        sp.run()
    sampler = ClockSampler(local_rank)
    sampler.start()
    torch.cuda.synchronize()
    ms = [sp.run() for _ in range(steps)]      # inputs (1.6 GB at 40 views) >> L2 (126 MB): no flush needed between launches
    torch.cuda.synchronize()
    clocks = sampler.stop()
    res = sp.result()
    dev_ms = float(sum(ms)) / steps
    vals = V * H
    peak, peak_src = peaks()
    abytes = vals * 52 + 16 * stats["n_inliers"] + 32 * int(prob["view_off"][-1])
    ach = abytes / (dev_ms * 1e-3) / 1e9
    out = {"workload": f"cfg5: support-test stress, {V} validation views x {shape['edges_per_view']} edges/view, {H} pre-formed hypothesis rows in "
                       f"{shape['n_groups']} groups, delta_theta {shape['delta_theta_deg']} deg",
           "bound": "hbm", "kernel": "k_validate<true,false>", "kernel_ms": dev_ms, "validations_per_launch": vals, "value": vals / (dev_ms * 1e-3), "unit": UNIT,
           "algorithmic_bytes": abytes, "achieved": ach, "peak": peak, "peak_source": peak_src, "frac": ach / peak,
           "bytes_formula": "validations x 52 + 16 x wedge-intersection members + 32 x staged edges (SURVEY.md 8d)",
           "traffic": traffic("k_validate<true,false>", V), "ncu": ncu_notes("k_validate<true,false>"),
           "mean_intersection": stats["n_inliers"] / vals, "entries_visited_per_validation": stats["n_visited"] / vals,
           "clocks": clocks, "h2d_bytes": h2d, "create_s": create_s, "stats": stats}
    # bounded CPU sample of the same arrays (view 0, the first groups): exact comparison of the support indices
    from oracle import oracle_py
    oracle_py.build(fast=False)
    threads = cpu_threads()
    E = shape["edges_per_view"]
    gs = 256 if quiet else 1024
    hs = int(prob["group_off"][gs])
    sub = dict(group_off=prob["group_off"][:gs + 1], view_off=prob["view_off"][:2], ang31=prob["ang31"][:E], ang32=prob["ang32"][:E],
               cs=prob["cs"][:E], sn=prob["sn"][:E], wedge31=prob["wedge31"][:1, :gs], wedge32=prob["wedge32"][:1, :hs],
               tgt=prob["tgt"][:1, :hs], thr=prob["thr"], parallel_deg=prob["parallel_deg"])
    t0 = time.perf_counter()
    exp = oracle_py.support_test(threads=threads, **sub)
    dt = time.perf_counter() - t0
    out["gpu_matches_on_sample"] = bool(np.array_equal(exp[0], res[0, :hs]))
    out["cpu_sample"] = {"value": hs / dt, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"view 0, groups [0,{gs}) ({hs} validations, {dt:.1f} s), orc_support_test (Theta(E3) scans per validation as the reference)"}
    sp.close(); g.close()
    return out


def run_support_stress(args, rank, world, local_rank):
    """--workload cfg5: the stand-alone support test as the bench line (one problem per GPU, weak scaling, no collective)."""
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    use_dist = world > 1
    if use_dist:
        import torch.distributed as tdist
        tdist.init_process_group("gloo")
    r = support_problem_bench(args, rank, world, local_rank, args.cfg5_views or CFG5["n_views"], args.steps)
    vals, ms = r["validations_per_launch"], r["kernel_ms"]
    if use_dist:
        allr = [None] * world
        tdist.all_gather_object(allr, (vals, ms))
        vals, ms = sum(x[0] for x in allr), max(x[1] for x in allr)
    if rank == 0:
        line = {"metric": METRIC, "value": vals / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": 3, "ms_per_step": ms,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": r["workload"]}, "clocks": r["clocks"],
                "e2e": {"value": None, "unit": UNIT, "h2d_bytes_per_step": r["h2d_bytes"], "d2h_bytes_per_step": 4 * r["validations_per_launch"]},
                "gpu_launches": args.steps,
                "roofline": {k: r[k] for k in ("bound", "kernel", "achieved", "peak", "frac", "traffic", "kernel_ms", "algorithmic_bytes", "bytes_formula",
                                               "mean_intersection", "entries_visited_per_validation", "peak_source", "ncu")},
                "cpu_baseline": r["cpu_sample"], "gpu_matches_on_sample": r["gpu_matches_on_sample"]}
        line["roofline"]["unit"] = "GB/s"
        print(json.dumps(line), flush=True)
    if use_dist:
        tdist.destroy_process_group()


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
        run_schedule(args, rank, world, local_rank)


if __name__ == "__main__":
    main()

"""world_size-2 gloo test of the multi-GPU host logic (3d_curvix_b200/dist.py): sharding helpers and the variable-length
gather of accepted pairs + 3-D edges to rank 0 in rank order.  The payload is produced by the CPU oracle on two H1 ranges,
so the gathered result must equal the unsharded pass."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, make_oracle, pkg


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, payloads, out_path):
    sys.path.insert(0, ROOT)
    import importlib
    distm = importlib.import_module("3d_curvix_b200.dist")
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    local = {k: torch.from_numpy(v) for k, v in payloads[rank].items()}
    got, counts = distm.gather_results(local)
    # steady-state form: fixed capacity, one collective; must deliver the same rows
    fields = {k: (tuple(v.shape[1:]), v.dtype) for k, v in local.items()}
    cap = torch.tensor([local["h1"].shape[0]], dtype=torch.int64)
    dist.all_reduce(cap, op=dist.ReduceOp.MAX)
    pg = distm.PackedGather(fields, int(cap.item()) + 3, torch.device("cpu"))
    for _ in range(2):      # reusable
        got2, counts2 = pg.run(local, concat=True)
    if rank == 0:
        assert counts2 == counts and all(torch.equal(got[k], got2[k]) for k in got)
        np.savez(out_path, counts=np.array(counts), **{k: v.numpy() for k, v in got.items()})
    else:
        assert got is None and got2 is None
    dist.barrier()
    dist.destroy_process_group()


def test_shard_helpers():
    d = pkg("dist")
    for n in (0, 1, 7, 10, 2817):
        for w in (1, 2, 3, 8):
            r = [d.shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n and all(r[k][1] == r[k + 1][0] for k in range(w - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1
    pairs = [(2 * k, 2 * k + 1) for k in range(5)]
    got = sorted(p for k in range(2) for p in d.shard_pairs(pairs, k, 2))
    assert got == pairs


def test_gather_two_ranks_matches_unsharded(oracle_mod, tiny_scene, tmp_path):
    scene, w = tiny_scene
    d = pkg("dist")
    o = make_oracle(oracle_mod, scene, max_support_views=w["max_support_views"])
    o.set_pair(0, 1)
    o.run(mode=1); pf = o.pairs(); ff = o.finalize()
    E1 = len(scene.edges[0])
    payloads = []
    for k in range(2):
        a, b = d.shard_range(E1, k, 2)
        o.run(a, b, mode=1); p = o.pairs(); f = o.finalize()
        payloads.append({"h1": p["h1"].copy(), "h2": p["h2rep"].copy(), "sup": p["supports"].copy(), "gw": f["gamma_world"].copy(),
                         "tw": f["tangent_world"].copy()})
    out = str(tmp_path / "gathered.npz")
    mp.spawn(_worker, args=(2, _free_port(), payloads, out), nprocs=2, join=True)
    g = np.load(out)
    assert g["counts"].sum() == len(pf["h1"]) and (g["counts"] > 0).all()
    assert np.array_equal(g["h1"], pf["h1"]) and np.array_equal(g["h2"], pf["h2rep"]) and np.array_equal(g["sup"], pf["supports"])
    assert np.array_equal(g["gw"], ff["gamma_world"]) and np.array_equal(g["tw"], ff["tangent_world"])


def test_unit_plan_is_balanced_and_deterministic():
    """LPT over measured unit costs (bench.py --gpus N, curvix_main --devices): every unit exactly once, loads within the LPT bound,
    units of one pair adjacent on a rank, the same plan on every call."""
    dist = pkg("dist")
    costs = [146.6, 118.4, 261.4, 118.4, 88.5, 157.6, 151.3, 153.3, 232.8, 248.2, 176.0, 141.5, 140.4, 104.2, 212.6, 122.0]   # cfg4, r02 probe (ms)
    pairs = [(2 * k, 2 * k + 1) for k in range(16)]
    for world in (1, 2, 4, 8):
        a, load = dist.plan_units(costs, pairs, world)
        assert sorted(i for r in a for i in r) == list(range(16))
        assert max(load) <= sum(costs) / world + max(costs) * (1 - 1 / world) + 1e-9      # Graham's bound for list scheduling
        assert dist.plan_units(costs, pairs, world) == (a, load)
    a8, load8 = dist.plan_units(costs, pairs, 8)
    assert sum(load8) / 8 / max(load8) > 0.9                                                # 16 whole pairs on 8 ranks: ~0.91
    # two parts per pair: finer units balance better, and both parts of a pair that land on one rank are adjacent
    c2 = [c / 2 for c in costs for _ in range(2)]
    p2 = [p for p in pairs for _ in range(2)]
    b, load = dist.plan_units(c2, p2, 8)
    assert sum(load) / 8 / max(load) > 0.95
    for r in b:
        seen = []
        for i in r:
            if p2[i] in seen:
                assert seen[-1] == p2[i]
            seen.append(p2[i])

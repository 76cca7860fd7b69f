"""Multi-GPU check of the C-ABI gather (cvx_comm.cu), run as one process per GPU:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 scripts/dist_check.py
torch.distributed is used for the rendezvous only (broadcast of the 128-byte NCCL id); every transfer goes through cvx_gather_*.
Checks, on rank 0, bit for bit:
  (1) pairs of a fixed schedule dealt to the ranks (unequal unit counts: some ranks post empty rounds) == the same pairs recomputed on rank 0;
  (2) ONE pair sharded by view-1 range with cvx_split_h1_by_rows == the unsharded pass (rows concatenated in rank order);
  (3) cvx_claims_allreduce: per-view claimed counts of the sharded pass == the unsharded pass.
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as tdist

api = importlib.import_module("3d_curvix_b200.api")
data = importlib.import_module("3d_curvix_b200.data")

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
tdist.init_process_group("gloo")
scene = data.make_synthetic_scene(n_views=12, edges_per_view=3000, rows=800, cols=800, seed=31, n_curves=10)
g = api.EdgeSketch(api.default_config(n_views=12, img_rows=800, img_cols=800, max_support_views=3), local)
g.set_cameras(scene.R, scene.T, scene.K)
g.set_all_view_edges(scene.edges)
ids = [api.EdgeSketch.comm_unique_id() if rank == 0 else None]
tdist.broadcast_object_list(ids, src=0)
g.comm_init(world, rank, ids[0])

FIELDS = ("h1", "h2rep", "count", "nan_flag", "supports", "gamma_world", "tangent_world", "gamma_cam", "h2_corrected")


def local_results():
    p, f = g.pairs(), g.finalize()
    return dict(h1=p["h1"], h2rep=p["h2rep"], count=p["count"], nan_flag=f["nan_flag"], supports=p["supports"], gamma_world=f["gamma_world"],
                tangent_world=f["tangent_world"], gamma_cam=f["gamma_cam"], h2_corrected=f["h2_corrected"])


# ---- (1) fixed schedule, 2 * world + 1 pairs dealt round-robin: the last round has one contributor
pairs = [(k % 12, (k + 5) % 12) for k in range(2 * world + 1)]
mine = [i for i in range(len(pairs)) if i % world == rank]
rounds = (len(pairs) + world - 1) // world
for step in range(2):           # twice: the second time reuses staging buffers, arena and round records
    g.gather_reset()
    for r in range(rounds):
        if r < len(mine):
            a, b = pairs[mine[r]]
            g.set_pair(a, b); g.run()
            g.gather_post(unit_id=mine[r])
        else:
            g.gather_post(empty=True)
    g.gather_wait(join_stream=True, host_sync=True)
    if rank == 0:
        seen = 0
        for r in range(rounds):
            for q in range(world):
                i = r * world + q
                u = g.gather_fetch(r, q)
                if i >= len(pairs):
                    assert u["meta"][0] == 0
                    continue
                a, b = pairs[i]
                assert u["meta"][1:3] == [a, b] and u["meta"][5] == i, (u["meta"], a, b, i)
                g.set_pair(a, b); g.run()
                want = local_results()
                assert u["meta"][0] == len(want["h1"]) > 0
                for k in FIELDS:
                    assert np.array_equal(u[k], want[k], equal_nan=True), (i, k)
                seen += 1
        assert seen == len(pairs)
        print(f"[dist_check] step {step}: {seen} units gathered from {world} ranks == recomputed on rank 0;", g.gather_stats(), flush=True)
tdist.barrier()

# ---- (2) + (3) one pair sharded by view-1 range, balanced by rows
a, b = 0, 1
g.set_pair(a, b)
bounds = g.split_h1_by_rows(world)
g.claims_reset()
g.gather_reset()
g.run(int(bounds[rank]), int(bounds[rank + 1]))
g.claims_accumulate()
g.gather_post(unit_id=rank)
g.gather_wait(join_stream=True, host_sync=True)
g.claims_allreduce()
claimed_sharded, _ = g.claims_counts()
if rank == 0:
    parts = [g.gather_fetch(0, q) for q in range(world)]
    assert [p["meta"][3:5] for p in parts] == [[int(bounds[q]), int(bounds[q + 1])] for q in range(world)]
    g.claims_reset()
    g.set_pair(a, b); g.run(); g.claims_accumulate()
    want = local_results()
    for k in FIELDS:
        got = np.concatenate([p[k] for p in parts])
        assert np.array_equal(got, want[k], equal_nan=True), k
    claimed_full, _ = g.claims_counts()
    assert np.array_equal(claimed_sharded, claimed_full), (claimed_sharded, claimed_full)
    rows = [int(p["meta"][0]) for p in parts]
    print(f"[dist_check] one pair sharded over {world} ranks by rows: bounds {bounds.tolist()}, accepted rows per rank {rows} == unsharded pass; "
          f"claims after cvx_claims_allreduce == unsharded ({int(claimed_full.sum())} edges)", flush=True)
tdist.barrier()
g.close()
tdist.destroy_process_group()
if rank == 0:
    print("[dist_check] OK", flush=True)

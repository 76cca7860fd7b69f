"""Multi-GPU layer of the edge-sketch path: one process per GPU, torch.distributed for the plumbing.

The path shards with NO data-path collective: inside a pass every view-1 edge is independent
(EdgeSketch_Core.cpp:209, disjoint output rows :515-523), so rank r processes a contiguous range of view-1 edges against
its own resident copy of all views; with a fixed schedule of hypothesis view pairs the pairs themselves are dealt
round-robin.  The only exchange is the final variable-length gather of accepted pairs + 3-D edges + tangents to rank 0,
concatenated in rank order = ascending view-1 index, i.e. the single-GPU / CPU row order that EdgeMapping insertion and
the output files expect.
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_range(n_items, rank, world):
    """Contiguous, balanced [begin, end) of rank `rank` among `world` ranks (first n_items % world ranks get one more)."""
    base, rem = divmod(int(n_items), int(world))
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def shard_pairs(pairs, rank, world):
    """Fixed hypothesis-pair schedule dealt round-robin (configs 3/4: the schedule is known up front)."""
    return [p for i, p in enumerate(pairs) if i % world == rank]


def lpt_assign(costs, world):
    """Longest-processing-time-first: units by decreasing cost, each to the least loaded rank.  -> (unit lists per rank, loads)."""
    order = sorted(range(len(costs)), key=lambda i: (-costs[i], i))
    load = [0.0] * world
    out = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda q: (load[q], q))
        out[r].append(i)
        load[r] += costs[i]
    return out, load


def plan_units(costs, pair_of, world):
    """Static plan of a fixed schedule cut into units (a hypothesis view pair, or a view-1 range of one): LPT over the measured unit
    costs, then inside every rank the units of one pair made adjacent (the pair is staged once), pairs by decreasing cost of their
    heaviest unit.  Every rank computes the same plan from the same inputs.  -> (unit lists per rank, planned loads)."""
    assign, load = lpt_assign(costs, world)
    out = []
    for a in assign:
        heaviest = {}
        for i in a:
            heaviest[pair_of[i]] = max(heaviest.get(pair_of[i], 0.0), costs[i])
        out.append(sorted(a, key=lambda i: (-heaviest[pair_of[i]], pair_of[i], i)))
    return out, load


class _DevArray:
    """Exposes a raw device pointer of the C-ABI (cvx_get_device_results) through __cuda_array_interface__."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 3, "strides": None}


def device_tensor(ptr, shape, dtype):
    """Zero-copy torch view of device memory owned by a cvx context (valid until its next run)."""
    n = int(np.prod(shape))
    if n == 0 or not ptr:
        return torch.empty(tuple(shape), dtype=dtype, device="cuda")
    ts = {torch.int32: "<i4", torch.float64: "<f8", torch.int64: "<i8"}[dtype]
    return torch.as_tensor(_DevArray(ptr, shape, ts), device="cuda")


def gather_varlen(t, group=None, dst=0):
    """Gather first-dimension-ragged tensors to rank `dst` (concatenated in rank order).  Works on gloo (CPU tensors)
    and nccl (CUDA tensors): all_gather of the row counts, then one padded all_gather (payload is KB-MB: latency-bound,
    not bandwidth-bound, so one fused collective beats per-rank send/recv)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    n = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    mx = max(max(counts), 1)
    pad = torch.zeros((mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    pad[: t.shape[0]] = t
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    if rank != dst:
        return None, counts
    return torch.cat([b[:c] for b, c in zip(bufs, counts)], 0), counts


def gather_results(local, group=None, dst=0):
    """local: dict name -> tensor with a common ragged first dimension (accepted pairs).  Returns (dict concatenated in rank
    order on dst | None elsewhere, per-rank row counts).  Two collectives in total: the row counts, then ONE all_gather of a
    byte buffer into which every field is packed (fields back to back, padded to the largest rank) -- the payload is KB-MB,
    so the exchange is latency-bound and the number of collectives is what matters."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    keys = sorted(local)
    first = local[keys[0]]
    dev = first.device
    n = torch.tensor([first.shape[0]], dtype=torch.int64, device=dev)
    cl = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(cl, n, group=group)
    counts = [int(c) for c in torch.cat(cl).tolist()]
    row_bytes = {k: int(np.prod(local[k].shape[1:], dtype=np.int64)) * local[k].element_size() for k in keys}
    per_row = sum(row_bytes.values())
    mx = max(max(counts), 1)
    # field k of a rank with c rows occupies bytes [c * prefix_k, c * prefix_k + c * row_bytes_k): 8-byte aligned when every
    # wider-than-4-byte field comes first
    order = sorted(keys, key=lambda k: (-local[k].element_size(), k))
    buf = torch.empty(mx * per_row, dtype=torch.uint8, device=dev)
    off = 0
    for k in order:
        nb = first.shape[0] * row_bytes[k]
        if nb:
            buf[off:off + nb] = local[k].contiguous().view(-1).view(torch.uint8)
        off += nb
    bufs = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(bufs, buf, group=group)
    if rank != dst:
        return None, counts
    out = {}
    for k in keys:
        parts = []
        for r in range(world):
            c = counts[r]
            o = sum(c * row_bytes[q] for q in order[:order.index(k)])
            parts.append(bufs[r][o:o + c * row_bytes[k]].view(local[k].dtype).view((c,) + tuple(local[k].shape[1:])))
        out[k] = torch.cat(parts, 0)
    return out, counts


class PackedGather:
    """Steady-state form of gather_results for a known row capacity: ONE collective per call and no host synchronisation on the
    sending ranks.  Every rank packs [int64 row count | fields back to back, widest element type first] into a fixed-size byte
    buffer (capacity `cap_rows`, agreed once, e.g. from an all_reduce MAX over a warm-up pass); one all_gather_into_tensor moves
    the buffers; rank `dst` reads the counts (its only synchronisation) and returns per-rank views in rank order = ascending
    pair / view-1 order.  A rank with more rows than the capacity raises (grow the capacity collectively and retry)."""

    HEADER = 16

    def __init__(self, fields, cap_rows, device, group=None, dst=0):
        """fields: dict name -> (row_shape tuple, torch dtype)."""
        self.group, self.dst, self.cap = group, dst, int(cap_rows)
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        esz = lambda dt: torch.empty((), dtype=dt).element_size()
        self.order = sorted(fields, key=lambda k: (-esz(fields[k][1]), k))
        self.fields = {k: (tuple(fields[k][0]), fields[k][1], int(np.prod(fields[k][0], dtype=np.int64)) * esz(fields[k][1])) for k in fields}
        per_row = sum(v[2] for v in self.fields.values())
        self.nbytes = self.HEADER + ((self.cap * per_row + 15) // 16) * 16
        self.send = torch.zeros(self.nbytes, dtype=torch.uint8, device=device)
        self.recv = torch.empty(self.world * self.nbytes, dtype=torch.uint8, device=device)
        self.hdr = torch.zeros(1, dtype=torch.int64, device=device)

    def run(self, local, concat=False):
        P = int(local[self.order[0]].shape[0])
        if P > self.cap:
            raise RuntimeError(f"PackedGather: {P} rows exceed the agreed capacity {self.cap}")
        self.hdr.fill_(P)
        self.send[:8].copy_(self.hdr.view(torch.uint8))
        off = self.HEADER
        for k in self.order:
            nb = P * self.fields[k][2]
            if nb:
                self.send[off:off + nb].copy_(local[k].contiguous().view(-1).view(torch.uint8))
            off += nb
        if dist.get_backend(self.group) == "gloo":
            parts = [self.recv[r * self.nbytes:(r + 1) * self.nbytes] for r in range(self.world)]
            dist.all_gather(parts, self.send, group=self.group)
        else:
            dist.all_gather_into_tensor(self.recv, self.send, group=self.group)
        if self.rank != self.dst:
            return None, None
        bufs = self.recv.view(self.world, self.nbytes)
        counts = [int(c) for c in bufs[:, :8].contiguous().view(torch.int64).view(-1).tolist()]
        out = {k: [] for k in self.fields}
        for r, c in enumerate(counts):
            off = self.HEADER
            for k in self.order:
                shape, dt, rb = self.fields[k]
                out[k].append(bufs[r, off:off + c * rb].view(dt).view((c,) + shape))
                off += c * rb
        if concat:
            out = {k: torch.cat(v, 0) for k, v in out.items()}
        return out, counts

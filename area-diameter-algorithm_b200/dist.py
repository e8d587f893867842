"""Multi-GPU plumbing: one process per GPU, shards of anchors / candidate pairs, one tiny all-gather at the end.

The data path has no collective: every rank solves its own shard (adk_*_partial).  The single exchange is the
32-byte (count, area, tie key, witness) record per rank, gathered with torch.distributed (NCCL over NVLink on GPUs,
gloo in the CPU tests) and merged lexicographically by adk_merge on every rank.
"""
import struct

import torch
import torch.distributed as dist

_FMT = "<iIdQII"     # adk_partial: found, count, area, tie_key, w0, w1  (32 bytes)
RECORD_BYTES = struct.calcsize(_FMT)


def pack_partial(p):
    return struct.pack(_FMT, p["found"], p["count"], p["area"], p["tie_key"], p["w0"], p["w1"])


def unpack_partial(b):
    f, c, a, k, w0, w1 = struct.unpack(_FMT, bytes(b))
    return {"found": f, "count": c, "area": a, "tie_key": k, "w0": w0, "w1": w1}


def merge_partials(parts):
    """Host-side lexicographic best: count desc, area asc, tie_key asc (same rule as adk_merge)."""
    best = None
    for p in parts:
        if not p["found"]:
            continue
        if best is None or p["count"] > best["count"] or (p["count"] == best["count"] and (
                p["area"] < best["area"] or (p["area"] == best["area"] and p["tie_key"] < best["tie_key"]))):
            best = p
    return best if best is not None else {"found": 0, "count": 0, "area": 0.0, "tie_key": 0, "w0": 0, "w1": 0}


def owner_of(best, parts):
    """The rank whose partial IS the merged best (a candidate pair / anchor belongs to exactly one shard, and tie_key names
    it): the winner's records are still resident on that rank, so it reconstructs without re-running anything."""
    if not best["found"]:
        return 0
    for r, p in enumerate(parts):
        if p["found"] and p["count"] == best["count"] and p["tie_key"] == best["tie_key"] and p["area"] == best["area"]:
            return r
    return 0


def share_result(result, owner, n, device=None, group=None):
    """The witness travels from the rank that reconstructed it to every rank: one broadcast of (n + 8) 8-byte words
    (found, count, area, diameter pair, hull).  `result` is the owner's result dict (ignored elsewhere)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return result
    dev = device if device is not None else torch.device("cpu")
    rank = dist.get_rank(group)
    if rank == owner:
        hull = [int(v) for v in result["hull"]]
        d = result["diam"]
        words = [int(result["found"]), int(result["count"]), struct.unpack("<q", struct.pack("<d", float(result["area"])))[0],
                 len(hull), 1 if d else 0, int(d[0]) if d else 0, int(d[1]) if d else 0, 0]
        t = torch.tensor(words + hull + [0] * (n - len(hull)), dtype=torch.int64).to(dev)
    else:
        t = torch.empty(n + 8, dtype=torch.int64, device=dev)
    dist.broadcast(t, src=dist.get_global_rank(group, owner) if group is not None else owner, group=group)
    if rank == owner:
        return result
    w = t.cpu().tolist()
    return {"found": w[0], "count": w[1], "area": struct.unpack("<d", struct.pack("<q", w[2]))[0], "hull": w[8:8 + w[3]],
            "diam": [w[5], w[6]] if w[4] else None}


def gather_partials(partial, device=None, group=None):
    """All-gather of one 32-byte record per rank; returns the list of every rank's partial."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return [partial]
    world = dist.get_world_size(group)
    dev = device if device is not None else torch.device("cpu")
    mine = torch.frombuffer(bytearray(pack_partial(partial)), dtype=torch.uint8).to(dev)
    out = torch.empty(world * RECORD_BYTES, dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(out, mine, group=group)
    raw = out.cpu().numpy().tobytes()
    return [unpack_partial(raw[r * RECORD_BYTES:(r + 1) * RECORD_BYTES]) for r in range(world)]


def gather_stats_sum(values, device=None, group=None):
    """Sum of a few per-rank counters (cells, states) and max of times, for reporting only."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return list(values)
    dev = device if device is not None else torch.device("cpu")
    t = torch.tensor(list(values), dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t.cpu().tolist()


def gather_batch_results(results, group=None):
    """Instance-level sharding (adk_solve_batch with shard = rank): every rank holds the results of the instances it owns
    (None elsewhere); returns the complete list on every rank.  A host gather of small records: no data-path collective."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return list(results)
    world = dist.get_world_size(group)
    parts = [None] * world
    dist.all_gather_object(parts, list(results), group=group)
    out = [None] * len(results)
    for part in parts:
        for i, r in enumerate(part):
            if r is not None:
                out[i] = r
    return out


def solve_batch_pair_sharded(solver, point_sets, kmax, amax, dmax, reconstruct=True, device=None, group=None):
    """A batch of area+diameter instances with EVERY instance's candidate pairs sharded over the ranks (instead of whole
    instances): the right split when one instance dominates the batch (the densest cell sample is 60 % of config 5).
    Per instance: adk_area_diameter_partial on every rank, one 32-byte all-gather, and the rank that owns the winning
    pair (owner_of: its records are still resident) reconstructs.  Returns (complete result list, my cell count)."""
    world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
    rank = dist.get_rank(group) if world > 1 else 0
    results, my_cells, my_dp_ms = [None] * len(point_sets), 0, 0.0
    for i, pts in enumerate(point_sets):
        x = [float(p[0]) for p in pts]
        y = [float(p[1]) for p in pts]
        part, st = solver.area_diameter_partial(x, y, kmax, amax, dmax, rank, world)
        my_cells += st["cells_executed"]
        my_dp_ms += st["dp_ms"]
        parts = gather_partials(part, device=device, group=group)
        best = merge_partials(parts)
        owner = owner_of(best, parts)
        if rank == owner:
            results[i] = solver.area_diameter_finish(x, y, kmax, amax, dmax, best, reconstruct)
    return gather_batch_results(results, group=group), my_cells, my_dp_ms

"""Sharded visibility pass (one process per GPU, torch.distributed).

The reference is single process; the north star shards
only the visibility pass (shen_partation_densify.py:932-936 is the loop being
split).  Default sharding: the (block, camera) pair list, in CSR order, is cut
into `world` CONTIGUOUS ranges of about equal cost (RangePlan / range_bounds);
rank r sweeps its range and the per-pair records are exchanged with ONE padded
all-gather of every rank's contiguous slice (exchange_ranges: NCCL over NVLink
on the GPU box, gloo in the CPU tests).  The older modulo sharding
(c % world == rank, shard_csr + gather_records) is kept for comparison.
Everything here is device agnostic torch; the compute callback is the CUDA
visibility pass and the pack / unpack kernels come from device.py.
"""
from __future__ import annotations

import numpy as np
import torch

# columns of the packed per-pair record (float64: every field is exact)
REC_N_HIT, REC_RATIO, REC_AREA, REC_KEEP_FLAGS = 0, 1, 2, 3
REC_COLS = 4


def shard_mask(cam_idx: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """Pairs (positions of the block->camera CSR) owned by `rank`."""
    return (cam_idx % world) == rank


def shard_csr(cam_off: torch.Tensor, cam_idx: torch.Tensor, rank: int, world: int):
    """CSR restricted to the pairs of `rank` (order kept) + their positions in
    the full pair list."""
    if world <= 1:
        pos = torch.arange(cam_idx.shape[0], device=cam_idx.device)
        return cam_off, cam_idx, pos
    L = cam_off.shape[0] - 1
    mine = shard_mask(cam_idx, rank, world)
    pos = torch.nonzero(mine).reshape(-1)
    # per block count of owned pairs: searchsorted of the block boundaries in pos
    new_off = torch.searchsorted(pos, cam_off[: L + 1].to(pos.dtype)).to(torch.int64)
    return new_off.contiguous(), cam_idx[pos].contiguous(), pos


def shard_csr_ranges(cam_off: torch.Tensor, cam_idx: torch.Tensor, rank: int, world: int,
                     block_cost: torch.Tensor):
    """Like shard_csr, but rank r owns a CONTIGUOUS range of the pair list whose cost
    (pair cost = the block's down-sampled point count, `block_cost[L]`) is 1/world of the
    total: whole blocks plus at most two partial camera lists per rank, so the per-tile work
    of the sweep (tile load, box, cull) is not repeated on every rank as it is with the
    modulo sharding.  Every rank evaluates the same expressions on the same data, so the
    ranges tile the pair list exactly."""
    P = cam_idx.shape[0]
    if world <= 1:
        return cam_off, cam_idx, torch.arange(P, device=cam_idx.device)
    lo, hi = range_bounds(cam_off.cpu().numpy(), block_cost.cpu().numpy(), rank, world)
    # the range is contiguous: the shard is a view of the pair list and a clamped offset table
    new_off = (cam_off.clamp(lo, hi) - lo).to(torch.int64)
    pos = torch.arange(lo, hi, device=cam_idx.device)
    return new_off.contiguous(), cam_idx[lo:hi], pos


def range_bounds(cam_off_h, block_cost_h, rank: int, world: int):
    """[lo, hi) of rank's pair range (host arithmetic on the L-sized tables, exact in integers):
    pair p of block b has exclusive cumulative cost  S_b + (p - cam_off[b]) * M_b;  rank r owns the
    pairs with  total * r <= cost * world < total * (r + 1)."""
    L = len(cam_off_h) - 1
    cnt = [int(cam_off_h[b + 1] - cam_off_h[b]) for b in range(L)]
    M = [int(block_cost_h[b]) for b in range(L)]
    total = sum(c * m for c, m in zip(cnt, M))

    def first_at_least(t):          # first pair whose exclusive cost * world >= t  (t = total * r)
        s = 0
        for b in range(L):
            if cnt[b] == 0:
                continue
            if M[b] == 0:
                if s * world >= t:
                    return int(cam_off_h[b])
                continue
            if (s + (cnt[b] - 1) * M[b]) * world >= t:
                need = t - s * world                        # k * M[b] * world >= need
                k = max(0, -(-need // (M[b] * world)))
                return int(cam_off_h[b]) + k
            s += cnt[b] * M[b]
        return int(cam_off_h[L])
    lo = first_at_least(total * rank)
    hi = first_at_least(total * (rank + 1)) if rank + 1 < world else int(cam_off_h[L])
    return lo, hi


def all_range_bounds(cam_off_h, block_cost_h, world: int) -> np.ndarray:
    """bounds[world + 1] with rank r owning pairs [bounds[r], bounds[r + 1]) -- the same ranges as
    range_bounds(), computed for every rank at once with exact int64 arithmetic on the L-sized host
    tables (vectorised: pair k of block b has exclusive cost E_b + k * M_b, and the first pair at or
    above a threshold lies in the first non-empty block whose LAST pair reaches it)."""
    co = np.asarray(cam_off_h, dtype=np.int64)
    L = len(co) - 1
    cnt = np.diff(co)
    M = np.asarray(block_cost_h, dtype=np.int64)[:L]
    P = int(co[L])
    world = max(int(world), 1)
    out = np.empty(world + 1, np.int64)
    out[0], out[world] = 0, P
    if world == 1:
        return out
    w = cnt * M
    E = np.cumsum(w) - w                                   # exclusive cost prefix per block
    total = int(w.sum())
    nz = np.nonzero(cnt > 0)[0]
    key = (E[nz] + (cnt[nz] - 1) * M[nz]) * world          # non-decreasing over the non-empty blocks
    for r in range(1, world):
        t = total * r
        j = int(np.searchsorted(key, t, side="left"))
        if j >= len(nz):
            out[r] = P
            continue
        b = int(nz[j])
        k = 0
        if M[b] > 0:
            need = t - int(E[b]) * world
            k = max(0, -(-need // (int(M[b]) * world)))
        out[r] = int(co[b]) + k
    return out


class RangePlan:
    """Static part of the range-sharded visibility pass for one block -> camera CSR.  The pair list is
    cut into world * chunks contiguous equal-cost ranges; rank r owns ranges r, r + world, ... (chunks = 1:
    one contiguous range per rank -- whole blocks plus two partial camera lists, what the entry points
    that move block DATA per rank use; chunks > 1 interleaves the ranks over the scene, which evens
    out what the cost model cannot know in advance: which pairs stay undecided and need sweep 2 +
    the exact hull).  Holds every range's bounds, this rank's offset table / camera list and the
    position of every range inside the gathered slots.  Built once per partition result (L-sized
    host arithmetic); the per-step work is then the sweep itself and one exchange."""

    def __init__(self, cam_off: torch.Tensor, cam_idx: torch.Tensor, cam_off_h, block_cost_h, rank: int,
                 world: int, chunks: int = 1):
        self.rank, self.world, self.chunks = int(rank), int(world), max(1, int(chunks))
        nr = self.world * self.chunks
        self.bounds = all_range_bounds(cam_off_h, block_cost_h, nr)           # range c = [bounds[c], bounds[c+1])
        self.P = int(self.bounds[-1])
        length = np.diff(self.bounds)
        owner = np.arange(nr) % self.world
        per_rank = np.array([int(length[owner == r].sum()) for r in range(self.world)], dtype=np.int64)
        self.slot = int(per_rank.max()) if self.world > 0 else self.P
        # record position of range c inside the gathered (world * slot) table
        self.base = np.zeros(nr, dtype=np.int64)
        fill = np.zeros(self.world, dtype=np.int64)
        for c in range(nr):
            r = int(owner[c])
            self.base[c] = r * max(self.slot, 1) + fill[r]
            fill[r] += length[c]
        self.mine = [c for c in range(nr) if owner[c] == self.rank]
        self.n_local = int(per_rank[self.rank])
        self.lo, self.hi = int(self.bounds[self.mine[0]]), int(self.bounds[self.mine[-1] + 1])   # chunks = 1: the range
        # a host-only plan (cam_off None) carries the bounds and the block span, no device tables
        if cam_off is not None:
            off = None
            for c in self.mine:
                lo, hi = int(self.bounds[c]), int(self.bounds[c + 1])
                t = cam_off.clamp(lo, hi) - lo
                off = t if off is None else off + t
            self.my_off = off.to(torch.int64).contiguous()
            parts = [cam_idx[int(self.bounds[c]):int(self.bounds[c + 1])] for c in self.mine]
            self.my_idx = parts[0] if len(parts) == 1 else torch.cat(parts)
        else:
            self.my_off = self.my_idx = None
        co = np.asarray(cam_off_h, dtype=np.int64)
        self._co = co
        if self.hi > self.lo:     # blocks the rank's pairs lie in: [bl0, bl1) (chunks = 1: exactly the touched ones)
            self.bl0 = int(np.searchsorted(co, self.lo, side="right")) - 1
            self.bl1 = int(np.searchsorted(co, self.hi, side="left"))
        else:
            self.bl0 = self.bl1 = 0

    def local_counts(self) -> np.ndarray:
        """Per block: how many of its pairs this rank owns."""
        cnt = np.zeros(len(self._co) - 1, dtype=np.int64)
        for c in self.mine:
            lo, hi = int(self.bounds[c]), int(self.bounds[c + 1])
            cl = np.clip(self._co, lo, hi)
            cnt += cl[1:] - cl[:-1]
        return cnt


def exchange_ranges(local, plan: "RangePlan", pack, unpack, group=None):
    """All-gather of the per-pair result vectors of a range-sharded pass.  `local` = this rank's
    (n_hit, ratio, area, keep, flags) over its ranges (in pair order); pack(local, slot) -> a
    contiguous (slot, k) send tensor whose first n_local rows are the rank's records;
    unpack(recv, plan) -> the full-length vectors from the (world * slot, k) gathered tensor
    (slot r holds rank r's ranges back to back; plan.base[c] = record position of range c).
    One collective, no zero fill, no reduction."""
    import torch.distributed as dist
    slot = max(plan.slot, 1)
    send = pack(local, slot)
    recv = torch.empty((plan.world * slot,) + tuple(send.shape[1:]), dtype=send.dtype, device=send.device)
    dist.all_gather_into_tensor(recv, send, group=group)
    return unpack(recv, plan)


def pack_slot_torch(local, slot: int) -> torch.Tensor:
    """Device-agnostic pack for exchange_ranges (CPU tests; the CUDA path uses ptgs_pack_records)."""
    rec = pack_records(*local)
    send = torch.zeros((slot, REC_COLS), dtype=torch.float64, device=rec.device)
    send[: rec.shape[0]] = rec
    return send


def unpack_slot_torch(recv: torch.Tensor, plan: "RangePlan"):
    parts = [recv[int(plan.base[c]): int(plan.base[c]) + int(plan.bounds[c + 1] - plan.bounds[c])]
             for c in range(len(plan.base))]
    return unpack_records(torch.cat(parts) if parts else recv[:0])


def pack_records(n_hit, ratio, area, keep, flags) -> torch.Tensor:
    rec = torch.empty((n_hit.shape[0], REC_COLS), dtype=torch.float64, device=n_hit.device)
    rec[:, REC_N_HIT] = n_hit.to(torch.float64)
    rec[:, REC_RATIO] = ratio
    rec[:, REC_AREA] = area
    rec[:, REC_KEEP_FLAGS] = keep.to(torch.float64) + 256.0 * flags.to(torch.float64)
    return rec


def unpack_records(rec: torch.Tensor):
    n_hit = rec[:, REC_N_HIT].to(torch.int32)
    kf = rec[:, REC_KEEP_FLAGS].to(torch.int64)
    return n_hit, rec[:, REC_RATIO].contiguous(), rec[:, REC_AREA].contiguous(), \
        (kf & 1).to(torch.uint8), (kf >> 8).to(torch.uint8)


def gather_records(local: torch.Tensor, cam_idx: torch.Tensor, world: int, group=None,
                   pos: torch.Tensor | None = None, rank: int | None = None) -> torch.Tensor:
    """All-gather of the per-pair records.  `local` holds this rank's pairs in
    shard order, `pos` their positions in the full pair list (shard_csr);
    returns the (P, REC_COLS) table in full pair order on every rank.

    Every rank sends a P-row table with its records at their global positions
    and zeros elsewhere; the gathered tables are summed (each pair has exactly
    one owner, x + 0.0 is exact).  P rows x 32 B x world is tiny (c4: ~13 MB over
    NVLink), and no count exchange / host synchronisation is needed."""
    import torch.distributed as dist
    P = cam_idx.shape[0]
    if world <= 1:
        return local
    if pos is None:
        if rank is None:
            rank = dist.get_rank(group)
        pos = torch.nonzero(shard_mask(cam_idx, rank, world)).reshape(-1)
    send = torch.zeros((max(P, 1), REC_COLS), dtype=torch.float64, device=local.device)
    send[pos] = local
    recv = torch.empty((world * max(P, 1), REC_COLS), dtype=torch.float64, device=local.device)
    dist.all_gather_into_tensor(recv, send, group=group)
    return recv.view(world, max(P, 1), REC_COLS).sum(dim=0)[:P]


# ---- coverage sweep (config c5): points replicated, cameras sharded in contiguous ranges ----
def camera_range(n_cameras: int, rank: int, world: int):
    """(first, count) of rank's cameras: ceil(C / world) consecutive cameras per rank."""
    per = (n_cameras + world - 1) // max(world, 1)
    first = min(rank * per, n_cameras)
    return first, max(0, min(per, n_cameras - first))


def gather_camera_counts(local: torch.Tensor, n_cameras: int, world: int, group=None) -> torch.Tensor:
    """All-gather of the per-camera hit counts of a camera-sharded coverage sweep (`local` = this
    rank's camera_range, int32) -> counts of all cameras on every rank.  64 KB at c5."""
    import torch.distributed as dist
    if world <= 1:
        return local
    per = (n_cameras + world - 1) // world
    send = torch.zeros(per, dtype=local.dtype, device=local.device)
    send[: local.shape[0]] = local
    recv = torch.empty(world * per, dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(recv, send, group=group)
    return recv[:n_cameras]


def merge_seen(seen: torch.Tensor, world: int, group=None) -> torch.Tensor:
    """Per-point "seen by any camera" flags of all ranks (uint8 max all-reduce, in place)."""
    import torch.distributed as dist
    if world > 1:
        dist.all_reduce(seen, op=dist.ReduceOp.MAX, group=group)
    return seen


def point_range(n_points: int, rank: int, world: int):
    """[p0, p1) of rank's slice of the cloud: equal slices, starts on multiples of 4 points so that
    the planar float32 slices stay 16-byte aligned."""
    per = ((n_points + max(world, 1) - 1) // max(world, 1) + 3) // 4 * 4
    p0 = min(rank * per, n_points)
    return p0, min(p0 + per, n_points)


def sum_camera_counts(n_hit: torch.Tensor, world: int, group=None) -> torch.Tensor:
    """Point-sharded coverage sweep: every rank counted its slice of the cloud for ALL cameras;
    the per-camera totals are one int32 sum all-reduce (64 KB at c5)."""
    import torch.distributed as dist
    if world > 1:
        dist.all_reduce(n_hit, op=dist.ReduceOp.SUM, group=group)
    return n_hit


def gather_seen(seen_local: torch.Tensor, n_points: int, world: int, group=None) -> torch.Tensor:
    """Per-point flags of the point-sharded sweep: slices (point_range) all-gathered in rank order."""
    import torch.distributed as dist
    if world <= 1:
        return seen_local
    per = ((n_points + world - 1) // world + 3) // 4 * 4
    send = torch.zeros(per, dtype=seen_local.dtype, device=seen_local.device)
    send[: seen_local.shape[0]] = seen_local
    recv = torch.empty(world * per, dtype=seen_local.dtype, device=seen_local.device)
    dist.all_gather_into_tensor(recv, send, group=group)
    return recv[:n_points]

"""Host side of the multi-GPU layer: contiguous row shards, the top-k exchange and the merge.

One process per GPU.  Rows are split into contiguous ranges (rank g holds
[g*ceil(N/G), (g+1)*ceil(N/G)) so that `shard base + local row` keeps the row-index tie-break
global); queries are replicated; every rank scans its shard and the sorted per-rank lists
(score f32, local row u32) are exchanged with ONE all-gather each (NCCL over NVLink on GPUs,
gloo in the CPU tests) and merged by (score, global row).  This is the device analogue of
CombinedVectorColumnIndexer::Search's concat + std::sort + truncate
(src/db/index/column/vector_column/combined_vector_column_indexer.cc:167-222).
torch.distributed is plumbing only; the scan and the merge on GPUs are libzvec_b200 kernels.
"""
import numpy as np

INVALID_INDEX = 0xFFFFFFFF


def shard_bounds(rows, world):
    """[(lo, hi)] per rank; later ranks may be empty when rows < world."""
    per = (rows + world - 1) // world if world > 0 else rows
    return [(min(rows, r * per), min(rows, (r + 1) * per)) for r in range(world)]


def shard_bases(rows, world):
    return [lo for lo, _ in shard_bounds(rows, world)]


def gather_topk(local_scores, local_index, world, group=None):
    """all-gather [q,k] score / index tensors -> [world,q,k] each (same device, same dtype)."""
    import torch
    import torch.distributed as dist
    q, k = local_scores.shape
    all_s = torch.empty((world, q, k), dtype=local_scores.dtype, device=local_scores.device)
    all_i = torch.empty((world, q, k), dtype=local_index.dtype, device=local_index.device)
    if world == 1:
        all_s[0], all_i[0] = local_scores, local_index
        return all_s, all_i
    # output viewed as [world*q, k]: the concatenation form both NCCL and gloo accept
    dist.all_gather_into_tensor(all_s.view(world * q, k), local_scores.contiguous(), group=group)
    dist.all_gather_into_tensor(all_i.view(world * q, k), local_index.contiguous(), group=group)
    return all_s, all_i


def merge_topk_host(all_scores, all_index, bases, k):
    """numpy statement of merge_shards_kernel (checker for it, and the merge of the gloo tests):
    [g,q,k] sorted lists with shard-local rows -> ([q,k] scores, [q,k] global rows u64), ordered by
    (score, global row); unused slots hold NaN / 2^64-1."""
    s = np.asarray(all_scores, np.float32)
    ix = np.asarray(all_index).astype(np.int64) & 0xFFFFFFFF
    g, q, kk = s.shape
    out_s = np.full((q, k), np.nan, np.float32)
    out_r = np.full((q, k), np.uint64(0xFFFFFFFFFFFFFFFF), np.uint64)
    base = np.asarray(bases, np.uint64).reshape(g, 1)
    for i in range(q):
        sc = s[:, i, :]
        li = ix[:, i, :]
        valid = (li != INVALID_INDEX) & ~np.isnan(sc)
        rows = (li.astype(np.uint64) + base)[valid]
        scv = sc[valid] + np.float32(0.0)              # -0 -> +0 like order_f32
        order = np.lexsort((rows, scv))[:k]
        out_s[i, :order.size] = scv[order]
        out_r[i, :order.size] = rows[order]
    return out_s, out_r


# ---- IVF: every rank holds a slice of EVERY posting list ---------------------------------------
def ivf_shard_plan(list_offsets, rank, world):
    """list_offsets u64[nlist+1] of the whole index (rows grouped by list) ->
    (lo, hi, local_offsets): rank `rank` holds rows [lo[l], hi[l]) of list l (global row numbers), stored
    list after list at local_offsets[l] .. local_offsets[l+1]."""
    off = np.asarray(list_offsets, np.uint64).astype(np.int64)
    sizes = off[1:] - off[:-1]
    lo = off[:-1] + sizes * rank // world
    hi = off[:-1] + sizes * (rank + 1) // world
    local = np.zeros(len(off), np.int64)
    local[1:] = np.cumsum(hi - lo)
    return lo, hi, local


def ivf_shard_rows(lo, hi):
    """global row numbers of the rank's rows, in local order"""
    if len(lo) == 0:
        return np.zeros(0, np.int64)
    return np.concatenate([np.arange(a, b, dtype=np.int64) for a, b in zip(lo, hi)]) if (hi > lo).any() \
        else np.zeros(0, np.int64)


def ivf_local_to_global(local_index, lo, local_offsets):
    """local row indices (any shape, INVALID_INDEX kept) -> global row numbers u64"""
    li = np.asarray(local_index).astype(np.int64) & 0xFFFFFFFF
    valid = li != INVALID_INDEX
    lst = np.searchsorted(local_offsets, np.where(valid, li, 0), side="right") - 1
    lst = np.clip(lst, 0, len(lo) - 1)
    g = lo[lst] + (np.where(valid, li, 0) - local_offsets[lst])
    return np.where(valid, g.astype(np.uint64), np.uint64(0xFFFFFFFFFFFFFFFF))


def merge_topk_global(all_scores, all_rows, k):
    """[g,q,k] per-rank lists with GLOBAL row numbers (u64, 2^64-1 = unused) -> [q,k] by (score, global row)"""
    s = np.asarray(all_scores, np.float32)
    r = np.asarray(all_rows, np.uint64)
    g, q, _ = s.shape
    out_s = np.full((q, k), np.nan, np.float32)
    out_r = np.full((q, k), np.uint64(0xFFFFFFFFFFFFFFFF), np.uint64)
    for i in range(q):
        sc, rw = s[:, i, :].ravel(), r[:, i, :].ravel()
        valid = (rw != np.uint64(0xFFFFFFFFFFFFFFFF)) & ~np.isnan(sc)
        scv, rwv = sc[valid] + np.float32(0.0), rw[valid]
        order = np.lexsort((rwv, scv))[:k]
        out_s[i, :order.size] = scv[order]
        out_r[i, :order.size] = rwv[order]
    return out_s, out_r

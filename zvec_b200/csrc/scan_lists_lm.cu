// K4b — IVF posting-list scan, LIST-major (query batches).
//
// Replaces the same loop as K4 (IVFSearcher::search_impl's per-list walk, src/core/algorithm/ivf/ivf_searcher.cc:
// 216-231 -> IVFEntity::search(list, ...), ivf_entity.cc:662-716), for a BATCH of queries: K4 lets every query walk
// its own lists, so a list probed by m queries of the batch is streamed from HBM m times (C4, batch 1024: 246 GB per
// batch out of a 61 GB index).  Here the (query, list) pairs of the batch are grouped by list; one CTA streams a
// list ONCE for up to QB of the queries that probe it (their vectors in shared memory, kU rows in flight per warp,
// the 1x1 kernels' arithmetic exactly as K1 / K4: same chunk order, same shuffle tree — scores are bit-identical to
// the query-major scan), keeps a warp top-k per query and writes one partial list per (query, probe slot).  The
// lists each query visits and the scan budget are decided by ivf_plan_kernel exactly as for K4.
//   lm_count   pairs per list
//   lm_scan    one CTA: pair offsets and work-item offsets per list (an item = one list x up to QB of its queries)
//   lm_fill    pairs grouped by list {query, slot}; items {list, first pair, pairs}
//   scan_lists_lm_kernel   persistent CTAs pull items from an atomic counter
#include <stdlib.h>

#include "chunk.cuh"
#include "internal.h"
#include "topk.cuh"

namespace zvb {

namespace {

// 16 warps of 4 row groups each: the same 64 rows in flight per CTA as 8 warps x 8, but twice the warps to hide the
// LDG / LDS / FFMA latencies behind (8 warps: issue slots 28 % busy, 16.1 ms per C4 batch)
constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr int kU = 4;

struct LmArgs {
  RowsView rv;
  QueryView qv;
  uint32_t k;
  float threshold;
  const uint32_t *filter;
  const uint32_t *list_off;   // [nlist + 1] first row of each list
  const uint32_t *pair_q, *pair_slot;
  const uint32_t *item_list, *item_first, *item_n;
  const uint32_t *n_items;    // device scalar
  uint32_t *counter;          // device scalar, zeroed before the launch
  u64 *gtau;                  // [q] smallest k-th best key any finished (query, list) item has published
  u64 *part;                  // [q][nslots][k]
  uint32_t nslots;
  int lpr_log2;
  uint32_t pf_dist;           // L2 prefetch distance, in iterations of a warp's chunk loop (0 = none)
};

// One (warp, query) top-k list lives entirely in shared memory: {cnt | qn << 32, flip} then list / alt / queue of
// WarpTopk.  The hot loop keeps only the bound (tau) in a register and tests one ballot per (row, query); everything
// behind an accepted candidate is OUT OF LINE.  (Inlined, 64 copies of WarpTopk::push and its merge made the QB = 8
// kernel 60 000 instructions = 970 KB of code: the scan stalled on instruction fetch, 1.6 TB/s.)
__device__ __forceinline__ WarpTopk lm_load(u64 *blk, int k, u64 tau) {
  WarpTopk t;
  const u64 h = blk[0];
  const bool flip = blk[1] != 0;
  t.list = blk + 2 + (flip ? k : 0);
  t.alt = blk + 2 + (flip ? 0 : k);
  t.queue = blk + 2 + 2 * k;
  t.k = k; t.cnt = (int)(uint32_t)h; t.qn = (int)(h >> 32); t.tau = tau;
  return t;
}
__device__ __forceinline__ void lm_store(u64 *blk, int k, const WarpTopk &t) {
  if (lane_id() == 0) {
    blk[0] = (u64)(uint32_t)t.cnt | ((u64)(uint32_t)t.qn << 32);
    blk[1] = t.list == blk + 2 ? 0ull : 1ull;
  }
  __syncwarp();
}
__device__ __noinline__ u64 lm_push(u64 *blk, int k, bool has, u64 key, u64 tau) {
  WarpTopk t = lm_load(blk, k, tau);
  t.push(has, key, kMaxKey);
  lm_store(blk, k, t);
  return t.tau;
}
__device__ __noinline__ u64 lm_finish(u64 *blk, int k, u64 tau) {
  WarpTopk t = lm_load(blk, k, tau);
  t.finish(kMaxKey);
  lm_store(blk, k, t);
  return t.tau;
}

__global__ void lm_count_kernel(const uint32_t *seg_list, const uint32_t *seg_count, uint32_t q, uint32_t cap,
                                uint32_t *cnt) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= q * cap) return;
  const uint32_t qi = i / cap, s = i % cap;
  if (s < seg_count[qi]) atomicAdd(cnt + seg_list[(size_t)qi * cap + s], 1u);
}

// exclusive scans over the lists of {pairs, ceil(pairs / QB)}; one CTA of 1024 threads
__global__ void lm_scan_kernel(const uint32_t *cnt, uint32_t nlist, uint32_t qb, uint32_t *pair_off,
                               uint32_t *item_off, uint32_t *n_items) {
  __shared__ uint2 wsum[32];
  __shared__ uint2 carry;
  if (threadIdx.x == 0) carry = make_uint2(0, 0);
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t base = 0; base < nlist; base += blockDim.x) {
    const uint32_t l = base + threadIdx.x;
    const uint32_t c = l < nlist ? cnt[l] : 0u;
    const uint2 v = make_uint2(c, (c + qb - 1) / qb);
    uint2 x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t a = __shfl_up_sync(0xffffffffu, x.x, o), b = __shfl_up_sync(0xffffffffu, x.y, o);
      if (lane >= o) { x.x += a; x.y += b; }
    }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    if (warp == 0) {
      uint2 w = wsum[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, w.x, o), b = __shfl_up_sync(0xffffffffu, w.y, o);
        if (lane >= o) { w.x += a; w.y += b; }
      }
      wsum[lane] = w;
    }
    __syncthreads();
    const uint2 before = warp ? wsum[warp - 1] : make_uint2(0, 0);
    const uint2 cy = carry;
    if (l < nlist) {
      pair_off[l] = cy.x + before.x + x.x - v.x;
      item_off[l] = cy.y + before.y + x.y - v.y;
    }
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = make_uint2(cy.x + before.x + x.x, cy.y + before.y + x.y);
    __syncthreads();
  }
  if (threadIdx.x == 0) *n_items = carry.y;
}

__global__ void lm_fill_pairs_kernel(const uint32_t *seg_list, const uint32_t *seg_count, uint32_t q, uint32_t cap,
                                     const uint32_t *pair_off, uint32_t *fill, uint32_t *pair_q, uint32_t *pair_slot) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= q * cap) return;
  const uint32_t qi = i / cap, s = i % cap;
  if (s >= seg_count[qi]) return;
  const uint32_t l = seg_list[(size_t)qi * cap + s];
  const uint32_t pos = pair_off[l] + atomicAdd(fill + l, 1u);
  pair_q[pos] = qi;
  pair_slot[pos] = s;
}

// Items that hold some query's NEAREST list (probe slot 0) run first: they publish that query's tightest bound, and
// every later item of the query starts from it.  Front items are numbered from 0 up, the others from the top down.
__global__ void lm_fill_items_kernel(const uint32_t *cnt, const uint32_t *pair_off, const uint32_t *item_off,
                                     const uint32_t *pair_slot, const uint32_t *n_items, uint32_t nlist, uint32_t qb,
                                     uint32_t *front, uint32_t *item_list, uint32_t *item_first, uint32_t *item_n) {
  const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= nlist) return;
  const uint32_t c = cnt[l], total = *n_items;
  uint32_t it = item_off[l];
  for (uint32_t b = 0; b < c; b += qb, ++it) {
    const uint32_t n = min(qb, c - b);
    bool first = false;
    for (uint32_t i = 0; i < n; ++i) first |= pair_slot[pair_off[l] + b + i] == 0;
    // a slot of its own in the run order: from the front, or from the back (both counters in front[0..1])
    const uint32_t pos = first ? atomicAdd(front, 1u) : total - 1u - atomicAdd(front + 1, 1u);
    item_list[pos] = l;
    item_first[pos] = pair_off[l] + b;
    item_n[pos] = n;
  }
}

struct LmItem {
  uint64_t vbegin, vend;
  const uint8_t *qs;
  size_t qbytes, tk_u64;
  const int *qq;
  u64 *tkwarp;   // this warp's QB top-k blocks
  int warp, lane;
};

// All tiles of one item for NQ live queries: kU rows in flight per warp, NQ accumulators per row.
template <int DT, bool L2, int QB, int NQ>
__device__ __forceinline__ void lm_scan_item(const LmArgs &a, const LmItem &it, u64 (&tau)[QB], const float4 (&qt)[QB]) {
  typedef Chunk<DT, L2> CK;
  typedef typename CK::Acc Acc;
  const uint32_t C = a.rv.chunks;
  const int k = a.k, lane = it.lane, warp = it.warp;
  const int lpr = 1 << a.lpr_log2, R = 32 >> a.lpr_log2;
  const int l = lane & (lpr - 1), sub = lane >> a.lpr_log2;
  const uint64_t rows_per_warp_step = (uint64_t)kU * R;
  const uint64_t rows_per_tile = rows_per_warp_step * kWarps;
  const int metric = a.rv.metric;
  const float dq = (float)a.rv.d;
  const uint64_t vbegin = it.vbegin, vend = it.vend;
  const uint8_t *qs = it.qs;
  const size_t qbytes = it.qbytes;
  // Once a row's sums are reduced, lane j of the row's lane group scores query j: NQ (row, query) pairs are handled
  // by NQ lanes at once instead of one pair at a time by lane 0 (rows narrower than NQ lanes: lane 0, pair by pair).
  const bool spread = lpr >= NQ;
  const uint64_t ntiles = (vend - vbegin + rows_per_tile - 1) / rows_per_tile;
  for (uint64_t t = 0; t < ntiles; ++t) {
    const uint64_t base = vbegin + t * rows_per_tile + (uint64_t)warp * rows_per_warp_step;
    if (base >= vend) continue;  // warp-uniform
    Acc acc[kU][NQ], aa[kU];
    const uint8_t *rp[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      uint64_t r = base + (uint64_t)u * R + sub;
      if (r >= vend) r = vend - 1;  // clamp: loaded, result discarded
      rp[u] = a.rv.rows + r * a.rv.row_stride;
      if (metric_is_qi8(metric) && l == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.rv.tails + r));
      aa[u] = 0;
#pragma unroll
      for (int j = 0; j < NQ; ++j) acc[u][j] = 0;
    }
    // Software prefetch into L2, kPfDist iterations ahead of this warp's LDG stream (one 128-byte line per 8 lanes;
    // past the end of these rows: the same warp's rows of the next tile).
    const size_t next_rows = (size_t)rows_per_tile * a.rv.row_stride;
    const bool has_next = base + rows_per_tile < vend && base + rows_per_tile + rows_per_warp_step <= a.rv.n;
    for (uint32_t c = l; c < C; c += lpr) {
      uint4 v[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) v[u] = ldg_stream(rp[u] + (size_t)c * 16);
      if (a.pf_dist && (l & 7) == 0) {
        const uint32_t pc = c + a.pf_dist * lpr;
        if (pc < C) {
#pragma unroll
          for (int u = 0; u < kU; ++u) asm volatile("prefetch.global.L2 [%0];" ::"l"(rp[u] + (size_t)pc * 16));
        } else if (has_next) {
          const uint32_t wc = pc - ((C - l + lpr - 1) / lpr) * lpr;  // the same lane's chunk, wrapped into the next tile
          if (wc < C) {
#pragma unroll
            for (int u = 0; u < kU; ++u)
              asm volatile("prefetch.global.L2 [%0];" ::"l"(rp[u] + next_rows + (size_t)wc * 16));
          }
        }
      }
      const uint8_t *qc = qs;
#pragma unroll
      for (int j = 0; j < NQ; ++j) {
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          Acc dummy = 0;
          CK::fma(acc[u][j], j == 0 ? aa[u] : dummy, v[u], qc, c);
        }
        qc += qbytes;
      }
    }
    for (int o = lpr >> 1; o > 0; o >>= 1) {
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        if (DT == ZVB200_DT_INT8 && L2) aa[u] += __shfl_xor_sync(0xffffffffu, aa[u], o);
#pragma unroll
        for (int j = 0; j < NQ; ++j) acc[u][j] += __shfl_xor_sync(0xffffffffu, acc[u][j], o);
      }
    }
    // epilogue: score, gate, push
    auto score_of = [&](Acc ab, Acc aa_u, int j, const float4 mt) -> float {
      if (DT == ZVB200_DT_INT8) {
        if (metric == ZVB200_METRIC_QI8_L2SQ) return qi8_l2sq((float)ab, mt, qt[j], dq);
        if (metric_is_qi8(metric)) return qi8_mip_ord((float)ab, mt, qt[j], dq, a.rv.query_first);
        if (L2) {
          const float s = (float)((int)aa_u - 2 * (int)ab + it.qq[j]);
          return metric == ZVB200_METRIC_L2 ? sqrtf(s) : s;
        }
        return -(float)ab;
      }
      const float f = (float)ab;
      return metric == ZVB200_METRIC_L2SQ ? f : metric == ZVB200_METRIC_L2 ? sqrtf(f) : metric == ZVB200_METRIC_IP ? -f : 1.0f - f;
    };
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const uint64_t r = base + (uint64_t)u * R + sub;
      bool valid = (spread ? l < NQ : l == 0) && r < vend;
      if (valid && a.rv.drop) valid = !bit_set(a.rv.drop, r);
      if (valid && a.filter) valid = !bit_set(a.filter, r);
      float4 mt = make_float4(0, 0, 0, 0);
      if (metric_is_qi8(metric) && valid) mt = ldg_tail(a.rv.tails + r);  // read-only, no L1 allocation (scan_gemv.cu)
      // the row's score scale (IVFEntity::search: distances[k] * norm_val, ivf_entity.cc:650); 1 for a plain store
      const float sc = a.rv.row_scale ? (valid ? a.rv.row_scale[r] : 1.f) : a.rv.score_scale;
      if (spread) {
        // this lane's query: j = l (a select chain over the NQ accumulators, every lane of the group holds all of them)
        Acc mine = acc[u][0];
        float4 myqt = qt[0];
        u64 mytau = tau[0];
#pragma unroll
        for (int j = 1; j < NQ; ++j) {
          if (l == j) { mine = acc[u][j]; myqt = qt[j]; mytau = tau[j]; }
        }
        float s;
        if (DT == ZVB200_DT_INT8 && metric_is_qi8(metric)) {
          s = metric == ZVB200_METRIC_QI8_L2SQ ? qi8_l2sq((float)mine, mt, myqt, dq)
                                               : qi8_mip_ord((float)mine, mt, myqt, dq, a.rv.query_first);
        } else if (DT == ZVB200_DT_INT8 && L2) {
          int myqq = it.qq[0];
#pragma unroll
          for (int j = 1; j < NQ; ++j) myqq = l == j ? it.qq[j] : myqq;
          s = (float)((int)aa[u] - 2 * (int)mine + myqq);
          if (metric == ZVB200_METRIC_L2) s = sqrtf(s);
        } else {
          s = score_of(mine, aa[u], 0, mt);
        }
        s *= sc;
        const u64 key = pack_key(s, (uint32_t)r);
        const bool has = valid && s <= a.threshold && key < mytau;
        uint32_t any = __ballot_sync(0xffffffffu, has);
        if (any) {
#pragma unroll
          for (int j = 0; j < NQ; ++j) {
            const bool hj = has && l == j;
            if (__any_sync(0xffffffffu, hj)) tau[j] = lm_push(it.tkwarp + (size_t)j * it.tk_u64, k, hj, key, tau[j]);
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < NQ; ++j) {
          const float s = score_of(acc[u][j], aa[u], j, mt) * sc;
          const u64 key = pack_key(s, (uint32_t)r);
          const bool has = valid && s <= a.threshold && key < tau[j];
          if (__any_sync(0xffffffffu, has)) tau[j] = lm_push(it.tkwarp + (size_t)j * it.tk_u64, k, has, key, tau[j]);
        }
      }
    }
  }
}

template <int DT, bool L2, int QB>
__global__ void __launch_bounds__(kThreads, 1) scan_lists_lm_kernel(const LmArgs a) {
  typedef Chunk<DT, L2> CK;
  typedef typename CK::Acc Acc;
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ uint32_t s_item, s_next;
  __shared__ uint32_t s_q[QB], s_slot[QB];
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  const uint32_t C = a.rv.chunks;
  const size_t qbytes = CK::qbytes(C);
  const int k = a.k;

  uint8_t *qs = smem;
  int *qq = reinterpret_cast<int *>(smem + (size_t)QB * qbytes);
  int2 *meta = reinterpret_cast<int2 *>(qq + QB + (QB & 1));
  u64 *tkbase = reinterpret_cast<u64 *>(meta + kWarps * QB);
  const size_t tk_u64 = (size_t)2 * k + 32 + 2;

  const int lpr = 1 << a.lpr_log2, R = 32 >> a.lpr_log2;
  const int l = lane & (lpr - 1), sub = lane >> a.lpr_log2;
  const uint64_t rows_per_warp_step = (uint64_t)kU * R;
  const uint64_t rows_per_tile = rows_per_warp_step * kWarps;
  const int metric = a.rv.metric;
  const float dq = (float)a.rv.d;
  const uint32_t n_items = *a.n_items;

  if (threadIdx.x == 0) s_next = atomicAdd(a.counter, 1u);
  for (;;) {
    __syncthreads();  // the previous item's shared state is no longer read
    if (threadIdx.x == 0) {  // one item of look-ahead: its first rows are prefetched while this one finishes
      s_item = s_next;
      s_next = s_item < n_items ? atomicAdd(a.counter, 1u) : n_items;
    }
    __syncthreads();
    const uint32_t item = s_item;
    if (item >= n_items) break;
    const uint32_t list = a.item_list[item], first = a.item_first[item];
    const int nq = (int)a.item_n[item];
    if (threadIdx.x < QB) {
      const bool live = (int)threadIdx.x < nq;
      s_q[threadIdx.x] = live ? a.pair_q[first + threadIdx.x] : 0u;
      s_slot[threadIdx.x] = live ? a.pair_slot[first + threadIdx.x] : 0u;
    }
    __syncthreads();
    // the item's queries (compute layout) into shared memory
    for (int j = 0; j < QB; ++j) {
      const bool live = j < nq;
      const uint8_t *src = (DT == ZVB200_DT_INT8)
                               ? a.qv.qraw + (size_t)s_q[j] * a.rv.row_stride
                               : reinterpret_cast<const uint8_t *>(a.qv.qf32) + (size_t)s_q[j] * qbytes;
      for (size_t o = (size_t)threadIdx.x * 16; o < qbytes; o += (size_t)kThreads * 16) {
        uint4 v = make_uint4(0, 0, 0, 0);
        if (live) v = *reinterpret_cast<const uint4 *>(src + o);
        *reinterpret_cast<uint4 *>(qs + (size_t)j * qbytes + o) = v;
      }
    }
    __syncthreads();
    if (DT == ZVB200_DT_INT8 && L2 && threadIdx.x < QB) {
      int s = 0;
      const int8_t *p = reinterpret_cast<const int8_t *>(qs + (size_t)threadIdx.x * qbytes);
      for (size_t i = 0; i < qbytes; ++i) s += (int)p[i] * (int)p[i];
      qq[threadIdx.x] = s;
    }
    __syncthreads();

    // The bound a list starts from: the smallest k-th best key any finished item of the SAME query has published.
    // Such a key is the k-th best of k actual rows of the query, so the query's final k-th best cannot be larger:
    // rows at or beyond it can be skipped, the merged result stays exact (and does not depend on the order in which
    // the items run).  Without it every (query, list) pair restarts from an empty list — 610 rows per list over 8
    // warps: 40 % of all rows were accepted into a warp's list, and the accepts cost more than the scan.
    u64 tau[QB];
#pragma unroll
    for (int j = 0; j < QB; ++j) {
      tau[j] = j < nq ? *reinterpret_cast<volatile u64 *>(a.gtau + s_q[j]) : 0ull;
      if (lane < 2) tkbase[(size_t)(warp * QB + j) * tk_u64 + lane] = 0ull;  // empty list, empty queue, list = first buffer
    }
    __syncwarp();
    float4 qt[QB];
    if (metric_is_qi8(metric)) {
#pragma unroll
      for (int j = 0; j < QB; ++j) qt[j] = j < nq ? a.qv.qtails[s_q[j]] : make_float4(0, 0, 0, 0);
    }

    // the scan proper, compiled once per number of live queries (no per-query guards in the hot loop)
    const uint64_t vbegin = a.list_off[list], vend = a.list_off[list + 1];
    LmItem it;
    it.vbegin = vbegin; it.vend = vend; it.qs = qs; it.qbytes = qbytes; it.qq = qq; it.tkwarp = tkbase + (size_t)warp * QB * tk_u64;
    it.tk_u64 = tk_u64; it.warp = warp; it.lane = lane;
    switch (nq) {
      case 1: lm_scan_item<DT, L2, QB, 1>(a, it, tau, qt); break;
      case 2: lm_scan_item<DT, L2, QB, 2>(a, it, tau, qt); break;
      case 3: lm_scan_item<DT, L2, QB, 3>(a, it, tau, qt); break;
      case 4: lm_scan_item<DT, L2, QB, 4>(a, it, tau, qt); break;
      case 5: lm_scan_item<DT, L2, QB, (QB > 4 ? 5 : 4)>(a, it, tau, qt); break;
      case 6: lm_scan_item<DT, L2, QB, (QB > 4 ? 6 : 4)>(a, it, tau, qt); break;
      case 7: lm_scan_item<DT, L2, QB, (QB > 4 ? 7 : 4)>(a, it, tau, qt); break;
      default: lm_scan_item<DT, L2, QB, (QB > 4 ? 8 : 4)>(a, it, tau, qt); break;
    }

    // merge the warps' lists inside the CTA, one owner warp per query; one partial list per (query, probe slot)
#pragma unroll
    for (int j = 0; j < QB; ++j) {
      if (j < nq) tau[j] = lm_finish(tkbase + (size_t)(warp * QB + j) * tk_u64, k, tau[j]);
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < QB; ++j) {
      if (warp != (j % kWarps) || j >= nq) continue;
      u64 *mine = tkbase + (size_t)(warp * QB + j) * tk_u64;
      for (int w = 0; w < kWarps; ++w) {
        if (w == warp) continue;
        const u64 *ob = tkbase + (size_t)(w * QB + j) * tk_u64;
        const int ocnt = (int)(uint32_t)ob[0];
        const u64 *src = ob + 2 + (ob[1] ? k : 0);
        for (int i0 = 0; i0 < ocnt; i0 += 32) {
          const int i = i0 + lane;
          const u64 e = i < ocnt ? src[i] : kMaxKey;
          const bool has = i < ocnt && e < tau[j];
          if (__any_sync(0xffffffffu, has)) tau[j] = lm_push(mine, k, has, e, tau[j]);
        }
      }
      tau[j] = lm_finish(mine, k, tau[j]);
      const int mcnt = (int)(uint32_t)mine[0];
      const u64 *mlist = mine + 2 + (mine[1] ? k : 0);
      u64 *dst = a.part + ((size_t)s_q[j] * a.nslots + s_slot[j]) * k;
      for (int i = lane; i < k; i += 32) dst[i] = i < mcnt ? mlist[i] : kMaxKey;
      if (lane == 0 && mcnt == k) atomicMin(reinterpret_cast<unsigned long long *>(a.gtau + s_q[j]), (unsigned long long)mlist[k - 1]);
    }
  }
}

size_t lm_smem(int dtype, uint32_t chunks, int qb, uint32_t k) {
  const size_t qbytes = dtype == ZVB200_DT_FP16 ? (size_t)chunks * 32 : (size_t)chunks * 16;
  return (size_t)qb * qbytes + sizeof(int) * (qb + (qb & 1)) + sizeof(int2) * kWarps * qb +
         (size_t)kWarps * qb * (2 * (size_t)k + 32 + 2) * sizeof(u64);
}

template <int DT, bool L2, int QB>
cudaError_t launch_lm(const LmArgs &a, int grid, size_t smem, cudaStream_t st) {
  auto kern = scan_lists_lm_kernel<DT, L2, QB>;
  // one constant for every launch of this function (see kMaxDynSmem); the kernel's static shared words come off the
  // per-CTA limit
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem - 1024);
  if (e != cudaSuccess) return e;
  kern<<<grid, kThreads, smem, st>>>(a);
  count_launch();
  return cudaGetLastError();
}

template <int DT, bool L2>
cudaError_t launch_lm_qb(const LmArgs &a, int qb, int grid, size_t smem, cudaStream_t st) {
  return qb == 8 ? launch_lm<DT, L2, 8>(a, grid, smem, st) : launch_lm<DT, L2, 4>(a, grid, smem, st);
}

}  // namespace

// queries per list pass: 8 when their vectors and the warps' top-k lists fit, else 4; 0 = not supported
int scan_lists_lm_qb(const RowsView &rv, uint32_t k) {
  for (int qb : {8, 4})
    if (lm_smem(rv.dtype, rv.chunks, qb, k) <= 200 * 1024) return qb;
  return 0;
}

// scratch words (uint32) the plan needs: cnt, fill, pair_off, item_off [nlist each] | pair_q, pair_slot, item_list,
// item_first, item_n [q * cap each] | n_items, counter
size_t scan_lists_lm_scratch_words(uint32_t q, uint32_t cap, uint32_t nlist) {
  return (size_t)2 * q + (size_t)4 * nlist + (size_t)5 * q * cap + 4;
}

cudaError_t launch_scan_lists_lm(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                                 const uint32_t *seg_list, const uint32_t *seg_count, uint32_t cap,
                                 const uint32_t *list_off, uint32_t nlist, uint32_t *scratch, u64 *part,
                                 uint32_t *nparts, int sm_count, cudaStream_t st) {
  const uint32_t q = qv.q;
  const int qb = scan_lists_lm_qb(rv, sp.k);
  if (qb == 0) return cudaErrorInvalidValue;
  u64 *gtau = reinterpret_cast<u64 *>(scratch);  // first: 8-byte aligned
  uint32_t *cnt = scratch + (size_t)2 * q, *fill = cnt + nlist, *pair_off = fill + nlist, *item_off = pair_off + nlist;
  uint32_t *pair_q = item_off + nlist, *pair_slot = pair_q + (size_t)q * cap, *item_list = pair_slot + (size_t)q * cap;
  uint32_t *item_first = item_list + (size_t)q * cap, *item_n = item_first + (size_t)q * cap;
  uint32_t *n_items = item_n + (size_t)q * cap, *counter = n_items + 1, *front = counter + 1;
  cudaError_t e = cudaMemsetAsync(cnt, 0, (size_t)2 * nlist * 4, st);  // cnt, fill
  if (e == cudaSuccess) e = cudaMemsetAsync(n_items, 0, 16, st);        // n_items, counter, front[2]
  if (e == cudaSuccess) e = cudaMemsetAsync(gtau, 0xff, (size_t)q * sizeof(u64), st);
  // slots a query does not visit stay empty lists
  if (e == cudaSuccess) e = cudaMemsetAsync(part, 0xff, (size_t)q * cap * sp.k * sizeof(u64), st);
  if (e != cudaSuccess) return e;
  const uint32_t pairs = q * cap;
  lm_count_kernel<<<(pairs + 255) / 256, 256, 0, st>>>(seg_list, seg_count, q, cap, cnt);
  lm_scan_kernel<<<1, 1024, 0, st>>>(cnt, nlist, (uint32_t)qb, pair_off, item_off, n_items);
  lm_fill_pairs_kernel<<<(pairs + 255) / 256, 256, 0, st>>>(seg_list, seg_count, q, cap, pair_off, fill, pair_q, pair_slot);
  lm_fill_items_kernel<<<(nlist + 127) / 128, 128, 0, st>>>(cnt, pair_off, item_off, pair_slot, n_items, nlist, (uint32_t)qb,
                                                          front, item_list, item_first, item_n);
  count_launch(4);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  LmArgs a;
  a.rv = rv; a.qv = qv; a.k = sp.k; a.threshold = sp.threshold; a.filter = sp.filter;
  a.list_off = list_off; a.pair_q = pair_q; a.pair_slot = pair_slot;
  a.item_list = item_list; a.item_first = item_first; a.item_n = item_n; a.n_items = n_items; a.counter = counter;
  a.part = part; a.nslots = cap; a.gtau = gtau;
  a.lpr_log2 = pick_lpr_log2(rv.chunks);
  {
    static const uint32_t pf = [] { const char *e = getenv("ZVB200_LM_PF"); return e ? (uint32_t)atoi(e) : 4u; }();
    a.pf_dist = pf;
  }
  *nparts = cap;
  const size_t smem = lm_smem(rv.dtype, rv.chunks, qb, sp.k);
  const int grid = sm_count;
  const bool l2 = metric_is_l2(rv.metric);
  if (rv.dtype == ZVB200_DT_FP32)
    return l2 ? launch_lm_qb<ZVB200_DT_FP32, true>(a, qb, grid, smem, st) : launch_lm_qb<ZVB200_DT_FP32, false>(a, qb, grid, smem, st);
  if (rv.dtype == ZVB200_DT_FP16)
    return l2 ? launch_lm_qb<ZVB200_DT_FP16, true>(a, qb, grid, smem, st) : launch_lm_qb<ZVB200_DT_FP16, false>(a, qb, grid, smem, st);
  return l2 ? launch_lm_qb<ZVB200_DT_INT8, true>(a, qb, grid, smem, st) : launch_lm_qb<ZVB200_DT_INT8, false>(a, qb, grid, smem, st);
}

}  // namespace zvb

// K1 — GEMV-style streaming scan with fused warp top-k (small query batches).
//
// Replaces the per-row loop of FlatStreamerEntity::search_block
// (src/core/algorithm/flat/flat_streamer_entity.cc:256-316) /
// FlatSearcherContext::search_row_nofilter (flat_searcher_context.h:847-921): for every
// stored row one ailego XxxDistanceMatrix<T,1,1>::Compute (euclidean_distance_matrix_fp32.cc:287,
// inner_product_matrix_fp16.cc:167-189, inner_product_matrix_int8.cc:374) and one
// heap->emplace.  HBM-bound: every stored byte is read exactly once with 128-bit
// L1-bypassing loads, the queries live in shared memory, nothing but k keys per CTA
// goes back to HBM.
#include <stdlib.h>

#include "chunk.cuh"
#include "internal.h"
#include "topk.cuh"

namespace zvb {

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kUmax = 8;  // row groups in flight per warp (memory-level parallelism); the kernels' default

struct GemvArgs {
  RowsView rv;
  QueryView qv;
  uint32_t q0;        // first query of this launch
  uint32_t k;
  float threshold;
  const uint32_t *filter;
  uint64_t row_begin, row_end;
  u64 *part;          // [q][nparts][k]
  uint32_t nparts;
  int lpr_log2;       // lanes per row = 1 << lpr_log2
  // SEG mode (IVF posting-list scan): query blockIdx.y scans the concatenation of its selected
  // lists; virtual row v of segment i (seg_prefix[i] <= v < seg_prefix[i+1]) is stored row
  // seg_row[i] + v - seg_prefix[i].  Arrays are [q][seg_cap(+1)].
  const uint32_t *seg_prefix, *seg_row, *seg_count;
  uint32_t seg_cap;
  // non-null: no top-k at all — every row's key (kMaxKey for a dropped / filtered / unscorable one) goes to
  // all_keys[query * keys_stride + row - row_begin]: a scan of FEW rows whose k takes most of them (the IVF coarse stage:
  // 55 rows per CTA, k = 64)
  u64 *all_keys;
  uint32_t keys_stride;
};


// PRE > 0: rows of at most PRE chunks per lane (short rows: 96-B int8 rows over 2 lanes = 3) are loaded WHOLE before
// the first multiply — one trip to HBM per warp step instead of one per chunk iteration.  A warp step of such rows
// is latency-bound (ncu, 40M x 96 int8: 0.41 IPC, 4.3 of 9 warp-cycles per issue on the long scoreboard, DRAM at
// 4.8 of 6.4 TB/s): the three dependent trips were the step.
template <int DT, bool L2, int QB, bool SEG, int kU = kUmax, int PRE = 0>
__global__ void __launch_bounds__(kThreads, QB == 1 ? 2 : 1) scan_gemv_kernel(const GemvArgs a) {
  typedef Chunk<DT, L2> CK;
  typedef typename CK::Acc Acc;
  extern __shared__ __align__(16) uint8_t smem[];
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  const uint32_t C = a.rv.chunks;
  const size_t qbytes = CK::qbytes(C);
  const uint32_t qbase = SEG ? a.q0 + blockIdx.y : a.q0;
  const int nq = min((uint32_t)QB, a.qv.q - qbase);  // live queries of this launch
  const int k = a.k;

  // ---- smem: QB queries | int qq[QB] | meta[kWarps*QB] | top-k storage ----
  uint8_t *qs = smem;
  int *qq = reinterpret_cast<int *>(smem + (size_t)QB * qbytes);
  int2 *meta = reinterpret_cast<int2 *>(qq + QB + (QB & 1));  // {list offset in u64s, cnt}
  u64 *tkbase = reinterpret_cast<u64 *>(meta + kWarps * QB);
  const size_t tk_u64 = (size_t)2 * k + 32;
  // SEG: this query's segment table behind the top-k storage
  uint32_t *sprefix = reinterpret_cast<uint32_t *>(tkbase + (size_t)kWarps * QB * tk_u64);
  uint32_t *srow = sprefix + a.seg_cap + 1;
  uint32_t nseg = 0;
  if (SEG) {
    nseg = a.seg_count[qbase];
    for (uint32_t i = threadIdx.x; i <= nseg; i += kThreads) sprefix[i] = a.seg_prefix[(size_t)qbase * (a.seg_cap + 1) + i];
    for (uint32_t i = threadIdx.x; i < nseg; i += kThreads) srow[i] = a.seg_row[(size_t)qbase * a.seg_cap + i];
  }

  // load queries (compute layout) into smem
  for (int j = 0; j < QB; ++j) {
    const bool live = j < nq;
    const uint8_t *src = (DT == ZVB200_DT_INT8)
                             ? a.qv.qraw + (size_t)(qbase + j) * a.rv.row_stride
                             : reinterpret_cast<const uint8_t *>(a.qv.qf32) + (size_t)(qbase + j) * qbytes;
    for (size_t o = (size_t)threadIdx.x * 16; o < qbytes; o += (size_t)kThreads * 16) {
      uint4 v = make_uint4(0, 0, 0, 0);
      if (live) v = *reinterpret_cast<const uint4 *>(src + o);
      *reinterpret_cast<uint4 *>(qs + (size_t)j * qbytes + o) = v;
    }
  }
  __syncthreads();
  if (DT == ZVB200_DT_INT8 && L2 && threadIdx.x < QB) {  // qq = sum q^2 (plain int8 L2)
    int s = 0;
    const int8_t *p = reinterpret_cast<const int8_t *>(qs + (size_t)threadIdx.x * qbytes);
    for (size_t i = 0; i < qbytes; ++i) s += (int)p[i] * (int)p[i];
    qq[threadIdx.x] = s;
  }
  __syncthreads();

  WarpTopk tk[QB];
#pragma unroll
  for (int j = 0; j < QB; ++j) {
    u64 *b = tkbase + (size_t)(warp * QB + j) * tk_u64;
    tk[j].init(b, b + k, b + 2 * k, k, kMaxKey);
  }

  const int lpr = 1 << a.lpr_log2, R = 32 >> a.lpr_log2;
  const int l = lane & (lpr - 1), sub = lane >> a.lpr_log2;
  const uint64_t rows_per_warp_step = (uint64_t)kU * R;
  const uint64_t rows_per_tile = rows_per_warp_step * kWarps;
  // row space walked by this CTA: the store's [row_begin, row_end), or the virtual rows of the query's lists
  const uint64_t vbegin = SEG ? 0 : a.row_begin;
  const uint64_t vend = SEG ? (uint64_t)sprefix[nseg] : a.row_end;
  const uint64_t nrows = vend - vbegin;
  const uint64_t ntiles = (nrows + rows_per_tile - 1) / rows_per_tile;
  auto to_row = [&](uint64_t v) -> uint64_t {
    if (!SEG) return v;
    uint32_t lo = 0, hi = nseg;  // last segment with prefix <= v
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (sprefix[mid] <= (uint32_t)v) lo = mid; else hi = mid;
    }
    return (uint64_t)srow[lo] + ((uint32_t)v - sprefix[lo]);
  };
  const int metric = a.rv.metric;
  const float dq = (float)a.rv.d;

  float4 qt[QB];
  if (metric_is_qi8(metric)) {
#pragma unroll
    for (int j = 0; j < QB; ++j) qt[j] = j < nq ? a.qv.qtails[qbase + j] : make_float4(0, 0, 0, 0);
  }

  for (uint64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const uint64_t base = vbegin + t * rows_per_tile + (uint64_t)warp * rows_per_warp_step;
    if (base >= vend) continue;  // warp-uniform
    Acc acc[kU][QB], aa[kU];
    const uint8_t *rp[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      uint64_t r = base + (uint64_t)u * R + sub;
      if (r >= vend) r = vend - 1;  // clamp: loaded, result discarded
      const uint64_t sr = to_row(r);
      rp[u] = a.rv.rows + sr * a.rv.row_stride;
      // QuantizedInteger: the row's tail is read only after the dot product; start its trip from HBM now (no
      // register held: the 128-register budget of 2 CTAs per SM is full) so that read finds it in L2
      // (one request per 64 B of tails = 4 rows; one per row measured the same)
      if (metric_is_qi8(metric) && l == 0 && (sr & 3) == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.rv.tails + sr));
      aa[u] = 0;
#pragma unroll
      for (int j = 0; j < QB; ++j) acc[u][j] = 0;
    }
    if (PRE > 0) {
      for (uint32_t c0 = l; c0 < C; c0 += PRE * lpr) {
        uint4 v[PRE > 0 ? PRE : 1][kU];
#pragma unroll
        for (int i = 0; i < PRE; ++i) {
          const uint32_t c = c0 + i * lpr;
#pragma unroll
          for (int u = 0; u < kU; ++u) v[i][u] = c < C ? ldg_stream(rp[u] + (size_t)c * 16) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int i = 0; i < PRE; ++i) {
          const uint32_t c = c0 + i * lpr;
          if (c < C) {
#pragma unroll
            for (int j = 0; j < QB; ++j) {
#pragma unroll
              for (int u = 0; u < kU; ++u) {
                Acc dummy = 0;
                CK::fma(acc[u][j], j == 0 ? aa[u] : dummy, v[i][u], qs + (size_t)j * qbytes, c);
              }
            }
          }
        }
      }
    } else
    for (uint32_t c = l; c < C; c += lpr) {
      uint4 v[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) v[u] = ldg_stream(rp[u] + (size_t)c * 16);
#pragma unroll
      for (int j = 0; j < QB; ++j) {
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          Acc dummy = 0;
          CK::fma(acc[u][j], j == 0 ? aa[u] : dummy, v[u], qs + (size_t)j * qbytes, c);
        }
      }
    }
    // reduce across the lanes of a row
    for (int o = lpr >> 1; o > 0; o >>= 1) {
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        if (DT == ZVB200_DT_INT8 && L2) aa[u] += __shfl_xor_sync(0xffffffffu, aa[u], o);
#pragma unroll
        for (int j = 0; j < QB; ++j) acc[u][j] += __shfl_xor_sync(0xffffffffu, acc[u][j], o);
      }
    }
    // epilogue: score, gate, push
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const uint64_t rv_ = base + (uint64_t)u * R + sub;
      bool valid = (l == 0) && rv_ < vend;
      const uint64_t r = SEG ? (uint64_t)((rp[u] - a.rv.rows) / a.rv.row_stride) : rv_;
      if (valid && a.rv.drop) valid = !bit_set(a.rv.drop, r);
      if (valid && a.filter) valid = !bit_set(a.filter, r);
      float4 mt = make_float4(0, 0, 0, 0);
      // through the read-only, no-L1-allocate path like the rows: a plain ld.global of the tails costs the C3 scan
      // 14 % (4.19 -> 4.77 TB/s at 100M x 96; ncu: 1.43x -> 1.20x the algorithmic DRAM bytes)
      if (metric_is_qi8(metric) && valid) mt = ldg_tail(a.rv.tails + r);
      // the row's score scale (IVFEntity::search: distances[k] * norm_val, ivf_entity.cc:650); 1 for a plain store
      const float sc = a.rv.row_scale ? (valid ? a.rv.row_scale[r] : 1.f) : a.rv.score_scale;
#pragma unroll
      for (int j = 0; j < QB; ++j) {
        float s;
        if (DT == ZVB200_DT_INT8) {
          if (metric == ZVB200_METRIC_QI8_L2SQ) s = qi8_l2sq((float)acc[u][j], mt, qt[j], dq);
          else if (metric_is_qi8(metric)) s = qi8_mip_ord((float)acc[u][j], mt, qt[j], dq, a.rv.query_first);
          else if (L2) {
            s = (float)((int)aa[u] - 2 * (int)acc[u][j] + qq[j]);
            if (metric == ZVB200_METRIC_L2) s = sqrtf(s);
          } else s = -(float)acc[u][j];
        } else {
          const float f = (float)acc[u][j];
          s = metric == ZVB200_METRIC_L2SQ ? f
              : metric == ZVB200_METRIC_L2 ? sqrtf(f)
              : metric == ZVB200_METRIC_IP ? -f : 1.0f - f;
        }
        s *= sc;
        const u64 key = pack_key(s, (uint32_t)r);
        if (!SEG && a.all_keys) {  // every row of the range gets a word: its key, or "none"
          if (l == 0 && rv_ < vend && j < nq)
            a.all_keys[(size_t)(qbase + j) * a.keys_stride + (r - a.row_begin)] = (valid && s <= a.threshold) ? key : kMaxKey;
          continue;
        }
        const bool has = valid && j < nq && s <= a.threshold && key < tk[j].tau;
        tk[j].push(has, key, kMaxKey);
      }
    }
  }

  if (!SEG && a.all_keys) return;
  // ---- merge the warps' lists inside the CTA, one owner warp per query ----
#pragma unroll
  for (int j = 0; j < QB; ++j) {
    tk[j].finish(kMaxKey);
    if (lane == 0) meta[warp * QB + j] = make_int2((int)(tk[j].list - tkbase), tk[j].cnt);
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < QB; ++j) {
    if (warp != (j % kWarps) || j >= nq) continue;
    for (int w = 0; w < kWarps; ++w) {
      if (w == warp) continue;
      const int2 m = meta[w * QB + j];
      const u64 *src = tkbase + m.x;
      for (int i0 = 0; i0 < m.y; i0 += 32) {
        const int i = i0 + lane;
        const u64 e = i < m.y ? src[i] : kMaxKey;
        tk[j].push(i < m.y && e < tk[j].tau, e, kMaxKey);
      }
    }
    tk[j].finish(kMaxKey);
    u64 *dst = a.part + ((size_t)(qbase + j) * a.nparts + blockIdx.x) * k;
    for (int i = lane; i < k; i += 32) dst[i] = i < tk[j].cnt ? tk[j].list[i] : kMaxKey;
  }
}

template <int DT, bool L2, int QB, bool SEG = false, int KU = kUmax, int PRE = 0>
cudaError_t launch_one(const GemvArgs &a, int grid, size_t smem, cudaStream_t st, int grid_y = 1) {
  auto kern = scan_gemv_kernel<DT, L2, QB, SEG, KU, PRE>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess && smem > (size_t)kMaxDynSmem) e = cudaErrorInvalidValue;
  if (e != cudaSuccess) return e;
  kern<<<dim3(grid, grid_y), kThreads, smem, st>>>(a);
  count_launch();
  return cudaGetLastError();
}

// whole-row preload (PRE): int8 rows of at most kPre chunks per lane, kPreU row groups in flight (12 x 16 B per lane)
constexpr int kPre = 3, kPreU = 4;
// Where it pays (B200, batch 1 unless noted): 96-B int8 rows 5.0 -> 5.5 TB/s, QuantizedInteger 96-B rows 4.9 -> 5.1,
// QuantizedInteger d = 768 5.6 -> 6.2, fp32 d = 1536 7.0 -> 7.3, fp16 d = 768 with 4 queries 8.9 -> 8.1 ms; where it does not:
// 512-B fp32 rows (8 chunks per lane = trips of 3, 3, 2: 6.9 -> 6.6 TB/s) and scans of a few tiles per CTA (the 16 384
// centroids of C4: 64-row tiles are one round of the grid, 32-row tiles two).
static bool pre_applies(const RowsView &rv, int lpr_log2, uint64_t rows, int grid) {
  static const int mode = [] { const char *e = getenv("ZVB200_GEMV_PRE"); return e ? atoi(e) : 1; }();
  if (mode != 1) return mode == 2;
  const uint32_t per_lane = (rv.chunks + (1u << lpr_log2) - 1) >> lpr_log2;
  const bool short_int8 = rv.dtype == ZVB200_DT_INT8 && per_lane <= (uint32_t)kPre;
  const bool long_rows = rv.row_stride >= 768 && per_lane % kPre == 0;
  const uint64_t tile = (uint64_t)kPreU * (32 >> lpr_log2) * kWarps;
  return (short_int8 || long_rows) && rows >= 4 * tile * (uint64_t)grid;
}

template <int DT, bool L2>
cudaError_t launch_qb(const GemvArgs &a, int qb, int grid, size_t smem, cudaStream_t st, bool pre = false) {
  if (pre) {
    switch (qb) {
      case 1: return launch_one<DT, L2, 1, false, kPreU, kPre>(a, grid, smem, st);
      case 2: return launch_one<DT, L2, 2, false, kPreU, kPre>(a, grid, smem, st);
      default: return launch_one<DT, L2, 4, false, kPreU, kPre>(a, grid, smem, st);
    }
  }
  switch (qb) {
    case 1: return launch_one<DT, L2, 1>(a, grid, smem, st);
    case 2: return launch_one<DT, L2, 2>(a, grid, smem, st);
    default: return launch_one<DT, L2, 4>(a, grid, smem, st);
  }
}

size_t gemv_smem(int dtype, uint32_t chunks, int qb, uint32_t k, uint32_t seg_cap = 0) {
  size_t qbytes = dtype == ZVB200_DT_FP16 ? (size_t)chunks * 32 : (size_t)chunks * 16;
  return (size_t)qb * qbytes + sizeof(int) * (qb + (qb & 1)) + sizeof(int2) * kWarps * qb +
         (size_t)kWarps * qb * (2 * (size_t)k + 32) * sizeof(u64) + (seg_cap ? (2 * (size_t)seg_cap + 2) * 4 : 0);
}

}  // namespace

size_t scan_gemv_max_parts(int sm_count) { return (size_t)sm_count * 2; }

cudaError_t launch_scan_gemv(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                             uint64_t row_begin, uint64_t row_end, u64 *part, uint32_t *nparts,
                             int sm_count, cudaStream_t st, u64 *all_keys, uint32_t keys_stride) {
  GemvArgs a;
  a.all_keys = all_keys; a.keys_stride = keys_stride;
  a.seg_prefix = a.seg_row = a.seg_count = nullptr; a.seg_cap = 0;
  a.rv = rv; a.qv = qv; a.k = sp.k; a.threshold = sp.threshold; a.filter = sp.filter;
  a.row_begin = row_begin; a.row_end = row_end; a.part = part;
  const int lg = pick_lpr_log2(rv.chunks);
  a.lpr_log2 = lg;
  const uint64_t nrows0 = row_end > row_begin ? row_end - row_begin : 0;
  const bool pre = pre_applies(rv, lg, nrows0, sm_count * 2);
  const uint64_t rows_per_tile = (uint64_t)(pre ? kPreU : kUmax) * (32 >> lg) * kWarps;
  const uint64_t nrows = row_end > row_begin ? row_end - row_begin : 0;
  uint64_t ntiles = (nrows + rows_per_tile - 1) / rows_per_tile;
  int grid = (int)std::min<uint64_t>(std::max<uint64_t>(ntiles, 1), (uint64_t)sm_count * 2);
  a.nparts = grid;
  *nparts = grid;
  // queries per pass: as many as fit 100 KB of smem (2 CTAs/SM), at most 4
  int qb = qv.q >= 4 ? 4 : (qv.q >= 2 ? 2 : 1);
  {
    static const int force = [] { const char *e = getenv("ZVB200_GEMV_QB"); return e ? atoi(e) : 0; }();
    if (force == 1 || force == 2 || force == 4) qb = std::min(force, qb);
  }
  while (qb > 1 && gemv_smem(rv.dtype, rv.chunks, qb, sp.k) > 100 * 1024) qb >>= 1;
  const size_t smem = gemv_smem(rv.dtype, rv.chunks, qb, sp.k);
  if (smem > 220 * 1024) return cudaErrorInvalidValue;
  const bool l2 = metric_is_l2(rv.metric);
  for (uint32_t q0 = 0; q0 < qv.q; q0 += qb) {
    a.q0 = q0;
    cudaError_t e;
    if (rv.dtype == ZVB200_DT_FP32)
      e = l2 ? launch_qb<ZVB200_DT_FP32, true>(a, qb, grid, smem, st, pre)
             : launch_qb<ZVB200_DT_FP32, false>(a, qb, grid, smem, st, pre);
    else if (rv.dtype == ZVB200_DT_FP16)
      e = l2 ? launch_qb<ZVB200_DT_FP16, true>(a, qb, grid, smem, st, pre)
             : launch_qb<ZVB200_DT_FP16, false>(a, qb, grid, smem, st, pre);
    else
      e = l2 ? launch_qb<ZVB200_DT_INT8, true>(a, qb, grid, smem, st, pre)
             : launch_qb<ZVB200_DT_INT8, false>(a, qb, grid, smem, st, pre);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

// K4 — IVF posting-list scan: grid.y = query, each query walks ITS selected lists.
// Replaces the per-list loop of IVFSearcher::search_impl (src/core/algorithm/ivf/
// ivf_searcher.cc:216-231) -> IVFEntity::search(list, query, ...) (ivf_entity.cc:662-716).
size_t scan_lists_parts(uint32_t q, int sm_count) {
  const uint32_t target = (uint32_t)sm_count * 8;  // CTAs wanted in flight (2 resident per SM, 4 waves)
  return std::max<uint32_t>(1, std::min<uint32_t>((target + q - 1) / q, (uint32_t)sm_count * 2));
}

// Row groups per warp step for a SHORT scan: the CTAs walk ceil(rows / (KU * R * 8)) tiles round-robin, so the scan
// lasts rounds * KU row-times; 39 063 rows of C4 over 296 CTAs are 2.06 rounds of 64-row tiles (3 rounds: 24) but
// 2.77 rounds of 48-row tiles (3 rounds: 18).
static int pick_ku(uint64_t rows_hint, int lpr_log2, int grid) {
  if (rows_hint == 0) return kUmax;
  int best = kUmax;
  uint64_t best_cost = ~0ull;
  for (int ku : {8, 6, 4}) {
    const uint64_t tile = (uint64_t)ku * (32 >> lpr_log2) * kWarps;
    const uint64_t rounds = ((rows_hint + tile - 1) / tile + grid - 1) / grid;
    const uint64_t cost = rounds * ku * 16 + (8 - ku);  // ties go to the deeper pipeline
    if (cost < best_cost) { best_cost = cost; best = ku; }
  }
  return best;
}

template <int DT, bool L2>
static cudaError_t launch_lists_ku(const GemvArgs &a, int ku, int grid, size_t smem, cudaStream_t st, int gy) {
  // the preload form of a query's lists: 3 row groups x 4 chunk iterations in flight (12 loads per lane) — C4's 39 063
  // rows: 4 x 3 64 -> 61 us, 6 x 2 60, 3 x 5 59, 3 x 4 and 2 x 6 57-58 (24-row tiles: 5.5 rounds of the grid, not 4.1)
  if (ku == -1) return launch_one<DT, L2, 1, true, 3, 4>(a, grid, smem, st, gy);
  switch (ku) {
    case 4: return launch_one<DT, L2, 1, true, 4>(a, grid, smem, st, gy);
    case 6: return launch_one<DT, L2, 1, true, 6>(a, grid, smem, st, gy);
    default: return launch_one<DT, L2, 1, true, 8>(a, grid, smem, st, gy);
  }
}

cudaError_t launch_scan_lists(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                              const uint32_t *seg_prefix, const uint32_t *seg_row, const uint32_t *seg_count,
                              uint32_t seg_cap, u64 *part, uint32_t *nparts, int sm_count, cudaStream_t st,
                              uint64_t rows_hint) {
  GemvArgs a;
  a.all_keys = nullptr; a.keys_stride = 0;
  a.rv = rv; a.qv = qv; a.k = sp.k; a.threshold = sp.threshold; a.filter = sp.filter;
  a.row_begin = 0; a.row_end = rv.n; a.part = part; a.q0 = 0;
  a.seg_prefix = seg_prefix; a.seg_row = seg_row; a.seg_count = seg_count; a.seg_cap = seg_cap;
  a.lpr_log2 = pick_lpr_log2(rv.chunks);
  const int grid = (int)scan_lists_parts(qv.q, sm_count);
  a.nparts = grid;
  *nparts = grid;
  // a query's lists are a short scan by nature: the finer tiles of the preload form balance it better, no size rule
  const int ku = pre_applies(rv, a.lpr_log2, ~0ull >> 8, 1) ? -1 : pick_ku(rows_hint, a.lpr_log2, grid);
  const size_t smem = gemv_smem(rv.dtype, rv.chunks, 1, sp.k, seg_cap);
  if (smem > 220 * 1024) return cudaErrorInvalidValue;
  const bool l2 = metric_is_l2(rv.metric);
  const uint32_t ymax = 65535;
  for (uint32_t q0 = 0; q0 < qv.q; q0 += ymax) {
    a.q0 = q0;
    const int gy = (int)std::min<uint32_t>(ymax, qv.q - q0);
    cudaError_t e;
    if (rv.dtype == ZVB200_DT_FP32)
      e = l2 ? launch_lists_ku<ZVB200_DT_FP32, true>(a, ku, grid, smem, st, gy)
             : launch_lists_ku<ZVB200_DT_FP32, false>(a, ku, grid, smem, st, gy);
    else if (rv.dtype == ZVB200_DT_FP16)
      e = l2 ? launch_lists_ku<ZVB200_DT_FP16, true>(a, ku, grid, smem, st, gy)
             : launch_lists_ku<ZVB200_DT_FP16, false>(a, ku, grid, smem, st, gy);
    else
      e = l2 ? launch_lists_ku<ZVB200_DT_INT8, true>(a, ku, grid, smem, st, gy)
             : launch_lists_ku<ZVB200_DT_INT8, false>(a, ku, grid, smem, st, gy);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

}  // namespace zvb

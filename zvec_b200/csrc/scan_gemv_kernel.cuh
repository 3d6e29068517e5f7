// The GEMV scan's kernel and its per-(data type) launch tables (scan_gemv.cu holds the host side).  The kernel is
// instantiated for 3 data types x 2 metric forms x query counts x tile shapes; one translation unit per data type
// (scan_gemv_fp32.cu / _fp16.cu / _int8.cu define ZVB_GEMV_DT before including this) keeps the build parallel.
#pragma once
#include <stdlib.h>

#include "chunk.cuh"
#include "internal.h"
#include "topk.cuh"

namespace zvb {
namespace gemv {

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
constexpr int kPre = 3, kPreU = 4;

// the kernels of one data type: flat scan (qb queries per pass, optional whole-row preload) and posting lists
template <int DT>
cudaError_t launch_flat(const GemvArgs &a, bool l2, int qb, int grid, size_t smem, cudaStream_t st, bool pre);
template <int DT>
cudaError_t launch_lists(const GemvArgs &a, bool l2, int ku, int grid, size_t smem, cudaStream_t st, int gy);

#ifdef ZVB_GEMV_DT
template <int DT, bool L2, int QB, bool SEG, int kU = kUmax, int PRE = 0, bool TE = false, bool AK = false>
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
    // short QuantizedInteger rows, one query: the tails travel WITH the rows (a register each) instead of after them
    // (TE: its own instantiation — the four registers cost plain int8 rows 10 %, 5.5 -> 5.0 TB/s, and gain the
    // QuantizedInteger ones 4 %, 5.12 -> 5.32)
    constexpr bool kTailEarly = TE;
    float4 mte[kTailEarly ? kU : 1];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      uint64_t r = base + (uint64_t)u * R + sub;
      if (r >= vend) r = vend - 1;  // clamp: loaded, result discarded
      const uint64_t sr = to_row(r);
      rp[u] = a.rv.rows + sr * a.rv.row_stride;
      // QuantizedInteger: the row's tail is read only after the dot product; start its trip from HBM now (no
      // register held: the 128-register budget of 2 CTAs per SM is full) so that read finds it in L2
      // (one request per 64 B of tails = 4 rows; one per row measured the same)
      if (kTailEarly) {
        mte[u] = ldg_tail(a.rv.tails + sr);
      } else if (metric_is_qi8(metric) && l == 0 && (sr & 3) == 0) {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a.rv.tails + sr));
      }
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
      if (kTailEarly) mt = mte[u];
      else if (metric_is_qi8(metric) && valid) mt = ldg_tail(a.rv.tails + r);
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
        if (AK) {  // (its own instantiation: as a run-time branch it cost the short-row scans 10 %) every row of the
                   // range gets a word: its key, or "none"
          if (l == 0 && rv_ < vend && j < nq)
            a.all_keys[(size_t)(qbase + j) * a.keys_stride + (r - a.row_begin)] = (valid && s <= a.threshold) ? key : kMaxKey;
          continue;
        }
        const bool has = valid && j < nq && s <= a.threshold && key < tk[j].tau;
        tk[j].push(has, key, kMaxKey);
      }
    }
  }

  if (AK) return;
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

template <int DT, bool L2, int QB, bool SEG = false, int KU = kUmax, int PRE = 0, bool TE = false, bool AK = false>
cudaError_t launch_one(const GemvArgs &a, int grid, size_t smem, cudaStream_t st, int grid_y = 1) {
  auto kern = scan_gemv_kernel<DT, L2, QB, SEG, KU, PRE, TE, AK>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess && smem > (size_t)kMaxDynSmem) e = cudaErrorInvalidValue;
  if (e != cudaSuccess) return e;
  kern<<<dim3(grid, grid_y), kThreads, smem, st>>>(a);
  count_launch();
  return cudaGetLastError();
}

// whole-row preload (PRE): int8 rows of at most kPre chunks per lane, kPreU row groups in flight (12 x 16 B per lane)
template <int DT, bool L2>
cudaError_t launch_qb(const GemvArgs &a, int qb, int grid, size_t smem, cudaStream_t st, bool pre = false) {
  if (a.all_keys) return launch_one<DT, L2, 1, false, kUmax, 0, false, true>(a, grid, smem, st);  // one query per pass
  if (pre) {
    if (DT == ZVB200_DT_INT8 && qb == 1 && metric_is_qi8(a.rv.metric) && a.rv.chunks <= (uint32_t)(kPre << a.lpr_log2))
      return launch_one<DT, L2, 1, false, kPreU, kPre, DT == ZVB200_DT_INT8>(a, grid, smem, st);
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

template <int DT, bool L2>
cudaError_t launch_lists_ku(const GemvArgs &a, int ku, int grid, size_t smem, cudaStream_t st, int gy) {
  // the preload form of a query's lists: 3 row groups x 4 chunk iterations in flight (12 loads per lane) — C4's 39 063
  // rows: 4 x 3 64 -> 61 us, 6 x 2 60, 3 x 5 59, 3 x 4 and 2 x 6 57-58 (24-row tiles: 5.5 rounds of the grid, not 4.1)
  if (ku == -1) return launch_one<DT, L2, 1, true, 3, 4>(a, grid, smem, st, gy);
  switch (ku) {
    case 4: return launch_one<DT, L2, 1, true, 4>(a, grid, smem, st, gy);
    case 6: return launch_one<DT, L2, 1, true, 6>(a, grid, smem, st, gy);
    default: return launch_one<DT, L2, 1, true, 8>(a, grid, smem, st, gy);
  }
}


template <>
cudaError_t launch_flat<ZVB_GEMV_DT>(const GemvArgs &a, bool l2, int qb, int grid, size_t smem, cudaStream_t st, bool pre) {
  return l2 ? launch_qb<ZVB_GEMV_DT, true>(a, qb, grid, smem, st, pre) : launch_qb<ZVB_GEMV_DT, false>(a, qb, grid, smem, st, pre);
}
template <>
cudaError_t launch_lists<ZVB_GEMV_DT>(const GemvArgs &a, bool l2, int ku, int grid, size_t smem, cudaStream_t st, int gy) {
  return l2 ? launch_lists_ku<ZVB_GEMV_DT, true>(a, ku, grid, smem, st, gy)
            : launch_lists_ku<ZVB_GEMV_DT, false>(a, ku, grid, smem, st, gy);
}
#endif

}  // namespace gemv
}  // namespace zvb

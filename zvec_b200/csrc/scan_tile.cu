// K2c — tiled CUDA-core contraction (128 rows x 64 queries per CTA tile) with a fused
// block-level top-k epilogue.  Exact reference arithmetic: L2 is evaluated as
// sum (x-q)^2 like SquaredEuclideanDistanceMatrix (euclidean_distance_matrix.h:33-52), not
// through norms, so it is safe for clustered data; int8 accumulates in int32.
//
// Replaces FlatSearcherContext::batch_search_* (src/core/algorithm/flat/
// flat_searcher_context.h:681-752,847-921): the XxxMatrix<T,32,N> tile calls
// (inner_product_matrix.h:56-133) and the 32*N heap inserts per tile
// (batch_enqueue_nofilter, flat_searcher_context.h:236-261).  The score tile never leaves
// registers; per-query candidate queues and sorted lists live in shared memory.
#include "internal.h"
#include "topk.cuh"

namespace zvb {

namespace {

constexpr int BM = 128, BN = 64, kThreads = 256, kWarps = 8;
constexpr int QC = 64;        // candidate queue slots per query
constexpr int kMaxK = 128;    // lists for 64 queries must fit in shared memory

struct TileArgs {
  RowsView rv;
  QueryView qv;
  uint32_t k;
  float threshold;
  const uint32_t *filter;
  uint64_t row_begin, row_end;
  u64 *part;
  uint32_t nparts;  // = row splits
  uint32_t qtiles;
};

template <int DT> struct TileCfg;
template <> struct TileCfg<ZVB200_DT_FP32> { static constexpr int KK = 16, QCH = 1; };  // 16 floats / stage
template <> struct TileCfg<ZVB200_DT_FP16> { static constexpr int KK = 32, QCH = 2; };  // 32 halves -> floats
template <> struct TileCfg<ZVB200_DT_INT8> { static constexpr int KK = 16, QCH = 1; };  // 16 words of 4 codes

__device__ __forceinline__ bool bit_set(const uint32_t *bm, uint64_t i) {
  return (bm[i >> 5] >> (i & 31)) & 1u;
}

template <int DT, bool L2>
__global__ void __launch_bounds__(kThreads) scan_tile_kernel(const TileArgs a) {
  typedef TileCfg<DT> CFG;
  constexpr int KK = CFG::KK;
  constexpr bool I8 = DT == ZVB200_DT_INT8;
  extern __shared__ __align__(16) uint8_t smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tx = tid & 15, ty = tid >> 4;
  const int k = a.k;
  const uint32_t qt = blockIdx.x % a.qtiles, rs = blockIdx.x / a.qtiles;
  const uint32_t q0 = qt * BN;
  const int nq = min((uint32_t)BN, a.qv.q - q0);

  // ---- shared memory carve-up ----
  uint32_t *Xs = reinterpret_cast<uint32_t *>(smem);             // [2][KK][BM]
  uint32_t *Qs = Xs + 2 * KK * BM;                               // [2][KK][BN]
  u64 *lists = reinterpret_cast<u64 *>(Qs + 2 * KK * BN);        // [BN][2k]
  u64 *queue = lists + (size_t)BN * 2 * k;                       // [BN][QC]
  u64 *tau = queue + (size_t)BN * QC;                            // [BN]
  int *qn = reinterpret_cast<int *>(tau + BN);                   // [BN]
  int *cnt = qn + BN;                                            // [BN]
  int *cur = cnt + BN;                                           // [BN] 0/1: which half is live
  for (int i = tid; i < BN; i += kThreads) { tau[i] = kMaxKey; qn[i] = 0; cnt[i] = 0; cur[i] = 0; }
  __syncthreads();

  const uint32_t stride = a.rv.row_stride;
  const int nstages = (stride + 63) / 64;
  const uint64_t nrows = a.row_end - a.row_begin;
  const uint64_t ntiles = (nrows + BM - 1) / BM;
  const int metric = a.rv.metric;
  const float dq = (float)a.rv.d;
  // query side loads are identical for every row tile
  const uint8_t *qbase;
  uint32_t qstride_bytes;
  if (I8) { qbase = a.qv.qraw; qstride_bytes = stride; }
  else { qbase = reinterpret_cast<const uint8_t *>(a.qv.qf32); qstride_bytes = a.rv.chunks * 16 / (DT == ZVB200_DT_FP32 ? 4 : 2) * 4; }
  const int qstage_bytes = KK * 4;  // bytes of one query per stage in the compute layout

  float4 qtl[4];
  if (metric_is_qi8(metric)) {
#pragma unroll
    for (int j = 0; j < 4; ++j) qtl[j] = (tx * 4 + j) < nq ? a.qv.qtails[q0 + tx * 4 + j] : make_float4(0, 0, 0, 0);
  }

  for (uint64_t t = rs; t < ntiles; t += a.nparts) {
    const uint64_t row0 = a.row_begin + t * BM;
    float accf[8][4];
    int acci[8][4], aai[8], qqi[4];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      aai[i] = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) { accf[i][j] = 0.f; acci[i][j] = 0; }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) qqi[j] = 0;

    uint4 xr[2], qr[CFG::QCH];
    auto prefetch = [&](int s) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int c = tid + h * kThreads;            // 0..511
        const int r = c & (BM - 1), kc = c >> 7;     // row in tile, 16-B chunk in stage
        const uint64_t row = row0 + r;
        const uint32_t off = (uint32_t)s * 64 + kc * 16;
        xr[h] = make_uint4(0, 0, 0, 0);
        if (row < a.row_end && off < stride)
          xr[h] = *reinterpret_cast<const uint4 *>(a.rv.rows + row * stride + off);
      }
#pragma unroll
      for (int h = 0; h < CFG::QCH; ++h) {
        const int c = tid + h * kThreads;
        const int ql = c & (BN - 1), kc = c >> 6;    // kc: 0..3 (fp32/int8) or 0..7 (fp16)
        const uint32_t off = (uint32_t)s * qstage_bytes + kc * 16;
        qr[h] = make_uint4(0, 0, 0, 0);
        if (ql < nq && off < qstride_bytes)
          qr[h] = *reinterpret_cast<const uint4 *>(qbase + (size_t)(q0 + ql) * qstride_bytes + off);
      }
    };
    auto stage_store = [&](int buf) {
      uint32_t *xs = Xs + buf * KK * BM;
      uint32_t *qs = Qs + buf * KK * BN;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int c = tid + h * kThreads;
        const int r = c & (BM - 1), kc = c >> 7;
        const uint32_t w[4] = {xr[h].x, xr[h].y, xr[h].z, xr[h].w};
        if (DT == ZVB200_DT_FP16) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&w[i]));
            xs[(kc * 8 + 2 * i) * BM + r] = __float_as_uint(f.x);
            xs[(kc * 8 + 2 * i + 1) * BM + r] = __float_as_uint(f.y);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 4; ++i) xs[(kc * 4 + i) * BM + r] = w[i];
        }
      }
#pragma unroll
      for (int h = 0; h < CFG::QCH; ++h) {
        const int c = tid + h * kThreads;
        const int ql = c & (BN - 1), kc = c >> 6;
        const uint32_t w[4] = {qr[h].x, qr[h].y, qr[h].z, qr[h].w};
#pragma unroll
        for (int i = 0; i < 4; ++i) qs[(kc * 4 + i) * BN + ql] = w[i];
      }
    };

    prefetch(0);
    int buf = 0;
    for (int s = 0; s < nstages; ++s) {
      stage_store(buf);
      __syncthreads();
      if (s + 1 < nstages) prefetch(s + 1);
      const uint32_t *xs = Xs + buf * KK * BM + ty * 8;
      const uint32_t *qs = Qs + buf * KK * BN + tx * 4;
#pragma unroll
      for (int kk = 0; kk < KK; ++kk) {
        const uint4 a0 = *reinterpret_cast<const uint4 *>(xs + kk * BM);
        const uint4 a1 = *reinterpret_cast<const uint4 *>(xs + kk * BM + 4);
        const uint4 b0 = *reinterpret_cast<const uint4 *>(qs + kk * BN);
        const uint32_t av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const uint32_t bv[4] = {b0.x, b0.y, b0.z, b0.w};
        if (I8) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            if (L2) aai[i] = __dp4a((int)av[i], (int)av[i], aai[i]);
#pragma unroll
            for (int j = 0; j < 4; ++j) acci[i][j] = __dp4a((int)av[i], (int)bv[j], acci[i][j]);
          }
          if (L2) {
#pragma unroll
            for (int j = 0; j < 4; ++j) qqi[j] = __dp4a((int)bv[j], (int)bv[j], qqi[j]);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float x = __uint_as_float(av[i]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float y = __uint_as_float(bv[j]);
              if (L2) { const float d = x - y; accf[i][j] = fmaf(d, d, accf[i][j]); }
              else accf[i][j] = fmaf(x, y, accf[i][j]);
            }
          }
        }
      }
      buf ^= 1;
    }

    // ---- epilogue: scores -> gate -> per-query queues ----
    u64 tq[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) tq[j] = tau[tx * 4 + j];
    uint32_t pend = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint64_t row = row0 + ty * 8 + i;
      bool rvalid = row < a.row_end;
      if (rvalid && a.rv.drop) rvalid = !bit_set(a.rv.drop, row);
      if (rvalid && a.filter) rvalid = !bit_set(a.filter, row);
      float4 mt = make_float4(0, 0, 0, 0);
      if (metric_is_qi8(metric) && rvalid) mt = a.rv.tails[row];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float s;
        if (I8) {
          if (metric == ZVB200_METRIC_QI8_L2SQ) s = qi8_l2sq((float)acci[i][j], mt, qtl[j], dq);
          else if (metric_is_qi8(metric)) s = qi8_mip_ord((float)acci[i][j], mt, qtl[j], dq, a.rv.query_first);
          else if (L2) { s = (float)(aai[i] - 2 * acci[i][j] + qqi[j]); if (metric == ZVB200_METRIC_L2) s = sqrtf(s); }
          else s = -(float)acci[i][j];
        } else {
          const float f = accf[i][j];
          s = metric == ZVB200_METRIC_L2SQ ? f : metric == ZVB200_METRIC_L2 ? sqrtf(f)
              : metric == ZVB200_METRIC_IP ? -f : 1.0f - f;
        }
        const u64 key = pack_key(s, (uint32_t)row);
        accf[i][j] = s;  // keep the score; the key is re-packed on demand
        if (rvalid && (tx * 4 + j) < nq && s <= a.threshold && key < tq[j]) pend |= 1u << (i * 4 + j);
      }
    }
    for (;;) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (pend & (1u << (i * 4 + j))) {
            const int ql = tx * 4 + j;
            const int pos = atomicAdd(&qn[ql], 1);
            if (pos < QC) {
              queue[(size_t)ql * QC + pos] = pack_key(accf[i][j], (uint32_t)(row0 + ty * 8 + i));
              pend &= ~(1u << (i * 4 + j));
            }
          }
        }
      }
      const int any = __syncthreads_or(pend != 0);
      for (int ql = warp; ql < BN; ql += kWarps) {
        const int n = min(qn[ql], QC);
        if (n >= 32 || (any && n > 0)) {
          WarpTopk tk;
          u64 *l0 = lists + (size_t)ql * 2 * k;
          const int cu = cur[ql];
          tk.init(l0 + cu * k, l0 + (cu ^ 1) * k, nullptr, k, tau[ql]);
          tk.cnt = cnt[ql];
          for (int c0 = 0; c0 < n; c0 += 32) tk.merge_chunk(queue + (size_t)ql * QC + c0, min(32, n - c0), kMaxKey);
          __syncwarp();
          if (lane == 0) { cnt[ql] = tk.cnt; cur[ql] = (tk.list == l0) ? 0 : 1; tau[ql] = tk.tau; qn[ql] = 0; }
        }
      }
      __syncthreads();
      if (!any) break;
#pragma unroll
      for (int j = 0; j < 4; ++j) tq[j] = tau[tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if ((pend & (1u << (i * 4 + j))) && !(pack_key(accf[i][j], (uint32_t)(row0 + ty * 8 + i)) < tq[j]))
            pend &= ~(1u << (i * 4 + j));
    }
  }

  // ---- final flush and write-out ----
  __syncthreads();
  for (int ql = warp; ql < BN; ql += kWarps) {
    if (ql >= nq) continue;
    const int n = min(qn[ql], QC);
    WarpTopk tk;
    u64 *l0 = lists + (size_t)ql * 2 * k;
    const int cu = cur[ql];
    tk.init(l0 + cu * k, l0 + (cu ^ 1) * k, nullptr, k, tau[ql]);
    tk.cnt = cnt[ql];
    for (int c0 = 0; c0 < n; c0 += 32) tk.merge_chunk(queue + (size_t)ql * QC + c0, min(32, n - c0), kMaxKey);
    __syncwarp();
    u64 *dst = a.part + ((size_t)(q0 + ql) * a.nparts + rs) * k;
    for (int i = lane; i < k; i += 32) dst[i] = i < tk.cnt ? tk.list[i] : kMaxKey;
  }
}

size_t tile_smem(int dtype, uint32_t k) {
  const int KK = dtype == ZVB200_DT_FP16 ? 32 : 16;
  return (size_t)2 * KK * (BM + BN) * 4 + (size_t)BN * (2 * (size_t)k + QC) * sizeof(u64) +
         (size_t)BN * sizeof(u64) + 3 * (size_t)BN * sizeof(int);
}

template <int DT, bool L2>
cudaError_t launch_t(const TileArgs &a, int grid, size_t smem, cudaStream_t st) {
  auto kern = scan_tile_kernel<DT, L2>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess && smem > (size_t)kMaxDynSmem) e = cudaErrorInvalidValue;
  if (e != cudaSuccess) return e;
  kern<<<grid, kThreads, smem, st>>>(a);
  count_launch();
  return cudaGetLastError();
}

}  // namespace

size_t scan_tile_max_parts(int sm_count) { return (size_t)sm_count * 2; }
bool scan_tile_supports(uint32_t k) { return k <= (uint32_t)kMaxK; }

cudaError_t launch_scan_tile(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                             uint64_t row_begin, uint64_t row_end, u64 *part, uint32_t *nparts,
                             int sm_count, cudaStream_t st) {
  if (sp.k > (uint32_t)kMaxK) return cudaErrorInvalidValue;
  TileArgs a;
  a.rv = rv; a.qv = qv; a.k = sp.k; a.threshold = sp.threshold; a.filter = sp.filter;
  a.row_begin = row_begin; a.row_end = row_end; a.part = part;
  a.qtiles = (qv.q + BN - 1) / BN;
  const uint64_t nrows = row_end > row_begin ? row_end - row_begin : 0;
  const uint64_t ntiles = std::max<uint64_t>(1, (nrows + BM - 1) / BM);
  uint64_t rs = ((uint64_t)sm_count * 2 + a.qtiles - 1) / a.qtiles;
  rs = std::max<uint64_t>(1, std::min<uint64_t>(rs, ntiles));
  a.nparts = (uint32_t)rs;
  *nparts = a.nparts;
  const int grid = (int)(rs * a.qtiles);
  const size_t smem = tile_smem(rv.dtype, sp.k);
  const bool l2 = metric_is_l2(rv.metric);
  if (rv.dtype == ZVB200_DT_FP32)
    return l2 ? launch_t<ZVB200_DT_FP32, true>(a, grid, smem, st) : launch_t<ZVB200_DT_FP32, false>(a, grid, smem, st);
  if (rv.dtype == ZVB200_DT_FP16)
    return l2 ? launch_t<ZVB200_DT_FP16, true>(a, grid, smem, st) : launch_t<ZVB200_DT_FP16, false>(a, grid, smem, st);
  return l2 ? launch_t<ZVB200_DT_INT8, true>(a, grid, smem, st) : launch_t<ZVB200_DT_INT8, false>(a, grid, smem, st);
}

}  // namespace zvb

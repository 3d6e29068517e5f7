// K1s — bulk-async streaming scan with fused warp top-k (small query batches, every dtype/metric,
// exact reference arithmetic) and K4 — the IVF posting-list scan built on the same pipeline.
//
// Replaces the per-row loop of FlatStreamerEntity::search_block
// (src/core/algorithm/flat/flat_streamer_entity.cc:256-316) / FlatSearcherContext::search_row_*
// (flat_searcher_context.h:847-921) and, in SEG mode, IVFEntity::search(list, query, ...)
// (src/core/algorithm/ivf/ivf_entity.cc:662-716): one XxxDistanceMatrix<T,1,1>::Compute and one
// heap->emplace per stored row.
//
// HBM-bound by construction, so the design is about bytes in flight: a producer warp keeps a ring
// of row tiles (16-32 KB each, up to ~190 KB per SM) filled with cp.async.bulk copies completing
// on mbarriers — no registers are tied up by loads in flight, unlike the LDG-based K1 (scan_gemv.cu)
// — and the consumer warps compute from shared memory with the same per-chunk arithmetic as K1.
// Stored rows are contiguous (row_stride = padded row bytes), so a tile is ONE 1-D bulk copy; in SEG
// mode a tile is one copy per posting list it touches.
#include <stdlib.h>

#include "chunk.cuh"
#include "internal.h"
#include "topk.cuh"

namespace zvb {

namespace {

constexpr int kU = 4;            // row groups per warp step (ILP on the shared-memory loads)
constexpr int kMaxStages = 8;

struct StreamArgs {
  RowsView rv;
  QueryView qv;
  uint32_t q0;           // query of CTA (x, y) = q0 + y
  uint32_t k;
  float threshold;
  const uint32_t *filter;
  uint64_t row_begin, row_end;
  u64 *part;             // [q][nparts][k]
  uint32_t nparts;
  int lpr_log2;          // lanes per row = 1 << lpr_log2
  uint32_t tile_rows, nstage, cwarps;
  int evict_first;
  const uint32_t *seg_prefix, *seg_row, *seg_count;
  uint32_t seg_cap;
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 1023u) == 0 && clock64() - t0 > 4000000000ll) __trap();  // never hang the GPU
  }
}
// 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar, bool evict_first) {
  if (evict_first) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
        ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(0x12F0000000000000ull)
        : "memory");
  } else {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
  }
}
__device__ __forceinline__ uint4 lds128(const uint8_t *p) { return *reinterpret_cast<const uint4 *>(p); }

template <int DT, bool L2, bool SEG>
__global__ void __launch_bounds__(544, 1) scan_stream_kernel(const StreamArgs a) {
  typedef Chunk<DT, L2> CK;
  typedef typename CK::Acc Acc;
  extern __shared__ __align__(128) uint8_t smem[];
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  const int W = (int)a.cwarps;                   // consumer warps; warp W is the producer
  const uint32_t C = a.rv.chunks, stride = a.rv.row_stride;
  const size_t qbytes = CK::qbytes(C);
  const uint32_t qi = a.q0 + blockIdx.y;
  const int k = a.k;
  const int metric = a.rv.metric;
  const bool qi8 = metric_is_qi8(metric);
  const uint32_t TR = a.tile_rows, NS = a.nstage;
  const uint32_t tile_bytes = TR * stride;

  // ---- shared memory: row stages | tail stages | query | top-k | segment table | barriers ----
  uint8_t *stages = smem;
  float4 *tail_stages = reinterpret_cast<float4 *>(stages + (size_t)NS * tile_bytes);
  uint8_t *qs = reinterpret_cast<uint8_t *>(tail_stages + (qi8 ? (size_t)NS * TR : 0));
  int *qq = reinterpret_cast<int *>(qs + qbytes);
  int2 *meta = reinterpret_cast<int2 *>(qq + 2);
  u64 *tkbase = reinterpret_cast<u64 *>(meta + W);
  const size_t tk_u64 = (size_t)2 * k + 32;
  uint32_t *sprefix = reinterpret_cast<uint32_t *>(tkbase + (size_t)W * tk_u64);
  uint32_t *srow = sprefix + (SEG ? a.seg_cap + 1 : 0);
  uint64_t *bars = reinterpret_cast<uint64_t *>(
      (reinterpret_cast<uintptr_t>(srow + (SEG ? a.seg_cap : 0)) + 7) & ~(uintptr_t)7);
  const uint32_t bar_full = smem_u32(bars), bar_empty = bar_full + 8 * NS;

  uint32_t nseg = 0;
  if (SEG) {
    nseg = a.seg_count[qi];
    for (uint32_t i = threadIdx.x; i <= nseg; i += blockDim.x) sprefix[i] = a.seg_prefix[(size_t)qi * (a.seg_cap + 1) + i];
    for (uint32_t i = threadIdx.x; i < nseg; i += blockDim.x) srow[i] = a.seg_row[(size_t)qi * a.seg_cap + i];
  }
  {  // the query in compute layout
    const uint8_t *src = (DT == ZVB200_DT_INT8) ? a.qv.qraw + (size_t)qi * stride
                                                : reinterpret_cast<const uint8_t *>(a.qv.qf32) + (size_t)qi * qbytes;
    for (size_t o = (size_t)threadIdx.x * 16; o < qbytes; o += (size_t)blockDim.x * 16)
      *reinterpret_cast<uint4 *>(qs + o) = *reinterpret_cast<const uint4 *>(src + o);
  }
  if (threadIdx.x == 0) {
    for (uint32_t s = 0; s < NS; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, (uint32_t)W);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (DT == ZVB200_DT_INT8 && L2 && threadIdx.x == 0) {  // qq = sum q^2 (plain int8 L2)
    int s = 0;
    const int8_t *p = reinterpret_cast<const int8_t *>(qs);
    for (size_t i = 0; i < qbytes; ++i) s += (int)p[i] * (int)p[i];
    qq[0] = s;
  }
  __syncthreads();

  // row space of this CTA: store rows [row_begin,row_end) or the virtual rows of the query's lists
  const uint64_t vbegin = SEG ? 0 : a.row_begin;
  const uint64_t vend = SEG ? (uint64_t)sprefix[nseg] : a.row_end;
  const uint64_t nrows = vend - vbegin;
  const uint64_t ntiles = (nrows + TR - 1) / TR;
  auto seg_of = [&](uint32_t v) -> uint32_t {  // last segment with prefix <= v
    uint32_t lo = 0, hi = nseg;
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (sprefix[mid] <= v) lo = mid; else hi = mid;
    }
    return lo;
  };

  if (warp == W) {
    // ================= producer: keep the ring full =================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (uint64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const uint64_t v0 = vbegin + t * TR;
        const uint32_t n = (uint32_t)min((uint64_t)TR, vend - v0);
        mbar_wait(bar_empty + 8 * stage, phase ^ 1);
        const uint32_t bar = bar_full + 8 * stage;
        mbar_expect_tx(bar, n * stride + (qi8 ? n * 16u : 0u));
        const uint32_t dst = smem_u32(stages + (size_t)stage * tile_bytes);
        const uint32_t tdst = smem_u32(tail_stages + (size_t)stage * TR);
        if (!SEG) {
          bulk_g2s(dst, a.rv.rows + v0 * stride, n * stride, bar, a.evict_first);
          if (qi8) bulk_g2s(tdst, a.rv.tails + v0, n * 16u, bar, a.evict_first);
        } else {
          uint32_t v = (uint32_t)v0, left = n, off = 0;
          uint32_t sg = seg_of(v);
          while (left) {
            const uint32_t run = min(left, sprefix[sg + 1] - v);
            const uint64_t r = (uint64_t)srow[sg] + (v - sprefix[sg]);
            bulk_g2s(dst + off * stride, a.rv.rows + r * stride, run * stride, bar, a.evict_first);
            if (qi8) bulk_g2s(tdst + off * 16u, a.rv.tails + r, run * 16u, bar, a.evict_first);
            v += run; off += run; left -= run; ++sg;
          }
        }
        if (++stage == NS) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ================= consumers: dot products from shared memory + warp top-k =================
    WarpTopk tk;
    {
      u64 *b = tkbase + (size_t)warp * tk_u64;
      tk.init(b, b + k, b + 2 * k, k, kMaxKey);
    }
    const int lpr = 1 << a.lpr_log2, R = 32 >> a.lpr_log2;
    const int l = lane & (lpr - 1), sub = lane >> a.lpr_log2;
    const uint32_t step = (uint32_t)(kU * R);
    const float dq = (float)a.rv.d;
    const float4 qt = qi8 ? a.qv.qtails[qi] : make_float4(0, 0, 0, 0);
    uint32_t stage = 0, phase = 0;
    for (uint64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
      const uint64_t v0 = vbegin + t * TR;
      const uint32_t n = (uint32_t)min((uint64_t)TR, vend - v0);
      mbar_wait(bar_full + 8 * stage, phase);
      const uint8_t *tile = stages + (size_t)stage * tile_bytes;
      const float4 *ttile = tail_stages + (size_t)stage * TR;
      for (uint32_t base = (uint32_t)warp * step; base < n; base += (uint32_t)W * step) {
        Acc acc[kU], aa[kU];
        const uint8_t *rp[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          uint32_t r = base + (uint32_t)(u * R + sub);
          if (r >= n) r = n - 1;  // clamp: computed, result discarded
          rp[u] = tile + (size_t)r * stride;
          acc[u] = 0; aa[u] = 0;
        }
        for (uint32_t c = l; c < C; c += lpr) {
          uint4 v[kU];
#pragma unroll
          for (int u = 0; u < kU; ++u) v[u] = lds128(rp[u] + (size_t)c * 16);
#pragma unroll
          for (int u = 0; u < kU; ++u) CK::fma(acc[u], aa[u], v[u], qs, c);
        }
        for (int o = lpr >> 1; o > 0; o >>= 1) {
#pragma unroll
          for (int u = 0; u < kU; ++u) {
            if (DT == ZVB200_DT_INT8 && L2) aa[u] += __shfl_xor_sync(0xffffffffu, aa[u], o);
            acc[u] += __shfl_xor_sync(0xffffffffu, acc[u], o);
          }
        }
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          const uint32_t rl = base + (uint32_t)(u * R + sub);
          bool valid = (l == 0) && rl < n;
          uint64_t r = v0 + rl;  // global row of the store
          if (SEG && valid) {
            const uint32_t sg = seg_of((uint32_t)r);
            r = (uint64_t)srow[sg] + ((uint32_t)r - sprefix[sg]);
          }
          if (valid && a.rv.drop) valid = !bit_set(a.rv.drop, r);
          if (valid && a.filter) valid = !bit_set(a.filter, r);
          float s;
          if (DT == ZVB200_DT_INT8) {
            const float4 mt = (qi8 && valid) ? ttile[rl] : make_float4(0, 0, 0, 0);
            if (metric == ZVB200_METRIC_QI8_L2SQ) s = qi8_l2sq((float)acc[u], mt, qt, dq);
            else if (qi8) s = qi8_mip_ord((float)acc[u], mt, qt, dq, a.rv.query_first);
            else if (L2) {
              s = (float)((int)aa[u] - 2 * (int)acc[u] + qq[0]);
              if (metric == ZVB200_METRIC_L2) s = sqrtf(s);
            } else s = -(float)acc[u];
          } else {
            const float f = (float)acc[u];
            s = metric == ZVB200_METRIC_L2SQ ? f : metric == ZVB200_METRIC_L2 ? sqrtf(f)
                : metric == ZVB200_METRIC_IP ? -f : 1.0f - f;
          }
          const u64 key = pack_key(s, (uint32_t)r);
          const bool has = valid && s <= a.threshold && key < tk.tau;
          tk.push(has, key, kMaxKey);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_empty + 8 * stage);  // this warp is done reading the stage
      if (++stage == NS) { stage = 0; phase ^= 1; }
    }
    tk.finish(kMaxKey);
    if (lane == 0) meta[warp] = make_int2((int)(tk.list - tkbase), tk.cnt);
    __syncthreads();
    if (warp == 0) {  // merge the warps' lists, write this CTA's partial list
      for (int w = 1; w < W; ++w) {
        const int2 m = meta[w];
        const u64 *src = tkbase + m.x;
        for (int i0 = 0; i0 < m.y; i0 += 32) {
          const int i = i0 + lane;
          const u64 e = i < m.y ? src[i] : kMaxKey;
          tk.push(i < m.y && e < tk.tau, e, kMaxKey);
        }
      }
      tk.finish(kMaxKey);
      u64 *dst = a.part + ((size_t)qi * a.nparts + blockIdx.x) * k;
      for (int i = lane; i < k; i += 32) dst[i] = i < tk.cnt ? tk.list[i] : kMaxKey;
    }
    return;
  }
  __syncthreads();  // producer warp: matches the consumers' barrier before the list merge
}

struct StreamCfg {
  uint32_t tile_rows, nstage, cwarps;
  size_t smem;
};

// tile size, ring depth and consumer warps that fit 227 KB next to the top-k storage
bool stream_config(const RowsView &rv, uint32_t k, uint32_t seg_cap, StreamCfg *cfg) {
  const size_t qbytes = rv.dtype == ZVB200_DT_FP16 ? (size_t)rv.chunks * 32 : (size_t)rv.chunks * 16;
  const bool qi8 = metric_is_qi8(rv.metric);
  const size_t per_row = rv.row_stride + (qi8 ? 16 : 0);
  for (uint32_t w : {16u, 8u, 4u}) {
    const size_t fixed = qbytes + 8 + w * sizeof(int2) + (size_t)w * (2 * (size_t)k + 32) * 8 +
                         (seg_cap ? (2 * (size_t)seg_cap + 1) * 4 : 0) + 8 + 2 * kMaxStages * 8 + 256;
    if (fixed + 2 * 8 * per_row > 225 * 1024) continue;
    const size_t room = 225 * 1024 - fixed;
    // rows per tile: ~32 KB, at least one warp step per consumer warp is not required (ragged is fine)
    uint32_t tr = (uint32_t)std::max<size_t>(8, std::min<size_t>(1024, (32 * 1024) / per_row));
    uint32_t ns = (uint32_t)std::min<size_t>(kMaxStages, room / (tr * per_row));
    while (ns < 3 && tr > 8) { tr = std::max<uint32_t>(8, tr / 2); ns = (uint32_t)std::min<size_t>(kMaxStages, room / (tr * per_row)); }
    if (ns < 2) continue;
    cfg->tile_rows = tr; cfg->nstage = ns; cfg->cwarps = w;
    cfg->smem = fixed + (size_t)ns * tr * per_row;
    return true;
  }
  return false;
}

template <int DT, bool L2, bool SEG>
cudaError_t launch_s(const StreamArgs &a, dim3 grid, size_t smem, cudaStream_t st) {
  auto kern = scan_stream_kernel<DT, L2, SEG>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e == cudaSuccess && smem > (size_t)kMaxDynSmem) e = cudaErrorInvalidValue;
  if (e != cudaSuccess) return e;
  kern<<<grid, (a.cwarps + 1) * 32, smem, st>>>(a);
  count_launch();
  return cudaGetLastError();
}

template <bool SEG>
cudaError_t dispatch(const StreamArgs &a, dim3 grid, size_t smem, cudaStream_t st) {
  const bool l2 = metric_is_l2(a.rv.metric);
  if (a.rv.dtype == ZVB200_DT_FP32)
    return l2 ? launch_s<ZVB200_DT_FP32, true, SEG>(a, grid, smem, st) : launch_s<ZVB200_DT_FP32, false, SEG>(a, grid, smem, st);
  if (a.rv.dtype == ZVB200_DT_FP16)
    return l2 ? launch_s<ZVB200_DT_FP16, true, SEG>(a, grid, smem, st) : launch_s<ZVB200_DT_FP16, false, SEG>(a, grid, smem, st);
  return l2 ? launch_s<ZVB200_DT_INT8, true, SEG>(a, grid, smem, st) : launch_s<ZVB200_DT_INT8, false, SEG>(a, grid, smem, st);
}

void fill_common(StreamArgs &a, const RowsView &rv, const QueryView &qv, const SearchParams &sp, const StreamCfg &cfg) {
  a.rv = rv; a.qv = qv; a.q0 = 0; a.k = sp.k; a.threshold = sp.threshold; a.filter = sp.filter;
  // shared-memory reads: narrower row groups than K1 cost bank conflicts, measured best at >= 8 lanes
  a.lpr_log2 = std::max(pick_lpr_log2(rv.chunks), std::min(3, max_lpr_log2(rv.chunks)));
  a.tile_rows = cfg.tile_rows; a.nstage = cfg.nstage; a.cwarps = cfg.cwarps;
  a.seg_prefix = a.seg_row = a.seg_count = nullptr; a.seg_cap = 0;
}

}  // namespace

bool scan_stream_supports(const RowsView &rv, uint32_t k, uint32_t seg_cap) {
  StreamCfg cfg;
  return stream_config(rv, k, seg_cap, &cfg);
}

size_t scan_stream_max_parts(int sm_count) { return (size_t)sm_count; }

// flat scan: every query of the (small) batch streams the rows once, grid.y = query
cudaError_t launch_scan_stream(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                               uint64_t row_begin, uint64_t row_end, u64 *part, uint32_t *nparts,
                               int sm_count, cudaStream_t st) {
  StreamCfg cfg;
  if (!stream_config(rv, sp.k, 0, &cfg)) return cudaErrorInvalidValue;
  StreamArgs a;
  fill_common(a, rv, qv, sp, cfg);
  a.row_begin = row_begin; a.row_end = row_end; a.part = part;
  const uint64_t nrows = row_end > row_begin ? row_end - row_begin : 0;
  const uint64_t ntiles = (nrows + cfg.tile_rows - 1) / cfg.tile_rows;
  const uint32_t gx = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(ntiles, (uint64_t)sm_count));
  a.nparts = gx; *nparts = gx;
  a.evict_first = qv.q == 1;
  return dispatch<false>(a, dim3(gx, qv.q), cfg.smem, st);
}

// K4 — IVF posting-list scan: grid.y = query, each query walks ITS selected lists
size_t scan_lists_stream_parts(uint32_t q, int sm_count) {
  const uint32_t target = (uint32_t)sm_count * 2;  // CTAs wanted (one resident per SM, two waves)
  return std::max<uint32_t>(1, std::min<uint32_t>((target + q - 1) / q, (uint32_t)sm_count));
}

cudaError_t launch_scan_lists_stream(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                              const uint32_t *seg_prefix, const uint32_t *seg_row, const uint32_t *seg_count,
                              uint32_t seg_cap, u64 *part, uint32_t *nparts, int sm_count, cudaStream_t st) {
  StreamCfg cfg;
  if (!stream_config(rv, sp.k, seg_cap, &cfg)) return cudaErrorInvalidValue;
  StreamArgs a;
  fill_common(a, rv, qv, sp, cfg);
  a.row_begin = 0; a.row_end = rv.n; a.part = part;
  a.seg_prefix = seg_prefix; a.seg_row = seg_row; a.seg_count = seg_count; a.seg_cap = seg_cap;
  const uint32_t gx = (uint32_t)scan_lists_stream_parts(qv.q, sm_count);
  a.nparts = gx; *nparts = gx;
  a.evict_first = 0;
  const uint32_t ymax = 65535;
  for (uint32_t q0 = 0; q0 < qv.q; q0 += ymax) {
    a.q0 = q0;
    cudaError_t e = dispatch<true>(a, dim3(gx, std::min<uint32_t>(ymax, qv.q - q0)), cfg.smem, st);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

}  // namespace zvb

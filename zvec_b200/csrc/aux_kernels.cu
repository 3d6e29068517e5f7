// Small kernels around the scan: query preparation, row repacking on ingest, the final
// per-query merge of partial lists (K5a) and the multi-GPU shard merge (K5b).
#include <cuda_bf16.h>
#include <math_constants.h>

#include "internal.h"
#include "topk.cuh"

namespace zvb {

namespace {

// ---- queries: stored layout -> compute layout ---------------------------------------
// fp32 / fp16: strip the cosine norm slot, widen to fp32, zero-pad to whole 16-B chunks;
// int8: codes zero-padded to row_stride, QuantizedInteger tail split off
// (tail = 4 floats right after the d codes, record_quantizer.h:72-80).
__global__ void prep_queries_kernel(RowsView rv, const uint8_t *qin, uint32_t elem_bytes,
                                    float *qf32, uint8_t *qraw, float4 *qtails, float *qnorms) {
  const uint32_t qi = blockIdx.x;
  const uint8_t *src = qin + (size_t)qi * elem_bytes;
  const uint32_t d = rv.d;
  __shared__ float red[32];
  float nrm = 0.f;
  if (rv.dtype == ZVB200_DT_INT8) {
    for (uint32_t i = threadIdx.x; i < rv.row_stride; i += blockDim.x)
      qraw[(size_t)qi * rv.row_stride + i] = i < d ? src[i] : 0;
    if (metric_is_qi8(rv.metric) && threadIdx.x < 4) {
      float v;
      memcpy(&v, src + d + threadIdx.x * 4, 4);
      reinterpret_cast<float *>(qtails + qi)[threadIdx.x] = v;
    }
  } else {
    const uint32_t esz = rv.dtype == ZVB200_DT_FP32 ? 4 : 2;
    const uint32_t padded = rv.chunks * 16 / esz;
    for (uint32_t i = threadIdx.x; i < padded; i += blockDim.x) {
      float v = 0.f;
      if (i < d) {
        if (esz == 4) memcpy(&v, src + (size_t)i * 4, 4);
        else { __half h; memcpy(&h, src + (size_t)i * 2, 2); v = __half2float(h); }
      }
      qf32[(size_t)qi * padded + i] = v;
      nrm = fmaf(v, v, nrm);
      if (esz == 2 && qraw) {
        uint16_t hv = 0;
        if (i < d) memcpy(&hv, src + (size_t)i * 2, 2);
        reinterpret_cast<uint16_t *>(qraw + (size_t)qi * rv.row_stride)[i] = hv;
      }
    }
  }
  // |q|^2
  for (int o = 16; o > 0; o >>= 1) nrm += __shfl_xor_sync(0xffffffffu, nrm, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = nrm;
  __syncthreads();
  if (threadIdx.x == 0 && qnorms) {
    float s = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
    qnorms[qi] = s;
  }
}

// ---- ingest: reference element layout -> store layout --------------------------------
__global__ void repack_rows_kernel(const uint8_t *src, uint32_t elem_bytes, uint64_t n,
                                   uint8_t *dst, uint32_t row_stride, uint32_t vec_bytes,
                                   float4 *tails, uint32_t tail_off) {
  const uint32_t cpr = row_stride / 16;
  const uint64_t total = n * cpr;
  for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t r = t / cpr;
    const uint32_t c = (uint32_t)(t % cpr);
    const uint8_t *s = src + r * elem_bytes + (size_t)c * 16;
    uint4 v = make_uint4(0, 0, 0, 0);
    const uint32_t off = c * 16;
    if (off + 16 <= vec_bytes && (elem_bytes & 15) == 0) {
      v = *reinterpret_cast<const uint4 *>(s);
    } else {
      uint8_t *b = reinterpret_cast<uint8_t *>(&v);
      for (int i = 0; i < 16; ++i)
        if (off + i < vec_bytes) b[i] = s[i];
    }
    *reinterpret_cast<uint4 *>(dst + r * row_stride + (size_t)c * 16) = v;
    if (tails && c == 0) {
      float4 tv;
      memcpy(&tv, src + r * elem_bytes + tail_off, 16);
      tails[r] = tv;
    }
  }
}

// |x|^2 per stored row (fp32 / fp16), one warp per row
__global__ void row_norms_kernel(RowsView rv, uint64_t row_begin, uint64_t row_end, float *norms) {
  const uint64_t r = row_begin + ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) / 32;
  if (r >= row_end) return;
  const int lane = threadIdx.x & 31;
  const uint8_t *p = rv.rows + r * rv.row_stride;
  float s = 0.f;
  for (uint32_t c = lane; c < rv.chunks; c += 32) {
    const uint4 v = *reinterpret_cast<const uint4 *>(p + (size_t)c * 16);
    if (rv.dtype == ZVB200_DT_INT8) {  // exact: the sum stays below 2^24 for d <= 1040
      int t = __dp4a((int)v.x, (int)v.x, 0);
      t = __dp4a((int)v.y, (int)v.y, t); t = __dp4a((int)v.z, (int)v.z, t); t = __dp4a((int)v.w, (int)v.w, t);
      s += (float)t;
    } else if (rv.dtype == ZVB200_DT_FP32) {
      const float x0 = __uint_as_float(v.x), x1 = __uint_as_float(v.y), x2 = __uint_as_float(v.z),
                  x3 = __uint_as_float(v.w);
      s = fmaf(x0, x0, s); s = fmaf(x1, x1, s); s = fmaf(x2, x2, s); s = fmaf(x3, x3, s);
    } else {
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&w[i]));
        s = fmaf(f.x, f.x, s); s = fmaf(f.y, f.y, s);
      }
    }
  }
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) norms[r] = s;
}

// ---- QuantizedInteger L2 on tensor cores: the "shadow" operands ------------------------------------
// The score (quantized_integer_metric_matrix.h:36-60) in real arithmetic is
//   s = R_r + Q0_q - [ ma*nQa*ip + mb*nQ1 + (ms*ma)*nQ2 ],   R_r = ma^2 ms2 + mb^2 d + 2 mb ms ma,
//   nQa = 2 qa, nQ1 = 2 (qb d + qa qs), nQ2 = 2 qb,  Q0_q = qa^2 qs2 + qb^2 d + 2 qb qa qs
// so (R_r - s) / nQa = sum_i (ma c_i) qc_i + mb (nQ1/nQa) + (ms ma)(nQ2/nQa) is BILINEAR in (row, query):
// one contraction over d + 9 bf16 columns yields g = (Q0 - s) / nQa up to the margin qi8_shadow_eps, and the
// epilogue compares g with a per-query register — no per-column constants (which cost the int8 kernel four
// shared-memory wavefronts per column).  Row side: bf16(ma c_i), then {u_hi, u_hi, u_lo} for u = -(R_r - margin),
// mb, ms*ma; query side: the codes (exact in bf16), then {v_hi, v_lo, v_hi} for v = 1/nQa, nQ1/nQa, nQ2/nQa.
// The filter only nominates; candidates get the exact reference-order score from the int8 codes.
__device__ __forceinline__ void bf16_split(float x, __nv_bfloat16 *hi, __nv_bfloat16 *lo) {
  *hi = __float2bfloat16_rn(x);
  *lo = __float2bfloat16_rn(x - __bfloat162float(*hi));
}

__global__ void qi8_shadow_rows_kernel(RowsView rv, uint64_t row_begin, uint64_t row_end, uint8_t *shadow,
                                       uint32_t shadow_k, float *mrow_max) {
  const uint32_t cpr = shadow_k / 8;  // 16-byte output chunks per row
  const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t r = row_begin + idx / cpr;
  if (r >= row_end) return;
  const uint32_t c = (uint32_t)(idx % cpr);
  const int8_t *codes = reinterpret_cast<const int8_t *>(rv.rows + r * rv.row_stride);
  const float4 m = rv.tails[r];
  const uint32_t d = rv.d;
  __align__(16) __nv_bfloat16 o[8];
  __nv_bfloat16 ex[9];
  bool have_ex = false;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t i = c * 8 + j;
    if (i < d) {
      o[j] = __float2bfloat16_rn(m.x * (float)codes[i]);
    } else if (i < d + 9) {
      if (!have_ex) {
        const float dq = (float)d;
        const float t1 = m.x * m.x * m.w, mbd = m.y * m.y * dq, msma = m.z * m.x;
        const float c0 = t1 + mbd + 2.0f * m.y * msma;
        const float mrow = qi8_shadow_eps(d) * (t1 + mbd);
        const float u1 = -(c0 - mrow);
        // the largest margin of any row: non-negative floats order like their bit patterns
        if (i == d && mrow == mrow) atomicMax(reinterpret_cast<int *>(mrow_max), __float_as_int(mrow));
        bf16_split(u1, &ex[0], &ex[2]); ex[1] = ex[0];
        bf16_split(m.y, &ex[3], &ex[5]); ex[4] = ex[3];
        bf16_split(msma, &ex[6], &ex[8]); ex[7] = ex[6];
        have_ex = true;
      }
      __nv_bfloat16 v = ex[0];
#pragma unroll
      for (int e = 1; e < 9; ++e) v = (i - d) == (uint32_t)e ? ex[e] : v;
      o[j] = v;
    } else {
      o[j] = __float2bfloat16_rn(0.f);
    }
  }
  *reinterpret_cast<uint4 *>(shadow + (r * shadow_k + (size_t)c * 8) * 2) = *reinterpret_cast<const uint4 *>(o);
}

__global__ void qi8_shadow_queries_kernel(const uint8_t *qraw, uint32_t row_stride, const float4 *qtails, uint32_t q,
                                          uint32_t d, uint8_t *out, uint32_t shadow_k) {
  const uint32_t cpr = shadow_k / 8;
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t qi = idx / cpr, c = idx % cpr;
  if (qi >= q) return;
  const int8_t *codes = reinterpret_cast<const int8_t *>(qraw + (size_t)qi * row_stride);
  const float4 t = qtails[qi];
  __align__(16) __nv_bfloat16 o[8];
  __nv_bfloat16 ex[9];
  bool have_ex = false;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t i = c * 8 + j;
    if (i < d) {
      o[j] = __float2bfloat16_rn((float)codes[i]);  // |code| <= 128: exact
    } else if (i < d + 9) {
      if (!have_ex) {
        const float dq = (float)d;
        const float nQa = 2.0f * t.x, nQ1 = 2.0f * (t.y * dq + t.x * t.z), nQ2 = 2.0f * t.y;
        const float inv = 1.0f / nQa;
        bf16_split(inv, &ex[0], &ex[1]); ex[2] = ex[0];
        bf16_split(nQ1 * inv, &ex[3], &ex[4]); ex[5] = ex[3];
        bf16_split(nQ2 * inv, &ex[6], &ex[7]); ex[8] = ex[6];
        have_ex = true;
      }
      __nv_bfloat16 v = ex[0];
#pragma unroll
      for (int e = 1; e < 9; ++e) v = (i - d) == (uint32_t)e ? ex[e] : v;
      o[j] = v;
    } else {
      o[j] = __float2bfloat16_rn(0.f);
    }
  }
  *reinterpret_cast<uint4 *>(out + ((size_t)qi * shadow_k + (size_t)c * 8) * 2) = *reinterpret_cast<const uint4 *>(o);
}

// ---- K5a: part[q][nparts][k] -> final q*k ---------------------------------------------
constexpr int kMergeWarps = 8;

// SLICED: blockIdx.y merges `per_cta` consecutive partial lists of its query and writes the k best keys back over the
// first of them (in place: a CTA only touches its own slice); the final pass then reads one list per slice
// (`stride` = per_cta lists apart).  A single CTA per query took 57 us for the 256 x 64 keys of an IVF coarse search.
template <bool SLICED>
__global__ void __launch_bounds__(kMergeWarps * 32) merge_parts_kernel(
    u64 *part, uint32_t nparts, uint32_t ninputs, uint32_t stride, uint32_t per_cta, uint32_t k, float *out_scores,
    uint32_t *out_index, u64 *out_keys) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int lane = lane_id(), warp = threadIdx.x >> 5;
  const uint32_t qi = blockIdx.x;
  const uint32_t in0 = SLICED ? blockIdx.y * per_cta : 0;
  const uint32_t in1 = SLICED ? min(ninputs, in0 + per_cta) : ninputs;
  const size_t total = in1 > in0 ? (size_t)(in1 - in0) * k : 0;
  u64 *src = part + ((size_t)qi * nparts + (size_t)in0 * stride) * k;
  int2 *meta = reinterpret_cast<int2 *>(smem);
  u64 *tkbase = reinterpret_cast<u64 *>(smem + sizeof(int2) * kMergeWarps);
  const size_t tk_u64 = (size_t)2 * k + 32;
  WarpTopk tk;
  u64 *b = tkbase + (size_t)warp * tk_u64;
  tk.init(b, b + k, b + 2 * k, (int)k, kMaxKey);
  for (size_t i0 = (size_t)warp * 32; i0 < total; i0 += (size_t)kMergeWarps * 32) {
    const size_t i = i0 + lane;
    u64 e = kMaxKey;
    if (i < total) {
      if (stride == 1) e = src[i];
      else { const size_t li = i / k; e = src[li * stride * k + (i - li * k)]; }
    }
    tk.push(e < tk.tau, e, kMaxKey);
  }
  tk.finish(kMaxKey);
  if (lane == 0) meta[warp] = make_int2((int)(tk.list - tkbase), tk.cnt);
  __syncthreads();
  if (warp != 0) return;
  for (int w = 1; w < kMergeWarps; ++w) {
    const int2 m = meta[w];
    const u64 *s = tkbase + m.x;
    for (int i0 = 0; i0 < m.y; i0 += 32) {
      const int i = i0 + lane;
      const u64 e = i < m.y ? s[i] : kMaxKey;
      tk.push(i < m.y && e < tk.tau, e, kMaxKey);
    }
  }
  tk.finish(kMaxKey);
  for (uint32_t i = lane; i < k; i += 32) {
    const bool ok = (int)i < tk.cnt;
    const u64 e = ok ? tk.list[i] : kMaxKey;
    if (SLICED) {
      if (total) src[i] = e;
    } else {
      if (out_scores) out_scores[(size_t)qi * k + i] = ok ? unorder_f32((uint32_t)(e >> 32)) : CUDART_NAN_F;
      if (out_index) out_index[(size_t)qi * k + i] = ok ? (uint32_t)e : ZVB200_INVALID_INDEX;
      if (out_keys) out_keys[(size_t)qi * k + i] = e;  // packed (score, row): what the multi-GPU exchange ships
    }
  }
}

// K5a', few queries with MANY partial keys (one query's IVF coarse search leaves 256 lists x 64 keys; the two-pass
// merge above took 28 us of a 160-us call): one CTA per query finds the k-th smallest key by radix select — six
// 11-bit digits of the u64 key, each thread's keys held in registers across the passes — compacts the keys at or
// below it and sorts those k.  Same result as merge_parts_kernel: ascending by (score, row), padding last.
constexpr int kSelPerThread = 16;  // keys a thread keeps in registers
constexpr int kSelMaxThreads = 1024;  // up to 16384 keys per query
}  // namespace
__device__ void ivf_plan_warp(const uint32_t *ci, uint32_t qi, uint32_t nprobe, const uint32_t *list_off,
                              const uint32_t *budget_sizes, uint64_t max_scan_count, uint32_t *seg_prefix,
                              uint32_t *seg_row, uint32_t *seg_count, uint32_t *seg_list);
namespace {

template <int NT>
__global__ void __launch_bounds__(NT) merge_select_kernel(const u64 *part, uint32_t total, uint32_t k,
                                                          float *out_scores, uint32_t *out_index, u64 *out_keys,
                                                          const IvfPlanArgs plan = IvfPlanArgs()) {
  constexpr int BPT = 2048 / NT;  // histogram bins per thread in the scan
  __shared__ uint32_t hist[2048];
  __shared__ uint32_t wsum[32];
  __shared__ uint32_t s_bin, s_need, s_all, s_cnt;
  __shared__ u64 sel[1024];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const u64 *keys = part + (size_t)blockIdx.x * total;
  u64 mine[kSelPerThread];
#pragma unroll
  for (int i = 0; i < kSelPerThread; ++i) {
    const uint32_t idx = tid + i * NT;
    mine[i] = idx < total ? keys[idx] : kMaxKey;
  }
  // the k-th smallest key, digit by digit from the top; stops as soon as the digit's bin is needed WHOLE (with
  // distinct scores that is after the 33 score bits at the latest: three passes, not six)
  u64 prefix = 0;
  uint32_t need = k;
  bool whole = false;
  int sel_shift = 0;  // the result = the keys whose bits above sel_shift are <= prefix's
  for (int shift = 53; shift >= -2; shift -= 11) {
    const int sh = shift < 0 ? 0 : shift;
    const int bits = shift < 0 ? 9 : 11;  // 64 = 5 * 11 + 9
    const u64 himask = bits + sh >= 64 ? 0ull : ~0ull << (bits + sh);
    for (int b = tid; b < 2048; b += NT) hist[b] = 0;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kSelPerThread; ++i)
      if ((mine[i] & himask) == prefix) atomicAdd(&hist[(uint32_t)(mine[i] >> sh) & ((1u << bits) - 1u)], 1u);
    __syncthreads();
    uint32_t h[BPT], x = 0;
#pragma unroll
    for (int b = 0; b < BPT; ++b) { h[b] = hist[BPT * tid + b]; x += h[b]; }
    const uint32_t own = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    if (warp == 0) {
      uint32_t w = lane < NT / 32 ? wsum[lane] : 0u;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
      }
      wsum[lane] = w;
    }
    __syncthreads();
    uint32_t before = (warp ? wsum[warp - 1] : 0u) + x - own;  // keys in the bins below this thread's
#pragma unroll
    for (int b = 0; b < BPT; ++b) {
      if (before < need && need <= before + h[b]) { s_bin = BPT * tid + b; s_need = need - before; s_all = h[b]; }
      before += h[b];
    }
    __syncthreads();
    prefix |= (u64)s_bin << sh;
    need = s_need;
    sel_shift = sh;
    whole = s_all == s_need;
    __syncthreads();
    if (whole) break;
  }
  // whole: the bins below plus this bin are exactly k keys.  Otherwise the loop ran out of digits: the bin is ONE key
  // value held more often than needed — kMaxKey padding when a query has fewer than k candidates (a tight threshold:
  // thousands of copies), or a candidate row listed several times — and only `need` copies of it belong to the result.
  // Padding is never selected: taking "everything <= the k-th" would push real keys out of the 1024 slots.
  if (tid == 0) { s_cnt = 0; s_all = 0; }
  __syncthreads();
  const u64 lim = prefix >> sel_shift;
#pragma unroll
  for (int i = 0; i < kSelPerThread; ++i) {
    if (mine[i] == kMaxKey) continue;
    const u64 tkey = mine[i] >> sel_shift;
    if (tkey < lim || (tkey == lim && whole)) {
      const uint32_t pos = atomicAdd(&s_cnt, 1u);
      if (pos < 1024) sel[pos] = mine[i];
    } else if (tkey == lim) {  // a copy of the k-th key itself: at most `need` of them, behind the smaller keys
      const uint32_t cpy = atomicAdd(&s_all, 1u);
      if (cpy < need) sel[k - need + cpy] = mine[i];
    }
  }
  __syncthreads();
  const uint32_t got = whole ? min(s_cnt, 1024u) : min(k, (k - need) + min(s_all, need));
  uint32_t p2 = 1;
  while (p2 < max(got, k)) p2 <<= 1;
  for (uint32_t i = got + tid; i < p2; i += NT) sel[i] = kMaxKey;
  __syncthreads();
  for (uint32_t size = 2; size <= p2; size <<= 1) {
    for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
      for (uint32_t t = tid; t < (p2 >> 1); t += NT) {
        const uint32_t lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
        const bool up = (lo & size) == 0;
        const u64 a = sel[lo], b = sel[hi];
        if ((a > b) == up) { sel[lo] = b; sel[hi] = a; }
      }
      __syncthreads();
    }
  }
  for (uint32_t i = tid; i < k; i += NT) {
    const u64 e = sel[i];
    const bool ok = e != kMaxKey;
    if (out_scores) out_scores[(size_t)blockIdx.x * k + i] = ok ? unorder_f32((uint32_t)(e >> 32)) : CUDART_NAN_F;
    if (out_index) out_index[(size_t)blockIdx.x * k + i] = ok ? (uint32_t)e : ZVB200_INVALID_INDEX;
    if (out_keys) out_keys[(size_t)blockIdx.x * k + i] = e;
    if (plan.list_off) hist[i] = ok ? (uint32_t)e : ZVB200_INVALID_INDEX;  // k <= 1024 ids in the 2048-word histogram
  }
  // the selected keys are a query's coarse ids: its posting-list plan right here, one launch less (IVF at batch 1)
  if (plan.list_off) {
    __syncthreads();
    if (warp == 0)
      ivf_plan_warp(hist, blockIdx.x, k, plan.list_off, plan.budget_sizes, plan.max_scan_count, plan.seg_prefix, plan.seg_row,
                    plan.seg_count, plan.seg_list);
  }
}

// ---- K5b: G sorted shard lists -> one list by (score, global row) ---------------------
// Replaces concat + std::sort + resize(topk) of CombinedVectorColumnIndexer::Search
// (combined_vector_column_indexer.cc:167-222).  Rank-based: candidate (g,i) lands at
// i + sum_{g'<g} #{s' <= s} + sum_{g'>g} #{s' < s}; shards are contiguous row ranges in
// shard order, so (score, g, i) order == (score, global row) order.
__device__ __forceinline__ uint32_t ord_or_max(float s) {
  return (s != s) ? 0xffffffffu : order_f32(s);
}

__global__ void merge_shards_kernel(const float *scores, const uint32_t *index, uint32_t g,
                                    const uint64_t *shard_base, uint32_t q, uint32_t k,
                                    float *out_scores, uint64_t *out_rows) {
  const uint32_t qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (qi >= q) return;
  const int lane = threadIdx.x & 31;
  for (uint32_t i = lane; i < k; i += 32) {
    out_scores[(size_t)qi * k + i] = CUDART_NAN_F;
    out_rows[(size_t)qi * k + i] = ZVB200_INVALID_KEY;
  }
  __syncwarp();
  const size_t gstride = (size_t)q * k;
  for (uint32_t c = lane; c < g * k; c += 32) {
    const uint32_t gi = c / k, i = c % k;
    const float s = scores[gi * gstride + (size_t)qi * k + i];
    const uint32_t idx = index[gi * gstride + (size_t)qi * k + i];
    if (idx == ZVB200_INVALID_INDEX || s != s) continue;
    const uint32_t so = order_f32(s);
    uint32_t rank = i;
    for (uint32_t g2 = 0; g2 < g; ++g2) {
      if (g2 == gi) continue;
      const float *l2 = scores + g2 * gstride + (size_t)qi * k;
      uint32_t lo = 0, hi = k;
      while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        const uint32_t mo = ord_or_max(l2[mid]);
        const bool before = g2 < gi ? (mo <= so) : (mo < so);
        if (before) lo = mid + 1; else hi = mid;
      }
      rank += lo;
    }
    if (rank < k) {
      out_scores[(size_t)qi * k + rank] = s;
      out_rows[(size_t)qi * k + rank] = shard_base[gi] + idx;
    }
  }
}

// K5c: the same merge for lists that carry explicit rows of the whole index (sharded IVF: a shard's local rows
// map to global rows list by list).  Every list is ascending by (score, row); rank by that pair.
__global__ void merge_rows_kernel(const float *scores, const uint64_t *rows, uint32_t g, uint32_t q, uint32_t k,
                                  float *out_scores, uint64_t *out_rows) {
  const uint32_t qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (qi >= q) return;
  const int lane = threadIdx.x & 31;
  for (uint32_t i = lane; i < k; i += 32) {
    out_scores[(size_t)qi * k + i] = CUDART_NAN_F;
    out_rows[(size_t)qi * k + i] = ZVB200_INVALID_KEY;
  }
  __syncwarp();
  const size_t gstride = (size_t)q * k;
  for (uint32_t c = lane; c < g * k; c += 32) {
    const uint32_t gi = c / k, i = c % k;
    const float s = scores[gi * gstride + (size_t)qi * k + i];
    const uint64_t row = rows[gi * gstride + (size_t)qi * k + i];
    if (row == ZVB200_INVALID_KEY || s != s) continue;
    const uint32_t so = order_f32(s);
    uint32_t rank = i;
    for (uint32_t g2 = 0; g2 < g; ++g2) {
      if (g2 == gi) continue;
      const float *l2 = scores + g2 * gstride + (size_t)qi * k;
      const uint64_t *r2 = rows + g2 * gstride + (size_t)qi * k;
      uint32_t lo = 0, hi = k;
      while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        const uint32_t mo = r2[mid] == ZVB200_INVALID_KEY ? 0xffffffffu : ord_or_max(l2[mid]);
        const bool before = mo < so || (mo == so && r2[mid] < row);
        if (before) lo = mid + 1; else hi = mid;
      }
      rank += lo;
    }
    if (rank < k) {
      out_scores[(size_t)qi * k + rank] = s;
      out_rows[(size_t)qi * k + rank] = row;
    }
  }
}

// local row of a list-grouped store -> row of the whole index: gbase[list] + offset inside the list's local slice
__global__ void ivf_global_rows_kernel(const uint32_t *index, uint64_t count, const uint32_t *list_off,
                                       const uint64_t *gbase, uint32_t nlist, uint64_t *out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const uint32_t r = index[i];
  if (r == ZVB200_INVALID_INDEX || r >= list_off[nlist]) { out[i] = ZVB200_INVALID_KEY; return; }
  uint32_t lo = 0, hi = nlist;  // last list with list_off <= r (empty lists share an offset: the LAST of them holds r)
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (list_off[mid] <= r) lo = mid; else hi = mid;
  }
  out[i] = gbase[lo] + (r - list_off[lo]);
}

// ---- distance matrix: out[j*n + i] = dist(row i, query j) ----------------------------
// The ailego <T,1,1>::Compute contract (IndexMetric::distance(), euclidean_metric.cc:785-851)
// one warp per row, queries looped.
__global__ void distance_matrix_kernel(RowsView rv, QueryView qv, float *out) {
  const uint64_t r = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) / 32;
  if (r >= rv.n) return;
  const int lane = threadIdx.x & 31;
  const uint8_t *p = rv.rows + r * rv.row_stride;
  const bool l2 = metric_is_l2(rv.metric);
  for (uint32_t j = 0; j < qv.q; ++j) {
    float fs = 0.f;
    int ab = 0, aa = 0, qq = 0;
    for (uint32_t c = lane; c < rv.chunks; c += 32) {
      const uint4 v = *reinterpret_cast<const uint4 *>(p + (size_t)c * 16);
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
      if (rv.dtype == ZVB200_DT_INT8) {
        const uint4 qv4 = *reinterpret_cast<const uint4 *>(qv.qraw + (size_t)j * rv.row_stride + (size_t)c * 16);
        const uint32_t qw[4] = {qv4.x, qv4.y, qv4.z, qv4.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          ab = __dp4a((int)w[i], (int)qw[i], ab);
          aa = __dp4a((int)w[i], (int)w[i], aa);
          qq = __dp4a((int)qw[i], (int)qw[i], qq);
        }
      } else if (rv.dtype == ZVB200_DT_FP32) {
        const float *qp = qv.qf32 + (size_t)j * rv.chunks * 4 + (size_t)c * 4;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float x = __uint_as_float(w[i]);
          if (l2) { const float d = x - qp[i]; fs = fmaf(d, d, fs); } else fs = fmaf(x, qp[i], fs);
        }
      } else {
        const float *qp = qv.qf32 + (size_t)j * rv.chunks * 8 + (size_t)c * 8;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&w[i]));
          if (l2) {
            float d = f.x - qp[2 * i]; fs = fmaf(d, d, fs);
            d = f.y - qp[2 * i + 1]; fs = fmaf(d, d, fs);
          } else { fs = fmaf(f.x, qp[2 * i], fs); fs = fmaf(f.y, qp[2 * i + 1], fs); }
        }
      }
    }
    for (int o = 16; o > 0; o >>= 1) {
      fs += __shfl_xor_sync(0xffffffffu, fs, o);
      ab += __shfl_xor_sync(0xffffffffu, ab, o);
      aa += __shfl_xor_sync(0xffffffffu, aa, o);
      qq += __shfl_xor_sync(0xffffffffu, qq, o);
    }
    if (lane == 0) {
      float s;
      const int m = rv.metric;
      if (rv.dtype == ZVB200_DT_INT8) {
        if (m == ZVB200_METRIC_QI8_L2SQ) s = qi8_l2sq((float)ab, rv.tails[r], qv.qtails[j], (float)rv.d);
        else if (metric_is_qi8(m)) s = qi8_mip_ord((float)ab, rv.tails[r], qv.qtails[j], (float)rv.d, rv.query_first);
        else if (l2) { s = (float)(aa - 2 * ab + qq); if (m == ZVB200_METRIC_L2) s = sqrtf(s); }
        else s = -(float)ab;
      } else {
        s = m == ZVB200_METRIC_L2SQ ? fs : m == ZVB200_METRIC_L2 ? sqrtf(fs)
            : m == ZVB200_METRIC_IP ? -fs : 1.0f - fs;
      }
      out[(size_t)j * rv.n + r] = s;
    }
  }
}

}  // namespace

cudaError_t launch_prep_queries(const RowsView &rv, const void *queries_dev, uint32_t q,
                                uint32_t elem_bytes, float *qf32, uint8_t *qraw, float4 *qtails,
                                float *qnorms, cudaStream_t st) {
  if (q == 0) return cudaSuccess;
  // one element per thread and pass: a query of 1536 floats on 128 threads was 12 dependent load round trips (7.5 us
  // of a 140-us IVF call at batch 1)
  const uint32_t elems = rv.dtype == ZVB200_DT_INT8 ? rv.row_stride : rv.chunks * 16 / (rv.dtype == ZVB200_DT_FP32 ? 4 : 2);
  const unsigned threads = elems > 512 ? 1024 : elems > 128 ? 512 : 128;
  prep_queries_kernel<<<q, threads, 0, st>>>(rv, (const uint8_t *)queries_dev, elem_bytes, qf32, qraw,
                                            qtails, qnorms);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_repack_rows(const uint8_t *src, uint32_t elem_bytes, uint64_t n, uint8_t *dst,
                               uint32_t row_stride, uint32_t vec_bytes, float4 *tails,
                               uint32_t tail_off, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  const uint64_t total = n * (row_stride / 16);
  const int grid = (int)std::min<uint64_t>((total + 255) / 256, 148 * 16);
  repack_rows_kernel<<<grid, 256, 0, st>>>(src, elem_bytes, n, dst, row_stride, vec_bytes, tails, tail_off);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_row_norms(const RowsView &rv, uint64_t row_begin, uint64_t row_end,
                             float *norms, cudaStream_t st) {
  if (row_end <= row_begin) return cudaSuccess;
  const uint64_t rows = row_end - row_begin;
  const uint64_t blocks = (rows + 7) / 8;
  row_norms_kernel<<<(unsigned)blocks, 256, 0, st>>>(rv, row_begin, row_end, norms);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_qi8_shadow_rows(const RowsView &rv, uint64_t row_begin, uint64_t row_end, uint8_t *shadow,
                                   uint32_t shadow_k, float *mrow_max, cudaStream_t st) {
  if (row_end <= row_begin) return cudaSuccess;
  const uint64_t cpr = shadow_k / 8;
  const uint64_t per_launch = ((uint64_t)1 << 31) / 256 * 256 / cpr;  // rows per launch: grid.x stays below 2^31
  for (uint64_t r0 = row_begin; r0 < row_end; r0 += per_launch) {
    const uint64_t r1 = std::min(row_end, r0 + per_launch);
    const uint64_t threads = (r1 - r0) * cpr;
    qi8_shadow_rows_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(rv, r0, r1, shadow, shadow_k, mrow_max);
    count_launch();
  }
  return cudaGetLastError();
}

cudaError_t launch_qi8_shadow_queries(const RowsView &rv, const uint8_t *qraw, const float4 *qtails, uint32_t q,
                                      uint8_t *out, uint32_t shadow_k, cudaStream_t st) {
  if (q == 0) return cudaSuccess;
  const uint32_t threads = q * (shadow_k / 8);
  qi8_shadow_queries_kernel<<<(threads + 255) / 256, 256, 0, st>>>(qraw, rv.row_stride, qtails, q, rv.d, out, shadow_k);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_merge_parts(const u64 *part, uint32_t nparts, uint32_t q, uint32_t k,
                               float *out_scores, uint32_t *out_index, u64 *out_keys, cudaStream_t st) {
  if (q == 0) return cudaSuccess;
  const size_t smem = sizeof(int2) * kMergeWarps + (size_t)kMergeWarps * (2 * (size_t)k + 32) * sizeof(u64);
  {  // a constant, every time: the attribute is per function AND per device, and shared by all host threads
    cudaError_t e = cudaFuncSetAttribute(merge_parts_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(merge_parts_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
  }
  u64 *p = const_cast<u64 *>(part);  // the sliced pass reuses the scratch lists in place
  // few queries, up to 16384 keys each: radix select + sort of the k survivors, one CTA per query
  {
    static const bool off = [] { const char *e = getenv("ZVB200_MERGE_SELECT"); return e && atoi(e) == 0; }();
    const size_t total = (size_t)nparts * k;
    if (!off && q <= 64 && k <= 1024 && total >= 1024 && total <= (size_t)kSelMaxThreads * kSelPerThread) {
      if (total <= 256 * kSelPerThread)
        merge_select_kernel<256><<<q, 256, 0, st>>>(part, (uint32_t)total, k, out_scores, out_index, out_keys);
      else if (total <= 512 * kSelPerThread)
        merge_select_kernel<512><<<q, 512, 0, st>>>(part, (uint32_t)total, k, out_scores, out_index, out_keys);
      else
        merge_select_kernel<1024><<<q, 1024, 0, st>>>(part, (uint32_t)total, k, out_scores, out_index, out_keys);
      count_launch();
      return cudaGetLastError();
    }
  }
  // few queries, many keys each: slices of the lists first (grid.y), then one list per slice
  if ((size_t)nparts * k >= 4096 && q <= 64 && nparts >= 32) {
    const uint32_t per_cta = 16, slices = (nparts + per_cta - 1) / per_cta;
    merge_parts_kernel<true><<<dim3(q, slices), kMergeWarps * 32, smem, st>>>(p, nparts, nparts, 1, per_cta, k, nullptr, nullptr, nullptr);
    count_launch();
    merge_parts_kernel<false><<<q, kMergeWarps * 32, smem, st>>>(p, nparts, slices, per_cta, 0, k, out_scores, out_index, out_keys);
    count_launch();
    return cudaGetLastError();
  }
  merge_parts_kernel<false><<<q, kMergeWarps * 32, smem, st>>>(p, nparts, nparts, 1, 0, k, out_scores, out_index, out_keys);
  count_launch();
  return cudaGetLastError();
}

uint32_t select_topk_max() { return (uint32_t)kSelMaxThreads * kSelPerThread; }

cudaError_t launch_select_topk(const u64 *keys, uint32_t total, uint32_t q, uint32_t k, float *out_scores,
                               uint32_t *out_index, cudaStream_t st, const IvfPlanArgs *plan) {
  if (q == 0) return cudaSuccess;
  if (total == 0 || total > select_topk_max() || k == 0 || k > 1024) return cudaErrorInvalidValue;
  const IvfPlanArgs pa = plan ? *plan : IvfPlanArgs();
  if (total <= 256 * kSelPerThread)
    merge_select_kernel<256><<<q, 256, 0, st>>>(keys, total, k, out_scores, out_index, nullptr, pa);
  else if (total <= 512 * kSelPerThread)
    merge_select_kernel<512><<<q, 512, 0, st>>>(keys, total, k, out_scores, out_index, nullptr, pa);
  else
    merge_select_kernel<1024><<<q, 1024, 0, st>>>(keys, total, k, out_scores, out_index, nullptr, pa);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_merge_shards(const float *scores, const uint32_t *index, uint32_t g,
                                const uint64_t *shard_base_dev, uint32_t q, uint32_t k,
                                float *out_scores, uint64_t *out_rows, cudaStream_t st) {
  if (q == 0) return cudaSuccess;
  const int wpb = 4;
  merge_shards_kernel<<<(q + wpb - 1) / wpb, wpb * 32, 0, st>>>(scores, index, g, shard_base_dev, q,
                                                               k, out_scores, out_rows);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_merge_rows(const float *scores, const uint64_t *rows, uint32_t g, uint32_t q, uint32_t k,
                              float *out_scores, uint64_t *out_rows, cudaStream_t st) {
  if (q == 0) return cudaSuccess;
  const int wpb = 4;
  merge_rows_kernel<<<(q + wpb - 1) / wpb, wpb * 32, 0, st>>>(scores, rows, g, q, k, out_scores, out_rows);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_ivf_global_rows(const uint32_t *index, uint64_t count, const uint32_t *list_off,
                                   const uint64_t *gbase, uint32_t nlist, uint64_t *out, cudaStream_t st) {
  if (count == 0) return cudaSuccess;
  ivf_global_rows_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(index, count, list_off, gbase, nlist, out);
  count_launch();
  return cudaGetLastError();
}

// per-list score scales -> one scale per row (the rows of a list are contiguous): what the scans' epilogues read
namespace {
__global__ void expand_list_scale_kernel(const uint32_t *list_off, uint32_t nlist, const float *scales, uint64_t n,
                                         float *row_scale) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  uint32_t lo = 0, hi = nlist;  // last list with list_off <= r (empty lists share an offset: the last of them wins,
  while (hi - lo > 1) {         // and it is the one that holds the row)
    const uint32_t mid = (lo + hi) >> 1;
    if (list_off[mid] <= (uint32_t)r) lo = mid; else hi = mid;
  }
  row_scale[r] = scales[lo];
}
}  // namespace

cudaError_t launch_expand_list_scale(const uint32_t *list_off, uint32_t nlist, const float *scales, uint64_t n,
                                     float *row_scale, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  expand_list_scale_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(list_off, nlist, scales, n, row_scale);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_distance_matrix(const RowsView &rv, const QueryView &qv, float *out,
                                   cudaStream_t st) {
  if (rv.n == 0 || qv.q == 0) return cudaSuccess;
  const uint64_t blocks = (rv.n + 7) / 8;
  distance_matrix_kernel<<<(unsigned)blocks, 256, 0, st>>>(rv, qv, out);
  count_launch();
  return cudaGetLastError();
}

// ---- full sort of a distance row: the coarse search of wide IVF probing (nprobe > 1024) -----------------------
// One CTA per query: (score, row) keys of all n distances, padded to a power of two, bitonic-sorted in the CTA's
// slice of global scratch (L2-resident: 16384 centroids = 128 KB), the first `take` written out.  NaN distances
// sort last and are never returned (the scans' rule, index_document.h:250-261).
namespace {
__global__ void __launch_bounds__(1024) sort_rows_kernel(const float *dist, uint32_t n, uint32_t p2, uint32_t take,
                                                         u64 *scratch, float *out_scores, uint32_t *out_index) {
  const uint32_t qi = blockIdx.x;
  u64 *keys = scratch + (size_t)qi * p2;
  const float *d = dist + (size_t)qi * n;
  for (uint32_t i = threadIdx.x; i < p2; i += blockDim.x) {
    u64 kv = kMaxKey;
    if (i < n) { const float s = d[i]; if (s == s) kv = pack_key(s, i); }
    keys[i] = kv;
  }
  __syncthreads();
  for (uint32_t size = 2; size <= p2; size <<= 1) {
    for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
      for (uint32_t t = threadIdx.x; t < (p2 >> 1); t += blockDim.x) {
        const uint32_t lo = 2 * t - (t & (stride - 1));  // index with bit `stride` clear
        const uint32_t hi = lo + stride;
        const bool up = (lo & size) == 0;
        const u64 a = keys[lo], b = keys[hi];
        if ((a > b) == up) { keys[lo] = b; keys[hi] = a; }
      }
      __syncthreads();
    }
  }
  for (uint32_t i = threadIdx.x; i < take; i += blockDim.x) {
    const u64 e = i < p2 ? keys[i] : kMaxKey;
    const bool ok = e != kMaxKey;
    out_scores[(size_t)qi * take + i] = ok ? unorder_f32((uint32_t)(e >> 32)) : CUDART_NAN_F;
    out_index[(size_t)qi * take + i] = ok ? (uint32_t)e : ZVB200_INVALID_INDEX;
  }
}
}  // namespace

cudaError_t launch_sort_rows(const float *dist, uint32_t q, uint32_t n, uint32_t p2, uint32_t take, u64 *scratch,
                             float *out_scores, uint32_t *out_index, cudaStream_t st) {
  if (q == 0) return cudaSuccess;
  const unsigned threads = (unsigned)std::min<uint32_t>(1024u, std::max<uint32_t>(32u, p2 >> 1));
  sort_rows_kernel<<<q, threads, 0, st>>>(dist, n, p2, take, scratch, out_scores, out_index);
  count_launch();
  return cudaGetLastError();
}

// ---- IVF: which posting lists a query scans -------------------------------------------------
// The loop header of IVFSearcher::search_impl (src/core/algorithm/ivf/ivf_searcher.cc:216-219):
// lists are visited in coarse order while the rows scanned so far are < max_scan_count (a list
// is always scanned whole).  One thread per query; nprobe <= 1024.
namespace {
__global__ void ivf_plan_kernel(const uint32_t *coarse_index, uint32_t q, uint32_t nprobe, const uint32_t *list_off,
                                const uint32_t *budget_sizes, uint64_t max_scan_count, uint32_t *seg_prefix,
                                uint32_t *seg_row, uint32_t *seg_count, uint32_t *seg_list) {
  // one WARP per query, 32 probes at a time (a single thread walking 64 probes through two dependent global loads
  // each cost 35 us at batch 1 — half of the posting-list scan it prepares)
  const uint32_t qi = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (qi >= q) return;
  ivf_plan_warp(coarse_index + (size_t)qi * nprobe, qi, nprobe, list_off, budget_sizes, max_scan_count, seg_prefix, seg_row,
                seg_count, seg_list);
}
}  // namespace

// the plan of ONE query by one warp; ci = its nprobe coarse ids in coarse order (global or shared memory)
__device__ void ivf_plan_warp(const uint32_t *ci, uint32_t qi, uint32_t nprobe, const uint32_t *list_off,
                              const uint32_t *budget_sizes, uint64_t max_scan_count, uint32_t *seg_prefix,
                              uint32_t *seg_row, uint32_t *seg_count, uint32_t *seg_list) {
  const int lane = threadIdx.x & 31;
  const uint32_t lt = (1u << lane) - 1u;
  uint32_t *pre = seg_prefix + (size_t)qi * (nprobe + 1), *row = seg_row + (size_t)qi * nprobe;
  // `total` is what the budget counts: the list's size in the WHOLE index (budget_sizes, when this store holds
  // one shard of every list) — every shard then visits exactly the lists the unsharded index would
  uint64_t total = 0;
  uint32_t n = 0, local = 0;
  if (lane == 0) pre[0] = 0;
  for (uint32_t c0 = 0; c0 < nprobe; c0 += 32) {
    const uint32_t i = c0 + lane;
    const uint32_t cid = i < nprobe ? ci[i] : ZVB200_INVALID_INDEX;
    const bool valid = cid != ZVB200_INVALID_INDEX;
    uint32_t b = 0, e = 0, bud = 0;
    if (valid) {
      b = list_off[cid]; e = list_off[cid + 1];
      bud = budget_sizes ? budget_sizes[cid] : e - b;
    }
    // the loop of IVFSearcher::search_impl stops at the first missing centroid and once the budget is spent
    const uint32_t dead = __ballot_sync(0xffffffffu, !valid);
    const bool alive = dead == 0 || lane < __ffs(dead) - 1;
    uint64_t inc = alive ? bud : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint64_t v = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += v;
    }
    const uint64_t excl = total + inc - (alive ? bud : 0);
    const bool visit = alive && excl < max_scan_count;
    const bool emit = visit && e != b;
    const uint32_t em = __ballot_sync(0xffffffffu, emit);
    uint32_t linc = emit ? e - b : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, linc, o);
      if (lane >= o) linc += v;
    }
    if (emit) {
      const uint32_t pos = n + __popc(em & lt);
      row[pos] = b;
      pre[pos + 1] = local + linc;
      if (seg_list) seg_list[(size_t)qi * nprobe + pos] = cid;
    }
    const uint32_t vis = __ballot_sync(0xffffffffu, visit);
    total += __shfl_sync(0xffffffffu, visit ? inc : 0, vis ? 31 - __clz(vis) : 0);
    n += __popc(em);
    local += __shfl_sync(0xffffffffu, linc, 31);
    if (vis != 0xffffffffu) break;  // a probe was not visited: nothing after it is
  }
  if (lane == 0) seg_count[qi] = n;
}
namespace {
}  // namespace

cudaError_t launch_ivf_plan(const uint32_t *coarse_index, uint32_t q, uint32_t nprobe, const uint32_t *list_off,
                            const uint32_t *budget_sizes, uint64_t max_scan_count, uint32_t *seg_prefix,
                            uint32_t *seg_row, uint32_t *seg_count, cudaStream_t st, uint32_t *seg_list) {
  ivf_plan_kernel<<<(q + 3) / 4, 128, 0, st>>>(coarse_index, q, nprobe, list_off, budget_sizes, max_scan_count,
                                                  seg_prefix, seg_row, seg_count, seg_list);
  count_launch();
  return cudaGetLastError();
}

}  // namespace zvb

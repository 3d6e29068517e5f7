// K7 — query reformers on the device, and the candidate plan of the refiner.
//
// Reformers: the per-query CPU work the reference does in front of every search of a converted index
// (IndexReformer::transform, called from core_interface::Index::_prepare_for_search):
//   fp16 stores            HalfFloatReformer       fp32 -> fp16, round to nearest even
//                          (src/core/quantizer/half_float_reformer.cc, record_quantizer.h:27-28)
//   cosine fp32 / fp16     CosineReformer          x / |x| (IEEE division by the fp32 norm, only when the norm
//                          is > 0), stored type, the fp32 norm appended
//                          (src/core/quantizer/cosine_reformer.cc:66-112, ailego/math/normalizer.h:45-51)
//   QuantizedInteger int8  IntegerStreamingReformer  per-record min/max quantisation + 16-byte tail
//                          (src/core/quantizer/integer_quantizer_reformer.cc:308-370 ->
//                          RecordQuantizer::quantize_record, record_quantizer.h:24-83)
// One warp per query.  The int8 record sums are the reference's SERIAL fp32 sums (sum += v; sq += v*v with
// the multiply-add contracted, as the reference's build does; DESIGN.md, "reformers"): lane 0 adds the d values in order, everything order-free
// (min/max, codes) is done by the whole warp.  The cosine norm is sqrt of an fp64 sum of squares rounded
// to fp32, like the restatement; the reference's SIMD fp32 sum can differ from it in the last place.
//
// Refiner (core_interface::Index::Search with a refiner_param, src/core/interface/index.cc:408-445: coarse
// search for floor(topk * scale_factor) on the quantised index, then FlatStreamer::search_bf_by_p_keys_impl
// on the exact index, flat_streamer.cc:347-389): rows_plan_kernel turns each query's candidate rows into
// one-row segments for the posting-list scan kernel (scan_gemv.cu, K4), which scores them exactly.
#include <cuda_fp16.h>
#include <float.h>

#include "internal.h"

namespace zvb {

namespace {

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__global__ void __launch_bounds__(32) reform_kernel(const float *in, uint32_t q, uint32_t dim, int kind, uint8_t *out,
                                                    uint32_t elem_bytes) {
  extern __shared__ float sv[];
  const uint32_t qi = blockIdx.x;
  const int lane = threadIdx.x;
  const float *x = in + (size_t)qi * dim;
  uint8_t *o = out + (size_t)qi * elem_bytes;
  if (kind == RF_FP16) {
    __half *oh = reinterpret_cast<__half *>(o);
    for (uint32_t i = lane; i < dim; i += 32) oh[i] = __float2half_rn(x[i]);
    return;
  }
  if (kind == RF_COS_FP32 || kind == RF_COS_FP16) {
    double acc = 0.0;
    for (uint32_t i = lane; i < dim; i += 32) {
      const float v = x[i];
      sv[i] = v;
      acc = fma((double)v, (double)v, acc);
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
    const float norm = (float)sqrt(acc);
    if (kind == RF_COS_FP32) {
      float *of = reinterpret_cast<float *>(o);
      for (uint32_t i = lane; i < dim; i += 32) of[i] = norm > 0.0f ? __fdiv_rn(sv[i], norm) : sv[i];
      if (lane == 0) of[dim] = norm;
    } else {
      __half *oh = reinterpret_cast<__half *>(o);
      for (uint32_t i = lane; i < dim; i += 32) oh[i] = __float2half_rn(norm > 0.0f ? __fdiv_rn(sv[i], norm) : sv[i]);
      if (lane == 0) {  // the fp32 norm in the last two fp16 slots (may be 2-byte aligned only)
        const uint32_t nb = __float_as_uint(norm);
        reinterpret_cast<uint16_t *>(o)[dim] = (uint16_t)(nb & 0xffffu);
        reinterpret_cast<uint16_t *>(o)[dim + 1] = (uint16_t)(nb >> 16);
      }
    }
    return;
  }
  // RF_QI8_L2 / RF_QI8_IP: record_quantizer.h:29-51,72-81
  float mn = FLT_MAX, mx = -FLT_MAX;
  for (uint32_t i = lane; i < dim; i += 32) {
    const float v = x[i];
    sv[i] = v;
    mn = fminf(mn, v);
    mx = fmaxf(mx, v);
  }
  mn = warp_min(mn);
  mx = warp_max(mx);
  const float scale = __fdiv_rn(254.0f, fmaxf(__fsub_rn(mx, mn), FLT_EPSILON));
  const float bias = __fmaf_rn(-mn, scale, -127.0f);
  int8_t *oc = reinterpret_cast<int8_t *>(o);
  __syncwarp();
  for (uint32_t i = lane; i < dim; i += 32) {
    const float v = __fmaf_rn(sv[i], scale, bias);
    sv[i] = v;
    oc[i] = (int8_t)(int)roundf(v);
  }
  __syncwarp();
  if (lane == 0) {
    float sum = 0.0f, sq = 0.0f;
    int isum = 0;
    for (uint32_t i = 0; i < dim; ++i) {
      const float v = sv[i];
      sq = __fmaf_rn(v, v, sq);
      sum = __fadd_rn(sum, v);
      isum += (int)(int8_t)(int)roundf(v);
    }
    const float ex0 = __fdiv_rn(1.0f, scale), ex1 = __fdiv_rn(-bias, scale);
    const uint32_t w[4] = {__float_as_uint(ex0), __float_as_uint(ex1), __float_as_uint(sum),
                           kind == RF_QI8_L2 ? __float_as_uint(sq) : (uint32_t)isum};
    uint8_t *t = o + dim;  // byte-aligned only when dim is not a multiple of 4
    for (int j = 0; j < 16; ++j) t[j] = (uint8_t)(w[j >> 2] >> (8 * (j & 3)));
  }
}

// candidate rows [q][m] -> one-row segments; rows that are invalid, out of range, deleted or filtered are
// skipped here (the list scan applies drop / filter again, harmlessly)
__global__ void rows_plan_kernel(const uint32_t *rows, uint32_t q, uint32_t m, uint64_t n, uint32_t *seg_prefix,
                                 uint32_t *seg_row, uint32_t *seg_count) {
  const uint32_t qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= q) return;
  const uint32_t *r = rows + (size_t)qi * m;
  uint32_t *pre = seg_prefix + (size_t)qi * (m + 1), *row = seg_row + (size_t)qi * m;
  uint32_t c = 0;
  pre[0] = 0;
  for (uint32_t i = 0; i < m; ++i) {
    const uint32_t v = r[i];
    if (v == ZVB200_INVALID_INDEX || v >= n) continue;
    row[c] = v;
    ++c;
    pre[c] = c;
  }
  seg_count[qi] = c;
}

}  // namespace

int reform_kind(int dtype, int metric) {
  if (metric_is_qi8(metric)) {
    if (dtype != ZVB200_DT_INT8) return -1;
    return metric == ZVB200_METRIC_QI8_L2SQ ? RF_QI8_L2 : metric == ZVB200_METRIC_QI8_IP ? RF_QI8_IP : -1;
  }
  if (dtype == ZVB200_DT_INT8) return -1;  // plain int8 stores take int8 queries: nothing to reform from fp32
  if (metric == ZVB200_METRIC_COSINE) return dtype == ZVB200_DT_FP32 ? RF_COS_FP32 : RF_COS_FP16;
  return dtype == ZVB200_DT_FP16 ? RF_FP16 : RF_NONE;
}

cudaError_t launch_reform(const float *in, uint32_t q, uint32_t dim, int kind, uint8_t *out, uint32_t elem_bytes,
                          cudaStream_t st) {
  if (q == 0) return cudaSuccess;
  if (kind == RF_NONE) return cudaMemcpyAsync(out, in, (size_t)q * dim * 4, cudaMemcpyDeviceToDevice, st);
  const size_t smem = kind == RF_FP16 ? 0 : (size_t)dim * 4;
  if (smem > 200 * 1024) return cudaErrorInvalidValue;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(reform_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
  }
  reform_kernel<<<q, 32, smem, st>>>(in, q, dim, kind, out, elem_bytes);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_rows_plan(const uint32_t *rows, uint32_t q, uint32_t m, uint64_t n, uint32_t *seg_prefix,
                             uint32_t *seg_row, uint32_t *seg_count, cudaStream_t st) {
  rows_plan_kernel<<<(q + 127) / 128, 128, 0, st>>>(rows, q, m, n, seg_prefix, seg_row, seg_count);
  count_launch();
  return cudaGetLastError();
}

}  // namespace zvb

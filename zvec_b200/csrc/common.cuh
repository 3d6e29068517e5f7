// Shared device/host helpers for the zvec_b200 scan kernels (sm_100a only).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/zvec_b200.h"

namespace zvb {

typedef unsigned long long u64;

constexpr u64 kMaxKey = ~0ull;  // sorts after every real (score,row) key
constexpr int kWarp = 32;

// ---- (score,row) -> one sortable u64: high word = order-preserving image of the
// float, low word = row.  u64 '<' == (score asc, row asc): the documented tie-break.
__host__ __device__ __forceinline__ uint32_t order_f32(float f) {
#ifdef __CUDA_ARCH__
  uint32_t b = __float_as_uint(f + 0.0f);  // -0 -> +0
#else
  union { float f; uint32_t u; } c; c.f = f + 0.0f; uint32_t b = c.u;
#endif
  return b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u);
}
__host__ __device__ __forceinline__ float unorder_f32(uint32_t o) {
  uint32_t b = o ^ ((o >> 31) ? 0x80000000u : 0xffffffffu);
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  union { float f; uint32_t u; } c; c.u = b; return c.f;
#endif
}
__host__ __device__ __forceinline__ u64 pack_key(float score, uint32_t row) {
  return ((u64)order_f32(score) << 32) | row;
}

// metric classes the kernels specialise on
__host__ __device__ __forceinline__ bool metric_is_l2(int m) {
  return m == ZVB200_METRIC_L2SQ || m == ZVB200_METRIC_L2;
}
__host__ __device__ __forceinline__ bool metric_is_qi8(int m) {
  return m >= ZVB200_METRIC_QI8_L2SQ;
}

// QuantizedInteger float epilogues, evaluated operation by operation, left to right,
// exactly as written at src/core/metric/quantized_integer_metric_matrix.h:36-60 (L2)
// and :94-112 (-IP); no FMA contraction (the oracle TU is built -ffp-contract=off).
// m/q tails = {a = 1/scale, b = -bias/scale, s = sum, s2 = squared sum | int code sum}.
__device__ __forceinline__ float qi8_l2sq(float ipf, float4 m, float4 q, float d) {
  const float sum = __fmul_rn(q.x, q.z);                      // qa * qs
  const float sum2 = __fmul_rn(__fmul_rn(q.x, q.x), q.w);     // qa*qa*qs2
  float t1 = __fmul_rn(__fmul_rn(m.x, m.x), m.w);             // ma*ma*ms2
  float r = __fadd_rn(t1, sum2);
  float t2 = __fmul_rn(__fmul_rn(__fmul_rn(2.0f, m.x), q.x), ipf);  // 2*ma*qa*ip
  r = __fsub_rn(r, t2);
  float db = __fsub_rn(m.y, q.y);                                   // mb-qb
  float t3 = __fmul_rn(__fmul_rn(db, db), d);
  r = __fadd_rn(r, t3);
  float t4 = __fmul_rn(__fmul_rn(2.0f, db), __fsub_rn(__fmul_rn(m.z, m.x), sum));
  return __fadd_rn(r, t4);
}
__device__ __forceinline__ float qi8_mip(float ipf, float4 m, float4 q, float d) {
  float r = __fmul_rn(__fmul_rn(m.x, q.x), ipf);                    // ma*qa*ip
  r = __fadd_rn(r, __fmul_rn(__fmul_rn(m.y, q.x), q.z));            // + mb*qa*qs
  r = __fadd_rn(r, __fmul_rn(__fmul_rn(q.y, m.x), m.z));            // + qb*ma*ms
  r = __fadd_rn(r, __fmul_rn(__fmul_rn(d, q.y), m.y));              // + d*qb*mb
  return -r;
}

// The reference hands its 1x1 distance function (feature, query) from FlatSearcher (flat_distance_matrix.h:53-55) but
// (query, feature) from FlatStreamer and the IVF list scan (flat_streamer_entity.h:155-162,
// ivf_distance_calculator.h:174-183).  The L2 form is symmetric bit for bit; -IP sums its middle terms in the other
// order and differs in the last place for ~15 % of the rows, so the store says which operand came first.
__device__ __forceinline__ float qi8_mip_ord(float ipf, float4 m, float4 q, float d, bool query_first) {
  return query_first ? qi8_mip(ipf, q, m, d) : qi8_mip(ipf, m, q, d);
}

// Relative margin of the bf16 "shadow" filter for QuantizedInteger L2 (scan_mma.cu KIND 3, aux_kernels.cu):
// |filter value - real-arithmetic score| <= eps * [(ma^2 ms2 + mb^2 d) + (qa^2 qs2 + qb^2 d)].
//   2^-8           scaled codes rounded to bf16 (8-bit significand), query codes exact; Cauchy-Schwarz on the sum
//   4.1 * 4.6e-5   the three bilinear extra terms as two-piece bf16 splits without the low*low product
//                  (3 * 2^-16 each), their absolute values total <= 4.1 [row terms] + 3 [query terms]
//   3.1e-5         fp32 evaluation of the constants and of the reference-order score (the int8 filter's margin)
//   (d+9) * 2.5e-6 fp32 accumulation of d + 9 exact products inside the tensor core (4.8e-7 per term, the bound the
//                  other filter modes use) over absolute terms that total <= 5.1 [row + query terms]
__host__ __device__ __forceinline__ float qi8_shadow_eps(uint32_t d) {
  return 4.2e-3f + (float)(d + 9u) * 2.5e-6f;
}

// 128-bit streaming load: read-only path, do not allocate in L1 (each stored byte is
// read once per scan).
__device__ __forceinline__ uint4 ldg_stream(const void *p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// a QuantizedInteger row tail {a, b, s, s2} through the same path
__device__ __forceinline__ float4 ldg_tail(const float4 *p) {
  const uint4 w = ldg_stream(p);
  return make_float4(__uint_as_float(w.x), __uint_as_float(w.y), __uint_as_float(w.z), __uint_as_float(w.w));
}

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

}  // namespace zvb

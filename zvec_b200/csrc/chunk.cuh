// Per-16-byte-chunk arithmetic of the streaming scans (shared by scan_gemv.cu and scan_stream.cu):
// one stored chunk against the matching query chunk, accumulated like the reference's 1x1 kernels.
#pragma once
#include "common.cuh"

namespace zvb {

__device__ __forceinline__ bool bit_set(const uint32_t *bm, uint64_t i) {
  return (bm[i >> 5] >> (i & 31)) & 1u;
}

// accumulate one 16-byte chunk of a stored row against the query chunk in smem
template <int DT, bool L2>
struct Chunk;

template <bool L2>
struct Chunk<ZVB200_DT_FP32, L2> {
  typedef float Acc;
  static __device__ __forceinline__ void fma(Acc &acc, Acc &, const uint4 &v, const uint8_t *qs, int c) {
    const float4 q = *reinterpret_cast<const float4 *>(qs + (size_t)c * 16);
    const float x0 = __uint_as_float(v.x), x1 = __uint_as_float(v.y), x2 = __uint_as_float(v.z),
                x3 = __uint_as_float(v.w);
    if (L2) {
      float d0 = x0 - q.x, d1 = x1 - q.y, d2 = x2 - q.z, d3 = x3 - q.w;
      acc = fmaf(d0, d0, acc); acc = fmaf(d1, d1, acc); acc = fmaf(d2, d2, acc); acc = fmaf(d3, d3, acc);
    } else {
      acc = fmaf(x0, q.x, acc); acc = fmaf(x1, q.y, acc); acc = fmaf(x2, q.z, acc); acc = fmaf(x3, q.w, acc);
    }
  }
  static __host__ __device__ size_t qbytes(uint32_t chunks) { return (size_t)chunks * 16; }
};

template <bool L2>
struct Chunk<ZVB200_DT_FP16, L2> {
  typedef float Acc;
  // query kept as fp32 in smem (32 B per 16-B chunk of halves): widen once, like the
  // reference's F16C path (inner_product_matrix_fp16.cc:144-189), accumulate in fp32.
  static __device__ __forceinline__ void fma(Acc &acc, Acc &, const uint4 &v, const uint8_t *qs, int c) {
    const float4 qa = *reinterpret_cast<const float4 *>(qs + (size_t)c * 32);
    const float4 qb = *reinterpret_cast<const float4 *>(qs + (size_t)c * 32 + 16);
    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&v.x));
    const float2 b = __half22float2(*reinterpret_cast<const __half2 *>(&v.y));
    const float2 e = __half22float2(*reinterpret_cast<const __half2 *>(&v.z));
    const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&v.w));
    if (L2) {
      float d;
      d = a.x - qa.x; acc = fmaf(d, d, acc); d = a.y - qa.y; acc = fmaf(d, d, acc);
      d = b.x - qa.z; acc = fmaf(d, d, acc); d = b.y - qa.w; acc = fmaf(d, d, acc);
      d = e.x - qb.x; acc = fmaf(d, d, acc); d = e.y - qb.y; acc = fmaf(d, d, acc);
      d = f.x - qb.z; acc = fmaf(d, d, acc); d = f.y - qb.w; acc = fmaf(d, d, acc);
    } else {
      acc = fmaf(a.x, qa.x, acc); acc = fmaf(a.y, qa.y, acc); acc = fmaf(b.x, qa.z, acc);
      acc = fmaf(b.y, qa.w, acc); acc = fmaf(e.x, qb.x, acc); acc = fmaf(e.y, qb.y, acc);
      acc = fmaf(f.x, qb.z, acc); acc = fmaf(f.y, qb.w, acc);
    }
  }
  static __host__ __device__ size_t qbytes(uint32_t chunks) { return (size_t)chunks * 32; }
};

template <bool L2>
struct Chunk<ZVB200_DT_INT8, L2> {
  typedef int Acc;
  // exact int32 accumulation as euclidean_distance_matrix_int8.cc:33-120 /
  // inner_product_matrix_int8.cc:319-368; L2 = aa - 2ab + qq with aa gathered here.
  static __device__ __forceinline__ void fma(Acc &ab, Acc &aa, const uint4 &v, const uint8_t *qs, int c) {
    const uint4 q = *reinterpret_cast<const uint4 *>(qs + (size_t)c * 16);
    ab = __dp4a((int)v.x, (int)q.x, ab); ab = __dp4a((int)v.y, (int)q.y, ab);
    ab = __dp4a((int)v.z, (int)q.z, ab); ab = __dp4a((int)v.w, (int)q.w, ab);
    if (L2) {
      aa = __dp4a((int)v.x, (int)v.x, aa); aa = __dp4a((int)v.y, (int)v.y, aa);
      aa = __dp4a((int)v.z, (int)v.z, aa); aa = __dp4a((int)v.w, (int)v.w, aa);
    }
  }
  static __host__ __device__ size_t qbytes(uint32_t chunks) { return (size_t)chunks * 16; }
};

}  // namespace zvb

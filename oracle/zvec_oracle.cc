// TEST INFRASTRUCTURE ONLY — plain scalar CPU restatement of zvec's exhaustive-scan
// hot path (distance + bounded-heap top-k + Flat / IVF scan drivers).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it;
// the product (zvec_b200/, include/) never links, loads or calls it.
//
// PARITY PINNED: checked (tests/test_oracle.py) against the reference's own
// known-answer vectors (tests/ailego/math/*_test.cc constants committed under
// tests/golden/) and against the reference itself compiled from its sources
// (oracle/_ref, built by oracle/build_ref.py) on seeded inputs.
//
// Every function cites the reference file:line it restates (paths relative to
// /root/reference).  fp32/fp16 dots are accumulated in double and rounded once:
// the reference's SIMD lane order is ISA-dependent (SSE/AVX2/AVX-512 variants in
// src/ailego/math/*_fp32.cc), the contract is "within 1e-4 relative".
// int8 arithmetic is exact (int32 accumulate), as in the reference.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

namespace {

// ---- fp16 <-> fp32 (src/ailego/utility/float_helper.cc:1228-1239; F16C
// _mm256_cvtps_ph(x, _MM_FROUND_NO_EXC) = round-to-nearest-even) ---------------
inline float h2f(uint16_t h) {
  uint32_t s = (h >> 15) & 1, e = (h >> 10) & 31, m = h & 1023, u;
  if (e == 0) {
    if (m == 0) u = s << 31;
    else {
      int sh = 0;
      while (!(m & 1024)) { m <<= 1; ++sh; }
      m &= 1023;
      u = (s << 31) | ((uint32_t)(127 - 15 - sh + 1) << 23) | (m << 13);
    }
  } else if (e == 31) u = (s << 31) | 0x7f800000u | (m << 13);
  else u = (s << 31) | ((e + 112) << 23) | (m << 13);
  float f; std::memcpy(&f, &u, 4); return f;
}

inline uint16_t f2h(float f) {
  uint32_t x; std::memcpy(&x, &f, 4);
  uint32_t s = (x >> 16) & 0x8000u; x &= 0x7fffffffu;
  if (x >= 0x7f800000u) return (uint16_t)(s | 0x7c00u | (x > 0x7f800000u ? 0x200u | ((x >> 13) & 1023) : 0));
  if (x >= 0x477ff000u) return (uint16_t)(s | 0x7c00u);          // rounds to inf
  if (x < 0x33000001u) return (uint16_t)s;                        // rounds to zero
  int e = (int)(x >> 23) - 127;
  uint32_t m = (x & 0x7fffffu) | 0x800000u;
  int shift = (e < -14) ? (13 + (-14 - e)) : 13;
  uint32_t hm = m >> shift, rem = m & ((1u << shift) - 1), half = 1u << (shift - 1);
  if (rem > half || (rem == half && (hm & 1))) ++hm;
  uint32_t he = (e < -14) ? 0 : (uint32_t)(e + 15);
  // hm carries the implicit bit for normals; adding lets a mantissa carry bump e
  uint32_t out = (e < -14) ? hm : ((he << 10) + (hm - 1024));
  return (uint16_t)(s | out);
}

}  // namespace

extern "C" {

// metric ids shared with include/zvec_b200.h (ZVB200_METRIC_*)
enum { M_L2SQ = 0, M_L2 = 1, M_IP = 2, M_COSINE = 3, M_QI8_L2SQ = 4, M_QI8_IP = 5, M_QI8_COSINE = 6 };
// dtype ids = reference IndexMeta::DataType (index_meta.h:31-40)
enum { DT_FP16 = 1, DT_FP32 = 2, DT_INT8 = 4 };

void zo_fp32_to_fp16(const float *in, uint64_t n, uint16_t *out) {
  for (uint64_t i = 0; i < n; ++i) out[i] = f2h(in[i]);
}
void zo_fp16_to_fp32(const uint16_t *in, uint64_t n, float *out) {
  for (uint64_t i = 0; i < n; ++i) out[i] = h2f(in[i]);
}

// bytes of one stored element for (metric, dtype, dim) where `dim` is the
// reference's stored dimension INCLUDING tails: cosine fp32 dim+1, cosine fp16
// dim+2 (cosine_converter.cc:204-217), quantized int8 +16 (+20 for cosine)
// (integer_quantizer_converter.cc:502-519, record_quantizer.h:72-80).
uint32_t zo_element_size(int dtype, uint32_t dim) {
  return dtype == DT_FP32 ? dim * 4 : dtype == DT_FP16 ? dim * 2 : dim;
}

// One distance, internal "smaller is better" convention:
//  L2SQ   src/ailego/math/euclidean_distance_matrix.h:33-52   sum (m-q)^2
//  L2     sqrt of it (EuclideanDistanceMatrix)
//  IP     src/ailego/math/inner_product_matrix.h (MinusInnerProductMatrix) -sum m*q
//  COSINE src/ailego/math/cosine_distance_matrix.h:30-49      1 - ip over dim-extra
//  QI8_*  src/core/metric/quantized_integer_metric_matrix.h:36-60,94-112,167-174
float zo_distance(int metric, int dtype, const void *m, const void *q, uint32_t dim) {
  if (metric == M_QI8_L2SQ || metric == M_QI8_IP || metric == M_QI8_COSINE) {
    const int8_t *a = (const int8_t *)m, *b = (const int8_t *)q;
    const size_t d = dim - (metric == M_QI8_COSINE ? 20 : 16);
    int32_t acc = 0;
    for (size_t i = 0; i < d; ++i) acc += (int32_t)a[i] * (int32_t)b[i];
    volatile float ip = (float)acc;  // InnerProductMatrix<int8_t,1,1> (inner_product_matrix_int8.cc:374)
    float ma, mb, ms, ms2, qa, qb, qs, qs2;
    std::memcpy(&ma, a + d, 4); std::memcpy(&mb, a + d + 4, 4);
    std::memcpy(&ms, a + d + 8, 4); std::memcpy(&ms2, a + d + 12, 4);
    std::memcpy(&qa, b + d, 4); std::memcpy(&qb, b + d + 4, 4);
    std::memcpy(&qs, b + d + 8, 4); std::memcpy(&qs2, b + d + 12, 4);
    // every product/sum rounded separately, left to right as written in the
    // reference (built with -ffp-contract=off; this TU too, see oracle/Makefile)
    if (metric == M_QI8_L2SQ) {
      const float sum = qa * qs;
      const float sum2 = qa * qa * qs2;
      return ma * ma * ms2 + sum2 - 2 * ma * qa * ip + (mb - qb) * (mb - qb) * d +
             2 * (mb - qb) * (ms * ma - sum);
    }
    return -(ma * qa * ip + mb * qa * qs + qb * ma * ms + d * qb * mb);
  }
  if (dtype == DT_INT8) {  // plain int8: exact int32 then one cvt (euclidean_distance_matrix_int8.cc:396)
    const int8_t *a = (const int8_t *)m, *b = (const int8_t *)q;
    int64_t acc = 0;
    if (metric == M_L2SQ || metric == M_L2) {
      for (uint32_t i = 0; i < dim; ++i) { int32_t t = (int32_t)a[i] - b[i]; acc += t * t; }
      float r = (float)acc;
      return metric == M_L2 ? std::sqrt(r) : r;
    }
    for (uint32_t i = 0; i < dim; ++i) acc += (int32_t)a[i] * b[i];
    return -(float)acc;
  }
  uint32_t d = dim;
  if (metric == M_COSINE) d = dim - (dtype == DT_FP32 ? 1 : 2);
  double acc = 0.0;
  for (uint32_t i = 0; i < d; ++i) {
    double x = dtype == DT_FP32 ? ((const float *)m)[i] : h2f(((const uint16_t *)m)[i]);
    double y = dtype == DT_FP32 ? ((const float *)q)[i] : h2f(((const uint16_t *)q)[i]);
    if (metric == M_L2SQ || metric == M_L2) { double t = x - y; acc += t * t; }
    else acc += x * y;
  }
  float r = (float)acc;
  switch (metric) {
    case M_L2SQ: return r;
    case M_L2: return std::sqrt(r);
    case M_IP: return -r;
    default: return 1 - r;  // M_COSINE
  }
}

// IndexMetric::normalize (inner_product_metric.cc:377-379, cosine_metric.cc,
// quantized_integer_metric.cc:231-239): back to the user-facing score.
float zo_normalize_score(int metric, float s) {
  if (metric == M_IP || metric == M_QI8_IP) return -s;
  if (metric == M_QI8_COSINE) return 1.0f + s;
  return s;
}

// RecordQuantizer::quantize_record, DT_INT8 branch
// (src/core/quantizer/record_quantizer.h:24-83).  out = dim codes + 16 B tail
// [1/scale, -bias/scale, sum, squared_sum | int code sum].
void zo_quantize_record_int8(const float *vec, uint32_t dim, int is_euclidean, void *out) {
  float mn = std::numeric_limits<float>::max(), mx = std::numeric_limits<float>::lowest();
  for (uint32_t i = 0; i < dim; ++i) { mn = std::min(mn, vec[i]); mx = std::max(mx, vec[i]); }
  const float eps = std::numeric_limits<float>::epsilon();
  float scale = 254 / std::max(mx - mn, eps);
  // the reference TU is built with GCC's default -ffp-contract=fast on an FMA
  // target: a*b+c contracts to one fma (verified bit-exact against oracle/_ref)
  float bias = std::fmaf(-mn, scale, -127.0f);
  float sum = 0.0f, sq = 0.0f; int isum = 0;
  int8_t *o = (int8_t *)out;
  for (uint32_t i = 0; i < dim; ++i) {
    float v = std::fmaf(vec[i], scale, bias);
    sq = std::fmaf(v, v, sq); sum += v;
    o[i] = (int8_t)std::round(v);
    isum += o[i];
  }
  float ex[4] = {1.0f / scale, -bias / scale, sum, sq};
  if (!is_euclidean) std::memcpy(&ex[3], &isum, 4);
  std::memcpy(o + dim, ex, 16);
}

// CosineNormalizeConverter / CosineReformer layout for fp32 input -> fp32:
// L2-normalise (arr[i] /= norm when norm > 0), then append the norm
// (src/core/quantizer/cosine_converter.cc:116-127, ailego/math/normalizer.h:45-51).
// out has dim+1 floats.
void zo_cosine_normalize_fp32(const float *vec, uint32_t dim, float *out) {
  double acc = 0.0;
  for (uint32_t i = 0; i < dim; ++i) acc += (double)vec[i] * vec[i];
  float norm = (float)std::sqrt(acc);
  for (uint32_t i = 0; i < dim; ++i) out[i] = norm > 0.0f ? vec[i] / norm : vec[i];
  out[dim] = norm;
}

}  // extern "C"

namespace {

struct Doc { uint64_t key; float score; uint32_t index; };
struct DocLess { bool operator()(const Doc &a, const Doc &b) const { return a.score < b.score; } };

// ailego::Heap<IndexDocument> (src/include/zvec/ailego/container/heap.h:103-115,
// 180-206) + IndexDocumentHeap::emplace threshold gate
// (src/include/zvec/core/framework/index_document.h:250-261).
struct RefHeap {
  std::vector<Doc> v; size_t limit; float threshold; DocLess cmp;
  RefHeap(size_t k, float th) : limit(std::max<size_t>(k, 1)), threshold(th) { v.reserve(limit); }
  void emplace(uint64_t key, float score, uint32_t index) {
    if (!(score <= threshold)) return;
    Doc val{key, score, index};
    if (v.size() == limit) {
      if (cmp(val, v[0])) replace_top(val);
    } else {
      v.push_back(val);
      std::push_heap(v.begin(), v.end(), cmp);
    }
  }
  void replace_top(const Doc &val) {  // heap.h:180-206
    ptrdiff_t hole = 0, count = (ptrdiff_t)v.size();
    if (count > 1) {
      ptrdiff_t child = 1;
      while (child < count) {
        ptrdiff_t rc = child + 1;
        if (rc < count && cmp(v[child], v[rc])) child = rc;
        if (!cmp(val, v[child])) break;
        v[hole] = v[child]; hole = child; child = (hole << 1) + 1;
      }
    }
    v[hole] = val;
  }
  void sort() { std::sort(v.begin(), v.end(), cmp); }  // heap.h:173-175
};

}  // namespace

extern "C" {

// Flat exhaustive scan, row-major, no filter: FlatSearcherContext::search_row_nofilter /
// single_enqueue_nofilter (src/core/algorithm/flat/flat_searcher_context.h:334-370,
// 847-921) == per row: 1x1 distance, heap.emplace(key, dist, row); then
// heap.sort(); topk_to_result stops at score > threshold (flat_streamer_context.h:104-128).
// rows: n stored elements of `elem` bytes (stride `elem`); keys may be NULL (= row).
// filter (may be NULL): byte per row, non-zero = drop (IndexFilter true = filtered).
// Outputs are q*k, padded with index=0xffffffff / key=~0 / score=NaN.
int zo_flat_search(int metric, int dtype, const void *rows, uint64_t n, uint32_t dim,
                   uint32_t elem, const uint64_t *keys, const uint8_t *filter,
                   const void *queries, uint32_t q, uint32_t k, float threshold,
                   uint64_t *out_keys, float *out_scores, uint32_t *out_index,
                   uint32_t *out_counts) {
  for (uint32_t qi = 0; qi < q; ++qi) {
    const char *qp = (const char *)queries + (size_t)qi * elem;
    RefHeap heap(k, threshold);
    for (uint64_t i = 0; i < n; ++i) {
      if (filter && filter[i]) continue;
      if (keys && keys[i] == ~0ull) continue;  // kInvalidKey holes (flat_streamer_entity.cc:269)
      float d = zo_distance(metric, dtype, (const char *)rows + i * elem, qp, dim);
      heap.emplace(keys ? keys[i] : i, d, (uint32_t)i);
    }
    heap.sort();
    uint32_t c = 0;
    for (; c < heap.v.size() && c < k; ++c) {
      if (heap.v[c].score > threshold) break;
      out_keys[(size_t)qi * k + c] = heap.v[c].key;
      out_scores[(size_t)qi * k + c] = heap.v[c].score;
      out_index[(size_t)qi * k + c] = heap.v[c].index;
    }
    if (out_counts) out_counts[qi] = c;
    for (uint32_t j = c; j < k; ++j) {
      out_keys[(size_t)qi * k + j] = ~0ull;
      out_scores[(size_t)qi * k + j] = std::numeric_limits<float>::quiet_NaN();
      out_index[(size_t)qi * k + j] = 0xffffffffu;
    }
  }
  return 0;
}

// All n distances of one query (the checker's brute-force table: lets the parity
// harness recompute the full candidate set at the k-th score).
void zo_distances(int metric, int dtype, const void *rows, uint64_t n, uint32_t dim,
                  uint32_t elem, const void *query, float *out) {
  for (uint64_t i = 0; i < n; ++i)
    out[i] = zo_distance(metric, dtype, (const char *)rows + i * elem, query, dim);
}

}  // extern "C"

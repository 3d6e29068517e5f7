// K2 — tcgen05 / TMEM contraction with a fused thread-per-query top-k epilogue (query batches).
//
// Replaces FlatSearcherContext::batch_search_* (src/core/algorithm/flat/
// flat_searcher_context.h:681-752,847-921): the XxxMatrix<T,32,N> tile calls
// (inner_product_matrix_fp16.cc:492,547, euclidean_distance_matrix_fp32.cc:521,601, the int8 siblings) and
// the 32*N heap inserts per tile (batch_enqueue_nofilter, flat_searcher_context.h:236-261).
//
// Shape of the contraction (one CTA, one SM), per 128-byte k-block:
//   A = 128 queries x {64 halves | 32 floats | 128 int8}   (UMMA M = 128 -> the 128 TMEM lanes)
//   B = 256 rows    x the same                              (UMMA N = 256 stored rows per tile)
//   D[query][row] accumulates in TMEM (fp32 or int32), two tiles deep (2 x 256 columns = all of TMEM)
// Operand kinds (template KIND): 0 kind::f16 on fp16 rows, 1 kind::tf32 on fp32 rows, 2 kind::i8 on int8 rows,
// 3 kind::f16 on the bf16 shadow operands of a QuantizedInteger L2 store.
// Both operands are K-major, staged by TMA into 128-byte-swizzled shared memory; one elected thread issues
// tcgen05.mma; four epilogue warps read the accumulator with tcgen05.ld (32x32b: one TMEM lane = one query per
// thread, 32 rows per load), so every thread filters ITS query's scores against ITS running bound held in a
// register.  Where the accumulator IS the score (fp16 InnerProduct / Cosine, plain int8) survivors enter the
// query's top-k pool directly; elsewhere it only NOMINATES rows with a sound margin and the exact score is
// recomputed (fp32 FMA over the stored row, or dp4a + the reference's float epilogue) before the pool sees it.
// Pools: k <= 16 in registers, larger k unsorted in XOR-swizzled shared memory with a warp-cooperative rescan.
// Warp 3 produces the per-column constants of the filter modes through an mbarrier ring.  CTAs that scan the same
// query publish their j-th best score; the m-th smallest published word bounds the global k-th best (union bound).
// The score matrix never leaves the SM; per CTA only 128 x k keys go back to HBM.
//
// CTAs are persistent: CTA c works on query tile c % QT for a contiguous range of row
// tiles, and the QT CTAs of a group walk the SAME rows at the same time so that HBM sees
// every stored byte once per batch (the other QT-1 reads hit L2).
#include <cuda.h>
#include <stdlib.h>
#include <stdio.h>

#include "chunk.cuh"
#include "internal.h"
#include "topk.cuh"

namespace zvb {

namespace {

constexpr int kQM = 128;       // queries per CTA tile
constexpr int kBK = 64;        // halves per k-block (128 B = one swizzle row)
constexpr int kBN = 256;       // stored rows per tile
constexpr int kThreads = 256;  // warp 0 TMA, 1 MMA, 2 TMEM alloc, 3 per-column constants, 4-7 epilogue
constexpr uint32_t kStageA = kQM * kBK * 2;
constexpr uint32_t kStageB = kBN * kBK * 2;
constexpr uint32_t kStage = kStageA + kStageB;
constexpr uint32_t kTmemCols = 512;
constexpr unsigned FULL = 0xffffffffu;

struct MmaArgs {
  uint32_t nq;          // queries of this launch
  uint32_t qtiles;      // ceil(nq / 128)
  uint32_t groups;      // row-range groups (= partial lists per query)
  uint32_t ntiles;      // row tiles of kBN
  uint32_t nkb;         // k-blocks of 64 halves
  uint32_t k, nstage;
  uint32_t pool_stride;  // query slots per pool entry row: 128, or 64 when the launch has at most 64 queries (half the
                         // shared memory: one more pipeline stage next to 100-entry pools)
  uint32_t row_begin, row_end;  // global row numbers (tensor map B starts at row_begin)
  int mode;             // 0: s = -dot (InnerProduct), 1: s = 1 - dot (Cosine) straight from the accumulator;
                        // 2/3/4: L2 / InnerProduct / Cosine with the accumulator as a FILTER and an exact re-score
  int metric;           // ZVB200_METRIC_* (the exact re-score's formula)
  const uint8_t *rows;  // store rows (absolute row 0), row_stride bytes apart, `chunks` 16-byte chunks each
  uint32_t row_stride, chunks;
  const float *qf32;    // queries widened to fp32, qstride floats apart
  uint32_t qstride;
  const float *norms;   // |x|^2 per stored row (absolute row index)
  const float *qnorms;  // |q|^2 per query
  float eps;            // |accumulator - exact dot| <= eps * (|x|^2 + |q|^2) / 2
  // int8 stores (kind::i8, exact int32 accumulators): mode 5 plain -IP, 6 plain L2, 7 QuantizedInteger L2,
  // 8 QuantizedInteger -IP; the score formulas are evaluated per column exactly as K1 does
  const float4 *tails;  // QuantizedInteger row tails {a, b, s, s2} (absolute row index)
  const float4 *qtails; // query tails
  const uint8_t *qraw;  // int8 query codes, row_stride apart
  uint32_t d;           // elements of the dot product
  float threshold;
  bool query_first;     // QuantizedInteger -IP operand order (common.cuh)
  const uint32_t *drop, *filter;
  u64 *part;            // [nq][groups][k]
  uint32_t *pub;        // [groups][nq] ordered image of each CTA's pub_j-th best score per query
  uint32_t pub_j;       // which list entry a CTA publishes (1-based)
  uint32_t pub_groups;  // groups whose published words a CTA reads for the union bound: min(groups, 64)
  uint32_t pub_m;       // pub_m * pub_j >= k: the pub_m-th smallest published score bounds the global k-th best
  float sh_mrow_max;    // KIND 3: largest row share of the filter margin in the store (bootstrap bound)
  uint32_t sh_qcap;     // KIND 3: nominated rows a warp queues before it re-scores them together (<= 16)
  uint32_t sh_kk_last;  // KIND 3: 32-byte K slices of the LAST k-block that hold operand columns (the rest is zero padding)
  int dbg;              // ZVB200_MMA_DEBUG probes (perf triage only): 1 no epilogue, 2 no TMA, 4 A only, 8 B only, 16 no slow path
};

// ---- PTX wrappers -----------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
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
// A wait that can never hang the GPU: a pipeline bug traps after ~2 s instead of spinning.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 1023u) == 0 && clock64() - t0 > 4000000000ll) __trap();
  }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *tm, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_hint(uint32_t dst, const CUtensorMap *tm, uint32_t bar, int c0, int c1,
                                                 uint64_t hint) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, "
      "%4}], [%2], %5;" ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "l"(hint)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(adesc), "l"(bdesc),
      "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(adesc), "l"(bdesc),
      "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_i8(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                        uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(adesc), "l"(bdesc),
      "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// the same wait, carrying the 32 destination registers of an earlier tmem_ld32 through it: nothing that
// reads them can be scheduled ahead of the wait (needed when another chunk is processed in between)
__device__ __forceinline__ void tmem_ld_wait_on(uint32_t (&v)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
                 "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]),
                 "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]),
                 "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
               :
               : "memory");
}
// max of 32 score words as floats, as a depth-5 tree (a running max is a 31-deep dependent chain)
// (three-input FMNMX3: 16 instructions instead of 31)
__device__ __forceinline__ float max3f(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
__device__ __forceinline__ float max32(const uint32_t (&v)[32]) {
  float m[12];
#pragma unroll
  for (int i = 0; i < 10; ++i)
    m[i] = max3f(__uint_as_float(v[3 * i]), __uint_as_float(v[3 * i + 1]), __uint_as_float(v[3 * i + 2]));
  m[10] = __uint_as_float(v[30]);
  m[11] = __uint_as_float(v[31]);
#pragma unroll
  for (int i = 0; i < 4; ++i) m[i] = max3f(m[3 * i], m[3 * i + 1], m[3 * i + 2]);
  return fmaxf(max3f(m[0], m[1], m[2]), m[3]);
}
__device__ __forceinline__ uint32_t tmem_ld1(uint32_t taddr) {  // one column of this warp's 32 lanes, waited for
  uint32_t r;
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];\n\t"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(r)
      : "r"(taddr)
      : "memory");
  return r;
}

// Shared-memory matrix descriptor, K-major, 128-byte swizzle: 8-row groups are 1024 B apart
// (SBO), one swizzle atom along K (LBO unused), descriptor version 1 (sm_100).
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t saddr) {
  uint64_t d = (uint64_t)((saddr & 0x3ffffu) >> 4);
  d |= (uint64_t)(1024u >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// Instruction descriptor, kind::f16: fp16 x fp16 -> fp32, A and B K-major.
__host__ __device__ constexpr uint32_t make_idesc_f16(int M, int N) {
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// kind::f16 with bf16 operands (the QuantizedInteger L2 shadow operands): bf16 x bf16 -> fp32.
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N) {
  return make_idesc_f16(M, N) | (1u << 7) | (1u << 10);
}

// kind::tf32: fp32 words in shared memory, the tensor core reads the top 19 bits; fp32 accumulate.
__host__ __device__ constexpr uint32_t make_idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// kind::i8: signed 8-bit operands, int32 accumulate (exact).
__host__ __device__ constexpr uint32_t make_idesc_i8(int M, int N) {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// Exact score of (row, query) by the whole warp, same arithmetic as K1 (scan_gemv.cu) with 32 lanes per
// row: fp32 FMA over 16-byte chunks, xor-shuffle tree.  Result on every lane.
template <int DT>
__device__ __forceinline__ float exact_score(const MmaArgs &a, uint32_t row, uint32_t qidx, int lane) {
  const uint8_t *rp = a.rows + (size_t)row * a.row_stride;
  const uint8_t *qp = reinterpret_cast<const uint8_t *>(a.qf32 + (size_t)qidx * a.qstride);
  float acc = 0.f, aa = 0.f;
  if (metric_is_l2(a.metric)) {
    for (uint32_t c = lane; c < a.chunks; c += 32)
      Chunk<DT, true>::fma(acc, aa, __ldg(reinterpret_cast<const uint4 *>(rp) + c), qp, c);
  } else {
    for (uint32_t c = lane; c < a.chunks; c += 32)
      Chunk<DT, false>::fma(acc, aa, __ldg(reinterpret_cast<const uint4 *>(rp) + c), qp, c);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(FULL, acc, o);
  return a.metric == ZVB200_METRIC_L2SQ ? acc : a.metric == ZVB200_METRIC_L2 ? sqrtf(acc)
         : a.metric == ZVB200_METRIC_IP ? -acc : 1.0f - acc;
}

// Exact QuantizedInteger L2 score of (row, query) by ONE lane: the int8 dot product with dp4a, then the reference's
// float epilogue in its own order — what K1 computes per row (scan_gemv.cu).
__device__ __forceinline__ float exact_qi8_l2(const MmaArgs &a, uint32_t row, uint32_t qidx, const float4 qt, float dq) {
  const uint4 *rp = reinterpret_cast<const uint4 *>(a.rows + (size_t)row * a.row_stride);
  const uint4 *qp = reinterpret_cast<const uint4 *>(a.qraw + (size_t)qidx * a.row_stride);
  int ip = 0;
  for (uint32_t c = 0; c < a.chunks; ++c) {
    const uint4 x = __ldg(rp + c), y = __ldg(qp + c);
    ip = __dp4a((int)x.x, (int)y.x, ip); ip = __dp4a((int)x.y, (int)y.y, ip);
    ip = __dp4a((int)x.z, (int)y.z, ip); ip = __dp4a((int)x.w, (int)y.w, ip);
  }
  return qi8_l2sq((float)ip, a.tails[row], qt, dq);
}

// KR > 0: the epilogue keeps each query's top-k in KR REGISTER slots (k <= KR); KR == 0: in shared memory.
// While the tensor pipe runs, tcgen05.mma operand fetch takes ~3/4 of the SM's shared-memory bandwidth
// (48 KB per 128x256x64 block every ~520 cycles), and an epilogue LDS/STS queues behind it for hundreds
// of cycles (measured: ~1000 cycles per accepted candidate with a shared-memory pool, tools/mma_probe.py
// ZVB200_MMA_DEBUG=128) — so for small k the pool never touches shared memory.
// KIND 0: fp16 rows, kind::f16 (64 halves per 128-byte k-block); KIND 1: fp32 rows, kind::tf32 (32 floats);
// KIND 2: int8 rows, kind::i8 (128 codes).
// KIND 3: QuantizedInteger L2 through the bf16 shadow operands (aux_kernels.cu: qi8_shadow_*): kind::f16 on bf16,
// the accumulator is g = (Q0 - s) / nQa up to the margin qi8_shadow_eps — compared with a per-query register like
// an fp16 inner product, no per-column constants; nominated rows get the exact score from the int8 codes.
// IM: the int8 score form (5 plain -IP, 6 plain L2, 7 QuantizedInteger L2, 8 QuantizedInteger -IP) as a
// compile-time constant — as a run-time value the per-column code kept a branch tree and one waited-for load
// per column (27k-44k cycles per tile); 0 for the other kinds, whose mode stays a run-time argument.
template <int KIND, int KR, int IM>
__global__ void __launch_bounds__(kThreads, 1)
scan_mma_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmX, const MmaArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta = blockIdx.x;
  if (cta >= a.groups * a.qtiles) return;
  const uint32_t qt = cta % a.qtiles, grp = cta / a.qtiles;
  const uint32_t q0 = qt * kQM;
  // balanced contiguous tile ranges per group
  const uint32_t t_begin = (uint32_t)(((uint64_t)a.ntiles * grp) / a.groups);
  const uint32_t t_end = (uint32_t)(((uint64_t)a.ntiles * (grp + 1)) / a.groups);
  const uint32_t ntl = t_end - t_begin;

  // ---- shared memory carve-up (stage buffers need 1024-B alignment for the 128-B swizzle)
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *sbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t k = a.k, nstage = a.nstage;
  u64 *lists = reinterpret_cast<u64 *>(sbase + (size_t)nstage * kStage);      // [k][128]: entry i of query s at i*128+s
  uint64_t *bars = reinterpret_cast<uint64_t *>(lists + (size_t)a.pool_stride * k);     // full[ns] empty[ns] tfull[2] tempty[2] cfull[4] cempty[4]
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * nstage + 12);
  float *colc = reinterpret_cast<float *>((reinterpret_cast<uintptr_t>(tmem_slot + 4) + 15) & ~(uintptr_t)15);  // [2][kBN] per-column constants (float, or float4 for int8)
  constexpr int kBKe = (KIND == 0 || KIND == 3) ? 64 : KIND == 1 ? 32 : 128;   // elements per 128-byte k-block
  const uint32_t bar_full = smem_u32(bars), bar_empty = bar_full + 8 * nstage;
  const uint32_t bar_tfull = bar_empty + 8 * nstage, bar_tempty = bar_tfull + 16;
  const uint32_t bar_cfull = bar_tempty + 16, bar_cempty = bar_cfull + 32;
  // ---- per-column constants (one float, or a float4 for the int8 forms): produced one or more tiles ahead by
  // warp 3 into a ring of kCB buffers, consumed by the epilogue warps through cfull / cempty mbarriers.  (The
  // epilogue threads used to fetch and store them themselves behind a bar.sync per tile: every tile then cost the
  // slowest of the four warps plus their LDG / STS issue — 2200 cycles per tile of 5000 for fp16 L2 at d = 96.)
  const int mode = (KIND == 2 || KIND == 3) ? 0 : a.mode;
  constexpr int imode = IM;
  constexpr bool ic4 = KIND == 2 && IM != 5;  // {|x|^2 as int} (plain int8 L2), the row tail or its filter form (QuantizedInteger)
  const bool resc = mode >= 2;
  constexpr uint32_t kCB = ic4 ? 2 : 4;
  float4 *colc4 = reinterpret_cast<float4 *>(colc);
  const float dq = (float)a.d;
  constexpr float kEpsQ = 3.1e-5f;  // see the QuantizedInteger L2 filter below
  auto cval4 = [&](uint32_t row) -> float4 {
    if (row >= a.row_end) {
      const float nan = __int_as_float(0x7fc00000);
      return imode <= 6 ? make_float4(0, 0, 0, 0) : make_float4(nan, nan, nan, nan);  // NaN score: never a candidate
    }
    if (imode == 5) return make_float4(0, 0, 0, 0);
    if (imode == 6) return make_float4(__int_as_float((int)a.norms[row]), 0, 0, 0);
    const float4 m = a.tails[row];
    if (imode == 7) {
      const float t1 = m.x * m.x * m.w, mbd = m.y * m.y * dq, msma = m.z * m.x;
      const float c0 = t1 + mbd + 2.0f * m.y * msma;
      return make_float4(-(c0 - kEpsQ * (t1 + mbd)), m.y, msma, m.x);
    }
    return m;
  };
  auto cval = [&](uint32_t row) -> float {
    if (row >= a.row_end) return __int_as_float(0xff800000);  // never a candidate
    const float hx = 0.5f * a.norms[row];
    return mode == 2 ? -(1.0f - a.eps) * hx : a.eps * hx;
  };

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmQ) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmX) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (uint32_t s = 0; s < nstage; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (uint32_t s = 0; s < 2; ++s) {
      mbar_init(bar_tfull + 8 * s, 1);
      mbar_init(bar_tempty + 8 * s, 4);  // one arrival per epilogue warp
    }
    for (uint32_t s = 0; s < 4; ++s) {
      mbar_init(bar_cfull + 8 * s, 1);
      mbar_init(bar_cempty + 8 * s, 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (uint32_t t = 0; t < ntl; ++t) {
        const int row_c = (int)((t_begin + t) * kBN);  // row coordinate inside tensor map X
        for (uint32_t kb = 0; kb < a.nkb; ++kb) {
          mbar_wait(bar_empty + 8 * stage, phase ^ 1);
          const uint32_t sa = base + stage * kStage, sb = sa + kStageA;
          if (a.dbg & 14) {
            if (a.dbg & 2) { mbar_arrive(bar_full + 8 * stage); }
            else if (a.dbg & 4) { mbar_expect_tx(bar_full + 8 * stage, kStageA); tma_load_2d_hint(sa, &tmQ, bar_full + 8 * stage, (int)(kb * kBKe), (int)q0, 0x14F0000000000000ull); }
            else { mbar_expect_tx(bar_full + 8 * stage, kStageB); tma_load_2d(sb, &tmX, bar_full + 8 * stage, (int)(kb * kBKe), row_c); }
            if (++stage == nstage) { stage = 0; phase ^= 1; }
            continue;
          }
          mbar_expect_tx(bar_full + 8 * stage, kStage);
          tma_load_2d_hint(sa, &tmQ, bar_full + 8 * stage, (int)(kb * kBKe), (int)q0, 0x14F0000000000000ull);
          tma_load_2d(sb, &tmX, bar_full + 8 * stage, (int)(kb * kBKe), row_c);
          if (++stage == nstage) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      constexpr uint32_t idesc = KIND == 0 ? make_idesc_f16(kQM, kBN) : KIND == 1 ? make_idesc_tf32(kQM, kBN)
                                 : KIND == 2 ? make_idesc_i8(kQM, kBN) : make_idesc_bf16(kQM, kBN);
      uint32_t stage = 0, phase = 0;
      for (uint32_t t = 0; t < ntl; ++t) {
        const uint32_t as = t & 1, aphase = (t >> 1) & 1;
        mbar_wait(bar_tempty + 8 * as, aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * kBN;
        for (uint32_t kb = 0; kb < a.nkb; ++kb) {
          mbar_wait(bar_full + 8 * stage, phase);
          tc_fence_after();
          const uint32_t sa = base + stage * kStage, sb = sa + kStageA;
          const uint64_t da = make_desc_sw128(sa), db = make_desc_sw128(sb);
#pragma unroll
          for (uint32_t kk = 0; kk < 4; ++kk) {  // four 32-byte K slices per 128-byte k-block
            if (KIND == 3 && kb + 1 == a.nkb && kk >= a.sh_kk_last) break;  // zero padding: nothing to add
            if (KIND == 0 || KIND == 3) umma_f16(d_tmem, da + 2 * kk, db + 2 * kk, idesc, (kb | kk) != 0);
            else if (KIND == 1) umma_tf32(d_tmem, da + 2 * kk, db + 2 * kk, idesc, (kb | kk) != 0);
            else umma_i8(d_tmem, da + 2 * kk, db + 2 * kk, idesc, (kb | kk) != 0);
          }
          umma_commit(bar_empty + 8 * stage);  // frees the stage when these MMAs retire
          if (++stage == nstage) { stage = 0; phase ^= 1; }
        }
        umma_commit(bar_tfull + 8 * as);       // accumulator tile complete
      }
    }
  } else if (warp == 3) {
    // ================= per-column constants producer =================
    if (resc || ic4) {
      for (uint32_t t = 0; t < ntl; ++t) {
        const uint32_t cb = t % kCB;
        if (t >= kCB) mbar_wait(bar_cempty + 8 * cb, ((t / kCB) - 1) & 1);
        const uint32_t row0 = a.row_begin + (t_begin + t) * kBN;
        if (ic4) {
          float4 v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = cval4(row0 + lane + 32 * i);
#pragma unroll
          for (int i = 0; i < 8; ++i) colc4[cb * kBN + lane + 32 * i] = v[i];
        } else {
          float v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = cval(row0 + lane + 32 * i);
#pragma unroll
          for (int i = 0; i < 8; ++i) colc[cb * kBN + lane + 32 * i] = v[i];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_cfull + 8 * cb);
      }
    }
  } else if (warp >= 4) {
    // ================= epilogue: thread-per-query filter + private top-k pool =================
    const int ew = warp - 4;
    const int slot = ew * 32 + lane;
    const bool qvalid = q0 + slot < a.nq;
    // entry i of the query in lane l of this warp lives at lists[i * 128 + ew * 32 + (l ^ (i & 31))]: for a fixed i the
    // 32 lanes (one query each) hit 32 different banks, and so do 32 consecutive entries of ONE query — the whole warp
    // can rescan a single query's pool (accept_coop below) as well as every lane its own
    u64 *Lw = lists + ew * 32;
    const uint32_t pstride = a.pool_stride;  // (with 64, warps 2 and 3 hold no live query: they never store, see accept_coop)
#define ZVB_POOL(i) Lw[(size_t)(i) * pstride + (lane ^ ((i) & 31))]
    // int8: the column value is MINUS the exact score, so the rest of the epilogue runs as mode 0 (s = -value)
    constexpr bool SH = KIND == 3;
    const int K = (int)k;
    // The pool is UNSORTED: the k smallest keys accepted so far, with the position of the largest
    // cached.  Accepting a key replaces the largest and rescans the pool for the new one — k
    // independent shared-memory loads (spread over the warp, accept_coop), instead of the dependent
    // load/compare/store chain of a sorted insertion (ailego Heap::emplace, heap.h:103-115, keeps a heap
    // for the same reason).
    // The partial lists leave unsorted; merge_parts does not need order.
    int cnt = 0, maxpos = 0;
    // KR > 0: R[] ascending; the KR - k slots in front hold 0 (never displaced), the rest start at kMaxKey.
    u64 R[KR > 0 ? KR : 1];
    const int rfirst = KR - K;
    if (KR > 0) {
#pragma unroll
      for (int i = 0; i < KR; ++i) R[i] = i < rfirst ? 0ull : kMaxKey;
    }
    // Accept a key only if it is < tau_key = min(threshold gate, this CTA's k-th best, union bound).
    // Union bound: every CTA scanning this query publishes the score of its pub_j-th best row
    // (one word, only ever decreasing).  CTAs hold disjoint rows, so the pub_m-th smallest of the
    // published words has pub_m * pub_j >= k rows at or below it somewhere: the global k-th best
    // cannot be worse.  With G groups and k <= G this tracks the best-of-G minima, i.e. almost the
    // true global k-th best, so a CTA stops inserting rows that only look good locally.  Sharing
    // the bound only removes work, never results: the final lists stay exact and deterministic.
    u64 tau_key = qvalid ? pack_key(a.threshold, 0xffffffffu) : 0ull;
    const uint32_t qcol = qvalid ? q0 + slot : q0;
    volatile uint32_t *my_pub = a.pub + (size_t)grp * a.nq + qcol;
    uint32_t pub_val = 0xffffffffu;
    const int pubj = (int)a.pub_j;  // 1-based; pubj == K publishes the pool's largest once it is full
    // Fast filter: column value f = accumulator + colc[column]; a column is a candidate iff f >= dthr.
    //   mode 0/1: colc = 0, dthr from the score bound directly (conservative for cosine).
    //   mode 2 (L2): exact s = |x|^2 + |q|^2 - 2 x.q <= T  <=>  x.q - hx >= hq - T/2 (hx = |x|^2/2, hq = |q|^2/2);
    //     the accumulator is within eps (hx + hq) of x.q, so colc = -(1 - eps) hx and dthr = hq - T/2 - eps hq.
    //   mode 3/4 (IP / cosine through tf32): colc = +eps hx, dthr = (-T | 1 - T) - eps hq.
    // Modes 2-4 never use f as a score: a candidate is re-scored exactly (exact_score) before it is offered.
    // ---- int8: per-thread query constants and the per-column score, the formulas of K1 (scan_gemv.cu) ----
    const float4 qt = (((KIND == 2 && imode >= 7) || SH) && qvalid) ? a.qtails[q0 + slot] : make_float4(0, 0, 0, 0);
    int qq_i = 0;
    if (KIND == 2 && imode == 6 && qvalid) {
      const int8_t *qp = reinterpret_cast<const int8_t *>(a.qraw + (size_t)(q0 + slot) * a.row_stride);
      for (uint32_t i = 0; i < a.d; ++i) qq_i += (int)qp[i] * (int)qp[i];
    }
    // QuantizedInteger L2 filter constants of this query; kEpsQ bounds the difference between the filter's
    // value and the reference-order score: both are <= 12 roundings of sums whose absolute terms total
    // <= 8 [(ma^2 ms2 + mb^2 d) + (qa^2 qs2 + qb^2 d)] (Cauchy-Schwarz on ip, ms, qs), so
    // 2 * 12 * 2^-24 * 8 = 1.2e-5 < kEpsQ
    const bool small_ip = a.d <= 260;
    const float qsum = qt.x * qt.z, qsum2 = qt.x * qt.x * qt.w;  // qa*qs, qa*qa*qs2
    const float nQ1 = 2.0f * (qt.y * dq + qsum), nQ2 = 2.0f * qt.y, nQa = 2.0f * qt.x;
    const float Q0 = qsum2 + qt.y * qt.y * dq + 2.0f * qt.y * qsum;
    const float Mthr = (SH ? qi8_shadow_eps(a.d) : kEpsQ) * (qsum2 + qt.y * qt.y * dq);
    const float inv_nQa = 1.0f / nQa;
    auto ival = [&](uint32_t raw, const float4 c4) -> float {  // = -score
      const int ip = (int)raw;
      if (imode == 5) return (float)ip;
      if (imode == 6) {
        // minus the SQUARED distance for both L2 metrics (exact integer): the root is taken only for the
        // few candidates (root6 below), a per-column sqrt would cost more than the rest of the filter
        return -(float)(__float_as_int(c4.x) - 2 * ip + qq_i);
      }
      if (imode == 7) {
        // FILTER value for QuantizedInteger L2: minus the score in its expanded, linear-in-ip form
        //   s = [ma^2 ms2 + mb^2 d + 2 mb ms ma] + [qa^2 qs2 + qb^2 d + 2 qb qa qs] - 2 mb (qb d + qa qs)
        //       - 2 (ms ma) qb - 2 ma qa ip
        // with the row bracket (and the row's share of the rounding margin) staged as c4.x = -(c0 - Mrow),
        // c4 = {.., mb, ms*ma, ma}; 1 mul + 3 fma per column instead of the 13 operations of the exact form.
        // Candidates are re-scored with the exact formula (qi8_l2sq) before they are offered.
        // int -> float without I2F (a quarter-rate instruction, the hot loop's bottleneck): exact for
        // |ip| < 2^22, i.e. d <= 260 at +-127 codes
        const float ipf = small_ip ? __int_as_float(ip + 0x4B400000) - 12582912.0f : (float)ip;
        return fmaf(__fmul_rn(c4.w, ipf), nQa, fmaf(c4.z, nQ2, fmaf(c4.y, nQ1, c4.x)));
      }
      return -qi8_mip_ord((float)ip, c4, qt, dq, a.query_first);
    };
    const bool root6 = KIND == 2 && imode == 6 && a.metric == ZVB200_METRIC_L2;
    const float hq = (resc && qvalid) ? 0.5f * a.qnorms[q0 + slot] : 0.f;
    auto dthr_of = [&](u64 tk) -> float {
      if (!qvalid) return __int_as_float(0x7f800000);
      const float tau_f = unorder_f32((uint32_t)(tk >> 32));
      if (KIND == 2 && imode == 7) return Q0 - tau_f - Mthr - (fabsf(Q0) + fabsf(tau_f)) * 4.8e-7f;
      if (SH) {  // the accumulator is (Q0 - s) / nQa within the margin: a candidate iff it reaches (Q0 - tau - margin) / nQa
        const float x = Q0 - tau_f - Mthr - (fabsf(Q0) + fabsf(tau_f) + Mthr) * 4.8e-7f;
        const float y = x * inv_nQa;
        return y - fabsf(y) * 4.8e-7f;
      }
      if (root6) return -(tau_f * tau_f * 1.000001f);  // a superset; the candidate's own root decides
      if (mode == 0) return -tau_f;
      if (mode == 1) {
        const float t = 1.0f - tau_f;
        return t - (fabsf(t) + fabsf(tau_f) + 1.0f) * 2.4e-7f;
      }
      if (mode == 2) {
        const float T = a.metric == ZVB200_METRIC_L2 ? tau_f * tau_f * 1.000001f : tau_f;
        const float hT = 0.5f * T;
        return hq - hT - a.eps * hq - (hq + fabsf(hT)) * 4.8e-7f;
      }
      const float b = mode == 3 ? -tau_f : 1.0f - tau_f;
      return b - a.eps * hq - (fabsf(b) + fabsf(tau_f) + 1.0f) * 4.8e-7f;
    };
    float dthr = dthr_of(tau_key);
    auto accept = [&](u64 key) {
      if (KR > 0) {
        // position-based insertion, every slot decided independently: slot i takes its lower neighbour
        // if the key goes below that one, the key itself if it lands here, else keeps its entry
#pragma unroll
        for (int i = KR - 1; i >= 1; --i) R[i] = key < R[i - 1] ? R[i - 1] : (key < R[i] ? key : R[i]);
        R[0] = key < R[0] ? key : R[0];
        const u64 nt = R[KR - 1];
        if (nt < tau_key) { tau_key = nt; dthr = dthr_of(nt); }
        return;
      }
    };
    // k > 16: the pool is in shared memory.  Warp-collective: `p` lanes accept their `key` one after the other; the
    // accepting lane stores (appends, or replaces its pool's largest), then ALL 32 lanes rescan that one pool for the
    // new largest — 4 loads per lane and a shuffle arg-max instead of k dependent loads by a single lane while 31 idle
    // (measured: 1900 cycles per accepted candidate at k = 100 before, the largest share of the k = 100 epilogue).
    auto accept_coop = [&](bool p, u64 key) {
      // filling a pool needs no rescan until its last slot: those lanes append on their own, in parallel (in the
      // first tiles every lane accepts every column; one at a time that alone cost 4M cycles)
      if (p && cnt + 1 < K) {
        ZVB_POOL(cnt) = key;
        ++cnt;
        p = false;
      }
      uint32_t pb = __ballot_sync(FULL, p);
      if (__popc(pb) >= 4) {
        // many lanes at once (early tiles, loose bounds): every lane rescans its own pool, all of them together
        if (p) {
          if (cnt < K) { ZVB_POOL(cnt) = key; ++cnt; }  // (the macro evaluates its argument twice)
          else ZVB_POOL(maxpos) = key;
          u64 m4[4] = {0, 0, 0, 0};
          int p4[4] = {0, 0, 0, 0};
          int i = 0;
#pragma unroll 4
          for (; i + 3 < K; i += 4) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const u64 x = ZVB_POOL(i + u);
              if (x > m4[u]) { m4[u] = x; p4[u] = i + u; }
            }
          }
          for (; i < K; ++i) {
            const u64 x = ZVB_POOL(i);
            if (x > m4[0]) { m4[0] = x; p4[0] = i; }
          }
          if (m4[1] > m4[0]) { m4[0] = m4[1]; p4[0] = p4[1]; }
          if (m4[3] > m4[2]) { m4[2] = m4[3]; p4[2] = p4[3]; }
          if (m4[2] > m4[0]) { m4[0] = m4[2]; p4[0] = p4[2]; }
          maxpos = p4[0];
          if (m4[0] < tau_key) { tau_key = m4[0]; dthr = dthr_of(m4[0]); }
        }
        return;
      }
      while (pb) {
        const int src = __ffs(pb) - 1;
        pb &= pb - 1;
        bool rescan = false;
        if (lane == src) {
          if (cnt < K) {
            ZVB_POOL(cnt) = key;
            rescan = ++cnt == K;
          } else {
            ZVB_POOL(maxpos) = key;
            rescan = true;
          }
        }
        rescan = __shfl_sync(FULL, (int)rescan, src) != 0;
        if (!rescan) continue;
        __syncwarp();
        u64 mx = 0;
        int mp = 0;
        for (int i = lane; i < K; i += 32) {  // i & 31 == lane: entry i of lane src's query
          const u64 x = Lw[(size_t)i * pstride + (src ^ lane)];
          if (x > mx) { mx = x; mp = i; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const u64 ox = __shfl_xor_sync(FULL, mx, o);
          const int op = __shfl_xor_sync(FULL, mp, o);
          if (ox > mx) { mx = ox; mp = op; }
        }
        if (lane == src) {
          maxpos = mp;
          if (mx < tau_key) { tau_key = mx; dthr = dthr_of(mx); }
        }
      }
    };
    auto publish = [&]() {  // score word of this CTA's pub_j-th best, once it has that many rows
      uint32_t pv = 0xffffffffu;
      if (KR > 0) {
        // masked OR, not a select chain: the compiler turns the latter into R[dynamic] and R into local memory
        uint32_t acc = 0;
#pragma unroll
        for (int i = 0; i < KR; ++i) acc |= (uint32_t)(R[i] >> 32) & (0u - (uint32_t)(i == rfirst + pubj - 1));
        pv = acc;
      } else if (cnt < pubj) {
        return;
      } else if (pubj == K) {
        pv = (uint32_t)(ZVB_POOL(maxpos) >> 32);
      } else {           // pub_j <= 8: the pub_j-th smallest score word of the pool
        uint32_t sel[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) sel[i] = 0xffffffffu;
        for (int e = 0; e < cnt; ++e) {
          uint32_t x = (uint32_t)(ZVB_POOL(e) >> 32);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t lo = min(sel[i], x);
            x = max(sel[i], x);
            sel[i] = lo;
          }
        }
        pv = sel[0];
#pragma unroll
        for (int i = 1; i < 8; ++i) pv = pubj == i + 1 ? sel[i] : pv;
      }
      if (pv < pub_val) { pub_val = pv; *my_pub = pv; }
    };
    auto union_bound = [&]() -> bool {
      uint32_t best[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) best[i] = 0xffffffffu;
      // Any SUBSET of the groups gives a sound bound too (its pub_m-th smallest word still has pub_m * pub_j >= k rows at
      // or below it), only a looser one.  A launch of one query tile has up to 148 groups: 19 dependent rounds of
      // loads, 13 us a call, ~100 us over a short scan (1.25M x 768 at batch 1: 0.36 ms where the rows take 0.27).
      // Such a launch looks at the first 64 groups (its own best is in tau_key already).
      for (uint32_t g0 = 0; g0 < a.pub_groups; g0 += 8) {
        uint32_t x[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)  // the loads first, so their latencies overlap
          x[u] = g0 + u < a.pub_groups ? *reinterpret_cast<volatile uint32_t *>(a.pub + (size_t)(g0 + u) * a.nq + qcol)
                                : 0xffffffffu;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          uint32_t y = x[u];
#pragma unroll
          for (int i = 0; i < 16; ++i) {  // sorted insert: best[] keeps the 16 smallest, ascending
            const uint32_t lo = min(best[i], y);
            y = max(best[i], y);
            best[i] = lo;
          }
        }
      }
      uint32_t b = best[0];
#pragma unroll
      for (int i = 1; i < 16; ++i) b = (int)a.pub_m == i + 1 ? best[i] : b;
      const u64 gk = ((u64)b << 32) | 0xffffffffu;
      if (gk < tau_key) { tau_key = gk; dthr = dthr_of(tau_key); }
      return b != 0xffffffffu;
    };
    uint32_t next_eval = 1;
    // KIND 3 — nominated rows are QUEUED (row, mask of the lanes that nominated it) and re-scored together: the rows of
    // a whole queue are staged into shared memory with all their loads in flight at once, then every nominating lane
    // scores its own query against the staged bytes.  One HBM round trip per queue instead of one per nominated row
    // (the exact re-score used to cost ~2000 cycles per row, 24 % of the k = 100 epilogue).
    const uint32_t sh_rowb = a.chunks * 16 + 16;                       // staged bytes per row: the codes + the tail
    uint8_t *sh_base = reinterpret_cast<uint8_t *>(colc) + (size_t)ew * (128 + (size_t)a.sh_qcap * sh_rowb);
    uint32_t *sh_qrow = reinterpret_cast<uint32_t *>(sh_base), *sh_qmask = sh_qrow + 16;
    uint8_t *sh_stage = sh_base + 128;
    uint32_t sh_qn = 0;                                                 // warp-uniform
    auto sh_drain = [&]() {
      if (!SH || sh_qn == 0) return;
      __syncwarp();
      const uint32_t P = a.chunks + 1, total = sh_qn * P;
      uint4 buf[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const uint32_t item = lane + 32 * r;
        buf[r] = make_uint4(0, 0, 0, 0);
        if (item < total) {
          const uint32_t e = item / P, p = item - e * P, row = sh_qrow[e];
          buf[r] = p < a.chunks ? __ldg(reinterpret_cast<const uint4 *>(a.rows + (size_t)row * a.row_stride) + p)
                                : __ldg(reinterpret_cast<const uint4 *>(a.tails + row));
        }
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const uint32_t item = lane + 32 * r;
        if (item < total) {
          const uint32_t e = item / P, p = item - e * P;
          *reinterpret_cast<uint4 *>(sh_stage + (size_t)e * sh_rowb + (size_t)p * 16) = buf[r];
        }
      }
      __syncwarp();
      const uint4 *qp = reinterpret_cast<const uint4 *>(a.qraw + (size_t)(q0 + slot) * a.row_stride);
      for (uint32_t e = 0; e < sh_qn; ++e) {
        const uint32_t row = sh_qrow[e];
        const bool cand = (sh_qmask[e] >> lane) & 1u;
        float s = 0.f;
        if (cand) {
          const uint4 *rp = reinterpret_cast<const uint4 *>(sh_stage + (size_t)e * sh_rowb);
          int ip = 0;
          for (uint32_t c = 0; c < a.chunks; ++c) {
            const uint4 x = rp[c], y = __ldg(qp + c);
            ip = __dp4a((int)x.x, (int)y.x, ip); ip = __dp4a((int)x.y, (int)y.y, ip);
            ip = __dp4a((int)x.z, (int)y.z, ip); ip = __dp4a((int)x.w, (int)y.w, ip);
          }
          const float4 mt = *reinterpret_cast<const float4 *>(rp + a.chunks);
          s = qi8_l2sq((float)ip, mt, qt, dq);
        }
        const u64 key = pack_key(s, row);
        bool p = cand && key < tau_key;
        if (p && a.drop) p = !bit_set(a.drop, row);
        if (p && a.filter) p = !bit_set(a.filter, row);
        if (KR > 0) { if (p) accept(key); }
        else accept_coop(p, key);
        __syncwarp();
      }
      sh_qn = 0;
    };
    // the first-tile bootstrap needs every row of that tile to be a live row
    const bool boot = !(KIND == 2 && imode == 7) && a.groups > 1 && pubj <= 8 && !a.drop && !a.filter && ntl > 1 &&
                      a.row_begin + (t_begin + 1) * kBN <= a.row_end && !(a.dbg & (1 | 16 | 32 | 64));
    const bool prof = (a.dbg & 128) != 0;  // triage: where the epilogue's cycles go (printed by CTA 0)
    long long c_wait = 0, c_slow = 0, c_ub = 0, c_t0 = 0, c_first4 = 0, c_tiles = 0, c_ex = 0, c_acc = 0, c_ld1 = 0;
    uint32_t n_slow = 0, n_ev = 0;
    const long long c_begin = prof ? clock64() : 0;

    for (uint32_t t = 0; t < ntl; ++t) {
      const uint32_t as = t & 1, aphase = (t >> 1) & 1;
      const uint32_t row0 = a.row_begin + (t_begin + t) * kBN;
      const uint32_t cb = t % kCB;
      const float *cc = colc + cb * kBN;
      const float4 *cc4 = colc4 + cb * kBN;
      const long long c0w = prof ? clock64() : 0;
      if (resc || ic4) mbar_wait(bar_cfull + 8 * cb, (t / kCB) & 1);  // this tile's constants (warp 3, tiles ahead)
      mbar_wait(bar_tfull + 8 * as, aphase);
      if (prof) c_wait += clock64() - c0w;
      const long long c_tile = prof ? clock64() : 0;
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + as * kBN;
      if (t == 0 && boot) {
        // Bootstrap.  With empty pools every score of the first tile is a candidate and the warp pays one
        // accept per COLUMN (256 of them, ~10% of a 1.25M-row shard's run time).  Instead: a first pass
        // over the tile takes only the best score of each 32-column chunk (8 distinct rows), publishes the
        // pub_j-th best of those, and waits briefly for the other CTAs scanning this query to do the same;
        // the union bound over G x 8 rows then lets the exact pass below skip almost every column.
        // The wait is bounded and only about speed: a late CTA just starts with a looser bound.
        uint32_t cm[8];  // ascending: the 8 chunk bests as score words
#pragma unroll
        for (int i = 0; i < 8; ++i) cm[i] = 0xffffffffu;
#pragma unroll 1
        for (int c = 0; c < kBN; c += 32) {
          uint32_t v[32];
          tmem_ld32(taddr + c, v);
          tmem_ld_wait();
          if (KIND == 2) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(ival(v[j], cc4[c + j]));
          }
          if (resc) {  // pessimistic side of the error bound: g = acc - (1 + eps) hx (L2), acc - eps hx (IP, cosine)
            const float cb = mode == 2 ? (1.0f + a.eps) / (1.0f - a.eps) * 1.000001f : -1.0f;
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(fmaf(cc[c + j], cb, __uint_as_float(v[j])));
          }
          const float mx = max32(v);
          float su;  // an upper bound of the exact score of the chunk's best row
          if (SH) {
            // g nQa = (Q0 - s) + Mrow +- (Mrow + Mthr)  =>  s <= Q0 - g nQa + 2 Mrow + Mthr, Mrow <= the store's largest
            su = Q0 - nQa * mx + 2.0f * a.sh_mrow_max + Mthr;
            su += (fabsf(Q0) + fabsf(nQa * mx) + a.sh_mrow_max + Mthr) * 1e-6f;
          } else if (mode == 0) su = root6 ? sqrtf(-mx) : -mx;
          else if (mode == 1) su = 1.0f - mx;
          else {
            su = mode == 2 ? 2.0f * ((1.0f + a.eps) * hq - mx) : (mode == 3 ? -mx : 1.0f - mx) + a.eps * hq;
            su += (fabsf(su) + fabsf(mx) + hq + 1.0f) * 1e-6f;
            if (a.metric == ZVB200_METRIC_L2) su = sqrtf(fmaxf(su, 0.f)) * 1.000001f;
          }
          uint32_t x = order_f32(su);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t lo = min(cm[i], x);
            x = max(cm[i], x);
            cm[i] = lo;
          }
        }
        uint32_t pv = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) pv |= cm[i] & (0u - (uint32_t)(i == pubj - 1));
        if (qvalid && pv < pub_val) { pub_val = pv; *my_pub = pv; }
        for (int tries = 0; tries < 8; ++tries) {
          const bool ready = union_bound() || !qvalid;
          if (__all_sync(FULL, ready)) break;
          __nanosleep(256);
        }
      }
      // one 32-column chunk of this thread's query: filter by the chunk's best score, rarely more
      auto chunk = [&](uint32_t (&v)[32], const int c0) {
        if (KIND == 2) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(ival(v[j], cc4[c0 + j]));
        }
        if (resc) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 c4 = *reinterpret_cast<const float4 *>(cc + c0 + j);  // broadcast read
            v[j] = __float_as_uint(__uint_as_float(v[j]) + c4.x);
            v[j + 1] = __float_as_uint(__uint_as_float(v[j + 1]) + c4.y);
            v[j + 2] = __float_as_uint(__uint_as_float(v[j + 2]) + c4.z);
            v[j + 3] = __float_as_uint(__uint_as_float(v[j + 3]) + c4.w);
          }
        }
        const float mx = max32(v);
        if (__any_sync(FULL, mx >= dthr) && !(a.dbg & 16)) {
          // rare once the bounds are warm: some query of this warp has a candidate among its 32 scores.
          // The warp walks the UNION of its lanes' candidate columns (one copy of the accept code,
          // a warp-uniform column index), each lane acting only on its own candidates.
          const long long c0s = prof ? clock64() : 0;
          if (prof) ++n_slow;
          uint32_t m = 0;
#pragma unroll
          for (int j = 0; j < 32; ++j) m |= (__uint_as_float(v[j]) >= dthr ? 1u : 0u) << j;
          uint32_t wm = __reduce_or_sync(FULL, m);
          if (SH && !a.sh_qcap) {  // (ZVB200_SH_DEFER=0: re-score every nominated row at once, the round-1 path)
            for (uint32_t pm = wm; pm; pm &= pm - 1) {
              const uint32_t prow = row0 + c0 + (__ffs(pm) - 1);
              if (prow < a.row_end && (uint32_t)lane < a.chunks)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(a.rows + (size_t)prow * a.row_stride + (size_t)lane * 16));
              if (prow < a.row_end && lane == 31) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.tails + prow));
            }
          }
          if (SH && a.sh_qcap) {
            // queue the nominated columns (their rows start their trip from HBM now); sh_drain re-scores them together
            for (uint32_t pm = wm; pm; pm &= pm - 1) {
              const int j = __ffs(pm) - 1;
              const uint32_t prow = row0 + c0 + j;
              const uint32_t lm = __ballot_sync(FULL, (m >> j) & 1u);
              if (prow >= a.row_end) continue;  // warp-uniform
              if ((uint32_t)lane < a.chunks)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(a.rows + (size_t)prow * a.row_stride + (size_t)lane * 16));
              if (lane == 31) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.tails + prow));
              if (lane == 0) { sh_qrow[sh_qn] = prow; sh_qmask[sh_qn] = lm; }
              if (prof) ++n_ev;
              if (++sh_qn == a.sh_qcap) sh_drain();
            }
            if (prof) c_slow += clock64() - c0s;
            return;
          }
          while (wm) {
            const int j = __ffs(wm) - 1;
            wm &= wm - 1;
            if (prof) ++n_ev;
            // the column again, straight from TMEM (collective, warp-uniform address): indexing v[] with a
            // run-time j would cost a branch tree or local memory
            const long long c0l = prof ? clock64() : 0;
            const uint32_t raw = tmem_ld1(taddr + c0 + j);
            if (prof) c_ld1 += clock64() - c0l;
            const uint32_t row = row0 + c0 + j;
            const float dv = KIND == 2 ? ival(raw, cc4[c0 + j])
                                       : __uint_as_float(raw) + (resc ? cc[c0 + j] : 0.f);
            bool cand = ((m >> j) & 1u) && dv >= dthr;  // dthr may have tightened since the mask was built
            float s = mode == 0 ? -dv : 1.0f - dv;
            if (root6) s = sqrtf(s);
            if (KIND == 2 && imode == 7) {  // the filter nominated it: the exact reference-order score
              if (cand) s = qi8_l2sq((float)(int)raw, a.tails[row], qt, dq);
            }
            const long long c0e = prof ? clock64() : 0;
            if (SH) {
              // the exact score from the int8 codes.  The row is the same for every nominating lane: with few of
              // them the warp fetches it ONCE (lane c holds chunk c: one HBM round trip instead of a serial walk by a
              // single lane) and scores one query at a time; with many (the first tiles) each lane scores its own.
              uint32_t cb = __ballot_sync(FULL, cand);
              if (__popc(cb) >= 8) {
                if (cand) s = exact_qi8_l2(a, row, q0 + slot, qt, dq);
              } else if (cb) {
                const uint4 *rp = reinterpret_cast<const uint4 *>(a.rows + (size_t)row * a.row_stride);
                uint4 rx[3];
#pragma unroll
                for (int u = 0; u < 3; ++u) {
                  const uint32_t c = lane + 32 * u;
                  rx[u] = c < a.chunks ? __ldg(rp + c) : make_uint4(0, 0, 0, 0);
                }
                const float4 mt = a.tails[row];
                while (cb) {
                  const int src = __ffs(cb) - 1;
                  cb &= cb - 1;
                  const uint4 *qp = reinterpret_cast<const uint4 *>(a.qraw + (size_t)(q0 + ew * 32 + src) * a.row_stride);
                  int ip = 0;
#pragma unroll
                  for (int u = 0; u < 3; ++u) {
                    const uint32_t c = lane + 32 * u;
                    if (c < a.chunks) {
                      const uint4 y = __ldg(qp + c);
                      ip = __dp4a((int)rx[u].x, (int)y.x, ip); ip = __dp4a((int)rx[u].y, (int)y.y, ip);
                      ip = __dp4a((int)rx[u].z, (int)y.z, ip); ip = __dp4a((int)rx[u].w, (int)y.w, ip);
                    }
                  }
#pragma unroll
                  for (int o = 16; o > 0; o >>= 1) ip += __shfl_xor_sync(FULL, ip, o);
                  if (lane == src) s = qi8_l2sq((float)ip, mt, qt, dq);
                }
              }
            }
            if (resc) {
              // the accumulator only nominated the row: its score is recomputed exactly, by the whole warp,
              // one nominating lane (query) at a time
              uint32_t cb = __ballot_sync(FULL, cand);
              while (cb) {
                const int src = __ffs(cb) - 1;
                cb &= cb - 1;
                const float e = exact_score<KIND == 1 ? ZVB200_DT_FP32 : ZVB200_DT_FP16>(a, row, q0 + ew * 32 + src, lane);
                if (lane == src) s = e;
              }
            }
            const long long c0a = prof ? clock64() : 0;
            if (prof) c_ex += c0a - c0e;
            {
              const u64 key = pack_key(s, row);
              bool p = cand && key < tau_key && row < a.row_end;
              if (p && a.drop) p = !bit_set(a.drop, row);
              if (p && a.filter) p = !bit_set(a.filter, row);
              if (KR > 0) { if (p) accept(key); }
              else accept_coop(p, key);
            }
            __syncwarp();
            if (prof) c_acc += clock64() - c0a;
          }
          if (prof) c_slow += clock64() - c0s;
        }
      };
      // the TMEM read of the next chunk is in flight while this one is filtered
      {
        const int cend = (a.dbg & 1) ? 0 : kBN;
        uint32_t va[32], vb[32];
        if (cend) tmem_ld32(taddr, va);
#pragma unroll 1
        for (int c0 = 0; c0 < cend; c0 += 64) {
          tmem_ld_wait_on(va);
          tmem_ld32(taddr + c0 + 32, vb);
          chunk(va, c0);
          tmem_ld_wait_on(vb);
          if (c0 + 64 < cend) tmem_ld32(taddr + c0 + 64, va);
          chunk(vb, c0 + 32);
        }
      }
      if (prof) {
        const long long dt = clock64() - c_tile;
        if (t == 0) c_t0 = dt;
        if (t < 4) c_first4 += dt;
        c_tiles += dt;
      }
      // accumulator stage drained: hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tempty + 8 * as);
      if ((resc || ic4) && lane == 0) mbar_arrive(bar_cempty + 8 * cb);  // (after the __syncwarp above)
      if (SH && (2 * sh_qn >= a.sh_qcap || t + 1 == ntl)) {  // off the MMA's critical path, like the union bound below
        const long long c0d = prof ? clock64() : 0;
        sh_drain();
        if (prof) c_ex += clock64() - c0d;
      }
      // off the MMA's critical path: refresh the union bound at tiles 1,2,3,4,6,9,13,19,... of this CTA
      if (a.groups > 1 && t + 1 == next_eval && !(a.dbg & 32)) {
        const long long c0u = prof ? clock64() : 0;
        publish();
        union_bound();
        if (prof) c_ub += clock64() - c0u;
        next_eval += max(1u, next_eval >> 1);
      }
    }
    if (prof && blockIdx.x == 0 && lane == 0)
      printf("cta0 warp%d: total %lld  wait %lld  slow %lld (%u chunks, %u column events: tmem re-read %lld, exact %lld, accept %lld)  union %lld  tile0 %lld  first4 %lld  all tiles %lld (%u)\n",
             warp, clock64() - c_begin, c_wait, c_slow, n_slow, n_ev, c_ld1, c_ex, c_acc, c_ub, c_t0, c_first4, c_tiles, ntl);
    // write this CTA's partial lists (unsorted): part[q][grp][0..k)
    if (qvalid) {
      u64 *dst = a.part + ((size_t)(q0 + slot) * a.groups + grp) * k;
      if (KR > 0) {
#pragma unroll
        for (int i = 0; i < KR; ++i)
          if (i >= rfirst) dst[i - rfirst] = R[i];
      } else {
        for (int i = 0; i < K; ++i) dst[i] = i < cnt ? ZVB_POOL(i) : kMaxKey;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
  }
}

// ---- host side ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// 2-D tensor map over `rows` rows of `d` elements (fp16 or fp32), `stride` bytes apart; box = one
// 128-byte k-block x box_rows, 128-byte swizzle; out-of-range elements read as zero (ragged d and
// the last row tile).
bool make_map(CUtensorMap *tm, const void *ptr, int kind, uint32_t d, uint64_t rows, uint32_t stride, uint32_t box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  cuuint64_t dims[2] = {d, rows};
  cuuint64_t strides[1] = {stride};
  cuuint32_t box[2] = {(cuuint32_t)(kind == 0 ? 64 : kind == 1 ? 32 : 128), box_rows};
  cuuint32_t estr[2] = {1, 1};
  const CUtensorMapDataType dt = kind == 0 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                 : kind == 1 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_UINT8;
  return fn(tm, dt, 2, const_cast<void *>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

constexpr size_t kSmemMax = 232448;  // 227 KB

// pipeline stages that fit next to the 128 x k lists; false when not even 2 stages fit
bool mma_config(uint32_t k, size_t colc_bytes, uint32_t *nstage, size_t *smem, uint32_t pool_stride = kQM) {
  for (uint32_t ns = 4; ns >= 2; --ns) {
    const size_t need = 1024 + (size_t)ns * kStage + (size_t)pool_stride * k * 8 + (2 * ns + 12) * 8 + 32 + colc_bytes;
    if (need <= kSmemMax) { *nstage = ns; *smem = need; return true; }
  }
  return false;
}

#undef ZVB_POOL

}  // namespace

// bytes of per-column constants the epilogue stages in shared memory for this store type
static bool use_shadow(const RowsView &rv) {
  static const bool off = [] { const char *e = getenv("ZVB200_QI8_SHADOW"); return e && atoi(e) == 0; }();
  return !off && rv.dtype == ZVB200_DT_INT8 && rv.metric == ZVB200_METRIC_QI8_L2SQ && rv.shadow && rv.tails;
}

// KIND 3: nominated rows a warp queues (and stages: codes + tail) before re-scoring them together
static uint32_t shadow_qcap(const RowsView &rv) {
  const uint32_t rowb = rv.chunks * 16 + 16;
  return std::max(1u, std::min(16u, 2048u / rowb));
}

static size_t colc_bytes_for(const RowsView &rv, bool shadow = false) {
  if (shadow) return 4 * (128 + (size_t)shadow_qcap(rv) * (rv.chunks * 16 + 16));  // per epilogue warp: queue + staged rows
  if (rv.dtype == ZVB200_DT_INT8) return 2 * kBN * sizeof(float4);
  const bool direct = rv.dtype == ZVB200_DT_FP16 && !metric_is_l2(rv.metric);
  return direct ? 0 : 4 * kBN * sizeof(float);  // ring of 4 tiles
}

bool scan_mma_supports(const RowsView &rv, uint32_t k) {
  uint32_t ns; size_t smem;
  if (rv.dtype == ZVB200_DT_INT8) {
    // exact while the int32 sums convert to float exactly, like the reference's own int8 kernels
    if (rv.d > 1040) return false;
    if (metric_is_qi8(rv.metric)) { if (!rv.tails) return false; }
    else if (rv.metric == ZVB200_METRIC_L2SQ || rv.metric == ZVB200_METRIC_L2) { if (!rv.norms) return false; }
    else if (rv.metric != ZVB200_METRIC_IP) return false;
  } else {
    if (rv.dtype != ZVB200_DT_FP16 && rv.dtype != ZVB200_DT_FP32) return false;
    if (rv.metric != ZVB200_METRIC_IP && rv.metric != ZVB200_METRIC_COSINE && rv.metric != ZVB200_METRIC_L2SQ &&
        rv.metric != ZVB200_METRIC_L2)
      return false;
    const bool direct = rv.dtype == ZVB200_DT_FP16 && !metric_is_l2(rv.metric);
    if (!direct && !rv.norms) return false;  // the filter + re-score modes need |x|^2 (computed at seal)
  }
  if (k > 128) return false;
  return mma_config(k, colc_bytes_for(rv), &ns, &smem) && encode_fn() != nullptr;
}

size_t scan_mma_max_parts(int sm_count) { return (size_t)sm_count; }

cudaError_t launch_scan_mma(const RowsView &rv, const QueryView &qv, const SearchParams &sp,
                            uint64_t row_begin, uint64_t row_end, u64 *part, uint32_t *nparts,
                            uint32_t *pub, int sm_count, cudaStream_t st, int *kind_out) {
  uint32_t ns; size_t smem;
  const bool sh = use_shadow(rv) && qv.qshadow && !sp.no_shadow;
  // at most 64 queries: half-size pools (measured at k = 100: 2 -> 3 pipeline stages)
  static const bool half_off = [] { const char *e = getenv("ZVB200_MMA_HALF_POOL"); return e && atoi(e) == 0; }();
  const uint32_t pool_stride = (!half_off && qv.q <= 64) ? 64 : kQM;
  if (!mma_config(sp.k, colc_bytes_for(rv, sh), &ns, &smem, pool_stride)) return cudaErrorInvalidValue;
  const uint64_t nrows = row_end > row_begin ? row_end - row_begin : 0;
  MmaArgs a;
  a.pool_stride = pool_stride;
  a.nq = qv.q;
  a.qtiles = (qv.q + kQM - 1) / kQM;
  if ((int)a.qtiles > sm_count) return cudaErrorInvalidValue;  // caller splits the batch
  a.ntiles = (uint32_t)((nrows + kBN - 1) / kBN);
  a.groups = std::max<uint32_t>(1, std::min<uint32_t>((uint32_t)sm_count / a.qtiles, std::max<uint32_t>(a.ntiles, 1)));
  const bool f32 = rv.dtype == ZVB200_DT_FP32, i8 = rv.dtype == ZVB200_DT_INT8;
  const int kind = f32 ? 1 : sh ? 3 : i8 ? 2 : 0;
  const uint32_t bke = f32 ? 32 : sh ? 64 : i8 ? 128 : 64;
  a.nkb = sh ? rv.shadow_k / 64 : (rv.d + bke - 1) / bke;
  a.k = sp.k; a.nstage = ns;
  a.row_begin = (uint32_t)row_begin; a.row_end = (uint32_t)row_end;
  const bool l2 = metric_is_l2(rv.metric);
  if (i8) a.mode = metric_is_qi8(rv.metric) ? (rv.metric == ZVB200_METRIC_QI8_L2SQ ? 7 : 8) : (l2 ? 6 : 5);
  else if (!f32 && !l2) a.mode = rv.metric == ZVB200_METRIC_IP ? 0 : 1;
  else a.mode = l2 ? 2 : rv.metric == ZVB200_METRIC_IP ? 3 : 4;
  a.tails = rv.tails; a.qtails = qv.qtails; a.qraw = qv.qraw; a.d = rv.d;
  a.metric = rv.metric;
  a.rows = rv.rows; a.row_stride = rv.row_stride; a.chunks = rv.chunks;
  a.qf32 = qv.qf32; a.qstride = rv.chunks * (f32 ? 4 : 8);
  a.norms = rv.norms; a.qnorms = qv.qnorms;
  // error bound of the accumulator against the exact dot product, relative to (|x|^2 + |q|^2) / 2 >= |x||q|:
  // fp32 accumulation of d exact products (fp16 inputs), plus two 10-bit-mantissa operands for tf32
  // (2x headroom on the accumulation term: the tensor core's internal alignment of partial sums is not IEEE)
  a.eps = (float)rv.d * 4.8e-7f + (f32 ? 2.2e-3f : 0.f);
  a.sh_mrow_max = rv.shadow_mrow_max;
  {
    static const bool defer = [] { const char *e = getenv("ZVB200_SH_DEFER"); return e && atoi(e) != 0; }();  // measured slower: off
    a.sh_qcap = defer ? shadow_qcap(rv) : 0;
  }
  {  // operand columns of the last k-block: d + 9 of shadow_k, in 16-half (32-byte) MMA slices
    const uint32_t used = rv.d + 9, last = used - std::min(used, (rv.shadow_k / 64 - 1) * 64);
    a.sh_kk_last = sh ? std::max(1u, std::min(4u, (last + 15) / 16)) : 4;
  }
  a.threshold = sp.threshold;
  a.query_first = rv.query_first;
  a.drop = rv.drop; a.filter = sp.filter;
  a.part = part;
  a.pub = pub;
  {
    // publish the j-th best so that the m-th smallest of the G published words covers k rows (m <= min(G,16))
    a.pub_groups = std::min<uint32_t>(a.groups, 64u);
    const uint32_t mmax = std::min<uint32_t>(a.groups, 16u);
    a.pub_j = (sp.k + mmax - 1) / mmax;
    a.pub_m = (sp.k + a.pub_j - 1) / a.pub_j;
    if (a.pub_j > 8) { a.pub_j = sp.k; a.pub_m = 1; }  // few groups, large k: share the pools' k-th bests
  }
  {
    static const int dbg = [] { const char *e = getenv("ZVB200_MMA_DEBUG"); return e ? atoi(e) : 0; }();
    a.dbg = dbg;
  }
  *nparts = a.groups;
  if (kind_out) *kind_out = kind;
  CUtensorMap tmQ, tmX;
  // the caller pads qraw with zero rows up to a whole tile when this is the last group of the batch
  const uint64_t qrows = qv.qpad_rows ? (uint64_t)a.qtiles * kQM : qv.q;
  if (sh) {
    if (!make_map(&tmQ, qv.qshadow, 0, rv.shadow_k, qrows, rv.shadow_k * 2, kQM)) return cudaErrorInvalidValue;
  } else if (f32) {
    if (!make_map(&tmQ, qv.qf32, 1, rv.d, qrows, a.qstride * 4, kQM)) return cudaErrorInvalidValue;
  } else {
    if (!make_map(&tmQ, qv.qraw, kind, rv.d, qrows, rv.row_stride, kQM)) return cudaErrorInvalidValue;
  }
  if (sh) {
    if (!make_map(&tmX, rv.shadow + row_begin * rv.shadow_k * 2, 0, rv.shadow_k, std::max<uint64_t>(nrows, 1),
                  rv.shadow_k * 2, kBN))
      return cudaErrorInvalidValue;
  } else if (!make_map(&tmX, rv.rows + row_begin * rv.row_stride, kind, rv.d, std::max<uint64_t>(nrows, 1), rv.row_stride, kBN))
    return cudaErrorInvalidValue;
  typedef void (*KernFn)(const CUtensorMap, const CUtensorMap, const MmaArgs);
#define ZVB_MMA_KR(KIND, IM) \
  (sp.k <= 8 ? (KernFn)scan_mma_kernel<KIND, 8, IM> : sp.k <= 16 ? (KernFn)scan_mma_kernel<KIND, 16, IM> : (KernFn)scan_mma_kernel<KIND, 0, IM>)
  KernFn kern = f32 ? ZVB_MMA_KR(1, 0)
                : sh ? ZVB_MMA_KR(3, 0)
                : !i8 ? ZVB_MMA_KR(0, 0)
                : a.mode == 5 ? ZVB_MMA_KR(2, 5)
                : a.mode == 6 ? ZVB_MMA_KR(2, 6)
                : a.mode == 7 ? ZVB_MMA_KR(2, 7)
                              : ZVB_MMA_KR(2, 8);
#undef ZVB_MMA_KR
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
  if (e != cudaSuccess) return e;
  if (nrows == 0) {
    // nothing to scan: every partial list is empty
    return cudaMemsetAsync(part, 0xff, (size_t)qv.q * a.groups * sp.k * sizeof(u64), st);
  }
  e = cudaMemsetAsync(pub, 0xff, (size_t)qv.q * a.groups * sizeof(uint32_t), st);
  if (e != cudaSuccess) return e;
  kern<<<a.groups * a.qtiles, kThreads, smem, st>>>(tmQ, tmX, a);
  count_launch();
  return cudaGetLastError();
}

}  // namespace zvb

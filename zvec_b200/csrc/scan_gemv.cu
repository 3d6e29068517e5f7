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

#include "scan_gemv_kernel.cuh"

namespace zvb {

using namespace gemv;

namespace {

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
  const bool pre = !all_keys && pre_applies(rv, lg, nrows0, sm_count * 2);
  const uint64_t rows_per_tile = (uint64_t)(pre ? kPreU : kUmax) * (32 >> lg) * kWarps;
  const uint64_t nrows = row_end > row_begin ? row_end - row_begin : 0;
  uint64_t ntiles = (nrows + rows_per_tile - 1) / rows_per_tile;
  int grid = (int)std::min<uint64_t>(std::max<uint64_t>(ntiles, 1), (uint64_t)sm_count * 2);
  a.nparts = grid;
  *nparts = grid;
  // queries per pass: as many as fit 100 KB of smem (2 CTAs/SM), at most 4
  int qb = all_keys ? 1 : qv.q >= 4 ? 4 : (qv.q >= 2 ? 2 : 1);  // the all-keys form is built for one query per pass
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
      e = launch_flat<ZVB200_DT_FP32>(a, l2, qb, grid, smem, st, pre);
    else if (rv.dtype == ZVB200_DT_FP16)
      e = launch_flat<ZVB200_DT_FP16>(a, l2, qb, grid, smem, st, pre);
    else
      e = launch_flat<ZVB200_DT_INT8>(a, l2, qb, grid, smem, st, pre);
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
      e = launch_lists<ZVB200_DT_FP32>(a, l2, ku, grid, smem, st, gy);
    else if (rv.dtype == ZVB200_DT_FP16)
      e = launch_lists<ZVB200_DT_FP16>(a, l2, ku, grid, smem, st, gy);
    else
      e = launch_lists<ZVB200_DT_INT8>(a, l2, ku, grid, smem, st, gy);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

}  // namespace zvb

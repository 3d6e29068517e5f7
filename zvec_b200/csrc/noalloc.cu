// zvb200_search_device_noalloc: the flat scan for callers that must not allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same kernels as zvb200_search_device's small-batch path (prep_queries -> the GEMV streaming scan ->
// merge_parts); every scratch buffer those kernels need — the prepared queries and the per-CTA partial top-k lists —
// lives in ONE file-scope device array of ZVB200_NOALLOC_SCRATCH_BYTES, which every GPU that loads the library
// holds until the process exits.
//
// The array's name is a host-side stand-in in host code, not a device address: the launchers get the pointer
// cudaGetSymbolAddress returns, resolved once per device with that device current.  Device code of this file names
// the array directly (noalloc_fill_kernel).
//
// Every call shares the array and finds the previous call's contents in it: the partial-list region is refilled
// with empty lists (kMaxKey) on the call's stream before the scan writes it, the prepared queries are rewritten
// whole by prep_queries.  Calls that overlap on the GPU (two streams, two host threads) race on it — serialise them.
#include <mutex>

#include "host.h"

using namespace zvb;

namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[ZVB200_NOALLOC_SCRATCH_BYTES];

__global__ void noalloc_fill_kernel(size_t first_word, size_t words) {
  u64 *p = reinterpret_cast<u64 *>(g_noalloc_scratch) + first_word;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (size_t)gridDim.x * blockDim.x)
    p[i] = kMaxKey;
}

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char *g_scratch_dev[kMaxDevices] = {};  // device address of g_noalloc_scratch on each device

// the current device must be `device`
int scratch_of(int device, unsigned char **out) {
  if (device < 0 || device >= kMaxDevices) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "noalloc: device index");
  std::lock_guard<std::mutex> g(g_mu);
  if (!g_scratch_dev[device]) {
    void *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_scratch)");
    if (!p) return set_err(ZVB200_ERR_RUNTIME, "noalloc: null scratch address");
    g_scratch_dev[device] = (unsigned char *)p;
  }
  *out = g_scratch_dev[device];
  return 0;
}

inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace

extern "C" int zvb200_search_device_noalloc(zvb200_ctx *c, const void *queries_dev, uint32_t q, uint32_t k,
                                            float threshold, const uint32_t *filter_dev, float *out_scores_dev,
                                            uint32_t *out_index_dev) {
  int r = zvb_check_search_args(c, queries_dev, q, k);
  if (r != 0) return r;
  if (!out_scores_dev || !out_index_dev) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "noalloc: null output");
  zvb200_store *s = c->s;
  if (!s->shards.empty()) return set_err(ZVB200_ERR_UNSUPPORTED, "noalloc: plain (one-GPU) stores only");
  if (q > ZVB200_NOALLOC_MAX_QUERIES || k > ZVB200_NOALLOC_MAX_TOPK)
    return set_err(ZVB200_ERR_INVALID_ARGUMENT, "noalloc: at most 8 queries and topk 128 per call");
  if (q == 0) return 0;
  // the layout prep_core / zvb_search_core use, without the 128-query padding only the tensor-core scan needs
  const uint32_t esz = s->dtype == ZVB200_DT_FP32 ? 4 : s->dtype == ZVB200_DT_FP16 ? 2 : 1;
  const size_t qf32_row = (size_t)(s->chunks * 16 / esz) * 4;
  const size_t b_qf32 = up256(s->dtype == ZVB200_DT_INT8 ? 0 : (size_t)q * qf32_row);
  const size_t b_qraw = up256(s->dtype == ZVB200_DT_FP32 ? 0 : (size_t)q * s->row_stride);
  const size_t b_tails = up256((size_t)q * sizeof(float4)), b_norms = up256((size_t)q * sizeof(float));
  const size_t maxparts = scan_gemv_max_parts(s->sm_count);
  const size_t b_part = (size_t)q * maxparts * k * sizeof(u64);
  const size_t o_qraw = b_qf32, o_tails = o_qraw + b_qraw, o_norms = o_tails + b_tails, o_part = o_norms + b_norms;
  if (o_part + b_part > (size_t)ZVB200_NOALLOC_SCRATCH_BYTES)
    return set_err(ZVB200_ERR_INVALID_ARGUMENT, "noalloc: queries * row bytes + queries * topk exceed the 4 MiB static scratch");
  CU(cudaSetDevice(s->device));
  unsigned char *base = nullptr;
  if ((r = scratch_of(s->device, &base)) != 0) return r;

  RowsView rv = zvb_view_for(c, s);
  const uint64_t rows = c->row_limit ? std::min<uint64_t>(c->row_limit, s->n) : s->n;
  QueryView qv;
  qv.qf32 = (const float *)base; qv.qraw = base + o_qraw;
  qv.qtails = (const float4 *)(base + o_tails); qv.qnorms = (const float *)(base + o_norms);
  qv.q = q;
  CU(launch_prep_queries(rv, queries_dev, q, s->elem_bytes, (float *)base, base + o_qraw, (float4 *)(base + o_tails),
                         (float *)(base + o_norms), c->stream));
  // empty partial lists: nothing of the previous call's lists can reach this call's merge
  const size_t words = b_part / sizeof(u64);
  noalloc_fill_kernel<<<(unsigned)std::min<size_t>((words + 255) / 256, 592), 256, 0, c->stream>>>(o_part / sizeof(u64), words);
  CU(cudaGetLastError());
  count_launch();
  SearchParams sp;
  sp.k = k; sp.threshold = threshold; sp.filter = filter_dev;
  uint32_t nparts = 0;
  CU(launch_scan_gemv(rv, qv, sp, 0, rows, (u64 *)(base + o_part), &nparts, s->sm_count, c->stream));
  CU(launch_merge_parts((const u64 *)(base + o_part), nparts, q, k, out_scores_dev, out_index_dev, nullptr, c->stream));
  c->last_kernel = "scan_gemv";
  return 0;
}

// C ABI of libzvec_b200.so (include/zvec_b200.h): device-resident vector store, search
// contexts and the search / merge / distance entry points.  No CPU fallback: every
// compute path launches the CUDA kernels in this directory or fails.
#include "host.h"

using namespace zvb;

namespace zvb {
static std::atomic<uint64_t> g_launches{0};
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }
}  // namespace zvb

static thread_local std::string g_err;

int zvb_set_err(int code, const char *what, const char *detail) {
  g_err = std::string(what) + (detail && *detail ? ": " : "") + (detail ? detail : "");
  return code;
}
int zvb_cuda_fail(cudaError_t e, const char *what) {
  cudaGetLastError();  // clear sticky-less error state
  return set_err(e == cudaErrorMemoryAllocation ? ZVB200_ERR_NOMEMORY : ZVB200_ERR_RUNTIME, what,
                 cudaGetErrorString(e));
}

static int layout_for(int dtype, int metric, uint32_t dim, zvb200_store *s) {
  s->has_tail = false;
  uint32_t esz;
  switch (dtype) {
    case ZVB200_DT_FP32: esz = 4; break;
    case ZVB200_DT_FP16: esz = 2; break;
    case ZVB200_DT_INT8: esz = 1; break;
    default: return set_err(ZVB200_ERR_UNSUPPORTED, "unsupported data type");
  }
  if (metric < ZVB200_METRIC_L2SQ || metric > ZVB200_METRIC_QI8_COSINE)
    return set_err(ZVB200_ERR_UNSUPPORTED, "unsupported metric");
  uint32_t strip = 0;
  if (metric_is_qi8(metric)) {
    // QuantizedIntegerMetric::init accepts DT_INT8 / DT_INT4 only (quantized_integer_metric.cc:33-38)
    if (dtype != ZVB200_DT_INT8) return set_err(ZVB200_ERR_UNSUPPORTED, "QuantizedInteger needs int8");
    strip = metric == ZVB200_METRIC_QI8_COSINE ? 20 : 16;
    s->has_tail = true;
  } else if (metric == ZVB200_METRIC_COSINE) {
    // CosineMetric: fp32 / fp16 only (cosine_metric.cc:150-160); trailing fp32 norm slot
    if (dtype == ZVB200_DT_INT8) return set_err(ZVB200_ERR_UNSUPPORTED, "Cosine needs fp32/fp16");
    strip = 4 / esz;
  }
  if (dim <= strip) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "dimension too small");
  s->dtype = dtype; s->metric = metric; s->dim = dim;
  s->elem_bytes = dim * esz;
  s->d = dim - strip;
  s->vec_bytes = s->d * esz;
  s->row_stride = (s->vec_bytes + 15u) & ~15u;
  s->chunks = s->row_stride / 16;
  s->tail_off = s->vec_bytes;
  return 0;
}

static int store_grow(zvb200_store *s, uint64_t want) {
  if (want <= s->cap) return 0;
  uint64_t ncap = std::max<uint64_t>(want, s->cap + s->cap / 2);
  ncap = std::max<uint64_t>(ncap, 1024);
  uint8_t *nr = nullptr;
  float4 *nt = nullptr;
  cudaError_t e = cudaMalloc((void **)&nr, ncap * s->row_stride);
  if (e == cudaSuccess && s->n) e = cudaMemcpyAsync(nr, s->rows, s->n * s->row_stride, cudaMemcpyDeviceToDevice, s->stream);
  if (e == cudaSuccess && s->has_tail) {
    e = cudaMalloc((void **)&nt, ncap * sizeof(float4));
    if (e == cudaSuccess && s->n) e = cudaMemcpyAsync(nt, s->tails, s->n * sizeof(float4), cudaMemcpyDeviceToDevice, s->stream);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
  if (e != cudaSuccess) {  // the old arrays stay in charge
    if (nr) cudaFree(nr);
    if (nt) cudaFree(nt);
    return cuda_fail(e, "store_grow");
  }
  if (s->rows) cudaFree(s->rows);
  if (s->tails) cudaFree(s->tails);
  if (s->norms) { cudaFree(s->norms); s->norms = nullptr; s->normed = 0; }
  if (s->shadow) { cudaFree(s->shadow); s->shadow = nullptr; s->shadowed = 0; }
  s->rows = nr; s->tails = nt; s->cap = ncap;
  return 0;
}

// A store addresses its rows with 32-bit indices (result rows, list offsets, the packed (score, row) keys):
// at most 2^32 - 2 rows, checked wherever rows are added or reserved.
int zvb_check_row_limit(uint64_t rows) {
  if (rows >= 0xffffffffull) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "a store holds at most 2^32 - 2 rows");
  return 0;
}

extern "C" {

const char *zvb200_version(void) { return "zvec_b200 0.1 (sm_100a)"; }
const char *zvb200_last_error(void) { return g_err.c_str(); }
uint64_t zvb200_kernel_launches(void) { return g_launches.load(); }

int zvb200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int zvb200_store_create(int device, int dtype, int metric, uint32_t dim, zvb200_store **out) {
  if (!out) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null out");
  *out = nullptr;
  auto s = new zvb200_store();
  int r = layout_for(dtype, metric, dim, s);
  if (r != 0) { delete s; return r; }
  int ndev = zvb200_device_count();
  if (ndev == 0) { delete s; return set_err(ZVB200_ERR_RUNTIME, "no CUDA device (there is no CPU fallback)"); }
  if (device < 0 || device >= ndev) { delete s; return set_err(ZVB200_ERR_INVALID_ARGUMENT, "bad device"); }
  s->device = device;
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, device);
  if (e != cudaSuccess) { delete s; return cuda_fail(e, "store_create"); }
  *out = s;
  return 0;
}

int zvb200_store_destroy(zvb200_store *s) {
  if (!s) return 0;
  if (!s->shards.empty() || s->shard_base_dev) zvb_sharded_destroy(s);
  cudaSetDevice(s->device);
  if (s->centroids) zvb200_store_destroy(s->centroids);
  if (s->rows) cudaFree(s->rows);
  if (s->tails) cudaFree(s->tails);
  if (s->norms) cudaFree(s->norms);
  if (s->shadow) cudaFree(s->shadow);
  if (s->shadow_mmax_dev) cudaFree(s->shadow_mmax_dev);
  if (s->drop) cudaFree(s->drop);
  if (s->list_off_dev) cudaFree(s->list_off_dev);
  if (s->budget_dev) cudaFree(s->budget_dev);
  if (s->gbase_dev) cudaFree(s->gbase_dev);
  if (s->row_scale) cudaFree(s->row_scale);
  if (s->staging) cudaFree(s->staging);
  if (s->stream) cudaStreamDestroy(s->stream);
  delete s;
  return 0;
}

int zvb200_store_reserve(zvb200_store *s, uint64_t rows) {
  if (!s) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null store");
  zvb200_store::Write g(s);
  if (int r = zvb_check_row_limit(rows)) return r;
  if (!s->shards.empty()) return zvb_sharded_reserve(s, rows);
  CU(cudaSetDevice(s->device));
  if (rows <= s->cap) return 0;
  uint64_t keep = s->cap;
  s->cap = 0;  // force exact size
  int r = store_grow(s, rows);
  if (r != 0) s->cap = keep;
  return r;
}

// rows [at, at + n) of the device arrays <- n elements in the reference layout (device memory)
static int append_common(zvb200_store *s, const uint8_t *src_dev, uint64_t at, uint64_t n) {
  CU(launch_repack_rows(src_dev, s->elem_bytes, n, s->rows + at * s->row_stride, s->row_stride,
                        s->vec_bytes, s->has_tail ? s->tails + at : nullptr, s->tail_off, s->stream));
  return 0;
}

int zvb200_store_append(zvb200_store *s, const void *rows, uint64_t n, const uint64_t *keys) {
  if (!s || (!rows && n)) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  zvb200_store::Write g(s);
  if (int r = zvb_check_row_limit(s->n + n)) return r;
  if (s->row_scale && n) return set_err(ZVB200_ERR_UNSUPPORTED, "per-list score scales cover the rows that were attached: no append");
  if (!s->shards.empty()) return zvb_sharded_append(s, rows, n, keys, false);
  CU(cudaSetDevice(s->device));
  int r = store_grow(s, s->n + n);
  if (r != 0) return r;
  const size_t chunk_rows = std::max<size_t>(1, ((size_t)64 << 20) / s->elem_bytes);
  if (!s->staging) {
    s->staging_bytes = chunk_rows * s->elem_bytes;
    CU(cudaMalloc((void **)&s->staging, s->staging_bytes));
  }
  // the row count, the keys and the hole bits are committed together, after every chunk has arrived: a CUDA
  // failure half way leaves the store as it was (rows beyond n are never read)
  const uint64_t n0 = s->n;
  for (uint64_t off = 0; off < n; off += chunk_rows) {
    const uint64_t m = std::min<uint64_t>(chunk_rows, n - off);
    CU(cudaMemcpyAsync(s->staging, (const uint8_t *)rows + off * s->elem_bytes, m * s->elem_bytes,
                       cudaMemcpyHostToDevice, s->stream));
    r = append_common(s, s->staging, n0 + off, m);
    if (r != 0) return r;
    CU(cudaStreamSynchronize(s->stream));
  }
  if (keys || !s->keys.empty()) {
    if (s->keys.empty()) { s->keys.resize(n0); for (uint64_t i = 0; i < n0; ++i) s->keys[i] = i; }
    s->keys.resize(n0 + n);
    for (uint64_t i = 0; i < n; ++i) {
      const uint64_t kv = keys ? keys[i] : n0 + i;
      s->keys[n0 + i] = kv;
      if (kv == ZVB200_INVALID_KEY) {
        s->drop_host.resize((n0 + n + 31) / 32, 0u);
        s->drop_host[(n0 + i) >> 5] |= 1u << ((n0 + i) & 31);
        s->has_drop = true; s->drop_dirty = true;
      }
    }
  }
  s->n = n0 + n;
  s->sealed = false;
  return 0;
}

int zvb200_store_append_device(zvb200_store *s, const void *rows_dev, uint64_t n) {
  if (!s || (!rows_dev && n)) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  zvb200_store::Write g(s);
  if (int r = zvb_check_row_limit(s->n + n)) return r;
  if (s->row_scale && n) return set_err(ZVB200_ERR_UNSUPPORTED, "per-list score scales cover the rows that were attached: no append");
  if (!s->shards.empty()) return zvb_sharded_append(s, rows_dev, n, nullptr, true);
  CU(cudaSetDevice(s->device));
  int r = store_grow(s, s->n + n);
  if (r != 0) return r;
  r = append_common(s, (const uint8_t *)rows_dev, s->n, n);
  if (r != 0) return r;
  CU(cudaStreamSynchronize(s->stream));
  if (!s->keys.empty()) { uint64_t n0 = s->n; s->keys.resize(n0 + n); for (uint64_t i = 0; i < n; ++i) s->keys[n0 + i] = n0 + i; }
  s->n += n;
  s->sealed = false;
  return 0;
}

// Device copy of the drop bitmap.  It must cover every row a scan can touch: rows appended after the last upload
// grow the bitmap too (their bits are clear), not only new holes / deletions.
static int sync_drop(zvb200_store *s) {
  if (!s->has_drop) return 0;
  const uint64_t words = (s->n + 31) / 32;
  if (!s->drop_dirty && words <= s->drop_synced) return 0;
  s->drop_host.resize(words, 0u);
  if (words > s->drop_words) {
    // sized for the store's capacity, so that appends within it only re-upload
    const uint64_t want = std::max<uint64_t>(std::max<uint64_t>(words, (s->cap + 31) / 32), 64);
    uint32_t *nd = nullptr;
    CU(cudaMalloc((void **)&nd, want * 4));
    if (s->drop) cudaFree(s->drop);
    s->drop = nd;
    s->drop_words = want;
  }
  if (words) CU(cudaMemcpyAsync(s->drop, s->drop_host.data(), words * 4, cudaMemcpyHostToDevice, s->stream));
  if (s->drop_words > words)
    CU(cudaMemsetAsync(s->drop + words, 0, (s->drop_words - words) * 4, s->stream));
  CU(cudaStreamSynchronize(s->stream));
  s->drop_dirty = false;
  s->drop_synced = words;
  return 0;
}

int zvb200_store_seal(zvb200_store *s) {
  if (!s) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null store");
  zvb200_store::Write g(s);
  if (!s->shards.empty()) return zvb_sharded_seal(s);
  CU(cudaSetDevice(s->device));
  int r = sync_drop(s);
  if (r != 0) return r;
  // |x|^2 per row: the tensor-core L2 / tf32 paths (fp16, fp32) and plain int8 L2
  if ((s->dtype != ZVB200_DT_INT8 || (!s->has_tail && s->chunks * 16 <= 1040)) && s->n > 0) {
    if (!s->norms) { CU(cudaMalloc((void **)&s->norms, s->cap * sizeof(float))); s->normed = 0; }
    if (s->normed < s->n) {
      RowsView v = s->view();
      CU(launch_row_norms(v, s->normed, s->n, s->norms, s->stream));
      CU(cudaStreamSynchronize(s->stream));
      s->normed = s->n;
    }
  }
  // QuantizedInteger L2: the bf16 shadow operands of the tensor-core filter (scan_mma.cu KIND 3).  Extra HBM
  // (2 * round_up(d + 9, 64) bytes per row); when it does not fit the int8 kind::i8 scan stays in charge.
  static const bool shadow_off = [] { const char *e = getenv("ZVB200_QI8_SHADOW"); return e && atoi(e) == 0; }();
  // Worth it while the epilogue, not the tensor pipe, bounds the int8 form: measured 3.1x at d = 96 and 0.7x at
  // d = 768 (bf16 contracts at half the int8 rate), so only up to ZVB200_QI8_SHADOW_MAXD dimensions (default 320).
  static const uint32_t shadow_maxd = [] { const char *e = getenv("ZVB200_QI8_SHADOW_MAXD"); return e ? (uint32_t)atoi(e) : 320u; }();
  if (s->metric == ZVB200_METRIC_QI8_L2SQ && s->d <= std::min(1040u, shadow_maxd) && s->n > 0 && !shadow_off) {
    if (!s->shadow) {
      s->shadow_k = qi8_shadow_k(s->d);
      s->shadowed = 0;
      if (cudaMalloc((void **)&s->shadow, s->cap * (size_t)s->shadow_k * 2) != cudaSuccess) {
        cudaGetLastError();
        s->shadow = nullptr;
      }
    }
    if (s->shadow && !s->shadow_mmax_dev) {
      CU(cudaMalloc((void **)&s->shadow_mmax_dev, sizeof(float)));
      CU(cudaMemset(s->shadow_mmax_dev, 0, sizeof(float)));
    }
    if (s->shadow && s->shadowed < s->n) {
      RowsView v = s->view();
      CU(launch_qi8_shadow_rows(v, s->shadowed, s->n, s->shadow, s->shadow_k, s->shadow_mmax_dev, s->stream));
      CU(cudaMemcpyAsync(&s->shadow_mrow_max, s->shadow_mmax_dev, sizeof(float), cudaMemcpyDeviceToHost, s->stream));
      CU(cudaStreamSynchronize(s->stream));
      s->shadowed = s->n;
    }
  }
  s->sealed = true;
  return 0;
}

uint64_t zvb200_store_count(const zvb200_store *s) { return s ? s->n : 0; }

uint64_t zvb200_store_scan_bytes(const zvb200_store *s) {
  if (!s) return 0;
  return s->n * ((uint64_t)s->vec_bytes + (s->has_tail ? 16 : 0));
}

int zvb200_store_set_operand_order(zvb200_store *s, int query_first) {
  if (!s) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null store");
  zvb200_store::Write g(s);
  s->query_first = query_first != 0;
  for (auto *sh : s->shards) sh->query_first = s->query_first;
  // the IVF centroid index is a nested FlatSearcher (ivf_centroid_index.cc:273-295): feature first, left as it is
  return 0;
}

int zvb200_store_set_deleted(zvb200_store *s, uint64_t row, int deleted) {
  if (!s) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null store");
  zvb200_store::Write g(s);
  if (!s->shards.empty()) return zvb_sharded_set_deleted(s, row, deleted);
  if (row >= s->n) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "bad row");
  s->drop_host.resize((s->n + 31) / 32, 0u);
  if (deleted) s->drop_host[row >> 5] |= 1u << (row & 31);
  else s->drop_host[row >> 5] &= ~(1u << (row & 31));
  s->has_drop = true; s->drop_dirty = true;
  CU(cudaSetDevice(s->device));
  return sync_drop(s);
}

int zvb200_store_write(zvb200_store *s, uint64_t row, const void *vec, uint64_t key) {
  if (!s || !vec) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "bad argument");
  zvb200_store::Write g(s);
  if (!s->shards.empty()) return zvb_sharded_write(s, row, vec, key);
  if (row >= s->n) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "bad row");
  CU(cudaSetDevice(s->device));
  if (!s->staging) {
    s->staging_bytes = std::max<size_t>(1, ((size_t)64 << 20) / s->elem_bytes) * s->elem_bytes;
    CU(cudaMalloc((void **)&s->staging, s->staging_bytes));
  }
  CU(cudaMemcpyAsync(s->staging, vec, s->elem_bytes, cudaMemcpyHostToDevice, s->stream));
  CU(launch_repack_rows(s->staging, s->elem_bytes, 1, s->rows + row * s->row_stride, s->row_stride, s->vec_bytes,
                        s->has_tail ? s->tails + row : nullptr, s->tail_off, s->stream));
  if (s->norms && row < s->normed) {
    RowsView v = s->view();
    CU(launch_row_norms(v, row, row + 1, s->norms, s->stream));
  }
  if (s->shadow && row < s->shadowed) {
    RowsView v = s->view();
    CU(launch_qi8_shadow_rows(v, row, row + 1, s->shadow, s->shadow_k, s->shadow_mmax_dev, s->stream));
    CU(cudaMemcpyAsync(&s->shadow_mrow_max, s->shadow_mmax_dev, sizeof(float), cudaMemcpyDeviceToHost, s->stream));
  }
  CU(cudaStreamSynchronize(s->stream));
  if (s->keys.empty() && key != row) { s->keys.resize(s->n); for (uint64_t i = 0; i < s->n; ++i) s->keys[i] = i; }
  if (!s->keys.empty()) s->keys[row] = key;
  const bool hole = key == ZVB200_INVALID_KEY;
  const bool was = s->has_drop && (row >> 5) < s->drop_host.size() && ((s->drop_host[row >> 5] >> (row & 31)) & 1u);
  if (hole != was) {
    s->drop_host.resize((s->n + 31) / 32, 0u);
    if (hole) s->drop_host[row >> 5] |= 1u << (row & 31);
    else s->drop_host[row >> 5] &= ~(1u << (row & 31));
    s->has_drop = true; s->drop_dirty = true;
    return sync_drop(s);
  }
  return 0;
}

int zvb200_store_get_vector(const zvb200_store *s, uint64_t row, void *out, size_t cap) {
  if (!s || !out || row >= s->n) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "bad argument");
  if (cap < s->elem_bytes) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "buffer too small");
  if (!s->shards.empty()) return zvb_sharded_get_vector(s, row, out, cap);
  CU(cudaSetDevice(s->device));
  memset(out, 0, s->elem_bytes);
  CU(cudaMemcpy(out, s->rows + row * s->row_stride, s->vec_bytes, cudaMemcpyDeviceToHost));
  if (s->has_tail)
    CU(cudaMemcpy((uint8_t *)out + s->tail_off, s->tails + row, 16, cudaMemcpyDeviceToHost));
  return 0;
}

int zvb200_ctx_create(zvb200_store *s, zvb200_ctx **out) {
  if (!s || !out) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  CU(cudaSetDevice(s->device));
  auto c = new zvb200_ctx();
  c->s = s;
  c->device = s->device;
  cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreate(&c->ev_scan0);
  if (e == cudaSuccess) e = cudaEventCreate(&c->ev_scan1);
  if (e != cudaSuccess) { delete c; return cuda_fail(e, "ctx_create"); }
  if (!s->shards.empty()) {
    const int r = zvb_sharded_ctx_create(s, c);
    if (r != 0) { zvb200_ctx_destroy(c); return r; }
  }
  *out = c;
  return 0;
}

int zvb200_ctx_destroy(zvb200_ctx *c) {
  if (!c) return 0;
  zvb_sharded_ctx_destroy(c);
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  c->rg.reset();
  c->xkeys.release(); c->gather.release(); c->out_rows.release(); c->h_rows.release();
  DevBuf *d[] = {&c->qin, &c->qin2, &c->qf32, &c->qraw, &c->qtails, &c->qnorms, &c->part, &c->out_scores,
                 &c->out_index, &c->filter, &c->misc, &c->gtau, &c->coarse_scores, &c->coarse_index,
                 &c->seg_prefix, &c->seg_row, &c->seg_count, &c->seg_list, &c->lm, &c->qsrc, &c->cand, &c->qshadow};
  for (auto b : d) b->release();
  c->h_q.release(); c->h_q2.release(); c->h_scores.release(); c->h_index.release();
  if (c->ev_scan0) cudaEventDestroy(c->ev_scan0);
  if (c->ev_scan1) cudaEventDestroy(c->ev_scan1);
  cudaStreamDestroy(c->stream);
  delete c;
  return 0;
}

int zvb200_ctx_sync(zvb200_ctx *c) {
  if (!c) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null ctx");
  CU(cudaSetDevice(c->device));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}
int zvb200_ctx_last_scan_ms(zvb200_ctx *c, float *ms) {
  if (!c || !ms) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  if (!c->scan_timed) return set_err(ZVB200_ERR_NOREADY, "no search yet");
  if (!c->sub.empty()) {  // sharded: the shards scan side by side; the longest one is the step's scan time
    float mx = 0.f;
    for (auto *sc : c->sub) {
      if (!sc->scan_timed) continue;
      float t = 0.f;
      const int r = zvb200_ctx_last_scan_ms(sc, &t);
      if (r != 0) return r;
      mx = std::max(mx, t);
    }
    *ms = mx;
    return 0;
  }
  CU(cudaSetDevice(c->device));
  CU(cudaEventSynchronize(c->ev_scan1));
  CU(cudaEventElapsedTime(ms, c->ev_scan0, c->ev_scan1));
  return 0;
}
void *zvb200_ctx_stream(zvb200_ctx *c) { return c ? (void *)c->stream : nullptr; }
int zvb200_ctx_set_kernel(zvb200_ctx *c, int which) {
  if (!c || which < 0 || which > 5) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "bad kernel id");
  c->force_kernel = which;
  return 0;
}
const char *zvb200_ctx_last_kernel(const zvb200_ctx *c) { return c ? c->last_kernel : "none"; }
int zvb200_ctx_set_operand_order(zvb200_ctx *c, int order) {
  if (!c || order < -1 || order > 1) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "operand order is -1, 0 or 1");
  c->operand_order = order;
  for (auto *sc : c->sub) sc->operand_order = order;
  return 0;
}
int zvb200_ctx_set_row_limit(zvb200_ctx *c, uint64_t rows) {
  if (!c) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null ctx");
  c->row_limit = rows;
  return 0;
}
int zvb200_ctx_last_mma_kind(const zvb200_ctx *c) { return c ? c->last_mma_kind : -1; }

}  // extern "C"

// ======================================================================================
// search
// ======================================================================================

// queries (stored layout, device) -> compute layout in the context's buffers
static int prep_core(zvb200_ctx *c, zvb200_store *s, const void *queries_dev, uint32_t q, QueryView *out,
                     bool pad_tile = true /* zero rows up to a whole 128-query tile: only the tensor-core scan reads them */) {
  RowsView rv = s->view();
  const uint32_t esz = s->dtype == ZVB200_DT_FP32 ? 4 : s->dtype == ZVB200_DT_FP16 ? 2 : 1;
  const size_t qf32_row = (size_t)(s->chunks * 16 / esz) * 4;
  const size_t qf32_bytes = s->dtype == ZVB200_DT_INT8 ? 0 : (size_t)((q + 127u) & ~127u) * qf32_row;
  // raw queries are padded with zero rows to a whole number of 128-query tiles: the tensor-core
  // scan's TMA box then never leaves the tensor (out-of-range boxes are filled much more slowly)
  const uint32_t qpad = (q + 127u) & ~127u;
  const size_t qraw_bytes = s->dtype == ZVB200_DT_FP32 ? 0 : (size_t)qpad * s->row_stride;
  CU(c->qf32.ensure(qf32_bytes));
  CU(c->qraw.ensure(qraw_bytes));
  CU(c->qtails.ensure((size_t)q * sizeof(float4)));
  CU(c->qnorms.ensure((size_t)q * sizeof(float)));
  if (pad_tile && qraw_bytes && qpad > q)
    CU(cudaMemsetAsync((uint8_t *)c->qraw.p + (size_t)q * s->row_stride, 0, (size_t)(qpad - q) * s->row_stride, c->stream));
  if (pad_tile && s->dtype == ZVB200_DT_FP32 && qpad > q)
    CU(cudaMemsetAsync((uint8_t *)c->qf32.p + (size_t)q * qf32_row, 0, (size_t)(qpad - q) * qf32_row, c->stream));
  CU(launch_prep_queries(rv, queries_dev, q, s->elem_bytes, (float *)c->qf32.p, (uint8_t *)c->qraw.p,
                         (float4 *)c->qtails.p, (float *)c->qnorms.p, c->stream));
  QueryView qv;
  qv.qf32 = (const float *)c->qf32.p; qv.qraw = (const uint8_t *)c->qraw.p;
  qv.qtails = (const float4 *)c->qtails.p; qv.qnorms = (const float *)c->qnorms.p;
  if (rv.shadow) {
    const size_t srow = (size_t)rv.shadow_k * 2;
    CU(c->qshadow.ensure((size_t)qpad * srow));
    if (pad_tile && qpad > q) CU(cudaMemsetAsync((uint8_t *)c->qshadow.p + (size_t)q * srow, 0, (size_t)(qpad - q) * srow, c->stream));
    CU(launch_qi8_shadow_queries(rv, (const uint8_t *)c->qraw.p, (const float4 *)c->qtails.p, q, (uint8_t *)c->qshadow.p,
                                 rv.shadow_k, c->stream));
    qv.qshadow = (const uint8_t *)c->qshadow.p;
  }
  qv.q = q;
  qv.qpad_rows = pad_tile;
  c->last_qv = qv;  // the IVF list scan reuses the prepared queries
  *out = qv;
  return 0;
}

// Prepare queries and run scan + final merge for queries already on the device.
int zvb_search_core(zvb200_ctx *c, zvb200_store *s, const void *queries_dev, uint32_t q, uint32_t k,
                    float threshold, const uint32_t *filter_dev, float *out_scores_dev,
                    uint32_t *out_index_dev, u64 *out_keys_dev, uint64_t row_begin, uint64_t row_end) {
  RowsView rv = zvb_view_for(c, s);
  const uint32_t esz = s->dtype == ZVB200_DT_FP32 ? 4 : s->dtype == ZVB200_DT_FP16 ? 2 : 1;
  SearchParams sp;
  sp.k = k; sp.threshold = threshold; sp.filter = filter_dev;

  int which = c->force_kernel;
  if (which == 5) { which = 3; sp.no_shadow = true; }  // the tcgen05 scan on the int8 codes themselves
  const bool mma_ok = scan_mma_supports(rv, k);
  // batch 1 streams best through the GEMV scan; from 2 queries on the tensor-core scan wins where it
  // applies (it is HBM-bound below ~200 queries and reads the rows once for the whole batch)
  const bool stream_ok = scan_stream_supports(rv, k, 0);
  // tensor cores: fp16 InnerProduct / Cosine at any batch (scores straight from the accumulator); fp16 L2 and
  // fp32 (tf32 filter + exact re-score) from 5 queries on — below that the streaming scan is as fast and simpler
  // (one query with k > 80: the pools leave the tensor-core scan only 2 pipeline stages, 5.7 TB/s; the GEMV scan
  // streams the same rows at 6.6 TB/s — measured at 10M x 768, k = 100)
  // (measured and NOT done: one query over a short range — 1.25M x 768 fp16, a shard of a store split over 8 GPUs — is
  // 0.317 ms on the GEMV scan against 0.352 ms here, but its fp32-FMA sums differ from the tensor core's in the last
  // place: a query would score differently alone and inside a batch, on one GPU and on eight)
  const bool mma_direct = mma_ok && s->dtype == ZVB200_DT_FP16 && !metric_is_l2(s->metric) && !(q == 1 && k > 80);
  // QuantizedInteger L2 with shadow operands: the streaming scan costs 2.5 / 8.2 ms at 1 / 4 queries (100M x 96), the
  // tensor-core scan 5.2 ms for any batch up to 128 — it takes over from 3 queries
  const bool mma_early = mma_ok && rv.shadow && q > 2;
  if (which == 0) which = (mma_direct || (mma_ok && q > 4) || mma_early) ? 3 : (q <= 4) ? 1 : scan_tile_supports(k) ? 2 : 1;
  if (which == 4 && !stream_ok) which = 1;
  if (rv.scaled()) which = 1;  // score scales live in the GEMV scan's epilogue (zvb200_store_set_score_scale)
  if (which == 3 && !mma_ok) return set_err(ZVB200_ERR_UNSUPPORTED, "tcgen05 scan: fp16 / fp32 stores, k <= 128 only");
  if (which == 2 && !scan_tile_supports(k)) which = 1;
  QueryView qv;
  {  // queries -> compute layout; only the tensor-core scan needs them padded to a whole tile
    const int r = prep_core(c, s, queries_dev, q, &qv, which == 3);
    if (r != 0) return r;
  }
  const size_t maxparts = which == 1 ? scan_gemv_max_parts(s->sm_count)
                          : which == 2 ? scan_tile_max_parts(s->sm_count)
                          : which == 3 ? scan_mma_max_parts(s->sm_count) : scan_stream_max_parts(s->sm_count);
  // bound the partial-list scratch to 256 MiB: process the queries in groups
  const size_t per_query = maxparts * (size_t)k * sizeof(u64);
  uint32_t group = (uint32_t)std::max<size_t>(1, ((size_t)256 << 20) / per_query);
  // tensor-core scan: 8 query tiles x 18 row groups per launch fill the 148 SMs and keep the union bound strong
  // (one launch of 79 query tiles would leave 1 row group and 69 idle SMs); the rows are re-read per launch,
  // which stays far below the launch's MMA time
  if (which == 3) {
    static const uint32_t cap = [] { const char *e = getenv("ZVB200_MMA_GROUP"); return e ? (uint32_t)atoi(e) : 1024u; }();
    const uint32_t want = c->mma_group_cap ? c->mma_group_cap : cap;
    group = std::min<uint32_t>(group, std::min<uint32_t>(std::max(want, 128u), (uint32_t)s->sm_count * 128u));
  }
  if (group >= 128) group &= ~127u; else if (group >= 64) group &= ~63u; else if (group >= 4) group &= ~3u;
  group = std::min(group, q);
  CU(c->part.ensure((size_t)group * per_query));
  CU(zvb_scan_event(c, c->ev_scan0));
  c->scan_timed = true;
  for (uint32_t q0 = 0; q0 < q; q0 += group) {
    const uint32_t qn = std::min(group, q - q0);
    QueryView g = qv;
    g.q = qn;
    if (g.qf32) g.qf32 += (size_t)q0 * (s->chunks * 16 / esz);
    if (g.qraw) g.qraw += (size_t)q0 * s->row_stride;
    if (g.qshadow) g.qshadow += (size_t)q0 * rv.shadow_k * 2;
    g.qtails += q0; g.qnorms += q0;
    uint32_t nparts = 0;
    if (which == 1) {
      CU(launch_scan_gemv(rv, g, sp, row_begin, row_end, (u64 *)c->part.p, &nparts, s->sm_count, c->stream));
      c->last_kernel = "scan_gemv";
    } else if (which == 2) {
      CU(launch_scan_tile(rv, g, sp, row_begin, row_end, (u64 *)c->part.p, &nparts, s->sm_count, c->stream));
      c->last_kernel = "scan_tile";
    } else if (which == 4) {
      CU(launch_scan_stream(rv, g, sp, row_begin, row_end, (u64 *)c->part.p, &nparts, s->sm_count, c->stream));
      c->last_kernel = "scan_stream";
    } else {
      CU(c->gtau.ensure((size_t)qn * scan_mma_max_parts(s->sm_count) * sizeof(uint32_t)));
      CU(launch_scan_mma(rv, g, sp, row_begin, row_end, (u64 *)c->part.p, &nparts, (uint32_t *)c->gtau.p,
                         s->sm_count, c->stream, &c->last_mma_kind));
      c->last_kernel = "scan_mma";
    }
    if (q0 + group >= q) CU(zvb_scan_event(c, c->ev_scan1));
    CU(launch_merge_parts((const u64 *)c->part.p, nparts, qn, k, out_scores_dev ? out_scores_dev + (size_t)q0 * k : nullptr,
                          out_index_dev ? out_index_dev + (size_t)q0 * k : nullptr,
                          out_keys_dev ? out_keys_dev + (size_t)q0 * k : nullptr, c->stream));
  }
  return 0;
}

// the searches of this file: plain (score, row) outputs; the context's row limit (zvb200_ctx_set_row_limit) applies
static int search_core(zvb200_ctx *c, zvb200_store *s, const void *queries_dev, uint32_t q, uint32_t k, float threshold,
                       const uint32_t *filter_dev, float *out_scores_dev, uint32_t *out_index_dev, uint64_t row_begin,
                       uint64_t row_end) {
  return zvb_search_core(c, s, queries_dev, q, k, threshold, filter_dev, out_scores_dev, out_index_dev, nullptr,
                         row_begin, row_end);
}

// ---- small searches as one CUDA graph (host.h: ReplayGraph) -------------------------------------------------
static bool graphs_enabled() {
  static const bool on = [] { const char *e = getenv("ZVB200_GRAPH"); return !e || atoi(e) != 0; }();
  return on;
}
constexpr uint32_t kGraphMaxQueries = 8;  // beyond, the kernels are long and the launches overlap them anyway

static GraphKey graph_key(const zvb200_ctx *c, int fn, uint32_t q, uint32_t k, float threshold, uint64_t rows,
                          uint32_t nprobe = 0, uint64_t max_scan = 0) {
  GraphKey key;
  memset(&key, 0, sizeof key);
  key.fn = fn; key.store = c->s; key.gen = c->s->gen; key.epoch = c->buffers_signature();
  key.rows = rows; key.max_scan = max_scan; key.q = q; key.k = k; key.nprobe = nprobe;
  memcpy(&key.thr_bits, &threshold, 4);
  key.mma_group_cap = c->mma_group_cap; key.force_kernel = c->force_kernel; key.operand_order = c->operand_order;
  return key;
}

// enqueue() puts the whole search on the context's stream (copies included) and returns 0; this runs it directly,
// records it (third call with one key) or replays the recording.  The caller synchronises afterwards.
template <class F>
static int run_replayable(zvb200_ctx *c, const GraphKey &key, bool eligible, F &&enqueue) {
  ReplayGraph &g = c->rg;
  if (!eligible || !graphs_enabled() || !c->sub.empty()) return enqueue();
  const bool same = g.seen > 0 && memcmp(&g.key, &key, sizeof key) == 0;
  if (same && g.exec) {
    CU(cudaGraphLaunch(g.exec, c->stream));
    count_launch((int)g.launches);
    ++g.replays;
    c->last_kernel = g.last_kernel; c->last_mma_kind = g.last_mma_kind;
    return 0;
  }
  if (g.failed) return enqueue();
  if (!same) {  // first call with these arguments: run it, every buffer gets its size
    g.reset();
    g.key = key; g.seen = 1;
    const int r = enqueue();
    g.key.epoch = c->buffers_signature();  // what this very call allocated is part of the state
    return r;
  }
  if (g.seen < 2) {  // second call: still direct (a recording has to be replayed a few times to pay for itself)
    ++g.seen;
    return enqueue();
  }
  const uint64_t l0 = g_launches.load();
  if (cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
    cudaGetLastError();
    g.failed = true;
    return enqueue();
  }
  c->capturing = true;
  const int r = enqueue();
  c->capturing = false;
  cudaGraph_t graph = nullptr;
  cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
  const uint32_t launched = (uint32_t)(g_launches.load() - l0);
  if (r == 0 && e == cudaSuccess && graph && c->buffers_signature() == key.epoch)
    e = cudaGraphInstantiate(&g.exec, graph, 0);
  else if (e == cudaSuccess)
    e = cudaErrorUnknown;
  if (graph) cudaGraphDestroy(graph);
  if (e != cudaSuccess || !g.exec) {  // nothing ran: the captured work was only recorded
    cudaGetLastError();
    if (g.exec) { cudaGraphExecDestroy(g.exec); g.exec = nullptr; }
    g.failed = true;
    if (r != 0) return r;
    return enqueue();
  }
  g.launches = launched; g.last_kernel = c->last_kernel; g.last_mma_kind = c->last_mma_kind;
  count_launch(-(int)launched);  // counted while recording, not launched yet
  CU(cudaGraphLaunch(g.exec, c->stream));
  count_launch((int)launched);
  return 0;
}

// device results of the context -> host outputs, document keys and per-query counts
static int queue_result_copies(zvb200_ctx *c, uint32_t q, uint32_t k) {
  const size_t nres = (size_t)q * k;
  CU(cudaMemcpyAsync(c->h_scores.p, c->out_scores.p, nres * 4, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(c->h_index.p, c->out_index.p, nres * 4, cudaMemcpyDeviceToHost, c->stream));
  return 0;
}
static int finish_results(zvb200_ctx *c, zvb200_store *s, uint32_t q, uint32_t k, uint64_t *out_keys, float *out_scores,
                          uint32_t *out_index, uint32_t *out_counts, bool copies_queued = false) {
  const size_t nres = (size_t)q * k;
  if (!copies_queued) {
    CU(c->h_scores.ensure(nres * 4));
    CU(c->h_index.ensure(nres * 4));
    if (int r = queue_result_copies(c, q, k)) return r;
  }
  CU(cudaStreamSynchronize(c->stream));
  const float *hs = (const float *)c->h_scores.p;
  const uint32_t *hi = (const uint32_t *)c->h_index.p;
  memcpy(out_scores, hs, nres * 4);
  if (out_index) memcpy(out_index, hi, nres * 4);
  for (uint32_t i = 0; i < q; ++i) {
    uint32_t cnt = 0;
    for (uint32_t j = 0; j < k; ++j) {
      const uint32_t ix = hi[(size_t)i * k + j];
      if (ix != ZVB200_INVALID_INDEX) ++cnt;
      if (out_keys)
        out_keys[(size_t)i * k + j] = ix == ZVB200_INVALID_INDEX ? ZVB200_INVALID_KEY
                                      : (s->keys.empty() ? (uint64_t)ix : s->keys[ix]);
    }
    if (out_counts) out_counts[i] = cnt;
  }
  return 0;
}

// The refiner's second stage on the device: exact scores of each query's candidate rows, best k
// (FlatStreamer::search_bf_by_p_keys_impl, flat_streamer.cc:347-389, with rows for keys).
int zvb_rows_core(zvb200_ctx *c, zvb200_store *s, const void *queries_dev, uint32_t q, const uint32_t *rows_dev,
                  uint32_t m, uint32_t k, float threshold, const uint32_t *filter_dev, float *out_scores_dev,
                  uint32_t *out_index_dev, u64 *out_keys_dev) {
  QueryView qv;
  int r = prep_core(c, s, queries_dev, q, &qv);
  if (r != 0) return r;
  CU(c->seg_prefix.ensure((size_t)q * (m + 1) * 4));
  CU(c->seg_row.ensure((size_t)q * m * 4));
  CU(c->seg_count.ensure((size_t)q * 4));
  CU(launch_rows_plan(rows_dev, q, m, s->n, (uint32_t *)c->seg_prefix.p, (uint32_t *)c->seg_row.p,
                      (uint32_t *)c->seg_count.p, c->stream));
  RowsView rv = zvb_view_for(c, s);
  SearchParams sp;
  sp.k = k; sp.threshold = threshold; sp.filter = filter_dev;
  const size_t parts = scan_lists_parts(q, s->sm_count);
  CU(c->part.ensure((size_t)q * parts * k * sizeof(u64)));
  uint32_t nparts = 0;
  CU(zvb_scan_event(c, c->ev_scan0));
  c->scan_timed = true;
  CU(launch_scan_lists(rv, qv, sp, (const uint32_t *)c->seg_prefix.p, (const uint32_t *)c->seg_row.p,
                       (const uint32_t *)c->seg_count.p, m, (u64 *)c->part.p, &nparts, s->sm_count, c->stream));
  CU(zvb_scan_event(c, c->ev_scan1));
  c->last_kernel = "scan_rows";
  CU(launch_merge_parts((const u64 *)c->part.p, nparts, q, k, out_scores_dev, out_index_dev, out_keys_dev, c->stream));
  return 0;
}

static int rows_core(zvb200_ctx *c, zvb200_store *s, const void *queries_dev, uint32_t q, const uint32_t *rows_dev,
                     uint32_t m, uint32_t k, float threshold, const uint32_t *filter_dev, float *out_scores_dev,
                     uint32_t *out_index_dev) {
  return zvb_rows_core(c, s, queries_dev, q, rows_dev, m, k, threshold, filter_dev, out_scores_dev, out_index_dev, nullptr);
}

// fp32 queries on the device -> the store's query layout in c->qin (the reformer the store's type implies)
static int reform_core(zvb200_ctx *c, zvb200_store *s, const float *in_dev, uint32_t q, uint32_t dim) {
  const int kind = reform_kind(s->dtype, s->metric);
  if (kind < 0) return set_err(ZVB200_ERR_UNSUPPORTED, "no fp32 query reformer for this store type");
  if (dim != s->d) return set_err(ZVB200_ERR_MISMATCH, "query dimension does not match the store");
  CU(c->qin.ensure((size_t)q * s->elem_bytes));
  cudaError_t e = launch_reform(in_dev, q, dim, kind, (uint8_t *)c->qin.p, s->elem_bytes, c->stream);
  if (e == cudaErrorInvalidValue) return set_err(ZVB200_ERR_UNSUPPORTED, "reformer: dimension too large");
  CU(e);
  return 0;
}

int zvb_check_search_args(zvb200_ctx *c, const void *queries, uint32_t q, uint32_t k) {
  cudaGetLastError();  // a stale non-sticky error of an unrelated call must not fail this search
  if (!c || (!queries && q)) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  if (k == 0 || k > ZVB200_MAX_TOPK) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "topk must be 1..1024");
  if (!c->s->sealed) return set_err(ZVB200_ERR_NOREADY, "store not sealed");
  return 0;
}

extern "C" {

int zvb200_search_device(zvb200_ctx *c, const void *queries_dev, uint32_t q, uint32_t k,
                         float threshold, const uint32_t *filter_dev, float *out_scores_dev,
                         uint32_t *out_index_dev) {
  int r = zvb_check_search_args(c, queries_dev, q, k);
  if (r != 0) return r;
  if (q == 0) return 0;
  if (!c->s->shards.empty())
    return zvb_sharded_search(c, queries_dev, false, q, k, threshold, filter_dev, nullptr, 0, out_scores_dev, out_index_dev);
  CU(cudaSetDevice(c->s->device));
  const uint64_t rows = c->row_limit ? std::min<uint64_t>(c->row_limit, c->s->n) : c->s->n;
  GraphKey key = graph_key(c, 4, q, k, threshold, rows);
  key.p0 = queries_dev; key.p1 = filter_dev; key.p2 = out_scores_dev; key.p3 = out_index_dev;
  return run_replayable(c, key, q <= kGraphMaxQueries, [&]() -> int {
    return search_core(c, c->s, queries_dev, q, k, threshold, filter_dev, out_scores_dev, out_index_dev, 0, rows);
  });
}

int zvb200_search(zvb200_ctx *c, const void *queries, uint32_t q, uint32_t k, float threshold,
                  const uint32_t *filter, uint64_t *out_keys, float *out_scores,
                  uint32_t *out_index, uint32_t *out_counts) {
  if (!c) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  std::shared_lock<std::shared_mutex> rl(c->s->mu);
  int r = zvb_check_search_args(c, queries, q, k);
  if (r != 0) return r;
  if (!out_scores) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null out_scores");
  if (q == 0) return 0;
  zvb200_store *s = c->s;
  CU(cudaSetDevice(s->device));
  const size_t qbytes = (size_t)q * s->elem_bytes, nres = (size_t)q * k;
  CU(c->out_scores.ensure(nres * 4));
  CU(c->out_index.ensure(nres * 4));
  CU(c->h_scores.ensure(nres * 4));
  CU(c->h_index.ensure(nres * 4));
  if (!s->shards.empty()) {
    r = zvb_sharded_search(c, queries, true, q, k, threshold, filter, nullptr, 0, (float *)c->out_scores.p,
                           (uint32_t *)c->out_index.p);
    if (r != 0) return r;
    return finish_results(c, s, q, k, out_keys, out_scores, out_index, out_counts);
  }
  // rows this search sees: all of them, or the snapshot the caller took (zvb200_ctx_set_row_limit); a filter
  // bitmap covers exactly those rows
  const uint64_t rows = c->row_limit ? std::min<uint64_t>(c->row_limit, s->n) : s->n;
  CU(c->h_q.ensure(qbytes));
  CU(c->qin.ensure(qbytes));
  memcpy(c->h_q.p, queries, qbytes);
  r = run_replayable(c, graph_key(c, 1, q, k, threshold, rows), !filter && q <= kGraphMaxQueries, [&]() -> int {
    CU(cudaMemcpyAsync(c->qin.p, c->h_q.p, qbytes, cudaMemcpyHostToDevice, c->stream));
    const uint32_t *fdev = nullptr;
    if (filter) {
      const size_t fb = ((rows + 31) / 32) * 4;
      CU(c->filter.ensure(fb));
      CU(cudaMemcpyAsync(c->filter.p, filter, fb, cudaMemcpyHostToDevice, c->stream));
      fdev = (const uint32_t *)c->filter.p;
    }
    const int rr = search_core(c, s, c->qin.p, q, k, threshold, fdev, (float *)c->out_scores.p,
                               (uint32_t *)c->out_index.p, 0, rows);
    return rr != 0 ? rr : queue_result_copies(c, q, k);
  });
  if (r != 0) return r;
  return finish_results(c, s, q, k, out_keys, out_scores, out_index, out_counts, true);
}

int zvb200_merge_topk_device(zvb200_ctx *c, const float *scores_dev, const uint32_t *index_dev,
                             uint32_t g, const uint64_t *shard_base, uint32_t q, uint32_t k,
                             float *out_scores_dev, uint64_t *out_rows_dev) {
  if (!c || !scores_dev || !index_dev || !shard_base || g == 0 || k == 0)
    return set_err(ZVB200_ERR_INVALID_ARGUMENT, "bad argument");
  CU(cudaSetDevice(c->s->device));
  CU(c->misc.ensure((size_t)g * 8));
  CU(cudaMemcpyAsync(c->misc.p, shard_base, (size_t)g * 8, cudaMemcpyHostToDevice, c->stream));
  CU(launch_merge_shards(scores_dev, index_dev, g, (const uint64_t *)c->misc.p, q, k, out_scores_dev,
                         out_rows_dev, c->stream));
  return 0;
}

// fusion: H2D of the lists, one kernel, D2H of the fused lists
static int fuse_common(zvb200_ctx *c, const uint64_t *keys, const float *scores, const uint32_t *counts,
                       uint32_t F, uint32_t q, uint32_t k, uint32_t topn, uint32_t rank_constant, int mode,
                       const double *weights, int negate, uint64_t *out_keys, double *out_scores,
                       uint32_t *out_counts) {
  if (!c || !keys || !counts || !out_keys || !out_scores || !out_counts || F == 0 || k == 0 || topn == 0)
    return set_err(ZVB200_ERR_INVALID_ARGUMENT, "bad argument");
  if (mode != 0 && (!scores || !weights)) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "weighted fusion needs scores and weights");
  if ((size_t)F * k > fuse_max_entries()) return set_err(ZVB200_ERR_UNSUPPORTED, "fusion: nfields * k > 1536");
  if (q == 0) return 0;
  CU(cudaSetDevice(c->device));
  const size_t nin = (size_t)F * q * k, nout = (size_t)q * topn;
  const size_t off_scores = nin * 8, off_counts = off_scores + nin * 4, off_w = (off_counts + (size_t)F * q * 4 + 7) & ~(size_t)7;
  const size_t off_ok = off_w + (size_t)F * 8, off_os = off_ok + nout * 8, off_oc = off_os + nout * 8;
  CU(c->misc.ensure(off_oc + (size_t)q * 4));
  uint8_t *d = (uint8_t *)c->misc.p;
  CU(cudaMemcpyAsync(d, keys, nin * 8, cudaMemcpyHostToDevice, c->stream));
  if (mode != 0) {
    CU(cudaMemcpyAsync(d + off_scores, scores, nin * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d + off_w, weights, (size_t)F * 8, cudaMemcpyHostToDevice, c->stream));
  }
  CU(cudaMemcpyAsync(d + off_counts, counts, (size_t)F * q * 4, cudaMemcpyHostToDevice, c->stream));
  CU(launch_fuse((const uint64_t *)d, (const float *)(d + off_scores), (const uint32_t *)(d + off_counts),
                 (const double *)(d + off_w), F, q, k, topn, rank_constant, mode, negate, (uint64_t *)(d + off_ok),
                 (double *)(d + off_os), (uint32_t *)(d + off_oc), c->stream));
  CU(cudaMemcpyAsync(out_keys, d + off_ok, nout * 8, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(out_scores, d + off_os, nout * 8, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(out_counts, d + off_oc, (size_t)q * 4, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  c->last_kernel = "fuse";
  return 0;
}

int zvb200_fuse_rrf(zvb200_ctx *c, const uint64_t *keys, const uint32_t *counts, uint32_t nfields, uint32_t q,
                    uint32_t k, uint32_t topn, uint32_t rank_constant, uint64_t *out_keys, double *out_scores,
                    uint32_t *out_counts) {
  return fuse_common(c, keys, nullptr, counts, nfields, q, k, topn, rank_constant, 0, nullptr, 0, out_keys,
                     out_scores, out_counts);
}

int zvb200_fuse_weighted(zvb200_ctx *c, const uint64_t *keys, const float *scores, const uint32_t *counts,
                         uint32_t nfields, uint32_t q, uint32_t k, uint32_t topn, int fuse_metric,
                         const double *weights, int negate, uint64_t *out_keys, double *out_scores,
                         uint32_t *out_counts) {
  int mode;
  if (fuse_metric == ZVB200_METRIC_L2 || fuse_metric == ZVB200_METRIC_L2SQ) mode = 1;
  else if (fuse_metric == ZVB200_METRIC_IP) mode = 2;
  else if (fuse_metric == ZVB200_METRIC_COSINE) mode = 3;
  else return set_err(ZVB200_ERR_UNSUPPORTED, "weighted fusion: L2, InnerProduct or Cosine only");  // ValueError in the reference
  return fuse_common(c, keys, scores, counts, nfields, q, k, topn, 0, mode, weights, negate, out_keys, out_scores,
                     out_counts);
}

int zvb200_distance_matrix(int device, int metric, int dtype, uint32_t dim, const void *m,
                           uint64_t n, const void *q, uint32_t nq, float *out) {
  if (!m || !q || !out) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  zvb200_store *s = nullptr;
  int r = zvb200_store_create(device, dtype, metric, dim, &s);
  if (r != 0) return r;
  zvb200_ctx *c = nullptr;
  r = zvb200_store_append(s, m, n, nullptr);
  if (r == 0) r = zvb200_ctx_create(s, &c);
  if (r == 0 && n && nq) {
    auto body = [&]() -> int {
      const uint32_t esz = dtype == ZVB200_DT_FP32 ? 4 : dtype == ZVB200_DT_FP16 ? 2 : 1;
      CU(c->qin.ensure((size_t)nq * s->elem_bytes));
      CU(c->qf32.ensure((size_t)nq * (s->chunks * 16 / esz) * 4));
      CU(c->qraw.ensure((size_t)nq * s->row_stride));
      CU(c->qtails.ensure((size_t)nq * sizeof(float4)));
      CU(c->qnorms.ensure((size_t)nq * 4));
      CU(c->misc.ensure((size_t)nq * n * 4));
      CU(cudaMemcpyAsync(c->qin.p, q, (size_t)nq * s->elem_bytes, cudaMemcpyHostToDevice, c->stream));
      RowsView rv = s->view();
      CU(launch_prep_queries(rv, c->qin.p, nq, s->elem_bytes, (float *)c->qf32.p, (uint8_t *)c->qraw.p,
                             (float4 *)c->qtails.p, (float *)c->qnorms.p, c->stream));
      QueryView qv;
      qv.qf32 = (const float *)c->qf32.p; qv.qraw = (const uint8_t *)c->qraw.p;
      qv.qtails = (const float4 *)c->qtails.p; qv.qnorms = (const float *)c->qnorms.p; qv.q = nq;
      CU(launch_distance_matrix(rv, qv, (float *)c->misc.p, c->stream));
      CU(cudaMemcpyAsync(out, c->misc.p, (size_t)nq * n * 4, cudaMemcpyDeviceToHost, c->stream));
      CU(cudaStreamSynchronize(c->stream));
      return 0;
    };
    r = body();
  }
  if (c) zvb200_ctx_destroy(c);
  zvb200_store_destroy(s);
  return r;
}

int zvb200_ivf_attach(zvb200_store *s, const void *centroids, uint32_t nlist, const uint64_t *list_offsets) {
  if (!s) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  return zvb200_ivf_attach_typed(s, centroids, nlist, list_offsets, s->dtype, s->metric, s->dim);
}

int zvb200_ivf_attach_typed(zvb200_store *s, const void *centroids, uint32_t nlist, const uint64_t *list_offsets,
                            int c_dtype, int c_metric, uint32_t c_dim) {
  if (!s || !centroids || !list_offsets || nlist == 0) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  if (!s->shards.empty()) return set_err(ZVB200_ERR_UNSUPPORTED, "IVF over a sharded store: shard the lists per process (zvb200_ivf_set_budget_sizes)");
  if (!s->sealed) return set_err(ZVB200_ERR_NOREADY, "store not sealed");
  if (list_offsets[0] != 0 || list_offsets[nlist] != s->n) return set_err(ZVB200_ERR_MISMATCH, "list offsets do not cover the rows");
  for (uint32_t l = 0; l < nlist; ++l)
    if (list_offsets[l] > list_offsets[l + 1]) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "list offsets not ascending");
  // list offsets travel as u32 on the device: the store's own row limit (check_row_limit) keeps them exact
  if (int r = zvb_check_row_limit(list_offsets[nlist])) return r;
  zvb200_store *cs = nullptr;
  int r = zvb200_store_create(s->device, c_dtype, c_metric, c_dim, &cs);
  if (r == 0) r = zvb200_store_append(cs, centroids, nlist, nullptr);
  if (r == 0) r = zvb200_store_seal(cs);
  uint32_t *d = nullptr;
  if (r == 0) {
    std::vector<uint32_t> off32(nlist + 1);
    for (uint32_t l = 0; l <= nlist; ++l) off32[l] = (uint32_t)list_offsets[l];
    cudaError_t e = cudaSetDevice(s->device);
    if (e == cudaSuccess) e = cudaMalloc((void **)&d, (size_t)(nlist + 1) * 4);
    if (e == cudaSuccess) e = cudaMemcpy(d, off32.data(), (size_t)(nlist + 1) * 4, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) r = cuda_fail(e, "ivf_attach");
  }
  if (r != 0) {  // nothing attached, nothing leaked
    if (d) cudaFree(d);
    if (cs) zvb200_store_destroy(cs);
    return r;
  }
  zvb200_store::Write g(s);
  if (s->centroids) zvb200_store_destroy(s->centroids);
  if (s->list_off_dev) cudaFree(s->list_off_dev);
  if (s->budget_dev) { cudaFree(s->budget_dev); s->budget_dev = nullptr; }
  if (s->gbase_dev) { cudaFree(s->gbase_dev); s->gbase_dev = nullptr; }
  if (s->row_scale) { cudaFree(s->row_scale); s->row_scale = nullptr; }  // per-list scales belong to the old lists
  s->centroids = cs; s->list_off_dev = d; s->nlist = nlist;
  s->coarse_typed = c_dtype != s->dtype || c_metric != s->metric || c_dim != s->dim;
  s->list_offsets.assign(list_offsets, list_offsets + nlist + 1);
  return 0;
}

int zvb200_ivf_set_budget_sizes(zvb200_store *s, const uint64_t *list_sizes) {
  if (!s || !s->centroids) return set_err(ZVB200_ERR_NOREADY, "no IVF structure attached");
  zvb200_store::Write g(s);
  CU(cudaSetDevice(s->device));
  if (s->budget_dev) { cudaFree(s->budget_dev); s->budget_dev = nullptr; }
  if (!list_sizes) return 0;
  std::vector<uint32_t> b(s->nlist);
  for (uint32_t l = 0; l < s->nlist; ++l) b[l] = (uint32_t)list_sizes[l];
  CU(cudaMalloc((void **)&s->budget_dev, (size_t)s->nlist * 4));
  CU(cudaMemcpy(s->budget_dev, b.data(), (size_t)s->nlist * 4, cudaMemcpyHostToDevice));
  return 0;
}

// top-nprobe centroids for nprobe > ZVB200_MAX_TOPK: distance matrix (the 1x1 kernels' arithmetic) + one CTA-wide
// bitonic sort of (score, centroid) keys per query; leaves coarse_scores / coarse_index and the prepared queries
// (c->last_qv) exactly as search_core would
static int coarse_wide(zvb200_ctx *c, zvb200_store *cs, const void *queries_dev, uint32_t q, uint32_t nprobe) {
  QueryView qv;
  int r = prep_core(c, cs, queries_dev, q, &qv);
  if (r != 0) return r;
  RowsView rv = cs->view();
  const uint64_t n = cs->n;
  uint64_t p2 = 1;
  while (p2 < n) p2 <<= 1;
  // queries in groups so that the distance matrix + sort scratch stay below 512 MiB
  const size_t per_query = (size_t)n * 4 + (size_t)p2 * 8;
  const uint32_t group = (uint32_t)std::max<size_t>(1, std::min<size_t>(q, ((size_t)512 << 20) / per_query));
  CU(c->misc.ensure((size_t)group * per_query));
  float *dist = (float *)((uint8_t *)c->misc.p + (size_t)group * p2 * 8);
  for (uint32_t q0 = 0; q0 < q; q0 += group) {
    const uint32_t qn = std::min(group, q - q0);
    QueryView g = qv;
    g.q = qn;
    const uint32_t esz = cs->dtype == ZVB200_DT_FP32 ? 4 : cs->dtype == ZVB200_DT_FP16 ? 2 : 1;
    if (g.qf32) g.qf32 += (size_t)q0 * (cs->chunks * 16 / esz);
    if (g.qraw) g.qraw += (size_t)q0 * cs->row_stride;
    g.qtails += q0; g.qnorms += q0;
    CU(launch_distance_matrix(rv, g, dist, c->stream));
    CU(launch_sort_rows(dist, qn, (uint32_t)n, (uint32_t)p2, nprobe, (u64 *)c->misc.p,
                        (float *)c->coarse_scores.p + (size_t)q0 * nprobe,
                        (uint32_t *)c->coarse_index.p + (size_t)q0 * nprobe, c->stream));
  }
  c->last_kernel = "coarse_wide";
  return 0;
}

// top-nprobe centroids for a FEW queries: the streaming scan writes every centroid's key, one select kernel per query
// picks and sorts the nprobe best; same arithmetic, same (score, row) order as search_core
static int coarse_direct(zvb200_ctx *c, zvb200_store *cs, const void *queries_dev, uint32_t q, uint32_t nprobe,
                         const IvfPlanArgs *plan) {
  QueryView qv;
  int r = prep_core(c, cs, queries_dev, q, &qv, false);
  if (r != 0) return r;
  RowsView rv = zvb_view_for(c, cs);
  const uint32_t total = (uint32_t)cs->n;
  CU(c->part.ensure((size_t)q * total * sizeof(u64)));
  SearchParams sp;
  sp.k = nprobe; sp.threshold = FLT_MAX; sp.filter = nullptr;
  uint32_t nparts = 0;
  CU(launch_scan_gemv(rv, qv, sp, 0, cs->n, nullptr, &nparts, cs->sm_count, c->stream, (u64 *)c->part.p, total));
  CU(launch_select_topk((const u64 *)c->part.p, total, q, nprobe, (float *)c->coarse_scores.p,
                        (uint32_t *)c->coarse_index.p, c->stream, plan));
  c->last_kernel = "scan_gemv";
  return 0;
}

// coarse search -> list plan -> posting-list scan -> merge, everything on the context's stream; queries, filter and
// outputs on the device
static int ivf_core(zvb200_ctx *c, const void *queries_dev, uint32_t q, uint32_t k, uint32_t nprobe,
                    uint64_t max_scan_count, float threshold, const uint32_t *filter_dev, float *out_scores_dev,
                    uint32_t *out_index_dev, const void *list_queries_dev = nullptr) {
  zvb200_store *s = c->s;
  if (s->coarse_typed && !list_queries_dev)
    return set_err(ZVB200_ERR_MISMATCH, "centroids in a layout of their own: search with zvb200_ivf_search_split");
  CU(c->coarse_scores.ensure((size_t)q * nprobe * 4));
  CU(c->coarse_index.ensure((size_t)q * nprobe * 4));
  CU(c->seg_prefix.ensure((size_t)q * (nprobe + 1) * 4));
  CU(c->seg_row.ensure((size_t)q * nprobe * 4));
  CU(c->seg_count.ensure((size_t)q * 4));
  // 1. coarse: top-nprobe centroids per query (IVFCentroidIndex::search, ivf_centroid_index.cc:273-295)
  const int keep = c->force_kernel;
  if (keep == 3 || keep == 5) c->force_kernel = 0;
  int r;
  bool planned = false;
  static const bool direct_off = [] { const char *e = getenv("ZVB200_COARSE_DIRECT"); return e && atoi(e) == 0; }();
  if (!direct_off && q <= 4 && keep == 0 && nprobe <= ZVB200_MAX_TOPK && s->centroids->n <= select_topk_max() &&
      !s->centroids->has_drop) {
    // few queries over few centroids: a CTA of the streaming scan sees ~55 rows and k = nprobe takes them all, so the
    // in-kernel top-k is pure overhead — every centroid's key goes straight to the select kernel, which also writes
    // the query's posting-list plan (q <= 4 is query-major: no seg_list)
    IvfPlanArgs pa;
    pa.list_off = s->list_off_dev; pa.budget_sizes = s->budget_dev; pa.max_scan_count = max_scan_count;
    pa.seg_prefix = (uint32_t *)c->seg_prefix.p; pa.seg_row = (uint32_t *)c->seg_row.p; pa.seg_count = (uint32_t *)c->seg_count.p;
    r = coarse_direct(c, s->centroids, queries_dev, q, nprobe, &pa);
    planned = true;
  } else if (nprobe <= ZVB200_MAX_TOPK) {
    r = search_core(c, s->centroids, queries_dev, q, nprobe, FLT_MAX, nullptr, (float *)c->coarse_scores.p,
                    (uint32_t *)c->coarse_index.p, 0, s->centroids->n);
  } else {
    // wide probing (the reference has no cap: nprobe = round(nlist * scan_ratio), ivf_searcher_context.h:61-79;
    // nlist 16384 at the default ratio 0.1 = 1638 lists): every centroid's distance, then a full sort per query
    r = coarse_wide(c, s->centroids, queries_dev, q, nprobe);
  }
  c->force_kernel = keep;
  if (r != 0) return r;
  // 2. which lists each query scans under the budget (ivf_searcher.cc:216-231)
  RowsView rv = zvb_view_for(c, s);
  // batches scan LIST-major: a list probed by several queries of the batch is streamed once for up to 8 of them
  // (query-major re-reads it per query: C4 at batch 1024 streamed 246 GB per batch out of a 61 GB index)
  static const uint32_t lm_from = [] { const char *e = getenv("ZVB200_IVF_LM_FROM"); return e ? (uint32_t)atoi(e) : 8u; }();
  const bool list_major = c->force_kernel != 1 && c->force_kernel != 4 && q >= lm_from && scan_lists_lm_qb(rv, k) > 0 &&
                          (size_t)q * nprobe * k * sizeof(u64) <= ((size_t)1 << 30);
  if (list_major) CU(c->seg_list.ensure((size_t)q * nprobe * 4));
  if (!planned || list_major)
  CU(launch_ivf_plan((const uint32_t *)c->coarse_index.p, q, nprobe, s->list_off_dev, s->budget_dev, max_scan_count,
                     (uint32_t *)c->seg_prefix.p, (uint32_t *)c->seg_row.p, (uint32_t *)c->seg_count.p, c->stream,
                     list_major ? (uint32_t *)c->seg_list.p : nullptr));
  // 3. posting-list scan, queries already in compute layout (same dtype / metric / dim as the centroids) — or, for
  // split queries, prepared now in the lists' layout (the coarse search is done with the context's query buffers)
  if (list_queries_dev) {
    QueryView lqv;
    const int pr = prep_core(c, s, list_queries_dev, q, &lqv, false);
    if (pr != 0) return pr;
  }
  SearchParams sp;
  sp.k = k; sp.threshold = threshold; sp.filter = filter_dev;
  if (list_major) {
    CU(c->lm.ensure(scan_lists_lm_scratch_words(q, nprobe, s->nlist) * 4));
    CU(c->part.ensure((size_t)q * nprobe * k * sizeof(u64)));
    uint32_t nparts = 0;
    CU(zvb_scan_event(c, c->ev_scan0));
    c->scan_timed = true;
    CU(launch_scan_lists_lm(rv, c->last_qv, sp, (const uint32_t *)c->seg_list.p, (const uint32_t *)c->seg_count.p, nprobe,
                            s->list_off_dev, s->nlist, (uint32_t *)c->lm.p, (u64 *)c->part.p, &nparts, s->sm_count,
                            c->stream));
    CU(zvb_scan_event(c, c->ev_scan1));
    c->last_kernel = "scan_lists_lm";
    CU(launch_merge_parts((const u64 *)c->part.p, nparts, q, k, out_scores_dev, out_index_dev, nullptr, c->stream));
    return 0;
  }
  const bool lists_stream = c->force_kernel == 4 && !rv.scaled() && scan_stream_supports(rv, k, nprobe);
  const size_t parts = lists_stream ? scan_lists_stream_parts(q, s->sm_count) : scan_lists_parts(q, s->sm_count);
  CU(c->part.ensure((size_t)q * parts * k * sizeof(u64)));
  uint32_t nparts = 0;
  CU(zvb_scan_event(c, c->ev_scan0));
  c->scan_timed = true;
  if (lists_stream)
    CU(launch_scan_lists_stream(rv, c->last_qv, sp, (const uint32_t *)c->seg_prefix.p, (const uint32_t *)c->seg_row.p,
                                (const uint32_t *)c->seg_count.p, nprobe, (u64 *)c->part.p, &nparts, s->sm_count,
                                c->stream));
  else
    CU(launch_scan_lists(rv, c->last_qv, sp, (const uint32_t *)c->seg_prefix.p, (const uint32_t *)c->seg_row.p,
                         (const uint32_t *)c->seg_count.p, nprobe, (u64 *)c->part.p, &nparts, s->sm_count, c->stream,
                         // what one query scans: the budget plus, on average, half of the list that crosses it
                         std::min<uint64_t>(s->n, max_scan_count + s->n / std::max<uint32_t>(1, s->nlist) / 2)));
  CU(zvb_scan_event(c, c->ev_scan1));
  c->last_kernel = lists_stream ? "scan_lists_stream" : "scan_lists";
  CU(launch_merge_parts((const u64 *)c->part.p, nparts, q, k, out_scores_dev, out_index_dev, nullptr, c->stream));
  return 0;
}

static int ivf_check(zvb200_ctx *c, const void *queries, uint32_t q, uint32_t k, uint32_t *nprobe) {
  int r = zvb_check_search_args(c, queries, q, k);
  if (r != 0) return r;
  if (!c->s->centroids) return set_err(ZVB200_ERR_NOREADY, "no IVF structure attached");
  if (*nprobe == 0) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "nprobe must be >= 1");
  *nprobe = std::min(*nprobe, c->s->nlist);  // no upper cap, like IVFSearcherContext::update (ivf_searcher_context.h:61-79)
  return 0;
}

int zvb200_ivf_search(zvb200_ctx *c, const void *queries, uint32_t q, uint32_t k, uint32_t nprobe,
                      uint64_t max_scan_count, float threshold, const uint32_t *filter_bitmap, uint64_t *out_keys,
                      float *out_scores, uint32_t *out_index, uint32_t *out_counts) {
  if (!c) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  std::shared_lock<std::shared_mutex> rl(c->s->mu);
  int r = ivf_check(c, queries, q, k, &nprobe);
  if (r != 0) return r;
  if (!out_scores) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null out_scores");
  zvb200_store *s = c->s;
  if (q == 0) return 0;
  CU(cudaSetDevice(s->device));
  const size_t qbytes = (size_t)q * s->elem_bytes, nres = (size_t)q * k;
  CU(c->h_q.ensure(qbytes));
  CU(c->qin.ensure(qbytes));
  CU(c->out_scores.ensure(nres * 4));
  CU(c->out_index.ensure(nres * 4));
  CU(c->h_scores.ensure(nres * 4));
  CU(c->h_index.ensure(nres * 4));
  memcpy(c->h_q.p, queries, qbytes);
  r = run_replayable(c, graph_key(c, 2, q, k, threshold, s->n, nprobe, max_scan_count),
                     !filter_bitmap && q <= kGraphMaxQueries, [&]() -> int {
    CU(cudaMemcpyAsync(c->qin.p, c->h_q.p, qbytes, cudaMemcpyHostToDevice, c->stream));
    const uint32_t *fdev = nullptr;
    if (filter_bitmap) {
      const size_t fb = (size_t)((s->n + 31) / 32) * 4;
      CU(c->filter.ensure(fb));
      CU(cudaMemcpyAsync(c->filter.p, filter_bitmap, fb, cudaMemcpyHostToDevice, c->stream));
      fdev = (const uint32_t *)c->filter.p;
    }
    const int rr = ivf_core(c, c->qin.p, q, k, nprobe, max_scan_count, threshold, fdev, (float *)c->out_scores.p,
                            (uint32_t *)c->out_index.p);
    return rr != 0 ? rr : queue_result_copies(c, q, k);
  });
  if (r != 0) return r;
  return finish_results(c, s, q, k, out_keys, out_scores, out_index, out_counts, true);
}

int zvb200_ivf_search_device(zvb200_ctx *c, const void *queries_dev, uint32_t q, uint32_t k, uint32_t nprobe,
                             uint64_t max_scan_count, float threshold, const uint32_t *filter_dev,
                             float *out_scores_dev, uint32_t *out_index_dev) {
  if (!c) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  int r = ivf_check(c, queries_dev, q, k, &nprobe);
  if (r != 0) return r;
  if (!out_scores_dev || !out_index_dev) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null output");
  if (q == 0) return 0;
  CU(cudaSetDevice(c->s->device));
  GraphKey key = graph_key(c, 5, q, k, threshold, c->s->n, nprobe, max_scan_count);
  key.p0 = queries_dev; key.p1 = filter_dev; key.p2 = out_scores_dev; key.p3 = out_index_dev;
  return run_replayable(c, key, q <= kGraphMaxQueries, [&]() -> int {
    return ivf_core(c, queries_dev, q, k, nprobe, max_scan_count, threshold, filter_dev, out_scores_dev, out_index_dev);
  });
}

static const uint32_t *upload_filter(zvb200_ctx *c, zvb200_store *s, const uint32_t *filter, int *err);

int zvb200_ivf_search_split(zvb200_ctx *c, const void *coarse_queries, const void *list_queries, uint32_t q, uint32_t k,
                            uint32_t nprobe, uint64_t max_scan_count, float threshold, const uint32_t *filter_bitmap,
                            uint64_t *out_keys, float *out_scores, uint32_t *out_index, uint32_t *out_counts) {
  if (!c) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  std::shared_lock<std::shared_mutex> rl(c->s->mu);
  int r = ivf_check(c, list_queries, q, k, &nprobe);
  if (r != 0) return r;
  if (!out_scores || (!coarse_queries && q)) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  zvb200_store *s = c->s;
  if (q == 0) return 0;
  CU(cudaSetDevice(s->device));
  const size_t cbytes = (size_t)q * s->centroids->elem_bytes, lbytes = (size_t)q * s->elem_bytes, nres = (size_t)q * k;
  CU(c->h_q.ensure(cbytes));
  CU(c->h_q2.ensure(lbytes));
  CU(c->qin.ensure(cbytes));
  CU(c->qin2.ensure(lbytes));
  CU(c->out_scores.ensure(nres * 4));
  CU(c->out_index.ensure(nres * 4));
  CU(c->h_scores.ensure(nres * 4));
  CU(c->h_index.ensure(nres * 4));
  memcpy(c->h_q.p, coarse_queries, cbytes);
  memcpy(c->h_q2.p, list_queries, lbytes);
  r = run_replayable(c, graph_key(c, 3, q, k, threshold, s->n, nprobe, max_scan_count),
                     !filter_bitmap && q <= kGraphMaxQueries, [&]() -> int {
    CU(cudaMemcpyAsync(c->qin.p, c->h_q.p, cbytes, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->qin2.p, c->h_q2.p, lbytes, cudaMemcpyHostToDevice, c->stream));
    int fe = 0;
    const uint32_t *fdev = upload_filter(c, s, filter_bitmap, &fe);
    if (fe != 0) return fe;
    const int rr = ivf_core(c, c->qin.p, q, k, nprobe, max_scan_count, threshold, fdev, (float *)c->out_scores.p,
                            (uint32_t *)c->out_index.p, c->qin2.p);
    return rr != 0 ? rr : queue_result_copies(c, q, k);
  });
  if (r != 0) return r;
  return finish_results(c, s, q, k, out_keys, out_scores, out_index, out_counts, true);
}

int zvb200_store_set_score_scale(zvb200_store *s, float scale, const float *list_scales) {
  if (!s) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null store");
  if (!s->shards.empty()) return set_err(ZVB200_ERR_UNSUPPORTED, "not available on a sharded store");
  if (!(scale == scale) || (list_scales && !s->centroids))
    return set_err(list_scales ? ZVB200_ERR_NOREADY : ZVB200_ERR_INVALID_ARGUMENT,
                   list_scales ? "per-list scales need an attached IVF structure" : "scale is NaN");
  zvb200_store::Write g(s);
  CU(cudaSetDevice(s->device));
  if (s->row_scale) { cudaFree(s->row_scale); s->row_scale = nullptr; }
  s->score_scale = scale;
  if (!list_scales || s->n == 0) return 0;
  float *ls = nullptr, *rs = nullptr;
  cudaError_t e = cudaMalloc((void **)&ls, (size_t)s->nlist * 4);
  if (e == cudaSuccess) e = cudaMalloc((void **)&rs, (size_t)s->n * 4);
  if (e == cudaSuccess) e = cudaMemcpy(ls, list_scales, (size_t)s->nlist * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = launch_expand_list_scale(s->list_off_dev, s->nlist, ls, s->n, rs, 0);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (ls) cudaFree(ls);
  if (e != cudaSuccess) {
    if (rs) cudaFree(rs);
    return cuda_fail(e, "set_score_scale");
  }
  s->row_scale = rs;
  return 0;
}

int zvb200_ivf_set_global_bases(zvb200_store *s, const uint64_t *list_global_base) {
  if (!s || !s->centroids) return set_err(ZVB200_ERR_NOREADY, "no IVF structure attached");
  zvb200_store::Write g(s);
  CU(cudaSetDevice(s->device));
  if (s->gbase_dev) { cudaFree(s->gbase_dev); s->gbase_dev = nullptr; }
  if (!list_global_base) return 0;
  CU(cudaMalloc((void **)&s->gbase_dev, (size_t)s->nlist * 8));
  CU(cudaMemcpy(s->gbase_dev, list_global_base, (size_t)s->nlist * 8, cudaMemcpyHostToDevice));
  return 0;
}

int zvb200_ivf_global_rows_device(zvb200_ctx *c, const uint32_t *index_dev, uint64_t count, uint64_t *out_rows_dev) {
  if (!c || !index_dev || !out_rows_dev) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  zvb200_store *s = c->s;
  if (!s->centroids || !s->gbase_dev) return set_err(ZVB200_ERR_NOREADY, "no global list bases set");
  CU(cudaSetDevice(s->device));
  CU(launch_ivf_global_rows(index_dev, count, s->list_off_dev, s->gbase_dev, s->nlist, out_rows_dev, c->stream));
  return 0;
}

int zvb200_merge_topk_rows_device(zvb200_ctx *c, const float *scores_dev, const uint64_t *rows_dev, uint32_t g,
                                  uint32_t q, uint32_t k, float *out_scores_dev, uint64_t *out_rows_dev) {
  if (!c || !scores_dev || !rows_dev || !out_scores_dev || !out_rows_dev || g == 0 || k == 0)
    return set_err(ZVB200_ERR_INVALID_ARGUMENT, "bad argument");
  CU(cudaSetDevice(c->s->device));
  CU(launch_merge_rows(scores_dev, rows_dev, g, q, k, out_scores_dev, out_rows_dev, c->stream));
  return 0;
}

// ---- query reformers and the refiner --------------------------------------------------

static int upload_fp32_queries(zvb200_ctx *c, const float *queries, uint32_t q, uint32_t dim) {
  const size_t b = (size_t)q * dim * 4;
  CU(c->h_q.ensure(b));
  CU(c->qsrc.ensure(b));
  memcpy(c->h_q.p, queries, b);
  CU(cudaMemcpyAsync(c->qsrc.p, c->h_q.p, b, cudaMemcpyHostToDevice, c->stream));
  return 0;
}

static const uint32_t *upload_filter(zvb200_ctx *c, zvb200_store *s, const uint32_t *filter, int *err) {
  *err = 0;
  if (!filter) return nullptr;
  const size_t fb = ((s->n + 31) / 32) * 4;
  cudaError_t e = c->filter.ensure(fb);
  if (e == cudaSuccess) e = cudaMemcpyAsync(c->filter.p, filter, fb, cudaMemcpyHostToDevice, c->stream);
  if (e != cudaSuccess) { *err = cuda_fail(e, "filter upload"); return nullptr; }
  return (const uint32_t *)c->filter.p;
}

int zvb200_reform_queries(zvb200_ctx *c, const float *queries_fp32, uint32_t q, uint32_t dim, void *out) {
  cudaGetLastError();
  if (!c || (!queries_fp32 && q) || (!out && q)) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  if (q == 0) return 0;
  zvb200_store *s = c->s;
  if (!s->shards.empty()) return set_err(ZVB200_ERR_UNSUPPORTED, "not available on a sharded store");
  CU(cudaSetDevice(s->device));
  int r = upload_fp32_queries(c, queries_fp32, q, dim);
  if (r == 0) r = reform_core(c, s, (const float *)c->qsrc.p, q, dim);
  if (r != 0) return r;
  const size_t ob = (size_t)q * s->elem_bytes;
  CU(c->h_scores.ensure(ob));
  CU(cudaMemcpyAsync(c->h_scores.p, c->qin.p, ob, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  memcpy(out, c->h_scores.p, ob);
  c->last_kernel = "reform";
  return 0;
}

int zvb200_reform_queries_device(zvb200_ctx *c, const float *queries_fp32_dev, uint32_t q, uint32_t dim, void *out_dev) {
  cudaGetLastError();
  if (!c || (!queries_fp32_dev && q) || (!out_dev && q)) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  if (q == 0) return 0;
  zvb200_store *s = c->s;
  if (!s->shards.empty()) return set_err(ZVB200_ERR_UNSUPPORTED, "not available on a sharded store");
  CU(cudaSetDevice(s->device));
  const int kind = reform_kind(s->dtype, s->metric);
  if (kind < 0) return set_err(ZVB200_ERR_UNSUPPORTED, "no fp32 query reformer for this store type");
  if (dim != s->d) return set_err(ZVB200_ERR_MISMATCH, "query dimension does not match the store");
  cudaError_t e = launch_reform(queries_fp32_dev, q, dim, kind, (uint8_t *)out_dev, s->elem_bytes, c->stream);
  if (e == cudaErrorInvalidValue) return set_err(ZVB200_ERR_UNSUPPORTED, "reformer: dimension too large");
  CU(e);
  c->last_kernel = "reform";
  return 0;
}

int zvb200_search_fp32(zvb200_ctx *c, const float *queries_fp32, uint32_t q, uint32_t dim, uint32_t k, float threshold,
                       const uint32_t *filter, uint64_t *out_keys, float *out_scores, uint32_t *out_index,
                       uint32_t *out_counts) {
  if (!c) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  std::shared_lock<std::shared_mutex> rl(c->s->mu);
  int r = zvb_check_search_args(c, queries_fp32, q, k);
  if (r != 0) return r;
  if (!out_scores) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null out_scores");
  if (q == 0) return 0;
  zvb200_store *s = c->s;
  if (!s->shards.empty()) return set_err(ZVB200_ERR_UNSUPPORTED, "not available on a sharded store");
  CU(cudaSetDevice(s->device));
  const size_t nres = (size_t)q * k;
  CU(c->out_scores.ensure(nres * 4));
  CU(c->out_index.ensure(nres * 4));
  r = upload_fp32_queries(c, queries_fp32, q, dim);
  if (r == 0) r = reform_core(c, s, (const float *)c->qsrc.p, q, dim);
  if (r != 0) return r;
  const uint32_t *fdev = upload_filter(c, s, filter, &r);
  if (r != 0) return r;
  r = search_core(c, s, c->qin.p, q, k, threshold, fdev, (float *)c->out_scores.p, (uint32_t *)c->out_index.p, 0, s->n);
  if (r != 0) return r;
  return finish_results(c, s, q, k, out_keys, out_scores, out_index, out_counts);
}

int zvb200_search_rows(zvb200_ctx *c, const void *queries, uint32_t q, const uint32_t *rows, uint32_t m, uint32_t k,
                       float threshold, const uint32_t *filter, uint64_t *out_keys, float *out_scores,
                       uint32_t *out_index, uint32_t *out_counts) {
  if (!c) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  std::shared_lock<std::shared_mutex> rl(c->s->mu);
  int r = zvb_check_search_args(c, queries, q, k);
  if (r != 0) return r;
  if (!out_scores || (!rows && q && m)) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  if (m == 0 || m > ZVB200_MAX_CANDIDATES) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "candidates per query must be 1..8192");
  if (q == 0) return 0;
  zvb200_store *s = c->s;
  if (!s->shards.empty()) {
    const size_t nr = (size_t)q * k;
    CU(cudaSetDevice(s->device));
    CU(c->out_scores.ensure(nr * 4));
    CU(c->out_index.ensure(nr * 4));
    r = zvb_sharded_search(c, queries, true, q, k, threshold, filter, rows, m, (float *)c->out_scores.p,
                           (uint32_t *)c->out_index.p);
    if (r != 0) return r;
    return finish_results(c, s, q, k, out_keys, out_scores, out_index, out_counts);
  }
  CU(cudaSetDevice(s->device));
  const size_t qbytes = (size_t)q * s->elem_bytes, nres = (size_t)q * k, rbytes = (size_t)q * m * 4;
  CU(c->h_q.ensure(std::max(qbytes, rbytes)));
  CU(c->qin.ensure(qbytes));
  CU(c->cand.ensure(rbytes));
  CU(c->out_scores.ensure(nres * 4));
  CU(c->out_index.ensure(nres * 4));
  memcpy(c->h_q.p, queries, qbytes);
  CU(cudaMemcpyAsync(c->qin.p, c->h_q.p, qbytes, cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));  // h_q is reused for the candidate rows
  memcpy(c->h_q.p, rows, rbytes);
  CU(cudaMemcpyAsync(c->cand.p, c->h_q.p, rbytes, cudaMemcpyHostToDevice, c->stream));
  const uint32_t *fdev = upload_filter(c, s, filter, &r);
  if (r != 0) return r;
  r = rows_core(c, s, c->qin.p, q, (const uint32_t *)c->cand.p, m, k, threshold, fdev, (float *)c->out_scores.p,
                (uint32_t *)c->out_index.p);
  if (r != 0) return r;
  return finish_results(c, s, q, k, out_keys, out_scores, out_index, out_counts);
}

int zvb200_search_refined(zvb200_ctx *coarse, zvb200_ctx *exact, const float *queries_fp32, uint32_t q, uint32_t dim,
                          uint32_t k, float scale_factor, const uint32_t *filter, uint64_t *out_keys, float *out_scores,
                          uint32_t *out_index, uint32_t *out_counts) {
  if (!coarse || !exact) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  std::shared_lock<std::shared_mutex> rl1(coarse->s->mu);
  std::shared_lock<std::shared_mutex> rl2;
  if (exact->s != coarse->s) rl2 = std::shared_lock<std::shared_mutex>(exact->s->mu);
  int r = zvb_check_search_args(coarse, queries_fp32, q, k);
  if (r == 0) r = zvb_check_search_args(exact, queries_fp32, q, k);
  if (r != 0) return r;
  if (!out_scores) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null out_scores");
  zvb200_store *sc = coarse->s, *se = exact->s;
  if (!sc->shards.empty() || !se->shards.empty()) return set_err(ZVB200_ERR_UNSUPPORTED, "not available on a sharded store");
  if (sc->device != se->device) return set_err(ZVB200_ERR_MISMATCH, "refiner: both stores must be on one device");
  if (sc->n != se->n) return set_err(ZVB200_ERR_MISMATCH, "refiner: the stores must hold the same rows");
  // Index::_get_coarse_search_topk (src/core/interface/index.cc:794-801)
  if (scale_factor == 0.f) scale_factor = 1.f;
  const double kc = floor((double)k * (double)scale_factor);
  if (!(kc >= 1.0) || kc > (double)ZVB200_MAX_TOPK) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "refiner: coarse topk must be 1..1024");
  const uint32_t m = (uint32_t)kc;
  if (q == 0) return 0;
  CU(cudaSetDevice(sc->device));
  const size_t nres = (size_t)q * k;
  // everything runs on the coarse context's stream; the exact context lends its buffers
  const cudaStream_t keep = exact->stream;
  exact->stream = coarse->stream;
  auto body = [&]() -> int {
    CU(coarse->out_scores.ensure((size_t)q * m * 4));
    CU(coarse->out_index.ensure((size_t)q * m * 4));
    CU(exact->out_scores.ensure(nres * 4));
    CU(exact->out_index.ensure(nres * 4));
    int e = upload_fp32_queries(coarse, queries_fp32, q, dim);
    if (e == 0) e = reform_core(coarse, sc, (const float *)coarse->qsrc.p, q, dim);
    if (e != 0) return e;
    const uint32_t *fdev = upload_filter(coarse, sc, filter, &e);
    if (e != 0) return e;
    // 1. coarse: floor(topk * scale_factor) rows per query from the quantised store
    e = search_core(coarse, sc, coarse->qin.p, q, m, FLT_MAX, fdev, (float *)coarse->out_scores.p,
                    (uint32_t *)coarse->out_index.p, 0, sc->n);
    if (e != 0) return e;
    // 2. exact scores of those rows in the reference store, best topk
    e = reform_core(exact, se, (const float *)coarse->qsrc.p, q, dim);
    if (e != 0) return e;
    e = rows_core(exact, se, exact->qin.p, q, (const uint32_t *)coarse->out_index.p, m, k, FLT_MAX, fdev,
                  (float *)exact->out_scores.p, (uint32_t *)exact->out_index.p);
    if (e != 0) return e;
    return finish_results(exact, se, q, k, out_keys, out_scores, out_index, out_counts);
  };
  r = body();
  exact->stream = keep;
  return r;
}

}  // extern "C"

// ---- k-means on the device: KmeansCluster::cluster / label ---------------------------------
// (src/core/algorithm/cluster/kmeans_cluster.cc:469-535, 571-610).  A Lloyd iteration = the scan with k = 1
// over a store of the current centroids (queries = the rows, 18 944 per tcgen05 launch) + the update kernels.

static void store_reset(zvb200_store *s) {
  zvb200_store::Write g(s);
  s->n = 0; s->normed = 0; s->shadowed = 0;
  s->keys.clear(); s->drop_host.clear();
  s->has_drop = false; s->drop_dirty = false; s->drop_synced = 0; s->sealed = false;
}

struct KmeansScratch {
  zvb200_store *cs = nullptr;
  zvb200_ctx *ctx = nullptr;
  float *cent = nullptr, *dist = nullptr, *rows_own = nullptr;
  uint32_t *lab = nullptr;
  double *sums = nullptr, *scores = nullptr, *sse = nullptr;
  unsigned long long *counts = nullptr;
  ~KmeansScratch() {
    if (ctx) zvb200_ctx_destroy(ctx);
    if (cs) zvb200_store_destroy(cs);
    void *p[] = {cent, dist, rows_own, lab, sums, scores, sse, counts};
    for (void *x : p) if (x) cudaFree(x);
  }
};

static int kmeans_core(int device, int metric, const float *rows_dev, const float *rows_host, uint64_t n, uint32_t dim,
                       float *centroids, uint32_t k, uint32_t max_iterations, double epsilon, uint32_t *labels,
                       uint64_t *follows, double *sse_out, uint32_t *iterations_out) {
  if ((!rows_dev && !rows_host) || !centroids || n == 0 || dim == 0 || k == 0)
    return set_err(ZVB200_ERR_INVALID_ARGUMENT, "kmeans: null or empty argument");
  if (n >= 0xffffffffull || k >= 0xffffffffu) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "kmeans: too many rows");
  if (metric != ZVB200_METRIC_L2SQ && metric != ZVB200_METRIC_L2 && metric != ZVB200_METRIC_IP)
    return set_err(ZVB200_ERR_UNSUPPORTED, "kmeans: fp32 rows with SquaredEuclidean / Euclidean / InnerProduct");
  KmeansScratch w;
  int r = zvb200_store_create(device, ZVB200_DT_FP32, metric, dim, &w.cs);
  if (r == 0) r = zvb200_store_reserve(w.cs, k);
  if (r == 0) r = zvb200_ctx_create(w.cs, &w.ctx);
  if (r != 0) return r;
  CU(cudaSetDevice(device));
  const size_t cbytes = (size_t)k * dim * sizeof(float);
  CU(cudaMalloc((void **)&w.cent, cbytes));
  CU(cudaMalloc((void **)&w.dist, n * sizeof(float)));
  CU(cudaMalloc((void **)&w.lab, n * sizeof(uint32_t)));
  CU(cudaMalloc((void **)&w.sums, (size_t)k * dim * sizeof(double)));
  CU(cudaMalloc((void **)&w.scores, (size_t)k * sizeof(double)));
  CU(cudaMalloc((void **)&w.counts, (size_t)k * sizeof(unsigned long long)));
  CU(cudaMalloc((void **)&w.sse, sizeof(double)));
  if (!rows_dev) {
    CU(cudaMalloc((void **)&w.rows_own, n * dim * sizeof(float)));
    CU(cudaMemcpy(w.rows_own, rows_host, n * dim * sizeof(float), cudaMemcpyHostToDevice));
    rows_dev = w.rows_own;
  }
  CU(cudaMemcpy(w.cent, centroids, cbytes, cudaMemcpyHostToDevice));
  zvb200_ctx *c = w.ctx;
  c->mma_group_cap = (uint32_t)w.cs->sm_count * 128u;  // one query tile per SM, every CTA walks all the centroids
  const uint64_t chunk = (uint64_t)w.cs->sm_count * 128u * 4u;
  auto assign = [&]() -> int {  // nearest centroid of every row: LinearSeeker::seek (linear_seeker.cc:20-42)
    store_reset(w.cs);
    int e = zvb200_store_append_device(w.cs, w.cent, k);
    if (e == 0) e = zvb200_store_seal(w.cs);
    for (uint64_t off = 0; e == 0 && off < n; off += chunk) {
      const uint32_t qn = (uint32_t)std::min<uint64_t>(chunk, n - off);
      e = search_core(c, w.cs, rows_dev + off * dim, qn, 1, FLT_MAX, nullptr, w.dist + off, w.lab + off, 0, k);
    }
    if (e == 0) CU(launch_kmeans_fix_labels(w.lab, w.dist, n, c->stream));
    return e;
  };
  // max_iterations == 0: nothing follows any centroid and the cost is 0 (KmeansCluster::cluster's loop does not run)
  CU(cudaMemsetAsync(w.counts, 0, (size_t)k * sizeof(unsigned long long), c->stream));
  double cost = 0.0, new_cost = 0.0;
  uint32_t it = 0;
  for (; it < max_iterations; ++it) {
    r = assign();
    if (r != 0) return r;
    CU(cudaMemsetAsync(w.sums, 0, (size_t)k * dim * sizeof(double), c->stream));
    CU(cudaMemsetAsync(w.scores, 0, (size_t)k * sizeof(double), c->stream));
    CU(cudaMemsetAsync(w.counts, 0, (size_t)k * sizeof(unsigned long long), c->stream));
    CU(launch_kmeans_accumulate(rows_dev, n, dim, w.lab, w.dist, w.sums, w.counts, w.scores, c->stream));
    CU(launch_kmeans_finalize(w.sums, w.counts, w.scores, k, dim, w.cent, w.counts, w.sse, c->stream));
    CU(cudaMemcpyAsync(&new_cost, w.sse, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    // the stopping rule of KmeansCluster::cluster (kmeans_cluster.cc:519-529)
    if (fabs(new_cost - cost) < epsilon) { ++it; break; }
    cost = new_cost;
  }
  // cluster() ends with PurgeCentroids(cents, purge_empty_) (kmeans_cluster.cc:363-383, 533): even when nothing is
  // cut, every centroid nobody follows is swapped with the last one not yet moved — the empty clusters end up at
  // the tail, and label() then runs against the centroids in THAT order
  static_assert(sizeof(unsigned long long) == sizeof(uint64_t), "follows layout");
  std::vector<uint64_t> fol(k, 0);
  CU(cudaMemcpyAsync(centroids, w.cent, cbytes, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(fol.data(), w.counts, (size_t)k * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  {  // also after zero passes: nobody follows anything then, and the reference rotates the list it was given
    bool moved = false;
    std::vector<float> tmp(dim);
    for (size_t index = 0, tamp = k; index < tamp;) {
      if (fol[index] == 0) {
        const size_t last = tamp - 1;
        if (index != last) {
          std::swap(fol[index], fol[last]);
          float *a = centroids + index * dim, *b = centroids + last * dim;
          memcpy(tmp.data(), a, dim * sizeof(float)); memcpy(a, b, dim * sizeof(float)); memcpy(b, tmp.data(), dim * sizeof(float));
          moved = true;
        }
        tamp = last;
        continue;
      }
      ++index;
    }
    if (moved) CU(cudaMemcpyAsync(w.cent, centroids, cbytes, cudaMemcpyHostToDevice, c->stream));
  }
  if (follows) memcpy(follows, fol.data(), (size_t)k * sizeof(uint64_t));
  if (labels) {  // KmeansCluster::label: against the centroids as returned
    r = assign();
    if (r != 0) return r;
    CU(cudaMemcpyAsync(labels, w.lab, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
  }
  if (sse_out) *sse_out = new_cost;
  if (iterations_out) *iterations_out = it;
  return 0;
}

extern "C" {

int zvb200_kmeans(int device, int metric, const float *rows, uint64_t n, uint32_t dim, float *centroids, uint32_t k,
                  uint32_t max_iterations, double epsilon, uint32_t *labels, uint64_t *follows, double *sse,
                  uint32_t *iterations) {
  return kmeans_core(device, metric, nullptr, rows, n, dim, centroids, k, max_iterations, epsilon, labels, follows, sse,
                     iterations);
}

int zvb200_kmeans_device(int device, int metric, const float *rows_dev, uint64_t n, uint32_t dim, float *centroids,
                         uint32_t k, uint32_t max_iterations, double epsilon, uint32_t *labels, uint64_t *follows,
                         double *sse, uint32_t *iterations) {
  return kmeans_core(device, metric, rows_dev, nullptr, n, dim, centroids, k, max_iterations, epsilon, labels, follows,
                     sse, iterations);
}

}  // extern "C"

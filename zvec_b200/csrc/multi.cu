// The multi-GPU shard / merge layer of the C ABI (SURVEY 8b "Thin C ABI", 8e; north_star: "a multi-GPU shard/merge
// layer ... Each GPU produces a local top-k, and these are merged with an NCCL gather over NVLink plus a final k-way
// merge kernel").  Replaces the per-block search + concat + std::sort + truncate of
// CombinedVectorColumnIndexer::Search (src/db/index/column/vector_column/combined_vector_column_indexer.cc:98-232).
//
// Two deployments, one exchange format — per-shard sorted lists of PACKED keys (order(score) << 32 | local row),
// merged by (score, shard, local row) = (score, global row) because shards are contiguous ascending row ranges:
//  * one process, several devices (a zvec process that loads the plugin with proxima.gpu.devices = "0,1,..."):
//    zvb200_store_create_sharded -> an ordinary zvb200_store handle whose rows live on G devices; a search launches
//    the scan on every device, each device copies its k keys per query into the root's gather buffer over NVLink
//    (peer DMA, ordered by events), the root runs merge_keys_kernel.
//  * one process per GPU (bench.py under torchrun, any multi-process server): zvb200_comm_* — rank-local store,
//    zvb200_search_sharded_device = scan -> exchange -> merge on the context's stream.  Exchange transport:
//    peer mailboxes (each rank's keys stored straight into every peer's CUDA-IPC-mapped buffer by push_keys_kernel,
//    a release flag per (slot, source); merge_keys_kernel acquires the flags — no collective launch, SURVEY §5's
//    batch-1 fast path, used at every batch size that fits the mailbox), else ONE ncclAllGather of the u64 keys.
#include <dlfcn.h>
#include <nccl.h>

#include "host.h"

using namespace zvb;

// ======================================================================================================
// kernels
// ======================================================================================================
namespace {

constexpr int kMaxPeers = 16;

struct PeerTable {
  u64 *data[kMaxPeers];  // where THIS rank's keys go on each peer (slot and source offset applied)
  u64 *flag[kMaxPeers];  // the matching arrival flag on each peer
};

__device__ __forceinline__ void st_release_sys(u64 *p, u64 v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ u64 ld_acquire_sys(const u64 *p) {
  u64 v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ u64 globaltimer_ns() {
  u64 t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// This rank's `count` keys -> every peer's mailbox (its own included), then ONE release flag per peer once every CTA
// of the grid has stored its share (last-block pattern: fence, count, the last CTA publishes).
__global__ void push_keys_kernel(const u64 *local, uint32_t count, PeerTable peers, uint32_t world, u64 seq,
                                 unsigned int *done) {
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t d = 0; d < world; ++d) {
    u64 *dst = peers.data[d];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) dst[i] = local[i];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(done, 1u);
    if (prev == gridDim.x - 1) {
      *done = 0;
      __threadfence_system();
      for (uint32_t d = 0; d < world; ++d) st_release_sys(peers.flag[d], seq);
    }
  }
}

// G sorted lists of packed keys (list g at keys + g * gstride, q * k keys, kMaxKey padded) -> per query the k best by
// (score, g, local row); rows come out as bases[g] + local row.  Rank-based like merge_shards_kernel: candidate (g, i)
// lands at i + sum_{g' < g} #{ord' <= ord} + sum_{g' > g} #{ord' < ord}.
// flags != null: wait until every source's flag has reached `seq` (the mailboxes of this slot are complete).
__global__ void merge_keys_kernel(const u64 *keys, size_t gstride, uint32_t G, const uint64_t *bases, uint32_t q,
                                  uint32_t k, float *out_scores, uint64_t *out_rows, uint32_t *out_index,
                                  const u64 *flags, u64 seq, uint32_t *status) {
  if (flags) {
    if (threadIdx.x < G) {
      const u64 t0 = globaltimer_ns();
      while (ld_acquire_sys(flags + threadIdx.x) < seq) {
        if (globaltimer_ns() - t0 > 20000000000ull) {  // 20 s: a peer died; report, do not hang the GPU
          atomicExch(status, 1u);
          break;
        }
        __nanosleep(100);
      }
    }
    __syncthreads();
  }
  const uint32_t qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (qi >= q) return;
  const int lane = threadIdx.x & 31;
  for (uint32_t i = lane; i < k; i += 32) {
    if (out_scores) out_scores[(size_t)qi * k + i] = __int_as_float(0x7fffffff);  // CUDART_NAN_F, the padding of every other merge
    if (out_rows) out_rows[(size_t)qi * k + i] = ZVB200_INVALID_KEY;
    if (out_index) out_index[(size_t)qi * k + i] = ZVB200_INVALID_INDEX;
  }
  __syncwarp();
  for (uint32_t c = lane; c < G * k; c += 32) {
    const uint32_t gi = c / k, i = c % k;
    const u64 e = keys[gi * gstride + (size_t)qi * k + i];
    if (e == kMaxKey) continue;
    const uint32_t so = (uint32_t)(e >> 32);
    uint32_t rank = i;
    for (uint32_t g2 = 0; g2 < G && rank < k; ++g2) {
      if (g2 == gi) continue;
      const u64 *l2 = keys + g2 * gstride + (size_t)qi * k;
      uint32_t lo = 0, hi = k;
      while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        const u64 m = l2[mid];
        const uint32_t mo = m == kMaxKey ? 0xffffffffu : (uint32_t)(m >> 32);
        const bool before = m != kMaxKey && (g2 < gi ? (mo <= so) : (mo < so));
        if (before) lo = mid + 1; else hi = mid;
      }
      rank += lo;
    }
    if (rank < k) {
      const uint64_t row = bases[gi] + (uint32_t)e;
      if (out_scores) out_scores[(size_t)qi * k + rank] = unorder_f32(so);
      if (out_rows) out_rows[(size_t)qi * k + rank] = row;
      if (out_index) out_index[(size_t)qi * k + rank] = (uint32_t)row;
    }
  }
}

cudaError_t launch_merge_keys(const u64 *keys, size_t gstride, uint32_t G, const uint64_t *bases, uint32_t q, uint32_t k,
                              float *out_scores, uint64_t *out_rows, uint32_t *out_index, const u64 *flags, u64 seq,
                              uint32_t *status, cudaStream_t st) {
  if (q == 0) return cudaSuccess;
  const int wpb = 4;
  merge_keys_kernel<<<(q + wpb - 1) / wpb, wpb * 32, 0, st>>>(keys, gstride, G, bases, q, k, out_scores, out_rows,
                                                             out_index, flags, seq, status);
  count_launch();
  return cudaGetLastError();
}

}  // namespace

// ======================================================================================================
// one process, several devices: the sharded parent store
// ======================================================================================================
static int enable_peers(const int *devs, uint32_t n) {
  for (uint32_t a = 0; a < n; ++a)
    for (uint32_t b = 0; b < n; ++b) {
      if (a == b || devs[a] == devs[b]) continue;
      int can = 0;
      CU(cudaDeviceCanAccessPeer(&can, devs[a], devs[b]));
      if (!can) return set_err(ZVB200_ERR_UNSUPPORTED, "sharded store: the devices cannot access each other's memory");
      CU(cudaSetDevice(devs[a]));
      const cudaError_t e = cudaDeviceEnablePeerAccess(devs[b], 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(e, "cudaDeviceEnablePeerAccess");
      cudaGetLastError();
    }
  return 0;
}

extern "C" int zvb200_store_create_sharded(const int *devices, uint32_t ndev, int dtype, int metric, uint32_t dim,
                                           zvb200_store **out) {
  if (!out || !devices || ndev == 0) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  *out = nullptr;
  if (ndev == 1) return zvb200_store_create(devices[0], dtype, metric, dim, out);
  if (ndev > kMaxPeers) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "at most 16 devices");
  for (uint32_t a = 0; a < ndev; ++a)
    for (uint32_t b = a + 1; b < ndev; ++b)
      if (devices[a] == devices[b]) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "a device is listed twice");
  zvb200_store *p = nullptr;
  int r = zvb200_store_create(devices[0], dtype, metric, dim, &p);  // the parent: layout + root device + stream
  if (r != 0) return r;
  for (uint32_t g = 0; r == 0 && g < ndev; ++g) {
    zvb200_store *sh = nullptr;
    r = zvb200_store_create(devices[g], dtype, metric, dim, &sh);
    if (r == 0) p->shards.push_back(sh);
  }
  if (r == 0) r = enable_peers(devices, ndev);
  if (r == 0) {
    cudaError_t e = cudaSetDevice(p->device);
    if (e == cudaSuccess) e = cudaMalloc((void **)&p->shard_base_dev, ndev * sizeof(uint64_t));
    if (e != cudaSuccess) r = cuda_fail(e, "create_sharded");
  }
  if (r != 0) { zvb200_store_destroy(p); return r; }
  *out = p;
  return 0;
}

int zvb_sharded_destroy(zvb200_store *s) {
  for (auto *sh : s->shards) zvb200_store_destroy(sh);
  s->shards.clear();
  if (s->shard_base_dev) { cudaSetDevice(s->device); cudaFree(s->shard_base_dev); s->shard_base_dev = nullptr; }
  return 0;
}

// The first reserve fixes the shard size: contiguous ranges need the total up front (GPU g holds rows
// [g * ceil(N / G), (g + 1) * ceil(N / G)); rounded up to 1024 rows so that bitmap words never straddle two shards).
int zvb_sharded_reserve(zvb200_store *s, uint64_t rows) {
  const uint64_t G = s->shards.size();
  if (s->shard_rows != 0) {
    if (rows > s->shard_rows * G) return set_err(ZVB200_ERR_UNSUPPORTED, "sharded store: the first reserve() fixed its size");
    return 0;
  }
  if (rows == 0) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "sharded store: reserve the total row count first");
  const uint64_t per = (((rows + G - 1) / G) + 1023) & ~(uint64_t)1023;
  std::vector<uint64_t> bases(G);
  for (uint64_t g = 0; g < G; ++g) {
    bases[g] = g * per;
    const uint64_t lo = std::min(rows, g * per), hi = std::min(rows, (g + 1) * per);
    if (hi > lo) {
      const int r = zvb200_store_reserve(s->shards[g], hi - lo);
      if (r != 0) return r;
    }
  }
  CU(cudaSetDevice(s->device));
  CU(cudaMemcpy(s->shard_base_dev, bases.data(), G * sizeof(uint64_t), cudaMemcpyHostToDevice));
  s->shard_rows = per;
  s->cap = per * G;
  return 0;
}

int zvb_sharded_append(zvb200_store *s, const void *rows, uint64_t n, const uint64_t *keys, bool on_device) {
  if (s->shard_rows == 0) return set_err(ZVB200_ERR_NOREADY, "sharded store: reserve the total row count first");
  if (s->n + n > s->cap) return set_err(ZVB200_ERR_UNSUPPORTED, "sharded store: more rows than reserved");
  const uint64_t n0 = s->n;
  uint64_t off = 0;
  while (off < n) {
    const uint64_t r = n0 + off, g = r / s->shard_rows;
    const uint64_t m = std::min(n - off, (g + 1) * s->shard_rows - r);
    const uint8_t *src = (const uint8_t *)rows + off * s->elem_bytes;
    // holes travel to the shard as keys (its drop bitmap); the document keys themselves stay with the parent
    const int rc = on_device ? zvb200_store_append_device(s->shards[g], src, m)
                             : zvb200_store_append(s->shards[g], src, m, keys ? keys + off : nullptr);
    if (rc != 0) return rc;
    off += m;
  }
  if (keys || !s->keys.empty()) {
    if (s->keys.empty()) { s->keys.resize(n0); for (uint64_t i = 0; i < n0; ++i) s->keys[i] = i; }
    s->keys.resize(n0 + n);
    for (uint64_t i = 0; i < n; ++i) s->keys[n0 + i] = keys ? keys[i] : n0 + i;
  }
  s->n = n0 + n;
  s->sealed = false;
  return 0;
}

int zvb_sharded_seal(zvb200_store *s) {
  for (auto *sh : s->shards) {
    const int r = zvb200_store_seal(sh);
    if (r != 0) return r;
  }
  s->sealed = true;
  return 0;
}

int zvb_sharded_set_deleted(zvb200_store *s, uint64_t row, int deleted) {
  if (row >= s->n) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "bad row");
  const uint64_t g = row / s->shard_rows;
  return zvb200_store_set_deleted(s->shards[g], row - g * s->shard_rows, deleted);
}

int zvb_sharded_write(zvb200_store *s, uint64_t row, const void *vec, uint64_t key) {
  if (row >= s->n) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "bad row");
  const uint64_t g = row / s->shard_rows;
  const int r = zvb200_store_write(s->shards[g], row - g * s->shard_rows, vec, key);
  if (r != 0) return r;
  if (s->keys.empty() && key != row) { s->keys.resize(s->n); for (uint64_t i = 0; i < s->n; ++i) s->keys[i] = i; }
  if (!s->keys.empty()) s->keys[row] = key;
  return 0;
}

int zvb_sharded_get_vector(const zvb200_store *s, uint64_t row, void *out, size_t cap) {
  if (row >= s->n) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "bad row");
  const uint64_t g = row / s->shard_rows;
  return zvb200_store_get_vector(s->shards[g], row - g * s->shard_rows, out, cap);
}

int zvb_sharded_ctx_create(zvb200_store *s, zvb200_ctx *c) {
  for (auto *sh : s->shards) {
    zvb200_ctx *sc = nullptr;
    const int r = zvb200_ctx_create(sh, &sc);
    if (r != 0) return r;
    c->sub.push_back(sc);
    cudaEvent_t ev = nullptr;
    CU(cudaSetDevice(sh->device));
    CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    c->sub_done.push_back(ev);
  }
  CU(cudaSetDevice(s->device));
  CU(cudaEventCreateWithFlags(&c->ev_ready, cudaEventDisableTiming));
  return 0;
}

void zvb_sharded_ctx_destroy(zvb200_ctx *c) {
  for (size_t g = 0; g < c->sub.size(); ++g) {
    if (g < c->sub_done.size() && c->sub_done[g]) { cudaSetDevice(c->sub[g]->device); cudaEventDestroy(c->sub_done[g]); }
    zvb200_ctx_destroy(c->sub[g]);
  }
  c->sub.clear(); c->sub_done.clear();
  if (c->ev_ready) { cudaSetDevice(c->device); cudaEventDestroy(c->ev_ready); c->ev_ready = nullptr; }
}

// scan on every shard -> packed keys -> the root's gather buffer (peer DMA) -> merge on the root
int zvb_sharded_search(zvb200_ctx *c, const void *queries, bool on_host, uint32_t q, uint32_t k, float threshold,
                       const uint32_t *filter, const uint32_t *rows, uint32_t m, float *out_scores_dev,
                       uint32_t *out_index_dev) {
  zvb200_store *s = c->s;
  const uint32_t G = (uint32_t)s->shards.size();
  const size_t qbytes = (size_t)q * s->elem_bytes, nkeys = (size_t)q * k;
  const uint64_t limit = c->row_limit ? std::min<uint64_t>(c->row_limit, s->n) : s->n;
  CU(cudaSetDevice(s->device));
  CU(c->gather.ensure((size_t)G * nkeys * sizeof(u64)));
  if (on_host) {
    CU(c->h_q.ensure(qbytes));
    memcpy(c->h_q.p, queries, qbytes);
  } else {
    CU(cudaEventRecord(c->ev_ready, c->stream));  // the caller's queries / filter are ready at this point of its stream
  }
  std::vector<uint32_t> cand;
  c->scan_timed = true;
  for (uint32_t g = 0; g < G; ++g) {
    zvb200_store *sh = s->shards[g];
    zvb200_ctx *sc = c->sub[g];
    sc->force_kernel = c->force_kernel;
    const uint64_t base = (uint64_t)g * s->shard_rows;
    const uint64_t rows_here = limit > base ? std::min<uint64_t>(limit - base, sh->n) : 0;
    CU(cudaSetDevice(sh->device));
    CU(sc->xkeys.ensure(nkeys * sizeof(u64)));
    if (rows_here == 0) {
      CU(cudaMemsetAsync(sc->xkeys.p, 0xff, nkeys * sizeof(u64), sc->stream));  // kMaxKey
    } else {
      const void *qdev = queries;
      if (on_host) {
        CU(sc->qin.ensure(qbytes));
        CU(cudaMemcpyAsync(sc->qin.p, c->h_q.p, qbytes, cudaMemcpyHostToDevice, sc->stream));
        qdev = sc->qin.p;
      } else {
        CU(cudaStreamWaitEvent(sc->stream, c->ev_ready, 0));  // read through NVLink from the root's memory
      }
      const uint32_t *fdev = nullptr;
      if (filter) {
        const uint32_t *fsrc = filter + base / 32;  // shard_rows is a multiple of 1024: whole words
        if (on_host) {
          const size_t fb = (size_t)((rows_here + 31) / 32) * 4;
          CU(sc->filter.ensure(fb));
          CU(cudaMemcpyAsync(sc->filter.p, fsrc, fb, cudaMemcpyHostToDevice, sc->stream));
          fdev = (const uint32_t *)sc->filter.p;
        } else {
          fdev = fsrc;
        }
      }
      int r;
      if (rows) {  // search_rows: this shard's candidates, local row numbers (others -> skipped)
        cand.assign((size_t)q * m, ZVB200_INVALID_INDEX);
        for (size_t i = 0; i < (size_t)q * m; ++i) {
          const uint64_t gr = rows[i];
          if (gr != ZVB200_INVALID_INDEX && gr >= base && gr < base + rows_here) cand[i] = (uint32_t)(gr - base);
        }
        CU(sc->cand.ensure(cand.size() * 4));
        CU(cudaMemcpyAsync(sc->cand.p, cand.data(), cand.size() * 4, cudaMemcpyHostToDevice, sc->stream));
        CU(cudaStreamSynchronize(sc->stream));  // `cand` is reused for the next shard
        r = zvb_rows_core(sc, sh, qdev, q, (const uint32_t *)sc->cand.p, m, k, threshold, fdev, nullptr, nullptr,
                          (u64 *)sc->xkeys.p);
      } else {
        r = zvb_search_core(sc, sh, qdev, q, k, threshold, fdev, nullptr, nullptr, (u64 *)sc->xkeys.p, 0, rows_here);
      }
      if (r != 0) return r;
    }
    u64 *dst = (u64 *)c->gather.p + (size_t)g * nkeys;
    if (sh->device == s->device)
      CU(cudaMemcpyAsync(dst, sc->xkeys.p, nkeys * sizeof(u64), cudaMemcpyDeviceToDevice, sc->stream));
    else
      CU(cudaMemcpyPeerAsync(dst, s->device, sc->xkeys.p, sh->device, nkeys * sizeof(u64), sc->stream));
    CU(cudaEventRecord(c->sub_done[g], sc->stream));
    c->last_kernel = sc->last_kernel;
    c->last_mma_kind = sc->last_mma_kind;
  }
  CU(cudaSetDevice(s->device));
  for (uint32_t g = 0; g < G; ++g) CU(cudaStreamWaitEvent(c->stream, c->sub_done[g], 0));
  CU(launch_merge_keys((const u64 *)c->gather.p, nkeys, G, s->shard_base_dev, q, k, out_scores_dev, nullptr,
                       out_index_dev, nullptr, 0, nullptr, c->stream));
  return 0;
}

// ======================================================================================================
// one process per GPU: communicator
// ======================================================================================================
namespace {

struct NcclApi {
  void *h = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};

// NCCL is bound at run time: a process that already carries one (PyTorch's bundled libnccl.so.2) shares it, anything
// else gets the system library; a box without NCCL can still use everything but zvb200_comm_*.
NcclApi *nccl() {
  static NcclApi api = [] {
    NcclApi a;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *n : names)
      if ((a.h = dlopen(n, RTLD_NOW | RTLD_GLOBAL)) != nullptr) break;
    if (!a.h) return a;
    a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(a.h, "ncclGetUniqueId");
    a.CommInitRank = (decltype(a.CommInitRank))dlsym(a.h, "ncclCommInitRank");
    a.CommDestroy = (decltype(a.CommDestroy))dlsym(a.h, "ncclCommDestroy");
    a.AllGather = (decltype(a.AllGather))dlsym(a.h, "ncclAllGather");
    a.GetErrorString = (decltype(a.GetErrorString))dlsym(a.h, "ncclGetErrorString");
    a.ok = a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllGather && a.GetErrorString;
    return a;
  }();
  return &api;
}

int nccl_fail(ncclResult_t r, const char *what) {
  return set_err(ZVB200_ERR_RUNTIME, what, nccl()->GetErrorString ? nccl()->GetErrorString(r) : "NCCL error");
}
#define NC(x)                                              \
  do {                                                     \
    ncclResult_t _r = (x);                                 \
    if (_r != ncclSuccess) return nccl_fail(_r, #x);       \
  } while (0)

constexpr size_t kMailKeys = 128 * 1024;  // keys per (slot, source): 1 MiB — 1024 queries x top-128

}  // namespace

struct zvb200_comm {
  int rank = 0, world = 1, device = 0;
  ncclComm_t nc = nullptr;
  cudaStream_t stream = nullptr;   // the owning context's stream
  uint64_t *bases_dev = nullptr;   // [world] first global row of every rank's shard
  bool bases_set = false;
  // peer mailboxes (CUDA IPC): mail = [2 slots][world sources][kMailKeys] keys, then [2][world] flags
  u64 *mail = nullptr;
  void *peer_mail[kMaxPeers] = {nullptr};  // peer_mail[r] = rank r's mailbox mapped here (own = mail)
  bool mailboxes = false;
  unsigned int *push_done = nullptr;
  uint32_t *status = nullptr;      // device word: 1 = a peer's keys never arrived
  u64 seq = 0;
  DevBuf gathered;
  u64 *slot_data(void *base, int slot, int src) const { return (u64 *)base + ((size_t)slot * world + src) * kMailKeys; }
  u64 *slot_flag(void *base, int slot, int src) const { return (u64 *)base + (size_t)2 * world * kMailKeys + (size_t)slot * world + src; }
  size_t mail_bytes() const { return ((size_t)2 * world * kMailKeys + (size_t)2 * world) * sizeof(u64); }
};

extern "C" {

int zvb200_comm_unique_id(void *id128) {
  if (!id128) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  if (!nccl()->ok) return set_err(ZVB200_ERR_RUNTIME, "NCCL is not available (libnccl.so.2)");
  static_assert(sizeof(ncclUniqueId) == ZVB200_COMM_ID_BYTES, "unique id size");
  ncclUniqueId id;
  NC(nccl()->GetUniqueId(&id));
  memcpy(id128, &id, sizeof id);
  return 0;
}

int zvb200_comm_destroy(zvb200_comm *cm) {
  if (!cm) return 0;
  cudaSetDevice(cm->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < cm->world; ++r)
    if (r != cm->rank && cm->peer_mail[r]) cudaIpcCloseMemHandle(cm->peer_mail[r]);
  if (cm->nc) nccl()->CommDestroy(cm->nc);
  if (cm->mail) cudaFree(cm->mail);
  if (cm->bases_dev) cudaFree(cm->bases_dev);
  if (cm->push_done) cudaFree(cm->push_done);
  if (cm->status) cudaFree(cm->status);
  cm->gathered.release();
  cudaGetLastError();
  delete cm;
  return 0;
}

int zvb200_comm_init(zvb200_ctx *c, int rank, int world, const void *id128, zvb200_comm **out) {
  if (!c || !out || !id128 || world < 1 || rank < 0 || rank >= world) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "bad argument");
  *out = nullptr;
  if (world > kMaxPeers) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "at most 16 ranks");
  if (!nccl()->ok) return set_err(ZVB200_ERR_RUNTIME, "NCCL is not available (libnccl.so.2)");
  auto cm = new zvb200_comm();
  cm->rank = rank; cm->world = world; cm->device = c->device; cm->stream = c->stream;
  auto body = [&]() -> int {
    CU(cudaSetDevice(cm->device));
    ncclUniqueId id;
    memcpy(&id, id128, sizeof id);
    NC(nccl()->CommInitRank(&cm->nc, world, id, rank));
    CU(cudaMalloc((void **)&cm->bases_dev, world * sizeof(uint64_t)));
    CU(cudaMemset(cm->bases_dev, 0, world * sizeof(uint64_t)));
    CU(cudaMalloc((void **)&cm->push_done, sizeof(unsigned int)));
    CU(cudaMemset(cm->push_done, 0, sizeof(unsigned int)));
    CU(cudaMalloc((void **)&cm->status, sizeof(uint32_t)));
    CU(cudaMemset(cm->status, 0, sizeof(uint32_t)));
    // mailboxes: every rank maps every peer's buffer.  The handles travel through NCCL itself.
    static const bool no_mail = [] { const char *e = getenv("ZVB200_COMM_MAILBOX"); return e && atoi(e) == 0; }();
    CU(cudaMalloc((void **)&cm->mail, cm->mail_bytes()));
    CU(cudaMemset(cm->mail, 0, cm->mail_bytes()));
    CU(cudaDeviceSynchronize());  // flags are zero before any peer can learn this buffer's handle
    struct Slot { cudaIpcMemHandle_t h; int ok; int pad[3]; };
    Slot mine;
    memset(&mine, 0, sizeof mine);
    mine.ok = !no_mail && cudaIpcGetMemHandle(&mine.h, cm->mail) == cudaSuccess;
    cudaGetLastError();
    Slot *dev_slots = nullptr;
    CU(cudaMalloc((void **)&dev_slots, sizeof(Slot) * (world + 1)));
    CU(cudaMemcpyAsync(dev_slots + world, &mine, sizeof mine, cudaMemcpyHostToDevice, cm->stream));
    ncclResult_t nr = nccl()->AllGather(dev_slots + world, dev_slots, sizeof(Slot), ncclUint8, cm->nc, cm->stream);
    std::vector<Slot> all(world);
    cudaError_t e = cudaMemcpyAsync(all.data(), dev_slots, sizeof(Slot) * world, cudaMemcpyDeviceToHost, cm->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(cm->stream);
    cudaFree(dev_slots);
    if (nr != ncclSuccess) return nccl_fail(nr, "ncclAllGather (mailbox handles)");
    CU(e);
    bool good = true;
    for (int r = 0; r < world; ++r) good = good && all[r].ok;
    if (good) {
      for (int r = 0; r < world && good; ++r) {
        if (r == rank) { cm->peer_mail[r] = cm->mail; continue; }
        if (cudaIpcOpenMemHandle(&cm->peer_mail[r], all[r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
          cudaGetLastError();
          cm->peer_mail[r] = nullptr;
          good = false;
        }
      }
    }
    // every rank must take the same transport: agree on the minimum
    int *agree = nullptr;
    CU(cudaMalloc((void **)&agree, sizeof(int) * (world + 1)));
    const int mine_ok = good ? 1 : 0;
    CU(cudaMemcpyAsync(agree + world, &mine_ok, sizeof(int), cudaMemcpyHostToDevice, cm->stream));
    nr = nccl()->AllGather(agree + world, agree, sizeof(int), ncclUint8, cm->nc, cm->stream);
    std::vector<int> oks(world);
    e = cudaMemcpyAsync(oks.data(), agree, sizeof(int) * world, cudaMemcpyDeviceToHost, cm->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(cm->stream);
    cudaFree(agree);
    if (nr != ncclSuccess) return nccl_fail(nr, "ncclAllGather (transport agreement)");
    CU(e);
    cm->mailboxes = true;
    for (int r = 0; r < world; ++r) cm->mailboxes = cm->mailboxes && oks[r] == 1;
    return 0;
  };
  const int r = body();
  if (r != 0) { zvb200_comm_destroy(cm); return r; }
  *out = cm;
  return 0;
}

int zvb200_comm_set_shard_bases(zvb200_comm *cm, const uint64_t *bases) {
  if (!cm || !bases) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  CU(cudaSetDevice(cm->device));
  CU(cudaMemcpy(cm->bases_dev, bases, cm->world * sizeof(uint64_t), cudaMemcpyHostToDevice));
  cm->bases_set = true;
  return 0;
}

int zvb200_comm_transport(const zvb200_comm *cm) { return cm && cm->mailboxes ? 1 : 0; }

int zvb200_comm_status(zvb200_comm *cm) {
  if (!cm) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  uint32_t st = 0;
  CU(cudaSetDevice(cm->device));
  CU(cudaMemcpy(&st, cm->status, sizeof st, cudaMemcpyDeviceToHost));
  if (st != 0) return set_err(ZVB200_ERR_RUNTIME, "a peer's top-k never arrived (exchange timed out)");
  return 0;
}

int zvb200_search_sharded_device(zvb200_ctx *c, zvb200_comm *cm, const void *queries_dev, uint32_t q, uint32_t k,
                                 float threshold, const uint32_t *filter_dev, float *out_scores_dev,
                                 uint64_t *out_rows_dev) {
  int r = zvb_check_search_args(c, queries_dev, q, k);
  if (r != 0) return r;
  if (!cm || !out_scores_dev || !out_rows_dev) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "null argument");
  if (!c->s->shards.empty()) return set_err(ZVB200_ERR_UNSUPPORTED, "a communicator shards plain stores");
  if (cm->device != c->device || cm->stream != c->stream) return set_err(ZVB200_ERR_MISMATCH, "communicator belongs to another context");
  if (!cm->bases_set) return set_err(ZVB200_ERR_NOREADY, "zvb200_comm_set_shard_bases first");
  if (q == 0) return 0;
  zvb200_store *s = c->s;
  CU(cudaSetDevice(s->device));
  const size_t nkeys = (size_t)q * k;
  CU(c->xkeys.ensure(nkeys * sizeof(u64)));
  const uint64_t limit = c->row_limit ? std::min<uint64_t>(c->row_limit, s->n) : s->n;
  if (limit == 0) {
    CU(cudaMemsetAsync(c->xkeys.p, 0xff, nkeys * sizeof(u64), c->stream));
  } else {
    r = zvb_search_core(c, s, queries_dev, q, k, threshold, filter_dev, nullptr, nullptr, (u64 *)c->xkeys.p, 0, limit);
    if (r != 0) return r;
  }
  const uint32_t W = (uint32_t)cm->world;
  if (cm->mailboxes && nkeys <= kMailKeys) {
    const u64 seq = ++cm->seq;
    const int slot = (int)(seq & 1);
    PeerTable pt;
    for (uint32_t d = 0; d < W; ++d) {
      pt.data[d] = cm->slot_data(cm->peer_mail[d], slot, cm->rank);
      pt.flag[d] = cm->slot_flag(cm->peer_mail[d], slot, cm->rank);
    }
    const int threads = 256;
    const int blocks = (int)std::min<size_t>(32, (nkeys + threads - 1) / threads);
    push_keys_kernel<<<blocks, threads, 0, c->stream>>>((const u64 *)c->xkeys.p, (uint32_t)nkeys, pt, W, seq, cm->push_done);
    count_launch();
    CU(cudaGetLastError());
    CU(launch_merge_keys(cm->slot_data(cm->mail, slot, 0), kMailKeys, W, cm->bases_dev, q, k, out_scores_dev,
                         out_rows_dev, nullptr, cm->slot_flag(cm->mail, slot, 0), seq, cm->status, c->stream));
  } else {
    CU(cm->gathered.ensure((size_t)W * nkeys * sizeof(u64)));
    NC(nccl()->AllGather(c->xkeys.p, cm->gathered.p, nkeys, ncclUint64, cm->nc, c->stream));
    CU(launch_merge_keys((const u64 *)cm->gathered.p, nkeys, W, cm->bases_dev, q, k, out_scores_dev, out_rows_dev,
                         nullptr, nullptr, 0, nullptr, c->stream));
  }
  return 0;
}

}  // extern "C"

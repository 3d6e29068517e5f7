// IVF build on the device, second half: after the centroids are trained (kmeans.cu / zvb200_kmeans),
//   zvb200_ivf_assign  labels every row with its nearest centroid — IVFBuilder::label ->
//                      IVFCentroidIndex::search_nearest_centroid (src/core/algorithm/ivf/ivf_builder.h:243-262,
//                      ivf_centroid_index.cc:297-318: a top-1 search over the centroids) — as one batched scan with
//                      k = 1 over a store of the centroids (the tcgen05 scan, the rows being the queries);
//   zvb200_ivf_pack    orders the rows the way IVFBuilder::dump_index writes them (ivf_builder.cc:672-683): posting
//                      list by posting list, ascending row number inside a list (labels_[i] in insertion order),
//                      and returns the list offsets (InvertedListMeta::id_offset / vector_count,
//                      ivf_index_format.h:41-47).
// The host adaptor (integration/gpu_ivf_builder.cc) feeds the order to the reference's own IVFDumper, so the file is
// the reference's on-disk format by construction and the stock IVFSearcher loads it.
#include "host.h"

using namespace zvb;

namespace {

constexpr uint32_t kSegments = 1024;  // the rows are cut into this many contiguous segments

// count[s][l] = rows of segment s with label l (one CTA per segment; atomics stay inside the segment's row)
__global__ void pack_count_kernel(const uint32_t *labels, uint64_t n, uint32_t nlist, uint64_t seg_rows,
                                  uint32_t *count) {
  const uint32_t s = blockIdx.x;
  const uint64_t lo = (uint64_t)s * seg_rows, hi = min(n, lo + seg_rows);
  uint32_t *c = count + (size_t)s * nlist;
  for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    const uint32_t l = labels[i];
    if (l < nlist) atomicAdd(c + l, 1u);
  }
}

// per list: exclusive running sum over the segments (count[s][l] becomes the rank of segment s's first row of list
// l inside the list), total[l] = rows of the list
__global__ void pack_scan_segments_kernel(uint32_t *count, uint32_t nlist, uint32_t nseg, uint32_t *total) {
  const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= nlist) return;
  uint32_t run = 0;
  for (uint32_t s = 0; s < nseg; ++s) {
    const uint32_t c = count[(size_t)s * nlist + l];
    count[(size_t)s * nlist + l] = run;
    run += c;
  }
  total[l] = run;
}

// list_offsets[l] = rows of the lists before l; one CTA, nlist in blockDim-sized strips with a carried prefix
__global__ void pack_scan_lists_kernel(const uint32_t *total, uint32_t nlist, unsigned long long *offsets) {
  __shared__ unsigned long long warp_sum[32];
  __shared__ unsigned long long carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t base = 0; base < nlist; base += blockDim.x) {
    const uint32_t l = base + threadIdx.x;
    const unsigned long long v = l < nlist ? total[l] : 0ull;
    unsigned long long x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) warp_sum[warp] = x;
    __syncthreads();
    if (warp == 0) {
      unsigned long long w = lane < (int)(blockDim.x >> 5) ? warp_sum[lane] : 0ull;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long y = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += y;
      }
      warp_sum[lane] = w;  // inclusive over the warps
    }
    __syncthreads();
    const unsigned long long before = carry + (warp ? warp_sum[warp - 1] : 0ull) + (x - v);
    if (l < nlist) offsets[l] = before;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) offsets[nlist] = carry;
}

// order[offsets[l] + rank] = row: one THREAD per segment walks its rows in order (the segment's ranks are private
// to it, so the walk needs no atomics and the result is the stable order: ascending row number inside a list)
__global__ void pack_scatter_kernel(const uint32_t *labels, uint64_t n, uint32_t nlist, uint64_t seg_rows,
                                    uint32_t nseg, uint32_t *rank, const unsigned long long *offsets,
                                    uint32_t *order) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nseg) return;
  const uint64_t lo = (uint64_t)s * seg_rows, hi = min(n, lo + seg_rows);
  uint32_t *r = rank + (size_t)s * nlist;
  for (uint64_t i = lo; i < hi; ++i) {
    const uint32_t l = labels[i];
    if (l >= nlist) continue;
    order[offsets[l] + r[l]++] = (uint32_t)i;
  }
}

struct Scratch {
  void *p[6] = {};
  ~Scratch() { for (void *x : p) if (x) cudaFree(x); }
};

}  // namespace

extern "C" {

int zvb200_ivf_pack(int device, const uint32_t *labels, uint64_t n, uint32_t nlist, uint32_t *order,
                    uint64_t *list_offsets) {
  if (!labels || !order || !list_offsets || nlist == 0) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "ivf_pack: null argument");
  if (n >= 0xffffffffull) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "ivf_pack: too many rows");
  if (n == 0) { memset(list_offsets, 0, (size_t)(nlist + 1) * 8); return 0; }
  CU(cudaSetDevice(device));
  const uint32_t nseg = (uint32_t)std::min<uint64_t>(kSegments, (n + 255) / 256);
  const uint64_t seg_rows = (n + nseg - 1) / nseg;
  Scratch w;
  uint32_t *d_lab, *d_cnt, *d_tot, *d_ord;
  unsigned long long *d_off;
  CU(cudaMalloc(&w.p[0], n * 4)); d_lab = (uint32_t *)w.p[0];
  CU(cudaMalloc(&w.p[1], (size_t)nseg * nlist * 4)); d_cnt = (uint32_t *)w.p[1];
  CU(cudaMalloc(&w.p[2], (size_t)nlist * 4)); d_tot = (uint32_t *)w.p[2];
  CU(cudaMalloc(&w.p[3], (size_t)(nlist + 1) * 8)); d_off = (unsigned long long *)w.p[3];
  CU(cudaMalloc(&w.p[4], n * 4)); d_ord = (uint32_t *)w.p[4];
  CU(cudaMemcpy(d_lab, labels, n * 4, cudaMemcpyHostToDevice));
  CU(cudaMemset(d_cnt, 0, (size_t)nseg * nlist * 4));
  CU(cudaMemset(d_ord, 0xff, n * 4));
  pack_count_kernel<<<nseg, 256>>>(d_lab, n, nlist, seg_rows, d_cnt);
  pack_scan_segments_kernel<<<(nlist + 127) / 128, 128>>>(d_cnt, nlist, nseg, d_tot);
  pack_scan_lists_kernel<<<1, 1024>>>(d_tot, nlist, d_off);
  pack_scatter_kernel<<<(nseg + 63) / 64, 64>>>(d_lab, n, nlist, seg_rows, nseg, d_cnt, d_off, d_ord);
  count_launch(4);
  CU(cudaGetLastError());
  CU(cudaMemcpy(order, d_ord, n * 4, cudaMemcpyDeviceToHost));
  static_assert(sizeof(unsigned long long) == sizeof(uint64_t), "offset layout");
  CU(cudaMemcpy(list_offsets, d_off, (size_t)(nlist + 1) * 8, cudaMemcpyDeviceToHost));
  if (list_offsets[nlist] != n) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "ivf_pack: a label is not below nlist");
  return 0;
}

int zvb200_ivf_assign(int device, int dtype, int metric, uint32_t dim, const void *centroids, uint32_t nlist,
                      const void *rows, uint64_t n, uint32_t *labels) {
  if (!centroids || !rows || !labels || nlist == 0) return set_err(ZVB200_ERR_INVALID_ARGUMENT, "ivf_assign: null argument");
  zvb200_store *cs = nullptr;
  zvb200_ctx *c = nullptr;
  int r = zvb200_store_create(device, dtype, metric, dim, &cs);
  if (r == 0) r = zvb200_store_append(cs, centroids, nlist, nullptr);
  if (r == 0) r = zvb200_store_seal(cs);
  if (r == 0) r = zvb200_ctx_create(cs, &c);
  if (r == 0) {
    c->mma_group_cap = (uint32_t)cs->sm_count * 128u;  // one query tile per SM, every CTA walks all the centroids
    const uint64_t chunk = (uint64_t)cs->sm_count * 128u * 4u;
    std::vector<float> scores(chunk);
    for (uint64_t off = 0; r == 0 && off < n; off += chunk) {
      const uint32_t qn = (uint32_t)std::min<uint64_t>(chunk, n - off);
      r = zvb200_search(c, (const uint8_t *)rows + off * cs->elem_bytes, qn, 1, FLT_MAX, nullptr, nullptr, scores.data(),
                        labels + off, nullptr);
      // LinearSeeker / the heap start at entry 0: a row no centroid can score (NaN everywhere) follows centroid 0
      for (uint32_t i = 0; i < qn; ++i)
        if (labels[off + i] == ZVB200_INVALID_INDEX) labels[off + i] = 0;
    }
  }
  if (c) zvb200_ctx_destroy(c);
  if (cs) zvb200_store_destroy(cs);
  return r;
}

}  // extern "C"

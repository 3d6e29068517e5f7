// K8 — the update half of a Lloyd iteration (the assignment half is the scan itself, k = 1).
//
// Replaces KmeansCluster::split_clusters_thread / update_centroid_thread
// (src/core/algorithm/cluster/kmeans_cluster.cc:812-856): per-cluster sums in DOUBLE
// (NumericalVectorMean::plus, vector_mean.h:204-214), the cluster's cost = sum of its rows' distances,
// mean = (float)(sum / count) — 0/0 = NaN for an empty cluster, exactly like vector_mean.h:217-226.
#include "internal.h"

namespace zvb {

namespace {

// one warp per row: red.add.f64 into the row's cluster (B200 resolves fp64 atomics in L2)
__global__ void kmeans_accumulate_kernel(const float *rows, uint64_t n, uint32_t dim, const uint32_t *labels,
                                         const float *dist, double *sums, unsigned long long *counts, double *scores) {
  const uint64_t r = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= n) return;
  const int lane = threadIdx.x & 31;
  const uint32_t l = labels[r];  // never INVALID: kmeans_fix_labels ran
  const float *x = rows + r * dim;
  double *s = sums + (size_t)l * dim;
  for (uint32_t i = lane; i < dim; i += 32) atomicAdd(s + i, (double)x[i]);
  if (lane == 0) {
    atomicAdd(counts + l, 1ull);
    atomicAdd(scores + l, (double)dist[r]);
  }
}

// A row none of whose distances compares below FLT_MAX (a NaN row, or every centroid NaN) keeps LinearSeeker::seek's
// initial answer: column 0 with score FLT_MAX (linear_seeker.cc:25-41) -- it joins cluster 0, its count and its cost.
__global__ void kmeans_fix_labels_kernel(uint32_t *labels, float *dist, uint64_t n) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (labels[i] == ZVB200_INVALID_INDEX) { labels[i] = 0; dist[i] = 3.402823466e+38f; }
}

__global__ void kmeans_finalize_kernel(const double *sums, const unsigned long long *counts, const double *scores,
                                       uint32_t k, uint32_t dim, float *centroids, unsigned long long *follows,
                                       double *sse) {
  const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < (uint64_t)k * dim) {
    const uint32_t c = (uint32_t)(idx / dim);
    centroids[idx] = (float)(sums[idx] / (double)counts[c]);
  }
  if (idx < k) follows[idx] = counts[idx];
  if (idx == 0) {  // CalculateSSE (kmeans_cluster.cc:355-361): in centroid order
    double a = 0.0;
    for (uint32_t c = 0; c < k; ++c) a += scores[c];
    *sse = a;
  }
}

}  // namespace

cudaError_t launch_kmeans_accumulate(const float *rows, uint64_t n, uint32_t dim, const uint32_t *labels,
                                     const float *dist, double *sums, unsigned long long *counts, double *scores,
                                     cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  const uint64_t blocks = (n + 7) / 8;
  kmeans_accumulate_kernel<<<(unsigned)blocks, 256, 0, st>>>(rows, n, dim, labels, dist, sums, counts, scores);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_kmeans_fix_labels(uint32_t *labels, float *dist, uint64_t n, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  kmeans_fix_labels_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(labels, dist, n);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_kmeans_finalize(const double *sums, const unsigned long long *counts, const double *scores,
                                   uint32_t k, uint32_t dim, float *centroids, unsigned long long *follows,
                                   double *sse, cudaStream_t st) {
  const uint64_t total = std::max<uint64_t>((uint64_t)k * dim, 1);
  kmeans_finalize_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(sums, counts, scores, k, dim, centroids,
                                                                         follows, sse);
  count_launch();
  return cudaGetLastError();
}

}  // namespace zvb

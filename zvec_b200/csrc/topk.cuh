// Warp-level bounded top-k over packed (score,row) keys, kept in shared memory.
//
// Replaces ailego::Heap<IndexDocument>::emplace + sort()
// (src/include/zvec/ailego/container/heap.h:103-115,173-175) and the threshold gate of
// IndexDocumentHeap::emplace (src/include/zvec/core/framework/index_document.h:250-261).
// Like the CPU heap, almost every candidate is rejected by one compare against the
// current k-th best (`tau`); the survivors are queued (32 per warp) and merged into the
// sorted list by rank: every element computes its final position with a binary search,
// no sorting network over the list, nothing ever goes to HBM.
#pragma once
#include "common.cuh"

namespace zvb {

__device__ __forceinline__ u64 shfl_xor_u64(u64 v, int m) {
  return __shfl_xor_sync(0xffffffffu, v, m);
}

// ascending bitonic sort of one u64 per lane
__device__ __forceinline__ u64 warp_sort32(u64 v) {
  const int lane = lane_id();
#pragma unroll
  for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      u64 o = shfl_xor_u64(v, j);
      bool up = ((lane & k) == 0);
      bool lower = ((lane & j) == 0);
      bool take_min = (lower == up);
      u64 mn = v < o ? v : o, mx = v < o ? o : v;
      v = take_min ? mn : mx;
    }
  }
  return v;
}

// One sorted list of at most k keys for one (warp, query).  All members are
// warp-uniform; every method must be called by all 32 lanes.
struct WarpTopk {
  u64 *list;   // k entries, sorted ascending, `cnt` valid
  u64 *alt;    // k entries, ping-pong target of a merge
  u64 *queue;  // >= 32 entries of pending candidates
  int k, cnt, qn;
  u64 tau;     // accept only keys < tau

  __device__ __forceinline__ void init(u64 *l, u64 *a, u64 *q, int kk, u64 tau0) {
    list = l; alt = a; queue = q; k = kk; cnt = 0; qn = 0; tau = tau0;
  }
  // smem bytes one instance needs
  __host__ __device__ static inline size_t bytes(int k, int qcap = 32) {
    return (size_t)(2 * k + qcap) * sizeof(u64);
  }

  // Merge chunk[0..m) (m <= 32, unsorted) into the list.
  __device__ __forceinline__ void merge_chunk(u64 *chunk, int m, u64 tau_cap) {
    const int lane = lane_id();
    __syncwarp();
    u64 v = lane < m ? chunk[lane] : kMaxKey;
    v = warp_sort32(v);
    __syncwarp();
    chunk[lane] = v;  // sorted, padded with kMaxKey (chunk always has 32 slots)
    __syncwarp();
    for (int i = lane; i < cnt; i += 32) {
      u64 e = list[i];
      int c = 0;
#pragma unroll
      for (int s = 16; s > 0; s >>= 1)
        if (chunk[c + s - 1] < e) c += s;
      if (chunk[c] < e) ++c;
      int np = i + c;
      if (np < k) alt[np] = e;
    }
    if (lane < m) {
      // number of list entries <= v.  Keys are unique per row, so '<=' only matters when a caller lists a row twice
      // (zvb200_search_rows scores it twice, like FlatStreamer::search_bf_by_p_keys_impl): the list's copy goes first,
      // the chunk's copy right behind it — with '<' both landed in one slot and the next slot kept a stale key
      int lo = 0, hi = cnt;
      while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (list[mid] <= v) lo = mid + 1; else hi = mid;
      }
      int np = lane + lo;
      if (np < k) alt[np] = v;
    }
    cnt = min(cnt + m, k);
    u64 *t = list; list = alt; alt = t;
    __syncwarp();
    tau = cnt == k ? list[k - 1] : tau_cap;
    if (tau > tau_cap) tau = tau_cap;
  }

  // Offer one candidate per lane (has = this lane has one, already tested < tau).
  __device__ __forceinline__ void push(bool has, u64 key, u64 tau_cap) {
    const unsigned mask = __ballot_sync(0xffffffffu, has);
    if (mask == 0) return;
    const int lane = lane_id();
    const int n = __popc(mask);
    const int pos = qn + __popc(mask & ((1u << lane) - 1u));
    if (has && pos < 32) queue[pos] = key;
    if (qn + n >= 32) {
      merge_chunk(queue, 32, tau_cap);
      __syncwarp();
      if (has && pos >= 32) queue[pos - 32] = key;
      qn = qn + n - 32;
    } else {
      qn += n;
    }
  }

  __device__ __forceinline__ void finish(u64 tau_cap) {
    if (qn > 0) merge_chunk(queue, qn, tau_cap);
    qn = 0;
    __syncwarp();
  }
};

}  // namespace zvb

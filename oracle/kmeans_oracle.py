"""TEST INFRASTRUCTURE ONLY — numpy restatement of the reference's k-means pass.

Follows KmeansCluster::clustering (src/core/algorithm/cluster/kmeans_cluster.cc:612-624):
split_clusters_thread (:812-830) seeks the nearest centroid of every row with LinearSeeker::seek
(linear_seeker.cc:20-42: strict '<', so the lower centroid wins an exact tie; a NaN distance never wins),
adds the row to that cluster's accumulator in DOUBLE (NumericalVectorMean::plus, vector_mean.h:204-214) and
its distance to the cluster's score; cluster() ends with PurgeCentroids (:363-383), which moves the empty clusters
to the tail even when it cuts nothing; update_centroid_thread (:832-856) sets centroid = (float)(sum / count)
(vector_mean.h:217-226: 0/0 = NaN for an empty cluster) and cost = sum of the scores (CalculateSSE, :355-361).
Only tests/ may import this; pinned against the compiled reference (oracle/_ref: zref_kmeans) in
tests/test_kmeans.py::test_oracle_kmeans_equals_the_reference.
"""
import numpy as np


def distances(rows, centroids, metric="l2sq"):
    """[n, k] float32 distances the way the fp32 1x1 kernels see them (float64 here: callers exclude near-ties)."""
    x = rows.astype(np.float64)
    c = centroids.astype(np.float64)
    if metric == "ip":
        return -(x @ c.T)
    d = (x * x).sum(1)[:, None] - 2.0 * (x @ c.T) + (c * c).sum(1)[None, :]
    return np.maximum(d, 0.0)


def assign(rows, centroids, metric="l2sq"):
    d = distances(rows, centroids, metric)
    d = np.where(np.isnan(d), np.inf, d)
    lab = d.argmin(1).astype(np.uint32)          # first minimum = strict '<'
    return lab, d[np.arange(len(rows)), lab], d


def lloyd_pass(rows, centroids, metric="l2sq"):
    """one clustering() pass -> (new centroids f32 [k, dim], follows u64[k], cost, labels used, distance matrix)"""
    k = len(centroids)
    lab, best, d = assign(rows, centroids, metric)
    sums = np.zeros((k, rows.shape[1]), np.float64)
    np.add.at(sums, lab, rows.astype(np.float64))
    cnt = np.bincount(lab, minlength=k).astype(np.uint64)
    with np.errstate(invalid="ignore", divide="ignore"):
        new = (sums / cnt[:, None].astype(np.float64)).astype(np.float32)
    return new, cnt, float(best.sum()), lab, d


def purge_order(follows):
    """PurgeCentroids(cents, cutting = false) (kmeans_cluster.cc:363-383), run at the end of cluster() even when
    nothing is cut: every centroid nobody follows is swapped with the last one not yet moved, so the empty
    clusters end up at the tail.  -> the permutation: new position -> old position."""
    order = list(range(len(follows)))
    index, tamp = 0, len(order)
    while index < tamp:
        if follows[order[index]] == 0:
            last = tamp - 1
            if index != last:
                order[index], order[last] = order[last], order[index]
            tamp = last
            continue
        index += 1
    return np.array(order, np.int64)


def kmeans(rows, centroids, iterations, metric="l2sq"):
    c = np.ascontiguousarray(centroids, np.float32).copy()
    cnt = cost = None
    for _ in range(iterations):
        c, cnt, cost, _, _ = lloyd_pass(rows, c, metric)
    if cnt is None:                      # zero passes: nobody follows anything, the purge still reorders the list
        cnt = np.zeros(len(c), np.uint64)
    order = purge_order(cnt)
    c, cnt = c[order], cnt[order]
    lab, _, d = assign(rows, c, metric)
    return c, lab, cnt, cost, d

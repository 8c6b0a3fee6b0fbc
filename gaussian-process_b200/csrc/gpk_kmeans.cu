// Lloyd iteration of the KMeans subsample of find_len_scales.len_scale_opt (find_len_scales.py:25-47), SURVEY 8f row N4.
//
// The reference calls sklearn.cluster.KMeans(n_clusters=100, n_init='auto', random_state=0).fit_predict on the data points
// before the swarm optimisation whenever there are more than 100 of them; which 100 points survive fixes everything that
// follows, so the result has to be sklearn's.  Its two O(N k) steps per iteration run here, with sklearn's single-threaded
// arithmetic (the reference pins OMP / BLAS to one thread):
//   * E-step (lloyd_iter_chunked_dense -> _update_chunk_dense): label_i = first j minimising  |c_j|^2 - 2 <x_i, c_j>  (the
//     point's own norm is left out, as sklearn does); the inner product is the BLAS chain fma(x_k, c_k, .) over k from a
//     rounded first product, the distance one fused multiply-add of -2;
//   * M-step sums: centre_j = sum of its members in SAMPLE ORDER (one accumulator per cluster and dimension, as the
//     single-threaded chunk loop adds them), count_j.
// Everything on the centres' side (squared norms, empty clusters, averaging, centre shift, tolerance test, k-means++
// seeding) stays in NumPy / sklearn's own function on the host, with sklearn's expressions: gaussian-process_b200/kmeans.py.
#include "gpk_common.cuh"

namespace gpk {

constexpr int KM_MAX_K = 1024;          // clusters (centres live in shared memory: k (d + 1) doubles)

// labels[i] = argmin_j (csq_j - 2 <x_i, c_j>), first minimum; *changed |= (label differs from the previous one)
__global__ void __launch_bounds__(256)
k_kmeans_assign(const double* __restrict__ X, int N, int d, const double* __restrict__ centers, const double* __restrict__ csq,
                int k, int* __restrict__ labels, int* __restrict__ changed) {
    extern __shared__ double s_c[];                       // [k][d] centres, then [k] squared norms
    double* s_q = s_c + (size_t)k * d;
    for (int t = threadIdx.x; t < k * d; t += blockDim.x) s_c[t] = centers[t];
    for (int t = threadIdx.x; t < k; t += blockDim.x) s_q[t] = csq[t];
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    double x[MAX_D];
    for (int m = 0; m < d; ++m) x[m] = X[(size_t)i * d + m];
    int best = 0;
    double best_d = 0.0;
    for (int j = 0; j < k; ++j) {
        const double* c = s_c + (size_t)j * d;
        double dot = __dmul_rn(x[0], c[0]);
        for (int m = 1; m < d; ++m) dot = fma(x[m], c[m], dot);
        const double dist = fma(-2.0, dot, s_q[j]);
        if (j == 0 || dist < best_d) { best = j; best_d = dist; }
    }
    if (labels[i] != best) {
        labels[i] = best;
        *changed = 1;                                     // every writer stores the same value
    }
}

// sums[j][m] = sum over the members of cluster j of X[i][m], members taken in ascending i; counts[j] = number of members.
// One warp per cluster: 32 labels are tested at a time, the members among them are added one after the other.
__global__ void __launch_bounds__(32)
k_kmeans_sums(const double* __restrict__ X, int N, int d, const int* __restrict__ labels, double* __restrict__ sums,
              double* __restrict__ counts) {
    const int j = blockIdx.x, lane = threadIdx.x;
    double acc = 0.0;
    int cnt = 0;
    for (int i0 = 0; i0 < N; i0 += 32) {
        const int i = i0 + lane;
        unsigned m = __ballot_sync(0xffffffffu, i < N && labels[i] == j);
        cnt += __popc(m);
        while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            if (lane < d) acc = __dadd_rn(acc, X[(size_t)(i0 + b) * d + lane]);
        }
    }
    if (lane < d) sums[(size_t)j * d + lane] = acc;
    if (lane == 0) counts[j] = (double)cnt;
}

int launch_kmeans_step(const double* X, int N, int d, const double* centers, const double* csq, int k, int* labels,
                       int* changed, double* sums, double* counts, cudaStream_t st) {
    if (k < 1 || k > KM_MAX_K || d < 1 || d > MAX_D || N < 1) { set_error("kmeans: unsupported sizes N=%d d=%d k=%d (k <= %d, d <= %d)", N, d, k, KM_MAX_K, MAX_D); return 2; }
    const size_t smem = (size_t)k * (d + 1) * sizeof(double);
    static size_t configured = 48 * 1024;
    if (smem > configured) {
        cudaError_t er = cudaFuncSetAttribute(k_kmeans_assign, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (er != cudaSuccess) { set_error("k_kmeans_assign smem attr: %s", cudaGetErrorString(er)); return 1; }
        configured = smem;
    }
    k_kmeans_assign<<<(N + 255) / 256, 256, smem, st>>>(X, N, d, centers, csq, k, labels, changed);
    if (check_launch("k_kmeans_assign")) return 1;
    if (sums != nullptr) {
        k_kmeans_sums<<<k, 32, 0, st>>>(X, N, d, labels, sums, counts);
        if (check_launch("k_kmeans_sums")) return 1;
    }
    return 0;
}

}  // namespace gpk

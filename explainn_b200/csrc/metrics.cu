// Validation metrics on the device: the reference hands the full prediction matrix to
// sklearn / scipy on the host (architectures/__init__.py:458-461 get_metrics; selene/__init__.py:359-366 one call
// on the flattened arrays; pwm-scoring.py:115-120 one call per PWM column, 750 of them).  Here one block per column
// works on device-resident, already sorted columns; the sort itself is torch.sort (library plumbing, like a cuBLAS
// call) on the host side of the C ABI.
//   * AUROC  = Mann-Whitney U with average ranks over ties (= sklearn.metrics.roc_auc_score's trapezoid over
//     distinct thresholds); every partial sum is a multiple of 1/2 below 2^53, so the result does not depend on
//     the summation order.
//   * AUPR   = sklearn.metrics.average_precision_score: sum over distinct thresholds of (R_k - R_{k-1}) P_k
//            = (1/n_pos) * sum over positives i of tp[end of i's tie run] / (end of i's tie run + 1).
//   * Pearson r in fp64, two passes (means, then centred sums) as scipy.stats.pearsonr; Spearman = Pearson of the
//     average ranks (scipy.stats.spearmanr / rankdata "average").
// Sums are reduced in a fixed order (thread-strided partials -> warp tree -> warp 0), so results are run-to-run
// bit-identical.
#include "common.cuh"

namespace xnn {

#define MET_THREADS 1024

__device__ __forceinline__ double block_sum(double v, double* s_red) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) s_red[warp] = v;
    __syncthreads();
    double t = (threadIdx.x < (blockDim.x >> 5)) ? s_red[threadIdx.x] : 0.0;
    if (warp == 0) t = warp_sum(t);
    if (threadIdx.x == 0) s_red[32] = t;
    __syncthreads();
    return s_red[32];
}

// first index in [0,N) of the descending array v with v[idx] <= x  ... (run start of x when x is present)
__device__ __forceinline__ long long desc_lower(const float* __restrict__ v, long long N, float x) {
    long long lo = 0, hi = N;
    while (lo < hi) {
        const long long mid = (lo + hi) >> 1;
        if (v[mid] > x) lo = mid + 1; else hi = mid;
    }
    return lo;
}
// first index with v[idx] < x (one past the run of x)
__device__ __forceinline__ long long desc_upper(const float* __restrict__ v, long long N, float x) {
    long long lo = 0, hi = N;
    while (lo < hi) {
        const long long mid = (lo + hi) >> 1;
        if (v[mid] >= x) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// pass A: tp[i] = number of positives among the first i+1 entries of the column in sorted order
__global__ void __launch_bounds__(MET_THREADS)
auc_cumsum_kernel(const long long* __restrict__ perm, const uint8_t* __restrict__ labels, long long label_stride,
                  long long N, int* __restrict__ tp) {
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    const int g = blockIdx.x;
    const long long* p = perm + (long long)g * N;
    const uint8_t* y = labels + (long long)g * label_stride;
    int* out = tp + (long long)g * N;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (long long base = 0; base < N; base += MET_THREADS) {
        const long long i = base + threadIdx.x;
        int v = (i < N) ? (y[p[i]] != 0) : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += t;
        }
        if (lane == 31) s_warp[warp] = v;
        __syncthreads();
        if (warp == 0) {
            int w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += t;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        const int incl = v + (warp > 0 ? s_warp[warp - 1] : 0) + s_carry;
        if (i < N) out[i] = incl;
        __syncthreads();
        if (threadIdx.x == MET_THREADS - 1) s_carry = incl;
        __syncthreads();
    }
}

// pass B: rank sum of the positives (AUROC) and the precision sum (AUPR)
__global__ void __launch_bounds__(MET_THREADS)
auc_reduce_kernel(const float* __restrict__ sorted_scores, const long long* __restrict__ perm,
                  const uint8_t* __restrict__ labels, long long label_stride, long long N,
                  const int* __restrict__ tp, double* __restrict__ auroc, double* __restrict__ aupr) {
    __shared__ double s_red[33];
    const int g = blockIdx.x;
    const float* v = sorted_scores + (long long)g * N;
    const long long* p = perm + (long long)g * N;
    const uint8_t* y = labels + (long long)g * label_stride;
    const int* t = tp + (long long)g * N;
    double rank2 = 0.0;        // twice the ascending average rank, summed over positives (exact in fp64)
    double prec = 0.0;
    for (long long i = threadIdx.x; i < N; i += MET_THREADS) {
        if (y[p[i]] == 0) continue;
        const float x = v[i];
        // most scores are distinct: look at the neighbours before searching
        long long lb = i, ub = i + 1;
        if (i > 0 && v[i - 1] == x) lb = desc_lower(v, N, x);
        if (i + 1 < N && v[i + 1] == x) ub = desc_upper(v, N, x);
        rank2 += (double)((N - ub + 1) + (N - lb));
        prec += (double)t[ub - 1] / (double)ub;
    }
    const double r2 = block_sum(rank2, s_red);
    const double pr = block_sum(prec, s_red);
    if (threadIdx.x == 0) {
        const double npos = (double)t[N - 1], nneg = (double)N - npos;
        const double u = 0.5 * r2 - 0.5 * npos * (npos + 1.0);
        auroc[g] = (npos > 0.0 && nneg > 0.0) ? u / (npos * nneg) : __longlong_as_double(0x7ff8000000000000LL);
        aupr[g] = npos > 0.0 ? pr / npos : __longlong_as_double(0x7ff8000000000000LL);
    }
}

// average 1-based ranks (ties share the mean of their positions) of an ASCENDING sorted column, scattered back to
// the original order through perm
__global__ void __launch_bounds__(256)
rank_avg_kernel(const float* __restrict__ sorted_vals, const long long* __restrict__ perm, long long N,
                double* __restrict__ ranks) {
    const int g = blockIdx.y;
    const float* v = sorted_vals + (long long)g * N;
    const long long* p = perm + (long long)g * N;
    double* out = ranks + (long long)g * N;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long long)gridDim.x * blockDim.x) {
        const float x = v[i];
        long long lb = i, ub = i + 1;
        if (i > 0 && v[i - 1] == x) {            // ascending: first index with v >= x
            long long lo = 0, hi = i;
            while (lo < hi) { const long long mid = (lo + hi) >> 1; if (v[mid] < x) lo = mid + 1; else hi = mid; }
            lb = lo;
        }
        if (i + 1 < N && v[i + 1] == x) {        // first index with v > x
            long long lo = i + 1, hi = N;
            while (lo < hi) { const long long mid = (lo + hi) >> 1; if (v[mid] <= x) lo = mid + 1; else hi = mid; }
            ub = lo;
        }
        out[p[i]] = 0.5 * (double)(lb + ub + 1);
    }
}

template <typename T>
__global__ void __launch_bounds__(MET_THREADS)
pearson_kernel(const T* __restrict__ x, const T* __restrict__ y, long long y_stride, long long N,
               double* __restrict__ r) {
    __shared__ double s_red[33];
    const int g = blockIdx.x;
    const T* xc = x + (long long)g * N;
    const T* yc = y + (long long)g * y_stride;
    double sx = 0.0, sy = 0.0;
    for (long long i = threadIdx.x; i < N; i += MET_THREADS) { sx += (double)xc[i]; sy += (double)yc[i]; }
    const double mx = block_sum(sx, s_red) / (double)N;
    const double my = block_sum(sy, s_red) / (double)N;
    double sxx = 0.0, syy = 0.0, sxy = 0.0;
    for (long long i = threadIdx.x; i < N; i += MET_THREADS) {
        const double dx = (double)xc[i] - mx, dy = (double)yc[i] - my;
        sxx += dx * dx; syy += dy * dy; sxy += dx * dy;
    }
    const double a = block_sum(sxx, s_red);
    const double b = block_sum(syy, s_red);
    const double c = block_sum(sxy, s_red);
    if (threadIdx.x == 0) {
        // a constant column has no correlation: scipy returns nan (ConstantInputWarning)
        double v = (a > 0.0 && b > 0.0) ? c / (sqrt(a) * sqrt(b)) : __longlong_as_double(0x7ff8000000000000LL);
        if (v > 1.0) v = 1.0;
        if (v < -1.0) v = -1.0;
        r[g] = v;
    }
}

}  // namespace xnn
using namespace xnn;

extern "C" size_t xnn_metric_auc_work_bytes(long long N, int G) {
    if (N < 1 || G < 1) return 0;
    return (size_t)N * (size_t)G * sizeof(int);
}

extern "C" int xnn_metric_auc(const float* sorted_scores, const long long* perm, const uint8_t* labels,
                              long long label_stride, long long N, int G, double* auroc, double* aupr, void* work,
                              xnn_stream_t s) {
    XNN_CHECK_ARG(sorted_scores && perm && labels && auroc && aupr && work, "NULL pointer");
    XNN_CHECK_ARG(N >= 1 && N < (1LL << 31) && G >= 1, "bad sizes");
    XNN_CHECK_ARG(label_stride == 0 || label_stride >= N, "label_stride must be 0 (shared labels) or >= N");
    cudaStream_t st = to_stream(s);
    auc_cumsum_kernel<<<G, MET_THREADS, 0, st>>>(perm, labels, label_stride, N, (int*)work);
    XNN_CHECK_LAUNCH();
    auc_reduce_kernel<<<G, MET_THREADS, 0, st>>>(sorted_scores, perm, labels, label_stride, N, (const int*)work, auroc,
                                                 aupr);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

extern "C" int xnn_metric_ranks(const float* sorted_vals, const long long* perm, long long N, int G, double* ranks,
                                xnn_stream_t s) {
    XNN_CHECK_ARG(sorted_vals && perm && ranks, "NULL pointer");
    XNN_CHECK_ARG(N >= 1 && G >= 1 && G <= 65535, "bad sizes");
    long long bx = (N + 255) / 256;
    const long long cap = (long long)num_sms() * 8;
    if (bx > cap) bx = cap;
    rank_avg_kernel<<<dim3((unsigned)bx, (unsigned)G), 256, 0, to_stream(s)>>>(sorted_vals, perm, N, ranks);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

extern "C" int xnn_metric_pearson_f32(const float* x, const float* y, long long y_stride, long long N, int G,
                                      double* r, xnn_stream_t s) {
    XNN_CHECK_ARG(x && y && r, "NULL pointer");
    XNN_CHECK_ARG(N >= 2 && G >= 1, "bad sizes");
    XNN_CHECK_ARG(y_stride == 0 || y_stride >= N, "y_stride must be 0 (shared column) or >= N");
    pearson_kernel<float><<<G, MET_THREADS, 0, to_stream(s)>>>(x, y, y_stride, N, r);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

extern "C" int xnn_metric_pearson_f64(const double* x, const double* y, long long y_stride, long long N, int G,
                                      double* r, xnn_stream_t s) {
    XNN_CHECK_ARG(x && y && r, "NULL pointer");
    XNN_CHECK_ARG(N >= 2 && G >= 1, "bad sizes");
    XNN_CHECK_ARG(y_stride == 0 || y_stride >= N, "y_stride must be 0 (shared column) or >= N");
    pearson_kernel<double><<<G, MET_THREADS, 0, to_stream(s)>>>(x, y, y_stride, N, r);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

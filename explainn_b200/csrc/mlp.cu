// a4 / a5 / a6 -- each unit's Linear->BN->ReLU->Dropout->Linear->BN->ReLU MLP, the
// final Linear, and their backward (explainn/architectures/__init__.py:84-104,114).
// fp32 FMA kernels; every cross-batch reduction is two-stage with fixed order
// (deterministic), statistics are accumulated in fp64.
#include "common.cuh"
using namespace xnn;

// ---------------------------------------------------------------------------
// strided batched SGEMM: C[bt] = epi(A[bt] (MxK) . B[bt] (KxN))
// 64x64 tile, 16-deep K step, 256 threads, 4x4 outputs per thread.
// ---------------------------------------------------------------------------
struct GemmArgs {
    const float* A; long long a_bs, a_rs, a_cs;
    const float* Bm; long long b_bs, b_rs, b_cs;
    float* C; long long c_bs, c_rs, c_cs;
    int M, N, K, batch;
    const float* bias; long long bias_bs;   // per-column bias (index j), may be NULL
    const float* mul;                       // same layout as C, may be NULL
    int mul_mode;                           // 1: C *= mul; 2: C = mul > 0 ? C : 0
    int splits;                             // split-K factor (>1: partials to `part`)
    float* part;                            // [splits][batch][M][N]
};

#define GT 64
#define GK 16
__global__ void __launch_bounds__(256) sgemm_kernel(GemmArgs g) {
    __shared__ float As[GK][GT + 4];
    __shared__ float Bs[GK][GT + 4];
    const int bt = blockIdx.z / g.splits, sp = blockIdx.z % g.splits;
    const int m0 = blockIdx.y * GT, n0 = blockIdx.x * GT;
    const float* A = g.A + bt * g.a_bs;
    const float* Bm = g.Bm + bt * g.b_bs;
    const int kchunk = ((g.K + g.splits - 1) / g.splits + GK - 1) / GK * GK;
    const int kbeg = sp * kchunk, kend = min(g.K, kbeg + kchunk);
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
    // element -> thread mapping for the tile loads follows the unit stride
    const bool a_kfast = (g.a_cs == 1), b_nfast = (g.b_cs == 1);
    for (int k0 = kbeg; k0 < kend; k0 += GK) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int idx = threadIdx.x + e * 256;
            int mm, kk;
            if (a_kfast) { kk = idx & (GK - 1); mm = idx >> 4; } else { mm = idx & (GT - 1); kk = idx >> 6; }
            const int gm = m0 + mm, gk = k0 + kk;
            As[kk][mm] = (gm < g.M && gk < kend) ? A[gm * g.a_rs + gk * g.a_cs] : 0.0f;
            int nn, kb;
            if (b_nfast) { nn = idx & (GT - 1); kb = idx >> 6; } else { kb = idx & (GK - 1); nn = idx >> 4; }
            const int gn = n0 + nn, gkb = k0 + kb;
            Bs[kb][nn] = (gn < g.N && gkb < kend) ? Bm[gkb * g.b_rs + gn * g.b_cs] : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < GK; ++kk) {
            float a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gm = m0 + ty * 4 + i;
        if (gm >= g.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gn = n0 + tx * 4 + j;
            if (gn >= g.N) continue;
            float v = acc[i][j];
            if (g.splits > 1) {
                g.part[(((long long)sp * g.batch + bt) * g.M + gm) * g.N + gn] = v;
            } else {
                if (g.bias) v += g.bias[bt * g.bias_bs + gn];
                const long long ci = bt * g.c_bs + gm * g.c_rs + gn * g.c_cs;
                if (g.mul_mode == 1) v *= g.mul[ci];
                else if (g.mul_mode == 2) v = (g.mul[ci] > 0.0f) ? v : 0.0f;
                g.C[ci] = v;
            }
        }
    }
}

__global__ void __launch_bounds__(256) sgemm_splitk_reduce_kernel(GemmArgs g) {
    const long long total = (long long)g.batch * g.M * g.N;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        float v = 0.0f;
        for (int sp = 0; sp < g.splits; ++sp) v += g.part[(long long)sp * total + i];
        const int gn = (int)(i % g.N), gm = (int)((i / g.N) % g.M), bt = (int)(i / ((long long)g.M * g.N));
        if (g.bias) v += g.bias[bt * g.bias_bs + gn];
        const long long ci = bt * g.c_bs + gm * g.c_rs + gn * g.c_cs;
        if (g.mul_mode == 1) v *= g.mul[ci];
        else if (g.mul_mode == 2) v = (g.mul[ci] > 0.0f) ? v : 0.0f;
        g.C[ci] = v;
    }
}

static int launch_gemm(const GemmArgs& g, cudaStream_t st) {
    dim3 grid(ceil_div(g.N, GT), ceil_div(g.M, GT), g.batch * g.splits);
    sgemm_kernel<<<grid, 256, 0, st>>>(g);
    if (cudaGetLastError() != cudaSuccess) return 1;
    if (g.splits > 1) {
        const long long total = (long long)g.batch * g.M * g.N;
        long long blocks = (total + 255) / 256;
        if (blocks > (long long)num_sms() * 8) blocks = (long long)num_sms() * 8;
        sgemm_splitk_reduce_kernel<<<(int)blocks, 256, 0, st>>>(g);
        if (cudaGetLastError() != cudaSuccess) return 1;
    }
    return 0;
}

// final Linear weight gradient (M = F, N = U are one handful of tiles): split the batch finely enough to fill the SMs
static int splitk_wide(int K) {
    int s = (K + 127) / 128;
    if (s > 32) s = 32;
    if (s < 1) s = 1;
    return s;
}
static int splitk_for(int K) {
    int s = (K + 511) / 512;
    if (s > 32) s = 32;
    if (s < 1) s = 1;
    return s;
}

// ---------------------------------------------------------------------------
// column reductions over the batch: S1[c] = sum_b x[b,c], S2[c] = sum_b x[b,c]*y[b,c]
// (y == NULL: S2 = sum x^2).  grid (C/32, tiles); lanes = channels (coalesced).
// ---------------------------------------------------------------------------
#define CS_ROWS_PER_TILE 256
static int colstats_tiles(int B) {
    int t = (B + CS_ROWS_PER_TILE - 1) / CS_ROWS_PER_TILE;
    if (t > 64) t = 64;
    if (t < 1) t = 1;
    return t;
}

__global__ void __launch_bounds__(256)
colsums_kernel(const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ relu_out,
               int B, int C, int tiles, double* __restrict__ part /* [tiles][2][C] */) {
    __shared__ double s1[8][32], s2[8][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + lane;
    const int rows_per = (B + tiles - 1) / tiles;
    const int r0 = blockIdx.y * rows_per, r1 = min(B, r0 + rows_per);
    double a1 = 0.0, a2 = 0.0;
    if (c < C) {
#pragma unroll 4                                       // four rows' loads in flight (the sums keep their order)
        for (int r = r0 + warp; r < r1; r += 8) {
            float xv = x[(long long)r * C + c];
            if (relu_out && !(relu_out[(long long)r * C + c] > 0.0f)) xv = 0.0f;
            const float yv = y ? y[(long long)r * C + c] : xv;
            a1 += (double)xv;
            a2 += (double)xv * (double)yv;
        }
    }
    s1[warp][lane] = a1; s2[warp][lane] = a2;
    __syncthreads();
    if (warp == 0 && c < C) {
        double t1 = 0.0, t2 = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) { t1 += s1[w][lane]; t2 += s2[w][lane]; }
        part[((long long)blockIdx.y * 2 + 0) * C + c] = t1;
        part[((long long)blockIdx.y * 2 + 1) * C + c] = t2;
    }
}

extern "C" size_t xnn_bn_work_bytes(int B, int C) {
    // [tiles][2][C] partial sums + [C][2] reduced coefficients
    return ((size_t)colstats_tiles(B) * 2 * C + 2 * (size_t)C) * sizeof(double);
}

__global__ void __launch_bounds__(256)
bn_train_finalize_kernel(const double* __restrict__ part, int tiles, int B, int C, const float* __restrict__ gamma,
                         const float* __restrict__ beta, float eps, float momentum,
                         float* __restrict__ running_mean, float* __restrict__ running_var,
                         float* __restrict__ scale, float* __restrict__ shift, double* __restrict__ stats) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double s1 = 0.0, s2 = 0.0;
    for (int t = 0; t < tiles; ++t) { s1 += part[((long long)t * 2) * C + c]; s2 += part[((long long)t * 2 + 1) * C + c]; }
    const double mean = s1 / B;
    double var = s2 / B - mean * mean;
    if (var < 0.0) var = 0.0;
    const double invstd = 1.0 / sqrt(var + (double)eps);
    const double sc = (double)gamma[c] * invstd;
    scale[c] = (float)sc;
    shift[c] = (float)((double)beta[c] - mean * sc);
    stats[4 * (long long)c + 0] = mean; stats[4 * (long long)c + 1] = var;
    stats[4 * (long long)c + 2] = invstd; stats[4 * (long long)c + 3] = 0.0;
    if (running_mean && running_var) {
        const double unb = (B > 1) ? var * (double)B / (double)(B - 1) : var;
        running_mean[c] = (float)((1.0 - momentum) * (double)running_mean[c] + momentum * mean);
        running_var[c] = (float)((1.0 - momentum) * (double)running_var[c] + momentum * unb);
    }
}

namespace xnn {
// part: [tiles][2][C] partial (sum, sum of squares) in fp64 -> scale/shift/stats (+ running stats)
int launch_bn_train_finalize(const double* part, int tiles, int B, int C, const float* gamma, const float* beta,
                             float eps, float momentum, float* running_mean, float* running_var, float* scale,
                             float* shift, double* stats, cudaStream_t st) {
    bn_train_finalize_kernel<<<ceil_div(C, 256), 256, 0, st>>>(part, tiles, B, C, gamma, beta, eps, momentum,
                                                              running_mean, running_var, scale, shift, stats);
    return cudaGetLastError() != cudaSuccess;
}
}  // namespace xnn

extern "C" size_t xnn_bn_partial_bytes(int B, int C) {
    return (size_t)colstats_tiles(B) * 2 * C * sizeof(double);
}

extern "C" int xnn_bn_train_prepare(const float* x, int B, int C, const float* gamma, const float* beta,
                                    float eps, float momentum, float* running_mean, float* running_var,
                                    float* scale, float* shift, double* stats, void* work, int phase,
                                    long long count, xnn_stream_t s) {
    XNN_CHECK_ARG(x && gamma && beta && scale && shift && stats && work, "NULL pointer");
    XNN_CHECK_ARG(phase >= 0 && phase <= 2, "phase: 0 = whole call, 1 = partial sums, 2 = finalize");
    if (phase == 0) count = B;
    XNN_CHECK_ARG(B >= 1 && C >= 1 && count >= 2 && count < (1LL << 31),
                  "train-mode BatchNorm needs more than 1 value per channel");
    const int tiles = colstats_tiles(B);
    if (phase != 2) {
        dim3 grid(ceil_div(C, 32), tiles);
        colsums_kernel<<<grid, 256, 0, to_stream(s)>>>(x, nullptr, nullptr, B, C, tiles, (double*)work);
        XNN_CHECK_LAUNCH();
    }
    if (phase != 1) {
        bn_train_finalize_kernel<<<ceil_div(C, 256), 256, 0, to_stream(s)>>>(
            (const double*)work, tiles, (int)count, C, gamma, beta, eps, momentum, running_mean, running_var, scale,
            shift, stats);
        XNN_CHECK_LAUNCH();
    }
    return XNN_OK;
}

// dgamma/dbeta from the partial sums, then dx (elementwise)
__global__ void __launch_bounds__(256)
bn_bwd_finalize_kernel(const double* __restrict__ part, int tiles, int C, const double* __restrict__ stats,
                       float* __restrict__ dgamma, float* __restrict__ dbeta, double* __restrict__ coef /* [C][2] */,
                       float out_scale) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double s1 = 0.0, s2 = 0.0;
    for (int t = 0; t < tiles; ++t) { s1 += part[((long long)t * 2) * C + c]; s2 += part[((long long)t * 2 + 1) * C + c]; }
    const double mean = stats[4 * (long long)c], invstd = stats[4 * (long long)c + 2];
    const double dg = invstd * (s2 - mean * s1);   // sum g * xhat
    dgamma[c] = (float)(dg * (double)out_scale);
    dbeta[c] = (float)(s1 * (double)out_scale);
    coef[2 * (long long)c] = s1; coef[2 * (long long)c + 1] = dg;
}

__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const float* g, const float* __restrict__ relu_out, const float* __restrict__ xpre,
                    const double* __restrict__ stats,
                    const float* __restrict__ gamma, const double* __restrict__ coef, int B, int C, long long count,
                    float* dx /* may alias g */) {
    const long long total = (long long)B * C;
    const float invB = 1.0f / (float)count;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const float mean = (float)stats[4 * (long long)c], invstd = (float)stats[4 * (long long)c + 2];
        const float db = (float)coef[2 * (long long)c], dg = (float)coef[2 * (long long)c + 1];
        const float xhat = (xpre[i] - mean) * invstd;
        float gi = g[i];
        if (relu_out && !(relu_out[i] > 0.0f)) gi = 0.0f;
        dx[i] = gamma[c] * invstd * (gi - db * invB - xhat * dg * invB);
    }
}

extern "C" int xnn_bn_bwd(const float* g, const float* relu_out, const float* xpre, const double* stats,
                          const float* gamma, int B, int C, float* dgamma, float* dbeta, float* dx, void* work,
                          int phase, long long count, float out_scale, xnn_stream_t s) {
    XNN_CHECK_ARG(g && xpre && stats && gamma && dgamma && dbeta && dx && work, "NULL pointer");
    XNN_CHECK_ARG(phase >= 0 && phase <= 2, "phase: 0 = whole call, 1 = partial sums, 2 = finalize + apply");
    if (phase == 0) { count = B; out_scale = 1.0f; }
    XNN_CHECK_ARG(B >= 1 && C >= 1 && count >= 2, "bad sizes");
    const int tiles = colstats_tiles(B);
    double* part = (double*)work;
    if (phase != 2) {
        dim3 grid(ceil_div(C, 32), tiles);
        colsums_kernel<<<grid, 256, 0, to_stream(s)>>>(g, xpre, relu_out, B, C, tiles, part);
        XNN_CHECK_LAUNCH();
    }
    if (phase != 1) {
        double* coef = part + (size_t)tiles * 2 * C;
        bn_bwd_finalize_kernel<<<ceil_div(C, 256), 256, 0, to_stream(s)>>>(part, tiles, C, stats, dgamma, dbeta, coef,
                                                                          out_scale);
        XNN_CHECK_LAUNCH();
        long long blocks = ((long long)B * C + 255) / 256;
        if (blocks > (long long)num_sms() * 16) blocks = (long long)num_sms() * 16;
        bn_bwd_apply_kernel<<<(int)blocks, 256, 0, to_stream(s)>>>(g, relu_out, xpre, stats, gamma, coef, B, C, count,
                                                                   dx);
        XNN_CHECK_LAUNCH();
    }
    return XNN_OK;
}

// ---------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------
extern "C" int xnn_fc1_fwd(const float* pool, const float* w1, const float* b1, int B, int U, int n, float* hpre,
                           xnn_stream_t s) {
    XNN_CHECK_ARG(pool && w1 && b1 && hpre, "NULL pointer");
    XNN_CHECK_ARG(B >= 0 && U >= 1 && n >= 1, "bad sizes");
    if (B == 0) return XNN_OK;
    GemmArgs g{};
    const int H = XNN_HIDDEN;
    g.A = pool; g.a_bs = n; g.a_rs = (long long)U * n; g.a_cs = 1;
    g.Bm = w1; g.b_bs = (long long)H * n; g.b_rs = 1; g.b_cs = n;      // B[q][m] = W1[(100u+m)*n + q]
    g.C = hpre; g.c_bs = H; g.c_rs = (long long)U * H; g.c_cs = 1;
    g.M = B; g.N = H; g.K = n; g.batch = U;
    g.bias = b1; g.bias_bs = H; g.mul = nullptr; g.mul_mode = 0; g.splits = 1; g.part = nullptr;
    if (launch_gemm(g, to_stream(s))) { set_error("xnn_fc1_fwd: launch failed"); return XNN_ERR_CUDA; }
    return XNN_OK;
}

// one warp per (b,u): lanes stride over the 100 hidden channels
__global__ void __launch_bounds__(256)
fc2_fwd_kernel(const float* __restrict__ hpre, const float* __restrict__ scale2, const float* __restrict__ shift2,
               const float* __restrict__ w2, const float* __restrict__ b2, long long BU, int U, float p,
               unsigned long long seed, const uint8_t* __restrict__ mask, float* __restrict__ opre) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const int H = XNN_HIDDEN;
    for (long long bu = warp; bu < BU; bu += nwarps) {
        const int u = (int)(bu % U);
        float acc = 0.0f;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int m = lane + 32 * r;
            if (m < H) {
                const int ch = u * H + m;
                float h = fmaxf(fmaf(hpre[bu * H + m], scale2[ch], shift2[ch]), 0.0f);
                h *= dropout_scale(p, seed, (unsigned long long)(bu * H + m), mask);
                acc = fmaf(h, w2[ch], acc);
            }
        }
        acc = warp_sum(acc);
        if (lane == 0) opre[bu] = acc + b2[u];
    }
}

extern "C" int xnn_fc2_fwd(const float* hpre, const float* scale2, const float* shift2, const float* w2,
                           const float* b2, int B, int U, float drop_p, unsigned long long seed,
                           const uint8_t* mask, float* opre, xnn_stream_t s) {
    XNN_CHECK_ARG(hpre && scale2 && shift2 && w2 && b2 && opre, "NULL pointer");
    XNN_CHECK_ARG(B >= 0 && U >= 1 && drop_p >= 0.0f && drop_p < 1.0f, "bad sizes");
    if (B == 0) return XNN_OK;
    const long long BU = (long long)B * U;
    long long blocks = (BU + 7) / 8;
    if (blocks > (long long)num_sms() * 16) blocks = (long long)num_sms() * 16;
    fc2_fwd_kernel<<<(int)blocks, 256, 0, to_stream(s)>>>(hpre, scale2, shift2, w2, b2, BU, U, drop_p, seed, mask,
                                                          opre);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// Final Linear forward for F <= 64 outputs: sgemm_kernel's 64 x 64 tiles give B/64 blocks (64 at B = 4096, on 148
// SMs) that each waste the columns beyond F.  Here: 32 batch rows per block, W^T and the x tile in shared memory,
// warp = 4 rows, lane = 2 outputs; per 4 values of u four 16-byte loads of x (broadcast), four 8-byte loads of W^T
// and 32 FMAs.  Sum over u ascending in one accumulator per output, as sgemm_kernel does.
#define LF_ROWS 32
#define LF_FS 64
#define LF_WS 66      // row stride of W^T in shared memory (floats)
__global__ void __launch_bounds__(256)
linear_fwd_small_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b, int B,
                        int U, int F, float* __restrict__ y) {
    extern __shared__ __align__(16) float s_lf[];
    const int Up = (U + 3) & ~3, S = Up + 4;
    float* s_w = s_lf;                      // [Up][LF_WS]  (W^T, zero beyond U and F)
    float* s_x = s_lf + (size_t)Up * LF_WS;  // [LF_ROWS][S]
    const int r0 = blockIdx.x * LF_ROWS, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < LF_FS * Up; i += 256) {
        const int f = i / Up, u = i - f * Up;                  // lanes along u: coalesced reads, 2-way conflicts at stride 66
        s_w[u * LF_WS + f] = (f < F && u < U) ? __ldg(w + (long long)f * U + u) : 0.0f;
    }
    for (int i = threadIdx.x; i < LF_ROWS * Up; i += 256) {
        const int r = i / Up, u = i - r * Up;
        s_x[r * S + u] = (r0 + r < B && u < U) ? x[(long long)(r0 + r) * U + u] : 0.0f;
    }
    __syncthreads();
    float acc[4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i][0] = acc[i][1] = 0.0f;
    const float* xr = s_x + (warp * 4) * S;
    const float* wc = s_w + 2 * lane;
    for (int u = 0; u < Up; u += 4) {
        float4 xv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) xv[i] = *reinterpret_cast<const float4*>(xr + i * S + u);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float2 wv = *reinterpret_cast<const float2*>(wc + (u + k) * LF_WS);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float xs = k == 0 ? xv[i].x : k == 1 ? xv[i].y : k == 2 ? xv[i].z : xv[i].w;
                acc[i][0] = fmaf(xs, wv.x, acc[i][0]);
                acc[i][1] = fmaf(xs, wv.y, acc[i][1]);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = r0 + warp * 4 + i;
        if (r >= B) continue;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int f = 2 * lane + j;
            if (f < F) y[(long long)r * F + f] = acc[i][j] + (b ? b[f] : 0.0f);
        }
    }
}

static int linear_fwd_launch(const float* x, const float* w, const float* b, int B, int U, int F, float* y,
                             cudaStream_t st) {
    const int Up = (U + 3) & ~3;
    const size_t lf_smem = ((size_t)Up * LF_WS + (size_t)LF_ROWS * (Up + 4)) * sizeof(float);
    if (F <= LF_FS && lf_smem <= 200 * 1024) {
        if (cudaFuncSetAttribute(linear_fwd_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lf_smem) !=
            cudaSuccess)
            return 1;
        linear_fwd_small_kernel<<<ceil_div(B, LF_ROWS), 256, lf_smem, st>>>(x, w, b, B, U, F, y);
        return cudaGetLastError() != cudaSuccess;
    }
    GemmArgs g{};
    g.A = x; g.a_bs = 0; g.a_rs = U; g.a_cs = 1;
    g.Bm = w; g.b_bs = 0; g.b_rs = 1; g.b_cs = U;            // B[u][f] = W[f*U + u]
    g.C = y; g.c_bs = 0; g.c_rs = F; g.c_cs = 1;
    g.M = B; g.N = F; g.K = U; g.batch = 1;
    g.bias = b; g.bias_bs = 0; g.mul = nullptr; g.mul_mode = 0; g.splits = 1; g.part = nullptr;
    return launch_gemm(g, st);
}

__global__ void __launch_bounds__(256)
affine_relu_kernel(const float* __restrict__ x, const float* __restrict__ scale, const float* __restrict__ shift,
                   long long total, int C, float* __restrict__ out) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        out[i] = fmaxf(fmaf(x[i], scale[c], shift[c]), 0.0f);
    }
}

extern "C" int xnn_head_fwd(const float* opre, const float* scale3, const float* shift3, const float* wf,
                            const float* bf, int B, int U, int F, float* o, float* logits, xnn_stream_t s) {
    XNN_CHECK_ARG(opre && scale3 && shift3 && o, "NULL pointer");
    XNN_CHECK_ARG(B >= 0 && U >= 1 && F >= 1, "bad sizes");
    if (B == 0) return XNN_OK;
    const long long total = (long long)B * U;
    long long blocks = (total + 255) / 256;
    if (blocks > (long long)num_sms() * 16) blocks = (long long)num_sms() * 16;
    affine_relu_kernel<<<(int)blocks, 256, 0, to_stream(s)>>>(opre, scale3, shift3, total, U, o);
    XNN_CHECK_LAUNCH();
    if (logits) {
        XNN_CHECK_ARG(wf && bf, "NULL pointer");
        if (linear_fwd_launch(o, wf, bf, B, U, F, logits, to_stream(s))) { set_error("xnn_head_fwd: launch failed"); return XNN_ERR_CUDA; }
    }
    return XNN_OK;
}

// ---------------------------------------------------------------------------
// backward
// ---------------------------------------------------------------------------
extern "C" int xnn_linear_fwd(const float* x, const float* w, const float* b, int B, int U, int F, float* y,
                              xnn_stream_t s) {
    XNN_CHECK_ARG(x && w && b && y, "NULL pointer");
    XNN_CHECK_ARG(B >= 0 && U >= 1 && F >= 1, "bad sizes");
    if (B == 0) return XNN_OK;
    if (linear_fwd_launch(x, w, b, B, U, F, y, to_stream(s))) { set_error("xnn_linear_fwd: launch failed"); return XNN_ERR_CUDA; }
    return XNN_OK;
}

extern "C" size_t xnn_linear_bwd_work_bytes(int B, int U, int F) {
    const size_t splitk = (size_t)splitk_wide(B) * F * U * sizeof(float);
    const size_t cs = xnn_bn_work_bytes(B, F);
    return splitk + cs + 256;
}

__global__ void colsum_finalize_kernel(const double* __restrict__ part, int tiles, int C, float* __restrict__ out) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double s1 = 0.0;
    for (int t = 0; t < tiles; ++t) s1 += part[((long long)t * 2) * C + c];
    out[c] = (float)s1;
}

extern "C" int xnn_linear_bwd(const float* dlogits, const float* wf, const float* o, int B, int U, int F, float* g3,
                              float* dwf, float* dbf, void* work, xnn_stream_t s) {
    XNN_CHECK_ARG(dlogits && wf && o && g3 && dwf && dbf && work, "NULL pointer");
    XNN_CHECK_ARG(B >= 1 && U >= 1 && F >= 1, "bad sizes");
    cudaStream_t st = to_stream(s);
    // dx = dy . W
    GemmArgs g{};
    g.A = dlogits; g.a_bs = 0; g.a_rs = F; g.a_cs = 1;
    g.Bm = wf; g.b_bs = 0; g.b_rs = U; g.b_cs = 1;
    g.C = g3; g.c_bs = 0; g.c_rs = U; g.c_cs = 1;
    g.M = B; g.N = U; g.K = F; g.batch = 1;
    g.bias = nullptr; g.bias_bs = 0; g.mul = nullptr; g.mul_mode = 0; g.splits = 1; g.part = nullptr;
    if (launch_gemm(g, st)) { set_error("xnn_linear_bwd: launch failed"); return XNN_ERR_CUDA; }
    // dWf = dlogits^T . o   (F x B) . (B x U), split over the batch
    GemmArgs h{};
    h.A = dlogits; h.a_bs = 0; h.a_rs = 1; h.a_cs = F;
    h.Bm = o; h.b_bs = 0; h.b_rs = U; h.b_cs = 1;
    h.C = dwf; h.c_bs = 0; h.c_rs = U; h.c_cs = 1;
    h.M = F; h.N = U; h.K = B; h.batch = 1;
    h.bias = nullptr; h.bias_bs = 0; h.mul = nullptr; h.mul_mode = 0;
    h.splits = splitk_wide(B); h.part = (float*)work;
    if (launch_gemm(h, st)) { set_error("xnn_linear_bwd: launch failed"); return XNN_ERR_CUDA; }
    // dbf = column sums of dlogits
    size_t off = ((size_t)h.splits * F * U * sizeof(float) + 255) & ~(size_t)255;
    double* part = (double*)((char*)work + off);
    const int tiles = colstats_tiles(B);
    dim3 grid(ceil_div(F, 32), tiles);
    colsums_kernel<<<grid, 256, 0, st>>>(dlogits, nullptr, nullptr, B, F, tiles, part);
    XNN_CHECK_LAUNCH();
    colsum_finalize_kernel<<<ceil_div(F, 256), 256, 0, st>>>(part, tiles, F, dbf);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// FC2 backward: thread per hidden channel (u,m) of a 32-channel group, walking a
// tile of batch rows: writes g2 and accumulates dW2 partials; db2 comes from the
// m == 0 threads.
__global__ void __launch_bounds__(256)
fc2_bwd_kernel(const float* __restrict__ dopre, const float* __restrict__ hpre, const float* __restrict__ scale2,
               const float* __restrict__ shift2, const float* __restrict__ w2, int B, int U, int tiles, float p,
               unsigned long long seed, const uint8_t* __restrict__ mask, float* __restrict__ g2,
               double* __restrict__ part /* [tiles][2][C] : dW2 partial, db2 partial (m==0 slots) */) {
    __shared__ double s1[8][32], s2[8][32];
    const int H = XNN_HIDDEN, C = U * H;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ch = blockIdx.x * 32 + lane;
    const int rows_per = (B + tiles - 1) / tiles;
    const int r0 = blockIdx.y * rows_per, r1 = min(B, r0 + rows_per);
    double a1 = 0.0, a2 = 0.0;
    if (ch < C) {
        const int u = ch / H;
        const float sc = scale2[ch], sh = shift2[ch], w = w2[ch];
        for (int r = r0 + warp; r < r1; r += 8) {
            const long long i = (long long)r * C + ch;
            const float pre = fmaf(hpre[i], sc, sh);
            const float ds = dropout_scale(p, seed, (unsigned long long)i, mask);
            const float d = dopre[(long long)r * U + u];
            const float hact = fmaxf(pre, 0.0f) * ds;
            g2[i] = (pre > 0.0f) ? d * w * ds : 0.0f;
            a1 += (double)d * (double)hact;
            a2 += (double)d;
        }
    }
    s1[warp][lane] = a1; s2[warp][lane] = a2;
    __syncthreads();
    if (warp == 0 && ch < C) {
        double t1 = 0.0, t2 = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) { t1 += s1[w][lane]; t2 += s2[w][lane]; }
        part[((long long)blockIdx.y * 2 + 0) * C + ch] = t1;
        part[((long long)blockIdx.y * 2 + 1) * C + ch] = t2;
    }
}

__global__ void __launch_bounds__(256)
fc2_bwd_finalize_kernel(const double* __restrict__ part, int tiles, int U, float* __restrict__ dw2,
                        float* __restrict__ db2) {
    const int H = XNN_HIDDEN, C = U * H;
    const int ch = blockIdx.x * blockDim.x + threadIdx.x;
    if (ch >= C) return;
    double t1 = 0.0, t2 = 0.0;
    for (int t = 0; t < tiles; ++t) { t1 += part[((long long)t * 2) * C + ch]; t2 += part[((long long)t * 2 + 1) * C + ch]; }
    dw2[ch] = (float)t1;
    if (ch % H == 0) db2[ch / H] = (float)t2;
}

extern "C" size_t xnn_fc2_bwd_work_bytes(int B, int U) { return xnn_bn_work_bytes(B, U * XNN_HIDDEN); }

extern "C" int xnn_fc2_bwd(const float* dopre, const float* hpre, const float* scale2, const float* shift2,
                           const float* w2, int B, int U, float drop_p, unsigned long long seed,
                           const uint8_t* mask, float* dw2, float* db2, float* g2, void* work, xnn_stream_t s) {
    XNN_CHECK_ARG(dopre && hpre && scale2 && shift2 && w2 && dw2 && db2 && g2 && work, "NULL pointer");
    XNN_CHECK_ARG(B >= 1 && U >= 1 && drop_p >= 0.0f && drop_p < 1.0f, "bad sizes");
    const int C = U * XNN_HIDDEN, tiles = colstats_tiles(B);
    dim3 grid(ceil_div(C, 32), tiles);
    fc2_bwd_kernel<<<grid, 256, 0, to_stream(s)>>>(dopre, hpre, scale2, shift2, w2, B, U, tiles, drop_p, seed, mask,
                                                   g2, (double*)work);
    XNN_CHECK_LAUNCH();
    fc2_bwd_finalize_kernel<<<ceil_div(C, 256), 256, 0, to_stream(s)>>>((const double*)work, tiles, U, dw2, db2);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

extern "C" size_t xnn_fc1_bwd_work_bytes(int B, int U, int n) {
    const size_t splitk = (((size_t)splitk_for(B) * U * XNN_HIDDEN * n * sizeof(float)) + 255) & ~(size_t)255;
    return splitk + xnn_bn_work_bytes(B, U * XNN_HIDDEN) + 256;
}

extern "C" int xnn_fc1_bwd(const float* dhpre, const float* pool, const float* w1, int B, int U, int n,
                           float* dw1, float* db1, float* dy, void* work, xnn_stream_t s) {
    XNN_CHECK_ARG(dhpre && pool && w1 && dw1 && db1 && dy && work, "NULL pointer");
    XNN_CHECK_ARG(B >= 1 && U >= 1 && n >= 1, "bad sizes");
    cudaStream_t st = to_stream(s);
    const int H = XNN_HIDDEN;
    // dW1_u (100 x n) = dhpre_u^T (100 x B) . pool_u (B x n)
    GemmArgs g{};
    g.A = dhpre; g.a_bs = H; g.a_rs = 1; g.a_cs = (long long)U * H;
    g.Bm = pool; g.b_bs = n; g.b_rs = (long long)U * n; g.b_cs = 1;
    g.C = dw1; g.c_bs = (long long)H * n; g.c_rs = n; g.c_cs = 1;
    g.M = H; g.N = n; g.K = B; g.batch = U;
    g.bias = nullptr; g.bias_bs = 0; g.mul = nullptr; g.mul_mode = 0;
    g.splits = splitk_for(B); g.part = (float*)work;
    if (launch_gemm(g, st)) { set_error("xnn_fc1_bwd: launch failed"); return XNN_ERR_CUDA; }
    // dy_u (B x n) = pool_u * (dhpre_u (B x 100) . W1_u (100 x n))
    GemmArgs h{};
    h.A = dhpre; h.a_bs = H; h.a_rs = (long long)U * H; h.a_cs = 1;
    h.Bm = w1; h.b_bs = (long long)H * n; h.b_rs = n; h.b_cs = 1;
    h.C = dy; h.c_bs = n; h.c_rs = (long long)U * n; h.c_cs = 1;
    h.M = B; h.N = n; h.K = H; h.batch = U;
    h.bias = nullptr; h.bias_bs = 0; h.mul = pool; h.mul_mode = 1; h.splits = 1; h.part = nullptr;
    if (launch_gemm(h, st)) { set_error("xnn_fc1_bwd: launch failed"); return XNN_ERR_CUDA; }
    // db1 = column sums of dhpre (mathematically 0 in front of a train-mode BatchNorm)
    const size_t off = (((size_t)g.splits * U * H * n * sizeof(float)) + 255) & ~(size_t)255;
    double* part = (double*)((char*)work + off);
    const int C = U * H, tiles = colstats_tiles(B);
    dim3 grid(ceil_div(C, 32), tiles);
    colsums_kernel<<<grid, 256, 0, st>>>(dhpre, nullptr, nullptr, B, C, tiles, part);
    XNN_CHECK_LAUNCH();
    colsum_finalize_kernel<<<ceil_div(C, 256), 256, 0, st>>>(part, tiles, C, db1);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// ---------------------------------------------------------------------------
// Adam (torch.optim.Adam defaults: no amsgrad, L2 weight decay added to grad)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
            long long numel, float lr, float b1, float b2, float eps, float wd, float bc1, float bc2_sqrt,
            float grad_scale) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < numel;
         i += (long long)gridDim.x * blockDim.x) {
        float gi = g[i] * grad_scale;
        const float pi = p[i];
        if (wd != 0.0f) gi = fmaf(wd, pi, gi);
        const float mi = m[i] + (1.0f - b1) * (gi - m[i]);        // lerp, as torch's single-tensor Adam
        const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
        m[i] = mi; v[i] = vi;
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        p[i] = pi - (lr / bc1) * (mi / denom);
    }
}

extern "C" int xnn_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long numel,
                             float lr, float beta1, float beta2, float eps, float weight_decay, long long step,
                             float grad_scale, xnn_stream_t s) {
    XNN_CHECK_ARG(param && grad && exp_avg && exp_avg_sq, "NULL pointer");
    XNN_CHECK_ARG(numel >= 0 && step >= 1, "bad sizes");
    if (numel == 0) return XNN_OK;
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    long long blocks = (numel + 255) / 256;
    if (blocks > (long long)num_sms() * 16) blocks = (long long)num_sms() * 16;
    adam_kernel<<<(int)blocks, 256, 0, to_stream(s)>>>(param, grad, exp_avg, exp_avg_sq, numel, lr, beta1, beta2, eps,
                                                       weight_decay, (float)bc1, (float)sqrt(bc2), grad_scale);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// ---------------------------------------------------------------------------
// The same update over a LIST of tensors in one launch (the reference's optimizer.step() walks 14 parameters; 14
// launches of a few microseconds each cost more in gaps than in work).  Element arithmetic is adam_kernel's, so the
// results are bit-identical to per-tensor calls.
// ---------------------------------------------------------------------------
#define ADAM_MAX_TENSORS 32
#define ADAM_CHUNK 2048              // elements per block
struct AdamList {
    float* p[ADAM_MAX_TENSORS];
    const float* g[ADAM_MAX_TENSORS];
    float* m[ADAM_MAX_TENSORS];
    float* v[ADAM_MAX_TENSORS];
    long long numel[ADAM_MAX_TENSORS];
    int first_block[ADAM_MAX_TENSORS + 1];
    int n;
};

__global__ void __launch_bounds__(256)
adam_multi_kernel(const __grid_constant__ AdamList L, float lr, float b1, float b2, float eps, float wd, float bc1,
                  float bc2_sqrt, float grad_scale, const long long* __restrict__ step_dev) {
    if (step_dev != nullptr) {
        // device-resident step count (CUDA-graph replays): the same fp64 bias corrections the host computes
        const double st = (double)*step_dev;
        bc1 = (float)(1.0 - pow((double)b1, st));
        bc2_sqrt = (float)sqrt(1.0 - pow((double)b2, st));
    }
    int t = 0;
    while (t + 1 < L.n && (int)blockIdx.x >= L.first_block[t + 1]) ++t;
    const long long base = (long long)((int)blockIdx.x - L.first_block[t]) * ADAM_CHUNK;
    const long long end = min(L.numel[t], base + ADAM_CHUNK);
    float* __restrict__ p = L.p[t];
    const float* __restrict__ g = L.g[t];
    float* __restrict__ m = L.m[t];
    float* __restrict__ v = L.v[t];
    for (long long i = base + threadIdx.x; i < end; i += 256) {
        float gi = g[i] * grad_scale;
        const float pi = p[i];
        if (wd != 0.0f) gi = fmaf(wd, pi, gi);
        const float mi = m[i] + (1.0f - b1) * (gi - m[i]);
        const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
        m[i] = mi; v[i] = vi;
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        p[i] = pi - (lr / bc1) * (mi / denom);
    }
}

extern "C" int xnn_adam_step_multi(int n_tensors, float* const* params, const float* const* grads,
                                   float* const* exp_avg, float* const* exp_avg_sq, const long long* numels, float lr,
                                   float beta1, float beta2, float eps, float weight_decay, long long step,
                                   const long long* step_dev, float grad_scale, xnn_stream_t s) {
    XNN_CHECK_ARG(n_tensors >= 0 && (step >= 1 || step_dev != nullptr), "bad sizes");
    if (step < 1) step = 1;
    XNN_CHECK_ARG(n_tensors == 0 || (params && grads && exp_avg && exp_avg_sq && numels), "NULL pointer");
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    for (int t0 = 0; t0 < n_tensors; t0 += ADAM_MAX_TENSORS) {
        AdamList L{};
        int blocks = 0;
        for (int t = t0; t < n_tensors && t < t0 + ADAM_MAX_TENSORS; ++t) {
            XNN_CHECK_ARG(numels[t] >= 0, "negative numel");
            if (numels[t] == 0) continue;
            XNN_CHECK_ARG(params[t] && grads[t] && exp_avg[t] && exp_avg_sq[t], "NULL tensor pointer");
            const long long nb = (numels[t] + ADAM_CHUNK - 1) / ADAM_CHUNK;
            XNN_CHECK_ARG(nb + blocks < (1LL << 30), "too many elements for one launch");
            L.p[L.n] = params[t]; L.g[L.n] = grads[t]; L.m[L.n] = exp_avg[t]; L.v[L.n] = exp_avg_sq[t];
            L.numel[L.n] = numels[t]; L.first_block[L.n] = blocks;
            blocks += (int)nb;
            ++L.n;
        }
        if (L.n == 0) continue;
        L.first_block[L.n] = blocks;
        adam_multi_kernel<<<blocks, 256, 0, to_stream(s)>>>(L, lr, beta1, beta2, eps, weight_decay, (float)bc1,
                                                            (float)sqrt(bc2), grad_scale, step_dev);
        XNN_CHECK_LAUNCH();
    }
    return XNN_OK;
}

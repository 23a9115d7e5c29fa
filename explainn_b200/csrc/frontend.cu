// a3 / a6 -- the conv front end of ExplaiNN on packed one-hot DNA:
//   z = conv(x), y = BN1(z), a = exp(y), pool = MaxPool1d(7,7)(a)
// (explainn/architectures/__init__.py:73-83,109-111) and its backward.
//
// One-hot input turns the convolution into a gather-sum over filter weights:
//   z[b,u,p] = sum_j W[u, base(b,p+j), j]          (N bases contribute 0)
// and makes the train-mode BatchNorm statistics of z a closed form of integer
// co-occurrence counts of the im2col features (xnn_cooc_counts): with
// phi[b,p] in {0,1}^{4k}, s = sum phi, S = sum phi phi^T, N = B*Lc,
//   mean_u = cb_u + w_u.s/N,  var_u = w_u^T (S/N - s s^T/N^2) w_u,
// so neither a statistics pass nor a dense weight-gradient pass over (B,U,Lc)
// is needed: dW_u = (gamma/sigma) [G_u - (dbeta/N) s - (dgamma/sigma) Cov w_u]
// where G_u is a sparse scatter of dy at the max-pool arg-max positions.
#include "common.cuh"
using namespace xnn;

#define XNN_KMAX 32  // largest kernel_size supported by the front end

// ---------------------------------------------------------------------------
// co-occurrence counts (bit-parallel over the packed planes)
// thread t -> (d, c, c') = (t/16, (t%16)/4, t%4): counts pairs base(i)=c,
// base(i+d)=c' over the windows i in [j, j+Lc) for every j in [0, k-d).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(16 * XNN_KMAX)
cooc_kernel(const uint32_t* __restrict__ planes, int B, int L, int Wp, int k,
            unsigned long long* __restrict__ cnt1, unsigned long long* __restrict__ cnt2) {
    extern __shared__ uint32_t sm_bits[];  // [4][Wp+1]
    const int Lc = L - k + 1;
    const int t = threadIdx.x;
    const int d = t >> 4, c = (t >> 2) & 3, c2 = t & 3;
    const bool active = d < k;
    int acc[XNN_KMAX];
#pragma unroll
    for (int j = 0; j < XNN_KMAX; ++j) acc[j] = 0;
    const int Ws = Wp + 1;

    for (int b = blockIdx.x; b < B; b += gridDim.x) {
        const uint32_t* sp = planes + (long long)b * 3 * Wp;
        for (int i = t; i < 4 * Ws; i += blockDim.x) {
            const int cc = i / Ws, w = i % Ws;
            uint32_t v = 0;
            if (w < Wp) {
                const uint32_t lo = sp[w], hi = sp[Wp + w], nm = sp[2 * Wp + w];
                const uint32_t mlo = (cc & 1) ? lo : ~lo, mhi = (cc & 2) ? hi : ~hi;
                v = mlo & mhi & ~nm;
            }
            sm_bits[i] = v;
        }
        __syncthreads();
        if (active) {
            const uint32_t* P = sm_bits + c * Ws;
            const uint32_t* Q = sm_bits + c2 * Ws;
            // A0 = #{i in [0,Lc): P[i] & Q[i+d]}
            int a = 0;
            const int full = Lc >> 5, rem = Lc & 31;
            for (int w = 0; w <= full; ++w) {
                if (w == full && rem == 0) break;
                uint32_t m = P[w] & __funnelshift_r(Q[w], Q[w + 1], d);
                if (w == full) m &= (1u << rem) - 1u;
                a += __popc(m);
            }
#pragma unroll
            for (int j = 0; j < XNN_KMAX; ++j) {
                if (j < k - d) {
                    acc[j] += a;
                    // slide the window: drop i = j, add i = j + Lc
                    const int i0 = j, i1 = j + Lc;
                    const int drop = ((P[i0 >> 5] >> (i0 & 31)) & 1u) & ((Q[(i0 + d) >> 5] >> ((i0 + d) & 31)) & 1u);
                    int add = 0;
                    if (i1 + d < L)
                        add = ((P[i1 >> 5] >> (i1 & 31)) & 1u) & ((Q[(i1 + d) >> 5] >> ((i1 + d) & 31)) & 1u);
                    a += add - drop;
                }
            }
        }
        __syncthreads();
    }
    if (active) {
        const int K4 = 4 * k;
#pragma unroll
        for (int j = 0; j < XNN_KMAX; ++j) {
            if (j < k - d && acc[j] != 0) {
                const int f = c * k + j, f2 = c2 * k + j + d;
                atomicAdd(&cnt2[(long long)f * K4 + f2], (unsigned long long)acc[j]);
                if (d > 0) atomicAdd(&cnt2[(long long)f2 * K4 + f], (unsigned long long)acc[j]);
                if (d == 0 && c == c2) atomicAdd(&cnt1[f], (unsigned long long)acc[j]);
            }
        }
    }
}

extern "C" int xnn_cooc_counts(const uint32_t* planes, int B, int L, int k, long long* cnt1, long long* cnt2,
                               xnn_stream_t s) {
    XNN_CHECK_ARG(planes && cnt1 && cnt2, "NULL pointer");
    XNN_CHECK_ARG(k >= 1 && k <= XNN_KMAX && L >= k && B >= 0, "bad sizes (kernel_size <= 32)");
    // the kernel accumulates with atomics: the counts are cleared here (memset nodes, no library fill kernels)
    XNN_CUDA(cudaMemsetAsync(cnt1, 0, (size_t)4 * k * sizeof(long long), to_stream(s)));
    XNN_CUDA(cudaMemsetAsync(cnt2, 0, (size_t)16 * k * k * sizeof(long long), to_stream(s)));
    if (B == 0) return XNN_OK;
    const int Wp = planes_words(L);
    const int threads = ((16 * k + 31) / 32) * 32;
    // two blocks per SM: every block ends with 2 * 16 k^2 / 2 global atomics on the same addresses, so the tail grows with
    // the grid (measured at cfg3, B = 4096: 1 block per SM 0.115 ms, 2: 0.082, 4: 0.098, 6: 0.110, 12: 0.158)
    int grid = num_sms() * 2;
    if (grid > B) grid = B;
    const size_t smem = (size_t)4 * (Wp + 1) * sizeof(uint32_t);
    XNN_CHECK_ARG(smem <= 48 * 1024, "sequence too long for cooc kernel");
    cooc_kernel<<<grid, threads, smem, to_stream(s)>>>(planes, B, L, Wp, k, (unsigned long long*)cnt1,
                                                       (unsigned long long*)cnt2);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// ---------------------------------------------------------------------------
// BN1 train-mode statistics from the counts (fp64), one block per unit
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
bn1_prepare_kernel(const long long* __restrict__ cnt1, const long long* __restrict__ cnt2, long long n_pos,
                   const float* __restrict__ conv_w, const float* __restrict__ conv_b,
                   const float* __restrict__ gamma, const float* __restrict__ beta, float eps, float momentum,
                   int U, int k, float* __restrict__ running_mean, float* __restrict__ running_var,
                   float* __restrict__ scale, float* __restrict__ shift, double* __restrict__ stats,
                   double* __restrict__ cov_w) {
    __shared__ double sw[4 * XNN_KMAX];
    __shared__ double red[2][4];
    const int u = blockIdx.x, f = threadIdx.x, K4 = 4 * k;
    const double N = (double)n_pos;
    if (f < K4) sw[f] = (double)conv_w[(long long)u * K4 + f];
    __syncthreads();
    double cw = 0.0, ws_part = 0.0;
    if (f < K4) {
        // cnt2 is symmetric: row g read across the threads (coalesced); one reciprocal instead of 2 K4 fp64 divisions
        // per thread (the kernel took 36 us for 300 units, most of it in the division sequences)
        const double invN = 1.0 / N;
        const double sf = (double)cnt1[f] * invN;
#pragma unroll 4
        for (int g = 0; g < K4; ++g) {
            const double cov = (double)cnt2[(long long)g * K4 + f] * invN - sf * ((double)cnt1[g] * invN);
            cw += cov * sw[g];
        }
        cov_w[(long long)u * K4 + f] = cw;
        ws_part = sw[f] * sf;
    }
    double var_part = (f < K4) ? cw * sw[f] : 0.0;
    var_part = warp_sum(var_part);
    ws_part = warp_sum(ws_part);
    if ((f & 31) == 0) { red[0][f >> 5] = var_part; red[1][f >> 5] = ws_part; }
    __syncthreads();
    if (f == 0) {
        double var = 0.0, ws = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { var += red[0][w]; ws += red[1][w]; }
        if (var < 0.0) var = 0.0;
        const double cb = conv_b ? (double)conv_b[u] : 0.0;
        const double mean = cb + ws;
        const double invstd = 1.0 / sqrt(var + (double)eps);
        const double sc = (double)gamma[u] * invstd;
        scale[u] = (float)sc;
        shift[u] = (float)((double)beta[u] - ws * sc);
        stats[4 * u + 0] = mean; stats[4 * u + 1] = var; stats[4 * u + 2] = invstd; stats[4 * u + 3] = ws;
        if (running_mean && running_var) {
            const double unb = (n_pos > 1) ? var * N / (N - 1.0) : var;
            running_mean[u] = (float)((1.0 - momentum) * (double)running_mean[u] + momentum * mean);
            running_var[u] = (float)((1.0 - momentum) * (double)running_var[u] + momentum * unb);
        }
    }
}

extern "C" int xnn_bn1_train_prepare(const long long* cnt1, const long long* cnt2, long long n_pos,
                                     const float* conv_w, const float* conv_b, const float* gamma,
                                     const float* beta, float eps, float momentum, int U, int k,
                                     float* running_mean, float* running_var, float* scale, float* shift,
                                     double* stats, double* cov_w, xnn_stream_t s) {
    XNN_CHECK_ARG(cnt1 && cnt2 && conv_w && gamma && beta && scale && shift && stats && cov_w, "NULL pointer");
    XNN_CHECK_ARG(k >= 1 && k <= XNN_KMAX && U >= 1 && n_pos >= 1, "bad sizes");
    bn1_prepare_kernel<<<U, 128, 0, to_stream(s)>>>(cnt1, cnt2, n_pos, conv_w, conv_b, gamma, beta, eps, momentum,
                                                    U, k, running_mean, running_var, scale, shift, stats, cov_w);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

__global__ void bn_fold_eval_kernel(const float* __restrict__ gamma, const float* __restrict__ beta,
                                    const float* __restrict__ rm, const float* __restrict__ rv,
                                    const float* __restrict__ pre_bias, float eps, int C,
                                    float* __restrict__ scale, float* __restrict__ shift) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const double invstd = 1.0 / sqrt((double)rv[c] + (double)eps);
    const double sc = (double)gamma[c] * invstd;
    const double pb = pre_bias ? (double)pre_bias[c] : 0.0;
    scale[c] = (float)sc;
    shift[c] = (float)((double)beta[c] + (pb - (double)rm[c]) * sc);
}

extern "C" int xnn_bn_fold_eval(const float* gamma, const float* beta, const float* running_mean,
                                const float* running_var, const float* pre_bias, float eps, int C, float* scale,
                                float* shift, xnn_stream_t s) {
    XNN_CHECK_ARG(gamma && beta && running_mean && running_var && scale && shift, "NULL pointer");
    XNN_CHECK_ARG(C >= 1, "bad sizes");
    bn_fold_eval_kernel<<<ceil_div(C, 256), 256, 0, to_stream(s)>>>(gamma, beta, running_mean, running_var,
                                                                    pre_bias, eps, C, scale, shift);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// ---------------------------------------------------------------------------
// fused forward: gather-sum conv + affine + max over 7 (+ exp on the maximum)
// lanes = 32 consecutive units; a warp takes one pool window at a time.
// LUT layout in shared memory: lut[c][j][lane] (c = 0..4, row 4 all zero), so
// for a given base the 7 window outputs read lut at compile-time offsets.
// ---------------------------------------------------------------------------
template <int KT>
__global__ void __launch_bounds__(256)
frontend_fwd_kernel(const uint32_t* __restrict__ planes, int B, int L, int Wp, int U, int k_rt, int n,
                    const float* __restrict__ conv_w, const float* __restrict__ scale,
                    const float* __restrict__ shift, float* __restrict__ pool, float* __restrict__ zmax,
                    uint8_t* __restrict__ argmax) {
    const int k = KT > 0 ? KT : k_rt;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int ns = n | 1;  // odd stride: conflict-free staging
    float* lut = reinterpret_cast<float*>(smem_raw);              // [5][k][32]
    float* s_pool = lut + 5 * k * 32;                             // [32][ns]
    float* s_z = s_pool + 32 * ns;                                // [32][ns]
    uint8_t* s_arg = reinterpret_cast<uint8_t*>(s_z + 32 * ns);   // [32][ns]
    uint8_t* codes = s_arg + 32 * ns;                             // [Lpad]
    const int Lpad = (L + 7 + 15) & ~15;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int u0 = blockIdx.x * 32;
    const int u = u0 + lane;
    const int nu = min(32, U - u0);

    for (int i = threadIdx.x; i < 5 * k * 32; i += blockDim.x) {
        const int ul = i & 31, j = (i >> 5) % k, c = (i >> 5) / k;
        float v = 0.0f;
        if (c < 4 && u0 + ul < U) v = conv_w[((long long)(u0 + ul) * 4 + c) * k + j];
        lut[i] = v;
    }
    const float sc = (u < U) ? scale[u] : 0.0f;
    const float sh = (u < U) ? shift[u] : 0.0f;
    const bool train_out = (zmax != nullptr);

    for (int b = blockIdx.y; b < B; b += gridDim.y) {
        __syncthreads();  // previous copy-out finished; LUT visible on first trip
        expand_codes(planes + (long long)b * 3 * Wp, Wp, L, Lpad, codes);
        __syncthreads();
        for (int q = warp; q < n; q += nwarps) {
            float acc[XNN_POOL];
#pragma unroll
            for (int r = 0; r < XNN_POOL; ++r) acc[r] = 0.0f;
            const uint8_t* cw = codes + q * XNN_POOL;
            if (KT > 0) {
#pragma unroll
                for (int t = 0; t < KT + XNN_POOL - 1; ++t) {
                    const float* row = lut + (int)cw[t] * (KT * 32) + lane;
#pragma unroll
                    for (int r = 0; r < XNN_POOL; ++r) {
                        const int j = t - r;
                        if (j >= 0 && j < KT) acc[r] += row[j * 32];
                    }
                }
            } else {
                for (int t = 0; t < k + XNN_POOL - 1; ++t) {
                    const float* row = lut + (int)cw[t] * (k * 32) + lane;
#pragma unroll
                    for (int r = 0; r < XNN_POOL; ++r) {
                        const int j = t - r;
                        if (j >= 0 && j < k) acc[r] += row[j * 32];
                    }
                }
            }
            float best = fmaf(sc, acc[0], sh), zb = acc[0];
            int arg = 0;
#pragma unroll
            for (int r = 1; r < XNN_POOL; ++r) {
                const float y = fmaf(sc, acc[r], sh);
                if (y > best || y != y) { best = y; zb = acc[r]; arg = r; }  // first max wins; NaN propagates
            }
            s_pool[lane * ns + q] = expf(best);
            if (train_out) { s_z[lane * ns + q] = zb; s_arg[lane * ns + q] = (uint8_t)arg; }
        }
        __syncthreads();
        const long long base = ((long long)b * U + u0) * n;
        const int cnt = nu * n;
        for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
            const int ul = i / n, q = i - ul * n;
            pool[base + i] = s_pool[ul * ns + q];
            if (train_out) {
                zmax[base + i] = s_z[ul * ns + q];
                argmax[base + i] = s_arg[ul * ns + q];
            }
        }
    }
}

static size_t frontend_fwd_smem(int k, int n, int L) {
    const int ns = n | 1;
    const int Lpad = (L + 7 + 15) & ~15;
    return (size_t)5 * k * 32 * 4 + (size_t)32 * ns * (4 + 4 + 1) + Lpad + 16;
}

extern "C" int xnn_frontend_fwd_lut(const uint32_t* planes, int B, int L, int U, int k, const float* conv_w,
                                const float* scale, const float* shift, float* pool, float* zmax,
                                uint8_t* argmax, xnn_stream_t s) {
    XNN_CHECK_ARG(planes && conv_w && scale && shift && pool, "NULL pointer");
    XNN_CHECK_ARG((zmax == nullptr) == (argmax == nullptr), "zmax and argmax go together");
    XNN_CHECK_ARG(k >= 1 && k <= XNN_KMAX && L >= k && U >= 1 && B >= 0, "bad sizes (kernel_size <= 32)");
    const int n = (L - k + 1) / XNN_POOL;
    if (B == 0 || n == 0) return XNN_OK;
    const int Wp = planes_words(L);
    const size_t smem = frontend_fwd_smem(k, n, L);
    XNN_CHECK_ARG(smem <= 227 * 1024, "sequence too long for the front-end staging buffers");
    const int chunks = ceil_div(U, 32);
    int per_sm = (int)((220 * 1024) / smem);
    if (per_sm > 8) per_sm = 8;
    if (per_sm < 1) per_sm = 1;
    int gy = ceil_div((long long)num_sms() * per_sm, chunks);
    if (gy > B) gy = B;
    if (gy > 65535) gy = 65535;
    dim3 grid(chunks, gy);
    if (k == 19) {
        XNN_CUDA(cudaFuncSetAttribute(frontend_fwd_kernel<19>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        frontend_fwd_kernel<19><<<grid, 256, smem, to_stream(s)>>>(planes, B, L, Wp, U, k, n, conv_w, scale, shift,
                                                                  pool, zmax, argmax);
    } else {
        XNN_CUDA(cudaFuncSetAttribute(frontend_fwd_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        frontend_fwd_kernel<0><<<grid, 256, smem, to_stream(s)>>>(planes, B, L, Wp, U, k, n, conv_w, scale, shift,
                                                                 pool, zmax, argmax);
    }
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// ---------------------------------------------------------------------------
// per-position activations (linears[:3], test.py:159-160); lanes = positions.
// LUT layout lut[ul][c][j] so the four bases of a tap fall in distinct banks.
// ---------------------------------------------------------------------------
struct ActsLut {
    // loads the weights of `nu` units starting at u0 into lut[ul][5][k]
    __device__ static void load(float* lut, const float* __restrict__ conv_w, int u0, int nu, int k) {
        const int per = 5 * k;
        for (int i = threadIdx.x; i < nu * per; i += blockDim.x) {
            const int ul = i / per, r = i - ul * per, c = r / k, j = r - c * k;
            lut[i] = (c < 4) ? conv_w[((long long)(u0 + ul) * 4 + c) * k + j] : 0.0f;
        }
    }
};

// z for one (unit, position): codes and lut in shared memory
__device__ __forceinline__ float gather_z(const float* __restrict__ lut_u, const uint8_t* __restrict__ codes,
                                          int p, int k) {
    float z = 0.0f;
    for (int j = 0; j < k; ++j) z += lut_u[(int)codes[p + j] * k + j];
    return z;
}

#define ACTS_UNITS 32
__global__ void __launch_bounds__(256)
frontend_acts_kernel(const uint32_t* __restrict__ planes, int B, int L, int Wp, int U, int k,
                     const float* __restrict__ conv_w, const float* __restrict__ scale,
                     const float* __restrict__ shift, float* __restrict__ act_f32, __half* __restrict__ act_f16) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* lut = reinterpret_cast<float*>(smem_raw);  // [ACTS_UNITS][5][k]
    uint8_t* codes = reinterpret_cast<uint8_t*>(lut + ACTS_UNITS * 5 * k);
    const int Lc = L - k + 1, Lpad = (L + 15) & ~15;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int u0 = blockIdx.x * ACTS_UNITS, nu = min(ACTS_UNITS, U - u0);
    ActsLut::load(lut, conv_w, u0, nu, k);
    for (int b = blockIdx.y; b < B; b += gridDim.y) {
        __syncthreads();
        expand_codes(planes + (long long)b * 3 * Wp, Wp, L, Lpad, codes);
        __syncthreads();
        for (int ul = warp; ul < nu; ul += nwarps) {
            const float sc = scale[u0 + ul], sh = shift[u0 + ul];
            const float* lu = lut + ul * 5 * k;
            const long long base = ((long long)b * U + u0 + ul) * Lc;
            for (int p = lane; p < Lc; p += 32) {
                const float a = expf(fmaf(sc, gather_z(lu, codes, p, k), sh));
                if (act_f32) act_f32[base + p] = a;
                if (act_f16) act_f16[base + p] = __float2half_rn(a);
            }
        }
    }
}

static void acts_launch_dims(int B, int U, int k, int L, dim3* grid, size_t* smem) {
    *smem = (size_t)ACTS_UNITS * 5 * k * 4 + ((L + 15) & ~15) + 16;
    const int chunks = ceil_div(U, ACTS_UNITS);
    int gy = ceil_div((long long)num_sms() * 6, chunks);
    if (gy > B) gy = B;
    if (gy > 65535) gy = 65535;
    if (gy < 1) gy = 1;
    *grid = dim3(chunks, gy);
}

extern "C" int xnn_frontend_acts(const uint32_t* planes, int B, int L, int U, int k, const float* conv_w,
                                 const float* scale, const float* shift, float* act_f32, uint16_t* act_f16,
                                 xnn_stream_t s) {
    XNN_CHECK_ARG(planes && conv_w && scale && shift && (act_f32 || act_f16), "NULL pointer");
    XNN_CHECK_ARG(k >= 1 && k <= XNN_KMAX && L >= k && U >= 1 && B >= 0, "bad sizes (kernel_size <= 32)");
    if (B == 0) return XNN_OK;
    dim3 grid; size_t smem;
    acts_launch_dims(B, U, k, L, &grid, &smem);
    XNN_CHECK_ARG(smem <= 48 * 1024, "sequence too long");
    frontend_acts_kernel<<<grid, 256, smem, to_stream(s)>>>(planes, B, L, planes_words(L), U, k, conv_w, scale,
                                                            shift, act_f32, reinterpret_cast<__half*>(act_f16));
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// ---------------------------------------------------------------------------
// backward: sparse scatter of dy at the arg-max windows + closed-form BN terms
// thread = (unit ul, tap j) of a 32-unit chunk; 4 fp64 accumulators per thread
// (one per base).  A block walks its slice of the batch; per-slice partials are
// reduced in fixed order by the finalize kernel (deterministic).
// ---------------------------------------------------------------------------
#define BWD_SLICES_MAX 64
static int bwd_slices(int B) { return B < BWD_SLICES_MAX ? B : BWD_SLICES_MAX; }

__global__ void __launch_bounds__(32 * XNN_KMAX)
frontend_bwd_scatter_kernel(const uint32_t* __restrict__ planes, int B, int L, int Wp, int U, int k, int n,
                            const float* __restrict__ dy, const float* __restrict__ zmax,
                            const uint8_t* __restrict__ argmax, const double* __restrict__ stats,
                            double* __restrict__ part /* [slices][U][4k+2] */) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* s_dy = reinterpret_cast<float*>(smem_raw);            // [32][n]
    float* s_z = s_dy + 32 * n;                                  // [32][n]
    uint8_t* s_arg = reinterpret_cast<uint8_t*>(s_z + 32 * n);   // [32][n]
    uint8_t* codes = s_arg + ((32 * n + 15) & ~15);              // [Lpad]
    const int Lpad = (L + 15) & ~15;
    const int u0 = blockIdx.x * 32, nu = min(32, U - u0);
    const int t = threadIdx.x, ul = t / k, j = t - ul * k;
    const bool active = ul < nu;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0, sb = 0.0, sg = 0.0;
    double ws = 0.0, invstd = 0.0;
    if (active) { invstd = stats[4 * (u0 + ul) + 2]; ws = stats[4 * (u0 + ul) + 3]; }
    for (int b = blockIdx.y; b < B; b += gridDim.y) {
        __syncthreads();
        expand_codes(planes + (long long)b * 3 * Wp, Wp, L, Lpad, codes);
        const long long base = ((long long)b * U + u0) * n;
        for (int i = t; i < nu * n; i += blockDim.x) {
            s_dy[i] = dy[base + i]; s_z[i] = zmax[base + i]; s_arg[i] = argmax[base + i];
        }
        __syncthreads();
        if (active) {
            const float* d_row = s_dy + ul * n;
            const uint8_t* a_row = s_arg + ul * n;
            for (int q = 0; q < n; ++q) {
                const double d = (double)d_row[q];
                const int c = codes[q * XNN_POOL + (int)a_row[q] + j];
                a0 += (c == 0) ? d : 0.0;
                a1 += (c == 1) ? d : 0.0;
                a2 += (c == 2) ? d : 0.0;
                a3 += (c == 3) ? d : 0.0;
            }
            if (j == 0) {
                const float* z_row = s_z + ul * n;
                for (int q = 0; q < n; ++q) {
                    const double d = (double)d_row[q];
                    sb += d;
                    sg += d * (((double)z_row[q] - ws) * invstd);
                }
            }
        }
    }
    if (active) {
        const int K4 = 4 * k, stride = K4 + 2;
        double* o = part + ((long long)blockIdx.y * U + (u0 + ul)) * stride;
        o[0 * k + j] = a0; o[1 * k + j] = a1; o[2 * k + j] = a2; o[3 * k + j] = a3;
        if (j == 0) { o[K4] = sb; o[K4 + 1] = sg; }
    }
}

__global__ void __launch_bounds__(160)      // 4k + 2 <= 130 columns
frontend_bwd_finalize_kernel(const double* __restrict__ part, int slices, int U, int k,
                             const float* __restrict__ gamma, const double* __restrict__ stats,
                             const double* __restrict__ cov_w, const long long* __restrict__ cnt1, long long n_pos,
                             float* __restrict__ dconv_w, float* __restrict__ dconv_b, float* __restrict__ dgamma,
                             float* __restrict__ dbeta) {
    const int u = blockIdx.x, f = threadIdx.x, K4 = 4 * k, stride = K4 + 2;
    __shared__ double s_bg[2];
    // every thread f < 4k + 2 sums its column of the slice partials (same order as before: bit-identical), the loads of
    // eight slices in flight; the two BatchNorm sums are columns 4k and 4k + 1 (they used to be a second serial chain of
    // `slices` dependent loads in front of the barrier)
    double G = 0.0;
    if (f < stride) {
        const double* p = part + (long long)u * stride + f;
        const long long step = (long long)U * stride;
#pragma unroll 8
        for (int sl = 0; sl < slices; ++sl) G += p[sl * step];
        if (f >= K4) s_bg[f - K4] = G;
    }
    __syncthreads();
    const double db = s_bg[0], dg = s_bg[1];
    const double invstd = stats[4 * u + 2];
    const double N = (double)n_pos;
    if (f < K4) {
        const double v = (double)gamma[u] * invstd *
                         (G - (db / N) * (double)cnt1[f] - dg * invstd * cov_w[(long long)u * K4 + f]);
        dconv_w[(long long)u * K4 + f] = (float)v;
    }
    if (f == 0) {
        dgamma[u] = (float)dg;
        dbeta[u] = (float)db;
        if (dconv_b) dconv_b[u] = 0.0f;  // a bias in front of a train-mode BatchNorm has zero gradient
    }
}

namespace xnn {
int frontend_bwd_simt_slices(int B) { return bwd_slices(B); }
int launch_frontend_bwd_finalize(const double* part, int slices, int U, int k, const float* gamma,
                                 const double* stats, const double* cov_w, const long long* cnt1, long long n_pos,
                                 float* dconv_w, float* dconv_b, float* dgamma, float* dbeta, cudaStream_t st) {
    frontend_bwd_finalize_kernel<<<U, 160, 0, st>>>(part, slices, U, k, gamma, stats, cov_w, cnt1, n_pos, dconv_w,
                                                    dconv_b, dgamma, dbeta);
    return cudaGetLastError() != cudaSuccess;
}
}  // namespace xnn

// debugging cross-check only: the same contract on the FP64 pipe (sparse scatter, no tensor cores);
// work: xnn_frontend_bwd_work_bytes
extern "C" int xnn_frontend_bwd_simt(const uint32_t* planes, int B, int L, int U, int k, const float* dy,
                                const float* zmax, const uint8_t* argmax, const float* gamma,
                                const double* stats, const double* cov_w, const long long* cnt1, long long n_pos,
                                float* dconv_w, float* dconv_b, float* dgamma, float* dbeta, void* work,
                                xnn_stream_t s) {
    XNN_CHECK_ARG(planes && dy && zmax && argmax && gamma && stats && cov_w && cnt1 && dconv_w && dgamma && dbeta && work,
                  "NULL pointer");
    XNN_CHECK_ARG(k >= 1 && k <= XNN_KMAX && L >= k && U >= 1 && B >= 1 && n_pos >= 1, "bad sizes");
    const int n = (L - k + 1) / XNN_POOL;
    XNN_CHECK_ARG(n >= 1, "sequence shorter than one pool window");
    const int slices = bwd_slices(B);
    const int Lpad = (L + 15) & ~15;
    const size_t smem = (size_t)32 * n * 8 + ((32 * n + 15) & ~15) + Lpad + 16;
    XNN_CHECK_ARG(smem <= 227 * 1024, "sequence too long");
    XNN_CUDA(cudaFuncSetAttribute(frontend_bwd_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(ceil_div(U, 32), slices);
    frontend_bwd_scatter_kernel<<<grid, 32 * k, smem, to_stream(s)>>>(planes, B, L, planes_words(L), U, k, n, dy,
                                                                      zmax, argmax, stats, (double*)work);
    XNN_CHECK_LAUNCH();
    frontend_bwd_finalize_kernel<<<U, 160, 0, to_stream(s)>>>((const double*)work, slices, U, k, gamma, stats,
                                                              cov_w, cnt1, n_pos, dconv_w, dconv_b, dgamma, dbeta);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

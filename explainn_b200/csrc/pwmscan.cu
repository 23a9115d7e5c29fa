// a11 / a13 / a14 -- PWM scans.
//   xnn_pwm_scan       : the `PWM` torch module (architectures/__init__.py:119-170): fp32 log-odds
//                        correlation on both strands, max or sum over all positions.
//   xnn_lpm_scan_*     : PWMScan `pwm_scoring` letter-probability mode (pwmscan/pwm_scoring.c:340-475), fp64.
//   xnn_ipwm_scan_best : PWMScan integer mode (pwm_scoring.c:477-573), int32, bit-exact.
#include "common.cuh"
#include <limits.h>
using namespace xnn;

// ---------------------------------------------------------------------------
// PWM module.  Reverse strand without flipping the sequence: the reverse-
// complement score at RC position p' equals the correlation of the forward
// window starting at p = L-w-p' with Wrc[c][j] = W[3-c][w-1-j].
// lanes = window starts, a warp owns one PWM of a 32-PWM chunk.
// ---------------------------------------------------------------------------
#define PG 32
__global__ void __launch_bounds__(256)
pwm_scan_kernel(const uint32_t* __restrict__ planes, int B, int L, int Wp, int G, int w,
                const float* __restrict__ pwm_w, int mode, float* __restrict__ out, int32_t* __restrict__ argmax) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* lut_f = reinterpret_cast<float*>(smem_raw);   // [PG][5][w]
    float* lut_r = lut_f + PG * 5 * w;                   // [PG][5][w]
    uint8_t* codes = reinterpret_cast<uint8_t*>(lut_r + PG * 5 * w);
    const int Lw = L - w + 1, Lpad = (L + 15) & ~15;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int g0 = blockIdx.x * PG, ng = min(PG, G - g0);
    const int per = 5 * w;
    for (int i = threadIdx.x; i < ng * per; i += blockDim.x) {
        const int gl = i / per, r = i - gl * per, c = r / w, j = r - c * w;
        const float* W = pwm_w + (long long)(g0 + gl) * 4 * w;
        lut_f[i] = (c < 4) ? W[c * w + j] : 0.0f;
        lut_r[i] = (c < 4) ? W[(3 - c) * w + (w - 1 - j)] : 0.0f;
    }
    for (int b = blockIdx.y; b < B; b += gridDim.y) {
        __syncthreads();
        expand_codes(planes + (long long)b * 3 * Wp, Wp, L, Lpad, codes);
        __syncthreads();
        for (int gl = warp; gl < ng; gl += nwarps) {
            const float* lf = lut_f + gl * per;
            const float* lr = lut_r + gl * per;
            float acc = 0.0f;            // sum mode
            float best = -INFINITY;      // max mode: value and index on the [fwd | rev] axis
            int bidx = 0x7fffffff;
            for (int p = lane; p < Lw; p += 32) {
                // sr is summed from the window's last base backwards: the same order in which the
                // forward score of the reverse-complemented sequence is summed, so scores are
                // bit-identical under reverse complement of the input
                float sf = 0.0f, sr = 0.0f;
                for (int j = 0; j < w; ++j) {
                    sf += lf[(int)codes[p + j] * w + j];
                    sr += lr[(int)codes[p + w - 1 - j] * w + (w - 1 - j)];
                }
                if (mode == 0) {
                    acc += sf + sr;
                } else {
                    const int ir = Lw + (L - w - p);
                    if (sf > best || (sf == best && p < bidx)) { best = sf; bidx = p; }
                    if (sr > best || (sr == best && ir < bidx)) { best = sr; bidx = ir; }
                }
            }
            const long long o = (long long)b * G + g0 + gl;
            if (mode == 0) {
                acc = warp_sum(acc);
                if (lane == 0) out[o] = acc;
            } else {
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) {
                    const float ov = __shfl_xor_sync(0xffffffffu, best, off);
                    const int oi = __shfl_xor_sync(0xffffffffu, bidx, off);
                    if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
                }
                if (lane == 0) { out[o] = best; if (argmax) argmax[o] = bidx; }
            }
        }
    }
}

extern "C" int xnn_pwm_scan(const uint32_t* planes, int B, int L, int G, int w, const float* pwm_w, int mode,
                            float* out, int32_t* argmax, xnn_stream_t s) {
    XNN_CHECK_ARG(planes && pwm_w && out, "NULL pointer");
    XNN_CHECK_ARG(w >= 1 && L >= w && G >= 1 && B >= 0 && (mode == 0 || mode == 1), "bad sizes");
    if (B == 0) return XNN_OK;
    const size_t smem = (size_t)2 * PG * 5 * w * 4 + ((L + 15) & ~15) + 16;
    XNN_CHECK_ARG(smem <= 200 * 1024, "PWM too wide / sequence too long");
    XNN_CUDA(cudaFuncSetAttribute(pwm_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int chunks = ceil_div(G, PG);
    int per_sm = (int)((200 * 1024) / smem);
    if (per_sm > 6) per_sm = 6;
    if (per_sm < 1) per_sm = 1;
    int gy = ceil_div((long long)num_sms() * per_sm, chunks);
    if (gy > B) gy = B;
    if (gy > 65535) gy = 65535;
    if (gy < 1) gy = 1;
    dim3 grid(chunks, gy);
    pwm_scan_kernel<<<grid, 256, smem, to_stream(s)>>>(planes, B, L, planes_words(L), G, w, pwm_w, mode, out, argmax);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// ---------------------------------------------------------------------------
// PWMScan letter-probability mode: one block per FASTA record.
// prod = prod * lpm[c][j] / bg[c] in the reference's operation order, so each
// window product is bit-identical to the C tool; only the order of the final
// sum over windows differs (sum mode).
// ---------------------------------------------------------------------------
#define LPM_WMAX 64
struct LpmTab {
    double l[5][LPM_WMAX];
    double bg[5];
};

__device__ __forceinline__ void lpm_window(const int8_t* __restrict__ sq, int w, const LpmTab& T, bool fwd_only,
                                           double& prod, double& prod_rc) {
    prod = 1.0; prod_rc = 1.0;
    for (int j = 0; j < w; ++j) {
        const int c = sq[j];
        prod = prod * T.l[c][j] / T.bg[c];
        if (!fwd_only) {
            const int r = (c == 4) ? 4 : 3 - c;
            prod_rc = prod_rc * T.l[r][w - j - 1] / T.bg[r];
        }
    }
}

__global__ void __launch_bounds__(256)
lpm_sum_kernel(const int8_t* __restrict__ codes, const long long* __restrict__ offsets, int R, int w,
               const double* __restrict__ lpm, const double* __restrict__ bg, int bg_per_record, int fwd_only,
               double* __restrict__ out_sum) {
    __shared__ LpmTab T;
    __shared__ double red[8];
    for (int i = threadIdx.x; i < 5 * w; i += blockDim.x) T.l[i / w][i % w] = lpm[i];
    if (threadIdx.x < 5) T.bg[threadIdx.x] = bg[threadIdx.x];
    __syncthreads();
    for (int r = blockIdx.x; r < R; r += gridDim.x) {
        if (bg_per_record) {
            if (threadIdx.x < 5) T.bg[threadIdx.x] = bg[(long long)r * 5 + threadIdx.x];
            __syncthreads();
        }
        const long long beg = offsets[r], len = offsets[r + 1] - beg;
        const long long nwin = len - w + 1;
        double acc = 0.0;
        for (long long i = threadIdx.x; i < nwin; i += blockDim.x) {
            double p, pr;
            lpm_window(codes + beg + i, w, T, fwd_only != 0, p, pr);
            acc += fwd_only ? p : (p + pr);
        }
        acc = warp_sum(acc);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int i = 0; i < 8; ++i) t += red[i];
            out_sum[r] = t;
        }
        __syncthreads();
    }
}

// best mode: pass A finds (best value, first window), pass B lists later ties in order
__global__ void __launch_bounds__(256)
lpm_best_kernel(const int8_t* __restrict__ codes, const long long* __restrict__ offsets, int R, int w,
                const double* __restrict__ lpm, const double* __restrict__ bg, int bg_per_record, int fwd_only,
                double* __restrict__ best_out, int32_t* __restrict__ pos_out, int8_t* __restrict__ strand_out,
                int32_t* __restrict__ nties_out, int32_t* __restrict__ tie_pos, int tie_cap) {
    __shared__ LpmTab T;
    __shared__ double s_val[8];
    __shared__ long long s_idx[8];
    __shared__ int s_cnt[8];
    __shared__ double s_best;
    __shared__ long long s_first;
    for (int i = threadIdx.x; i < 5 * w; i += blockDim.x) T.l[i / w][i % w] = lpm[i];
    if (threadIdx.x < 5) T.bg[threadIdx.x] = bg[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int r = blockIdx.x; r < R; r += gridDim.x) {
        if (bg_per_record) {
            if (threadIdx.x < 5) T.bg[threadIdx.x] = bg[(long long)r * 5 + threadIdx.x];
            __syncthreads();
        }
        const long long beg = offsets[r], len = offsets[r + 1] - beg;
        const long long nwin = len - w + 1;
        // pass A
        double bv = 0.0;                 // the reference starts from best_score = 0.0 and needs max > best
        long long bi = LLONG_MAX;
        for (long long i = threadIdx.x; i < nwin; i += blockDim.x) {
            double p, pr;
            lpm_window(codes + beg + i, w, T, fwd_only != 0, p, pr);
            const double m = fwd_only ? p : (p > pr ? p : pr);
            if (m > bv || (m == bv && m != 0.0 && i < bi)) { bv = m; bi = i; }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, bv, off);
            const long long oi = __shfl_xor_sync(0xffffffffu, bi, off);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { s_val[warp] = bv; s_idx[warp] = bi; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double v = s_val[0]; long long ix = s_idx[0];
            for (int i = 1; i < 8; ++i)
                if (s_val[i] > v || (s_val[i] == v && s_idx[i] < ix)) { v = s_val[i]; ix = s_idx[i]; }
            s_best = v; s_first = ix;
        }
        __syncthreads();
        const double best = s_best;
        const long long first = s_first;
        if (best == 0.0 || first == LLONG_MAX) {
            // nothing beat 0.0: the reference prints position "0", strand '+'
            if (threadIdx.x == 0) { best_out[r] = 0.0; pos_out[r] = 0; strand_out[r] = '+'; nties_out[r] = 0; }
            __syncthreads();
            continue;
        }
        // pass B: later windows whose max equals best, in ascending order
        int total = 0;
        for (long long i0 = first; i0 < nwin; i0 += blockDim.x) {
            const long long i = i0 + threadIdx.x;
            bool hit = false; int enc = 0; bool plus = true;
            if (i < nwin) {
                double p, pr;
                lpm_window(codes + beg + i, w, T, fwd_only != 0, p, pr);
                const double m = fwd_only ? p : (p > pr ? p : pr);
                plus = fwd_only ? true : (m == p);
                enc = plus ? (int)i : (int)(i + w);
                hit = (m == best);
                if (i == first) {
                    hit = false;
                    best_out[r] = best; pos_out[r] = enc; strand_out[r] = plus ? '+' : '-';
                }
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, hit);
            if (lane == 0) s_cnt[warp] = __popc(bal);
            __syncthreads();
            int base = total;
            for (int i2 = 0; i2 < warp; ++i2) base += s_cnt[i2];
            int tile_total = 0;
            for (int i2 = 0; i2 < 8; ++i2) tile_total += s_cnt[i2];
            if (hit) {
                const int my = base + __popc(bal & ((1u << lane) - 1u));
                if (tie_pos && my < tie_cap) tie_pos[(long long)r * tie_cap + my] = enc;
            }
            total += tile_total;
            __syncthreads();
        }
        if (threadIdx.x == 0) nties_out[r] = total;
        __syncthreads();
    }
}

static int scan_grid(int R) {
    int g = num_sms() * 8;
    if (g > R) g = R;
    if (g < 1) g = 1;
    return g;
}

extern "C" int xnn_lpm_scan_sum(const int8_t* codes, const long long* offsets, int R, int w, const double* lpm,
                                const double* bg, int bg_per_record, int forward_only, double* out_sum,
                                xnn_stream_t s) {
    XNN_CHECK_ARG(codes && offsets && lpm && bg && out_sum, "NULL pointer");
    XNN_CHECK_ARG(R >= 0 && w >= 1 && w <= LPM_WMAX, "bad sizes (matrix length <= 64)");
    if (R == 0) return XNN_OK;
    lpm_sum_kernel<<<scan_grid(R), 256, 0, to_stream(s)>>>(codes, offsets, R, w, lpm, bg, bg_per_record, forward_only, out_sum);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

extern "C" int xnn_lpm_scan_best(const int8_t* codes, const long long* offsets, int R, int w, const double* lpm,
                                 const double* bg, int bg_per_record, int forward_only, double* best,
                                 int32_t* best_pos,
                                 int8_t* best_strand, int32_t* n_ties, int32_t* tie_pos, int tie_cap,
                                 xnn_stream_t s) {
    XNN_CHECK_ARG(codes && offsets && lpm && bg && best && best_pos && best_strand && n_ties, "NULL pointer");
    XNN_CHECK_ARG(R >= 0 && w >= 1 && w <= LPM_WMAX && tie_cap >= 0, "bad sizes (matrix length <= 64)");
    if (R == 0) return XNN_OK;
    lpm_best_kernel<<<scan_grid(R), 256, 0, to_stream(s)>>>(codes, offsets, R, w, lpm, bg, bg_per_record, forward_only, best,
                                                            best_pos, best_strand, n_ties, tie_pos, tie_cap);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// ---------------------------------------------------------------------------
// integer PWM mode: best hit, first maximum wins, '+' wins strand ties
// (the reference's branch-free max: k = sign bit of score - rev_score).
// Arithmetic is done in uint32 so that INT_MIN entries wrap like the C tool.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
ipwm_best_kernel(const int8_t* __restrict__ codes, const long long* __restrict__ offsets, int R, int w,
                 const int32_t* __restrict__ pwm, int fwd_only, int32_t* __restrict__ best_out,
                 int32_t* __restrict__ pos_out, int8_t* __restrict__ strand_out) {
    __shared__ int32_t Tm[5][LPM_WMAX];
    __shared__ int s_val[8];
    __shared__ long long s_idx[8];
    __shared__ int s_str[8];
    for (int i = threadIdx.x; i < 5 * w; i += blockDim.x) Tm[i / w][i % w] = pwm[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int r = blockIdx.x; r < R; r += gridDim.x) {
        const long long beg = offsets[r], len = offsets[r + 1] - beg;
        const long long nwin = len - w + 1;
        int bv = INT_MIN, bs = 0;
        long long bi = LLONG_MAX;
        for (long long i = threadIdx.x; i < nwin; i += blockDim.x) {
            const int8_t* sq = codes + beg + i;
            uint32_t sc = 0, rv = 0;
            for (int j = 0; j < w; ++j) {
                const int c = sq[j];
                sc += (uint32_t)Tm[c][j];
                if (!fwd_only) rv += (uint32_t)Tm[(c == 4) ? 4 : 3 - c][w - j - 1];
            }
            int m, kk = 0;
            if (fwd_only) {
                m = (int)sc;
            } else {
                const uint32_t cdiff = sc - rv;
                kk = (int)(cdiff >> 31);
                m = (int)(sc - (uint32_t)kk * cdiff);
            }
            // strict '>' against INT_MIN start: a window scoring exactly INT_MIN never becomes best
            if (m > bv || (m == bv && bi != LLONG_MAX && i < bi)) { bv = m; bi = i; bs = kk; }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const int ov = __shfl_xor_sync(0xffffffffu, bv, off);
            const long long oi = __shfl_xor_sync(0xffffffffu, bi, off);
            const int os = __shfl_xor_sync(0xffffffffu, bs, off);
            if (oi != LLONG_MAX && (ov > bv || bi == LLONG_MAX || (ov == bv && oi < bi))) { bv = ov; bi = oi; bs = os; }
        }
        if (lane == 0) { s_val[warp] = bv; s_idx[warp] = bi; s_str[warp] = bs; }
        __syncthreads();
        if (threadIdx.x == 0) {
            int v = INT_MIN, st = 0; long long ix = LLONG_MAX;
            for (int i = 0; i < 8; ++i)
                if (s_idx[i] != LLONG_MAX && (ix == LLONG_MAX || s_val[i] > v || (s_val[i] == v && s_idx[i] < ix))) {
                    v = s_val[i]; ix = s_idx[i]; st = s_str[i];
                }
            if (nwin <= 0) { best_out[r] = INT_MIN; pos_out[r] = -1; strand_out[r] = 0; }        // "NOTAG"
            else if (ix == LLONG_MAX) { best_out[r] = INT_MIN; pos_out[r] = 0; strand_out[r] = 0; }  // nothing > INT_MIN
            else { best_out[r] = v; pos_out[r] = (int32_t)ix; strand_out[r] = (int8_t)st; }
        }
        __syncthreads();
    }
}

extern "C" int xnn_ipwm_scan_best(const int8_t* codes, const long long* offsets, int R, int w, const int32_t* pwm,
                                  int forward_only, int32_t* best, int32_t* pos, int8_t* strand, xnn_stream_t s) {
    XNN_CHECK_ARG(codes && offsets && pwm && best && pos && strand, "NULL pointer");
    XNN_CHECK_ARG(R >= 0 && w >= 1 && w <= LPM_WMAX, "bad sizes (matrix length <= 64)");
    if (R == 0) return XNN_OK;
    ipwm_best_kernel<<<scan_grid(R), 256, 0, to_stream(s)>>>(codes, offsets, R, w, pwm, forward_only, best, pos,
                                                             strand);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// ---------------------------------------------------------------------------
// Long records (a chromosome is ONE FASTA record): the kernels above give a record to one block.  The
// `_parts` entry points split every record's windows into P contiguous parts (P depends on R only, so the
// summation order is fixed), one block per (record, part), partial results in a caller-provided work buffer,
// and a second kernel combines the parts in ascending order: sums are deterministic, the integer best hit keeps
// the tool's first-maximum rule.
// ---------------------------------------------------------------------------
static int scan_parts(int R) {
    const int want = num_sms() * 8;
    if (R >= want) return 1;
    int P = (want + R - 1) / R;
    return P > 1024 ? 1024 : P;
}
extern "C" int xnn_scan_parts(int R) { return R < 1 ? 1 : scan_parts(R); }
extern "C" size_t xnn_scan_work_bytes(int R) { return R < 1 ? 0 : (size_t)R * scan_parts(R) * 24; }

__global__ void __launch_bounds__(256)
lpm_sum_parts_kernel(const int8_t* __restrict__ codes, const long long* __restrict__ offsets, int R, int w, int P,
                     const double* __restrict__ lpm, const double* __restrict__ bg, int bg_per_record, int fwd_only,
                     double* __restrict__ part_sum) {
    __shared__ LpmTab T;
    __shared__ double red[8];
    for (int i = threadIdx.x; i < 5 * w; i += blockDim.x) T.l[i / w][i % w] = lpm[i];
    if (threadIdx.x < 5) T.bg[threadIdx.x] = bg[threadIdx.x];
    __syncthreads();
    const int part = blockIdx.y;
    for (int r = blockIdx.x; r < R; r += gridDim.x) {
        if (bg_per_record) {
            if (threadIdx.x < 5) T.bg[threadIdx.x] = bg[(long long)r * 5 + threadIdx.x];
            __syncthreads();
        }
        const long long beg = offsets[r], len = offsets[r + 1] - beg;
        const long long nwin = len - w + 1;
        const long long per = nwin > 0 ? (nwin + P - 1) / P : 0;
        const long long i0 = (long long)part * per, i1 = min(nwin, i0 + per);
        double acc = 0.0;
        for (long long i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
            double p, pr;
            lpm_window(codes + beg + i, w, T, fwd_only != 0, p, pr);
            acc += fwd_only ? p : (p + pr);
        }
        acc = warp_sum(acc);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int i = 0; i < 8; ++i) t += red[i];
            part_sum[(long long)r * P + part] = t;
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256)
lpm_sum_combine_kernel(const double* __restrict__ part_sum, int R, int P, double* __restrict__ out_sum) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    double t = 0.0;
    for (int p = 0; p < P; ++p) t += part_sum[(long long)r * P + p];
    out_sum[r] = t;
}

extern "C" int xnn_lpm_scan_sum_parts(const int8_t* codes, const long long* offsets, int R, int w, const double* lpm,
                                      const double* bg, int bg_per_record, int forward_only, double* out_sum,
                                      void* work, xnn_stream_t s) {
    XNN_CHECK_ARG(codes && offsets && lpm && bg && out_sum && work, "NULL pointer");
    XNN_CHECK_ARG(R >= 0 && w >= 1 && w <= LPM_WMAX, "bad sizes (matrix length <= 64)");
    if (R == 0) return XNN_OK;
    const int P = scan_parts(R);
    cudaStream_t st = to_stream(s);
    dim3 grid(R < 65535 ? R : 65535, P);
    if ((long long)grid.x * P > (long long)num_sms() * 64) grid.x = (unsigned)max(1, num_sms() * 64 / P);
    lpm_sum_parts_kernel<<<grid, 256, 0, st>>>(codes, offsets, R, w, P, lpm, bg, bg_per_record, forward_only,
                                               (double*)work);
    XNN_CHECK_LAUNCH();
    lpm_sum_combine_kernel<<<ceil_div(R, 256), 256, 0, st>>>((const double*)work, R, P, out_sum);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

struct IpwmPart { int v; int st; long long ix; };

__global__ void __launch_bounds__(256)
ipwm_best_parts_kernel(const int8_t* __restrict__ codes, const long long* __restrict__ offsets, int R, int w, int P,
                       const int32_t* __restrict__ pwm, int fwd_only, IpwmPart* __restrict__ parts) {
    __shared__ int32_t Tm[5][LPM_WMAX];
    __shared__ int s_val[8];
    __shared__ long long s_idx[8];
    __shared__ int s_str[8];
    for (int i = threadIdx.x; i < 5 * w; i += blockDim.x) Tm[i / w][i % w] = pwm[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, part = blockIdx.y;
    for (int r = blockIdx.x; r < R; r += gridDim.x) {
        const long long beg = offsets[r], len = offsets[r + 1] - beg;
        const long long nwin = len - w + 1;
        const long long per = nwin > 0 ? (nwin + P - 1) / P : 0;
        const long long i0 = (long long)part * per, i1 = min(nwin, i0 + per);
        int bv = INT_MIN, bs = 0;
        long long bi = LLONG_MAX;
        for (long long i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
            const int8_t* sq = codes + beg + i;
            uint32_t sc = 0, rv = 0;
            for (int j = 0; j < w; ++j) {
                const int c = sq[j];
                sc += (uint32_t)Tm[c][j];
                if (!fwd_only) rv += (uint32_t)Tm[(c == 4) ? 4 : 3 - c][w - j - 1];
            }
            int m, kk = 0;
            if (fwd_only) {
                m = (int)sc;
            } else {
                const uint32_t cdiff = sc - rv;
                kk = (int)(cdiff >> 31);
                m = (int)(sc - (uint32_t)kk * cdiff);
            }
            if (m > bv || (m == bv && bi != LLONG_MAX && i < bi)) { bv = m; bi = i; bs = kk; }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const int ov = __shfl_xor_sync(0xffffffffu, bv, off);
            const long long oi = __shfl_xor_sync(0xffffffffu, bi, off);
            const int os = __shfl_xor_sync(0xffffffffu, bs, off);
            if (oi != LLONG_MAX && (ov > bv || bi == LLONG_MAX || (ov == bv && oi < bi))) { bv = ov; bi = oi; bs = os; }
        }
        if (lane == 0) { s_val[warp] = bv; s_idx[warp] = bi; s_str[warp] = bs; }
        __syncthreads();
        if (threadIdx.x == 0) {
            int v = INT_MIN, st = 0; long long ix = LLONG_MAX;
            for (int i = 0; i < 8; ++i)
                if (s_idx[i] != LLONG_MAX && (ix == LLONG_MAX || s_val[i] > v || (s_val[i] == v && s_idx[i] < ix))) {
                    v = s_val[i]; ix = s_idx[i]; st = s_str[i];
                }
            IpwmPart o; o.v = v; o.st = st; o.ix = ix;
            parts[(long long)r * P + part] = o;
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256)
ipwm_best_combine_kernel(const IpwmPart* __restrict__ parts, const long long* __restrict__ offsets, int R, int w, int P,
                         int32_t* __restrict__ best_out, int32_t* __restrict__ pos_out, int8_t* __restrict__ strand_out) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    const long long nwin = offsets[r + 1] - offsets[r] - w + 1;
    int v = INT_MIN, st = 0; long long ix = LLONG_MAX;
    for (int p = 0; p < P; ++p) {
        const IpwmPart q = parts[(long long)r * P + p];
        if (q.ix != LLONG_MAX && (ix == LLONG_MAX || q.v > v || (q.v == v && q.ix < ix))) { v = q.v; ix = q.ix; st = q.st; }
    }
    if (nwin <= 0) { best_out[r] = INT_MIN; pos_out[r] = -1; strand_out[r] = 0; }        // "NOTAG"
    else if (ix == LLONG_MAX) { best_out[r] = INT_MIN; pos_out[r] = 0; strand_out[r] = 0; }  // nothing > INT_MIN
    else { best_out[r] = v; pos_out[r] = (int32_t)ix; strand_out[r] = (int8_t)st; }
}

extern "C" int xnn_ipwm_scan_best_parts(const int8_t* codes, const long long* offsets, int R, int w,
                                        const int32_t* pwm, int forward_only, int32_t* best, int32_t* pos,
                                        int8_t* strand, void* work, xnn_stream_t s) {
    XNN_CHECK_ARG(codes && offsets && pwm && best && pos && strand && work, "NULL pointer");
    XNN_CHECK_ARG(R >= 0 && w >= 1 && w <= LPM_WMAX, "bad sizes (matrix length <= 64)");
    XNN_CHECK_ARG(((uintptr_t)work & 7) == 0, "work must be 8-byte aligned");
    if (R == 0) return XNN_OK;
    const int P = scan_parts(R);
    cudaStream_t st = to_stream(s);
    dim3 grid(R < 65535 ? R : 65535, P);
    if ((long long)grid.x * P > (long long)num_sms() * 64) grid.x = (unsigned)max(1, num_sms() * 64 / P);
    ipwm_best_parts_kernel<<<grid, 256, 0, st>>>(codes, offsets, R, w, P, pwm, forward_only, (IpwmPart*)work);
    XNN_CHECK_LAUNCH();
    ipwm_best_combine_kernel<<<ceil_div(R, 256), 256, 0, st>>>((const IpwmPart*)work, offsets, R, w, P, best, pos, strand);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// ---- LPM best-hit mode over record parts --------------------------------------------------------------
// A: per (record, part) best value and first window;  B: per record best / first (kept in the record's part 0);
// C: per part, the number of later windows equal to the best (and the hit itself: position, strand);
// D: per record, exclusive prefix of those counts;  E: per part, the tied positions in ascending order.
struct LpmPart { double bv; long long bi; int ties; int off; };

__device__ __forceinline__ void lpm_load_tab(LpmTab& T, const double* lpm, const double* bg, int w) {
    for (int i = threadIdx.x; i < 5 * w; i += blockDim.x) T.l[i / w][i % w] = lpm[i];
    if (threadIdx.x < 5) T.bg[threadIdx.x] = bg[threadIdx.x];
    __syncthreads();
}

__global__ void __launch_bounds__(256)
lpm_best_partsA_kernel(const int8_t* __restrict__ codes, const long long* __restrict__ offsets, int R, int w, int P,
                       const double* __restrict__ lpm, const double* __restrict__ bg, int bg_per_record, int fwd_only,
                       LpmPart* __restrict__ parts) {
    __shared__ LpmTab T;
    __shared__ double s_val[8];
    __shared__ long long s_idx[8];
    lpm_load_tab(T, lpm, bg, w);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, part = blockIdx.y;
    for (int r = blockIdx.x; r < R; r += gridDim.x) {
        if (bg_per_record) {
            if (threadIdx.x < 5) T.bg[threadIdx.x] = bg[(long long)r * 5 + threadIdx.x];
            __syncthreads();
        }
        const long long beg = offsets[r], nwin = offsets[r + 1] - beg - w + 1;
        const long long per = nwin > 0 ? (nwin + P - 1) / P : 0;
        const long long i0 = (long long)part * per, i1 = min(nwin, i0 + per);
        double bv = 0.0;                 // the reference starts from best_score = 0.0 and needs max > best
        long long bi = LLONG_MAX;
        for (long long i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
            double p, pr;
            lpm_window(codes + beg + i, w, T, fwd_only != 0, p, pr);
            const double m = fwd_only ? p : (p > pr ? p : pr);
            if (m > bv || (m == bv && m != 0.0 && i < bi)) { bv = m; bi = i; }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, bv, off);
            const long long oi = __shfl_xor_sync(0xffffffffu, bi, off);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { s_val[warp] = bv; s_idx[warp] = bi; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double v = s_val[0]; long long ix = s_idx[0];
            for (int i = 1; i < 8; ++i)
                if (s_val[i] > v || (s_val[i] == v && s_idx[i] < ix)) { v = s_val[i]; ix = s_idx[i]; }
            LpmPart o; o.bv = v; o.bi = ix; o.ties = 0; o.off = 0;
            parts[(long long)r * P + part] = o;
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256)
lpm_best_partsB_kernel(LpmPart* __restrict__ parts, int R, int P, double* __restrict__ best_out,
                       int32_t* __restrict__ pos_out, int8_t* __restrict__ strand_out, int32_t* __restrict__ nties_out) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    double v = 0.0; long long ix = LLONG_MAX;
    for (int p = 0; p < P; ++p) {
        const LpmPart q = parts[(long long)r * P + p];
        if (q.bv > v || (q.bv == v && q.bi < ix)) { v = q.bv; ix = q.bi; }
    }
    parts[(long long)r * P].bv = v;
    parts[(long long)r * P].bi = ix;
    if (v == 0.0 || ix == LLONG_MAX) {
        // nothing beat 0.0: the reference prints position "0", strand '+'
        best_out[r] = 0.0; pos_out[r] = 0; strand_out[r] = '+'; nties_out[r] = 0;
        parts[(long long)r * P].bi = LLONG_MAX;
    }
}

// MODE 0: count ties per part (and write the hit); MODE 1: list them at the part's offset
template <int MODE>
__global__ void __launch_bounds__(256)
lpm_best_partsC_kernel(const int8_t* __restrict__ codes, const long long* __restrict__ offsets, int R, int w, int P,
                       const double* __restrict__ lpm, const double* __restrict__ bg, int bg_per_record, int fwd_only,
                       LpmPart* __restrict__ parts, double* __restrict__ best_out, int32_t* __restrict__ pos_out,
                       int8_t* __restrict__ strand_out, int32_t* __restrict__ tie_pos, int tie_cap) {
    __shared__ LpmTab T;
    __shared__ int s_cnt[8];
    lpm_load_tab(T, lpm, bg, w);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, part = blockIdx.y;
    for (int r = blockIdx.x; r < R; r += gridDim.x) {
        const double best = parts[(long long)r * P].bv;
        const long long first = parts[(long long)r * P].bi;
        if (first == LLONG_MAX) continue;
        LpmPart* me = parts + (long long)r * P + part;
        if (MODE == 1 && me->ties == 0) continue;
        if (bg_per_record) {
            __syncthreads();
            if (threadIdx.x < 5) T.bg[threadIdx.x] = bg[(long long)r * 5 + threadIdx.x];
            __syncthreads();
        }
        const long long beg = offsets[r], nwin = offsets[r + 1] - beg - w + 1;
        const long long per = nwin > 0 ? (nwin + P - 1) / P : 0;
        const long long i0 = max((long long)part * per, first), i1 = min(nwin, (long long)part * per + per);
        int total = 0;
        const int base0 = MODE == 1 ? me->off : 0;
        for (long long ib = i0; ib < i1; ib += blockDim.x) {
            const long long i = ib + threadIdx.x;
            bool hit = false; int enc = 0;
            if (i < i1) {
                double p, pr;
                lpm_window(codes + beg + i, w, T, fwd_only != 0, p, pr);
                const double m = fwd_only ? p : (p > pr ? p : pr);
                const bool plus = fwd_only ? true : (m == p);
                enc = plus ? (int)i : (int)(i + w);
                hit = (m == best);
                if (i == first) {
                    hit = false;
                    if (MODE == 0) { best_out[r] = best; pos_out[r] = enc; strand_out[r] = plus ? '+' : '-'; }
                }
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, hit);
            __syncthreads();
            if (lane == 0) s_cnt[warp] = __popc(bal);
            __syncthreads();
            int before = 0, tile_total = 0;
            for (int i2 = 0; i2 < 8; ++i2) { if (i2 < warp) before += s_cnt[i2]; tile_total += s_cnt[i2]; }
            if (MODE == 1 && hit) {
                const int my = base0 + total + before + __popc(bal & ((1u << lane) - 1u));
                if (my < tie_cap) tie_pos[(long long)r * tie_cap + my] = enc;
            }
            total += tile_total;
        }
        if (MODE == 0 && threadIdx.x == 0) me->ties = total;
    }
}

__global__ void __launch_bounds__(256)
lpm_best_partsD_kernel(LpmPart* __restrict__ parts, int R, int P, int32_t* __restrict__ nties_out) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R || parts[(long long)r * P].bi == LLONG_MAX) return;
    int run = 0;
    for (int p = 0; p < P; ++p) {
        LpmPart* q = parts + (long long)r * P + p;
        q->off = run;
        run += q->ties;
    }
    nties_out[r] = run;
}

extern "C" int xnn_lpm_scan_best_parts(const int8_t* codes, const long long* offsets, int R, int w, const double* lpm,
                                       const double* bg, int bg_per_record, int forward_only, double* best,
                                       int32_t* best_pos, int8_t* best_strand, int32_t* n_ties, int32_t* tie_pos,
                                       int tie_cap, void* work, xnn_stream_t s) {
    XNN_CHECK_ARG(codes && offsets && lpm && bg && best && best_pos && best_strand && n_ties && work, "NULL pointer");
    XNN_CHECK_ARG(R >= 0 && w >= 1 && w <= LPM_WMAX && tie_cap >= 0, "bad sizes (matrix length <= 64)");
    XNN_CHECK_ARG(((uintptr_t)work & 7) == 0, "work must be 8-byte aligned");
    if (R == 0) return XNN_OK;
    const int P = scan_parts(R);
    cudaStream_t st = to_stream(s);
    dim3 grid(R < 65535 ? R : 65535, P);
    if ((long long)grid.x * P > (long long)num_sms() * 64) grid.x = (unsigned)max(1, num_sms() * 64 / P);
    LpmPart* parts = (LpmPart*)work;
    lpm_best_partsA_kernel<<<grid, 256, 0, st>>>(codes, offsets, R, w, P, lpm, bg, bg_per_record, forward_only, parts);
    XNN_CHECK_LAUNCH();
    lpm_best_partsB_kernel<<<ceil_div(R, 256), 256, 0, st>>>(parts, R, P, best, best_pos, best_strand, n_ties);
    XNN_CHECK_LAUNCH();
    lpm_best_partsC_kernel<0><<<grid, 256, 0, st>>>(codes, offsets, R, w, P, lpm, bg, bg_per_record, forward_only,
                                                    parts, best, best_pos, best_strand, tie_pos, tie_cap);
    XNN_CHECK_LAUNCH();
    lpm_best_partsD_kernel<<<ceil_div(R, 256), 256, 0, st>>>(parts, R, P, n_ties);
    XNN_CHECK_LAUNCH();
    if (tie_pos && tie_cap > 0) {
        lpm_best_partsC_kernel<1><<<grid, 256, 0, st>>>(codes, offsets, R, w, P, lpm, bg, bg_per_record, forward_only,
                                                        parts, best, best_pos, best_strand, tie_pos, tie_cap);
        XNN_CHECK_LAUNCH();
    }
    return XNN_OK;
}

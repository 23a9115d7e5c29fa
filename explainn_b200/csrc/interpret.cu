// a7-a10 -- the interpret path without the (N,U,Lc) fp16 activation array:
//   test.py:128-166      activations = exp(BN1(conv(x))) stored as np.float16
//   interpret.py:363-373 threshold_u = 0.5 * max activation over the selected sequences
//   interpret.py:375-421 sites = (sequence, start) with activation > threshold, in
//                        (listed sequence, ascending start) order, capped at 1e6 per unit
//   interpret.py:423-454 PFM = per-column A/C/G/T counts of the emitted k-mers
// Every pass recomputes the activation from the packed input and rounds it to
// fp16 (round-to-nearest-even) before any comparison, exactly as the reference
// compares its stored fp16 values.  lanes = positions; a warp owns one unit.
#include "common.cuh"
using namespace xnn;

#define XNN_KMAX 32
#define IU 32  // units per block

__device__ __forceinline__ void load_lut(float* lut, const float* __restrict__ conv_w, int u0, int nu, int k) {
    const int per = 5 * k;  // lut[ul][c][j], row c = 4 (N) all zero
    for (int i = threadIdx.x; i < nu * per; i += blockDim.x) {
        const int ul = i / per, r = i - ul * per, c = r / k, j = r - c * k;
        lut[i] = (c < 4) ? conv_w[((long long)(u0 + ul) * 4 + c) * k + j] : 0.0f;
    }
}

// fp16-rounded activation of (unit, position) as a float
__device__ __forceinline__ float act16(const float* __restrict__ lu, const uint8_t* __restrict__ codes, int p, int k,
                                       float sc, float sh) {
    float z = 0.0f;
    for (int j = 0; j < k; ++j) z += lu[(int)codes[p + j] * k + j];
    return __half2float(__float2half_rn(expf(fmaf(sc, z, sh))));
}

// mode 0: per-unit max (amax_bits); mode 1: counts[u][i]; mode 2: emit
template <int MODE>
__global__ void __launch_bounds__(256)
interpret_kernel(const uint32_t* __restrict__ planes, int B, int L, int Wp, int U, int k,
                 const float* __restrict__ conv_w, const float* __restrict__ scale, const float* __restrict__ shift,
                 const float* __restrict__ thr, const int32_t* __restrict__ idx, int n_idx,
                 uint32_t* __restrict__ amax_bits, int32_t* __restrict__ counts,
                 const long long* __restrict__ offsets, long long cap, int32_t* __restrict__ pfm,
                 int32_t* __restrict__ sites, long long site_cap) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* lut = reinterpret_cast<float*>(smem_raw);                 // [IU][5][k]
    int* s_pfm = reinterpret_cast<int*>(lut + IU * 5 * k);           // [IU][4][k] (mode 2)
    uint8_t* codes = reinterpret_cast<uint8_t*>(s_pfm + (MODE == 2 ? IU * 4 * k : 0));
    const int Lc = L - k + 1, Lpad = (L + 15) & ~15;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int u0 = blockIdx.x * IU, nu = min(IU, U - u0);
    load_lut(lut, conv_w, u0, nu, k);
    if (MODE == 2)
        for (int i = threadIdx.x; i < IU * 4 * k; i += blockDim.x) s_pfm[i] = 0;
    float wmax[IU / 8];  // per-warp running max for the units this warp owns (mode 0)
#pragma unroll
    for (int i = 0; i < IU / 8; ++i) wmax[i] = 0.0f;

    for (int ii = blockIdx.y; ii < n_idx; ii += gridDim.y) {
        const int b = idx ? idx[ii] : ii;
        __syncthreads();
        expand_codes(planes + (long long)b * 3 * Wp, Wp, L, Lpad, codes);
        __syncthreads();
        int slot = 0;
        for (int ul = warp; ul < nu; ul += nwarps, ++slot) {
            const int u = u0 + ul;
            const float sc = scale[u], sh = shift[u];
            const float* lu = lut + ul * 5 * k;
            if (MODE == 0) {
                float m = 0.0f;
                for (int p = lane; p < Lc; p += 32) m = fmaxf(m, act16(lu, codes, p, k, sc, sh));
                wmax[slot] = fmaxf(wmax[slot], m);
            } else {
                const float t = thr[u];
                long long rank = (MODE == 2) ? offsets[(long long)u * n_idx + ii] : 0;
                int cnt = 0;
                for (int p0 = 0; p0 < Lc; p0 += 32) {
                    const int p = p0 + lane;
                    const bool hit = (p < Lc) && (act16(lu, codes, p, k, sc, sh) > t);
                    const uint32_t bal = __ballot_sync(0xffffffffu, hit);
                    if (MODE == 1) {
                        cnt += __popc(bal);
                    } else if (bal) {
                        const long long my = rank + __popc(bal & ((1u << lane) - 1u));
                        if (hit && my < cap) {
                            if (sites && my < site_cap) {
                                sites[((long long)u * site_cap + my) * 2 + 0] = ii;
                                sites[((long long)u * site_cap + my) * 2 + 1] = p;
                            }
                            for (int j = 0; j < k; ++j) {
                                const int c = codes[p + j];
                                if (c < 4) atomicAdd(&s_pfm[(ul * 4 + c) * k + j], 1);
                            }
                        }
                        rank += __popc(bal);
                    }
                }
                if (MODE == 1 && lane == 0) counts[(long long)u * n_idx + ii] = cnt;
            }
        }
    }
    if (MODE == 0) {
        int slot = 0;
        for (int ul = warp; ul < nu; ul += nwarps, ++slot) {
            const float m = warp_max(wmax[slot]);
            // activations are >= 0 (or +inf / NaN): uint order == float order for non-negative floats
            if (lane == 0) atomicMax(&amax_bits[u0 + ul], __float_as_uint(m));
        }
    }
    if (MODE == 2) {
        __syncthreads();
        for (int i = threadIdx.x; i < nu * 4 * k; i += blockDim.x)
            if (s_pfm[i]) atomicAdd(&pfm[(long long)u0 * 4 * k + i], s_pfm[i]);
    }
}

template <int MODE>
static int launch_interpret(const uint32_t* planes, int B, int L, int U, int k, const float* conv_w,
                            const float* scale, const float* shift, const float* thr, const int32_t* idx, int n_idx,
                            uint32_t* amax_bits, int32_t* counts, const long long* offsets, long long cap,
                            int32_t* pfm, int32_t* sites, long long site_cap, cudaStream_t st) {
    const size_t smem = (size_t)IU * 5 * k * 4 + (MODE == 2 ? (size_t)IU * 4 * k * 4 : 0) + ((L + 15) & ~15) + 16;
    if (smem > 48 * 1024) return -1;
    const int chunks = ceil_div(U, IU);
    int gy = ceil_div((long long)num_sms() * 6, chunks);
    if (gy > n_idx) gy = n_idx;
    if (gy > 65535) gy = 65535;
    if (gy < 1) gy = 1;
    dim3 grid(chunks, gy);
    interpret_kernel<MODE><<<grid, 256, smem, st>>>(planes, B, L, planes_words(L), U, k, conv_w, scale, shift, thr,
                                                    idx, n_idx, amax_bits, counts, offsets, cap, pfm, sites,
                                                    site_cap);
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

#define INTERP_COMMON_CHECKS()                                                                          \
    XNN_CHECK_ARG(planes && conv_w && scale && shift, "NULL pointer");                                  \
    XNN_CHECK_ARG(k >= 1 && k <= XNN_KMAX && L >= k && U >= 1 && B >= 0 && n_idx >= 0, "bad sizes");    \
    XNN_CHECK_ARG(idx != nullptr || n_idx == B, "idx == NULL means all B sequences (n_idx must equal B)"); \
    if (n_idx == 0) return XNN_OK;

extern "C" int xnn_act_max(const uint32_t* planes, int B, int L, int U, int k, const float* conv_w,
                           const float* scale, const float* shift, const int32_t* idx, int n_idx,
                           uint32_t* amax_bits, xnn_stream_t s) {
    INTERP_COMMON_CHECKS();
    XNN_CHECK_ARG(amax_bits, "NULL pointer");
    const int r = launch_interpret<0>(planes, B, L, U, k, conv_w, scale, shift, nullptr, idx, n_idx, amax_bits,
                                      nullptr, nullptr, 0, nullptr, nullptr, 0, to_stream(s));
    if (r) { set_error("xnn_act_max: %s", r < 0 ? "sequence too long" : "launch failed"); return r < 0 ? XNN_ERR_ARG : XNN_ERR_CUDA; }
    return XNN_OK;
}

extern "C" int xnn_site_counts(const uint32_t* planes, int B, int L, int U, int k, const float* conv_w,
                               const float* scale, const float* shift, const float* thr, const int32_t* idx,
                               int n_idx, int32_t* counts, xnn_stream_t s) {
    INTERP_COMMON_CHECKS();
    XNN_CHECK_ARG(thr && counts, "NULL pointer");
    const int r = launch_interpret<1>(planes, B, L, U, k, conv_w, scale, shift, thr, idx, n_idx, nullptr, counts,
                                      nullptr, 0, nullptr, nullptr, 0, to_stream(s));
    if (r) { set_error("xnn_site_counts: %s", r < 0 ? "sequence too long" : "launch failed"); return r < 0 ? XNN_ERR_ARG : XNN_ERR_CUDA; }
    return XNN_OK;
}

extern "C" int xnn_site_emit(const uint32_t* planes, int B, int L, int U, int k, const float* conv_w,
                             const float* scale, const float* shift, const float* thr, const int32_t* idx,
                             int n_idx, const long long* offsets, long long cap, int32_t* pfm, int32_t* sites,
                             long long site_cap, xnn_stream_t s) {
    INTERP_COMMON_CHECKS();
    XNN_CHECK_ARG(thr && offsets && pfm, "NULL pointer");
    XNN_CHECK_ARG(cap >= 0 && (sites == nullptr || site_cap >= 0), "bad cap");
    const int r = launch_interpret<2>(planes, B, L, U, k, conv_w, scale, shift, thr, idx, n_idx, nullptr, nullptr,
                                      offsets, cap, pfm, sites, site_cap, to_stream(s));
    if (r) { set_error("xnn_site_emit: %s", r < 0 ? "sequence too long" : "launch failed"); return r < 0 ? XNN_ERR_ARG : XNN_ERR_CUDA; }
    return XNN_OK;
}


// ---------------------------------------------------------------------------
// Between passes 2a and 2b (interpret.py:414-417: the first 1e6 sites of a unit, in sequence order): the rank of
// the first site of (unit, listed sequence) = carry[u] + the sites of the unit at earlier listed sequences.
// One block per unit: exclusive scan of the row of int32 counts into int64 offsets, carry[u] += the row total.
// (Was five ATen launches per chunk -- int64 copy, cumsum, subtract, add, slice-add -- over 8-byte temporaries.)
// ---------------------------------------------------------------------------
#define OFS_THREADS 1024
#define OFS_PER 8                                          // counts per thread and tile
__global__ void __launch_bounds__(OFS_THREADS)
site_offsets_kernel(const int32_t* __restrict__ counts, int n, long long* __restrict__ carry,
                    long long* __restrict__ offsets) {
    __shared__ long long s_warp[2][OFS_THREADS / 32];
    const int u = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int32_t* row = counts + (long long)u * n;
    long long* out = offsets + (long long)u * n;
    const bool vec = ((n & 3) == 0);                       // rows 16-byte aligned: vector loads / stores
    long long run = carry[u];                              // rank of the first site of the current tile
    // tiles of 8 x 1024 consecutive counts, eight per thread, scanned through shuffles; the next tile's counts are
    // loaded before this tile's scan (one barrier per tile: the warp totals are double buffered)
    int c[OFS_PER], nx[OFS_PER];
    auto load = [&](int i0, int (&v)[OFS_PER]) {
        const int i = i0 + OFS_PER * tid;
        if (vec && i + OFS_PER <= n) {
            const int4 a = __ldg(reinterpret_cast<const int4*>(row + i));
            const int4 b = __ldg(reinterpret_cast<const int4*>(row + i + 4));
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
        } else {
#pragma unroll
            for (int j = 0; j < OFS_PER; ++j) v[j] = (i + j < n) ? __ldg(row + i + j) : 0;
        }
    };
    load(0, c);
    int buf = 0;
    for (int i0 = 0; i0 < n; i0 += OFS_PER * OFS_THREADS, buf ^= 1) {
        if (i0 + OFS_PER * OFS_THREADS < n) load(i0 + OFS_PER * OFS_THREADS, nx);
        long long sum = 0;
#pragma unroll
        for (int j = 0; j < OFS_PER; ++j) sum += c[j];
        long long inc = sum;                               // inclusive scan over the lanes
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[buf][warp] = inc;
        __syncthreads();
        // every warp scans the 32 warp totals itself: no second barrier
        long long w = s_warp[buf][lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += t;
        }
        const long long tile_total = __shfl_sync(0xffffffffu, w, 31);
        const long long before_warp = __shfl_sync(0xffffffffu, w - s_warp[buf][lane], warp);
        long long r = run + before_warp + inc - sum;
        const int i = i0 + OFS_PER * tid;
        if (vec && i + OFS_PER <= n) {
#pragma unroll
            for (int j = 0; j < OFS_PER; j += 2) {
                const long long r0 = r, r1 = r + c[j];
                *reinterpret_cast<longlong2*>(out + i + j) = make_longlong2(r0, r1);
                r = r1 + c[j + 1];
            }
        } else {
#pragma unroll
            for (int j = 0; j < OFS_PER; ++j) {
                if (i + j < n) out[i + j] = r;
                r += c[j];
            }
        }
        run += tile_total;
#pragma unroll
        for (int j = 0; j < OFS_PER; ++j) c[j] = nx[j];
    }
    if (tid == 0) carry[u] = run;
}

extern "C" int xnn_site_offsets(const int32_t* counts, int U, int n, long long* carry, long long* offsets,
                                xnn_stream_t s) {
    XNN_CHECK_ARG(counts && carry && offsets, "NULL pointer");
    XNN_CHECK_ARG(U >= 1 && n >= 0, "bad sizes");
    if (n == 0) return XNN_OK;
    site_offsets_kernel<<<U, OFS_THREADS, 0, to_stream(s)>>>(counts, n, carry, offsets);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

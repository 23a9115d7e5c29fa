// a4 / a6 on the tensor cores -- each unit's Linear(n->100) -> BN -> ReLU -> Dropout -> Linear(100->1)
// (explainn/architectures/__init__.py:84-98) as a grouped GEMM on tcgen05, forward and backward, with the
// (B, 100*U) hidden activations never written to HBM: every pass recomputes its FC1 tile
//   hpre_u[b, m] = b1[u,m] + sum_q W1[100u+m, q] * pool[b, u, q]
// in TMEM (it costs ~1/10 of the HBM traffic it saves) and fuses what follows into the epilogue:
//   pass F1  BatchNorm statistics of hpre                       (thread = hidden channel, columns = batch)
//   pass F2  BN + ReLU + dropout + Linear(100->1) -> opre        (thread = batch row, columns = hidden)
//   pass B1  BN-backward reductions sum g, sum g*x, dW2, db2     (thread = hidden channel)
//   pass B2  dhpre -> dW1 (accumulated in TMEM), dpool -> dy     (mlp_tc_bwd.cu)
//
// fp32 parity on tensor cores: kind::tf32 with an error-compensated split of BOTH operands,
// x = hi + lo (each tf32, round-to-nearest), x*y ~= hi*hi' + hi*lo' + lo*hi' accumulated in fp32 (TMEM):
// relative error ~2^-22 per product, i.e. fp32 round-off level.
//
// Operand staging (shared memory, no swizzle): an [R rows][K] fp32 matrix is kept as float4 [K/4][R]
// ("chunk-major").  As a K-major UMMA operand that is LBO = 16R (next 4 K), SBO = 128 (next 8 rows);
// the SAME bytes read as an MN-major operand (LBO = 128, SBO = 16R) are the transposed matrix, which
// is how the backward pass gets W1^T, pool^T and dhpre^T without a second copy.
#include "common.cuh"
#include "umma.cuh"
#include <cuda_bf16.h>
using namespace xnn;

#define MT_THREADS 512
#define MT_ROWS 128
#define MT_H XNN_HIDDEN          // 100
#define MT_NMAX 96               // largest pool width n handled on this path
#define MT_PHILOX_CALLS 13       // 8 sixteen-bit draws per call -> 104 >= 100 hidden channels

namespace xnn {

__device__ __forceinline__ uint4 philox4(unsigned long long seed, unsigned long long idx) {
    uint32_t c0 = (uint32_t)idx, c1 = (uint32_t)(idx >> 32), c2 = 0x9E3779B9u, c3 = 0xBB67AE85u;
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

// keep-bits of one (sequence, unit) row of the hidden layer: bit m of kb[m/32] = keep channel m.
// Injected byte mask (tests), or Philox draws (16 bits each, keep if draw >= p*65536), or all ones.
__device__ __forceinline__ void dropout_keep_bits(float p, unsigned long long seed, long long bu,
                                                  const uint8_t* __restrict__ mask_u8, uint32_t kb[4]) {
    kb[0] = kb[1] = kb[2] = kb[3] = 0u;
    if (mask_u8) {
        const uint8_t* mr = mask_u8 + bu * MT_H;
#pragma unroll
        for (int m = 0; m < MT_H; ++m) kb[m >> 5] |= (mr[m] ? 1u : 0u) << (m & 31);
    } else if (p > 0.0f) {
        const uint32_t thr = (uint32_t)(p * 65536.0f + 0.5f);
#pragma unroll
        for (int i = 0; i < MT_PHILOX_CALLS; ++i) {
            const uint4 r = philox4(seed, (unsigned long long)bu * 16ull + (unsigned long long)i);
            const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int s = 0; s < 8; ++s) {
                const int m = 8 * i + s;
                const uint32_t d16 = (w[s >> 1] >> (16 * (s & 1))) & 0xFFFFu;
                if (m < MT_H) kb[m >> 5] |= (d16 >= thr ? 1u : 0u) << (m & 31);
            }
        }
    } else {
        kb[0] = kb[1] = kb[2] = 0xFFFFFFFFu; kb[3] = 0xFu;
    }
}

// stage one fp32 row (n values, rest zero up to 4*chunks) as tf32 hi / lo float4 chunks:
// dst_hi[c*R + row], dst_lo[c*R + row]
__device__ __forceinline__ void stage_row_split(const float* __restrict__ src, int n, int chunks, int R, int row,
                                                float4* __restrict__ dst_hi, float4* __restrict__ dst_lo) {
#pragma unroll 4
    for (int c = 0; c < chunks; ++c) {
        float v[4], h[4], l[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int q = 4 * c + e;
            v[e] = (src != nullptr && q < n) ? __ldg(src + q) : 0.0f;
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) umma::split_tf32(v[e], h[e], l[e]);
        dst_hi[c * R + row] = make_float4(h[0], h[1], h[2], h[3]);
        dst_lo[c * R + row] = make_float4(l[0], l[1], l[2], l[3]);
    }
}

}  // namespace xnn

struct MlpArgs {
    const float* pool;      // (B,U,n)
    const float* w1;        // (100U,n)
    const float* b1;        // (100U)
    const float* scale2;    // (100U)   F2 / B1
    const float* shift2;    // (100U)
    const float* w2;        // (U,100)
    const float* b2;        // (U)      F2
    const float* dopre;     // (B,U)    B1
    const uint8_t* mask_u8; // (B,U,100) injected dropout mask (F2) or NULL
    uint32_t* maskbits;     // (B,U,4)  active bits (kept by dropout AND BN2 output > 0): written by F2, read by B1/B2
    float* opre;            // (B,U)    F2
    double* part;           // [slots][nstat][C] partial sums (F1: 2, B1: 4)
    int B, U, n, T, slots;
    float drop_p;
    unsigned long long seed;
    const unsigned long long* seed_dev;   // optional device-resident draw counter mixed into the seed (graph replays)
    unsigned char* image_out;             // F1: write the staged pool tiles here (mlp_image layout), or NULL
    const unsigned char* image;           // F2 / B1 image kernels: pool tile images
    const unsigned char* w1_image;        //                        W1 images [U][3][chunksQ][112 rows] uint4
};

enum { MLP_STATS = 0, MLP_APPLY = 1, MLP_BWD_STATS = 2 };

struct MtShared {
    uint64_t bar;
    uint32_t tmem_slot;
    uint32_t pad;
};

// ---------------------------------------------------------------------------
// FC1 tile on tcgen05 + fused epilogue, warp-specialised.  One CTA per SM walks a contiguous range of the
// flattened (unit, batch tile) space, so W1 of a unit is staged once per CTA (at most ~3 units per CTA).
// Measured on the lock-step version (every warp staging, then waiting, then in the epilogue): global loads
// 0.10 ms + staging stores 0.055 + MMAs 0.065 + epilogue 0.04 (statistics) .. 0.155 (apply) + loop 0.07 were
// ADDITIVE.  Here three roles run on different tiles at the same time:
//   warps 0-15  producers: prefetch tile t+1 of `pool` into registers, split it (three bf16 pieces, kind::f16
//               with the six significant products = fp32-level accuracy at 6 instead of 8 bytes per element)
//               into pool stage t%2 as soon as the MMAs of tile t-2 released it, restage W1 at a unit
//               boundary (W1 is kept in a 104-row layout, 60 KB, so that two 74 KB pool stages fit); in the
//               backward-statistics pass also dopre and the active bits of the tile (four-deep ring)
//   warp  24    issuer: waits for "staged" and for the accumulator stage, issues the tile's MMAs
//   warps 16-23 epilogue of tile t out of TMEM stage t%2 while tile t+1 is staged and multiplied:
//               warp e owns TMEM lanes 32*(e%4).. and the column half e/4
// mbarriers: staged[2] (16 producer warps -> issuer), mma_done[2] (tcgen05.commit -> epilogue + producers),
// acc_free[2] (8 epilogue warps -> issuer).
// ---------------------------------------------------------------------------
#define MT_CMAX 24   // staging chunks (of 4 floats) per row: MT_NMAX / 4
#define WS_P_MAX 16     // producer warps: 16 for the statistics pass (one row each), 8 for apply (two rows each)
#define WS_E_WARPS 8

struct WsShared {
    uint64_t staged[2];
    uint64_t mma_done[2];
    uint64_t acc_free[2];
    uint32_t tmem_slot;
    uint32_t pad;
};

__device__ __forceinline__ void ws_epi_barrier() {
    asm volatile("bar.sync 1, %0;" ::"n"(32 * WS_E_WARPS) : "memory");
}

// keep-bit word `w` (hidden channels 32w .. 32w+31) of one (sequence, unit) row
__device__ __forceinline__ uint32_t dropout_keep_word(float p, unsigned long long seed, long long bu,
                                                      const uint8_t* __restrict__ mask_u8, int w) {
    uint32_t kb = 0u;
    if (mask_u8) {
        const uint8_t* mr = mask_u8 + bu * MT_H;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const int m = 32 * w + j;
            if (m < MT_H) kb |= (mr[m] ? 1u : 0u) << j;
        }
    } else if (p > 0.0f) {
        const uint32_t thr = (uint32_t)(p * 65536.0f + 0.5f);
#pragma unroll
        for (int ii = 0; ii < 4; ++ii) {
            const int i = 4 * w + ii;                      // Philox call index: channels 8i .. 8i+7
            if (i < MT_PHILOX_CALLS) {
                const uint4 r = philox4(seed, (unsigned long long)bu * 16ull + (unsigned long long)i);
                const uint32_t wd[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
                for (int s = 0; s < 8; ++s) {
                    const uint32_t d16 = (wd[s >> 1] >> (16 * (s & 1))) & 0xFFFFu;
                    if (8 * i + s < MT_H) kb |= (d16 >= thr ? 1u : 0u) << (8 * ii + s);
                }
            }
        }
    } else {
        kb = (w < 3) ? 0xFFFFFFFFu : 0xFu;
    }
    return kb;
}

// two floats -> one packed bf16 pair per piece (hi, mid, lo).  Each piece is the TRUNCATED upper half of the
// running remainder (a mask and an exact subtraction instead of three round-to-nearest conversions):
// x = hi + mid + lo + e with |e| < 2^-23 |x|, every piece exactly representable in bf16.
__device__ __forceinline__ void split2_bf16x3(float x0, float x1, uint32_t& w0, uint32_t& w1, uint32_t& w2) {
    const uint32_t h0 = __float_as_uint(x0) & 0xFFFF0000u, h1 = __float_as_uint(x1) & 0xFFFF0000u;
    const float r0 = x0 - __uint_as_float(h0), r1 = x1 - __uint_as_float(h1);
    const uint32_t m0 = __float_as_uint(r0) & 0xFFFF0000u, m1 = __float_as_uint(r1) & 0xFFFF0000u;
    const float q0 = r0 - __uint_as_float(m0), q1 = r1 - __uint_as_float(m1);
    w0 = __byte_perm(h0, h1, 0x7632);                       // (upper half of x0) | (upper half of x1) << 16
    w1 = __byte_perm(m0, m1, 0x7632);
    w2 = __byte_perm(__float_as_uint(q0), __float_as_uint(q1), 0x7632);
}

// The same three pieces with ROUND-TO-NEAREST conversions (cvt.rn.bf16x2.f32 packs the pair): hi + mid + lo = x to
// 2^-25 |x|, and hi + mid alone = x to 2^-18 |x| with an UNBIASED error, which is what lets the backward pass run
// on two pieces (three products) inside its 1e-4 gradient tolerance.  Used for everything that goes into the tile
// images (mlp_image layout).
__device__ __forceinline__ uint32_t cvt_bf16x2_rn(float lo_elem, float hi_elem) {
    uint32_t w;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w) : "f"(hi_elem), "f"(lo_elem));
    return w;
}
__device__ __forceinline__ void split2_bf16x3_rn(float x0, float x1, uint32_t& w0, uint32_t& w1, uint32_t& w2) {
    w0 = cvt_bf16x2_rn(x0, x1);
    const float r0 = x0 - __uint_as_float(w0 << 16), r1 = x1 - __uint_as_float(w0 & 0xFFFF0000u);
    w1 = cvt_bf16x2_rn(r0, r1);
    const float q0 = r0 - __uint_as_float(w1 << 16), q1 = r1 - __uint_as_float(w1 & 0xFFFF0000u);
    w2 = cvt_bf16x2_rn(q0, q1);
}

#define WS_WROWS 104   // rows of the W1 operand layout (100 hidden channels, padded to a multiple of 8)
#define WS_QMAX 6      // 16-byte chunks per producer lane: 2 columns of each of <= 12 chunks -> pv[2][12][2]

template <int MODE, int WS_P_WARPS>
__global__ void __launch_bounds__(32 * (WS_P_WARPS + WS_E_WARPS + 1), 1)
mlp_fc1_kernel(MlpArgs a) {
    constexpr int WS_THREADS = 32 * (WS_P_WARPS + WS_E_WARPS + 1);
    constexpr int HS = 16 / WS_P_WARPS;          // rows of the tile per producer thread
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int n = a.n, U = a.U, B = a.B, T = a.T;
    const int Nq = (n + 15) & ~15, chunksQ = Nq >> 3, kstepsQ = Nq >> 4;       // K in 16-byte chunks of 8 bf16
    const uint32_t wlbo = WS_WROWS * 16u;                                     // W1: [piece][chunk][104 rows] uint4
    const uint32_t wpiece = (uint32_t)chunksQ * wlbo;
    const uint32_t ppiece = (uint32_t)chunksQ * 2048u;                        // pool: [stage][piece][chunk][128 rows] uint4
    const uint32_t pstage = 3u * ppiece;
    uint4* w1s = reinterpret_cast<uint4*>(smem_raw);
    uint4* ps = reinterpret_cast<uint4*>(smem_raw + 3 * wpiece);
    unsigned char* tail = smem_raw + 3 * wpiece + 2 * pstage;
    float4* s_c4 = reinterpret_cast<float4*>(tail);                  // [128] (b1, scale2, shift2, w2) per channel
    float* s_dopre_all = reinterpret_cast<float*>(tail + 2048);      // [4][128]
    uint32_t* s_mask_all = reinterpret_cast<uint32_t*>(tail + 2048 + 2048);   // [4][128][4]
    float* s_acc = reinterpret_cast<float*>(tail + 2048 + 2048 + 8192);       // [2][2][128]
    WsShared* sh = reinterpret_cast<WsShared*>(tail + 2048 + 2048 + 8192 + 2048);
    uint2* s_tab = reinterpret_cast<uint2*>(tail + 2048 + 2048 + 8192 + 2048 + 64);   // [6 * kstepsQ] descriptor words

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    {
        uint4* z = reinterpret_cast<uint4*>(smem_raw);
        for (uint32_t i = tid; i < (3 * wpiece + 2 * pstage) / 16; i += WS_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    if (warp == 0) umma::tmem_alloc(&sh->tmem_slot, 256);            // two accumulator stages of 128 columns
    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            umma::mbar_init(&sh->staged[s], WS_P_WARPS);
            umma::mbar_init(&sh->mma_done[s], 1);
            umma::mbar_init(&sh->acc_free[s], WS_E_WARPS);
        }
    }
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = sh->tmem_slot;
    const uint32_t w_addr = umma::smem_u32(w1s), p_addr = umma::smem_u32(ps);
    // F1 / B1: D[hidden][batch] (A = W1, N = 128 batch rows); F2: D[batch][hidden] (A = pool tile, N = 112)
    constexpr int NDIM = (MODE == MLP_APPLY) ? 112 : 128;
    const uint32_t idesc = umma::idesc_bf16_f32(128, NDIM);
    // MMA sequence of a tile: the six significant products of the 3-way split, smallest first
    // (piece of W1, piece of pool): lo*hi, hi*lo, mid*mid, mid*hi, hi*mid, hi*hi   (0 = hi, 1 = mid, 2 = lo)
    if (tid < 6 * kstepsQ) {
        const int pass = tid / kstepsQ, ks = tid - pass * kstepsQ;
        const uint32_t pw = pass == 0 ? 2u : pass == 1 ? 0u : pass == 2 ? 1u : pass == 3 ? 1u : 0u;
        const uint32_t pp = pass == 0 ? 0u : pass == 1 ? 2u : pass == 2 ? 1u : pass == 3 ? 0u : pass == 4 ? 1u : 0u;
        const uint32_t wd = umma::desc_lo(w_addr + pw * wpiece + (uint32_t)(2 * ks) * wlbo, wlbo);
        const uint32_t pd2 = umma::desc_lo(p_addr + pp * ppiece + (uint32_t)(2 * ks) * 2048u, 2048u);
        s_tab[tid] = (MODE == MLP_APPLY) ? make_uint2(pd2, wd) : make_uint2(wd, pd2);   // (A, B)
    }
    __syncthreads();

    const long long tiles_total = (long long)U * T;
    const long long lo = (long long)blockIdx.x * tiles_total / gridDim.x;
    const long long hi = (long long)(blockIdx.x + 1) * tiles_total / gridDim.x;
    const int ntl = (int)(hi - lo);
    const float inv_keep = 1.0f / (1.0f - a.drop_p);
    const int C = U * MT_H;
    const unsigned long long seed = a.seed + (a.seed_dev ? *a.seed_dev * 0x9E3779B97F4A7C15ull : 0ull);

    if (warp < WS_P_WARPS) {
        // ================= producers =================
        // staging map (coalesced): lane = (row & 7) * 4 + word, word = two consecutive pool columns of a 16-byte
        // (8 x bf16) chunk: one warp-wide access covers 8 rows x 32 B of global memory and 32 distinct banks;
        // warp p stages rows 8p .. 8p+7 (and 64+8p .. with 8 producer warps), all chunks
        const int swd = lane & 3;
        const int srow0 = warp * 8 + (lane >> 2);
        float pv[HS][12][2];
        float pd = 0.0f;
        uint4 pm = make_uint4(0u, 0u, 0u, 0u);
        auto prefetch = [&](int i) {
            const bool live = i < ntl;
            const long long tile = lo + i;
            const int u = live ? (int)(tile / T) : 0, t = live ? (int)(tile % T) : 0;
            const int b0 = t * MT_ROWS;
#pragma unroll
            for (int h = 0; h < HS; ++h) {
                const int srow = srow0 + 8 * WS_P_WARPS * h;
                const float* src = (live && b0 + srow < B) ? a.pool + ((long long)(b0 + srow) * U + u) * n : nullptr;
#pragma unroll
                for (int c = 0; c < 12; ++c) {
                    const int q = 8 * c + 2 * swd;
                    pv[h][c][0] = (src != nullptr && q < n) ? __ldg(src + q) : 0.0f;
                    pv[h][c][1] = (src != nullptr && q + 1 < n) ? __ldg(src + q + 1) : 0.0f;
                }
            }
            if (MODE == MLP_BWD_STATS && tid < MT_ROWS) {
                const bool ok = live && b0 + tid < B;
                const long long bu = (long long)(b0 + tid) * U + u;
                pd = ok ? __ldg(a.dopre + bu) : 0.0f;
                pm = ok ? __ldg(reinterpret_cast<const uint4*>(a.maskbits + bu * 4)) : make_uint4(0u, 0u, 0u, 0u);
            }
            // L2 prefetch of the tile after the next two: the register prefetch above runs one tile ahead;
            // thread (row, part) touches one 128-byte line of the row
            if (i + 2 < ntl) {
                const long long tile2 = lo + i + 2;
                const int u2 = (int)(tile2 / T), r2 = (int)(tile2 % T) * MT_ROWS + (tid & 127);
                const char* p2 = reinterpret_cast<const char*>(a.pool + ((long long)r2 * U + u2) * n);
#pragma unroll
                for (int part = tid >> 7; part < 4; part += WS_P_WARPS / 4) {
                    const int off = part < 3 ? 128 * part : 4 * n - 1;
                    if (r2 < B && off < 4 * n) asm volatile("prefetch.global.L2 [%0];" ::"l"(p2 + off));
                }
            }
        };
        int cur_u = -1;
        prefetch(0);
        for (int i = 0; i < ntl; ++i) {
            const int u = (int)((lo + i) / T);
            const uint32_t st = (uint32_t)i & 1u;
            if (i >= 2) umma::mbar_wait(&sh->mma_done[st], (uint32_t)((i >> 1) - 1) & 1u);   // pool stage free
            if (u != cur_u) {
                cur_u = u;
                // every MMA reads W1: wait for the previous tile's too, then restage (rows >= 100 are never read
                // into a used accumulator lane / column)
                if (i >= 1) umma::mbar_wait(&sh->mma_done[st ^ 1u], (uint32_t)((i - 1) >> 1) & 1u);
                uint32_t* ww = reinterpret_cast<uint32_t*>(w1s);
#pragma unroll
                for (int h = 0; h < HS; ++h) {
                    const int srow = srow0 + 8 * WS_P_WARPS * h;
                    if (srow < MT_H) {
                        const float* src = a.w1 + ((long long)u * MT_H + srow) * n;
                        for (int c = 0; c < chunksQ; ++c) {
                            const int q = 8 * c + 2 * swd;
                            uint32_t w0, w1v, w2v;
                            split2_bf16x3(q < n ? __ldg(src + q) : 0.0f, q + 1 < n ? __ldg(src + q + 1) : 0.0f, w0, w1v, w2v);
                            const int o = (c * WS_WROWS + srow) * 4 + swd;
                            ww[o] = w0; ww[wpiece / 4 + o] = w1v; ww[2 * (wpiece / 4) + o] = w2v;
                        }
                    }
                }
            }
            uint32_t* pw = reinterpret_cast<uint32_t*>(ps) + st * (pstage / 4);
#pragma unroll
            for (int h = 0; h < HS; ++h) {
                const int srow = srow0 + 8 * WS_P_WARPS * h;
#pragma unroll
                for (int c = 0; c < 12; ++c) {
                    if (c < chunksQ) {
                        uint32_t w0, w1v, w2v;
                        if (MODE == MLP_STATS) split2_bf16x3_rn(pv[h][c][0], pv[h][c][1], w0, w1v, w2v);
                        else split2_bf16x3(pv[h][c][0], pv[h][c][1], w0, w1v, w2v);
                        const int o = (c * 128 + srow) * 4 + swd;
                        pw[o] = w0; pw[ppiece / 4 + o] = w1v; pw[2 * (ppiece / 4) + o] = w2v;
                    }
                }
            }
            if (MODE == MLP_BWD_STATS && tid < MT_ROWS) {
                s_dopre_all[(i & 3) * 128 + tid] = pd;
                *reinterpret_cast<uint4*>(s_mask_all + (i & 3) * 512 + 4 * tid) = pm;
            }
            umma::fence_proxy_async();
            __syncwarp();
            if (lane == 0) umma::mbar_arrive(&sh->staged[st]);
            prefetch(i + 1);                    // global loads of the next tile fly under the MMAs
        }
    } else if (warp == WS_P_WARPS + WS_E_WARPS) {
        // ================= issuer: one elected lane per tile, descriptors in the uniform datapath =================
        const uint32_t w_base = umma::warp_uniform(umma::desc_lo(w_addr, wlbo));
        const uint32_t p_base = umma::warp_uniform(umma::desc_lo(p_addr, 2048u));
        const uint32_t tmem_u = umma::warp_uniform(tmem);
        for (int i = 0; i < ntl; ++i) {
            const uint32_t st = (uint32_t)i & 1u;
            umma::mbar_wait(&sh->staged[st], ((uint32_t)i >> 1) & 1u);
            umma::mbar_wait(&sh->acc_free[st], (((uint32_t)i >> 1) & 1u) ^ 1u);
            umma::fence_after_sync();
            if (umma::elect_one()) {
                const uint32_t pst = p_base + st * (pstage >> 4);    // pool stage, in descriptor units
                const bool img = (MODE == MLP_STATS) && a.image_out != nullptr;
                // the staged tile IS its image (mlp_image layout): one bulk store, no LSU work; the later passes of
                // the step fetch their operand stages from it with one bulk load each
                if (img) {
                    umma::bulk_store(a.image_out + (size_t)(lo + i) * pstage,
                                     reinterpret_cast<unsigned char*>(ps) + (size_t)st * pstage, pstage);
                    umma::bulk_commit_group();
                }
                if (MODE == MLP_APPLY)      // A = pool tile, B = W1
                    umma::issue_split6(pst, ppiece >> 4, 256u, umma::desc_hi(128u), w_base, wpiece >> 4, (2u * wlbo) >> 4,
                                       umma::desc_hi(128u), kstepsQ, tmem_u + st * 128u, idesc);
                else                        // A = W1, B = pool tile
                    umma::issue_split6(w_base, wpiece >> 4, (2u * wlbo) >> 4, umma::desc_hi(128u), pst, ppiece >> 4, 256u,
                                       umma::desc_hi(128u), kstepsQ, tmem_u + st * 128u, idesc);
                // mma_done[st] releases the stage to the producers: the store must have read it by then
                if (img) umma::bulk_wait_read_all();
                umma::commit(&sh->mma_done[st]);
            }
            __syncwarp();
        }
    } else {
        // ================= epilogue =================
        const int e = warp - WS_P_WARPS;
        const int quarter = e & 3, cg = e >> 2;      // TMEM lane quarter (= warp % 4) and column half
        const int row = quarter * 32 + lane;
        int cur_u = -1;
        float b1m = 0.f, scm = 0.f, shm = 0.f, w2m = 0.f;   // per-channel constants (thread = channel modes)
        double s1 = 0.0, s2 = 0.0, s3 = 0.0, s4 = 0.0;
        auto flush = [&](int u) {
            if (MODE == MLP_APPLY) return;
            int j = blockIdx.x;
            const long long start = (long long)u * T;
            while (j > 0 && ((long long)j * tiles_total) / gridDim.x > start) --j;
            const int slot = (blockIdx.x - j) * 4 + cg;
            if (row < MT_H && slot < a.slots) {
                const int c = u * MT_H + row;
                constexpr int NS = (MODE == MLP_STATS) ? 2 : 4;
                double* o = a.part + (long long)slot * NS * C + c;
                o[0] = s1; o[(long long)C] = s2;
                if (MODE == MLP_BWD_STATS) { o[2LL * C] = s3; o[3LL * C] = s4; }
            }
            s1 = s2 = s3 = s4 = 0.0;
        };
        for (int i = 0; i < ntl; ++i) {
            const long long tile = lo + i;
            const int u = (int)(tile / T), t = (int)(tile % T);
            const int b0 = t * MT_ROWS, rows = min(MT_ROWS, B - b0);
            const uint32_t st = (uint32_t)i & 1u;
            if (u != cur_u) {
                if (cur_u >= 0) flush(cur_u);
                cur_u = u;
                const int ch = u * MT_H + row;
                const bool ok = row < MT_H;
                b1m = ok ? a.b1[ch] : 0.f;
                if (MODE != MLP_STATS) { scm = ok ? a.scale2[ch] : 0.f; shm = ok ? a.shift2[ch] : 0.f; w2m = ok ? a.w2[ch] : 0.f; }
                if (MODE == MLP_APPLY) {
                    ws_epi_barrier();           // every epilogue warp is done with the previous unit's constants
                    if (cg == 0) s_c4[row] = make_float4(b1m, scm, shm, w2m);
                    ws_epi_barrier();
                }
            }
            umma::mbar_wait(&sh->mma_done[st], ((uint32_t)i >> 1) & 1u);
            umma::fence_after_sync();
            const uint32_t trow = tmem + st * 128u + ((uint32_t)(quarter * 32) << 16);
            if (MODE == MLP_STATS || MODE == MLP_BWD_STATS) {
                // thread = (hidden channel, half of the tile's batch rows)
                const float* s_dopre = s_dopre_all + (i & 3) * 128;
                const uint32_t* s_mask = s_mask_all + (i & 3) * 512;
#pragma unroll
                for (int hh = 0; hh < 4; ++hh) {
                    const int c0 = 64 * cg + 16 * hh;
                    if (c0 < rows) {
                        uint32_t v[16];
                        umma::tmem_ld16(trow + c0, v);
                        umma::tmem_ld_wait();
                        float f1 = 0.0f, f2 = 0.0f, f3 = 0.0f, f4 = 0.0f;   // backward: fp32 partials per chunk
#pragma unroll
                        for (int k2 = 0; k2 < 16; ++k2) {
                            if (c0 + k2 < rows) {
                                const float x = __uint_as_float(v[k2]) + b1m;
                                if (MODE == MLP_STATS) {
                                    s1 += (double)x;
                                    s2 += (double)x * (double)x;
                                } else {
                                    const float pre = fmaf(x, scm, shm);
                                    const uint32_t act = (s_mask[4 * (c0 + k2) + (row >> 5)] >> (row & 31)) & 1u;
                                    const float d = s_dopre[c0 + k2];
                                    const float hact = act ? pre * inv_keep : 0.0f;
                                    const float g2 = act ? d * w2m * inv_keep : 0.0f;
                                    f1 += g2;
                                    f2 = fmaf(g2, x, f2);
                                    f3 = fmaf(d, hact, f3);
                                    f4 += d;
                                }
                            }
                        }
                        if (MODE == MLP_BWD_STATS) { s1 += (double)f1; s2 += (double)f2; s3 += (double)f3; s4 += (double)f4; }
                    }
                }
                umma::fence_before_sync();
                __syncwarp();
                if (lane == 0) umma::mbar_arrive(&sh->acc_free[st]);
            } else {
                // thread = (batch row, 64 hidden channels = keep-bit words 2cg, 2cg+1)
                const long long bu = (long long)(b0 + row) * U + u;
                uint32_t kbw[2], actw[2] = {0u, 0u};
                kbw[0] = (row < rows) ? dropout_keep_word(a.drop_p, seed, bu, a.mask_u8, 2 * cg) : 0u;
                kbw[1] = (row < rows) ? dropout_keep_word(a.drop_p, seed, bu, a.mask_u8, 2 * cg + 1) : 0u;
                float accv = 0.0f;
#pragma unroll
                for (int hh = 0; hh < 4; ++hh) {
                    const int c0 = 64 * cg + 16 * hh;
                    if (c0 < 112) {
                        uint32_t v[16];
                        umma::tmem_ld16(trow + c0, v);
                        umma::tmem_ld_wait();
#pragma unroll
                        for (int k2 = 0; k2 < 16; ++k2) {
                            const int m = c0 + k2;
                            if (m < MT_H) {
                                const float4 c4 = s_c4[m];
                                const float pre = fmaf(__uint_as_float(v[k2]) + c4.x, c4.y, c4.z);
                                const int bit = 16 * (hh & 1) + k2;
                                const uint32_t act = ((kbw[hh >> 1] >> bit) & 1u) & (pre > 0.0f ? 1u : 0u);
                                actw[hh >> 1] |= act << bit;   // keep AND (BN2 output > 0): the backward reuses THIS decision
                                const float h = act ? pre * inv_keep : 0.0f;
                                accv = fmaf(h, c4.w, accv);
                            }
                        }
                    }
                }
                umma::fence_before_sync();
                __syncwarp();
                if (lane == 0) umma::mbar_arrive(&sh->acc_free[st]);
                float* acc2 = s_acc + st * 256;              // [tile parity][column half][row]
                acc2[cg * 128 + row] = accv;
                if (row < rows && a.maskbits)
                    *reinterpret_cast<uint2*>(a.maskbits + bu * 4 + 2 * cg) = make_uint2(actw[0], actw[1]);
                ws_epi_barrier();
                if (cg == 0 && row < rows) a.opre[bu] = (acc2[row] + acc2[128 + row]) + a.b2[u];
            }
        }
        if (cur_u >= 0) flush(cur_u);
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, 256);
}


// ---------------------------------------------------------------------------
// The lock-step version of the same tile loop (512 threads: every warp stages, waits for the MMAs, then runs
// the epilogue; kind::tf32 hi/lo split).  Kept for the backward-statistics pass, whose fp64-heavy epilogue is
// faster on 16 epilogue warps than on the 8 of the warp-specialised kernel (0.43 vs 0.58 ms at cfg3 B=4096).
// ---------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(MT_THREADS, 1)
mlp_fc1_lockstep_kernel(MlpArgs a) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int n = a.n, U = a.U, B = a.B, T = a.T;
    const int Kp = (n + 7) & ~7, chunks = Kp >> 2, ksteps = Kp >> 3;
    const uint32_t piece_bytes = (uint32_t)chunks * 2048u;           // [chunks][128 rows] float4
    float4* w1s = reinterpret_cast<float4*>(smem_raw);               // [2][chunks][128]
    float4* ps = reinterpret_cast<float4*>(smem_raw + 2 * piece_bytes);   // [2][chunks][128]
    unsigned char* tail = smem_raw + 4 * piece_bytes;
    float4* s_c4 = reinterpret_cast<float4*>(tail);                  // [128] (b1, scale2, shift2, w2) per channel
    float* s_dopre = reinterpret_cast<float*>(tail + 2048);          // [128]
    uint32_t* s_mask = reinterpret_cast<uint32_t*>(tail + 2048 + 512);   // [128][4]
    float* s_acc = reinterpret_cast<float*>(tail + 2048 + 512 + 2048);   // [4][128]
    MtShared* sh = reinterpret_cast<MtShared*>(tail + 2048 + 512 + 2048 + 2048);
    uint2* s_tab = reinterpret_cast<uint2*>(tail + 2048 + 512 + 2048 + 2048 + 32);   // [3 * ksteps] descriptor words

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int row = (warp & 3) * 32 + lane;   // TMEM lane = staging row
    const int cg = warp >> 2;                 // column group / staging quarter
    {
        uint4* z = reinterpret_cast<uint4*>(smem_raw);
        for (uint32_t i = tid; i < 4 * piece_bytes / 16; i += MT_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    if (warp == 0) umma::tmem_alloc(&sh->tmem_slot, 128);
    if (tid == 0) umma::mbar_init(&sh->bar, 1);
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = sh->tmem_slot;
    const uint32_t w_addr = umma::smem_u32(w1s), p_addr = umma::smem_u32(ps);
    // F1 / B1: D[hidden][batch] (A = W1, N = 128 batch rows); F2: D[batch][hidden] (A = pool tile, N = 112)
    constexpr int NDIM = (MODE == MLP_APPLY) ? 112 : 128;
    const uint32_t idesc = umma::idesc_tf32_f32(128, NDIM, 0, 0);
    const uint32_t a_addr = (MODE == MLP_APPLY) ? p_addr : w_addr;
    const uint32_t b_addr = (MODE == MLP_APPLY) ? w_addr : p_addr;
    // MMA sequence of a tile (the operand buffers never move): small terms first, lo*hi, hi*lo, then hi*hi
    if (tid < 3 * ksteps) {
        const int pass = tid / ksteps, ks = tid - pass * ksteps;
        const uint32_t ap = (pass == 0) ? 1u : 0u, bp = (pass == 1) ? 1u : 0u;   // piece 0 = hi, 1 = lo
        s_tab[tid] = make_uint2(umma::desc_lo(a_addr + ap * piece_bytes + (uint32_t)(2 * ks) * 2048u, 2048u),
                                umma::desc_lo(b_addr + bp * piece_bytes + (uint32_t)(2 * ks) * 2048u, 2048u));
    }
    __syncthreads();

    const long long tiles_total = (long long)U * T;
    const long long lo = (long long)blockIdx.x * tiles_total / gridDim.x;
    const long long hi = (long long)(blockIdx.x + 1) * tiles_total / gridDim.x;
    const float inv_keep = 1.0f / (1.0f - a.drop_p);
    const int C = U * MT_H;
    // staging map (coalesced): warp w stages rows 8w..8w+7, lane = (row & 7) * 4 + element of the 4-float chunk,
    // so one warp-wide access covers 8 rows x 16 B in global memory and 32 distinct banks in shared memory
    const int srow = warp * 8 + (lane >> 2), se = lane & 3;

    int cur_u = -1;
    uint32_t phase = 0;
    float b1m = 0.f, scm = 0.f, shm = 0.f, w2m = 0.f;   // per-channel constants (thread = channel modes)
    double s1 = 0.0, s2 = 0.0, s3 = 0.0, s4 = 0.0;
    float pv[MT_CMAX];
    float pd = 0.0f;                                     // B1: prefetched dopre / active bits of row `row`
    uint4 pm = make_uint4(0u, 0u, 0u, 0u);

    auto prefetch = [&](long long tile) {
        const bool live = tile < hi;
        const int u = live ? (int)(tile / T) : 0, t = live ? (int)(tile % T) : 0;
        const int b0 = t * MT_ROWS;
        const float* src = (live && b0 + srow < B) ? a.pool + ((long long)(b0 + srow) * U + u) * n : nullptr;
#pragma unroll
        for (int c = 0; c < MT_CMAX; ++c) {
            const int q = 4 * c + se;
            pv[c] = (src != nullptr && q < n) ? __ldg(src + q) : 0.0f;
        }
        if (MODE == MLP_BWD_STATS && cg == 0) {
            const bool ok = live && b0 + row < B;
            const long long bu = (long long)(b0 + row) * U + u;
            pd = ok ? __ldg(a.dopre + bu) : 0.0f;
            pm = ok ? __ldg(reinterpret_cast<const uint4*>(a.maskbits + bu * 4)) : make_uint4(0u, 0u, 0u, 0u);
        }
    };
    auto store_tile = [&]() {
        float* hi_p = reinterpret_cast<float*>(ps) + srow * 4 + se;
        float* lo_p = hi_p + chunks * 512;
#pragma unroll
        for (int c = 0; c < MT_CMAX; ++c) {
            if (c < chunks) {
                float h, l;
                umma::split_tf32(pv[c], h, l);
                hi_p[c * 512] = h;
                lo_p[c * 512] = l;
            }
        }
        if (MODE == MLP_BWD_STATS && cg == 0) {
            s_dopre[row] = pd;
            *reinterpret_cast<uint4*>(s_mask + 4 * row) = pm;
        }
    };
    auto flush = [&](int u) {
        if (MODE == MLP_APPLY) return;
        int j = blockIdx.x;
        const long long start = (long long)u * T;
        while (j > 0 && ((long long)j * tiles_total) / gridDim.x > start) --j;
        const int slot = (blockIdx.x - j) * 4 + cg;
        if (row < MT_H && slot < a.slots) {
            const int c = u * MT_H + row;
            constexpr int NS = (MODE == MLP_STATS) ? 2 : 4;
            double* o = a.part + (long long)slot * NS * C + c;
            o[0] = s1; o[(long long)C] = s2;
            if (MODE == MLP_BWD_STATS) { o[2LL * C] = s3; o[3LL * C] = s4; }
        }
        s1 = s2 = s3 = s4 = 0.0;
    };

    prefetch(lo);
    for (long long tile = lo; tile < hi; ++tile) {
        const int u = (int)(tile / T), t = (int)(tile % T);
        const int b0 = t * MT_ROWS, rows = min(MT_ROWS, B - b0);
        if (u != cur_u) {
            if (cur_u >= 0) flush(cur_u);
            cur_u = u;
            // W1 of this unit, same coalesced map (rows >= 100 stay zero)
            if (srow < MT_H) {
                const float* src = a.w1 + ((long long)u * MT_H + srow) * n;
                float* hi_p = reinterpret_cast<float*>(w1s) + srow * 4 + se;
                float* lo_p = hi_p + chunks * 512;
#pragma unroll 8
                for (int c = 0; c < chunks; ++c) {
                    const int q = 4 * c + se;
                    float h, l;
                    umma::split_tf32(q < n ? __ldg(src + q) : 0.0f, h, l);
                    hi_p[c * 512] = h;
                    lo_p[c * 512] = l;
                }
            }
            const int ch = u * MT_H + row;
            const bool ok = row < MT_H;
            b1m = ok ? a.b1[ch] : 0.f;
            if (MODE != MLP_STATS) {
                scm = ok ? a.scale2[ch] : 0.f; shm = ok ? a.shift2[ch] : 0.f; w2m = ok ? a.w2[ch] : 0.f;
                if (MODE == MLP_APPLY && cg == 0) s_c4[row] = make_float4(b1m, scm, shm, w2m);
            }
        }
        store_tile();
        umma::fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
            umma::fence_after_sync();
            umma::issue_table<true>(s_tab, 3 * ksteps, umma::desc_hi(128u), umma::desc_hi(128u), 0u, tmem, idesc, 0u);
            umma::commit(&sh->bar);
        }
        prefetch(tile + 1);                 // global loads of the next tile fly under the MMAs
        umma::mbar_wait(&sh->bar, phase);
        phase ^= 1u;
        umma::fence_after_sync();
        const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16);

        if (MODE == MLP_STATS || MODE == MLP_BWD_STATS) {
            // thread = (hidden channel, quarter of the tile's batch rows)
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                const int c0 = 32 * cg + 16 * hh;
                if (c0 < rows) {
                    uint32_t v[16];
                    umma::tmem_ld16(trow + c0, v);
                    umma::tmem_ld_wait();
                    // backward reductions: fp32 partial sums over the 16 columns, folded into the fp64 running sums
                    // once per chunk (these sums have no large cancelling terms, unlike E[x^2] - mean^2)
                    float f1 = 0.0f, f2 = 0.0f, f3 = 0.0f, f4 = 0.0f;
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        if (c0 + i < rows) {
                            const float x = __uint_as_float(v[i]) + b1m;
                            if (MODE == MLP_STATS) {
                                s1 += (double)x;
                                s2 += (double)x * (double)x;
                            } else {
                                const float pre = fmaf(x, scm, shm);
                                const uint32_t act = (s_mask[4 * (c0 + i) + (row >> 5)] >> (row & 31)) & 1u;
                                const float d = s_dopre[c0 + i];
                                const float hact = act ? pre * inv_keep : 0.0f;
                                const float g2 = act ? d * w2m * inv_keep : 0.0f;
                                f1 += g2;
                                f2 = fmaf(g2, x, f2);
                                f3 = fmaf(d, hact, f3);
                                f4 += d;
                            }
                        }
                    }
                    if (MODE == MLP_BWD_STATS) { s1 += (double)f1; s2 += (double)f2; s3 += (double)f3; s4 += (double)f4; }
                }
            }
        } else {
            // thread = (batch row, 32 hidden channels = keep-bit word cg)
            const long long bu = (long long)(b0 + row) * U + u;
            const uint32_t kbw = (row < rows) ? dropout_keep_word(a.drop_p, a.seed, bu, a.mask_u8, cg) : 0u;
            uint32_t actw = 0u;   // keep AND (BN2 output > 0): the backward passes reuse THIS decision
            float accv = 0.0f;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                const int c0 = 32 * cg + 16 * hh;
                if (c0 < 112) {
                    uint32_t v[16];
                    umma::tmem_ld16(trow + c0, v);
                    umma::tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const int m = c0 + i;
                        if (m < MT_H) {
                            const float4 c4 = s_c4[m];
                            const float pre = fmaf(__uint_as_float(v[i]) + c4.x, c4.y, c4.z);
                            const uint32_t act = ((kbw >> (16 * hh + i)) & 1u) & (pre > 0.0f ? 1u : 0u);
                            actw |= act << (16 * hh + i);
                            const float h = act ? pre * inv_keep : 0.0f;
                            accv = fmaf(h, c4.w, accv);
                        }
                    }
                }
            }
            s_acc[cg * 128 + row] = accv;
            if (row < rows && a.maskbits) a.maskbits[bu * 4 + cg] = actw;
            __syncthreads();
            if (cg == 0 && row < rows)
                a.opre[bu] = ((s_acc[row] + s_acc[128 + row]) + (s_acc[256 + row] + s_acc[384 + row])) + a.b2[u];
        }
        umma::fence_before_sync();
        __syncthreads();   // accumulator drained, staging buffers free
    }
    if (cur_u >= 0) flush(cur_u);
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, 128);
}

// ---------------------------------------------------------------------------
// Tile images ("mlp_image" layout).  The operand stage of one (unit, 128-row batch tile) in shared memory is
//   [piece 3][chunk chunksQ][row 128] x 16 bytes    (8 bf16 of K = pool columns 8c .. 8c+7 per 16 bytes)
// = 3 * chunksQ * 2048 bytes, contiguous.  The statistics pass F1 -- the first pass over `pool` in a training step --
// builds exactly these bytes for its own MMAs and writes them out with ONE bulk store per tile:
//   image[(u * T + t)]  ,  T = ceil(B / 128)
// Every later pass of the step (F2 apply, B1 backward reductions, B2 dW1 / dy) then fetches a stage with ONE bulk
// load (cp.async.bulk, the TMA engine) that lands on an mbarrier: no global loads, no split arithmetic and no
// shared-memory stores on the LSU path, which the round-1 ablations identified as the bound of these kernels
// (HBM at 16-21 %).  It costs bytes: 6 * Nq / (4 * n) = 1.73x the fp32 tile at n = 83.
// W1 gets the same treatment once per step: [U][3][chunksQ][112 rows] x 16 bytes (w1_image_kernel).
// ---------------------------------------------------------------------------
#define IMG_WROWS 112
#define IMG_E_WARPS 16
#define IMG_THREADS (32 * (IMG_E_WARPS + 2))

__global__ void __launch_bounds__(256)
w1_image_kernel(const float* __restrict__ w1, int U, int n, int chunksQ, uint4* __restrict__ out) {
    const long long total = (long long)U * chunksQ * IMG_WROWS;
    const long long piece = (long long)chunksQ * IMG_WROWS;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(i % IMG_WROWS);
        const int c = (int)((i / IMG_WROWS) % chunksQ);
        const int u = (int)(i / piece);
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int q = 8 * c + e;
            v[e] = (r < MT_H && q < n) ? __ldg(w1 + ((long long)u * MT_H + r) * n + q) : 0.0f;
        }
        uint32_t a[4], b[4], d[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) split2_bf16x3_rn(v[2 * e], v[2 * e + 1], a[e], b[e], d[e]);
        uint4* o = out + (long long)u * 3 * piece + (long long)c * IMG_WROWS + r;
        o[0] = make_uint4(a[0], a[1], a[2], a[3]);
        o[piece] = make_uint4(b[0], b[1], b[2], b[3]);
        o[2 * piece] = make_uint4(d[0], d[1], d[2], d[3]);
    }
}

struct ImgShared {
    uint64_t full[2];        // loader -> issuer / epilogue: the tile's stage (and W1 at a unit boundary) has landed
    uint64_t mma_done[2];    // tcgen05.commit -> epilogue + loader (stage free)
    uint64_t acc_free[2];    // 16 epilogue warps -> issuer
    uint64_t aux_full[4];    // loader -> epilogue (B1): dopre / active bits of the tile are in their ring slot
    uint32_t tmem_slot;
    uint32_t pad;
};

__device__ __forceinline__ void img_epi_barrier() {
    asm volatile("bar.sync 1, %0;" ::"n"(32 * IMG_E_WARPS) : "memory");
}

// F2 (MODE = MLP_APPLY) and B1 (MODE = MLP_BWD_STATS) on tile images:
//   warps 0-15  epilogue (warp e: TMEM lanes 32 (e % 4) .., column group e / 4)
//   warp  16    loader: one elected lane issues the bulk loads two tiles ahead; in B1 the warp also brings the tile's
//               dopre / active bits into a four-deep ring
//   warp  17    issuer
template <int MODE>
__global__ void __launch_bounds__(IMG_THREADS, 1)
mlp_img_kernel(MlpArgs a) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int n = a.n, U = a.U, B = a.B, T = a.T;
    const int Nq = (n + 15) & ~15, chunksQ = Nq >> 3, kstepsQ = Nq >> 4;
    const uint32_t wlbo = IMG_WROWS * 16u;
    const uint32_t wpiece = (uint32_t)chunksQ * wlbo, wbytes = 3u * wpiece;
    const uint32_t ppiece = (uint32_t)chunksQ * 2048u, pstage = 3u * ppiece;
    unsigned char* w1s = smem_raw;
    unsigned char* ps = smem_raw + wbytes;
    unsigned char* tail = smem_raw + wbytes + 2 * pstage;
    float4* s_c4 = reinterpret_cast<float4*>(tail);                              // [128] (b1, scale2, shift2, w2)
    float* s_dopre_all = reinterpret_cast<float*>(tail + 2048);                  // [4][128]
    uint32_t* s_mask_all = reinterpret_cast<uint32_t*>(tail + 2048 + 2048);      // [4][128][4]
    float* s_acc = reinterpret_cast<float*>(tail + 2048 + 2048 + 8192);          // [2][4][128]
    ImgShared* sh = reinterpret_cast<ImgShared*>(tail + 2048 + 2048 + 8192 + 4096);
    uint2* s_tab = reinterpret_cast<uint2*>(tail + 2048 + 2048 + 8192 + 4096 + 128);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (warp == 0) umma::tmem_alloc(&sh->tmem_slot, 256);
    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            umma::mbar_init(&sh->full[s], 1);
            umma::mbar_init(&sh->mma_done[s], 1);
            umma::mbar_init(&sh->acc_free[s], IMG_E_WARPS);
        }
        for (int s = 0; s < 4; ++s) umma::mbar_init(&sh->aux_full[s], 1);
    }
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = sh->tmem_slot;
    const uint32_t w_addr = umma::smem_u32(w1s), p_addr = umma::smem_u32(ps);
    constexpr int NDIM = (MODE == MLP_APPLY) ? 112 : 128;
    const uint32_t idesc = umma::idesc_bf16_f32(128, NDIM);
    if (tid < 6 * kstepsQ) {
        const int pass = tid / kstepsQ, ks = tid - pass * kstepsQ;
        const uint32_t pw = pass == 0 ? 2u : pass == 1 ? 0u : pass == 2 ? 1u : pass == 3 ? 1u : 0u;
        const uint32_t pp = pass == 0 ? 0u : pass == 1 ? 2u : pass == 2 ? 1u : pass == 3 ? 0u : pass == 4 ? 1u : 0u;
        const uint32_t wd = umma::desc_lo(w_addr + pw * wpiece + (uint32_t)(2 * ks) * wlbo, wlbo);
        const uint32_t pd2 = umma::desc_lo(p_addr + pp * ppiece + (uint32_t)(2 * ks) * 2048u, 2048u);
        s_tab[tid] = (MODE == MLP_APPLY) ? make_uint2(pd2, wd) : make_uint2(wd, pd2);   // (A, B)
    }
    __syncthreads();

    const long long tiles_total = (long long)U * T;
    const long long lo = (long long)blockIdx.x * tiles_total / gridDim.x;
    const long long hi = (long long)(blockIdx.x + 1) * tiles_total / gridDim.x;
    const int ntl = (int)(hi - lo);
    const float inv_keep = 1.0f / (1.0f - a.drop_p);
    const int C = U * MT_H;
    const unsigned long long seed = a.seed + (a.seed_dev ? *a.seed_dev * 0x9E3779B97F4A7C15ull : 0ull);

    if (warp == IMG_E_WARPS) {
        // ================= loader =================
        int cur_u = -1;
        for (int i = 0; i < ntl; ++i) {
            const long long tile = lo + i;
            const int u = (int)(tile / T), t = (int)(tile % T);
            const uint32_t st = (uint32_t)i & 1u;
            if (i >= 2) umma::mbar_wait(&sh->mma_done[st], (uint32_t)((i >> 1) - 1) & 1u);   // pool stage free
            const bool new_u = (u != cur_u);
            if (new_u && i >= 1) umma::mbar_wait(&sh->mma_done[st ^ 1u], (uint32_t)((i - 1) >> 1) & 1u);   // W1 free
            cur_u = u;
            if (MODE == MLP_BWD_STATS) {
                // dopre / active bits of the tile's 128 batch rows: 4 rows per lane
                const int b0 = t * MT_ROWS;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int r = lane + 32 * j;
                    const bool ok = b0 + r < B;
                    const long long bu = (long long)(b0 + r) * U + u;
                    s_dopre_all[(i & 3) * 128 + r] = ok ? __ldg(a.dopre + bu) : 0.0f;
                    *reinterpret_cast<uint4*>(s_mask_all + (i & 3) * 512 + 4 * r) =
                        ok ? __ldg(reinterpret_cast<const uint4*>(a.maskbits + bu * 4)) : make_uint4(0u, 0u, 0u, 0u);
                }
                __syncwarp();
                // ring slot i & 3 is rewritten for tile i + 4, whose load waits for the MMAs of tile i + 2, which
                // wait for the epilogue of tile i to hand its accumulator stage back: never more than one phase ahead
                if (lane == 0) umma::mbar_arrive(&sh->aux_full[i & 3]);
            }
            if (lane == 0) {
                umma::mbar_arrive_expect_tx(&sh->full[st], pstage + (new_u ? wbytes : 0u));
                if (new_u) umma::bulk_load(w1s, a.w1_image + (size_t)u * wbytes, wbytes, &sh->full[st]);
                umma::bulk_load(ps + (size_t)st * pstage, a.image + (size_t)tile * pstage, pstage, &sh->full[st]);
            }
            __syncwarp();
        }
    } else if (warp == IMG_E_WARPS + 1) {
        // ================= issuer: one elected lane per tile, descriptors in the uniform datapath =================
        const uint32_t w_base = umma::warp_uniform(umma::desc_lo(w_addr, wlbo));
        const uint32_t p_base = umma::warp_uniform(umma::desc_lo(p_addr, 2048u));
        const uint32_t tmem_u = umma::warp_uniform(tmem);
        for (int i = 0; i < ntl; ++i) {
            const uint32_t st = (uint32_t)i & 1u;
            umma::mbar_wait(&sh->full[st], ((uint32_t)i >> 1) & 1u);
            umma::mbar_wait(&sh->acc_free[st], (((uint32_t)i >> 1) & 1u) ^ 1u);
            umma::fence_after_sync();
            if (umma::elect_one()) {
                const uint32_t pst = p_base + st * (pstage >> 4);
                if (MODE == MLP_APPLY)      // A = pool tile, B = W1
                    umma::issue_split6(pst, ppiece >> 4, 256u, umma::desc_hi(128u), w_base, wpiece >> 4, (2u * wlbo) >> 4,
                                       umma::desc_hi(128u), kstepsQ, tmem_u + st * 128u, idesc);
                else                        // A = W1, B = pool tile
                    umma::issue_split6(w_base, wpiece >> 4, (2u * wlbo) >> 4, umma::desc_hi(128u), pst, ppiece >> 4, 256u,
                                       umma::desc_hi(128u), kstepsQ, tmem_u + st * 128u, idesc);
                umma::commit(&sh->mma_done[st]);
            }
            __syncwarp();
        }
    } else {
        // ================= epilogue (16 warps) =================
        const int quarter = warp & 3, cg = warp >> 2;
        const int row = quarter * 32 + lane;
        int cur_u = -1;
        float b1m = 0.f, scm = 0.f, shm = 0.f, w2m = 0.f;
        double s1 = 0.0, s2 = 0.0, s3 = 0.0, s4 = 0.0;
        auto flush = [&](int u) {
            if (MODE == MLP_APPLY) return;
            int j = blockIdx.x;
            const long long start = (long long)u * T;
            while (j > 0 && ((long long)j * tiles_total) / gridDim.x > start) --j;
            const int slot = (blockIdx.x - j) * 4 + cg;
            if (row < MT_H && slot < a.slots) {
                const int c = u * MT_H + row;
                double* o = a.part + (long long)slot * 4 * C + c;
                o[0] = s1; o[(long long)C] = s2; o[2LL * C] = s3; o[3LL * C] = s4;
            }
            s1 = s2 = s3 = s4 = 0.0;
        };
        for (int i = 0; i < ntl; ++i) {
            const long long tile = lo + i;
            const int u = (int)(tile / T), t = (int)(tile % T);
            const int b0 = t * MT_ROWS, rows = min(MT_ROWS, B - b0);
            const uint32_t st = (uint32_t)i & 1u;
            if (u != cur_u) {
                if (cur_u >= 0) flush(cur_u);
                cur_u = u;
                const int ch = u * MT_H + row;
                const bool ok = row < MT_H;
                b1m = ok ? a.b1[ch] : 0.f;
                const float sc = ok ? a.scale2[ch] : 0.f, shf = ok ? a.shift2[ch] : 0.f;
                // folded per-channel constants: BN2(acc + b1) / keep = acc * scm + shm;  w2m = w2 / keep
                scm = sc * inv_keep; shm = fmaf(b1m, sc, shf) * inv_keep; w2m = ok ? a.w2[ch] * inv_keep : 0.f;
                if (MODE == MLP_APPLY) {
                    img_epi_barrier();          // every epilogue warp is done with the previous unit's constants
                    // (scale, shift') of the UNSCALED pre-activation (its sign is the ReLU decision), w2 / keep
                    if (cg == 0) s_c4[row] = make_float4(sc, fmaf(b1m, sc, shf), w2m, 0.f);
                    img_epi_barrier();
                }
            }
            if (MODE == MLP_BWD_STATS) umma::mbar_wait(&sh->aux_full[i & 3], ((uint32_t)i >> 2) & 1u);   // the loader's ring writes
            umma::mbar_wait(&sh->mma_done[st], ((uint32_t)i >> 1) & 1u);
            umma::fence_after_sync();
            const uint32_t trow = tmem + st * 128u + ((uint32_t)(quarter * 32) << 16);
            if (MODE == MLP_BWD_STATS) {
                // thread = (hidden channel, quarter of the tile's batch rows)
                const float* s_dopre = s_dopre_all + (i & 3) * 128;
                const uint32_t* s_mask = s_mask_all + (i & 3) * 512;
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    const int c0 = 32 * cg + 16 * hh;
                    if (c0 < rows) {
                        uint32_t v[16];
                        umma::tmem_ld16(trow + c0, v);
                        umma::tmem_ld_wait();
                        // t = dopre where the channel was active: sum g2 = w2m * sum t, sum g2 * x = w2m * (sum t * acc +
                        // b1 * sum t), sum dopre * h = sum t * (acc * scm + shm)
                        float f1 = 0.0f, f2 = 0.0f, f3 = 0.0f, f4 = 0.0f;
                        const uint32_t bit = 1u << (row & 31);
                        const int wsel = row >> 5;
#pragma unroll
                        for (int k2 = 0; k2 < 16; ++k2) {
                            if (rows == MT_ROWS || c0 + k2 < rows) {
                                const float acc = __uint_as_float(v[k2]);
                                const float d = s_dopre[c0 + k2];
                                const float t = (s_mask[4 * (c0 + k2) + wsel] & bit) ? d : 0.0f;
                                f1 += t;
                                f2 = fmaf(t, acc, f2);
                                f3 = fmaf(t, fmaf(acc, scm, shm), f3);
                                f4 += d;
                            }
                        }
                        s1 += (double)(w2m * f1); s2 += (double)(w2m * fmaf(b1m, f1, f2)); s3 += (double)f3; s4 += (double)f4;
                    }
                }
                umma::fence_before_sync();
                __syncwarp();
                if (lane == 0) umma::mbar_arrive(&sh->acc_free[st]);
            } else {
                // thread = (batch row, 32 hidden channels = keep-bit word cg)
                const long long bu = (long long)(b0 + row) * U + u;
                const uint32_t kbw = (row < rows) ? dropout_keep_word(a.drop_p, seed, bu, a.mask_u8, cg) : 0u;
                uint32_t actw = 0u;   // keep AND (BN2 output > 0): the backward passes reuse THIS decision
                float accv = 0.0f;
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    const int c0 = 32 * cg + 16 * hh;
                    if (c0 < 112) {
                        uint32_t v[16];
                        umma::tmem_ld16(trow + c0, v);
                        umma::tmem_ld_wait();
#pragma unroll
                        for (int k2 = 0; k2 < 16; ++k2) {
                            const int m = c0 + k2;
                            if (m < MT_H) {
                                const float4 c4 = s_c4[m];
                                const float pre = fmaf(__uint_as_float(v[k2]), c4.x, c4.y);
                                const uint32_t act = ((kbw >> (16 * hh + k2)) & 1u) & (pre > 0.0f ? 1u : 0u);
                                actw |= act << (16 * hh + k2);
                                accv = fmaf(act ? pre : 0.0f, c4.z, accv);
                            }
                        }
                    }
                }
                umma::fence_before_sync();
                __syncwarp();
                if (lane == 0) umma::mbar_arrive(&sh->acc_free[st]);
                float* acc4 = s_acc + st * 512;              // [tile parity][column group][row]
                acc4[cg * 128 + row] = accv;
                if (row < rows && a.maskbits) a.maskbits[bu * 4 + cg] = actw;
                img_epi_barrier();
                if (cg == 0 && row < rows)
                    a.opre[bu] = ((acc4[row] + acc4[128 + row]) + (acc4[256 + row] + acc4[384 + row])) + a.b2[u];
            }
        }
        if (cur_u >= 0) flush(cur_u);
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, 256);
}

static size_t img_smem(int n) {
    const size_t chunksQ = (size_t)(((n + 15) & ~15) >> 3);
    return 3 * chunksQ * IMG_WROWS * 16 + 2 * 3 * chunksQ * 2048 + 2048 + 2048 + 8192 + 4096 + 128 + 6 * 6 * 8 + 16;
}

namespace xnn {
int launch_bn_train_finalize(const double* part, int tiles, int B, int C, const float* gamma, const float* beta,
                             float eps, float momentum, float* running_mean, float* running_var, float* scale,
                             float* shift, double* stats, cudaStream_t st);
}

// ---- launch geometry shared by the passes ---------------------------------------------------------
static int mt_tiles(int B) { return (B + MT_ROWS - 1) / MT_ROWS; }
static int mt_grid(int B, int U) {
    const long long tiles = (long long)U * mt_tiles(B);
    const int sms = num_sms();
    return (int)(tiles < sms ? tiles : sms);
}
// upper bound on the number of CTAs whose range touches one unit
static int mt_slots(int B, int U) {
    const int T = mt_tiles(B), G = mt_grid(B, U);
    const long long per = ((long long)U * T) / G;   // >= 1
    return 4 * ((int)((T + per - 1) / per) + 1);    // x4: one partial per column group of the CTA
}
static size_t mt_smem(int n) {
    const size_t chunksQ = (size_t)(((n + 15) & ~15) >> 3);
    return 3 * chunksQ * WS_WROWS * 16 + 2 * 3 * chunksQ * 2048 + 2048 + 2048 + 8192 + 2048 + 64 + 6 * 6 * 8 + 16;
}

static size_t mt_smem_lockstep(int n) {
    const int chunks = ((n + 7) & ~7) >> 2;
    return (size_t)4 * chunks * 2048 + 2048 + 512 + 2048 + 2048 + 32 + 3 * 12 * 8 + 16;
}

template <int MODE>
static int mt_launch(const MlpArgs& a, cudaStream_t st) {
    if (MODE != MLP_STATS && a.image != nullptr) {
        // tile images present (training step): operand stages by bulk load, 16 epilogue warps
        constexpr int M2 = (MODE == MLP_STATS) ? MLP_APPLY : MODE;
        const size_t smem = img_smem(a.n);
        if (cudaFuncSetAttribute(mlp_img_kernel<M2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return 1;
        mlp_img_kernel<M2><<<mt_grid(a.B, a.U), IMG_THREADS, smem, st>>>(a);
        return cudaGetLastError() != cudaSuccess;
    }
    if (MODE == MLP_BWD_STATS) {      // measured with the warp-specialised kernel instead: 1.46 vs 1.40 ms for xnn_mlp_bwd
        const size_t smem = mt_smem_lockstep(a.n);
        if (cudaFuncSetAttribute(mlp_fc1_lockstep_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return 1;
        mlp_fc1_lockstep_kernel<MODE><<<mt_grid(a.B, a.U), MT_THREADS, smem, st>>>(a);
        return cudaGetLastError() != cudaSuccess;
    }
    const size_t smem = mt_smem(a.n);
    constexpr int PW = (MODE == MLP_APPLY) ? 8 : 16;       // measured: apply 0.34 (8) vs 0.39 ms (16); statistics 0.288 vs 0.278
    if (cudaFuncSetAttribute(mlp_fc1_kernel<MODE, PW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
        return 1;
    mlp_fc1_kernel<MODE, PW><<<mt_grid(a.B, a.U), 32 * (PW + WS_E_WARPS + 1), smem, st>>>(a);
    return cudaGetLastError() != cudaSuccess;
}

extern "C" int xnn_mlp_supported(int n) { return n >= 1 && n <= MT_NMAX; }

// tile images of one training step: [U][T] pool stages, then [U] W1 images (mlp_image layout, see mlp_img_kernel)
static size_t img_stage_bytes(int n) { return (size_t)3 * (((n + 15) & ~15) >> 3) * 2048; }
static size_t img_w1_bytes(int n) { return (size_t)3 * (((n + 15) & ~15) >> 3) * IMG_WROWS * 16; }
extern "C" size_t xnn_mlp_image_bytes(int B, int U, int n) {
    if (B < 1 || U < 1 || !xnn_mlp_supported(n)) return 0;
    return (size_t)U * mt_tiles(B) * img_stage_bytes(n) + (size_t)U * img_w1_bytes(n);
}
static const unsigned char* img_w1_ptr(const void* image, int B, int U, int n) {
    return (const unsigned char*)image + (size_t)U * mt_tiles(B) * img_stage_bytes(n);
}

extern "C" size_t xnn_mlp_bn2_work_bytes(int B, int U) {
    if (B < 1 || U < 1) return 0;
    return (size_t)mt_slots(B, U) * 2 * U * MT_H * sizeof(double);
}

extern "C" int xnn_mlp_bn2_train_prepare(const float* pool, const float* w1, const float* b1, int B, int U, int n,
                                         const float* gamma, const float* beta, float eps, float momentum,
                                         float* running_mean, float* running_var, float* scale, float* shift,
                                         double* stats, void* work, void* image, int phase, long long count,
                                         xnn_stream_t s) {
    XNN_CHECK_ARG(pool && w1 && b1 && gamma && beta && scale && shift && stats && work, "NULL pointer");
    XNN_CHECK_ARG(((uintptr_t)image & 127) == 0, "image must be 128-byte aligned");
    XNN_CHECK_ARG(phase >= 0 && phase <= 2, "phase: 0 = whole call, 1 = partial sums, 2 = finalize");
    if (phase == 0) count = B;
    XNN_CHECK_ARG(B >= 1 && U >= 1 && count >= 2 && count < (1LL << 31),
                  "train-mode BatchNorm needs more than 1 value per channel");
    XNN_CHECK_ARG(xnn_mlp_supported(n), "pool width outside the tensor-core path (n <= 96)");
    cudaStream_t st = to_stream(s);
    MlpArgs a{};
    a.pool = pool; a.w1 = w1; a.b1 = b1; a.B = B; a.U = U; a.n = n; a.T = mt_tiles(B); a.slots = mt_slots(B, U);
    a.part = (double*)work;
    a.image_out = (unsigned char*)image;
    if (phase != 2) {
        XNN_CUDA(cudaMemsetAsync(work, 0, xnn_mlp_bn2_work_bytes(B, U), st));
        if (image != nullptr) {
            const int chunksQ = ((n + 15) & ~15) >> 3;
            const long long items = (long long)U * chunksQ * IMG_WROWS;
            w1_image_kernel<<<ceil_div(items, 256), 256, 0, st>>>(w1, U, n, chunksQ,
                                                                  (uint4*)const_cast<unsigned char*>(img_w1_ptr(image, B, U, n)));
            XNN_CHECK_LAUNCH();
        }
        if (mt_launch<MLP_STATS>(a, st)) { set_error("xnn_mlp_bn2_train_prepare: launch failed: %s", cudaGetErrorString(cudaGetLastError())); return XNN_ERR_CUDA; }
    }
    if (phase != 1 && launch_bn_train_finalize(a.part, a.slots, (int)count, U * MT_H, gamma, beta, eps, momentum,
                                               running_mean, running_var, scale, shift, stats, st)) {
        set_error("xnn_mlp_bn2_train_prepare: finalize launch failed");
        return XNN_ERR_CUDA;
    }
    return XNN_OK;
}

extern "C" int xnn_mlp_fwd(const float* pool, const float* w1, const float* b1, const float* scale2,
                           const float* shift2, const float* w2, const float* b2, int B, int U, int n, float drop_p,
                           unsigned long long seed, const unsigned long long* seed_dev, const uint8_t* mask,
                           uint32_t* maskbits, float* opre, const void* image, xnn_stream_t s) {
    XNN_CHECK_ARG(pool && w1 && b1 && scale2 && shift2 && w2 && b2 && opre, "NULL pointer");
    XNN_CHECK_ARG(B >= 0 && U >= 1 && drop_p >= 0.0f && drop_p < 1.0f, "bad sizes");
    XNN_CHECK_ARG(xnn_mlp_supported(n), "pool width outside the tensor-core path (n <= 96)");
    XNN_CHECK_ARG(maskbits == nullptr || ((uintptr_t)maskbits & 15) == 0, "maskbits must be 16-byte aligned");
    if (B == 0) return XNN_OK;
    MlpArgs a{};
    a.pool = pool; a.w1 = w1; a.b1 = b1; a.scale2 = scale2; a.shift2 = shift2; a.w2 = w2; a.b2 = b2;
    a.B = B; a.U = U; a.n = n; a.T = mt_tiles(B); a.slots = 0;
    a.drop_p = drop_p; a.seed = seed; a.seed_dev = seed_dev; a.mask_u8 = mask; a.maskbits = maskbits; a.opre = opre;
    XNN_CHECK_ARG(((uintptr_t)image & 127) == 0, "image must be 128-byte aligned");
    if (image != nullptr) { a.image = (const unsigned char*)image; a.w1_image = img_w1_ptr(image, B, U, n); }
    if (mt_launch<MLP_APPLY>(a, to_stream(s))) { set_error("xnn_mlp_fwd: launch failed: %s", cudaGetErrorString(cudaGetLastError())); return XNN_ERR_CUDA; }
    return XNN_OK;
}

// ---------------------------------------------------------------------------
// backward, pass B1: reductions of the BatchNorm (linears.7) backward and dW2 / db2
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
mlp_bwd_finalize_kernel(const double* __restrict__ part, int slots, long long B, int U,
                        const double* __restrict__ stats, float* __restrict__ dgamma, float* __restrict__ dbeta,
                        float* __restrict__ dw2, float* __restrict__ db2, float2* __restrict__ coef, float out_scale) {
    const int C = U * MT_H;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double s1 = 0.0, s2 = 0.0, s3 = 0.0, s4 = 0.0;
    for (int t = 0; t < slots; ++t) {
        const double* p = part + (long long)t * 4 * C + c;
        s1 += p[0]; s2 += p[(long long)C]; s3 += p[2LL * C]; s4 += p[3LL * C];
    }
    const double mean = stats[4LL * c], invstd = stats[4LL * c + 2];
    const double dg = invstd * (s2 - mean * s1);       // sum g * xhat
    const double os = (double)out_scale;
    dgamma[c] = (float)(dg * os);
    dbeta[c] = (float)(s1 * os);
    dw2[c] = (float)(s3 * os);
    if (c % MT_H == 0) db2[c / MT_H] = (float)(s4 * os);
    coef[c] = make_float2((float)(s1 / (double)B), (float)(dg / (double)B));
}

// ---------------------------------------------------------------------------
// backward, pass B2 (thread = hidden channel, 128 batch rows per tile -- N = 128 MMAs: the shared-memory operand
// fetch of the 126 MMAs per tile bounds this kernel, and 64-row tiles fetched a third more per row):
//   D1[m][b]  = W1 . pool^T                 -> dh[m][b] = gamma*invstd*(g2 - c1 - xhat*c2)   (registers -> smem)
//   D2[q][b]  = W1^T . dh                   -> dy[b,u,q] = pool[b,u,q] * D2[q][b]
//   D3[m][q] += dh . pool                   (accumulated in TMEM over the CTA's tiles of the unit) -> dW1
// The transposed operands are MN-major views of the same staging buffers, which tcgen05 offers without
// swizzle only for 16-bit types: this pass therefore runs kind::f16 with a THREE-way bf16 split of both
// operands (x = x1+x2+x3, 8 mantissa bits each) and the six products x1y1 x1y2 x2y1 x1y3 x2y2 x3y1 --
// the same fp32-level accuracy as 3xTF32 at the same tensor time, and 6 instead of 8 bytes per element.
// Staging: an [R rows][K] matrix as uint4 [K/8][R] (8 bf16 per 16 bytes), one buffer per piece; W1 and dh use
// R = 112 (100 hidden channels, zero rows up to the 7th K step of 16), the pool tile R = 128.
// ---------------------------------------------------------------------------
#define MB_ROWS 128
#define MB_WROWS 112      // rows per 16-byte chunk column of the W1 / dh staging: 100 hidden channels, 7 K steps of 16

struct MbShared {
    uint64_t bar;
    uint32_t tmem_slot;
    uint32_t pad[5];
};

struct MlpBwdArgs {
    const float* pool; const float* w1; const float* b1; const float* scale2; const float* shift2;
    const float* w2; const float* gamma2; const double* stats2; const float2* coef;
    const float* dopre; const uint32_t* maskbits;
    float* dy;            // (B,U,n)
    float* part_w1;       // [slots][100U][n]
    double* part_db1;     // [slots][100U]
    int B, U, n, T, slots;
    int P;                // unit interleave (mb_interleave): CTAs that run side by side work on adjacent units
    float drop_p;
};

// 8 floats -> three uint4 of packed bf16 pieces (hi, mid, lo), truncating split (split2_bf16x3)
__device__ __forceinline__ void split8_bf16x3(const float* v, uint4& p0, uint4& p1, uint4& p2) {
    uint32_t w0[4], w1[4], w2[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) split2_bf16x3(v[2 * i], v[2 * i + 1], w0[i], w1[i], w2[i]);
    p0 = make_uint4(w0[0], w0[1], w0[2], w0[3]);
    p1 = make_uint4(w1[0], w1[1], w1[2], w1[3]);
    p2 = make_uint4(w2[0], w2[1], w2[2], w2[3]);
}

// stage chunks [cb, ce) of one fp32 row (n values, zero beyond) into the three piece buffers
__device__ __forceinline__ void stage_row_bf16x3(const float* __restrict__ src, int n, int cb, int ce, int R, int row,
                                                 uint4* __restrict__ d0, uint32_t piece_u4) {
#pragma unroll 2
    for (int c = cb; c < ce; ++c) {
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int q = 8 * c + e;
            v[e] = (src != nullptr && q < n) ? __ldg(src + q) : 0.0f;
        }
        uint4 p0, p1, p2;
        split8_bf16x3(v, p0, p1, p2);
        d0[c * R + row] = p0;
        d0[piece_u4 + c * R + row] = p1;
        d0[2 * piece_u4 + c * R + row] = p2;
    }
}

__global__ void __launch_bounds__(MT_THREADS, 1)
mlp_bwd_kernel(MlpBwdArgs a) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int n = a.n, U = a.U, B = a.B, T = a.T;
    const int Nq = (n + 15) & ~15, chunksQ = Nq >> 3, kstepsQ = Nq >> 4;
    const uint32_t wcol = MB_WROWS * 16u;                  // one 16-byte chunk column of W1 / dh: 112 rows
    const uint32_t wpiece = (uint32_t)chunksQ * wcol;      // [chunksQ][112 rows] uint4
    const uint32_t pcol = MB_ROWS * 16u;                   // one chunk column of the pool tile: 128 batch rows
    const uint32_t ppiece = (uint32_t)chunksQ * pcol;      // [chunksQ][128 rows] uint4
    const uint32_t tpiece = (MB_ROWS / 8) * wcol;          // [16][112 rows] uint4 (K = 128 batch rows)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    uint4* w1s = reinterpret_cast<uint4*>(smem_raw);
    uint4* ps = reinterpret_cast<uint4*>(smem_raw + 3 * wpiece);
    uint4* ts = reinterpret_cast<uint4*>(smem_raw + 3 * wpiece + 3 * ppiece);
    // (the MN-major view of W1 with M = 128 reads 16 chunk columns from each piece: 2 wpiece + 16 wcol stays
    //  inside 3 wpiece + 3 ppiece + 3 tpiece for every n)
    unsigned char* tail = smem_raw + 3 * wpiece + 3 * ppiece + 3 * tpiece;
    float* s_dopre = reinterpret_cast<float*>(tail);                                     // [128]
    uint32_t* s_mask = reinterpret_cast<uint32_t*>(tail + 512);                          // [4 words][128 batch rows]
    MbShared* sh = reinterpret_cast<MbShared*>(tail + 512 + 2048);
    uint2* s_tab = reinterpret_cast<uint2*>(tail + 512 + 2048 + 32);                     // [36 + 42 + 48] descriptor words

    const int row = (warp & 3) * 32 + lane;   // TMEM lane: hidden channel m (D1, D3) / pool column q (D2)
    const int cg = warp >> 2;                 // column group: 32 of the tile's 128 batch rows
    {
        uint4* z = reinterpret_cast<uint4*>(smem_raw);
        const uint32_t tot = (uint32_t)(tail - smem_raw) / 16;
        for (uint32_t i = tid; i < tot; i += MT_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    if (warp == 0) umma::tmem_alloc(&sh->tmem_slot, 512);
    if (tid == 0) umma::mbar_init(&sh->bar, 1);
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = sh->tmem_slot;
    const uint32_t idesc1 = umma::idesc_bf16_f32(128, MB_ROWS);                          // W1 x pool, both K-major
    const uint32_t idesc2 = umma::idesc_bf16_f32(128, MB_ROWS) | (1u << 15) | (1u << 16); // W1^T x dh^T (MN views)
    const uint32_t idesc3 = umma::idesc_bf16_f32(128, Nq) | (1u << 16);                  // dh x pool^T (MN view)
    const uint32_t D1 = tmem, D2 = tmem + MB_ROWS, D3 = tmem + 2 * MB_ROWS;
    // the six products of the 3-way split, smallest first: (piece of A, piece of B), 0 = hi, 1 = mid, 2 = lo
    // MMA sequences of a tile (the operand buffers never move):
    //   T1 [6*kstepsQ]  D1 = W1 (K-major) x pool (K-major)
    //   T3 [42]         D2 = W1^T (MN view) x dh^T (MN view), K = hidden rows, 16 per MMA (7 steps = 112 rows)
    //   T4 [48]         D3 += dh (K-major, K = batch rows) x pool^T (MN view)
    const int n1 = 6 * kstepsQ, n4 = 6 * (MB_ROWS / 16), ntab = n1 + 42 + n4;
    if (tid < ntab) {
        const int ti = tid;
        const uint32_t w_addr = umma::smem_u32(smem_raw);
        const uint32_t p_addr = w_addr + 3 * wpiece, t_addr = p_addr + 3 * ppiece;
        const int seq = ti < n1 ? 0 : (ti < n1 + 42 ? 1 : 2);
        const int i = ti - (seq == 0 ? 0 : (seq == 1 ? n1 : n1 + 42));
        const int per = seq == 0 ? kstepsQ : (seq == 1 ? 7 : MB_ROWS / 16);
        const int pass = i / per, ks = i - pass * per;
        const uint32_t pa = pass == 0 ? 2u : pass == 1 ? 0u : pass == 2 ? 1u : pass == 3 ? 1u : 0u;
        const uint32_t pb = pass == 0 ? 0u : pass == 1 ? 2u : pass == 2 ? 1u : pass == 3 ? 0u : pass == 4 ? 1u : 0u;
        uint2 e;
        if (seq == 0)
            e = make_uint2(umma::desc_lo(w_addr + pa * wpiece + (uint32_t)(2 * ks) * wcol, wcol),
                           umma::desc_lo(p_addr + pb * ppiece + (uint32_t)(2 * ks) * pcol, pcol));
        else if (seq == 1)
            e = make_uint2(umma::desc_lo(w_addr + pa * wpiece + (uint32_t)ks * 256u, 128u),
                           umma::desc_lo(t_addr + pb * tpiece + (uint32_t)ks * 256u, 128u));
        else
            e = make_uint2(umma::desc_lo(t_addr + pa * tpiece + (uint32_t)(2 * ks) * wcol, wcol),
                           umma::desc_lo(p_addr + pb * ppiece + (uint32_t)ks * 256u, 128u));
        s_tab[ti] = e;
    }
    const uint2* tab = s_tab;
    uint64_t* bar = &sh->bar;

    const long long tiles_total = (long long)U * T;
    const long long lo = (long long)blockIdx.x * tiles_total / gridDim.x;
    const long long hi = (long long)(blockIdx.x + 1) * tiles_total / gridDim.x;
    const float inv_keep = 1.0f / (1.0f - a.drop_p);
    const int C = U * MT_H;
    const uint32_t trow = (uint32_t)((warp & 3) * 32) << 16;
    // staging map (coalesced): a warp-wide access covers 8 rows x 32 B of global memory and 32 distinct banks:
    // lane = (row & 7) * 4 + word, word = two consecutive pool columns of a 16-byte (8 x bf16) chunk.
    // pool tile: warp w -> rows 8 w .. 8 w + 7, every chunk
    const int srow = warp * 8 + (lane >> 2), swd = lane & 3;

    uint32_t phase = 0, d3_started = 0;
    float b1m = 0.f, w2m = 0.f, meanm = 0.f, invstdm = 0.f, gism = 0.f, c1m = 0.f, c2m = 0.f;
    double sdb1 = 0.0;
    float pv[MT_NMAX / 8][2];
    float pd = 0.0f;                                   // prefetched dopre / active bits of batch row tid (< 128)
    uint4 pm = make_uint4(0u, 0u, 0u, 0u);

    auto prefetch = [&](long long tile, int u, bool live) {
        const int t = live ? (int)(tile % T) : 0;
        const int b0 = t * MB_ROWS;
        const float* src = (live && b0 + srow < B) ? a.pool + ((long long)(b0 + srow) * U + u) * n : nullptr;
#pragma unroll
        for (int i = 0; i < MT_NMAX / 8; ++i) {
            const int q = 8 * i + 2 * swd;
            pv[i][0] = (src != nullptr && q < n) ? __ldg(src + q) : 0.0f;
            pv[i][1] = (src != nullptr && q + 1 < n) ? __ldg(src + q + 1) : 0.0f;
        }
        if (tid < MB_ROWS) {
            const bool ok = live && b0 + tid < B;
            const long long bu = (long long)(b0 + tid) * U + u;
            pd = ok ? __ldg(a.dopre + bu) : 0.0f;
            pm = ok ? __ldg(reinterpret_cast<const uint4*>(a.maskbits + bu * 4)) : make_uint4(0u, 0u, 0u, 0u);
        }
    };
    auto store_tile = [&]() {
        uint32_t* pw = reinterpret_cast<uint32_t*>(ps);
#pragma unroll
        for (int i = 0; i < MT_NMAX / 8; ++i) {
            if (i < chunksQ) {
                uint32_t w0, w1, w2;
                split2_bf16x3(pv[i][0], pv[i][1], w0, w1, w2);
                const int o = (i * MB_ROWS + srow) * 4 + swd;
                pw[o] = w0; pw[ppiece / 4 + o] = w1; pw[2 * (ppiece / 4) + o] = w2;
            }
        }
        if (tid < MB_ROWS) {
            s_dopre[tid] = pd;
            s_mask[tid] = pm.x; s_mask[MB_ROWS + tid] = pm.y; s_mask[2 * MB_ROWS + tid] = pm.z; s_mask[3 * MB_ROWS + tid] = pm.w;
        }
    };
    auto flush = [&](int v, int u) {
        // every MMA into D3 has completed (the tile loop waits on each commit)
        int j = blockIdx.x;
        const long long start = (long long)v * T;
        while (j > 0 && ((long long)j * tiles_total) / gridDim.x > start) --j;
        const int slot = blockIdx.x - j;
        const bool ok = row < MT_H && slot < a.slots;
        float* o = a.part_w1 + ((long long)slot * C + (long long)u * MT_H + row) * n;
        for (int c0 = 16 * cg; c0 < Nq; c0 += 64) {
            uint32_t v[16];
            umma::tmem_ld16(D3 + trow + c0, v);
            umma::tmem_ld_wait();
            if (ok) {
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    if (c0 + i < n) o[c0 + i] = d3_started ? __uint_as_float(v[i]) : 0.0f;
            }
        }
        if (ok) a.part_db1[((long long)slot * 4 + cg) * C + (long long)u * MT_H + row] = sdb1;
        sdb1 = 0.0;
        d3_started = 0;
        umma::fence_before_sync();
    };

    for (long long seg = lo; seg < hi;) {
        // position v in the CTA order -> unit u: class v % P holds a contiguous run of units, so the CTAs that are
        // at the same batch rows at the same time read and write neighbouring 4n-byte segments of those rows
        const int v = (int)(seg / T);
        int u = v / a.P;
        for (int c = 0; c < v % a.P; ++c) u += (U - c + a.P - 1) / a.P;
        const long long seg_end = min(hi, (long long)(v + 1) * T);
        __syncthreads();            // the previous unit's flush is done
        {
            // W1 of this unit, same coalesced map: items (row group of 8, chunk) dealt round-robin to the warps
            uint32_t* ww = reinterpret_cast<uint32_t*>(w1s);
            for (int it = warp; it < 13 * chunksQ; it += MT_THREADS / 32) {
                const int g8 = it / chunksQ, c = it - g8 * chunksQ;
                const int r = g8 * 8 + (lane >> 2);
                if (r < MT_H) {
                    const float* src = a.w1 + ((long long)u * MT_H + r) * n;
                    const int q = 8 * c + 2 * swd;
                    uint32_t w0, w1v, w2v;
                    split2_bf16x3(q < n ? __ldg(src + q) : 0.0f, q + 1 < n ? __ldg(src + q + 1) : 0.0f, w0, w1v, w2v);
                    const int o = (c * MB_WROWS + r) * 4 + swd;
                    ww[o] = w0; ww[wpiece / 4 + o] = w1v; ww[2 * (wpiece / 4) + o] = w2v;
                }
            }
            const int ch = u * MT_H + row;
            const bool ok = row < MT_H;
            b1m = ok ? a.b1[ch] : 0.f;
            w2m = ok ? a.w2[ch] : 0.f;
            meanm = ok ? (float)a.stats2[4LL * ch] : 0.f; invstdm = ok ? (float)a.stats2[4LL * ch + 2] : 0.f;
            gism = ok ? (float)((double)a.gamma2[ch] * a.stats2[4LL * ch + 2]) : 0.f;
            const float2 cf = ok ? a.coef[ch] : make_float2(0.f, 0.f);
            c1m = cf.x; c2m = cf.y;
        }
        prefetch(seg, u, true);
        umma::fence_proxy_async();
        __syncthreads();

        for (long long tile = seg; tile < seg_end; ++tile) {
            const int t = (int)(tile % T);
            const int b0 = t * MB_ROWS, rows = min(MB_ROWS, B - b0);
            store_tile();
            prefetch(tile + 1, u, tile + 1 < seg_end);
            umma::fence_proxy_async();
            __syncthreads();
            if (tid == 0) {
                umma::fence_after_sync();
                umma::issue_table<false>(tab, n1, umma::desc_hi(128u), umma::desc_hi(128u), 0u, D1, idesc1, 0u);
                umma::commit(bar);
            }
            umma::mbar_wait(bar, phase);
            phase ^= 1u;
            umma::fence_after_sync();

            // ---- epilogue 1: dh for (channel row, batch rows 32cg..32cg+31) -> four 16-byte chunks per piece
            {
                uint32_t v[2][16];
                umma::tmem_ld16(D1 + trow + 32 * cg, v[0]);
                umma::tmem_ld16(D1 + trow + 32 * cg + 16, v[1]);
                umma::tmem_ld_wait();
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int c0 = 32 * cg + 16 * h;
                    float dh[16];
                    float part = 0.0f;
                    // mask words of this warp's 32 channels and dopre for 16 batch rows: four 16-byte loads each
                    uint32_t mk[16];
                    float dp[16];
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
                        const uint4 mq = *reinterpret_cast<const uint4*>(s_mask + (row >> 5) * MB_ROWS + c0 + 4 * q4);
                        const float4 dq = *reinterpret_cast<const float4*>(s_dopre + c0 + 4 * q4);
                        mk[4 * q4] = mq.x; mk[4 * q4 + 1] = mq.y; mk[4 * q4 + 2] = mq.z; mk[4 * q4 + 3] = mq.w;
                        dp[4 * q4] = dq.x; dp[4 * q4 + 1] = dq.y; dp[4 * q4 + 2] = dq.z; dp[4 * q4 + 3] = dq.w;
                    }
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const int c = c0 + i;
                        const float x = __uint_as_float(v[h][i]) + b1m;
                        const uint32_t act = (mk[i] >> (row & 31)) & 1u;
                        const float g2 = act ? dp[i] * w2m * inv_keep : 0.0f;
                        const float xhat = (x - meanm) * invstdm;
                        float d = gism * (g2 - c1m - xhat * c2m);
                        if (c >= rows || row >= MT_H) d = 0.0f;
                        dh[i] = d;
                        part += d;
                    }
                    sdb1 += (double)part;
                    if (row < MB_WROWS) {
#pragma unroll
                        for (int gg = 0; gg < 2; ++gg) {
                            uint4 p0, p1, p2;
                            split8_bf16x3(dh + 8 * gg, p0, p1, p2);
                            const int chunk = (c0 >> 3) + gg;
                            ts[chunk * MB_WROWS + row] = p0;
                            ts[tpiece / 16 + chunk * MB_WROWS + row] = p1;
                            ts[2 * (tpiece / 16) + chunk * MB_WROWS + row] = p2;
                        }
                    }
                }
            }
            umma::fence_proxy_async();
            umma::fence_before_sync();
            __syncthreads();
            if (tid == 0) {
                umma::fence_after_sync();
                umma::issue_table<false>(tab + n1, 42, umma::desc_hi(wcol), umma::desc_hi(wcol), 0u, D2, idesc2, 0u);
                umma::issue_table<false>(tab + n1 + 42, n4, umma::desc_hi(128u), umma::desc_hi(pcol), 0u, D3, idesc3,
                                         d3_started);
                umma::commit(bar);
            }
            d3_started = 1;
            // ---- epilogue 2: thread = (pool column q, 32 batch rows); dy[b,u,q] = pool[b,u,q] * D2[q][b];
            //      the pool values come back from L2 (read one tile ago), requested while the MMAs run
            const long long e2 = ((long long)(b0 + 32 * cg) * U + u) * n + row;
            const long long bstride = (long long)U * n;
            float x[2][16];
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int c = 32 * cg + 16 * h + i;
                    x[h][i] = (row < n && c < rows) ? __ldg(a.pool + e2 + (long long)(16 * h + i) * bstride) : 0.0f;
                }
            umma::mbar_wait(bar, phase);
            phase ^= 1u;
            umma::fence_after_sync();
            if (32 * cg < rows) {
                uint32_t v[2][16];
                umma::tmem_ld16(D2 + trow + 32 * cg, v[0]);
                umma::tmem_ld16(D2 + trow + 32 * cg + 16, v[1]);
                umma::tmem_ld_wait();
                if (row < n) {
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (32 * cg + 16 * h + i < rows)
                                a.dy[e2 + (long long)(16 * h + i) * bstride] = x[h][i] * __uint_as_float(v[h][i]);
                }
            }
            // no barrier here: the next tile's two barriers order these TMEM reads before the MMAs that
            // overwrite D1 / D2, and every MMA that reads the staging buffers has completed
        }
        flush(v, u);
        seg = seg_end;
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(sh->tmem_slot, 512);
}

// ---------------------------------------------------------------------------
// B2 on tile images, pipelined.  The same three GEMMs as mlp_bwd_kernel (D1 = W1 . pool^T -> dh, D2 = W1^T . dh ->
// dpool, D3 += dh . pool -> dW1), with
//   * the pool tile and W1 fetched by bulk load (TMA) from the images F1 wrote -- no global loads, no split, no
//     staging stores for the two big operands; the pool stage is double-buffered;
//   * TWO bf16 pieces per operand (hi, mid of the round-to-nearest split: x to 2^-18, unbiased) and the three
//     products hi*hi, hi*mid, mid*hi: gradients are compared at 1e-4, not at the forward's 1e-5, and the ReLU /
//     dropout decisions are the forward's stored bits, so nothing here can flip; half the MMAs (63 instead of 126
//     per tile), half the shared-memory operand fetch, two thirds of the dh staging;
//   * the phases of neighbouring tiles overlapped (a lock-step version of this kernel -- every warp in every phase:
//     T1 -> dh epilogue -> T3/T4 -> dy epilogue -- measured 0.53 ms with 24 % tensor-pipe active, this one 0.38 ms
//     with 37 %):
//   warps 0-11   E1: dh of tile g out of D1 (three warps per TMEM lane quarter, 16-column chunks k, k+3, k+6),
//                staged as two bf16 pieces; the unit's dW1 read-out (D3) at a unit boundary
//   warps 12-15  E2: dpool of tile g-1 out of D2[(g-1) & 1] (thread = pool column, all 128 batch rows in four
//                passes); no loads: the factor pool of dy = dpool * pool is applied by xnn_frontend_bwd
//   warp  16     issuer: T1(g) as soon as D1 is free, T3/T4(g) when dh is staged -- T1(g+1) therefore queues right
//                behind T3/T4(g) and E1(g+1) starts while E2(g) is still writing dy
//   warp  17     loader: bulk loads of the pool stage two tiles ahead and of W1 at a unit boundary, dopre / active
//                bits of the tile into a four-deep ring
// TMEM: D1 128 | D2[2] 2 x 128 | D3 96 columns.  Critical chain per tile: T3/T4(g) -> T1(g+1) -> E1(g+1) -> T3/T4(g+1).
// ---------------------------------------------------------------------------
#define MBP_E1_WARPS 12
#define MBP_E2_WARPS 4
#define MBP_THREADS (32 * (MBP_E1_WARPS + MBP_E2_WARPS + 2))

struct MbpShared {
    uint64_t full[2];        // pool stage landed (tx)
    uint64_t wfull;          // W1 landed (tx)
    uint64_t aux_full[4];    // dopre / active bits ring slot written
    uint64_t d1_full;        // T1 done (commit)
    uint64_t d1_free;        // 12 E1 warps have D1 in registers
    uint64_t dh_full;        // 12 E1 warps staged dh
    uint64_t d2_full[2];     // T3 / T4 done (commit): dy accumulator ready, pool stage and dh staging free
    uint64_t d2_free[2];     // 4 E2 warps have read D2[s]
    uint64_t d3_free;        // 12 E1 warps have read D3 (unit boundary)
    uint32_t tmem_slot;
    uint32_t pad;
};

__global__ void __launch_bounds__(MBP_THREADS, 1)
mlp_bwd_pipe_kernel(MlpBwdArgs a, const unsigned char* __restrict__ image, const unsigned char* __restrict__ w1_image) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int n = a.n, U = a.U, B = a.B, T = a.T;
    const int Nq = (n + 15) & ~15, chunksQ = Nq >> 3, kstepsQ = Nq >> 4;
    const uint32_t wcol = MB_WROWS * 16u;
    const uint32_t wpiece = (uint32_t)chunksQ * wcol;
    const uint32_t pcol = MB_ROWS * 16u;
    const uint32_t ppiece = (uint32_t)chunksQ * pcol;
    const uint32_t tpiece = (MB_ROWS / 8) * wcol;
    const uint32_t pstage2 = 2u * ppiece;
    const size_t img_stage = (size_t)3 * ppiece, img_w = (size_t)3 * wpiece;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    unsigned char* w1s = smem_raw;
    unsigned char* ps = smem_raw + 2 * wpiece;
    uint4* ts = reinterpret_cast<uint4*>(smem_raw + 2 * wpiece + 2 * pstage2);
    unsigned char* tail = smem_raw + 2 * wpiece + 2 * pstage2 + 2 * tpiece;
    float* s_dopre_all = reinterpret_cast<float*>(tail);                                 // [4][128]
    uint32_t* s_mask_all = reinterpret_cast<uint32_t*>(tail + 2048);                     // [4][4 words][128 rows]
    MbpShared* sh = reinterpret_cast<MbpShared*>(tail + 2048 + 8192);
    uint2* s_tab = reinterpret_cast<uint2*>(tail + 2048 + 8192 + 128);

    {
        uint4* z = reinterpret_cast<uint4*>(smem_raw);
        const uint32_t tot = (uint32_t)(tail - smem_raw) / 16;
        for (uint32_t i = tid; i < tot; i += MBP_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    if (warp == 0) umma::tmem_alloc(&sh->tmem_slot, 512);
    if (tid == 0) {
        umma::mbar_init(&sh->full[0], 1); umma::mbar_init(&sh->full[1], 1);
        umma::mbar_init(&sh->wfull, 1);
        for (int s = 0; s < 4; ++s) umma::mbar_init(&sh->aux_full[s], 1);
        umma::mbar_init(&sh->d1_full, 1);
        umma::mbar_init(&sh->d1_free, MBP_E1_WARPS);
        umma::mbar_init(&sh->dh_full, MBP_E1_WARPS);
        umma::mbar_init(&sh->d2_full[0], 1); umma::mbar_init(&sh->d2_full[1], 1);
        umma::mbar_init(&sh->d2_free[0], MBP_E2_WARPS); umma::mbar_init(&sh->d2_free[1], MBP_E2_WARPS);
        umma::mbar_init(&sh->d3_free, MBP_E1_WARPS);
    }
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = sh->tmem_slot;
    const uint32_t D1 = tmem, D2 = tmem + MB_ROWS, D3 = tmem + 3 * MB_ROWS;
    const int n1 = 3 * kstepsQ, n3 = 21, n4 = 3 * (MB_ROWS / 16), ntab = n1 + n3 + n4;
    if (tid < ntab) {
        const int ti = tid;
        const uint32_t w_addr = umma::smem_u32(w1s);
        const uint32_t p_addr = umma::smem_u32(ps), t_addr = umma::smem_u32(ts);
        const int seq = ti < n1 ? 0 : (ti < n1 + n3 ? 1 : 2);
        const int i = ti - (seq == 0 ? 0 : (seq == 1 ? n1 : n1 + n3));
        const int per = seq == 0 ? kstepsQ : (seq == 1 ? 7 : MB_ROWS / 16);
        const int pass = i / per, ks = i - pass * per;
        const uint32_t pa = pass == 0 ? 1u : 0u, pb = pass == 1 ? 1u : 0u;      // mid*hi, hi*mid, hi*hi
        uint2 e;
        if (seq == 0)
            e = make_uint2(umma::desc_lo(w_addr + pa * wpiece + (uint32_t)(2 * ks) * wcol, wcol),
                           umma::desc_lo(p_addr + pb * ppiece + (uint32_t)(2 * ks) * pcol, pcol));
        else if (seq == 1)
            e = make_uint2(umma::desc_lo(w_addr + pa * wpiece + (uint32_t)ks * 256u, 128u),
                           umma::desc_lo(t_addr + pb * tpiece + (uint32_t)ks * 256u, 128u));
        else
            e = make_uint2(umma::desc_lo(t_addr + pa * tpiece + (uint32_t)(2 * ks) * wcol, wcol),
                           umma::desc_lo(p_addr + pb * ppiece + (uint32_t)ks * 256u, 128u));
        s_tab[ti] = e;
    }
    __syncthreads();

    const long long tiles_total = (long long)U * T;
    const long long lo = (long long)blockIdx.x * tiles_total / gridDim.x;
    const long long hi = (long long)(blockIdx.x + 1) * tiles_total / gridDim.x;
    const int ntl = (int)(hi - lo);
    const float inv_keep = 1.0f / (1.0f - a.drop_p);
    const int C = U * MT_H;
    auto unit_of = [&](int v) {
        int u = v / a.P;
        for (int c = 0; c < v % a.P; ++c) u += (U - c + a.P - 1) / a.P;
        return u;
    };
    // tile g of this CTA: position v in the CTA order, unit, batch tile, first / last tile of its unit IN THIS CTA.
    // Every role walks its tiles in order, so (v, t, u) advance incrementally (no 64-bit divisions per tile).
    const int v_lo = (int)(lo / T), t_lo = (int)(lo % T);
    int cur_v = v_lo, cur_t = t_lo - 1, cur_u = unit_of(v_lo);
    auto tile_info = [&](int g, int& v, int& t, bool& first, bool& last) {
        if (++cur_t == T) { cur_t = 0; ++cur_v; cur_u = unit_of(cur_v); }
        v = cur_v; t = cur_t;
        first = (g == 0) || t == 0;
        last = (g == ntl - 1) || t == T - 1;
    };

    if (warp == MBP_E1_WARPS + MBP_E2_WARPS + 1) {
        // ================= loader =================
        for (int g = 0; g < ntl; ++g) {
            int v, t; bool first, last;
            tile_info(g, v, t, first, last);
            const int u = cur_u;
            const uint32_t st = (uint32_t)g & 1u;
            if (g >= 2) umma::mbar_wait(&sh->d2_full[st], (uint32_t)((g >> 1) - 1) & 1u);        // stage free
            if (first) {
                if (g >= 1) umma::mbar_wait(&sh->d2_full[st ^ 1u], (uint32_t)((g - 1) >> 1) & 1u);   // every MMA of the previous unit
                if (lane == 0) {
                    umma::mbar_arrive_expect_tx(&sh->wfull, 2u * wpiece);
                    umma::bulk_load(w1s, w1_image + (size_t)u * img_w, 2u * wpiece, &sh->wfull);
                }
            }
            // the big transfer first: the ring values below come through the LSU with a global-load latency of their own
            if (lane == 0) {
                umma::mbar_arrive_expect_tx(&sh->full[st], pstage2);
                umma::bulk_load(ps + (size_t)st * pstage2, image + (size_t)((long long)u * T + t) * img_stage, pstage2,
                                &sh->full[st]);
            }
            {
                const int b0 = t * MB_ROWS;
                float* sd = s_dopre_all + (g & 3) * 128;
                uint32_t* sm = s_mask_all + (g & 3) * 512;
                float dv[4];
                uint4 pm[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int r = lane + 32 * j;
                    const bool ok = b0 + r < B;
                    const long long bu = (long long)(b0 + r) * U + u;
                    dv[j] = ok ? __ldg(a.dopre + bu) : 0.0f;
                    pm[j] = ok ? __ldg(reinterpret_cast<const uint4*>(a.maskbits + bu * 4)) : make_uint4(0u, 0u, 0u, 0u);
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int r = lane + 32 * j;
                    sd[r] = dv[j];
                    sm[r] = pm[j].x; sm[128 + r] = pm[j].y; sm[256 + r] = pm[j].z; sm[384 + r] = pm[j].w;
                }
                __syncwarp();
                if (lane == 0) umma::mbar_arrive(&sh->aux_full[g & 3]);
            }
            __syncwarp();
        }
    } else if (warp == MBP_E1_WARPS + MBP_E2_WARPS) {
        // ================= issuer: one elected lane per MMA group, descriptors in the uniform datapath =================
        const uint32_t idesc1 = umma::idesc_bf16_f32(128, MB_ROWS);
        const uint32_t idesc2 = umma::idesc_bf16_f32(128, MB_ROWS) | (1u << 15) | (1u << 16);
        const uint32_t idesc3 = umma::idesc_bf16_f32(128, Nq) | (1u << 16);
        const uint32_t w_a = umma::warp_uniform(umma::desc_lo(umma::smem_u32(w1s), 0u)),
                       p_a = umma::warp_uniform(umma::desc_lo(umma::smem_u32(ps), 0u)),
                       t_a = umma::warp_uniform(umma::desc_lo(umma::smem_u32(ts), 0u));
        const uint32_t tm = umma::warp_uniform(tmem);
        const uint32_t uD1 = tm, uD2 = tm + MB_ROWS, uD3 = tm + 3 * MB_ROWS;
        const uint32_t lb_w = ((wcol >> 4) & 0x3FFFu) << 16, lb_p = ((pcol >> 4) & 0x3FFFu) << 16, lb_128 = 8u << 16;
        int ui = 0;                                                   // units started by this CTA
        for (int g = 0; g < ntl; ++g) {
            int v, t; bool first, last;
            tile_info(g, v, t, first, last);
            const uint32_t st = (uint32_t)g & 1u;
            if (first) { umma::mbar_wait(&sh->wfull, (uint32_t)ui & 1u); }
            umma::mbar_wait(&sh->full[st], (uint32_t)(g >> 1) & 1u);
            umma::mbar_wait(&sh->d1_free, ((uint32_t)g & 1u) ^ 1u);
            umma::fence_after_sync();
            const uint32_t p_st = p_a + st * (pstage2 >> 4);
            if (umma::elect_one()) {
                // D1[hidden][batch] = W1 . pool^T: A = W1 (K chunk pitch wcol), B = pool stage (pitch pcol)
                umma::issue_split3(w_a | lb_w, wpiece >> 4, (2u * wcol) >> 4, umma::desc_hi(128u),
                                   p_st | lb_p, ppiece >> 4, (2u * pcol) >> 4, umma::desc_hi(128u), kstepsQ, uD1, idesc1, 0u);
                umma::commit(&sh->d1_full);
            }
            __syncwarp();
            umma::mbar_wait(&sh->dh_full, (uint32_t)g & 1u);
            umma::mbar_wait(&sh->d2_free[st], (((uint32_t)g >> 1) & 1u) ^ 1u);
            if (first && ui > 0) umma::mbar_wait(&sh->d3_free, (uint32_t)(ui - 1) & 1u);
            umma::fence_after_sync();
            if (umma::elect_one()) {
                // D2[pool dim][batch] = W1^T . dh^T (both MN-major views), D3[hidden][pool dim] += dh . pool
                umma::issue_split3(w_a | lb_128, wpiece >> 4, 16u, umma::desc_hi(wcol), t_a | lb_128, tpiece >> 4, 16u,
                                   umma::desc_hi(wcol), 7, uD2 + st * MB_ROWS, idesc2, 0u);
                umma::issue_split3(t_a | lb_w, tpiece >> 4, (2u * wcol) >> 4, umma::desc_hi(128u), p_st | lb_128, ppiece >> 4,
                                   16u, umma::desc_hi(pcol), MB_ROWS / 16, uD3, idesc3, first ? 0u : 1u);
                umma::commit(&sh->d2_full[st]);
            }
            __syncwarp();
            if (first) ++ui;
        }
    } else if (warp < MBP_E1_WARPS) {
        // ================= E1: dh, and dW1 at unit boundaries =================
        const int quarter = warp & 3, k3 = warp >> 2;                 // k3 = 0..2: 16-column chunks k3, k3 + 3, k3 + 6
        const int row = quarter * 32 + lane;
        const uint32_t trow = (uint32_t)(quarter * 32) << 16;
        const uint32_t bit = 1u << (row & 31);
        const int wsel = row >> 5;
        float k0m = 0.f, k1m = 0.f, k2m = 0.f;
        double sdb1 = 0.0;
        for (int g = 0; g < ntl; ++g) {
            int v, t; bool first, last;
            tile_info(g, v, t, first, last);
            const int u = cur_u;
            const int rows = min(MB_ROWS, B - t * MB_ROWS);
            if (first) {
                const int ch = u * MT_H + row;
                if (row < MT_H) {
                    const float b1v = a.b1[ch], w2v = a.w2[ch];
                    const float mean = (float)a.stats2[4LL * ch], invstd = (float)a.stats2[4LL * ch + 2];
                    const float gis = (float)((double)a.gamma2[ch] * a.stats2[4LL * ch + 2]);
                    const float2 cf = a.coef[ch];
                    k1m = gis * w2v * inv_keep;
                    k2m = -gis * cf.y * invstd;
                    k0m = -gis * cf.x + k2m * (b1v - mean);
                } else {
                    k0m = k1m = k2m = 0.f;
                }
            }
            umma::mbar_wait(&sh->aux_full[g & 3], ((uint32_t)g >> 2) & 1u);
            umma::mbar_wait(&sh->d1_full, (uint32_t)g & 1u);
            umma::fence_after_sync();
            uint32_t vv[3][16];
#pragma unroll
            for (int m = 0; m < 3; ++m)
                if (k3 + 3 * m < 8) umma::tmem_ld16(D1 + trow + 16 * (k3 + 3 * m), vv[m]);
            umma::tmem_ld_wait();
            umma::fence_before_sync();
            __syncwarp();
            if (lane == 0) umma::mbar_arrive(&sh->d1_free);
            const float* s_dopre = s_dopre_all + (g & 3) * 128;
            const uint32_t* s_mask = s_mask_all + (g & 3) * 512 + wsel * 128;
#pragma unroll
            for (int m = 0; m < 3; ++m) {
                if (k3 + 3 * m < 8) {
                    const int c0 = 16 * (k3 + 3 * m);
                    float dh[16];
                    float part = 0.0f;
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
                        const uint4 mq = *reinterpret_cast<const uint4*>(s_mask + c0 + 4 * q4);
                        const float4 dq = *reinterpret_cast<const float4*>(s_dopre + c0 + 4 * q4);
                        const uint32_t mk[4] = {mq.x, mq.y, mq.z, mq.w};
                        const float dp[4] = {dq.x, dq.y, dq.z, dq.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int i = 4 * q4 + e;
                            const float base = fmaf(k2m, __uint_as_float(vv[m][i]), k0m);
                            float d = (mk[e] & bit) ? fmaf(k1m, dp[e], base) : base;
                            if (c0 + i >= rows) d = 0.0f;
                            dh[i] = d;
                            part += d;
                        }
                    }
                    sdb1 += (double)part;
                    if (row < MB_WROWS) {
#pragma unroll
                        for (int gg = 0; gg < 2; ++gg) {
                            uint32_t p0[4], p1[4];
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                const float x0 = dh[8 * gg + 2 * e], x1 = dh[8 * gg + 2 * e + 1];
                                p0[e] = cvt_bf16x2_rn(x0, x1);
                                p1[e] = cvt_bf16x2_rn(x0 - __uint_as_float(p0[e] << 16),
                                                      x1 - __uint_as_float(p0[e] & 0xFFFF0000u));
                            }
                            const int chunk = (c0 >> 3) + gg;
                            ts[chunk * MB_WROWS + row] = make_uint4(p0[0], p0[1], p0[2], p0[3]);
                            ts[tpiece / 16 + chunk * MB_WROWS + row] = make_uint4(p1[0], p1[1], p1[2], p1[3]);
                        }
                    }
                }
            }
            umma::fence_proxy_async();
            __syncwarp();
            if (lane == 0) umma::mbar_arrive(&sh->dh_full);
            if (last) {
                // dW1 of the unit: every T4 has completed when the last tile's commit arrives
                umma::mbar_wait(&sh->d2_full[g & 1], ((uint32_t)g >> 1) & 1u);
                umma::fence_after_sync();
                int j = blockIdx.x;
                const long long start = (long long)v * T;
                while (j > 0 && ((long long)j * tiles_total) / gridDim.x > start) --j;
                const int slot = blockIdx.x - j;
                const bool ok = row < MT_H && slot < a.slots;
                float* o = a.part_w1 + ((long long)slot * C + (long long)u * MT_H + row) * n;
                for (int c0 = 16 * k3; c0 < Nq; c0 += 48) {
                    uint32_t w[16];
                    umma::tmem_ld16(D3 + trow + c0, w);
                    umma::tmem_ld_wait();
                    if (ok) {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (c0 + i < n) o[c0 + i] = __uint_as_float(w[i]);
                    }
                }
                if (ok) a.part_db1[((long long)slot * 4 + k3) * C + (long long)u * MT_H + row] = sdb1;
                sdb1 = 0.0;
                umma::fence_before_sync();
                __syncwarp();
                if (lane == 0) umma::mbar_arrive(&sh->d3_free);
            }
        }
    } else {
        // ================= E2: dpool[b,u,q] = D2[q][b] =================
        // (the factor pool[b,u,q] of dy = dpool * pool -- the derivative of exp at the pooled maximum -- is applied
        //  by the conv weight gradient kernel, which recomputes pool from the zmax it reads anyway: this pass
        //  then has no global loads at all)
        const int quarter = warp & 3;
        const int row = quarter * 32 + lane;                          // pool column q
        const uint32_t trow = (uint32_t)(quarter * 32) << 16;
        const long long bstride = (long long)U * n;
        for (int g = 0; g < ntl; ++g) {
            int v, t; bool first, last;
            tile_info(g, v, t, first, last);
            const int u = cur_u;
            const int b0 = t * MB_ROWS, rows = min(MB_ROWS, B - b0);
            const uint32_t st = (uint32_t)g & 1u;
            umma::mbar_wait(&sh->d2_full[st], ((uint32_t)g >> 1) & 1u);
            umma::fence_after_sync();
#pragma unroll 1
            for (int pass = 0; pass < 4; ++pass) {
                uint32_t w[2][16];
                umma::tmem_ld16(D2 + st * MB_ROWS + trow + 32 * pass, w[0]);
                umma::tmem_ld16(D2 + st * MB_ROWS + trow + 32 * pass + 16, w[1]);
                umma::tmem_ld_wait();
                if (pass == 3) {
                    umma::fence_before_sync();
                    __syncwarp();
                    if (lane == 0) umma::mbar_arrive(&sh->d2_free[st]);
                }
                if (row < n && 32 * pass < rows) {
                    float* dp2 = a.dy + ((long long)(b0 + 32 * pass) * U + u) * n + row;
                    if (32 * pass + 32 <= rows) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) { *dp2 = __uint_as_float(w[i >> 4][i & 15]); dp2 += bstride; }
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            if (32 * pass + i < rows) *dp2 = __uint_as_float(w[i >> 4][i & 15]);
                            dp2 += bstride;
                        }
                    }
                }
            }
        }
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(sh->tmem_slot, 512);
}

static size_t mbp_smem(int n) {
    const size_t chunksQ = (size_t)(((n + 15) & ~15) >> 3);
    const size_t wcol = MB_WROWS * 16;
    return 2 * chunksQ * wcol + 2 * 2 * chunksQ * MB_ROWS * 16 + 2 * (MB_ROWS / 8) * wcol + 2048 + 8192 + 128 + 64 * 8 + 16;
}

__global__ void __launch_bounds__(256)
mlp_bwd_reduce_kernel(const float* __restrict__ part_w1, const double* __restrict__ part_db1, int slots,
                      long long nw, int C, float* __restrict__ dw1, float* __restrict__ db1) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nw;
         i += (long long)gridDim.x * blockDim.x) {
        float v = 0.0f;
        for (int s = 0; s < slots; ++s) v += part_w1[(long long)s * nw + i];
        dw1[i] = v;
        if (i < C) {
            double d = 0.0;
            for (int s = 0; s < 4 * slots; ++s) d += part_db1[(long long)s * C + i];
            db1[i] = (float)d;
        }
    }
}

static int mb_tiles(int B) { return (B + MB_ROWS - 1) / MB_ROWS; }
static int mb_grid(int B, int U) {
    const long long tiles = (long long)U * mb_tiles(B);
    const int sms = num_sms();
    return (int)(tiles < sms ? tiles : sms);
}
static int mb_interleave(int B, int U) {
    const int G = mb_grid(B, U);
    const int P = (U + G / 2) / G;      // units per CTA, rounded
    return P < 1 ? 1 : P;
}
static int mb_slots(int B, int U) {
    const int T = mb_tiles(B), G = mb_grid(B, U);
    const long long per = ((long long)U * T) / G;
    return (int)((T + per - 1) / per) + 1;      // CTAs that can touch one unit
}
static size_t mb_smem(int n) {
    const size_t chunksQ = (size_t)(((n + 15) & ~15) >> 3);
    const size_t wcol = MB_WROWS * 16;
    return 3 * chunksQ * wcol + 3 * chunksQ * MB_ROWS * 16 + 3 * (MB_ROWS / 8) * wcol + 512 + 2048 + 32 + 126 * 8 + 16;
}
static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct MbWork { size_t part1, coef, part_w1, part_db1, total; };
static MbWork mb_work(int B, int U, int n) {
    const size_t C = (size_t)U * MT_H;
    MbWork w;
    w.part1 = 0;
    w.coef = align256((size_t)mt_slots(B, U) * 4 * C * sizeof(double));
    w.part_w1 = w.coef + align256(C * sizeof(float2));
    w.part_db1 = w.part_w1 + align256((size_t)mb_slots(B, U) * C * n * sizeof(float));
    w.total = w.part_db1 + align256((size_t)4 * mb_slots(B, U) * C * sizeof(double));
    return w;
}

// bytes at the start of `work` that hold the partial BN-backward sums after phase 1 (summed over ranks between
// the phases under synchronised BatchNorm)
extern "C" size_t xnn_mlp_bwd_partial_bytes(int B, int U) {
    if (B < 1 || U < 1) return 0;
    return (size_t)mt_slots(B, U) * 4 * U * MT_H * sizeof(double);
}

extern "C" size_t xnn_mlp_bwd_work_bytes(int B, int U, int n) {
    if (B < 1 || U < 1 || n < 1) return 0;
    return mb_work(B, U, n).total;
}

extern "C" int xnn_mlp_bwd(const float* dopre, const float* pool, const float* w1, const float* b1,
                           const float* scale2, const float* shift2, const double* stats2, const float* gamma2,
                           const float* w2, const uint32_t* maskbits, float drop_p, int B, int U, int n,
                           float* dw1, float* db1, float* dgamma2, float* dbeta2, float* dw2, float* db2,
                           float* dy, void* work, const void* image, int phase, long long count, float out_scale,
                           xnn_stream_t s) {
    XNN_CHECK_ARG(dopre && pool && w1 && b1 && scale2 && shift2 && stats2 && gamma2 && w2 && dw1 && db1 && dgamma2 &&
                  dbeta2 && dw2 && db2 && dy && work, "NULL pointer");
    XNN_CHECK_ARG(phase >= 0 && phase <= 2, "phase: 0 = whole call, 1 = reductions, 2 = the rest");
    if (phase == 0) { count = B; out_scale = 1.0f; }
    XNN_CHECK_ARG(B >= 1 && count >= 2 && U >= 1 && drop_p >= 0.0f && drop_p < 1.0f, "bad sizes");
    XNN_CHECK_ARG(xnn_mlp_supported(n), "pool width outside the tensor-core path (n <= 96)");
    XNN_CHECK_ARG(((uintptr_t)work & 255) == 0, "work must be 256-byte aligned");
    XNN_CHECK_ARG(maskbits != nullptr && ((uintptr_t)maskbits & 15) == 0, "maskbits: required, 16-byte aligned");
    cudaStream_t st = to_stream(s);
    const MbWork w = mb_work(B, U, n);
    const int C = U * MT_H;
    char* base = (char*)work;
    if (phase != 2) XNN_CUDA(cudaMemsetAsync(work, 0, w.total, st));
    // B1: reductions
    MlpArgs a{};
    a.pool = pool; a.w1 = w1; a.b1 = b1; a.scale2 = scale2; a.shift2 = shift2; a.w2 = w2; a.dopre = dopre;
    a.maskbits = const_cast<uint32_t*>(maskbits);
    a.B = B; a.U = U; a.n = n; a.T = mt_tiles(B); a.slots = mt_slots(B, U); a.drop_p = drop_p;
    a.part = (double*)(base + w.part1);
    XNN_CHECK_ARG(((uintptr_t)image & 127) == 0, "image must be 128-byte aligned");
    if (image != nullptr) { a.image = (const unsigned char*)image; a.w1_image = img_w1_ptr(image, B, U, n); }
    if (phase != 2 && mt_launch<MLP_BWD_STATS>(a, st)) { set_error("xnn_mlp_bwd: B1 launch failed: %s", cudaGetErrorString(cudaGetLastError())); return XNN_ERR_CUDA; }
    if (phase == 1) return XNN_OK;
    float2* coef = (float2*)(base + w.coef);
    mlp_bwd_finalize_kernel<<<ceil_div(C, 256), 256, 0, st>>>(a.part, a.slots, count, U, stats2, dgamma2, dbeta2, dw2,
                                                              db2, coef, out_scale);
    XNN_CHECK_LAUNCH();
    // B2: dh -> dW1, dy
    MlpBwdArgs g{};
    g.pool = pool; g.w1 = w1; g.b1 = b1; g.scale2 = scale2; g.shift2 = shift2; g.w2 = w2; g.gamma2 = gamma2;
    g.stats2 = stats2; g.coef = coef; g.dopre = dopre; g.maskbits = maskbits; g.dy = dy;
    g.part_w1 = (float*)(base + w.part_w1); g.part_db1 = (double*)(base + w.part_db1);
    g.B = B; g.U = U; g.n = n; g.T = mb_tiles(B); g.slots = mb_slots(B, U); g.drop_p = drop_p;
    g.P = mb_interleave(B, U);
    if (image != nullptr) {
        const size_t smem = mbp_smem(n);
        XNN_CUDA(cudaFuncSetAttribute(mlp_bwd_pipe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        mlp_bwd_pipe_kernel<<<mb_grid(B, U), MBP_THREADS, smem, st>>>(g, a.image, a.w1_image);
    } else {
        const size_t smem = mb_smem(n);
        XNN_CUDA(cudaFuncSetAttribute(mlp_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        mlp_bwd_kernel<<<mb_grid(B, U), MT_THREADS, smem, st>>>(g);
    }
    XNN_CHECK_LAUNCH();
    const long long nw = (long long)C * n;
    long long blocks = (nw + 255) / 256;
    if (blocks > (long long)num_sms() * 8) blocks = (long long)num_sms() * 8;
    mlp_bwd_reduce_kernel<<<(int)blocks, 256, 0, st>>>(g.part_w1, g.part_db1, g.slots, nw, C, dw1, db1);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// a6 on the tensor cores -- the conv weight gradient as a Toeplitz GEMM.
//
//   G_u[c,j] = sum_{b,q} dy[b,u,q] * x[b, c, 7q + arg[b,u,q] + j]
//
// (explainn/architectures/__init__.py:73-83 differentiated; selene/__init__.py:291).  dy is
// the gradient at the max-pool arg-max, so with g[u,p] = dy at p = 7q+arg and 0 elsewhere
//
//   D[unit, f = 4j+c] (fp32, TMEM) = A[unit, p] (g, bf16 x2 split) . B[p, f] (one-hot window)
//
// B is never built.  The sequence sits in shared memory as bf16 one-hot PAIRS, 16 bytes per
// position: P2[p] = (x[.,p], x[.,p+1]).  Element (p, f) of the im2col matrix is then at byte
// 16p + 2f, and an MN-major UMMA operand without swizzle, which addresses
//   K row r, 16-byte N chunk c  ->  start + 16*(r%8) + LBO*(r/8) + SBO*c,
// is the Toeplitz view with SBO = 32 (chunk c = taps 2c, 2c+1 = P2[p+2c]).  LBO = 112 = 7*16
// makes every group of 8 K rows start at a pool window: K rows 8g..8g+7 <-> positions
// 7g..7g+7, so ONE 16-byte K chunk of A per pool window holds that window's single non-zero
// at slot arg (slot 7, the next window's first position, is always zero).
//
// A (K-major): each producer lane owns one (unit row, window): it loads dy / zmax / argmax
// coalesced from HBM, splits dy into two bf16 pieces (round to nearest) and writes one 16-byte chunk per
// piece -- no transposition through shared memory, no cross-thread scatter.  Chunk
// pitch LBO_A = 144 B (128 + 16) keeps those stores bank-conflict free.
// The accumulator stays in TMEM for the CTA's whole slice of the batch and is read once.
//
// Roles (one CTA of 18 warps per SM): two producer groups of 8 warps convert alternate tiles (a warp's
// serial path per tile is longer than the tensor time of a tile); warp 16 issues every tcgen05.mma
// (elect.sync, uniform descriptors); warp 17 expands the next sequence's planes into the pair buffer.  Stages of 8 windows x 128
// rows cycle through full[] (one arrival per producer warp) / empty[] (tcgen05.commit) mbarriers; there is
// no CTA-wide barrier in the loop, and each producer prefetches its next tile while it converts this one.
#include "common.cuh"
#include "umma.cuh"
#include <cuda_bf16.h>
using namespace xnn;

#define BT_KMAX 32
#define BT_ROWS 128                 // units per CTA = UMMA M
#define BT_WIN 8                    // pool windows per tile (stage)
#define BT_SPLITS 2                 // bf16 pieces of dy (round-to-nearest: hi + mid = dy to 2^-18, unbiased)
#define BT_LBO_A 144                // bytes between K chunks (windows) of one row group
#define BT_SBO_A (BT_WIN * BT_LBO_A)         // bytes between 8-row groups (1152)
#define BT_PIECE_BYTES (16 * BT_SBO_A)       // one bf16 piece of a stage (18432)
#define BT_STAGE_BYTES (BT_SPLITS * BT_PIECE_BYTES)   // 55296
#define BT_MAX_STAGES 4
#define BT_GROUPS 2                 // producer groups; group g converts the tiles with global index = g mod 2
#define BT_GROUP_WARPS 8
#define BT_PRODUCERS (BT_GROUPS * BT_GROUP_WARPS * 32)   // 512
#define BT_THREADS (BT_PRODUCERS + 64)                   // + issuer warp + one-hot warp
#define BT_RGS 4                    // row groups of 4 rows per producer warp (16 rows per warp)

static inline int bt_npad(int k) { return ((4 * k + 15) / 16) * 16; }
static inline int bt_mtiles(int U) { return (U + BT_ROWS - 1) / BT_ROWS; }

// batch slices (CTAs) per M tile: the cost of a tile is the same for every M tile (UMMA M = 128)
static int bt_slices(int U, int B) {
    int s = num_sms() / bt_mtiles(U);
    if (s < 1) s = 1;
    if (s > B) s = B;
    return s;
}

struct BtShared {
    uint64_t full[BT_MAX_STAGES];    // producer group -> issuer: stage written and fenced (8 warp arrivals)
    uint64_t empty[BT_MAX_STAGES];   // issuer -> producers: the MMAs that read the stage completed
    uint64_t oh_full[2];             // one-hot warp -> issuer: pair buffer of a sequence written
    uint64_t oh_empty[2];            // issuer -> one-hot warp: the MMAs of the sequence completed
    uint32_t tmem_slot;
    uint32_t pad;
};

__device__ __forceinline__ void bt_producer_sync() { asm volatile("bar.sync 1, 512;" ::: "memory"); }

struct BtVals {   // one tile's values of a producer lane: 4 row groups
    float d[BT_RGS], z[BT_RGS];
    int a[BT_RGS];
};

__global__ void __launch_bounds__(BT_THREADS, 1)
frontend_bwd_tc_kernel(const uint32_t* __restrict__ planes, int B, int L, int Wp, int U, int k, int n, int npad,
                       int Lbuf, int nstages,
                       const float* __restrict__ dy, const float* __restrict__ zmax,
                       const uint8_t* __restrict__ argmax, const double* __restrict__ stats,
                       uint32_t b_lbo, uint32_t b_sbo,
                       const float* __restrict__ act_scale, const float* __restrict__ act_shift,
                       double* __restrict__ part /* [gridDim.y][U][4k+2] */) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* g_smem = smem_raw;                                                  // [nstages][3][16][8][144]
    uint2* oh = reinterpret_cast<uint2*>(smem_raw + (size_t)nstages * BT_STAGE_BYTES); // [2][Lbuf] pairs (16 B)
    double* s_red = reinterpret_cast<double*>(oh + 4 * (size_t)Lbuf);                  // [2 groups][128][2]
    uint32_t* s_pl = reinterpret_cast<uint32_t*>(s_red + 2 * 2 * BT_ROWS);             // [3 * Wp] plane words
    BtShared* sh = reinterpret_cast<BtShared*>(s_pl + ((3 * Wp + 3) & ~3));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int u0 = blockIdx.x * BT_ROWS;
    const int nu = min(BT_ROWS, U - u0);
    const int K4 = 4 * k, pstride = K4 + 2;
    const int ns = gridDim.y;

    // ---- one-time setup
    if (warp == 0) umma::tmem_alloc(&sh->tmem_slot, 128);
    if (tid == 0) {
        for (int s = 0; s < BT_MAX_STAGES; ++s) {
            umma::mbar_init(&sh->full[s], BT_GROUP_WARPS);
            umma::mbar_init(&sh->empty[s], 1);
        }
        for (int s = 0; s < 2; ++s) { umma::mbar_init(&sh->oh_full[s], 1); umma::mbar_init(&sh->oh_empty[s], 1); }
    }
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = sh->tmem_slot;
    const int ntiles = (n + BT_WIN - 1) / BT_WIN;
    const int nseq = (B - 1 - (int)blockIdx.y) / ns + 1;
    const int gtotal = nseq * ntiles;

    if (warp == BT_PRODUCERS / 32) {
        // ================= issuer warp =================
        // every lane runs the loop and waits on the barriers; the MMAs of a tile are issued by ONE elected lane from
        // operands made warp-uniform up front (shfl from lane 0), so the descriptor arithmetic runs in the uniform
        // datapath: ~5 instructions per MMA.  (Per-MMA elect.sync on descriptors derived from threadIdx cost ~30
        // instructions and ~170 cycles per MMA in R2UR / VOTEU chains; under a plain `if (lane == 0)` the compiler
        // wraps every MMA in a waterfall loop.)
        const uint32_t idesc = umma::idesc_bf16_f32(BT_ROWS, npad) | (1u << 16);   // B is MN-major
        const uint32_t a_lo = umma::warp_uniform(umma::desc_lo(umma::smem_u32(g_smem), BT_LBO_A));
        const uint32_t b_lo = umma::warp_uniform(umma::desc_lo(umma::smem_u32(oh), b_lbo));
        const uint32_t tmem_u = umma::warp_uniform(tmem);
        const uint32_t a_hi = umma::desc_hi(BT_SBO_A), b_hi = umma::desc_hi(b_sbo);
        uint32_t stage = 0, use = 0, any_mma = 0;
        for (int sq = 0; sq < nseq; ++sq) {
            const uint32_t b_sq = b_lo + (uint32_t)(sq & 1) * (uint32_t)Lbuf;      // Lbuf * 16 B >> 4
            umma::mbar_wait(&sh->oh_full[sq & 1], (uint32_t)(sq >> 1) & 1u);
            for (int t = 0; t < ntiles; ++t) {
                umma::mbar_wait(&sh->full[stage], use & 1u);
                umma::fence_after_sync();
                if (umma::elect_one()) {
                    const uint32_t a_st = a_lo + stage * (BT_STAGE_BYTES >> 4);
                    const uint32_t b_t = b_sq + 56u * (uint32_t)t;                 // 8 windows * 7 positions
                    const int ksteps = (min(BT_WIN, n - t * BT_WIN) + 1) >> 1;
                    uint32_t acc = any_mma;
#pragma unroll
                    for (int s3 = 0; s3 < BT_SPLITS; ++s3) {
                        uint32_t a_k = a_st + (uint32_t)(s3 * BT_PIECE_BYTES) / 16u, b_k = b_t;
#pragma unroll 4
                        for (int ks = 0; ks < ksteps; ++ks) {
                            umma::mma_bf16(tmem_u, ((uint64_t)a_hi << 32) | (uint64_t)a_k, ((uint64_t)b_hi << 32) | (uint64_t)b_k,
                                           idesc, acc);
                            acc = 1;
                            a_k += (uint32_t)(2 * BT_LBO_A) / 16u;
                            b_k += 14u;
                        }
                    }
                    umma::commit(&sh->empty[stage]);
                    if (t == ntiles - 1) umma::commit(&sh->oh_empty[sq & 1]);      // pair buffer may be rewritten
                }
                any_mma = 1;
                __syncwarp();
                if (++stage == (uint32_t)nstages) { stage = 0; ++use; }
            }
        }
    } else if (warp == BT_PRODUCERS / 32 + 1) {
        // ================= one-hot warp: planes -> bf16 pair buffer of each sequence =================
        for (int sq = 0; sq < nseq; ++sq) {
            const int b = blockIdx.y + sq * ns;
            const uint32_t* sp = planes + (long long)b * 3 * Wp;
            for (int i = lane; i < 3 * Wp; i += 32) s_pl[i] = __ldg(sp + i);
            if (sq >= 2) {
                if (lane == 0) umma::mbar_wait(&sh->oh_empty[sq & 1], (uint32_t)((sq >> 1) - 1) & 1u);
            }
            __syncwarp();
            uint2* o = oh + (size_t)(sq & 1) * 2 * Lbuf;
            for (int i = lane; i <= Lbuf; i += 32) {
                uint2 v = make_uint2(0u, 0u);
                if (i < L) {
                    const int w = i >> 5, bt = i & 31;
                    const uint32_t lo = (s_pl[w] >> bt) & 1u, hi = (s_pl[Wp + w] >> bt) & 1u,
                                   nm = (s_pl[2 * Wp + w] >> bt) & 1u;
                    const uint32_t half = nm ? 0u : (lo ? 0x3F800000u : 0x00003F80u);
                    if (hi) v.y = half; else v.x = half;
                }
                if (i < Lbuf) o[2 * i] = v;
                if (i > 0) o[2 * i - 1] = v;
            }
            umma::fence_proxy_async();
            __syncwarp();
            if (lane == 0) umma::mbar_arrive(&sh->oh_full[sq & 1]);
        }
    } else {
        // ================= producers =================
        // group grp converts the tiles whose global index (sequence-major) is = grp mod 2; inside a group warp
        // wg owns unit rows 16wg..16wg+15; lane = (rsub = lane/8, qsub = lane%8); row group rg -> row 16wg+4rg+rsub
        const int grp = warp / BT_GROUP_WARPS, wg = warp % BT_GROUP_WARPS;
        const int rsub = lane >> 3, qsub = lane & 7;
        int row[BT_RGS];
        int rowd[BT_RGS];                                // element offset of the row from the warp's first row
        float wsr[BT_RGS], scr[BT_RGS], shr[BT_RGS];
        const bool dpool_in = (act_scale != nullptr);   // dy holds dL/dpool: multiply by pool = exp(scale * zmax + shift)
        // rows 16wg + 4rg + rsub: (row >> 3) = 2wg + (rg >> 1), (row & 7) = 4 (rg & 1) + rsub, so the stage offset of
        // row group rg is ONE base plus the compile-time delta (rg >> 1) * BT_SBO_A + (rg & 1) * 64
        const uint32_t soff0 = (uint32_t)(2 * wg * BT_SBO_A + rsub * 16 + qsub * BT_LBO_A);
        const int row0 = min(16 * wg, nu - 1);
        const long long rowbase = ((long long)blockIdx.y * U + u0 + row0) * n;
#pragma unroll
        for (int rg = 0; rg < BT_RGS; ++rg) {
            row[rg] = 16 * wg + 4 * rg + rsub;
            const int lrow = min(row[rg], nu - 1);       // rows past the tile load a valid row, then zero it
            rowd[rg] = (lrow - row0) * n;
            wsr[rg] = (float)stats[4 * (u0 + lrow) + 3];
            scr[rg] = dpool_in ? act_scale[u0 + lrow] : 0.0f;
            shr[rg] = dpool_in ? act_shift[u0 + lrow] : 0.0f;
        }
        double sbd[BT_RGS], sgd[BT_RGS];
        float a_d[BT_RGS], a_dz[BT_RGS];
#pragma unroll
        for (int rg = 0; rg < BT_RGS; ++rg) { sbd[rg] = 0.0; sgd[rg] = 0.0; a_d[rg] = 0.0f; a_dz[rg] = 0.0f; }
        const long long seq_step = (long long)ns * U * n;

        // 8 lanes share a row: fold the per-lane float partial sums of one sequence into the fp64 row sums
        auto flush = [&]() {
#pragma unroll
            for (int rg = 0; rg < BT_RGS; ++rg) {
                float t1 = a_d[rg], t2 = a_dz[rg];
#pragma unroll
                for (int o = 1; o < 8; o <<= 1) {
                    t1 += __shfl_xor_sync(0xffffffffu, t1, o);
                    t2 += __shfl_xor_sync(0xffffffffu, t2, o);
                }
                sbd[rg] += (double)t1;
                sgd[rg] += (double)t2;
                a_d[rg] = 0.0f; a_dz[rg] = 0.0f;
            }
        };

        // this group's tile list: g = grp, grp + 2, ...;  (f_sq, f_t) runs one tile ahead of (c_sq, c_t)
        int f_sq = 0, f_t = grp;
        while (f_t >= ntiles) { f_t -= ntiles; ++f_sq; }
        long long f_seq = (long long)f_sq * seq_step;
        auto fetch = [&](BtVals& v) {
            if (f_sq < nseq) {
                const int q = min(f_t * BT_WIN + qsub, n - 1);     // windows past n load a valid one, then zero it
                const long long g0 = rowbase + f_seq + q;          // one 64-bit base, 32-bit row deltas
                const float* pd = dy + g0;
                const float* pz = zmax + g0;
                const uint8_t* pa = argmax + g0;
#pragma unroll
                for (int rg = 0; rg < BT_RGS; ++rg) {
                    v.d[rg] = __ldg(pd + rowd[rg]);
                    v.z[rg] = __ldg(pz + rowd[rg]);
                    v.a[rg] = (int)__ldg(pa + rowd[rg]);
                }
            }
            f_t += BT_GROUPS;
            while (f_t >= ntiles) { f_t -= ntiles; ++f_sq; f_seq += seq_step; }
        };

        // L2 prefetch of the NEXT sequence's rows of this warp (16 rows are one contiguous block of each array):
        // the register prefetch below runs only one tile ahead, far less than the HBM latency under load
        auto prefetch_seq = [&](int sq) {
            if (sq >= nseq) return;
            const int r0 = 16 * wg;
            if (r0 >= nu) return;
            const long long e0 = ((long long)(blockIdx.y + (long long)sq * ns) * U + u0 + r0) * n;
            const long long ne = (long long)min(16, nu - r0) * n;
            if (grp == 0) {
                const char* p = reinterpret_cast<const char*>(dy + e0);
                for (long long o = (long long)lane * 128; o < ne * 4; o += 32 * 128)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(p + o));
            } else {
                const char* p = reinterpret_cast<const char*>(zmax + e0);
                for (long long o = (long long)lane * 128; o < ne * 4; o += 32 * 128)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(p + o));
                const char* pa = reinterpret_cast<const char*>(argmax + e0);
                for (long long o = (long long)lane * 128; o < ne; o += 32 * 128)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(pa + o));
            }
        };

        BtVals cur, nxt;
#pragma unroll
        for (int rg = 0; rg < BT_RGS; ++rg) { cur.d[rg] = nxt.d[rg] = 0.0f; cur.z[rg] = nxt.z[rg] = 0.0f; cur.a[rg] = nxt.a[rg] = 0; }
        int c_sq = f_sq, c_t = f_t;
        fetch(cur);
        prefetch_seq(c_sq + 1);
        uint32_t stage = (uint32_t)grp % (uint32_t)nstages, use = (uint32_t)grp / (uint32_t)nstages;
        for (int g = grp; g < gtotal; g += BT_GROUPS) {
            const int n_sq = f_sq, n_t = f_t;               // coordinates of the tile fetched next
            if (use) {                                      // lane 0 polls, the warp follows
                if (lane == 0) umma::mbar_wait(&sh->empty[stage], (use - 1) & 1u);
                __syncwarp();
            }
            // the next tile's loads fly under this tile's conversion (issued after the wait loop: before it,
            // the register allocator copies the first loaded value at once and the warp stalls on the load)
            fetch(nxt);
            unsigned char* st = g_smem + (size_t)stage * BT_STAGE_BYTES;
            const bool q_ok = (c_t * BT_WIN + qsub) < n;
#pragma unroll
            for (int rg = 0; rg < BT_RGS; ++rg) {
                // pool is recomputed with the forward's own expression (|s| * (sgn(s) z) == s * z exactly), so the
                // factor is bit-identical to the pooled activation the forward wrote
                float dv = (q_ok && row[rg] < nu) ? cur.d[rg] : 0.0f;
                if (dpool_in) dv *= expf(fmaf(scr[rg], cur.z[rg], shr[rg]));
                // two-way round-to-nearest split: dv = hi + mid to 2^-18 |dv| with an unbiased error; the one-hot
                // operand is exact, so the weight gradient (a sum over B * n such terms, compared at 1e-4) keeps
                // fp32-class accuracy with two thirds of the MMAs and staging stores of a three-way split
                const __nv_bfloat16 hi = __float2bfloat16_rn(dv);
                const __nv_bfloat16 mid = __float2bfloat16_rn(dv - __bfloat162float(hi));
                unsigned char* e = st + soff0 + (rg >> 1) * BT_SBO_A + (rg & 1) * 64;
                // zero the window's 16-byte chunk of each piece, then drop the value at slot arg (same thread,
                // program order); piece order mid, hi: the small piece enters the fp32 accumulator first.
                // (Building the chunk in registers for ONE store per piece was measured slower: 0.685 -> 0.711 ms,
                // eight selects cost this instruction-bound loop more than the second store.)
                const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
                *reinterpret_cast<uint4*>(e + 0 * BT_PIECE_BYTES) = zero4;
                *reinterpret_cast<uint4*>(e + 1 * BT_PIECE_BYTES) = zero4;
                unsigned char* e2 = e + 2 * cur.a[rg];
                *reinterpret_cast<__nv_bfloat16*>(e2 + 0 * BT_PIECE_BYTES) = mid;
                *reinterpret_cast<__nv_bfloat16*>(e2 + 1 * BT_PIECE_BYTES) = hi;
                a_d[rg] += dv;
                a_dz[rg] = fmaf(dv, cur.z[rg] - wsr[rg], a_dz[rg]);
            }
            umma::fence_proxy_async();                      // this thread's stage writes -> async proxy
            __syncwarp();
            if (lane == 0) umma::mbar_arrive(&sh->full[stage]);
            stage += BT_GROUPS;
            while (stage >= (uint32_t)nstages) { stage -= (uint32_t)nstages; ++use; }
            if (n_sq != c_sq) { flush(); prefetch_seq(n_sq + 1); }   // sequence boundary: float partials -> fp64
            c_sq = n_sq; c_t = n_t;
            cur = nxt;
        }
        flush();
        // all MMAs complete (they finish in order: wait for the last tile)
        {
            const int g2 = gtotal - 1;
            umma::mbar_wait(&sh->empty[g2 % nstages], (uint32_t)(g2 / nstages) & 1u);
        }
        umma::fence_after_sync();
        if (qsub == 0) {
#pragma unroll
            for (int rg = 0; rg < BT_RGS; ++rg) {
                if (row[rg] < nu) {
                    s_red[(grp * BT_ROWS + row[rg]) * 2] = sbd[rg];
                    s_red[(grp * BT_ROWS + row[rg]) * 2 + 1] = sgd[rg] * stats[4 * (u0 + row[rg]) + 2];   // * invstd
                }
            }
        }
        bt_producer_sync();

        // ---- read the accumulator once: warps 0-3, thread = unit (TMEM lane), columns f = 4j + c
        if (warp < 4) {
            const int r = tid;                      // TMEM lane = unit row
            const bool row_ok = r < nu;
            double* o = part + ((long long)blockIdx.y * U + u0 + r) * pstride;
            const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);
            for (int c0 = 0; c0 < npad; c0 += 16) {
                uint32_t v[16];
                umma::tmem_ld16(trow + c0, v);
                umma::tmem_ld_wait();
                if (row_ok) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const int f = c0 + i;
                        if (f < K4) o[(f & 3) * k + (f >> 2)] = (double)__uint_as_float(v[i]);
                    }
                }
            }
            if (row_ok) {
                o[K4] = s_red[r * 2] + s_red[(BT_ROWS + r) * 2];
                o[K4 + 1] = s_red[r * 2 + 1] + s_red[(BT_ROWS + r) * 2 + 1];
            }
        }
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, 128);
}

namespace xnn {
int launch_frontend_bwd_finalize(const double* part, int slices, int U, int k, const float* gamma,
                                 const double* stats, const double* cov_w, const long long* cnt1, long long n_pos,
                                 float* dconv_w, float* dconv_b, float* dgamma, float* dbeta, cudaStream_t st);
int frontend_bwd_simt_slices(int B);
}

// shared memory of one CTA; picks the number of A stages (3, or 2 for long sequences)
static size_t bt_smem_bytes(int n, int* Lbuf_out, int* nstages_out) {
    const int ntiles = (n + BT_WIN - 1) / BT_WIN;
    const int Lbuf = 56 * ntiles + 24;     // last tile reads P2 up to 56*ntiles + 18
    // pair buffers, fp64 row sums of both groups, plane words (<= 3 * ceil(L/32), L <= 7n + 38), barriers
    const size_t fixed = (size_t)2 * Lbuf * 16 + (size_t)2 * 2 * BT_ROWS * 8 + (size_t)3 * 4 * ((7 * n + 38 + 31) / 32 + 2) +
                         sizeof(BtShared) + 32;
    int nstages = BT_MAX_STAGES;
    while (nstages > 2 && (size_t)nstages * BT_STAGE_BYTES + fixed > (size_t)227 * 1024) --nstages;
    if (Lbuf_out) *Lbuf_out = Lbuf;
    if (nstages_out) *nstages_out = nstages;
    return (size_t)nstages * BT_STAGE_BYTES + fixed;
}

extern "C" size_t xnn_frontend_bwd_work_bytes(int B, int U, int k) {
    if (B < 1 || U < 1 || k < 1) return 0;
    int slices = bt_slices(U, B);
    const int s2 = frontend_bwd_simt_slices(B);
    if (s2 > slices) slices = s2;
    return (size_t)slices * U * (4 * k + 2) * sizeof(double);
}

extern "C" int xnn_frontend_bwd(const uint32_t* planes, int B, int L, int U, int k, const float* dy,
                                const float* zmax, const uint8_t* argmax, const float* gamma,
                                const double* stats, const double* cov_w, const long long* cnt1, long long n_pos,
                                const float* act_scale, const float* act_shift,
                                float* dconv_w, float* dconv_b, float* dgamma, float* dbeta, void* work,
                                xnn_stream_t s) {
    XNN_CHECK_ARG(planes && dy && zmax && argmax && gamma && stats && cov_w && cnt1 && dconv_w && dgamma && dbeta && work,
                  "NULL pointer");
    XNN_CHECK_ARG(k >= 1 && k <= BT_KMAX && L >= k && U >= 1 && B >= 1 && n_pos >= 1, "bad sizes (kernel_size <= 32)");
    XNN_CHECK_ARG((act_scale == nullptr) == (act_shift == nullptr), "act_scale and act_shift go together");
    const int n = (L - k + 1) / XNN_POOL;
    XNN_CHECK_ARG(n >= 1, "sequence shorter than one pool window");
    int Lbuf = 0, nstages = 0;
    const size_t smem = bt_smem_bytes(n, &Lbuf, &nstages);
    XNN_CHECK_ARG(smem <= 227 * 1024, "sequence too long for the front-end backward staging buffers");
    cudaStream_t st = to_stream(s);
    // equal batch slices per M tile.  Measured alternatives (cfg3, B = 4096): half the slices for the partly filled
    // last tile (44 of 128 units) 0.65 -> 1.07 ms -- a tile costs the same whatever its fill, the issue / MMA chain
    // bounds it, not the producers' rows; skipping the dead row groups in the producers 0.65 -> 0.74 ms (the
    // extra predicates cost the instruction-bound loop more than the skipped stores save)
    const int slices = bt_slices(U, B);
    // MN-major Toeplitz view of the one-hot pair buffer: K-group (pool window) stride 112 B, N-chunk
    // stride 32 B (pinned by test_frontend_bwd_tensor_core_vs_brute_force).
    const uint32_t lbo = 112u, sbo = 32u;
    XNN_CUDA(cudaFuncSetAttribute(frontend_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(bt_mtiles(U), slices);
    frontend_bwd_tc_kernel<<<grid, BT_THREADS, smem, st>>>(planes, B, L, planes_words(L), U, k, n, bt_npad(k), Lbuf,
                                                          nstages, dy, zmax, argmax, stats, lbo, sbo, act_scale,
                                                          act_shift, (double*)work);
    XNN_CHECK_LAUNCH();
    if (launch_frontend_bwd_finalize((const double*)work, slices, U, k, gamma, stats, cov_w, cnt1, n_pos, dconv_w,
                                     dconv_b, dgamma, dbeta, st)) {
        set_error("xnn_frontend_bwd: finalize launch failed");
        return XNN_ERR_CUDA;
    }
    return XNN_OK;
}

// a3 on the 5th-generation tensor cores -- the conv front end as a Toeplitz GEMM.
//
//   z[u,p] = sum_{j<k, c<4} W[u,c,j] * x[c,p+j]        (explainn/architectures/__init__.py:73-78,111)
//
// With the sequence held in shared memory position-major as bf16 one-hot,
// xs[4*p + c] = x[c,p] (8 bytes per position, an N base is 8 zero bytes), row p of
// the im2col matrix is the 4k contiguous elements starting at xs[4*p]: the matrix
// never has to be built.  A K-major UMMA operand without swizzle addresses
//   row r, 16-byte K chunk c  ->  start + 16*r + LBO*c      (8-row groups SBO = 128 B apart)
// so with LBO = 16 B the descriptor *is* the Toeplitz view: row r <-> position
// P + 2r (two positions per 16 bytes), chunk c <-> taps 2c, 2c+1.  Even and odd
// positions are two MMAs (the odd one reads a copy of xs shifted by one position so
// that its start stays 16-byte aligned) into two TMEM column ranges.
//
//   D[unit, position] (fp32, TMEM) = A[unit, 4j+c] (weights, bf16 x3 split) . B[position, 4j+c]^T (one-hot)
//
// fp32 parity: the one-hot operand is exact in bf16; each weight is split into
// hi + mid + lo bf16 pieces (3 x 8 mantissa bits = fp32's 24), every product is
// exact and accumulation is fp32 in TMEM -> z matches an fp32 gather-sum to rounding.
// Units sit on TMEM lanes (one thread = one unit), positions on columns, so
// BatchNorm-affine + max over 7 + exp (K3-K5) run in registers straight from
// tcgen05.ld: since exp and a positive affine are monotone, the pooled value is
// exp(|s|*max(sgn(s) z) + t) -- one exp per window; sgn(s) is folded into A.
#include "common.cuh"
#include "umma.cuh"
#include <cuda_bf16.h>
using namespace xnn;

#define TC_KMAX 32          // largest kernel_size
#define TC_ROWS 128         // units per CTA = UMMA M
#define TC_WIN 32           // pool windows per tile
#define TC_TILE_POS (TC_WIN * XNN_POOL)   // 224 positions per tile
#define TC_NMAX (TC_TILE_POS / 2)         // UMMA N per parity (112)
#define TC_ODD_COL 128      // TMEM column of the odd-position accumulator
#define TC_STAGE_COLS 256   // TMEM columns of one accumulator stage (even + odd)
#define TC_TMEM_COLS 512    // two accumulator stages
#define TC_GROUP_WARPS 8
#define TC_THREADS (32 * (2 * TC_GROUP_WARPS + 2))   // two epilogue groups + issuer warp + one-hot warp

static inline int tc_kpad(int k) { return ((4 * k + 15) / 16) * 16; }
static inline int tc_mtiles(int U) { return (U + TC_ROWS - 1) / TC_ROWS; }
static inline size_t tc_split_bytes(int k) { return (size_t)3 * (tc_kpad(k) / 8) * 2048; }  // per M tile

extern "C" size_t xnn_frontend_work_bytes(int U, int k) {
    if (U < 1 || k < 1 || k > TC_KMAX) return 0;
    return (size_t)tc_mtiles(U) * tc_split_bytes(k);
}

// ---------------------------------------------------------------------------
// weights -> three bf16 pieces in the UMMA K-major core-matrix layout:
// element (row r, feature f = 4j+c) of piece s at
//   ((s*chunks + f/8) * 16 + r/8) * 64 + (r%8)*8 + f%8            (bf16 units, per M tile)
// i.e. 8x8 core matrices of 128 contiguous bytes, SBO = 128 B, LBO = 2048 B.
// The sign of the BatchNorm scale is folded in (see header).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
tc_split_weights_kernel(const float* __restrict__ conv_w, const float* __restrict__ scale, int U, int k, int kpad,
                        __nv_bfloat16* __restrict__ out) {
    const int chunks = kpad / 8;
    const long long per_tile = (long long)3 * chunks * 1024;  // bf16 elements
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)TC_ROWS * kpad;
         i += (long long)gridDim.x * blockDim.x) {
        const int mt = blockIdx.y;
        const int r = (int)(i / kpad), f = (int)(i % kpad);
        const int u = mt * TC_ROWS + r, j = f >> 2, c = f & 3;
        float w = 0.0f;
        if (u < U && j < k) {
            w = conv_w[((long long)u * 4 + c) * k + j];
            if (scale[u] < 0.0f) w = -w;
        }
        const __nv_bfloat16 hi = __float2bfloat16_rn(w);
        const float r1 = w - __bfloat162float(hi);
        const __nv_bfloat16 mid = __float2bfloat16_rn(r1);
        const float r2 = r1 - __bfloat162float(mid);
        const __nv_bfloat16 lo = __float2bfloat16_rn(r2);
        __nv_bfloat16* o = out + mt * per_tile;
        const long long e = ((long long)(f >> 3) * 16 + (r >> 3)) * 64 + (r & 7) * 8 + (f & 7);
        // piece order lo, mid, hi: the small pieces enter the fp32 accumulator first
        o[(long long)0 * chunks * 1024 + e] = lo;
        o[(long long)1 * chunks * 1024 + e] = mid;
        o[(long long)2 * chunks * 1024 + e] = hi;
    }
}

// one-hot bf16 x4 of position i of a sequence held as planes (lo, hi, N mask)
__device__ __forceinline__ uint2 onehot_bf16x4(const uint32_t* __restrict__ sp, int Wp, int L, int i) {
    uint2 v = make_uint2(0u, 0u);
    if (i < L) {
        const int w = i >> 5, b = i & 31;
        const uint32_t lo = (sp[w] >> b) & 1u, hi = (sp[Wp + w] >> b) & 1u, nm = (sp[2 * Wp + w] >> b) & 1u;
        if (!nm) {
            const uint32_t half = lo ? 0x3F800000u : 0x00003F80u;  // bf16 1.0 in the upper / lower half
            if (hi) v.y = half; else v.x = half;
        }
    }
    return v;
}

struct TcShared {
    uint64_t acc_full[2];    // issuer -> epilogue group: the tile's MMAs completed (tcgen05.commit)
    uint64_t acc_empty[2];   // epilogue group -> issuer: accumulator stage read out (8 warp arrivals)
    uint64_t oh_full[2];     // one-hot warp -> issuer: a sequence's bf16 one-hot buffers written
    uint64_t oh_empty[2];    // issuer -> one-hot warp: the MMAs of that sequence completed
    uint32_t tmem_slot;
    uint32_t pad;
};

__device__ __forceinline__ void group_barrier(int grp) {
    asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "r"(32 * TC_GROUP_WARPS) : "memory");
}

// ---------------------------------------------------------------------------
// fused forward: conv (tcgen05) + BN affine + max over 7 (+ exp on the maximum).
// One CTA per SM, grid (M tiles, batch slices), warp-specialised over mbarriers:
//   warps 0-15  two epilogue groups of 8; group g drains TMEM accumulator stage g (tiles alternate).
//               Inside a group warp wi owns TMEM lanes 32(wi%4).. (one thread = one unit) and the
//               window half wi/4 of the tile; 8 windows (28 even + 28 odd columns) are read at a time,
//               max-of-7 is a tree, one expf per window; results are transposed through shared memory
//               so that each global store is 32 consecutive windows of one unit.
//   warp 16     issuer: every lane runs the loop, elect.sync issues the MMAs (uniform descriptors)
//   warp 17     expands the next sequence's planes into the bf16 one-hot buffers (double buffered)
// The tensor core works on tile g+1 while group g%2 is still in the epilogue of tile g.
// ---------------------------------------------------------------------------
template <bool TRAIN>
__global__ void __launch_bounds__(TC_THREADS, 1)
frontend_tc_kernel(const uint32_t* __restrict__ planes, int B, int L, int Wp, int U, int n, int kpad, int Lbuf,
                   int noh, const __nv_bfloat16* __restrict__ wsplit, const float* __restrict__ scale,
                   const float* __restrict__ shift, float* __restrict__ pool, float* __restrict__ zmax,
                   uint8_t* __restrict__ argmax) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int chunks = kpad >> 3, ksteps = kpad >> 4;
    const uint32_t a_bytes = 3u * chunks * 2048u;
    constexpr uint32_t STG = TRAIN ? (2u * TC_ROWS * 33u * 4u + TC_ROWS * 36u) : (TC_ROWS * 33u * 4u);
    unsigned char* a_smem = smem_raw;                                        // [3][chunks][16][8][8] bf16
    uint2* oh = reinterpret_cast<uint2*>(smem_raw + a_bytes);                // [noh][2][Lbuf]: position i / i+1
    unsigned char* stg = reinterpret_cast<unsigned char*>(oh + (size_t)noh * 2 * Lbuf);   // [2 groups][STG]
    uint32_t* s_pl = reinterpret_cast<uint32_t*>(stg + 2 * STG);             // [3 * Wp] planes of one sequence
    TcShared* sh = reinterpret_cast<TcShared*>(reinterpret_cast<unsigned char*>(s_pl) + ((3 * Wp * 4 + 15) & ~15));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int u0 = blockIdx.x * TC_ROWS;
    const int nu = min(TC_ROWS, U - u0);
    const int ns = gridDim.y;

    // ---- one-time setup: weights -> smem, TMEM, barriers
    {
        const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const unsigned char*>(wsplit) +
                                                          (size_t)blockIdx.x * a_bytes);
        uint4* dst = reinterpret_cast<uint4*>(a_smem);
        for (uint32_t i = tid; i < a_bytes / 16; i += TC_THREADS) dst[i] = src[i];
    }
    if (warp == 0) umma::tmem_alloc(&sh->tmem_slot, TC_TMEM_COLS);
    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            umma::mbar_init(&sh->acc_full[s], 1);
            umma::mbar_init(&sh->acc_empty[s], TC_GROUP_WARPS);
            umma::mbar_init(&sh->oh_full[s], 1);
            umma::mbar_init(&sh->oh_empty[s], 1);
        }
    }
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = sh->tmem_slot;
    const int ntiles = (n + TC_WIN - 1) / TC_WIN;
    const int nseq = (B - 1 - (int)blockIdx.y) / ns + 1;
    const int gtotal = nseq * ntiles;

    if (warp == 2 * TC_GROUP_WARPS) {
        // ================= issuer warp =================
        const uint32_t a_lo = umma::desc_lo(umma::smem_u32(a_smem), 2048u);
        const uint32_t b_lo = umma::desc_lo(umma::smem_u32(oh), 16u);
        const uint32_t hi128 = umma::desc_hi(128u);
        int g = 0;
        for (int sq = 0; sq < nseq; ++sq) {
            const int so = sq % noh;
            umma::mbar_wait(&sh->oh_full[so], (uint32_t)(sq / noh) & 1u);
            const uint32_t b_sq = b_lo + (uint32_t)so * (uint32_t)Lbuf;          // 2 * Lbuf * 8 B >> 4
            for (int t = 0; t < ntiles; ++t, ++g) {
                const uint32_t stage = (uint32_t)g & 1u;
                umma::mbar_wait(&sh->acc_empty[stage], (((uint32_t)g >> 1) & 1u) ^ 1u);
                umma::fence_after_sync();
                const int wins = min(TC_WIN, n - t * TC_WIN);
                const int N = ((((wins * XNN_POOL + 1) >> 1) + 15) & ~15);
                const uint32_t idesc = umma::idesc_bf16_f32(TC_ROWS, N);
#pragma unroll
                for (int par = 0; par < 2; ++par) {
                    const uint32_t b_t = b_sq + (uint32_t)par * (uint32_t)(Lbuf >> 1) + 112u * (uint32_t)t;
                    const uint32_t d = tmem + stage * TC_STAGE_COLS + (par ? TC_ODD_COL : 0);
                    uint32_t acc = 0;
#pragma unroll
                    for (int s3 = 0; s3 < 3; ++s3) {
#pragma unroll
                        for (int ks = 0; ks < TC_KMAX / 4; ++ks) {
                            if (ks < ksteps) {
                                const uint64_t ad = ((uint64_t)hi128 << 32) |
                                                    (uint64_t)(a_lo + (uint32_t)(s3 * chunks + 2 * ks) * 128u);
                                const uint64_t bd = ((uint64_t)hi128 << 32) | (uint64_t)(b_t + 2u * ks);
                                umma::mma_bf16_elect(d, ad, bd, idesc, acc);
                                acc = 1;
                            }
                        }
                    }
                }
                umma::commit_elect(&sh->acc_full[stage]);
                if (t == ntiles - 1) umma::commit_elect(&sh->oh_empty[so]);
                __syncwarp();
            }
        }
    } else if (warp == 2 * TC_GROUP_WARPS + 1) {
        // ================= one-hot warp =================
        for (int sq = 0; sq < nseq; ++sq) {
            const int so = sq % noh;
            const uint32_t* sp = planes + (long long)(blockIdx.y + (long long)sq * ns) * 3 * Wp;
            __syncwarp();
            for (int i = lane; i < 3 * Wp; i += 32) s_pl[i] = __ldg(sp + i);
            if (sq >= noh) umma::mbar_wait(&sh->oh_empty[so], (uint32_t)(sq / noh - 1) & 1u);
            __syncwarp();
            uint2* buf0 = oh + (size_t)so * 2 * Lbuf;
            uint2* buf1 = buf0 + Lbuf;
            for (int i = lane; i < Lbuf; i += 32) {
                const uint2 v = onehot_bf16x4(s_pl, Wp, L, i);
                buf0[i] = v;
                if (i > 0) buf1[i - 1] = v;
                if (i == Lbuf - 1) buf1[i] = make_uint2(0u, 0u);
            }
            umma::fence_proxy_async();
            __syncwarp();
            if (lane == 0) umma::mbar_arrive(&sh->oh_full[so]);
        }
    } else {
        // ================= epilogue groups =================
        const int grp = warp / TC_GROUP_WARPS, wi = warp % TC_GROUP_WARPS;
        const int quarter = wi & 3, half = wi >> 2;
        const int row = quarter * 32 + lane;
        const int u = u0 + row;
        const bool rows_live = quarter * 32 < nu;
        float asc = 0.0f, shf = 0.0f, sgn = 1.0f;
        if (u < U) {
            const float sc = scale[u];
            asc = fabsf(sc); shf = shift[u]; sgn = (sc < 0.0f) ? -1.0f : 1.0f;
        }
        float* s_pool = reinterpret_cast<float*>(stg + (size_t)grp * STG);            // [128][33]
        float* s_z = s_pool + TC_ROWS * 33;                                           // [128][33] (train)
        uint8_t* s_arg = reinterpret_cast<uint8_t*>(s_pool + 2 * TC_ROWS * 33);       // [128][36] (train)
        const uint32_t trow = tmem + (uint32_t)grp * TC_STAGE_COLS + ((uint32_t)(quarter * 32) << 16);
        int sq = 0, t = grp;
        while (t >= ntiles) { t -= ntiles; ++sq; }
        for (int g = grp; g < gtotal; g += 2) {
            const int q0 = t * TC_WIN;
            const int wins = min(TC_WIN, n - q0);
            umma::mbar_wait(&sh->acc_full[grp], ((uint32_t)g >> 1) & 1u);
            umma::fence_after_sync();
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                const int c = 2 * half + cc;                       // 8 windows = columns 28c .. 28c+27
                const bool live = rows_live && 8 * c < wins;
                uint32_t E[28], O[28];
                if (live) {
                    umma::tmem_ld28(trow + 28 * c, E);
                    umma::tmem_ld28(trow + TC_ODD_COL + 28 * c, O);
                    umma::tmem_ld_wait();
                }
                if (cc == 1) {
                    // this warp's share of the accumulator is in registers: hand the stage back
                    umma::fence_before_sync();
                    __syncwarp();
                    if (lane == 0) umma::mbar_arrive(&sh->acc_empty[grp]);
                }
                if (live) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        float z[7];
#pragma unroll
                        for (int r = 0; r < 7; ++r) {
                            const int p = 7 * j + r;
                            z[r] = __uint_as_float((p & 1) ? O[p >> 1] : E[p >> 1]);
                        }
                        const float best = fmaxf(fmaxf(fmaxf(z[0], z[1]), fmaxf(z[2], z[3])),
                                                 fmaxf(fmaxf(z[4], z[5]), z[6]));
                        const int q = 8 * c + j;
                        s_pool[row * 33 + q] = expf(fmaf(asc, best, shf));
                        if (TRAIN) {
                            int arg = 6;                           // first maximum wins
#pragma unroll
                            for (int r = 5; r >= 0; --r) arg = (z[r] == best) ? r : arg;
                            s_z[row * 33 + q] = sgn * best;
                            s_arg[row * 36 + q] = (uint8_t)arg;
                        }
                    }
                }
            }
            group_barrier(grp);                                    // tile staged
            {
                const long long b = blockIdx.y + (long long)sq * ns;
                const bool lq = lane < wins;
#pragma unroll 4
                for (int rr = 0; rr < 16; ++rr) {
                    const int r = 16 * wi + rr;
                    if (r < nu && lq) {
                        const long long gi = (b * U + u0 + r) * n + q0 + lane;
                        pool[gi] = s_pool[r * 33 + lane];
                        if (TRAIN) { zmax[gi] = s_z[r * 33 + lane]; argmax[gi] = s_arg[r * 36 + lane]; }
                    }
                }
            }
            group_barrier(grp);                                    // staging free again
            t += 2;
            while (t >= ntiles) { t -= ntiles; ++sq; }
        }
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, TC_TMEM_COLS);
}

static size_t tc_smem_bytes(int k, int L, int n, bool train, int noh, int* Lbuf_out) {
    const int ntiles = (n + TC_WIN - 1) / TC_WIN;
    const int Lbuf = ntiles * TC_TILE_POS + 64;   // last tile reads <= P + 2*112 + 2*chunks + 2 positions
    if (Lbuf_out) *Lbuf_out = Lbuf;
    const size_t stg = train ? (size_t)2 * TC_ROWS * 33 * 4 + (size_t)TC_ROWS * 36 : (size_t)TC_ROWS * 33 * 4;
    const int Wp = (L + 31) / 32;
    return tc_split_bytes(k) + (size_t)noh * 2 * Lbuf * 8 + 2 * stg + (size_t)((3 * Wp * 4 + 15) & ~15) +
           sizeof(TcShared) + 16;
}

extern "C" int xnn_frontend_fwd(const uint32_t* planes, int B, int L, int U, int k, const float* conv_w,
                                const float* scale, const float* shift, float* pool, float* zmax,
                                uint8_t* argmax, void* work, xnn_stream_t s) {
    XNN_CHECK_ARG(planes && conv_w && scale && shift && pool && work, "NULL pointer");
    XNN_CHECK_ARG((zmax == nullptr) == (argmax == nullptr), "zmax and argmax go together");
    XNN_CHECK_ARG(k >= 1 && k <= TC_KMAX && L >= k && U >= 1 && B >= 0, "bad sizes (kernel_size <= 32)");
    XNN_CHECK_ARG(((uintptr_t)work & 15) == 0, "work must be 16-byte aligned");
    const int n = (L - k + 1) / XNN_POOL;
    if (B == 0 || n == 0) return XNN_OK;
    const int kpad = tc_kpad(k), mt = tc_mtiles(U);
    const bool train = (zmax != nullptr);
    int Lbuf = 0, noh = 2;
    size_t smem = tc_smem_bytes(k, L, n, train, noh, &Lbuf);
    if (smem > 227 * 1024) { noh = 1; smem = tc_smem_bytes(k, L, n, train, noh, &Lbuf); }
    XNN_CHECK_ARG(smem <= 227 * 1024, "sequence too long for the front-end staging buffers");
    cudaStream_t st = to_stream(s);
    tc_split_weights_kernel<<<dim3(ceil_div((long long)TC_ROWS * kpad, 256), mt), 256, 0, st>>>(
        conv_w, scale, U, k, kpad, reinterpret_cast<__nv_bfloat16*>(work));
    XNN_CHECK_LAUNCH();
    int slices = num_sms() / mt;                  // one CTA per SM: each holds all 512 TMEM columns
    if (slices < 1) slices = 1;
    if (slices > B) slices = B;
    dim3 grid(mt, slices);
    if (train) {
        XNN_CUDA(cudaFuncSetAttribute(frontend_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        frontend_tc_kernel<true><<<grid, TC_THREADS, smem, st>>>(planes, B, L, planes_words(L), U, n, kpad, Lbuf, noh,
                                                                 reinterpret_cast<const __nv_bfloat16*>(work), scale,
                                                                 shift, pool, zmax, argmax);
    } else {
        XNN_CUDA(cudaFuncSetAttribute(frontend_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        frontend_tc_kernel<false><<<grid, TC_THREADS, smem, st>>>(planes, B, L, planes_words(L), U, n, kpad, Lbuf, noh,
                                                                  reinterpret_cast<const __nv_bfloat16*>(work), scale,
                                                                  shift, pool, zmax, argmax);
    }
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

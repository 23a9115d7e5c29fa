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
#define TC_TMEM_COLS 256
#define TC_THREADS 128

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
    uint64_t bar;
    uint32_t tmem_slot;
    uint32_t pad;
};

// ---------------------------------------------------------------------------
// fused forward: conv (tcgen05) + BN affine + max over 7 (+ exp on the maximum)
// grid (M tiles, batch slices); 128 threads = 4 warps, warp w owns TMEM lanes 32w..32w+31
// ---------------------------------------------------------------------------
template <bool TRAIN>
__global__ void __launch_bounds__(TC_THREADS)
frontend_tc_kernel(const uint32_t* __restrict__ planes, int B, int L, int Wp, int U, int n, int kpad, int Lbuf,
                   const __nv_bfloat16* __restrict__ wsplit, const float* __restrict__ scale,
                   const float* __restrict__ shift, float* __restrict__ pool, float* __restrict__ zmax,
                   uint8_t* __restrict__ argmax) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int chunks = kpad >> 3, ksteps = kpad >> 4;
    const uint32_t a_bytes = 3u * chunks * 2048u;
    unsigned char* a_smem = smem_raw;                                        // [3][chunks][16][8][8] bf16
    uint2* buf0 = reinterpret_cast<uint2*>(smem_raw + a_bytes);              // [Lbuf] one-hot of position i
    uint2* buf1 = buf0 + Lbuf;                                               // [Lbuf] one-hot of position i+1
    float* s_pool = reinterpret_cast<float*>(buf1 + Lbuf);                   // [128][33]
    float* s_z = s_pool + TC_ROWS * 33;                                      // [128][33] (train)
    uint8_t* s_arg = reinterpret_cast<uint8_t*>(s_z + (TRAIN ? TC_ROWS * 33 : 0));  // [128][36] (train)
    TcShared* sh = reinterpret_cast<TcShared*>(s_arg + (TRAIN ? TC_ROWS * 36 : 0));

    const int tid = threadIdx.x, warp = tid >> 5;
    const int u0 = blockIdx.x * TC_ROWS;
    const int nu = min(TC_ROWS, U - u0);
    const int u = u0 + tid;

    // ---- one-time setup: weights -> smem, TMEM, barrier
    {
        const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const unsigned char*>(wsplit) +
                                                          (size_t)blockIdx.x * a_bytes);
        uint4* dst = reinterpret_cast<uint4*>(a_smem);
        for (uint32_t i = tid; i < a_bytes / 16; i += TC_THREADS) dst[i] = src[i];
    }
    if (warp == 0) umma::tmem_alloc(&sh->tmem_slot, TC_TMEM_COLS);
    if (tid == 0) umma::mbar_init(&sh->bar, 1);
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = sh->tmem_slot;
    const uint32_t a_addr = umma::smem_u32(a_smem);
    const uint32_t b0_addr = umma::smem_u32(buf0), b1_addr = umma::smem_u32(buf1);

    float asc = 0.0f, shf = 0.0f, sgn = 1.0f;
    if (u < U) {
        const float sc = scale[u];
        asc = fabsf(sc); shf = shift[u]; sgn = (sc < 0.0f) ? -1.0f : 1.0f;
    }
    const int ntiles = (n + TC_WIN - 1) / TC_WIN;
    uint32_t phase = 0;

    auto issue_tile = [&](int t) {
        const int wins = min(TC_WIN, n - t * TC_WIN);
        const int rows = (wins * XNN_POOL + 1) >> 1;
        const int N = (rows + 15) & ~15;
        const uint32_t idesc = umma::idesc_bf16_f32(TC_ROWS, N);
        const uint32_t boff = 8u * (uint32_t)(t * TC_TILE_POS);
        umma::fence_after_sync();
#pragma unroll 1
        for (int par = 0; par < 2; ++par) {
            const uint32_t bbase = (par ? b1_addr : b0_addr) + boff;
            const uint32_t d = tmem + (par ? TC_ODD_COL : 0);
            uint32_t acc = 0;
#pragma unroll 1
            for (int s3 = 0; s3 < 3; ++s3) {
#pragma unroll 1
                for (int ks = 0; ks < ksteps; ++ks) {
                    const uint64_t ad = umma::smem_desc(a_addr + (uint32_t)(s3 * chunks + 2 * ks) * 2048u, 2048u, 128u);
                    const uint64_t bd = umma::smem_desc(bbase + 32u * ks, 16u, 128u);
                    umma::mma_bf16(d, ad, bd, idesc, acc);
                    acc = 1;
                }
            }
        }
        umma::commit(&sh->bar);
    };

    for (int b = blockIdx.y; b < B; b += gridDim.y) {
        // previous sequence: every MMA was waited for and every copy-out finished (trailing barrier)
        const uint32_t* sp = planes + (long long)b * 3 * Wp;
        for (int i = tid; i < Lbuf; i += TC_THREADS) {
            const uint2 v = onehot_bf16x4(sp, Wp, L, i);
            buf0[i] = v;
            if (i > 0) buf1[i - 1] = v;
            if (i == Lbuf - 1) buf1[i] = make_uint2(0u, 0u);
        }
        umma::fence_proxy_async();
        __syncthreads();
        if (tid == 0) issue_tile(0);

        for (int t = 0; t < ntiles; ++t) {
            const int q0 = t * TC_WIN;
            const int wins = min(TC_WIN, n - q0);
            const int steps = ((((wins * XNN_POOL + 1) >> 1) + 15) & ~15) >> 4;
            umma::mbar_wait(&sh->bar, phase);
            phase ^= 1u;
            umma::fence_after_sync();

            const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);
            float best = 0.0f;
            int arg = 0;
#pragma unroll
            for (int s = 0; s < 7; ++s) {
                if (s < steps) {
                    uint32_t E[16], O[16];
                    umma::tmem_ld16(trow + s * 16, E);
                    umma::tmem_ld16(trow + TC_ODD_COL + s * 16, O);
                    umma::tmem_ld_wait();
#pragma unroll
                    for (int m = 0; m < 32; ++m) {
                        const int tt = s * 32 + m, r = tt % XNN_POOL, q = tt / XNN_POOL;
                        const float z = __uint_as_float((m & 1) ? O[m >> 1] : E[m >> 1]);
                        if (r == 0) { best = z; arg = 0; }
                        else if (z > best) { best = z; arg = r; }   // first maximum wins
                        if (r == XNN_POOL - 1) {
                            s_pool[tid * 33 + q] = expf(fmaf(asc, best, shf));
                            if (TRAIN) { s_z[tid * 33 + q] = sgn * best; s_arg[tid * 36 + q] = (uint8_t)arg; }
                        }
                    }
                }
            }
            umma::fence_before_sync();
            __syncthreads();                       // staging complete, accumulator drained
            if (tid == 0 && t + 1 < ntiles) issue_tile(t + 1);   // tensor core runs under the copy-out

            for (int e = tid; e < nu * wins; e += TC_THREADS) {
                const int row = e / wins, q = e - row * wins;
                const long long g = ((long long)b * U + u0 + row) * n + q0 + q;
                pool[g] = s_pool[row * 33 + q];
                if (TRAIN) { zmax[g] = s_z[row * 33 + q]; argmax[g] = s_arg[row * 36 + q]; }
            }
            __syncthreads();                       // staging free again
        }
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, TC_TMEM_COLS);
}

static size_t tc_smem_bytes(int k, int n, bool train, int* Lbuf_out) {
    const int ntiles = (n + TC_WIN - 1) / TC_WIN;
    const int Lbuf = ntiles * TC_TILE_POS + 64;   // last tile reads <= P + 2*112 + 2*chunks + 2 positions
    if (Lbuf_out) *Lbuf_out = Lbuf;
    size_t s = tc_split_bytes(k) + (size_t)2 * Lbuf * 8 + (size_t)TC_ROWS * 33 * 4;
    if (train) s += (size_t)TC_ROWS * 33 * 4 + (size_t)TC_ROWS * 36;
    s += sizeof(TcShared) + 16;
    // at most two CTAs per SM: each holds TC_TMEM_COLS of the SM's 512 TMEM columns
    if (s < 80 * 1024) s = 80 * 1024;
    return s;
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
    int Lbuf = 0;
    const size_t smem = tc_smem_bytes(k, n, train, &Lbuf);
    XNN_CHECK_ARG(smem <= 227 * 1024, "sequence too long for the front-end staging buffers");
    cudaStream_t st = to_stream(s);
    tc_split_weights_kernel<<<dim3(ceil_div((long long)TC_ROWS * kpad, 256), mt), 256, 0, st>>>(
        conv_w, scale, U, k, kpad, reinterpret_cast<__nv_bfloat16*>(work));
    XNN_CHECK_LAUNCH();
    const int per_sm = (smem <= 113 * 1024) ? 2 : 1;
    int slices = (num_sms() * per_sm) / mt;
    if (slices < 1) slices = 1;
    if (slices > B) slices = B;
    dim3 grid(mt, slices);
    if (train) {
        XNN_CUDA(cudaFuncSetAttribute(frontend_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        frontend_tc_kernel<true><<<grid, TC_THREADS, smem, st>>>(planes, B, L, planes_words(L), U, n, kpad, Lbuf,
                                                                 reinterpret_cast<const __nv_bfloat16*>(work), scale,
                                                                 shift, pool, zmax, argmax);
    } else {
        XNN_CUDA(cudaFuncSetAttribute(frontend_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        frontend_tc_kernel<false><<<grid, TC_THREADS, smem, st>>>(planes, B, L, planes_words(L), U, n, kpad, Lbuf,
                                                                  reinterpret_cast<const __nv_bfloat16*>(work), scale,
                                                                  shift, pool, zmax, argmax);
    }
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// a6 on the tensor cores -- the conv weight gradient as a Toeplitz GEMM.
//
//   G_u[c,j] = sum_{b,q} dy[b,u,q] * x[b, c, 7q + arg[b,u,q] + j]
//
// (explainn/architectures/__init__.py:73-83 differentiated; selene/__init__.py:291).  dy is
// the gradient at the max-pool arg-max, so with g[u,p] = dy at p = 7q+arg and 0 elsewhere
//
//   D[unit, f = 4j+c] (fp32, TMEM) = A[unit, p] (g, bf16 x3 split) . B[p, f] (one-hot window)
//
// B is never built: with the sequence position-major in shared memory as bf16 one-hot
// (8 bytes per position, the forward kernel's layout) element (p, f) sits at byte 8p + 2f.
// An MN-major UMMA operand without swizzle addresses
//   K row r, 16-byte N chunk c  ->  start + 16*(r%8) + LBO*(r/8) + SBO*c
// so LBO = 128, SBO = 16 is the Toeplitz view with K row r <-> position P + 2r; odd
// positions are a second MMA on a copy of the buffer shifted by one position.
// The accumulator stays in TMEM for the CTA's whole slice of the batch and is read once.
//
// A is the sparse g (one non-zero per pool window) scattered into a zeroed K-major tile of
// 4 windows = 28 positions = 14 (padded to 16) K rows per parity, three bf16 pieces each.
#include "common.cuh"
#include "umma.cuh"
#include <cuda_bf16.h>
using namespace xnn;

#define BT_KMAX 32
#define BT_ROWS 128                 // units per CTA = UMMA M
#define BT_WIN 4                    // pool windows per tile
#define BT_TILE_POS (BT_WIN * XNN_POOL)   // 28 positions
#define BT_SPLITS 3
#define BT_BLOCK_BYTES 4096         // one (piece, parity) A block: K = 16 -> 2 chunks of 2048 B
#define BT_STAGE_BYTES (BT_SPLITS * 2 * BT_BLOCK_BYTES)
#define BT_THREADS 128
#define BT_RG 4                     // rows per staging group (loads in flight)

static inline int bt_npad(int k) { return ((4 * k + 15) / 16) * 16; }
static inline int bt_mtiles(int U) { return (U + BT_ROWS - 1) / BT_ROWS; }

static int bt_slices(int B, int U) {
    int s = (num_sms() * 2) / bt_mtiles(U);
    if (s < 1) s = 1;
    if (s > B) s = B;
    return s;
}

struct BtShared {
    uint64_t bar[2];
    uint32_t tmem_slot;
    uint32_t pad;
};

__device__ __forceinline__ uint2 bt_onehot_bf16x4(const uint32_t* __restrict__ sp, int Wp, int L, int i) {
    uint2 v = make_uint2(0u, 0u);
    if (i < L) {
        const int w = i >> 5, b = i & 31;
        const uint32_t lo = (sp[w] >> b) & 1u, hi = (sp[Wp + w] >> b) & 1u, nm = (sp[2 * Wp + w] >> b) & 1u;
        if (!nm) {
            const uint32_t half = lo ? 0x3F800000u : 0x00003F80u;
            if (hi) v.y = half; else v.x = half;
        }
    }
    return v;
}

__global__ void __launch_bounds__(BT_THREADS)
frontend_bwd_tc_kernel(const uint32_t* __restrict__ planes, int B, int L, int Wp, int U, int k, int n, int npad,
                       int Lbuf, int dstride, int astride,
                       const float* __restrict__ dy, const float* __restrict__ zmax,
                       const uint8_t* __restrict__ argmax, const double* __restrict__ stats,
                       uint32_t b_lbo, uint32_t b_sbo,
                       double* __restrict__ part /* [slices][U][4k+2] */) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* g_smem = smem_raw;                                              // [2 stages][pieces][2][4096]
    uint2* buf0 = reinterpret_cast<uint2*>(smem_raw + 2 * BT_STAGE_BYTES);         // [Lbuf]
    uint2* buf1 = buf0 + Lbuf;                                                     // [Lbuf] shifted by one
    float* s_d = reinterpret_cast<float*>(buf1 + Lbuf);                            // [128][dstride]
    float* s_sb = s_d + BT_ROWS * dstride;                                         // [128]
    float* s_sg = s_sb + BT_ROWS;                                                  // [128]
    float* s_ws = s_sg + BT_ROWS;                                                  // [128]
    uint8_t* s_arg = reinterpret_cast<uint8_t*>(s_ws + BT_ROWS);                   // [128][astride]
    BtShared* sh = reinterpret_cast<BtShared*>(s_arg + ((BT_ROWS * astride + 15) & ~15));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int u0 = blockIdx.x * BT_ROWS;
    const int nu = min(BT_ROWS, U - u0);
    const int u = u0 + tid;
    const bool row_ok = tid < nu;

    // ---- one-time setup
    {
        uint4* z = reinterpret_cast<uint4*>(g_smem);
        for (int i = tid; i < 2 * BT_STAGE_BYTES / 16; i += BT_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    s_ws[tid] = row_ok ? (float)stats[4 * u + 3] : 0.0f;
    if (warp == 0) umma::tmem_alloc(&sh->tmem_slot, 128);
    if (tid == 0) { umma::mbar_init(&sh->bar[0], 1); umma::mbar_init(&sh->bar[1], 1); }
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = sh->tmem_slot;
    const uint32_t g_addr = umma::smem_u32(g_smem);
    const uint32_t b0_addr = umma::smem_u32(buf0), b1_addr = umma::smem_u32(buf1);
    const uint32_t idesc = umma::idesc_bf16_f32(BT_ROWS, npad) | (1u << 16);   // B is MN-major

    const int ntiles = (n + BT_WIN - 1) / BT_WIN;
    uint32_t uses0 = 0, uses1 = 0, waited0 = 0, waited1 = 0;   // commits issued / completions observed per stage
    uint32_t any_mma = 0;
    double sb = 0.0, sg = 0.0;
    const double invstd = row_ok ? stats[4 * u + 2] : 0.0;

    for (int b = blockIdx.y; b < B; b += gridDim.y) {
        // every MMA of the previous sequence must be done before its one-hot buffers are rewritten
        while (waited0 < uses0) { umma::mbar_wait(&sh->bar[0], waited0 & 1u); ++waited0; }
        while (waited1 < uses1) { umma::mbar_wait(&sh->bar[1], waited1 & 1u); ++waited1; }
        umma::fence_after_sync();
        __syncthreads();   // all threads finished reading s_d / s_arg of the previous sequence

        const uint32_t* sp = planes + (long long)b * 3 * Wp;
        for (int i = tid; i < Lbuf; i += BT_THREADS) {
            const uint2 v = bt_onehot_bf16x4(sp, Wp, L, i);
            buf0[i] = v;
            if (i > 0) buf1[i - 1] = v;
            if (i == Lbuf - 1) buf1[i] = make_uint2(0u, 0u);
        }
        // stage dy / argmax rows (a warp per row, BT_RG rows of loads in flight), and the per-row
        // BatchNorm reductions  sum dy  and  sum dy * (z - mean_nobias)
        for (int row0 = warp * BT_RG; row0 < nu; row0 += 4 * BT_RG) {
            float a_d[BT_RG], a_dz[BT_RG];
#pragma unroll
            for (int r = 0; r < BT_RG; ++r) { a_d[r] = 0.0f; a_dz[r] = 0.0f; }
            for (int q0 = 0; q0 < n; q0 += 96) {
                float d[BT_RG][3], z[BT_RG][3];
                uint8_t a[BT_RG][3];
#pragma unroll
                for (int r = 0; r < BT_RG; ++r) {
                    const long long g = ((long long)b * U + u0 + row0 + r) * n;
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const int q = q0 + 32 * c + lane;
                        const bool ok = (row0 + r < nu) && (q < n);
                        d[r][c] = ok ? __ldg(dy + g + q) : 0.0f;
                        z[r][c] = ok ? __ldg(zmax + g + q) : 0.0f;
                        a[r][c] = ok ? __ldg(argmax + g + q) : (uint8_t)0;
                    }
                }
#pragma unroll
                for (int r = 0; r < BT_RG; ++r) {
                    const float wsr = s_ws[min(row0 + r, BT_ROWS - 1)];
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const int q = q0 + 32 * c + lane;
                        if ((row0 + r < nu) && (q < n)) {
                            s_d[(row0 + r) * dstride + q] = d[r][c];
                            s_arg[(row0 + r) * astride + q] = a[r][c];
                            a_d[r] += d[r][c];
                            a_dz[r] = fmaf(d[r][c], z[r][c] - wsr, a_dz[r]);
                        }
                    }
                }
            }
#pragma unroll
            for (int r = 0; r < BT_RG; ++r) {
                const float t1 = warp_sum(a_d[r]), t2 = warp_sum(a_dz[r]);
                if (lane == 0 && row0 + r < nu) { s_sb[row0 + r] = t1; s_sg[row0 + r] = t2; }
            }
        }
        umma::fence_proxy_async();
        __syncthreads();
        if (row_ok) { sb += (double)s_sb[tid]; sg += (double)s_sg[tid] * invstd; }

        for (int t = 0; t < ntiles; ++t) {
            const int stage = t & 1;
            if (stage == 0) { while (waited0 < uses0) { umma::mbar_wait(&sh->bar[0], waited0 & 1u); ++waited0; } }
            else            { while (waited1 < uses1) { umma::mbar_wait(&sh->bar[1], waited1 & 1u); ++waited1; } }
            umma::fence_after_sync();
            unsigned char* st = g_smem + stage * BT_STAGE_BYTES;
            // clear this thread's row of the stage, then scatter its (up to) 4 windows
#pragma unroll
            for (int c = 0; c < BT_SPLITS * 2 * 2; ++c)
                *reinterpret_cast<uint4*>(st + c * 2048 + 16 * tid) = make_uint4(0u, 0u, 0u, 0u);
            if (row_ok) {
#pragma unroll
                for (int w = 0; w < BT_WIN; ++w) {
                    const int q = t * BT_WIN + w;
                    if (q < n) {
                        const float dv = s_d[tid * dstride + q];
                        const int pl = w * XNN_POOL + (int)s_arg[tid * astride + q];
                        const int par = pl & 1, kk = pl >> 1;
                        const __nv_bfloat16 hi = __float2bfloat16_rn(dv);
                        const float r1 = dv - __bfloat162float(hi);
                        const __nv_bfloat16 mid = __float2bfloat16_rn(r1);
                        const __nv_bfloat16 lo = __float2bfloat16_rn(r1 - __bfloat162float(mid));
                        unsigned char* e = st + par * BT_BLOCK_BYTES + (kk >> 3) * 2048 + 16 * tid + (kk & 7) * 2;
                        *reinterpret_cast<__nv_bfloat16*>(e + 0 * 2 * BT_BLOCK_BYTES) = lo;
                        *reinterpret_cast<__nv_bfloat16*>(e + 1 * 2 * BT_BLOCK_BYTES) = mid;
                        *reinterpret_cast<__nv_bfloat16*>(e + 2 * 2 * BT_BLOCK_BYTES) = hi;
                    }
                }
            }
            umma::fence_proxy_async();
            __syncthreads();
            if (tid == 0) {
                umma::fence_after_sync();
                const uint32_t boff = 8u * (uint32_t)(t * BT_TILE_POS);
#pragma unroll 1
                for (int s3 = 0; s3 < BT_SPLITS; ++s3) {
#pragma unroll 1
                    for (int par = 0; par < 2; ++par) {
                        const uint64_t ad = umma::smem_desc(
                            g_addr + (uint32_t)(stage * BT_STAGE_BYTES + (s3 * 2 + par) * BT_BLOCK_BYTES), 2048u, 128u);
                        const uint64_t bd = umma::smem_desc((par ? b1_addr : b0_addr) + boff, b_lbo, b_sbo);
                        umma::mma_bf16(tmem, ad, bd, idesc, any_mma);
                        any_mma = 1;
                    }
                }
                umma::commit(&sh->bar[stage]);
            }
            if (stage == 0) ++uses0; else ++uses1;
        }
    }
    while (waited0 < uses0) { umma::mbar_wait(&sh->bar[0], waited0 & 1u); ++waited0; }
    while (waited1 < uses1) { umma::mbar_wait(&sh->bar[1], waited1 & 1u); ++waited1; }
    umma::fence_after_sync();

    // ---- read the accumulator once: thread = unit (TMEM lane), columns f = 4j + c
    const int K4 = 4 * k, stride = K4 + 2;
    double* o = part + ((long long)blockIdx.y * U + u) * stride;
    const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);
    const bool have = (uses0 + uses1) > 0;
    for (int c0 = 0; c0 < npad; c0 += 16) {
        uint32_t v[16];
        umma::tmem_ld16(trow + c0, v);
        umma::tmem_ld_wait();
        if (row_ok) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int f = c0 + i;
                if (f < K4) o[(f & 3) * k + (f >> 2)] = have ? (double)__uint_as_float(v[i]) : 0.0;
            }
        }
    }
    if (row_ok) { o[K4] = sb; o[K4 + 1] = sg; }
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

static size_t bt_smem_bytes(int n, int* Lbuf_out, int* dstride_out, int* astride_out) {
    const int ntiles = (n + BT_WIN - 1) / BT_WIN;
    const int Lbuf = ntiles * BT_TILE_POS + 64;            // even; last tile reads < P0 + 64 positions
    const int dstride = n | 1;
    int a4 = (n + 3) / 4;
    if ((a4 & 1) == 0) ++a4;
    const int astride = 4 * a4;
    if (Lbuf_out) *Lbuf_out = Lbuf;
    if (dstride_out) *dstride_out = dstride;
    if (astride_out) *astride_out = astride;
    return (size_t)2 * BT_STAGE_BYTES + (size_t)2 * Lbuf * 8 + (size_t)BT_ROWS * dstride * 4 + 3 * BT_ROWS * 4 +
           (((size_t)BT_ROWS * astride + 15) & ~(size_t)15) + sizeof(BtShared) + 16;
}

extern "C" size_t xnn_frontend_bwd_work_bytes(int B, int U, int k) {
    if (B < 1 || U < 1 || k < 1) return 0;
    int slices = bt_slices(B, U);
    const int s2 = frontend_bwd_simt_slices(B);
    if (s2 > slices) slices = s2;
    return (size_t)slices * U * (4 * k + 2) * sizeof(double);
}

extern "C" int xnn_frontend_bwd(const uint32_t* planes, int B, int L, int U, int k, const float* dy,
                                const float* zmax, const uint8_t* argmax, const float* gamma,
                                const double* stats, const double* cov_w, const long long* cnt1, long long n_pos,
                                float* dconv_w, float* dconv_b, float* dgamma, float* dbeta, void* work,
                                xnn_stream_t s) {
    XNN_CHECK_ARG(planes && dy && zmax && argmax && gamma && stats && cov_w && cnt1 && dconv_w && dgamma && dbeta && work,
                  "NULL pointer");
    XNN_CHECK_ARG(k >= 1 && k <= BT_KMAX && L >= k && U >= 1 && B >= 1 && n_pos >= 1, "bad sizes (kernel_size <= 32)");
    const int n = (L - k + 1) / XNN_POOL;
    XNN_CHECK_ARG(n >= 1, "sequence shorter than one pool window");
    int Lbuf = 0, dstride = 0, astride = 0;
    const size_t smem = bt_smem_bytes(n, &Lbuf, &dstride, &astride);
    XNN_CHECK_ARG(smem <= 227 * 1024, "sequence too long for the front-end backward staging buffers");
    cudaStream_t st = to_stream(s);
    const int slices = bt_slices(B, U);
    // MN-major Toeplitz view of the one-hot buffer: K-group stride 128 B, N-chunk stride 16 B.
    // XNN_BWD_DESC_SWAP=1 exchanges the two fields (bring-up aid for the descriptor convention).
    static const int swap = [] { const char* e = getenv("XNN_BWD_DESC_SWAP"); return e && e[0] == '1'; }();
    const uint32_t lbo = swap ? 16u : 128u, sbo = swap ? 128u : 16u;
    XNN_CUDA(cudaFuncSetAttribute(frontend_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(bt_mtiles(U), slices);
    frontend_bwd_tc_kernel<<<grid, BT_THREADS, smem, st>>>(planes, B, L, planes_words(L), U, k, n, bt_npad(k), Lbuf,
                                                          dstride, astride, dy, zmax, argmax, stats, lbo, sbo,
                                                          (double*)work);
    XNN_CHECK_LAUNCH();
    if (launch_frontend_bwd_finalize((const double*)work, slices, U, k, gamma, stats, cov_w, cnt1, n_pos, dconv_w,
                                     dconv_b, dgamma, dbeta, st)) {
        set_error("xnn_frontend_bwd: finalize launch failed");
        return XNN_ERR_CUDA;
    }
    return XNN_OK;
}

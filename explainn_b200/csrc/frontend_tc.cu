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

#define TC_KMAX 48          // largest kernel_size / PWM width
#define TC_ROWS 128         // units per CTA = UMMA M
#define TC_WIN 32           // pool windows per tile
#define TC_TILE_POS (TC_WIN * XNN_POOL)   // 224 positions per tile
#define TC_NMAX (TC_TILE_POS / 2)         // UMMA N per parity (112)
#define TC_ODD_COL 128      // TMEM column of the odd-position accumulator
#define TC_STAGE_COLS 256   // TMEM columns of one accumulator stage (even + odd)
#define TC_TMEM_COLS 512    // two accumulator stages
#define TC_EPI_WARPS 16
#define TC_THREADS (32 * (TC_EPI_WARPS + 3))   // epilogue warps + two issuer warps + one-hot warp

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
            if (scale != nullptr && scale[u] < 0.0f) w = -w;
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

__device__ __forceinline__ void epi_barrier() {
    asm volatile("bar.sync 1, %0;" ::"n"(32 * TC_EPI_WARPS) : "memory");
}

// contiguous copy shared -> global of `cnt` elements by ONE warp, asynchronously: the shared copy starts at the
// same 16-byte phase as the global destination (sbase is 16-byte aligned, element i sits at sbase[phase + i]),
// so everything between a scalar head and tail (< 16 bytes each, written by the lanes) is one bulk-async
// store (cp.async.bulk.global.shared::cta) issued by lane 0; the caller commits the group.
template <typename T>
__device__ __forceinline__ void bulk_copy_out(T* __restrict__ gdst, const T* __restrict__ sbase, int phase, int cnt,
                                              int lane) {
    constexpr int V = 16 / (int)sizeof(T);
    const T* src = sbase + phase;
    const int head = min((V - phase) & (V - 1), cnt);
    const int nvec = (cnt - head) / V;
    const int tail0 = head + nvec * V;
    if (lane < head) gdst[lane] = src[lane];
    if (tail0 + lane < cnt) gdst[tail0 + lane] = src[tail0 + lane];
    if (lane == 0 && nvec > 0)
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst + head),
                     "r"(umma::smem_u32(src + head)), "r"(nvec * 16)
                     : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// the same contiguous copy by the 512 epilogue threads with 16-byte vector loads / stores (tid = 0..511): the staging
// is free again as soon as every thread's shared-memory loads have returned (the next barrier), while a bulk store
// holds it until the engine has drained it -- measured ~3500 cycles per 95 KB sequence block, one quarter of the
// kernel, with every epilogue warp waiting at the barrier behind `cp.async.bulk.wait_group.read`
template <typename T>
__device__ __forceinline__ void vector_copy_out(T* __restrict__ gdst, const T* __restrict__ sbase, int phase, int cnt,
                                                int tid) {
    constexpr int V = 16 / (int)sizeof(T);
    const T* src = sbase + phase;
    const int head = min((V - phase) & (V - 1), cnt);
    const int nvec = (cnt - head) / V;
    const int tail0 = head + nvec * V;
    if (tid < head) gdst[tid] = src[tid];
    if (tail0 + tid < cnt) gdst[tail0 + tid] = src[tail0 + tid];
    const uint4* s4 = reinterpret_cast<const uint4*>(src + head);
    uint4* g4 = reinterpret_cast<uint4*>(gdst + head);
    for (int v = tid; v < nvec; v += 32 * TC_EPI_WARPS) g4[v] = s4[v];
}

// ---------------------------------------------------------------------------
// fused forward: conv (tcgen05) + BN affine + max over 7 (+ exp on the maximum).
// One CTA per SM, grid (M tiles, batch slices), warp-specialised over mbarriers:
//   warps 0-15  epilogue: warp w owns TMEM lanes 32(w%4).. (one thread = one unit) and windows 8(w/4)..+7 of
//               the tile (28 even + 28 odd accumulator columns), max-of-7 as a tree, one expf per window.
//               Results are staged in shared memory in the layout of the output itself: the (units of this
//               M tile) x n block of one sequence is ONE contiguous range of pool / zmax / argmax, so the
//               copy-out is one bulk-async store per array and sequence (`linear`); sequences too long for
//               that (or n a multiple of 4, whose row stride would serialise the staging stores on a few
//               banks) are staged and written tile by tile, one 32-window row segment per store.
//   warps 16-17 issuers (even / odd positions): every lane runs the loop, elect.sync issues the MMAs
//   warp 18     expands the next sequence's planes into the bf16 one-hot buffers (double buffered)
// Two TMEM accumulator stages: the tensor core works on tile g+1 during the epilogue of tile g.
// ---------------------------------------------------------------------------
template <bool TRAIN>
__global__ void __launch_bounds__(TC_THREADS, 1)
frontend_tc_kernel(const uint32_t* __restrict__ planes, int B, int L, int Wp, int U, int n, int kpad, int Lbuf,
                   int noh, int linear, const __nv_bfloat16* __restrict__ wsplit, const float* __restrict__ scale,
                   const float* __restrict__ shift, float* __restrict__ pool, float* __restrict__ zmax,
                   uint8_t* __restrict__ argmax, const __grid_constant__ TileMap tmap) {
    int bx, by, ns;
    tile_of_cta(tmap, bx, by, ns);
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int chunks = kpad >> 3, ksteps = kpad >> 4;
    const uint32_t a_bytes = 3u * chunks * 2048u;
    // staging: floats [rows][sstride] (+ 4 elements of phase slack), bytes [rows][bstride] (+ 16)
    const int sstride = linear ? n : 33, bstride = linear ? n : 36;
    const uint32_t f_bytes = (uint32_t)((TC_ROWS * sstride + 4) * 4 + 15) & ~15u;
    const uint32_t b_bytes = (uint32_t)(TC_ROWS * bstride + 16 + 15) & ~15u;
    unsigned char* a_smem = smem_raw;                                        // [3][chunks][16][8][8] bf16
    uint2* oh = reinterpret_cast<uint2*>(smem_raw + a_bytes);                // [noh][2][Lbuf]: position i / i+1
    unsigned char* stg = reinterpret_cast<unsigned char*>(oh + (size_t)noh * 2 * Lbuf);
    float* s_pool = reinterpret_cast<float*>(stg);
    float* s_z = reinterpret_cast<float*>(stg + f_bytes);                    // (train)
    uint8_t* s_arg = stg + 2 * f_bytes;                                      // (train)
    uint32_t* s_pl = reinterpret_cast<uint32_t*>(stg + (TRAIN ? 2 * f_bytes + b_bytes : f_bytes));
    TcShared* sh = reinterpret_cast<TcShared*>(reinterpret_cast<unsigned char*>(s_pl) + ((3 * Wp * 4 + 15) & ~15));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int u0 = bx * TC_ROWS;
    const int nu = min(TC_ROWS, U - u0);

    // ---- one-time setup: weights -> smem, TMEM, barriers
    {
        const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const unsigned char*>(wsplit) +
                                                          (size_t)bx * a_bytes);
        uint4* dst = reinterpret_cast<uint4*>(a_smem);
        for (uint32_t i = tid; i < a_bytes / 16; i += TC_THREADS) dst[i] = src[i];
    }
    if (warp == 0) umma::tmem_alloc(&sh->tmem_slot, TC_TMEM_COLS);
    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            umma::mbar_init(&sh->acc_full[s], 2);       // one commit per issuer warp
            umma::mbar_init(&sh->acc_empty[s], TC_EPI_WARPS);
            umma::mbar_init(&sh->oh_full[s], 1);
            umma::mbar_init(&sh->oh_empty[s], 2);
        }
    }
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = sh->tmem_slot;
    const int ntiles = (n + TC_WIN - 1) / TC_WIN;
    const int nseq = (B - 1 - by) / ns + 1;

    if (warp == TC_EPI_WARPS || warp == TC_EPI_WARPS + 1) {
        // ================= issuer warps =================
        // The even- and odd-position accumulators are independent chains with their own issuer warp (on different
        // schedulers); both commit to acc_full.  Every operand is made warp-uniform up front and ONE elected lane issues
        // the whole tile: five uniform-datapath instructions per MMA (per-MMA elect.sync on descriptors derived from
        // threadIdx cost ~30 instructions and ~170 cycles each -- R2UR / VOTEU chains -- three times the MMA itself).
        const uint32_t par = umma::warp_uniform((uint32_t)(warp - TC_EPI_WARPS));
        const uint32_t a_lo = umma::warp_uniform(umma::desc_lo(umma::smem_u32(a_smem), 2048u));
        const uint32_t b_lo = umma::warp_uniform(umma::desc_lo(umma::smem_u32(oh), 16u)) + par * (uint32_t)(Lbuf >> 1);
        const uint32_t d_par = umma::warp_uniform(tmem) + (par ? TC_ODD_COL : 0);
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
                if (umma::elect_one()) {
                    const int wins = min(TC_WIN, n - t * TC_WIN);
                    const int N = ((((wins * XNN_POOL + 1) >> 1) + 15) & ~15);
                    const uint32_t idesc = umma::idesc_bf16_f32(TC_ROWS, N);
                    const uint32_t b_t = b_sq + 112u * (uint32_t)t;
                    const uint32_t d = d_par + stage * TC_STAGE_COLS;
                    uint32_t acc = 0;
                    for (int s3 = 0; s3 < 3; ++s3) {
                        uint32_t a_k = a_lo + (uint32_t)(s3 * chunks) * 128u, b_k = b_t;
#pragma unroll 4
                        for (int ks = 0; ks < ksteps; ++ks) {
                            umma::mma_bf16(d, ((uint64_t)hi128 << 32) | (uint64_t)a_k, ((uint64_t)hi128 << 32) | (uint64_t)b_k,
                                           idesc, acc);
                            acc = 1;
                            a_k += 256u;
                            b_k += 2u;
                        }
                    }
                    umma::commit(&sh->acc_full[stage]);
                    if (t == ntiles - 1) umma::commit(&sh->oh_empty[so]);
                }
                __syncwarp();
            }
        }
    } else if (warp == TC_EPI_WARPS + 2) {
        // ================= one-hot warp =================
        // the planes of sequence sq+1 are loaded into registers before sequence sq is expanded, so the global
        // latency never sits between two sequences (up to 128 words; longer sequences load the rest directly)
        uint32_t nx[4];
        auto load_planes = [&](int sq) {
            const uint32_t* sp = planes + (long long)(by + (long long)sq * ns) * 3 * Wp;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int i = lane + 32 * j;
                nx[j] = (sq < nseq && i < 3 * Wp) ? __ldg(sp + i) : 0u;
            }
        };
        load_planes(0);
        for (int sq = 0; sq < nseq; ++sq) {
            const int so = sq % noh;
            const uint32_t* sp = planes + (long long)(by + (long long)sq * ns) * 3 * Wp;
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (lane + 32 * j < 3 * Wp) s_pl[lane + 32 * j] = nx[j];
            for (int i = lane + 128; i < 3 * Wp; i += 32) s_pl[i] = __ldg(sp + i);
            load_planes(sq + 1);
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
        // ================= epilogue warps =================
        const int quarter = warp & 3, wq = warp >> 2;
        const int row = quarter * 32 + lane;
        const int u = u0 + row;
        const bool rows_live = quarter * 32 < nu;
        float asc = 0.0f, shf = 0.0f, sgn = 1.0f;
        if (u < U) {
            const float sc = scale[u];
            asc = fabsf(sc); shf = shift[u]; sgn = (sc < 0.0f) ? -1.0f : 1.0f;
        }
        const uint32_t tlane = tmem + ((uint32_t)(quarter * 32) << 16);
        int g = 0;
        for (int sq = 0; sq < nseq; ++sq) {
            const long long b = by + (long long)sq * ns;
            const long long gbase = (b * U + u0) * n;              // first element of this CTA's block
            const int ph4 = linear ? (int)(gbase & 3) : 0, ph16 = linear ? (int)(gbase & 15) : 0;
            for (int t = 0; t < ntiles; ++t, ++g) {
                const int q0 = t * TC_WIN;
                const int wins = min(TC_WIN, n - q0);
                const uint32_t stage = (uint32_t)g & 1u;
                umma::mbar_wait(&sh->acc_full[stage], ((uint32_t)g >> 1) & 1u);
                umma::fence_after_sync();
                const bool live = rows_live && 8 * wq < wins;
                uint32_t E[28], O[28];
                if (live) {
                    const uint32_t trow = tlane + stage * TC_STAGE_COLS + 28u * wq;
                    umma::tmem_ld28(trow, E);
                    umma::tmem_ld28(trow + TC_ODD_COL, O);
                    umma::tmem_ld_wait();
                }
                // this warp's share of the accumulator is in registers: hand the stage back
                umma::fence_before_sync();
                __syncwarp();
                if (lane == 0) umma::mbar_arrive(&sh->acc_empty[stage]);
                if (linear && t == 0 && sq > 0) epi_barrier();     // previous sequence read out of the staging
                if (live) {
                    const int qs = (linear ? q0 : 0) + 8 * wq;
                    float* pp = s_pool + ph4 + row * sstride + qs;
                    float* pz = s_z + ph4 + row * sstride + qs;
                    uint8_t* pa = s_arg + ph16 + row * bstride + qs;
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
                        if (8 * wq + j < wins) {                   // windows past the tile are never staged
                            pp[j] = expf(fmaf(asc, best, shf));
                            if (TRAIN) {
                                int arg = 6;                       // first maximum wins
#pragma unroll
                                for (int r = 5; r >= 0; --r) arg = (z[r] == best) ? r : arg;
                                pz[j] = sgn * best;
                                pa[j] = (uint8_t)arg;
                            }
                        }
                    }
                }
                if (!linear) {
                    epi_barrier();                                 // tile staged
                    float* gp = pool + gbase + (long long)(8 * warp) * n + q0 + lane;
                    float* gz = TRAIN ? zmax + gbase + (long long)(8 * warp) * n + q0 + lane : nullptr;
                    uint8_t* ga = TRAIN ? argmax + gbase + (long long)(8 * warp) * n + q0 + lane : nullptr;
                    if (lane < wins) {
#pragma unroll
                        for (int rr = 0; rr < 8; ++rr) {
                            const int r = 8 * warp + rr;
                            if (r < nu) {
                                gp[rr * n] = s_pool[r * 33 + lane];
                                if (TRAIN) { gz[rr * n] = s_z[r * 33 + lane]; ga[rr * n] = s_arg[r * 36 + lane]; }
                            }
                        }
                    }
                    epi_barrier();                                 // staging free again
                }
            }
            if (linear) {
                epi_barrier();                                     // sequence staged
                const int cnt = nu * n, etid = warp * 32 + lane;
                vector_copy_out<float>(pool + gbase, s_pool, ph4, cnt, etid);
                if (TRAIN) {
                    vector_copy_out<float>(zmax + gbase, s_z, ph4, cnt, etid);
                    vector_copy_out<uint8_t>(argmax + gbase, s_arg, ph16, cnt, etid);
                }
            }
        }
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, TC_TMEM_COLS);
}

static size_t tc_smem_bytes(int k, int L, int n, bool train, int noh, bool linear, int* Lbuf_out) {
    const int ntiles = (n + TC_WIN - 1) / TC_WIN;
    const int Lbuf = ntiles * TC_TILE_POS + 64;   // last tile reads <= P + 2*112 + 2*chunks + 2 positions
    if (Lbuf_out) *Lbuf_out = Lbuf;
    const size_t sstride = linear ? n : 33, bstride = linear ? n : 36;
    const size_t f_bytes = ((TC_ROWS * sstride + 4) * 4 + 15) & ~(size_t)15;
    const size_t b_bytes = (TC_ROWS * bstride + 16 + 15) & ~(size_t)15;
    const size_t stg = train ? 2 * f_bytes + b_bytes : f_bytes;
    const int Wp = (L + 31) / 32;
    return tc_split_bytes(k) + (size_t)noh * 2 * Lbuf * 8 + stg + (size_t)((3 * Wp * 4 + 15) & ~15) +
           sizeof(TcShared) + 16;
}

extern "C" int xnn_frontend_fwd(const uint32_t* planes, int B, int L, int U, int k, const float* conv_w,
                                const float* scale, const float* shift, float* pool, float* zmax,
                                uint8_t* argmax, void* work, xnn_stream_t s) {
    XNN_CHECK_ARG(planes && conv_w && scale && shift && pool && work, "NULL pointer");
    XNN_CHECK_ARG((zmax == nullptr) == (argmax == nullptr), "zmax and argmax go together");
    XNN_CHECK_ARG(k >= 1 && k <= TC_KMAX && L >= k && U >= 1 && B >= 0, "bad sizes (kernel_size <= 48)");
    XNN_CHECK_ARG(((uintptr_t)work & 15) == 0, "work must be 16-byte aligned");
    const int n = (L - k + 1) / XNN_POOL;
    if (B == 0 || n == 0) return XNN_OK;
    const int kpad = tc_kpad(k), mt = tc_mtiles(U);
    const bool train = (zmax != nullptr);
    // staging choice: whole sequence in output layout (one vector memcpy per sequence) when it fits and the row
    // stride n keeps the staging stores off a handful of banks; else tile by tile.  One-hot buffers: 2, else 1.
    int Lbuf = 0, noh = 2;
    bool linear = (n % 4) != 0;
    size_t smem = tc_smem_bytes(k, L, n, train, noh, linear, &Lbuf);
    if (smem > 227 * 1024 && linear) { noh = 1; smem = tc_smem_bytes(k, L, n, train, noh, linear, &Lbuf); }
    if (smem > 227 * 1024) { noh = 2; linear = false; smem = tc_smem_bytes(k, L, n, train, noh, linear, &Lbuf); }
    if (smem > 227 * 1024) { noh = 1; smem = tc_smem_bytes(k, L, n, train, noh, linear, &Lbuf); }
    XNN_CHECK_ARG(smem <= 227 * 1024, "sequence too long for the front-end staging buffers");
    cudaStream_t st = to_stream(s);
    tc_split_weights_kernel<<<dim3(ceil_div((long long)TC_ROWS * kpad, 256), mt), 256, 0, st>>>(
        conv_w, scale, U, k, kpad, reinterpret_cast<__nv_bfloat16*>(work));
    XNN_CHECK_LAUNCH();
    int slices = num_sms() / mt;                  // one CTA per SM: each holds all 512 TMEM columns
    if (slices < 1) slices = 1;
    if (slices > B) slices = B;
    dim3 grid(mt, slices);
    // slices per M tile from a TileMap.  Weight floor 1.0 = equal weights (the 148th SM gets a CTA too): giving the
    // partly filled last tile (44 of 128 units at U = 300) HALF the slices was measured SLOWER -- 0.55 -> 0.66 ms --
    // i.e. a tile costs the same whatever its fill: the MMAs and the TMEM read-out bound it, not the live rows
    const TileMap tmap = make_tile_map(U, TC_ROWS, num_sms(), B, 1.0f);
    if (tmap.n) grid = dim3(tmap.first[tmap.n], 1);
    if (train) {
        XNN_CUDA(cudaFuncSetAttribute(frontend_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        frontend_tc_kernel<true><<<grid, TC_THREADS, smem, st>>>(planes, B, L, planes_words(L), U, n, kpad, Lbuf, noh, (int)linear,
                                                                 reinterpret_cast<const __nv_bfloat16*>(work), scale,
                                                                 shift, pool, zmax, argmax, tmap);
    } else {
        XNN_CUDA(cudaFuncSetAttribute(frontend_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        frontend_tc_kernel<false><<<grid, TC_THREADS, smem, st>>>(planes, B, L, planes_words(L), U, n, kpad, Lbuf, noh, (int)linear,
                                                                  reinterpret_cast<const __nv_bfloat16*>(work), scale,
                                                                  shift, pool, zmax, argmax, tmap);
    }
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}


// ===========================================================================
// a11 on the tensor cores: the `PWM` module (architectures/__init__.py:116-170) as the same Toeplitz GEMM.
//   o[b, g, :] = [ conv(x_b, W_g) | conv(flip(x_b), W_g) ],  out[b, g] = max / sum over that axis
// flip(x) (channel and position reversed, :387-404) is the reverse complement, so every sequence is run as
// two "strands" with the SAME A operand (the G matrices) and the one-hot of x, then of RC(x), as B: the rev
// half is by construction bit-identical to the forward half of the reverse-complemented input.  Rows of an
// M tile are 128 matrices, columns are window starts; each epilogue thread keeps the running maximum (and
// its index on the concatenated axis, first maximum wins) or the running sum of its 56-column share of
// every tile, and the four column shares of a row are merged through shared memory once per sequence.
// ===========================================================================
// The scan kernels run SCAN_OH_WARPS one-hot warps (sequence sn is expanded by warp sn % SCAN_OH_WARPS): one warp
// expanding every sequence bounded the whole pipeline (ncu: tensor pipe 40 % active in pass 1, whose epilogue is
// 56 fmaxf per thread).
#define SCAN_OH_WARPS 2
#define SCAN_NOH_MAX 4       // one-hot ring: deep enough to hide the expansion latency of a sequence (~1 us) behind the MMAs
#define SCAN_THREADS (32 * (TC_EPI_WARPS + 2 + SCAN_OH_WARPS))

// both one-hot buffers of a sequence (x at buf0[i], x shifted by one position at buf1[i]) by one warp; only the
// L positions of the sequence are written, the zero tails of the ring are laid down once at kernel start
__device__ __forceinline__ void build_onehot_pair(uint2* __restrict__ buf0, uint2* __restrict__ buf1,
                                                  const uint32_t* __restrict__ pl, int Wp, int L, int lane, bool rc);

struct PwShared {
    uint64_t acc_full[2], acc_empty[2], oh_full[SCAN_NOH_MAX], oh_empty[SCAN_NOH_MAX];
    uint64_t codes_full[8];        // pass 2b: the one-hot warp wrote the base codes of a sequence (ring of 8)
    uint32_t tmem_slot, pad;
    float val[2][4][TC_ROWS];      // [sequence parity][column share][row]
    int idx[2][4][TC_ROWS];
};

// one-hot bf16 x4 of position i of the reverse complement of a sequence held as planes
__device__ __forceinline__ uint2 onehot_rc_bf16x4(const uint32_t* __restrict__ sp, int Wp, int L, int i) {
    uint2 v = make_uint2(0u, 0u);
    if (i < L) {
        const int s = L - 1 - i, w = s >> 5, b = s & 31;
        const uint32_t lo = ((sp[w] >> b) & 1u) ^ 1u, hi = ((sp[Wp + w] >> b) & 1u) ^ 1u, nm = (sp[2 * Wp + w] >> b) & 1u;
        if (!nm) {
            const uint32_t half = lo ? 0x3F800000u : 0x00003F80u;
            if (hi) v.y = half; else v.x = half;
        }
    }
    return v;
}

__device__ __forceinline__ void build_onehot_pair(uint2* __restrict__ buf0, uint2* __restrict__ buf1,
                                                  const uint32_t* __restrict__ pl, int Wp, int L, int lane, bool rc) {
    if (!rc) {
        // position 32 w + lane is bit `lane` of word w of each plane (lo, hi, N mask): three broadcast loads, a
        // handful of bit operations and two stores per position
#pragma unroll 4
        for (int w = 0; w < Wp; ++w) {
            const int i = 32 * w + lane;
            const uint32_t lo = (pl[w] >> lane) & 1u, hi = (pl[Wp + w] >> lane) & 1u, nm = (pl[2 * Wp + w] >> lane) & 1u;
            const uint32_t half = nm ? 0u : (0x3F80u << (16u * lo));          // bf16 1.0 in the lower / upper half
            const uint2 x = make_uint2(hi ? 0u : half, hi ? half : 0u);
            if (i < L) {
                buf0[i] = x;
                if (i > 0) buf1[i - 1] = x;
            }
        }
    } else {
        for (int i = lane; i < L; i += 32) {
            const uint2 x = onehot_rc_bf16x4(pl, Wp, L, i);
            buf0[i] = x;
            if (i > 0) buf1[i - 1] = x;
        }
    }
}

// The same pipeline serves the interpret passes 1 and 2a (test.py:128-166, interpret.py:363-421): rows are
// units (sign of the BatchNorm scale folded into A), one strand, and the epilogue keeps per unit
//   SCAN_ACT_MAX    the maximum of D = sgn(s) z over every listed sequence; the activation
//                   fp16(exp(|s| D + t)) is monotone in D, so its maximum is the activation of max D;
//   SCAN_SITE_COUNT the number of starts whose fp16 activation exceeds thr_u, per listed sequence.  The
//                   predicate is monotone in D as well: each thread bisects (on the float order) the first D
//                   for which it holds and afterwards evaluates exp only for D within 64 ulps of that
//                   point -- one or two compares per position instead of an exp and two conversions;
//   SCAN_SITE_EMIT  pass 2b: the same predicate; every site gets its rank (caller's exclusive offset of the
//                   (unit, sequence) pair + sites at earlier starts of the sequence), and while rank < cap its
//                   k-mer is counted into the unit's PFM (shared-memory atomics, flushed once per CTA) and,
//                   optionally, (list position, start) is written to sites[unit][rank].
enum { SCAN_SUM = 0, SCAN_MAX = 1, SCAN_MAX_ARG = 2, SCAN_ACT_MAX = 3, SCAN_SITE_COUNT = 4, SCAN_SITE_EMIT = 5 };

struct ScanArgs {
    const uint32_t* planes;
    int n_items, L, Wp, R, Lw, kpad, Lbuf, noh, nstr;
    const int32_t* idx;            // listed sequences (interpret) or NULL
    const __nv_bfloat16* wsplit;
    const float* scale; const float* shift; const float* thr;   // interpret
    float* out; int32_t* argmax;                                // PWM
    uint32_t* amax_bits; int32_t* counts;                       // interpret
    const long long* offsets; long long cap; int32_t* pfm; int32_t* sites; long long site_cap; int k;   // emit
};

__device__ __forceinline__ void cp_async4(uint32_t saddr, const void* g) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(saddr), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ float act16_of(float asc, float D, float shf) {
    return __half2float(__float2half_rn(expf(fmaf(asc, D, shf))));
}
__device__ __forceinline__ uint32_t float_key(float x) {
    const uint32_t b = __float_as_uint(x);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key_float(uint32_t k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
}

// The site predicate fp16(exp(asc D + shf)) > thr as compares on D.  It is monotone in D up to the last-ulp
// behaviour of expf, so: bisect (on the float order) the first D for which it holds, then evaluate it at all
// 128 floats within 64 ulps of that point.  If it is a clean step there (false below, true from the crossing
// on), the predicate IS "D >= Dhi" and the function returns true with Dlo the float just below Dhi.  Otherwise
// (never seen; a threshold on +-0 is the one constructible case) it returns false with the band
//   D >= Dhi: a site,  D <= Dlo: not a site,  in between: evaluate exactly,
// which is also what the callers use when any row of the CTA is not exact.
// No D satisfies the predicate: Dlo = Dhi = NaN (every compare false), also the "row switched off" value.
__device__ __forceinline__ bool site_thresholds(float asc, float shf, float thr, float& Dlo, float& Dhi) {
    const float qnan = __uint_as_float(0x7FC00000u);
    if (act16_of(asc, -INFINITY, shf) > thr) { Dlo = Dhi = -INFINITY; return true; }    // every start is a site
    if (!(act16_of(asc, INFINITY, shf) > thr)) { Dlo = Dhi = qnan; return true; }       // no start is
    uint32_t lo = float_key(-INFINITY), hi = float_key(INFINITY);                       // predicate false at lo, true at hi
    while (hi - lo > 1u) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (act16_of(asc, key_float(mid), shf) > thr) hi = mid; else lo = mid;
    }
    const uint32_t kmin = float_key(-INFINITY), kmax = float_key(INFINITY);
    const uint32_t k0 = hi - kmin > 64u ? hi - 64u : kmin, k1 = kmax - hi > 64u ? hi + 64u : kmax;
    bool step = true;
    for (uint32_t kk = k0; kk <= k1; ++kk) step = step && ((act16_of(asc, key_float(kk), shf) > thr) == (kk >= hi));
    Dhi = key_float(step ? hi : k1);
    Dlo = key_float(step ? hi - 1u : k0);
    return step;
}

// Site predicate of one thread's 2 NH window starts (E = even, O = odd local starts of its share) as a bit mask,
// bit pl <-> local start pl.  EXACT (every row of the CTA has a clean step): one compare and one predicated OR
// per start.  Otherwise two of each, without a branch, and the starts inside the band are evaluated exactly
// afterwards -- a warp almost never has one.
template <bool EXACT, int NH>
__device__ __forceinline__ unsigned long long site_mask(const uint32_t (&E)[NH], const uint32_t (&O)[NH], int limit,
                                                        float Dlo, float Dhi, float asc, float shf, float thr) {
    constexpr int NP = 2 * NH;
    uint32_t h0 = 0u, h1 = 0u, n0 = 0u, n1 = 0u;               // starts 0..31 / 32..NP-1
#pragma unroll
    for (int pl = 0; pl < NP; ++pl) {
        const float z = __uint_as_float((pl & 1) ? O[pl >> 1] : E[pl >> 1]);
        if (pl < 32) {
            if (z >= Dhi) h0 |= 1u << pl;
            if (!EXACT && z > Dlo) n0 |= 1u << pl;
        } else {
            if (z >= Dhi) h1 |= 1u << (pl - 32);
            if (!EXACT && z > Dlo) n1 |= 1u << (pl - 32);
        }
    }
    if (!EXACT) {
        n0 &= ~h0; n1 &= ~h1;
        if ((n0 | n1) != 0u) {
#pragma unroll
            for (int pl = 0; pl < NP; ++pl) {
                const uint32_t bit = (pl < 32) ? (n0 >> pl) & 1u : (n1 >> (pl - 32)) & 1u;
                if (bit) {
                    const float z = __uint_as_float((pl & 1) ? O[pl >> 1] : E[pl >> 1]);
                    if (act16_of(asc, z, shf) > thr) { if (pl < 32) h0 |= 1u << pl; else h1 |= 1u << (pl - 32); }
                }
            }
        }
    }
    unsigned long long hit = ((unsigned long long)h1 << 32) | h0;
    if (limit < NP) hit &= limit > 0 ? ((1ull << limit) - 1ull) : 0ull;
    return hit;
}

template <int MODE>
__global__ void __launch_bounds__(SCAN_THREADS, 1)
scan_tc_kernel(ScanArgs a) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const uint32_t* __restrict__ planes = a.planes;
    const int L = a.L, Wp = a.Wp, Lw = a.Lw, kpad = a.kpad, Lbuf = a.Lbuf, noh = a.noh, nstr = a.nstr;
    const int chunks = kpad >> 3, ksteps = kpad >> 4;
    const uint32_t a_bytes = 3u * chunks * 2048u;
    unsigned char* a_smem = smem_raw;
    uint2* oh = reinterpret_cast<uint2*>(smem_raw + a_bytes);                // [noh][2][Lbuf]
    uint32_t* s_pl = reinterpret_cast<uint32_t*>(oh + (size_t)noh * 2 * Lbuf);       // [one-hot warp][2][pl_words]
    const int pl_words = (3 * Wp + 3) & ~3;
    PwShared* sh = reinterpret_cast<PwShared*>(s_pl + SCAN_OH_WARPS * 2 * pl_words);
    const int code_bytes = (L + 15) & ~15;
    uint8_t* s_codes = reinterpret_cast<uint8_t*>(sh + 1);                   // [8][L] (emit): base codes 0..3 / 4 of the
    int* s_pfm = reinterpret_cast<int*>(s_codes + 8 * code_bytes);           // sequence in the epilogue; [128][4][k]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g0 = blockIdx.x * TC_ROWS;
    const int ng = min(TC_ROWS, a.R - g0);
    if (MODE == SCAN_SITE_EMIT)
        for (int i = tid; i < TC_ROWS * (4 * a.k + 1); i += SCAN_THREADS) s_pfm[i] = 0;
    const int ns = gridDim.y;
    {
        const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const unsigned char*>(a.wsplit) +
                                                          (size_t)blockIdx.x * a_bytes);
        uint4* dst = reinterpret_cast<uint4*>(a_smem);
        for (uint32_t i = tid; i < a_bytes / 16; i += SCAN_THREADS) dst[i] = src[i];
        for (int i = tid; i < noh * 2 * Lbuf; i += SCAN_THREADS) oh[i] = make_uint2(0u, 0u);   // the zero tails stay
    }
    if (warp == 0) umma::tmem_alloc(&sh->tmem_slot, TC_TMEM_COLS);
    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            umma::mbar_init(&sh->acc_full[s], 2);
            umma::mbar_init(&sh->acc_empty[s], TC_EPI_WARPS);
        }
        for (int s = 0; s < SCAN_NOH_MAX; ++s) {
            umma::mbar_init(&sh->oh_full[s], 1);
            umma::mbar_init(&sh->oh_empty[s], 2);
        }
        for (int s = 0; s < 8; ++s) umma::mbar_init(&sh->codes_full[s], 1);
        sh->pad = 0u;                                          // set by a row whose site predicate is not a clean step
    }
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = sh->tmem_slot;
    const int ntiles = (Lw + TC_TILE_POS - 1) / TC_TILE_POS;
    const int nseq = (a.n_items - 1 - (int)blockIdx.y) / ns + 1;
    const int nvs = nstr * nseq;                                // (sequence, strand) pairs of this CTA

    if (warp == TC_EPI_WARPS || warp == TC_EPI_WARPS + 1) {
        // ================= issuer warps (even / odd window starts) =================
        // Every operand of the MMAs is made warp-uniform up front (shfl from lane 0) and the whole tile is issued by
        // one elected lane: per-MMA elect.sync with descriptors built from threadIdx-derived values cost ~30
        // instructions and ~170 cycles per MMA (R2UR / VOTEU chains), three times the tensor time of the MMA --
        // measured with clock64: the issue loop, not the MMAs, the one-hot expansion or the epilogue, bounded all
        // three interpret passes.
        const uint32_t par = umma::warp_uniform((uint32_t)(warp - TC_EPI_WARPS));
        const uint32_t a_lo = umma::warp_uniform(umma::desc_lo(umma::smem_u32(a_smem), 2048u));
        const uint32_t b_lo = umma::warp_uniform(umma::desc_lo(umma::smem_u32(oh), 16u)) + par * (uint32_t)(Lbuf >> 1);
        const uint32_t d_par = umma::warp_uniform(tmem) + (par ? TC_ODD_COL : 0);
        const uint32_t hi128 = umma::desc_hi(128u);
        int g = 0;
        for (int v = 0; v < nvs; ++v) {
            const int so = v % noh;
            umma::mbar_wait(&sh->oh_full[so], (uint32_t)(v / noh) & 1u);
            const uint32_t b_sq = b_lo + (uint32_t)so * (uint32_t)Lbuf;
            for (int t = 0; t < ntiles; ++t, ++g) {
                const uint32_t stage = (uint32_t)g & 1u;
                umma::mbar_wait(&sh->acc_empty[stage], (((uint32_t)g >> 1) & 1u) ^ 1u);
                umma::fence_after_sync();
                if (umma::elect_one()) {
                    const int cnt = min(TC_TILE_POS, Lw - t * TC_TILE_POS);
                    const int N = (((cnt + 1) >> 1) + 15) & ~15;
                    const uint32_t idesc = umma::idesc_bf16_f32(TC_ROWS, N);
                    const uint32_t b_t = b_sq + 112u * (uint32_t)t;
                    const uint32_t d = d_par + stage * TC_STAGE_COLS;
                    uint32_t acc = 0;
                    for (int s3 = 0; s3 < 3; ++s3) {
                        uint32_t a_k = a_lo + (uint32_t)(s3 * chunks) * 128u, b_k = b_t;
#pragma unroll 4
                        for (int ks = 0; ks < ksteps; ++ks) {
                            umma::mma_bf16(d, ((uint64_t)hi128 << 32) | (uint64_t)a_k, ((uint64_t)hi128 << 32) | (uint64_t)b_k,
                                           idesc, acc);
                            acc = 1;
                            a_k += 256u;
                            b_k += 2u;
                        }
                    }
                    umma::commit(&sh->acc_full[stage]);
                    if (t == ntiles - 1) umma::commit(&sh->oh_empty[so]);
                }
                __syncwarp();
            }
        }
    } else if (warp >= TC_EPI_WARPS + 2) {
        // ================= one-hot warps: x or RC(x) of the (sequence, strand) pairs v = ow, ow + nw, ... =================
        // The planes of the warp's NEXT pair are in flight (cp.async) while this one is expanded.  Pairs are dealt by v,
        // and nw <= noh: a warp then never reaches the wait for ring slot v % noh before the slot's phase of pair
        // v - 2 noh has completed (its previous wait covered pair v - nw - noh, and pairs complete in order), so the
        // parity wait cannot alias.
        const int ow = warp - (TC_EPI_WARPS + 2);
        const int nw = min(SCAN_OH_WARPS, noh);
        uint32_t* my_pl = s_pl + ow * 2 * pl_words;
        auto fetch = [&](int v, int q) {
            const int item = blockIdx.y + (nstr == 2 ? (v >> 1) : v) * ns;
            const long long b = a.idx ? (long long)__ldg(a.idx + item) : (long long)item;
            const uint32_t* sp = planes + b * 3 * Wp;
            const uint32_t dst = umma::smem_u32(my_pl + (q & 1) * pl_words);
            for (int i = lane; i < 3 * Wp; i += 32) cp_async4(dst + 4u * i, sp + i);
            cp_async_commit();
        };
        if (ow < nw && ow < nvs) fetch(ow, 0);
        int q = 0;
        for (int v = ow; ow < nw && v < nvs; v += nw, ++q) {
            const int so = v % noh, strand = (nstr == 2) ? (v & 1) : 0;
            const uint32_t* pl = my_pl + (q & 1) * pl_words;
            cp_async_wait_all();
            __syncwarp();
            if (v + nw < nvs) fetch(v + nw, q + 1);
            if (v >= noh) umma::mbar_wait(&sh->oh_empty[so], (uint32_t)(v / noh - 1) & 1u);
            __syncwarp();
            if (MODE == SCAN_SITE_EMIT) {
                // base codes of the sequence for the epilogue's k-mer counts (one strand in this mode: v is the
                // sequence).  Ring of 8: past the wait above the epilogue is at most 6 sequences behind (<= 4 one-hot
                // buffers + 2 accumulator stages), so slot v & 7 is free and its barrier at most one phase ahead.
                uint8_t* codes = s_codes + (v & 7) * code_bytes;
                for (int i = lane; i < L; i += 32) {
                    const int w = i >> 5;
                    const uint32_t lo = (pl[w] >> lane) & 1u, hi = (pl[Wp + w] >> lane) & 1u,
                                   nm = (pl[2 * Wp + w] >> lane) & 1u;
                    codes[i] = (uint8_t)(nm ? 4u : (lo | (hi << 1)));
                }
                __syncwarp();
                if (lane == 0) umma::mbar_arrive(&sh->codes_full[v & 7]);
            }
            uint2* buf0 = oh + (size_t)so * 2 * Lbuf;
            build_onehot_pair(buf0, buf0 + Lbuf, pl, Wp, L, lane, strand != 0);
            umma::fence_proxy_async();
            __syncwarp();
            if (lane == 0) umma::mbar_arrive(&sh->oh_full[so]);
        }
    } else {
        // ================= epilogue warps =================
        const int quarter = warp & 3, wq = warp >> 2;
        const int row = quarter * 32 + lane;
        const bool rows_live = quarter * 32 < ng;
        const uint32_t tlane = tmem + ((uint32_t)(quarter * 32) << 16) + 28u * wq;
        constexpr bool INTERP = (MODE >= SCAN_ACT_MAX);
        float asc = 0.0f, shf = 0.0f, thr = 0.0f;
        float Dlo = __uint_as_float(0x7FC00000u), Dhi = Dlo;   // site passes: see site_thresholds (NaN: no site)
        bool all_exact = true;
        if (INTERP && row < ng) {
            asc = fabsf(a.scale[g0 + row]);
            shf = a.shift[g0 + row];
            if (MODE == SCAN_SITE_COUNT || MODE == SCAN_SITE_EMIT) {
                thr = a.thr[g0 + row];
                if (!site_thresholds(asc, shf, thr, Dlo, Dhi)) sh->pad = 1u;
            }
        }
        if (MODE == SCAN_SITE_COUNT || MODE == SCAN_SITE_EMIT) {
            epi_barrier();
            all_exact = sh->pad == 0u;                         // CTA-uniform: one compare per start
        }
        float best = -INFINITY, acc = 0.0f;
        int bidx = 0x7fffffff, cnt = 0;
        int g = 0;
        for (int sq = 0; sq < nseq; ++sq) {
            // pass 2b: the exclusive site offset of (unit, sequence) is fetched at the START of the sequence, under the
            // MMAs -- loaded at its first use it was a dependent global load inside the emission loop of every tile
            long long site_off = 0;
            if (MODE == SCAN_SITE_EMIT && row < ng)
                site_off = __ldg(a.offsets + (long long)(g0 + row) * a.n_items + (blockIdx.y + sq * ns));
            for (int strand = 0; strand < nstr; ++strand) {
                for (int t = 0; t < ntiles; ++t, ++g) {
                    const uint32_t stage = (uint32_t)g & 1u;
                    umma::mbar_wait(&sh->acc_full[stage], ((uint32_t)g >> 1) & 1u);
                    umma::fence_after_sync();
                    const int p0 = t * TC_TILE_POS + 56 * wq;      // first window start of this warp's share
                    const int limit = Lw - p0;                     // local starts 0 .. limit-1 exist
                    const bool live = rows_live && limit > 0;
                    uint32_t E[28], O[28];
                    if (live) {
                        umma::tmem_ld28(tlane + stage * TC_STAGE_COLS, E);
                        umma::tmem_ld28(tlane + stage * TC_STAGE_COLS + TC_ODD_COL, O);
                        umma::tmem_ld_wait();
                    }
                    umma::fence_before_sync();
                    __syncwarp();
                    if (lane == 0) umma::mbar_arrive(&sh->acc_empty[stage]);
                    if (MODE == SCAN_SITE_EMIT) {
                        // hit mask of this thread's 56 starts, ascending
                        unsigned long long mask = !live ? 0ull
                                                  : all_exact ? site_mask<true, 28>(E, O, limit, Dlo, Dhi, asc, shf, thr)
                                                              : site_mask<false, 28>(E, O, limit, Dlo, Dhi, asc, shf, thr);
                        const int tp = g & 1;
                        const int item = blockIdx.y + sq * ns;
                        uint8_t* codes = s_codes + (sq & 7) * code_bytes;
                        sh->idx[tp][wq][row] = __popcll(mask);
                        if (t == 0 && strand == 0) umma::mbar_wait(&sh->codes_full[sq & 7], ((uint32_t)sq >> 3) & 1u);
                        epi_barrier();
                        int before = 0, total = 0;
#pragma unroll
                        for (int s2 = 0; s2 < 4; ++s2) {
                            const int c2 = sh->idx[tp][s2][row];
                            total += c2;
                            if (s2 < wq) before += c2;
                        }
                        if (mask != 0ull && row < ng) {
                            const int u = g0 + row, k = a.k;
                            long long rank = site_off + cnt + before;
                            while (mask != 0ull && rank < a.cap) {
                                const int pl = __ffsll((long long)mask) - 1;
                                mask &= mask - 1ull;
                                const int p = p0 + pl;
                                if (a.sites && rank < a.site_cap) {
                                    a.sites[((long long)u * a.site_cap + rank) * 2 + 0] = item;
                                    a.sites[((long long)u * a.site_cap + rank) * 2 + 1] = p;
                                }
                                const uint8_t* cp = codes + p;
                                int* pr = s_pfm + row * (4 * k + 1);           // odd row pitch: rows on distinct banks
#pragma unroll 4
                                for (int j = 0; j < k; ++j) {
                                    const int c = cp[j];
                                    if (c < 4) atomicAdd(&pr[c * k + j], 1);
                                }
                                ++rank;
                            }
                        }
                        cnt += total;                              // sites at earlier starts of this sequence
                    } else if (MODE == SCAN_SITE_COUNT) {
                        if (live)
                            cnt += __popcll(all_exact ? site_mask<true, 28>(E, O, limit, Dlo, Dhi, asc, shf, thr)
                                                      : site_mask<false, 28>(E, O, limit, Dlo, Dhi, asc, shf, thr));
                    } else if (live) {
                        const int base = strand * Lw + p0;         // index on the [fwd | rev] axis
                        const bool full = limit >= 56;
#pragma unroll
                        for (int pl = 0; pl < 56; ++pl) {
                            const float z = __uint_as_float((pl & 1) ? O[pl >> 1] : E[pl >> 1]);
                            if (full || pl < limit) {
                                if (MODE == SCAN_SUM) acc += z;
                                else if (MODE == SCAN_MAX_ARG) { if (z > best) { best = z; bidx = base + pl; } }
                                else if (MODE == SCAN_MAX || MODE == SCAN_ACT_MAX) best = fmaxf(best, z);
                            }
                        }
                    }
                }
            }
            if (MODE == SCAN_ACT_MAX) continue;                    // one maximum per unit over the whole list
            if (MODE == SCAN_SITE_EMIT) { cnt = 0; continue; }
            // merge the four column shares of every row (first maximum wins: smaller index on ties)
            const int par = sq & 1;
            sh->val[par][wq][row] = MODE == SCAN_SUM ? acc : best;
            sh->idx[par][wq][row] = MODE == SCAN_SITE_COUNT ? cnt : bidx;
            epi_barrier();
            if (wq == 0 && row < ng) {
                float bv = sh->val[par][0][row];
                int bi = sh->idx[par][0][row];
#pragma unroll
                for (int s = 1; s < 4; ++s) {
                    const float v = sh->val[par][s][row];
                    const int i = sh->idx[par][s][row];
                    if (MODE == SCAN_SUM) bv += v;
                    else if (MODE == SCAN_SITE_COUNT) bi += i;
                    else if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }
                }
                const long long item = blockIdx.y + (long long)sq * ns;
                if (MODE == SCAN_SITE_COUNT) {
                    a.counts[(long long)(g0 + row) * a.n_items + item] = bi;
                } else {
                    a.out[item * a.R + g0 + row] = bv;
                    if (MODE == SCAN_MAX_ARG) a.argmax[item * a.R + g0 + row] = bi;
                }
            }
            best = -INFINITY; acc = 0.0f; bidx = 0x7fffffff; cnt = 0;
        }
        if (MODE == SCAN_ACT_MAX) {
            sh->val[0][wq][row] = best;
            epi_barrier();
            if (wq == 0 && row < ng) {
                const float m = fmaxf(fmaxf(sh->val[0][0][row], sh->val[0][1][row]),
                                      fmaxf(sh->val[0][2][row], sh->val[0][3][row]));
                // activations are >= 0 (or +inf): uint order == float order for non-negative floats
                if (m > -INFINITY) atomicMax(&a.amax_bits[g0 + row], __float_as_uint(act16_of(asc, m, shf)));
            }
        }
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, TC_TMEM_COLS);
    if (MODE == SCAN_SITE_EMIT)
        for (int i = tid; i < ng * 4 * a.k; i += SCAN_THREADS) {
            const int r = i / (4 * a.k), e = i - r * 4 * a.k;
            const int v = s_pfm[r * (4 * a.k + 1) + e];
            if (v) atomicAdd(&a.pfm[(long long)g0 * 4 * a.k + i], v);
        }
}

// ===========================================================================
// Pass 2b without site records (interpret.py:423-454 keeps only the PFMs): the k-mer counts are a SECOND GEMM
// on the one-hot buffers the scan already holds in shared memory,
//
//   PFM[u, 4j+c] = sum over the sites p of unit u of onehot[p+j][c] = sum_p H[u,p] . X[p, 4j+c],
//
// H[u,p] = 1 where start p is a site of unit u whose rank is below the cap.  X is never built: position q sits
// 8 bytes after position q-1 in the one-hot buffer, so the even starts of a tile read through an MN-major view
// (K row r <-> start 2r, 16 bytes apart; N chunk c <-> taps 2c, 2c+1, 16 bytes apart) ARE the Toeplitz
// matrix, and the odd starts are the same view of the buffer shifted by one position (the pair the scan uses
// for its odd window starts).  The epilogue turns the scan's accumulators into H -- bf16 0 / 1, one compare
// and half a byte-permute per start, written K-major, 16 bytes per store -- and issuer warp 0 runs the PFM MMAs
// of tile g-1 between the scan MMAs of tiles g and g+1.  The counts stay in TMEM (fp32, exact below 2^24, and
// a unit emits at most `cap` sites) for the CTA's whole share of the list and are added to the PFM once.
// No shared-memory atomics, no per-site loop, no CTA-wide barrier in the loop.
//
// Tiles are 192 window starts here (96 even + 96 odd accumulator columns), not 224: two scan stages of 192
// columns leave 128 TMEM columns for the PFM accumulator (k <= 32; longer kernels take the site-record kernel).
// With ONE scan stage the commit -> epilogue tcgen05.ld -> arrive -> issue round trip (~2 us, measured with
// everything else switched off) sat between any two tiles; two stages hide it behind the other tile's MMAs.
//
// The cap: a row whose offset has reached the cap contributes nothing (its thresholds become NaN); a row with
// room for every start of the sequence needs no ranks at all.  Only the ONE sequence per unit that crosses the
// cap needs the rank of each site: the warps holding such a row read the other column shares of the tile
// from TMEM as well, count the sites before their own share and trim their mask to what is left.
// Sequences after the point where every unit of the CTA is capped are not visited at all (offsets are
// exclusive prefix sums, so non-decreasing along the list; found by bisection at kernel start).
// ===========================================================================
#define PF_TILE 192                                      // window starts per tile
#define PF_HALF (PF_TILE / 2)                            // accumulator columns per parity = UMMA N max
#define PF_SHARE (PF_TILE / 4)                           // starts per epilogue thread
#define PF_NH (PF_SHARE / 2)                             // ... per parity (24)
#define PF_PIECES (PF_SHARE / 8)                         // 8-byte pieces of H per thread and parity (6)
#define PF_HCHUNKS (PF_TILE / 8)                         // 16-byte K chunks of H per tile: 12 even + 12 odd
#define PF_HBYTES (PF_HCHUNKS * TC_ROWS * 16)            // 49152
#define PF_COL (2 * PF_TILE)                             // first TMEM column of the PFM accumulator (384)
#define PF_KMAX 32                                       // 4 k <= 128 accumulator columns

struct PfShared {
    uint64_t acc_full[2], acc_empty[2];    // scan accumulator stages: MMAs completed / read out (16 warp arrivals)
    uint64_t oh_full[SCAN_NOH_MAX], oh_empty[SCAN_NOH_MAX];   // one-hot ring: written / both issuers done with it
    uint64_t h_full[2], h_empty[2];        // H buffers: written (16 warp arrivals) / PFM MMAs of the tile completed
    uint64_t pfm_done;                     // the last PFM MMA completed
    uint32_t tmem_slot, pad;
};

// 24 consecutive TMEM columns of this warp's lanes as 16 + 8
__device__ __forceinline__ void tmem_ld24(uint32_t taddr, uint32_t (&r)[PF_NH]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    umma::tmem_ld8(taddr + 16, &r[16]);
}

// bf16 pair (lo = a is a site, hi = b is a site); !EXACT: also the "needs the exact predicate" accumulator
template <bool EXACT>
__device__ __forceinline__ uint32_t site_pair(uint32_t a, uint32_t b, float Dlo, float Dhi, uint32_t& near) {
    const float za = __uint_as_float(a), zb = __uint_as_float(b);
    const uint32_t ha = __float_as_uint(za >= Dhi ? 1.0f : 0.0f), hb = __float_as_uint(zb >= Dhi ? 1.0f : 0.0f);
    if (!EXACT) {
        const uint32_t na = __float_as_uint(za > Dlo ? 1.0f : 0.0f), nb = __float_as_uint(zb > Dlo ? 1.0f : 0.0f);
        near |= (ha ^ na) | (hb ^ nb);
    }
    return __byte_perm(ha, hb, 0x7632);                  // 1.0f >> 16 == bf16 1.0
}

// piece m (four even and four odd starts, 8 m .. 8 m + 7) of one thread's share of H as bf16 pairs
template <bool EXACT>
__device__ __forceinline__ void pfm_piece(int m, const uint32_t (&E)[PF_NH], const uint32_t (&O)[PF_NH], int limit,
                                          float dlo, float dhi, uint32_t& near, uint32_t (&w)[4]) {
    if (limit >= 8 * m + 8) {
        w[0] = site_pair<EXACT>(E[4 * m], E[4 * m + 1], dlo, dhi, near);
        w[1] = site_pair<EXACT>(E[4 * m + 2], E[4 * m + 3], dlo, dhi, near);
        w[2] = site_pair<EXACT>(O[4 * m], O[4 * m + 1], dlo, dhi, near);
        w[3] = site_pair<EXACT>(O[4 * m + 2], O[4 * m + 3], dlo, dhi, near);
    } else if (limit > 8 * m) {                          // the piece straddles Lw: starts >= limit compare as NaN
        uint32_t x[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t z = (i & 1) ? O[4 * m + (i >> 1)] : E[4 * m + (i >> 1)];
            x[i] = (8 * m + i < limit) ? z : 0x7FC00000u;
        }
        w[0] = site_pair<EXACT>(x[0], x[2], dlo, dhi, near);
        w[1] = site_pair<EXACT>(x[4], x[6], dlo, dhi, near);
        w[2] = site_pair<EXACT>(x[1], x[3], dlo, dhi, near);
        w[3] = site_pair<EXACT>(x[5], x[7], dlo, dhi, near);
    } else {
        w[0] = w[1] = w[2] = w[3] = 0u;
    }
}

// The six pieces (per parity) of one thread's share of H.  Piece m of column share wq is piece 6 wq + m of the
// tile, 8 bytes at chunk (6 wq + m) / 2, half m & 1: two pieces fill one 16-byte chunk and go out as ONE store
// (8-byte stores would leave every 16-byte slot of the row group half-filled: twice the shared-memory
// wavefronts, on the pipe the MMAs read their operands through).
//   hp: this thread's row of chunk 3 wq, even starts; the odd starts sit PF_HCHUNKS / 2 chunks further.
template <bool EXACT>
__device__ __forceinline__ void pfm_pieces(const uint32_t (&E)[PF_NH], const uint32_t (&O)[PF_NH], int limit, float dlo,
                                           float dhi, uint32_t& near, unsigned char* hp) {
#pragma unroll
    for (int j = 0; j < PF_PIECES / 2; ++j) {
        uint32_t a[4], b[4];
        pfm_piece<EXACT>(2 * j, E, O, limit, dlo, dhi, near, a);
        pfm_piece<EXACT>(2 * j + 1, E, O, limit, dlo, dhi, near, b);
        *reinterpret_cast<uint4*>(hp + j * 2048) = make_uint4(a[0], a[1], b[0], b[1]);
        *reinterpret_cast<uint4*>(hp + (j + PF_HCHUNKS / 2) * 2048) = make_uint4(a[2], a[3], b[2], b[3]);
    }
}

__global__ void __launch_bounds__(SCAN_THREADS, 1)
site_pfm_tc_kernel(ScanArgs a, int nh) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const uint32_t* __restrict__ planes = a.planes;
    const int L = a.L, Wp = a.Wp, Lw = a.Lw, kpad = a.kpad, Lbuf = a.Lbuf, noh = a.noh;
    const int chunks = kpad >> 3, ksteps = kpad >> 4;
    const uint32_t a_bytes = 3u * chunks * 2048u;
    unsigned char* a_smem = smem_raw;
    unsigned char* h_smem = smem_raw + a_bytes;                              // [nh][24][128] x 16 B
    uint2* oh = reinterpret_cast<uint2*>(h_smem + (size_t)nh * PF_HBYTES);   // [noh][2][Lbuf]
    const int pl_words = (3 * Wp + 3) & ~3;
    uint32_t* s_pl = reinterpret_cast<uint32_t*>(oh + (size_t)noh * 2 * Lbuf);   // [one-hot warp][2][pl_words]
    PfShared* sh = reinterpret_cast<PfShared*>(s_pl + SCAN_OH_WARPS * 2 * pl_words);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g0 = blockIdx.x * TC_ROWS;
    const int ng = min(TC_ROWS, a.R - g0);
    const int ns = gridDim.y;
    const int ntiles = (Lw + PF_TILE - 1) / PF_TILE;
    int nseq = (a.n_items - 1 - (int)blockIdx.y) / ns + 1;
    {
        // sequences from the first one at which every unit of this CTA is capped contribute nothing
        auto all_capped = [&](int sq) -> bool {
            bool c = true;
            if (tid < ng)
                c = __ldg(a.offsets + (long long)(g0 + tid) * a.n_items + (blockIdx.y + (long long)sq * ns)) >= a.cap;
            return __syncthreads_and(c) != 0;
        };
        if (all_capped(nseq - 1)) {
            if (all_capped(0)) return;
            int lo = 0, hi = nseq - 1;                                       // all_capped: false at lo, true at hi
            while (hi - lo > 1) {
                const int mid = lo + ((hi - lo) >> 1);
                if (all_capped(mid)) hi = mid; else lo = mid;
            }
            nseq = hi;
        }
    }
    {
        const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const unsigned char*>(a.wsplit) +
                                                          (size_t)blockIdx.x * a_bytes);
        uint4* dst = reinterpret_cast<uint4*>(a_smem);
        for (uint32_t i = tid; i < a_bytes / 16; i += SCAN_THREADS) dst[i] = src[i];
        for (int i = tid; i < noh * 2 * Lbuf; i += SCAN_THREADS) oh[i] = make_uint2(0u, 0u);   // the zero tails stay
    }
    if (warp == 0) umma::tmem_alloc(&sh->tmem_slot, TC_TMEM_COLS);
    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            umma::mbar_init(&sh->acc_full[s], 2);
            umma::mbar_init(&sh->acc_empty[s], TC_EPI_WARPS);
            umma::mbar_init(&sh->h_full[s], TC_EPI_WARPS);
            umma::mbar_init(&sh->h_empty[s], 1);
        }
        for (int s = 0; s < SCAN_NOH_MAX; ++s) {
            umma::mbar_init(&sh->oh_full[s], 1);
            umma::mbar_init(&sh->oh_empty[s], 2);
        }
        umma::mbar_init(&sh->pfm_done, 1);
        sh->pad = 0u;                                          // set by a row whose site predicate is not a clean step
    }
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = sh->tmem_slot;
    const int npfm = (4 * a.k + 15) & ~15;                                   // UMMA N of the PFM GEMM (<= 128)

    if (warp == TC_EPI_WARPS || warp == TC_EPI_WARPS + 1) {
        // ================= issuer warps: scan MMAs of the even / odd window starts; warp 0 also the PFM GEMM ==========
        // (operands warp-uniform, one elected lane issues a whole tile: see scan_tc_kernel)
        const uint32_t par = umma::warp_uniform((uint32_t)(warp - TC_EPI_WARPS));
        const uint32_t tmem_u = umma::warp_uniform(tmem);
        const uint32_t a_lo = umma::warp_uniform(umma::desc_lo(umma::smem_u32(a_smem), 2048u));
        const uint32_t b_lo = umma::warp_uniform(umma::desc_lo(umma::smem_u32(oh), 16u)) + par * (uint32_t)(Lbuf >> 1);
        const uint32_t hi128 = umma::desc_hi(128u);
        const uint32_t h_lo = umma::warp_uniform(umma::desc_lo(umma::smem_u32(h_smem), 2048u));   // H: K-major, chunk pitch 2048, SBO 128
        const uint32_t x_lo = umma::warp_uniform(umma::desc_lo(umma::smem_u32(oh), 128u));        // X: MN-major, K group pitch 128,
        const uint32_t hi16 = umma::desc_hi(16u);                                                 //    N chunk pitch 16
        const uint32_t idesc_p = umma::idesc_bf16_f32(TC_ROWS, npfm) | (1u << 16);
        uint32_t pacc = 0;
        // the PFM MMAs of tile tp of the sequence in one-hot slot sp, H in buffer gp % nh
        auto issue_pfm = [&](int gp, int sp, int tp) {
            const int hb = gp % nh;
            umma::mbar_wait(&sh->h_full[hb], (uint32_t)(gp / nh) & 1u);
            umma::fence_after_sync();
            if (umma::elect_one()) {
                const int cnt = min(PF_TILE, Lw - tp * PF_TILE);
                const uint32_t x_t = x_lo + (uint32_t)sp * (uint32_t)Lbuf + (uint32_t)(PF_TILE / 2) * (uint32_t)tp;
                const uint32_t h_b = h_lo + (uint32_t)hb * (uint32_t)(PF_HBYTES >> 4);
                for (int p2 = 0; p2 < 2; ++p2) {
                    const int nk = (((cnt + 1 - p2) >> 1) + 15) >> 4;            // K steps of 16 even / odd starts
                    uint32_t h_k = h_b + (uint32_t)(p2 * (PF_HCHUNKS / 2)) * 128u,
                             x_k = x_t + (uint32_t)p2 * (uint32_t)(Lbuf >> 1);
#pragma unroll 3
                    for (int ks = 0; ks < nk; ++ks) {
                        umma::mma_bf16(tmem_u + PF_COL, ((uint64_t)hi128 << 32) | (uint64_t)h_k,
                                       ((uint64_t)hi16 << 32) | (uint64_t)x_k, idesc_p, pacc);
                        pacc = 1;
                        h_k += 256u;
                        x_k += 16u;
                    }
                }
                umma::commit(&sh->h_empty[hb]);
                if (tp == ntiles - 1) umma::commit(&sh->oh_empty[sp]);
            }
            pacc = 1;
            __syncwarp();
        };
        int g = 0, pso = 0, pt = 0;
        for (int v = 0; v < nseq; ++v) {
            const int so = v % noh;
            umma::mbar_wait(&sh->oh_full[so], (uint32_t)(v / noh) & 1u);
            const uint32_t b_sq = b_lo + (uint32_t)so * (uint32_t)Lbuf;
            for (int t = 0; t < ntiles; ++t, ++g) {
                const uint32_t stage = (uint32_t)g & 1u;
                umma::mbar_wait(&sh->acc_empty[stage], (((uint32_t)g >> 1) & 1u) ^ 1u);
                umma::fence_after_sync();
                if (umma::elect_one()) {
                    const int cnt = min(PF_TILE, Lw - t * PF_TILE);
                    const int N = (((cnt + 1) >> 1) + 15) & ~15;
                    const uint32_t idesc = umma::idesc_bf16_f32(TC_ROWS, N);
                    const uint32_t b_t = b_sq + (uint32_t)(PF_TILE / 2) * (uint32_t)t;
                    const uint32_t d = tmem_u + stage * PF_TILE + (par ? PF_HALF : 0);
                    uint32_t acc = 0;
                    for (int s3 = 0; s3 < 3; ++s3) {
                        uint32_t a_k = a_lo + (uint32_t)(s3 * chunks) * 128u, b_k = b_t;
#pragma unroll 4
                        for (int ks = 0; ks < ksteps; ++ks) {
                            umma::mma_bf16(d, ((uint64_t)hi128 << 32) | (uint64_t)a_k, ((uint64_t)hi128 << 32) | (uint64_t)b_k,
                                           idesc, acc);
                            acc = 1;
                            a_k += 256u;
                            b_k += 2u;
                        }
                    }
                    umma::commit(&sh->acc_full[stage]);
                    if (par == 1 && t == ntiles - 1) umma::commit(&sh->oh_empty[so]);
                }
                __syncwarp();
                if (par == 0 && g > 0) issue_pfm(g - 1, pso, pt);
                pso = so; pt = t;
            }
        }
        if (par == 0) {
            if (g > 0) issue_pfm(g - 1, pso, pt);
            if (umma::elect_one()) umma::commit(&sh->pfm_done);
            __syncwarp();
        }
    } else if (warp >= TC_EPI_WARPS + 2) {
        // ================= one-hot warps: sequences v = ow, ow + 2, ...; the planes of the warp's next sequence are in
        // flight (cp.async) while this one is expanded (2 warps <= noh: the parity waits cannot alias, see scan_tc_kernel)
        const int ow = warp - (TC_EPI_WARPS + 2);
        uint32_t* my_pl = s_pl + ow * 2 * pl_words;
        auto fetch = [&](int v) {
            const int item = blockIdx.y + v * ns;
            const long long b = a.idx ? (long long)__ldg(a.idx + item) : (long long)item;
            const uint32_t* sp = planes + b * 3 * Wp;
            const uint32_t dst = umma::smem_u32(my_pl + ((v / SCAN_OH_WARPS) & 1) * pl_words);
            for (int i = lane; i < 3 * Wp; i += 32) cp_async4(dst + 4u * i, sp + i);
            cp_async_commit();
        };
        if (ow < nseq) fetch(ow);
        for (int v = ow; v < nseq; v += SCAN_OH_WARPS) {
            const int so = v % noh;
            const uint32_t* pl = my_pl + ((v / SCAN_OH_WARPS) & 1) * pl_words;
            cp_async_wait_all();
            __syncwarp();
            if (v + SCAN_OH_WARPS < nseq) fetch(v + SCAN_OH_WARPS);
            if (v >= noh) umma::mbar_wait(&sh->oh_empty[so], (uint32_t)(v / noh - 1) & 1u);
            __syncwarp();
            uint2* buf0 = oh + (size_t)so * 2 * Lbuf;
            build_onehot_pair(buf0, buf0 + Lbuf, pl, Wp, L, lane, false);
            umma::fence_proxy_async();
            __syncwarp();
            if (lane == 0) umma::mbar_arrive(&sh->oh_full[so]);
        }
    } else {
        // ================= epilogue warps: accumulators -> H =================
        const int quarter = warp & 3, wq = warp >> 2;
        const int row = quarter * 32 + lane;
        const bool rows_live = quarter * 32 < ng;
        const uint32_t tq = tmem + ((uint32_t)(quarter * 32) << 16);
        // the exact predicate's constants are reloaded where it is evaluated (rare): three registers less in the loop
        auto exact_consts = [&](float& asc, float& shf, float& thr) {
            asc = shf = thr = 0.0f;
            if (row < ng) { asc = fabsf(__ldg(a.scale + g0 + row)); shf = __ldg(a.shift + g0 + row); thr = __ldg(a.thr + g0 + row); }
        };
        const float qnan = __uint_as_float(0x7FC00000u);
        float Dlo = qnan, Dhi = qnan;
        if (row < ng) {
            float asc, shf, thr;
            exact_consts(asc, shf, thr);
            if (!site_thresholds(asc, shf, thr, Dlo, Dhi)) sh->pad = 1u;
        }
        epi_barrier();
        const bool all_exact = sh->pad == 0u;                    // CTA-uniform: one compare per start, no band
        // this thread's row of H; its pieces start at chunk 3 wq of each parity
        unsigned char* hrow = h_smem + (size_t)row * 16 + (size_t)(3 * wq) * 2048;
        int g = 0;
        // the offset of (row, next sequence) is in flight during this sequence: fetched at its first use, the load
        // sat on the epilogue's critical path once per sequence
        const long long* offp = a.offsets + (long long)(g0 + min(row, ng - 1)) * a.n_items + blockIdx.y;
        long long off_next = nseq > 0 ? __ldg(offp) : 0ll;
        for (int sq = 0; sq < nseq; ++sq) {
            int rem = 0;                                         // sites this row may still emit, clamped to 0 .. Lw
            const long long off = off_next;
            if (sq + 1 < nseq) off_next = __ldg(offp + (long long)(sq + 1) * ns);
            if (row < ng) rem = (int)max(0ll, min(a.cap - off, (long long)Lw));
            const bool crossing = rem > 0 && rem < Lw;           // the cap may fall inside this sequence
            const bool any_crossing = __any_sync(0xffffffffu, crossing);
            const float dlo = rem > 0 ? Dlo : qnan, dhi = rem > 0 ? Dhi : qnan;
            int seen = 0;                                        // crossing rows: sites at earlier tiles of the sequence
            for (int t = 0; t < ntiles; ++t, ++g) {
                const uint32_t stage = (uint32_t)g & 1u;
                const uint32_t tst = tq + stage * PF_TILE;
                umma::mbar_wait(&sh->acc_full[stage], ((uint32_t)g >> 1) & 1u);
                umma::fence_after_sync();
                const int p0 = t * PF_TILE + PF_SHARE * wq;
                const int limit = Lw - p0;
                const bool live = rows_live && limit > 0;
                int before = 0, total = 0;
                if (any_crossing && rows_live) {
                    // ranks are needed: the sites of the other column shares of this tile, from TMEM
                    float asc, shf, thr;
                    exact_consts(asc, shf, thr);
#pragma unroll 1
                    for (int s2 = 0; s2 < 4; ++s2) {
                        const int lim2 = Lw - (t * PF_TILE + PF_SHARE * s2);
                        if (s2 == wq || lim2 <= 0) continue;
                        uint32_t E2[PF_NH], O2[PF_NH];
                        tmem_ld24(tst + (uint32_t)(PF_NH * s2), E2);
                        tmem_ld24(tst + (uint32_t)(PF_NH * s2) + PF_HALF, O2);
                        umma::tmem_ld_wait();
                        const int c2 = __popcll(site_mask<false, PF_NH>(E2, O2, lim2, dlo, dhi, asc, shf, thr));
                        total += c2;
                        if (s2 < wq) before += c2;
                    }
                }
                uint32_t E[PF_NH], O[PF_NH];
                if (live) {
                    tmem_ld24(tst + (uint32_t)(PF_NH * wq), E);
                    tmem_ld24(tst + (uint32_t)(PF_NH * wq) + PF_HALF, O);
                    umma::tmem_ld_wait();
                }
                umma::fence_before_sync();
                __syncwarp();
                if (lane == 0) umma::mbar_arrive(&sh->acc_empty[stage]);

                const int hb = g % nh;
                if (g >= nh) umma::mbar_wait(&sh->h_empty[hb], (uint32_t)(g / nh - 1) & 1u);
                unsigned char* hp = hrow + (size_t)hb * PF_HBYTES;
                uint32_t near = 0u;
                if (live) {
                    if (all_exact) pfm_pieces<true>(E, O, limit, dlo, dhi, near, hp);
                    else pfm_pieces<false>(E, O, limit, dlo, dhi, near, hp);
                } else if (rows_live) {
#pragma unroll
                    for (int j = 0; j < PF_PIECES; ++j)
                        *reinterpret_cast<uint4*>(hp + (j % (PF_PIECES / 2) + (j / (PF_PIECES / 2)) * (PF_HCHUNKS / 2)) * 2048) =
                            make_uint4(0u, 0u, 0u, 0u);
                }
                if (__any_sync(0xffffffffu, near != 0u || (crossing && live))) {
                    // rare: exact predicate and / or the cap, from the bit mask; the pieces are written again
                    float asc, shf, thr;
                    exact_consts(asc, shf, thr);
                    unsigned long long mask = live ? site_mask<false, PF_NH>(E, O, limit, dlo, dhi, asc, shf, thr) : 0ull;
                    if (crossing) {
                        const int own = __popcll(mask);
                        const int room = max(0, rem - seen - before);
                        if (own > room) {
                            unsigned long long keep = 0ull;
                            for (int i = 0; i < room; ++i) {
                                const unsigned long long b = mask & (0ull - mask);
                                keep |= b;
                                mask ^= b;
                            }
                            mask = keep;
                        }
                        seen += total + own;
                    }
                    if (live) {
#pragma unroll
                        for (int j = 0; j < PF_PIECES / 2; ++j) {
                            uint32_t w8[8];                      // [piece 2j, 2j+1][parity * 2 + half]
#pragma unroll
                            for (int q = 0; q < 8; ++q) {
                                const int m = 2 * j + (q >> 2), par2 = (q >> 1) & 1, half = q & 1;
                                const int pl0 = 8 * m + 4 * half + par2;
                                const uint32_t b0 = (uint32_t)(mask >> pl0) & 1u, b1 = (uint32_t)(mask >> (pl0 + 2)) & 1u;
                                w8[q] = b0 * 0x3F80u + b1 * 0x3F800000u;
                            }
                            *reinterpret_cast<uint4*>(hp + j * 2048) = make_uint4(w8[0], w8[1], w8[4], w8[5]);
                            *reinterpret_cast<uint4*>(hp + (j + PF_HCHUNKS / 2) * 2048) = make_uint4(w8[2], w8[3], w8[6], w8[7]);
                        }
                    }
                }
                umma::fence_proxy_async();
                __syncwarp();
                if (lane == 0) umma::mbar_arrive(&sh->h_full[hb]);
            }
        }
        // ---- the CTA's counts: TMEM -> global PFM (U, 4, k) ----
        umma::mbar_wait(&sh->pfm_done, 0u);
        umma::fence_after_sync();
        if (rows_live && nseq > 0) {
            for (int c0 = 16 * wq; c0 < npfm; c0 += 64) {
                uint32_t r[16];
                umma::tmem_ld16(tq + PF_COL + (uint32_t)c0, r);
                umma::tmem_ld_wait();
                if (row < ng) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const int n = c0 + i, j = n >> 2, c = n & 3;
                        const int v = (int)__uint_as_float(r[i]);
                        if (j < a.k && v != 0) atomicAdd(&a.pfm[((long long)(g0 + row) * 4 + c) * a.k + j], v);
                    }
                }
            }
        }
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, TC_TMEM_COLS);
}

static bool pwm_tc_fits(int L, int w, int* noh_out, int* Lbuf_out, size_t* smem_out, bool emit = false) {
    if (w < 1 || w > TC_KMAX || L < w) return false;
    const int Lw = L - w + 1, kpad = tc_kpad(w);
    const int ntiles = (Lw + TC_TILE_POS - 1) / TC_TILE_POS;
    const int Lbuf = ntiles * TC_TILE_POS + 2 * (kpad / 4) + 64;   // last tile reads up to start + 224 + kpad/4 + 2
    const int Wp = (L + 31) / 32;
    for (int noh = SCAN_NOH_MAX; noh >= 1; --noh) {
        const size_t plb = (size_t)SCAN_OH_WARPS * 2 * ((3 * Wp + 3) & ~3) * 4;
        const size_t smem = tc_split_bytes(w) + (size_t)noh * 2 * Lbuf * 8 + plb + sizeof(PwShared) + 16 +
                            (emit ? 8 * (size_t)((L + 15) & ~15) + (size_t)TC_ROWS * (4 * w + 1) * 4 : 0);
        if (smem <= 227 * 1024) {
            if (noh_out) *noh_out = noh;
            if (Lbuf_out) *Lbuf_out = Lbuf;
            if (smem_out) *smem_out = smem;
            return true;
        }
    }
    return false;
}

// shared memory of site_pfm_tc_kernel: prefers two H buffers and a one-hot ring of four (the one-hot buffer of a
// sequence lives until the PFM MMAs of its last tile, which are issued after the scan MMAs of the NEXT tile; a
// ring of one would deadlock)
static bool site_pfm_fits(int L, int w, int* noh_out, int* nh_out, int* Lbuf_out, size_t* smem_out) {
    if (w < 1 || w > PF_KMAX || L < w) return false;
    const int Lw = L - w + 1, kpad = tc_kpad(w);
    const int ntiles = (Lw + PF_TILE - 1) / PF_TILE;
    const int Lbuf = ntiles * PF_TILE + 2 * (kpad / 4) + 64;     // last tile reads up to start + 192 + kpad/4 + 2
    const int Wp = (L + 31) / 32, pl_words = (3 * Wp + 3) & ~3;
    for (int nh = 2; nh >= 1; --nh)
        for (int noh = SCAN_NOH_MAX; noh >= 2; --noh) {
            const size_t smem = tc_split_bytes(w) + (size_t)nh * PF_HBYTES + (size_t)noh * 2 * Lbuf * 8 +
                                (size_t)SCAN_OH_WARPS * 2 * pl_words * 4 + sizeof(PfShared) + 16;
            if (smem <= 227 * 1024) {
                *noh_out = noh; *nh_out = nh; *Lbuf_out = Lbuf; *smem_out = smem;
                return true;
            }
        }
    return false;
}

extern "C" int xnn_pwm_scan_tc_supported(int L, int w) { return pwm_tc_fits(L, w, nullptr, nullptr, nullptr) ? 1 : 0; }
extern "C" size_t xnn_pwm_scan_tc_work_bytes(int G, int w) {
    if (G < 1 || w < 1 || w > TC_KMAX) return 0;
    return (size_t)tc_mtiles(G) * tc_split_bytes(w);
}

template <int MODE>
static int launch_scan(const ScanArgs& a, dim3 grid, size_t smem, cudaStream_t st) {
    if (cudaFuncSetAttribute(scan_tc_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
        return -1;
    scan_tc_kernel<MODE><<<grid, SCAN_THREADS, smem, st>>>(a);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// mode: SCAN_*; returns 0 when launched, 1 when the shape is outside this path, -1 on a CUDA error
static int launch_scan_tc(int mode, const uint32_t* planes, int n_items, const int32_t* idx, int L, int R, int w,
                          const float* weights, const float* scale, const float* shift, const float* thr, float* out,
                          int32_t* argmax, uint32_t* amax_bits, int32_t* counts, void* work, cudaStream_t st,
                          const long long* offsets = nullptr, long long cap = 0, int32_t* pfm = nullptr,
                          int32_t* sites = nullptr, long long site_cap = 0) {
    int noh = 0, Lbuf = 0, nh = 0;
    size_t smem = 0;
    // pass 2b without site records: the PFM as a second GEMM (site_pfm_tc_kernel)
    const bool pfm_gemm = mode == SCAN_SITE_EMIT && sites == nullptr && site_pfm_fits(L, w, &noh, &nh, &Lbuf, &smem);
    if (!pfm_gemm && !pwm_tc_fits(L, w, &noh, &Lbuf, &smem, mode == SCAN_SITE_EMIT)) return 1;
    const int kpad = tc_kpad(w), mt = tc_mtiles(R);
    tc_split_weights_kernel<<<dim3(ceil_div((long long)TC_ROWS * kpad, 256), mt), 256, 0, st>>>(
        weights, scale, R, w, kpad, reinterpret_cast<__nv_bfloat16*>(work));
    if (cudaGetLastError() != cudaSuccess) return -1;
    int slices = num_sms() / mt;
    if (slices < 1) slices = 1;
    if (slices > n_items) slices = n_items;
    dim3 grid(mt, slices);
    ScanArgs a{};
    a.planes = planes; a.n_items = n_items; a.L = L; a.Wp = (L + 31) / 32; a.R = R; a.Lw = L - w + 1; a.kpad = kpad;
    a.Lbuf = Lbuf; a.noh = noh; a.nstr = (mode <= SCAN_MAX_ARG) ? 2 : 1; a.idx = idx;
    a.wsplit = reinterpret_cast<const __nv_bfloat16*>(work); a.scale = scale; a.shift = shift; a.thr = thr;
    a.out = out; a.argmax = argmax; a.amax_bits = amax_bits; a.counts = counts;
    a.offsets = offsets; a.cap = cap; a.pfm = pfm; a.sites = sites; a.site_cap = site_cap; a.k = w;
    if (pfm_gemm) {
        if (cudaFuncSetAttribute(site_pfm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return -1;
        site_pfm_tc_kernel<<<grid, SCAN_THREADS, smem, st>>>(a, nh);
        return cudaGetLastError() == cudaSuccess ? 0 : -1;
    }
    switch (mode) {
        case SCAN_SUM: return launch_scan<SCAN_SUM>(a, grid, smem, st);
        case SCAN_MAX: return launch_scan<SCAN_MAX>(a, grid, smem, st);
        case SCAN_MAX_ARG: return launch_scan<SCAN_MAX_ARG>(a, grid, smem, st);
        case SCAN_ACT_MAX: return launch_scan<SCAN_ACT_MAX>(a, grid, smem, st);
        case SCAN_SITE_EMIT: return launch_scan<SCAN_SITE_EMIT>(a, grid, smem, st);
        default: return launch_scan<SCAN_SITE_COUNT>(a, grid, smem, st);
    }
}

extern "C" int xnn_pwm_scan_tc(const uint32_t* planes, int B, int L, int G, int w, const float* pwm_w, int mode,
                               float* out, int32_t* argmax, void* work, xnn_stream_t s) {
    XNN_CHECK_ARG(planes && pwm_w && out && work, "NULL pointer");
    XNN_CHECK_ARG(w >= 1 && L >= w && G >= 1 && B >= 0 && (mode == 0 || mode == 1), "bad sizes");
    XNN_CHECK_ARG(((uintptr_t)work & 15) == 0, "work must be 16-byte aligned");
    XNN_CHECK_ARG(argmax == nullptr || mode == 1, "argmax goes with mode 1 (max)");
    if (B == 0) return XNN_OK;
    const int m = mode == 0 ? SCAN_SUM : (argmax ? SCAN_MAX_ARG : SCAN_MAX);
    const int r = launch_scan_tc(m, planes, B, nullptr, L, G, w, pwm_w, nullptr, nullptr, nullptr, out, argmax, nullptr,
                                 nullptr, work, to_stream(s));
    if (r == 1) { set_error("xnn_pwm_scan_tc: shape outside the tensor-core path (see xnn_pwm_scan_tc_supported)"); return XNN_ERR_ARG; }
    if (r != 0) { set_error("xnn_pwm_scan_tc: launch failed: %s", cudaGetErrorString(cudaGetLastError())); return XNN_ERR_CUDA; }
    return XNN_OK;
}

#define INTERP_TC_CHECKS(fn)                                                                               \
    XNN_CHECK_ARG(planes && conv_w && scale && shift && work, "NULL pointer");                             \
    XNN_CHECK_ARG(k >= 1 && L >= k && U >= 1 && B >= 0 && n_idx >= 0, "bad sizes");                        \
    XNN_CHECK_ARG(idx != nullptr || n_idx == B, "idx == NULL means all B sequences (n_idx must equal B)"); \
    XNN_CHECK_ARG(((uintptr_t)work & 15) == 0, "work must be 16-byte aligned");                            \
    if (n_idx == 0) return XNN_OK;

extern "C" int xnn_act_max_tc(const uint32_t* planes, int B, int L, int U, int k, const float* conv_w,
                              const float* scale, const float* shift, const int32_t* idx, int n_idx,
                              uint32_t* amax_bits, void* work, xnn_stream_t s) {
    INTERP_TC_CHECKS();
    XNN_CHECK_ARG(amax_bits, "NULL pointer");
    const int r = launch_scan_tc(SCAN_ACT_MAX, planes, n_idx, idx, L, U, k, conv_w, scale, shift, nullptr, nullptr,
                                 nullptr, amax_bits, nullptr, work, to_stream(s));
    if (r) { set_error("xnn_act_max_tc: %s", r == 1 ? "shape outside the tensor-core path" : "launch failed"); return r == 1 ? XNN_ERR_ARG : XNN_ERR_CUDA; }
    return XNN_OK;
}

extern "C" int xnn_site_counts_tc(const uint32_t* planes, int B, int L, int U, int k, const float* conv_w,
                                  const float* scale, const float* shift, const float* thr, const int32_t* idx,
                                  int n_idx, int32_t* counts, void* work, xnn_stream_t s) {
    INTERP_TC_CHECKS();
    XNN_CHECK_ARG(thr && counts, "NULL pointer");
    const int r = launch_scan_tc(SCAN_SITE_COUNT, planes, n_idx, idx, L, U, k, conv_w, scale, shift, thr, nullptr,
                                 nullptr, nullptr, counts, work, to_stream(s));
    if (r) { set_error("xnn_site_counts_tc: %s", r == 1 ? "shape outside the tensor-core path" : "launch failed"); return r == 1 ? XNN_ERR_ARG : XNN_ERR_CUDA; }
    return XNN_OK;
}

extern "C" int xnn_interpret_tc_supported(int L, int k) { return pwm_tc_fits(L, k, nullptr, nullptr, nullptr, true) ? 1 : 0; }

extern "C" int xnn_site_emit_tc(const uint32_t* planes, int B, int L, int U, int k, const float* conv_w,
                                const float* scale, const float* shift, const float* thr, const int32_t* idx,
                                int n_idx, const long long* offsets, long long cap, int32_t* pfm, int32_t* sites,
                                long long site_cap, void* work, xnn_stream_t s) {
    INTERP_TC_CHECKS();
    XNN_CHECK_ARG(thr && offsets && pfm, "NULL pointer");
    XNN_CHECK_ARG(cap >= 0 && (sites == nullptr || site_cap >= 0), "bad cap");
    const int r = launch_scan_tc(SCAN_SITE_EMIT, planes, n_idx, idx, L, U, k, conv_w, scale, shift, thr, nullptr,
                                 nullptr, nullptr, nullptr, work, to_stream(s), offsets, cap, pfm, sites, site_cap);
    if (r) { set_error("xnn_site_emit_tc: %s", r == 1 ? "shape outside the tensor-core path" : "launch failed"); return r == 1 ? XNN_ERR_ARG : XNN_ERR_CUDA; }
    return XNN_OK;
}

// ===========================================================================
// xnn_pwm_scan_tc_noalloc: xnn_pwm_scan_tc for callers that cannot provide (or allocate) the work buffer -- code under
// CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
// The scratch (the three bf16 pieces of the G matrices in UMMA core-matrix
// order) lives in a file-scope device array of XNN_PWM_SCAN_NOALLOC_BYTES = 1 MiB, which every GPU that loads the
// library holds until exit.  In host code `pwm_scan_static_work` is only a host-side stand-in for the symbol: the
// device address comes from cudaGetSymbolAddress, once per device, with that device current.
// Every call on a device shares the one array: calls that overlap on the GPU (different streams) race.  Nothing in
// it has to start at zero -- tc_split_weights_kernel rewrites every byte the scan reads (padding rows and taps
// included) at the start of each call -- so the contents the previous call left behind are harmless.
// ===========================================================================
#define XNN_PWM_SCAN_NOALLOC_BYTES (1u << 20)
static __device__ __align__(256) unsigned char pwm_scan_static_work[XNN_PWM_SCAN_NOALLOC_BYTES];

#include <mutex>
static void* g_static_work_ptr[64];       // device address of pwm_scan_static_work, per device ordinal
static std::mutex g_static_work_mutex;

// device address of the static scratch on the CURRENT device (looked up on first use per device); nullptr + error
static void* pwm_scan_static_work_ptr() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        set_error("xnn_pwm_scan_tc_noalloc: cudaGetDevice failed or device ordinal >= 64");
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(g_static_work_mutex);
    if (g_static_work_ptr[dev] == nullptr) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, pwm_scan_static_work);
        if (e != cudaSuccess || p == nullptr) {
            set_error("xnn_pwm_scan_tc_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
            return nullptr;
        }
        g_static_work_ptr[dev] = p;
    }
    return g_static_work_ptr[dev];
}

extern "C" size_t xnn_pwm_scan_tc_noalloc_max_bytes(void) { return XNN_PWM_SCAN_NOALLOC_BYTES; }

extern "C" int xnn_pwm_scan_tc_noalloc(const uint32_t* planes, int B, int L, int G, int w, const float* pwm_w, int mode,
                                       float* out, int32_t* argmax, xnn_stream_t s) {
    XNN_CHECK_ARG(planes && pwm_w && out, "NULL pointer");
    XNN_CHECK_ARG(w >= 1 && L >= w && G >= 1 && B >= 0 && (mode == 0 || mode == 1), "bad sizes");
    XNN_CHECK_ARG(argmax == nullptr || mode == 1, "argmax goes with mode 1 (max)");
    const size_t need = xnn_pwm_scan_tc_work_bytes(G, w);
    if (need == 0 || !pwm_tc_fits(L, w, nullptr, nullptr, nullptr)) {
        set_error("xnn_pwm_scan_tc_noalloc: shape outside the tensor-core path (see xnn_pwm_scan_tc_supported)");
        return XNN_ERR_ARG;
    }
    if (need > (size_t)XNN_PWM_SCAN_NOALLOC_BYTES) {
        set_error("xnn_pwm_scan_tc_noalloc: G = %d matrices of width %d need %zu bytes of scratch, the static array "
                  "holds %u; use xnn_pwm_scan_tc with a caller-provided work buffer", G, w, need,
                  (unsigned)XNN_PWM_SCAN_NOALLOC_BYTES);
        return XNN_ERR_ARG;
    }
    void* work = pwm_scan_static_work_ptr();
    if (work == nullptr) return XNN_ERR_CUDA;
    if (B == 0) return XNN_OK;
    const int m = mode == 0 ? SCAN_SUM : (argmax ? SCAN_MAX_ARG : SCAN_MAX);
    const int r = launch_scan_tc(m, planes, B, nullptr, L, G, w, pwm_w, nullptr, nullptr, nullptr, out, argmax, nullptr,
                                 nullptr, work, to_stream(s));
    if (r != 0) { set_error("xnn_pwm_scan_tc_noalloc: launch failed: %s", cudaGetErrorString(cudaGetLastError())); return XNN_ERR_CUDA; }
    return XNN_OK;
}

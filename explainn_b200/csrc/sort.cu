// Segmented sort for the validation metrics: every column of a (G, N) float matrix sorted on
// its own, ascending or descending, with the permutation -- what `torch.sort(scores, dim=1)` did for 84 % of the
// metrics step in round 1.  sklearn's roc_auc_score / average_precision_score and scipy's rankdata sort each column
// on the host (explainn/architectures/__init__.py:458-461 via selene/__init__.py:359-366, pwm-scoring.py:115-120).
//
// One block per column, least-significant-digit radix sort with 8-bit digits over order-preserving 32-bit keys
// (4 passes, stable).  A pass: (1) 256-bin histogram of the column in shared memory, (2) exclusive scan of the bins,
// (3) the column again in order, 1024 elements at a time: every warp ranks its lanes inside their digit with
// __match_any_sync, a per-(warp, digit) count table and a scan over the warps give each element its stable position.
// Keys and indices ping-pong between two global buffers per column (work); the last pass writes the float values
// and int64 indices.  Traffic: 4 x (8 B read + 8 B written) per element; N < 2^31.
#include "common.cuh"
using namespace xnn;

#define SORT_THREADS 1024
#define SORT_WARPS (SORT_THREADS / 32)

__device__ __forceinline__ uint32_t sort_key(float x, bool desc) {
    uint32_t u = __float_as_uint(x);
    u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);      // ascending order of the floats = ascending unsigned order
    return desc ? ~u : u;
}
__device__ __forceinline__ float sort_unkey(uint32_t u, bool desc) {
    if (desc) u = ~u;
    u = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
    return __uint_as_float(u);
}

__global__ void __launch_bounds__(SORT_THREADS)
sort_columns_kernel(const float* __restrict__ vals, long long N, int desc, float* __restrict__ sorted,
                    long long* __restrict__ perm, uint32_t* __restrict__ work) {
    __shared__ uint32_t s_hist[256];
    __shared__ uint32_t s_base[256];                     // running start of every digit for the current chunk
    __shared__ uint16_t s_cnt[SORT_WARPS][256];          // per-(warp, digit) counts of the chunk
    const int g = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const float* v = vals + (long long)g * N;
    uint32_t* k0 = work + (long long)g * 4 * N;          // [keys A][idx A][keys B][idx B]
    uint32_t* i0 = k0 + N;
    uint32_t* k1 = i0 + N;
    uint32_t* i1 = k1 + N;
    const bool d = desc != 0;
    for (int pass = 0; pass < 4; ++pass) {
        const uint32_t* kin = (pass & 1) ? k1 : k0;
        const uint32_t* iin = (pass & 1) ? i1 : i0;
        uint32_t* kout = (pass & 1) ? k0 : k1;
        uint32_t* iout = (pass & 1) ? i0 : i1;
        const int shift = 8 * pass;
        if (tid < 256) s_hist[tid] = 0u;
        __syncthreads();
        for (long long i = tid; i < N; i += SORT_THREADS) {
            const uint32_t key = pass == 0 ? sort_key(v[i], d) : kin[i];
            atomicAdd(&s_hist[(key >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (warp == 0) {                                  // exclusive scan of the 256 bins by one warp (8 per lane)
            uint32_t loc[8], sum = 0u;
#pragma unroll
            for (int j = 0; j < 8; ++j) { loc[j] = s_hist[8 * lane + j]; sum += loc[j]; }
            uint32_t incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            uint32_t run = incl - sum;
#pragma unroll
            for (int j = 0; j < 8; ++j) { s_base[8 * lane + j] = run; run += loc[j]; }
        }
        __syncthreads();
        for (long long c0 = 0; c0 < N; c0 += SORT_THREADS) {
            for (int j = tid; j < SORT_WARPS * 256; j += SORT_THREADS) (&s_cnt[0][0])[j] = 0;
            __syncthreads();
            const long long i = c0 + tid;
            const bool live = i < N;
            uint32_t key = 0u, idx = 0u, dig = 0u, rank_in_warp = 0u;
            if (live) {
                key = pass == 0 ? sort_key(v[i], d) : kin[i];
                idx = pass == 0 ? (uint32_t)i : iin[i];
                dig = (key >> shift) & 255u;
            }
            // lanes of the warp with the same digit (dead lanes get a digit of their own: 256 + lane never matches)
            const uint32_t same = __match_any_sync(0xffffffffu, live ? dig : 256u + (uint32_t)lane);
            rank_in_warp = __popc(same & ((1u << lane) - 1u));
            if (live && rank_in_warp == 0u) s_cnt[warp][dig] = (uint16_t)__popc(same);
            __syncthreads();
            uint32_t chunk_total = 0u;
            if (tid < 256) {                              // exclusive scan over the warps: offset of every (warp, digit) group
                uint32_t run = 0u;
#pragma unroll 8
                for (int w = 0; w < SORT_WARPS; ++w) {
                    const uint32_t c = s_cnt[w][tid];
                    s_cnt[w][tid] = (uint16_t)run;        // at most 1024 elements per chunk: fits 16 bits
                    run += c;
                }
                chunk_total = run;
            }
            __syncthreads();
            if (live) {
                const uint32_t pos = s_base[dig] + (uint32_t)s_cnt[warp][dig] + rank_in_warp;
                if (pass == 3) {
                    sorted[(long long)g * N + pos] = sort_unkey(key, d);
                    perm[(long long)g * N + pos] = (long long)idx;
                } else {
                    kout[pos] = key;
                    iout[pos] = idx;
                }
            }
            __syncthreads();
            if (tid < 256) s_base[tid] += chunk_total;    // digit starts move past this chunk
        }
    }
}

extern "C" size_t xnn_sort_columns_work_bytes(long long N, int G) {
    if (N < 1 || G < 1) return 0;
    return (size_t)G * 4 * (size_t)N * sizeof(uint32_t);
}

extern "C" int xnn_sort_columns(const float* vals, long long N, int G, int descending, float* sorted, long long* perm,
                                void* work, xnn_stream_t s) {
    XNN_CHECK_ARG(vals && sorted && perm && work, "NULL pointer");
    XNN_CHECK_ARG(N >= 1 && N < (1LL << 31) && G >= 1, "bad sizes");
    XNN_CHECK_ARG(vals != sorted, "in-place sort is not supported");
    sort_columns_kernel<<<G, SORT_THREADS, 0, to_stream(s)>>>(vals, N, descending, sorted, perm, (uint32_t*)work);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

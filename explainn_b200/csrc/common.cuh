// Shared helpers for the explainn_b200 C-ABI library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/explainn_b200.h"

namespace xnn {

void set_error(const char* fmt, ...);
int num_sms();

#define XNN_CHECK_ARG(cond, msg)                                        \
    do {                                                                \
        if (!(cond)) {                                                  \
            ::xnn::set_error("%s: %s", __func__, msg);                  \
            return XNN_ERR_ARG;                                         \
        }                                                               \
    } while (0)

#define XNN_CHECK_LAUNCH()                                              \
    do {                                                                \
        cudaError_t e__ = cudaGetLastError();                           \
        if (e__ != cudaSuccess) {                                       \
            ::xnn::set_error("%s: launch failed: %s", __func__,         \
                             cudaGetErrorString(e__));                  \
            return XNN_ERR_CUDA;                                        \
        }                                                               \
    } while (0)

#define XNN_CUDA(call)                                                  \
    do {                                                                \
        cudaError_t e__ = (call);                                       \
        if (e__ != cudaSuccess) {                                       \
            ::xnn::set_error("%s: %s: %s", __func__, #call,             \
                             cudaGetErrorString(e__));                  \
            return XNN_ERR_CUDA;                                        \
        }                                                               \
    } while (0)

static inline cudaStream_t to_stream(xnn_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }
static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }
static inline int planes_words(int L) { return (L + 31) / 32; }

// ---- unequal batch slices per M tile ---------------------------------------
// The front-end kernels tile the units in blocks of 128 (UMMA M) and split the batch into slices, one CTA per
// (tile, slice).  With U = 300 the third tile holds 44 units: its CTAs have a third of the epilogue / producer work
// of a full tile, and with equal slice counts they idle while the full tiles finish.  A TileMap gives every tile a
// slice count proportional to its cost; CTA c of a 1-D grid finds (tile, slice, slices of that tile) from it.
#define XNN_TILEMAP_MAX 16
struct TileMap {
    int first[XNN_TILEMAP_MAX + 1];   // first CTA of tile t; first[n] = grid size
    int n;                            // 0: plain 2-D grid (tile = blockIdx.x, slice = blockIdx.y)
};
// weight of a tile with `live` of `rows` unit rows in use: a floor for the work that does not shrink with the
// rows (MMAs, TMEM read-out), whole warps of 32 rows otherwise
static inline TileMap make_tile_map(int U, int rows, int ctas, long long B, float floor_w) {
    TileMap m{};
    const int mt = (U + rows - 1) / rows;
    if (mt < 2 || mt > XNN_TILEMAP_MAX || ctas < 2 * mt || (U % rows) == 0) return m;   // n = 0
    float w[XNN_TILEMAP_MAX], tot = 0.0f;
    for (int t = 0; t < mt; ++t) {
        const int live = (t == mt - 1) ? U - t * rows : rows;
        const float f = (float)(((live + 31) / 32) * 32) / (float)rows;
        w[t] = f > floor_w ? f : floor_w;
        tot += w[t];
    }
    int ns[XNN_TILEMAP_MAX], used = 0;
    for (int t = 0; t < mt; ++t) {
        ns[t] = (int)((float)ctas * w[t] / tot);
        if (ns[t] < 1) ns[t] = 1;
        if ((long long)ns[t] > B) ns[t] = (int)B;
        used += ns[t];
    }
    // leftovers go round-robin to the full tiles
    for (int rem = ctas - used, t = 0, guard = 0; rem > 0 && guard < 4 * ctas; ++guard, t = (t + 1) % (mt - 1))
        if ((long long)ns[t] < B) { ++ns[t]; --rem; }
    m.n = mt;
    m.first[0] = 0;
    for (int t = 0; t < mt; ++t) m.first[t + 1] = m.first[t] + ns[t];
    return m;
}
static inline int tile_map_max_slices(const TileMap& m) {
    int mx = 0;
    for (int t = 0; t < m.n; ++t) mx = (m.first[t + 1] - m.first[t] > mx) ? m.first[t + 1] - m.first[t] : mx;
    return mx;
}
__device__ __forceinline__ void tile_of_cta(const TileMap& m, int& tile, int& slice, int& ns) {
    if (m.n == 0) { tile = blockIdx.x; slice = blockIdx.y; ns = gridDim.y; return; }
    const int c = blockIdx.x;
    int t = 0;
    while (t + 1 < m.n && c >= m.first[t + 1]) ++t;
    tile = t; slice = c - m.first[t]; ns = m.first[t + 1] - m.first[t];
}

// ---- device helpers -------------------------------------------------------
// base code at position i of one sequence's planes (lo,hi,nmask each Wp words):
// 0..3 = A,C,G,T, 4 = N (all-zero one-hot column)
__device__ __forceinline__ int plane_code(const uint32_t* __restrict__ seq_planes, int Wp, int i) {
    const int w = i >> 5, b = i & 31;
    const uint32_t lo = seq_planes[w], hi = seq_planes[Wp + w], nm = seq_planes[2 * Wp + w];
    const int c = ((lo >> b) & 1u) | (((hi >> b) & 1u) << 1);
    return ((nm >> b) & 1u) ? 4 : c;
}

// expand one sequence's planes into byte codes in shared memory; positions
// [L, Lpad) are filled with 4 (N).  All threads of the block participate.
__device__ __forceinline__ void expand_codes(const uint32_t* __restrict__ seq_planes, int Wp, int L,
                                             int Lpad, uint8_t* __restrict__ codes) {
    for (int i = threadIdx.x; i < Lpad; i += blockDim.x)
        codes[i] = (i < L) ? (uint8_t)plane_code(seq_planes, Wp, i) : (uint8_t)4;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// counter-based dropout RNG (Philox-4x32-10 on (seed, element index)).  It
// cannot reproduce torch's CPU Philox stream (architectures/__init__.py:92);
// parity tests inject a mask or run with p = 0.
__device__ __forceinline__ uint32_t philox_u32(unsigned long long seed, unsigned long long idx) {
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
    return c0;
}
// keep-scale of dropout for one element: 0 or 1/(1-p)
__device__ __forceinline__ float dropout_scale(float p, unsigned long long seed, unsigned long long idx,
                                               const uint8_t* __restrict__ mask) {
    if (mask) return mask[idx] ? 1.0f / (1.0f - p) : 0.0f;
    if (p <= 0.0f) return 1.0f;
    const float u = (float)(philox_u32(seed, idx) >> 8) * (1.0f / 16777216.0f);
    return (u >= p) ? 1.0f / (1.0f - p) : 0.0f;
}

}  // namespace xnn

// a16 -- the input contract of the hot path (explainn/sequence/__init__.py:8-29,58-60):
// float one-hot / ASCII <-> packed planes, reverse complement in the packed domain.
#include "common.cuh"
using namespace xnn;

// one warp packs 32 consecutive positions of one sequence: each lane classifies
// its column, three ballots give the lo / hi / N words.
__global__ void __launch_bounds__(256)
pack_onehot_kernel(const float* __restrict__ x, long long bs, long long cs, int B, int L, int Wp,
                   uint32_t* __restrict__ planes, int* __restrict__ bad) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const long long total = (long long)B * Wp;
    for (long long item = warp; item < total; item += nwarps) {
        const int b = (int)(item / Wp), w = (int)(item % Wp);
        const int pos = w * 32 + lane;
        int code = 4;
        bool isbad = false;
        if (pos < L) {
            const float* p = x + (long long)b * bs + pos;
            const float v0 = __ldg(p), v1 = __ldg(p + cs), v2 = __ldg(p + 2 * cs), v3 = __ldg(p + 3 * cs);
            const int ones = (v0 == 1.0f) + (v1 == 1.0f) + (v2 == 1.0f) + (v3 == 1.0f);
            const int zeros = (v0 == 0.0f) + (v1 == 0.0f) + (v2 == 0.0f) + (v3 == 0.0f);
            if (ones == 1 && zeros == 3) code = (v1 == 1.0f) ? 1 : (v2 == 1.0f) ? 2 : (v3 == 1.0f) ? 3 : 0;
            else if (zeros == 4) code = 4;
            else { code = 4; isbad = true; }
        }
        const uint32_t lo = __ballot_sync(0xffffffffu, code != 4 && (code & 1));
        const uint32_t hi = __ballot_sync(0xffffffffu, code != 4 && (code & 2));
        const uint32_t nm = __ballot_sync(0xffffffffu, code == 4);
        const uint32_t anybad = __ballot_sync(0xffffffffu, isbad);
        if (lane == 0) {
            uint32_t* o = planes + (long long)b * 3 * Wp;
            o[w] = lo; o[Wp + w] = hi; o[2 * Wp + w] = nm;
            if (anybad && bad) atomicExch(bad, 1);
        }
    }
}

__global__ void __launch_bounds__(256)
pack_ascii_kernel(const uint8_t* __restrict__ seqs, long long rs, int B, int L, int Wp,
                  uint32_t* __restrict__ planes) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const long long total = (long long)B * Wp;
    for (long long item = warp; item < total; item += nwarps) {
        const int b = (int)(item / Wp), w = (int)(item % Wp);
        const int pos = w * 32 + lane;
        int code = 4;
        if (pos < L) {
            const uint8_t ch = seqs[(long long)b * rs + pos] & 0xDF;  // upper-case
            code = (ch == 'A') ? 0 : (ch == 'C') ? 1 : (ch == 'G') ? 2 : (ch == 'T') ? 3 : 4;
        }
        const uint32_t lo = __ballot_sync(0xffffffffu, code != 4 && (code & 1));
        const uint32_t hi = __ballot_sync(0xffffffffu, code != 4 && (code & 2));
        const uint32_t nm = __ballot_sync(0xffffffffu, code == 4);
        if (lane == 0) {
            uint32_t* o = planes + (long long)b * 3 * Wp;
            o[w] = lo; o[Wp + w] = hi; o[2 * Wp + w] = nm;
        }
    }
}

__global__ void __launch_bounds__(256)
unpack_onehot_kernel(const uint32_t* __restrict__ planes, int B, int L, int Wp, float* __restrict__ x) {
    const long long total = (long long)B * 4 * L;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int pos = (int)(i % L);
        const int c = (int)((i / L) % 4);
        const int b = (int)(i / (4LL * L));
        const int code = plane_code(planes + (long long)b * 3 * Wp, Wp, pos);
        x[i] = (code == c) ? 1.0f : 0.0f;
    }
}

// out position i takes the complement of in position L-1-i
__global__ void __launch_bounds__(256)
revcomp_kernel(const uint32_t* __restrict__ in, int B, int L, int Wp, uint32_t* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const long long total = (long long)B * Wp;
    for (long long item = warp; item < total; item += nwarps) {
        const int b = (int)(item / Wp), w = (int)(item % Wp);
        const int pos = w * 32 + lane;
        int code = 4;
        if (pos < L) {
            const int c = plane_code(in + (long long)b * 3 * Wp, Wp, L - 1 - pos);
            code = (c == 4) ? 4 : 3 - c;
        }
        const uint32_t lo = __ballot_sync(0xffffffffu, code != 4 && (code & 1));
        const uint32_t hi = __ballot_sync(0xffffffffu, code != 4 && (code & 2));
        const uint32_t nm = __ballot_sync(0xffffffffu, code == 4);
        if (lane == 0) {
            uint32_t* o = out + (long long)b * 3 * Wp;
            o[w] = lo; o[Wp + w] = hi; o[2 * Wp + w] = nm;
        }
    }
}

static int grid_for_warps(long long warps) {
    long long blocks = (warps + 7) / 8;
    const long long cap = (long long)num_sms() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

extern "C" int xnn_pack_onehot_f32(const float* x, int64_t batch_stride, int64_t chan_stride, int B, int L,
                                   uint32_t* planes, int* bad_flag, xnn_stream_t s) {
    XNN_CHECK_ARG(x && planes, "NULL pointer");
    XNN_CHECK_ARG(B >= 0 && L > 0, "bad sizes");
    if (B == 0) return XNN_OK;
    const int Wp = planes_words(L);
    pack_onehot_kernel<<<grid_for_warps((long long)B * Wp), 256, 0, to_stream(s)>>>(
        x, batch_stride, chan_stride, B, L, Wp, planes, bad_flag);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

extern "C" int xnn_pack_ascii(const uint8_t* seqs, int64_t row_stride, int B, int L, uint32_t* planes,
                              xnn_stream_t s) {
    XNN_CHECK_ARG(seqs && planes, "NULL pointer");
    XNN_CHECK_ARG(B >= 0 && L > 0, "bad sizes");
    if (B == 0) return XNN_OK;
    const int Wp = planes_words(L);
    pack_ascii_kernel<<<grid_for_warps((long long)B * Wp), 256, 0, to_stream(s)>>>(seqs, row_stride, B, L, Wp,
                                                                                  planes);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

extern "C" int xnn_unpack_onehot_f32(const uint32_t* planes, int B, int L, float* x, xnn_stream_t s) {
    XNN_CHECK_ARG(x && planes, "NULL pointer");
    XNN_CHECK_ARG(B >= 0 && L > 0, "bad sizes");
    if (B == 0) return XNN_OK;
    const long long total = (long long)B * 4 * L;
    long long blocks = (total + 255) / 256;
    const long long cap = (long long)num_sms() * 16;
    if (blocks > cap) blocks = cap;
    unpack_onehot_kernel<<<(int)blocks, 256, 0, to_stream(s)>>>(planes, B, L, planes_words(L), x);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

extern "C" int xnn_revcomp_planes(const uint32_t* planes, int B, int L, uint32_t* out, xnn_stream_t s) {
    XNN_CHECK_ARG(out && planes && out != planes, "NULL or aliased pointer");
    XNN_CHECK_ARG(B >= 0 && L > 0, "bad sizes");
    if (B == 0) return XNN_OK;
    const int Wp = planes_words(L);
    revcomp_kernel<<<grid_for_warps((long long)B * Wp), 256, 0, to_stream(s)>>>(planes, B, L, Wp, out);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

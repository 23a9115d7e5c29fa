// Step-level pieces that keep the training step free of host scalars and of library kernels, so that the whole
// step (forward, loss, backward, all-reduce, Adam) can be captured into ONE CUDA graph:
//   xnn_counters_advance   device-resident step counters (Adam's bias-correction step, the dropout draw index)
//   xnn_loss_fwd / _bwd    mean MSE / BCE-with-logits and their gradient (explainn/architectures/__init__.py:446-456,
//                          nn.MSELoss() / nn.BCEWithLogitsLoss() with the default 'mean' reduction)
#include "common.cuh"
using namespace xnn;

__global__ void counters_advance_kernel(long long* __restrict__ c, int n) {
    if (threadIdx.x < n) c[threadIdx.x] += 1;
}

extern "C" int xnn_counters_advance(long long* counters, int n, xnn_stream_t s) {
    XNN_CHECK_ARG(counters && n >= 1 && n <= 32, "counters: 1..32 int64 values");
    counters_advance_kernel<<<1, 32, 0, to_stream(s)>>>(counters, n);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

struct CounterList { long long* p[8]; int n; };
__global__ void counters_advance_list_kernel(const __grid_constant__ CounterList L) {
    if ((int)threadIdx.x < L.n) *L.p[threadIdx.x] += 1;
}

// the same for up to 8 counters that live in separate allocations (the three num_batches_tracked buffers of
// linears.1 / .7 / .11, explainn/architectures/__init__.py:79,90,99): one launch instead of three library adds
extern "C" int xnn_counters_advance_list(int n, long long* const* ptrs, xnn_stream_t s) {
    XNN_CHECK_ARG(ptrs && n >= 1 && n <= 8, "1..8 counters");
    CounterList L{};
    L.n = n;
    for (int i = 0; i < n; ++i) { XNN_CHECK_ARG(ptrs[i] != nullptr, "NULL counter"); L.p[i] = ptrs[i]; }
    counters_advance_list_kernel<<<1, 32, 0, to_stream(s)>>>(L);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

// ---------------------------------------------------------------------------
// running statistics of up to 4 train-mode BatchNorm layers + their num_batches_tracked, in one launch, from the
// `stats` rows (mean, biased variance, ...) the statistics passes produced: nn.BatchNorm1d's update
//   running = (1 - momentum) * running + momentum * (mean | variance * count / (count - 1))
// (explainn/architectures/__init__.py:79,90,99; default momentum 0.1).  Kept apart from the statistics passes so
// that the training forward is a FUNCTIONAL op (torch.library.register_autograd requires one).
// ---------------------------------------------------------------------------
struct BnUpdList {
    const double* stats[4];
    float* rm[4];
    float* rv[4];
    long long* nbt[4];
    int C[4];
    float momentum[4];
    double count[4];
    int first_block[5];
    int n;
};

__global__ void __launch_bounds__(256) bn_running_update_kernel(const __grid_constant__ BnUpdList L) {
    int t = 0;
    while (t + 1 < L.n && (int)blockIdx.x >= L.first_block[t + 1]) ++t;
    const int c = ((int)blockIdx.x - L.first_block[t]) * 256 + (int)threadIdx.x;
    if (c == 0 && L.nbt[t] != nullptr) *L.nbt[t] += 1;
    if (c >= L.C[t]) return;
    const double mean = L.stats[t][4LL * c], var = L.stats[t][4LL * c + 1];
    const double N = L.count[t], mom = (double)L.momentum[t];
    const double unb = (N > 1.0) ? var * N / (N - 1.0) : var;
    L.rm[t][c] = (float)((1.0 - mom) * (double)L.rm[t][c] + mom * mean);
    L.rv[t][c] = (float)((1.0 - mom) * (double)L.rv[t][c] + mom * unb);
}

extern "C" int xnn_bn_running_update(int n_layers, const double* const* stats, float* const* running_mean,
                                     float* const* running_var, long long* const* num_batches_tracked,
                                     const int* channels, const float* momentum, const double* count,
                                     xnn_stream_t s) {
    XNN_CHECK_ARG(n_layers >= 1 && n_layers <= 4, "1..4 layers");
    XNN_CHECK_ARG(stats && running_mean && running_var && channels && momentum && count, "NULL pointer");
    BnUpdList L{};
    int blocks = 0;
    for (int i = 0; i < n_layers; ++i) {
        XNN_CHECK_ARG(stats[i] && running_mean[i] && running_var[i] && channels[i] >= 1 && count[i] >= 1.0, "bad layer");
        L.stats[i] = stats[i]; L.rm[i] = running_mean[i]; L.rv[i] = running_var[i];
        L.nbt[i] = num_batches_tracked ? num_batches_tracked[i] : nullptr;
        L.C[i] = channels[i]; L.momentum[i] = momentum[i]; L.count[i] = count[i];
        L.first_block[i] = blocks;
        blocks += (channels[i] + 255) / 256;
    }
    L.first_block[n_layers] = blocks;
    L.n = n_layers;
    bn_running_update_kernel<<<blocks, 256, 0, to_stream(s)>>>(L);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

#define LOSS_THREADS 256
#define LOSS_MAX_BLOCKS 256

__device__ __forceinline__ float loss_term(int kind, float x, float y) {
    if (kind == 0) { const float d = x - y; return d * d; }
    // BCE with logits, the numerically stable form torch uses: max(x,0) - x*y + log(1 + exp(-|x|))
    return fmaxf(x, 0.0f) - x * y + log1pf(expf(-fabsf(x)));
}

// deterministic two-level reduction in one launch: every block writes an fp64 partial, the last block to finish
// (atomic ticket) adds the partials in block order
__global__ void __launch_bounds__(LOSS_THREADS)
loss_fwd_kernel(const float* __restrict__ x, const float* __restrict__ y, long long n, int kind,
                double* __restrict__ part, unsigned int* __restrict__ ticket, float* __restrict__ loss) {
    __shared__ double sh[LOSS_THREADS / 32];
    __shared__ bool last;
    double acc = 0.0;
    for (long long i = (long long)blockIdx.x * LOSS_THREADS + threadIdx.x; i < n; i += (long long)gridDim.x * LOSS_THREADS)
        acc += (double)loss_term(kind, x[i], y[i]);
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < LOSS_THREADS / 32; ++w) t += sh[w];
        part[blockIdx.x] = t;
        __threadfence();
        last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (last) {
        // the last block adds the partials as a fixed tree (thread b holds block b's partial: gridDim.x <= LOSS_THREADS):
        // deterministic whatever the order in which the blocks finished; a serial walk over 256 partials by one thread
        // was most of the kernel's 13 us
        __threadfence();
        double t = threadIdx.x < gridDim.x ? ((volatile double*)part)[threadIdx.x] : 0.0;
        t = warp_sum(t);
        __syncthreads();                 // sh[] of the first reduction has been read by thread 0
        if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = t;
        __syncthreads();
        if (threadIdx.x == 0) {
            double tot = 0.0;
#pragma unroll
            for (int w = 0; w < LOSS_THREADS / 32; ++w) tot += sh[w];
            *loss = (float)(tot / (double)n);
            *ticket = 0u;                // ready for the next launch (graph replays reuse the work buffer)
        }
    }
}

__global__ void __launch_bounds__(LOSS_THREADS)
loss_bwd_kernel(const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ gout, long long n,
                int kind, float* __restrict__ dx) {
    const float g = (gout ? *gout : 1.0f) / (float)n;
    for (long long i = (long long)blockIdx.x * LOSS_THREADS + threadIdx.x; i < n; i += (long long)gridDim.x * LOSS_THREADS) {
        const float xi = x[i], yi = y[i];
        float d;
        if (kind == 0) d = 2.0f * (xi - yi);
        else d = 1.0f / (1.0f + expf(-xi)) - yi;
        dx[i] = g * d;
    }
}

static int loss_blocks(long long n) {
    long long b = (n + LOSS_THREADS * 4 - 1) / (LOSS_THREADS * 4);
    if (b > LOSS_MAX_BLOCKS) b = LOSS_MAX_BLOCKS;
    return b < 1 ? 1 : (int)b;
}

extern "C" size_t xnn_loss_work_bytes(void) { return LOSS_MAX_BLOCKS * sizeof(double) + 16; }

extern "C" int xnn_loss_fwd(const float* logits, const float* target, long long n, int kind, float* loss, void* work,
                            xnn_stream_t s) {
    XNN_CHECK_ARG(logits && target && loss && work, "NULL pointer");
    XNN_CHECK_ARG(n >= 1 && (kind == 0 || kind == 1), "kind: 0 = MSE, 1 = BCE with logits");
    XNN_CHECK_ARG(((uintptr_t)work & 7) == 0, "work must be 8-byte aligned");
    double* part = (double*)work;
    unsigned int* ticket = (unsigned int*)(part + LOSS_MAX_BLOCKS);
    loss_fwd_kernel<<<loss_blocks(n), LOSS_THREADS, 0, to_stream(s)>>>(logits, target, n, kind, part, ticket, loss);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

extern "C" int xnn_loss_bwd(const float* logits, const float* target, const float* grad_out, long long n, int kind,
                            float* dlogits, xnn_stream_t s) {
    XNN_CHECK_ARG(logits && target && dlogits, "NULL pointer");
    XNN_CHECK_ARG(n >= 1 && (kind == 0 || kind == 1), "kind: 0 = MSE, 1 = BCE with logits");
    long long b = (n + LOSS_THREADS - 1) / LOSS_THREADS;
    if (b > (long long)num_sms() * 8) b = (long long)num_sms() * 8;
    loss_bwd_kernel<<<(int)b, LOSS_THREADS, 0, to_stream(s)>>>(logits, target, grad_out, n, kind, dlogits);
    XNN_CHECK_LAUNCH();
    return XNN_OK;
}

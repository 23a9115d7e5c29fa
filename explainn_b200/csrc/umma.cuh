// Thin inline-PTX layer over the sm_100a tensor-core path used by the Toeplitz
// front-end kernels: tcgen05.mma (kind::f16, bf16 x bf16 -> fp32 in TMEM),
// tcgen05.alloc/dealloc/ld/commit/fence, mbarrier wait, async-proxy fence.
// Bit layouts of the shared-memory and instruction descriptors follow the PTX
// ISA "tcgen05 matrix descriptor" / "instruction descriptor" tables.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace xnn {
namespace umma {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

// K-major, no swizzle.  In 16-byte units the operand is ((8,m),2):((1,SBO),LBO):
//   row r, 16-byte K chunk c  ->  start + (r%8)*16 + (r/8)*SBO + c*LBO   bytes
// version field (bits 46-47) = 1 on sm_100.
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ULL << 46);
}

// kind::f16 instruction descriptor: D = fp32, A = B = bf16, both K-major, dense.
__host__ __device__ constexpr uint32_t idesc_bf16_f32(int M, int N) {
    return (1u << 4)                      // c_format  = F32
           | (1u << 7)                    // a_format  = BF16
           | (1u << 10)                   // b_format  = BF16
           | ((uint32_t)(N >> 3) << 17)   // n_dim
           | ((uint32_t)(M >> 4) << 24);  // m_dim
}

__device__ __forceinline__ void mma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// The same MMA issued by ONE elected lane of a converged warp: every lane executes the statement with
// warp-uniform operands (so the compiler keeps the descriptors in uniform registers and emits the UTCHMMA
// without a per-lane waterfall loop); elect.sync picks the issuing lane.
__device__ __forceinline__ void mma_bf16_elect(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                               uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// one lane of the (converged) warp: the body of `if (elect_one())` is a single-thread region in which
// tcgen05.mma / tcgen05.commit issue once
__device__ __forceinline__ bool elect_one() {
    uint32_t p;
    asm volatile(
        "{\n\t"
        ".reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, q;\n\t"
        "}\n"
        : "=r"(p));
    return p != 0u;
}
// a value every lane holds identically, in a form the compiler can keep in a uniform register
__device__ __forceinline__ uint32_t warp_uniform(uint32_t v) { return __shfl_sync(0xffffffffu, v, 0); }

__device__ __forceinline__ void commit_elect(uint64_t* bar) {
    asm volatile(
        "{\n\t"
        ".reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t"
        "}\n" ::"r"(smem_u32(bar))
        : "memory");
}

// kind::tf32 instruction descriptor: D = fp32, A = B = tf32 (32-bit containers in shared memory);
// a_mn / b_mn = 1 selects the MN-major (transposed) view of that operand.
__host__ __device__ constexpr uint32_t idesc_tf32_f32(int M, int N, int a_mn, int b_mn) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// x = hi + lo with hi, lo both exactly representable in tf32 up to 2^-22 |x| (round to nearest)
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
    uint32_t h, l;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
    hi = __uint_as_float(h);
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(x - hi));
    lo = __uint_as_float(l);
}

// ---- descriptor tables ---------------------------------------------------------------------------
// The issuing thread is a single lane: every integer instruction it spends building descriptors sits on
// the tensor core's critical path (a K = 16 step of a 128 x 64 tile is only 32 cycles of math).  The
// operand addresses of a tile are known up front, so each kernel precomputes the low descriptor words
// (start address >> 4 | LBO << 16) of its MMA sequence into shared memory once; the issue loop is then
// one 8-byte load per MMA.  The high word (SBO | version) is a per-operand constant.
__device__ __forceinline__ uint32_t desc_lo(uint32_t saddr, uint32_t lbo_bytes) {
    return ((saddr & 0x3FFFFu) >> 4) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
}
__host__ __device__ constexpr uint32_t desc_hi(uint32_t sbo_bytes) { return ((sbo_bytes >> 4) & 0x3FFFu) | (1u << 14); }

// issue `count` MMAs D (+)= A_i . B_i from a table of (A low word, B low word); add_b is added to every
// B low word (a per-tile start offset >> 4); the first MMA uses `first_acc`, the rest accumulate.
template <bool TF32>
__device__ __forceinline__ void issue_table(const uint2* __restrict__ tab, int count, uint32_t hi_a, uint32_t hi_b,
                                            uint32_t add_b, uint32_t d_tmem, uint32_t idesc, uint32_t first_acc) {
    // batches of 6: the table loads of a batch are issued together, ahead of its MMAs
    for (int i0 = 0; i0 < count; i0 += 6) {
        uint2 e[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) e[j] = tab[min(i0 + j, count - 1)];
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            if (i0 + j < count) {
                const uint64_t ad = ((uint64_t)hi_a << 32) | (uint64_t)e[j].x;
                const uint64_t bd = ((uint64_t)hi_b << 32) | (uint64_t)(e[j].y + add_b);
                const uint32_t acc = (i0 + j == 0) ? first_acc : 1u;
                if (TF32) mma_tf32(d_tmem, ad, bd, idesc, acc);
                else mma_bf16(d_tmem, ad, bd, idesc, acc);
            }
        }
    }
}

// kind::f16 variant with a per-tile offset for BOTH operands (double-buffered operand stages)
__device__ __forceinline__ void issue_table2(const uint2* __restrict__ tab, int count, uint32_t hi_a, uint32_t hi_b,
                                             uint32_t add_a, uint32_t add_b, uint32_t d_tmem, uint32_t idesc) {
    for (int i0 = 0; i0 < count; i0 += 6) {
        uint2 e[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) e[j] = tab[min(i0 + j, count - 1)];
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            if (i0 + j < count) {
                const uint64_t ad = ((uint64_t)hi_a << 32) | (uint64_t)(e[j].x + add_a);
                const uint64_t bd = ((uint64_t)hi_b << 32) | (uint64_t)(e[j].y + add_b);
                mma_bf16(d_tmem, ad, bd, idesc, (i0 + j == 0) ? 0u : 1u);
            }
        }
    }
}

// arrive on an mbarrier when every previously issued tcgen05.mma of this thread has completed
__device__ __forceinline__ void commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy shared-memory writes -> visible to the async proxy (tcgen05.mma operand reads)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// one warp: allocate `cols` TMEM columns (power of two >= 32); the base address lands in *slot (shared)
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}

// 32 lanes x 16 consecutive 32-bit columns: thread t of the warp gets lane (taddr.lane + t)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld2(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(taddr) : "memory");
}
// 14 consecutive columns (4 pool windows of one position parity) as 8 + 4 + 2; asynchronous until tmem_ld_wait
__device__ __forceinline__ void tmem_ld14(uint32_t taddr, uint32_t (&r)[14]) {
    tmem_ld8(taddr, &r[0]);
    tmem_ld4(taddr + 8, &r[8]);
    tmem_ld2(taddr + 12, &r[12]);
}
// wait for the outstanding tcgen05.ld's; the loaded registers are tied to the statement so that no use of them
// can be scheduled above the wait
__device__ __forceinline__ void tmem_ld_wait14(uint32_t (&a)[14], uint32_t (&b)[14]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7]), "+r"(a[8]), "+r"(a[9]), "+r"(a[10]), "+r"(a[11]), "+r"(a[12]), "+r"(a[13]), "+r"(b[0]), "+r"(b[1]), "+r"(b[2]), "+r"(b[3]), "+r"(b[4]), "+r"(b[5]), "+r"(b[6]), "+r"(b[7]), "+r"(b[8]), "+r"(b[9]), "+r"(b[10]), "+r"(b[11]), "+r"(b[12]), "+r"(b[13]) : : "memory");
}
// 28 consecutive columns (a quarter of a 112-column pool tile) as 16 + 8 + 4
__device__ __forceinline__ void tmem_ld28(uint32_t taddr, uint32_t (&r)[28]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    tmem_ld8(taddr + 16, &r[16]);
    tmem_ld4(taddr + 24, &r[24]);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// Wait for the phase of parity `parity` to complete.  try_wait WITHOUT a suspend-time hint: with the 10 ms hint used
// before, a failed first check compiled to NANOSLEEP.SYNCS 0x989680 and the wake-up after the arrival cost several
// hundred cycles per wait -- on every hop of the producer / issuer / epilogue chains (measured with clock64: no wait
// section ever took less than ~400 cycles, even with nothing to wait for).
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    uint32_t done = 0, spins = 0;
    while (!done) {
        if (++spins > (1u << 26)) __trap();   // a lost arrival must fail loudly, never hang the GPU
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}\n"
            : "=r"(done)
            : "r"(a), "r"(parity)
            : "memory");
    }
}

// The six significant products of a three-way bf16 split (A piece, B piece), smallest first:
// lo*hi, hi*lo, mid*mid, mid*hi, hi*mid, hi*hi  (0 = hi, 1 = mid, 2 = lo); each runs `ksteps` K steps.
// To be called by ONE elected lane (`if (elect_one())`) with warp-uniform bases (warp_uniform()): the descriptor
// arithmetic then stays in the uniform datapath, ~5 instructions per MMA.  (Descriptor tables read under
// `if (lane == 0)` cost an election / R2UR waterfall of ~30 instructions around every MMA.)
//   a_base / b_base: desc_lo of piece 0, K step 0 (LBO folded in);  *_piece, *_kstep: strides in 16-byte units
__device__ __forceinline__ void issue_split6(uint32_t a_base, uint32_t a_piece, uint32_t a_kstep, uint32_t hi_a,
                                             uint32_t b_base, uint32_t b_piece, uint32_t b_kstep, uint32_t hi_b,
                                             int ksteps, uint32_t d_tmem, uint32_t idesc) {
    uint32_t acc = 0;
#pragma unroll
    for (int pass = 0; pass < 6; ++pass) {
        const uint32_t pa = pass == 0 ? 2u : pass == 1 ? 0u : pass == 2 ? 1u : pass == 3 ? 1u : 0u;
        const uint32_t pb = pass == 0 ? 0u : pass == 1 ? 2u : pass == 2 ? 1u : pass == 3 ? 0u : pass == 4 ? 1u : 0u;
        uint32_t ad = a_base + pa * a_piece, bd = b_base + pb * b_piece;
#pragma unroll 2
        for (int ks = 0; ks < ksteps; ++ks) {
            mma_bf16(d_tmem, ((uint64_t)hi_a << 32) | (uint64_t)ad, ((uint64_t)hi_b << 32) | (uint64_t)bd, idesc, acc);
            acc = 1;
            ad += a_kstep;
            bd += b_kstep;
        }
    }
}

// The three significant products of a two-way bf16 split (A piece, B piece), smallest first: mid*hi, hi*mid, hi*hi
// (0 = hi, 1 = mid).  Same calling convention as issue_split6; first_acc = 0 overwrites the accumulator.
__device__ __forceinline__ void issue_split3(uint32_t a_base, uint32_t a_piece, uint32_t a_kstep, uint32_t hi_a,
                                             uint32_t b_base, uint32_t b_piece, uint32_t b_kstep, uint32_t hi_b,
                                             int ksteps, uint32_t d_tmem, uint32_t idesc, uint32_t first_acc) {
    uint32_t acc = first_acc;
#pragma unroll
    for (int pass = 0; pass < 3; ++pass) {
        uint32_t ad = a_base + (pass == 0 ? a_piece : 0u), bd = b_base + (pass == 1 ? b_piece : 0u);
#pragma unroll 2
        for (int ks = 0; ks < ksteps; ++ks) {
            mma_bf16(d_tmem, ((uint64_t)hi_a << 32) | (uint64_t)ad, ((uint64_t)hi_b << 32) | (uint64_t)bd, idesc, acc);
            acc = 1;
            ad += a_kstep;
            bd += b_kstep;
        }
    }
}

// ---- bulk asynchronous copies (the TMA engine's non-tensor mode; SASS: UBLKCP) --------------------------------
// global -> shared, completion counted in bytes on an mbarrier: the operand stage of a tile whose image in global
// memory already IS the shared-memory layout arrives without a single LSU instruction
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_load(void* sdst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(sdst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// shared -> global, tracked by the issuing thread's bulk groups
__device__ __forceinline__ void bulk_store(void* gdst, const void* ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit_group() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// the shared-memory source of every committed bulk store of this thread has been read (the buffer may be rewritten)
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

}  // namespace umma
}  // namespace xnn

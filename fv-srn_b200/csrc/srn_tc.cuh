// tcgen05 / TMEM core of the fused SRN evaluation for sm_100a (fast mode).
//
// Stands in for reference renderer/renderer_volume_tensorcores.cuh:839-1164 (the layer chain of
// VolumeInterpolationTensorcores::eval). One CTA of 128 threads evaluates 128 samples per call:
// thread t owns sample t = row t of every MMA tile = TMEM lane t.
//   * A operand (activations) lives in TMEM, written by the owning thread with tcgen05.st (fp16x2 per 32-bit column,
//     8 columns per K=16 step) - activations never touch shared memory;
//   * B operand (weights) lives in shared memory as K-major no-swizzle UMMA tiles (8 rows x 16 B core matrices),
//     pinned for the CTA lifetime;
//   * D accumulates in fp32 in TMEM; one elected thread issues tcgen05.mma.cta_group::1.kind::f16 (M=128, N=H, K=16)
//     and tcgen05.commit's to an mbarrier every thread waits on;
//   * the bias is folded into the MMA: a constant "ones" A chunk [1,0,...,0] multiplies a weight chunk that holds the
//     bias in its first K column (for layer 0 of networks without a time input the bias sits in the unused input
//     channel 3 instead, which the row builder sets to 1);
//   * epilogue per layer: tcgen05.ld 32 columns -> activation in fp32 -> cvt.rn[.relu].f16x2.f32 -> tcgen05.st as the
//     next layer's A operand.
#pragma once
#include "srn_core.cuh"

namespace fvsrn {
namespace tc {

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" :: "r"(smem_dst), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" :: "r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" :: "r"(bar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok;
}
// -DFVSRN_BOUNDED_WAIT: bring-up builds turn a protocol bug into a launch failure (trap) instead of a hung device.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
#ifdef FVSRN_BOUNDED_WAIT
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > (1u << 24)) __trap();
    }
#else
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "WAIT_%=:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"   // suspends up to the hint: no busy polling
                 "@p bra DONE_%=;\n\t"
                 "bra WAIT_%=;\n\t"
                 "DONE_%=:\n\t}\n" :: "r"(bar), "r"(parity), "r"(0x989680u) : "memory");
#endif
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" :: "r"(bar) : "memory");
}
// converged warp: one elected lane arrives (no %tid read, no divergent branch)
__device__ __forceinline__ void mbar_arrive_elect(uint32_t bar) {
    asm volatile("{\n\t.reg .pred q;\n\t"
                 "elect.sync _|q, 0xffffffff;\n\t"
                 "@q mbarrier.arrive.shared::cta.b64 _, [%0];\n\t}\n" :: "r"(bar) : "memory");
}
// arrive on the mbarrier once all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void mma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" :: "r"(bar) : "memory");
}
// same with the descriptor given as its two 32-bit halves (the K loop only advances the low word)
__device__ __forceinline__ void mma_ts2(uint32_t d_tmem, uint32_t a_tmem, uint32_t desc_lo, uint32_t desc_hi, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\t.reg .b64 d;\n\t"
                 "setp.ne.b32 p, %5, 0;\n\t"
                 "mov.b64 d, {%2, %3};\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], d, %4, p;\n\t}\n"
                 :: "r"(d_tmem), "r"(a_tmem), "r"(desc_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
}
// Warp-convergent variants: every lane of a converged warp executes them, one elected lane issues (no divergent
// branch around the uniform-datapath instruction, so the operands go to uniform registers without an election loop).
__device__ __forceinline__ void mma_ts2_elect(uint32_t d_tmem, uint32_t a_tmem, uint32_t desc_lo, uint32_t desc_hi, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p, q;\n\t.reg .b64 d;\n\t"
                 "elect.sync _|q, 0xffffffff;\n\t"
                 "setp.ne.b32 p, %5, 0;\n\t"
                 "mov.b64 d, {%2, %3};\n\t"
                 "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], d, %4, p;\n\t}\n"
                 :: "r"(d_tmem), "r"(a_tmem), "r"(desc_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
}
// SS mode: A and B both by shared-memory descriptor
__device__ __forceinline__ void mma_ss2_elect(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p, q;\n\t.reg .b64 da, db;\n\t"
                 "elect.sync _|q, 0xffffffff;\n\t"
                 "setp.ne.b32 p, %6, 0;\n\t"
                 "mov.b64 da, {%1, %2};\n\t"
                 "mov.b64 db, {%3, %4};\n\t"
                 "@q tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}\n"
                 :: "r"(d_tmem), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_commit_elect(uint32_t bar) {
    asm volatile("{\n\t.reg .pred q;\n\t"
                 "elect.sync _|q, 0xffffffff;\n\t"
                 "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}\n" :: "r"(bar) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]^T, kind::f16, issued by ONE thread
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "setp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
                 :: "r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// K-major, no swizzle: core matrix = 8 rows x 16 B (128 B contiguous); SBO = byte distance between 8-row groups
// (along N), LBO = byte distance between the two 8-element K chunks of one K=16 step; version 1 (Blackwell).
__device__ __forceinline__ uint64_t make_b_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return uint64_t((smem_addr >> 4) & 0x3FFF) | (uint64_t((lbo_bytes >> 4) & 0x3FFF) << 16) |
           (uint64_t((sbo_bytes >> 4) & 0x3FFF) << 32) | (uint64_t(1) << 46);
}
// instruction descriptor: D = F32 (or F16 when acc16), A = B = F16, both K-major, dense
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, bool acc16 = false) {
    return (acc16 ? 0u : (1u << 4)) | (uint32_t(N >> 3) << 17) | (uint32_t(M >> 4) << 24);
}

#define FV_R8(r, o)  "=r"(r[o+0]), "=r"(r[o+1]), "=r"(r[o+2]), "=r"(r[o+3]), "=r"(r[o+4]), "=r"(r[o+5]), "=r"(r[o+6]), "=r"(r[o+7])
#define FV_W8(r, o)  "r"(r[o+0]), "r"(r[o+1]), "r"(r[o+2]), "r"(r[o+3]), "r"(r[o+4]), "r"(r[o+5]), "r"(r[o+6]), "r"(r[o+7])

// this thread's TMEM lane, 32 / 16 / 4 / 1 consecutive 32-bit columns
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n"
                 : FV_R8(r, 0), FV_R8(r, 8), FV_R8(r, 16), FV_R8(r, 24) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
                 : FV_R8(r, 0), FV_R8(r, 8) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld1(uint32_t taddr, uint32_t& r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];\n" : "=r"(r) : "r"(taddr) : "memory");
}
// fp16 accumulators (one per 32-bit column): 32 / 16 consecutive columns packed two per register
__device__ __forceinline__ void tmem_ld16_pack(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.pack::16b.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
                 : FV_R8(r, 0), FV_R8(r, 8) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld8_pack(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.pack::16b.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n"
                 : FV_R8(r, 0) : "r"(taddr) : "memory");
}
template <int N>
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[N], int o) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};\n"
                 :: "r"(taddr), FV_W8(r, o) : "memory");
}
template <int N>
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[N], int o) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};\n"
                 :: "r"(taddr), FV_W8(r, o), FV_W8(r, o + 8) : "memory");
}

__device__ __forceinline__ uint32_t pack_relu(float lo, float hi) {   // (half(max(lo,0)), half(max(hi,0)))
    uint32_t r; asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;\n" : "=r"(r) : "f"(hi), "f"(lo)); return r;
}
// filtered fetch of a half4 3-D texture with the result left in fp16 (two f16x2 registers): no fp32 round trip
__device__ __forceinline__ void tex3d_h4(cudaTextureObject_t t, float x, float y, float z, uint32_t& xy, uint32_t& zw) {
    asm("tex.3d.v2.f16x2.f32 {%0, %1}, [%2, {%3, %4, %5, %5}];" : "=r"(xy), "=r"(zw) : "l"(t), "f"(x), "f"(y), "f"(z));
}

__device__ __forceinline__ uint32_t pack_h2(float lo, float hi) {
    uint32_t r; asm("cvt.rn.f16x2.f32 %0, %1, %2;\n" : "=r"(r) : "f"(hi), "f"(lo)); return r;
}

// hidden activation on two fp32 accumulators -> packed fp16x2 (reference renderer_activations.cuh:39-341, evaluated in fp32)
struct ActF {
    float p, q, inv_q, inv_q2;   // Sine/Snake: p = param; Snake: q = 1/param; SnakeAlt: q = 2*param, inv_q = 1/q, inv_q2 = 1/q^2
    int act;                     // the activation code itself, for the instantiations that take it at run time (ACT = -1)
};
// D_ACT_* plus: SnakeAlt on accumulators that already hold q*v (weights and biases of the producing layer pre-multiplied
// by q = 2*param when that is a power of two - exact in fp16; see build_tc_blob)
constexpr int D_ACT_SNAKEALT_PRE = 6;
__device__ __forceinline__ ActF make_actf(int act, float param) {
    ActF a; a.p = param; a.act = act;
    a.q = (act == D_ACT_SNAKE) ? 1.0f / param : 2.0f * param;
    a.inv_q = 1.0f / a.q;
    a.inv_q2 = a.inv_q * a.inv_q;
    return a;
}
// SnakeAlt (v + 1 - cos(q v)) / q as fma(-cos(q v), 1/q, fma(v, 1/q, 1/q)): 5 issue slots per element instead of 6; on
// pre-scaled accumulators z = q v: fma(-cos z, 1/q, fma(z, 1/q^2, 1/q)): 4.
// (Packed-fp32 FMUL2 / FFMA2 forms of these activations were measured slower on B200: 14.57 vs 13.58 ms on C3-snake, r02c.)
// ACT = -1: the activation code is a.act, uniform at run time (generic instantiations: Sine, Snake, SnakeAlt with a
// parameter that is no power of two). Measured on C3-snake (r02c, one B200): this three-way run-time chain 12.65 ms; with
// the pre-scaled form added to the chain 21.0 ms (the compiler stops hoisting it out of the element loop); a switch over
// per-activation instantiations 13.32 ms; the pre-scaled SnakeAlt as its own kernel instantiation (VARIANT bit 2) 12.1 ms.
template <int ACT>
__device__ __forceinline__ float act1(const ActF& a, float v) {
    if constexpr (ACT == -1) {
        if (a.act == D_ACT_SINE) return __sinf(v * a.p);
        if (a.act == D_ACT_SNAKE) { const float s = __sinf(a.p * v); return fmaf(a.q * s, s, v); }
        /* SnakeAlt */ return fmaf(-__cosf(a.q * v), a.inv_q, fmaf(v, a.inv_q, a.inv_q));
    }
    else if constexpr (ACT == D_ACT_SINE) return __sinf(v * a.p);
    else if constexpr (ACT == D_ACT_SNAKE) { const float s = __sinf(a.p * v); return fmaf(a.q * s, s, v); }
    else if constexpr (ACT == D_ACT_SNAKEALT) return fmaf(-__cosf(a.q * v), a.inv_q, fmaf(v, a.inv_q, a.inv_q));
    else if constexpr (ACT == D_ACT_SNAKEALT_PRE) return fmaf(-__cosf(v), a.inv_q, fmaf(v, a.inv_q2, a.inv_q));
    else return v;
}
template <int ACT>
__device__ __forceinline__ uint32_t act_pack(const ActF& a, uint32_t lo, uint32_t hi) {
    const float x = __uint_as_float(lo), y = __uint_as_float(hi);
    if constexpr (ACT == D_ACT_RELU) return pack_relu(x, y);
    else if constexpr (ACT == -1) {
        if (a.act == D_ACT_RELU) return pack_relu(x, y);
        if (a.act == D_ACT_NONE) return pack_h2(x, y);
        return pack_h2(act1<ACT>(a, x), act1<ACT>(a, y));
    }
    else return pack_h2(act1<ACT>(a, x), act1<ACT>(a, y));
}
// the same on a packed fp16x2 accumulator pair (FVSRN_TC_ACC16)
template <int ACT>
__device__ __forceinline__ uint32_t act_h2(const ActF& a, uint32_t v) {
    const __half2 h = u2h(v);
    if constexpr (ACT == D_ACT_RELU) return h2u(__hmax2(h, __float2half2_rn(0.f)));
    else if constexpr (ACT == D_ACT_NONE) return v;
    else {
        if constexpr (ACT == -1) { if (a.act == D_ACT_RELU) return h2u(__hmax2(h, __float2half2_rn(0.f))); if (a.act == D_ACT_NONE) return v; }
        const float2 f = __half22float2(h); return pack_h2(act1<ACT>(a, f.x), act1<ACT>(a, f.y));
    }
}
__device__ __forceinline__ uint32_t act_pack(int act, const ActF& a, uint32_t lo, uint32_t hi) {   // run-time form (experiments)
    ActF b = a; b.act = act;
    return act_pack<-1>(b, lo, hi);
}

} // namespace tc
} // namespace fvsrn

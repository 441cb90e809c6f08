// Shared device helpers of the row-streaming likelihood kernels (K4 nb_likelihood_fast.cu, K5
// vmf_likelihood_fast.cu): MUFU approximations, packed fp32 pairs, the cp.async.bulk row ring and the
// st.async cluster exchange.
#pragma once
#include <cstdint>

namespace mmv {
namespace rowstream {


__device__ __forceinline__ float
rcp_f(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

constexpr float kLog2e = 1.4426950408889634f, kLn2 = 0.6931471805599453f;

__device__ __forceinline__ float
ex2_f(float x)
{
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

__device__ __forceinline__ float
lg2_f(float x)
{
    float r;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}


// ---- packed fp32 pairs (sm_100 FFMA2 / FMUL2 / FADD2): the dense sweep is issue-bound, and every
//      thread owns adjacent gene pairs, so the FMA-pipe arithmetic is issued two lanes at a time
typedef unsigned long long f2_t;

__device__ __forceinline__ f2_t
pk2(float a, float b)
{
    f2_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}

__device__ __forceinline__ f2_t
bc2(float a)
{
    return pk2(a, a);
}

__device__ __forceinline__ void
up2(f2_t r, float &a, float &b)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(r));
}

__device__ __forceinline__ f2_t
fma2(f2_t a, f2_t b, f2_t c)
{
    f2_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}

__device__ __forceinline__ f2_t
mul2(f2_t a, f2_t b)
{
    f2_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

__device__ __forceinline__ f2_t
add2(f2_t a, f2_t b)
{
    f2_t r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}


__device__ __forceinline__ uint32_t
smem_addr(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void
bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

__device__ __forceinline__ void
bar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "NBW_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra NBW_DONE;\n"
        "bra NBW_LOOP;\n"
        "NBW_DONE:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}

// The same with a suspend-time hint (ns): the warp sleeps in hardware until the phase completes or the hint runs out,
// instead of re-issuing try_wait back to back.  For waits that are long by construction (the cluster hop of the previous
// row: ncu r02 step 4 counted 74 polls per wait, 18.5 % of K4's issued instructions, taken from the CTA sharing the SM).
__device__ __forceinline__ void
bar_wait_hint(uint32_t bar, uint32_t parity, uint32_t ns)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "NBH_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
        "@p bra NBH_DONE;\n"
        "bra NBH_LOOP;\n"
        "NBH_DONE:\n"
        "}\n" ::"r"(bar), "r"(parity), "r"(ns) : "memory");
}

// One thread publishes (a, b) into slot[my_rank] of every CTA of the cluster and signals that CTA's
// mbarrier with the 8 bytes written (st.async ... complete_tx): the data and the signal travel together.
__device__ __forceinline__ void
xchg_send(uint32_t slot_local, uint32_t bar_local, int cs, float a, float b)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_local), "r"(8u * (uint32_t)cs) : "memory");
    for (int r = 0; r < cs; ++r) {
        uint32_t rslot, rbar;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rslot) : "r"(slot_local), "r"(r));
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rbar) : "r"(bar_local), "r"(r));
        asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];"
                     ::"r"(rslot), "f"(a), "f"(b), "r"(rbar) : "memory");
    }
}


} // namespace rowstream

// CSR offsets at which each gene slice of a minibatch row begins: bounds[row * (cs + 1) + c]
// (defined in nb_likelihood_fast.cu; one thread per (row, boundary), bisection of the sorted gene ids)
int launch_slice_bounds(const int64_t *indptr, const int32_t *indices, const int32_t *cells, const int *unsorted, int B, int cs,
                        int Ds, int64_t *bounds, cudaStream_t stream);

} // namespace mmv

// common.cuh — shared device helpers for the B200 (sm_100a) MUL_MAT kernels.
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ggml_b200.h"

namespace b200 {

// ---- block formats: byte sizes only, the kernels address raw bytes (ggml/src/ggml-common.h:170-175, 218-238, 291-305, 322-328)
constexpr int QK      = 32;
constexpr int QK_K    = 256;
constexpr int SZ_Q4_0 = 18;
constexpr int SZ_Q8_0 = 34;
constexpr int SZ_Q8_1 = 36;
constexpr int SZ_Q4_K = 144;
constexpr int SZ_Q8_K = 292;
constexpr int SZ_Q5_K = 176;  // half d, dmin; uint8 scales[12]; uint8 qh[32]; uint8 qs[128] (ggml-common.h:312-323): 11 x 16 B
constexpr int SZ_Q6_K = 210;  // uint8 ql[128]; uint8 qh[64]; int8 scales[16]; half d (ggml-common.h:335-343): only 2-byte aligned

// ---- "unit": 256 consecutive weights of one row. 8 q4_0 blocks and one q4_K block are both 144 bytes = 9 x 16 B,
// 8 q8_0 blocks are 272 bytes = 17 x 16 B, so a unit is always a whole number of 128-bit words and a stride of
// 144 B / 272 B between lanes is bank-conflict free for 128-bit shared-memory loads (36 resp. 68 words == 4 mod 32).
constexpr int UNIT_ELEMS = 256;
// A q6_K unit is its 210-byte block: NOT a multiple of 16 - units of a row are only 2-byte aligned in shared memory (the dot product
// realigns them with funnel shifts), and a row slice is bulk-copyable only when it holds a multiple of 8 units (K % 2048 == 0).
__host__ __device__ constexpr int unit_bytes(int type) { return type == B200_TYPE_Q8_0 ? 272 : type == B200_TYPE_Q6_K ? 210 : type == B200_TYPE_Q5_K ? 176 : 144; }
__host__ __device__ constexpr bool type_is_k(int type) { return type == B200_TYPE_Q4_K || type == B200_TYPE_Q5_K || type == B200_TYPE_Q6_K; }
__host__ __device__ constexpr int type_block_elems(int type) { return type_is_k(type) ? 256 : 32; }
__host__ __device__ constexpr int type_block_bytes(int type) {
    return type == B200_TYPE_Q4_0 ? SZ_Q4_0 : type == B200_TYPE_Q8_0 ? SZ_Q8_0 : type == B200_TYPE_Q6_K ? SZ_Q6_K : type == B200_TYPE_Q5_K ? SZ_Q5_K : SZ_Q4_K;
}

// ---- activation unit record (kernel-internal layout, written by quantize.cu, read by mmvq.cu / mmq)
//   main (272 B): int8 q[256] ; then 16 B: kind 0 (q8_0-valued): half d[8]       (one per 32-block)
//                                          kind 1 (q8_K-valued): int16 bsum32[8] (sum of q per 32-block)
//   aux  (16 B):  kind 0: int16 sum[8] (sum of q per 32-block) ; kind 1: float d, 12 B zero
constexpr int ACT_MAIN = 272;
constexpr int ACT_AUX  = 16;
constexpr int ACT_KIND_Q8_0 = 0;
constexpr int ACT_KIND_Q8_K = 1;
__host__ __device__ constexpr int act_kind_for(int type) { return type_is_k(type) ? ACT_KIND_Q8_K : ACT_KIND_Q8_0; }

// programmatic dependent launch: the next kernel of the stream may be scheduled now (it parks in its own griddepcontrol.wait), and
// this one must not touch memory before the previous kernel has completed. No-ops when the launch did not carry the attribute.
__device__ __forceinline__ void pdl_prologue() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
// The canonical sum of squares of a row whose length is a multiple of 256 (RMS_NORM): per 256-element unit, lane l adds the squares
// of its 8 consecutive values by fmaf in order, the 32 lane sums go through an xor-shuffle tree (every lane gets the unit's sum), and
// the unit sums are added by a second xor-shuffle tree (lane u holds units u and u + 32). rms_norm_kernel (ops.cu) and the mat-vec kernels that normalize their activations themselves
// (mmvq.cu, MvNorm) both follow it, so a fused launch gives the bits of the stand-alone node.
__device__ __forceinline__ float unit_sumsq(const float (&v)[8]) {
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; i++) s = fmaf(v[i], v[i], s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    return s;
}
__device__ __forceinline__ float rms_scale(const float * unit_sums, int n_units, float eps) {   // whole warps call this (n_units <= 64)
    const int lane = threadIdx.x & 31;
    float tot = (lane < n_units ? unit_sums[lane] : 0.0f) + (lane + 32 < n_units ? unit_sums[lane + 32] : 0.0f);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    return 1.0f / sqrtf(tot / (float) (n_units * 256) + eps);
}
__device__ __forceinline__ unsigned long long global_ns_c() {   // %globaltimer, for the debug timelines (b200_ctx_trace)
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ float half_bits_to_float(uint16_t h) { return __half2float(__ushort_as_half(h)); }

__device__ __forceinline__ int dp4a_ss(int a, int b, int c) { return __dp4a(a, b, c); }
// unsigned 8-bit a (0..255) times signed 8-bit b
__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// two signed 16-bit halves of a times the low / high two unsigned bytes of b
__device__ __forceinline__ int dp2a_lo_su(int a, uint32_t b, int c) {
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp2a_hi_su(int a, uint32_t b, int c) {
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ uint2 lds64(uint32_t addr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
// same with a compile-time byte offset folded into the instruction (no address arithmetic per load)
template <int OFF>
__device__ __forceinline__ uint4 lds128o(uint32_t base) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4+%5];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(base), "n"(OFF));
    return v;
}
template <int OFF>
__device__ __forceinline__ uint32_t lds32o(uint32_t base) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(base), "n"(OFF));
    return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

// ---- mbarrier + bulk async copy (TMA 1-D, SASS UBLKCP) ----
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {}
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
// global -> shared bulk copy, completion reported as transaction bytes on `bar`. dst/src 16-B aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void * src, uint32_t bytes, uint32_t bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
        "l"(src), "r"(bytes), "r"(bar), "l"(policy)
        : "memory");
}
// asynchronous prefetch of [src, src + bytes) into L2 (no destination, no completion); src 16-B aligned, bytes % 16 == 0
__device__ __forceinline__ void bulk_prefetch_l2(const void * src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}

} // namespace b200

// quant_core.cuh — the arithmetic of the activation quantizer, shared by the stand-alone K1 kernels (quantize.cu) and the
// mat-vec kernel's fused prologue (mmvq.cu), so that both produce the same bits as the reference's
// quantize_row_q8_0_ref / _q8_1_ref / _q8_K_ref (/root/reference/ggml/src/ggml-quants.c:199-222, 225-258, 2555-2592).
//
// One warp quantizes one 256-element unit; each lane owns 8 consecutive floats, so a 32-element block is a 4-lane
// shuffle group and a q8_K super-block is the whole warp. Every floating-point step uses a round-to-nearest intrinsic
// (__fdiv_rn, __fmul_rn, __fadd_rn): neither -use_fast_math nor FMA contraction can change a bit.
#pragma once

#include "common.cuh"

namespace b200 {

__device__ __forceinline__ int nearest_int_rne(float v) {
    // round-half-even through the 1.5*2^23 magic add (ggml-quants.c:444-449)
    const float t = __fadd_rn(v, 12582912.f);
    return (__float_as_int(t) & 0x007fffff) - 0x00400000;
}

__device__ __forceinline__ uint32_t pack4(int a, int b, int c, int d) {
    return (uint32_t) (a & 0xff) | ((uint32_t) (b & 0xff) << 8) | ((uint32_t) (c & 0xff) << 16) | ((uint32_t) (d & 0xff) << 24);
}

// q8_0 / q8_1 values of this lane's 8 floats. All 32 lanes must call (4-lane shuffles).
//   d   : block scale amax/127 (fp32; stored as fp16)     sum : integer sum of the block's 32 quants
struct Q80Lane {
    uint32_t w0, w1;
    float    d;
    int      sum;
};
__device__ __forceinline__ Q80Lane quant_lane_q8_0(const float (&v)[8]) {
    float amax = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v[i]));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    int q[8];
#pragma unroll
    for (int i = 0; i < 8; i++) q[i] = (int) roundf(__fmul_rn(v[i], id));
    Q80Lane r;
    r.w0  = pack4(q[0], q[1], q[2], q[3]);
    r.w1  = pack4(q[4], q[5], q[6], q[7]);
    int sum = __dp4a((int) r.w0, 0x01010101, __dp4a((int) r.w1, 0x01010101, 0));  // sum of the 8 signed bytes
    sum += __shfl_xor_sync(0xffffffffu, sum, 1);
    sum += __shfl_xor_sync(0xffffffffu, sum, 2);
    r.d   = d;
    r.sum = sum;
    return r;
}

// q8_K values: the warp is one 256-element super-block.
//   d : 1/iscale (fp32)     s16 : sum of this lane pair's 16 quants (bsums entry)     s32 : sum over the 4-lane group
struct Q8KLane {
    uint32_t w0, w1;
    float    d;
    int      s16, s32;
};
__device__ __forceinline__ Q8KLane quant_lane_q8_K(const float (&v)[8], int lane) {
    // max |v| of the super-block and the FIRST element that attains it (ggml-quants.c:2564-2569: `if (ax > amax)` keeps the
    // earliest). |v| >= 0, so the fp32 bit patterns order like unsigned integers and one REDUX gives the warp maximum; a second
    // one gives the smallest index among the elements equal to it.
    float amax_l = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; i++) amax_l = fmaxf(amax_l, fabsf(v[i]));
    const uint32_t amax_b = __reduce_max_sync(0xffffffffu, __float_as_uint(amax_l));
    uint32_t idx_l = 0xffffffffu;
    float    cand  = 0.0f;
#pragma unroll
    for (int i = 7; i >= 0; i--) {
        if (__float_as_uint(fabsf(v[i])) == amax_b) { idx_l = (uint32_t) (lane * 8 + i); cand = v[i]; }
    }
    const uint32_t idx  = __reduce_min_sync(0xffffffffu, idx_l);
    const float    vmax = __shfl_sync(0xffffffffu, cand, (int) (idx >> 3) & 31);
    int   q[8];
    float d = 0.0f;
    if (amax_b != 0u) {
        const float iscale = __fdiv_rn(-127.f, vmax);
#pragma unroll
        for (int i = 0; i < 8; i++) q[i] = min(127, nearest_int_rne(__fmul_rn(iscale, v[i])));
        d = __fdiv_rn(1.0f, iscale);
    } else {
#pragma unroll
        for (int i = 0; i < 8; i++) q[i] = 0;
    }
    Q8KLane r;
    r.w0 = pack4(q[0], q[1], q[2], q[3]);
    r.w1 = pack4(q[4], q[5], q[6], q[7]);
    const int s8 = __dp4a((int) r.w0, 0x01010101, __dp4a((int) r.w1, 0x01010101, 0));  // sum of the 8 signed bytes
    r.s16 = s8 + __shfl_xor_sync(0xffffffffu, s8, 1);
    r.s32 = r.s16 + __shfl_xor_sync(0xffffffffu, r.s16, 2);
    r.d   = d;
    return r;
}

// writes this lane's share of one unit record (common.cuh layout) to generic memory (global or shared)
template <int KIND>
__device__ __forceinline__ void quant_unit_store(const float (&v)[8], int lane, uint8_t * m, uint8_t * x) {
    if (KIND == ACT_KIND_Q8_0) {
        const Q80Lane r = quant_lane_q8_0(v);
        *(uint2 *) (m + lane * 8) = make_uint2(r.w0, r.w1);
        if ((lane & 3) == 0) {
            *(uint16_t *) (m + 256 + 2 * (lane >> 2)) = __half_as_ushort(__float2half_rn(r.d));
            *(int16_t *) (x + 2 * (lane >> 2))        = (int16_t) r.sum;
        }
    } else {
        const Q8KLane r = quant_lane_q8_K(v, lane);
        *(uint2 *) (m + lane * 8) = make_uint2(r.w0, r.w1);
        if ((lane & 3) == 0) *(int16_t *) (m + 256 + 2 * (lane >> 2)) = (int16_t) r.s32;
        if (lane == 0) *(float4 *) x = make_float4(r.d, 0.f, 0.f, 0.f);
    }
}

} // namespace b200

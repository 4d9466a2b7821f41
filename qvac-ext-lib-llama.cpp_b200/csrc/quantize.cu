// quantize.cu — K1: F32 activation rows -> q8_0 / q8_1 / q8_K valued blocks, bit-exact to the reference's
// quantize_row_q8_0_ref / _q8_1_ref / _q8_K_ref (/root/reference/ggml/src/ggml-quants.c:199-222, 225-258, 2555-2592).
// Replaces quantize_q8_1 / quantize_mmq_q8_1 (/root/reference/ggml/src/ggml-cuda/quantize.cu:4-48, 50-147), which are
// NOT bit-exact to the CPU reference (they store s = sum(x) and divide by d instead of multiplying by 1/d).
//
// One warp quantizes one 256-element unit; each lane owns 8 consecutive floats (two 128-bit loads), so a 32-element
// block is a 4-lane shuffle group and a q8_K super-block is the whole warp. All floating-point steps use the
// round-to-nearest intrinsics (__fdiv_rn, __fmul_rn, __fadd_rn) so that neither -use_fast_math nor FMA contraction can
// change a bit. Two output layouts from the same arithmetic:
//   LAYOUT_GGML  — the reference's block structs (what the C-ABI quantizers return and the parity tests memcmp)
//   LAYOUT_UNITS — the kernels' internal unit records (common.cuh), 16-byte aligned for 128-bit shared loads.

#include "quant_core.cuh"
#include "kernels.h"

namespace b200 {

enum { VAL_Q8_0 = 0, VAL_Q8_K = 1, VAL_Q8_1 = 2 };
enum { LAYOUT_GGML = 0, LAYOUT_UNITS = 1 };

template <int VAL, int LAYOUT>
__global__ void __launch_bounds__(128) quantize_rows_kernel(const float * __restrict__ x, int64_t k, int64_t nrows, size_t x_row_stride,
                                                            uint8_t * __restrict__ y_main, uint8_t * __restrict__ y_aux, int vec_ok) {
    // programmatic dependent launch (no-ops without the launch attribute): let the consumer of the records start its weight
    // prologue now; wait for the producer of x - and for the previous reader of the record scratch - before touching memory
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int     lane  = threadIdx.x & 31;
    const int64_t unit  = (int64_t) blockIdx.x * 4 + (threadIdx.x >> 5);
    const int64_t row   = blockIdx.y;
    const int64_t units = (k + UNIT_ELEMS - 1) / UNIT_ELEMS;
    if (unit >= units || row >= nrows) {
        return;
    }
    const int64_t e0     = unit * UNIT_ELEMS + lane * 8;
    const bool    active = e0 < k;  // k % 32 == 0, so a lane is either fully inside or fully outside the row

    float v[8];
    const float * xr = (const float *) ((const char *) x + row * x_row_stride) + e0;
    if (active) {
        if (vec_ok) {
            const float4 a = __ldg((const float4 *) xr);
            const float4 b = __ldg((const float4 *) xr + 1);
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
            v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = __ldg(xr + i);
        }
    } else {
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = 0.0f;
    }

    if (VAL == VAL_Q8_0 || VAL == VAL_Q8_1) {
        const Q80Lane  r  = quant_lane_q8_0(v);
        const uint16_t dh = __half_as_ushort(__float2half_rn(r.d));
        if (!active) {
            return;
        }
        const int blk = lane >> 2, sub = lane & 3;
        if (LAYOUT == LAYOUT_UNITS) {
            uint8_t * m = y_main + (row * units + unit) * ACT_MAIN;
            *(uint2 *) (m + lane * 8) = make_uint2(r.w0, r.w1);
            if (sub == 0) {
                *(uint16_t *) (m + 256 + 2 * blk) = dh;
                *(int16_t *) (y_aux + (row * units + unit) * ACT_AUX + 2 * blk) = (int16_t) r.sum;
            }
        } else if (VAL == VAL_Q8_0) {
            uint8_t * b = y_main + row * (size_t) (k / QK * SZ_Q8_0) + (unit * 8 + blk) * SZ_Q8_0;
            if (sub == 0) *(uint16_t *) b = dh;
            uint16_t * qs = (uint16_t *) (b + 2 + sub * 8);  // only 2-byte aligned
            qs[0] = (uint16_t) r.w0; qs[1] = (uint16_t) (r.w0 >> 16);
            qs[2] = (uint16_t) r.w1; qs[3] = (uint16_t) (r.w1 >> 16);
        } else {
            uint8_t * b = y_main + row * (size_t) (k / QK * SZ_Q8_1) + (unit * 8 + blk) * SZ_Q8_1;
            if (sub == 0) {
                const uint16_t sh = __half_as_ushort(__float2half_rn(__fmul_rn((float) r.sum, r.d)));
                *(uint32_t *) b = (uint32_t) dh | ((uint32_t) sh << 16);
            }
            uint32_t * qs = (uint32_t *) (b + 4 + sub * 8);
            qs[0] = r.w0; qs[1] = r.w1;
        }
    } else {  // VAL_Q8_K: the warp is one 256-element super-block (k % 256 == 0 checked on the host)
        const Q8KLane r = quant_lane_q8_K(v, lane);
        if (LAYOUT == LAYOUT_UNITS) {
            uint8_t * m = y_main + (row * units + unit) * ACT_MAIN;
            *(uint2 *) (m + lane * 8) = make_uint2(r.w0, r.w1);
            if ((lane & 3) == 0) *(int16_t *) (m + 256 + 2 * (lane >> 2)) = (int16_t) r.s32;
            if (lane == 0) *(float4 *) (y_aux + (row * units + unit) * ACT_AUX) = make_float4(r.d, 0.f, 0.f, 0.f);
        } else {
            uint8_t * b = y_main + row * (size_t) (k / QK_K * SZ_Q8_K) + unit * SZ_Q8_K;
            if (lane == 0) *(float *) b = r.d;
            uint32_t * qs = (uint32_t *) (b + 4 + lane * 8);
            qs[0] = r.w0; qs[1] = r.w1;
            if ((lane & 1) == 0) *(int16_t *) (b + 4 + QK_K + 2 * (lane >> 1)) = (int16_t) r.s16;
        }
    }
}

template <int VAL, int LAYOUT>
static cudaError_t launch(const float * x, int64_t k, int64_t nrows, size_t stride, void * y_main, void * y_aux, cudaStream_t st, int pdl = 0) {
    if (nrows == 0 || k == 0) return cudaSuccess;
    const int64_t units  = (k + UNIT_ELEMS - 1) / UNIT_ELEMS;
    const int     vec_ok = ((uintptr_t) x % 16 == 0) && (stride % 16 == 0);
    // grid.y is limited to 65535 rows per launch
    for (int64_t r0 = 0; r0 < nrows; r0 += 65535) {
        const int64_t nr = nrows - r0 < 65535 ? nrows - r0 : 65535;
        dim3 grid((unsigned) ((units + 3) / 4), (unsigned) nr);
        size_t main_row = 0, aux_row = 0;
        if (LAYOUT == LAYOUT_UNITS) { main_row = units * ACT_MAIN; aux_row = units * ACT_AUX; }
        else if (VAL == VAL_Q8_0)   { main_row = k / QK * SZ_Q8_0; }
        else if (VAL == VAL_Q8_1)   { main_row = k / QK * SZ_Q8_1; }
        else                        { main_row = k / QK_K * SZ_Q8_K; }
        cudaLaunchConfig_t  cfg = {};
        cudaLaunchAttribute attr[1];
        attr[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.gridDim  = grid;
        cfg.blockDim = dim3(128);
        cfg.stream   = st;
        cfg.attrs    = attr;
        cfg.numAttrs = pdl ? 1 : 0;
        const cudaError_t e = cudaLaunchKernelEx(&cfg, quantize_rows_kernel<VAL, LAYOUT>, (const float *) ((const char *) x + r0 * stride), k,
                                                 nr, stride, (uint8_t *) y_main + r0 * main_row,
                                                 y_aux ? (uint8_t *) y_aux + r0 * aux_row : (uint8_t *) nullptr, vec_ok);
        if (e != cudaSuccess) return e;
    }
    return cudaGetLastError();
}

cudaError_t quantize_ggml(int val, const float * x, int64_t k, int64_t nrows, size_t stride, void * y, cudaStream_t st, int * launches) {
    const int n = (int) ((nrows + 65534) / 65535);
    if (launches) *launches += n;
    switch (val) {
        case VAL_Q8_0: return launch<VAL_Q8_0, LAYOUT_GGML>(x, k, nrows, stride, y, nullptr, st);
        case VAL_Q8_1: return launch<VAL_Q8_1, LAYOUT_GGML>(x, k, nrows, stride, y, nullptr, st);
        case VAL_Q8_K: return launch<VAL_Q8_K, LAYOUT_GGML>(x, k, nrows, stride, y, nullptr, st);
    }
    return cudaErrorInvalidValue;
}

cudaError_t quantize_units(int kind, const float * x, int64_t k, int64_t nrows, size_t stride, void * main, void * aux, cudaStream_t st,
                           int * launches, int pdl) {
    const int n = (int) ((nrows + 65534) / 65535);
    if (launches) *launches += n;
    if (kind == ACT_KIND_Q8_0) return launch<VAL_Q8_0, LAYOUT_UNITS>(x, k, nrows, stride, main, aux, st, pdl);
    return launch<VAL_Q8_K, LAYOUT_UNITS>(x, k, nrows, stride, main, aux, st, pdl);
}

} // namespace b200

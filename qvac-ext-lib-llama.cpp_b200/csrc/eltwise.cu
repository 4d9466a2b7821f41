// K7 companions: the element-wise nodes around a MUL_MAT that the backend has to own before it can fuse
// them - GGML_OP_GLU, GGML_OP_ADD (bias / residual) and GGML_OP_ADD_ID (per-expert bias) on F32. They run stand-alone whenever
// the fused mat-vec epilogue does not apply (more than one column, a layout the TMA-staged kernel does not take, a GLU whose
// inputs are not two MUL_MATs). HBM-trivial: 12 bytes per element, one thread per element, consecutive threads on consecutive
// floats. Semantics: ggml-cpu/ops.cpp (GLU: 2934-3290; ADD_ID: 701-750) and binary-ops.cpp (ADD with ggml broadcasting: every
// src1 dimension divides the src0 dimension, index = i % ne1x).
#include "internal.h"
#include "epilogue.cuh"

namespace b200 {

__global__ void __launch_bounds__(256) glu_kernel(int op, float alpha, float limit, const float * __restrict__ x, int64_t x_rs,
                                                  const float * __restrict__ g, int64_t g_rs, float * __restrict__ dst, int64_t d_rs, int64_t nc,
                                                  int64_t total) {
    pdl_prologue();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t r = i / nc, c = i - r * nc;
        dst[r * d_rs + c] = glu_apply(op, x[r * x_rs + c], g[r * g_rs + c], alpha, limit);
    }
}

// the shapes a prompt batch produces in bulk: all three tensors contiguous and 16-byte aligned, src1 either the same shape (residual
// add) or one row repeated (norm weight, bias): float4 per thread, no index arithmetic (the general kernel divides 64-bit indices per element)
template <bool MUL>
__global__ void __launch_bounds__(256) add_flat_kernel(const float4 * a, const float4 * b, float4 * dst, int64_t n4, int64_t period4) {
    pdl_prologue();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t) gridDim.x * blockDim.x) {
        const float4 x = a[i], y = b[period4 ? i % period4 : i];
        dst[i] = MUL ? make_float4(x.x * y.x, x.y * y.y, x.z * y.z, x.w * y.w) : make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
    }
}
__global__ void __launch_bounds__(256) glu_flat_kernel(int op, float alpha, float limit, const float4 * x, const float4 * g, float4 * dst, int64_t n4) {
    pdl_prologue();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t) gridDim.x * blockDim.x) {
        const float4 a = x[i], b = g[i];
        dst[i] = make_float4(glu_apply(op, a.x, b.x, alpha, limit), glu_apply(op, a.y, b.y, alpha, limit), glu_apply(op, a.z, b.z, alpha, limit),
                             glu_apply(op, a.w, b.w, alpha, limit));
    }
}

struct AddParams {
    int64_t ne[4];                // dst / src0 extents
    int64_t ne1[4];               // src1 extents (each divides ne[i])
    int64_t s0[4], s1[4], sd[4];  // strides in floats; dimension 0 is contiguous (stride 1)
};

// no __restrict__: ggml_add_inplace makes dst alias src0 (element for element, which this kernel tolerates)
template <bool MUL>
__global__ void __launch_bounds__(256) add_kernel(const AddParams p, const float * a, const float * b, float * dst, int64_t total) {
    pdl_prologue();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t r = i;
        const int64_t i0 = r % p.ne[0]; r /= p.ne[0];
        const int64_t i1 = r % p.ne[1]; r /= p.ne[1];
        const int64_t i2 = r % p.ne[2];
        const int64_t i3 = r / p.ne[2];
        const float va = a[i0 + i1 * p.s0[1] + i2 * p.s0[2] + i3 * p.s0[3]];
        const float vb = b[i0 % p.ne1[0] + (i1 % p.ne1[1]) * p.s1[1] + (i2 % p.ne1[2]) * p.s1[2] + (i3 % p.ne1[3]) * p.s1[3]];
        dst[i0 + i1 * p.sd[1] + i2 * p.sd[2] + i3 * p.sd[3]] = MUL ? va * vb : va + vb;
    }
}

// dst[:, i, t] = a[:, i, t] + bias[:, ids[i, t]]
__global__ void __launch_bounds__(256) add_id_kernel(const float * a, int64_t a_s1, int64_t a_s2, const float * __restrict__ bias, int64_t b_s1,
                                                     int64_t n_as, const int32_t * __restrict__ ids, int64_t ids_s1, float * dst, int64_t d_s1,
                                                     int64_t d_s2, int64_t M, int64_t n_used, int64_t total) {
    pdl_prologue();
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t r = i;
        const int64_t m = r % M; r /= M;
        const int64_t s = r % n_used;
        const int64_t t = r / n_used;
        const int e = ids[t * ids_s1 + s];
        const float vb = (e >= 0 && e < n_as) ? bias[(int64_t) e * b_s1 + m] : 0.0f;  // ggml-cpu asserts on a bad id (ops.cpp:743); never read out of bounds
        dst[m + s * d_s1 + t * d_s2] = a[m + s * a_s1 + t * a_s2] + vb;
    }
}

static int grid_for(const b200_ctx * ctx, int64_t total) {
    const int64_t blocks = (total + 255) / 256;
    const int64_t cap    = (int64_t) ctx->sm_count * 8;
    return (int) (blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

}  // namespace b200

using namespace b200;

#define FAIL_CUDA(e) b200_fail(B200_ERR_CUDA, "%s:%d: %s", __FILE__, __LINE__, cudaGetErrorString(e))

b200_status b200_glu(b200_ctx * ctx, int glu_op, float alpha, float limit, const float * gate, size_t gate_nb1, const float * up, size_t up_nb1,
                     float * dst, size_t dst_nb1, int64_t nc, int64_t nrows) {
    if (!ctx) return b200_fail(B200_ERR_ARG, "null context");
    if (glu_op < 0 || glu_op >= GLU_COUNT) return b200_fail(B200_ERR_UNSUPPORTED, "GLU op %d", glu_op);
    if (nc < 0 || nrows < 0 || gate_nb1 % 4 || up_nb1 % 4 || dst_nb1 % 4) return b200_fail(B200_ERR_ARG, "GLU geometry");
    if (nc == 0 || nrows == 0) return B200_OK;
    if (!gate || !up || !dst) return b200_fail(B200_ERR_ARG, "null argument");
    cudaError_t e = cudaSetDevice(ctx->device);
    if (e != cudaSuccess) return FAIL_CUDA(e);
    const int64_t total = nc * nrows;
    if (b200_status fst = b200_split_flush(ctx)) return fst;  // the inputs may be a dst whose remote slices are still in flight
    if (nc % 4 == 0 && gate_nb1 == (size_t) nc * 4 && up_nb1 == (size_t) nc * 4 && dst_nb1 == (size_t) nc * 4 &&
        ((uintptr_t) gate | (uintptr_t) up | (uintptr_t) dst) % 16 == 0) {
        e = b200_launch(ctx, glu_flat_kernel, dim3(grid_for(ctx, total / 4)), dim3(256), 0, glu_op, alpha, limit, (const float4 *) gate, (const float4 *) up,
                        (float4 *) dst, total / 4);
        if (e != cudaSuccess) return FAIL_CUDA(e);
        ctx->launches++;
        return B200_OK;
    }
    e = b200_launch(ctx, glu_kernel, dim3(grid_for(ctx, total)), dim3(256), 0, glu_op, alpha, limit, gate, (int64_t) (gate_nb1 / 4), up, (int64_t) (up_nb1 / 4), dst,
                    (int64_t) (dst_nb1 / 4), nc, total);
    if (e != cudaSuccess) return FAIL_CUDA(e);
    ctx->launches++;
    return B200_OK;
}

static b200_status binary_bcast(b200_ctx * ctx, bool mul, const int64_t * ne, const float * a, const size_t * nb_a, const float * b, const int64_t * ne_b,
                                const size_t * nb_b, float * dst, const size_t * nb_d) {
    if (!ctx || !ne || !nb_a || !ne_b || !nb_b || !nb_d) return b200_fail(B200_ERR_ARG, "null argument");
    AddParams p;
    int64_t total = 1;
    for (int i = 0; i < 4; i++) {
        if (ne[i] < 0 || ne_b[i] <= 0 || ne[i] % ne_b[i] != 0) return b200_fail(B200_ERR_UNSUPPORTED, "ADD / MUL: src1 does not broadcast onto src0");
        if (nb_a[i] % 4 || nb_b[i] % 4 || nb_d[i] % 4) return b200_fail(B200_ERR_ARG, "ADD / MUL strides");
        p.ne[i] = ne[i]; p.ne1[i] = ne_b[i];
        p.s0[i] = (int64_t) (nb_a[i] / 4); p.s1[i] = (int64_t) (nb_b[i] / 4); p.sd[i] = (int64_t) (nb_d[i] / 4);
        total *= ne[i];
    }
    if (nb_a[0] != 4 || nb_b[0] != 4 || nb_d[0] != 4) return b200_fail(B200_ERR_UNSUPPORTED, "ADD / MUL: dimension 0 must be contiguous F32");
    if (total == 0) return B200_OK;
    if (!a || !b || !dst) return b200_fail(B200_ERR_ARG, "null argument");
    cudaError_t e = cudaSetDevice(ctx->device);
    if (e != cudaSuccess) return FAIL_CUDA(e);
    if (b200_status fst = b200_split_flush(ctx)) return fst;
    {
        auto contiguous = [](const int64_t * n, const size_t * nb) {
            size_t want = 4;
            for (int i = 0; i < 4; i++) { if (n[i] != 1 && nb[i] != want) return false; want *= (size_t) n[i]; }
            return true;
        };
        const bool same = ne_b[0] == ne[0] && ne_b[1] == ne[1] && ne_b[2] == ne[2] && ne_b[3] == ne[3];
        const bool row  = ne_b[0] == ne[0] && ne_b[1] == 1 && ne_b[2] == 1 && ne_b[3] == 1;
        if ((same || row) && ne[0] % 4 == 0 && total >= 4096 && contiguous(ne, nb_a) && contiguous(ne, nb_d) && contiguous(ne_b, nb_b) &&
            ((uintptr_t) a | (uintptr_t) b | (uintptr_t) dst) % 16 == 0) {
            const int64_t n4 = total / 4, per = same ? 0 : ne[0] / 4;
            e = mul ? b200_launch(ctx, add_flat_kernel<true>, dim3(grid_for(ctx, n4)), dim3(256), 0, (const float4 *) a, (const float4 *) b, (float4 *) dst, n4, per)
                    : b200_launch(ctx, add_flat_kernel<false>, dim3(grid_for(ctx, n4)), dim3(256), 0, (const float4 *) a, (const float4 *) b, (float4 *) dst, n4, per);
            if (e != cudaSuccess) return FAIL_CUDA(e);
            ctx->launches++;
            return B200_OK;
        }
    }
    e = mul ? b200_launch(ctx, add_kernel<true>, dim3(grid_for(ctx, total)), dim3(256), 0, p, a, b, dst, total)
            : b200_launch(ctx, add_kernel<false>, dim3(grid_for(ctx, total)), dim3(256), 0, p, a, b, dst, total);
    if (e != cudaSuccess) return FAIL_CUDA(e);
    ctx->launches++;
    return B200_OK;
}

b200_status b200_add(b200_ctx * ctx, const int64_t * ne, const float * a, const size_t * nb_a, const float * b, const int64_t * ne_b,
                     const size_t * nb_b, float * dst, const size_t * nb_d) {
    return binary_bcast(ctx, false, ne, a, nb_a, b, ne_b, nb_b, dst, nb_d);
}

b200_status b200_mul(b200_ctx * ctx, const int64_t * ne, const float * a, const size_t * nb_a, const float * b, const int64_t * ne_b,
                     const size_t * nb_b, float * dst, const size_t * nb_d) {
    return binary_bcast(ctx, true, ne, a, nb_a, b, ne_b, nb_b, dst, nb_d);
}

b200_status b200_add_id(b200_ctx * ctx, const float * a, size_t a_nb1, size_t a_nb2, const float * bias, size_t bias_nb1, int64_t n_as,
                        const int32_t * ids, size_t ids_nb1, float * dst, size_t nb1, size_t nb2, int64_t M, int64_t n_used, int64_t n_tok) {
    if (!ctx) return b200_fail(B200_ERR_ARG, "null context");
    if (M < 0 || n_used < 0 || n_tok < 0 || n_as <= 0 || a_nb1 % 4 || a_nb2 % 4 || bias_nb1 % 4 || ids_nb1 % 4 || nb1 % 4 || nb2 % 4)
        return b200_fail(B200_ERR_ARG, "ADD_ID geometry");
    const int64_t total = M * n_used * n_tok;
    if (total == 0) return B200_OK;
    if (!a || !bias || !ids || !dst) return b200_fail(B200_ERR_ARG, "null argument");
    cudaError_t e = cudaSetDevice(ctx->device);
    if (e != cudaSuccess) return FAIL_CUDA(e);
    if (b200_status fst = b200_split_flush(ctx)) return fst;
    e = b200_launch(ctx, add_id_kernel, dim3(grid_for(ctx, total)), dim3(256), 0, a, (int64_t) (a_nb1 / 4), (int64_t) (a_nb2 / 4), bias, (int64_t) (bias_nb1 / 4), n_as,
                    ids, (int64_t) (ids_nb1 / 4), dst, (int64_t) (nb1 / 4), (int64_t) (nb2 / 4), M, n_used, total);
    if (e != cudaSuccess) return FAIL_CUDA(e);
    ctx->launches++;
    return B200_OK;
}

// ops.cu — SURVEY 8(f) rank 3, first slice: the ops BETWEEN the matmuls of a Llama decode graph, so that the graph stays on the
// device (no scheduler split per layer): GET_ROWS, RMS_NORM, ROPE, SET_ROWS, CPY / CONT / DUP and a FLASH_ATTN_EXT for F16 K/V.
// Semantics are ggml-cpu's (/root/reference/ggml/src/ggml-cpu/ops.cpp: rms_norm :3646-3695, rope :5525-5780, set_rows :4828-4876,
// get_rows :4540-4780, flash_attn_ext :7860-8140, dup/cpy :20-700); what the reference's CUDA backend does in norm.cu, rope.cu,
// set-rows.cu, getrows.cu, cpy.cu, fattn*.cu. These are the small ops of a decode step (a few KB each): plain, coalesced,
// one-pass kernels - the bytes of a token live in the weight matrices, which is where the tuned kernels are.

#include "internal.h"

#include <algorithm>

using namespace b200;
#define fail b200_fail

namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

struct T4 {   // a ggml tensor's geometry: elements and byte strides
    int64_t ne[4];
    int64_t nb[4];
};
T4 t4(const b200_tensor * t) {
    T4 r;
    for (int i = 0; i < 4; i++) { r.ne[i] = t->ne[i]; r.nb[i] = (int64_t) t->nb[i]; }
    return r;
}

// ------------------------------------------------------------------------------------------------ RMS_NORM (+ MUL)
// y = x / sqrt(mean(x^2) + eps) [* w], one CTA per row (ggml-cpu sums in double; fp32 tree sums here: NMSE ~1e-13 apart). The MUL that
// follows every norm of a transformer graph (the norm weight, broadcast over the rows: build_norm, src/llama-graph.cpp) rides along:
// (x * scale) * w with both products rounded, bit for bit what the two nodes give one after the other.
struct NormP {
    T4 s, d, w;
    const char * x, * wd;   // wd == nullptr: no MUL
    char * y;
    float eps;
};
__device__ __forceinline__ float block_sum(float v, float * part) {   // all threads get the sum; blockDim a multiple of 32, <= 1024
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = v;
    __syncthreads();
    const int nw = blockDim.x >> 5;
    float t = (int) (threadIdx.x & 31) < nw ? part[threadIdx.x & 31] : 0.0f;
    return warp_sum(t);
}
// rows of 256 .. 16384 elements, a multiple of 256, 16-byte aligned (every transformer width): the canonical summation order of
// common.cuh unit_sumsq / rms_scale, which the mat-vec kernels reproduce when they normalize their own activations (mmvq.cu MvNorm).
// blockDim = 32 * min(32, units): a warp holds at most two units.
__global__ void __launch_bounds__(1024) rms_norm_units_kernel(const NormP p) {
    __shared__ float part[64];
    pdl_prologue();
    const int64_t row = blockIdx.x;
    const int64_t i1 = row % p.s.ne[1], i2 = (row / p.s.ne[1]) % p.s.ne[2], i3 = row / (p.s.ne[1] * p.s.ne[2]);
    const char * xr = p.x + i1 * p.s.nb[1] + i2 * p.s.nb[2] + i3 * p.s.nb[3];
    char *       yr = p.y + i1 * p.d.nb[1] + i2 * p.d.nb[2] + i3 * p.d.nb[3];
    const char * wr = p.wd ? p.wd + (i1 % p.w.ne[1]) * p.w.nb[1] + (i2 % p.w.ne[2]) * p.w.nb[2] + (i3 % p.w.ne[3]) * p.w.nb[3] : nullptr;
    const int U = (int) (p.s.ne[0] >> 8), nw = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float v[2][8], wv[2][8];
#pragma unroll
    for (int b = 0; b < 2; b++) {
        const int u = warp + b * nw;
        if (u < U) {
            const float4 a = ((const float4 *) xr)[u * 64 + lane * 2], c = ((const float4 *) xr)[u * 64 + lane * 2 + 1];
            v[b][0] = a.x; v[b][1] = a.y; v[b][2] = a.z; v[b][3] = a.w; v[b][4] = c.x; v[b][5] = c.y; v[b][6] = c.z; v[b][7] = c.w;
            if (wr) {
                const float4 e = ((const float4 *) wr)[u * 64 + lane * 2], f = ((const float4 *) wr)[u * 64 + lane * 2 + 1];
                wv[b][0] = e.x; wv[b][1] = e.y; wv[b][2] = e.z; wv[b][3] = e.w; wv[b][4] = f.x; wv[b][5] = f.y; wv[b][6] = f.z; wv[b][7] = f.w;
            }
            const float sq = unit_sumsq(v[b]);
            if (lane == 0) part[u] = sq;
        }
    }
    __syncthreads();
    const float scale = rms_scale(part, U, p.eps);
#pragma unroll
    for (int b = 0; b < 2; b++) {
        const int u = warp + b * nw;
        if (u < U) {
            float o[8];
#pragma unroll
            for (int i = 0; i < 8; i++) { o[i] = v[b][i] * scale; if (wr) o[i] *= wv[b][i]; }
            ((float4 *) yr)[u * 64 + lane * 2]     = make_float4(o[0], o[1], o[2], o[3]);
            ((float4 *) yr)[u * 64 + lane * 2 + 1] = make_float4(o[4], o[5], o[6], o[7]);
        }
    }
}

template <bool VEC>   // VEC: rows are 16-byte aligned, ne0 % 4 == 0 and a row fits the CTA's registers (4 float4 per thread)
__global__ void __launch_bounds__(1024) rms_norm_kernel(const NormP p) {
    __shared__ float part[32];
    pdl_prologue();
    const int64_t row = blockIdx.x;
    const int64_t i1 = row % p.s.ne[1], i2 = (row / p.s.ne[1]) % p.s.ne[2], i3 = row / (p.s.ne[1] * p.s.ne[2]);
    const char * xr = p.x + i1 * p.s.nb[1] + i2 * p.s.nb[2] + i3 * p.s.nb[3];
    char *       yr = p.y + i1 * p.d.nb[1] + i2 * p.d.nb[2] + i3 * p.d.nb[3];
    const char * wr = p.wd ? p.wd + (i1 % p.w.ne[1]) * p.w.nb[1] + (i2 % p.w.ne[2]) * p.w.nb[2] + (i3 % p.w.ne[3]) * p.w.nb[3] : nullptr;
    const int64_t n = p.s.ne[0];
    if constexpr (VEC) {
        const int64_t n4 = n >> 2;
        float4 v[4];
        float sum = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int64_t i = threadIdx.x + (int64_t) j * blockDim.x;
            if (i < n4) { v[j] = ((const float4 *) xr)[i]; sum += v[j].x * v[j].x + v[j].y * v[j].y + v[j].z * v[j].z + v[j].w * v[j].w; }
        }
        const float scale = 1.0f / sqrtf(block_sum(sum, part) / (float) n + p.eps);
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int64_t i = threadIdx.x + (int64_t) j * blockDim.x;
            if (i < n4) {
                float4 o = make_float4(v[j].x * scale, v[j].y * scale, v[j].z * scale, v[j].w * scale);
                if (wr) { const float4 w = ((const float4 *) wr)[i]; o.x *= w.x; o.y *= w.y; o.z *= w.z; o.w *= w.w; }
                ((float4 *) yr)[i] = o;
            }
        }
    } else {
        float sum = 0.0f;
        for (int64_t i = threadIdx.x; i < n; i += blockDim.x) { const float v = ((const float *) xr)[i]; sum += v * v; }
        const float scale = 1.0f / sqrtf(block_sum(sum, part) / (float) n + p.eps);
        for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
            float o = ((const float *) xr)[i] * scale;
            if (wr) o *= ((const float *) wr)[i % p.w.ne[0]];
            ((float *) yr)[i] = o;
        }
    }
}

// ------------------------------------------------------------------------------------------------ dequantization of one element
// the arithmetic of dequantize_row_<type> (ggml-quants.c:307-325, 400-420, 1352-1373, 1536-1560, 1762-1791), element by element
__device__ __forceinline__ void k4_scale_min(int j, const uint8_t * q, int & sc, int & m) {
    if (j < 4) { sc = q[j] & 63; m = q[j + 4] & 63; }
    else { sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4); m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}
__device__ __forceinline__ float dequant_at(int type, const uint8_t * row, int64_t k) {
    switch (type) {
        case B200_TYPE_Q4_0: {
            const uint8_t * b = row + (k >> 5) * SZ_Q4_0;
            const int j = (int) (k & 31);
            const int q = j < 16 ? (b[2 + j] & 0xF) : (b[2 + j - 16] >> 4);
            return (float) (q - 8) * half_bits_to_float(*(const uint16_t *) b);
        }
        case B200_TYPE_Q8_0: {
            const uint8_t * b = row + (k >> 5) * SZ_Q8_0;
            return (float) ((const int8_t *) (b + 2))[k & 31] * half_bits_to_float(*(const uint16_t *) b);
        }
        case B200_TYPE_Q4_K:
        case B200_TYPE_Q5_K: {
            const bool q5 = type == B200_TYPE_Q5_K;
            const uint8_t * b = row + (k >> 8) * (q5 ? SZ_Q5_K : SZ_Q4_K);
            const int e = (int) (k & 255), j = e >> 5, l = e & 31;
            int sc, m;
            k4_scale_min(j, b + 4, sc, m);
            const float d = half_bits_to_float(*(const uint16_t *) b), dmin = half_bits_to_float(*(const uint16_t *) (b + 2));
            const uint8_t qb = b[(q5 ? 48 : 16) + 32 * (j >> 1) + l];
            int q = (j & 1) ? (qb >> 4) : (qb & 0xF);
            if (q5) q += ((b[16 + l] >> j) & 1) ? 16 : 0;
            // d1 = d * sc; m1 = dmin * m; y = d1 * q - m1, each rounded (no contraction: ggml-cpu's x86-64 baseline build has no FMA)
            return __fsub_rn(__fmul_rn(__fmul_rn(d, (float) sc), (float) q), __fmul_rn(dmin, (float) m));
        }
        case B200_TYPE_Q6_K: {
            const uint8_t * b = row + (k >> 8) * SZ_Q6_K;
            const int e = (int) (k & 255), h = e >> 7, r = e & 127, quarter = r >> 5, l = r & 31;
            const uint8_t * ql = b + 64 * h, * qh = b + 128 + 32 * h;
            const int8_t *  sc = (const int8_t *) (b + 192);
            const int lo = quarter == 0 ? (ql[l] & 0xF) : quarter == 1 ? (ql[l + 32] & 0xF) : quarter == 2 ? (ql[l] >> 4) : (ql[l + 32] >> 4);
            const int q  = (lo | (((qh[l] >> (2 * quarter)) & 3) << 4)) - 32;
            return __fmul_rn(__fmul_rn(half_bits_to_float(*(const uint16_t *) (b + 208)), (float) sc[e >> 4]), (float) q);
        }
    }
    return 0.0f;
}

// ------------------------------------------------------------------------------------------------ GET_ROWS
// dst[:, i10, i11, i12] = row ids[i10, i11, i12] of src0[:, :, i11, i12], as F32
__global__ void __launch_bounds__(256) get_rows_kernel(int type, const char * src0, int64_t ne00, int64_t ne01, int64_t nb01, int64_t nb02, int64_t nb03,
                                                       const char * ids, T4 is, char * dst, int64_t nb1, int64_t nb2, int64_t nb3) {
    pdl_prologue();
    const int64_t r = blockIdx.x;
    const int64_t i10 = r % is.ne[0], i11 = (r / is.ne[0]) % is.ne[1], i12 = r / (is.ne[0] * is.ne[1]);
    const int32_t row = *(const int32_t *) (ids + i10 * is.nb[0] + i11 * is.nb[1] + i12 * is.nb[2]);
    float * out = (float *) (dst + i10 * nb1 + i11 * nb2 + i12 * nb3);
    if (row < 0 || row >= ne01) {   // ggml-cpu asserts; never read out of bounds
        for (int64_t k = threadIdx.x; k < ne00; k += blockDim.x) out[k] = 0.0f;
        return;
    }
    const char * in = src0 + row * nb01 + i11 * nb02 + i12 * nb03;
    if (type == B200_TYPE_F32) {
        for (int64_t k = threadIdx.x; k < ne00; k += blockDim.x) out[k] = ((const float *) in)[k];
    } else if (type == B200_TYPE_F16) {
        for (int64_t k = threadIdx.x; k < ne00; k += blockDim.x) out[k] = __half2float(((const __half *) in)[k]);
    } else {
        for (int64_t k = threadIdx.x; k < ne00; k += blockDim.x) out[k] = dequant_at(type, (const uint8_t *) in, k);
    }
}

// ------------------------------------------------------------------------------------------------ ROPE (normal and neox pairs)
// where a rotated K row also goes (attn_prep_kernel): row idx[token] of the KV cache, at this head's channels, as F16 or F32 -
// the SET_ROWS that follows ROPE(k) in build_attn (src/llama-graph.cpp:1604, llama_kv_cache::cpy_k)
struct RowStore {
    char *       base;      // nullptr: none
    const char * idx;
    int64_t      idx_nb0, row_nb, n_rows;
    int          f16, i64;
};
struct RopeJob {
    T4 s, d;
    const char * x;
    char * y;
    const int32_t * pos;
    const float * ff;
    RowStore st;
};
template <typename OUT>
__device__ __forceinline__ OUT cvt_out(float v);
template <> __device__ __forceinline__ float cvt_out<float>(float v) { return v; }
template <> __device__ __forceinline__ __half cvt_out<__half>(float v) { return __float2half_rn(v); }

// one CTA rotates row `row` (i1 = head, i2 = position, i3 = batch): one thread per pair, or per two pass-through channels beyond n_dims
__device__ __forceinline__ void rope_row(const RopeJob & j, const RopeP & p, int64_t row) {
    const int64_t i1 = row % j.s.ne[1], i2 = (row / j.s.ne[1]) % j.s.ne[2], i3 = row / (j.s.ne[1] * j.s.ne[2]);
    const float * xr = (const float *) (j.x + i1 * j.s.nb[1] + i2 * j.s.nb[2] + i3 * j.s.nb[3]);
    float *       yr = (float *) (j.y + i1 * j.d.nb[1] + i2 * j.d.nb[2] + i3 * j.d.nb[3]);
    const float   base = (float) j.pos[i2];
    char * sr = nullptr;   // the cache row of this token, this head's first channel
    if (j.st.base) {
        const int64_t r = load_idx(j.st.idx + i2 * j.st.idx_nb0, j.st.i64);
        if (r >= 0 && r < j.st.n_rows) sr = j.st.base + r * j.st.row_nb + i1 * j.s.ne[0] * (j.st.f16 ? 2 : 4);
    }
    auto put = [&](int64_t i, float v) {
        yr[i] = v;
        if (sr) { if (j.st.f16) ((__half *) sr)[i] = __float2half_rn(v); else ((float *) sr)[i] = v; }
    };
    for (int64_t i0 = 2 * threadIdx.x; i0 < j.s.ne[0]; i0 += 2 * blockDim.x) {
        if (i0 >= p.n_dims) {   // channels beyond n_dims are copied
            put(i0, xr[i0]); put(i0 + 1, xr[i0 + 1]);
            continue;
        }
        float c, sn;
        rope_cs(p, j.ff, base, (int) (i0 / 2), c, sn);
        const int64_t ia = p.neox ? i0 / 2 : i0, ib = p.neox ? i0 / 2 + p.n_dims / 2 : i0 + 1;
        const float x0 = xr[ia], x1 = xr[ib];
        put(ia, rope_lo(x0, x1, c, sn));   // separately rounded products, as ggml-cpu's x86-64 baseline build (no FMA) computes them
        put(ib, rope_hi(x0, x1, c, sn));
    }
}
__global__ void __launch_bounds__(256) rope_kernel(const RopeJob j, const RopeP p) {
    pdl_prologue();
    rope_row(j, p, blockIdx.x);
}

// ------------------------------------------------------------------------------------------------ SET_ROWS (F32 -> F32 / F16 rows by index)
struct SetRowsJob {
    T4 s, is, d;
    const char * src, * idx;
    char * dst;
    int f16, i64;
};
__device__ __forceinline__ void set_rows_row(const SetRowsJob & j, int64_t r) {
    const int64_t i = r % j.s.ne[1], i02 = (r / j.s.ne[1]) % j.s.ne[2], i03 = r / (j.s.ne[1] * j.s.ne[2]);
    const int64_t i1 = load_idx(j.idx + i * j.is.nb[0] + (i02 % j.is.ne[1]) * j.is.nb[1] + (i03 % j.is.ne[2]) * j.is.nb[2], j.i64);
    if (i1 < 0 || i1 >= j.d.ne[1]) return;   // ggml-cpu asserts
    const float * in = (const float *) (j.src + i * j.s.nb[1] + i02 * j.s.nb[2] + i03 * j.s.nb[3]);
    char * out = j.dst + i1 * j.d.nb[1] + i02 * j.d.nb[2] + i03 * j.d.nb[3];
    if (j.f16) { for (int64_t k = threadIdx.x; k < j.s.ne[0]; k += blockDim.x) ((__half *) out)[k] = __float2half_rn(in[k]); }
    else       { for (int64_t k = threadIdx.x; k < j.s.ne[0]; k += blockDim.x) ((float *) out)[k] = in[k]; }
}
__global__ void __launch_bounds__(256) set_rows_kernel(const SetRowsJob j) {
    pdl_prologue();
    set_rows_row(j, blockIdx.x);
}

// ------------------------------------------------------------------------------------------------ attention inputs in one launch
// ROPE(q), ROPE(k) + SET_ROWS(k cache), SET_ROWS(v cache): four nodes of every layer of a decode graph, independent of each other
// once ROPE(k) stores its rows into the cache itself. CTAs [0, nq) rotate q rows, [nq, nq + nk) k rows, the rest copy v rows.
__global__ void __launch_bounds__(256) attn_prep_kernel(const RopeJob q, const RopeJob k, const SetRowsJob v, const RopeP p, int nq, int nk) {
    pdl_prologue();
    const int b = blockIdx.x;
    if (b < nq) rope_row(q, p, b);
    else if (b < nq + nk) rope_row(k, p, b - nq);
    else set_rows_row(v, b - nq - nk);
}

// The same for a batch of tokens: one CTA per token computes the cos / sin of its position once (the chain of rounded products and
// the sinf / cosf are per pair, not per head) and applies them to every q and k head of the token; then its v row.
__device__ __forceinline__ void rope_heads_tab(const RopeJob & j, const RopeP & p, int64_t i2, const float * cs, const float * sn) {
    const int64_t half = j.s.ne[0] / 2, n = j.s.ne[1] * half;
    char * srow = nullptr;
    if (j.st.base) {
        const int64_t r = load_idx(j.st.idx + i2 * j.st.idx_nb0, j.st.i64);
        if (r >= 0 && r < j.st.n_rows) srow = j.st.base + r * j.st.row_nb;
    }
    for (int64_t t = threadIdx.x; t < n; t += blockDim.x) {
        const int64_t i1 = t / half, pr = t - i1 * half, i0 = 2 * pr;
        const float * xr = (const float *) (j.x + i1 * j.s.nb[1] + i2 * j.s.nb[2]);
        float *       yr = (float *) (j.y + i1 * j.d.nb[1] + i2 * j.d.nb[2]);
        char * sr = srow ? srow + i1 * j.s.ne[0] * (j.st.f16 ? 2 : 4) : nullptr;
        auto put = [&](int64_t i, float v) {
            yr[i] = v;
            if (sr) { if (j.st.f16) ((__half *) sr)[i] = __float2half_rn(v); else ((float *) sr)[i] = v; }
        };
        if (i0 >= p.n_dims) { put(i0, xr[i0]); put(i0 + 1, xr[i0 + 1]); continue; }
        const int64_t ia = p.neox ? pr : i0, ib = p.neox ? pr + p.n_dims / 2 : i0 + 1;
        const float x0 = xr[ia], x1 = xr[ib];
        put(ia, rope_lo(x0, x1, cs[pr], sn[pr]));
        put(ib, rope_hi(x0, x1, cs[pr], sn[pr]));
    }
}
__global__ void __launch_bounds__(256) attn_prep_tok_kernel(const RopeJob q, const RopeJob k, const SetRowsJob v, const RopeP p) {
    __shared__ float cs[256], sn[256];
    pdl_prologue();
    const int64_t tok = blockIdx.x;
    const float base = (float) q.pos[tok];
    for (int ic = threadIdx.x; ic < p.n_dims / 2; ic += blockDim.x) rope_cs(p, q.ff, base, ic, cs[ic], sn[ic]);
    __syncthreads();
    rope_heads_tab(q, p, tok, cs, sn);
    rope_heads_tab(k, p, tok, cs, sn);
    set_rows_row(v, tok);
}

// ------------------------------------------------------------------------------------------------ CPY / CONT / DUP
// element e (in the order of the SOURCE's logical index) goes to the same linear index of the destination's shape
template <typename IN> __device__ __forceinline__ float cvt_in(IN v);
template <> __device__ __forceinline__ float cvt_in<float>(float v) { return v; }
template <> __device__ __forceinline__ float cvt_in<__half>(__half v) { return __half2float(v); }

template <typename IN, typename OUT>
__global__ void __launch_bounds__(256) cpy_kernel(const char * src, T4 s, char * dst, T4 d, int64_t n) {
    pdl_prologue();
    const int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    int64_t r = e;
    const int64_t s0 = r % s.ne[0]; r /= s.ne[0];
    const int64_t s1 = r % s.ne[1]; r /= s.ne[1];
    const int64_t s2 = r % s.ne[2]; const int64_t s3 = r / s.ne[2];
    r = e;
    const int64_t d0 = r % d.ne[0]; r /= d.ne[0];
    const int64_t d1 = r % d.ne[1]; r /= d.ne[1];
    const int64_t d2 = r % d.ne[2]; const int64_t d3 = r / d.ne[2];
    const IN v = *(const IN *) (src + s0 * s.nb[0] + s1 * s.nb[1] + s2 * s.nb[2] + s3 * s.nb[3]);
    *(OUT *) (dst + d0 * d.nb[0] + d1 * d.nb[1] + d2 * d.nb[2] + d3 * d.nb[3]) = cvt_out<OUT>(cvt_in<IN>(v));
}

// ------------------------------------------------------------------------------------------------ FLASH_ATTN_EXT (F16 K / V)
// One CTA per (query row, head, batch) x KV split; 8 warps stride over the KV positions of the split, four at a time (their K and V
// rows are in flight together); a lane owns E consecutive channels of the head (DK = DV = 32 E), so a K or V row is one coalesced
// 64 E-byte read per warp. Online softmax per warp in fp32 (ggml-cpu keeps the F16-V accumulator in fp16: ops.cpp:8102-8123; this
// one is the more accurate side), warps merged through shared memory, KV splits merged by whichever CTA of a row finishes last. The heads of one GQA
// group read the same K/V rows out of L2. Decode-shaped: a wide prefill batch re-reads K/V per query row.
// decode fusion (b200_attn_decode): the kernel rotates its q head and the new token's K row itself (the arithmetic of rope_row),
// uses them - and the new V row - instead of the cache row they belong in, and the first CTA of a KV head writes that row
struct FaDec {
    int on;
    const float * q_raw, * k_raw, * v_raw;   // the projections: [D * n_head], [D * n_head_kv], [D * n_head_kv]
    const int32_t * pos;
    const float * ff;
    RopeP rp;
    const char * k_idx, * v_idx;             // the cache slot of the token (I64 / I32)
    int k_i64, v_i64;
    char * kc, * vc;                          // row 0 of the K / V cache (F16 rows of kc_nb1 / vc_nb1 bytes)
    int64_t kc_nb1, vc_nb1, kc_rows, vc_rows;
    float * q_dst, * k_dst;                   // ROPE(q), ROPE(k) at their graph addresses
};
struct FaP {
    T4 q, k, v, m, d;
    const char * qd, * kd, * vd, * md;
    const float * sinks;
    char * dd;
    float  scale, softcap, max_bias, m0, m1;
    uint32_t n_head_log2;
    int      nsplit;
    int64_t  chunk;
    float *  ws;       // partials of the KV splits: [row][split][2 + D] = running maximum, sum, accumulator
    unsigned * cnt;    // arrivals per row; the CTA that completes a row merges its splits and puts the counter back to 0
    const char * early_idx;   // != nullptr: the only rows of k / v that launches still in flight on the stream may write are rows
    int      early_i64;       // early_idx[0] (the token's cache slot): every other row may be read before the stream dependency resolves
    int      cluster;  // the nsplit CTAs of a row are one thread-block cluster: partials merged through distributed shared memory
    MvNext   next;     // n_mats > 0: weights to pull into L2 meanwhile (kernels.h mv_prefetch_next)
    unsigned long long * trace;   // b200_ctx_trace slot of this launch (8 x u64 per CTA): [0] CTA start, [3] past the stream dependency, [4] first
                                  // round's rows used, [5] rounds done, [6] done, [1] warps merged, [2] partial published / arrival counted
    FaDec    dec;
};
constexpr int FA_MAX_WARPS = 16;   // warps per CTA: 8 or 16 (launch parameter)
constexpr int FAQ_WARPS = 8, FAQ_U = 4;   // fa_rows_kernel: 8 warps, 4 KV positions per warp and round
// a warp keeps FA_U KV positions in flight, K and V rows as packed halves (E registers per position): one round of a CTA covers 8 x FA_U positions
__host__ __device__ constexpr int fa_u(int E) { return 8; }   // measured on the 8B decode step: 8 beats 4 by 2 % and 16 (128 registers, spills) loses 13 %

template <int E> struct HPack { uint32_t w[E / 2]; };
template <int E>
__device__ __forceinline__ HPack<E> load_pack(const char * p) {   // E halves, E * 2 bytes aligned
    HPack<E> r;
    if constexpr (E == 2) {
        r.w[0] = *(const uint32_t *) p;
    } else if constexpr (E == 4) {
        const uint2 u = *(const uint2 *) p;
        r.w[0] = u.x; r.w[1] = u.y;
    } else {
        const uint4 u = *(const uint4 *) p;
        r.w[0] = u.x; r.w[1] = u.y; r.w[2] = u.z; r.w[3] = u.w;
    }
    return r;
}
template <int E>
__device__ __forceinline__ void unpack(const HPack<E> & h, float * out) {
#pragma unroll
    for (int i = 0; i < E / 2; i++) {
        const __half2 v = *(const __half2 *) &h.w[i];
        out[2 * i] = __low2float(v); out[2 * i + 1] = __high2float(v);
    }
}

__device__ __forceinline__ float fa_slope(const FaP & p, uint32_t h) {
    if (!(p.max_bias > 0.0f)) return 1.0f;
    return h < p.n_head_log2 ? powf(p.m0, (float) (h + 1)) : powf(p.m1, (float) (2 * (h - p.n_head_log2) + 1));
}

// dst row of (query iq1, head iq2, batch iq3): dst is [DV, n_head, n_q, n_batch] (ops.cpp:8217)
__device__ __forceinline__ float * fa_dst_row(const FaP & p, int64_t iq1, int64_t iq2, int64_t iq3) {
    return (float *) (p.dd + (iq3 * p.d.ne[2] * p.d.ne[1] + iq2 + iq1 * p.d.ne[1]) * p.d.nb[1]);
}

// decode fusion: qraw / knew / vnew hold the raw projections of this lane's channels; on return the rotated q (F32), and the new K / V
// rows as the cache will hold them (rounded to F16). Warp 0 of the row's first split writes what the graph's ROPE / SET_ROWS nodes would.
template <int E>
__device__ __forceinline__ void fa_dec_rotate(const FaP & p, float (&qraw)[E], float (&knew)[E], float (&vnew)[E], int64_t iq2, int64_t ik2, int64_t iv2,
                                              int warp, int lane) {
    constexpr int D = 32 * E;
    const FaDec & dc = p.dec;
    const float base = (float) dc.pos[0];
    float qr[E], kr[E];
    if (!dc.rp.neox) {   // pairs (2m, 2m + 1): both in this lane
#pragma unroll
        for (int m = 0; m < E / 2; m++) {
            float c, sn;
            rope_cs(dc.rp, dc.ff, base, lane * (E / 2) + m, c, sn);
            qr[2 * m] = rope_lo(qraw[2 * m], qraw[2 * m + 1], c, sn); qr[2 * m + 1] = rope_hi(qraw[2 * m], qraw[2 * m + 1], c, sn);
            kr[2 * m] = rope_lo(knew[2 * m], knew[2 * m + 1], c, sn); kr[2 * m + 1] = rope_hi(knew[2 * m], knew[2 * m + 1], c, sn);
        }
    } else {             // pairs (i, i + D/2): the partner is the same slot of lane ^ 16
#pragma unroll
        for (int e = 0; e < E; e++) {
            float c, sn;
            rope_cs(dc.rp, dc.ff, base, (lane & 15) * E + e, c, sn);
            const float pq_ = __shfl_xor_sync(0xffffffffu, qraw[e], 16), pk_ = __shfl_xor_sync(0xffffffffu, knew[e], 16);
            qr[e] = lane < 16 ? rope_lo(qraw[e], pq_, c, sn) : rope_hi(pq_, qraw[e], c, sn);
            kr[e] = lane < 16 ? rope_lo(knew[e], pk_, c, sn) : rope_hi(pk_, knew[e], c, sn);
        }
    }
    if (blockIdx.y == 0 && warp == 0) {
        const int64_t slot = load_idx(dc.k_idx, dc.k_i64), slot_v = load_idx(dc.v_idx, dc.v_i64);
#pragma unroll
        for (int e = 0; e < E; e++) dc.q_dst[iq2 * D + lane * E + e] = qr[e];
        if (iq2 % (p.q.ne[2] / p.k.ne[2]) == 0) {
#pragma unroll
            for (int e = 0; e < E; e++) dc.k_dst[ik2 * D + lane * E + e] = kr[e];
            if (slot >= 0 && slot < dc.kc_rows) {
                __half * o = (__half *) (dc.kc + slot * dc.kc_nb1) + ik2 * D + lane * E;
#pragma unroll
                for (int e = 0; e < E; e++) o[e] = __float2half_rn(kr[e]);
            }
        }
        if (iq2 % (p.q.ne[2] / p.v.ne[2]) == 0 && slot_v >= 0 && slot_v < dc.vc_rows) {
            __half * o = (__half *) (dc.vc + slot_v * dc.vc_nb1) + iv2 * D + lane * E;
#pragma unroll
            for (int e = 0; e < E; e++) o[e] = __float2half_rn(vnew[e]);
        }
    }
#pragma unroll
    for (int e = 0; e < E; e++) { qraw[e] = qr[e]; knew[e] = __half2float(__float2half_rn(kr[e])); vnew[e] = __half2float(__float2half_rn(vnew[e])); }
}

#define FA_TRACE(i) do { if (p.trace && threadIdx.x == 0) atomicMax(p.trace + (size_t) ((blockIdx.y * gridDim.x + blockIdx.x) % 148) * 8 + (i), global_ns_c()); } while (0)   // CTAs beyond 148 share slots: the latest stamp wins
template <int E, int FA_U, bool DEC = false>
__global__ void __launch_bounds__(FA_MAX_WARPS * 32, 1) fa_kernel(const FaP p) {
    constexpr int D = 32 * E;
    const int FA_WARPS = blockDim.x >> 5;
    FA_TRACE(0);
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int64_t row = blockIdx.x;
    const int64_t iq1 = row % p.q.ne[1], iq2 = (row / p.q.ne[1]) % p.q.ne[2], iq3 = row / (p.q.ne[1] * p.q.ne[2]);
    const int64_t ik2 = iq2 / (p.q.ne[2] / p.k.ne[2]), ik3 = iq3 / (p.q.ne[3] / p.k.ne[3]);
    const int64_t iv2 = iq2 / (p.q.ne[2] / p.v.ne[2]), iv3 = iq3 / (p.q.ne[3] / p.v.ne[3]);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // everything the first round needs is requested before anything is used: q, the mask values and the K / V rows are independent
    // loads (a masked position's rows are read anyway - they are valid cache memory - so that the K / V loads do not wait for the mask)
    const __half * mp = p.md ? (const __half *) (p.md + iq1 * p.m.nb[1] + (iq2 % p.m.ne[2]) * p.m.nb[2] + (iq3 % p.m.ne[3]) * p.m.nb[3]) : nullptr;
    const char * kb = p.kd + ik2 * p.k.nb[2] + ik3 * p.k.nb[3] + lane * E * 2;
    const char * vb = p.vd + iv2 * p.v.nb[2] + iv3 * p.v.nb[3] + lane * E * 2;
    const int64_t c0 = (int64_t) blockIdx.y * p.chunk, c1 = min(c0 + p.chunk, p.k.ne[1]);

    // Decode behind the layer's own KV-cache write (early_idx): the K / V rows of the first round are requested BEFORE the stream dependency
    // resolves - this CTA is resident while the q/k/v launch is still running, and the only row that launch writes is the token's own
    // slot, which is read again afterwards. The device timeline showed these rows arriving 1.7 us after q otherwise: they come from HBM
    // at the moment the next mat-vec launch starts streaming its weights.
    __half mh[FA_U];
    HPack<E> kp[FA_U], vp[FA_U];
    const int64_t base0 = c0 + warp * FA_U;
    const bool early = !DEC && p.early_idx != nullptr && base0 < c1;
    if (early) {
#pragma unroll
        for (int u = 0; u < FA_U; u++) {
            if (base0 + u < c1) {
                kp[u] = load_pack<E>(kb + (base0 + u) * p.k.nb[1]);
                vp[u] = load_pack<E>(vb + (base0 + u) * p.v.nb[1]);
            }
        }
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");
    FA_TRACE(3);
    float qraw[E], knew[E], vnew[E];
    int64_t slot = -1;
    if (early) slot = load_idx(p.early_idx, p.early_i64);
    if constexpr (DEC) {   // the projections of this lane's channels; rotated below, while the first round's K / V rows are on their way
#pragma unroll
        for (int e = 0; e < E; e++) {
            qraw[e] = p.dec.q_raw[iq2 * D + lane * E + e];
            knew[e] = p.dec.k_raw[ik2 * D + lane * E + e];
            vnew[e] = p.dec.v_raw[iv2 * D + lane * E + e];
        }
        slot = load_idx(p.dec.k_idx, p.dec.k_i64);
    } else {
        const float * pq = (const float *) (p.qd + iq1 * p.q.nb[1] + iq2 * p.q.nb[2] + iq3 * p.q.nb[3]) + lane * E;
#pragma unroll
        for (int e = 0; e < E; e++) qraw[e] = pq[e];
    }
    if (p.next.n_mats > 0) mv_prefetch_next(p.next, (int) (blockIdx.y * gridDim.x + blockIdx.x) * FA_WARPS + warp, (int) (gridDim.x * gridDim.y) * FA_WARPS, lane);
    float M = -INFINITY, S = 0.0f, acc[E], qf[E];
#pragma unroll
    for (int e = 0; e < E; e++) acc[e] = 0.0f;
    const float slope = fa_slope(p, (uint32_t) iq2);
    bool q_ready = false;

    for (int64_t base = c0 + warp * FA_U; base < c1; base += FA_WARPS * FA_U) {
        float s[FA_U];
        const bool have = early && base == base0;   // this round's rows were requested ahead of the dependency
#pragma unroll
        for (int u = 0; u < FA_U; u++) {
            if (base + u < c1) {   // warp-uniform
                if (mp) mh[u] = mp[base + u];
                if (!have || base + u == slot) {
                    kp[u] = load_pack<E>(kb + (base + u) * p.k.nb[1]);
                    vp[u] = load_pack<E>(vb + (base + u) * p.v.nb[1]);
                }
            }
        }
        if (!q_ready) {
            if constexpr (DEC) fa_dec_rotate<E>(p, qraw, knew, vnew, iq2, ik2, iv2, warp, lane);
#pragma unroll
            for (int e = 0; e < E; e++) qf[e] = __half2float(__float2half_rn(qraw[e])) * p.scale;   // ggml-cpu converts q to the K type's vec_dot type: fp16
            q_ready = true;
            if (p.trace) { asm volatile("" : "+f"(qf[0])); FA_TRACE(4); }
        }
#pragma unroll
        for (int u = 0; u < FA_U; u++) {
            float t = 0.0f;
            if (base + u < c1) {
                float kf[E];
                unpack<E>(kp[u], kf);
                if (DEC && base + u == slot) {   // the token's own row is not in the cache yet (or is being written right now)
#pragma unroll
                    for (int e = 0; e < E; e++) kf[e] = knew[e];
                }
#pragma unroll
                for (int e = 0; e < E; e++) t += qf[e] * kf[e];
            }
            s[u] = t;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {   // the FA_U reductions side by side
#pragma unroll
            for (int u = 0; u < FA_U; u++) s[u] += __shfl_xor_sync(0xffffffffu, s[u], o);
        }
        float mx = -INFINITY;
#pragma unroll
        for (int u = 0; u < FA_U; u++) {
            const float mv = base + u < c1 ? (mp ? slope * __half2float(mh[u]) : 0.0f) : -INFINITY;
            if (mv != -INFINITY) {
                if (p.softcap != 0.0f) s[u] = p.softcap * tanhf(s[u]);
                s[u] += mv;
            } else {
                s[u] = -INFINITY;
            }
            mx = fmaxf(mx, s[u]);
        }
        if (mx == -INFINITY) continue;
        const float Mn = fmaxf(M, mx), ms = expf(M - Mn);   // M = -inf: expf(-inf) = 0
        M = Mn;
        S *= ms;
#pragma unroll
        for (int e = 0; e < E; e++) acc[e] *= ms;
#pragma unroll
        for (int u = 0; u < FA_U; u++) {
            if (s[u] != -INFINITY) {
                const float pr = expf(s[u] - Mn);
                float vf[E];
                unpack<E>(vp[u], vf);
                if (DEC && base + u == slot) {
#pragma unroll
                    for (int e = 0; e < E; e++) vf[e] = vnew[e];
                }
                S += pr;
#pragma unroll
                for (int e = 0; e < E; e++) acc[e] += pr * vf[e];
            }
        }
    }

    if (p.trace) { asm volatile("" : "+f"(acc[0])); FA_TRACE(5); }
    __shared__ float sh_acc[FA_MAX_WARPS][D];
    __shared__ float sh_m[FA_MAX_WARPS], sh_s[FA_MAX_WARPS];
#pragma unroll
    for (int e = 0; e < E; e++) sh_acc[warp][lane * E + e] = acc[e];
    if (lane == 0) { sh_m[warp] = M; sh_s[warp] = S; }
    __syncthreads();
    FA_TRACE(1);
    float Mx = -INFINITY;
    for (int w = 0; w < FA_WARPS; w++) Mx = fmaxf(Mx, sh_m[w]);
    float St = 0.0f;
    float f[FA_MAX_WARPS];
#pragma unroll
    for (int w = 0; w < FA_MAX_WARPS; w++) { f[w] = w >= FA_WARPS || sh_m[w] == -INFINITY ? 0.0f : expf(sh_m[w] - Mx); St += w < FA_WARPS ? sh_s[w] * f[w] : 0.0f; }
    // KV splits of a row as one cluster (nsplit <= 8): every CTA stores its (max, sum, accumulator) into the shared memory of the
    // cluster's first CTA, one cluster barrier, and that CTA merges them in split order - no global round trips (partials, fence,
    // atomic, reload: 4 - 5 us on the device timeline)
    __shared__ float cl_acc[8][D];
    __shared__ float cl_m[8], cl_s[8];
    if (p.cluster) {
        const uint32_t rank = blockIdx.y;   // cluster dims (1, nsplit, 1): the rank is the split index
        for (int t = threadIdx.x; t < D; t += blockDim.x) {
            float a = 0.0f;
#pragma unroll
            for (int w = 0; w < FA_MAX_WARPS; w++) a += w < FA_WARPS ? sh_acc[w][t] * f[w] : 0.0f;
            uint32_t dst;
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(dst) : "r"((uint32_t) __cvta_generic_to_shared(&cl_acc[rank][t])), "r"(0u));
            asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(dst), "f"(a) : "memory");
        }
        if (threadIdx.x == 0) {
            uint32_t dm, ds;
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(dm) : "r"((uint32_t) __cvta_generic_to_shared(&cl_m[rank])), "r"(0u));
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ds) : "r"((uint32_t) __cvta_generic_to_shared(&cl_s[rank])), "r"(0u));
            asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(dm), "f"(Mx) : "memory");
            asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(ds), "f"(St) : "memory");
        }
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
        FA_TRACE(2);
        if (rank != 0) { FA_TRACE(6); return; }
        float Mg = -INFINITY;
        for (int s = 0; s < p.nsplit; s++) Mg = fmaxf(Mg, cl_m[s]);
        float Sg = 0.0f;
        for (int s = 0; s < p.nsplit; s++) Sg += cl_m[s] == -INFINITY ? 0.0f : cl_s[s] * expf(cl_m[s] - Mg);
        float r = 1.0f;
        if (p.sinks) {
            const float sk = p.sinks[iq2];
            if (sk > Mg) { r = expf(Mg - sk); Sg = Sg * r + 1.0f; } else { Sg += expf(sk - Mg); }
        }
        for (int t = threadIdx.x; t < D; t += blockDim.x) {
            float a = 0.0f;
            for (int s = 0; s < p.nsplit; s++) { if (cl_m[s] != -INFINITY) a += cl_acc[s][t] * expf(cl_m[s] - Mg); }
            fa_dst_row(p, iq1, iq2, iq3)[t] = Sg == 0.0f ? 0.0f : a * r / Sg;
        }
        FA_TRACE(6);
        return;
    }
    for (int t = threadIdx.x; t < D; t += blockDim.x) {
        float a = 0.0f;
#pragma unroll
        for (int w = 0; w < FA_MAX_WARPS; w++) a += w < FA_WARPS ? sh_acc[w][t] * f[w] : 0.0f;
        if (p.nsplit > 1) {
            p.ws[(row * p.nsplit + blockIdx.y) * (D + 2) + 2 + t] = a;
        } else {
            float Sf = St, af = a;
            if (p.sinks) {   // ops.cpp:8190-8204
                const float sk = p.sinks[iq2];
                if (sk > Mx) { const float r = expf(Mx - sk); af *= r; Sf = Sf * r + 1.0f; } else { Sf += expf(sk - Mx); }
            }
            fa_dst_row(p, iq1, iq2, iq3)[t] = Sf == 0.0f ? 0.0f : af / Sf;
        }
    }
    if (p.nsplit == 1) { FA_TRACE(6); return; }
    if (threadIdx.x == 0) {
        float * w = p.ws + (row * p.nsplit + blockIdx.y) * (D + 2);
        w[0] = Mx; w[1] = St;
    }
    // the last CTA of this row to get here merges the splits. Publication: the CTA's partials are ordered before the barrier, thread 0's
    // fence is cumulative over what it has observed through the barrier, then it counts the arrival (the grid-sync pattern: one fence
    // per CTA, not one per thread)
    __shared__ unsigned last;
    __shared__ float cm[64], cs[64];
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned n = atomicAdd(p.cnt + row, 1u);
        last = n == (unsigned) p.nsplit - 1;
        if (last) { p.cnt[row] = 0; __threadfence(); }   // acquire: the other CTAs' partials, for the whole CTA through the barrier below
    }
    __syncthreads();
    FA_TRACE(2);
    if (!last) { FA_TRACE(6); return; }
    const float * w = p.ws + row * p.nsplit * (D + 2);
    // every load of the merge is issued before any is used: maxima and sums by one thread per split, and the first eight splits'
    // accumulators of this thread's channel - one L2 round trip
    float v0[8];
    {
        const int t = threadIdx.x < D ? threadIdx.x : 0;
#pragma unroll
        for (int j = 0; j < 8; j++) v0[j] = j < p.nsplit ? __ldcg(w + j * (D + 2) + 2 + t) : 0.0f;
    }
    if ((int) threadIdx.x < p.nsplit) {
        cm[threadIdx.x] = __ldcg(w + threadIdx.x * (D + 2));
        cs[threadIdx.x] = __ldcg(w + threadIdx.x * (D + 2) + 1);
    }
    __syncthreads();
    float Mg = -INFINITY;
    for (int s = 0; s < p.nsplit; s++) Mg = fmaxf(Mg, cm[s]);
    float Sg = 0.0f;
    for (int s = 0; s < p.nsplit; s++) Sg += cm[s] == -INFINITY ? 0.0f : cs[s] * expf(cm[s] - Mg);
    float r = 1.0f;
    if (p.sinks) {
        const float sk = p.sinks[iq2];
        if (sk > Mg) { r = expf(Mg - sk); Sg = Sg * r + 1.0f; } else { Sg += expf(sk - Mg); }
    }
    for (int t = threadIdx.x; t < D; t += blockDim.x) {
        float a = 0.0f;
        for (int s0 = 0; s0 < p.nsplit; s0 += 8) {
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = s0 == 0 && t == (int) threadIdx.x ? v0[j] : (s0 + j < p.nsplit ? __ldcg(w + (s0 + j) * (D + 2) + 2 + t) : 0.0f);
#pragma unroll
            for (int j = 0; j < 8; j++) {
                if (s0 + j < p.nsplit && cm[s0 + j] != -INFINITY) a += v[j] * expf(cm[s0 + j] - Mg);
            }
        }
        fa_dst_row(p, iq1, iq2, iq3)[t] = Sg == 0.0f ? 0.0f : a * r / Sg;
    }
    FA_TRACE(6);
}

// Wide batches (prompt processing): a CTA takes NQ consecutive query rows of one head, so a K / V row is loaded once for all of them
// (the decode kernel above would re-read the whole K / V per query row). Same arithmetic per (query, position) pair; a round whose
// positions are masked for every row of the block - the upper triangle of a causal prompt - is skipped before its K / V loads.
template <int E, int NQ>
__global__ void __launch_bounds__(FAQ_WARPS * 32, 1) fa_rows_kernel(const FaP p) {
    constexpr int D = 32 * E, U = FAQ_U;
    pdl_prologue();
    const int64_t nqb = (p.q.ne[1] + NQ - 1) / NQ;
    const int64_t blk = blockIdx.x;
    const int64_t iq1b = blk % nqb, iq2 = (blk / nqb) % p.q.ne[2], iq3 = blk / (nqb * p.q.ne[2]);
    const int64_t q0 = iq1b * NQ;
    const int     nq = (int) min((int64_t) NQ, p.q.ne[1] - q0);
    const int64_t ik2 = iq2 / (p.q.ne[2] / p.k.ne[2]), ik3 = iq3 / (p.q.ne[3] / p.k.ne[3]);
    const int64_t iv2 = iq2 / (p.q.ne[2] / p.v.ne[2]), iv3 = iq3 / (p.q.ne[3] / p.v.ne[3]);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t n_kv = p.k.ne[1];

    float qf[NQ][E], acc[NQ][E], M[NQ], S[NQ];
    const __half * mp[NQ];
#pragma unroll
    for (int r = 0; r < NQ; r++) {
        const int64_t iq1 = q0 + min(r, nq - 1);   // rows past the end repeat the last one and are never stored
        const float * pq = (const float *) (p.qd + iq1 * p.q.nb[1] + iq2 * p.q.nb[2] + iq3 * p.q.nb[3]) + lane * E;
#pragma unroll
        for (int e = 0; e < E; e++) { qf[r][e] = __half2float(__float2half_rn(pq[e])) * p.scale; acc[r][e] = 0.0f; }
        mp[r] = p.md ? (const __half *) (p.md + iq1 * p.m.nb[1] + (iq2 % p.m.ne[2]) * p.m.nb[2] + (iq3 % p.m.ne[3]) * p.m.nb[3]) : nullptr;
        M[r] = -INFINITY; S[r] = 0.0f;
    }
    const float slope = fa_slope(p, (uint32_t) iq2);
    const char * kb = p.kd + ik2 * p.k.nb[2] + ik3 * p.k.nb[3] + lane * E * 2;
    const char * vb = p.vd + iv2 * p.v.nb[2] + iv3 * p.v.nb[3] + lane * E * 2;

    for (int64_t base = warp * U; base < n_kv; base += FAQ_WARPS * U) {
        float s[NQ][U];
        bool any = false;
#pragma unroll
        for (int r = 0; r < NQ; r++) {
#pragma unroll
            for (int u = 0; u < U; u++) {
                const float mv = r < nq && base + u < n_kv ? (mp[r] ? slope * __half2float(mp[r][base + u]) : 0.0f) : -INFINITY;
                s[r][u] = mv;   // the mask value until the score replaces it
                any = any || mv != -INFINITY;
            }
        }
        if (!any) continue;   // warp-uniform: every lane reads the same mask values
        HPack<E> kp[U], vp[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (base + u < n_kv) {
                kp[u] = load_pack<E>(kb + (base + u) * p.k.nb[1]);
                vp[u] = load_pack<E>(vb + (base + u) * p.v.nb[1]);
            }
        }
        float d[NQ][U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            float kf[E];
            if (base + u < n_kv) unpack<E>(kp[u], kf);
#pragma unroll
            for (int r = 0; r < NQ; r++) {
                float t = 0.0f;
                if (base + u < n_kv) {
#pragma unroll
                    for (int e = 0; e < E; e++) t += qf[r][e] * kf[e];
                }
                d[r][u] = t;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int r = 0; r < NQ; r++) {
#pragma unroll
                for (int u = 0; u < U; u++) d[r][u] += __shfl_xor_sync(0xffffffffu, d[r][u], o);
            }
        }
#pragma unroll
        for (int r = 0; r < NQ; r++) {
            float mx = -INFINITY;
#pragma unroll
            for (int u = 0; u < U; u++) {
                if (s[r][u] != -INFINITY) {
                    float t = d[r][u];
                    if (p.softcap != 0.0f) t = p.softcap * tanhf(t);
                    s[r][u] += t;
                }
                mx = fmaxf(mx, s[r][u]);
            }
            if (mx == -INFINITY) {
#pragma unroll
                for (int u = 0; u < U; u++) s[r][u] = 0.0f;   // no weight for this row in this round
                continue;
            }
            const float Mn = fmaxf(M[r], mx), ms = expf(M[r] - Mn);
            M[r] = Mn;
            S[r] *= ms;
#pragma unroll
            for (int e = 0; e < E; e++) acc[r][e] *= ms;
#pragma unroll
            for (int u = 0; u < U; u++) {
                const float pr = s[r][u] == -INFINITY ? 0.0f : expf(s[r][u] - Mn);
                S[r] += pr;
                s[r][u] = pr;
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (base + u < n_kv) {
                float vf[E];
                unpack<E>(vp[u], vf);
#pragma unroll
                for (int r = 0; r < NQ; r++) {
#pragma unroll
                    for (int e = 0; e < E; e++) acc[r][e] += s[r][u] * vf[e];
                }
            }
        }
    }

    __shared__ float sh_acc[FAQ_WARPS][NQ][D];
    __shared__ float sh_m[FAQ_WARPS][NQ], sh_s[FAQ_WARPS][NQ];
#pragma unroll
    for (int r = 0; r < NQ; r++) {
#pragma unroll
        for (int e = 0; e < E; e++) sh_acc[warp][r][lane * E + e] = acc[r][e];
        if (lane == 0) { sh_m[warp][r] = M[r]; sh_s[warp][r] = S[r]; }
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < nq * D; idx += blockDim.x) {
        const int r = idx / D, t = idx - r * D;
        float Mx = -INFINITY;
#pragma unroll
        for (int w = 0; w < FAQ_WARPS; w++) Mx = fmaxf(Mx, sh_m[w][r]);
        float St = 0.0f, a = 0.0f;
#pragma unroll
        for (int w = 0; w < FAQ_WARPS; w++) {
            const float f = sh_m[w][r] == -INFINITY ? 0.0f : expf(sh_m[w][r] - Mx);
            St += sh_s[w][r] * f;
            a  += sh_acc[w][r][t] * f;
        }
        if (p.sinks) {
            const float sk = p.sinks[iq2];
            if (sk > Mx) { const float rr = expf(Mx - sk); a *= rr; St = St * rr + 1.0f; } else { St += expf(sk - Mx); }
        }
        fa_dst_row(p, q0 + r, iq2, iq3)[t] = St == 0.0f ? 0.0f : a / St;
    }
}

// ------------------------------------------------------------------------------------------------ FLASH_ATTN_EXT, prompt batches
// Tensor-core tiles (warp-level mma.sync m16n8k16, fp16 x fp16 -> fp32; tcgen05 would want 128-row tiles and TMEM round trips for the
// softmax - at 64-row blocks the attention of a 512-token prompt is ~2 % of its MUL_MATs either way). A CTA of 4 warps takes 64
// query rows of one head (16 per warp: Q lives in registers as A fragments, fp16 as ggml-cpu's own q conversion) and walks the KV
// cache in tiles of 64 positions staged in shared memory (rows padded to D + 8 halves: the fragment loads are bank-conflict free):
// S = Q K^T (fp32), scale / soft-cap / mask, online softmax in fp32, P rounded to fp16 as the A operand of O += P V (fp32).
// A tile that is masked for all 64 rows (the upper triangle of a causal prompt) is skipped before it is loaded.
constexpr int FT_BQ = 64, FT_BK = 64, FT_WARPS = 4;

__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_half2(float lo, float hi) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *(const uint32_t *) &h;
}

template <int D>
__global__ void __launch_bounds__(FT_WARPS * 32) fa_tile_kernel(const FaP p) {
    constexpr int LD = D + 8, KS = D / 16, NTD = D / 8, NTS = FT_BK / 8;
    __shared__ __align__(16) __half Ks[FT_BK][LD];
    __shared__ __align__(16) __half Vs[FT_BK][LD];
    pdl_prologue();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int64_t iq2 = blockIdx.y, iq3 = blockIdx.z;
    const int64_t q0 = (int64_t) blockIdx.x * FT_BQ + warp * 16;
    const int64_t n_q = p.q.ne[1], n_kv = p.k.ne[1];
    const int64_t ik2 = iq2 / (p.q.ne[2] / p.k.ne[2]), ik3 = iq3 / (p.q.ne[3] / p.k.ne[3]);
    const int64_t iv2 = iq2 / (p.q.ne[2] / p.v.ne[2]), iv3 = iq3 / (p.q.ne[3] / p.v.ne[3]);
    const int64_t r0 = q0 + g, r1 = q0 + g + 8;   // this thread's two query rows
    const bool ok0 = r0 < n_q, ok1 = r1 < n_q;

    uint32_t qa[KS][4];   // A fragments of the warp's 16 x D block of Q
    {
        const char * qb = p.qd + iq2 * p.q.nb[2] + iq3 * p.q.nb[3];
        const float * q0p = (const float *) (qb + (ok0 ? r0 : 0) * p.q.nb[1]), * q1p = (const float *) (qb + (ok1 ? r1 : 0) * p.q.nb[1]);
#pragma unroll
        for (int ks = 0; ks < KS; ks++) {
            const int c = ks * 16 + 2 * t;
            qa[ks][0] = ok0 ? pack_half2(q0p[c], q0p[c + 1]) : 0u;
            qa[ks][1] = ok1 ? pack_half2(q1p[c], q1p[c + 1]) : 0u;
            qa[ks][2] = ok0 ? pack_half2(q0p[c + 8], q0p[c + 9]) : 0u;
            qa[ks][3] = ok1 ? pack_half2(q1p[c + 8], q1p[c + 9]) : 0u;
        }
    }
    const char * mb = p.md ? p.md + (iq2 % p.m.ne[2]) * p.m.nb[2] + (iq3 % p.m.ne[3]) * p.m.nb[3] : nullptr;
    const __half * m0p = mb && ok0 ? (const __half *) (mb + r0 * p.m.nb[1]) : nullptr, * m1p = mb && ok1 ? (const __half *) (mb + r1 * p.m.nb[1]) : nullptr;
    const float slope = fa_slope(p, (uint32_t) iq2);
    const char * kb = p.kd + ik2 * p.k.nb[2] + ik3 * p.k.nb[3];
    const char * vb = p.vd + iv2 * p.v.nb[2] + iv3 * p.v.nb[3];

    float o[NTD][4];
#pragma unroll
    for (int n = 0; n < NTD; n++) { o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.0f; }
    float mrow[2] = { -INFINITY, -INFINITY }, lrow[2] = { 0.0f, 0.0f };

    for (int64_t j0 = 0; j0 < n_kv; j0 += FT_BK) {
        // mask values of this thread's (row, position) pairs; -inf: masked, out of range, or a row past the batch
        float mv[NTS][4];
        bool any = false;
#pragma unroll
        for (int n = 0; n < NTS; n++) {
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int64_t pos = j0 + n * 8 + 2 * t + (i & 1);
                const bool okr = i < 2 ? ok0 : ok1;
                const __half * mp = i < 2 ? m0p : m1p;
                float v = okr && pos < n_kv ? (mb ? slope * __half2float(mp[pos]) : 0.0f) : -INFINITY;
                mv[n][i] = v;
                any = any || v != -INFINITY;
            }
        }
        if (!__syncthreads_or(any)) continue;   // also orders the previous tile's shared-memory reads before this tile's writes
        for (int c = threadIdx.x; c < FT_BK * (D / 8); c += FT_WARPS * 32) {
            const int row = c / (D / 8), col = (c % (D / 8)) * 8;
            const int64_t pos = j0 + row;
            uint4 kk = make_uint4(0, 0, 0, 0), vv = make_uint4(0, 0, 0, 0);
            if (pos < n_kv) {
                kk = *(const uint4 *) (kb + pos * p.k.nb[1] + col * 2);
                vv = *(const uint4 *) (vb + pos * p.v.nb[1] + col * 2);
            }
            *(uint4 *) &Ks[row][col] = kk;
            *(uint4 *) &Vs[row][col] = vv;
        }
        __syncthreads();

        float sc[NTS][4];
#pragma unroll
        for (int n = 0; n < NTS; n++) {
            sc[n][0] = sc[n][1] = sc[n][2] = sc[n][3] = 0.0f;
#pragma unroll
            for (int ks = 0; ks < KS; ks++) {
                const uint32_t b0 = *(const uint32_t *) &Ks[n * 8 + g][ks * 16 + 2 * t], b1 = *(const uint32_t *) &Ks[n * 8 + g][ks * 16 + 2 * t + 8];
                mma_16816(sc[n], qa[ks], b0, b1);
            }
        }
        float mx[2] = { -INFINITY, -INFINITY };
#pragma unroll
        for (int n = 0; n < NTS; n++) {
#pragma unroll
            for (int i = 0; i < 4; i++) {
                float v = sc[n][i] * p.scale;
                if (p.softcap != 0.0f) v = p.softcap * tanhf(v);
                v = mv[n][i] == -INFINITY ? -INFINITY : v + mv[n][i];
                sc[n][i] = v;
                mx[i >> 1] = fmaxf(mx[i >> 1], v);
            }
        }
        float alpha[2], msafe[2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
            mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
            mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 2));
            const float mn = fmaxf(mrow[h], mx[h]);
            msafe[h] = mn == -INFINITY ? 0.0f : mn;
            alpha[h] = expf(mrow[h] - msafe[h]);   // mrow = -inf: 0
            mrow[h]  = mn;
            lrow[h] *= alpha[h];
        }
#pragma unroll
        for (int n = 0; n < NTD; n++) { o[n][0] *= alpha[0]; o[n][1] *= alpha[0]; o[n][2] *= alpha[1]; o[n][3] *= alpha[1]; }
        uint32_t pa[NTS / 2][4];   // P as A fragments: positions kk * 16 .. + 15
#pragma unroll
        for (int n = 0; n < NTS; n++) {
            float e[4];
#pragma unroll
            for (int i = 0; i < 4; i++) { e[i] = expf(sc[n][i] - msafe[i >> 1]); lrow[i >> 1] += e[i]; }
            pa[n >> 1][(n & 1) * 2 + 0] = pack_half2(e[0], e[1]);
            pa[n >> 1][(n & 1) * 2 + 1] = pack_half2(e[2], e[3]);
        }
#pragma unroll
        for (int n = 0; n < NTD; n++) {
#pragma unroll
            for (int kk = 0; kk < NTS / 2; kk++) {
                const __half2 h0 = __halves2half2(Vs[kk * 16 + 2 * t][n * 8 + g], Vs[kk * 16 + 2 * t + 1][n * 8 + g]);
                const __half2 h1 = __halves2half2(Vs[kk * 16 + 2 * t + 8][n * 8 + g], Vs[kk * 16 + 2 * t + 9][n * 8 + g]);
                mma_16816(o[n], pa[kk], *(const uint32_t *) &h0, *(const uint32_t *) &h1);
            }
        }
    }
    // the row sums live spread over the four lanes of a quad
#pragma unroll
    for (int h = 0; h < 2; h++) {
        lrow[h] += __shfl_xor_sync(0xffffffffu, lrow[h], 1);
        lrow[h] += __shfl_xor_sync(0xffffffffu, lrow[h], 2);
    }
    float rs[2] = { 1.0f, 1.0f };
    if (p.sinks) {
        const float sk = p.sinks[iq2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
            if (sk > mrow[h]) { rs[h] = expf(mrow[h] - sk); lrow[h] = lrow[h] * rs[h] + 1.0f; } else { lrow[h] += expf(sk - mrow[h]); }
        }
    }
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int64_t r = h ? r1 : r0;
        if (!(h ? ok1 : ok0)) continue;
        float * dr = fa_dst_row(p, r, iq2, iq3);
        const float inv = lrow[h] == 0.0f ? 0.0f : rs[h] / lrow[h];
#pragma unroll
        for (int n = 0; n < NTD; n++) *(float2 *) (dr + n * 8 + 2 * t) = make_float2(o[n][2 * h] * inv, o[n][2 * h + 1] * inv);
    }
}

bool rows_ok(const b200_tensor * t, size_t elem) { return t && t->nb[0] == elem; }
int64_t nelem(const b200_tensor * t) { return t->ne[0] * t->ne[1] * t->ne[2] * t->ne[3]; }
bool nonneg(const b200_tensor * t) { return t->ne[0] >= 0 && t->ne[1] >= 0 && t->ne[2] >= 0 && t->ne[3] >= 0; }

#define OPS_PROLOGUE(ctx)                                                                      \
    if (!(ctx)) return fail(B200_ERR_ARG, "null context");                                     \
    CU(cudaSetDevice((ctx)->device));                                                          \
    if (b200_status fst_ = b200_split_flush(ctx)) return fst_

// the decode attention kernel: programmatic dependent launch, and - when the KV splits of a row merge through distributed shared
// memory - the nsplit CTAs of a row as one cluster
template <typename K>
static cudaError_t fa_launch(const b200_ctx * ctx, K kernel, dim3 grid, dim3 block, const FaP & p) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (ctx->pdl) { attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization; attr[na].val.programmaticStreamSerializationAllowed = 1; na++; }
    if (p.cluster) { attr[na].id = cudaLaunchAttributeClusterDimension; attr[na].val.clusterDim.x = 1; attr[na].val.clusterDim.y = (unsigned) p.nsplit; attr[na].val.clusterDim.z = 1; na++; }
    cfg.attrs = attr; cfg.numAttrs = na;
    return cudaLaunchKernelEx(&cfg, kernel, p);
}

} // namespace

extern "C" {

int b200_rms_norm_supported(const b200_tensor * x, const b200_tensor * dst) {
    return x && dst && x->type == B200_TYPE_F32 && dst->type == B200_TYPE_F32 && rows_ok(x, 4) && rows_ok(dst, 4) && nonneg(x) &&
           x->ne[0] == dst->ne[0] && x->ne[1] == dst->ne[1] && x->ne[2] == dst->ne[2] && x->ne[3] == dst->ne[3] &&
           x->ne[1] * x->ne[2] * x->ne[3] < (1ll << 31);
}

int b200_rms_norm_mul_supported(const b200_tensor * x, const b200_tensor * w, const b200_tensor * dst) {
    if (!b200_rms_norm_supported(x, dst) || !w || w->type != B200_TYPE_F32 || w->nb[0] != 4) return 0;
    if (w->ne[0] != x->ne[0]) return 0;   // per-channel weights; rows, planes and batches broadcast
    for (int i = 1; i < 4; i++) if (w->ne[i] <= 0 || x->ne[i] % w->ne[i]) return 0;
    return 1;
}

static b200_status rms_norm_run(b200_ctx * ctx, const b200_tensor * x, const b200_tensor * w, const b200_tensor * dst, float eps) {
    const int64_t rows = x->ne[1] * x->ne[2] * x->ne[3], n = x->ne[0];
    if (rows == 0 || n == 0) return B200_OK;
    if (!x->data || !dst->data || (w && !w->data)) return fail(B200_ERR_ARG, "null argument");
    NormP p;
    p.s = t4(x); p.d = t4(dst); p.w = w ? t4(w) : T4{};
    p.x = (const char *) x->data; p.y = (char *) dst->data; p.wd = w ? (const char *) w->data : nullptr;
    p.eps = eps;
    bool vec = n % 4 == 0 && n <= 16384 && (uintptr_t) x->data % 16 == 0 && (uintptr_t) dst->data % 16 == 0 && (!w || (uintptr_t) w->data % 16 == 0);
    for (int i = 1; i < 4 && vec; i++) vec = x->nb[i] % 16 == 0 && dst->nb[i] % 16 == 0 && (!w || w->nb[i] % 16 == 0);
    const int threads = (int) std::min<int64_t>(1024, std::max<int64_t>(32, ((vec ? n / 4 : n) + 31) / 32 * 32));
    OPS_PROLOGUE(ctx);
    if (vec && n % 256 == 0) {   // the canonical order (n <= 16384: at most 64 units)
        CU(b200_launch(ctx, rms_norm_units_kernel, dim3((unsigned) rows), dim3(32 * (unsigned) std::min<int64_t>(32, n / 256)), 0, p));
        ctx->launches++;
        return B200_OK;
    }
    CU(vec ? b200_launch(ctx, rms_norm_kernel<true>, dim3((unsigned) rows), dim3(threads), 0, p)
           : b200_launch(ctx, rms_norm_kernel<false>, dim3((unsigned) rows), dim3(threads), 0, p));
    ctx->launches++;
    return B200_OK;
}

b200_status b200_rms_norm(b200_ctx * ctx, const b200_tensor * x, const b200_tensor * dst, float eps) {
    if (!b200_rms_norm_supported(x, dst)) return fail(B200_ERR_UNSUPPORTED, "RMS_NORM: F32 rows, same shapes");
    return rms_norm_run(ctx, x, nullptr, dst, eps);
}

b200_status b200_rms_norm_mul(b200_ctx * ctx, const b200_tensor * x, const b200_tensor * w, const b200_tensor * dst, float eps) {
    if (!b200_rms_norm_mul_supported(x, w, dst)) return fail(B200_ERR_UNSUPPORTED, "RMS_NORM + MUL: F32 rows, per-channel weights broadcast over rows");
    return rms_norm_run(ctx, x, w, dst, eps);
}

int b200_get_rows_supported(const b200_tensor * src0, const b200_tensor * ids, const b200_tensor * dst) {
    if (!src0 || !ids || !dst || dst->type != B200_TYPE_F32 || ids->type != B200_TYPE_I32 || !rows_ok(dst, 4)) return 0;
    if (!nonneg(src0) || !nonneg(ids) || ids->ne[3] != 1) return 0;
    switch (src0->type) {
        case B200_TYPE_F32: if (src0->nb[0] != 4) return 0; break;
        case B200_TYPE_F16: if (src0->nb[0] != 2) return 0; break;
        case B200_TYPE_Q4_0: case B200_TYPE_Q8_0: if (src0->ne[0] % 32) return 0; break;
        case B200_TYPE_Q4_K: case B200_TYPE_Q5_K: case B200_TYPE_Q6_K: if (src0->ne[0] % 256) return 0; break;
        default: return 0;
    }
    if (ids->ne[1] != src0->ne[2] || ids->ne[2] != src0->ne[3]) return 0;
    return dst->ne[0] == src0->ne[0] && dst->ne[1] == ids->ne[0] && dst->ne[2] == ids->ne[1] && dst->ne[3] == ids->ne[2] &&
           ids->ne[0] * ids->ne[1] * ids->ne[2] < (1ll << 31);
}

b200_status b200_get_rows(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * ids, const b200_tensor * dst) {
    if (!b200_get_rows_supported(src0, ids, dst)) return fail(B200_ERR_UNSUPPORTED, "GET_ROWS: F32/F16/q4_0/q8_0/q4_K/q5_K/q6_K rows by I32 index into F32");
    const int64_t rows = ids->ne[0] * ids->ne[1] * ids->ne[2];
    if (rows == 0 || src0->ne[0] == 0) return B200_OK;
    if (!src0->data || !ids->data || !dst->data) return fail(B200_ERR_ARG, "null argument");
    OPS_PROLOGUE(ctx);
    CU(b200_launch(ctx, get_rows_kernel, dim3((unsigned) rows), dim3(256), 0, (int) src0->type, (const char *) src0->data, src0->ne[0], src0->ne[1],
                   (int64_t) src0->nb[1], (int64_t) src0->nb[2], (int64_t) src0->nb[3], (const char *) ids->data, t4(ids), (char *) dst->data,
                   (int64_t) dst->nb[1], (int64_t) dst->nb[2], (int64_t) dst->nb[3]));
    ctx->launches++;
    return B200_OK;
}

int b200_rope_supported(const b200_tensor * x, const b200_tensor * pos, const b200_tensor * freq_factors, const b200_tensor * dst, const b200_rope_params * rp) {
    if (!x || !pos || !dst || !rp || x->type != B200_TYPE_F32 || dst->type != B200_TYPE_F32 || pos->type != B200_TYPE_I32) return 0;
    if (!rows_ok(x, 4) || !rows_ok(dst, 4) || !nonneg(x) || pos->nb[0] != 4 || pos->ne[0] < x->ne[2]) return 0;
    for (int i = 0; i < 4; i++) if (x->ne[i] != dst->ne[i]) return 0;
    if (rp->mode != B200_ROPE_NORMAL && rp->mode != B200_ROPE_NEOX) return 0;   // mrope / vision: not on the Llama decode path
    if (rp->n_dims <= 0 || rp->n_dims % 2 || rp->n_dims > x->ne[0] || x->ne[0] % 2) return 0;
    if (freq_factors && (freq_factors->type != B200_TYPE_F32 || freq_factors->nb[0] != 4 || freq_factors->ne[0] < rp->n_dims / 2)) return 0;
    return x->ne[1] * x->ne[2] * x->ne[3] < (1ll << 31);
}

static RopeP rope_consts(const b200_rope_params * rp) { return b200_rope_consts(rp); }
} // extern "C"
b200::RopeP b200_rope_consts(const b200_rope_params * rp) {
    // ggml_rope_yarn_corr_dims (ggml.c: ggml_rope_yarn_corr_dim): the channel range the YaRN ramp covers
    RopeP p;
    p.n_dims = rp->n_dims; p.neox = rp->mode == B200_ROPE_NEOX;
    p.theta_scale = powf(rp->freq_base, -2.0f / (float) rp->n_dims);
    p.freq_scale = rp->freq_scale; p.ext_factor = rp->ext_factor; p.attn_factor = rp->attn_factor;
    auto corr_dim = [&](float n_rot) { return (float) rp->n_dims * logf((float) rp->n_ctx_orig / (n_rot * 2.0f * 3.14159265358979323846f)) / (2.0f * logf(rp->freq_base)); };
    p.corr0 = fmaxf(0.0f, floorf(corr_dim(rp->beta_fast)));
    p.corr1 = fminf((float) rp->n_dims - 1.0f, ceilf(corr_dim(rp->beta_slow)));
    return p;
}
extern "C" {
static RopeJob rope_job(const b200_tensor * x, const b200_tensor * pos, const b200_tensor * ff, const b200_tensor * dst) {
    RopeJob j;
    j.s = t4(x); j.d = t4(dst);
    j.x = (const char *) x->data; j.y = (char *) dst->data;
    j.pos = (const int32_t *) pos->data;
    j.ff = ff ? (const float *) ff->data : nullptr;
    j.st = RowStore{};
    return j;
}
static int rope_threads(int64_t ne0) { return (int) std::min<int64_t>(256, std::max<int64_t>(32, (ne0 / 2 + 31) / 32 * 32)); }

b200_status b200_rope(b200_ctx * ctx, const b200_tensor * x, const b200_tensor * pos, const b200_tensor * freq_factors, const b200_tensor * dst,
                      const b200_rope_params * rp) {
    if (!b200_rope_supported(x, pos, freq_factors, dst, rp)) return fail(B200_ERR_UNSUPPORTED, "ROPE: F32, normal or neox mode");
    const int64_t rows = x->ne[1] * x->ne[2] * x->ne[3];
    if (rows == 0 || x->ne[0] == 0) return B200_OK;
    if (!x->data || !pos->data || !dst->data || (freq_factors && !freq_factors->data)) return fail(B200_ERR_ARG, "null argument");
    OPS_PROLOGUE(ctx);
    CU(b200_launch(ctx, rope_kernel, dim3((unsigned) rows), dim3(rope_threads(x->ne[0])), 0, rope_job(x, pos, freq_factors, dst), rope_consts(rp)));
    ctx->launches++;
    return B200_OK;
}

int b200_set_rows_supported(const b200_tensor * src, const b200_tensor * idx, const b200_tensor * dst) {
    if (!src || !idx || !dst || src->type != B200_TYPE_F32 || !rows_ok(src, 4) || !nonneg(src) || !nonneg(idx)) return 0;
    if (idx->type != B200_TYPE_I64 && idx->type != B200_TYPE_I32) return 0;
    if (!((dst->type == B200_TYPE_F32 && dst->nb[0] == 4) || (dst->type == B200_TYPE_F16 && dst->nb[0] == 2))) return 0;
    if (dst->ne[0] != src->ne[0] || dst->ne[2] != src->ne[2] || dst->ne[3] != src->ne[3] || idx->ne[0] != src->ne[1] || idx->ne[3] != 1) return 0;
    if (idx->ne[1] <= 0 || idx->ne[2] <= 0 || src->ne[2] % idx->ne[1] || src->ne[3] % idx->ne[2]) return 0;
    return src->ne[1] * src->ne[2] * src->ne[3] < (1ll << 31);
}

static SetRowsJob set_rows_job(const b200_tensor * src, const b200_tensor * idx, const b200_tensor * dst) {
    SetRowsJob j;
    j.s = t4(src); j.is = t4(idx); j.d = t4(dst);
    j.src = (const char *) src->data; j.idx = (const char *) idx->data; j.dst = (char *) dst->data;
    j.f16 = dst->type == B200_TYPE_F16; j.i64 = idx->type == B200_TYPE_I64;
    return j;
}

b200_status b200_set_rows(b200_ctx * ctx, const b200_tensor * src, const b200_tensor * idx, const b200_tensor * dst) {
    if (!b200_set_rows_supported(src, idx, dst)) return fail(B200_ERR_UNSUPPORTED, "SET_ROWS: F32 rows into F32 / F16 by I64 / I32 index");
    const int64_t rows = src->ne[1] * src->ne[2] * src->ne[3];
    if (rows == 0 || src->ne[0] == 0) return B200_OK;
    if (!src->data || !idx->data || !dst->data) return fail(B200_ERR_ARG, "null argument");
    OPS_PROLOGUE(ctx);
    CU(b200_launch(ctx, set_rows_kernel, dim3((unsigned) rows), dim3(256), 0, set_rows_job(src, idx, dst)));
    ctx->launches++;
    return B200_OK;
}

// ROPE(q), ROPE(k), SET_ROWS(k_cache <- rope(k) seen as [n_embd_k_gqa, n_tokens]) and SET_ROWS(v_cache <- v) as one launch
int b200_attn_prep_supported(const b200_tensor * q, const b200_tensor * q_dst, const b200_tensor * k, const b200_tensor * k_dst, const b200_tensor * pos,
                             const b200_tensor * freq_factors, const b200_rope_params * rp, const b200_tensor * k_idx, const b200_tensor * k_cache,
                             const b200_tensor * v, const b200_tensor * v_idx, const b200_tensor * v_cache) {
    if (!b200_rope_supported(q, pos, freq_factors, q_dst, rp) || !b200_rope_supported(k, pos, freq_factors, k_dst, rp)) return 0;
    if (k->ne[3] != 1 || k_dst->nb[1] != (size_t) k->ne[0] * 4) return 0;   // the heads of a token are contiguous: one cache row
    // the SET_ROWS of k: src = k_dst as [ne0 * ne1, ne2] rows
    b200_tensor ks = *k_dst;
    ks.ne[0] = k->ne[0] * k->ne[1]; ks.ne[1] = k->ne[2]; ks.ne[2] = 1; ks.ne[3] = 1;
    ks.nb[1] = k_dst->nb[2]; ks.nb[2] = ks.nb[3] = ks.nb[1] * (size_t) ks.ne[1];
    if (!b200_set_rows_supported(&ks, k_idx, k_cache) || k_idx->ne[1] != 1 || k_idx->ne[2] != 1 || k_cache->ne[2] != 1 || k_cache->ne[3] != 1) return 0;
    if (!b200_set_rows_supported(v, v_idx, v_cache)) return 0;
    return q->ne[1] * q->ne[2] * q->ne[3] + k->ne[1] * k->ne[2] + v->ne[1] * v->ne[2] * v->ne[3] < (1ll << 31);
}

b200_status b200_attn_prep(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * q_dst, const b200_tensor * k, const b200_tensor * k_dst,
                           const b200_tensor * pos, const b200_tensor * freq_factors, const b200_rope_params * rp, const b200_tensor * k_idx,
                           const b200_tensor * k_cache, const b200_tensor * v, const b200_tensor * v_idx, const b200_tensor * v_cache) {
    if (!b200_attn_prep_supported(q, q_dst, k, k_dst, pos, freq_factors, rp, k_idx, k_cache, v, v_idx, v_cache))
        return fail(B200_ERR_UNSUPPORTED, "attn_prep: ROPE(q), ROPE(k) + SET_ROWS(k), SET_ROWS(v) of one layer");
    const int64_t nq = q->ne[1] * q->ne[2] * q->ne[3], nk = k->ne[1] * k->ne[2], nv = v->ne[1] * v->ne[2] * v->ne[3];
    if (nq + nk + nv == 0 || q->ne[0] == 0) return B200_OK;
    for (const b200_tensor * t : { q, q_dst, k, k_dst, pos, k_idx, k_cache, v, v_idx, v_cache }) if (!t->data) return fail(B200_ERR_ARG, "null argument");
    if (freq_factors && !freq_factors->data) return fail(B200_ERR_ARG, "null argument");
    RopeJob jq = rope_job(q, pos, freq_factors, q_dst), jk = rope_job(k, pos, freq_factors, k_dst);
    jk.st.base = (char *) k_cache->data;
    jk.st.idx = (const char *) k_idx->data; jk.st.idx_nb0 = (int64_t) k_idx->nb[0];
    jk.st.row_nb = (int64_t) k_cache->nb[1]; jk.st.n_rows = k_cache->ne[1];
    jk.st.f16 = k_cache->type == B200_TYPE_F16; jk.st.i64 = k_idx->type == B200_TYPE_I64;
    OPS_PROLOGUE(ctx);
    const int64_t n_tok = q->ne[2];
    if (n_tok >= 8 && q->ne[3] == 1 && k->ne[2] == n_tok && v->ne[1] == n_tok && v->ne[2] == 1 && v->ne[3] == 1 && rp->n_dims <= 512) {
        // a batch: one CTA per token, the cos / sin of the token's position computed once for all its heads
        CU(b200_launch(ctx, attn_prep_tok_kernel, dim3((unsigned) n_tok), dim3(256), 0, jq, jk, set_rows_job(v, v_idx, v_cache), rope_consts(rp)));
        ctx->launches++;
        return B200_OK;
    }
    CU(b200_launch(ctx, attn_prep_kernel, dim3((unsigned) (nq + nk + nv)), dim3(256), 0, jq, jk, set_rows_job(v, v_idx, v_cache), rope_consts(rp), (int) nq, (int) nk));
    ctx->launches++;
    return B200_OK;
}

int b200_cpy_supported(const b200_tensor * src, const b200_tensor * dst) {
    if (!src || !dst || !nonneg(src) || !nonneg(dst) || nelem(src) != nelem(dst)) return 0;
    const bool s_ok = src->type == B200_TYPE_F32 || src->type == B200_TYPE_F16, d_ok = dst->type == B200_TYPE_F32 || dst->type == B200_TYPE_F16;
    return s_ok && d_ok;
}

b200_status b200_cpy(b200_ctx * ctx, const b200_tensor * src, const b200_tensor * dst) {
    if (!b200_cpy_supported(src, dst)) return fail(B200_ERR_UNSUPPORTED, "CPY: F32 / F16 to F32 / F16 with the same number of elements");
    const int64_t n = nelem(src);
    if (n == 0) return B200_OK;
    if (!src->data || !dst->data) return fail(B200_ERR_ARG, "null argument");
    OPS_PROLOGUE(ctx);
    const bool sf = src->type == B200_TYPE_F32, df = dst->type == B200_TYPE_F32;
    const dim3 g((unsigned) ((n + 255) / 256)), b(256);
    const char * s = (const char *) src->data;
    char * d = (char *) dst->data;
    if (sf && df)       CU(b200_launch(ctx, cpy_kernel<float, float>, g, b, 0, s, t4(src), d, t4(dst), n));
    else if (sf)        CU(b200_launch(ctx, cpy_kernel<float, __half>, g, b, 0, s, t4(src), d, t4(dst), n));
    else if (df)        CU(b200_launch(ctx, cpy_kernel<__half, float>, g, b, 0, s, t4(src), d, t4(dst), n));
    else                CU(b200_launch(ctx, cpy_kernel<__half, __half>, g, b, 0, s, t4(src), d, t4(dst), n));
    ctx->launches++;
    return B200_OK;
}

int b200_flash_attn_ext_supported(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask, const b200_tensor * sinks,
                                  const b200_tensor * dst) {
    if (!q || !k || !v || !dst) return 0;
    if (q->type != B200_TYPE_F32 || k->type != B200_TYPE_F16 || v->type != B200_TYPE_F16 || dst->type != B200_TYPE_F32) return 0;
    if (!rows_ok(q, 4) || !rows_ok(k, 2) || !rows_ok(v, 2) || !rows_ok(dst, 4) || !nonneg(q) || !nonneg(k)) return 0;
    const int64_t D = q->ne[0];
    if (!(D == 64 || D == 128 || D == 256) || k->ne[0] != D || v->ne[0] != D || dst->ne[0] != D) return 0;
    if (k->ne[1] != v->ne[1] || k->ne[2] <= 0 || k->ne[3] <= 0 || v->ne[2] <= 0 || v->ne[3] <= 0) return 0;
    if (q->ne[2] % k->ne[2] || q->ne[3] % k->ne[3] || q->ne[2] % v->ne[2] || q->ne[3] % v->ne[3]) return 0;
    if (dst->ne[1] != q->ne[2] || dst->ne[2] != q->ne[1] || dst->ne[3] != q->ne[3]) return 0;
    const size_t al = (size_t) D / 16;   // bytes one lane reads of a K / V row
    for (int i = 1; i < 4; i++) if (k->nb[i] % al || v->nb[i] % al || q->nb[i] % 4) return 0;
    if (dst->nb[1] != (size_t) D * 4 || dst->nb[2] != dst->nb[1] * dst->ne[1] || dst->nb[3] != dst->nb[2] * dst->ne[2]) return 0;
    if (mask) {
        if (mask->type != B200_TYPE_F16 || mask->nb[0] != 2 || mask->ne[0] < k->ne[1] || mask->ne[1] < q->ne[1] || mask->ne[2] <= 0 || mask->ne[3] <= 0) return 0;
    }
    if (sinks && (sinks->type != B200_TYPE_F32 || sinks->nb[0] != 4 || sinks->ne[0] < q->ne[2])) return 0;
    return q->ne[1] * q->ne[2] * q->ne[3] < (1ll << 31);
}

constexpr size_t FA_CNT_BYTES = 4096;   // arrival counters at the head of the split workspace (splits only run with < 2 x SM-count rows)

static b200_status fa_run(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask,
                          const b200_tensor * sinks, const b200_tensor * dst, float scale, float max_bias, float logit_softcap, const FaDec * dec,
                          const b200_tensor * written_idx = nullptr) {
    if (!b200_flash_attn_ext_supported(q, k, v, mask, sinks, dst))
        return fail(B200_ERR_UNSUPPORTED, "FLASH_ATTN_EXT: F32 q, F16 k / v of head size 64 / 128 / 256, F16 mask");
    if (!ctx) return fail(B200_ERR_ARG, "null context");
    const int64_t rows = q->ne[1] * q->ne[2] * q->ne[3];
    if (rows == 0) return B200_OK;
    if ((!dec && !q->data) || !dst->data || (k->ne[1] > 0 && (!k->data || !v->data)) || (mask && !mask->data) || (sinks && !sinks->data)) return fail(B200_ERR_ARG, "null argument");
    const int D = (int) q->ne[0];
    if ((uintptr_t) k->data % (D / 16) || (uintptr_t) v->data % (D / 16)) return fail(B200_ERR_UNSUPPORTED, "FLASH_ATTN_EXT: K / V rows must be %d-byte aligned", D / 16);
    FaP p;
    p.q = t4(q); p.k = t4(k); p.v = t4(v); p.d = t4(dst);
    p.m = mask ? t4(mask) : T4{};
    p.qd = (const char *) q->data; p.kd = (const char *) k->data; p.vd = (const char *) v->data; p.md = mask ? (const char *) mask->data : nullptr;
    p.sinks = sinks ? (const float *) sinks->data : nullptr;
    p.dd = (char *) dst->data;
    p.scale = logit_softcap != 0.0f ? scale / logit_softcap : scale;
    p.softcap = logit_softcap; p.max_bias = max_bias;
    const uint32_t n_head = (uint32_t) q->ne[2];
    p.n_head_log2 = 1u << (uint32_t) floorf(log2f((float) n_head));
    p.m0 = powf(2.0f, -max_bias / (float) p.n_head_log2);
    p.m1 = powf(2.0f, -(max_bias / 2.0f) / (float) p.n_head_log2);
    // KV splits (decode: a few dozen rows, one per head): about two CTAs per SM, whole rounds (what the 8 warps of a CTA have in flight
    // at a time: one memory round trip); the splits of a row are merged by the last of its CTAs to finish
    int nsplit = 1;
    static const int u_env = [] { const char * e = getenv("GGML_B200_FA_U"); return e ? atoi(e) : 0; }();   // experiments: 4 / 8 / 16 positions per warp in flight
    const int U = !dec && (u_env == 4 || u_env == 8 || u_env == 16) && D <= 128 ? u_env : fa_u(D / 32);
    static const int w_env = [] { const char * e = getenv("GGML_B200_FA_WARPS"); return e ? atoi(e) : 0; }();
    const int FA_WARPS = w_env == 16 ? 16 : 8;
    const int64_t n_kv = k->ne[1], round = FA_WARPS * U;   // positions one CTA has in flight at a time
    if (rows < 2 * ctx->sm_count && rows <= (int64_t) (FA_CNT_BYTES / 4) && n_kv >= 2 * round) {
        nsplit = (int) std::min<int64_t>(std::min<int64_t>((2 * (int64_t) ctx->sm_count + rows - 1) / rows, (n_kv + round - 1) / round), 64);
        if (nsplit < 1) nsplit = 1;
    }
    static const int max_split = [] { const char * e = getenv("GGML_B200_FA_MAX_SPLIT"); return e ? atoi(e) : 64; }();   // experiments
    if (nsplit > max_split) nsplit = std::max(1, max_split);
    static const bool cluster_on = [] { const char * e = getenv("GGML_B200_FA_CLUSTER"); return !(e && atoi(e) == 0); }();
    if (cluster_on && nsplit > 8) nsplit = 8;   // the splits of a row form one thread-block cluster (portable size: 8)
    p.chunk  = nsplit == 1 ? std::max<int64_t>(n_kv, 1) : ((n_kv + nsplit - 1) / nsplit + round - 1) / round * round;
    if (nsplit > 1) nsplit = (int) ((n_kv + p.chunk - 1) / p.chunk);   // no empty splits
    p.nsplit  = nsplit;
    p.cluster = cluster_on && nsplit > 1 && nsplit <= 8;
    p.ws      = nullptr;
    p.cnt     = nullptr;
    OPS_PROLOGUE(ctx);
    if (nsplit > 1 && !p.cluster) {
        const void * before = ctx->fa_ws;
        if (b200_status st = b200_ensure(ctx, &ctx->fa_ws, &ctx->fa_ws_size, FA_CNT_BYTES + (size_t) rows * nsplit * (D + 2) * sizeof(float))) return st;
        if (ctx->fa_ws != before) CU(cudaMemsetAsync(ctx->fa_ws, 0, FA_CNT_BYTES, ctx->stream));   // (re)allocated: the counters start at zero
        p.cnt = (unsigned *) ctx->fa_ws;
        p.ws  = (float *) ((char *) ctx->fa_ws + FA_CNT_BYTES);
    }
    p.trace = nullptr;
    // opt-in (GGML_B200_FA_EARLY=1): measured no faster at 512 cached positions and 10 % slower at 4096 - memory operations issued ahead of
    // griddepcontrol.wait delay the wait itself (device timeline: the dependency resolves 0.7 us after the previous kernel without
    // them, 0.7 - 3.5 us with them)
    static const bool early_on = [] { const char * e = getenv("GGML_B200_FA_EARLY"); return e && atoi(e) != 0; }();
    const bool early = early_on && written_idx && written_idx->data && written_idx->ne[0] == 1 && q->ne[1] == 1 &&
                       (written_idx->type == B200_TYPE_I64 || written_idx->type == B200_TYPE_I32);
    p.early_idx = early ? (const char *) written_idx->data : nullptr;
    p.early_i64 = early && written_idx->type == B200_TYPE_I64;
    memset(&p.dec, 0, sizeof(p.dec));
    if (dec) p.dec = *dec;
    memset(&p.next, 0, sizeof(p.next));
    if (ctx->fa_prefetch_kb > 0 && ctx->prefetch_kb > 0 && ctx->last_key && ctx->mem_gen == b200_mem_generation()) {
        // the mat-vec launch behind this kernel has been staged by the one in front of it; this kernel's idle HBM time goes to the one after
        auto it1 = ctx->next_of.find(ctx->last_key);
        if (it1 != ctx->next_of.end() && it1->second.n_mats > 0) {
            auto it2 = ctx->next_of.find(it1->second.w[0]);
            if (it2 != ctx->next_of.end() && it2->second.n_mats > 0) {
                p.next = it2->second;
                const int64_t gb = (int64_t) p.next.rpg * p.next.row_bytes * (p.next.glu ? 2 : 1);
                p.next.groups_per_cta = (int) (((int64_t) ctx->fa_prefetch_kb * 1024 + gb - 1) / gb);
            }
        }
    }
    if (dec) {   // one token: the decode kernel with the ROPEs and the cache rows folded in (8 positions per warp in flight)
        const dim3 g((unsigned) rows, (unsigned) nsplit), b(FA_WARPS * 32);
        if (D == 64)       CU(fa_launch(ctx, fa_kernel<2, 8, true>, g, b, p));
        else if (D == 128) CU(fa_launch(ctx, fa_kernel<4, 8, true>, g, b, p));
        else               CU(fa_launch(ctx, fa_kernel<8, 8, true>, g, b, p));
        ctx->launches++;
        return B200_OK;
    }
    static const int tile_min = [] { const char * e = getenv("GGML_B200_FA_TILE_MIN"); return e ? atoi(e) : 16; }();   // 0: never
    if (tile_min > 0 && q->ne[1] >= tile_min && D <= 128 && q->ne[2] <= 65535 && q->ne[3] <= 65535 && (uintptr_t) k->data % 16 == 0 &&
        (uintptr_t) v->data % 16 == 0 && k->nb[1] % 16 == 0 && k->nb[2] % 16 == 0 && k->nb[3] % 16 == 0 && v->nb[1] % 16 == 0 && v->nb[2] % 16 == 0 &&
        v->nb[3] % 16 == 0 && (uintptr_t) dst->data % 8 == 0) {
        // a prompt batch on the tensor cores: 64 query rows of a head per CTA
        p.nsplit = 1;
        p.chunk  = std::max<int64_t>(n_kv, 1);
        const dim3 g((unsigned) ((q->ne[1] + FT_BQ - 1) / FT_BQ), (unsigned) q->ne[2], (unsigned) q->ne[3]), b(FT_WARPS * 32);
        if (D == 64) CU(b200_launch(ctx, fa_tile_kernel<64>, g, b, 0, p));
        else         CU(b200_launch(ctx, fa_tile_kernel<128>, g, b, 0, p));
        ctx->launches++;
        return B200_OK;
    }
    if (q->ne[1] >= 4) {   // a batch: several query rows per CTA
        const int NQ = D == 256 ? 4 : 8;
        const int64_t blocks = (q->ne[1] + NQ - 1) / NQ * q->ne[2] * q->ne[3];
        p.nsplit = 1;
        p.chunk  = std::max<int64_t>(n_kv, 1);
        const dim3 g((unsigned) blocks), b(FAQ_WARPS * 32);
        if (D == 64)       CU(b200_launch(ctx, fa_rows_kernel<2, 8>, g, b, 0, p));
        else if (D == 128) CU(b200_launch(ctx, fa_rows_kernel<4, 8>, g, b, 0, p));
        else               CU(b200_launch(ctx, fa_rows_kernel<8, 4>, g, b, 0, p));
        ctx->launches++;
        return B200_OK;
    }
    if (ctx->trace && ctx->sm_count >= 148) p.trace = ctx->trace + (size_t) (ctx->launches % ctx->trace_slots) * ctx->sm_count * 8;
    const dim3 grid((unsigned) rows, (unsigned) nsplit), block(FA_WARPS * 32);
    if (D == 64)       CU(U == 4 ? fa_launch(ctx, fa_kernel<2, 4>, grid, block, p) : U == 8 ? fa_launch(ctx, fa_kernel<2, 8>, grid, block, p) : fa_launch(ctx, fa_kernel<2, 16>, grid, block, p));
    else if (D == 128) CU(U == 4 ? fa_launch(ctx, fa_kernel<4, 4>, grid, block, p) : U == 8 ? fa_launch(ctx, fa_kernel<4, 8>, grid, block, p) : fa_launch(ctx, fa_kernel<4, 16>, grid, block, p));
    else               CU(fa_launch(ctx, fa_kernel<8, 8>, grid, block, p));
    ctx->launches++;
    return B200_OK;
}

b200_status b200_flash_attn_ext(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask,
                                const b200_tensor * sinks, const b200_tensor * dst, float scale, float max_bias, float logit_softcap) {
    return fa_run(ctx, q, k, v, mask, sinks, dst, scale, max_bias, logit_softcap, nullptr);
}

b200_status b200_flash_attn_ext_decode(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask,
                                       const b200_tensor * sinks, const b200_tensor * dst, float scale, float max_bias, float logit_softcap,
                                       const b200_tensor * written_idx) {
    return fa_run(ctx, q, k, v, mask, sinks, dst, scale, max_bias, logit_softcap, nullptr, written_idx);
}

// One token: ROPE(q), ROPE(k), SET_ROWS(k cache), SET_ROWS(v cache) and FLASH_ATTN_EXT of a layer as ONE launch. q / k / v are the
// projections ([D, n_head, 1], [D, n_head_kv, 1], [D * n_head_kv, 1], contiguous F32); q_dst / k_dst the ROPE nodes' outputs; fa_q the
// FLASH_ATTN_EXT node's q operand (a view of q_dst: shape only), fa_k / fa_v its views of the caches, which must address the row
// k_idx / v_idx of k_cache / v_cache exactly where SET_ROWS puts it.
int b200_attn_decode_supported(const b200_tensor * q, const b200_tensor * q_dst, const b200_tensor * k, const b200_tensor * k_dst, const b200_tensor * pos,
                               const b200_tensor * freq_factors, const b200_rope_params * rp, const b200_tensor * k_idx, const b200_tensor * k_cache,
                               const b200_tensor * v, const b200_tensor * v_idx, const b200_tensor * v_cache, const b200_tensor * fa_q,
                               const b200_tensor * fa_k, const b200_tensor * fa_v, const b200_tensor * mask, const b200_tensor * sinks, const b200_tensor * dst) {
    if (!b200_attn_prep_supported(q, q_dst, k, k_dst, pos, freq_factors, rp, k_idx, k_cache, v, v_idx, v_cache)) return 0;
    if (!b200_flash_attn_ext_supported(fa_q, fa_k, fa_v, mask, sinks, dst)) return 0;
    const int64_t D = q->ne[0], nh = q->ne[1], nkv = k->ne[1];
    if (q->ne[2] != 1 || q->ne[3] != 1 || k->ne[2] != 1 || rp->n_dims != D || fa_q->ne[1] != 1 || fa_q->ne[2] != nh || fa_q->ne[3] != 1) return 0;
    if (fa_k->ne[2] != nkv || fa_v->ne[2] != nkv || fa_k->ne[3] != 1 || fa_v->ne[3] != 1 || v->ne[0] != D * nkv || v->ne[1] != 1) return 0;
    for (const b200_tensor * t : { q, q_dst, k, k_dst }) if (t->nb[1] != (size_t) D * 4) return 0;   // heads back to back
    if (k_cache->type != B200_TYPE_F16 || v_cache->type != B200_TYPE_F16 || k_cache->ne[0] != D * nkv || v_cache->ne[0] != D * nkv) return 0;
    // the FLASH_ATTN_EXT views see cache row r, head h at base + r * nb1 + h * D * 2: the address SET_ROWS writes
    if (fa_k->data != k_cache->data || fa_k->nb[1] != k_cache->nb[1] || fa_k->nb[2] != (size_t) D * 2 || fa_k->ne[1] > k_cache->ne[1]) return 0;
    if (fa_v->data != v_cache->data || fa_v->nb[1] != v_cache->nb[1] || fa_v->nb[2] != (size_t) D * 2 || fa_v->ne[1] > v_cache->ne[1]) return 0;
    if (k_idx->ne[0] != 1 || v_idx->ne[0] != 1) return 0;
    return 1;
}

b200_status b200_attn_decode(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * q_dst, const b200_tensor * k, const b200_tensor * k_dst,
                             const b200_tensor * pos, const b200_tensor * freq_factors, const b200_rope_params * rp, const b200_tensor * k_idx,
                             const b200_tensor * k_cache, const b200_tensor * v, const b200_tensor * v_idx, const b200_tensor * v_cache,
                             const b200_tensor * fa_q, const b200_tensor * fa_k, const b200_tensor * fa_v, const b200_tensor * mask,
                             const b200_tensor * sinks, const b200_tensor * dst, float scale, float max_bias, float logit_softcap) {
    if (!b200_attn_decode_supported(q, q_dst, k, k_dst, pos, freq_factors, rp, k_idx, k_cache, v, v_idx, v_cache, fa_q, fa_k, fa_v, mask, sinks, dst))
        return fail(B200_ERR_UNSUPPORTED, "attn_decode: one token, rope over the whole head, F16 caches seen by FLASH_ATTN_EXT where SET_ROWS writes them");
    for (const b200_tensor * t : { q, q_dst, k, k_dst, pos, k_idx, k_cache, v, v_idx, v_cache }) if (!t->data) return fail(B200_ERR_ARG, "null argument");
    if (freq_factors && !freq_factors->data) return fail(B200_ERR_ARG, "null argument");
    FaDec dc;
    memset(&dc, 0, sizeof(dc));
    dc.on = 1;
    dc.q_raw = (const float *) q->data; dc.k_raw = (const float *) k->data; dc.v_raw = (const float *) v->data;
    dc.pos = (const int32_t *) pos->data;
    dc.ff = freq_factors ? (const float *) freq_factors->data : nullptr;
    dc.rp = rope_consts(rp);
    dc.k_idx = (const char *) k_idx->data; dc.v_idx = (const char *) v_idx->data;
    dc.k_i64 = k_idx->type == B200_TYPE_I64; dc.v_i64 = v_idx->type == B200_TYPE_I64;
    dc.kc = (char *) k_cache->data; dc.vc = (char *) v_cache->data;
    dc.kc_nb1 = (int64_t) k_cache->nb[1]; dc.vc_nb1 = (int64_t) v_cache->nb[1];
    dc.kc_rows = k_cache->ne[1]; dc.vc_rows = v_cache->ne[1];
    dc.q_dst = (float *) q_dst->data; dc.k_dst = (float *) k_dst->data;
    return fa_run(ctx, fa_q, fa_k, fa_v, mask, sinks, dst, scale, max_bias, logit_softcap, &dc);
}

} // extern "C"

// mmid.cu — K5/K6: GGML_OP_MUL_MAT_ID (MoE expert mat-vec / mat-mul).
//
// Replaces ggml_cuda_mul_mat_id (/root/reference/ggml/src/ggml-cuda/ggml-cuda.cu:2263-2409) and mm_ids_helper
// (/root/reference/ggml/src/ggml-cuda/mmid.cu:26-164). Semantics follow ggml-cpu (ggml-cpu.c:1476-1649):
//   dst[:, i, t] = as[ids[i, t]] . b[:, i % ne11, t]      for every slot i < n_used and token t < n_tok.
// There is no host synchronisation anywhere on this path (the reference's fallback copies ids to the host,
// ggml-cuda.cu:2322-2345): the expert id of a (slot, token) pair is read on the device by the CTAs that work on it.
//
// Routing of the work:
//   * every b row is quantized once (K1, unit records) — ne11*n_tok rows, not n_used*n_tok;
//   * each (slot, token) pair is one mat-vec of the TMA-staged kernel (mmvq.cu) with blockIdx.y = pair; the CTAs of a
//     pair split the expert's rows. With n_pairs >= #SMs a pair gets one CTA.
//   * b200_mm_ids_sort (counting sort by expert, stable) is the building block of the grouped tcgen05 path.

#include "internal.h"

#include <vector>

using namespace b200;

#define fail b200_fail

// ------------------------------------------------------------------------------------------------ K5: sort pairs by expert

// one CTA; warp w handles experts w, w+32, ...; pairs are scanned in order so the output is stable in p = t*n_used + i
__global__ void __launch_bounds__(1024) mm_ids_sort_kernel(const int32_t * __restrict__ ids, int n_used, int n_tok, int64_t ids_rs, int n_as,
                                                           int32_t * __restrict__ sorted_pairs, int32_t * __restrict__ expert_bounds) {
    extern __shared__ int counts[];  // [n_as + 1]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, n_warps = blockDim.x >> 5;
    const int n_pairs = n_used * n_tok;
    for (int e = warp; e < n_as; e += n_warps) {
        int cnt = 0;
        for (int p0 = 0; p0 < n_pairs; p0 += 32) {
            const int p = p0 + lane;
            bool hit = false;
            if (p < n_pairs) {
                const int t = p / n_used, i = p - t * n_used;
                hit = ids[(int64_t) t * ids_rs + i] == e;
            }
            cnt += __popc(__ballot_sync(0xffffffffu, hit));
        }
        if (lane == 0) counts[e] = cnt;
    }
    __syncthreads();
    if (threadIdx.x == 0) {  // exclusive scan; n_as is at most a few hundred
        int run = 0;
        for (int e = 0; e < n_as; e++) {
            const int c = counts[e];
            counts[e] = run;
            expert_bounds[e] = run;
            run += c;
        }
        counts[n_as] = run;
        expert_bounds[n_as] = run;
    }
    __syncthreads();
    for (int e = warp; e < n_as; e += n_warps) {
        int pos = counts[e];
        for (int p0 = 0; p0 < n_pairs; p0 += 32) {
            const int p = p0 + lane;
            bool hit = false;
            if (p < n_pairs) {
                const int t = p / n_used, i = p - t * n_used;
                hit = ids[(int64_t) t * ids_rs + i] == e;
            }
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (hit) sorted_pairs[pos + __popc(m & ((1u << lane) - 1))] = p;
            pos += __popc(m);
        }
    }
}

b200_status b200_mm_ids_sort(b200_ctx * ctx, const int32_t * ids, int64_t n_used, int64_t n_tok, size_t ids_nb1, int64_t n_as,
                             int32_t * sorted_pairs, int32_t * expert_bounds) {
    if (!ctx || !ids || !sorted_pairs || !expert_bounds) return fail(B200_ERR_ARG, "null argument");
    if (n_as <= 0 || n_as > 8192 || ids_nb1 % 4 != 0 || n_used * n_tok > INT32_MAX) return fail(B200_ERR_UNSUPPORTED, "bad MUL_MAT_ID index geometry");
    CU(cudaSetDevice(ctx->device));
    if (b200_status fst = b200_split_flush(ctx)) return fst;
    mm_ids_sort_kernel<<<1, 1024, (size_t) (n_as + 1) * sizeof(int), ctx->stream>>>(ids, (int) n_used, (int) n_tok, (int64_t) (ids_nb1 / 4), (int) n_as,
                                                                                  sorted_pairs, expert_bounds);
    CU(cudaGetLastError());
    ctx->launches++;
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------ K6: per-expert tile GEMMs
//
// Enough pairs per expert (prefill): what the reference does with ids helper -> gathered quantize -> mul_mat_q per expert window ->
// un-permuting store (ggml-cuda/mmq.cu:148-206, mmq.cuh:3425-3455, 3006-3053). Here: device-side sort, ONE small read-back of the
// expert bounds (the launches are sized on the host), the pairs' activation rows gathered expert by expert into a contiguous
// buffer, one MUL_MAT per expert that has pairs - through the ordinary dispatcher, so every expert gets the kernel its column count
// calls for (tcgen05 GEMM, tcgen05 mat-vec, mat-vec) and its weights are read ONCE - and a scatter of the result rows to their
// (slot, token) places in dst.

__global__ void __launch_bounds__(256) moe_gather_rows_kernel(const float * __restrict__ b, int64_t x_is, int64_t x_ts, int n_used, int ne11,
                                                              const int32_t * __restrict__ sp, float * __restrict__ xg, int64_t K) {
    const int pr = sp[blockIdx.x];
    const int t = pr / n_used, i = pr - t * n_used;
    const float * src = (const float *) ((const char *) b + (int64_t) t * x_ts + (int64_t) (i % ne11) * x_is);
    float * dst = xg + (int64_t) blockIdx.x * K;
    if (((uintptr_t) src % 16) == 0 && K % 4 == 0) {
        for (int64_t k = threadIdx.x; k < K / 4; k += blockDim.x) ((float4 *) dst)[k] = __ldg((const float4 *) src + k);
    } else {
        for (int64_t k = threadIdx.x; k < K; k += blockDim.x) dst[k] = __ldg(src + k);
    }
}

__global__ void __launch_bounds__(256) moe_scatter_rows_kernel(const float * __restrict__ yg, const int32_t * __restrict__ sp, int n_used,
                                                               int64_t dst_is, int64_t dst_ts, float * __restrict__ dst, int64_t M) {
    const int pr = sp[blockIdx.x];
    const int t = pr / n_used, i = pr - t * n_used;
    const float * src = yg + (int64_t) blockIdx.x * M;
    float * out = dst + (int64_t) t * dst_ts + (int64_t) i * dst_is;
    for (int64_t m = threadIdx.x; m < M; m += blockDim.x) out[m] = src[m];
}

static b200_status mul_mat_id_tiled(b200_ctx * ctx, const b200_mul_mat_id_desc * d, const void * as, const float * b, const int32_t * ids, float * dst,
                                    const MvIds & mi) {
    const int64_t K = d->ne00, M = d->ne01, n_pairs = d->n_used * d->n_tok;
    const size_t sp_b = (size_t) ((n_pairs + 63) & ~(int64_t) 63) * 4, eb_b = (size_t) ((d->n_as + 1 + 63) & ~(int64_t) 63) * 4;
    const size_t xg_b = (size_t) n_pairs * K * 4, yg_b = (size_t) n_pairs * M * 4;
    if (b200_status st = b200_ensure(ctx, &ctx->io, &ctx->io_size, sp_b + eb_b + xg_b + yg_b + 256)) return st;
    int32_t * sp = (int32_t *) ctx->io;
    int32_t * eb = (int32_t *) ((uint8_t *) ctx->io + sp_b);
    float *   xg = (float *) ((uint8_t *) ctx->io + sp_b + eb_b);
    float *   yg = (float *) ((uint8_t *) xg + xg_b);
    if (b200_status st = b200_mm_ids_sort(ctx, ids, d->n_used, d->n_tok, d->ids_nb1, d->n_as, sp, eb)) return st;
    std::vector<int32_t> bounds((size_t) d->n_as + 1);
    CU(cudaMemcpyAsync(bounds.data(), eb, bounds.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
    moe_gather_rows_kernel<<<(unsigned) n_pairs, 256, 0, ctx->stream>>>(b, mi.x_is, mi.x_ts, mi.n_used, mi.ne11, sp, xg, K);
    CU(cudaGetLastError());
    ctx->launches++;
    CU(cudaStreamSynchronize(ctx->stream));   // the expert bounds size the launches
    if (bounds[(size_t) d->n_as] != (int32_t) n_pairs) return fail(B200_ERR_ARG, "MUL_MAT_ID: expert ids out of range");  // ggml-cpu asserts (ggml-cpu.c:1574)
    for (int64_t e = 0; e < d->n_as; e++) {
        const int64_t n0 = bounds[(size_t) e], cnt = bounds[(size_t) e + 1] - n0;
        if (cnt <= 0) continue;
        b200_mul_mat_desc md;
        memset(&md, 0, sizeof(md));
        md.type = d->type;
        md.ne00 = K; md.ne01 = M; md.ne02 = 1; md.ne03 = 1;
        md.nb01 = d->nb01; md.nb02 = d->nb02; md.nb03 = d->nb02;
        md.ne11 = cnt; md.ne12 = 1; md.ne13 = 1;
        md.nb11 = (size_t) K * 4; md.nb12 = md.nb11 * (size_t) cnt; md.nb13 = md.nb12;
        md.nb1 = (size_t) M * 4; md.nb2 = md.nb1 * (size_t) cnt; md.nb3 = md.nb2;
        if (b200_status st = b200_mul_mat(ctx, &md, (const uint8_t *) as + e * d->nb02, xg + n0 * K, yg + n0 * M)) return st;
    }
    moe_scatter_rows_kernel<<<(unsigned) n_pairs, 256, 0, ctx->stream>>>(yg, sp, mi.n_used, mi.dst_is, mi.dst_ts, dst, M);
    CU(cudaGetLastError());
    ctx->launches++;
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------ K6: dispatch

int b200_mul_mat_id_supported(const b200_mul_mat_id_desc * d) {
    if (!d) return 0;
    if (d->type != B200_TYPE_Q4_0 && d->type != B200_TYPE_Q8_0 && d->type != B200_TYPE_Q4_K && d->type != B200_TYPE_Q5_K && d->type != B200_TYPE_Q6_K) return 0;
    if (d->ne00 <= 0 || d->ne00 % type_block_elems(d->type) != 0) return 0;
    if (d->ne01 < 0 || d->n_as <= 0 || d->ne11 <= 0 || d->n_tok < 0 || d->n_used <= 0) return 0;
    if (d->nb01 < (size_t) (d->ne00 / type_block_elems(d->type)) * type_block_bytes(d->type)) return 0;
    if (d->nb11 % 4 || d->nb12 % 4 || d->nb1 % 4 || d->nb2 % 4 || d->ids_nb1 % 4) return 0;
    if (d->ne01 > INT32_MAX || d->ne00 > INT32_MAX || d->n_used * d->n_tok > INT32_MAX) return 0;
    return 1;
}

static bool mmid_fast_layout(const b200_ctx * ctx, const b200_mul_mat_id_desc * d, const void * as) {
    const int64_t K = d->ne00;
    return ctx->force_path != B200_PATH_SIMPLE && K % 256 == 0 && (uintptr_t) as % 16 == 0 && d->nb01 % 16 == 0 && d->nb02 % 16 == 0 &&
           (d->type != B200_TYPE_Q6_K || K % 2048 == 0);  // 210-byte units: bulk-copyable 8 at a time
}

// fz != nullptr: K7 - one token, the epilogue of b200_fusion on the TMA-staged kernel (gate_as != nullptr iff a GLU is fused)
static b200_status mul_mat_id_impl(b200_ctx * ctx, const b200_mul_mat_id_desc * d, const void * as, const void * gate_as, const float * b,
                                   const int32_t * ids, float * dst, const b200_fusion * fz) {
    if (!ctx || !d || !as || !b || !ids || !dst) return fail(B200_ERR_ARG, "null argument");
    if (!b200_mul_mat_id_supported(d)) return fail(B200_ERR_UNSUPPORTED, "MUL_MAT_ID geometry/type not supported");
    const int64_t K = d->ne00, M = d->ne01;
    const int64_t n_pairs = d->n_used * d->n_tok;
    if (n_pairs == 0 || M == 0) return B200_OK;
    CU(cudaSetDevice(ctx->device));
    if (b200_status fst = b200_split_flush(ctx)) return fst;

    MvIds mi;
    mi.ids    = ids;
    mi.ids_rs = (int64_t) (d->ids_nb1 / 4);
    mi.nb02   = (int64_t) d->nb02;
    mi.dst_is = (int64_t) (d->nb1 / 4);
    mi.dst_ts = (int64_t) (d->nb2 / 4);
    mi.n_used = (int) d->n_used;
    mi.ne11   = (int) d->ne11;
    mi.n_as   = (int) d->n_as;

    const bool fast = mmid_fast_layout(ctx, d, as) && (!gate_as || mmid_fast_layout(ctx, d, gate_as));
    const int64_t n_rows = d->ne11 * d->n_tok;
    if (fz && (!fast || d->n_tok != 1 || ctx->split.n_ranks > 1)) return fail(B200_ERR_UNSUPPORTED, "fused MUL_MAT_ID: one token on the TMA-staged kernel only");

    mi.x_is = (int64_t) d->nb11;
    mi.x_ts = (int64_t) d->nb12;
    // prefill-sized batches: one tile GEMM per expert (no host sync is possible while the stream is being captured: then the paths below)
    if (!fz && ctx->moe_tile_min > 0 && n_pairs >= (int64_t) ctx->moe_tile_min * d->n_as && n_pairs <= 65535 && ctx->force_path == B200_PATH_AUTO &&
        ctx->split.n_ranks <= 1 && !ctx->exact) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing(ctx->stream, &cs);
        if (cs == cudaStreamCaptureStatusNone) return mul_mat_id_tiled(ctx, d, as, b, ids, dst, mi);
    }
    if (fast) {
        const int  U    = (int) (K / 256);
        // enough pairs per expert: sort the pairs by expert and let every expert multiply its tokens 8 columns at a time on the
        // tensor-core mat-vec kernel (mvmma.cu) - an expert's weights are then read once per 8 tokens instead of once per token
        const bool grouped = !fz && ctx->moe_group_min > 0 && n_pairs >= (int64_t) ctx->moe_group_min * d->n_as && n_pairs >= 8 && mv_mma_usable(d->type) && mv_mma_fits(d->type, K) &&
                             d->nb12 == d->nb11 * (size_t) d->ne11;
        const bool fuse = !grouped && ctx->fuse_quant && U <= 64;  // every pair's CTAs quantize that pair's activation row themselves
        uint8_t * am = nullptr, * ax = nullptr;
        if (!fuse) {
            const size_t main_b = (size_t) n_rows * U * ACT_MAIN, aux_b = (size_t) n_rows * U * ACT_AUX;
            const size_t aux_o  = (main_b + 255) & ~(size_t) 255;
            b200_status  st     = b200_ensure(ctx, &ctx->scratch, &ctx->scratch_size, aux_o + aux_b);
            if (st != B200_OK) return st;
            am = (uint8_t *) ctx->scratch;
            ax = am + aux_o;
            if (d->nb12 == d->nb11 * (size_t) d->ne11) {  // rows of all tokens are equally spaced: one launch
                int l = 0;
                CU(quantize_units(act_kind_for(d->type), b, K, n_rows, d->nb11, am, ax, ctx->stream, &l, ctx->pdl));
                ctx->launches += l;
            } else {
                for (int64_t t = 0; t < d->n_tok; t++) {
                    int l = 0;
                    CU(quantize_units(act_kind_for(d->type), (const float *) ((const char *) b + t * d->nb12), K, d->ne11, d->nb11,
                                      am + (size_t) t * d->ne11 * U * ACT_MAIN, ax + (size_t) t * d->ne11 * U * ACT_AUX, ctx->stream, &l, ctx->pdl));
                    ctx->launches += l;
                }
            }
        }
        mi.x_is = (int64_t) d->nb11;
        mi.x_ts = (int64_t) d->nb12;
        if (grouped) {
            b200_status st = b200_ensure(ctx, &ctx->scratch2, &ctx->scratch2_size, (size_t) (((n_pairs + 63) & ~(int64_t) 63) + d->n_as + 1) * 4 + 256);
            if (st != B200_OK) return st;
            int32_t * sp = (int32_t *) ctx->scratch2;
            int32_t * eb = sp + ((n_pairs + 63) & ~(int64_t) 63);
            st = b200_mm_ids_sort(ctx, ids, d->n_used, d->n_tok, d->ids_nb1, d->n_as, sp, eb);
            if (st != B200_OK) return st;
            return b200_mv_mma_run_ids(ctx, d->type, K, (const uint8_t *) as, d->nb01, d->nb02, M, d->n_as, dst, mi, sp, eb, am, ax);
        }
        MvParams p;
        memset(&p, 0, sizeof(p));
        p.U = U;
        if (U <= 32) { int L = 1; while (L < U) L <<= 1; p.L = L; p.rpg = 32 / L; p.slices = 1; }
        else { p.L = 32; p.rpg = 1; p.slices = (U + 31) / 32; }
        p.n_mats        = 1;
        p.mat[0].w      = (const uint8_t *) as;
        p.mat[0].nb01   = (int64_t) d->nb01;
        p.mat[0].dst    = dst;
        p.mat[0].dst_cs = 0;
        p.mat[0].M      = (int) M;
        p.mat[0].group0 = 0;
        for (int i = 1; i < MV_MAX_MATS; i++) p.mat[i].group0 = INT32_MAX;
        if (fz) {
            p.mat[0].bias    = fz->up_bias;
            p.mat[0].bias_es = (int64_t) (fz->bias_nb1 / 4);
            if (gate_as) {  // mat[0] = gate, mat[1] = up: same rows, one after the other (kernels.h MvEpi)
                p.mat[1]      = p.mat[0];
                p.mat[0].w    = (const uint8_t *) gate_as;
                p.mat[0].bias = fz->gate_bias;
                p.n_mats      = 2;
                p.epi.glu     = fz->glu_op + 1;
                p.epi.alpha   = fz->alpha;
                p.epi.limit   = fz->limit;
            }
        }
        p.total_groups = (int) ((M + p.rpg - 1) / p.rpg);
        p.pdl          = ctx->pdl;
        p.split        = ctx->split;
        p.fuse_quant   = fuse ? 1 : 0;
        p.x_vec        = (uintptr_t) b % 16 == 0 && d->nb11 % 16 == 0 && d->nb12 % 16 == 0;
        p.ids          = mi;
        // shared-memory configuration for one column
        static const int cand[][2] = { { 16, 2 }, { 8, 3 }, { 8, 2 }, { 4, 3 }, { 4, 2 }, { 2, 2 }, { 1, 2 } };
        size_t smem = 0;
        for (auto & c : cand) {
            smem = mv_fast_smem(d->type, U, 1, c[0], c[1]);
            if (smem <= 227 * 1024) { p.n_cons = c[0]; p.n_stages = c[1]; break; }
            smem = 0;
        }
        if (!smem) return fail(B200_ERR_UNSUPPORTED, "K=%lld: activations do not fit in shared memory", (long long) K);
        // grid.y carries the pairs; chunk by whole tokens so that a chunk's pair index restarts at (slot 0, token t0)
        const int64_t tok_chunk = 32768 / d->n_used > 0 ? 32768 / d->n_used : 1;
        if (d->n_used > 65535) return fail(B200_ERR_UNSUPPORTED, "n_used > 65535");
        for (int64_t t0 = 0; t0 < d->n_tok; t0 += tok_chunk) {
            const int64_t nt = d->n_tok - t0 < tok_chunk ? d->n_tok - t0 : tok_chunk;
            const int     np = (int) (nt * d->n_used);
            int ctas = (ctx->sm_count + np - 1) / np;
            if (ctas > p.total_groups) ctas = p.total_groups;
            if (ctas < 1) ctas = 1;
            MvParams q = p;
            q.ids.ids    = ids + t0 * mi.ids_rs;
            q.mat[0].dst = dst + t0 * mi.dst_ts;
            q.mat[1].dst = q.mat[0].dst;
            if (fuse) {
                q.x = (const float *) ((const char *) b + t0 * d->nb12);
            } else {
                q.act_main = am + (size_t) t0 * d->ne11 * U * ACT_MAIN;
                q.act_aux  = ax + (size_t) t0 * d->ne11 * U * ACT_AUX;
            }
            CU(mv_fast_launch(d->type, 1, q, ctas, np, smem, ctx->stream));
            ctx->launches++;
        }
        return B200_OK;
    }

    // generic path
    const int    val    = type_is_k(d->type) ? 1 : 0;
    const size_t act_rs = type_is_k(d->type) ? (size_t) (K / QK_K) * SZ_Q8_K : (size_t) (K / QK) * SZ_Q8_0;
    b200_status  st     = b200_ensure(ctx, &ctx->scratch, &ctx->scratch_size, (size_t) n_rows * act_rs + 16);
    if (st != B200_OK) return st;
    uint8_t * act = (uint8_t *) ctx->scratch;
    for (int64_t t = 0; t < d->n_tok; t++) {
        int l = 0;
        CU(quantize_ggml(val, (const float *) ((const char *) b + t * d->nb12), K, d->ne11, d->nb11, act + (size_t) t * d->ne11 * act_rs,
                         ctx->stream, &l));
        ctx->launches += l;
    }
    if (n_pairs > 65535) return fail(B200_ERR_UNSUPPORTED, "more than 65535 (slot, token) pairs on the generic MUL_MAT_ID path");
    MvSimpleParams sp;
    memset(&sp, 0, sizeof(sp));
    sp.w      = (const uint8_t *) as;
    sp.nb01   = (int64_t) d->nb01;
    sp.act    = act;
    sp.act_rs = (int64_t) act_rs;
    sp.dst    = dst;
    sp.dst_cs = 0;
    sp.M      = (int) M;
    sp.K      = (int) K;
    sp.ncols  = (int) n_pairs;
    sp.ids    = mi;
    sp.exact  = ctx->exact;
    CU(mv_simple_launch(d->type, sp, ctx->stream));
    ctx->launches++;
    return B200_OK;
}

b200_status b200_mul_mat_id(b200_ctx * ctx, const b200_mul_mat_id_desc * d, const void * as, const float * b, const int32_t * ids, float * dst) {
    return mul_mat_id_impl(ctx, d, as, nullptr, b, ids, dst, nullptr);
}

// ------------------------------------------------------------------------------------------------ K7: one token, fused epilogue

int b200_mul_mat_id_fused_supported(const b200_ctx * ctx, const b200_mul_mat_id_desc * d) {
    if (!ctx || !d || !b200_mul_mat_id_supported(d)) return 0;
    if (d->n_tok != 1 || ctx->split.n_ranks > 1 || ctx->force_path == B200_PATH_SIMPLE) return 0;
    if (d->ne00 % 256 != 0 || d->nb01 % 16 != 0 || d->nb02 % 16 != 0 || (d->type == B200_TYPE_Q6_K && d->ne00 % 2048 != 0)) return 0;
    return mv_fast_smem(d->type, (int) (d->ne00 / 256), 1, 1, 2) <= 227 * 1024;
}

b200_status b200_mul_mat_id_fused(b200_ctx * ctx, const b200_mul_mat_id_desc * d, const void * up_as, const void * gate_as, const float * b,
                                  const int32_t * ids, const b200_fusion * f, float * dst) {
    if (!f) return fail(B200_ERR_ARG, "null argument");
    const bool glu = f->glu_op != B200_GLU_NONE;
    if (glu ? (f->glu_op < 0 || f->glu_op > B200_GLU_GEGLU_QUICK || !gate_as) : (gate_as != nullptr || f->gate_bias != nullptr))
        return fail(B200_ERR_ARG, "fused MUL_MAT_ID: glu_op / gate / bias arguments do not agree");
    if (f->bias_nb1 % 4) return fail(B200_ERR_ARG, "fused MUL_MAT_ID: bias stride");
    return mul_mat_id_impl(ctx, d, up_as, gate_as, b, ids, dst, f);
}

// internal.h — context struct and error helpers shared by the C-ABI translation units (api.cu, mmid.cu, mmq*.cu).
#pragma once

#include "common.cuh"
#include "kernels.h"

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <unordered_map>
#include <utility>

b200_status b200_fail(b200_status st, const char * fmt, ...);

#define CU(call)                                                                                                  \
    do {                                                                                                          \
        cudaError_t e_ = (call);                                                                                  \
        if (e_ != cudaSuccess) {                                                                                  \
            cudaGetLastError();                                                                                   \
            return b200_fail(e_ == cudaErrorMemoryAllocation ? B200_ERR_ALLOC : B200_ERR_CUDA, "%s:%d %s: %s",   \
                             __FILE__, __LINE__, #call, cudaGetErrorString(e_));                                  \
        }                                                                                                         \
    } while (0)

struct b200_ctx {
    int          device     = 0;
    cudaStream_t stream     = nullptr;
    bool         own_stream = false;
    int          sm_count   = 148;
    uint64_t     launches   = 0;
    uint64_t     capture_launches0 = 0;         // launch counter when b200_graph_capture_begin was called
    int          force_path = B200_PATH_AUTO;
    int          exact      = 0;                // b200_ctx_set_exact: every MUL_MAT(_ID) on mv_exact_kernel (bit-identical to ggml-cpu's generic build)
    // scratch: quantized activations (grow-only; stream ordered because one stream per context)
    void *       scratch      = nullptr;
    size_t       scratch_size = 0;
    // staging for the *_host entry points
    void *       io           = nullptr;
    size_t       io_size      = 0;
    // second scratch: MoE index tables, tcgen05 activation tiles
    void *       scratch2      = nullptr;
    size_t       scratch2_size = 0;
    void *       scratch3      = nullptr;      // stream-K workspace of the fp16 prefill GEMM: flags + partial accumulators
    size_t       scratch3_size = 0;
    int          mv_warps = 0, mv_stages = 0;  // 0 = auto; GGML_B200_MV_WARPS / GGML_B200_MV_STAGES
    int          pdl        = 1;               // programmatic dependent launch of the mat-vec kernels; GGML_B200_PDL=0 disables
    int          weight_cache = 0;             // keep the int8 expansion of a weight matrix across calls (owner must invalidate)
    int          mmq_min_cols = 0;             // 0 = auto; GGML_B200_MMQ_MIN
    b200::SplitParams split = {};              // row-split all-gather fused into the kernels (b200_ctx_set_row_split)
    int          ll_on = 0;                    // flagged-payload exchange configured (b200_ctx_set_row_split_ll)
    double       ll_tensor_split[8] = {};
    int          ll_rounding = 1;
    uint64_t     ll_region = 0;
    int          split_deferred = 0;           // mat-vec kernels publish their epoch and let the NEXT launch wait (b200_ctx_set_row_split_deferred)
    int          split_pending  = 0;           // a deferred kernel was launched and nothing has waited for the peers since
    int          fuse_quant = 1;               // quantize src1 inside the mat-vec kernel when small; GGML_B200_FUSE_QUANT=0 disables
    // cross-kernel weight prefetch (kernels.h MvNext): the launch that followed a weight tensor last time is assumed to follow
    // it again (decode repeats the same graph every token). Keyed by the first weight pointer of a launch.
    int          prefetch_kb = 0;              // per-CTA prefetch depth in KB; 0 = off (b200_ctx_set_prefetch / GGML_B200_PREFETCH_KB)
    std::unordered_map<const void *, b200::MvNext> next_of;
    const void * last_key   = nullptr;
    uint64_t     mem_gen    = 0;               // generation of the device allocations the predictions were learned under
    // debug timeline of the mat-vec launches (b200_ctx_trace): slot = launch counter % trace_slots, sm_count CTAs x 8 u64 each
    unsigned long long * trace = nullptr;
    int          trace_slots = 0;
    int          dbg = 0;                      // GGML_B200_DBG: experiment switches handed to the kernels
    int          mma_min_cols = 4;             // columns from which the IMMA mat-vec kernel (mvmma.cu) replaces the dp4a one; GGML_B200_MMA_MIN
    int          mma_fuse_units = 0;           // the IMMA kernel quantizes src1 itself up to this many (column, unit) pairs (0: never - measured no faster than the PDL-chained quantizer launch); GGML_B200_MMA_FUSE_UNITS
    int          mma_warps = 0;                // 0 = auto; GGML_B200_MMA_WARPS
    // the fp16 activation operand last built in scratch2 (kind::f16 prefill): valid for further matrices of the same shared-src1 group
    const void * act_x = nullptr; size_t act_rs = 0; int64_t act_N = 0, act_K = 0; const void * act_buf = nullptr; int act_kind = -1;
    int          mmq_f16 = 1;                  // prefill GEMM on kind::f16 with the block scales folded into fp16 operands (mmf16.cuh); 0: kind::i8
    // tcgen05 mat-vec from raw blocks (mvtc.cuh): q4_K, tc_min_cols..tc_max_cols columns (GGML_B200_TC_MIN / _MAX; min 0 = off). Measured on
    // the 8B stack: wins at 9..16 columns (62 vs 51 TOPS at 16), loses to the IMMA mat-vec below 9 and to the kind::f16 GEMM from 32
    int          tc_min_cols = 9, tc_max_cols = 16;
    uint32_t *   tc_turn = nullptr;            // per-tile turn counters of its split-K accumulation
    const void * tc_act_x = nullptr; size_t tc_act_rs = 0; int64_t tc_act_N = 0, tc_act_K = 0; const void * tc_act_buf = nullptr;
    b200::MvNorm norm = {};                     // b200_mul_mat_*_norm: the RMS_NORM (+ MUL) the next one-column mat-vec launch applies to its src1
    bool         norm_used = false;             // ... and whether a launch took it
    b200::MvRope rope = {};                     // b200_qkv_rope: the ROPE / KV-cache stores the next q/k/v mat-vec launch applies to its rows
    bool         rope_used = false;
    int          fa_prefetch_kb = 0;            // FLASH_ATTN_EXT pulls this much per CTA of the mat-vec launch AFTER the next one into L2 (GGML_B200_FA_PREFETCH_KB)
    void *       fa_ws = nullptr; size_t fa_ws_size = 0;   // FLASH_ATTN_EXT: partial (max, sum, accumulator) of the KV splits (ops.cu)
    int          moe_tile_min  = 48;           // MUL_MAT_ID: one tile GEMM per expert from this many pairs per expert on average (GGML_B200_MOE_TILE_MIN; 0 = never)
    int          moe_group_min = 2;            // MUL_MAT_ID: group by expert (tensor-core mat-vec) from this many pairs per expert on
                                               // average; 0 = never; GGML_B200_MOE_GROUP_MIN
};

// launch on the context's stream with programmatic stream serialization (ctx->pdl): kernels launched this way start with
// pdl_prologue(), so that launch latency overlaps the previous kernel instead of adding to the dependent chain of a token step
template <typename... KArgs, typename... Args>
inline cudaError_t b200_launch(const b200_ctx * ctx, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, Args &&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = grid;
    cfg.blockDim         = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream           = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs    = attr;
    cfg.numAttrs = ctx->pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

b200::RopeP b200_rope_consts(const b200_rope_params * rp);   // ops.cu: theta_scale and the YaRN ramp range of a ROPE node

// bumped whenever device memory the library may hold predictions / caches for is freed or rewritten
uint64_t b200_mem_generation();
void     b200_mem_generation_bump();

// row split, deferred mode: enqueue the wait for the peers (a one-warp kernel) if a deferred kernel is outstanding. Every entry
// point that reads a dst, synchronizes, or launches a kernel without a split_wait prologue calls this first.
b200_status b200_split_flush(b200_ctx * ctx);

// row split, flagged-payload exchange: fills `g` with every rank's row bounds of the launch's matrices (full row count = dst_cs[i],
// the column stride of the full dst) and returns true if the launch can use it (configured, not deferred, slices match the rule,
// every dst inside the symmetric region)
bool b200_split_ll_geom(const b200_ctx * ctx, int n_mats, const int64_t * Ms, float * const * dsts, const int64_t * dst_cs, int ncols, b200::SplitGeom * g);

// grow-only device buffer owned by the context (fails instead of reallocating while the stream is being captured)
b200_status b200_ensure(b200_ctx * ctx, void ** buf, size_t * cur, size_t need);

// mvmma.cu: the IMMA mat-vec kernel for 3..8 columns (q4_0, q4_K)
namespace b200 { bool mv_mma_usable(int type); bool mv_mma_fits(int type, int64_t K); }
b200_status b200_mv_mma_run(b200_ctx * ctx, int type, int64_t K, int n_ops, const uint8_t * const * ws, const size_t * nb01s, const int64_t * Ms,
                            float * const * dsts, const int64_t * dst_cs, int ncols, const uint8_t * act_main, const uint8_t * act_aux,
                            const float * x = nullptr, size_t x_cs = 0);  // x: F32 columns quantized inside the kernel (no quantizer launch)

b200_status b200_mv_mma_run_ids(b200_ctx * ctx, int type, int64_t K, const uint8_t * as, size_t nb01, size_t nb02, int64_t M, int64_t n_as,
                                float * dst, const b200::MvIds & ids, const int32_t * sp, const int32_t * eb, const uint8_t * act_main,
                                const uint8_t * act_aux);

// mvtc.cuh (in mmq.cu): tcgen05 mat-vec for 4..32 columns from raw q4_K blocks
bool        b200_mvtc_usable(const b200_ctx * ctx, int type, int64_t K, int64_t N, const void * w, size_t nb01);
b200_status b200_mvtc_run(b200_ctx * ctx, const void * w, size_t nb01, int64_t M, int64_t K, const float * x, size_t x_rs, int64_t N, float * dst,
                          int64_t ldc, bool act_shared);
// mmq.cu: the tcgen05 int8 tile kernel
bool        b200_mmq_usable(int type, int64_t K, const void * w, size_t nb01);
// several matrices sharing src1 in one launch (kind::f16 path, weight cache on, every expansion cached); B200_ERR_UNSUPPORTED = loop
b200_status b200_mmq_run_group(b200_ctx * ctx, int type, int n_ops, const void * const * ws, const size_t * nb01s, const int64_t * Ms, int64_t K,
                               const float * x, size_t x_rs, int64_t N, float * const * dsts, const int64_t * ldcs);
// act_shared: this call belongs to a group of MUL_MATs that read the same src1 (b200_mul_mat_shared_src1) and is not the group's first:
// the fp16 activation operand built by the previous call of the group is still in scratch and is used as it is
b200_status b200_mmq_run(b200_ctx * ctx, int type, const void * w, size_t nb01, int64_t M, int64_t K, const float * x, size_t x_rs, int64_t N,
                         float * dst, int64_t ldc, bool act_shared = false);

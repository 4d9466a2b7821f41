/* include/ggml_b200.h — the thin C-ABI between the ggml-backend host code and the sm_100a CUDA kernels.
 *
 * Everything the B200 backend does on the device goes through these entry points: plain pointers,
 * sizes and ggml-style byte strides — no ggml, torch or CUDA types in the signatures. The host side
 * (qvac-ext-lib-llama.cpp_b200/csrc/ggml_backend_b200.cpp) implements the reference's
 * ggml_backend_reg_i / ggml_backend_device_i / ggml_backend_i / ggml_backend_buffer(_type)_i vtables
 * (/root/reference/ggml/src/ggml-backend-impl.h:16-204) purely in terms of this header; Python tests
 * and bench.py bind the same symbols with ctypes.
 *
 * Each entry point names the reference interface it replaces (paths relative to /root/reference).
 * There is NO CPU fallback anywhere behind this API: without a CUDA device every call fails with
 * B200_ERR_CUDA, and an unsupported shape/type fails with B200_ERR_UNSUPPORTED.
 */
#ifndef GGML_B200_H
#define GGML_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_API __attribute__((visibility("default")))

typedef enum {
    B200_OK              = 0,
    B200_ERR_UNSUPPORTED = 1, /* type / layout outside the hot path (ggml-cuda.cu supports_op :4258-4618 analogue) */
    B200_ERR_CUDA        = 2, /* CUDA runtime error, no device, or kernel image missing */
    B200_ERR_ARG         = 3,
    B200_ERR_ALLOC       = 4  /* device OOM: alloc_buffer returns nullptr (ggml-cuda.cu:695-712) */
} b200_status;

/* enum ggml_type values of the three weight formats (ggml/include/ggml.h:378-395) */
#define B200_TYPE_F32  0    /* activations; with F16 / I32 / I64 only as operands of the f3 ops at the end of this header */
#define B200_TYPE_F16  1
#define B200_TYPE_I32  26
#define B200_TYPE_I64  27
#define B200_TYPE_Q4_0 2
#define B200_TYPE_Q8_0 8
#define B200_TYPE_Q4_K 12
#define B200_TYPE_Q5_K 13   /* the Q5_K tensors of Q5_K_M / Q5_K_S files (src/llama-quant.cpp); 176-byte blocks: q4_K plus a fifth bit */
#define B200_TYPE_Q6_K 14   /* the attn_v / ffn_down / output tensors of a Q4_K_M file (src/llama-quant.cpp:226,303,360-363) */

/* which kernel family a MUL_MAT is routed to (b200_set_force_path / B200 env GGML_B200_FORCE) */
#define B200_PATH_AUTO   0
#define B200_PATH_MMVQ   1   /* decode mat-vec kernels, any n (looped over 8-column groups) */
#define B200_PATH_MMQ    2   /* tcgen05 int8 tile kernels */
#define B200_PATH_SIMPLE 3   /* generic one-warp-per-row kernel (any K % 32, any alignment) */

typedef struct b200_ctx b200_ctx; /* one device + one stream + scratch pool: a ggml_backend_t "stream"
                                     (ggml_backend_cuda_context, ggml-cuda/common.cuh:1134-1196) */

/* ---------------------------------------------------------------- device / context ---- */

/* number of usable devices = CUDA devices with compute capability 10.0 (ggml_backend_cuda_get_device_count, ggml-cuda.cu:4821) */
B200_API int          b200_device_count(void);
/* name[<=128], PCI bus id[<=32] (device_id, ggml-cuda.cu:4814-4816), SM count */
B200_API b200_status  b200_device_info(int device, char * name, size_t name_len, char * pci_bus_id, size_t id_len,
                                       int * sm_count, int * cc_major, int * cc_minor);
/* get_memory (ggml-cuda.cu:4188-4193) */
B200_API b200_status  b200_device_memory(int device, size_t * free_bytes, size_t * total_bytes);

/* init_backend (ggml-cuda.cu:4711-4717). stream: an existing cudaStream_t (as void*) to launch on, or NULL to create one */
B200_API b200_status  b200_ctx_create(int device, void * stream, b200_ctx ** out);
B200_API void         b200_ctx_destroy(b200_ctx * ctx);
B200_API int          b200_ctx_device(const b200_ctx * ctx);
B200_API void *       b200_ctx_stream(const b200_ctx * ctx);                 /* the cudaStream_t kernels are launched on */
B200_API b200_status  b200_ctx_synchronize(b200_ctx * ctx);                  /* iface.synchronize (ggml-cuda.cu:2828-2834) */
B200_API const char * b200_last_error(void);                                 /* thread-local message of the last failure */
B200_API uint64_t     b200_ctx_launch_count(const b200_ctx * ctx);           /* kernels launched so far (bench.py gpu_launches) */
B200_API void         b200_ctx_set_force_path(b200_ctx * ctx, int path);     /* B200_PATH_* */
/* Exact mode (env GGML_B200_EXACT=1): every MUL_MAT / MUL_MAT_ID of the context runs on a one-thread-per-element kernel that
 * performs the fp32 operations of ggml-cpu's generic ggml_vec_dot_<type>_q8_* (ggml-cpu/quants.c:115-150, 305-333, 550-623,
 * 625-703, 705-758) in the same order with the same roundings: results are bit-identical to the reference's x86-64 baseline
 * build of ggml-cpu. For validation (free-running greedy token streams against ggml-cpu); one to two orders of magnitude
 * slower than the default kernels, whose integer block sums are the same but whose fp32 summation order is not. */
B200_API void         b200_ctx_set_exact(b200_ctx * ctx, int enable);

/* ---------------------------------------------------------------- memory ---- */

/* buffer type alloc_buffer / free_buffer (ggml-cuda.cu:564-712); 256-byte aligned, NULL + B200_ERR_ALLOC on OOM */
B200_API b200_status  b200_malloc(int device, size_t size, void ** ptr);
B200_API b200_status  b200_free(int device, void * ptr);
/* pinned host buffer type (ggml-cuda.cu:1104-1168) */
B200_API b200_status  b200_host_malloc(size_t size, void ** ptr);
B200_API b200_status  b200_host_free(void * ptr);
/* buffer set_tensor / get_tensor / memset_tensor / clear: synchronous on return (ggml-cuda.cu:614-662) */
B200_API b200_status  b200_memcpy_h2d(int device, void * dst, const void * src, size_t size);
B200_API b200_status  b200_memcpy_d2h(int device, void * dst, const void * src, size_t size);
B200_API b200_status  b200_memcpy_d2d(int device, void * dst, const void * src, size_t size);
B200_API b200_status  b200_memcpy_peer(int dst_device, void * dst, int src_device, const void * src, size_t size);
B200_API b200_status  b200_memset(int device, void * dst, int value, size_t size);
/* iface.set_tensor_async / get_tensor_async / cpy_tensor_async on the context's stream (ggml-cuda.cu:2755-2826) */
B200_API b200_status  b200_memcpy_h2d_async(b200_ctx * ctx, void * dst, const void * src, size_t size);
B200_API b200_status  b200_memcpy_d2h_async(b200_ctx * ctx, void * dst, const void * src, size_t size);
B200_API b200_status  b200_memcpy_d2d_async(b200_ctx * ctx, void * dst, const void * src, size_t size);
/* split tensors (ggml_cuda_op_mul_mat, ggml-cuda.cu:1676-1734): src1 to the other devices and their dst slices back to the main
 * device. `height` rows of `width` bytes between any two devices of the process, on the context's stream (the reference's
 * cudaMemcpy2DPeerAsync / cudaMemcpyPeerAsync); b200_enable_peer_access lets the copies and kernels of `device` reach `peer`'s
 * memory directly (best effort: without it the driver stages the copies) */
B200_API b200_status  b200_memcpy2d_async(b200_ctx * ctx, void * dst, size_t dpitch, const void * src, size_t spitch, size_t width, size_t height);
B200_API b200_status  b200_enable_peer_access(int device, int peer);
/* iface.cpy_tensor_async between two backends of this library (ggml_backend_cuda_cpy_tensor_async, ggml-cuda.cu:2773-2826): a copy
 * between (or within) devices on ctx's stream; the caller orders the destination backend's stream behind it with an event */
B200_API b200_status  b200_memcpy_peer_async(b200_ctx * ctx, void * dst, int dst_device, const void * src, int src_device, size_t size);

/* events: device event_new/free/synchronize, backend event_record/event_wait (ggml-cuda.cu:2836-2862, 4649-4677) */
B200_API b200_status  b200_event_create(int device, void ** event);
B200_API b200_status  b200_event_destroy(int device, void * event);
B200_API b200_status  b200_event_record(b200_ctx * ctx, void * event);
B200_API b200_status  b200_event_wait(b200_ctx * ctx, void * event);
B200_API b200_status  b200_event_synchronize(void * event);
B200_API b200_status  b200_event_elapsed_ms(void * start, void * stop, float * ms);

/* ---------------------------------------------------------------- CUDA graphs ---- */

/* graph_compute with CUDA-graph replay (ggml_backend_cuda_graph_compute, ggml-cuda.cu:3679-3754): everything the context launches
 * between capture_begin and capture_end is recorded instead of run; capture_end instantiates it. A call that would have to grow
 * a scratch buffer or build a weight expansion fails during capture (run the same sequence once uncaptured first); after a failed
 * call the caller must still call capture_end (which then fails too) and fall back to plain launches. b200_graph_launch replays
 * the recorded launches on the context's stream and adds their number to b200_ctx_launch_count. */
typedef struct b200_graph b200_graph;
B200_API b200_status  b200_graph_capture_begin(b200_ctx * ctx);
B200_API b200_status  b200_graph_capture_end(b200_ctx * ctx, b200_graph ** out);
B200_API b200_status  b200_graph_launch(b200_ctx * ctx, b200_graph * g);
B200_API void         b200_graph_destroy(b200_graph * g);
B200_API uint64_t     b200_graph_kernel_count(const b200_graph * g);

/* ---------------------------------------------------------------- K1: activation quantizer ---- */

/* F32 rows -> ggml blocks, bit-exact to quantize_row_q8_0_ref / _q8_1_ref / _q8_K_ref
 * (ggml/src/ggml-quants.c:199-222, 225-258, 2555-2592); replaces quantize_row_q8_1_cuda
 * (ggml-cuda/quantize.cu:148-162). x: device F32, row r at (char*)x + r*x_row_stride; y: device, rows of
 * k/32*34 (q8_0), k/32*36 (q8_1) or k/256*292 (q8_K) bytes, contiguous. k % 32 == 0 (q8_K: % 256). */
B200_API b200_status  b200_quantize_q8_0(b200_ctx * ctx, const float * x, int64_t k, int64_t nrows, size_t x_row_stride, void * y);
B200_API b200_status  b200_quantize_q8_1(b200_ctx * ctx, const float * x, int64_t k, int64_t nrows, size_t x_row_stride, void * y);
B200_API b200_status  b200_quantize_q8_K(b200_ctx * ctx, const float * x, int64_t k, int64_t nrows, size_t x_row_stride, void * y);

/* the kernels' internal activation layout ("unit records": 256 int8 + 16 B scales/sums per 256 elements, + 16 B aux),
 * same values as the functions above. kind: 0 = q8_0-valued (for q4_0/q8_0 weights), 1 = q8_K-valued (for q4_K).
 * main: nrows * k/256 * 272 bytes, aux: nrows * k/256 * 16 bytes. Exposed for the parity tests. k % 256 == 0. */
B200_API b200_status  b200_quantize_units(b200_ctx * ctx, int kind, const float * x, int64_t k, int64_t nrows, size_t x_row_stride,
                                          void * main, void * aux);

/* ---------------------------------------------------------------- K3/K4: MUL_MAT ---- */

/* tensor geometry exactly as struct ggml_tensor carries it (ggml/include/ggml.h:645-676): ne in elements, nb in bytes.
 * src0 = weights [ne00=K, ne01=M, ne02, ne03] of `type`, rows contiguous (nb00 == type size);
 * src1 = activations F32 [K, ne11=N, ne12, ne13], nb10 == 4; dst F32 [M, N, ne12, ne13], nb0 == 4.
 * src0 is broadcast over ne12/ne02 and ne13/ne03 (ggml-cpu.c:1133-1134). */
typedef struct {
    int32_t type;
    int64_t ne00, ne01, ne02, ne03;
    size_t  nb01, nb02, nb03;
    int64_t ne11, ne12, ne13;
    size_t  nb11, nb12, nb13;
    size_t  nb1, nb2, nb3;
} b200_mul_mat_desc;

/* supports_op for MUL_MAT (ggml-cuda.cu:4258-4330 analogue): 1 if b200_mul_mat accepts this geometry */
B200_API int          b200_mul_mat_supported(const b200_mul_mat_desc * d);

/* GGML_OP_MUL_MAT on device pointers; replaces ggml_cuda_mul_mat (ggml-cuda.cu:2178-2261) for the three types:
 * quantizes src1 (K1), then mat-vec kernels for N <= 8 (ggml_cuda_mul_mat_vec_q, mmvq.cu:605-732) or the tcgen05
 * int8 tile kernel otherwise (ggml_cuda_mul_mat_q, mmq.cu:70-236). Asynchronous on the context's stream. */
B200_API b200_status  b200_mul_mat(b200_ctx * ctx, const b200_mul_mat_desc * d, const void * src0, const float * src1, float * dst);

/* Several MUL_MATs that share src1 (q/k/v, gate/up) in one launch. All descs must have identical type, K, N, batch
 * dims 1 and src1 geometry. The ggml backend's graph_compute uses this when consecutive nodes share src1
 * (the reference re-quantizes per op: SURVEY §8a a15). */
B200_API b200_status  b200_mul_mat_shared_src1(b200_ctx * ctx, int n_ops, const b200_mul_mat_desc * descs,
                                               const void * const * src0s, const float * src1, float * const * dsts);

/* same op with HOST buffers: uploads src1 (and src0 when w_on_device == 0), runs, downloads dst, synchronizes.
 * This is the call bench.py's `e2e` leg times (host<->device copies inside). */
B200_API b200_status  b200_mul_mat_host(b200_ctx * ctx, const b200_mul_mat_desc * d, const void * src0, int w_on_device,
                                        const float * src1_host, float * dst_host);

/* GGML_OP_MUL_MAT for callers that may not allocate device memory (code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE. Same arithmetic and
 * kernels as b200_mul_mat's decode path (K1 quantizes src1 into unit records, the TMA-staged mat-vec kernels read them), but the
 * records live in a file-scope `static __device__` array of B200_NOALLOC_SCRATCH_BYTES that every GPU which loads the library holds
 * until exit; the call performs no cudaMalloc / cudaFree and calls no allocating library routine. Limits (B200_ERR_UNSUPPORTED
 * with a message otherwise): plain 2-D MUL_MAT (ne02 = ne03 = ne12 = ne13 = 1), 1..8 activation columns, K % 256 == 0,
 * columns x K/256 <= B200_NOALLOC_MAX_UNITS (e.g. 8 columns of K = 32768), 16-byte aligned weight rows, not in row-split mode.
 * All calls on a device share the array, which keeps the last call's records (every byte that is read is rewritten by the call
 * itself, nothing needs to start at zero): calls that overlap on the GPU - two streams or contexts of one device - race. */
#define B200_NOALLOC_MAX_UNITS     1024
#define B200_NOALLOC_SCRATCH_BYTES (B200_NOALLOC_MAX_UNITS * (272 + 16))
B200_API b200_status  b200_mul_mat_noalloc(b200_ctx * ctx, const b200_mul_mat_desc * d, const void * src0, const float * src1, float * dst);
B200_API size_t       b200_mul_mat_noalloc_max_units(void);

/* ---------------------------------------------------------------- K5/K6: MUL_MAT_ID ---- */

/* as = experts [K, M, n_as] of `type` (expert e at as + e*nb02); b F32 [K, ne11, n_tok] (nb11, nb12);
 * ids I32 [n_used, n_tok] (row stride ids_nb1); dst F32 [M, n_used, n_tok] (nb1, nb2):
 * dst[:, i, t] = as[ids[i,t]] . b[:, i % ne11, t]   (ggml-cpu.c:1476-1649; ggml_cuda_mul_mat_id, ggml-cuda.cu:2263-2409) */
typedef struct {
    int32_t type;
    int64_t ne00, ne01, n_as;
    size_t  nb01, nb02;
    int64_t ne11, n_tok;
    size_t  nb11, nb12;
    int64_t n_used;
    size_t  ids_nb1;
    size_t  nb1, nb2;
} b200_mul_mat_id_desc;

B200_API int          b200_mul_mat_id_supported(const b200_mul_mat_id_desc * d);
B200_API b200_status  b200_mul_mat_id(b200_ctx * ctx, const b200_mul_mat_id_desc * d, const void * as, const float * b,
                                      const int32_t * ids, float * dst);

/* K5 alone (mm_ids_helper, ggml-cuda/mmid.cu:26-164): counting sort of the (slot, token) pairs by expert.
 * pair p = t*n_used + i. Outputs (device): sorted_pairs[n_pairs] (pair index in expert order, stable in p),
 * expert_bounds[n_as + 1]. Exposed for the parity tests. */
B200_API b200_status  b200_mm_ids_sort(b200_ctx * ctx, const int32_t * ids, int64_t n_used, int64_t n_tok, size_t ids_nb1,
                                       int64_t n_as, int32_t * sorted_pairs, int32_t * expert_bounds);

/* ---------------------------------------------------------------- K7: fused epilogues of the one-column mat-vec ---- */

/* SURVEY 8(f) rank 1. The reference fuses, for one dst column only, {MUL_MAT(up), MUL_MAT(gate), GLU}, the same with an ADD
 * (ADD_ID for MUL_MAT_ID) of a bias row behind each MUL_MAT, and {MUL_MAT, ADD} (ggml_cuda_should_fuse_mul_mat + the
 * graph loop, ggml-cuda.cu:2020-2176, 3404-3600; epilogue of mul_mat_vec_q, mmvq.cu:181-328; test_mul_mat_vec_fusion,
 * tests/test-backend-ops.cpp:5120-5249). Here it is an epilogue of the TMA-staged mat-vec kernel: a warp runs the gate rows
 * of a row group, keeps the sums in registers, runs the same rows of the up matrix and stores glu(gate, up) - one launch,
 * one quantization of src1, no intermediate tensors. glu_op values are enum ggml_glu_op (ggml.h:596-605). */
#define B200_GLU_NONE        (-1)  /* dst = up . x + up_bias   ({MUL_MAT, ADD}) */
#define B200_GLU_REGLU       0
#define B200_GLU_GEGLU       1
#define B200_GLU_SWIGLU      2
#define B200_GLU_SWIGLU_OAI  3
#define B200_GLU_GEGLU_ERF   4
#define B200_GLU_GEGLU_QUICK 5

typedef struct {
    int32_t       glu_op;             /* B200_GLU_* */
    float         alpha, limit;       /* SWIGLU_OAI (op_params 2, 3 of the GLU node; ggml-cpu/ops.cpp:3236-3237) */
    const float * up_bias;            /* nullable. MUL_MAT: F32 of dst's shape [M, 1, ne12, ne13]; MUL_MAT_ID: [M, n_as] */
    const float * gate_bias;          /* nullable; only with a GLU */
    size_t        bias_nb1;           /* MUL_MAT_ID: bytes between the bias rows of two experts */
    size_t        bias_nb2, bias_nb3; /* MUL_MAT: batch strides of the bias tensors */
} b200_fusion;

/* 1 if b200_mul_mat_fused would run this geometry as ONE fused launch; 0: run the nodes one by one (b200_mul_mat, b200_add,
 * b200_glu). d describes the up MUL_MAT (ne11 must be 1); the gate matrix has the same type, shape and strides. */
B200_API int          b200_mul_mat_fused_supported(const b200_ctx * ctx, const b200_mul_mat_desc * d);
/* d->nb1..nb3 are the strides of `dst`, the output of the LAST fused node. gate == NULL iff f->glu_op == B200_GLU_NONE. */
B200_API b200_status  b200_mul_mat_fused(b200_ctx * ctx, const b200_mul_mat_desc * d, const void * up, const void * gate, const float * src1,
                                         const b200_fusion * f, float * dst);
/* same for MUL_MAT_ID with one token (d->n_tok == 1): dst[:, i] = glu(gate_as[ids[i]] . b + gate_bias[:, ids[i]], up_as[ids[i]] . b + ...) */
B200_API int          b200_mul_mat_id_fused_supported(const b200_ctx * ctx, const b200_mul_mat_id_desc * d);
B200_API b200_status  b200_mul_mat_id_fused(b200_ctx * ctx, const b200_mul_mat_id_desc * d, const void * up_as, const void * gate_as,
                                            const float * b, const int32_t * ids, const b200_fusion * f, float * dst);

/* The same nodes stand-alone (F32, dimension 0 contiguous), for when the pattern above does not apply.
 * GGML_OP_GLU (ggml-cpu/ops.cpp:2934-3290; ggml-cuda/unary.cu ggml_cuda_op_unary_gated): dst[r][c] = glu(gate[r][c], up[r][c]) over
 * nrows rows of nc values; the single-tensor and swapped forms are the caller's pointer arithmetic. */
B200_API b200_status  b200_glu(b200_ctx * ctx, int glu_op, float alpha, float limit, const float * gate, size_t gate_nb1, const float * up,
                               size_t up_nb1, float * dst, size_t dst_nb1, int64_t nc, int64_t nrows);
/* GGML_OP_ADD with ggml broadcasting (ggml-cpu/binary-ops.cpp; ggml-cuda/binbcast.cu): ne/nb_* as struct ggml_tensor carries them */
B200_API b200_status  b200_add(b200_ctx * ctx, const int64_t * ne, const float * a, const size_t * nb_a, const float * b, const int64_t * ne_b,
                               const size_t * nb_b, float * dst, const size_t * nb_d);
/* GGML_OP_MUL, same geometry rules (the expert-weight scaling behind a fused MoE FFN: test_mul_mat_vec_fusion, test-backend-ops.cpp:5227-5229) */
B200_API b200_status  b200_mul(b200_ctx * ctx, const int64_t * ne, const float * a, const size_t * nb_a, const float * b, const int64_t * ne_b,
                               const size_t * nb_b, float * dst, const size_t * nb_d);
/* GGML_OP_ADD_ID (ggml-cpu/ops.cpp:701-750; ggml-cuda/add-id.cu): dst[:, i, t] = a[:, i, t] + bias[:, ids[i, t]] */
B200_API b200_status  b200_add_id(b200_ctx * ctx, const float * a, size_t a_nb1, size_t a_nb2, const float * bias, size_t bias_nb1, int64_t n_as,
                                  const int32_t * ids, size_t ids_nb1, float * dst, size_t nb1, size_t nb2, int64_t M, int64_t n_used, int64_t n_tok);

/* ---------------------------------------------------------------- row-split tensor parallelism ---- */

/* Weight rows split over n_ranks GPUs of one box (ggml_backend_split_buffer_type + ggml_cuda_op_mul_mat,
 * ggml-cuda.cu:807-1102, 1447-1765): each rank holds rows [lo, hi) of every split matrix, src1 is replicated, and every rank
 * needs the whole dst. Instead of the reference's cudaMemcpy2DPeerAsync per device (:1721-1734) the kernels store each
 * result into the dst of every rank (peer memory over NVLink/NVSwitch) and exchange one flag per rank and op: the
 * all-gather is part of the MUL_MAT launch. While a context is in this mode, every MUL_MAT / MUL_MAT_ID it runs must have
 *   - desc.ne01 = the rank's row count, src0 = the rank's rows, dst = address of (row lo, column 0) inside the rank's FULL dst,
 *     nb1 = the full dst's column stride;
 *   - dst inside a region that exists on every rank such that  peer address = local address + dst_delta[r]  (symmetric
 *     allocation; single-process hosts enable peer access, one-process-per-GPU hosts map the peers' allocations);
 *   - the same sequence of calls on every rank (the flag protocol counts ops).
 * sync_words[r]: rank r's 16 zero-initialised uint32 words (also peer-mapped). */
typedef struct {
    int32_t n_ranks, rank;
    int64_t dst_delta[8];
    void *  sync_words[8];
} b200_row_split;
B200_API b200_status  b200_ctx_set_row_split(b200_ctx * ctx, const b200_row_split * rs);   /* rs == NULL or n_ranks <= 1: off */
/* Flagged-payload exchange for the decode mat-vec kernels (at most 8 activation columns, plain MUL_MAT). Instead of peer stores
 * into dst followed by a system-scope release and a flag round trip, every result travels as ONE aligned 8-byte {value, epoch}
 * store into a staging area on each peer, and every rank's kernel polls the staging records of the rows the other ranks own and
 * copies them into its dst: no fence, no flag and no counter on the data path. Call after b200_ctx_set_row_split (which clears it).
 *   stage[r]      rank r's staging area, peer-mapped like the sync words: 4 * region_bytes bytes, zero-initialised
 *   dst_base      local address of the symmetric dst region (every dst of a split MUL_MAT lies in [dst_base, dst_base + region_bytes))
 *   tensor_split  cumulative start fraction of every rank, rounding: the get_row_split rule (ggml-cuda.cu:786-799) that produced the
 *                 ranks' slices; with nb1 / 4 = rows of the full matrix the kernels derive who owns which rows from it. A MUL_MAT
 *                 whose ne01 does not match the rule, or whose dst lies outside the region, uses the flag protocol.
 * ll == NULL switches it off. */
typedef struct {
    void *       stage[8];
    const void * dst_base;
    uint64_t     region_bytes;
    double       tensor_split[8];
    int32_t      rounding;
} b200_row_split_ll;
B200_API b200_status  b200_ctx_set_row_split_ll(b200_ctx * ctx, const b200_row_split_ll * ll);
/* Deferred mode: a decode mat-vec kernel publishes its epoch to the peers and returns without waiting for theirs; the wait moves into
 * the prologue of the next kernel of the context (behind its weight prefetch) and into every entry point that reads results or
 * synchronizes (memcpy, event record, synchronize run b200_ctx_row_split_wait first). A caller that lets anything else consume a
 * dst on the stream must call b200_ctx_row_split_wait itself before that. */
B200_API b200_status  b200_ctx_set_row_split_deferred(b200_ctx * ctx, int enable);
B200_API b200_status  b200_ctx_row_split_wait(b200_ctx * ctx);

/* ---------------------------------------------------------------- prefill weight cache ---- */

/* The tcgen05 path consumes an int8, K-major, TMA-tileable expansion of a weight matrix plus per-32 float scales
 * (the load-time analogue of ggml-cpu's repack buffer type, ggml-cpu/repack.cpp). By default it is rebuilt on every
 * call in scratch memory. A caller that owns the weight memory and reports every write to it (the ggml backend's
 * buffer: set_tensor / memset_tensor / clear / cpy_tensor / free_buffer) may switch a context to keep one expansion
 * per weight matrix (keyed by device + src0 pointer) and must then invalidate overwritten or freed ranges. */
B200_API void         b200_ctx_set_weight_cache(b200_ctx * ctx, int enable);
B200_API void         b200_weight_cache_invalidate(int device, const void * base, size_t size);
B200_API void         b200_weight_cache_clear(int device);
B200_API size_t       b200_weight_cache_bytes(void);
/* Upper bound for the cached expansions, all devices together (default: 40 % of the first device's memory; env
 * GGML_B200_WEIGHT_CACHE_MB). Beyond it the least recently used expansions nobody is using are dropped; a matrix that still does not
 * fit - or whose allocation fails, or whose source keeps being rewritten - is expanded into scratch memory for the one call. */
B200_API void         b200_weight_cache_set_budget(size_t bytes);

/* Prefill arithmetic (more than 8 activation columns). enable = 1 (default; env GGML_B200_MMQ_F16): tcgen05 kind::f16 GEMM on
 * fp16 operands that carry the block scales (weights: dequantize_row_<type> rounded to fp16; activations: the bit-exact
 * q8_0 / q8_K quantization, dequantized, times a per-row power of two, rounded to fp16) with one fp32 accumulation over K:
 * NMSE against the CPU reference about 1e-7. enable = 0: tcgen05 kind::i8 with an exact int32 sum per 32-block scaled in
 * fp32 (the arithmetic of ggml_cuda_mul_mat_q, ggml-cuda/mmq.cu:70-236), about three times slower. */
B200_API void         b200_ctx_set_mmq_f16(b200_ctx * ctx, int enable);

/* Cross-kernel weight prefetch for decode. A token step is a dependent chain of short mat-vec kernels; at every kernel
 * boundary HBM idles for the launch / dependency / drain latency (~3 us, as long as a 4096x4096 q4 matrix takes to stream).
 * With prefetch on, a context remembers which launch followed each weight tensor the last time (decode replays the same
 * graph every token: ggml_backend_sched, ggml-backend.cpp:1439-1560) and lets every mat-vec kernel pull the first
 * `kb_per_cta` KB per SM of the predicted next kernel's weights into L2 once its own copies are in flight. Results are
 * unaffected; a wrong prediction only costs the prefetch traffic. Predictions hold raw device pointers: they are dropped
 * whenever b200_free is called, and a caller that frees weight memory by other means must switch prefetch off (0) first. */
B200_API void         b200_ctx_set_prefetch(b200_ctx * ctx, int kb_per_cta);

/* Debug: device-side timeline of the mat-vec launches of a context (how profiles/ explains kernel-boundary latency without a
 * profiler attached). n_slots > 0 allocates n_slots x sm_count x 8 uint64; launch number L of the context (see
 * b200_ctx_launch_count) records into slot L % n_slots, CTA b at [slot][b][0..7] = %globaltimer ns of: CTA start, first weight
 * copy issued, last weight copy issued, consumers past the stream dependency, activations quantized, first stage landed,
 * consumer warp 0 done; [7] = SM id. n_slots = 0 frees it. b200_ctx_trace_read synchronizes the stream and copies it out. */
B200_API b200_status  b200_ctx_trace(b200_ctx * ctx, int n_slots);
B200_API b200_status  b200_ctx_trace_read(b200_ctx * ctx, void * host, size_t bytes);

/* ---------------------------------------------------------------- f3: the ops between the matmuls of a decode graph ---- */

/* SURVEY 8(f) rank 3, first slice: with these a Llama token step (src/llama-graph.cpp build_attn_mha :1335-1398, src/models/llama.cpp)
 * runs on the device from the embedding lookup to the logits, instead of one scheduler split per MUL_MAT. Each replaces the
 * ggml-cuda op named beside it and follows ggml-cpu's semantics (ggml/src/ggml-cpu/ops.cpp). A tensor is passed as struct
 * ggml_tensor carries it (ggml/include/ggml.h:645-676): data pointer (device), ggml type, ne in elements, nb in bytes. Every
 * op has a *_supported predicate (what dev_supports_op answers, ggml-cuda.cu:4258-4618); the op itself returns
 * B200_ERR_UNSUPPORTED for anything else. All run on the context's stream; none allocates except FLASH_ATTN_EXT's split
 * workspace (grow-only, fails instead of growing during a graph capture). */
typedef struct {
    void *  data;
    int32_t type;     /* B200_TYPE_* (= enum ggml_type) */
    int64_t ne[4];
    size_t  nb[4];
} b200_tensor;

/* GGML_OP_RMS_NORM (ggml-cpu/ops.cpp:3646-3695; ggml-cuda/norm.cu rms_norm_f32): y = x / sqrt(mean(x^2) + eps) per F32 row */
B200_API int          b200_rms_norm_supported(const b200_tensor * x, const b200_tensor * dst);
B200_API b200_status  b200_rms_norm(b200_ctx * ctx, const b200_tensor * x, const b200_tensor * dst, float eps);

/* FLASH_ATTN_EXT of one token right behind the layer's own KV-cache write. written_idx (I64 / I32, one element): the caller guarantees
 * that the only rows of k / v which launches still in flight on the context's stream may write are rows written_idx[0] - what a
 * decode graph does (SET_ROWS(k), SET_ROWS(v) by the token's slot, then the attention; the previous token's launches completed
 * long ago). The kernel then requests the other K / V rows of its first round before the stream dependency resolves (it is resident
 * while the q/k/v launch still runs) and reads row written_idx[0] again afterwards. Same results as b200_flash_attn_ext. */
B200_API b200_status  b200_flash_attn_ext_decode(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                                                 const b200_tensor * mask, const b200_tensor * sinks, const b200_tensor * dst, float scale,
                                                 float max_bias, float logit_softcap, const b200_tensor * written_idx);

/* RMS_NORM followed by the MUL with the norm weight (build_norm, src/llama-graph.cpp; ggml-cuda fuses the same pair:
 * ggml_cuda_op_rms_norm_fused, ggml-cuda.cu:3360-3380): dst = (x * scale) * w, both products rounded as the two nodes would;
 * w F32 with w->ne[0] == x->ne[0], its other dimensions broadcast */
B200_API int          b200_rms_norm_mul_supported(const b200_tensor * x, const b200_tensor * w, const b200_tensor * dst);
B200_API b200_status  b200_rms_norm_mul(b200_ctx * ctx, const b200_tensor * x, const b200_tensor * w, const b200_tensor * dst, float eps);

/* RMS_NORM (+ MUL) folded into the MUL_MATs that read it (a decode step's attn_norm -> q/k/v and ffn_norm -> gate/up; the reference
 * keeps them apart: norm.cu + quantize.cu + mmvq.cu). x is the UN-normalized src1; with one activation column on the TMA-staged
 * mat-vec kernels the kernel normalizes the row in its own quantizer prologue, in the summation order of the stand-alone RMS_NORM
 * kernel (same bits); otherwise the norm runs first, into y ([K, N] F32, rows y_nb1 bytes apart: the memory of the graph's MUL
 * node), and the MUL_MATs read y. Either way the results are those of the nodes one after the other. w == NULL: no MUL. */
typedef struct {
    const float * w;      /* [K] per-channel norm weights (src of the MUL), 16-byte aligned, or NULL */
    float         eps;
    float *       y;      /* where RMS_NORM(x) * w goes when it has to be materialized */
    size_t        y_nb1;
} b200_norm;
B200_API b200_status  b200_mul_mat_shared_src1_norm(b200_ctx * ctx, int n_ops, const b200_mul_mat_desc * descs, const void * const * src0s,
                                                    const float * x, const b200_norm * norm, float * const * dsts);
B200_API b200_status  b200_mul_mat_fused_norm(b200_ctx * ctx, const b200_mul_mat_desc * d, const void * up, const void * gate, const float * x,
                                              const b200_norm * norm, const b200_fusion * f, float * dst);

/* GGML_OP_GET_ROWS (ggml-cpu/ops.cpp ggml_compute_forward_get_rows; ggml-cuda/getrows.cu): dst[:, i, j, k] = row ids[i, j, k] of
 * src0[:, :, j, k] as F32; src0 F32, F16 or one of the five quantized formats (dequantize_row_<type>, ggml-quants.c); ids I32 */
B200_API int          b200_get_rows_supported(const b200_tensor * src0, const b200_tensor * ids, const b200_tensor * dst);
B200_API b200_status  b200_get_rows(b200_ctx * ctx, const b200_tensor * src0, const b200_tensor * ids, const b200_tensor * dst);

/* GGML_OP_ROPE (ggml-cpu/ops.cpp:5525-5780; ggml-cuda/rope.cu): rotary embedding of F32 [n_embd_head, n_head, n_tokens, n_batch]
 * by I32 positions [n_tokens]; mode 0 rotates adjacent pairs, GGML_ROPE_TYPE_NEOX pairs (i, i + n_dims/2); optional per-channel
 * frequency factors (src[2]); YaRN through ext_factor / attn_factor / beta_*; the fields are op_params[1,2,4..10] of the node */
#define B200_ROPE_NORMAL 0
#define B200_ROPE_NEOX   2
typedef struct {
    int32_t n_dims, mode, n_ctx_orig;
    float   freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow;
} b200_rope_params;
B200_API int          b200_rope_supported(const b200_tensor * x, const b200_tensor * pos, const b200_tensor * freq_factors, const b200_tensor * dst,
                                          const b200_rope_params * rp);
B200_API b200_status  b200_rope(b200_ctx * ctx, const b200_tensor * x, const b200_tensor * pos, const b200_tensor * freq_factors, const b200_tensor * dst,
                                const b200_rope_params * rp);

/* GGML_OP_SET_ROWS (ggml-cpu/ops.cpp:4828-4876; ggml-cuda/set-rows.cu): dst[:, idx[i, j % ne11, k % ne12], j, k] = src[:, i, j, k];
 * src F32, dst F32 or F16 (the KV cache write of llama_kv_cache::cpy_k / cpy_v), idx I64 or I32 */
B200_API int          b200_set_rows_supported(const b200_tensor * src, const b200_tensor * idx, const b200_tensor * dst);
B200_API b200_status  b200_set_rows(b200_ctx * ctx, const b200_tensor * src, const b200_tensor * idx, const b200_tensor * dst);

/* The attention inputs of one layer in one launch (the reference fuses ROPE + VIEW + SET_ROWS: ggml-cuda.cu:3300-3330): ROPE(q) into
 * q_dst, ROPE(k) into k_dst AND - as SET_ROWS(k_cache, k_dst seen as [n_embd_k_gqa, n_tokens], k_idx) would - into the cache rows, and
 * SET_ROWS(v_cache, v, v_idx). Same results as the four nodes one after the other (build_attn, src/llama-graph.cpp:1593-1606). */
B200_API int          b200_attn_prep_supported(const b200_tensor * q, const b200_tensor * q_dst, const b200_tensor * k, const b200_tensor * k_dst,
                                               const b200_tensor * pos, const b200_tensor * freq_factors, const b200_rope_params * rp,
                                               const b200_tensor * k_idx, const b200_tensor * k_cache, const b200_tensor * v, const b200_tensor * v_idx,
                                               const b200_tensor * v_cache);
B200_API b200_status  b200_attn_prep(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * q_dst, const b200_tensor * k, const b200_tensor * k_dst,
                                     const b200_tensor * pos, const b200_tensor * freq_factors, const b200_rope_params * rp, const b200_tensor * k_idx,
                                     const b200_tensor * k_cache, const b200_tensor * v, const b200_tensor * v_idx, const b200_tensor * v_cache);

/* The three attention projections of a layer (src0s = wq, wk, wv against the shared activations x, optionally un-normalized: norm != NULL
 * as in b200_mul_mat_shared_src1_norm) into dsts[0..2], followed by b200_attn_prep on them (q / k / v describe dsts[0..2] as the ROPE /
 * SET_ROWS nodes see them). For one token on the lean mat-vec kernel (normal-mode rope over the whole head, F16 caches) the rotation
 * and the cache stores happen in the kernel's own store - a row group is an even number of consecutive rows, i.e. whole rope pairs - and
 * neither dsts[] nor a second launch are needed: the decode step's [norm + q/k/v + rope + KV store] is ONE launch. Same bits either way. */
B200_API b200_status  b200_qkv_rope(b200_ctx * ctx, const b200_mul_mat_desc * descs, const void * const * src0s, const float * x, const b200_norm * norm,
                                    float * const * dsts, const b200_tensor * q, const b200_tensor * q_dst, const b200_tensor * k, const b200_tensor * k_dst,
                                    const b200_tensor * pos, const b200_tensor * freq_factors, const b200_rope_params * rp, const b200_tensor * k_idx,
                                    const b200_tensor * k_cache, const b200_tensor * v, const b200_tensor * v_idx, const b200_tensor * v_cache);

/* One token (n_tokens == 1): the same four nodes AND the FLASH_ATTN_EXT behind them as ONE launch. Every CTA of the attention kernel
 * rotates its own q head and the new K row of its KV head (rope_row's arithmetic), uses them and the new V row instead of the
 * cache row they belong in, and the first CTA of each KV head writes that row (and q_dst / k_dst). fa_q / fa_k / fa_v: the operands of
 * the FLASH_ATTN_EXT node (views of q_dst and of the caches); the views must address cache row r, head h where SET_ROWS writes it.
 * Results equal to b200_attn_prep followed by b200_flash_attn_ext, bit for bit. */
B200_API int          b200_attn_decode_supported(const b200_tensor * q, const b200_tensor * q_dst, const b200_tensor * k, const b200_tensor * k_dst,
                                                 const b200_tensor * pos, const b200_tensor * freq_factors, const b200_rope_params * rp,
                                                 const b200_tensor * k_idx, const b200_tensor * k_cache, const b200_tensor * v, const b200_tensor * v_idx,
                                                 const b200_tensor * v_cache, const b200_tensor * fa_q, const b200_tensor * fa_k, const b200_tensor * fa_v,
                                                 const b200_tensor * mask, const b200_tensor * sinks, const b200_tensor * dst);
B200_API b200_status  b200_attn_decode(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * q_dst, const b200_tensor * k, const b200_tensor * k_dst,
                                       const b200_tensor * pos, const b200_tensor * freq_factors, const b200_rope_params * rp, const b200_tensor * k_idx,
                                       const b200_tensor * k_cache, const b200_tensor * v, const b200_tensor * v_idx, const b200_tensor * v_cache,
                                       const b200_tensor * fa_q, const b200_tensor * fa_k, const b200_tensor * fa_v, const b200_tensor * mask,
                                       const b200_tensor * sinks, const b200_tensor * dst, float scale, float max_bias, float logit_softcap);

/* GGML_OP_CPY / CONT / DUP (ggml-cpu/ops.cpp ggml_compute_forward_dup; ggml-cuda/cpy.cu): element e of src (any strides) to
 * element e of dst (any strides, same number of elements), F32 / F16 in any combination (F32 -> F16 rounds to nearest even) */
B200_API int          b200_cpy_supported(const b200_tensor * src, const b200_tensor * dst);
B200_API b200_status  b200_cpy(b200_ctx * ctx, const b200_tensor * src, const b200_tensor * dst);

/* GGML_OP_FLASH_ATTN_EXT (ggml-cpu/ops.cpp:7994-8218; ggml-cuda/fattn*.cu): q F32 [D, n_q, n_head, n_batch], k / v F16
 * [D, n_kv, n_head_kv, n_batch] (D = 64, 128 or 256; rows may be strided views of the KV cache), optional F16 mask
 * [n_kv, >= n_q, 1 | n_head, 1 | n_batch] (-inf = masked, times the ALiBi slope when max_bias > 0), optional F32 sinks
 * [n_head]; dst F32 [D, n_head, n_q, n_batch] = softmax(scale * q k^T (+ softcap) + mask) v. fp32 online softmax. */
B200_API int          b200_flash_attn_ext_supported(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask,
                                                    const b200_tensor * sinks, const b200_tensor * dst);
B200_API b200_status  b200_flash_attn_ext(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask,
                                          const b200_tensor * sinks, const b200_tensor * dst, float scale, float max_bias, float logit_softcap);

#ifdef __cplusplus
}
#endif
#endif /* GGML_B200_H */

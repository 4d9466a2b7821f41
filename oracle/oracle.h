/* oracle/oracle.h — TEST INFRASTRUCTURE ONLY (not product code).
 *
 * Plain-C CPU restatement of the reference's algorithm for the hot path
 * (GGML_OP_MUL_MAT / GGML_OP_MUL_MAT_ID on q4_0 / q8_0 / q4_K weights x F32
 * activations, as ggml-cpu computes it). Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline leg may load this; the product path never does.
 *
 * Parity: PINNED. tests/test_oracle.py checks every function below against the
 * reference's own compiled code (oracle/_ref/libggml-base.so + libggml-cpu-x64.so,
 * built from /root/reference by oracle/Makefile) and against committed fixtures in
 * tests/golden/ generated from that build by tests/golden/make_golden.py.
 *
 * All citations are into /root/reference.
 */
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum orc_type { ORC_Q4_0 = 2, ORC_Q8_0 = 8, ORC_Q4_K = 12, ORC_Q5_K = 13, ORC_Q6_K = 14 }; /* == enum ggml_type values, ggml/include/ggml.h:378-395 */

/* block sizes in bytes / elements (ggml/src/ggml-common.h:170-175, 218-238, 291-305, 322-328) */
#define ORC_QK      32
#define ORC_QK_K    256
#define ORC_SZ_Q4_0 18
#define ORC_SZ_Q8_0 34
#define ORC_SZ_Q8_1 36
#define ORC_SZ_Q4_K 144
#define ORC_SZ_Q8_K 292
#define ORC_SZ_Q5_K 176   /* fp16 d, dmin; uint8 scales[12]; uint8 qh[32]; uint8 qs[128] (ggml-common.h:312-323) */
#define ORC_SZ_Q6_K 210   /* uint8 ql[128]; uint8 qh[64]; int8 scales[16]; fp16 d (ggml-common.h:335-343) */

uint16_t orc_fp32_to_fp16(float f);   /* ggml/src/ggml-impl.h ggml_compute_fp32_to_fp16 (IEEE RNE) */
float    orc_fp16_to_fp32(uint16_t h);

size_t orc_row_size(int type, int64_t k);     /* ggml_row_size */

/* activation quantizers (the *_ref functions are the bit-exactness target) */
void orc_quantize_row_q8_0(const float *x, void *y, int64_t k);  /* ggml-quants.c:199-222 */
void orc_quantize_row_q8_1(const float *x, void *y, int64_t k);  /* ggml-quants.c:225-258 */
void orc_quantize_row_q8_K(const float *x, void *y, int64_t k);  /* ggml-quants.c:2555-2592 */

/* weight quantizer with a closed form (q4_K's search-based quantizer is taken from oracle/_ref) */
void orc_quantize_row_q4_0(const float *x, void *y, int64_t k);  /* ggml-quants.c:36-72 */

/* dequantizers */
void orc_dequantize_row_q4_0(const void *x, float *y, int64_t k); /* ggml-quants.c:307-325 */
void orc_dequantize_row_q8_0(const void *x, float *y, int64_t k); /* ggml-quants.c:370-384 */
void orc_dequantize_row_q4_K(const void *x, float *y, int64_t k); /* ggml-quants.c:1352-1373 */
void orc_dequantize_row_q5_K(const void *x, float *y, int64_t k); /* ggml-quants.c dequantize_row_q5_K */
void orc_dequantize_row_q6_K(const void *x, float *y, int64_t k); /* ggml-quants.c:1762-1791 */

/* dot products, generic (scalar) versions */
float orc_vec_dot_q4_0_q8_0(int64_t k, const void *vx, const void *vy); /* ggml-cpu/quants.c:115-150 */
float orc_vec_dot_q8_0_q8_0(int64_t k, const void *vx, const void *vy); /* ggml-cpu/quants.c:305-333 */
float orc_vec_dot_q4_K_q8_K(int64_t k, const void *vx, const void *vy); /* ggml-cpu/quants.c:550-623 */
float orc_vec_dot_q5_K_q8_K(int64_t k, const void *vx, const void *vy); /* ggml-cpu/quants.c:625-703 */
float orc_vec_dot_q6_K_q8_K(int64_t k, const void *vx, const void *vy); /* ggml-cpu/quants.c:705-758 */

/* GGML_OP_MUL_MAT as ggml-cpu computes it (ggml-cpu.c:1202-1394, 1112-1200):
 * W: ne02*ne03 matrices of M rows; row r of matrix (i2,i3) at W + r*nb01 + i2*nb02 + i3*nb03 (bytes).
 * X: F32, ne12*ne13 matrices of N rows of K floats, strides nb11/nb12/nb13 bytes.
 * dst: F32 contiguous [M, N, ne12, ne13]; src0 broadcast over ne12/ne02, ne13/ne03. */
void orc_mul_mat(int type, const void *W, int64_t M, int64_t K, int64_t ne02, int64_t ne03,
                 size_t nb01, size_t nb02, size_t nb03,
                 const float *X, int64_t N, int64_t ne12, int64_t ne13,
                 size_t nb11, size_t nb12, size_t nb13,
                 float *dst);

/* GGML_OP_MUL_MAT_ID (ggml-cpu.c:1476-1649, 1405-1466):
 * as: n_as expert matrices [M rows x K], expert e at W + e*nb02; b: F32 [K, ne11, n_tok];
 * ids: int32 [n_used, n_tok] (row stride ids_nb1 bytes); dst F32 contiguous [M, n_used, n_tok]:
 * dst[:, id, t] = as[ids[id,t]] . b[:, id % ne11, t]. */
void orc_mul_mat_id(int type, const void *W, int64_t M, int64_t K, int64_t n_as, size_t nb01, size_t nb02,
                    const float *B, int64_t ne11, int64_t n_tok, size_t nb11, size_t nb12,
                    const int32_t *ids, int64_t n_used, size_t ids_nb1,
                    float *dst);

/* K7 element-wise nodes around the MUL_MATs. glu_op = enum ggml_glu_op (ggml.h:596-605):
 * 0 REGLU, 1 GEGLU, 2 SWIGLU, 3 SWIGLU_OAI, 4 GEGLU_ERF, 5 GEGLU_QUICK. x = gate branch (src0), g = linear branch (src1).
 * Scalar float forms: ggml-cpu/vec.h:873-875 (gelu), 958-960 (silu), ops.cpp:3254-3259 (swiglu_oai), vec.h reglu /
 * geglu_erf / geglu_quick. REGLU, SWIGLU_OAI and GEGLU_ERF are bit-identical to the compiled reference; ggml-cpu's vector paths use
 * a polynomial expf for SWIGLU (ggml_v_silu, vec.cpp:376-410) and fp16 lookup tables for GEGLU / GEGLU_QUICK (vec.h:1326-1340), which
 * these float formulas are the definition of: tests/test_fusion_k7.py pins those three at NMSE <= 1e-6 (measured <= 7e-8). */
float orc_glu_single(int glu_op, float x, float g, float alpha, float limit);
void  orc_glu(int glu_op, float alpha, float limit, const float *gate, const float *up, float *dst, int64_t n);
/* dst[r][c] = a[r][c] + bias[c]  (GGML_OP_ADD of a [M] bias row onto [M, nrows]; ggml-cpu/binary-ops.cpp) */
void  orc_add_row(const float *a, const float *bias, float *dst, int64_t M, int64_t nrows);
/* GGML_OP_ADD_ID (ggml-cpu/ops.cpp:701-750): dst[:, i, t] = a[:, i, t] + bias[:, ids[i, t]]; all contiguous */
void  orc_add_id(const float *a, const float *bias, const int32_t *ids, float *dst, int64_t M, int64_t n_used, int64_t n_tok, int64_t n_as);

/* float64 reference of the same op on dequantized weights and UNquantized activations
 * (for NMSE context only; not the parity target) */
void orc_mul_mat_f64(int type, const void *W, int64_t M, int64_t K, size_t nb01,
                     const float *X, int64_t N, size_t nb11, double *dst);

/* NMSE as tests/test-backend-ops.cpp:274-287 : sum((a-b)^2)/sum(a^2), a = reference */
double orc_nmse(const float *a, const float *b, size_t n);

#ifdef __cplusplus
}
#endif

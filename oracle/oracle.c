/* oracle/oracle.c — TEST INFRASTRUCTURE ONLY (not product code). See oracle.h.
 *
 * Scalar restatement of what ggml-cpu computes for MUL_MAT / MUL_MAT_ID on
 * q4_0 / q8_0 / q4_K weights: quantize each activation row to the CPU's
 * vec_dot_type (q8_0 for q4_0/q8_0 weights, q8_K for q4_K weights;
 * ggml-cpu/ggml-cpu.c:209-218, 244-253, 279-288), then one integer block dot
 * product per (weight row, activation row) accumulated in fp32 in block order.
 *
 * Parity: PINNED against oracle/_ref (the reference's own compiled code) by
 * tests/test_oracle.py and against tests/golden/ fixtures.
 * Compile with -ffp-contract=off -fno-fast-math (see Makefile): the reference's
 * scalar code is not FMA-contracted in the x64 build either.
 */
#include "oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ---------- fp16 <-> fp32, IEEE-754 binary16, round-to-nearest-even ---------- */

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float    u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

uint16_t orc_fp32_to_fp16(float f) {
    const uint32_t u    = f2u(f);
    const uint16_t sign = (uint16_t)((u >> 16) & 0x8000u);
    const uint32_t a    = u & 0x7fffffffu;
    if (a > 0x7f800000u) return (uint16_t)(sign | 0x7e00u);        /* NaN (ggml-impl.h:424 returns 0x7E00) */
    if (a >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u);       /* >= 65520 rounds to inf */
    if (a < 0x33000001u) return sign;                              /* <= 2^-25 rounds to zero */
    int32_t  e = (int32_t)(a >> 23) - 127;
    uint32_t m = (a & 0x7fffffu) | 0x800000u;                      /* 24-bit significand */
    int shift;                                                     /* bits to drop */
    uint32_t base;
    if (e < -14) { shift = 13 + (-14 - e); base = 0; }             /* subnormal half */
    else         { shift = 13;             base = (uint32_t)(e + 15) << 10; m &= 0x7fffffu; }
    uint32_t q    = m >> shift;
    uint32_t rem  = m & ((1u << shift) - 1u);
    uint32_t half = 1u << (shift - 1);
    if (rem > half || (rem == half && (q & 1u))) q++;              /* RNE; carry propagates into exponent */
    return (uint16_t)(sign | (base + q));
}

float orc_fp16_to_fp32(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    const uint32_t e    = (h >> 10) & 0x1fu;
    const uint32_t m    = h & 0x3ffu;
    if (e == 0) {
        if (m == 0) return u2f(sign);
        float v = (float)m * 0x1.0p-24f;                           /* subnormal */
        return (sign ? -v : v);
    }
    if (e == 31) return u2f(sign | 0x7f800000u | (m << 13));
    return u2f(sign | ((e + 112u) << 23) | (m << 13));
}

size_t orc_row_size(int type, int64_t k) {
    switch (type) {
        case ORC_Q4_0: return (size_t)(k / ORC_QK) * ORC_SZ_Q4_0;
        case ORC_Q8_0: return (size_t)(k / ORC_QK) * ORC_SZ_Q8_0;
        case ORC_Q4_K: return (size_t)(k / ORC_QK_K) * ORC_SZ_Q4_K;
        case ORC_Q5_K: return (size_t)(k / ORC_QK_K) * ORC_SZ_Q5_K;
        case ORC_Q6_K: return (size_t)(k / ORC_QK_K) * ORC_SZ_Q6_K;
    }
    return 0;
}

/* round-half-to-even on |v| <= 4194303 via the 1.5*2^23 magic add (ggml-quants.c:444-449) */
static inline int nearest_int_magic(float v) {
    float t = v + 12582912.f;
    int32_t i; memcpy(&i, &t, 4);
    return (i & 0x007fffff) - 0x00400000;
}

/* ---------- activation quantizers ---------- */

/* block_q8_0 = { fp16 d; int8 qs[32] } ; d = amax/127, q = roundf(x * (1/d))  (ggml-quants.c:199-222) */
void orc_quantize_row_q8_0(const float *x, void *vy, int64_t k) {
    uint8_t *y = (uint8_t *)vy;
    for (int64_t b = 0; b < k / ORC_QK; b++, x += ORC_QK, y += ORC_SZ_Q8_0) {
        float amax = 0.0f;
        for (int j = 0; j < ORC_QK; j++) { float a = fabsf(x[j]); if (a > amax) amax = a; }
        const float d  = amax / 127;
        const float id = d ? 1.0f / d : 0.0f;
        uint16_t hd = orc_fp32_to_fp16(d);
        memcpy(y, &hd, 2);
        for (int j = 0; j < ORC_QK; j++) ((int8_t *)(y + 2))[j] = (int8_t)roundf(x[j] * id);
    }
}

/* block_q8_1 = { fp16 d; fp16 s; int8 qs[32] } ; s = fp16(sum(q) * d)  (ggml-quants.c:225-258) */
void orc_quantize_row_q8_1(const float *x, void *vy, int64_t k) {
    uint8_t *y = (uint8_t *)vy;
    for (int64_t b = 0; b < k / ORC_QK; b++, x += ORC_QK, y += ORC_SZ_Q8_1) {
        float amax = 0.0f;
        for (int j = 0; j < ORC_QK; j++) { float a = fabsf(x[j]); if (a > amax) amax = a; }
        const float d  = amax / 127;
        const float id = d ? 1.0f / d : 0.0f;
        int sum = 0;
        for (int j = 0; j < ORC_QK; j++) {
            int8_t q = (int8_t)roundf(x[j] * id);
            ((int8_t *)(y + 4))[j] = q;
            sum += q;
        }
        uint16_t hd = orc_fp32_to_fp16(d), hs = orc_fp32_to_fp16(sum * d);
        memcpy(y, &hd, 2);
        memcpy(y + 2, &hs, 2);
    }
}

/* block_q8_K = { float d; int8 qs[256]; int16 bsums[16] }  (ggml-quants.c:2555-2592)
 * max = the signed x of largest |x| (first one on ties), iscale = -127/max,
 * q = min(127, nearest_int(iscale*x)), d = 1/iscale. An all-zero block stores d=0, q=0 and
 * leaves bsums untouched in the reference; ggml-cpu zero-fills nothing there, so the bsums of such
 * a block are unspecified — we write zeros and the comparison masks them (tests/test_oracle.py). */
void orc_quantize_row_q8_K(const float *x, void *vy, int64_t k) {
    uint8_t *y = (uint8_t *)vy;
    for (int64_t b = 0; b < k / ORC_QK_K; b++, x += ORC_QK_K, y += ORC_SZ_Q8_K) {
        float  *pd  = (float *)y;
        int8_t *qs  = (int8_t *)(y + 4);
        int16_t bs[16];
        float max = 0, amax = 0;
        for (int j = 0; j < ORC_QK_K; j++) { float ax = fabsf(x[j]); if (ax > amax) { amax = ax; max = x[j]; } }
        if (!amax) {
            *pd = 0; memset(qs, 0, ORC_QK_K); memset(y + 4 + ORC_QK_K, 0, 32);
            continue;
        }
        const float iscale = -127.f / max;
        for (int j = 0; j < ORC_QK_K; j++) {
            int v = nearest_int_magic(iscale * x[j]);
            qs[j] = (int8_t)(v < 127 ? v : 127);
        }
        for (int j = 0; j < 16; j++) {
            int s = 0;
            for (int i = 0; i < 16; i++) s += qs[16 * j + i];
            bs[j] = (int16_t)s;
        }
        memcpy(y + 4 + ORC_QK_K, bs, 32);
        *pd = 1 / iscale;
    }
}

/* ---------- weight quantizer with a closed form ---------- */

/* block_q4_0 = { fp16 d; uint8 qs[16] } ; d = max/-8 with max the signed value of largest |x|
 * (ggml-quants.c:36-72) */
void orc_quantize_row_q4_0(const float *x, void *vy, int64_t k) {
    uint8_t *y = (uint8_t *)vy;
    for (int64_t b = 0; b < k / ORC_QK; b++, x += ORC_QK, y += ORC_SZ_Q4_0) {
        float amax = 0.0f, max = 0.0f;
        for (int j = 0; j < ORC_QK; j++) { if (amax < fabsf(x[j])) { amax = fabsf(x[j]); max = x[j]; } }
        const float d  = max / -8;
        const float id = d ? 1.0f / d : 0.0f;
        uint16_t hd = orc_fp32_to_fp16(d);
        memcpy(y, &hd, 2);
        for (int j = 0; j < 16; j++) {
            int a = (int8_t)(x[j] * id + 8.5f);       if (a > 15) a = 15;
            int c = (int8_t)(x[j + 16] * id + 8.5f);  if (c > 15) c = 15;
            y[2 + j] = (uint8_t)((a & 0xff) | ((c & 0xff) << 4));
        }
    }
}

/* ---------- dequantizers ---------- */

void orc_dequantize_row_q4_0(const void *vx, float *y, int64_t k) {
    const uint8_t *x = (const uint8_t *)vx;
    for (int64_t b = 0; b < k / ORC_QK; b++, x += ORC_SZ_Q4_0, y += ORC_QK) {
        uint16_t hd; memcpy(&hd, x, 2);
        const float d = orc_fp16_to_fp32(hd);
        for (int j = 0; j < 16; j++) {
            y[j]      = ((x[2 + j] & 0x0f) - 8) * d;
            y[j + 16] = ((x[2 + j] >> 4) - 8) * d;
        }
    }
}

void orc_dequantize_row_q8_0(const void *vx, float *y, int64_t k) {
    const uint8_t *x = (const uint8_t *)vx;
    for (int64_t b = 0; b < k / ORC_QK; b++, x += ORC_SZ_Q8_0, y += ORC_QK) {
        uint16_t hd; memcpy(&hd, x, 2);
        const float d = orc_fp16_to_fp32(hd);
        for (int j = 0; j < ORC_QK; j++) y[j] = ((const int8_t *)(x + 2))[j] * d;
    }
}

/* 6-bit (scale, min) pair j of a q4_K block's 12 packed bytes (ggml-quants.c:703-711) */
static inline void q4k_scale_min(int j, const uint8_t *q, uint8_t *sc, uint8_t *mn) {
    if (j < 4) { *sc = q[j] & 63;                             *mn = q[j + 4] & 63; }
    else       { *sc = (q[j + 4] & 0x0f) | ((q[j - 4] >> 6) << 4); *mn = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}

/* block_q4_K = { fp16 d; fp16 dmin; uint8 scales[12]; uint8 qs[128] }  (ggml-quants.c:1352-1373) */
void orc_dequantize_row_q4_K(const void *vx, float *y, int64_t k) {
    const uint8_t *x = (const uint8_t *)vx;
    for (int64_t b = 0; b < k / ORC_QK_K; b++, x += ORC_SZ_Q4_K) {
        uint16_t hd, hm; memcpy(&hd, x, 2); memcpy(&hm, x + 2, 2);
        const float d = orc_fp16_to_fp32(hd), dmin = orc_fp16_to_fp32(hm);
        const uint8_t *sc12 = x + 4, *q = x + 16;
        for (int g = 0; g < 4; g++, q += 32) {       /* 64 weights per group: low nibbles then high nibbles */
            uint8_t s, m;
            q4k_scale_min(2 * g, sc12, &s, &m);
            const float d1 = d * s, m1 = dmin * m;
            q4k_scale_min(2 * g + 1, sc12, &s, &m);
            const float d2 = d * s, m2 = dmin * m;
            for (int l = 0; l < 32; l++) *y++ = d1 * (q[l] & 0x0f) - m1;
            for (int l = 0; l < 32; l++) *y++ = d2 * (q[l] >> 4) - m2;
        }
    }
}

/* ---------- dot products (generic scalar semantics) ---------- */

float orc_vec_dot_q4_0_q8_0(int64_t k, const void *vx, const void *vy) {
    const uint8_t *x = (const uint8_t *)vx, *y = (const uint8_t *)vy;
    float sumf = 0;
    for (int64_t b = 0; b < k / ORC_QK; b++, x += ORC_SZ_Q4_0, y += ORC_SZ_Q8_0) {
        const int8_t *q8 = (const int8_t *)(y + 2);
        int s0 = 0, s1 = 0;
        for (int j = 0; j < 16; j++) {
            s0 += ((x[2 + j] & 0x0f) - 8) * q8[j];
            s1 += ((x[2 + j] >> 4) - 8) * q8[j + 16];
        }
        uint16_t hx, hy; memcpy(&hx, x, 2); memcpy(&hy, y, 2);
        /* left-to-right: (sumi*dx)*dy, ggml-cpu/quants.c:146 */
        sumf += (s0 + s1) * orc_fp16_to_fp32(hx) * orc_fp16_to_fp32(hy);
    }
    return sumf;
}

float orc_vec_dot_q8_0_q8_0(int64_t k, const void *vx, const void *vy) {
    const uint8_t *x = (const uint8_t *)vx, *y = (const uint8_t *)vy;
    float sumf = 0;
    for (int64_t b = 0; b < k / ORC_QK; b++, x += ORC_SZ_Q8_0, y += ORC_SZ_Q8_0) {
        const int8_t *a = (const int8_t *)(x + 2), *c = (const int8_t *)(y + 2);
        int s = 0;
        for (int j = 0; j < ORC_QK; j++) s += a[j] * c[j];
        uint16_t hx, hy; memcpy(&hx, x, 2); memcpy(&hy, y, 2);
        /* sumi*(dx*dy), ggml-cpu/quants.c:329 */
        sumf += s * (orc_fp16_to_fp32(hx) * orc_fp16_to_fp32(hy));
    }
    return sumf;
}

/* ggml-cpu/quants.c:550-623: per 256-block, 8 interleaved int32 lanes scaled by the 6-bit sub-block
 * scale, then sums[l] += (d*y.d)*aux32[l]; min term sumf -= (dmin*y.d) * sum_j bsums[j]*mins[j/2];
 * final sumf += sum_l sums[l]. The lane structure is kept so that fp32 rounding matches the generic code. */
float orc_vec_dot_q4_K_q8_K(int64_t k, const void *vx, const void *vy) {
    const uint8_t *x = (const uint8_t *)vx, *y = (const uint8_t *)vy;
    float sums[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float sumf = 0;
    for (int64_t b = 0; b < k / ORC_QK_K; b++, x += ORC_SZ_Q4_K, y += ORC_SZ_Q8_K) {
        float yd; memcpy(&yd, y, 4);
        const int8_t *q8 = (const int8_t *)(y + 4);
        int16_t bsums[16]; memcpy(bsums, y + 4 + ORC_QK_K, 32);
        uint16_t hd, hm; memcpy(&hd, x, 2); memcpy(&hm, x + 2, 2);
        const uint8_t *sc12 = x + 4, *q4 = x + 16;
        int32_t lane[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        int summin = 0;
        for (int s = 0; s < 8; s++) {                 /* sub-block s: 32 weights */
            uint8_t sc, mn;
            q4k_scale_min(s, sc12, &sc, &mn);
            summin += (bsums[2 * s] + bsums[2 * s + 1]) * mn;
            const uint8_t *qq = q4 + 32 * (s / 2);
            for (int l = 0; l < 32; l++) {
                int w = (s & 1) ? (qq[l] >> 4) : (qq[l] & 0x0f);
                lane[l & 7] += (int32_t)sc * (int16_t)(q8[32 * s + l] * w);
            }
        }
        const float d = orc_fp16_to_fp32(hd) * yd;
        for (int l = 0; l < 8; l++) sums[l] += d * lane[l];
        const float dmin = orc_fp16_to_fp32(hm) * yd;
        sumf -= dmin * summin;
    }
    for (int l = 0; l < 8; l++) sumf += sums[l];
    return sumf;
}

/* block_q5_K = { fp16 d, dmin; uint8 scales[12]; uint8 qh[32]; uint8 qs[128] } (ggml-common.h:312-323): q4_K's layout plus a fifth
 * bit: sub-block s (32 weights) takes the low (s even) / high (s odd) nibbles of qs[32*(s/2) + l] and bit s of qh[l]
 * (dequantize_row_q5_K, ggml-quants.c; vec_dot: ggml-cpu/quants.c:625-703, same lane structure as q4_K) */
void orc_dequantize_row_q5_K(const void *vx, float *y, int64_t k) {
    const uint8_t *x = (const uint8_t *)vx;
    for (int64_t b = 0; b < k / ORC_QK_K; b++, x += ORC_SZ_Q5_K) {
        uint16_t hd, hm; memcpy(&hd, x, 2); memcpy(&hm, x + 2, 2);
        const float d = orc_fp16_to_fp32(hd), dmin = orc_fp16_to_fp32(hm);
        const uint8_t *sc12 = x + 4, *qh = x + 16, *q = x + 48;
        for (int g = 0; g < 4; g++, q += 32) {
            uint8_t s, m;
            q4k_scale_min(2 * g, sc12, &s, &m);
            const float d1 = d * s, m1 = dmin * m;
            q4k_scale_min(2 * g + 1, sc12, &s, &m);
            const float d2 = d * s, m2 = dmin * m;
            for (int l = 0; l < 32; l++) *y++ = d1 * ((q[l] & 0x0f) + ((qh[l] >> (2 * g)) & 1 ? 16 : 0)) - m1;
            for (int l = 0; l < 32; l++) *y++ = d2 * ((q[l] >> 4) + ((qh[l] >> (2 * g + 1)) & 1 ? 16 : 0)) - m2;
        }
    }
}

float orc_vec_dot_q5_K_q8_K(int64_t k, const void *vx, const void *vy) {
    const uint8_t *x = (const uint8_t *)vx, *y = (const uint8_t *)vy;
    float sums[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float sumf = 0;
    for (int64_t b = 0; b < k / ORC_QK_K; b++, x += ORC_SZ_Q5_K, y += ORC_SZ_Q8_K) {
        float yd; memcpy(&yd, y, 4);
        const int8_t *q8 = (const int8_t *)(y + 4);
        int16_t bsums[16]; memcpy(bsums, y + 4 + ORC_QK_K, 32);
        uint16_t hd, hm; memcpy(&hd, x, 2); memcpy(&hm, x + 2, 2);
        const uint8_t *sc12 = x + 4, *qh = x + 16, *q4 = x + 48;
        int32_t lane[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        int summin = 0;
        for (int s = 0; s < 8; s++) {                 /* sub-block s: 32 weights */
            uint8_t sc, mn;
            q4k_scale_min(s, sc12, &sc, &mn);
            summin += (bsums[2 * s] + bsums[2 * s + 1]) * mn;
            const uint8_t *qq = q4 + 32 * (s / 2);
            for (int l = 0; l < 32; l++) {
                int w = ((s & 1) ? (qq[l] >> 4) : (qq[l] & 0x0f)) + (((qh[l] >> s) & 1) ? 16 : 0);
                lane[l & 7] += (int32_t)sc * (int16_t)(q8[32 * s + l] * w);
            }
        }
        const float d = orc_fp16_to_fp32(hd) * yd;
        for (int l = 0; l < 8; l++) sums[l] += d * lane[l];
        const float dmin = orc_fp16_to_fp32(hm) * yd;
        sumf -= dmin * summin;
    }
    for (int l = 0; l < 8; l++) sumf += sums[l];
    return sumf;
}

/* block_q6_K = { uint8 ql[128]; uint8 qh[64]; int8 scales[16]; fp16 d } (ggml-common.h:335-343). Weight l of a 128-half:
 * low nibble / high nibble of ql[l], ql[l+32] plus two bits of qh[l], minus 32; one int8 scale per 16 weights
 * (ggml-quants.c:1762-1791) */
static void q6k_unpack(const uint8_t *x, int8_t *a) {
    const uint8_t *ql = x, *qh = x + 128;
    for (int n = 0; n < ORC_QK_K; n += 128, a += 128, ql += 64, qh += 32) {
        for (int l = 0; l < 32; l++) {
            a[l +  0] = (int8_t)((ql[l +  0] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
            a[l + 32] = (int8_t)((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
            a[l + 64] = (int8_t)((ql[l +  0] >>  4) | (((qh[l] >> 4) & 3) << 4)) - 32;
            a[l + 96] = (int8_t)((ql[l + 32] >>  4) | (((qh[l] >> 6) & 3) << 4)) - 32;
        }
    }
}
void orc_dequantize_row_q6_K(const void *vx, float *y, int64_t k) {
    const uint8_t *x = (const uint8_t *)vx;
    for (int64_t b = 0; b < k / ORC_QK_K; b++, x += ORC_SZ_Q6_K, y += ORC_QK_K) {
        uint16_t hd; memcpy(&hd, x + 208, 2);
        const float d = orc_fp16_to_fp32(hd);
        const int8_t *sc = (const int8_t *)(x + 192);
        int8_t a[ORC_QK_K];
        q6k_unpack(x, a);
        for (int i = 0; i < ORC_QK_K; i++) y[i] = d * sc[i / 16] * a[i];   /* (d * sc) * q, left to right as :1780-1783 */
    }
}
/* ggml-cpu/quants.c:705-758: 8 interleaved int32 lanes, each 16-group scaled by its int8 scale, sums[l] += (d*y.d)*aux32[l],
 * sumf = sum_l sums[l] at the very end. The lane structure is kept so that fp32 rounding matches the generic code. */
float orc_vec_dot_q6_K_q8_K(int64_t k, const void *vx, const void *vy) {
    const uint8_t *x = (const uint8_t *)vx, *y = (const uint8_t *)vy;
    float sums[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int64_t b = 0; b < k / ORC_QK_K; b++, x += ORC_SZ_Q6_K, y += ORC_SZ_Q8_K) {
        float yd; memcpy(&yd, y, 4);
        const int8_t *q8 = (const int8_t *)(y + 4);
        const int8_t *sc = (const int8_t *)(x + 192);
        uint16_t hd; memcpy(&hd, x + 208, 2);
        int8_t a[ORC_QK_K];
        q6k_unpack(x, a);
        int32_t lane[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int i = 0; i < ORC_QK_K; i++) lane[i & 7] += (int32_t)sc[i / 16] * (int16_t)(q8[i] * a[i]);
        const float d = orc_fp16_to_fp32(hd) * yd;
        for (int l = 0; l < 8; l++) sums[l] += d * lane[l];
    }
    float sumf = 0;
    for (int l = 0; l < 8; l++) sumf += sums[l];
    return sumf;
}

/* ---------- the ops ---------- */

static size_t act_row_size(int type, int64_t k) {
    return (type == ORC_Q4_K || type == ORC_Q5_K || type == ORC_Q6_K) ? (size_t)(k / ORC_QK_K) * ORC_SZ_Q8_K : (size_t)(k / ORC_QK) * ORC_SZ_Q8_0;
}
static void act_quantize(int type, const float *x, void *y, int64_t k) {
    if (type == ORC_Q4_K || type == ORC_Q5_K || type == ORC_Q6_K) orc_quantize_row_q8_K(x, y, k); else orc_quantize_row_q8_0(x, y, k);
}
static float dot(int type, int64_t k, const void *w, const void *a) {
    switch (type) {
        case ORC_Q4_0: return orc_vec_dot_q4_0_q8_0(k, w, a);
        case ORC_Q8_0: return orc_vec_dot_q8_0_q8_0(k, w, a);
        case ORC_Q5_K: return orc_vec_dot_q5_K_q8_K(k, w, a);
        case ORC_Q6_K: return orc_vec_dot_q6_K_q8_K(k, w, a);
        default:       return orc_vec_dot_q4_K_q8_K(k, w, a);
    }
}

void orc_mul_mat(int type, const void *W, int64_t M, int64_t K, int64_t ne02, int64_t ne03,
                 size_t nb01, size_t nb02, size_t nb03,
                 const float *X, int64_t N, int64_t ne12, int64_t ne13,
                 size_t nb11, size_t nb12, size_t nb13,
                 float *dst) {
    const size_t rs = act_row_size(type, K);
    uint8_t *act = (uint8_t *)malloc(rs ? rs : 1);
    const int64_t r2 = ne12 / ne02, r3 = ne13 / ne03;        /* ggml-cpu.c:1133-1134 */
    for (int64_t i13 = 0; i13 < ne13; i13++)
    for (int64_t i12 = 0; i12 < ne12; i12++)
    for (int64_t i11 = 0; i11 < N; i11++) {
        const float *xrow = (const float *)((const char *)X + i11 * nb11 + i12 * nb12 + i13 * nb13);
        act_quantize(type, xrow, act, K);
        const char *w = (const char *)W + (i12 / r2) * nb02 + (i13 / r3) * nb03;
        float *out = dst + ((i13 * ne12 + i12) * N + i11) * M;
        for (int64_t r = 0; r < M; r++) out[r] = dot(type, K, w + r * nb01, act);
    }
    free(act);
}

void orc_mul_mat_id(int type, const void *W, int64_t M, int64_t K, int64_t n_as, size_t nb01, size_t nb02,
                    const float *B, int64_t ne11, int64_t n_tok, size_t nb11, size_t nb12,
                    const int32_t *ids, int64_t n_used, size_t ids_nb1,
                    float *dst) {
    const size_t rs = act_row_size(type, K);
    uint8_t *act = (uint8_t *)malloc(rs ? rs : 1);
    for (int64_t t = 0; t < n_tok; t++)
    for (int64_t id = 0; id < n_used; id++) {
        const int32_t e = *(const int32_t *)((const char *)ids + t * ids_nb1 + id * 4);
        if (e < 0 || e >= n_as) abort();                         /* ggml-cpu.c:1574 asserts */
        const float *brow = (const float *)((const char *)B + (id % ne11) * nb11 + t * nb12);
        act_quantize(type, brow, act, K);
        const char *w = (const char *)W + (size_t)e * nb02;
        float *out = dst + (t * n_used + id) * M;
        for (int64_t r = 0; r < M; r++) out[r] = dot(type, K, w + r * nb01, act);
    }
    free(act);
}

/* ---------------------------------------------------------------- K7: GLU / ADD / ADD_ID ---- */

float orc_glu_single(int op, float x, float g, float alpha, float limit) {
    switch (op) {
        case 0: return x > 0.0f ? x * g : 0.0f;                                       /* ggml_vec_reglu_f32 */
        case 1: return 0.5f * x * (1.0f + tanhf(0.79788456080286535587989211986876f * x * (1.0f + 0.044715f * x * x))) * g; /* vec.h:873-875 */
        case 2: return x / (1.0f + expf(-x)) * g;                                       /* vec.h:958-960 */
        case 3: {                                                                       /* ops.cpp:3254-3259 */
            const float xc = x < limit ? x : limit;
            const float gc = g < -limit ? -limit : (g > limit ? limit : g);
            return xc / (1.0f + expf(alpha * (-xc))) * (gc + 1.0f);
        }
        case 4: return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440084436210484f)) * g;
        default: return x * (1.0f / (1.0f + expf(-1.702f * x))) * g;
    }
}

void orc_glu(int op, float alpha, float limit, const float *gate, const float *up, float *dst, int64_t n) {
    for (int64_t i = 0; i < n; i++) dst[i] = orc_glu_single(op, gate[i], up[i], alpha, limit);
}

void orc_add_row(const float *a, const float *bias, float *dst, int64_t M, int64_t nrows) {
    for (int64_t r = 0; r < nrows; r++)
        for (int64_t c = 0; c < M; c++) dst[r * M + c] = a[r * M + c] + bias[c];
}

void orc_add_id(const float *a, const float *bias, const int32_t *ids, float *dst, int64_t M, int64_t n_used, int64_t n_tok, int64_t n_as) {
    for (int64_t t = 0; t < n_tok; t++)
        for (int64_t i = 0; i < n_used; i++) {
            const int32_t e = ids[t * n_used + i];
            if (e < 0 || e >= n_as) abort();                                            /* ops.cpp:743 asserts */
            for (int64_t c = 0; c < M; c++) dst[(t * n_used + i) * M + c] = a[(t * n_used + i) * M + c] + bias[(int64_t)e * M + c];
        }
}

void orc_mul_mat_f64(int type, const void *W, int64_t M, int64_t K, size_t nb01,
                     const float *X, int64_t N, size_t nb11, double *dst) {
    float *wf = (float *)malloc(sizeof(float) * (size_t)K);
    for (int64_t r = 0; r < M; r++) {
        const void *row = (const char *)W + r * nb01;
        if (type == ORC_Q4_0) orc_dequantize_row_q4_0(row, wf, K);
        else if (type == ORC_Q8_0) orc_dequantize_row_q8_0(row, wf, K);
        else if (type == ORC_Q5_K) orc_dequantize_row_q5_K(row, wf, K);
        else if (type == ORC_Q6_K) orc_dequantize_row_q6_K(row, wf, K);
        else orc_dequantize_row_q4_K(row, wf, K);
        for (int64_t n = 0; n < N; n++) {
            const float *x = (const float *)((const char *)X + n * nb11);
            double s = 0;
            for (int64_t i = 0; i < K; i++) s += (double)wf[i] * (double)x[i];
            dst[n * M + r] = s;
        }
    }
    free(wf);
}

double orc_nmse(const float *a, const float *b, size_t n) {
    double e = 0.0, r = 0.0;
    for (size_t i = 0; i < n; i++) {
        float d = a[i] - b[i];
        e += d * d;          /* float product promoted, as tests/test-backend-ops.cpp:282 */
        r += a[i] * a[i];
    }
    return e / r;
}

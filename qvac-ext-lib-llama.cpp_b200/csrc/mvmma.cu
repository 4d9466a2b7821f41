// mvmma.cu — K3b: batched decode (3..8 activation columns) on the warp-level int8 tensor-core path.
//
// The dp4a mat-vec kernel (mmvq.cu) spends ~100 instructions per 256 weights and column: beyond 2 columns it is instruction
// bound (n = 4: 2.1 TB/s, n = 8: 1.2 TB/s). Here one mma.sync.m16n8k32 (SASS IMMA.16832) forms 16 weight rows x 8 columns x 32
// weights at once, A fragments unpacked from the 4-bit stream in registers (one 32-bit shared load holds the nibbles of both
// k-halves a lane needs), so the instruction count per weight byte is independent of the column count and the kernel stays HBM
// bound. Replaces mul_mat_vec_q for ncols_dst 3..8 (/root/reference/ggml/src/ggml-cuda/mmvq.cu:141-335) and is the building block
// for small MUL_MAT_ID batches.
//
// Geometry: a CTA owns a contiguous range of 16-row tiles. A tile's K range is streamed in chunks of NWC units (a unit = 256
// weights = 144 B of q4_0 / q4_K); one 3-D TMA box [units][16 rows][144 B] per chunk lands so that consumer warp w finds "its"
// unit as a dense [16][144 B] sub-tile (36-word row pitch: the 8 rows x 4 words a fragment load touches hit 32 distinct banks).
// Warp w multiplies unit w of every chunk for all 16 rows; the NWC partial 16x8 results of a tile are summed in a fixed order
// through shared memory (bit-reproducible run to run). One producer warp issues the TMA loads through a ring of stages.
// Integer arithmetic is the CPU reference's: per 32-block integer dot products, 6-bit scales / block scales applied exactly as
// vec_dot_q4_K_q8_K / vec_dot_q4_0_q8_0 do (ggml-cpu/quants.c), activations quantized bit-exactly by quantize.cu.

#include "internal.h"
#include "tc.cuh"
#include "quant_core.cuh"

#include <cuda.h>

using namespace b200;

namespace b200 {

struct MmaMvParams {
    MvMat           mat[MV_MAX_MATS];   // group0 = first global tile id of the matrix
    int             n_mats, total_tiles;
    int             U, ncols, n_cons, n_stages, pdl;
    const uint8_t * act_main;           // [ncols][U][272]
    const uint8_t * act_aux;            // [ncols][U][16]
    const float *   x;                  // fuse_quant: F32 activations, column c at (char*)x + c*x_cs - quantized by the kernel itself
    int64_t         x_cs;
    int             fuse_quant, x_vec;
    SplitParams     split;
    SplitGeom       geom;               // split.ll: who owns which rows of mat[i]
    // MUL_MAT_ID mode (sp != nullptr): blockIdx.y = expert e; its (slot, token) pairs sp[eb[e] .. eb[e+1]) are the columns, 8 at a
    // time; pair p = t * n_used + i reads activation row t * ne11 + i % ne11 and writes dst + t * dst_ts + i * dst_is
    const int32_t * sp;                 // pairs sorted by expert (b200_mm_ids_sort)
    const int32_t * eb;                 // [n_as + 1] first sorted position of every expert
    MvIds           ids;
};

namespace {

constexpr int MMA_ROWS = 16;
constexpr int MMA_RT   = 2;   // 16-row tiles a warp accumulates before the CTA-wide reduction (one barrier per 64 rows)

__device__ __forceinline__ void mma_u8s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
                 : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0), "r"(0), "r"(0), "r"(0));
}

__device__ __forceinline__ void mma_s8s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
                 : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0), "r"(0), "r"(0), "r"(0));
}

__device__ __forceinline__ void tma_load_3d(uint32_t dst, const void * tmap, int c0, int c1, int c2, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
        "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
        : "memory");
}

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const void * tmap, int c0, int c1, int c2, int c3, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];" ::"r"(dst),
        "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(bar)
        : "memory");
}

// ---- one unit (256 weights) of 16 rows against 8 columns. wt = shared address of the [16][144 B] sub-tile, am / ax = shared
// addresses of the unit's activation records of column 0 (column stride cs_m / cs_x bytes). acc: (g, 2t) (g, 2t+1) (g+8, 2t) (g+8, 2t+1)

__device__ __forceinline__ void unit_mma_q4_K(uint32_t wt, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, int g, int t, float (&acc)[4]) {
    const uint32_t r0 = wt + g * 144, r1 = r0 + 8 * 144;
    const uint4 h0 = lds128(r0), h1 = lds128(r1);  // d | dmin, 12 bytes of 6-bit scales / mins (ggml-common.h:291-305)
    // 8 scales and 8 mins as bytes (ggml-cpu/quants.c:586-591)
    const uint32_t sc0[2] = { h0.y & 0x3f3f3f3fu, (h0.w & 0x0f0f0f0fu) | ((h0.y >> 2) & 0x30303030u) };
    const uint32_t mn0[2] = { h0.z & 0x3f3f3f3fu, ((h0.w >> 4) & 0x0f0f0f0fu) | ((h0.z >> 2) & 0x30303030u) };
    const uint32_t sc1[2] = { h1.y & 0x3f3f3f3fu, (h1.w & 0x0f0f0f0fu) | ((h1.y >> 2) & 0x30303030u) };
    const uint32_t mn1[2] = { h1.z & 0x3f3f3f3fu, ((h1.w >> 4) & 0x0f0f0f0fu) | ((h1.z >> 2) & 0x30303030u) };
    const uint32_t bcol = am + g * cs_m + 4 * t;  // B fragments: column g, k bytes 4t.. and 16 + 4t..
    int isum[4] = { 0, 0, 0, 0 };
#pragma unroll
    for (int i = 0; i < 4; i++) {  // 64 weights: low nibbles = sub-block 2i, high nibbles = sub-block 2i + 1
        const uint32_t wa = lds32(r0 + 16 + 32 * i + 4 * t), wb = lds32(r0 + 32 + 32 * i + 4 * t);
        const uint32_t wc = lds32(r1 + 16 + 32 * i + 4 * t), wd = lds32(r1 + 32 + 32 * i + 4 * t);
#pragma unroll
        for (int hn = 0; hn < 2; hn++) {
            const int j = 2 * i + hn;
            const uint32_t b0 = lds32(bcol + 32 * j), b1 = lds32(bcol + 32 * j + 16);
            int c[4];
            mma_u8s8(c, (wa >> (4 * hn)) & 0x0f0f0f0fu, (wc >> (4 * hn)) & 0x0f0f0f0fu, (wb >> (4 * hn)) & 0x0f0f0f0fu,
                     (wd >> (4 * hn)) & 0x0f0f0f0fu, b0, b1);
            const int s0 = (int) ((sc0[j >> 2] >> (8 * (j & 3))) & 0xff), s1 = (int) ((sc1[j >> 2] >> (8 * (j & 3))) & 0xff);
            isum[0] += s0 * c[0]; isum[1] += s0 * c[1];
            isum[2] += s1 * c[2]; isum[3] += s1 * c[3];
        }
    }
    // min term: sum_j m_j * bsum_j per (row, column) - int16 block sums at +256 of the column's record
    const uint4 bs0 = lds128(am + (2 * t) * cs_m + 256), bs1 = lds128(am + (2 * t + 1) * cs_m + 256);
    auto mdot = [](const uint4 & bs, const uint32_t (&mn)[2]) {
        int v = dp2a_lo_su((int) bs.x, mn[0], 0);
        v     = dp2a_hi_su((int) bs.y, mn[0], v);
        v     = dp2a_lo_su((int) bs.z, mn[1], v);
        return dp2a_hi_su((int) bs.w, mn[1], v);
    };
    const float dy0 = __uint_as_float(lds32(ax + (2 * t) * cs_x)), dy1 = __uint_as_float(lds32(ax + (2 * t + 1) * cs_x));
    const float d0 = half_bits_to_float((uint16_t) (h0.x & 0xffff)), m0 = half_bits_to_float((uint16_t) (h0.x >> 16));
    const float d1 = half_bits_to_float((uint16_t) (h1.x & 0xffff)), m1 = half_bits_to_float((uint16_t) (h1.x >> 16));
    acc[0] = fmaf(d0 * dy0, (float) isum[0], acc[0]); acc[0] = fmaf(-(m0 * dy0), (float) mdot(bs0, mn0), acc[0]);
    acc[1] = fmaf(d0 * dy1, (float) isum[1], acc[1]); acc[1] = fmaf(-(m0 * dy1), (float) mdot(bs1, mn0), acc[1]);
    acc[2] = fmaf(d1 * dy0, (float) isum[2], acc[2]); acc[2] = fmaf(-(m1 * dy0), (float) mdot(bs0, mn1), acc[2]);
    acc[3] = fmaf(d1 * dy1, (float) isum[3], acc[3]); acc[3] = fmaf(-(m1 * dy1), (float) mdot(bs1, mn1), acc[3]);
}

// q5_K: q4_K's unit with the fifth bits at +16 and the nibbles at +48 (176 B, row pitch 44 words == 12 mod 32: the 8 rows x 4 words of a
// fragment load still fall into 32 different banks). Sub-block j takes bit j of qh[l]: moved to bit 4 and OR-ed onto the nibble.
__device__ __forceinline__ void unit_mma_q5_K(uint32_t wt, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, int g, int t, float (&acc)[4]) {
    const uint32_t r0 = wt + g * 176, r1 = r0 + 8 * 176;
    const uint4 h0 = lds128(r0), h1 = lds128(r1);
    const uint32_t sc0[2] = { h0.y & 0x3f3f3f3fu, (h0.w & 0x0f0f0f0fu) | ((h0.y >> 2) & 0x30303030u) };
    const uint32_t mn0[2] = { h0.z & 0x3f3f3f3fu, ((h0.w >> 4) & 0x0f0f0f0fu) | ((h0.z >> 2) & 0x30303030u) };
    const uint32_t sc1[2] = { h1.y & 0x3f3f3f3fu, (h1.w & 0x0f0f0f0fu) | ((h1.y >> 2) & 0x30303030u) };
    const uint32_t mn1[2] = { h1.z & 0x3f3f3f3fu, ((h1.w >> 4) & 0x0f0f0f0fu) | ((h1.z >> 2) & 0x30303030u) };
    // fifth bits of l = 4t .. 4t + 3 (k-half 0) and l = 16 + 4t .. (k-half 1), rows g and g + 8
    const uint32_t qa0 = lds32(r0 + 16 + 4 * t), qb0 = lds32(r0 + 32 + 4 * t), qa1 = lds32(r1 + 16 + 4 * t), qb1 = lds32(r1 + 32 + 4 * t);
    const uint32_t bcol = am + g * cs_m + 4 * t;
    int isum[4] = { 0, 0, 0, 0 };
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const uint32_t wa = lds32(r0 + 48 + 32 * i + 4 * t), wb = lds32(r0 + 64 + 32 * i + 4 * t);
        const uint32_t wc = lds32(r1 + 48 + 32 * i + 4 * t), wd = lds32(r1 + 64 + 32 * i + 4 * t);
#pragma unroll
        for (int hn = 0; hn < 2; hn++) {
            const int j = 2 * i + hn;
            auto bit = [&](uint32_t q) { return (j < 4 ? q << (4 - j) : q >> (j - 4)) & 0x10101010u; };
            const uint32_t b0 = lds32(bcol + 32 * j), b1 = lds32(bcol + 32 * j + 16);
            int c[4];
            mma_u8s8(c, ((wa >> (4 * hn)) & 0x0f0f0f0fu) | bit(qa0), ((wc >> (4 * hn)) & 0x0f0f0f0fu) | bit(qa1),
                     ((wb >> (4 * hn)) & 0x0f0f0f0fu) | bit(qb0), ((wd >> (4 * hn)) & 0x0f0f0f0fu) | bit(qb1), b0, b1);
            const int s0 = (int) ((sc0[j >> 2] >> (8 * (j & 3))) & 0xff), s1 = (int) ((sc1[j >> 2] >> (8 * (j & 3))) & 0xff);
            isum[0] += s0 * c[0]; isum[1] += s0 * c[1];
            isum[2] += s1 * c[2]; isum[3] += s1 * c[3];
        }
    }
    const uint4 bs0 = lds128(am + (2 * t) * cs_m + 256), bs1 = lds128(am + (2 * t + 1) * cs_m + 256);
    auto mdot = [](const uint4 & bs, const uint32_t (&mn)[2]) {
        int v = dp2a_lo_su((int) bs.x, mn[0], 0);
        v     = dp2a_hi_su((int) bs.y, mn[0], v);
        v     = dp2a_lo_su((int) bs.z, mn[1], v);
        return dp2a_hi_su((int) bs.w, mn[1], v);
    };
    const float dy0 = __uint_as_float(lds32(ax + (2 * t) * cs_x)), dy1 = __uint_as_float(lds32(ax + (2 * t + 1) * cs_x));
    const float d0 = half_bits_to_float((uint16_t) (h0.x & 0xffff)), m0 = half_bits_to_float((uint16_t) (h0.x >> 16));
    const float d1 = half_bits_to_float((uint16_t) (h1.x & 0xffff)), m1 = half_bits_to_float((uint16_t) (h1.x >> 16));
    acc[0] = fmaf(d0 * dy0, (float) isum[0], acc[0]); acc[0] = fmaf(-(m0 * dy0), (float) mdot(bs0, mn0), acc[0]);
    acc[1] = fmaf(d0 * dy1, (float) isum[1], acc[1]); acc[1] = fmaf(-(m0 * dy1), (float) mdot(bs1, mn0), acc[1]);
    acc[2] = fmaf(d1 * dy0, (float) isum[2], acc[2]); acc[2] = fmaf(-(m1 * dy0), (float) mdot(bs0, mn1), acc[2]);
    acc[3] = fmaf(d1 * dy1, (float) isum[3], acc[3]); acc[3] = fmaf(-(m1 * dy1), (float) mdot(bs1, mn1), acc[3]);
}

// q4_0: 8 blocks of { half d; uint8 qs[16] } = 144 B per unit; element k = low nibble of qs[k], k + 16 = high nibble (one word
// holds both k-halves of a fragment). Blocks are 18 bytes: even blocks start on a word boundary (qs is shifted by 16 bits), odd
// blocks have qs word aligned. sum((q - 8) * y) = sum(q * y) - 8 * sum(y)  (ggml-cpu/quants.c:135-146)
__device__ __forceinline__ void unit_mma_q4_0(uint32_t wt, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, int g, int t, float (&acc)[4]) {
    const uint32_t r0 = wt + g * 144, r1 = r0 + 8 * 144;
    const uint32_t bcol = am + g * cs_m + 4 * t;
    // activation block scales (fp16 x 8 at +256 of the record) and block sums (int16 x 8 in the aux record) of columns 2t, 2t+1
    const uint4 dy0 = lds128(am + (2 * t) * cs_m + 256), dy1 = lds128(am + (2 * t + 1) * cs_m + 256);
    const uint4 ys0 = lds128(ax + (2 * t) * cs_x), ys1 = lds128(ax + (2 * t + 1) * cs_x);
#pragma unroll
    for (int bp = 0; bp < 4; bp++) {  // block pair = 36 bytes = 9 words
        const uint32_t p0 = r0 + 36 * bp, p1 = r1 + 36 * bp;
        // even block: d in the low half of word 0, qs word t = bytes 2 + 4t .. 5 + 4t = funnel(word t, word t + 1) >> 16
        const uint32_t e0a = lds32(p0 + 4 * t), e0b = lds32(p0 + 4 * t + 4), e1a = lds32(p1 + 4 * t), e1b = lds32(p1 + 4 * t + 4);
        const uint32_t hd0 = lds32(p0), hd1 = lds32(p1), od0 = lds32(p0 + 16), od1 = lds32(p1 + 16);  // d of the even / odd block
        const uint32_t o0 = lds32(p0 + 20 + 4 * t), o1 = lds32(p1 + 20 + 4 * t);                         // odd block: qs word aligned
#pragma unroll
        for (int ob = 0; ob < 2; ob++) {
            const int b = 2 * bp + ob;
            const uint32_t q0 = ob ? o0 : __funnelshift_r(e0a, e0b, 16), q1 = ob ? o1 : __funnelshift_r(e1a, e1b, 16);
            const float dw0 = half_bits_to_float((uint16_t) (ob ? od0 >> 16 : hd0 & 0xffff));
            const float dw1 = half_bits_to_float((uint16_t) (ob ? od1 >> 16 : hd1 & 0xffff));
            const uint32_t b0 = lds32(bcol + 32 * b), b1 = lds32(bcol + 32 * b + 16);
            int c[4];
            mma_u8s8(c, q0 & 0x0f0f0f0fu, q1 & 0x0f0f0f0fu, (q0 >> 4) & 0x0f0f0f0fu, (q1 >> 4) & 0x0f0f0f0fu, b0, b1);
            const uint32_t dwd0 = b < 2 ? dy0.x : b < 4 ? dy0.y : b < 6 ? dy0.z : dy0.w, dwd1 = b < 2 ? dy1.x : b < 4 ? dy1.y : b < 6 ? dy1.z : dy1.w;
            const uint32_t swd0 = b < 2 ? ys0.x : b < 4 ? ys0.y : b < 6 ? ys0.z : ys0.w, swd1 = b < 2 ? ys1.x : b < 4 ? ys1.y : b < 6 ? ys1.z : ys1.w;
            const float y0 = half_bits_to_float((uint16_t) ((b & 1) ? dwd0 >> 16 : dwd0 & 0xffff));
            const float y1 = half_bits_to_float((uint16_t) ((b & 1) ? dwd1 >> 16 : dwd1 & 0xffff));
            const int   s0 = (int) (int16_t) ((b & 1) ? swd0 >> 16 : swd0 & 0xffff), s1 = (int) (int16_t) ((b & 1) ? swd1 >> 16 : swd1 & 0xffff);
            acc[0] = fmaf((float) (c[0] - 8 * s0) * dw0, y0, acc[0]);
            acc[1] = fmaf((float) (c[1] - 8 * s1) * dw0, y1, acc[1]);
            acc[2] = fmaf((float) (c[2] - 8 * s0) * dw1, y0, acc[2]);
            acc[3] = fmaf((float) (c[3] - 8 * s1) * dw1, y1, acc[3]);
        }
    }
}

// q8_0: 8 blocks of { half d; int8 qs[32] } = 272 B per unit; a block pair is 68 bytes = 17 words: the even block's qs is shifted
// by 16 bits against the word grid, the odd block's is word aligned. sumi * (d_w * d_y) per block (ggml-cpu/quants.c:329)
__device__ __forceinline__ void unit_mma_q8_0(uint32_t wt, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, int g, int t, float (&acc)[4]) {
    (void) ax; (void) cs_x;
    const uint32_t r0 = wt + g * 272, r1 = r0 + 8 * 272;
    const uint32_t bcol = am + g * cs_m + 4 * t;
    const uint4 dy0 = lds128(am + (2 * t) * cs_m + 256), dy1 = lds128(am + (2 * t + 1) * cs_m + 256);
#pragma unroll
    for (int bp = 0; bp < 4; bp++) {
        const uint32_t p0 = r0 + 68 * bp, p1 = r1 + 68 * bp;
#pragma unroll
        for (int ob = 0; ob < 2; ob++) {
            const int b = 2 * bp + ob;
            uint32_t a0, a1, a2, a3;
            float dw0, dw1;
            if (ob == 0) {  // qs word k = bytes 2 + 4k .. 5 + 4k of the pair = funnel(word k, word k + 1) >> 16
                const uint32_t x0 = lds32(p0 + 4 * t), x1 = lds32(p0 + 4 * t + 4), x2 = lds32(p0 + 16 + 4 * t), x3 = lds32(p0 + 20 + 4 * t);
                const uint32_t y0 = lds32(p1 + 4 * t), y1 = lds32(p1 + 4 * t + 4), y2 = lds32(p1 + 16 + 4 * t), y3 = lds32(p1 + 20 + 4 * t);
                a0 = __funnelshift_r(x0, x1, 16); a2 = __funnelshift_r(x2, x3, 16);
                a1 = __funnelshift_r(y0, y1, 16); a3 = __funnelshift_r(y2, y3, 16);
                dw0 = half_bits_to_float((uint16_t) (lds32(p0) & 0xffff));
                dw1 = half_bits_to_float((uint16_t) (lds32(p1) & 0xffff));
            } else {        // odd block at +34: d in the high half of word 8, qs from word 9 on
                a0 = lds32(p0 + 36 + 4 * t); a2 = lds32(p0 + 52 + 4 * t);
                a1 = lds32(p1 + 36 + 4 * t); a3 = lds32(p1 + 52 + 4 * t);
                dw0 = half_bits_to_float((uint16_t) (lds32(p0 + 32) >> 16));
                dw1 = half_bits_to_float((uint16_t) (lds32(p1 + 32) >> 16));
            }
            const uint32_t b0 = lds32(bcol + 32 * b), b1 = lds32(bcol + 32 * b + 16);
            int c[4];
            mma_s8s8(c, a0, a1, a2, a3, b0, b1);
            const uint32_t dwd0 = b < 2 ? dy0.x : b < 4 ? dy0.y : b < 6 ? dy0.z : dy0.w, dwd1 = b < 2 ? dy1.x : b < 4 ? dy1.y : b < 6 ? dy1.z : dy1.w;
            const float y0f = half_bits_to_float((uint16_t) ((b & 1) ? dwd0 >> 16 : dwd0 & 0xffff));
            const float y1f = half_bits_to_float((uint16_t) ((b & 1) ? dwd1 >> 16 : dwd1 & 0xffff));
            acc[0] = fmaf((float) c[0], dw0 * y0f, acc[0]);
            acc[1] = fmaf((float) c[1], dw0 * y1f, acc[1]);
            acc[2] = fmaf((float) c[2], dw1 * y0f, acc[2]);
            acc[3] = fmaf((float) c[3], dw1 * y1f, acc[3]);
        }
    }
}

__device__ __forceinline__ void mma_s8s8_k16(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%7,%8,%9,%10};"
                 : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3])
                 : "r"(a0), "r"(a1), "r"(b0), "r"(0), "r"(0), "r"(0), "r"(0));
}

// q6_K: one block_q6_K = 210 B per unit (ql[128] | qh[64] | int8 scales[16] | half d, ggml-common.h:330-336). The stage holds
// [16 rows][8 units x 210 B] (row pitch Q6_ROW = 1680 B = 420 words: the 8 rows x 4 words of a fragment load fall into 32 different
// banks); warp w owns unit w of every row, which starts on a 2-byte phase for odd w: aligned 32-bit loads, realigned by a funnel shift
// (the phase is warp-uniform). The scales apply per 16 weights, so a 16-group is one m16n8k16 MMA (SASS IMMA.16816) on signed
// operands: ((lo + hi + 0x60606060) ^ 0x80808080) turns four 6-bit codes into int8 q - 32. isum += sc * c in int32, one fp32 FMA per
// 256 weights with d * d_y: the arithmetic of ggml_vec_dot_q6_K_q8_K_generic (ggml-cpu/quants.c:705-758).
constexpr int Q6_ROW = 8 * SZ_Q6_K;
template <bool ODD>  // ODD: the unit starts 2 bytes behind a word boundary (warp-uniform: own code path, no predicated second loads)
__device__ __forceinline__ void unit_mma_q6_K_t(uint32_t wt, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, int g, int t, float (&acc)[4]) {
    const uint32_t r0 = (wt & ~3u) + g * Q6_ROW, r1 = r0 + 8 * Q6_ROW;
    auto ldw = [&](uint32_t row, int word) {  // 32-bit word `word` of the unit
        const uint32_t a = row + 4 * word;
        return ODD ? __funnelshift_r(lds32(a), lds32(a + 4), 16) : lds32(a);
    };
    uint32_t sc0[4], sc1[4];
#pragma unroll
    for (int i = 0; i < 4; i++) { sc0[i] = ldw(r0, 48 + i); sc1[i] = ldw(r1, 48 + i); }
    const float d0 = half_bits_to_float((uint16_t) (ldw(r0, 52) & 0xffff)), d1 = half_bits_to_float((uint16_t) (ldw(r1, 52) & 0xffff));
    const uint32_t bcol = am + g * cs_m + 4 * t;  // B fragment: column g, k bytes 4t .. 4t + 3 of a 16-group
    int isum[4] = { 0, 0, 0, 0 };
#pragma unroll
    for (int h = 0; h < 2; h++) {
#pragma unroll
        for (int s = 0; s < 2; s++) {  // l = 16 s + 4 t .. + 3 of the 32 l values of a half
            const uint32_t A0 = ldw(r0, 16 * h + 4 * s + t), B0 = ldw(r0, 16 * h + 8 + 4 * s + t), H0 = ldw(r0, 32 + 8 * h + 4 * s + t);
            const uint32_t A1 = ldw(r1, 16 * h + 4 * s + t), B1 = ldw(r1, 16 * h + 8 + 4 * s + t), H1 = ldw(r1, 32 + 8 * h + 4 * s + t);
#pragma unroll
            for (int qv = 0; qv < 4; qv++) {  // elements 128 h + 32 qv + l
                const uint32_t l0 = ((qv & 1) ? B0 : A0) >> ((qv & 2) ? 4 : 0), l1 = ((qv & 1) ? B1 : A1) >> ((qv & 2) ? 4 : 0);
                const uint32_t h0 = qv == 0 ? H0 << 4 : qv == 1 ? H0 << 2 : qv == 2 ? H0 : H0 >> 2;
                const uint32_t h1 = qv == 0 ? H1 << 4 : qv == 1 ? H1 << 2 : qv == 2 ? H1 : H1 >> 2;
                // (lo | hi) + 0x60 never carries out of a byte; ^ 0x80 then gives q - 32 as int8. The two ANDs and the OR are one LOP3 pair
                const uint32_t a0 = (((l0 & 0x0F0F0F0Fu) | (h0 & 0x30303030u)) + 0x60606060u) ^ 0x80808080u;
                const uint32_t a1 = (((l1 & 0x0F0F0F0Fu) | (h1 & 0x30303030u)) + 0x60606060u) ^ 0x80808080u;
                int c[4];
                mma_s8s8_k16(c, a0, a1, lds32(bcol + 128 * h + 32 * qv + 16 * s));
                const int gi = 8 * h + 2 * qv + s;  // scale group (16 weights)
                const int s0 = (int) (int8_t) (sc0[gi >> 2] >> (8 * (gi & 3))), s1 = (int) (int8_t) (sc1[gi >> 2] >> (8 * (gi & 3)));
                isum[0] += s0 * c[0]; isum[1] += s0 * c[1];
                isum[2] += s1 * c[2]; isum[3] += s1 * c[3];
            }
        }
    }
    const float dy0 = __uint_as_float(lds32(ax + (2 * t) * cs_x)), dy1 = __uint_as_float(lds32(ax + (2 * t + 1) * cs_x));
    acc[0] = fmaf(d0 * dy0, (float) isum[0], acc[0]);
    acc[1] = fmaf(d0 * dy1, (float) isum[1], acc[1]);
    acc[2] = fmaf(d1 * dy0, (float) isum[2], acc[2]);
    acc[3] = fmaf(d1 * dy1, (float) isum[3], acc[3]);
}

__device__ __forceinline__ void unit_mma_q6_K(uint32_t wt, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, int g, int t, float (&acc)[4]) {
    if (wt & 2u) unit_mma_q6_K_t<true>(wt, am, ax, cs_m, cs_x, g, t, acc);
    else unit_mma_q6_K_t<false>(wt, am, ax, cs_m, cs_x, g, t, acc);
}

template <int TYPE>
__device__ __forceinline__ void unit_mma(uint32_t wt, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, int g, int t, float (&acc)[4]) {
    if (TYPE == B200_TYPE_Q6_K) unit_mma_q6_K(wt, am, ax, cs_m, cs_x, g, t, acc);
    else if (TYPE == B200_TYPE_Q5_K) unit_mma_q5_K(wt, am, ax, cs_m, cs_x, g, t, acc);
    else if (TYPE == B200_TYPE_Q4_K) unit_mma_q4_K(wt, am, ax, cs_m, cs_x, g, t, acc);
    else if (TYPE == B200_TYPE_Q8_0) unit_mma_q8_0(wt, am, ax, cs_m, cs_x, g, t, acc);
    else unit_mma_q4_0(wt, am, ax, cs_m, cs_x, g, t, acc);
}

__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void bar_named(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

struct MmaMaps {
    CUtensorMap m[MV_MAX_MATS];
};

// shared memory: [0,8) act barrier | full[NS] empty[NS] | act: 8 columns x U x (272 + 16) | red: 2 x NWC x MMA_RT x 128 floats | ring
template <int TYPE>
__global__ void __launch_bounds__(32 * 17, 1)
mv_mma_kernel(const __grid_constant__ MmaMaps maps, const __grid_constant__ MmaMvParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    constexpr int UB = unit_bytes(TYPE);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int NWC = p.n_cons, NS = p.n_stages, U = p.U;
    // q6_K: a chunk is 8 units of a row as one 1680-byte line (the 210-byte units are not 16-byte multiples): NWC == 8, U % 8 == 0
    // With 16 consumer warps the two groups of 8 take the chunks alternately (twice the warps in flight: the unpack chain of a unit is
    // long), and all 16 partial tiles meet in the reduction.
    constexpr bool LINE = TYPE == B200_TYPE_Q6_K;
    const int stage_bytes = LINE ? MMA_ROWS * Q6_ROW : NWC * MMA_ROWS * UB;
    const int n_chunks = LINE ? U / 8 : (U + NWC - 1) / NWC;

    uint64_t * act_bar = (uint64_t *) smem;
    uint64_t * full    = (uint64_t *) (smem + 64);
    uint64_t * empty   = full + NS;
    size_t     off     = (64 + 16 * (size_t) NS + 127) / 128 * 128;
    uint8_t *  act_m   = smem + off;
    const uint32_t cs_m = (uint32_t) U * ACT_MAIN + 16;  // +4 words: the 8 columns of a B-fragment load fall into 8 different bank groups
    const uint32_t cs_x = (uint32_t) U * ACT_AUX;
    uint8_t *  act_x   = act_m + (size_t) 8 * cs_m;
    off                = (off + (size_t) 8 * (cs_m + cs_x) + 127) / 128 * 128;
    float *    red     = (float *) (smem + off);
    off                = (off + (size_t) 2 * NWC * MMA_RT * 128 * 4 + 1023) / 1024 * 1024;
    uint8_t *  ring    = smem + off;

    const int t_begin = (int) ((int64_t) p.total_tiles * blockIdx.x / gridDim.x);
    const int t_end   = (int) ((int64_t) p.total_tiles * (blockIdx.x + 1) / gridDim.x);

    // MUL_MAT_ID: this CTA row works for expert e on the pairs routed to it (the sort ran earlier on the stream)
    const bool by_expert = p.sp != nullptr;
    int e = 0, pair0 = 0, n_pairs = p.ncols;
    if (by_expert) {
        pdl_wait();
        e       = blockIdx.y;
        pair0   = p.eb[e];
        n_pairs = p.eb[e + 1] - pair0;
        if (n_pairs <= 0 || t_begin >= t_end) {
            if (p.split.n_ranks > 1 && warp == 0) split_complete(p.split, gridDim.x * gridDim.y);
            return;
        }
    }
    const int n_passes = (n_pairs + 7) / 8;  // the weight stream is read once per 8 columns

    if (threadIdx.x == 0) {
        mbar_init(smem_u32(act_bar), 1);
        for (int i = 0; i < NS; i++) {
            mbar_init(smem_u32(full + i), 1);
            mbar_init(smem_u32(empty + i), TYPE == B200_TYPE_Q6_K ? 8 : NWC);  // q6_K: a stage (8 units) belongs to one group of 8 warps
        }
        fence_mbar_init();
    }
    __syncthreads();
    pdl_launch_dependents();

    if (warp == NWC) {
        // ================================================================ producer: one TMA box per (pass, tile, chunk)
        if (lane == 0) {
            int stage = 0; uint32_t phase = 1;
            for (int pass = 0; pass < n_passes; pass++) {
                for (int tile = t_begin; tile < t_end; tile++) {
                    int m = 0;
#pragma unroll
                    for (int i = 1; i < MV_MAX_MATS; i++) if (i < p.n_mats && tile >= p.mat[i].group0) m = i;
                    const int row0 = (tile - p.mat[m].group0) * MMA_ROWS;
                    for (int ch = 0; ch < n_chunks; ch++) {
                        mbar_wait(smem_u32(empty + stage), phase);
                        const uint32_t bar = smem_u32(full + stage);
                        mbar_expect_tx(bar, (uint32_t) stage_bytes);  // out-of-range rows / units are zero filled and still counted
                        const uint32_t dst = smem_u32(ring + (size_t) stage * stage_bytes);
                        const int c2 = LINE ? ch : ch * NWC;  // q6_K maps count dimension 2 in 8-unit lines
                        if (by_expert) tma_load_4d(dst, &maps.m[0], 0, row0, c2, e, bar);
                        else tma_load_3d(dst, &maps.m[m], 0, row0, c2, bar);
                        if (++stage == NS) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
        return;
    }
    if (warp > NWC) return;

    // ==================================================================== consumers
    const int w = warp, g = lane >> 2, t = lane & 3;
    if (!by_expert) pdl_wait();  // the activation records are written by the quantizer launch before this one
    const bool     ll = p.split.n_ranks > 1 && p.split.ll && !by_expert;
    const uint32_t ep = ll ? ll_epoch(p.split) : 0u;
    const uint32_t am0 = smem_u32(act_m), ax0 = smem_u32(act_x);
    int stage = 0; uint32_t phase = 0;
    int parity = 0;
    const int n_groups = LINE ? NWC / 8 : 1, my_group = LINE ? w >> 3 : 0, wu = LINE ? (w & 7) : w;
    int seq = 0;  // q6_K: running chunk number, chunk seq belongs to group seq % n_groups
    for (int pass = 0; pass < n_passes; pass++) {
        const int nc = min(8, n_pairs - 8 * pass);
        // activations: nc columns of U records (one bulk copy per column: the shared column stride is padded); columns nc..7 are
        // zero (the MMA always multiplies 8). The previous pass's readers are behind its last reduction barrier.
        for (int i = threadIdx.x; i < (8 - nc) * (int) (cs_m / 16); i += NWC * 32) ((uint4 *) (act_m + (size_t) nc * cs_m))[i] = make_uint4(0, 0, 0, 0);
        for (int i = threadIdx.x; i < (8 - nc) * (int) (cs_x / 16); i += NWC * 32) ((uint4 *) (act_x + (size_t) nc * cs_x))[i] = make_uint4(0, 0, 0, 0);
        if (p.fuse_quant) {
            // fused K1 (as in mv_fast_kernel): unit (c, u) of the activations, F32 global -> unit record in shared memory; up to 4
            // units per warp are loaded before the first is quantized, so the L2 round trips overlap. No quantizer launch.
            constexpr int KIND = act_kind_for(TYPE);
            const int n_tasks = nc * U;
            for (int t0 = w; t0 < n_tasks; t0 += 4 * NWC) {
                float v[4][8];
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    const int tk = t0 + b * NWC;
                    if (tk < n_tasks) {
                        const int c = tk / U, u = tk - c * U;
                        const float * xr = (const float *) ((const char *) p.x + (int64_t) c * p.x_cs) + (int64_t) u * UNIT_ELEMS + lane * 8;
                        if (p.x_vec) {
                            const float4 a = __ldg((const float4 *) xr), bb = __ldg((const float4 *) xr + 1);
                            v[b][0] = a.x; v[b][1] = a.y; v[b][2] = a.z; v[b][3] = a.w;
                            v[b][4] = bb.x; v[b][5] = bb.y; v[b][6] = bb.z; v[b][7] = bb.w;
                        } else {
#pragma unroll
                            for (int i = 0; i < 8; i++) v[b][i] = __ldg(xr + i);
                        }
                    }
                }
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    const int tk = t0 + b * NWC;
                    if (tk < n_tasks) {
                        const int c = tk / U, u = tk - c * U;
                        quant_unit_store<KIND>(v[b], lane, act_m + (size_t) c * cs_m + (size_t) u * ACT_MAIN, act_x + (size_t) c * cs_x + (size_t) u * ACT_AUX);
                    }
                }
            }
        } else if (w == 0 && lane == 0) {
            const uint64_t pol_keep = l2_policy_evict_last();
            const uint32_t bytes_m = (uint32_t) U * ACT_MAIN, bytes_x = (uint32_t) U * ACT_AUX;
            mbar_expect_tx(smem_u32(act_bar), (uint32_t) nc * (bytes_m + bytes_x));
            for (int c = 0; c < nc; c++) {
                int64_t arow = c;
                if (by_expert) {
                    const int pr = p.sp[pair0 + 8 * pass + c];
                    const int tk = pr / p.ids.n_used, sl = pr - tk * p.ids.n_used;
                    arow = (int64_t) tk * p.ids.ne11 + sl % p.ids.ne11;
                }
                bulk_g2s(smem_u32(act_m + (size_t) c * cs_m), p.act_main + (size_t) arow * bytes_m, bytes_m, smem_u32(act_bar), pol_keep);
                bulk_g2s(smem_u32(act_x + (size_t) c * cs_x), p.act_aux + (size_t) arow * bytes_x, bytes_x, smem_u32(act_bar), pol_keep);
            }
        }
        if (!p.fuse_quant) mbar_wait(smem_u32(act_bar), (uint32_t) (pass & 1));
        bar_named(1, NWC * 32);  // the zero fill of the unused columns (and the fused quantization) is visible

        for (int tile0 = t_begin; tile0 < t_end; tile0 += MMA_RT) {
            const int nt = min(MMA_RT, t_end - tile0);
            float acc[MMA_RT][4];
#pragma unroll
            for (int r = 0; r < MMA_RT; r++) { acc[r][0] = acc[r][1] = acc[r][2] = acc[r][3] = 0.0f; }
#pragma unroll
            for (int r = 0; r < MMA_RT; r++) {
                if (r < nt) {
                    for (int ch = 0; ch < n_chunks; ch++) {
                        const bool mine = !LINE || (seq++ % n_groups) == my_group;
                        if (mine) {
                            mbar_wait(smem_u32(full + stage), phase);
                            const int u = LINE ? ch * 8 + wu : ch * NWC + w;
                            if (u < U) {
                                const uint32_t wt = smem_u32(ring + (size_t) stage * stage_bytes) + (uint32_t) wu * (LINE ? UB : MMA_ROWS * UB);
                                unit_mma<TYPE>(wt, am0 + (uint32_t) u * ACT_MAIN, ax0 + (uint32_t) u * ACT_AUX, cs_m, cs_x, g, t, acc[r]);
                            }
                            __syncwarp();
                            if (lane == 0) mbar_arrive(smem_u32(empty + stage));
                        }
                        if (++stage == NS) { stage = 0; phase ^= 1; }
                    }
                }
            }
            // fixed-order sum of the NWC partial results: red[parity][w][tile r][(row, col)]
            float * rp = red + (size_t) parity * NWC * MMA_RT * 128;
#pragma unroll
            for (int r = 0; r < MMA_RT; r++) {
                float * q = rp + (w * MMA_RT + r) * 128;
                q[g * 8 + 2 * t]           = acc[r][0];
                q[g * 8 + 2 * t + 1]       = acc[r][1];
                q[(g + 8) * 8 + 2 * t]     = acc[r][2];
                q[(g + 8) * 8 + 2 * t + 1] = acc[r][3];
            }
            bar_named(1, NWC * 32);
            for (int o = threadIdx.x; o < nt * 128; o += NWC * 32) {
                const int r = o >> 7, el = o & 127;
                float v = 0.0f;
                for (int i = 0; i < NWC; i++) v += rp[(i * MMA_RT + r) * 128 + el];
                const int tile = tile0 + r;
                int m = 0;
#pragma unroll
                for (int i = 1; i < MV_MAX_MATS; i++) if (i < p.n_mats && tile >= p.mat[i].group0) m = i;
                const int row = (tile - p.mat[m].group0) * MMA_ROWS + (el >> 3), c = el & 7;
                if (row < p.mat[m].M && c < nc) {
                    int64_t col_off = (int64_t) c * p.mat[m].dst_cs;
                    if (by_expert) {
                        const int pr = p.sp[pair0 + 8 * pass + c];
                        const int tk = pr / p.ids.n_used, sl = pr - tk * p.ids.n_used;
                        col_off = (int64_t) tk * p.ids.dst_ts + (int64_t) sl * p.ids.dst_is;
                    }
                    if (ll) ll_store(p.split, ep, p.mat[m].dst + col_off + row, v);
                    else split_store(p.split, p.mat[m].dst + col_off + row, v);
                }
            }
            parity ^= 1;
        }
        bar_named(1, NWC * 32);  // every warp is done with this pass's activation columns
    }
    if (ll) {
        ll_receive(p.split, ep, p.geom, p.mat, p.n_mats, p.ncols, (int) blockIdx.x * NWC * 32 + w * 32 + lane, (int) gridDim.x * NWC * 32);
        bar_named(1, NWC * 32);
        if (w == 0 && lane == 0) ll_finish(p.split, ep, gridDim.x);
    } else if (p.split.n_ranks > 1) {
        if (w == 0) split_complete(p.split, gridDim.x * gridDim.y);
    }
}

typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
encode_tiled_fn mma_get_encode() {
    static encode_tiled_fn fn = [] {
        void * ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
            return (encode_tiled_fn) ptr;
        return (encode_tiled_fn) nullptr;
    }();
    return fn;
}

} // namespace

size_t mv_mma_smem(int type, int U, int n_cons, int n_stages) {
    size_t off = (64 + 16 * (size_t) n_stages + 127) / 128 * 128;
    off = (off + (size_t) 8 * ((size_t) U * (ACT_MAIN + ACT_AUX) + 16) + 127) / 128 * 128;
    off = (off + (size_t) 2 * n_cons * MMA_RT * 128 * 4 + 1023) / 1024 * 1024;
    return off + (size_t) n_stages * (type == B200_TYPE_Q6_K ? 8 : n_cons) * MMA_ROWS * unit_bytes(type) + 1024;
}

// the 8 activation columns of a K-long row plus a minimal ring must fit in shared memory (K = 28672 does not: 258 KB of activations)
bool mv_mma_fits(int type, int64_t K) {
    if (type == B200_TYPE_Q6_K) return K % 2048 == 0 && mv_mma_smem(type, (int) (K / 256), 8, 2) <= 227 * 1024;
    return mv_mma_smem(type, (int) (K / 256), 4, 2) <= 227 * 1024;
}

bool mv_mma_usable(int type) {
    return (type == B200_TYPE_Q4_K || type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0 || type == B200_TYPE_Q5_K || type == B200_TYPE_Q6_K) && mma_get_encode() != nullptr;
}

// consumer warps / ring stages: as many warps as leave room for >= 3 stages next to the 8 activation columns (K = 14336: 129 KB of
// them); q6_K takes 16 or 8 warps (groups of 8: one warp per unit of a 1680-byte line)
static bool mma_pick_cfg(const b200_ctx * ctx, int type, int U, int * n_cons, int * n_stages) {
    static const int cand[] = { 16, 12, 8, 6, 4 };
    const int forced = type == B200_TYPE_Q6_K ? (ctx->mma_warps == 8 ? 8 : 0) : ctx->mma_warps;
    for (int nc : cand) {
        if (forced > 0 && nc != forced) continue;
        if (type == B200_TYPE_Q6_K && nc != 16 && nc != 8) continue;  // groups of 8 warps: one per unit of a 1680-byte line
        for (int ns = 6; ns >= 2; ns--) {
            if (ns < 3 && nc > 4 && type != B200_TYPE_Q6_K) continue;
            if (mv_mma_smem(type, U, nc, ns) <= 227 * 1024) { *n_cons = nc; *n_stages = ns; return true; }
        }
    }
    if (forced > 0 && mv_mma_smem(type, U, forced, 2) <= 227 * 1024) { *n_cons = forced; *n_stages = 2; return true; }
    return false;
}

// the weight stream as a tensor map: [unit][row][unit bytes] (a box lands as NWC dense [16][unit bytes] sub-tiles) or, for q6_K,
// [8-unit line][row][1680 B] as 210 uint64 per line (inner box dimensions are limited to 256 elements); dimension 3 = experts
static CUresult mma_encode_map(CUtensorMap * map, int type, const void * w, int64_t M, int U, size_t nb01, int n_cons, int64_t n_as, size_t nb02) {
    const int  UB   = unit_bytes(type);
    const bool line = type == B200_TYPE_Q6_K;
    cuuint64_t dims[4]    = { (cuuint64_t) (line ? Q6_ROW / 8 : UB / 4), (cuuint64_t) M, (cuuint64_t) (line ? U / 8 : U), (cuuint64_t) (n_as > 0 ? n_as : 1) };
    cuuint64_t strides[3] = { (cuuint64_t) nb01, (cuuint64_t) (line ? Q6_ROW : UB), (cuuint64_t) nb02 };
    cuuint32_t box[4]     = { (cuuint32_t) dims[0], MMA_ROWS, (cuuint32_t) (line ? 1 : n_cons), 1 };
    cuuint32_t estr[4]    = { 1, 1, 1, 1 };
    return mma_get_encode()(map, line ? CU_TENSOR_MAP_DATA_TYPE_UINT64 : CU_TENSOR_MAP_DATA_TYPE_UINT32, n_as > 0 ? 4 : 3, (void *) w, dims, strides, box,
                            estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
}

} // namespace b200

template <int TYPE>
static b200_status mma_launch_t(b200_ctx * ctx, const cudaLaunchConfig_t & cfg, const MmaMaps & maps, const MmaMvParams & p) {
    static bool done[64] = {};
    if (ctx->device < 64 && !done[ctx->device]) {
        CU(cudaFuncSetAttribute(mv_mma_kernel<TYPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        done[ctx->device] = true;
    }
    CU(cudaLaunchKernelEx(&cfg, mv_mma_kernel<TYPE>, maps, p));
    ctx->launches++;
    return B200_OK;
}

static b200_status mma_launch(b200_ctx * ctx, int type, dim3 grid, size_t smem, const MmaMaps & maps, const MmaMvParams & p) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = grid;
    cfg.blockDim         = dim3(32 * (p.n_cons + 1));
    cfg.dynamicSmemBytes = smem;
    cfg.stream           = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs    = attr;
    cfg.numAttrs = p.pdl ? 1 : 0;
    switch (type) {
        case B200_TYPE_Q4_K: return mma_launch_t<B200_TYPE_Q4_K>(ctx, cfg, maps, p);
        case B200_TYPE_Q8_0: return mma_launch_t<B200_TYPE_Q8_0>(ctx, cfg, maps, p);
        case B200_TYPE_Q5_K: return mma_launch_t<B200_TYPE_Q5_K>(ctx, cfg, maps, p);
        case B200_TYPE_Q6_K: return mma_launch_t<B200_TYPE_Q6_K>(ctx, cfg, maps, p);
        default:             return mma_launch_t<B200_TYPE_Q4_0>(ctx, cfg, maps, p);
    }
}

// host: n_ops matrices (same type, K) against `ncols` pre-quantized activation columns
b200_status b200_mv_mma_run(b200_ctx * ctx, int type, int64_t K, int n_ops, const uint8_t * const * ws, const size_t * nb01s, const int64_t * Ms,
                            float * const * dsts, const int64_t * dst_cs, int ncols, const uint8_t * act_main, const uint8_t * act_aux,
                            const float * x, size_t x_cs) {
    const int U  = (int) (K / 256);
    int n_cons = 0, n_stages = 0;
    if (!mma_pick_cfg(ctx, type, U, &n_cons, &n_stages)) return b200_fail(B200_ERR_UNSUPPORTED, "K=%lld: activations do not fit in shared memory", (long long) K);
    const size_t smem = mv_mma_smem(type, U, n_cons, n_stages);
    for (int o0 = 0; o0 < n_ops; o0 += MV_MAX_MATS) {
        MmaMvParams p;
        MmaMaps     maps;
        memset(&p, 0, sizeof(p));
        memset(&maps, 0, sizeof(maps));
        p.n_mats = n_ops - o0 < MV_MAX_MATS ? n_ops - o0 : MV_MAX_MATS;
        int tiles = 0;
        for (int i = 0; i < p.n_mats; i++) {
            p.mat[i].w      = ws[o0 + i];
            p.mat[i].nb01   = (int64_t) nb01s[o0 + i];
            p.mat[i].dst    = dsts[o0 + i];
            p.mat[i].dst_cs = dst_cs[o0 + i];
            p.mat[i].M      = (int) Ms[o0 + i];
            p.mat[i].group0 = tiles;
            tiles += (int) ((Ms[o0 + i] + MMA_ROWS - 1) / MMA_ROWS);
            // (a 2-D box with long rows is no faster for the 144 / 272-byte units and needs an odd unit count to stay bank-conflict free:
            // measured 2441 vs 2612 GB/s)
            const CUresult cr = mma_encode_map(&maps.m[i], type, ws[o0 + i], Ms[o0 + i], U, nb01s[o0 + i], n_cons, 0, 0);
            if (cr != CUDA_SUCCESS) {
                return b200_fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled failed for the weight stream (M=%lld K=%lld nb01=%zu)", (long long) Ms[o0 + i],
                                 (long long) K, nb01s[o0 + i]);
            }
        }
        for (int i = p.n_mats; i < MV_MAX_MATS; i++) p.mat[i].group0 = INT32_MAX;
        if (tiles == 0) continue;
        p.total_tiles = tiles;
        p.U = U; p.ncols = ncols; p.n_cons = n_cons; p.n_stages = n_stages; p.pdl = ctx->pdl;
        p.act_main = act_main; p.act_aux = act_aux;
        if (x) {
            p.fuse_quant = 1;
            p.x = x; p.x_cs = (int64_t) x_cs;
            p.x_vec = (uintptr_t) x % 16 == 0 && x_cs % 16 == 0;
        }
        p.split = ctx->split;
        if (b200_split_ll_geom(ctx, p.n_mats, Ms + o0, dsts + o0, dst_cs + o0, ncols, &p.geom)) p.split.ll = 1;
        const int grid = tiles < ctx->sm_count ? tiles : ctx->sm_count;
        if (b200_status lst = mma_launch(ctx, type, dim3(grid), smem, maps, p)) return lst;
    }
    return B200_OK;
}

// host: MUL_MAT_ID grouped by expert. as = [n_as] expert matrices (nb02 apart), activations pre-quantized ([ne11 * n_tok] rows of
// unit records), sp / eb = the device arrays b200_mm_ids_sort produced for these ids.
b200_status b200_mv_mma_run_ids(b200_ctx * ctx, int type, int64_t K, const uint8_t * as, size_t nb01, size_t nb02, int64_t M, int64_t n_as,
                                float * dst, const MvIds & ids, const int32_t * sp, const int32_t * eb, const uint8_t * act_main,
                                const uint8_t * act_aux) {
    const int U  = (int) (K / 256);
    int n_cons = 0, n_stages = 0;
    if (!mma_pick_cfg(ctx, type, U, &n_cons, &n_stages)) return b200_fail(B200_ERR_UNSUPPORTED, "K=%lld: activations do not fit in shared memory", (long long) K);
    const size_t smem = mv_mma_smem(type, U, n_cons, n_stages);
    MmaMvParams p;
    MmaMaps     maps;
    memset(&p, 0, sizeof(p));
    memset(&maps, 0, sizeof(maps));
    if (mma_encode_map(&maps.m[0], type, as, M, U, nb01, n_cons, n_as, nb02) != CUDA_SUCCESS) {
        return b200_fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled failed for the expert weight stream (M=%lld K=%lld n_as=%lld)", (long long) M,
                         (long long) K, (long long) n_as);
    }
    p.n_mats        = 1;
    p.mat[0].w      = as;
    p.mat[0].nb01   = (int64_t) nb01;
    p.mat[0].dst    = dst;
    p.mat[0].dst_cs = 0;
    p.mat[0].M      = (int) M;
    p.mat[0].group0 = 0;
    for (int i = 1; i < MV_MAX_MATS; i++) p.mat[i].group0 = INT32_MAX;
    p.total_tiles = (int) ((M + MMA_ROWS - 1) / MMA_ROWS);
    p.U = U; p.ncols = 0; p.n_cons = n_cons; p.n_stages = n_stages; p.pdl = ctx->pdl;
    p.act_main = act_main; p.act_aux = act_aux;
    p.split = ctx->split;
    p.sp = sp; p.eb = eb; p.ids = ids;
    int per_expert = (int) (ctx->sm_count / n_as);
    if (per_expert < 1) per_expert = 1;
    if (per_expert > p.total_tiles) per_expert = p.total_tiles;
    return mma_launch(ctx, type, dim3(per_expert, (unsigned) n_as), smem, maps, p);
}

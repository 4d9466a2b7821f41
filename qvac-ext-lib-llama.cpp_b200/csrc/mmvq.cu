// mmvq.cu — K3: decode mat-vec kernels (1..8 activation columns) for q4_0 / q8_0 / q4_K weights on sm_100a.
//
// Replaces mul_mat_vec_q + vec_dot_{q4_0,q8_0,q4_K}_q8_1 (/root/reference/ggml/src/ggml-cuda/mmvq.cu:141-335,
// vecdotq.cuh:103-122, 231-243, 461-483). Not a port: the reference walks a row with 2-/4-byte scalar loads
// (vecdotq.cuh:7-30) because 18-/34-byte blocks are only 2-byte aligned; here
//   * the weight stream is moved by the TMA engine: cp.async.bulk (SASS UBLKCP) copies whole row slices
//     (>= 2.3 KB each) global -> shared into a per-warp ring of stages, completion through mbarriers, with an
//     L2 evict_first policy (weights are read once per token);
//   * each warp is its own producer/consumer pipeline (lane 0 issues the copies), so there is no CTA-wide
//     barrier in the steady state; a persistent grid of one CTA per SM takes contiguous row ranges and its warps
//     pull row groups from a shared counter;
//   * out of shared memory a lane owns one 256-weight "unit" (8 x q4_0 = 1 x q4_K = 144 B, 8 x q8_0 = 272 B):
//     every unit is a whole number of 128-bit words, lane strides of 144 / 272 bytes are bank-conflict free for
//     ld.shared.v4, and all sub-block offsets inside a unit are compile-time constants (the 2-byte misalignment
//     of q4_0/q8_0 blocks becomes a constant 16-bit funnel shift);
//   * integer dot products with dp4a (dp4a.u32.s32 on the un-shifted high nibbles, exact >>4 afterwards) and
//     dp2a for the q4_K min term, fp32 accumulation per lane, warp-shuffle reduction per row.
// The integer block sums are identical to ggml-cpu's vec_dot_q4_0_q8_0 / q8_0_q8_0 / q4_K_q8_K because the
// activations are quantized exactly like the CPU does (quantize.cu); only the fp32 summation order differs.

#include "quant_core.cuh"
#include "kernels.h"

#include <atomic>
#include "epilogue.cuh"

#include <type_traits>

namespace b200 {

// ------------------------------------------------------------------------------------------------
// per-unit dot products. `w` = shared address of the unit's weights, `am`/`ax` = shared addresses of the
// unit's activation record for column 0 (main / aux), `cs_m`/`cs_x` = byte strides between columns.
// ------------------------------------------------------------------------------------------------

// compile-time loop helper so that every shared-memory offset below is an immediate
template <int I, int N, typename F>
__device__ __forceinline__ void static_for(F && f) {
    if constexpr (I < N) {
        f(std::integral_constant<int, I>{});
        static_for<I + 1, N>(f);
    }
}

template <int NC>
__device__ __forceinline__ void unit_dot_q4_K(uint32_t w, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, float (&acc)[NC]) {
    const uint4 hdr = lds128o<0>(w);  // d | dmin, 12 bytes of 6-bit scales/mins (ggml-common.h:291-305)
    const float d    = half_bits_to_float((uint16_t) (hdr.x & 0xffff));
    const float dmin = half_bits_to_float((uint16_t) (hdr.x >> 16));
    // 8 scales and 8 mins as bytes (same bit surgery as ggml-cpu/quants.c:586-591)
    const uint32_t sc_lo = hdr.y & 0x3f3f3f3fu;
    const uint32_t mn_lo = hdr.z & 0x3f3f3f3fu;
    const uint32_t sc_hi = (hdr.w & 0x0f0f0f0fu) | ((hdr.y >> 2) & 0x30303030u);
    const uint32_t mn_hi = ((hdr.w >> 4) & 0x0f0f0f0fu) | ((hdr.z >> 2) & 0x30303030u);

    int      isum[NC];
    uint32_t ac[NC];
#pragma unroll
    for (int c = 0; c < NC; c++) {
        isum[c] = 0;
        ac[c]   = am + c * cs_m;
    }

    static_for<0, 4>([&](auto gi) {  // 64 weights: 32 bytes, low nibbles = sub-block 2g, high nibbles = sub-block 2g+1
        constexpr int g = decltype(gi)::value;
        const uint4 qa = lds128o<16 + 32 * g>(w);
        const uint4 qb = lds128o<32 + 32 * g>(w);
        const uint32_t q[8] = { qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w };
        const uint32_t scw  = g < 2 ? sc_lo : sc_hi;
        const int      s0   = (scw >> (16 * (g & 1))) & 0xff;
        const int      s1   = (scw >> (16 * (g & 1) + 8)) & 0xff;
#pragma unroll
        for (int c = 0; c < NC; c++) {
            const uint4 l0 = lds128o<64 * g>(ac[c]), l1 = lds128o<64 * g + 16>(ac[c]), h0 = lds128o<64 * g + 32>(ac[c]),
                        h1 = lds128o<64 * g + 48>(ac[c]);
            const int lo[8] = { (int) l0.x, (int) l0.y, (int) l0.z, (int) l0.w, (int) l1.x, (int) l1.y, (int) l1.z, (int) l1.w };
            const int hi[8] = { (int) h0.x, (int) h0.y, (int) h0.z, (int) h0.w, (int) h1.x, (int) h1.y, (int) h1.z, (int) h1.w };
            int sl = 0, sh = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                sl = dp4a_us(q[k] & 0x0f0f0f0fu, lo[k], sl);
                sh = dp4a_us(q[k] & 0xf0f0f0f0u, hi[k], sh);  // 16 x the high-nibble dot product
            }
            isum[c] += s0 * sl + s1 * (sh >> 4);
        }
    });
#pragma unroll
    for (int c = 0; c < NC; c++) {
        const uint4 bs = lds128o<256>(ac[c]);  // int16 sum of q per 32-block (x8)
        int imin = dp2a_lo_su((int) bs.x, mn_lo, 0);
        imin     = dp2a_hi_su((int) bs.y, mn_lo, imin);
        imin     = dp2a_lo_su((int) bs.z, mn_hi, imin);
        imin     = dp2a_hi_su((int) bs.w, mn_hi, imin);
        const float dy = __uint_as_float(lds32(ax + c * cs_x));
        acc[c] = fmaf(d * dy, (float) isum[c], acc[c]);
        acc[c] = fmaf(-(dmin * dy), (float) imin, acc[c]);
    }
}

template <int NC>
__device__ __forceinline__ void unit_dot_q4_0(uint32_t w, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, float (&acc)[NC]) {
    uint32_t ww[36];  // 8 blocks of { half d; uint8 qs[16] } = 144 bytes
    static_for<0, 9>([&](auto ii) {
        constexpr int i = decltype(ii)::value;
        const uint4 v = lds128o<16 * i>(w);
        ww[4 * i] = v.x; ww[4 * i + 1] = v.y; ww[4 * i + 2] = v.z; ww[4 * i + 3] = v.w;
    });
    uint4    dyh[NC], ys[NC];
    uint32_t ac[NC];
#pragma unroll
    for (int c = 0; c < NC; c++) {
        ac[c]  = am + c * cs_m;
        dyh[c] = lds128o<256>(ac[c]);      // half d[8]
        ys[c]  = lds128(ax + c * cs_x);    // int16 sum[8]
    }
    static_for<0, 8>([&](auto bi) {
        constexpr int b = decltype(bi)::value;
        constexpr int p = 9 * (b >> 1);  // word offset of the block pair
        float    dw;
        uint32_t q[4];
        if ((b & 1) == 0) {  // block starts on a word boundary: d in the low half, nibbles shifted by 16 bits
            dw = half_bits_to_float((uint16_t) (ww[p] & 0xffff));
#pragma unroll
            for (int k = 0; k < 4; k++) q[k] = __funnelshift_r(ww[p + k], ww[p + k + 1], 16);
        } else {  // block starts in the middle of word p+4: d in its high half, nibbles word aligned
            dw = half_bits_to_float((uint16_t) (ww[p + 4] >> 16));
#pragma unroll
            for (int k = 0; k < 4; k++) q[k] = ww[p + 5 + k];
        }
#pragma unroll
        for (int c = 0; c < NC; c++) {
            const uint4 l = lds128o<32 * b>(ac[c]), h = lds128o<32 * b + 16>(ac[c]);  // elements 0..15 (low nibbles), 16..31 (high nibbles)
            int sl = 0, sh = 0;
            sl = dp4a_us(q[0] & 0x0f0f0f0fu, (int) l.x, sl); sh = dp4a_us(q[0] & 0xf0f0f0f0u, (int) h.x, sh);
            sl = dp4a_us(q[1] & 0x0f0f0f0fu, (int) l.y, sl); sh = dp4a_us(q[1] & 0xf0f0f0f0u, (int) h.y, sh);
            sl = dp4a_us(q[2] & 0x0f0f0f0fu, (int) l.z, sl); sh = dp4a_us(q[2] & 0xf0f0f0f0u, (int) h.z, sh);
            sl = dp4a_us(q[3] & 0x0f0f0f0fu, (int) l.w, sl); sh = dp4a_us(q[3] & 0xf0f0f0f0u, (int) h.w, sh);
            const uint32_t dword = b < 2 ? dyh[c].x : b < 4 ? dyh[c].y : b < 6 ? dyh[c].z : dyh[c].w;
            const uint32_t sword = b < 2 ? ys[c].x : b < 4 ? ys[c].y : b < 6 ? ys[c].z : ys[c].w;
            const float    dy    = half_bits_to_float((uint16_t) ((b & 1) ? dword >> 16 : dword & 0xffff));
            const int      ysum  = (int) (int16_t) ((b & 1) ? sword >> 16 : sword & 0xffff);
            const int      sumi  = sl + (sh >> 4) - 8 * ysum;  // sum((q-8)*y), ggml-cpu/quants.c:135-146
            acc[c] = fmaf((float) sumi * dw, dy, acc[c]);
        }
    });
}

template <int NC>
__device__ __forceinline__ void unit_dot_q8_0(uint32_t w, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, float (&acc)[NC]) {
    (void) ax; (void) cs_x;
    uint32_t ww[68];  // 8 blocks of { half d; int8 qs[32] } = 272 bytes; 128-bit loads at a 272-byte lane stride are conflict free
    static_for<0, 17>([&](auto ii) {
        constexpr int i = decltype(ii)::value;
        const uint4 v = lds128o<16 * i>(w);
        ww[4 * i] = v.x; ww[4 * i + 1] = v.y; ww[4 * i + 2] = v.z; ww[4 * i + 3] = v.w;
    });
    uint4    dyh[NC];
    uint32_t ac[NC];
#pragma unroll
    for (int c = 0; c < NC; c++) {
        ac[c]  = am + c * cs_m;
        dyh[c] = lds128o<256>(ac[c]);
    }
    static_for<0, 8>([&](auto bi) {
        constexpr int b = decltype(bi)::value;
        constexpr int p = 17 * (b >> 1);  // a block pair is 68 bytes = 17 words starting on a word boundary
        float    dw;
        uint32_t q[8];
        if ((b & 1) == 0) {
            dw = half_bits_to_float((uint16_t) (ww[p] & 0xffff));
#pragma unroll
            for (int k = 0; k < 8; k++) q[k] = __funnelshift_r(ww[p + k], ww[p + k + 1], 16);
        } else {
            dw = half_bits_to_float((uint16_t) (ww[p + 8] >> 16));
#pragma unroll
            for (int k = 0; k < 8; k++) q[k] = ww[p + 9 + k];
        }
#pragma unroll
        for (int c = 0; c < NC; c++) {
            const uint4 y0 = lds128o<32 * b>(ac[c]), y1 = lds128o<32 * b + 16>(ac[c]);
            int s = 0;
            s = dp4a_ss((int) q[0], (int) y0.x, s); s = dp4a_ss((int) q[1], (int) y0.y, s);
            s = dp4a_ss((int) q[2], (int) y0.z, s); s = dp4a_ss((int) q[3], (int) y0.w, s);
            s = dp4a_ss((int) q[4], (int) y1.x, s); s = dp4a_ss((int) q[5], (int) y1.y, s);
            s = dp4a_ss((int) q[6], (int) y1.z, s); s = dp4a_ss((int) q[7], (int) y1.w, s);
            const uint32_t dword = b < 2 ? dyh[c].x : b < 4 ? dyh[c].y : b < 6 ? dyh[c].z : dyh[c].w;
            const float    dy    = half_bits_to_float((uint16_t) ((b & 1) ? dword >> 16 : dword & 0xffff));
            acc[c] = fmaf((float) s, dw * dy, acc[c]);  // sumi*(dx*dy), ggml-cpu/quants.c:329
        }
    });
}

// q5_K: q4_K's unit with 32 bytes of fifth bits between the header and the nibbles (176 B = 11 x 16 B, lane stride 44 words == 12 mod
// 32: conflict free for 128-bit loads). Sub-block 2g takes the low nibbles of qs[32 g + l] and bit 2g of qh[l], sub-block 2g + 1 the
// high nibbles and bit 2g + 1 (ggml-cpu/quants.c:655-669): the bit is moved to bit 4 of the byte and OR-ed in, u8 x s8 dp4a as for q4_K.
template <int NC>
__device__ __forceinline__ void unit_dot_q5_K(uint32_t w, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, float (&acc)[NC]) {
    const uint4 hdr = lds128o<0>(w);
    const float d    = half_bits_to_float((uint16_t) (hdr.x & 0xffff));
    const float dmin = half_bits_to_float((uint16_t) (hdr.x >> 16));
    const uint32_t sc_lo = hdr.y & 0x3f3f3f3fu;
    const uint32_t mn_lo = hdr.z & 0x3f3f3f3fu;
    const uint32_t sc_hi = (hdr.w & 0x0f0f0f0fu) | ((hdr.y >> 2) & 0x30303030u);
    const uint32_t mn_hi = ((hdr.w >> 4) & 0x0f0f0f0fu) | ((hdr.z >> 2) & 0x30303030u);
    const uint4 ha = lds128o<16>(w), hb = lds128o<32>(w);
    const uint32_t H[8] = { ha.x, ha.y, ha.z, ha.w, hb.x, hb.y, hb.z, hb.w };

    int      isum[NC];
    uint32_t ac[NC];
#pragma unroll
    for (int c = 0; c < NC; c++) {
        isum[c] = 0;
        ac[c]   = am + c * cs_m;
    }

    static_for<0, 4>([&](auto gi) {
        constexpr int g = decltype(gi)::value;
        const uint4 qa = lds128o<48 + 32 * g>(w);
        const uint4 qb = lds128o<64 + 32 * g>(w);
        const uint32_t q[8] = { qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w };
        uint32_t vl[8], vh[8];
#pragma unroll
        for (int k = 0; k < 8; k++) {  // bit 2g -> bit 4 of the low value, bit 2g + 1 -> bit 4 of the high value
            const uint32_t bl = g == 0 ? H[k] << 4 : g == 1 ? H[k] << 2 : g == 2 ? H[k] : H[k] >> 2;
            const uint32_t bh = g == 0 ? H[k] << 3 : g == 1 ? H[k] << 1 : g == 2 ? H[k] >> 1 : H[k] >> 3;
            vl[k] = (q[k] & 0x0f0f0f0fu) | (bl & 0x10101010u);
            vh[k] = ((q[k] >> 4) & 0x0f0f0f0fu) | (bh & 0x10101010u);
        }
        const uint32_t scw  = g < 2 ? sc_lo : sc_hi;
        const int      s0   = (scw >> (16 * (g & 1))) & 0xff;
        const int      s1   = (scw >> (16 * (g & 1) + 8)) & 0xff;
#pragma unroll
        for (int c = 0; c < NC; c++) {
            const uint4 l0 = lds128o<64 * g>(ac[c]), l1 = lds128o<64 * g + 16>(ac[c]), h0 = lds128o<64 * g + 32>(ac[c]),
                        h1 = lds128o<64 * g + 48>(ac[c]);
            const int lo[8] = { (int) l0.x, (int) l0.y, (int) l0.z, (int) l0.w, (int) l1.x, (int) l1.y, (int) l1.z, (int) l1.w };
            const int hi[8] = { (int) h0.x, (int) h0.y, (int) h0.z, (int) h0.w, (int) h1.x, (int) h1.y, (int) h1.z, (int) h1.w };
            int sl = 0, sh = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                sl = dp4a_us(vl[k], lo[k], sl);
                sh = dp4a_us(vh[k], hi[k], sh);
            }
            isum[c] += s0 * sl + s1 * sh;
        }
    });
#pragma unroll
    for (int c = 0; c < NC; c++) {
        const uint4 bs = lds128o<256>(ac[c]);  // int16 sum of q per 32-block (x8)
        int imin = dp2a_lo_su((int) bs.x, mn_lo, 0);
        imin     = dp2a_hi_su((int) bs.y, mn_lo, imin);
        imin     = dp2a_lo_su((int) bs.z, mn_hi, imin);
        imin     = dp2a_hi_su((int) bs.w, mn_hi, imin);
        const float dy = __uint_as_float(lds32(ax + c * cs_x));
        acc[c] = fmaf(d * dy, (float) isum[c], acc[c]);
        acc[c] = fmaf(-(dmin * dy), (float) imin, acc[c]);
    }
}

// q6_K: { uint8 ql[128]; uint8 qh[64]; int8 scales[16]; half d } = 210 bytes, so a lane's unit is only 2-byte aligned in shared
// memory: every run of words is read as aligned 32-bit words and realigned with a funnel shift by 0 or 16 bits (lane dependent).
// Weight l of a 128-half: nibbles of ql[l] / ql[l + 32] plus two bits of qh[l], minus 32 (ggml-quants.c:1762-1791). The two
// fields occupy disjoint bits, so  signed byte = ((low + high + 0x60) ^ 0x80)  per byte (v in 0..63: v + 0x60 never carries and the
// flipped top bit is v - 32 as int8): one 3-input add and one xor per four weights, shared by all columns; dp4a.s32.s32 against
// the q8_K activations, one int8 scale per 16 weights applied in int32 exactly like vec_dot_q6_K_q8_K (ggml-cpu/quants.c:743-754).
template <int NC>
__device__ __forceinline__ void unit_dot_q6_K(uint32_t w, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, float (&acc)[NC]) {
    const uint32_t base = w & ~3u, sh = (w & 2u) * 8u;
    auto run4 = [&](int word0, uint32_t (&o)[4]) {  // words word0 .. word0 + 3 of the unit
        uint32_t r[5];
#pragma unroll
        for (int i = 0; i < 5; i++) r[i] = lds32(base + 4 * (word0 + i));
#pragma unroll
        for (int i = 0; i < 4; i++) o[i] = __funnelshift_r(r[i], r[i + 1], sh);
    };
    uint32_t scw[4];
    run4(48, scw);  // 16 int8 scales
    const float d = half_bits_to_float((uint16_t) (__funnelshift_r(lds32(base + 208), lds32(base + 212), sh) & 0xffff));
    int      isum[NC];
    uint32_t ac[NC];
#pragma unroll
    for (int c = 0; c < NC; c++) {
        isum[c] = 0;
        ac[c]   = am + c * cs_m;
    }
#pragma unroll
    for (int h = 0; h < 2; h++) {
#pragma unroll
        for (int gq = 0; gq < 2; gq++) {  // 16 values of l: one scale group in each of the four 32-weight quarters
            uint32_t A[4], B[4], H[4], s1[4], s2[4], s3[4], s4[4];
            run4(16 * h + 4 * gq, A);
            run4(16 * h + 8 + 4 * gq, B);
            run4(32 + 8 * h + 4 * gq, H);
#pragma unroll
            for (int i = 0; i < 4; i++) {
                s1[i] = ((A[i] & 0x0F0F0F0Fu) + ((H[i] << 4) & 0x30303030u) + 0x60606060u) ^ 0x80808080u;
                s2[i] = ((B[i] & 0x0F0F0F0Fu) + ((H[i] << 2) & 0x30303030u) + 0x60606060u) ^ 0x80808080u;
                s3[i] = (((A[i] >> 4) & 0x0F0F0F0Fu) + (H[i] & 0x30303030u) + 0x60606060u) ^ 0x80808080u;
                s4[i] = (((B[i] >> 4) & 0x0F0F0F0Fu) + ((H[i] >> 2) & 0x30303030u) + 0x60606060u) ^ 0x80808080u;
            }
            const int g1 = 8 * h + gq;  // scale groups g1, g1 + 2, g1 + 4, g1 + 6
            auto scale = [&](int g) { return (int) (int8_t) (scw[g >> 2] >> (8 * (g & 3))); };
            const int c1 = scale(g1), c2 = scale(g1 + 2), c3 = scale(g1 + 4), c4 = scale(g1 + 6);
#pragma unroll
            for (int c = 0; c < NC; c++) {
                const uint32_t yb = ac[c] + 128 * h + 16 * gq;
                const uint4 y1 = lds128(yb), y2 = lds128(yb + 32), y3 = lds128(yb + 64), y4 = lds128(yb + 96);
                auto dot4 = [](const uint32_t (&q)[4], const uint4 & y) {
                    int t = dp4a_ss((int) q[0], (int) y.x, 0);
                    t     = dp4a_ss((int) q[1], (int) y.y, t);
                    t     = dp4a_ss((int) q[2], (int) y.z, t);
                    return dp4a_ss((int) q[3], (int) y.w, t);
                };
                isum[c] += c1 * dot4(s1, y1) + c2 * dot4(s2, y2) + c3 * dot4(s3, y3) + c4 * dot4(s4, y4);
            }
        }
    }
#pragma unroll
    for (int c = 0; c < NC; c++) {
        const float dy = __uint_as_float(lds32(ax + c * cs_x));
        acc[c] = fmaf(d * dy, (float) isum[c], acc[c]);
    }
}

template <int TYPE, int NC>
__device__ __forceinline__ void unit_dot(uint32_t w, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, float (&acc)[NC]) {
    if (TYPE == B200_TYPE_Q6_K) unit_dot_q6_K<NC>(w, am, ax, cs_m, cs_x, acc);
    else if (TYPE == B200_TYPE_Q5_K) unit_dot_q5_K<NC>(w, am, ax, cs_m, cs_x, acc);
    else if (TYPE == B200_TYPE_Q4_K) unit_dot_q4_K<NC>(w, am, ax, cs_m, cs_x, acc);
    else if (TYPE == B200_TYPE_Q4_0) unit_dot_q4_0<NC>(w, am, ax, cs_m, cs_x, acc);
    else unit_dot_q8_0<NC>(w, am, ax, cs_m, cs_x, acc);
}

// ------------------------------------------------------------------------------------------------
// the persistent, warp-specialised kernel
//
// CTA = n_cons consumer warps + 1 producer warp; grid = one CTA per SM (x pairs for MUL_MAT_ID). A CTA owns a contiguous
// range of row groups; consumer warp w takes groups w, w + n_cons, ... of that range, each group in `slices` stages
// (a stage = up to 32 units: rpg rows x U units when U <= 32, else a 32-unit slice of one row).
//   producer warp : lane w is the producer of consumer warp w. It polls the consumer's "empty" mbarriers and issues
//                   cp.async.bulk copies (L2 evict_first) into that consumer's ring; completion arrives on "full".
//                   Weights never depend on an earlier kernel of the stream, so the producer starts immediately.
//   consumer warps: griddepcontrol.wait (programmatic dependent launch: the previous kernel of the stream has completed
//                   and its writes are visible) -> quantize the activations straight from the F32 src1 into shared
//                   memory (fused K1, same arithmetic as quantize.cu) or bulk-load pre-quantized unit records ->
//                   wait full / unit dot products / warp-shuffle row reduction / store / arrive empty.
// griddepcontrol.launch_dependents is executed at kernel start and the default configuration keeps shared memory
// <= 113 KB and registers <= 112, so the NEXT mat-vec kernel of the stream becomes co-resident on every SM and streams
// its first stages while this one is still computing: HBM stays busy across kernel boundaries (a 4096x4096 q4 matrix
// is 1.4 us of HBM time - shorter than a launch + drain).
//
// shared memory map (dynamic, 128-byte aligned):
//   [0, 8)                      : act mbarrier (pre-quantized activation loads)
//   [64, 64 + 8*NWC*NS)         : full barriers ; then 8*NWC*NS empty barriers
//   [act_main: NC*U*272][act_aux: NC*U*16]
//   [ring: NWC*NS*stage_bytes]
// ------------------------------------------------------------------------------------------------

constexpr int MV_HDR_BYTES = 64 + 256;   // act barrier | 64 unit sums of a fused RMS_NORM (MvNorm)

__host__ __device__ inline size_t mv_align(size_t x, size_t a) { return (x + a - 1) / a * a; }

size_t mv_fast_smem(int type, int U, int ncols, int n_cons, int n_stages) {
    const size_t stage = 32 * (size_t) unit_bytes(type);  // a stage always has room for 32 units
    size_t s = MV_HDR_BYTES;
    s += (size_t) 16 * n_cons * n_stages;  // full + empty barriers
    s = mv_align(s, 128);
    s += (size_t) ncols * U * (ACT_MAIN + ACT_AUX);
    s = mv_align(s, 128);
    s += (size_t) n_cons * n_stages * stage;
    if (type == B200_TYPE_Q6_K) s += 16;  // the realigning loads of the last unit read up to 6 bytes past it
    return s;
}

__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// MvNorm inside the fused K1 prologue: v[b] = unit t0 + b * stride of the row (valid while < n_units). Every quantizing warp calls it
// exactly once (also a warp without a unit: the barrier counts it).
// this warp's share of the norm weights (they do not depend on the previous kernel: the lean kernel requests them before its
// griddepcontrol.wait - measured on the device timeline, the norm prologue took 2.2 - 3.2 us against 1.0 us without a norm, the
// difference being these loads queueing behind the weight stream)
template <int NB>
__device__ __forceinline__ void norm_load_w(float (&wv)[NB][8], const MvNorm & nm, int t0, int stride, int n_units, int lane) {
#pragma unroll
    for (int b = 0; b < NB; b++) {
        const int t = t0 + b * stride;
        if (t < n_units && nm.w) {
            const float4 e = __ldg((const float4 *) nm.w + t * 64 + lane * 2), f = __ldg((const float4 *) nm.w + t * 64 + lane * 2 + 1);
            wv[b][0] = e.x; wv[b][1] = e.y; wv[b][2] = e.z; wv[b][3] = e.w; wv[b][4] = f.x; wv[b][5] = f.y; wv[b][6] = f.z; wv[b][7] = f.w;
        }
    }
}
template <bool PRE, int NB>
__device__ __forceinline__ void norm_units(float (&v)[NB][8], const MvNorm & nm, int t0, int stride, int n_units, int lane, float * part, int n_threads,
                                           float (&wv)[NB][8]) {
    if (!PRE) norm_load_w<NB>(wv, nm, t0, stride, n_units, lane);   // requested before the row's values are waited for
#pragma unroll
    for (int b = 0; b < NB; b++) {
        const int t = t0 + b * stride;
        if (t < n_units) {
            const float sq = unit_sumsq(v[b]);
            if (lane == 0) part[t] = sq;
        }
    }
    asm volatile("bar.sync 3, %0;" ::"r"(n_threads) : "memory");
    const float scale = rms_scale(part, n_units, nm.eps);
#pragma unroll
    for (int b = 0; b < NB; b++) {
        if (t0 + b * stride < n_units) {
#pragma unroll
            for (int i = 0; i < 8; i++) { v[b][i] = v[b][i] * scale; if (nm.w) v[b][i] *= wv[b][i]; }
        }
    }
}
__device__ __forceinline__ void bar_sync_named(int id, int n_threads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n_threads) : "memory"); }
__device__ __forceinline__ bool mbar_test_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define MV_TRACE(i) do { if (p.trace && lane == 0) p.trace[(size_t) blockIdx.x * 8 + (i)] = global_ns(); } while (0)

__device__ __forceinline__ int mv_find_mat(const MvParams & p, int group) {
    int m = 0;
#pragma unroll
    for (int i = 1; i < MV_MAX_MATS; i++) if (i < p.n_mats && group >= p.mat[i].group0) m = i;
    return m;
}

// WIDE: up to 16 consumer warps and the SM to itself (<= 120 registers); otherwise up to 8 consumer warps and, for <= 4
// columns, two CTAs per SM (<= 112 registers) so that consecutive kernels of the stream overlap
// WCLASS (CTA shapes): 0 = up to 8 consumer warps, two CTAs per SM for <= 4 columns; 1 = up to 16 consumer warps, the SM to
// itself; 2 = up to 11 consumer warps and two CTAs per SM (<= 80 registers): consecutive kernels of a stream co-reside
constexpr int mv_class_warps(int c) { return c == 1 ? MV_MAX_CONS + 1 : c == 2 ? 12 : 9; }
// EPI: the K7 epilogues (bias rows, GLU of a gate / up matrix pair; kernels.h MvEpi) - one column only, own instantiation so that
// the plain mat-vec keeps its code
template <int TYPE, int NC, int WCLASS, bool EPI = false>
__global__ void __launch_bounds__(32 * mv_class_warps(WCLASS), (WCLASS == 2 || (WCLASS == 0 && NC <= 4)) ? 2 : 1)
mv_fast_kernel(const __grid_constant__ MvParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int UB    = unit_bytes(TYPE);
    constexpr int STAGE = 32 * UB;
    constexpr int KIND  = act_kind_for(TYPE);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int NWC = p.n_cons, NS = p.n_stages;
    const int U = p.U;

    uint64_t * act_bar = (uint64_t *) smem;
    uint64_t * full    = (uint64_t *) (smem + MV_HDR_BYTES);
    uint64_t * empty   = full + NWC * NS;
    size_t     off     = mv_align(MV_HDR_BYTES + (size_t) 16 * NWC * NS, 128);
    uint8_t *  act_m   = smem + off;
    uint8_t *  act_x   = act_m + (size_t) NC * U * ACT_MAIN;
    off                = mv_align(off + (size_t) NC * U * (ACT_MAIN + ACT_AUX), 128);
    uint8_t *  ring0   = smem + off;

    // MUL_MAT_ID: this CTA row (blockIdx.y) is one (slot, token) pair with its own expert, activation row and dst column
    int64_t w_off = 0, dst_off = 0, x_off = 0;
    int     act_col = 0, expert = 0;  // expert: row of the ADD_ID bias
    if (p.ids.ids) {
        const int pair = blockIdx.y;
        const int t = pair / p.ids.n_used, i = pair - t * p.ids.n_used;
        const int e = p.ids.ids[(int64_t) t * p.ids.ids_rs + i];
        if (e < 0 || e >= p.ids.n_as) {  // ggml-cpu asserts here (ggml-cpu.c:1574); never read out of bounds
            if (p.split.n_ranks > 1 && warp == 0) split_complete(p.split, gridDim.x * gridDim.y);
            return;
        }
        w_off   = (int64_t) e * p.ids.nb02;
        expert  = e;
        dst_off = (int64_t) t * p.ids.dst_ts + (int64_t) i * p.ids.dst_is;
        act_col = t * p.ids.ne11 + i % p.ids.ne11;
        x_off   = (int64_t) t * p.ids.x_ts + (int64_t) (i % p.ids.ne11) * p.ids.x_is;
    }

    // this CTA's contiguous range of row groups
    const int g_begin = (int) ((int64_t) p.total_groups * blockIdx.x / gridDim.x);
    const int g_end   = (int) ((int64_t) p.total_groups * (blockIdx.x + 1) / gridDim.x);

    // barrier set-up by the producer warp: lane w initialises the ring barriers of consumer warp w and starts copying at once;
    // the consumers only have to see them before their first wait (CTA barrier 0: the producer arrives, the consumers sync)
    if (warp == NWC) {
        if (lane == 0) mbar_init(smem_u32(act_bar), 1);
        if (lane < NWC) {
            for (int i = 0; i < NS; i++) {
                mbar_init(smem_u32(full + lane * NS + i), 1);
                mbar_init(smem_u32(empty + lane * NS + i), 1);
            }
        }
        fence_mbar_init();
        __syncwarp();
        asm volatile("bar.arrive 0, %0;" ::"r"(blockDim.x) : "memory");
    } else {
        asm volatile("bar.sync 0, %0;" ::"r"(blockDim.x) : "memory");
    }
    griddep_launch_dependents();
    if (p.trace && threadIdx.x == 0) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        p.trace[(size_t) blockIdx.x * 8 + 0] = global_ns();
        (void) smid;
    }

    const uint64_t pol_stream = l2_policy_evict_first();
    // one stage of consumer warp w_: row group j_ of the warp's list, 32-unit slice slice_ -> ring stage stage_
    // K7 GLU: a group is the gate rows followed by the same rows of the up matrix: SPG = 2 * slices stages per group
    const bool GLU = EPI && p.epi.glu != 0;
    const int  SPG = GLU ? 2 * p.slices : p.slices;
    auto issue = [&](int w_, int stage_, int j_, int slice_) {
        const int     group = g_begin + w_ + j_ * NWC;
        const int     half_ = GLU && slice_ >= p.slices ? 1 : 0;
        if (half_) slice_ -= p.slices;
        const MvMat & mat   = p.mat[GLU ? half_ : mv_find_mat(p, group)];
        const int     row0  = (group - mat.group0) * p.rpg;
        const int     nrow  = min(p.rpg, mat.M - row0);
        const int     u0    = slice_ * 32;
        const int     nu    = min(32, U - u0);
        const uint32_t bar  = smem_u32(full + w_ * NS + stage_);
        const uint32_t dst  = smem_u32(ring0 + ((size_t) w_ * NS + stage_) * STAGE);
        const uint32_t row_bytes = (uint32_t) nu * UB;
        mbar_expect_tx(bar, row_bytes * nrow);
        const uint8_t * src = mat.w + w_off + (int64_t) row0 * mat.nb01 + (int64_t) u0 * UB;
        if (nrow > 1 && mat.nb01 == (int64_t) U * UB) {
            bulk_g2s(dst, src, row_bytes * nrow, bar, pol_stream);  // rows are adjacent in memory: one copy
        } else {
            for (int r = 0; r < nrow; r++) bulk_g2s(dst + r * row_bytes, src + (int64_t) r * mat.nb01, row_bytes, bar, pol_stream);
        }
    };
    if (warp == NWC) {
        // ================================================================ producer warp: lane w feeds consumer warp w
        const int w = lane;
        int my_total = 0;
        if (w < NWC && g_end - g_begin > w) my_total = ((g_end - g_begin - w + NWC - 1) / NWC) * SPG;
        int issued = 0, stage = 0, j = 0, slice = 0;
        uint32_t phase = 1;  // parity of the "previous" phase of a fresh barrier: the first NS waits fall through
        while (__any_sync(0xffffffffu, issued < my_total)) {
            bool can = issued < my_total;
            if (can) can = mbar_test_wait(smem_u32(empty + w * NS + stage), phase);
            if (can) {
                if (p.trace && w == 0 && issued == 0) p.trace[(size_t) blockIdx.x * 8 + 1] = global_ns();
                issue(w, stage, j, slice);
                if (p.trace && w == 0 && issued == 0) p.trace[(size_t) blockIdx.x * 8 + 2] = global_ns();
                issued++;
                if (++stage == NS) { stage = 0; phase ^= 1; }
                if (++slice == SPG) { slice = 0; j++; }
            }
        }
        // every copy of this kernel is in flight: keep HBM busy across the kernel boundary with the head of the predicted
        // next kernel's weight stream (see MvNext)
        if (p.next.n_mats > 0) {
            mv_prefetch_next(p.next, (int) blockIdx.x, (int) gridDim.x, lane);
        }
        return;
    }
    if (warp > NWC) return;

    // ==================================================================== consumer warps
    const int w = warp;
    int my_total = 0;
    if (g_end - g_begin > w) my_total = ((g_end - g_begin - w + NWC - 1) / NWC) * SPG;

    griddep_wait();  // everything below reads src1 / writes dst, which earlier kernels of the stream may produce / still read

    if (w == 0) MV_TRACE(3);
    const bool     ll = p.split.n_ranks > 1 && p.split.ll;
    const uint32_t ep = ll ? ll_epoch(p.split) : 0u;
    if (p.split.n_ranks > 1 && p.split.deferred && !ll) split_wait(p.split);  // src1 may derive from a dst whose remote slices are still in flight
    if (p.fuse_quant) {
        // fused K1: unit (c, u) of the activations, F32 global -> unit record in shared memory. Up to 4 units per warp are
        // loaded before the first is quantized, so the L2 round trips overlap instead of chaining
        const int n_tasks = NC * U;
        const bool norm_on = EPI && p.norm.on;   // one column, n_tasks <= 4 * NWC (the launcher's condition): a single round
        for (int t0 = w; t0 < n_tasks || (norm_on && t0 == w); t0 += 4 * NWC) {
            float v[4][8];
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const int t = t0 + b * NWC;
                if (t < n_tasks) {
                    const int c = t / U, u = t - c * U;
                    const float * xr = (const float *) ((const char *) p.x + x_off + (int64_t) c * p.x_cs) + (int64_t) u * UNIT_ELEMS + lane * 8;
                    if (p.x_vec) {
                        const float4 a = __ldg((const float4 *) xr);
                        const float4 bb = __ldg((const float4 *) xr + 1);
                        v[b][0] = a.x; v[b][1] = a.y; v[b][2] = a.z; v[b][3] = a.w;
                        v[b][4] = bb.x; v[b][5] = bb.y; v[b][6] = bb.z; v[b][7] = bb.w;
                    } else {
#pragma unroll
                        for (int i = 0; i < 8; i++) v[b][i] = __ldg(xr + i);
                    }
                }
            }
            if (p.trace && w == 0 && t0 == w) {  // the first batch's loads have returned
                asm volatile("" : "+f"(v[0][0]), "+f"(v[0][7]));
                MV_TRACE(7);
            }
            if constexpr (EPI) { if (norm_on) { float wv[4][8]; norm_units<false, 4>(v, p.norm, t0, NWC, n_tasks, lane, (float *) (smem + 64), NWC * 32, wv); } }
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const int t = t0 + b * NWC;
                if (t < n_tasks) quant_unit_store<KIND>(v[b], lane, act_m + (size_t) t * ACT_MAIN, act_x + (size_t) t * ACT_AUX);
            }
        }
        bar_sync_named(1, NWC * 32);
    } else {
        if (w == 0 && lane == 0) {  // pre-quantized activations: NC columns of U records (main + aux)
            const uint64_t pol_keep = l2_policy_evict_last();
            const uint32_t bytes_m = (uint32_t) NC * U * ACT_MAIN, bytes_x = (uint32_t) NC * U * ACT_AUX;
            mbar_expect_tx(smem_u32(act_bar), bytes_m + bytes_x);
            bulk_g2s(smem_u32(act_m), p.act_main + (size_t) act_col * U * ACT_MAIN, bytes_m, smem_u32(act_bar), pol_keep);
            bulk_g2s(smem_u32(act_x), p.act_aux + (size_t) act_col * U * ACT_AUX, bytes_x, smem_u32(act_bar), pol_keep);
        }
        mbar_wait(smem_u32(act_bar), 0);
    }

    if (w == 0) MV_TRACE(4);
    const uint32_t am0  = smem_u32(act_m), ax0 = smem_u32(act_x);
    const uint32_t cs_m = (uint32_t) U * ACT_MAIN, cs_x = (uint32_t) U * ACT_AUX;
    const uint32_t ring = smem_u32(ring0 + (size_t) w * NS * STAGE);

    const int L   = p.L;
    const int sub = lane & (L - 1);  // unit index inside the row slice
    const int rr  = lane / L;        // row inside the group

    float acc[NC];
#pragma unroll
    for (int c = 0; c < NC; c++) acc[c] = 0.0f;
    float gate_v = 0.0f;  // K7: the finished gate row while the up row is being summed

    int stage = 0, j = 0, slice = 0;
    uint32_t phase = 0;
    for (int q = 0; q < my_total; q++) {
        mbar_wait(smem_u32(full + w * NS + stage), phase);
        if (w == 0 && q == 0) MV_TRACE(5);
        const int group = g_begin + w + j * NWC;
        const int half  = GLU && slice >= p.slices ? 1 : 0;
        const int sl    = slice - half * p.slices;
        const int m     = GLU ? half : mv_find_mat(p, group);
        const int row0  = (group - p.mat[m].group0) * p.rpg;
        const int M     = p.mat[m].M;
        const int u0    = sl * 32;
        const int nu    = min(32, U - u0);
        const int row   = row0 + rr;
        const int u     = u0 + sub;
        const bool mine = sub < nu && row < M && rr < p.rpg;

        if (mine) {
            const uint32_t wp = ring + (uint32_t) stage * STAGE + (uint32_t) (rr * nu + sub) * UB;
            unit_dot<TYPE, NC>(wp, am0 + (uint32_t) u * ACT_MAIN, ax0 + (uint32_t) u * ACT_AUX, cs_m, cs_x, acc);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(empty + w * NS + stage));  // the stage's bytes are in registers: hand it back

        if (sl == p.slices - 1) {  // row finished: reduce the L lanes of each row and store
#pragma unroll
            for (int c = 0; c < NC; c++) {
                float v = acc[c];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const float t = __shfl_xor_sync(0xffffffffu, v, o);
                    if (o < L) v += t;
                }
                if (sub == 0 && row < M && rr < p.rpg) {
                    if (EPI && p.mat[m].bias) v += p.mat[m].bias[(int64_t) expert * p.mat[m].bias_es + row];  // ADD / ADD_ID of the bias row
                    if (GLU) {
                        if (half == 0) gate_v = v;
                        else p.mat[0].dst[dst_off + row] = glu_apply(p.epi.glu - 1, gate_v, v, p.epi.alpha, p.epi.limit);
                    } else {
                        float * a = p.mat[m].dst + dst_off + (int64_t) c * p.mat[m].dst_cs + row;
                        if (ll) ll_store(p.split, ep, a, v);
                        else split_store(p.split, a, v);
                    }
                }
                acc[c] = 0.0f;
            }
        }
        if (++stage == NS) { stage = 0; phase ^= 1; }
        if (++slice == SPG) { slice = 0; j++; }
    }
    if (w == 0) MV_TRACE(6);
    if (ll) {  // flagged-payload exchange: this CTA's rows are on their way to every rank; collect this CTA's share of the peers' rows
        ll_receive(p.split, ep, p.geom, p.mat, p.n_mats, NC, (int) blockIdx.x * NWC * 32 + w * 32 + lane, (int) gridDim.x * NWC * 32);
        bar_sync_named(1, NWC * 32);
        if (w == 0 && lane == 0) ll_finish(p.split, ep, gridDim.x);
    } else if (p.split.n_ranks > 1) {  // fused all-gather: this CTA's slice is on its way to every rank; the last CTA exchanges flags
        bar_sync_named(1, NWC * 32);
        if (w == 0) split_complete(p.split, gridDim.x * gridDim.y);
    }
}


// ------------------------------------------------------------------------------------------------
// the lean one-column kernel (round 2): the same ring / producer / fused-K1 structure as mv_fast_kernel for the case a decode
// token is made of - one activation column, plain MUL_MAT (no expert routing, no row-split exchange, no K7 epilogue) - with
// the consumer loop rebuilt around what the round-1 profile showed: the kernel was ISSUE-bound at ~370 instructions per
// 256 weights (ncu: issue active 47 %, 23.6 M warp instructions for lm_head), half of them bookkeeping. Here
//   * a warp consumes its row groups in PAIRS: the two stages of a pair hold the same unit indices of different rows, so a lane
//     loads its activation unit from shared memory once for two weight units (27 -> 18 ld.shared.v4 per unit) and the two
//     independent dot-product chains give the scheduler twice the instruction-level parallelism;
//   * everything that does not change per stage (matrix lookup, row numbers, unit addresses) is computed once per pair,
//     and the store is a plain st.global.
// The ring order follows the pairs: (group A slice 0, group B slice 0, A slice 1, B slice 1, ...), an even number of stages.
// ------------------------------------------------------------------------------------------------

// two weight units (w0, w1: same unit index, different rows) against one activation unit
template <int NC>
__device__ __forceinline__ void unit_dot2_q4_K(uint32_t w0, uint32_t w1, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, float (&acc0)[NC],
                                               float (&acc1)[NC]) {
    const uint4 hA = lds128o<0>(w0), hB = lds128o<0>(w1);
    const uint32_t scA_lo = hA.y & 0x3f3f3f3fu, scA_hi = (hA.w & 0x0f0f0f0fu) | ((hA.y >> 2) & 0x30303030u);
    const uint32_t scB_lo = hB.y & 0x3f3f3f3fu, scB_hi = (hB.w & 0x0f0f0f0fu) | ((hB.y >> 2) & 0x30303030u);
    int isA[NC], isB[NC];
    uint32_t ac[NC];
#pragma unroll
    for (int c = 0; c < NC; c++) { isA[c] = 0; isB[c] = 0; ac[c] = am + c * cs_m; }
    static_for<0, 4>([&](auto gi) {
        constexpr int g = decltype(gi)::value;
        const uint4 a0 = lds128o<16 + 32 * g>(w0), a1 = lds128o<32 + 32 * g>(w0);
        const uint4 b0 = lds128o<16 + 32 * g>(w1), b1 = lds128o<32 + 32 * g>(w1);
        const uint32_t qa[8] = { a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w };
        const uint32_t qb[8] = { b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w };
        const uint32_t wA = g < 2 ? scA_lo : scA_hi, wB = g < 2 ? scB_lo : scB_hi;
        const int sA0 = (int) ((wA >> (16 * (g & 1))) & 0xff), sA1 = (int) ((wA >> (16 * (g & 1) + 8)) & 0xff);
        const int sB0 = (int) ((wB >> (16 * (g & 1))) & 0xff), sB1 = (int) ((wB >> (16 * (g & 1) + 8)) & 0xff);
#pragma unroll
        for (int c = 0; c < NC; c++) {
            const uint4 l0 = lds128o<64 * g>(ac[c]), l1 = lds128o<64 * g + 16>(ac[c]), h0 = lds128o<64 * g + 32>(ac[c]), h1 = lds128o<64 * g + 48>(ac[c]);
            const int lo[8] = { (int) l0.x, (int) l0.y, (int) l0.z, (int) l0.w, (int) l1.x, (int) l1.y, (int) l1.z, (int) l1.w };
            const int hi[8] = { (int) h0.x, (int) h0.y, (int) h0.z, (int) h0.w, (int) h1.x, (int) h1.y, (int) h1.z, (int) h1.w };
            int sAl = 0, sAh = 0, sBl = 0, sBh = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                sAl = dp4a_us(qa[k] & 0x0f0f0f0fu, lo[k], sAl);
                sBl = dp4a_us(qb[k] & 0x0f0f0f0fu, lo[k], sBl);
                sAh = dp4a_us(qa[k] & 0xf0f0f0f0u, hi[k], sAh);  // 16 x the high-nibble dot product
                sBh = dp4a_us(qb[k] & 0xf0f0f0f0u, hi[k], sBh);
            }
            isA[c] += sA0 * sAl + sA1 * (sAh >> 4);
            isB[c] += sB0 * sBl + sB1 * (sBh >> 4);
        }
    });
    const uint32_t mnA_lo = hA.z & 0x3f3f3f3fu, mnA_hi = ((hA.w >> 4) & 0x0f0f0f0fu) | ((hA.z >> 2) & 0x30303030u);
    const uint32_t mnB_lo = hB.z & 0x3f3f3f3fu, mnB_hi = ((hB.w >> 4) & 0x0f0f0f0fu) | ((hB.z >> 2) & 0x30303030u);
    const float dA = half_bits_to_float((uint16_t) (hA.x & 0xffff)), mA_ = half_bits_to_float((uint16_t) (hA.x >> 16));
    const float dB = half_bits_to_float((uint16_t) (hB.x & 0xffff)), mB_ = half_bits_to_float((uint16_t) (hB.x >> 16));
#pragma unroll
    for (int c = 0; c < NC; c++) {
        const uint4 bs = lds128o<256>(ac[c]);  // int16 sum of q per 32-block (x8)
        const float dy = __uint_as_float(lds32(ax + c * cs_x));
        int mA = dp2a_lo_su((int) bs.x, mnA_lo, 0), mB = dp2a_lo_su((int) bs.x, mnB_lo, 0);
        mA = dp2a_hi_su((int) bs.y, mnA_lo, mA); mB = dp2a_hi_su((int) bs.y, mnB_lo, mB);
        mA = dp2a_lo_su((int) bs.z, mnA_hi, mA); mB = dp2a_lo_su((int) bs.z, mnB_hi, mB);
        mA = dp2a_hi_su((int) bs.w, mnA_hi, mA); mB = dp2a_hi_su((int) bs.w, mnB_hi, mB);
        // the same two fused multiply-adds per unit and column as unit_dot_q4_K: results are bit-identical to the one-unit path
        acc0[c] = fmaf(dA * dy, (float) isA[c], acc0[c]);
        acc0[c] = fmaf(-(mA_ * dy), (float) mA, acc0[c]);
        acc1[c] = fmaf(dB * dy, (float) isB[c], acc1[c]);
        acc1[c] = fmaf(-(mB_ * dy), (float) mB, acc1[c]);
    }
}

// q4_0 / q8_0 pairs: the unit is walked in two halves of four blocks, so that two rows' weights fit in registers next to the
// activation words they share. Same operations per block and row as unit_dot_q4_0 / unit_dot_q8_0 (bit-identical results).
template <int NC>
__device__ __forceinline__ void unit_dot2_q4_0(uint32_t w0, uint32_t w1, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, float (&acc0)[NC],
                                               float (&acc1)[NC]) {
    uint4    dyh[NC], ys[NC];
    uint32_t ac[NC];
#pragma unroll
    for (int c = 0; c < NC; c++) {
        ac[c]  = am + c * cs_m;
        dyh[c] = lds128o<256>(ac[c]);      // half d[8]
        ys[c]  = lds128(ax + c * cs_x);    // int16 sum[8]
    }
    static_for<0, 2>([&](auto hi_) {
        constexpr int H  = decltype(hi_)::value;
        constexpr int W0 = 16 * H;        // first word held: words W0 .. W0 + 19 cover block pairs 2H and 2H + 1 (words 18H .. 18H + 17)
        uint32_t a[20], b[20];
        static_for<0, 5>([&](auto ii) {
            constexpr int i = decltype(ii)::value;
            const uint4 v = lds128o<4 * W0 + 16 * i>(w0), u = lds128o<4 * W0 + 16 * i>(w1);
            a[4 * i] = v.x; a[4 * i + 1] = v.y; a[4 * i + 2] = v.z; a[4 * i + 3] = v.w;
            b[4 * i] = u.x; b[4 * i + 1] = u.y; b[4 * i + 2] = u.z; b[4 * i + 3] = u.w;
        });
        static_for<0, 4>([&](auto bi) {
            constexpr int blk = 4 * H + decltype(bi)::value;
            constexpr int p   = 9 * (blk >> 1) - W0;  // index of the block pair's first word in a[] / b[]
            float dA, dB;
            uint32_t qa[4], qb[4];
            if ((blk & 1) == 0) {
                dA = half_bits_to_float((uint16_t) (a[p] & 0xffff)); dB = half_bits_to_float((uint16_t) (b[p] & 0xffff));
#pragma unroll
                for (int k = 0; k < 4; k++) { qa[k] = __funnelshift_r(a[p + k], a[p + k + 1], 16); qb[k] = __funnelshift_r(b[p + k], b[p + k + 1], 16); }
            } else {
                dA = half_bits_to_float((uint16_t) (a[p + 4] >> 16)); dB = half_bits_to_float((uint16_t) (b[p + 4] >> 16));
#pragma unroll
                for (int k = 0; k < 4; k++) { qa[k] = a[p + 5 + k]; qb[k] = b[p + 5 + k]; }
            }
#pragma unroll
            for (int c = 0; c < NC; c++) {
                const uint4 l = lds128o<32 * blk>(ac[c]), h = lds128o<32 * blk + 16>(ac[c]);
                const int lo[4] = { (int) l.x, (int) l.y, (int) l.z, (int) l.w }, hi[4] = { (int) h.x, (int) h.y, (int) h.z, (int) h.w };
                const uint32_t dword = blk < 2 ? dyh[c].x : blk < 4 ? dyh[c].y : blk < 6 ? dyh[c].z : dyh[c].w;
                const uint32_t sword = blk < 2 ? ys[c].x : blk < 4 ? ys[c].y : blk < 6 ? ys[c].z : ys[c].w;
                const float    dy    = half_bits_to_float((uint16_t) ((blk & 1) ? dword >> 16 : dword & 0xffff));
                const int      ysum  = (int) (int16_t) ((blk & 1) ? sword >> 16 : sword & 0xffff);
                int sAl = 0, sAh = 0, sBl = 0, sBh = 0;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    sAl = dp4a_us(qa[k] & 0x0f0f0f0fu, lo[k], sAl); sAh = dp4a_us(qa[k] & 0xf0f0f0f0u, hi[k], sAh);
                    sBl = dp4a_us(qb[k] & 0x0f0f0f0fu, lo[k], sBl); sBh = dp4a_us(qb[k] & 0xf0f0f0f0u, hi[k], sBh);
                }
                acc0[c] = fmaf((float) (sAl + (sAh >> 4) - 8 * ysum) * dA, dy, acc0[c]);
                acc1[c] = fmaf((float) (sBl + (sBh >> 4) - 8 * ysum) * dB, dy, acc1[c]);
            }
        });
    });
}

template <int NC>
__device__ __forceinline__ void unit_dot2_q8_0(uint32_t w0, uint32_t w1, uint32_t am, uint32_t cs_m, float (&acc0)[NC], float (&acc1)[NC]) {
    uint4    dyh[NC];
    uint32_t ac[NC];
#pragma unroll
    for (int c = 0; c < NC; c++) { ac[c] = am + c * cs_m; dyh[c] = lds128o<256>(ac[c]); }
    static_for<0, 4>([&](auto pi) {
        constexpr int P = decltype(pi)::value;           // block pair: 68 bytes = words 17P .. 17P + 16
        constexpr int W0 = (17 * P) & ~3;                 // 16-byte aligned start; up to 3 words of slack in front
        constexpr int NW = ((17 * P + 17 - W0) + 3) / 4;  // 128-bit loads needed (5 or 6)
        uint32_t a[4 * NW], b[4 * NW];
        static_for<0, NW>([&](auto ii) {
            constexpr int i = decltype(ii)::value;
            if constexpr (4 * W0 + 16 * i < 272) {
                const uint4 v = lds128o<4 * W0 + 16 * i>(w0), u = lds128o<4 * W0 + 16 * i>(w1);
                a[4 * i] = v.x; a[4 * i + 1] = v.y; a[4 * i + 2] = v.z; a[4 * i + 3] = v.w;
                b[4 * i] = u.x; b[4 * i + 1] = u.y; b[4 * i + 2] = u.z; b[4 * i + 3] = u.w;
            }
        });
        constexpr int p = 17 * P - W0;
        static_for<0, 2>([&](auto ei) {
            constexpr int e = decltype(ei)::value, blk = 2 * P + e;
            float dA, dB;
            uint32_t qa[8], qb[8];
            if (e == 0) {
                dA = half_bits_to_float((uint16_t) (a[p] & 0xffff)); dB = half_bits_to_float((uint16_t) (b[p] & 0xffff));
#pragma unroll
                for (int k = 0; k < 8; k++) { qa[k] = __funnelshift_r(a[p + k], a[p + k + 1], 16); qb[k] = __funnelshift_r(b[p + k], b[p + k + 1], 16); }
            } else {
                dA = half_bits_to_float((uint16_t) (a[p + 8] >> 16)); dB = half_bits_to_float((uint16_t) (b[p + 8] >> 16));
#pragma unroll
                for (int k = 0; k < 8; k++) { qa[k] = a[p + 9 + k]; qb[k] = b[p + 9 + k]; }
            }
#pragma unroll
            for (int c = 0; c < NC; c++) {
                const uint4 y0 = lds128o<32 * blk>(ac[c]), y1 = lds128o<32 * blk + 16>(ac[c]);
                const int y[8] = { (int) y0.x, (int) y0.y, (int) y0.z, (int) y0.w, (int) y1.x, (int) y1.y, (int) y1.z, (int) y1.w };
                const uint32_t dword = blk < 2 ? dyh[c].x : blk < 4 ? dyh[c].y : blk < 6 ? dyh[c].z : dyh[c].w;
                const float    dy    = half_bits_to_float((uint16_t) ((blk & 1) ? dword >> 16 : dword & 0xffff));
                int sA = 0, sB = 0;
#pragma unroll
                for (int k = 0; k < 8; k++) { sA = dp4a_ss((int) qa[k], y[k], sA); sB = dp4a_ss((int) qb[k], y[k], sB); }
                acc0[c] = fmaf((float) sA, dA * dy, acc0[c]);
                acc1[c] = fmaf((float) sB, dB * dy, acc1[c]);
            }
        });
    });
}

template <int TYPE, int NC>
__device__ __forceinline__ void unit_dot2(uint32_t w0, uint32_t w1, uint32_t am, uint32_t ax, uint32_t cs_m, uint32_t cs_x, float (&acc0)[NC],
                                          float (&acc1)[NC]) {
    if (TYPE == B200_TYPE_Q4_K) {
        unit_dot2_q4_K<NC>(w0, w1, am, ax, cs_m, cs_x, acc0, acc1);
    } else if (TYPE == B200_TYPE_Q4_0) {
        unit_dot2_q4_0<NC>(w0, w1, am, ax, cs_m, cs_x, acc0, acc1);
    } else if (TYPE == B200_TYPE_Q8_0) {
        unit_dot2_q8_0<NC>(w0, w1, am, cs_m, acc0, acc1);
    } else {  // the other formats keep their one-unit dot products
        unit_dot<TYPE, NC>(w0, am, ax, cs_m, cs_x, acc0);
        unit_dot<TYPE, NC>(w1, am, ax, cs_m, cs_x, acc1);
    }
}

// LL: row split with the flagged-payload exchange (its own instantiation: the plain kernel keeps its registers). BIAS: the ADD of a
// K7 epilogue without GLU (attn_output / ffn_down + residual in a decode graph): mat.bias[row] is added before the store
// NORM: MvNorm - the RMS_NORM + MUL in front of the q/k/v projections happens in this kernel's K1 prologue.
// GLU: the K7 gate / up pair (MvEpi): the two stages of a pair are the gate rows and the up rows of ONE row group, and the pair's
// two sums become glu(gate + bias, up + bias) in mat[0].dst - the stage pairs this kernel is built around are exactly that.
// ROPE: MvRope - the launch is q / k / v of one token; the rows leave the kernel rotated / as KV-cache rows.
// warps a CTA of the lean kernel may have: 16 quantizing warps + the producer for one column; 12 + producer from two columns on - ptxas
// sizes registers for the bound (96 at 17 warps: the two- to four-column dot products spilled up to 136 bytes; 168 at 13 warps)
constexpr int mv_lean_max_cons(int nc) { return nc == 1 ? MV_MAX_CONS : 12; }
template <int TYPE, int NC, bool LL, bool BIAS = false, bool NORM = false, bool GLU = false, bool ROPE = false>
__global__ void __launch_bounds__(32 * (mv_lean_max_cons(NC) + 1), 1)
mv_lean_kernel(const __grid_constant__ MvParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int UB    = unit_bytes(TYPE);
    constexpr int STAGE = 32 * UB;
    constexpr int KIND  = act_kind_for(TYPE);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int NWC = p.n_cons, NS = p.n_stages;  // NS is even
    const int U = p.U, SL = p.slices;

    uint64_t * act_bar = (uint64_t *) smem;
    uint64_t * full    = (uint64_t *) (smem + MV_HDR_BYTES);
    uint64_t * empty   = full + NWC * NS;
    size_t     off     = mv_align(MV_HDR_BYTES + (size_t) 16 * NWC * NS, 128);
    uint8_t *  act_m   = smem + off;
    uint8_t *  act_x   = act_m + (size_t) NC * U * ACT_MAIN;
    off                = mv_align(off + (size_t) NC * U * (ACT_MAIN + ACT_AUX), 128);
    uint8_t *  ring0   = smem + off;

    const int g_begin = (int) ((int64_t) p.total_groups * blockIdx.x / gridDim.x);
    const int g_end   = (int) ((int64_t) p.total_groups * (blockIdx.x + 1) / gridDim.x);

    if (warp == NWC) {
        if (lane == 0) mbar_init(smem_u32(act_bar), 1);
        if (lane < NWC) {
            for (int i = 0; i < NS; i++) {
                mbar_init(smem_u32(full + lane * NS + i), 1);
                mbar_init(smem_u32(empty + lane * NS + i), 1);
            }
        }
        fence_mbar_init();
        __syncwarp();
        asm volatile("bar.arrive 0, %0;" ::"r"(blockDim.x) : "memory");
    } else {
        asm volatile("bar.sync 0, %0;" ::"r"(blockDim.x) : "memory");
    }
    griddep_launch_dependents();
    if (p.trace && threadIdx.x == 0) p.trace[(size_t) blockIdx.x * 8 + 0] = global_ns();

    if (warp == NWC) {
        // ================================================================ producer warp: lane w feeds consumer warp w
        const uint64_t pol_stream = l2_policy_evict_first();
        const int w = lane;
        int cnt = 0;
        if (w < NWC && g_end - g_begin > w) cnt = (g_end - g_begin - w + NWC - 1) / NWC;
        const int n_pair_stages = GLU ? cnt * 2 * SL : (cnt >> 1) * 2 * SL, my_total = GLU ? cnt * 2 * SL : cnt * SL;
        int issued = 0, stage = 0, pg = 0, r = 0;  // r: position inside the pair (slice = r >> 1, which = r & 1)
        uint32_t phase = 1;
        while (__any_sync(0xffffffffu, issued < my_total)) {
            bool can = issued < my_total;
            if (can) can = mbar_test_wait(smem_u32(empty + w * NS + stage), phase);
            if (can) {
                int j, slice;
                if (GLU)                         { j = pg; slice = r >> 1; }
                else if (issued < n_pair_stages) { j = 2 * pg + (r & 1); slice = r >> 1; }
                else                             { j = cnt - 1; slice = issued - n_pair_stages; }
                const int     group = g_begin + w + j * NWC;
                const MvMat & mat   = p.mat[GLU ? (r & 1) : mv_find_mat(p, group)];
                const int     row0  = (group - mat.group0) * p.rpg;
                const int     nrow  = min(p.rpg, mat.M - row0);
                const int     u0    = slice * 32;
                const int     nu    = min(32, U - u0);
                const uint32_t bar  = smem_u32(full + w * NS + stage);
                const uint32_t dst  = smem_u32(ring0 + ((size_t) w * NS + stage) * STAGE);
                const uint32_t row_bytes = (uint32_t) nu * UB;
                if (p.trace && w == 0 && issued == 0) p.trace[(size_t) blockIdx.x * 8 + 1] = global_ns();
                mbar_expect_tx(bar, row_bytes * nrow);
                const uint8_t * src = mat.w + (int64_t) row0 * mat.nb01 + (int64_t) u0 * UB;
                if (nrow > 1 && mat.nb01 == (int64_t) U * UB) {
                    bulk_g2s(dst, src, row_bytes * nrow, bar, pol_stream);
                } else {
                    for (int i = 0; i < nrow; i++) bulk_g2s(dst + i * row_bytes, src + (int64_t) i * mat.nb01, row_bytes, bar, pol_stream);
                }
                if (p.trace && w == 0 && issued == 0) p.trace[(size_t) blockIdx.x * 8 + 2] = global_ns();
                issued++;
                if (++stage == NS) { stage = 0; phase ^= 1; }
                if (++r == 2 * SL) { r = 0; pg++; }
            }
        }
        if (p.next.n_mats > 0) {  // head of the predicted next kernel's weight stream -> L2 (see MvNext)
            mv_prefetch_next(p.next, (int) blockIdx.x, (int) gridDim.x, lane);
        }
        return;
    }
    // ==================================================================== consumer warps (+ helper warps behind the producer warp:
    // they only take a share of the K1 prologue, which is issue-bound - 16 units over 11 warps are two rounds, over 16 warps one)
    const int w  = warp;
    const int NQ = NWC + p.n_helpers;            // warps that quantize
    const int qw = warp < NWC ? warp : warp - 1; // this warp's index among them
    int cnt = 0;
    if (w < NWC && g_end - g_begin > w) cnt = (g_end - g_begin - w + NWC - 1) / NWC;

    // units a quantizing warp holds at a time: a NORM launch keeps the norm weights of its units in registers as well, and two units
    // (rows up to 2 x 16 x 256 = 8192 elements) keep that free of spills - the launcher sends longer rows to the stand-alone norm
    constexpr int NB = NORM ? 2 : 4;
    float nwv[NORM ? NB : 1][8];
    if constexpr (NORM) norm_load_w<NB>(nwv, p.norm, qw, NQ, NC * U, lane);   // ahead of the stream dependency: weights never depend on it
    griddep_wait();
    if (w == 0) MV_TRACE(3);
    // row split, flagged-payload exchange (kernels.h ll_store / ll_receive): this launch's epoch is final once the stream dependency is
    constexpr bool ll = LL;                    // the launcher only sends split launches here when they use the flagged payload
    const uint32_t ep = ll ? ll_epoch(p.split) : 0u;
    if (warp > NWC && !p.fuse_quant) return;
    if (p.fuse_quant) {
        constexpr int n_tasks_per_col = 0; (void) n_tasks_per_col;
        const int n_tasks = NC * U;   // task t = (column t / U, unit t % U); the records of column c start at c * U
        for (int t0 = qw; t0 < n_tasks || (NORM && t0 == qw); t0 += NB * NQ) {   // NORM: n_tasks <= NB * NQ, every warp takes part once
            float v[NB][8];
#pragma unroll
            for (int b = 0; b < NB; b++) {
                const int t = t0 + b * NQ;
                if (t < n_tasks) {
                    const int c = NC == 1 ? 0 : t / U, u = t - c * U;
                    const float * xr = (const float *) ((const char *) p.x + (int64_t) c * p.x_cs) + (int64_t) u * UNIT_ELEMS + lane * 8;
                    if (p.x_vec) {
                        const float4 a = __ldg((const float4 *) xr);
                        const float4 bb = __ldg((const float4 *) xr + 1);
                        v[b][0] = a.x; v[b][1] = a.y; v[b][2] = a.z; v[b][3] = a.w;
                        v[b][4] = bb.x; v[b][5] = bb.y; v[b][6] = bb.z; v[b][7] = bb.w;
                    } else {
#pragma unroll
                        for (int i = 0; i < 8; i++) v[b][i] = __ldg(xr + i);
                    }
                }
            }
            if (p.trace && w == 0 && t0 == qw) {
                asm volatile("" : "+f"(v[0][0]), "+f"(v[0][7]));
                MV_TRACE(7);
            }
            if constexpr (NORM) norm_units<true, NB>(v, p.norm, t0, NQ, n_tasks, lane, (float *) (smem + 64), NQ * 32, nwv);
#pragma unroll
            for (int b = 0; b < NB; b++) {
                const int t = t0 + b * NQ;
                if (t < n_tasks) quant_unit_store<KIND>(v[b], lane, act_m + (size_t) t * ACT_MAIN, act_x + (size_t) t * ACT_AUX);
            }
        }
        bar_sync_named(1, NQ * 32);
        if (warp > NWC) return;   // helper warps are done
    } else {
        if (w == 0 && lane == 0) {
            const uint64_t pol_keep = l2_policy_evict_last();
            const uint32_t bytes_m = (uint32_t) NC * U * ACT_MAIN, bytes_x = (uint32_t) NC * U * ACT_AUX;
            mbar_expect_tx(smem_u32(act_bar), bytes_m + bytes_x);
            bulk_g2s(smem_u32(act_m), p.act_main, bytes_m, smem_u32(act_bar), pol_keep);
            bulk_g2s(smem_u32(act_x), p.act_aux, bytes_x, smem_u32(act_bar), pol_keep);
        }
        mbar_wait(smem_u32(act_bar), 0);
    }
    if (w == 0) MV_TRACE(4);

    const int L   = p.L;
    const int sub = lane & (L - 1);
    const int rr  = lane / L;
    const uint32_t ring  = smem_u32(ring0 + (size_t) w * NS * STAGE);
    const uint32_t fullw = smem_u32(full + w * NS), emptyw = smem_u32(empty + w * NS);
    const uint32_t am0 = smem_u32(act_m) + (uint32_t) sub * ACT_MAIN, ax0 = smem_u32(act_x) + (uint32_t) sub * ACT_AUX;
    const bool     lane_on = rr < p.rpg;
    const uint32_t cs_m = (uint32_t) U * ACT_MAIN, cs_x = (uint32_t) U * ACT_AUX;

    auto reduce_store = [&](float (&a)[NC], int group) {  // sum over the L lanes of a row, one store per row and column
        const MvMat & mat = p.mat[mv_find_mat(p, group)];
        const int     row = (group - mat.group0) * p.rpg + rr;
#pragma unroll
        for (int c = 0; c < NC; c++) {
            float v = a[c];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float t = __shfl_xor_sync(0xffffffffu, v, o);
                if (o < L) v += t;
            }
            if constexpr (ROPE) {   // rows 2m and 2m + 1 of a head sit in neighbouring row slots of this group: lanes rr * L and (rr ^ 1) * L
                const float other = __shfl_xor_sync(0xffffffffu, v, L);
                if (sub == 0 && lane_on && row < mat.M) {
                    const MvRope & r = p.rope;
                    const int m = (int) (&mat - p.mat);
                    if (m < 2) {
                        float cs, sn;
                        rope_cs(r.rp, r.ff, (float) r.pos[0], (row % r.head_dim) >> 1, cs, sn);
                        const float y = (row & 1) ? rope_hi(other, v, cs, sn) : rope_lo(v, other, cs, sn);
                        if (m == 0) {
                            r.q_dst[row] = y;
                        } else {
                            r.k_dst[row] = y;
                            const int64_t slot = load_idx(r.k_idx, r.k_i64);
                            if (slot >= 0 && slot < r.kc_rows) ((__half *) (r.kc + slot * r.kc_nb1))[row] = __float2half_rn(y);
                        }
                    } else {
                        const int64_t slot = load_idx(r.v_idx, r.v_i64);
                        if (slot >= 0 && slot < r.vc_rows) ((__half *) (r.vc + slot * r.vc_nb1))[row] = __float2half_rn(v);
                    }
                }
                continue;
            }
            if (sub == 0 && lane_on && row < mat.M) {
                float * a = mat.dst + (int64_t) c * mat.dst_cs + row;
                if constexpr (BIAS) { if (mat.bias) v += mat.bias[row]; }
                if (ll) ll_store(p.split, ep, a, v);
                else *a = v;
            }
        }
    };

    int stage = 0;
    uint32_t phase = 0;
    const int n_pairs = GLU ? cnt : cnt >> 1;
    for (int pg = 0; pg < n_pairs; pg++) {
        float accA[NC], accB[NC];
#pragma unroll
        for (int c = 0; c < NC; c++) { accA[c] = 0.0f; accB[c] = 0.0f; }
        for (int s = 0; s < SL; s++) {
            mbar_wait(fullw + 8 * stage, phase);
            mbar_wait(fullw + 8 * stage + 8, phase);
            if (w == 0 && pg == 0 && s == 0) MV_TRACE(5);
            const int nu = min(32, U - s * 32);
            if (lane_on && sub < nu) {
                const uint32_t wp = ring + (uint32_t) stage * STAGE + (uint32_t) (rr * nu + sub) * UB;
                unit_dot2<TYPE, NC>(wp, wp + STAGE, am0 + (uint32_t) s * 32 * ACT_MAIN, ax0 + (uint32_t) s * 32 * ACT_AUX, cs_m, cs_x, accA, accB);
            }
            __syncwarp();
            if (lane == 0) { mbar_arrive(emptyw + 8 * stage); mbar_arrive(emptyw + 8 * stage + 8); }
            stage += 2;
            if (stage == NS) { stage = 0; phase ^= 1; }
        }
        if constexpr (GLU) {   // one row group: accA = its gate rows, accB = its up rows
            const int row = (g_begin + w + pg * NWC) * p.rpg + rr;
            float vg = accA[0], vu = accB[0];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float tg = __shfl_xor_sync(0xffffffffu, vg, o), tu = __shfl_xor_sync(0xffffffffu, vu, o);
                if (o < L) { vg += tg; vu += tu; }
            }
            if (sub == 0 && lane_on && row < p.mat[0].M) {
                if (p.mat[0].bias) vg += p.mat[0].bias[row];
                if (p.mat[1].bias) vu += p.mat[1].bias[row];
                p.mat[0].dst[row] = glu_apply(p.epi.glu - 1, vg, vu, p.epi.alpha, p.epi.limit);
            }
            continue;
        }
        const int groupA = g_begin + w + 2 * pg * NWC;
        reduce_store(accA, groupA);
        reduce_store(accB, groupA + NWC);
    }
    if (!GLU && (cnt & 1)) {  // the odd group at the end of the warp's list: one stage per slice
        float acc[NC];
#pragma unroll
        for (int c = 0; c < NC; c++) acc[c] = 0.0f;
        for (int s = 0; s < SL; s++) {
            mbar_wait(fullw + 8 * stage, phase);
            if (w == 0 && n_pairs == 0 && s == 0) MV_TRACE(5);
            const int nu = min(32, U - s * 32);
            if (lane_on && sub < nu) {
                const uint32_t wp = ring + (uint32_t) stage * STAGE + (uint32_t) (rr * nu + sub) * UB;
                unit_dot<TYPE, NC>(wp, am0 + (uint32_t) s * 32 * ACT_MAIN, ax0 + (uint32_t) s * 32 * ACT_AUX, cs_m, cs_x, acc);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(emptyw + 8 * stage);
            if (++stage == NS) { stage = 0; phase ^= 1; }
        }
        reduce_store(acc, g_begin + w + (cnt - 1) * NWC);
    }
    if (w == 0) MV_TRACE(6);
    if (ll) {  // this CTA's rows are on their way to every rank; collect this CTA's share of the peers' rows
        ll_receive(p.split, ep, p.geom, p.mat, p.n_mats, NC, (int) blockIdx.x * NWC * 32 + w * 32 + lane, (int) gridDim.x * NWC * 32);
        bar_sync_named(2, NWC * 32);
        if (w == 0 && lane == 0) ll_finish(p.split, ep, gridDim.x);
    }
}

template <int TYPE, int NC, bool LL, bool BIAS = false, bool NORM = false, bool GLU = false, bool ROPE = false>
static cudaError_t mv_lean_launch_t(const MvParams & p_in, int n_ctas, size_t smem, cudaStream_t st) {
    MvParams p = p_in;
    p.n_helpers = 0;
    if (p.fuse_quant && !(p.dbg & 4)) {   // as many quantizing warps as src1 has units, up to the 16 + 1 warps the kernel is compiled for
        const int room = mv_lean_max_cons(NC) - p.n_cons, want = NC * p.U - p.n_cons;
        p.n_helpers = want < 0 ? 0 : (want < room ? want : room);
    }
    static std::atomic<bool> attr_done[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_done[dev]) {
        cudaError_t e = cudaFuncSetAttribute(mv_lean_kernel<TYPE, NC, LL, BIAS, NORM, GLU, ROPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(mv_lean_kernel<TYPE, NC, LL, BIAS, NORM, GLU, ROPE>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        if (e != cudaSuccess) return e;
        attr_done[dev] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = dim3(n_ctas);
    cfg.blockDim         = dim3(32 * (p.n_cons + 1 + p.n_helpers));
    cfg.dynamicSmemBytes = smem;
    cfg.stream           = st;
    cudaLaunchAttribute attr[1];
    int na = 0;
    if (p.pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    cfg.attrs    = attr;
    cfg.numAttrs = na;
    return cudaLaunchKernelEx(&cfg, mv_lean_kernel<TYPE, NC, LL, BIAS, NORM, GLU, ROPE>, p);
}

// the launch can take the lean kernel: one column, plain MUL_MAT, no exchange, no epilogue, an even ring
template <int TYPE>
static cudaError_t mv_lean_launch_n(int nc, const MvParams & p, int n_ctas, size_t smem, cudaStream_t st) {
    if (p.epi.glu) {   // the K7 pair (biases, if any, are looked at inside)
        if (nc != 1 || p.split.n_ranks > 1 || p.n_mats != 2) return cudaErrorInvalidValue;
        return p.norm.on ? mv_lean_launch_t<TYPE, 1, false, false, true, true>(p, n_ctas, smem, st) : mv_lean_launch_t<TYPE, 1, false, false, false, true>(p, n_ctas, smem, st);
    }
    for (int i = 0; i < p.n_mats; i++) {
        if (p.mat[i].bias) return nc == 1 && p.split.n_ranks <= 1 && !p.norm.on ? mv_lean_launch_t<TYPE, 1, false, true>(p, n_ctas, smem, st) : cudaErrorInvalidValue;
    }
    if (p.rope.on) {   // q / k / v of one token: rows leave rotated / as cache rows
        if (nc != 1 || p.split.n_ranks > 1 || p.n_mats != 3 || (p.rpg & 1)) return cudaErrorInvalidValue;
        return p.norm.on ? mv_lean_launch_t<TYPE, 1, false, false, true, false, true>(p, n_ctas, smem, st) : mv_lean_launch_t<TYPE, 1, false, false, false, false, true>(p, n_ctas, smem, st);
    }
    if (p.norm.on) return nc == 1 && p.split.n_ranks <= 1 && p.fuse_quant ? mv_lean_launch_t<TYPE, 1, false, false, true>(p, n_ctas, smem, st) : cudaErrorInvalidValue;
    if (p.split.n_ranks > 1) {
        switch (nc) {
            case 1: return mv_lean_launch_t<TYPE, 1, true>(p, n_ctas, smem, st);
            case 2: return mv_lean_launch_t<TYPE, 2, true>(p, n_ctas, smem, st);
            case 3: return mv_lean_launch_t<TYPE, 3, true>(p, n_ctas, smem, st);
            case 4: return mv_lean_launch_t<TYPE, 4, true>(p, n_ctas, smem, st);
        }
        return cudaErrorInvalidValue;
    }
    switch (nc) {
        case 1: return mv_lean_launch_t<TYPE, 1, false>(p, n_ctas, smem, st);
        case 2: return mv_lean_launch_t<TYPE, 2, false>(p, n_ctas, smem, st);
        case 3: return mv_lean_launch_t<TYPE, 3, false>(p, n_ctas, smem, st);
        case 4: return mv_lean_launch_t<TYPE, 4, false>(p, n_ctas, smem, st);
    }
    return cudaErrorInvalidValue;
}

bool mv_lean_type(int type) { return type == B200_TYPE_Q4_K || type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0; }  // formats with a two-unit dot product

static bool mv_lean_ok(int type, int ncols, const MvParams & p, int n_pairs);
bool mv_lean_takes(int type, int ncols, const MvParams & p, int n_pairs) { return mv_lean_ok(type, ncols, p, n_pairs); }

static bool mv_lean_ok(int type, int ncols, const MvParams & p, int n_pairs) {
    if (!mv_lean_type(type) || ncols < 1 || ncols > MV_LEAN_MAX_COLS || n_pairs != 1 || p.ids.ids || (p.split.n_ranks > 1 && !p.split.ll) || (p.n_stages & 1) || (p.dbg & 1)) return false;
    if (p.epi.glu && (ncols != 1 || p.split.n_ranks > 1 || p.n_mats != 2 || (p.dbg & 32))) return false;   // dbg 32: the K7 pair stays on mv_fast_kernel
    for (int i = 0; i < p.n_mats; i++) if (p.mat[i].bias && (ncols != 1 || p.split.n_ranks > 1)) return false;
    return true;
}

template <int TYPE, int NC, int WIDE, bool EPI = false>
static cudaError_t mv_fast_launch_w(const MvParams & p, int n_ctas, int n_pairs, size_t smem, cudaStream_t st) {
    static std::atomic<bool> attr_done[64];  // one attribute call per instantiation and device
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_done[dev]) {
        cudaError_t e = cudaFuncSetAttribute(mv_fast_kernel<TYPE, NC, WIDE, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        // always the largest shared-memory carve-out: a CTA of the NEXT kernel can only become co-resident if the SM does not
        // have to be reconfigured between the two launches
        e = cudaFuncSetAttribute(mv_fast_kernel<TYPE, NC, WIDE, EPI>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        if (e != cudaSuccess) return e;
        attr_done[dev] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = dim3(n_ctas, n_pairs);
    cfg.blockDim         = dim3(32 * (p.n_cons + 1));
    cfg.dynamicSmemBytes = smem;
    cfg.stream           = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs    = attr;
    cfg.numAttrs = p.pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, mv_fast_kernel<TYPE, NC, WIDE, EPI>, p);
}

template <int TYPE, int NC>
static cudaError_t mv_fast_launch_t(const MvParams & p, int n_ctas, int n_pairs, size_t smem, cudaStream_t st) {
    if (p.rope.on) return cudaErrorInvalidValue;   // MvRope: the lean kernel only (api.cu decides)
    bool epi = p.epi.glu != 0;
    for (int i = 0; i < p.n_mats; i++) epi = epi || p.mat[i].bias != nullptr;
    if (p.norm.on && !epi) return cudaErrorInvalidValue;   // MvNorm: the lean kernel's NORM instantiation or an EPI one (api.cu decides)
    if (epi) {  // K7: one column; the wide CTA shape (or the 8-warp one for rows that leave less shared memory)
        if constexpr (NC == 1) {
            if (p.n_cons > 8) return mv_fast_launch_w<TYPE, 1, 1, true>(p, n_ctas, n_pairs, smem, st);
            return mv_fast_launch_w<TYPE, 1, 0, true>(p, n_ctas, n_pairs, smem, st);
        }
        return cudaErrorInvalidValue;
    }
    if (p.n_cons > 8) {
        if constexpr (NC <= MV_WIDE_MAX_COLS) {
            if (p.n_cons <= 11) return mv_fast_launch_w<TYPE, NC, 2>(p, n_ctas, n_pairs, smem, st);
            return mv_fast_launch_w<TYPE, NC, 1>(p, n_ctas, n_pairs, smem, st);
        } else return cudaErrorInvalidValue;
    }
    return mv_fast_launch_w<TYPE, NC, 0>(p, n_ctas, n_pairs, smem, st);
}

template <int TYPE>
static cudaError_t mv_fast_launch_n(int nc, const MvParams & p, int n_ctas, int n_pairs, size_t smem, cudaStream_t st) {
    switch (nc) {
        case 1: return mv_fast_launch_t<TYPE, 1>(p, n_ctas, n_pairs, smem, st);
        case 2: return mv_fast_launch_t<TYPE, 2>(p, n_ctas, n_pairs, smem, st);
        case 3: return mv_fast_launch_t<TYPE, 3>(p, n_ctas, n_pairs, smem, st);
        case 4: return mv_fast_launch_t<TYPE, 4>(p, n_ctas, n_pairs, smem, st);
        case 5: return mv_fast_launch_t<TYPE, 5>(p, n_ctas, n_pairs, smem, st);
        case 6: return mv_fast_launch_t<TYPE, 6>(p, n_ctas, n_pairs, smem, st);
        case 7: return mv_fast_launch_t<TYPE, 7>(p, n_ctas, n_pairs, smem, st);
        case 8: return mv_fast_launch_t<TYPE, 8>(p, n_ctas, n_pairs, smem, st);
    }
    return cudaErrorInvalidValue;
}

cudaError_t mv_fast_launch(int type, int ncols, const MvParams & p, int n_ctas, int n_pairs, size_t smem, cudaStream_t st) {
    if (mv_lean_ok(type, ncols, p, n_pairs)) {
        switch (type) {
            case B200_TYPE_Q4_0: return mv_lean_launch_n<B200_TYPE_Q4_0>(ncols, p, n_ctas, smem, st);
            case B200_TYPE_Q8_0: return mv_lean_launch_n<B200_TYPE_Q8_0>(ncols, p, n_ctas, smem, st);
            case B200_TYPE_Q4_K: return mv_lean_launch_n<B200_TYPE_Q4_K>(ncols, p, n_ctas, smem, st);
        }
    }
    switch (type) {
        case B200_TYPE_Q4_0: return mv_fast_launch_n<B200_TYPE_Q4_0>(ncols, p, n_ctas, n_pairs, smem, st);
        case B200_TYPE_Q8_0: return mv_fast_launch_n<B200_TYPE_Q8_0>(ncols, p, n_ctas, n_pairs, smem, st);
        case B200_TYPE_Q4_K: return mv_fast_launch_n<B200_TYPE_Q4_K>(ncols, p, n_ctas, n_pairs, smem, st);
        case B200_TYPE_Q5_K: return mv_fast_launch_n<B200_TYPE_Q5_K>(ncols, p, n_ctas, n_pairs, smem, st);
        case B200_TYPE_Q6_K: return mv_fast_launch_n<B200_TYPE_Q6_K>(ncols, p, n_ctas, n_pairs, smem, st);
    }
    return cudaErrorInvalidValue;
}

// ------------------------------------------------------------------------------------------------
// generic kernel: one warp per weight row, one lane per block, 16-bit loads only (blocks are 2-byte aligned),
// activations in the reference's own q8_0 / q8_K block layout. Used for K % 256 != 0, unaligned views, and as an
// independent on-device cross-check of the fast kernel. Performance is not a goal here.
// ------------------------------------------------------------------------------------------------

__device__ __forceinline__ uint32_t ld16x2(const uint8_t * p) {  // 4 bytes from a 2-byte aligned address
    const uint16_t * h = (const uint16_t *) p;
    return (uint32_t) h[0] | ((uint32_t) h[1] << 16);
}

template <int TYPE>
__global__ void __launch_bounds__(128) mv_simple_kernel(const MvSimpleParams p) {
    const int lane = threadIdx.x & 31;
    const int row  = blockIdx.x * 4 + (threadIdx.x >> 5);
    int     col = blockIdx.y, act_row = blockIdx.y;
    int64_t w_off = 0, dst_off = 0;
    if (row >= p.M) return;
    if (p.ids.ids) {
        const int t = col / p.ids.n_used, i = col - t * p.ids.n_used;
        const int e = p.ids.ids[(int64_t) t * p.ids.ids_rs + i];
        if (e < 0 || e >= p.ids.n_as) return;
        w_off   = (int64_t) e * p.ids.nb02;
        dst_off = (int64_t) t * p.ids.dst_ts + (int64_t) i * p.ids.dst_is;
        act_row = t * p.ids.ne11 + i % p.ids.ne11;
        col     = 0;
    }
    const uint8_t * wr = p.w + w_off + (int64_t) row * p.nb01;
    const uint8_t * ar = p.act + (int64_t) act_row * p.act_rs;
    float acc = 0.0f;
    if (TYPE == B200_TYPE_Q6_K) {
        for (int b = lane; b < p.K / QK_K; b += 32) {
            const uint8_t * wb = wr + (int64_t) b * SZ_Q6_K;   // 2-byte aligned: byte and 16-bit loads only
            const uint8_t * ab = ar + (int64_t) b * SZ_Q8_K;
            const float     dy = *(const float *) ab;
            const int8_t *  q8 = (const int8_t *) (ab + 4);
            const int8_t *  sc = (const int8_t *) (wb + 192);
            const float     d  = half_bits_to_float(*(const uint16_t *) (wb + 208));
            int isum = 0;
            for (int h = 0; h < 2; h++) {
                const uint8_t * ql = wb + 64 * h, * qh = wb + 128 + 32 * h;
                for (int l = 0; l < 32; l++) {
                    const int is = 8 * h + l / 16;
                    const int q1 = (int) ((ql[l] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                    const int q2 = (int) ((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                    const int q3 = (int) ((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                    const int q4 = (int) ((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                    isum += sc[is] * q1 * q8[128 * h + l] + sc[is + 2] * q2 * q8[128 * h + 32 + l] + sc[is + 4] * q3 * q8[128 * h + 64 + l] +
                            sc[is + 6] * q4 * q8[128 * h + 96 + l];
                }
            }
            acc = fmaf(d * dy, (float) isum, acc);
        }
    } else if (TYPE == B200_TYPE_Q4_K || TYPE == B200_TYPE_Q5_K) {
        constexpr bool Q5 = TYPE == B200_TYPE_Q5_K;  // 32 bytes of fifth bits between the header and the nibbles
        for (int b = lane; b < p.K / QK_K; b += 32) {
            const uint8_t * wb = wr + (int64_t) b * (Q5 ? SZ_Q5_K : SZ_Q4_K);
            const uint8_t * ab = ar + (int64_t) b * SZ_Q8_K;
            const uint32_t  h0 = ld16x2(wb), u0 = ld16x2(wb + 4), u1 = ld16x2(wb + 8), u2 = ld16x2(wb + 12);
            const float d = half_bits_to_float((uint16_t) (h0 & 0xffff)), dmin = half_bits_to_float((uint16_t) (h0 >> 16));
            const uint32_t sc4[2] = { u0 & 0x3f3f3f3fu, (u2 & 0x0f0f0f0fu) | ((u0 >> 2) & 0x30303030u) };
            const uint32_t mn4[2] = { u1 & 0x3f3f3f3fu, ((u2 >> 4) & 0x0f0f0f0fu) | ((u1 >> 2) & 0x30303030u) };
            const float    dy     = *(const float *) ab;
            const int8_t *  q8    = (const int8_t *) (ab + 4);
            const int16_t * bs    = (const int16_t *) (ab + 4 + QK_K);
            int isum = 0, imin = 0;
            for (int s = 0; s < 8; s++) {
                const int sc = (sc4[s >> 2] >> (8 * (s & 3))) & 0xff, mn = (mn4[s >> 2] >> (8 * (s & 3))) & 0xff;
                imin += mn * (bs[2 * s] + bs[2 * s + 1]);
                const uint8_t * qq = wb + (Q5 ? 48 : 16) + 32 * (s >> 1);
                int dot = 0;
                for (int l = 0; l < 32; l += 2) {
                    const uint16_t two = *(const uint16_t *) (qq + l);
                    int w0 = (s & 1) ? (two >> 4) & 0xf : two & 0xf;
                    int w1 = (s & 1) ? (two >> 12) & 0xf : (two >> 8) & 0xf;
                    if (Q5) {
                        const uint16_t hb = *(const uint16_t *) (wb + 16 + l);
                        w0 += ((hb >> s) & 1) << 4;
                        w1 += ((hb >> (8 + s)) & 1) << 4;
                    }
                    dot += w0 * q8[32 * s + l] + w1 * q8[32 * s + l + 1];
                }
                isum += sc * dot;
            }
            acc = fmaf(d * dy, (float) isum, acc);
            acc = fmaf(-(dmin * dy), (float) imin, acc);
        }
    } else {
        constexpr int WB = TYPE == B200_TYPE_Q4_0 ? SZ_Q4_0 : SZ_Q8_0;
        for (int b = lane; b < p.K / QK; b += 32) {
            const uint8_t * wb = wr + (int64_t) b * WB;
            const uint8_t * ab = ar + (int64_t) b * SZ_Q8_0;
            const float     dw = half_bits_to_float(*(const uint16_t *) wb);
            const float     dy = half_bits_to_float(*(const uint16_t *) ab);
            const int8_t *  q8 = (const int8_t *) (ab + 2);
            int sumi = 0;
            if (TYPE == B200_TYPE_Q4_0) {
                for (int j = 0; j < 16; j += 2) {
                    const uint16_t two = *(const uint16_t *) (wb + 2 + j);
                    sumi += ((two & 0xf) - 8) * q8[j] + (((two >> 4) & 0xf) - 8) * q8[j + 16];
                    sumi += (((two >> 8) & 0xf) - 8) * q8[j + 1] + (((two >> 12) & 0xf) - 8) * q8[j + 17];
                }
                acc = fmaf((float) sumi * dw, dy, acc);
            } else {
                for (int j = 0; j < 32; j += 2) {
                    const uint16_t two = *(const uint16_t *) (wb + 2 + j);
                    sumi += (int) (int8_t) (two & 0xff) * q8[j] + (int) (int8_t) (two >> 8) * q8[j + 1];
                }
                acc = fmaf((float) sumi, dw * dy, acc);
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) p.dst[dst_off + (int64_t) col * p.dst_cs + row] = acc;
}

// ------------------------------------------------------------------------------------------------
// exact kernel (b200_ctx_set_exact / GGML_B200_EXACT=1): one thread per dst element, walking the row the way the generic scalar
// ggml_vec_dot_q4_0_q8_0 / q8_0_q8_0 / q4_K_q8_K / q5_K_q8_K / q6_K_q8_K do (ggml-cpu/quants.c:115-150, 305-333, 550-623, 625-703,
// 705-758): the same integer sums AND the same order and rounding of every fp32 operation (one accumulator walked block by block
// for q4_0 / q8_0; eight interleaved lane accumulators plus the min accumulator for the K-quants; __fmul_rn / __fadd_rn so that
// nothing is contracted into an FMA - the x86-64 baseline build of ggml-cpu has none). Results are BIT-IDENTICAL to ggml-cpu's
// generic build, which is what makes a free-running greedy token stream comparable at all: activation quantization turns any
// fp32 difference of an earlier layer into rounding flips, so two correct implementations with different summation orders
// (the reference's own x64 and AVX2 builds included) drift apart within a few dozen tokens. Slow by design; never the default.
// ------------------------------------------------------------------------------------------------

__device__ __forceinline__ float ld_f32_unaligned2(const uint8_t * p) { return __uint_as_float(ld16x2(p)); }

template <int TYPE>
__global__ void __launch_bounds__(128) mv_exact_kernel(const MvSimpleParams p) {
    const int row = blockIdx.x * 128 + threadIdx.x;
    int     col = blockIdx.y, act_row = blockIdx.y;
    int64_t w_off = 0, dst_off = 0;
    if (row >= p.M) return;
    if (p.ids.ids) {
        const int t = col / p.ids.n_used, i = col - t * p.ids.n_used;
        const int e = p.ids.ids[(int64_t) t * p.ids.ids_rs + i];
        if (e < 0 || e >= p.ids.n_as) return;
        w_off   = (int64_t) e * p.ids.nb02;
        dst_off = (int64_t) t * p.ids.dst_ts + (int64_t) i * p.ids.dst_is;
        act_row = t * p.ids.ne11 + i % p.ids.ne11;
        col     = 0;
    }
    const uint8_t * wr = p.w + w_off + (int64_t) row * p.nb01;
    const uint8_t * ar = p.act + (int64_t) act_row * p.act_rs;
    float sumf = 0.0f;
    if (TYPE == B200_TYPE_Q4_0 || TYPE == B200_TYPE_Q8_0) {
        constexpr int WB = TYPE == B200_TYPE_Q4_0 ? SZ_Q4_0 : SZ_Q8_0;
        for (int b = 0; b < p.K / QK; b++) {
            const uint8_t * wb = wr + (int64_t) b * WB;
            const uint8_t * ab = ar + (int64_t) b * SZ_Q8_0;
            const float     dx = half_bits_to_float(*(const uint16_t *) wb);
            const float     dy = half_bits_to_float(*(const uint16_t *) ab);
            const int8_t *  q8 = (const int8_t *) (ab + 2);
            int sumi = 0;
            if (TYPE == B200_TYPE_Q4_0) {
                for (int j = 0; j < 16; j++) {
                    const int v = wb[2 + j];
                    sumi += ((v & 0xf) - 8) * q8[j] + ((v >> 4) - 8) * q8[j + 16];
                }
                sumf = __fadd_rn(sumf, __fmul_rn(__fmul_rn((float) sumi, dx), dy));  // sumf += sumi*dx*dy, left to right (quants.c:146)
            } else {
                const int8_t * qw = (const int8_t *) (wb + 2);
                for (int j = 0; j < 32; j++) sumi += qw[j] * q8[j];
                sumf = __fadd_rn(sumf, __fmul_rn((float) sumi, __fmul_rn(dx, dy)));  // sumf += sumi*(dx*dy) (quants.c:329)
            }
        }
    } else {
        float sums[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
        for (int b = 0; b < p.K / QK_K; b++) {
            const uint8_t * ab = ar + (int64_t) b * SZ_Q8_K;
            const float     yd = *(const float *) ab;
            const int8_t *  q8 = (const int8_t *) (ab + 4);
            const int16_t * bs = (const int16_t *) (ab + 4 + QK_K);
            int lane[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
            if (TYPE == B200_TYPE_Q6_K) {
                const uint8_t * wb = wr + (int64_t) b * SZ_Q6_K;
                const int8_t *  sc = (const int8_t *) (wb + 192);
                const float     d  = __fmul_rn(half_bits_to_float(*(const uint16_t *) (wb + 208)), yd);
                for (int h = 0; h < 2; h++) {
                    const uint8_t * ql = wb + 64 * h, * qh = wb + 128 + 32 * h;
                    for (int l = 0; l < 32; l++) {
                        const int q1 = (int) ((ql[l] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                        const int q2 = (int) ((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                        const int q3 = (int) ((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                        const int q4 = (int) ((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                        const int i0 = 128 * h + l;
                        lane[l & 7] += sc[(i0) / 16] * (q8[i0] * q1) + sc[(i0 + 32) / 16] * (q8[i0 + 32] * q2) +
                                       sc[(i0 + 64) / 16] * (q8[i0 + 64] * q3) + sc[(i0 + 96) / 16] * (q8[i0 + 96] * q4);
                    }
                }
#pragma unroll
                for (int l = 0; l < 8; l++) sums[l] = __fadd_rn(sums[l], __fmul_rn(d, (float) lane[l]));
            } else {
                constexpr bool Q5 = TYPE == B200_TYPE_Q5_K;
                const uint8_t * wb = wr + (int64_t) b * (Q5 ? SZ_Q5_K : SZ_Q4_K);
                const uint32_t  h0 = ld16x2(wb), u0 = ld16x2(wb + 4), u1 = ld16x2(wb + 8), u2 = ld16x2(wb + 12);
                const uint32_t sc4[2] = { u0 & 0x3f3f3f3fu, (u2 & 0x0f0f0f0fu) | ((u0 >> 2) & 0x30303030u) };
                const uint32_t mn4[2] = { u1 & 0x3f3f3f3fu, ((u2 >> 4) & 0x0f0f0f0fu) | ((u1 >> 2) & 0x30303030u) };
                int summin = 0;
                for (int s = 0; s < 8; s++) {
                    const int sc = (sc4[s >> 2] >> (8 * (s & 3))) & 0xff, mn = (mn4[s >> 2] >> (8 * (s & 3))) & 0xff;
                    summin += (bs[2 * s] + bs[2 * s + 1]) * mn;
                    const uint8_t * qq = wb + (Q5 ? 48 : 16) + 32 * (s >> 1);
                    for (int l = 0; l < 32; l++) {
                        int w = (s & 1) ? (qq[l] >> 4) : (qq[l] & 0xf);
                        if (Q5) w += ((wb[16 + l] >> s) & 1) << 4;
                        lane[l & 7] += sc * (q8[32 * s + l] * w);
                    }
                }
                const float d    = __fmul_rn(half_bits_to_float((uint16_t) (h0 & 0xffff)), yd);
                const float dmin = __fmul_rn(half_bits_to_float((uint16_t) (h0 >> 16)), yd);
#pragma unroll
                for (int l = 0; l < 8; l++) sums[l] = __fadd_rn(sums[l], __fmul_rn(d, (float) lane[l]));
                sumf = __fsub_rn(sumf, __fmul_rn(dmin, (float) summin));
            }
        }
#pragma unroll
        for (int l = 0; l < 8; l++) sumf = __fadd_rn(sumf, sums[l]);
    }
    p.dst[dst_off + (int64_t) col * p.dst_cs + row] = sumf;
}

template <int TYPE>
static void mv_exact_launch_t(const MvSimpleParams & p, cudaStream_t st) {
    dim3 grid((unsigned) ((p.M + 127) / 128), (unsigned) p.ncols);
    mv_exact_kernel<TYPE><<<grid, 128, 0, st>>>(p);
}

cudaError_t mv_simple_launch(int type, const MvSimpleParams & p, cudaStream_t st) {
    if (p.M == 0 || p.ncols == 0) return cudaSuccess;
    if (p.exact) {
        switch (type) {
            case B200_TYPE_Q4_0: mv_exact_launch_t<B200_TYPE_Q4_0>(p, st); break;
            case B200_TYPE_Q8_0: mv_exact_launch_t<B200_TYPE_Q8_0>(p, st); break;
            case B200_TYPE_Q4_K: mv_exact_launch_t<B200_TYPE_Q4_K>(p, st); break;
            case B200_TYPE_Q5_K: mv_exact_launch_t<B200_TYPE_Q5_K>(p, st); break;
            case B200_TYPE_Q6_K: mv_exact_launch_t<B200_TYPE_Q6_K>(p, st); break;
            default: return cudaErrorInvalidValue;
        }
        return cudaGetLastError();
    }
    dim3 grid((unsigned) ((p.M + 3) / 4), (unsigned) p.ncols);
    switch (type) {
        case B200_TYPE_Q4_0: mv_simple_kernel<B200_TYPE_Q4_0><<<grid, 128, 0, st>>>(p); break;
        case B200_TYPE_Q8_0: mv_simple_kernel<B200_TYPE_Q8_0><<<grid, 128, 0, st>>>(p); break;
        case B200_TYPE_Q4_K: mv_simple_kernel<B200_TYPE_Q4_K><<<grid, 128, 0, st>>>(p); break;
        case B200_TYPE_Q5_K: mv_simple_kernel<B200_TYPE_Q5_K><<<grid, 128, 0, st>>>(p); break;
        case B200_TYPE_Q6_K: mv_simple_kernel<B200_TYPE_Q6_K><<<grid, 128, 0, st>>>(p); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

} // namespace b200

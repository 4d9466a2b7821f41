// mmq.cu — K2/K4: tiled int8 GEMM for prefill / batched decode (more than 8 activation columns) on the sm_100a
// 5th-generation tensor cores: tcgen05.mma kind::i8 (SASS UTCIMMA) with int32 accumulators in TMEM, operands staged by
// TMA (cp.async.bulk.tensor, SASS UTMALDG) into 128-byte-swizzled shared memory, per-32-block scales applied by the
// epilogue warps while they drain TMEM (tcgen05.ld, SASS LDTM).
//
// Replaces ggml_cuda_mul_mat_q / mul_mat_q + load_tiles_* + vec_dot_*_mma + mmq_write_back_mma + quantize_mmq_q8_1
// (/root/reference/ggml/src/ggml-cuda/mmq.cu:70-236, mmq.cuh:3287-3536, 277-336, 637-697, 1840-1947, 796-931, 965-1091,
// 3006-3053, quantize.cu:50-147). Not a port: the reference keeps 16x8 mma.sync fragments in registers and re-loads
// tiles synchronously (4 __syncthreads per 256-K step, mmq.cuh:3241-3275); here one thread issues 128xBNx32 MMAs, the
// copy engine feeds a 4-stage ring, and 16 epilogue warps do nothing but rescale (double-buffered TMEM drain, packed fp32 math).
//
// Arithmetic is the CPU reference's (ggml-cpu vec_dot_q4_0_q8_0 / q8_0_q8_0 / q4_K_q8_K): activations are quantized
// exactly like quantize_row_q8_0_ref / q8_K_ref, every 32-element block's integer dot product is formed exactly
// (one UTCIMMA with K = 32 and accumulate = 0 per block, because the fp16 block scales differ per block and per row),
// and scaled in fp32:   dst[n][m] = sum_kb  sW[kb][m] * sA[kb][n] * C_kb[m][n]   ( - mW[kb][m] * aS[kb][n]  for q4_K )
//   q4_0 / q8_0 : sW = d_w            sA = fp16(d_a)                     C = sum (q_w - 8 | q_w) * q_a
//   q4_K        : sW = d * sc_j       sA = d_a (per 256)   mW = dmin * m_j   aS = d_a * sum_32(q_a)   C = sum q_w * q_a
//
// Operands:
//   weights     : int8 [M][K] K-major + float scales [K/32][M]   ("expanded" form of the ggml blocks, built by
//                 expand_weights_kernel; cached per weight tensor when the owner of the memory allows it)
//   activations : int8 [N][K] K-major + float scales [K/32][N]   (quantize_tiles_kernel, bit-exact values)
// Tiling: CTA tile 128 (M) x 128 (N); K in stages of 128 (4 blocks of 32); TMEM: 4 slots of 128 columns, one slot per
// 32-block, recycled as soon as the epilogue has pulled it into registers.
// Warp roles (640 threads): warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer, warps 4..19 epilogue
// (warp % 4 = TMEM lane quadrant = 32 weight rows, (warp - 4) / 4 = quarter of the tile's 128 activation columns).

#include "internal.h"
#include "quant_core.cuh"
#include "tc.cuh"

#include <cuda.h>

#include <map>
#include <mutex>

using namespace b200;

#define fail b200_fail

namespace {

constexpr int BM = 128, BN = 128, BKB = 4;        // CTA tile; 32-blocks per stage
constexpr int STAGES = 4, SLOTS = 4;
constexpr int EPI_WARPS = 16, FIRST_EPI_WARP = 4;  // 4 warps per TMEM lane quadrant, 32 accumulator columns per thread: the
                                                   // epilogue is latency bound (TMEM round trips, LDS), so warps beat registers
constexpr int MMQ_THREADS = 32 * (FIRST_EPI_WARP + EPI_WARPS);
constexpr int A_BYTES = BM * 128, B_BYTES = BN * 128, S_BYTES = BKB * 128 * 4;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES + 4 * S_BYTES;  // A, B, sW, sA, mW, aS
constexpr int MMQ_SMEM = 1024 + STAGES * STAGE_BYTES + 256;

// ------------------------------------------------------------------------------------------------ weight expansion

__device__ __forceinline__ uint32_t ld16x2(const uint8_t * p) {  // 4 bytes from a 2-byte aligned address
    const uint16_t * h = (const uint16_t *) p;
    return (uint32_t) h[0] | ((uint32_t) h[1] << 16);
}
// 4 nibbles (low nibble of each byte of x) - 8, as 4 packed signed bytes
__device__ __forceinline__ uint32_t nib_minus8(uint32_t x) { return (((x & 0x0f0f0f0fu) | 0x80808080u) - 0x08080808u) ^ 0x80808080u; }

// one thread per 32-element block: ggml block -> 32 int8 at q[row][kb*32 ..] and float scale(s) at sc/mn[kb][row]
template <int TYPE>
__global__ void __launch_bounds__(256) expand_weights_kernel(const uint8_t * __restrict__ w, int64_t nb01, int64_t M, int64_t K,
                                                             int8_t * __restrict__ q, float * __restrict__ sc, float * __restrict__ mn,
                                                             int64_t Mp) {
    const int64_t KB  = K / 32;
    const int64_t idx = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= M * KB) return;
    const int64_t row = idx / KB, kb = idx - row * KB;
    const uint8_t * wr = w + row * nb01;
    uint32_t o[8];
    if (TYPE == B200_TYPE_Q4_0) {  // { half d; uint8 qs[16] }: element j = low nibble of qs[j], j + 16 = high nibble (ggml-quants.c:307-325)
        const uint8_t * b = wr + kb * SZ_Q4_0;
        const float d = half_bits_to_float(*(const uint16_t *) b);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t x = ld16x2(b + 2 + 4 * i);
            o[i]     = nib_minus8(x);
            o[4 + i] = nib_minus8(x >> 4);
        }
        sc[kb * Mp + row] = d;
    } else if (TYPE == B200_TYPE_Q8_0) {  // { half d; int8 qs[32] }
        const uint8_t * b = wr + kb * SZ_Q8_0;
        const float d = half_bits_to_float(*(const uint16_t *) b);
#pragma unroll
        for (int i = 0; i < 8; i++) o[i] = ld16x2(b + 2 + 4 * i);
        sc[kb * Mp + row] = d;
    } else {  // q4_K: { half d, dmin; uint8 scales[12]; uint8 qs[128] } (ggml-common.h:291-305), sub-block j of 8
        const uint8_t * b = wr + (kb >> 3) * SZ_Q4_K;
        const int j = (int) (kb & 7);
        const uint32_t hdr = *(const uint32_t *) b;
        const float d = half_bits_to_float((uint16_t) (hdr & 0xffff)), dmin = half_bits_to_float((uint16_t) (hdr >> 16));
        const uint8_t * s = b + 4;
        int scv, mv;  // get_scale_min_k4 (ggml-quants.c:703-711)
        if (j < 4) {
            scv = s[j] & 63;
            mv  = s[j + 4] & 63;
        } else {
            scv = (s[j + 4] & 0xF) | ((s[j - 4] >> 6) << 4);
            mv  = (s[j + 4] >> 4) | ((s[j] >> 6) << 4);
        }
        const uint4 * qs = (const uint4 *) (b + 16 + 32 * (j >> 1));  // 32 bytes: low nibbles = sub-block 2i, high = 2i+1
        const uint4 a = qs[0], c = qs[1];
        const uint32_t x[8] = { a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w };
        const int sh = (j & 1) ? 4 : 0;
#pragma unroll
        for (int i = 0; i < 8; i++) o[i] = (x[i] >> sh) & 0x0f0f0f0fu;
        sc[kb * Mp + row] = d * (float) scv;
        mn[kb * Mp + row] = dmin * (float) mv;
    }
    uint4 * dst = (uint4 *) (q + row * K + kb * 32);
    dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
    dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
}

// ------------------------------------------------------------------------------------------------ activation tiles (K2)

// one warp per 256-element unit of one activation row: same values as quantize_row_q8_0_ref (KIND 0) / q8_K_ref (KIND 1)
template <int KIND>
__global__ void __launch_bounds__(128) quantize_tiles_kernel(const float * __restrict__ x, int64_t K, int64_t N, size_t x_rs,
                                                             int8_t * __restrict__ q, float * __restrict__ sA, float * __restrict__ aS,
                                                             int64_t Np, int vec_ok, int16_t * __restrict__ bs = nullptr) {
    const int     lane  = threadIdx.x & 31;
    const int64_t unit  = (int64_t) blockIdx.x * 4 + (threadIdx.x >> 5);
    const int64_t n     = blockIdx.y;
    const int64_t units = (K + UNIT_ELEMS - 1) / UNIT_ELEMS;
    if (unit >= units || n >= N) return;
    const int64_t e0     = unit * UNIT_ELEMS + lane * 8;
    const bool    active = e0 < K;
    float v[8];
    const float * xr = (const float *) ((const char *) x + n * x_rs) + e0;
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
    const int64_t kb = unit * 8 + (lane >> 2);
    if (KIND == ACT_KIND_Q8_0) {
        const Q80Lane r = quant_lane_q8_0(v);
        if (!active) return;
        *(uint2 *) (q + n * K + e0) = make_uint2(r.w0, r.w1);
        if ((lane & 3) == 0) sA[kb * Np + n] = __half2float(__float2half_rn(r.d));  // the CPU reads the fp16-rounded d (quants.c:146)
    } else {
        const Q8KLane r = quant_lane_q8_K(v, lane);
        *(uint2 *) (q + n * K + e0) = make_uint2(r.w0, r.w1);
        if ((lane & 3) == 0) {
            sA[kb * Np + n] = r.d;
            if (bs) bs[(n * units + unit) * 8 + (lane >> 2)] = (int16_t) r.s32;  // q8_K bsums per 32 (the CPU sums its 16-wide pairs)
            else aS[kb * Np + n] = r.d * (float) r.s32;
        }
    }
}

// ------------------------------------------------------------------------------------------------ the GEMM (K4)

struct MmqParams {
    SplitParams split;
    float * dst;
    int64_t ldc;  // floats between dst columns (tokens)
    int     M, N, K;
    int     m_tiles, n_tiles;
};

template <bool HAS_MIN>
__global__ void __launch_bounds__(MMQ_THREADS, 1)
mmq_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const __grid_constant__ CUtensorMap map_sw,
           const __grid_constant__ CUtensorMap map_sa, const __grid_constant__ CUtensorMap map_mw, const __grid_constant__ CUtensorMap map_as,
           const MmqParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);  // 128B-swizzle atoms are 1024-byte aligned
    uint64_t * bars      = (uint64_t *) (smem + STAGES * STAGE_BYTES);
    uint64_t * full      = bars;               // [STAGES]  TMA -> MMA, epilogue
    uint64_t * empty     = bars + STAGES;      // [STAGES]  MMA commit + epilogue warps -> TMA
    uint64_t * tfull     = bars + 2 * STAGES;  // [SLOTS]   MMA commit -> epilogue
    uint64_t * tempty    = tfull + SLOTS;      // [SLOTS]   epilogue warps -> MMA
    uint32_t * tmem_slot = (uint32_t *) (tempty + SLOTS);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tiles_total = p.m_tiles * p.n_tiles;
    const int k_stages = (p.K + 127) / 128;

    if (threadIdx.x == 0) {
        for (int i = 0; i < STAGES; i++) {
            mbar_init(smem_u32(full + i), 1);
            mbar_init(smem_u32(empty + i), 1 + EPI_WARPS);
        }
        for (int i = 0; i < SLOTS; i++) {
            mbar_init(smem_u32(tfull + i), 1);
            mbar_init(smem_u32(tempty + i), EPI_WARPS);
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<512>(smem_u32(tmem_slot));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================================================================ TMA producer
        if (lane == 0) {
            tma_prefetch_desc(&map_a); tma_prefetch_desc(&map_b); tma_prefetch_desc(&map_sw); tma_prefetch_desc(&map_sa);
            int stage = 0; uint32_t phase = 1;
            for (int t = blockIdx.x; t < n_tiles_total; t += gridDim.x) {
                const int m0 = (t / p.n_tiles) * BM, n0 = (t % p.n_tiles) * BN;
                for (int ks = 0; ks < k_stages; ks++) {
                    mbar_wait(smem_u32(empty + stage), phase);
                    const uint32_t bar = smem_u32(full + stage);
                    uint8_t * st = smem + stage * STAGE_BYTES;
                    mbar_expect_tx(bar, A_BYTES + B_BYTES + (HAS_MIN ? 4 : 2) * S_BYTES);
                    tma_load_2d(smem_u32(st), &map_a, ks * 128, m0, bar);
                    tma_load_2d(smem_u32(st + A_BYTES), &map_b, ks * 128, n0, bar);
                    tma_load_2d(smem_u32(st + A_BYTES + B_BYTES), &map_sw, m0, ks * BKB, bar);
                    tma_load_2d(smem_u32(st + A_BYTES + B_BYTES + S_BYTES), &map_sa, n0, ks * BKB, bar);
                    if (HAS_MIN) {
                        tma_load_2d(smem_u32(st + A_BYTES + B_BYTES + 2 * S_BYTES), &map_mw, m0, ks * BKB, bar);
                        tma_load_2d(smem_u32(st + A_BYTES + B_BYTES + 3 * S_BYTES), &map_as, n0, ks * BKB, bar);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ================================================================ MMA issuer: one UTCIMMA (128 x 128 x 32) per 32-block
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_i8(BM, BN);
            int stage = 0, slot = 0; uint32_t phase = 0, tphase = 1;
            for (int t = blockIdx.x; t < n_tiles_total; t += gridDim.x) {
                for (int ks = 0; ks < k_stages; ks++) {
                    mbar_wait(smem_u32(full + stage), phase);
                    tc_fence_after();
                    const uint32_t a0 = smem_u32(smem + stage * STAGE_BYTES), b0 = a0 + A_BYTES;
#pragma unroll
                    for (int kb = 0; kb < BKB; kb++) {
                        mbar_wait(smem_u32(tempty + slot), tphase);
                        tc_fence_after();
                        umma_i8(tmem_base + slot * BN, umma_desc_sw128(a0 + kb * 32), umma_desc_sw128(b0 + kb * 32), idesc, 0);
                        umma_commit(smem_u32(tfull + slot));
                        if (++slot == SLOTS) { slot = 0; tphase ^= 1; }
                    }
                    umma_commit(smem_u32(empty + stage));  // the operands of this stage have been read
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp >= FIRST_EPI_WARP) {
        // ================================================================ epilogue: TMEM -> registers, scale, accumulate, store
        // A thread owns one weight row (TMEM lane) and 32 of the tile's 128 columns: per 32-block one tcgen05.ld.32x32b.x32, then
        // I2FP + FMUL2 + FFMA2 on the 32 values.
        const int quad = warp & 3, part = (warp - FIRST_EPI_WARP) >> 2;
        const int row_in_tile = quad * 32 + lane;
        const uint32_t t_lane = tmem_base + ((uint32_t) (quad * 32) << 16) + part * 32;
        int stage = 0, slot = 0; uint32_t phase = 0, tphase = 0;
        for (int t = blockIdx.x; t < n_tiles_total; t += gridDim.x) {
            const int m0 = (t / p.n_tiles) * BM, n0 = (t % p.n_tiles) * BN;
            float2 acc[16];
#pragma unroll
            for (int j = 0; j < 16; j++) acc[j] = make_float2(0.0f, 0.0f);
            const int n_blocks = k_stages * BKB;
            uint32_t c[32];

            auto math = [&](const uint32_t (&c)[32], int kb, uint32_t st) {
                const float sw = __uint_as_float(lds32(st + 4 * (kb * 128 + row_in_tile)));
                const float2 sw2 = make_float2(sw, sw);
                const uint32_t sa = st + S_BYTES + 4 * (kb * 128 + part * 32);
                float2 mw2 = make_float2(0.0f, 0.0f);
                if (HAS_MIN) {
                    const float mw = -__uint_as_float(lds32(st + 2 * S_BYTES + 4 * (kb * 128 + row_in_tile)));
                    mw2 = make_float2(mw, mw);
                }
#pragma unroll
                for (int j4 = 0; j4 < 8; j4++) {
                    const uint4 a = lds128(sa + 16 * j4);
                    const float2 p0 = fmul2(make_float2(__uint_as_float(a.x), __uint_as_float(a.y)), sw2);
                    const float2 p1 = fmul2(make_float2(__uint_as_float(a.z), __uint_as_float(a.w)), sw2);
                    float2 & r0 = acc[2 * j4], & r1 = acc[2 * j4 + 1];
                    r0 = ffma2(p0, make_float2((float) (int) c[4 * j4 + 0], (float) (int) c[4 * j4 + 1]), r0);
                    r1 = ffma2(p1, make_float2((float) (int) c[4 * j4 + 2], (float) (int) c[4 * j4 + 3]), r1);
                    if (HAS_MIN) {
                        const uint4 b = lds128(sa + 2 * S_BYTES + 16 * j4);
                        r0 = ffma2(mw2, make_float2(__uint_as_float(b.x), __uint_as_float(b.y)), r0);
                        r1 = ffma2(mw2, make_float2(__uint_as_float(b.z), __uint_as_float(b.w)), r1);
                    }
                }
            };
            // four warps per scheduler hide the TMEM round trip of one another: a single register buffer per thread is enough
            for (int blk = 0; blk < n_blocks; blk++) {
                const int kb = blk & (BKB - 1);
                mbar_wait(smem_u32(tfull + slot), tphase);
                tc_fence_after();
                tmem_ld_32x32(t_lane + slot * BN, c);
                if (kb == 0) mbar_wait(smem_u32(full + stage), phase);  // the scales of this stage have landed
                const uint32_t st = smem_u32(smem + stage * STAGE_BYTES + A_BYTES + B_BYTES);
                tc_wait_ld_regs(c);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(tempty + slot));  // the block is in registers: the next MMA may overwrite the slot
                if (++slot == SLOTS) { slot = 0; tphase ^= 1; }
                math(c, kb, st);
                if (kb == BKB - 1) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(smem_u32(empty + stage));  // done with this stage's scales
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
            const int m = m0 + row_in_tile;
            if (m < p.M) {
                float v[32];
#pragma unroll
                for (int j = 0; j < 16; j++) { v[2 * j] = acc[j].x; v[2 * j + 1] = acc[j].y; }
                const int nb = n0 + part * 32;
                split_store_cols32(p.split, p.dst + (int64_t) nb * p.ldc + m, p.ldc, v, p.N - nb);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (p.split.n_ranks > 1 && warp == 0) split_complete(p.split, gridDim.x);
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

// ------------------------------------------------------------------------------------------------ q4_K: exact integer scale folding
//
// q4_K's sub-block scales are 6-bit integers under one fp16 d per 256-weight super-block, and q8_K activations carry one scale
// per 256 elements, so the per-32 rescale can move into the integer domain:  sc = 8*a + b  (a, b in 0..7)  gives
//     sum_j sc_j * sum_{k in j} q_k * y_k  =  8 * sum_k (q_k * a_j(k)) * y_k  +  sum_k (q_k * b_j(k)) * y_k
// with q*a and q*b <= 105, i.e. two int8 weight planes. Per super-block the tensor core accumulates 8 + 8 UTCIMMAs (K = 256) into
// two int32 TMEM accumulators and the CUDA cores touch each output once per 256 instead of once per 32 elements:
//     dst[n][m] = sum_sb  dA[sb][n] * ( d[sb][m] * (8 * Ca + Cb)  -  dmin[sb][m] * sum_kb m_kb[m] * bsum_kb[n] )
// where the min term is an 8-term integer dot product (dp2a: int16 block sums x 6-bit mins) per output and super-block
// - the same integers ggml-cpu's vec_dot_q4_K_q8_K forms (ggml-cpu/quants.c:560-640), so parity stays in the 1e-13 class.

// one thread per 32-block: plane a = q * (sc >> 3), plane b = q * (sc & 7) (packed bytes: no carries, products <= 105)
__global__ void __launch_bounds__(256) expand_q4k_planes_kernel(const uint8_t * __restrict__ w, int64_t nb01, int64_t M, int64_t K,
                                                                int8_t * __restrict__ qa, int8_t * __restrict__ qb, float * __restrict__ dw,
                                                                float * __restrict__ dm, uint8_t * __restrict__ mb, int64_t Mp) {
    const int64_t KB  = K / 32;
    const int64_t idx = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= M * KB) return;
    const int64_t row = idx / KB, kb = idx - row * KB;
    const uint8_t * b = w + row * nb01 + (kb >> 3) * SZ_Q4_K;
    const int j = (int) (kb & 7);
    const uint32_t hdr = *(const uint32_t *) b;
    const float d = half_bits_to_float((uint16_t) (hdr & 0xffff)), dmin = half_bits_to_float((uint16_t) (hdr >> 16));
    const uint8_t * s = b + 4;
    int scv, mv;  // get_scale_min_k4 (ggml-quants.c:703-711)
    if (j < 4) {
        scv = s[j] & 63;
        mv  = s[j + 4] & 63;
    } else {
        scv = (s[j + 4] & 0xF) | ((s[j - 4] >> 6) << 4);
        mv  = (s[j + 4] >> 4) | ((s[j] >> 6) << 4);
    }
    const uint4 * qs = (const uint4 *) (b + 16 + 32 * (j >> 1));
    const uint4 v0 = qs[0], v1 = qs[1];
    const uint32_t x[8] = { v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w };
    const int sh = (j & 1) ? 4 : 0;
    const uint32_t fa = (uint32_t) (scv >> 3), fb = (uint32_t) (scv & 7);
    uint32_t oa[8], ob[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t q = (x[i] >> sh) & 0x0f0f0f0fu;
        oa[i] = q * fa;
        ob[i] = q * fb;
    }
    uint4 * da = (uint4 *) (qa + row * K + kb * 32);
    uint4 * db = (uint4 *) (qb + row * K + kb * 32);
    da[0] = make_uint4(oa[0], oa[1], oa[2], oa[3]); da[1] = make_uint4(oa[4], oa[5], oa[6], oa[7]);
    db[0] = make_uint4(ob[0], ob[1], ob[2], ob[3]); db[1] = make_uint4(ob[4], ob[5], ob[6], ob[7]);
    if (j == 0) {
        dw[(kb >> 3) * Mp + row] = d;
        dm[(kb >> 3) * Mp + row] = dmin;
    }
    uint8_t * mrow = mb + (row * (K / 256) + (kb >> 3)) * 16;  // 8 mins (+ 8 bytes of padding: TMA boxes are >= 16 bytes wide)
    mrow[j]     = (uint8_t) mv;
    mrow[8 + j] = 0;
}

constexpr int KQ_OSTAGES = 4, KQ_SSTAGES = 3, KQ_SLOTS = 2;   // operand ring (half super-blocks), scale ring (super-blocks), TMEM slots
constexpr int KQ_O_BYTES = 3 * BM * 128;                       // plane a, plane b, activations: 128 rows x 128 K-bytes each
constexpr int KQ_S_BYTES = 3 * 128 * 4 + 2 * 128 * 16;         // dW[128], dmin[128], dA[128], mins[128][16 B], bsums[128][8 x int16]
constexpr int KQ_SMEM    = 1024 + KQ_OSTAGES * KQ_O_BYTES + KQ_SSTAGES * KQ_S_BYTES + 256;

__global__ void __launch_bounds__(MMQ_THREADS, 1)
mmq_q4k_kernel(const __grid_constant__ CUtensorMap map_wa, const __grid_constant__ CUtensorMap map_wb, const __grid_constant__ CUtensorMap map_b,
               const __grid_constant__ CUtensorMap map_mb, const __grid_constant__ CUtensorMap map_bs, const __grid_constant__ CUtensorMap map_dw,
               const __grid_constant__ CUtensorMap map_dm, const __grid_constant__ CUtensorMap map_da, const MmqParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *  smem   = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
    uint8_t *  sring  = smem + KQ_OSTAGES * KQ_O_BYTES;
    uint64_t * bars   = (uint64_t *) (sring + KQ_SSTAGES * KQ_S_BYTES);
    uint64_t * ofull  = bars;                      // [OSTAGES] TMA -> MMA
    uint64_t * oempty = ofull + KQ_OSTAGES;        // [OSTAGES] MMA commit -> TMA
    uint64_t * sfull  = oempty + KQ_OSTAGES;       // [SSTAGES] TMA -> epilogue
    uint64_t * sempty = sfull + KQ_SSTAGES;        // [SSTAGES] epilogue warps -> TMA
    uint64_t * tfull  = sempty + KQ_SSTAGES;       // [SLOTS]   MMA commit -> epilogue
    uint64_t * tempty = tfull + KQ_SLOTS;          // [SLOTS]   epilogue warps -> MMA
    uint32_t * tmem_slot = (uint32_t *) (tempty + KQ_SLOTS);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tiles_total = p.m_tiles * p.n_tiles;
    const int n_sb = p.K / 256;

    if (threadIdx.x == 0) {
        for (int i = 0; i < KQ_OSTAGES; i++) { mbar_init(smem_u32(ofull + i), 1); mbar_init(smem_u32(oempty + i), 1); }
        for (int i = 0; i < KQ_SSTAGES; i++) { mbar_init(smem_u32(sfull + i), 1); mbar_init(smem_u32(sempty + i), EPI_WARPS); }
        for (int i = 0; i < KQ_SLOTS; i++) { mbar_init(smem_u32(tfull + i), 1); mbar_init(smem_u32(tempty + i), EPI_WARPS); }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<512>(smem_u32(tmem_slot));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================================================================ TMA producer
        if (lane == 0) {
            tma_prefetch_desc(&map_wa); tma_prefetch_desc(&map_wb); tma_prefetch_desc(&map_b);
            int os = 0, ss = 0; uint32_t ophase = 1, sphase = 1;
            for (int t = blockIdx.x; t < n_tiles_total; t += gridDim.x) {
                const int m0 = (t / p.n_tiles) * BM, n0 = (t % p.n_tiles) * BN;
                for (int sb = 0; sb < n_sb; sb++) {
                    mbar_wait(smem_u32(sempty + ss), sphase);
                    {
                        const uint32_t bar = smem_u32(sfull + ss);
                        const uint32_t st  = smem_u32(sring + ss * KQ_S_BYTES);
                        mbar_expect_tx(bar, KQ_S_BYTES);
                        tma_load_2d(st, &map_dw, m0, sb, bar);
                        tma_load_2d(st + 512, &map_dm, m0, sb, bar);
                        tma_load_2d(st + 1024, &map_da, n0, sb, bar);
                        tma_load_2d(st + 1536, &map_mb, sb * 16, m0, bar);
                        tma_load_2d(st + 1536 + 2048, &map_bs, sb * 8, n0, bar);
                    }
                    if (++ss == KQ_SSTAGES) { ss = 0; sphase ^= 1; }
                    for (int h = 0; h < 2; h++) {
                        mbar_wait(smem_u32(oempty + os), ophase);
                        const uint32_t bar = smem_u32(ofull + os);
                        const uint32_t st  = smem_u32(smem + os * KQ_O_BYTES);
                        mbar_expect_tx(bar, KQ_O_BYTES);
                        tma_load_2d(st, &map_wa, sb * 256 + h * 128, m0, bar);
                        tma_load_2d(st + BM * 128, &map_wb, sb * 256 + h * 128, m0, bar);
                        tma_load_2d(st + 2 * BM * 128, &map_b, sb * 256 + h * 128, n0, bar);
                        if (++os == KQ_OSTAGES) { os = 0; ophase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================================================================ MMA issuer: 8 + 8 UTCIMMAs (128 x 128 x 32) per super-block
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_i8(BM, BN);
            int os = 0, slot = 0; uint32_t ophase = 0, tphase = 1;
            for (int t = blockIdx.x; t < n_tiles_total; t += gridDim.x) {
                for (int sb = 0; sb < n_sb; sb++) {
                    mbar_wait(smem_u32(tempty + slot), tphase);
                    tc_fence_after();
                    const uint32_t d_a = tmem_base + slot * 2 * BN, d_b = d_a + BN;
                    for (int h = 0; h < 2; h++) {
                        mbar_wait(smem_u32(ofull + os), ophase);
                        tc_fence_after();
                        const uint32_t a0 = smem_u32(smem + os * KQ_O_BYTES), a1 = a0 + BM * 128, b0 = a1 + BM * 128;
#pragma unroll
                        for (int kb = 0; kb < 4; kb++) {
                            const uint32_t accumulate = (h | kb) != 0;
                            const uint64_t db = umma_desc_sw128(b0 + kb * 32);
                            umma_i8(d_a, umma_desc_sw128(a0 + kb * 32), db, idesc, accumulate);
                            umma_i8(d_b, umma_desc_sw128(a1 + kb * 32), db, idesc, accumulate);
                        }
                        umma_commit(smem_u32(oempty + os));  // the operands of this stage have been read
                        if (++os == KQ_OSTAGES) { os = 0; ophase ^= 1; }
                    }
                    umma_commit(smem_u32(tfull + slot));
                    if (++slot == KQ_SLOTS) { slot = 0; tphase ^= 1; }
                }
            }
        }
    } else if (warp >= FIRST_EPI_WARP) {
        // ================================================================ epilogue: one pass per super-block
        const int quad = warp & 3, part = (warp - FIRST_EPI_WARP) >> 2;
        const int row_in_tile = quad * 32 + lane;
        const uint32_t t_lane = tmem_base + ((uint32_t) (quad * 32) << 16) + part * 32;
        int ss = 0, slot = 0; uint32_t sphase = 0, tphase = 0;
        for (int t = blockIdx.x; t < n_tiles_total; t += gridDim.x) {
            const int m0 = (t / p.n_tiles) * BM, n0 = (t % p.n_tiles) * BN;
            float2 acc[16];
#pragma unroll
            for (int j = 0; j < 16; j++) acc[j] = make_float2(0.0f, 0.0f);
            for (int sb = 0; sb < n_sb; sb++) {
                mbar_wait(smem_u32(tfull + slot), tphase);
                tc_fence_after();
                uint32_t ca[16], cb[16];
                const uint32_t t_a = t_lane + slot * 2 * BN, t_b = t_a + BN;
                tmem_ld_32x16(t_a, ca);
                tmem_ld_32x16(t_b, cb);
                mbar_wait(smem_u32(sfull + ss), sphase);  // the scales of this super-block have landed
                const uint32_t st = smem_u32(sring + ss * KQ_S_BYTES);
                const float dw = __uint_as_float(lds32(st + 4 * row_in_tile));
                const float2 dw2 = make_float2(dw, dw);
                const uint32_t da = st + 1024 + 4 * (part * 32);
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    tc_wait_ld_regs2(ca, cb);
                    float2 v[8];
#pragma unroll
                    for (int j = 0; j < 8; j++)
                        v[j] = make_float2((float) (int) ((ca[2 * j] << 3) + cb[2 * j]), (float) (int) ((ca[2 * j + 1] << 3) + cb[2 * j + 1]));
                    if (h == 0) {  // second half of the columns in flight while the first is scaled
                        tmem_ld_32x16(t_a + 16, ca);
                        tmem_ld_32x16(t_b + 16, cb);
                    } else {
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(smem_u32(tempty + slot));  // both accumulators are in registers
                    }
#pragma unroll
                    for (int j4 = 0; j4 < 4; j4++) {
                        const uint4 a = lds128(da + 64 * h + 16 * j4);
                        const float2 s0 = fmul2(make_float2(__uint_as_float(a.x), __uint_as_float(a.y)), dw2);
                        const float2 s1 = fmul2(make_float2(__uint_as_float(a.z), __uint_as_float(a.w)), dw2);
                        acc[8 * h + 2 * j4]     = ffma2(s0, v[2 * j4], acc[8 * h + 2 * j4]);
                        acc[8 * h + 2 * j4 + 1] = ffma2(s1, v[2 * j4 + 1], acc[8 * h + 2 * j4 + 1]);
                    }
                }
                if (++slot == KQ_SLOTS) { slot = 0; tphase ^= 1; }
                // mins: acc -= dmin[row] * dA[col] * sum_kb m_kb[row] * bsum_kb[col]   (integer dot, ggml-cpu/quants.c:618-626)
                {
                    const float dm = -__uint_as_float(lds32(st + 512 + 4 * row_in_tile));
                    const float2 dm2 = make_float2(dm, dm);
                    const uint2 mm = lds64(st + 1536 + 16 * row_in_tile);
                    const uint32_t bs = st + 1536 + 2048 + 16 * (part * 32);
#pragma unroll
                    for (int j4 = 0; j4 < 8; j4++) {
                        const uint4 a = lds128(da + 16 * j4);
                        int iv[4];
#pragma unroll
                        for (int c = 0; c < 4; c++) {
                            const uint4 b = lds128(bs + 16 * (4 * j4 + c));
                            int i = dp2a_lo_su((int) b.x, mm.x, 0);
                            i     = dp2a_hi_su((int) b.y, mm.x, i);
                            i     = dp2a_lo_su((int) b.z, mm.y, i);
                            iv[c] = dp2a_hi_su((int) b.w, mm.y, i);
                        }
                        const float2 s0 = fmul2(make_float2(__uint_as_float(a.x), __uint_as_float(a.y)), dm2);
                        const float2 s1 = fmul2(make_float2(__uint_as_float(a.z), __uint_as_float(a.w)), dm2);
                        acc[2 * j4]     = ffma2(s0, make_float2((float) iv[0], (float) iv[1]), acc[2 * j4]);
                        acc[2 * j4 + 1] = ffma2(s1, make_float2((float) iv[2], (float) iv[3]), acc[2 * j4 + 1]);
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(sempty + ss));  // done with this super-block's scales
                if (++ss == KQ_SSTAGES) { ss = 0; sphase ^= 1; }
            }
            const int m = m0 + row_in_tile;
            if (m < p.M) {
                float v[32];
#pragma unroll
                for (int j = 0; j < 16; j++) { v[2 * j] = acc[j].x; v[2 * j + 1] = acc[j].y; }
                const int nb = n0 + part * 32;
                split_store_cols32(p.split, p.dst + (int64_t) nb * p.ldc + m, p.ldc, v, p.N - nb);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (p.split.n_ranks > 1 && warp == 0) split_complete(p.split, gridDim.x);
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

// ------------------------------------------------------------------------------------------------ host side

typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

encode_tiled_fn get_encode() {
    static encode_tiled_fn fn = nullptr;
    static std::once_flag  once;
    std::call_once(once, [] {
        void * p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess) {
            fn = (encode_tiled_fn) p;
        }
    });
    return fn;
}

// 2-D map over a row-major array: `inner` elements per row (contiguous), `outer` rows, `pitch` bytes between rows
bool make_map(CUtensorMap * m, CUtensorMapDataType dt, int esz, void * base, uint64_t inner, uint64_t outer, uint64_t pitch, uint32_t box_inner,
              uint32_t box_outer, bool swizzle128) {
    encode_tiled_fn enc = get_encode();
    if (!enc) return false;
    cuuint64_t dims[2]    = { inner, outer };
    cuuint64_t strides[1] = { pitch };
    cuuint32_t box[2]     = { box_inner, box_outer };
    cuuint32_t estr[2]    = { 1, 1 };
    (void) esz;
    return enc(m, dt, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

struct WEntry {  // expanded form of one weight matrix
    int       type = 0;
    int64_t   M = 0, K = 0, Mp = 0;
    size_t    nb01 = 0;
    uint8_t * buf = nullptr;  // [q: M*K int8][sc: KB*Mp float][mn: KB*Mp float]
    size_t    bytes = 0;
    CUtensorMap map_q, map_sc, map_mn;
    bool        planes = false;   // q4_K scale-folded form: [qa][qb][d: SB*Mp float][dmin: SB*Mp float][mins: M*SB*16 bytes];
                                  // map_sc = dmin, map_mn = mins
    CUtensorMap map_qb, map_dw;
    bool        f16 = false;      // fp16 [M][K] for the kind::f16 kernel (mmf16.cuh): map_q only
    bool        settled = false;  // the expansion kernel is known to have finished (set after a stream synchronize)
    // cache bookkeeping (under g_wmutex)
    size_t      extent = 0;       // bytes of the SOURCE tensor the expansion was made from: (M - 1) * nb01 + one row
    uint64_t    stamp  = 0;       // last use (LRU)
    int         refs   = 0;       // calls that are using the entry right now: never freed under them
    bool        dead   = false;   // invalidated while in use: the last user frees it
};

#include "mmf16.cuh"
#include "mvtc.cuh"

struct WKey {
    int          device;
    const void * src;
    bool operator<(const WKey & o) const { return device != o.device ? device < o.device : src < o.src; }
};

std::mutex                 g_wmutex;
std::map<WKey, WEntry *>   g_wcache;
size_t                     g_wbytes = 0;
uint64_t                   g_wclock = 0;
std::map<WKey, int>        g_wrewrites;   // how often a cached source was overwritten: tensors rewritten every graph are not cached
size_t                     g_wbudget = 0; // 0 = not set yet: 40 % of the device's memory (GGML_B200_WEIGHT_CACHE_MB overrides)

size_t wcache_budget(int device) {
    if (g_wbudget == 0) {
        if (const char * s = getenv("GGML_B200_WEIGHT_CACHE_MB")) g_wbudget = (size_t) atoll(s) << 20;
        if (g_wbudget == 0) {
            size_t fr = 0, tot = 0;
            cudaSetDevice(device);
            if (cudaMemGetInfo(&fr, &tot) != cudaSuccess) { cudaGetLastError(); tot = (size_t) 64 << 30; }
            g_wbudget = tot / 5 * 2;
        }
    }
    return g_wbudget;
}

void wentry_free(WEntry * e, int device) {  // g_wmutex held; cudaFree synchronizes with the device: no kernel still reads the buffer
    cudaSetDevice(device);
    cudaFree(e->buf);
    delete e;
}

// makes room for `need` more bytes on `device` by dropping least recently used entries nobody is using; g_wmutex held
bool wcache_make_room(int device, size_t need) {
    const size_t budget = wcache_budget(device);
    if (need > budget) return false;
    while (g_wbytes + need > budget) {
        auto victim = g_wcache.end();
        for (auto it = g_wcache.begin(); it != g_wcache.end(); ++it) {
            if (it->second->refs > 0) continue;
            if (victim == g_wcache.end() || it->second->stamp < victim->second->stamp) victim = it;
        }
        if (victim == g_wcache.end()) return false;
        g_wbytes -= victim->second->bytes;
        wentry_free(victim->second, victim->first.device);
        g_wcache.erase(victim);
    }
    return true;
}

void wcache_release(WEntry * e, int device) {
    std::lock_guard<std::mutex> lock(g_wmutex);
    if (--e->refs == 0 && e->dead) wentry_free(e, device);
}

bool q4k_planes_enabled() {
    static const bool on = [] { const char * s = getenv("GGML_B200_MMQ_Q4K_PLANES"); return !(s && atoi(s) == 0); }();
    return on;
}

size_t wentry_bytes(int type, int64_t M, int64_t K, int64_t Mp, bool f16) {
    if (f16) return ((size_t) M * K * 2 + 255) & ~(size_t) 255;
    const size_t qb = ((size_t) M * K + 255) & ~(size_t) 255;
    const size_t sb = (((size_t) (K / 32) * Mp * 4) + 255) & ~(size_t) 255;
    if (type == B200_TYPE_Q4_K && K % 256 == 0 && q4k_planes_enabled()) {
        const size_t db = (((size_t) (K / 256) * Mp * 4) + 255) & ~(size_t) 255;
        const size_t mb = (((size_t) M * (K / 256) * 16) + 255) & ~(size_t) 255;
        return 2 * qb + 2 * db + mb;
    }
    return qb + sb * (type == B200_TYPE_Q4_K ? 2 : 1);
}

b200_status wentry_fill(b200_ctx * ctx, WEntry * e, const void * src, uint8_t * buf) {
    const int64_t M = e->M, K = e->K, KB = K / 32;
    const size_t  qb = ((size_t) M * K + 255) & ~(size_t) 255;
    const size_t  sb = (((size_t) KB * e->Mp * 4) + 255) & ~(size_t) 255;
    int8_t * q  = (int8_t *) buf;
    float *  sc = (float *) (buf + qb);
    float *  mn = e->type == B200_TYPE_Q4_K ? (float *) (buf + qb + sb) : sc;
    e->buf = buf;
    const int64_t total = M * KB;
    const unsigned grid = (unsigned) ((total + 255) / 256);
    if (e->f16) {
        __half * h = (__half *) buf;
        switch (e->type) {
            case B200_TYPE_Q4_0: expand_f16_kernel<B200_TYPE_Q4_0><<<grid, 256, 0, ctx->stream>>>((const uint8_t *) src, (int64_t) e->nb01, M, K, h); break;
            case B200_TYPE_Q8_0: expand_f16_kernel<B200_TYPE_Q8_0><<<grid, 256, 0, ctx->stream>>>((const uint8_t *) src, (int64_t) e->nb01, M, K, h); break;
            case B200_TYPE_Q5_K: expand_f16_kernel<B200_TYPE_Q5_K><<<grid, 256, 0, ctx->stream>>>((const uint8_t *) src, (int64_t) e->nb01, M, K, h); break;
            case B200_TYPE_Q6_K: expand_f16_kernel<B200_TYPE_Q6_K><<<grid, 256, 0, ctx->stream>>>((const uint8_t *) src, (int64_t) e->nb01, M, K, h); break;
            default:             expand_f16_kernel<B200_TYPE_Q4_K><<<grid, 256, 0, ctx->stream>>>((const uint8_t *) src, (int64_t) e->nb01, M, K, h); break;
        }
        CU(cudaGetLastError());
        ctx->launches++;
        if (!make_map(&e->map_q, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, h, (uint64_t) K, (uint64_t) M, (uint64_t) K * 2, F_BK, F_BM, true)) {
            return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled failed for the fp16 weights (M=%lld K=%lld)", (long long) M, (long long) K);
        }
        return B200_OK;
    }
    if (e->type == B200_TYPE_Q4_K && K % 256 == 0 && q4k_planes_enabled()) {
        const size_t db = (((size_t) (K / 256) * e->Mp * 4) + 255) & ~(size_t) 255;
        int8_t * qa = (int8_t *) buf, * qbp = (int8_t *) (buf + qb);
        float *   dw = (float *) (buf + 2 * qb);
        float *   dm = (float *) (buf + 2 * qb + db);
        uint8_t * mb = buf + 2 * qb + 2 * db;
        expand_q4k_planes_kernel<<<grid, 256, 0, ctx->stream>>>((const uint8_t *) src, (int64_t) e->nb01, M, K, qa, qbp, dw, dm, mb, e->Mp);
        CU(cudaGetLastError());
        ctx->launches++;
        e->planes = true;
        if (!make_map(&e->map_q, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, qa, (uint64_t) K, (uint64_t) M, (uint64_t) K, 128, BM, true) ||
            !make_map(&e->map_qb, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, qbp, (uint64_t) K, (uint64_t) M, (uint64_t) K, 128, BM, true) ||
            !make_map(&e->map_dw, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, dw, (uint64_t) M, (uint64_t) (K / 256), (uint64_t) e->Mp * 4, 128, 1, false) ||
            !make_map(&e->map_sc, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, dm, (uint64_t) M, (uint64_t) (K / 256), (uint64_t) e->Mp * 4, 128, 1, false) ||
            !make_map(&e->map_mn, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, mb, (uint64_t) (K / 256) * 16, (uint64_t) M, (uint64_t) (K / 256) * 16, 16, BM, false)) {
            return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled failed for the q4_K planes (M=%lld K=%lld)", (long long) M, (long long) K);
        }
        return B200_OK;
    }
    switch (e->type) {
        case B200_TYPE_Q4_0: expand_weights_kernel<B200_TYPE_Q4_0><<<grid, 256, 0, ctx->stream>>>((const uint8_t *) src, (int64_t) e->nb01, M, K, q, sc, mn, e->Mp); break;
        case B200_TYPE_Q8_0: expand_weights_kernel<B200_TYPE_Q8_0><<<grid, 256, 0, ctx->stream>>>((const uint8_t *) src, (int64_t) e->nb01, M, K, q, sc, mn, e->Mp); break;
        default:             expand_weights_kernel<B200_TYPE_Q4_K><<<grid, 256, 0, ctx->stream>>>((const uint8_t *) src, (int64_t) e->nb01, M, K, q, sc, mn, e->Mp); break;
    }
    CU(cudaGetLastError());
    ctx->launches++;
    if (!make_map(&e->map_q, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, q, (uint64_t) K, (uint64_t) M, (uint64_t) K, 128, BM, true) ||
        !make_map(&e->map_sc, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, sc, (uint64_t) M, (uint64_t) KB, (uint64_t) e->Mp * 4, 128, BKB, false) ||
        !make_map(&e->map_mn, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, mn, (uint64_t) M, (uint64_t) KB, (uint64_t) e->Mp * 4, 128, BKB, false)) {
        return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled failed for the expanded weights (M=%lld K=%lld)", (long long) M, (long long) K);
    }
    return B200_OK;
}

} // namespace

// ------------------------------------------------------------------------------------------------ weight cache C-ABI

void b200_weight_cache_invalidate(int device, const void * base, size_t size) {
    std::lock_guard<std::mutex> lock(g_wmutex);
    const char * lo = (const char *) base, * hi = size > (size_t) (UINTPTR_MAX - (uintptr_t) lo) ? (const char *) UINTPTR_MAX : lo + size;
    for (auto it = g_wcache.begin(); it != g_wcache.end();) {
        const char * s = (const char *) it->first.src;
        // any overlap of the written range with the bytes the expansion was made from (a write through a view that starts behind
        // row 0 must drop the entry too)
        if (it->first.device == device && s < hi && s + it->second->extent > lo) {
            if (g_wrewrites.size() > 65536) g_wrewrites.clear();
            g_wrewrites[it->first]++;
            g_wbytes -= it->second->bytes;
            if (it->second->refs > 0) it->second->dead = true;  // a call on another thread is using it: that call frees it
            else wentry_free(it->second, device);
            it = g_wcache.erase(it);
        } else {
            ++it;
        }
    }
    if (size == SIZE_MAX) g_wrewrites.clear();
}
void b200_weight_cache_set_budget(size_t bytes) {
    std::lock_guard<std::mutex> lock(g_wmutex);
    g_wbudget = bytes;   // 0: back to the default (recomputed on next use)
}
void b200_weight_cache_clear(int device) { b200_weight_cache_invalidate(device, nullptr, SIZE_MAX); }
size_t b200_weight_cache_bytes(void) {
    std::lock_guard<std::mutex> lock(g_wmutex);
    return g_wbytes;
}

// ------------------------------------------------------------------------------------------------ dispatch

bool b200_mmq_usable(int type, int64_t K, const void * w, size_t nb01) {
    (void) type;
    if (K % 32 != 0 || K < 32) return false;
    if ((type == B200_TYPE_Q4_K || type == B200_TYPE_Q5_K) && ((uintptr_t) w % 16 != 0 || nb01 % 16 != 0)) return false;
    if (type_is_k(type) && K % 256 != 0) return false;
    if ((uintptr_t) w % 2 != 0 || nb01 % 2 != 0) return false;
    return get_encode() != nullptr;
}

// dst[n*ldc + m] = sum_k W[m][k] * X[n][k] for one weight matrix and N activation rows
b200_status b200_mmq_run(b200_ctx * ctx, int type, const void * w, size_t nb01, int64_t M, int64_t K, const float * x, size_t x_rs, int64_t N,
                         float * dst, int64_t ldc, bool act_shared) {
    if (M <= 0 || N <= 0) return B200_OK;
    if (M > INT32_MAX - BM || N > INT32_MAX - BN || K > INT32_MAX - 128) return fail(B200_ERR_UNSUPPORTED, "MMQ shape too large");
    const int64_t KB = K / 32;
    const int64_t Mp = (M + 3) & ~(int64_t) 3, Np = (N + 3) & ~(int64_t) 3;
    const bool    has_min = type == B200_TYPE_Q4_K;

    // ---- activations: int8 [N][K] + scales [KB][Np] (+ sums) in scratch2
    const size_t aq = ((size_t) N * K + 255) & ~(size_t) 255;
    const size_t as = (((size_t) KB * Np * 4) + 255) & ~(size_t) 255;
    const bool   f16 = ctx->mmq_f16 != 0 || type == B200_TYPE_Q6_K || type == B200_TYPE_Q5_K;  // q5_K / q6_K have no exact kind::i8 kernel: always the kind::f16 GEMM
    const size_t hb  = ((size_t) N * K * 2 + 255) & ~(size_t) 255;   // fp16 activations [N][K] + 2^e[n]
    const size_t rb  = ((size_t) N * 4 + 255) & ~(size_t) 255;
    const size_t act_bytes = f16 ? hb + rb : aq + as * (has_min ? 2 : 1);
    size_t need = act_bytes;
    // ---- weights: cached or transient (transient copies live behind the activations in scratch2)
    WEntry * e = nullptr;
    WEntry   tmp;
    bool     cached = false, held = false;
    const size_t wb = wentry_bytes(type, M, K, Mp, f16);
    const size_t src_extent = (size_t) (M - 1) * nb01 + (size_t) (K / type_block_elems(type)) * type_block_bytes(type);
    bool cacheable = ctx->weight_cache != 0;
    uint8_t * cbuf = nullptr;
    if (ctx->weight_cache) {
        std::lock_guard<std::mutex> lock(g_wmutex);
        const WKey key{ ctx->device, w };
        auto it = g_wcache.find(key);
        if (it != g_wcache.end() && it->second->type == type && it->second->M == M && it->second->K == K && it->second->nb01 == nb01 &&
            it->second->f16 == f16) {
            e = it->second;
            e->refs++;
            e->stamp = ++g_wclock;
            cached = held = true;
        } else {
            if (it != g_wcache.end()) {  // geometry changed under the same address: drop the old expansion
                g_wbytes -= it->second->bytes;
                if (it->second->refs > 0) it->second->dead = true; else wentry_free(it->second, ctx->device);
                g_wcache.erase(it);
            }
            // a tensor that keeps being rewritten (the scheduler copies weights of CPU buffers in for every graph) is expanded per
            // call; so is one that does not fit the budget even after evicting everything idle, or when the allocation fails
            auto rw = g_wrewrites.find(key);
            cacheable = !(rw != g_wrewrites.end() && rw->second >= 2) && wcache_make_room(ctx->device, wb);
            if (cacheable) {
                cudaSetDevice(ctx->device);
                if (cudaMalloc((void **) &cbuf, wb) != cudaSuccess) {
                    cudaGetLastError();
                    cbuf = nullptr;
                    cacheable = false;
                } else {
                    g_wbytes += wb;   // reserved now, so that concurrent calls see the budget
                }
            }
        }
    }
    if (!e && !cacheable) need += wb;
    b200_status st = b200_ensure(ctx, &ctx->scratch2, &ctx->scratch2_size, need);
    if (st != B200_OK) {
        if (cbuf) { std::lock_guard<std::mutex> lock(g_wmutex); g_wbytes -= wb; cudaFree(cbuf); }
        if (held) wcache_release(e, ctx->device);
        return st;
    }
    uint8_t * s2 = (uint8_t *) ctx->scratch2;

    if (!e) {
        WEntry * ne = cacheable ? new WEntry() : &tmp;
        ne->type = type; ne->M = M; ne->K = K; ne->Mp = Mp; ne->nb01 = nb01; ne->bytes = wb; ne->f16 = f16; ne->extent = src_extent;
        uint8_t * buf = cacheable ? cbuf : s2 + act_bytes;
        st = wentry_fill(ctx, ne, w, buf);
        if (st != B200_OK) {
            if (cacheable) { std::lock_guard<std::mutex> lock(g_wmutex); g_wbytes -= wb; cudaFree(buf); delete ne; }
            return st;
        }
        if (cacheable) {
            // filling the cache is a load-time step: wait for it once, so that later launches may read the expansion ahead of
            // their stream dependency (not possible while the stream is being captured: such an entry stays unsettled)
            cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
            cudaStreamIsCapturing(ctx->stream, &cs);
            if (cs == cudaStreamCaptureStatusNone) {
                CU(cudaStreamSynchronize(ctx->stream));
                ne->settled = true;
            }
            std::lock_guard<std::mutex> lock(g_wmutex);
            const WKey key{ ctx->device, w };
            auto it = g_wcache.find(key);
            if (it != g_wcache.end()) {  // another thread expanded the same tensor meanwhile: keep ours, retire theirs
                g_wbytes -= it->second->bytes;
                if (it->second->refs > 0) it->second->dead = true; else wentry_free(it->second, ctx->device);
                g_wcache.erase(it);
            }
            ne->refs  = 1;
            ne->stamp = ++g_wclock;
            g_wcache[key] = ne;
            held = true;
        }
        e = ne;
    }
    struct Release { WEntry * e; int dev; bool on; ~Release() { if (on) wcache_release(e, dev); } } release{ e, ctx->device, held };
    (void) cached;
    if (f16) {
        const int kind = act_kind_for(type);
        const bool have_act = act_shared && ctx->act_x == x && ctx->act_rs == x_rs && ctx->act_N == N && ctx->act_K == K && ctx->act_buf == s2 &&
                              ctx->act_kind == kind;
        ctx->act_x = x; ctx->act_rs = x_rs; ctx->act_N = N; ctx->act_K = K; ctx->act_buf = s2; ctx->act_kind = kind;
        WEntry * es[1] = { e };
        float *  ds[1] = { dst };
        return mmf_run(ctx, 1, es, cached && e->settled, x, x_rs, N, (__half *) s2, (float *) (s2 + hb), ds, &ldc, have_act);
    }
    ctx->act_x = nullptr;

    int8_t * xq = (int8_t *) s2;
    float *  sA = (float *) (s2 + aq);
    float *  aS = has_min ? (float *) (s2 + aq + as) : sA;
    int16_t * bsums = (int16_t *) aS;  // q4_K planes: int16 block sums [N][K/256][8] instead of the float products (N*K/16 <= as bytes)
    {
        const int64_t units  = (K + 255) / 256;
        const int     vec_ok = ((uintptr_t) x % 16 == 0) && (x_rs % 16 == 0);
        for (int64_t r0 = 0; r0 < N; r0 += 65535) {
            const int64_t nr = N - r0 < 65535 ? N - r0 : 65535;
            dim3 grid((unsigned) ((units + 3) / 4), (unsigned) nr);
            const float * xs = (const float *) ((const char *) x + r0 * x_rs);
            if (has_min) quantize_tiles_kernel<ACT_KIND_Q8_K><<<grid, 128, 0, ctx->stream>>>(xs, K, nr, x_rs, xq + r0 * K, sA + r0, aS + r0, Np, vec_ok,
                                                                                               e->planes ? bsums + r0 * (K / 256) * 8 : nullptr);
            else         quantize_tiles_kernel<ACT_KIND_Q8_0><<<grid, 128, 0, ctx->stream>>>(xs, K, nr, x_rs, xq + r0 * K, sA + r0, aS + r0, Np, vec_ok);
            ctx->launches++;
        }
        CU(cudaGetLastError());
    }
    CUtensorMap map_b, map_sa, map_as;
    if (!make_map(&map_b, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, xq, (uint64_t) K, (uint64_t) N, (uint64_t) K, 128, BN, true) ||
        !make_map(&map_sa, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, sA, (uint64_t) N, (uint64_t) KB, (uint64_t) Np * 4, 128, BKB, false) ||
        !make_map(&map_as, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, aS, (uint64_t) N, (uint64_t) KB, (uint64_t) Np * 4, 128, BKB, false)) {
        return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled failed for the activation tiles (N=%lld K=%lld)", (long long) N, (long long) K);
    }

    MmqParams p;
    p.split = ctx->split;
    p.dst = dst; p.ldc = ldc; p.M = (int) M; p.N = (int) N; p.K = (int) K;
    p.m_tiles = (int) ((M + BM - 1) / BM);
    p.n_tiles = (int) ((N + BN - 1) / BN);
    const int tiles = p.m_tiles * p.n_tiles;
    const int grid  = tiles < ctx->sm_count ? tiles : ctx->sm_count;
    if (e->planes) {  // q4_K with the sub-block scales folded into two int8 planes
        CUtensorMap map_bs, map_da;
        if (!make_map(&map_bs, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, bsums, (uint64_t) (K / 256) * 8, (uint64_t) N, (uint64_t) (K / 256) * 16, 8, BN, false) ||
            !make_map(&map_da, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, sA, (uint64_t) N, (uint64_t) (K / 256), (uint64_t) Np * 4 * 8, 128, 1, false)) {
            return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled failed for the q8_K scale tiles (N=%lld K=%lld)", (long long) N, (long long) K);
        }
        static std::atomic<bool> kattr_done[64];
        if (ctx->device < 64 && !kattr_done[ctx->device]) {
            CU(cudaFuncSetAttribute(mmq_q4k_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, KQ_SMEM));
            kattr_done[ctx->device] = true;
        }
        mmq_q4k_kernel<<<grid, MMQ_THREADS, KQ_SMEM, ctx->stream>>>(e->map_q, e->map_qb, map_b, e->map_mn, map_bs, e->map_dw, e->map_sc, map_da, p);
        CU(cudaGetLastError());
        ctx->launches++;
        return B200_OK;
    }
    static std::atomic<bool> attr_done[2][64];   // zero-initialised; several host threads may launch
    if (ctx->device < 64 && !attr_done[has_min][ctx->device]) {
        if (has_min) CU(cudaFuncSetAttribute(mmq_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, MMQ_SMEM));
        else         CU(cudaFuncSetAttribute(mmq_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, MMQ_SMEM));
        attr_done[has_min][ctx->device] = true;
    }
    if (has_min) mmq_kernel<true><<<grid, MMQ_THREADS, MMQ_SMEM, ctx->stream>>>(e->map_q, map_b, e->map_sc, map_sa, e->map_mn, map_as, p);
    else         mmq_kernel<false><<<grid, MMQ_THREADS, MMQ_SMEM, ctx->stream>>>(e->map_q, map_b, e->map_sc, map_sa, e->map_mn, map_as, p);
    CU(cudaGetLastError());
    ctx->launches++;
    return B200_OK;
}

// Several weight matrices (same type and K) against the same N activation rows in ONE launch of the kind::f16 GEMM: one activation
// conversion, one pipeline fill / drain and one stream-K fix-up for q/k/v or gate/up. Only with the weight cache on (every
// matrix's fp16 expansion must be alive at the same time); returns B200_ERR_UNSUPPORTED when the caller should loop instead.
b200_status b200_mmq_run_group(b200_ctx * ctx, int type, int n_ops, const void * const * ws, const size_t * nb01s, const int64_t * Ms, int64_t K,
                               const float * x, size_t x_rs, int64_t N, float * const * dsts, const int64_t * ldcs) {
    if ((!ctx->mmq_f16 && type != B200_TYPE_Q6_K && type != B200_TYPE_Q5_K) || !ctx->weight_cache || n_ops < 2 || n_ops > F_MAX_MATS || N <= 0) return B200_ERR_UNSUPPORTED;
    for (int i = 0; i < n_ops; i++) if (Ms[i] <= 0 || Ms[i] > INT32_MAX - BM || N > INT32_MAX - BN || K > INT32_MAX - 128) return B200_ERR_UNSUPPORTED;
    static const int env_group = mmf_env("GGML_B200_MMF_GROUP", 1);
    if (!env_group) return B200_ERR_UNSUPPORTED;
    WEntry * es[F_MAX_MATS];
    bool     settled = true;
    for (int i = 0; i < n_ops; i++) {
        WEntry * e = nullptr;
        {
            std::lock_guard<std::mutex> lock(g_wmutex);
            auto it = g_wcache.find(WKey{ ctx->device, ws[i] });
            if (it != g_wcache.end() && it->second->type == type && it->second->M == Ms[i] && it->second->K == K && it->second->nb01 == nb01s[i] &&
                it->second->f16) {
                e = it->second;
                e->refs++;
                e->stamp = ++g_wclock;
            }
        }
        if (!e) {  // not expanded yet (or evicted): the per-matrix path fills the cache
            for (int j = 0; j < i; j++) wcache_release(es[j], ctx->device);
            return B200_ERR_UNSUPPORTED;
        }
        settled = settled && e->settled;
        es[i] = e;
    }
    struct ReleaseAll { WEntry ** es; int n, dev; ~ReleaseAll() { for (int j = 0; j < n; j++) wcache_release(es[j], dev); } } release_all{ es, n_ops, ctx->device };
    const size_t hb = ((size_t) N * K * 2 + 255) & ~(size_t) 255, rb = ((size_t) N * 4 + 255) & ~(size_t) 255;
    if (b200_status st = b200_ensure(ctx, &ctx->scratch2, &ctx->scratch2_size, hb + rb)) return st;
    uint8_t * s2 = (uint8_t *) ctx->scratch2;
    ctx->act_x = nullptr;
    return mmf_run(ctx, n_ops, es, settled, x, x_rs, N, (__half *) s2, (float *) (s2 + hb), dsts, ldcs, false);
}

// ------------------------------------------------------------------------------------------------ mvtc.cuh: host side

bool b200_mvtc_usable(const b200_ctx * ctx, int type, int64_t K, int64_t N, const void * w, size_t nb01) {
    if (ctx->tc_min_cols <= 0 || N < ctx->tc_min_cols || N > ctx->tc_max_cols || N > 32 || type != B200_TYPE_Q4_K) return false;
    if (K % 256 != 0 || (uintptr_t) w % 16 != 0 || nb01 % 16 != 0 || K > INT32_MAX - 256) return false;
    return get_encode() != nullptr;
}

// dst[n * ldc + m] = sum_k W[m][k] * X[n][k], N <= 32 activation rows, weights read as raw q4_K blocks.
// act_shared: the q8_K activation tiles of the previous call (same x, N, K: a shared-src1 group) are still in scratch2
b200_status b200_mvtc_run(b200_ctx * ctx, const void * w, size_t nb01, int64_t M, int64_t K, const float * x, size_t x_rs, int64_t N, float * dst,
                          int64_t ldc, bool act_shared) {
    if (M <= 0 || N <= 0) return B200_OK;
    const int64_t KB = K / 32, n_sb = K / 256, Np = (N + 3) & ~(int64_t) 3;
    const size_t aq = ((size_t) N * K + 255) & ~(size_t) 255;
    const size_t as = (((size_t) KB * Np * 4) + 255) & ~(size_t) 255;
    const size_t ab = (((size_t) N * n_sb * 16) + 255) & ~(size_t) 255;
    if (b200_status st = b200_ensure(ctx, &ctx->scratch2, &ctx->scratch2_size, aq + as + ab)) return st;
    uint8_t * s2 = (uint8_t *) ctx->scratch2;
    int8_t *  xq = (int8_t *) s2;
    float *   sA = (float *) (s2 + aq);
    int16_t * bsums = (int16_t *) (s2 + aq + as);
    if (!ctx->tc_turn) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing(ctx->stream, &cs);
        if (cs != cudaStreamCaptureStatusNone) return fail(B200_ERR_ARG, "first use of the tcgen05 mat-vec kernel during stream capture; run once uncaptured first");
        CU(cudaMalloc((void **) &ctx->tc_turn, 16384 * sizeof(uint32_t)));
        CU(cudaMemsetAsync(ctx->tc_turn, 0, 16384 * sizeof(uint32_t), ctx->stream));
    }
    ctx->act_x = nullptr;   // scratch2 no longer holds the fp16 operand of the kind::f16 path
    if (!(act_shared && ctx->tc_act_x == x && ctx->tc_act_rs == x_rs && ctx->tc_act_N == N && ctx->tc_act_K == K && ctx->tc_act_buf == s2)) {
        const int64_t units  = K / 256;
        const int     vec_ok = ((uintptr_t) x % 16 == 0) && (x_rs % 16 == 0);
        dim3 grid((unsigned) ((units + 3) / 4), (unsigned) N);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = grid; cfg.blockDim = dim3(128); cfg.stream = ctx->stream;
        CU(cudaLaunchKernelEx(&cfg, quantize_tiles_kernel<ACT_KIND_Q8_K>, x, K, N, x_rs, xq, sA, sA, Np, vec_ok, bsums));
        ctx->launches++;
        ctx->tc_act_x = x; ctx->tc_act_rs = x_rs; ctx->tc_act_N = N; ctx->tc_act_K = K; ctx->tc_act_buf = s2;
    }
    const int BN = N <= 16 ? 16 : 32;
    CUtensorMap map_w, map_b, map_bs, map_da;
    if (!make_map(&map_w, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, (void *) w, (uint64_t) n_sb * SZ_Q4_K, (uint64_t) M, (uint64_t) nb01, SZ_Q4_K, TC_BM, false) ||
        !make_map(&map_b, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, xq, (uint64_t) K, (uint64_t) N, (uint64_t) K, 128, BN, true) ||
        !make_map(&map_bs, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, bsums, (uint64_t) n_sb * 8, (uint64_t) N, (uint64_t) n_sb * 16, 8, BN, false) ||
        !make_map(&map_da, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, sA, (uint64_t) N, (uint64_t) n_sb, (uint64_t) Np * 4 * 8, BN, 1, false)) {
        return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled failed for the tcgen05 mat-vec operands (M=%lld K=%lld N=%lld nb01=%zu)", (long long) M,
                    (long long) K, (long long) N, nb01);
    }
    TcParams p;
    p.dst = dst; p.ldc = ldc; p.M = (int) M; p.N = (int) N; p.K = (int) K;
    p.m_tiles = (int) ((M + TC_BM - 1) / TC_BM);
    // cut K so that every SM has work: waves x (super-blocks per part + ~3 of pipeline fill) is what a launch costs
    int best = 1;
    if (p.m_tiles <= 16384) {
        double best_cost = 1e30;
        for (int s = 1; s <= 8 && s <= n_sb; s++) {
            const int per = (int) ((n_sb + s - 1) / s);
            const int parts = (int) ((n_sb + per - 1) / per);
            if (parts != s) continue;
            const int waves = (p.m_tiles * s + ctx->sm_count - 1) / ctx->sm_count;
            const double cost = (double) waves * (per + 3);
            if (cost < best_cost - 1e-9) { best_cost = cost; best = s; }
        }
    }
    p.splits = best;
    p.sb_per_split = (int) ((n_sb + best - 1) / best);
    p.count = ctx->tc_turn;
    p.dbg   = ctx->dbg;
    p.ws    = nullptr;
    if (p.splits > 1) {
        const int NRc = N <= 8 ? 8 : (N <= 16 ? 16 : 32);
        // scratch3 is shared with the kind::f16 GEMM's stream-K workspace: its first 4 KB are flags that stay zero between launches
        const size_t need3 = 4096 + (size_t) p.m_tiles * p.splits * NRc * TC_BM * sizeof(float);
        if (ctx->scratch3_size < need3) {
            if (b200_status st = b200_ensure(ctx, &ctx->scratch3, &ctx->scratch3_size, need3)) return st;
            CU(cudaMemsetAsync(ctx->scratch3, 0, 4096, ctx->stream));
        }
        p.ws = (float *) ((uint8_t *) ctx->scratch3 + 4096);
    }
    const int items = p.m_tiles * p.splits;
    const int grid  = items < ctx->sm_count ? items : ctx->sm_count;
    if (BN == 16) {
        if (N <= 8) return mvtc_launch<16, 8>(ctx, map_w, map_b, map_bs, map_da, p, grid);
        return mvtc_launch<16, 16>(ctx, map_w, map_b, map_bs, map_da, p, grid);
    }
    return mvtc_launch<32, 32>(ctx, map_w, map_b, map_bs, map_da, p, grid);
}


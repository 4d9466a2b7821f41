// mmq.cu — K2/K4: tiled int8 GEMM for prefill / batched decode (more than 8 activation columns) on the sm_100a
// 5th-generation tensor cores: tcgen05.mma kind::i8 (SASS UTCIMMA) with int32 accumulators in TMEM, operands staged by
// TMA (cp.async.bulk.tensor, SASS UTMALDG) into 128-byte-swizzled shared memory, per-32-block scales applied by the
// epilogue warps while they drain TMEM (tcgen05.ld, SASS LDTM).
//
// Replaces ggml_cuda_mul_mat_q / mul_mat_q + load_tiles_* + vec_dot_*_mma + mmq_write_back_mma + quantize_mmq_q8_1
// (/root/reference/ggml/src/ggml-cuda/mmq.cu:70-236, mmq.cuh:3287-3536, 277-336, 637-697, 1840-1947, 796-931, 965-1091,
// 3006-3053, quantize.cu:50-147). Not a port: the reference keeps 16x8 mma.sync fragments in registers and re-loads
// tiles synchronously (4 __syncthreads per 256-K step, mmq.cuh:3241-3275); here one thread issues 128xBNx32 MMAs, the
// copy engine feeds a 4-stage ring, and 8 epilogue warps do nothing but rescale (double-buffered TMEM drain, packed fp32 math).
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
// Warp roles (384 threads): warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer, warps 4..11 epilogue
// (warp % 4 = TMEM lane quadrant = 32 weight rows, (warp - 4) / 4 = half of the tile's 128 activation columns).

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
constexpr int EPI_WARPS = 8, FIRST_EPI_WARP = 4;   // 2 warps per TMEM lane quadrant, 64 accumulator columns per thread
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
                                                             int64_t Np, int vec_ok) {
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
            aS[kb * Np + n] = r.d * (float) r.s32;
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
        // A thread owns one weight row (TMEM lane) and 64 of the tile's 128 columns, drained as chunks of 32 columns.
        // The tcgen05.ld of chunk i+1 is in flight while chunk i is converted and scaled (I2FP + FMUL2 + FFMA2).
        const int quad = warp & 3, half = (warp - FIRST_EPI_WARP) >> 2;
        const int row_in_tile = quad * 32 + lane;
        const uint32_t t_lane = tmem_base + ((uint32_t) (quad * 32) << 16) + half * 64;
        int stage = 0, slot = 0; uint32_t phase = 0, tphase = 0;
        for (int t = blockIdx.x; t < n_tiles_total; t += gridDim.x) {
            const int m0 = (t / p.n_tiles) * BM, n0 = (t % p.n_tiles) * BN;
            float2 acc[32];
#pragma unroll
            for (int j = 0; j < 32; j++) acc[j] = make_float2(0.0f, 0.0f);
            const int n_chunks = k_stages * BKB * 2;
            uint32_t ca[32], cb[32];
            // chunk 0
            mbar_wait(smem_u32(tfull + slot), tphase);
            tc_fence_after();
            tmem_ld_32x32(t_lane + slot * BN, ca);
            tc_wait_ld_regs(ca);

            auto math = [&](const uint32_t (&c)[32], int kb, int h, uint32_t st) {
                const float sw = __uint_as_float(lds32(st + 4 * (kb * 128 + row_in_tile)));
                const float2 sw2 = make_float2(sw, sw);
                const uint32_t sa = st + S_BYTES + 4 * (kb * 128 + half * 64 + h * 32);
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
                    float2 & r0 = acc[h * 16 + 2 * j4], & r1 = acc[h * 16 + 2 * j4 + 1];
                    r0 = ffma2(p0, make_float2((float) (int) c[4 * j4 + 0], (float) (int) c[4 * j4 + 1]), r0);
                    r1 = ffma2(p1, make_float2((float) (int) c[4 * j4 + 2], (float) (int) c[4 * j4 + 3]), r1);
                    if (HAS_MIN) {
                        const uint4 b = lds128(sa + 2 * S_BYTES + 16 * j4);
                        r0 = ffma2(mw2, make_float2(__uint_as_float(b.x), __uint_as_float(b.y)), r0);
                        r1 = ffma2(mw2, make_float2(__uint_as_float(b.z), __uint_as_float(b.w)), r1);
                    }
                }
            };

            // chunks are processed in pairs (h = 0, 1 of one 32-block) so that the two register buffers alternate statically
            for (int ck = 0; ck < n_chunks; ck += 2) {
                const int kb = (ck >> 1) & (BKB - 1);
                if (kb == 0) mbar_wait(smem_u32(full + stage), phase);  // the scales of this stage have landed
                const uint32_t st = smem_u32(smem + stage * STAGE_BYTES + A_BYTES + B_BYTES);
                // second half of this 32-block in flight while the first is scaled
                tmem_ld_32x32(t_lane + slot * BN + 32, cb);
                math(ca, kb, 0, st);
                tc_wait_ld_regs(cb);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(tempty + slot));  // both halves are in registers: the next MMA may overwrite the slot
                if (++slot == SLOTS) { slot = 0; tphase ^= 1; }
                // first half of the next 32-block in flight while the second half of this one is scaled
                const bool more = ck + 2 < n_chunks;
                if (more) {
                    mbar_wait(smem_u32(tfull + slot), tphase);
                    tc_fence_after();
                    tmem_ld_32x32(t_lane + slot * BN, ca);
                }
                math(cb, kb, 1, st);
                if (more) tc_wait_ld_regs(ca);
                if (kb == BKB - 1) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(smem_u32(empty + stage));  // done with this stage's scales
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
            const int m = m0 + row_in_tile;
            if (m < p.M) {
#pragma unroll
                for (int j = 0; j < 32; j++) {
                    const int n = n0 + half * 64 + 2 * j;
                    if (n < p.N) split_store(p.split, p.dst + (int64_t) n * p.ldc + m, acc[j].x);
                    if (n + 1 < p.N) split_store(p.split, p.dst + (int64_t) (n + 1) * p.ldc + m, acc[j].y);
                }
            }
        }
    }

    if (p.split.n_ranks > 1) __threadfence_system();
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
};

struct WKey {
    int          device;
    const void * src;
    bool operator<(const WKey & o) const { return device != o.device ? device < o.device : src < o.src; }
};

std::mutex                 g_wmutex;
std::map<WKey, WEntry *>   g_wcache;
size_t                     g_wbytes = 0;

size_t wentry_bytes(int type, int64_t M, int64_t K, int64_t Mp) {
    const size_t qb = ((size_t) M * K + 255) & ~(size_t) 255;
    const size_t sb = (((size_t) (K / 32) * Mp * 4) + 255) & ~(size_t) 255;
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
    const char * lo = (const char *) base, * hi = lo + size;
    for (auto it = g_wcache.begin(); it != g_wcache.end();) {
        const char * s = (const char *) it->first.src;
        if (it->first.device == device && s >= lo && s < hi) {
            cudaSetDevice(device);
            cudaFree(it->second->buf);  // synchronizes with the device: no kernel is still reading it
            g_wbytes -= it->second->bytes;
            delete it->second;
            it = g_wcache.erase(it);
        } else {
            ++it;
        }
    }
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
    if (type == B200_TYPE_Q4_K && ((uintptr_t) w % 16 != 0 || nb01 % 16 != 0)) return false;
    if ((uintptr_t) w % 2 != 0 || nb01 % 2 != 0) return false;
    return get_encode() != nullptr;
}

// dst[n*ldc + m] = sum_k W[m][k] * X[n][k] for one weight matrix and N activation rows
b200_status b200_mmq_run(b200_ctx * ctx, int type, const void * w, size_t nb01, int64_t M, int64_t K, const float * x, size_t x_rs, int64_t N,
                         float * dst, int64_t ldc) {
    if (M <= 0 || N <= 0) return B200_OK;
    if (M > INT32_MAX - BM || N > INT32_MAX - BN || K > INT32_MAX - 128) return fail(B200_ERR_UNSUPPORTED, "MMQ shape too large");
    const int64_t KB = K / 32;
    const int64_t Mp = (M + 3) & ~(int64_t) 3, Np = (N + 3) & ~(int64_t) 3;
    const bool    has_min = type == B200_TYPE_Q4_K;

    // ---- activations: int8 [N][K] + scales [KB][Np] (+ sums) in scratch2
    const size_t aq = ((size_t) N * K + 255) & ~(size_t) 255;
    const size_t as = (((size_t) KB * Np * 4) + 255) & ~(size_t) 255;
    size_t need = aq + as * (has_min ? 2 : 1);
    // ---- weights: cached or transient (transient copies live behind the activations in scratch2)
    WEntry * e = nullptr;
    WEntry   tmp;
    bool     cached = false;
    if (ctx->weight_cache) {
        std::lock_guard<std::mutex> lock(g_wmutex);
        auto it = g_wcache.find(WKey{ ctx->device, w });
        if (it != g_wcache.end() && it->second->type == type && it->second->M == M && it->second->K == K && it->second->nb01 == nb01) {
            e = it->second;
            cached = true;
        }
    }
    const size_t wb = wentry_bytes(type, M, K, Mp);
    if (!e && !ctx->weight_cache) need += wb;
    b200_status st = b200_ensure(ctx, &ctx->scratch2, &ctx->scratch2_size, need);
    if (st != B200_OK) return st;
    uint8_t * s2 = (uint8_t *) ctx->scratch2;

    if (!e) {
        WEntry * ne = ctx->weight_cache ? new WEntry() : &tmp;
        ne->type = type; ne->M = M; ne->K = K; ne->Mp = Mp; ne->nb01 = nb01; ne->bytes = wb;
        uint8_t * buf = nullptr;
        if (ctx->weight_cache) {
            cudaError_t ce = cudaMalloc((void **) &buf, wb);
            if (ce != cudaSuccess) {
                cudaGetLastError();
                delete ne;
                return fail(B200_ERR_ALLOC, "weight cache: cudaMalloc(%zu): %s", wb, cudaGetErrorString(ce));
            }
        } else {
            buf = s2 + aq + as * (has_min ? 2 : 1);
        }
        st = wentry_fill(ctx, ne, w, buf);
        if (st != B200_OK) {
            if (ctx->weight_cache) { cudaFree(buf); delete ne; }
            return st;
        }
        if (ctx->weight_cache) {
            std::lock_guard<std::mutex> lock(g_wmutex);
            auto it = g_wcache.find(WKey{ ctx->device, w });
            if (it != g_wcache.end()) {  // geometry changed under the same address: replace
                cudaFree(it->second->buf);
                g_wbytes -= it->second->bytes;
                delete it->second;
                g_wcache.erase(it);
            }
            g_wcache[WKey{ ctx->device, w }] = ne;
            g_wbytes += wb;
        }
        e = ne;
    }
    (void) cached;

    int8_t * xq = (int8_t *) s2;
    float *  sA = (float *) (s2 + aq);
    float *  aS = has_min ? (float *) (s2 + aq + as) : sA;
    {
        const int64_t units  = (K + 255) / 256;
        const int     vec_ok = ((uintptr_t) x % 16 == 0) && (x_rs % 16 == 0);
        for (int64_t r0 = 0; r0 < N; r0 += 65535) {
            const int64_t nr = N - r0 < 65535 ? N - r0 : 65535;
            dim3 grid((unsigned) ((units + 3) / 4), (unsigned) nr);
            const float * xs = (const float *) ((const char *) x + r0 * x_rs);
            if (has_min) quantize_tiles_kernel<ACT_KIND_Q8_K><<<grid, 128, 0, ctx->stream>>>(xs, K, nr, x_rs, xq + r0 * K, sA + r0, aS + r0, Np, vec_ok);
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
    static bool attr_done[2][64] = {};
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

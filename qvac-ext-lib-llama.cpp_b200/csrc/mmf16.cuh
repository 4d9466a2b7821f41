// mmf16.cuh — K5: prefill GEMM on tcgen05.mma kind::f16 (SASS UTCHMMA) with ONE fp32 accumulation over the whole K in TMEM.
// Included by mmq.cu (inside its anonymous namespace: shares make_map and the weight cache).
//
// Why a second tensor-core path: the kind::i8 kernels of mmq.cu keep the reference's per-32-block structure (an int32 dot
// product per block, scaled in fp32 by the epilogue warps). That epilogue costs M*N*K/32 I2F + FMUL + FFMA and is what bounds
// them (profiles/r01_pass7_mmq_q4k_full.md: imma pipe 26 %). Here the block scales are folded into the OPERANDS instead:
//   A[m][k] = fp16( dequantize_row_<type>(W)[m][k] )                     (ggml-quants.c:307-325, 359-372, 1352-1374)
//   B[n][k] = fp16( 2^-e[n] * d_a[n][k/32|k/256] * q_a[n][k] )           (q_a, d_a: bit-exact quantize_row_q8_0_ref / q8_K_ref)
//   dst[n][m] = 2^e[n] * sum_k A[m][k] * B[n][k]                         (fp32 accumulation in TMEM)
// which is the same sum the CPU reference forms (vec_dot_q4_0_q8_0 / q8_0_q8_0 / q4_K_q8_K factor the block scales out of it),
// with two roundings to 11 significant bits per product (relative 2^-12 each) in place of exact integer block sums: NMSE
// against the oracle is about 1e-7 (tests/test_gpu_parity.py states the bound), north_star allows 5e-4. The per-row power of
// two 2^e[n] brings every activation row's largest magnitude into [2^14, 2^15): no fp16 overflow or subnormals whatever the
// activation range, and a power of two commutes with the rounding. GGML_B200_MMQ_F16=0 (b200_ctx_set_mmq_f16) selects the
// exact-block-sum kind::i8 kernels instead.
//
// Tiling: a CTA PAIR (tcgen05.mma cta_group::2, SASS UTCHMMA.2CTA) owns a 256 (M) x 256 (N) tile: each CTA loads its 128 weight rows
// and HALF of the activation tile (16 + 16 KB per K stage of 64 halfs = one 128-byte swizzle row), the leader's elected thread issues
// four 256 x 256 x 16 MMAs per stage, both CTAs keep 128 lanes x 256 columns of the fp32 accumulator in their TMEM (2 slots: the
// epilogue of one work item overlaps the main loop of the next). Up to 128 columns (N <= 128) and with GGML_B200_MMF_CL=1 a single
// CTA owns a 128 x BN tile. The SM's operand ingest (~50 B/clk), not the L2 read rate, is what a single-CTA 128 x 256 tile is short
// of: a cluster that merely multicasts the activation tile is no faster, the pair that SPLITS it is (755 -> 805 TOPS).
// Scheduling is stream-K: the launch's tiles x K-stages (tiles of up to four weight matrices that share src1: q/k/v, gate/up) are
// cut into equal contiguous spans, one per CTA (pair),
// so every SM is busy for the same time whatever the tile count (wk/wv of Llama-3-8B at n = 512 are 8 tiles). A tile cut by
// span boundaries is finished deterministically in two phases: every CTA parks the partial accumulator of its part in a
// workspace and bumps a counter; at the end of its span each of the P CTAs that share the tile waits for the P parts and sums
// all of them in ascending K order for ITS 1/P of the tile's columns (parts fetched into the idle operand ring by bulk copies),
// scales and stores that share (no atomics on data: results do not depend on timing; the fix-up of a tile cut P ways runs on P SMs -
// a single owner summing P x 128 KB took 28 us on wk).
// Warp roles (384 threads): warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer (leader CTA of a pair), warps 4..11 epilogue.

#include <cuda_fp16.h>
#include <type_traits>

constexpr int F_BM = 128, F_BK = 64;
constexpr int F_EPI_WARPS = 8, F_FIRST_EPI = 4;
constexpr int F_THREADS = 32 * (F_FIRST_EPI + F_EPI_WARPS);
constexpr int F_A_BYTES = F_BM * 128;
__host__ __device__ constexpr int f_b_bytes(int bn) { return bn * 128; }
__host__ __device__ constexpr int f_stages(int bn, int cl = 1) { return bn / cl == 256 ? 4 : 6; }
__host__ __device__ constexpr int f_smem(int bn, int cl = 1) { return 1024 + f_stages(bn, cl) * (F_A_BYTES + f_b_bytes(bn / cl)) + 256; }

// ---- instruction descriptor for kind::f16: D = f32 ([4,6) = 1), A and B fp16 ([7,10) = [10,13) = 0), both K-major
__host__ __device__ constexpr uint32_t umma_idesc_f16(int M, int N) { return (1u << 4) | ((uint32_t) (N >> 3) << 17) | ((uint32_t) (M >> 4) << 24); }

__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// ---- CTA pair (cta_group::2): one tcgen05.mma spans two SMs - M = 256 (128 rows of A from each CTA), N = 256 with each CTA holding
// half of the B rows - so an SM ingests 16 + 16 KB per K stage instead of 16 + 32 KB for the same 128 x 256 x 64 MACs
template <int CL>
__device__ __forceinline__ void umma_f16_cl(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    if (CL == 1) {
        umma_f16(d_tmem, a_desc, b_desc, idesc, accumulate);
    } else {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
            "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
            : "memory");
    }
}
template <int CL>
__device__ __forceinline__ void umma_commit_cl(uint32_t bar) {  // CL == 2: arrives on the barrier at this offset in BOTH CTAs of the pair
    if (CL == 1) {
        umma_commit(bar);
    } else {
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"((uint16_t) 3)
                     : "memory");
    }
}
// TMA load whose completion bytes go to the barrier `bar` (a shared::cluster address: the LEADER CTA's barrier for a pair)
template <int CL>
__device__ __forceinline__ void tma_load_2d_cl(uint32_t dst, const void * tmap, int c0, int c1, uint32_t bar) {
    if (CL == 1) {
        tma_load_2d(dst, tmap, c0, c1, bar);
    } else {
        asm volatile(
            "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
            "l"(tmap), "r"(c0), "r"(c1), "r"(bar)
            : "memory");
    }
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t mapa_cta(uint32_t addr, uint32_t cta) {  // the same shared-memory offset in CTA `cta` of the cluster
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(cta));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
template <int NCOLS>
__device__ __forceinline__ void tmem_alloc2(uint32_t smem_dst) {  // one full warp of each CTA of the pair
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "n"(NCOLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int NCOLS>
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(NCOLS) : "memory");
}

__device__ __forceinline__ void tma_prefetch_l2_2d(const void * tmap, int c0, int c1) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(tmap), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void st_release_gpu(unsigned * p, unsigned v) { asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned * p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long mmf_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) :: "memory");
    return t;
}

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *(const uint32_t *) &h;
}

// ------------------------------------------------------------------------------------------------ weights -> fp16 [M][K]

// one thread per 32-element block: the values dequantize_row_q4_0 / q8_0 / q4_K produce (fp32), rounded to fp16
template <int TYPE>
__global__ void __launch_bounds__(256) expand_f16_kernel(const uint8_t * __restrict__ w, int64_t nb01, int64_t M, int64_t K, __half * __restrict__ h) {
    const int64_t KB  = K / 32;
    const int64_t idx = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= M * KB) return;
    const int64_t row = idx / KB, kb = idx - row * KB;
    const uint8_t * wr = w + row * nb01;
    float v[32];
    if (TYPE == B200_TYPE_Q4_0) {
        const uint8_t * b = wr + kb * SZ_Q4_0;
        const float d = half_bits_to_float(*(const uint16_t *) b);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t x = ld16x2(b + 2 + 4 * i);
#pragma unroll
            for (int j = 0; j < 4; j++) {
                v[4 * i + j]      = (float) ((int) ((x >> (8 * j)) & 0xf) - 8) * d;
                v[16 + 4 * i + j] = (float) ((int) ((x >> (8 * j + 4)) & 0xf) - 8) * d;
            }
        }
    } else if (TYPE == B200_TYPE_Q8_0) {
        const uint8_t * b = wr + kb * SZ_Q8_0;
        const float d = half_bits_to_float(*(const uint16_t *) b);
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const uint32_t x = ld16x2(b + 2 + 4 * i);
#pragma unroll
            for (int j = 0; j < 4; j++) v[4 * i + j] = (float) (int) (int8_t) (x >> (8 * j)) * d;
        }
    } else if (TYPE == B200_TYPE_Q6_K) {
        // 32-element block kb of the row = quarter (j & 3) of half (j >> 2) of a 256-weight q6_K block (2-byte aligned: byte loads)
        const uint8_t * b = wr + (kb >> 3) * SZ_Q6_K;
        const int j = (int) (kb & 7), hf = j >> 2, qv = j & 3;
        const uint8_t * ql = b + 64 * hf + 32 * (qv & 1), * qh = b + 128 + 32 * hf;
        const int8_t *  sc = (const int8_t *) (b + 192) + 8 * hf + 2 * qv;
        const float d = half_bits_to_float(*(const uint16_t *) (b + 208));
#pragma unroll
        for (int l = 0; l < 32; l++) {
            const int lo = (qv & 2) ? (ql[l] >> 4) : (ql[l] & 0xF);
            const int q  = (lo | (((qh[l] >> (2 * qv)) & 3) << 4)) - 32;
            v[l] = __fmul_rn(__fmul_rn(d, (float) (int) sc[l >> 4]), (float) q);  // d * sc * q, left to right (ggml-quants.c:1780-1783)
        }
    } else {
        constexpr bool Q5 = TYPE == B200_TYPE_Q5_K;  // q4_K's layout with 32 bytes of fifth bits behind the header (dequantize_row_q5_K)
        const uint8_t * b = wr + (kb >> 3) * (Q5 ? SZ_Q5_K : SZ_Q4_K);
        const int j = (int) (kb & 7);
        const uint32_t hdr = *(const uint32_t *) b;
        const float d = half_bits_to_float((uint16_t) (hdr & 0xffff)), dmin = half_bits_to_float((uint16_t) (hdr >> 16));
        const uint8_t * s = b + 4;
        int scv, mv;
        if (j < 4) {
            scv = s[j] & 63;
            mv  = s[j + 4] & 63;
        } else {
            scv = (s[j + 4] & 0xF) | ((s[j - 4] >> 6) << 4);
            mv  = (s[j + 4] >> 4) | ((s[j] >> 6) << 4);
        }
        const float d1 = __fmul_rn(d, (float) scv), m1 = __fmul_rn(dmin, (float) mv);  // dequantize_row_q4_K: y = d1 * q - m1
        const uint4 * qs = (const uint4 *) (b + (Q5 ? 48 : 16) + 32 * (j >> 1));
        const uint4 a = qs[0], c = qs[1];
        const uint32_t x[8] = { a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w };
        uint32_t hb[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
        if (Q5) {
            const uint4 ha = ((const uint4 *) (b + 16))[0], hc = ((const uint4 *) (b + 16))[1];
            const uint32_t hw[8] = { ha.x, ha.y, ha.z, ha.w, hc.x, hc.y, hc.z, hc.w };
#pragma unroll
            for (int i = 0; i < 8; i++) hb[i] = hw[i] >> j;  // bit j of qh[l] -> bit 0 of byte l
        }
        const int sh = (j & 1) ? 4 : 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
#pragma unroll
            for (int l = 0; l < 4; l++) {
                const int q = (int) ((x[i] >> (8 * l + sh)) & 0xf) + (int) (((hb[i] >> (8 * l)) & 1) << 4);
                v[4 * i + l] = __fsub_rn(__fmul_rn(d1, (float) q), m1);
            }
        }
    }
    uint4 * dst = (uint4 *) (h + row * K + kb * 32);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        dst[i] = make_uint4(pack_h2(v[8 * i], v[8 * i + 1]), pack_h2(v[8 * i + 2], v[8 * i + 3]), pack_h2(v[8 * i + 4], v[8 * i + 5]),
                            pack_h2(v[8 * i + 6], v[8 * i + 7]));
    }
}

// ------------------------------------------------------------------------------------------------ activations -> fp16 [N][K]

// one CTA per activation row: (1) the row's largest magnitude -> the power of two that normalises it, (2) per 256-element unit
// the bit-exact q8_0 / q8_K quantization, dequantized (q * d), pre-scaled and rounded to fp16. rs[n] = 2^e[n] for the epilogue.
// UPW > 0: a warp keeps its UPW units of the row in registers between the two steps (rows of up to 2048 * UPW floats: one
// trip to memory); UPW = 0: any row length, the row is read twice.
template <int KIND, int UPW>
__global__ void __launch_bounds__(256) act_f16_kernel(const float * __restrict__ x, int64_t K, size_t x_rs, __half * __restrict__ h,
                                                      float * __restrict__ rs, int vec_ok) {
    __shared__ float s_amax[8];
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // the GEMM's set-up may start; it waits for this grid before it reads h
    asm volatile("griddepcontrol.wait;" ::: "memory");               // x is the previous kernel's output; h / rs are still read by the previous GEMM
    const int     lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t n     = blockIdx.x;
    const int64_t units = (K + UNIT_ELEMS - 1) / UNIT_ELEMS;
    const float * xr = (const float *) ((const char *) x + n * x_rs);
    constexpr int NV = UPW > 0 ? UPW : 1;

    auto load = [&](int64_t unit, float (&v)[8]) -> bool {
        const int64_t e0 = unit * UNIT_ELEMS + lane * 8;
        if (unit >= units || e0 >= K) {
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = 0.0f;
            return false;
        }
        if (vec_ok) {
            const float4 a = __ldg((const float4 *) (xr + e0));
            const float4 b = __ldg((const float4 *) (xr + e0) + 1);
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
            v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = __ldg(xr + e0 + i);
        }
        return true;
    };

    float keep[NV][8];
    bool  act[NV];
    float amax = 0.0f;
    if (UPW > 0) {
#pragma unroll
        for (int i = 0; i < NV; i++) {
            act[i] = load(warp + 8 * i, keep[i]);
#pragma unroll
            for (int j = 0; j < 8; j++) amax = fmaxf(amax, fabsf(keep[i][j]));  // NaN-ignoring, like the quantizers' own amax
        }
    } else {
        for (int64_t u = warp; u < units; u += 8) {
            load(u, keep[0]);
#pragma unroll
            for (int j = 0; j < 8; j++) amax = fmaxf(amax, fabsf(keep[0][j]));
        }
    }
    amax = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(amax)));
    if (lane == 0) s_amax[warp] = amax;
    __syncthreads();
    amax = s_amax[0];
#pragma unroll
    for (int i = 1; i < 8; i++) amax = fmaxf(amax, s_amax[i]);
    // 2^e with amax * 2^-e in [2^14, 2^15); e clamped so that both powers stay normal floats
    int e = (int) ((__float_as_uint(amax) >> 23) & 0xff) - 127 - 14;
    if (amax == 0.0f || !(amax < INFINITY)) e = 0;
    e = max(-100, min(100, e));
    const float pre = __uint_as_float((uint32_t) (127 - e) << 23);
    if (threadIdx.x == 0) rs[n] = __uint_as_float((uint32_t) (127 + e) << 23);

    auto emit = [&](int64_t u, const float (&v)[8], bool active) {
        uint32_t w0, w1;
        float    d;
        if (KIND == ACT_KIND_Q8_0) {
            const Q80Lane r = quant_lane_q8_0(v);
            w0 = r.w0; w1 = r.w1;
            d  = __half2float(__float2half_rn(r.d));  // the dot product reads the stored fp16 scale
        } else {
            const Q8KLane r = quant_lane_q8_K(v, lane);
            w0 = r.w0; w1 = r.w1;
            d  = r.d;
        }
        if (!active) return;
        float f[8];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            f[i]     = __fmul_rn(__fmul_rn((float) (int) (int8_t) (w0 >> (8 * i)), d), pre);
            f[4 + i] = __fmul_rn(__fmul_rn((float) (int) (int8_t) (w1 >> (8 * i)), d), pre);
        }
        *(uint4 *) (h + n * K + u * UNIT_ELEMS + lane * 8) = make_uint4(pack_h2(f[0], f[1]), pack_h2(f[2], f[3]), pack_h2(f[4], f[5]), pack_h2(f[6], f[7]));
    };
    if (UPW > 0) {
#pragma unroll
        for (int i = 0; i < NV; i++) {
            if (warp + 8 * i < units) emit(warp + 8 * i, keep[i], act[i]);  // warp-uniform
        }
    } else {
        for (int64_t u = warp; u < units; u += 8) {
            const bool active = load(u, keep[0]);
            emit(u, keep[0], active);
        }
    }
}

// ------------------------------------------------------------------------------------------------ the GEMM (K5)

constexpr int F_MAX_MATS = 4;  // weight matrices of one launch (MUL_MATs that read the same src1: q/k/v, gate/up)
struct MmfMaps {
    CUtensorMap a[F_MAX_MATS];
};
struct MmfParams {
    SplitParams   split;
    float *       dst[F_MAX_MATS];
    int64_t       ldc[F_MAX_MATS];    // floats between dst columns (tokens)
    int           Mi[F_MAX_MATS];     // rows of matrix i
    int           tile0[F_MAX_MATS];  // first tile of matrix i in the launch's tile order (INT_MAX for unused slots)
    int           n_mats;
    const float * rs;   // [N] 2^e[n]
    int           N, K;
    int           n_tiles, k_stages;
    long long     units;   // tiles x k_stages
    float *       ws;      // [gridDim.x][2][BN][128] partial accumulators of cut tiles (slot 1: the K-range that starts at 0)
    unsigned *    flags;   // [h] parts of the cut tile headed by CTA h that are parked; [512 + h] CTAs that have summed their share of it
    unsigned long long * trace;  // debug timeline (b200_ctx_trace): 8 x u64 per CTA
    int           pf;      // L2 prefetch distance of the weight tiles in stages (0 = off)
    int           early_a;      // the weight tiles may be requested before the stream dependency resolves
    int           whole_tiles;  // spans end on tile boundaries (no cut tiles; GGML_B200_MMF_STREAMK=0)
};

#define MMF_TRACE(i) do { if (p.trace) p.trace[(size_t) blockIdx.x * 8 + (i)] = mmf_ns(); } while (0)

template <int BN, int CL>
__global__ void __launch_bounds__(F_THREADS, 1)
mmf_kernel(const __grid_constant__ MmfMaps maps, const __grid_constant__ CUtensorMap map_b, const __grid_constant__ MmfParams p) {
    constexpr int STAGES = f_stages(BN, CL), SLOTS = 2;
    constexpr int B_ROWS = BN / CL;  // rows of the activation tile this CTA loads (a pair: half each)
    constexpr int B_BYTES = f_b_bytes(B_ROWS), STAGE_BYTES = F_A_BYTES + B_BYTES;
    constexpr int TILE_M = F_BM * CL;
    const int       crank = CL == 2 ? (int) cluster_ctarank() : 0;                 // 0 = the CTA of a pair that issues the MMAs
    const int       bid   = CL == 2 ? (int) (blockIdx.x >> 1) : (int) blockIdx.x;  // this CTA's (pair's) place in the schedule
    const long long G     = CL == 2 ? (long long) (gridDim.x >> 1) : (long long) gridDim.x;
    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
    uint64_t * bars      = (uint64_t *) (smem + STAGES * STAGE_BYTES);
    uint64_t * full      = bars;               // [STAGES]  TMA -> MMA
    uint64_t * empty     = bars + STAGES;      // [STAGES]  MMA commit -> TMA
    uint64_t * tfull     = bars + 2 * STAGES;  // [SLOTS]   MMA commit -> epilogue
    uint64_t * tempty    = tfull + SLOTS;      // [SLOTS]   epilogue warps -> MMA
    uint32_t * tmem_slot = (uint32_t *) (tempty + SLOTS);
    uint64_t * fbar      = tempty + SLOTS + 1;  // parts of a cut tile -> shared memory (phase 2 of the fix-up)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KS = p.k_stages;
    // this CTA's span of the launch's tile-major (tile, K-stage) units
    const long long u_begin = p.whole_tiles ? (p.units / KS * bid / G) * KS : p.units * bid / G;
    const long long u_end   = p.whole_tiles ? (p.units / KS * (bid + 1) / G) * KS : p.units * (bid + 1) / G;

    if (threadIdx.x == 0) {
        MMF_TRACE(0);
        for (int i = 0; i < STAGES; i++) {
            mbar_init(smem_u32(full + i), 1);
            mbar_init(smem_u32(empty + i), 1);
        }
        for (int i = 0; i < SLOTS; i++) {
            mbar_init(smem_u32(tfull + i), 1);
            mbar_init(smem_u32(tempty + i), F_EPI_WARPS * CL);  // a pair: the leader's barrier counts the epilogue warps of both CTAs
        }
        mbar_init(smem_u32(fbar), 1);
        fence_mbar_init();
        for (int i = 0; i < p.n_mats; i++) tma_prefetch_desc(&maps.a[i]);
        tma_prefetch_desc(&map_b);
    }
    if (warp == 1) {
        if (CL == 2) tmem_alloc2<SLOTS * BN>(smem_u32(tmem_slot));
        else tmem_alloc<SLOTS * BN>(smem_u32(tmem_slot));
    }
    tc_fence_before();
    if (CL == 2) cluster_sync_all();  // the peer's barriers are initialised before anything arrives on them
    else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (threadIdx.x == 0) MMF_TRACE(1);

    // one (tile, K-stage) unit after the other of this CTA's span
    struct Cursor {
        long long u;
        int       ks, k1, m0, n0, mat;
    };
    auto mat_of = [&](int tile) {  // which matrix of the launch a tile belongs to
        int i = 0;
#pragma unroll
        for (int j = 1; j < F_MAX_MATS; j++) if (tile >= p.tile0[j]) i = j;
        return i;
    };
    auto open_item = [&](Cursor & c) {
        const int tile = (int) (c.u / KS);
        c.ks = (int) (c.u - (long long) tile * KS);
        c.k1 = (int) min((long long) KS, c.ks + (u_end - c.u));
        c.mat = mat_of(tile);
        const int lt = tile - p.tile0[c.mat];
        c.m0 = (lt / p.n_tiles) * TILE_M + crank * F_BM;  // this CTA's 128 weight rows of the tile
        c.n0 = (lt % p.n_tiles) * BN + crank * B_ROWS;    // and the activation rows it loads
    };
    auto step = [&](Cursor & c) {
        c.u++;
        if (++c.ks == c.k1 && c.u < u_end) open_item(c);
    };

    if (warp == 0) {
        // ================================================================ TMA producer
        if (lane == 0) {
            Cursor c;
            c.u = u_begin;
            if (c.u < u_end) open_item(c);
            // the weights do not depend on the stream: the first ring of weight tiles is requested before the dependency resolves
            if (!p.early_a) asm volatile("griddepcontrol.wait;" ::: "memory");  // the fp16 weights were written on this stream just now
            Cursor ca = c;
            int npre = 0;
            // completion bytes of both CTAs of a pair go to the leader's barrier, which expects the bytes of both
            auto full_bar = [&](int s_) { return CL == 2 ? mapa_cta(smem_u32(full + s_), 0) : smem_u32(full + s_); };
            for (; npre < STAGES && ca.u < u_end; npre++, step(ca)) {
                if (crank == 0) mbar_expect_tx(smem_u32(full + npre), CL * STAGE_BYTES);
                tma_load_2d_cl<CL>(smem_u32(smem + npre * STAGE_BYTES), &maps.a[ca.mat], ca.ks * F_BK, ca.m0, full_bar(npre));
            }
            if (p.early_a) asm volatile("griddepcontrol.wait;" ::: "memory");  // the activations are the previous kernel's output
            for (int i = 0; i < npre; i++, step(c)) {
                tma_load_2d_cl<CL>(smem_u32(smem + i * STAGE_BYTES) + F_A_BYTES, &map_b, c.ks * F_BK, c.n0, full_bar(i));
            }
            int stage = npre == STAGES ? 0 : npre; uint32_t phase = npre == STAGES ? 0 : 1;
            for (; c.u < u_end; step(c)) {
                if (p.pf > 0 && c.ks + p.pf < c.k1) tma_prefetch_l2_2d(&maps.a[c.mat], (c.ks + p.pf) * F_BK, c.m0);
                mbar_wait(smem_u32(empty + stage), phase);
                const uint32_t bar = full_bar(stage);
                const uint32_t st  = smem_u32(smem + stage * STAGE_BYTES);
                if (crank == 0) mbar_expect_tx(smem_u32(full + stage), CL * STAGE_BYTES);
                tma_load_2d_cl<CL>(st, &maps.a[c.mat], c.ks * F_BK, c.m0, bar);
                tma_load_2d_cl<CL>(st + F_A_BYTES, &map_b, c.ks * F_BK, c.n0, bar);
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ================================================================ MMA issuer: four UTCHMMA (128 x BN x 16) per stage
        if (lane == 0 && crank == 0) {
            constexpr uint32_t idesc = umma_idesc_f16(TILE_M, BN);
            int stage = 0, slot = 0; uint32_t phase = 0, tphase = 1;
            bool first = true;
            for (long long u = u_begin; u < u_end;) {
                const int tile = (int) (u / KS), k0 = (int) (u - (long long) tile * KS);
                const int k1 = (int) min((long long) KS, k0 + (u_end - u));
                mbar_wait(smem_u32(tempty + slot), tphase);  // the epilogue has drained this accumulator
                tc_fence_after();
                for (int ks = k0; ks < k1; ks++) {
                    mbar_wait(smem_u32(full + stage), phase);
                    tc_fence_after();
                    if (first && ks == k0) MMF_TRACE(3);
                    const uint32_t a0 = smem_u32(smem + stage * STAGE_BYTES), b0 = a0 + F_A_BYTES;
#pragma unroll
                    for (int k = 0; k < F_BK / 16; k++) {
                        umma_f16_cl<CL>(tmem_base + slot * BN, umma_desc_sw128(a0 + k * 32), umma_desc_sw128(b0 + k * 32), idesc, (uint32_t) ((ks - k0) | k));
                    }
                    umma_commit_cl<CL>(smem_u32(empty + stage));
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit_cl<CL>(smem_u32(tfull + slot));
                if (first) { MMF_TRACE(5); first = false; }
                if (++slot == SLOTS) { slot = 0; tphase ^= 1; }
                u += k1 - k0;
            }
        }
    } else if (warp >= F_FIRST_EPI) {
        // ================================================================ epilogue: TMEM -> registers -> * 2^e[n] -> dst, or -> workspace
        // a thread owns one weight row (TMEM lane); the two warps of a lane quadrant take half of the tile's columns each
        const int quad = warp & 3, half = (warp - F_FIRST_EPI) >> 2;
        const int row  = quad * 32 + lane;
        const int te   = (warp - F_FIRST_EPI) * 32 + lane;  // 0 .. 255
        // the accumulator is free again once the epilogue warps of both CTAs of a pair have drained their halves
        auto release_slot = [&](int slot_) {
            if (CL == 2) mbar_arrive_cluster(mapa_cta(smem_u32(tempty + slot_), 0));
            else mbar_arrive(smem_u32(tempty + slot_));
        };
        int slot = 0; uint32_t tphase = 0;
        bool first = true;
        int  cut_tile[2], n_cut = 0;  // the (at most two) tiles of this span that other CTAs share: finished in phase 2
        asm volatile("griddepcontrol.wait;" ::: "memory");  // rs, the workspace, its counters and dst belong to earlier kernels until here
        for (long long u = u_begin; u < u_end;) {
            const int tile = (int) (u / KS), k0 = (int) (u - (long long) tile * KS);
            const int k1 = (int) min((long long) KS, k0 + (u_end - u));
            const int mi = mat_of(tile), lt = tile - p.tile0[mi];
            const int m0 = (lt / p.n_tiles) * TILE_M + crank * F_BM, n0 = (lt % p.n_tiles) * BN;
            const int m  = m0 + row;
            mbar_wait(smem_u32(tfull + slot), tphase);
            tc_fence_after();
            const uint32_t t_lane = tmem_base + ((uint32_t) (quad * 32) << 16) + slot * BN + half * (BN / 2);
            if (k0 > 0 || k1 < KS) {
                // part of a tile that is cut by span boundaries: park the accumulator (slot 1 of this CTA for the K-range that starts
                // at 0, slot 0 otherwise - a later range is always the first item of its CTA) and count it on the tile's head CTA
                float * wsp = p.ws + (((size_t) (2 * blockIdx.x + (k0 == 0 ? 1 : 0)) * BN + half * (BN / 2)) * F_BM + row);
#pragma unroll 1
                for (int c0 = 0; c0 < BN / 2; c0 += 32) {
                    uint32_t c[32];
                    tmem_ld_32x32(t_lane + c0, c);
                    tc_wait_ld_regs(c);
#pragma unroll
                    for (int j = 0; j < 32; j++) __stcg(wsp + (size_t) (c0 + j) * F_BM, __uint_as_float(c[j]));
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) release_slot(slot);
                asm volatile("fence.proxy.async;" ::: "memory");  // the parts are read back with bulk copies (async proxy)
                __threadfence();
                asm volatile("bar.sync 1, %0;" ::"n"(32 * F_EPI_WARPS) : "memory");
                if (warp == F_FIRST_EPI && lane == 0) {
                    const int h = (int) ((((long long) tile * KS + 1) * G + p.units - 1) / p.units) - 1;  // the CTA (pair) whose span holds the tile's first stage
                    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p.flags + (CL * h + crank)) : "memory");
                }
                cut_tile[n_cut++] = tile;
            } else {
#pragma unroll 1
                for (int c0 = 0; c0 < BN / 2; c0 += 32) {
                    uint32_t c[32];
                    tmem_ld_32x32(t_lane + c0, c);
                    tc_wait_ld_regs(c);
                    const int   nb  = n0 + half * (BN / 2) + c0;
                    const float rsv = nb + lane < p.N ? __ldg(p.rs + nb + lane) : 0.0f;
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; j++) v[j] = __uint_as_float(c[j]) * __shfl_sync(0xffffffffu, rsv, j);
                    if (m < p.Mi[mi]) split_store_cols32(p.split, p.dst[mi] + (int64_t) nb * p.ldc[mi] + m, p.ldc[mi], v, p.N - nb);
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) release_slot(slot);
            }
            if (first) first = false;
            if (++slot == SLOTS) { slot = 0; tphase ^= 1; }
            u += k1 - k0;
        }
        if (warp == F_FIRST_EPI && lane == 0) MMF_TRACE(6);  // every item of the span drained (stored or parked)
        // ---- phase 2: every CTA that holds a part of a cut tile sums ALL of the tile's parts (ascending K: fixed order, no atomics on
        // data) for its own share of the tile's columns and stores that share - the fix-up of a tile cut P ways runs on P SMs
        for (int r = 0; r < n_cut; r++) {
            const int tile = cut_tile[r];
            const int mi = mat_of(tile), lt = tile - p.tile0[mi];
            const int m0 = (lt / p.n_tiles) * TILE_M + crank * F_BM, n0 = (lt % p.n_tiles) * BN;
            const int     M_ = p.Mi[mi];
            const int64_t ldc_ = p.ldc[mi];
            float * const dst_ = p.dst[mi];
            const int h  = (int) ((((long long) tile * KS + 1) * G + p.units - 1) / p.units) - 1;
            const int l  = (int) ((((long long) (tile + 1) * KS) * G + p.units - 1) / p.units) - 1;
            const int P  = l - h + 1, me = bid - h;
            const int hc = CL * h + crank;  // a pair: the two CTAs reduce their own 128 rows with the same-rank CTAs of the other pairs
            if (lane == 0) {
                long long spins = 0;
                while (ld_acquire_gpu(p.flags + hc) < (unsigned) P) {
                    if (++spins > (1ll << 24)) { printf("mmf_kernel: CTA %d timed out waiting for the parts of tile %d\n", blockIdx.x, tile); __trap(); }
                }
            }
            __syncwarp();
            if (r == 0 && warp == F_FIRST_EPI && lane == 0) MMF_TRACE(2);  // the parts of the first cut tile are all parked
            const int c_lo = BN * me / P, c_hi = BN * (me + 1) / P;
            const int n_pos = (c_hi - c_lo) * 32;  // positions of 4 consecutive rows (16 bytes) of one column
            const bool vec = p.split.n_ranks <= 1 && (ldc_ & 3) == 0 && ((uintptr_t) dst_ & 15) == 0;
            const size_t   slot_floats = (size_t) BN * F_BM;
            const uint32_t part_bytes  = (uint32_t) (c_hi - c_lo) * F_BM * 4;  // this CTA's columns of one part: contiguous in the workspace
            // The ring is idle by now (every TMA load has landed and every MMA that read it has completed): the P x part_bytes
            // (= one tile, 128 KB) are fetched into it with P bulk copies - the copy engine hides the L2 latency that bounded
            // this phase when the threads loaded the parts themselves (4 us per tile with 24 x 16 B in flight per thread)
            if (warp == F_FIRST_EPI && lane == 0) {
                asm volatile("fence.proxy.async;" ::: "memory");
                const uint64_t pol = l2_policy_evict_first();
                mbar_expect_tx(smem_u32(fbar), (uint32_t) P * part_bytes);
                for (int j = 0; j < P; j++) {
                    const int sl = j == 0 ? 2 * hc + 1 : 2 * (hc + CL * j);  // ascending K: the head's slot 1, then slot 0 of the followers
                    bulk_g2s(smem_u32(smem) + (uint32_t) j * part_bytes, p.ws + (size_t) sl * slot_floats + (size_t) c_lo * F_BM, part_bytes, smem_u32(fbar), pol);
                }
            }
            mbar_wait(smem_u32(fbar), (uint32_t) (r & 1));
            for (int pos = te; pos < n_pos; pos += 32 * F_EPI_WARPS) {
                const uint32_t sp = smem_u32(smem) + (uint32_t) pos * 16;
                float4 acc;
                {
                    const uint4 v = lds128(sp);
                    acc = make_float4(__uint_as_float(v.x), __uint_as_float(v.y), __uint_as_float(v.z), __uint_as_float(v.w));
                }
                for (int j = 1; j < P; j++) {
                    const uint4 v = lds128(sp + (uint32_t) j * part_bytes);
                    acc.x += __uint_as_float(v.x); acc.y += __uint_as_float(v.y); acc.z += __uint_as_float(v.z); acc.w += __uint_as_float(v.w);
                }
                const int n = n0 + c_lo + (pos >> 5), m = m0 + 4 * (pos & 31);
                if (n >= p.N || m >= M_) continue;
                const float rsv = __ldg(p.rs + n);
                acc.x *= rsv; acc.y *= rsv; acc.z *= rsv; acc.w *= rsv;
                float * d = dst_ + (int64_t) n * ldc_ + m;
                if (vec && m + 3 < M_) {
                    *(float4 *) d = acc;
                } else {
                    split_store(p.split, d, acc.x);
                    if (m + 1 < M_) split_store(p.split, d + 1, acc.y);
                    if (m + 2 < M_) split_store(p.split, d + 2, acc.z);
                    if (m + 3 < M_) split_store(p.split, d + 3, acc.w);
                }
            }
            // the last of the P readers lowers the counters for the next launch
            asm volatile("bar.sync 1, %0;" ::"n"(32 * F_EPI_WARPS) : "memory");
            if (warp == F_FIRST_EPI && lane == 0) {
                if (atomicAdd(p.flags + 512 + hc, 1u) == (unsigned) (P - 1)) {
                    p.flags[hc]       = 0u;
                    p.flags[512 + hc] = 0u;
                }
            }
        }
        if (warp == F_FIRST_EPI && lane == 0) MMF_TRACE(7);
    }

    tc_fence_before();
    if (CL == 2) cluster_sync_all();  // the leader's MMAs read the peer's shared memory and the peer arrives on the leader's barriers
    else __syncthreads();
    if (threadIdx.x == 0) MMF_TRACE(4);  // every role of this CTA is through its span
    if (p.split.n_ranks > 1 && warp == 0) split_complete(p.split, gridDim.x);
    if (warp == 1) {
        tc_fence_after();
        if (CL == 2) tmem_dealloc2<SLOTS * BN>(tmem_base);
        else tmem_dealloc<SLOTS * BN>(tmem_base);
    }
}

// ------------------------------------------------------------------------------------------------ host side

// ------------------------------------------------------------------------------------------------ row split: slice -> peers

// Row split with more than 8 columns: the GEMM stores this rank's row slice into the LOCAL dst only, and this kernel then copies
// the slice into every peer's dst with 16-byte stores, coalesced along the rows, before the flag exchange (split_complete).
// Peer stores issued from the GEMM epilogue itself (4-byte stores, 32 per thread and rank, from 8 of 12 warps) ran 8x below the
// NVLink rate on 8 GPUs (profiles/README.md, pass 10).
struct ScatterParams {
    SplitParams split;
    float *     dst[F_MAX_MATS];
    int64_t     ldc[F_MAX_MATS];
    int         M[F_MAX_MATS];
    int         n_mats, N;
};
template <bool VEC>
__global__ void __launch_bounds__(512) split_scatter_kernel(const ScatterParams p) {
    const long long tid = (long long) blockIdx.x * blockDim.x + threadIdx.x, T = (long long) gridDim.x * blockDim.x;
    for (int i = 0; i < p.n_mats; i++) {
        const int       per_row = VEC ? p.M[i] / 4 : p.M[i];
        const long long total   = (long long) per_row * p.N;
        for (long long e = tid; e < total; e += T) {
            const long long n = e / per_row;
            const int       c = (int) (e - n * per_row);
            if (VEC) {
                float4 * a = (float4 *) (p.dst[i] + n * p.ldc[i]) + c;
                const float4 v = *a;
                for (int r = 0; r < p.split.n_ranks; r++) if (r != p.split.rank) *(float4 *) ((char *) a + p.split.dst_delta[r]) = v;
            } else {
                float * a = p.dst[i] + n * p.ldc[i] + c;
                const float v = *a;
                for (int r = 0; r < p.split.n_ranks; r++) if (r != p.split.rank) *(float *) ((char *) a + p.split.dst_delta[r]) = v;
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < 32) split_complete(p.split, gridDim.x);
}

int mmf_env(const char * name, int dflt) {
    const char * s = getenv(name);
    return s && *s ? atoi(s) : dflt;
}

template <int BN, int CL>
b200_status mmf_launch(b200_ctx * ctx, const MmfMaps & maps, const CUtensorMap & map_b, MmfParams & p, int tiles) {
    static std::atomic<bool> attr_done[64];
    if (ctx->device < 64 && !attr_done[ctx->device]) {
        CU(cudaFuncSetAttribute(mmf_kernel<BN, CL>, cudaFuncAttributeMaxDynamicSharedMemorySize, f_smem(BN, CL)));
        attr_done[ctx->device] = true;
    }
    static const int env_sk = mmf_env("GGML_B200_MMF_STREAMK", 1), env_min = mmf_env("GGML_B200_MMF_MIN_SPAN", 8);
    p.units = (long long) tiles * p.k_stages;
    const int slots = ctx->sm_count / CL;  // CTAs (CL == 2: CTA pairs) that run at once
    int grid;
    if (env_sk) {  // equal spans of at least env_min stages
        const long long g = p.units / (env_min > 0 ? env_min : 1);
        grid = (int) (g < 1 ? 1 : g < slots ? g : slots);
    } else {       // whole tiles only
        grid = tiles < slots ? tiles : slots;
    }
    grid *= CL;
    if (grid > 512) grid = 512;  // the counters of the cut tiles are indexed by CTA
    p.whole_tiles = !env_sk;
    // workspace: [counters: 4 KB (parts parked per head CTA, readers done per head CTA)][2 slots per CTA][BN][128] floats
    const size_t need = 4096 + (size_t) 2 * ctx->sm_count * 256 * F_BM * 4;
    if (ctx->scratch3_size < need) {
        b200_status st = b200_ensure(ctx, &ctx->scratch3, &ctx->scratch3_size, need);
        if (st != B200_OK) return st;
        CU(cudaMemsetAsync(ctx->scratch3, 0, 4096, ctx->stream));
    }
    p.flags = (unsigned *) ctx->scratch3;
    p.ws    = (float *) ((uint8_t *) ctx->scratch3 + 4096);
    p.trace = ctx->trace ? ctx->trace + (size_t) (ctx->launches % ctx->trace_slots) * ctx->sm_count * 8 : nullptr;
    cudaLaunchConfig_t  cfg = {};
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (CL == 2) {
        attr[na].id               = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = 2;
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = 1;
        na++;
    }
    if (ctx->pdl) {
        attr[na].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        na++;
    }
    cfg.gridDim          = dim3((unsigned) grid);
    cfg.blockDim         = dim3(F_THREADS);
    cfg.dynamicSmemBytes = f_smem(BN, CL);
    cfg.stream           = ctx->stream;
    cfg.attrs            = attr;
    cfg.numAttrs         = na;
    CU(cudaLaunchKernelEx(&cfg, mmf_kernel<BN, CL>, maps, map_b, p));
    ctx->launches++;
    return B200_OK;
}

// dst_i[n*ldc_i + m] = sum_k W_i[m][k] * X[n][k] for the n_mats matrices es[i] (fp16 expansions, same type and K) in ONE launch:
// h / rs are scratch for the fp16 activations and 2^e[n]
// have_act: h / rs already hold this src1 (an earlier matrix of the same shared-src1 group built them): no quantizer launch
b200_status mmf_run(b200_ctx * ctx, int n_mats, WEntry * const * es, bool settled, const float * x, size_t x_rs, int64_t N, __half * h, float * rs,
                    float * const * dsts, const int64_t * ldcs, bool have_act) {
    const int64_t K = es[0]->K;
    const int     vec_ok = ((uintptr_t) x % 16 == 0) && (x_rs % 16 == 0);
    for (int64_t r0 = 0; r0 < N && !have_act; r0 += 1 << 30) {
        const int64_t nr = N - r0 < (1 << 30) ? N - r0 : (1 << 30);
        const float * xs = (const float *) ((const char *) x + r0 * x_rs);
        cudaLaunchConfig_t  cfg = {};
        cudaLaunchAttribute attr[1];
        attr[0].id                                         = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.gridDim  = dim3((unsigned) nr);
        cfg.blockDim = dim3(256);
        cfg.stream   = ctx->stream;
        cfg.attrs    = attr;
        cfg.numAttrs = ctx->pdl ? 1 : 0;
        const int64_t units = (K + 255) / 256;
        const bool    k8    = type_is_k(es[0]->type);
#define ACT_LAUNCH(UPW) (k8 ? cudaLaunchKernelEx(&cfg, act_f16_kernel<ACT_KIND_Q8_K, UPW>, xs, K, x_rs, h + r0 * K, rs + r0, vec_ok) \
                            : cudaLaunchKernelEx(&cfg, act_f16_kernel<ACT_KIND_Q8_0, UPW>, xs, K, x_rs, h + r0 * K, rs + r0, vec_ok))
        // rows kept in registers between the two steps up to `env_reg` units; longer rows are read twice (the second read hits L2):
        // 8 units per warp cost 117 registers - two CTAs per SM, two waves for 512 rows
        static const int env_reg = mmf_env("GGML_B200_MMF_ACT_REG", 64);
        if (units <= 16 && units <= env_reg)      CU(ACT_LAUNCH(2));
        else if (units <= 32 && units <= env_reg) CU(ACT_LAUNCH(4));
        else if (units <= 64 && units <= env_reg) CU(ACT_LAUNCH(8));
        else                                      CU(ACT_LAUNCH(0));
#undef ACT_LAUNCH
        ctx->launches++;
    }

    static const int env_bn = mmf_env("GGML_B200_MMF_BN", 0), env_pf = mmf_env("GGML_B200_MMF_PF", 0), env_cl = mmf_env("GGML_B200_MMF_CL", 2);
    const int bn = env_bn ? env_bn : (N <= 128 ? 128 : 256);
    const int cl = bn == 256 && env_cl == 2 ? 2 : 1;  // CTA pairs (cta_group::2): 256 x 256 tiles over two SMs
    MmfParams p;
    MmfMaps   maps;
    memset(&maps, 0, sizeof(maps));
    p.split = ctx->split;
    p.rs = rs; p.N = (int) N; p.K = (int) K;
    p.n_tiles  = (int) ((N + bn - 1) / bn);
    p.k_stages = (int) ((K + F_BK - 1) / F_BK);
    p.pf       = env_pf;
    p.early_a  = settled;
    p.n_mats   = n_mats;
    int tiles = 0;
    for (int i = 0; i < F_MAX_MATS; i++) {
        if (i < n_mats) {
            maps.a[i]  = es[i]->map_q;
            p.dst[i]   = dsts[i];
            p.ldc[i]   = ldcs[i];
            p.Mi[i]    = (int) es[i]->M;
            p.tile0[i] = tiles;
            tiles += (int) ((es[i]->M + F_BM * cl - 1) / (F_BM * cl)) * p.n_tiles;
        } else {
            p.dst[i] = nullptr; p.ldc[i] = 0; p.Mi[i] = 0; p.tile0[i] = INT32_MAX;
        }
    }
    CUtensorMap map_b;
    if (!make_map(&map_b, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, h, (uint64_t) K, (uint64_t) N, (uint64_t) K * 2, F_BK, (uint32_t) (bn / cl), true)) {
        return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled failed for the fp16 activations (N=%lld K=%lld)", (long long) N, (long long) K);
    }
    static const int env_scatter = mmf_env("GGML_B200_MMF_SCATTER", 1);
    const bool scatter = ctx->split.n_ranks > 1 && env_scatter;
    if (scatter) p.split = SplitParams{};  // the GEMM writes the local dst only
    b200_status st = cl == 2     ? mmf_launch<256, 2>(ctx, maps, map_b, p, tiles)
                     : bn == 256 ? mmf_launch<256, 1>(ctx, maps, map_b, p, tiles)
                                 : mmf_launch<128, 1>(ctx, maps, map_b, p, tiles);
    if (st != B200_OK || !scatter) return st;
    ScatterParams sp;
    sp.split = ctx->split;
    sp.n_mats = n_mats; sp.N = (int) N;
    bool vec = true;
    for (int i = 0; i < F_MAX_MATS; i++) {
        sp.dst[i] = i < n_mats ? dsts[i] : nullptr;
        sp.ldc[i] = i < n_mats ? ldcs[i] : 0;
        sp.M[i]   = i < n_mats ? (int) es[i]->M : 0;
        if (i < n_mats) vec = vec && es[i]->M % 4 == 0 && ldcs[i] % 4 == 0 && (uintptr_t) dsts[i] % 16 == 0;
    }
    for (int r = 0; r < ctx->split.n_ranks; r++) vec = vec && ctx->split.dst_delta[r] % 16 == 0;
    if (vec) split_scatter_kernel<true><<<ctx->sm_count, 512, 0, ctx->stream>>>(sp);
    else     split_scatter_kernel<false><<<ctx->sm_count, 512, 0, ctx->stream>>>(sp);
    CU(cudaGetLastError());
    ctx->launches++;
    return B200_OK;
}

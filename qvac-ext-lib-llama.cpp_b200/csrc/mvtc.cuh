// mvtc.cuh — batched decode / small prefill (4..32 activation columns) for q4_K on the 5th-generation tensor cores, straight from
// the RAW 144-byte blocks: no expanded copy of the weights anywhere in HBM. Included by mmq.cu (shares its TMA / tcgen05 plumbing
// and the q8_K activation tiles of quantize_tiles_kernel).
//
// Why: with 2..8 columns the warp-level mma.sync kernel (mvmma.cu) is issue-bound at ~20 warp instructions per 256 weights (0.44 of
// HBM), the dp4a kernels cost ~6 per column. tcgen05.mma does the dot products of 128 rows x 16..32 columns x 32 weights in ONE
// asynchronous instruction with the accumulator in TMEM, so what the CUDA cores still have to do per 256 weights is
//   * expand 144 raw bytes into 256 int8 (one ld.shared.v4 + masks + two st.shared.v4 per 32 weights)      ~5 warp instr.
//   * one integer multiply-add per (row, 32-block, column) while draining TMEM, two fp32 FMAs per 256       ~0.5 per column
// MEASURED (profiles/r02/mvtc_full.md): 3.8 TB/s on the 296 MB lm_head at 8 columns - the design is bound by SHARED-MEMORY bandwidth, not
// by instructions: per raw byte the SM moves ~5.6 bytes through shared memory (TMA write 1, expander read 1, int8 tile write 1.8,
// tensor-core operand read >= 1.8), 44 % LSU wavefront utilisation plus the async-proxy traffic. It beats the other paths at 9..16
// columns (no fp16 weight cache needed there either) and is used for those; see DESIGN.md 4.3c.
// The integers are ggml-cpu's (exact sub-block dot products, 6-bit scales applied
// in int32, block sums x mins with dp2a), the activations are the bit-exact q8_K values: NMSE vs the oracle stays in the 1e-13 class.
//
// CTA = 12 warps: warp 0 TMA producer (raw weight rows, activation tile, q8_K scales / block sums), warp 1 TMEM allocator + MMA
// issuer, warps 4..7 epilogue (TMEM lane quadrant = warp % 4; thread = weight row), warps 8..11 expanders (thread = weight row).
// Per 256-weight super-block and stage s (all rings have depth TC_STAGES and advance together):
//   raw[s]  128 rows x 144 B           TMA -> expander            rawfull / rawempty
//   exp[s]  2 x (128 rows x 128 B)     expander -> UMMA A operand (K-major, 128-byte swizzle)   expfull / expempty (commit)
//   b[s]    2 x (BN rows x 128 B)      TMA -> UMMA B operand                                    bfull / bempty (commit)
//   side[s] 128 x 16 B block headers (expander), BN floats d_a, BN x 8 int16 block sums (TMA)   sidefull / sideempty (epilogue)
//   TMEM slot t = sb % 2: 8 accumulators of BN columns, one per 32-block                        tfull (commit) / tempty (epilogue)
// Work item = (128-row tile, K range): small matrices are cut along K so that every SM has work; the parts of a tile are added into
// dst in part order (a per-tile turn counter), so results do not depend on timing.
#pragma once

// (included inside mmq.cu's anonymous namespace)

constexpr int TC_BM      = 128;
constexpr int TC_ES      = 2;                    // expanded-operand stages
constexpr int TC_THREADS = 32 * 12;
constexpr int TC_RAW     = TC_BM * SZ_Q4_K;      // 18432
constexpr int TC_EXP     = 2 * TC_BM * 128;      // 32768
__host__ __device__ constexpr int tc_b_bytes(int BN) { return 2 * BN * 128; }
// a raw stage: activation tile (1024-byte aligned for the 128-byte swizzle), raw weight rows, q8_K scales d_a (128 B), block sums
__host__ __device__ constexpr int tc_rstage_bytes(int BN) { return (tc_b_bytes(BN) + TC_RAW + 128 + BN * 16 + 1023) / 1024 * 1024; }
// the raw ring is what covers the HBM latency (a stage is recycled only when the epilogue is through with it: ~4000 cycles after
// its TMA was issued): as deep as shared memory allows
__host__ __device__ constexpr int tc_rstages(int BN) { return (226 * 1024 - 1024 - TC_ES * TC_EXP - 512) / tc_rstage_bytes(BN); }
__host__ __device__ constexpr int tc_smem(int BN) { return 1024 + TC_ES * TC_EXP + tc_rstages(BN) * tc_rstage_bytes(BN) + 512; }

struct TcParams {
    float *    dst;
    int64_t    ldc;          // floats between dst columns
    int        M, N, K;
    int        m_tiles, splits, sb_per_split;   // work items = m_tiles * splits; item i = (tile i / splits, part i % splits)
    uint32_t * count;        // [m_tiles] parts of a tile that have parked their partial result (self-resetting)
    float *    ws;           // [items][NR][128] partial results of split tiles
    int        dbg;          // timing experiments (GGML_B200_DBG): 8 = half of the MMAs, 16 = expander without its stores, 32 = epilogue without the scale loop
};

__device__ __forceinline__ void tmem_ld_32x8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// BN: MMA width (16 or 32 activation columns, zero rows beyond N come from the TMA out-of-bounds fill); NR: columns the epilogue
// actually drains (8, 16 or 32)
template <int BN, int NR>
__global__ void __launch_bounds__(TC_THREADS, 1)
mvtc_q4k_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_b, const __grid_constant__ CUtensorMap map_bs,
                const __grid_constant__ CUtensorMap map_da, const TcParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
    constexpr int RS      = tc_rstages(BN);
    constexpr int RSTAGE  = tc_rstage_bytes(BN);
    constexpr int OFF_RAW = tc_b_bytes(BN);          // inside a raw stage: [activation tile][raw rows][d_a][block sums]
    constexpr int OFF_DA  = OFF_RAW + TC_RAW;
    constexpr int OFF_BS  = OFF_DA + 128;
    uint8_t *  ering     = smem;                     // TC_ES expanded stages
    uint8_t *  rring     = smem + TC_ES * TC_EXP;    // RS raw stages
    uint64_t * bars      = (uint64_t *) (rring + RS * RSTAGE);
    uint64_t * rfull     = bars,               * rempty = bars + RS;
    uint64_t * efull     = bars + 2 * RS,      * eempty = efull + TC_ES;
    uint64_t * tfull     = eempty + TC_ES,     * tempty = tfull + 2;
    uint32_t * tmem_slot = (uint32_t *) (tempty + 2);
    uint32_t * last_flag = tmem_slot + 1;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_items = p.m_tiles * p.splits;
    const int n_sb    = p.K / 256;

    if (threadIdx.x == 0) {
        for (int i = 0; i < RS; i++) { mbar_init(smem_u32(rfull + i), 1); mbar_init(smem_u32(rempty + i), 9); }  // 4 expander + 4 epilogue warps + the MMA commit
        for (int i = 0; i < TC_ES; i++) { mbar_init(smem_u32(efull + i), 4); mbar_init(smem_u32(eempty + i), 1); }
        for (int i = 0; i < 2; i++) { mbar_init(smem_u32(tfull + i), 1); mbar_init(smem_u32(tempty + i), 4); }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<512>(smem_u32(tmem_slot));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    auto item_range = [&](int item, int & tile, int & part, int & sb0, int & sb1) {
        tile = item / p.splits;
        part = item - tile * p.splits;
        sb0  = part * p.sb_per_split;
        sb1  = min(n_sb, sb0 + p.sb_per_split);
    };

    if (warp == 0) {
        // ================================================================ TMA producer
        if (lane == 0) {
            tma_prefetch_desc(&map_w); tma_prefetch_desc(&map_b); tma_prefetch_desc(&map_bs); tma_prefetch_desc(&map_da);
            asm volatile("griddepcontrol.wait;" ::: "memory");  // the activation tiles are written by the quantizer launch in front of this one
            int rs = 0; uint32_t ph = 1;   // "empty" barriers: the first pass over the ring falls through
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                int tile, part, sb0, sb1;
                item_range(item, tile, part, sb0, sb1);
                for (int sb = sb0; sb < sb1; sb++) {
                    const uint32_t st = smem_u32(rring + rs * RSTAGE), bar = smem_u32(rfull + rs);
                    mbar_wait(smem_u32(rempty + rs), ph);
                    mbar_expect_tx(bar, TC_RAW + tc_b_bytes(BN) + BN * 4 + BN * 16);
                    tma_load_2d(st + OFF_RAW, &map_w, sb * SZ_Q4_K, tile * TC_BM, bar);
                    tma_load_2d(st, &map_b, sb * 256, 0, bar);
                    tma_load_2d(st + BN * 128, &map_b, sb * 256 + 128, 0, bar);
                    tma_load_2d(st + OFF_DA, &map_da, 0, sb, bar);
                    tma_load_2d(st + OFF_BS, &map_bs, sb * 8, 0, bar);
                    if (++rs == RS) { rs = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ================================================================ MMA issuer: 8 x UTCIMMA (128 x BN x 32) per super-block
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_i8(TC_BM, BN);
            int rs = 0, es = 0, t = 0; uint32_t rph = 0, eph = 0, tph = 1;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                int tile, part, sb0, sb1;
                item_range(item, tile, part, sb0, sb1);
                for (int sb = sb0; sb < sb1; sb++) {
                    mbar_wait(smem_u32(tempty + t), tph);
                    mbar_wait(smem_u32(rfull + rs), rph);
                    mbar_wait(smem_u32(efull + es), eph);
                    tc_fence_after();
                    const uint32_t a0 = smem_u32(ering + es * TC_EXP), b0 = smem_u32(rring + rs * RSTAGE);
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        if ((p.dbg & 8) && j >= 4) continue;
                        umma_i8(tmem_base + (uint32_t) (t * 8 * BN + j * BN), umma_desc_sw128(a0 + (j >> 2) * (TC_BM * 128) + (j & 3) * 32),
                                umma_desc_sw128(b0 + (j >> 2) * (BN * 128) + (j & 3) * 32), idesc, 0);
                    }
                    umma_commit(smem_u32(eempty + es));
                    umma_commit(smem_u32(rempty + rs));
                    umma_commit(smem_u32(tfull + t));
                    if (++rs == RS) { rs = 0; rph ^= 1; }
                    if (++es == TC_ES) { es = 0; eph ^= 1; }
                    if (++t == 2) { t = 0; tph ^= 1; }
                }
            }
        }
    } else if (warp >= 8) {
        // ================================================================ expanders: thread = weight row; 144 raw bytes -> 256 int8
        const int r = (warp - 8) * 32 + lane;
        const uint32_t sw = (uint32_t) (r & 7);
        int rs = 0, es = 0; uint32_t rph = 0, eph = 1;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int tile, part, sb0, sb1;
            item_range(item, tile, part, sb0, sb1);
            for (int sb = sb0; sb < sb1; sb++) {
                const uint32_t raw = smem_u32(rring + rs * RSTAGE) + OFF_RAW + (uint32_t) r * SZ_Q4_K;
                const uint32_t row = smem_u32(ering + es * TC_EXP) + (uint32_t) r * 128;
                mbar_wait(smem_u32(rfull + rs), rph);
                mbar_wait(smem_u32(eempty + es), eph);
#pragma unroll
                for (int i = 0; i < 4; i++) {  // 64 weights: bytes l = 0..31; low nibbles = elements 64 i + l, high nibbles = 64 i + 32 + l
                    if (p.dbg & 16) continue;
                    const uint4 qa = lds128(raw + 16 + 32 * i), qb = lds128(raw + 32 + 32 * i);
                    const uint32_t tl = row + (uint32_t) (i >> 1) * (TC_BM * 128);   // elements 0..127 -> first operand tile, 128..255 -> second
                    const uint32_t c0 = (uint32_t) ((4 * i) & 7);
                    sts128(tl + (((c0 + 0) ^ sw) << 4), qa.x & 0x0f0f0f0fu, qa.y & 0x0f0f0f0fu, qa.z & 0x0f0f0f0fu, qa.w & 0x0f0f0f0fu);
                    sts128(tl + (((c0 + 1) ^ sw) << 4), qb.x & 0x0f0f0f0fu, qb.y & 0x0f0f0f0fu, qb.z & 0x0f0f0f0fu, qb.w & 0x0f0f0f0fu);
                    sts128(tl + (((c0 + 2) ^ sw) << 4), (qa.x >> 4) & 0x0f0f0f0fu, (qa.y >> 4) & 0x0f0f0f0fu, (qa.z >> 4) & 0x0f0f0f0fu, (qa.w >> 4) & 0x0f0f0f0fu);
                    sts128(tl + (((c0 + 3) ^ sw) << 4), (qb.x >> 4) & 0x0f0f0f0fu, (qb.y >> 4) & 0x0f0f0f0fu, (qb.z >> 4) & 0x0f0f0f0fu, (qb.w >> 4) & 0x0f0f0f0fu);
                }
                fence_proxy_async();   // the operand tile is read by the tensor core through the async proxy
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(smem_u32(rempty + rs));
                    mbar_arrive(smem_u32(efull + es));
                }
                if (++rs == RS) { rs = 0; rph ^= 1; }
                if (++es == TC_ES) { es = 0; eph ^= 1; }
            }
        }
    } else if (warp >= 4) {
        // ================================================================ epilogue: thread = weight row (TMEM lane), NR columns
        const int quad = warp & 3, r = quad * 32 + lane;
        const uint32_t t_lane = tmem_base + ((uint32_t) (quad * 32) << 16);
        constexpr int G = NR >= 32 ? 2 : (NR >= 16 ? 4 : 8);   // 32-blocks drained per batch of tcgen05.ld (64 registers)
        int rs = 0, t = 0; uint32_t rph = 0, tph = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int tile, part, sb0, sb1;
            item_range(item, tile, part, sb0, sb1);
            float acc[NR];
#pragma unroll
            for (int n = 0; n < NR; n++) acc[n] = 0.0f;
            for (int sb = sb0; sb < sb1; sb++) {
                const uint32_t st = smem_u32(rring + rs * RSTAGE);
                mbar_wait(smem_u32(rfull + rs), rph);
                const uint4 hdr = lds128(st + OFF_RAW + (uint32_t) r * SZ_Q4_K);   // d | dmin, 12 bytes of 6-bit scales / mins
                const uint32_t scw[2] = { hdr.y & 0x3f3f3f3fu, (hdr.w & 0x0f0f0f0fu) | ((hdr.y >> 2) & 0x30303030u) };
                const uint32_t mn_lo = hdr.z & 0x3f3f3f3fu, mn_hi = ((hdr.w >> 4) & 0x0f0f0f0fu) | ((hdr.z >> 2) & 0x30303030u);
                int isum[NR];
#pragma unroll
                for (int n = 0; n < NR; n++) isum[n] = 0;
                mbar_wait(smem_u32(tfull + t), tph);
                tc_fence_after();
#pragma unroll
                for (int j0 = 0; j0 < 8; j0 += G) {
                    if (p.dbg & 32) continue;
                    uint32_t c[G][NR];
#pragma unroll
                    for (int g = 0; g < G; g++) {
                        const uint32_t ta = t_lane + (uint32_t) (t * 8 * BN + (j0 + g) * BN);
                        if constexpr (NR == 8) tmem_ld_32x8(ta, c[g]);
                        else if constexpr (NR == 16) tmem_ld_32x16(ta, c[g]);
                        else tmem_ld_32x32(ta, c[g]);
                    }
                    tc_wait_ld();
#pragma unroll
                    for (int g = 0; g < G; g++) {
                        const int j  = j0 + g;
                        const int sc = (int) ((scw[j >> 2] >> (8 * (j & 3))) & 0xff);
#pragma unroll
                        for (int n = 0; n < NR; n++) {
                            asm volatile("" : "+r"(c[g][n]));   // keeps the use below the wait
                            isum[n] += sc * (int) c[g][n];
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(tempty + t));   // the accumulators are in registers
                if (++t == 2) { t = 0; tph ^= 1; }
                const float d = half_bits_to_float((uint16_t) (hdr.x & 0xffff)), dmin = half_bits_to_float((uint16_t) (hdr.x >> 16));
#pragma unroll
                for (int n = 0; n < NR; n++) {
                    const float dy = __uint_as_float(lds32(st + OFF_DA + 4 * n));
                    const uint4 bs = lds128(st + OFF_BS + 16 * n);
                    int imin = dp2a_lo_su((int) bs.x, mn_lo, 0);
                    imin     = dp2a_hi_su((int) bs.y, mn_lo, imin);
                    imin     = dp2a_lo_su((int) bs.z, mn_hi, imin);
                    imin     = dp2a_hi_su((int) bs.w, mn_hi, imin);
                    acc[n] = fmaf(d * dy, (float) isum[n], acc[n]);
                    acc[n] = fmaf(-(dmin * dy), (float) imin, acc[n]);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(rempty + rs));
                if (++rs == RS) { rs = 0; rph ^= 1; }
            }
            // ---- the item's result. One part: store. Several: every part parks its partial sums; the part that arrives last adds
            // all of them in part order (fixed order: the result does not depend on timing) and stores
            const int row = tile * TC_BM + r;
            if (p.splits == 1) {
                if (row < p.M) {
#pragma unroll
                    for (int n = 0; n < NR; n++) if (n < p.N) p.dst[(int64_t) n * p.ldc + row] = acc[n];
                }
            } else {
                float * mine = p.ws + (size_t) item * (NR * TC_BM);
#pragma unroll
                for (int n = 0; n < NR; n++) mine[n * TC_BM + r] = acc[n];
                __threadfence();
                asm volatile("bar.sync 2, 128;" ::: "memory");
                if (warp == 4 && lane == 0) {
                    const uint32_t old = atomicAdd(p.count + tile, 1u);
                    *last_flag = old == (uint32_t) (p.splits - 1);
                    if (*last_flag) p.count[tile] = 0;
                }
                asm volatile("bar.sync 2, 128;" ::: "memory");
                if (*(volatile uint32_t *) last_flag) {
                    __threadfence();
                    const float * base = p.ws + (size_t) tile * p.splits * (NR * TC_BM);
                    if (row < p.M) {
#pragma unroll
                        for (int n = 0; n < NR; n++) {
                            if (n < p.N) {
                                float v = 0.0f;
                                for (int q = 0; q < p.splits; q++) v += __ldcg(base + (size_t) q * (NR * TC_BM) + n * TC_BM + r);
                                p.dst[(int64_t) n * p.ldc + row] = v;
                            }
                        }
                    }
                }
                asm volatile("bar.sync 2, 128;" ::: "memory");   // last_flag is reused by the next item
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

template <int BN, int NR>
b200_status mvtc_launch(b200_ctx * ctx, const CUtensorMap & map_w, const CUtensorMap & map_b, const CUtensorMap & map_bs, const CUtensorMap & map_da,
                        const TcParams & p, int grid) {
    static std::atomic<bool> attr_done[64];
    constexpr int SMEM = tc_smem(BN);
    if (ctx->device < 64 && !attr_done[ctx->device]) {
        CU(cudaFuncSetAttribute(mvtc_q4k_kernel<BN, NR>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        attr_done[ctx->device] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = dim3(grid);
    cfg.blockDim         = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = SMEM;
    cfg.stream           = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs    = attr;
    cfg.numAttrs = ctx->pdl ? 1 : 0;
    CU(cudaLaunchKernelEx(&cfg, mvtc_q4k_kernel<BN, NR>, map_w, map_b, map_bs, map_da, p));
    ctx->launches++;
    return B200_OK;
}


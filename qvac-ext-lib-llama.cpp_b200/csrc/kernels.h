// kernels.h — internal C++ interface between api.cu (dispatch) and the kernel translation units.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

// ---- quantize.cu (K1)
// val: 0 = q8_0, 1 = q8_K, 2 = q8_1 valued blocks in the reference's block layout
cudaError_t quantize_ggml(int val, const float * x, int64_t k, int64_t nrows, size_t stride, void * y, cudaStream_t st, int * launches);
// kind: ACT_KIND_Q8_0 / ACT_KIND_Q8_K, unit-record layout (common.cuh)
cudaError_t quantize_units(int kind, const float * x, int64_t k, int64_t nrows, size_t stride, void * main, void * aux, cudaStream_t st,
                           int * launches, int pdl = 0);

// ---- mmvq.cu (K3)
constexpr int MV_MAX_MATS = 4;
constexpr int MV_MAX_COLS = 8;
constexpr int MV_MAX_CONS = 16;  // consumer warps per CTA (+ 1 producer warp); more than 8 only up to MV_WIDE_MAX_COLS columns
constexpr int MV_WIDE_MAX_COLS = 4;

// expert routing for MUL_MAT_ID (ggml-cpu.c:1564-1578): pair (i, t) uses expert ids[t*ids_rs + i], activation row
// t*ne11 + i % ne11 of the quantized activations, and writes dst + t*dst_ts + i*dst_is (floats)
struct MvIds {
    const int32_t * ids;      // nullptr = plain MUL_MAT
    int64_t         ids_rs;   // int32 elements between token rows of ids
    int64_t         nb02;     // bytes between expert matrices
    int64_t         dst_is, dst_ts;
    int64_t         x_is, x_ts;   // bytes between b rows (nb11) / tokens (nb12) of the F32 activations (fused quantization)
    int             n_used, ne11, n_as;
};

// Row-split tensor parallelism: this rank computes a slice of dst's rows and the epilogue stores every value
// into the dst of ALL ranks (peer memory over NVLink / NVSwitch) - the all-gather is fused into the kernel. Completion:
// the last CTA to finish bumps the local epoch, writes it into every peer's flag slot for this rank (st.release.sys) and
// waits until every peer's flag has reached the same epoch (ld.acquire.sys); kernel completion therefore means "dst is
// complete on this rank". Epoch and CTA counter live in device memory, so CUDA-graph replays keep counting.
constexpr int SPLIT_MAX_RANKS = 8;
struct SplitParams {
    int        n_ranks;                       // 0 / 1 = off
    int        rank;
    int64_t    dst_delta[SPLIT_MAX_RANKS];    // bytes from a local dst address to the same element on rank r (0 for own rank)
    uint32_t * sync_local;                    // [0..7] flags written by the peers, [8] epoch, [9] CTA counter
    uint32_t * sync_peer[SPLIT_MAX_RANKS];    // rank r's sync words
    int        deferred;                      // 1: publish the epoch but do not wait for the peers here - the next kernel of the
                                              // stream (split_wait in its prologue) or b200_ctx_row_split_wait does
    // ---- flagged-payload exchange for the decode mat-vec kernels (b200_ctx_set_row_split_ll; see ll_store / ll_receive below)
    int          ll;                          // 1: this launch exchanges its rows as 8-byte {value, epoch} records
    char *       ll_stage[SPLIT_MAX_RANKS];   // rank r's staging area: two halves (epoch parity) of ll_half bytes
    const char * ll_base;                     // local address of the symmetric dst region the staging area mirrors (8 bytes per float)
    uint64_t     ll_half;
};

// waits until every peer has published the epoch this rank is at (all of their slices of every earlier dst have arrived).
// Called by one warp after the stream dependency is resolved (the local epoch is final then).
__device__ __forceinline__ void split_wait(const SplitParams & sp) {
    const int lane = threadIdx.x & 31;
    if (lane < sp.n_ranks && lane != sp.rank) {
        uint32_t e, v;
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(e) : "l"(sp.sync_local + 8) : "memory");
        uint64_t t0 = 0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (unsigned spins = 1;; spins++) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(sp.sync_local + lane) : "memory");
            if ((int32_t) (v - e) >= 0) break;
            if ((spins & 0xfff) == 0) {
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 5000000000ull) __trap();
            }
        }
    }
    __syncwarp();
}

__device__ __forceinline__ void split_store(const SplitParams & sp, float * addr, float v) {
    if (sp.n_ranks <= 1) {
        *addr = v;
    } else {
        for (int r = 0; r < sp.n_ranks; r++) *(float *) ((char *) addr + sp.dst_delta[r]) = v;
    }
}

// tile epilogues: a thread's 32 values for dst columns nb .. nb+31 (ldc floats apart) of one dst row, to every rank's dst.
// The rank loop is outside and rolled, the column loop inside: per-element split_store calls unroll into 32 x 8 predicated
// stores, and an epilogue of ten thousand instructions runs at instruction-fetch speed (16 us per 128 x 256 tile, measured).
__device__ __forceinline__ void split_store_cols32(const SplitParams & sp, float * base, int64_t ldc, const float (&v)[32], int nvalid) {
    if (nvalid <= 0) return;
    const int nr = sp.n_ranks <= 1 ? 1 : sp.n_ranks;
#pragma unroll 1
    for (int r = 0; r < nr; r++) {
        float * b = sp.n_ranks <= 1 ? base : (float *) ((char *) base + sp.dst_delta[r]);
        if (nvalid >= 32) {
#pragma unroll
            for (int j = 0; j < 32; j++) b[(int64_t) j * ldc] = v[j];
        } else {
#pragma unroll
            for (int j = 0; j < 32; j++) if (j < nvalid) b[(int64_t) j * ldc] = v[j];
        }
    }
}

// called by ONE warp of every CTA after a CTA barrier that follows all of the CTA's stores. membar.sys is expensive (microseconds,
// and worse when thousands run concurrently), so there is exactly ONE per kernel: every CTA orders its stores before the counter
// with a device-scope fence (the CTA barrier orders the other threads' stores before lane 0's fence; fences are cumulative), and
// the last CTA - which has observed every other CTA's counter increment - issues the system-scope fence before the release store
// of the flag: everything that happens-before that fence is visible to the peers before the flag.
__device__ __forceinline__ void split_complete(const SplitParams & sp, unsigned n_ctas) {
    const int lane = threadIdx.x & 31;
    unsigned last = 0;
    if (lane == 0) {
        __threadfence();  // device scope: orders this CTA's (peer) stores before the counter; the last CTA's system fence is cumulative
        last = atomicAdd(&sp.sync_local[9], 1u) == n_ctas - 1 ? 1u : 0u;
    }
    last = __shfl_sync(0xffffffffu, last, 0);
    if (!last) return;
    __threadfence();  // acquire side of the counter; the release.sys store below carries the (single) system-scope fence
    uint32_t e = 0;
    if (lane == 0) {
        e = sp.sync_local[8] + 1;
        sp.sync_local[8] = e;
        sp.sync_local[9] = 0;
    }
    e = __shfl_sync(0xffffffffu, e, 0);
    if (lane < sp.n_ranks && lane != sp.rank) {
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(sp.sync_peer[lane] + sp.rank), "r"(e) : "memory");
        uint32_t v;
        uint64_t t0 = 0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (unsigned spins = 1; !sp.deferred; spins++) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(sp.sync_local + lane) : "memory");
            if ((int32_t) (v - e) >= 0) break;
            if ((spins & 0xfff) == 0) {  // a peer that never arrives (crashed rank, mismatched launch sequence) must not hang the GPU
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 5000000000ull) __trap();
            }
        }
    }
    __syncwarp();
}

// ---- flagged-payload ("LL") exchange. The flag protocol above needs a system-scope release per kernel, which alone costs ~3 us
// on this machine and leaves 8-GPU decode at ~11 us of exchange per launch. Here a value and the launch's
// epoch travel in ONE aligned 8-byte store into a staging area on every peer (8-byte stores are single-copy atomic: no fence,
// no flag, no counter on the data path), and every rank's threads poll their share of the staging records of the rows the
// OTHER ranks own and copy the values into the local dst: kernel completion still means "the whole dst is here". The staging area
// has two halves used by odd / even epochs: a rank can be at most one launch ahead of a peer (it cannot finish launch L + 1
// before every peer has sent its rows of L + 1, i.e. has finished L), so records of launch L are never overwritten before they
// are read. Epochs are counted in device memory (sync word 8, bumped by the last CTA of every launch of either protocol), so
// CUDA-graph replays keep counting and all ranks agree as long as they run the same launch sequence.
struct SplitGeom {
    int lo[4][SPLIT_MAX_RANKS + 1];           // per matrix of a launch: first row of every rank's slice; [n_ranks] = rows of the full matrix
};

__device__ __forceinline__ uint32_t ll_epoch(const SplitParams & sp) {  // after the stream dependency is resolved
    uint32_t e;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(e) : "l"(sp.sync_local + 8) : "memory");
    return e + 1;
}

__device__ __forceinline__ void ll_store(const SplitParams & sp, uint32_t ep, float * addr, float v) {
    *addr = v;
    const uint64_t off = (uint64_t) ((const char *) addr - sp.ll_base) * 2 + (uint64_t) (ep & 1) * sp.ll_half;
    const uint64_t rec = (uint64_t) __float_as_uint(v) | ((uint64_t) ep << 32);
    for (int r = 0; r < sp.n_ranks; r++) {
        if (r != sp.rank) asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(sp.ll_stage[r] + off), "l"(rec) : "memory");
    }
}

// thread `tid` of `T` receiving threads of the launch: polls the records of every T-th element the peers own and stores the values
template <typename MAT>
__device__ __forceinline__ void ll_receive(const SplitParams & sp, uint32_t ep, const SplitGeom & g, const MAT * mat, int n_mats, int ncols, int tid, int T) {
    const char * stage = sp.ll_stage[sp.rank] + (uint64_t) (ep & 1) * sp.ll_half;
    int done = 0;
    for (int i = 0; i < n_mats; i++) {
        const int lo = g.lo[i][sp.rank], own = g.lo[i][sp.rank + 1] - lo, remote = g.lo[i][sp.n_ranks] - own;
        const int total = remote * ncols;
        for (int e = ((tid - done) % T + T) % T; e < total; e += T) {
            const int c = e / remote, j = e - c * remote;
            float * a = mat[i].dst + (int64_t) c * mat[i].dst_cs + ((j < lo ? j : j + own) - lo);
            const char * rp = stage + (uint64_t) ((const char *) a - sp.ll_base) * 2;
            uint64_t rec, t0 = 0, t1;
            for (unsigned spins = 1;; spins++) {
                asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(rec) : "l"(rp) : "memory");
                if ((uint32_t) (rec >> 32) == ep) break;
                if ((spins & 0x3ff) == 0) {  // a peer that never sends (crashed rank, mismatched launch sequence) must not hang the GPU
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                    if (t0 == 0) t0 = t1;
                    if (t1 - t0 > 5000000000ull) __trap();
                }
            }
            *a = __uint_as_float((uint32_t) rec);
        }
        done += total;
    }
}

// one thread per CTA, after the CTA's receiving threads are done: the last CTA of the launch publishes the epoch for the next one
__device__ __forceinline__ void ll_finish(const SplitParams & sp, uint32_t ep, unsigned n_ctas) {
    if (atomicAdd(&sp.sync_local[9], 1u) == n_ctas - 1) {
        sp.sync_local[9] = 0;
        sp.sync_local[8] = ep;
    }
}

struct MvMat {
    const uint8_t * w;        // row 0 of the weight matrix (16-byte aligned)
    int64_t         nb01;     // bytes between rows (multiple of 16)
    float *         dst;      // dst(row 0, col 0)
    int64_t         dst_cs;   // floats between dst columns
    int             M;        // rows
    int             group0;   // first global row-group id of this matrix
    const float *   bias;     // K7: added to the row's dot product before the GLU / store (one column only); nullptr = none
    int64_t         bias_es;  // MUL_MAT_ID (ADD_ID): floats between the bias rows of two experts
};

// K7 epilogue of the one-column mat-vec (ggml_cuda_mm_fusion_args, ggml-cuda/common.cuh; mmvq.cu:181-328 in the reference):
// glu != 0: mat[0] = gate and mat[1] = up matrix of the same shape, both with group0 = 0; a warp runs the gate rows of a group,
// keeps the result in registers, runs the up rows and stores glu(gate + gate_bias, up + up_bias) to mat[0].dst.
struct MvEpi {
    int   glu;            // 0 = none, else 1 + enum ggml_glu_op (ggml.h:596-605)
    float alpha, limit;   // SWIGLU_OAI
};

// Cross-kernel weight prefetch. A decode step is a dependent chain of small mat-vec kernels: every kernel boundary costs
// ~3 us of launch / dependency / drain latency during which HBM would idle. Weights never depend on earlier kernels, so the
// producer warp of THIS kernel, once its own copies are issued, pulls the head of the NEXT kernel's weight stream (the first
// `groups_per_cta` row groups of every CTA's range - what that kernel's rings will ask for first) into L2 with
// cp.async.bulk.prefetch.L2. The next kernel is predicted by the context (api.cu: the launch that followed these weights
// last time); a wrong prediction costs only the prefetch traffic.
struct MvNext {
    const uint8_t * w[MV_MAX_MATS];
    int64_t         nb01[MV_MAX_MATS];
    int             M[MV_MAX_MATS];
    int             group0[MV_MAX_MATS];
    int             n_mats;           // 0 = no prediction
    int             total_groups;
    int             n_ctas;           // grid of the predicted launch
    int             rpg;
    int             row_bytes;        // U * unit_bytes
    int             groups_per_cta;   // prefetch depth
    int             glu;              // the predicted launch is a K7 gate / up pair: a group is rows of w[0] AND the same rows of w[1]
};

// RMS_NORM (+ MUL by per-channel weights) of the ONE activation column, done by the K1 prologue of the mat-vec kernel before it
// quantizes: x is then the un-normalized row. Needs fuse_quant, one column, and every unit of the row in the prologue's single
// round (U <= 4 x quantizing warps). Summation order: common.cuh unit_sumsq / rms_scale - the bits of rms_norm_units_kernel.
struct MvNorm {
    const float * w;    // nullptr: no MUL
    float         eps;
    int           on;
};

// ROPE constants of a node (ops.cu rope_consts) and the rotation itself, shared by the stand-alone ROPE kernels (ops.cu) and the mat-vec
// kernel that rotates its q / k rows in its epilogue (MvRope): one set of functions, one set of bits
struct RopeP {
    int   n_dims, neox;
    float theta_scale, freq_scale, ext_factor, attn_factor, corr0, corr1;
};
#ifdef __CUDACC__
__device__ __forceinline__ int64_t load_idx(const char * p, int i64) { return i64 ? *(const int64_t *) p : (int64_t) *(const int32_t *) p; }

// cos / sin (times the YaRN magnitude) of pair `ic` at position `base`. theta_base as ggml_rope_cache_init forms it (ops.cpp:5549-5563):
// pos times theta_scale, ic times over, each product rounded - not powf, so that the angle is the reference's to the bit; then
// rope_yarn (ops.cpp:5532-5547)
__device__ __forceinline__ void rope_cs(const RopeP & p, const float * ff, float base, int ic, float & c, float & sn) {
    float tb = base;
    for (int k = 0; k < ic; k++) tb = __fmul_rn(tb, p.theta_scale);
    const float theta_extrap = tb / (ff ? ff[ic] : 1.0f);
    const float theta_interp = p.freq_scale * theta_extrap;
    float theta = theta_interp, mscale = p.attn_factor;
    if (p.ext_factor != 0.0f) {
        const float yv = ((float) ic - p.corr0) / fmaxf(0.001f, p.corr1 - p.corr0);
        const float ramp_mix = (1.0f - fminf(1.0f, fmaxf(0.0f, yv))) * p.ext_factor;
        theta = theta_interp * (1.0f - ramp_mix) + theta_extrap * ramp_mix;
        mscale *= 1.0f + 0.1f * logf(1.0f / p.freq_scale);
    }
    c = cosf(theta) * mscale; sn = sinf(theta) * mscale;
}
__device__ __forceinline__ float rope_lo(float x0, float x1, float c, float sn) { return __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, sn)); }
__device__ __forceinline__ float rope_hi(float x0, float x1, float c, float sn) { return __fadd_rn(__fmul_rn(x0, sn), __fmul_rn(x1, c)); }
#endif

#ifdef __CUDACC__
// L2 prefetch of the head (groups_per_cta row groups) of every CTA's range of the predicted launch nx, shared among `nb` warps of the
// calling grid (this is warp b). Used by the mat-vec kernels' producer warps for the launch behind them, and by FLASH_ATTN_EXT - whose
// own traffic is small and latency-bound - for the launch after that (the FFN's gate / up matrices), so that HBM keeps streaming
// weights while the attention of a decode step runs.
__device__ __forceinline__ void mv_prefetch_next(const MvNext & nx, int b, int nb, int lane) {
    for (int c = b; c < nx.n_ctas; c += nb) {
        const int g0 = (int) ((int64_t) nx.total_groups * c / nx.n_ctas);
        const int g1 = min((int) ((int64_t) nx.total_groups * (c + 1) / nx.n_ctas), g0 + nx.groups_per_cta);
        const int gs = nx.glu ? 2 : 1;  // a gate / up pair: lanes 2k and 2k + 1 take the two matrices of group k
        for (int gl = gs * g0 + lane; gl < gs * g1; gl += 32) {
            const int g = gl / gs;
            int m = nx.glu ? gl - 2 * g : 0;
#pragma unroll
            for (int i = 1; i < MV_MAX_MATS; i++) if (!nx.glu && i < nx.n_mats && g >= nx.group0[i]) m = i;
            const int row0 = (g - nx.group0[m]) * nx.rpg;
            const int nrow = min(nx.rpg, nx.M[m] - row0);
            const uint8_t * src = nx.w[m] + (int64_t) row0 * nx.nb01[m];
            if (nx.nb01[m] == (int64_t) nx.row_bytes) {
                bulk_prefetch_l2(src, (uint32_t) (nrow * nx.row_bytes));
            } else {
                for (int r = 0; r < nrow; r++) bulk_prefetch_l2(src + (int64_t) r * nx.nb01[m], (uint32_t) nx.row_bytes);
            }
        }
    }
}
#endif

// The q / k / v projections of ONE token with what build_attn does to them next folded into the store of the lean mat-vec kernel:
// mat[0] = wq, mat[1] = wk, mat[2] = wv; a row group holds an even number of consecutive rows, i.e. whole ROPE pairs (normal mode,
// n_dims = head size). q rows go rotated to q_dst, k rows rotated to k_dst and - as F16 - into row k_idx[0] of the K cache, v rows as
// F16 into row v_idx[0] of the V cache: ROPE(q), ROPE(k), SET_ROWS(k), SET_ROWS(v) without a launch of their own.
struct MvRope {
    int             on;
    int             head_dim;
    RopeP           rp;
    const int32_t * pos;
    const float *   ff;
    float *         q_dst, * k_dst;
    const char *    k_idx, * v_idx;
    int             k_i64, v_i64;
    char *          kc, * vc;
    int64_t         kc_nb1, vc_nb1, kc_rows, vc_rows;
};

struct MvParams {
    MvMat           mat[MV_MAX_MATS];
    int             n_mats;
    int             total_groups;
    int             U;        // 256-element units per row (K / 256)
    int             L;        // lanes per row: power of two >= min(U, 32)
    int             rpg;      // rows per group: 32 / L when U <= 32, else 1
    int             slices;   // ceil(U / 32)
    int             n_cons, n_stages; // consumer warps (<= MV_MAX_CONS), ring stages per consumer warp
    int             n_helpers;        // mv_lean_kernel: extra warps that only share the K1 prologue (set by the launcher)
    int             fuse_quant;       // 1: quantize src1 inside the kernel (x, x_cs); 0: load act_main / act_aux
    int             x_vec;            // x rows are 16-byte aligned
    int             pdl;              // launch with programmatic stream serialization
    const float *   x;                // F32 activations, column c at (char*)x + c*x_cs
    int64_t         x_cs;
    const uint8_t * act_main;         // [ncols][U][272]
    const uint8_t * act_aux;          // [ncols][U][16]
    // MUL_MAT_ID mode (ids != nullptr): blockIdx.y = pair p = t*n_used + i; one column per pair; mat[0] = expert 0
    MvIds           ids;
    SplitParams     split;
    SplitGeom       geom;     // split.ll: who owns which rows of mat[i]
    MvNext          next;
    MvEpi           epi;
    MvNorm          norm;
    MvRope          rope;
    // debug timeline (b200_ctx_trace): 8 x u64 per CTA at trace + blockIdx.x*8: globaltimer ns of [0] CTA start, [1] first and
    // [2] last weight copy issued, [3] consumer warp 0 past griddepcontrol.wait, [4] activations quantized, [5] first stage
    // landed, [6] consumer warp 0 done, [7] SM id. nullptr = off.
    unsigned long long * trace;
    int             dbg;      // experiment switches (GGML_B200_DBG)
};

// one-column plain launches of these formats run on mv_lean_kernel, which consumes ring stages in pairs (wants 4 stages)
bool mv_lean_type(int type);
bool mv_lean_takes(int type, int ncols, const MvParams & p, int n_pairs);   // mv_fast_launch would run this launch on mv_lean_kernel
constexpr int MV_LEAN_MAX_COLS = 4;
// smem bytes needed by the fast kernel for a configuration
size_t mv_fast_smem(int type, int U, int ncols, int n_cons, int n_stages);
// launches the persistent TMA-staged mat-vec kernel (type, ncols 1..8); grid = n_ctas
cudaError_t mv_fast_launch(int type, int ncols, const MvParams & p, int n_ctas, int n_pairs, size_t smem, cudaStream_t st);

// generic kernel: any K % block == 0, any (2-byte) alignment; activations in the reference's q8_0 / q8_K block layout
struct MvSimpleParams {
    const uint8_t * w;
    int64_t         nb01;
    const uint8_t * act;      // [ncols] rows of q8_0 (34 B / 32) or q8_K (292 B / 256) blocks
    int64_t         act_rs;   // bytes between activation rows
    float *         dst;
    int64_t         dst_cs;
    int             M, K, ncols;
    MvIds           ids;      // ids.ids != nullptr: blockIdx.y = pair, act row = t*ne11 + i % ne11
    int             exact;    // 1: mv_exact_kernel - fp32 operation order of ggml-cpu's generic vec_dots, bit-identical results
};
cudaError_t mv_simple_launch(int type, const MvSimpleParams & p, cudaStream_t st);

} // namespace b200

// internal.h — context struct and error helpers shared by the C-ABI translation units (api.cu, mmid.cu, mmq*.cu).
#pragma once

#include "common.cuh"
#include "kernels.h"

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

b200_status b200_fail(b200_status st, const char * fmt, ...);

#define CU(call)                                                                                                  \
    do {                                                                                                          \
        cudaError_t e_ = (call);                                                                                  \
        if (e_ != cudaSuccess) {                                                                                  \
            cudaGetLastError();                                                                                   \
            return b200_fail(e_ == cudaErrorMemoryAllocation ? B200_ERR_ALLOC : B200_ERR_CUDA, "%s:%d %s: %s",   \
                             __FILE__, __LINE__, #call, cudaGetErrorString(e_));                                  \
        }                                                                                                         \
    } while (0)

struct b200_ctx {
    int          device     = 0;
    cudaStream_t stream     = nullptr;
    bool         own_stream = false;
    int          sm_count   = 148;
    uint64_t     launches   = 0;
    int          force_path = B200_PATH_AUTO;
    // scratch: quantized activations (grow-only; stream ordered because one stream per context)
    void *       scratch      = nullptr;
    size_t       scratch_size = 0;
    // staging for the *_host entry points
    void *       io           = nullptr;
    size_t       io_size      = 0;
    // second scratch: MoE index tables, tcgen05 activation tiles
    void *       scratch2      = nullptr;
    size_t       scratch2_size = 0;
    int          mv_warps = 0, mv_stages = 0;  // 0 = auto; GGML_B200_MV_WARPS / GGML_B200_MV_STAGES
};

// grow-only device buffer owned by the context (fails instead of reallocating while the stream is being captured)
b200_status b200_ensure(b200_ctx * ctx, void ** buf, size_t * cur, size_t need);

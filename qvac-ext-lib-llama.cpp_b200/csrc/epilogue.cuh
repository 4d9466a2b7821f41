// K7: the gated-linear-unit epilogues of GGML_OP_GLU on single values, shared by the fused mat-vec store (mmvq.cu) and the
// standalone element-wise kernel (eltwise.cu). The arithmetic is ggml-cpu's scalar form (ggml-cpu/vec.h:873-875, 958-960,
// ops.cpp:3254-3259; the reference's CUDA forms are ggml-cuda/unary.cuh:92-110): libm-accurate expf / tanhf / erff, no fast-math.
// ggml-cpu itself evaluates GEGLU through an fp16 lookup table (vec.h:1326-1340, GGML_GELU_FP16) and SWIGLU through a polynomial
// expf (ggml_v_silu); the float formulas below are what both approximate, and what the reference's CUDA backend computes.
#pragma once

namespace b200 {

enum { GLU_REGLU = 0, GLU_GEGLU = 1, GLU_SWIGLU = 2, GLU_SWIGLU_OAI = 3, GLU_GEGLU_ERF = 4, GLU_GEGLU_QUICK = 5, GLU_COUNT = 6 };

// x = the gate branch (src0 of GGML_OP_GLU), g = the linear branch (src1)
__device__ __forceinline__ float glu_apply(int op, float x, float g, float alpha, float limit) {
    switch (op) {
        case GLU_REGLU: return (x > 0.0f ? x : 0.0f) * g;
        case GLU_GEGLU: return 0.5f * x * (1.0f + tanhf(0.79788456080286535587989211986876f * x * (1.0f + 0.044715f * x * x))) * g;
        case GLU_SWIGLU: return x / (1.0f + expf(-x)) * g;
        case GLU_SWIGLU_OAI: {
            x = fminf(x, limit);
            g = fmaxf(fminf(g, limit), -limit);
            return x / (1.0f + expf(alpha * (-x))) * (g + 1.0f);
        }
        case GLU_GEGLU_ERF: return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440084436210484f)) * g;
        default: return x * (1.0f / (1.0f + expf(-1.702f * x))) * g;  // GEGLU_QUICK
    }
}

}  // namespace b200

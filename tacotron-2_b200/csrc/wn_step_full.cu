// wn_step_full.cu - WaveNet.step (wavenet.py:650-721): the teacher-forced full-sequence pass used for logit
// equivalence checks (BASELINE.json config 5).  Default: the time-parallel dense-GEMM formulation - on the tensor cores (wn_step_umma.cu, tcgen05.mma) where the channel
// counts allow, else the fp32 FFMA2 kernels of wn_step_gemm.cu.
// Configurations outside its scope, or WN_STEP_SERIAL=1, use the loop engines in "inputs mode" (x[t] is the network input
// of step t, nothing is sampled), which evaluate the same causal dilated convolutions position by position and are
// bit-identical to incremental synthesis under teacher forcing.
#include "wn_common.cuh"

int wn_launch_step_full(wn_handle *h, const float *x, const float *c_up, const float *g_emb, int B, int T,
                        float *raw, cudaStream_t st) {
    if (wn_gemm_supported(h)) return wn_launch_step_gemm(h, x, c_up, B, T, raw, st);
    if (h->pipe && h->engine != WN_ENGINE_GENERIC) {
        return wn_launch_loop_pipe(h, c_up, nullptr, x, 1, B, T, nullptr, raw, st);
    }
    return wn_launch_loop_generic(h, c_up, g_emb, nullptr, x, 1, B, T, nullptr, raw, st);
}

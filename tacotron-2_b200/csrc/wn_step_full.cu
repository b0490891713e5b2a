// wn_step_full.cu - WaveNet.step (wavenet.py:650-721): the teacher-forced full-sequence pass used for logit
// equivalence checks (BASELINE.json config 5).  Round-1 implementation: the loop engines in "inputs mode" (x[t] is the
// network input of step t, nothing is sampled), which evaluate exactly the same causal dilated convolutions position
// by position: the pipelined persistent engine when the model fits it, else the generic kernel.  A time-parallel
// tensor-core kernel (dense GEMMs over time tiles) is the natural next step for this row.
#include "wn_common.cuh"

int wn_launch_step_full(wn_handle *h, const float *x, const float *c_up, const float *g_emb, int B, int T,
                        float *raw, cudaStream_t st) {
    if (h->engine == WN_ENGINE_PIPELINE) {
        return wn_launch_loop_pipe(h, c_up, nullptr, x, 1, B, T, nullptr, raw, st);
    }
    return wn_launch_loop_generic(h, c_up, g_emb, nullptr, x, 1, B, T, nullptr, raw, st);
}

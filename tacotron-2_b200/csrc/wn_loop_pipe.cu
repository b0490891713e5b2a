// wn_loop_pipe.cu - pipelined persistent engine (placeholder until the kernel lands: reports "unavailable").
#include "wn_common.cuh"

int wn_pipe_plan(wn_handle *h) { h->err = "pipelined engine not built yet"; return -1; }
void wn_pipe_free(wn_handle *) {}
void wn_pipe_fill_info(wn_handle *, wn_info *) {}
int wn_launch_loop_pipe(wn_handle *h, const float *, const float *, const float *, int, int, float *, float *,
                        cudaStream_t) {
    h->err = "pipelined engine not built yet";
    return WN_ERR_STATE;
}

// wn_common.cuh - internal types shared by the kernels and the C-ABI layer of libwavenet_b200.so.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/wavenet_b200.h"
#include "../../include/wn_math.h"

// Partial sums per "dot" reduction of the canonical fp32 order (DESIGN.md "Arithmetic"); oracle/c/wn_oracle.c: WNO_LANES.
#define WN_DOT_LANES 16

// Device-side view of the model in the "generic" layout.  Matrices reduced with dot16 (the gate pre-activation over
// the concatenated input [older taps | c | g | x(t)], the two head 1x1s) are stored TRANSPOSED, [N][Kpad] with Kpad a
// multiple of 4 and zero padding, so that the 16 lanes of one output read 256 contiguous bytes; the out / skip 1x1s
// (blocked sequential chains, one thread per output) stay [K][N] as TensorFlow stores them.  Biases always present
// (zeros when hparams.use_bias is off).  Residual layers are laid out with one common stride.
struct WnModelDev {
    int L, R, G, S, O, ksize, H, cin, gin, in_ch, NZ;
    int input_type, legacy, residual_legacy;
    int ksplit;                    // blocks of the out / skip reductions (DESIGN.md "Arithmetic")
    int Kcat, KcatPad, Spad;       // Kcat = ksize*R + cin + gin
    float log_scale_min, log_scale_min_gauss;
    int dil[WN_MAX_LAYERS];
    int qlen[WN_MAX_LAYERS];       // ring slots per layer = (k-1)*d + 1   (wavenet.py:815)
    long long qoff[WN_MAX_LAYERS]; // float offset of layer l's ring inside one utterance's ring block
    long long ring_per_utt;        // floats
    const float *first_w, *first_b;
    const float *layers;           // base of layer 0
    long long layer_stride;        // floats
    int off_cat_w, off_bsum, off_out_w, off_out_b, off_skip_w, off_skip_b;
    int off_os_t, Hpad;            // [skip ; out] transposed, [(S + R)][Hpad]: the copy the cluster engine stages in shared memory
    const float *h1_w, *h1_b, *h2_w, *h2_b;   // h1_w [S][Spad], h2_w [O][Spad] (transposed)
    const float *emb;              // [n_speakers][gin] or nullptr
};

struct WnUpLayerDev {
    const float *w, *b;
    int scale;
};

struct WnHostTensor {
    std::vector<float> data;
    std::vector<int64_t> shape;
};

struct WnPipePlan;  // wn_loop_pipe.cu
struct WnGemmPlan;  // wn_step_gemm.cu
struct WnUmmaPlan;  // wn_step_umma.cu
struct WnTcPlan;    // wn_loop_tc.cu
struct WnTc2Plan;   // wn_loop_tc2.cu

struct wn_handle {
    wn_config cfg;
    int device = 0, sm_count = 0;
    bool finalized = false;
    std::string err;
    std::map<std::string, WnHostTensor> host_w;
    WnModelDev m;
    float *d_blob = nullptr;  // all generic-layout weights
    size_t blob_floats = 0;
    WnUpLayerDev up[WN_MAX_UPSAMPLE];
    int n_up = 0;
    // scratch (grown on demand, owned by the handle)
    float *d_ring = nullptr;  size_t ring_floats = 0;
    float *d_cup = nullptr;   size_t cup_floats = 0;
    float *d_tmp0 = nullptr, *d_tmp1 = nullptr; size_t tmp_floats = 0, tmp1_floats = 0;
    float *d_gemb = nullptr;  size_t gemb_floats = 0;
    // host-call staging
    float *d_in_c = nullptr, *d_in_noise = nullptr, *d_in_teacher = nullptr, *d_out_audio = nullptr,
          *d_out_raw = nullptr;
    int32_t *d_in_g = nullptr;
    size_t in_c_n = 0, in_noise_n = 0, in_teacher_n = 0, out_audio_n = 0, out_raw_n = 0, in_g_n = 0;
    WnPipePlan *pipe = nullptr;
    WnGemmPlan *gemm = nullptr;
    WnUmmaPlan *umma = nullptr;
    WnTcPlan *tc = nullptr;
    WnTc2Plan *tc2 = nullptr;
    int engine = WN_ENGINE_GENERIC;
    int64_t launches = 0;
    cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};  // before upsample / before loop / after loop
    bool ev_valid = false;
    int64_t n_params = 0;
};

#define WN_CUDA_TRY(h, expr)                                                                   \
    do {                                                                                       \
        cudaError_t e__ = (expr);                                                              \
        if (e__ != cudaSuccess) {                                                              \
            (h)->err = std::string(#expr) + ": " + cudaGetErrorString(e__);                    \
            return WN_ERR_CUDA;                                                                \
        }                                                                                      \
    } while (0)

// kernels / launchers implemented in the other translation units --------------------------------------------
int wn_launch_upsample(wn_handle *h, const float *c, int B, int Tc, float *c_up, cudaStream_t st);
int wn_launch_loop_generic(wn_handle *h, const float *c_up, const float *g_emb, const float *noise,
                           const float *teacher, int inputs_mode, int B, int T, float *audio, float *raw,
                           cudaStream_t st);
int wn_launch_step_full(wn_handle *h, const float *x, const float *c_up, const float *g_emb, int B, int T,
                        float *raw, cudaStream_t st);
// pipelined persistent engine (wn_loop_pipe.cu)
int wn_pipe_plan(wn_handle *h);   // decides whether the model fits; builds packed weights; 0 = usable
void wn_pipe_free(wn_handle *h);
void wn_pipe_fill_info(wn_handle *h, wn_info *info);
int wn_pipe_check_abort(wn_handle *h);  // after a stream sync: WN_ERR_WATCHDOG if the persistent kernel bailed out
int wn_launch_loop_pipe(wn_handle *h, const float *c_up, const float *noise, const float *teacher, int inputs_mode,
                        int B, int T, float *audio, float *raw, cudaStream_t st);

// tensor-core loop engine (wn_loop_tc.cu): tcgen05.mma with the step-critical weights resident in tensor memory
bool wn_tc_supported(const wn_handle *h);
int wn_tc_plan(wn_handle *h);
void wn_tc_free(wn_handle *h);
int wn_tc_max_batch(const wn_handle *h);
void wn_tc_fill_info(wn_handle *h, wn_info *info);
int wn_tc_check_abort(wn_handle *h);
int wn_launch_loop_tc(wn_handle *h, const float *c_up, const float *noise, const float *teacher, int inputs_mode, int B, int T,
                      float *audio, float *raw, cudaStream_t st);
// the same for the 256-channel block of paper_hparams.py (wn_loop_tc2.cu: four CTAs per layer); reached through the wn_tc_* entry
// points above, which hand over to these when the model is that block
bool wn_tc2_supported(const wn_handle *h);
int wn_tc2_plan(wn_handle *h);
void wn_tc2_free(wn_handle *h);
void wn_tc2_fill_info(wn_handle *h, wn_info *info);
int wn_tc2_check_abort(wn_handle *h);
int wn_launch_loop_tc2(wn_handle *h, const float *c_up, const float *noise, const float *teacher, int inputs_mode, int B, int T,
                       float *audio, float *raw, cudaStream_t st);

// time-parallel WaveNet.step (wn_step_gemm.cu)
bool wn_gemm_supported(const wn_handle *h);
void wn_gemm_free(wn_handle *h);
int wn_launch_step_gemm(wn_handle *h, const float *x, const float *c_up, int B, int T, float *raw, cudaStream_t st);
// the two GEMMs of a residual layer on the tensor cores (wn_step_umma.cu: tcgen05.mma, bf16 hi/lo split, fp32 TMEM accumulators)
bool wn_umma_supported(const wn_handle *h);
void wn_umma_free(wn_handle *h);
bool wn_umma_head_supported(wn_handle *h);
int wn_umma_head(wn_handle *h, const float *skips, float *h1, float *raw, long long N, int T, cudaStream_t st);
int wn_umma_layer(wn_handle *h, int l, const float *x, const float *c_up, float *z, float *x_out, float *skips, long long N, int T,
                  cudaStream_t st);

int wn_ensure(wn_handle *h, float **p, size_t *cur, size_t need);

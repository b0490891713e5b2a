/* TEST INFRASTRUCTURE - plain-C restatement of the WaveNet incremental synthesis path (parity unpinned for the
 * TF numerics, see oracle/wavenet_oracle.py).  Same math as the numpy oracle, but every dot product is reduced in
 * one fixed order of float32 fma / add operations (wn_oracle.c: dot16, dot_blocked) and every transcendental comes
 * from include/wn_math.h, which is exactly what the fp32 CUDA kernels do - so kernel and checker can be compared
 * bit for bit on free-running audio.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.
 */
#ifndef WN_ORACLE_H_
#define WN_ORACLE_H_
#include <stdint.h>

#define WNO_MAX_LAYERS 64
#define WNO_MAX_UPSAMPLE 8

enum { WNO_INPUT_RAW = 0, WNO_INPUT_MULAW = 1, WNO_INPUT_MULAW_QUANTIZE = 2 };
enum { WNO_UP_SUBPIXEL = 0, WNO_UP_2D = 1, WNO_UP_1D = 2, WNO_UP_RESIZE = 3, WNO_UP_NEAREST = 4 };
enum { WNO_ACT_NONE = 0, WNO_ACT_RELU = 1, WNO_ACT_LEAKY = 2 };

typedef struct {
    int32_t layers, residual_channels, gate_channels, skip_channels, out_channels, kernel_size;
    int32_t cin_channels, gin_channels, in_channels; /* in_channels: 1 (scalar) or quantize_channels */
    int32_t input_type, legacy, residual_legacy, use_bias;
    float log_scale_min, log_scale_min_gauss;
    int32_t dilations[WNO_MAX_LAYERS];
    /* upsampling net */
    int32_t upsample_type, upsample_activation, n_upsample, freq_axis_kernel_size, hop_size;
    int32_t upsample_scales[WNO_MAX_UPSAMPLE];
    float leaky_alpha;
    int32_t ksplit; /* blocks of the out / skip 1x1 reductions (= CTAs per layer of the CUDA plan, wn_info.ksplit); <= 1: one block */
} wno_config;

/* TF layouts, C-contiguous: conv kernels [k*C_in][C_out] (row = tap*C_in + ci, tap 0 oldest; modules.py:238-251). */
typedef struct {
    const float *first_w, *first_b;
    const float *conv_w[WNO_MAX_LAYERS], *conv_b[WNO_MAX_LAYERS];
    const float *cin_w[WNO_MAX_LAYERS], *cin_b[WNO_MAX_LAYERS];
    const float *gin_w[WNO_MAX_LAYERS], *gin_b[WNO_MAX_LAYERS];
    const float *out_w[WNO_MAX_LAYERS], *out_b[WNO_MAX_LAYERS];
    const float *skip_w[WNO_MAX_LAYERS], *skip_b[WNO_MAX_LAYERS];
    const float *head1_w, *head1_b, *head2_w, *head2_b;
    const float *up_w[WNO_MAX_UPSAMPLE], *up_b[WNO_MAX_UPSAMPLE];
} wno_weights;

#ifdef __cplusplus
extern "C" {
#endif

/* c [B][Tc][cin] (normalised mels) -> c_up [B][Tc*hop][cin], time-major.  wavenet.py:781-803 + modules.py:524-770. */
int wno_upsample(const wno_config *cfg, const wno_weights *w, const float *c, int B, int Tc, float *c_up, int nthreads);

/* The loop, wavenet.py:805-911.  c_up [B][T][cin] or NULL; g_emb [B][gin] or NULL; noise [T][B][NZ]
 * (NZ = 1 Gaussian / categorical, K+1 mixture of logistics); teacher [B][T] or NULL (scalar value, or class id for
 * mulaw-quantize); audio [B][T] = decoded waveform (wavenet.py:450-456); samples [B][T] or NULL = the value fed back
 * (scalar sample or class id); raw [B][T][O] or NULL = per-step distribution parameters (wavenet.py:847). */
int wno_incremental(const wno_config *cfg, const wno_weights *w, const float *c_up, const float *g_emb,
                    const float *noise, const float *teacher, int B, int T, float *audio, float *samples, float *raw,
                    int nthreads);

/* Reference-faithful variant of the same loop: literal shift-and-concat queues (modules.py:285-292). Same results. */
int wno_incremental_shiftqueue(const wno_config *cfg, const wno_weights *w, const float *c_up, const float *g_emb,
                               const float *noise, const float *teacher, int B, int T, float *audio, float *samples,
                               float *raw, int nthreads);

/* elementwise checks of include/wn_math.h from Python: fn 0 exp, 1 log, 2 tanh, 3 sigmoid, 4 inv_mulaw */
void wno_math(int fn, const float *x, float *y, int64_t n);

#ifdef __cplusplus
}
#endif
#endif

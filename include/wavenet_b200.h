/* wavenet_b200.h - C ABI of libwavenet_b200.so: the B200 (sm_100a) implementation of the WaveNet vocoder's
 * incremental mel->audio synthesis path of mwhitehill/Tacotron-2.
 *
 * The reference has no FFI for this path: its boundary is the Python class
 *   wavenet_vocoder.synthesizer.Synthesizer.load / .synthesize            (code/wavenet_vocoder/synthesizer.py:15, :46)
 * over one tf.Session.run of
 *   WaveNet.initialize (synthesis branch)  -> WaveNet.incremental / .step (code/wavenet_vocoder/models/wavenet.py:407-465, :724-911, :650-721).
 * Each entry point below names the reference interface it replaces.  All pointers are plain C pointers; sizes are
 * explicit; nothing throws across the ABI; every function returns 0 or a negative wn_status and leaves a message
 * retrievable with wn_last_error().  One handle = one model on one GPU (the current CUDA device at wn_create).
 * Handles are not thread-safe; distinct handles are independent.  INTEGRATION.md shows the ctypes binding.
 */
#ifndef WAVENET_B200_H_
#define WAVENET_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WN_MAX_LAYERS 64
#define WN_MAX_UPSAMPLE 8
#define WN_ABI_VERSION 2

typedef enum {
    WN_OK = 0,
    WN_ERR_INVALID = -1,      /* bad argument / unsupported configuration (reference: ValueError / AssertionError) */
    WN_ERR_STATE = -2,        /* call order (weights missing, not finalized) */
    WN_ERR_CUDA = -3,         /* CUDA runtime failure; message holds cudaGetErrorString */
    WN_ERR_NOMEM = -4,
    WN_ERR_WATCHDOG = -5      /* persistent kernel gave up waiting on a peer CTA (never expected) */
} wn_status;

typedef enum { WN_INPUT_RAW = 0, WN_INPUT_MULAW = 1, WN_INPUT_MULAW_QUANTIZE = 2 } wn_input_type;   /* hparams.input_type */
typedef enum { WN_UP_SUBPIXEL = 0, WN_UP_2D = 1, WN_UP_1D = 2, WN_UP_RESIZE = 3, WN_UP_NEAREST = 4 } wn_upsample_type;
typedef enum { WN_ACT_NONE = 0, WN_ACT_RELU = 1, WN_ACT_LEAKY = 2 } wn_activation;
typedef enum { WN_MEM_DEVICE = 0, WN_MEM_HOST = 1 } wn_mem_kind;
typedef enum { WN_ENGINE_AUTO = 0, WN_ENGINE_GENERIC = 1, WN_ENGINE_PIPELINE = 2, WN_ENGINE_TC = 3 } wn_engine;

/* Mirrors the hparams keys the path reads (code/hparams.py:207-253). */
typedef struct wn_config {
    int32_t abi_version;          /* WN_ABI_VERSION */
    int32_t input_type;           /* wn_input_type */
    int32_t quantize_channels;
    int32_t out_channels;
    int32_t layers, stacks;
    int32_t residual_channels, gate_channels, skip_out_channels, kernel_size;
    int32_t use_bias, legacy, residual_legacy;
    float log_scale_min, log_scale_min_gauss;
    int32_t cin_channels;         /* <= 0 disables local conditioning */
    int32_t upsample_type;        /* wn_upsample_type */
    int32_t upsample_activation;  /* wn_activation */
    int32_t n_upsample_scales;
    int32_t upsample_scales[WN_MAX_UPSAMPLE];
    int32_t freq_axis_kernel_size;
    float leaky_alpha;
    int32_t nn_init;              /* hparams.NN_init: SubPixel without it tiles filter 0 (modules.py:584-592) */
    int32_t hop_size;
    int32_t gin_channels;         /* <= 0 disables global conditioning */
    int32_t use_speaker_embedding, n_speakers;
    int32_t engine;               /* wn_engine; AUTO: the tensor-core loop engine (WN_ENGINE_TC) for the blocks it is built for
                                     (R = S = G/2 = 128 or 256, kernel_size 3, scalar input, local conditioning only), else the
                                     pipelined fp32 engine when the model fits it, else the generic one */
    int32_t ksplit;               /* fp32 summation order of the out / skip 1x1s: number of contiguous blocks of the gated
                                     channels summed separately and then added up (DESIGN.md "Arithmetic").  0 = the
                                     pipelined engine's CTAs-per-layer (1 when only the generic engine applies).  Both
                                     engines honour it, so a fixed value makes results independent of the engine. */
} wn_config;

typedef struct wn_info {
    int32_t receptive_field;      /* wavenet.py:54-71 */
    int64_t n_params;
    int64_t flops_per_sample;     /* 2*MAC of one autoregressive step of one utterance */
    int64_t bytes_per_sample;     /* algorithmic bytes/sample with resident weights */
    int64_t weight_bytes;
    int32_t engine;               /* engine wn_synthesize will use */
    int32_t ksplit;               /* the value in effect (see wn_config.ksplit); a CPU checker must use the same */
    int32_t pipeline_ctas, pipeline_ctas_per_layer, pipeline_group_size;
    int32_t sm_count;
    int64_t kernel_launches;      /* kernels launched by this handle so far (for the bench's gpu_launches) */
    float last_upsample_ms;       /* CUDA-event duration of the upsampling net in the last completed wn_synthesize (-1: n/a) */
    float last_loop_ms;           /* CUDA-event duration of the loop kernel in the last completed wn_synthesize (-1: n/a) */
    int32_t aborted;              /* 1: the persistent kernel of the last call gave up waiting (WN_ERR_WATCHDOG); read after a stream sync */
} wn_info;

typedef struct wn_handle wn_handle;

/* create_model('WaveNet', hparams) + WaveNet.__init__   (models/__init__.py:5-14, wavenet.py:89-208) */
int wn_create(const wn_config *cfg, wn_handle **out);

/* Variable restore (synthesizer.py:33,44 -> train.py:67-86): one call per tensor, HOST pointer, TF layout
 * (Conv1D kernel [k, C_in, C_out], upsampling kernels [kh, kw, ., .], embedding [n_speakers, gin]).
 * Names: tacotron-2_b200/wavenet_vocoder/weights.py (e.g. "ResidualConv1DGLU_3/residual_block_causal_conv/kernel"). */
int wn_load_weights(wn_handle *h, const char *name, const float *data, const int64_t *shape, int rank);

/* End of graph construction: checks every tensor is present, repacks to kernel layouts, uploads. */
int wn_finalize(wn_handle *h);

/* Upsampling network  (wavenet.py:781-803 -> modules.py:524-770).
 * c [B][Tc][cin] normalised mels -> c_up [B][Tc*hop][cin] (time-major).  Device pointers; async on `stream`. */
int wn_upsample(wn_handle *h, const float *c, int B, int Tc, float *c_up, void *stream);

/* The synthesis branch of WaveNet.initialize + WaveNet.incremental (wavenet.py:407-456, :724-911): upsample, run
 * the autoregressive loop for T = Tc*hop steps, decode.  Device pointers; async on `stream`.
 *   c        [B][Tc][cin]   normalised mels (what Synthesizer.synthesize feeds, synthesizer.py:61-70)
 *   g_ids    [B] or NULL    speaker ids (synthesizer.py:72)
 *   noise    [T][B][NZ]     the random draws the reference takes inside the loop, made explicit:
 *                           Gaussian head NZ=1 standard normal (gaussian.py:49-50); mixture of logistics NZ=K+1 uniforms
 *                           in [1e-5, 1-1e-5], K for the component pick then 1 for the logistic (mixture.py:91,104);
 *                           mulaw-quantize NZ=1 uniform in [0,1) (replaces tf.multinomial, wavenet.py:865)
 *   teacher  [B][T] or NULL teacher forcing (test_inputs, wavenet.py:877-878): value fed back instead of the sample
 *   audio    [B][T]         decoded waveform (tower_y_hat, wavenet.py:450-456)
 *   raw      [B][T][O] or NULL  per-step distribution parameters (tower_y_hat_eval, wavenet.py:847, :904-908)
 *   c_up     [B][T][cin] or NULL upsampled conditioning (tower_synth_upsampled_local_features); NULL uses scratch */
int wn_synthesize(wn_handle *h, const float *c, const int32_t *g_ids, int B, int Tc, const float *noise,
                  const float *teacher, float *audio, float *raw, float *c_up, void *stream);

/* Same call with HOST buffers (what Synthesizer.synthesize does around session.run, synthesizer.py:97): copies the
 * inputs to the device, runs, copies audio (and raw if non-NULL) back, and synchronises the stream. */
int wn_synthesize_host(wn_handle *h, const float *c, const int32_t *g_ids, int B, int Tc, const float *noise,
                       const float *teacher, float *audio, float *raw, void *stream);

/* WaveNet.step (wavenet.py:650-721): teacher-forced full-sequence pass, dropout off.
 * x [B][T] scalar inputs (or class ids), c [B][Tc][cin] -> raw [B][T][O].  Device pointers. */
int wn_step_full(wn_handle *h, const float *x, const float *c, const int32_t *g_ids, int B, int Tc, float *raw,
                 void *stream);

/* The upsampling network without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments, kernels and bits as wn_upsample, but the intermediate layers live in a file-scope
 * __device__ array of WN_STATIC_SCRATCH_FLOATS floats (16 MiB, held by every GPU that loads the library until exit)
 * instead of scratch grown with cudaMalloc.  Limit: the input of the last layer, B * Tc * prod(upsample_scales[:-1]) *
 * cin floats, must fit the array (half of it for networks of three or more layers); larger calls return
 * WN_ERR_INVALID with a message.  All calls share the array and it keeps the last call's contents (nothing in it has
 * to start at zero); calls that overlap on the GPU race on it.  The first call on a device looks the array's device
 * address up (cudaGetSymbolAddress): make one call before starting a graph capture. */
#define WN_STATIC_SCRATCH_FLOATS (4u * 1024u * 1024u)
int wn_upsample_static(wn_handle *h, const float *c, int B, int Tc, float *c_up, void *stream);

int wn_query(wn_handle *h, wn_info *info);
const char *wn_last_error(wn_handle *h);   /* h may be NULL: error of the last failed wn_create on this thread */
void wn_destroy(wn_handle *h);

#ifdef __cplusplus
}
#endif
#endif /* WAVENET_B200_H_ */

// wn_api.cu - the C ABI of libwavenet_b200.so (include/wavenet_b200.h): model construction, weight restore,
// repacking, and the synthesis entry points that chain the sm_100a kernels on the caller's stream.
#include <cstdio>
#include <cstring>

#include "wn_common.cuh"

static thread_local std::string g_create_err;

static int fail(wn_handle *h, int code, const std::string &msg) {
    if (h) h->err = msg; else g_create_err = msg;
    return code;
}

int wn_ensure(wn_handle *h, float **p, size_t *cur, size_t need) {
    if (*p && *cur >= need) return 0;
    if (*p) { cudaFree(*p); *p = nullptr; *cur = 0; }
    if (need == 0) need = 1;
    cudaError_t e = cudaMalloc((void **)p, need * sizeof(float));
    if (e != cudaSuccess) {
        h->err = std::string("cudaMalloc(") + std::to_string(need * sizeof(float)) + " B): " + cudaGetErrorString(e);
        *p = nullptr;
        return WN_ERR_NOMEM;
    }
    *cur = need;
    return 0;
}

static bool scalar_input(const wn_config &c) { return c.input_type != WN_INPUT_MULAW_QUANTIZE; }

// Eager versions of the reference's constructor / graph asserts (models/__init__.py:6-9, wavenet.py:93-97, :800).
static int validate(const wn_config &c, std::string &why) {
    if (c.abi_version != WN_ABI_VERSION) { why = "wn_config.abi_version mismatch"; return WN_ERR_INVALID; }
    if (c.input_type < 0 || c.input_type > 2) { why = "input_type must be raw, mulaw or mulaw-quantize"; return WN_ERR_INVALID; }
    if (c.layers <= 0 || c.layers > WN_MAX_LAYERS) { why = "layers out of range (1..64)"; return WN_ERR_INVALID; }
    if (c.stacks <= 0 || c.layers % c.stacks != 0) { why = "layers must be a multiple of stacks"; return WN_ERR_INVALID; }
    if (c.residual_channels <= 0 || c.gate_channels <= 0 || c.skip_out_channels <= 0 || c.out_channels <= 0 ||
        c.kernel_size <= 0) { why = "channel counts and kernel_size must be positive"; return WN_ERR_INVALID; }
    if (c.gate_channels % 2) { why = "gate_channels must be even"; return WN_ERR_INVALID; }
    if (c.layers / c.stacks > 24) { why = "dilation 2^(layers/stacks-1) too large"; return WN_ERR_INVALID; }
    if (c.input_type == WN_INPUT_MULAW_QUANTIZE && c.out_channels != c.quantize_channels) {
        why = "out_channels must equal to quantize_chennels if input_type is 'mulaw-quantize'";
        return WN_ERR_INVALID;
    }
    if (scalar_input(c) && c.out_channels != 2 && c.out_channels % 3 != 0) {
        why = "scalar input needs out_channels == 2 (Gaussian) or a multiple of 3 (mixture of logistics)";
        return WN_ERR_INVALID;
    }
    if (c.cin_channels > 0) {
        if (c.upsample_type < 0 || c.upsample_type > 4) { why = "unknown upsample_type"; return WN_ERR_INVALID; }
        if (c.upsample_activation < 0 || c.upsample_activation > 2) { why = "unknown upsample_activation"; return WN_ERR_INVALID; }
        if (c.hop_size <= 0) { why = "hop_size must be positive"; return WN_ERR_INVALID; }
        if (c.upsample_type != WN_UP_NEAREST) {
            if (c.n_upsample_scales <= 0 || c.n_upsample_scales > WN_MAX_UPSAMPLE) { why = "1..8 upsample_scales"; return WN_ERR_INVALID; }
            long long prod = 1;
            for (int i = 0; i < c.n_upsample_scales; ++i) {
                if (c.upsample_scales[i] <= 0) { why = "upsample_scales must be positive"; return WN_ERR_INVALID; }
                prod *= c.upsample_scales[i];
            }
            if (prod != c.hop_size) {
                why = "prod(upsample_scales) must equal hop_size (the reference fails late at wavenet.py:800)";
                return WN_ERR_INVALID;
            }
            if (c.freq_axis_kernel_size <= 0 || c.freq_axis_kernel_size > 7) { why = "freq_axis_kernel_size 1..7"; return WN_ERR_INVALID; }
        }
    }
    if (c.gin_channels > 0 && c.use_speaker_embedding && c.n_speakers <= 0) { why = "n_speakers must be set"; return WN_ERR_INVALID; }
    return WN_OK;
}

extern "C" int wn_create(const wn_config *cfg, wn_handle **out) {
    if (!cfg || !out) return fail(nullptr, WN_ERR_INVALID, "wn_create: null argument");
    std::string why;
    int rc = validate(*cfg, why);
    if (rc) return fail(nullptr, rc, why);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, WN_ERR_CUDA, std::string("no CUDA device: this library has no CPU path (") +
                                              cudaGetErrorString(e) + ")");
    wn_handle *h = new wn_handle();
    h->cfg = *cfg;
    cudaGetDevice(&h->device);
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, h->device);
    if (e != cudaSuccess) { delete h; return fail(nullptr, WN_ERR_CUDA, cudaGetErrorString(e)); }
    if (prop.major < 10) {
        delete h;
        return fail(nullptr, WN_ERR_CUDA, "libwavenet_b200 is built for sm_100a only; found sm_" +
                                              std::to_string(prop.major) + std::to_string(prop.minor));
    }
    h->sm_count = prop.multiProcessorCount;
    memset(&h->m, 0, sizeof(h->m));
    *out = h;
    return WN_OK;
}

// name -> expected shape (TF layouts).  Mirrors tacotron-2_b200/wavenet_vocoder/weights.py:weight_specs.
static std::map<std::string, std::vector<int64_t>> expected_shapes(const wn_config &c) {
    std::map<std::string, std::vector<int64_t>> s;
    const int64_t R = c.residual_channels, G = c.gate_channels, S = c.skip_out_channels, O = c.out_channels,
                  k = c.kernel_size, in_ch = scalar_input(c) ? 1 : c.quantize_channels;
    auto conv = [&](const std::string &n, int64_t kw, int64_t ci, int64_t co, bool always_bias = false) {
        s[n + "/kernel"] = {kw, ci, co};
        if (c.use_bias || always_bias) s[n + "/bias"] = {co};
    };
    // hparams.use_bias reaches the residual blocks only (wavenet.py:118-133); the input and final 1x1s always have a bias
    // (wavenet.py:103-115, :136-149 build them without the argument)
    conv("input_convolution", 1, in_ch, R, true);
    for (int l = 0; l < c.layers; ++l) {
        std::string p = "ResidualConv1DGLU_" + std::to_string(l) + "/residual_block_";
        conv(p + "causal_conv", k, R, G);
        if (c.cin_channels > 0) conv(p + "cin_conv", 1, c.cin_channels, G);
        if (c.gin_channels > 0) conv(p + "gin_conv", 1, c.gin_channels, G);
        conv(p + "out_conv", 1, G / 2, R);
        conv(p + "skip_conv", 1, G / 2, S);
    }
    conv("final_convolution_1", 1, S, S, true);
    conv("final_convolution_2", 1, S, O, true);
    if (c.gin_channels > 0 && c.use_speaker_embedding) s["gc_embedding"] = {c.n_speakers, c.gin_channels};
    if (c.cin_channels > 0 && c.upsample_type != WN_UP_NEAREST) {
        for (int i = 0; i < c.n_upsample_scales; ++i) {
            const int64_t sc = c.upsample_scales[i], kh = c.freq_axis_kernel_size, C = c.cin_channels;
            std::string p = "local_conditioning_upsampling_" + std::to_string(i + 1);
            switch (c.upsample_type) {
                case WN_UP_SUBPIXEL: s[p + "/kernel"] = {kh, 3, 1, sc}; s[p + "/bias"] = {sc}; break;
                case WN_UP_1D: s[p + "/kernel"] = {1, sc, C, C}; s[p + "/bias"] = {C}; break;
                default: s[p + "/kernel"] = {kh, sc, 1, 1}; s[p + "/bias"] = {1}; break;  // 2D, Resize
            }
        }
    }
    return s;
}

extern "C" int wn_load_weights(wn_handle *h, const char *name, const float *data, const int64_t *shape, int rank) {
    if (!h) return WN_ERR_INVALID;
    if (!name || !data || !shape || rank <= 0 || rank > 4) return fail(h, WN_ERR_INVALID, "wn_load_weights: bad argument");
    if (h->finalized) return fail(h, WN_ERR_STATE, "wn_load_weights after wn_finalize");
    auto exp = expected_shapes(h->cfg);
    auto it = exp.find(name);
    if (it == exp.end()) return fail(h, WN_ERR_INVALID, std::string("unknown weight name: ") + name);
    std::vector<int64_t> shp(shape, shape + rank);
    if (shp != it->second) {
        std::string msg = std::string("weight ") + name + " has shape [";
        for (auto v : shp) msg += std::to_string(v) + ",";
        msg += "] expected [";
        for (auto v : it->second) msg += std::to_string(v) + ",";
        return fail(h, WN_ERR_INVALID, msg + "]");
    }
    size_t n = 1;
    for (auto v : shp) n *= (size_t)v;
    WnHostTensor t;
    t.data.assign(data, data + n);
    t.shape = shp;
    h->host_w[name] = std::move(t);
    return WN_OK;
}

namespace {
struct BlobBuilder {
    std::vector<float> v;
    size_t add(const float *p, size_t n) {   // returns float offset, 16-byte aligned
        while (v.size() % 4) v.push_back(0.f);
        size_t off = v.size();
        if (p) v.insert(v.end(), p, p + n); else v.insert(v.end(), n, 0.f);
        return off;
    }
};

__global__ void gather_embedding_kernel(const float *__restrict__ table, const int32_t *__restrict__ ids, int B,
                                        int gin, int n_speakers, float *__restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * gin) return;
    int b = i / gin, j = i - b * gin;
    int id = ids[b];
    id = id < 0 ? 0 : (id >= n_speakers ? n_speakers - 1 : id);
    out[i] = table[(long long)id * gin + j];   // tf.nn.embedding_lookup, modules.py:19-21
}
}  // namespace

extern "C" int wn_finalize(wn_handle *h) {
    if (!h) return WN_ERR_INVALID;
    if (h->finalized) return WN_OK;
    const wn_config &c = h->cfg;
    auto exp = expected_shapes(c);
    for (auto &kv : exp)
        if (!h->host_w.count(kv.first)) return fail(h, WN_ERR_STATE, "wn_finalize: missing weight " + kv.first);
    WnModelDev &m = h->m;
    m.L = c.layers; m.R = c.residual_channels; m.G = c.gate_channels; m.S = c.skip_out_channels; m.O = c.out_channels;
    m.ksize = c.kernel_size; m.H = c.gate_channels / 2; m.cin = c.cin_channels > 0 ? c.cin_channels : 0;
    m.gin = c.gin_channels > 0 ? c.gin_channels : 0;
    m.in_ch = scalar_input(c) ? 1 : c.quantize_channels;
    m.NZ = (scalar_input(c) && c.out_channels != 2) ? c.out_channels / 3 + 1 : 1;
    m.input_type = c.input_type; m.legacy = c.legacy; m.residual_legacy = c.residual_legacy;
    m.log_scale_min = c.log_scale_min; m.log_scale_min_gauss = c.log_scale_min_gauss;
    const int per_stack = c.layers / c.stacks;
    long long qo = 0;
    for (int l = 0; l < m.L; ++l) {
        m.dil[l] = 1 << (l % per_stack);                      // wavenet.py:125
        m.qlen[l] = (m.ksize - 1) * m.dil[l] + 1;             // wavenet.py:815
        m.qoff[l] = qo;
        qo += (long long)m.qlen[l] * m.R;
    }
    m.ring_per_utt = qo;

    auto W = [&](const std::string &n) -> const float * {
        auto it = h->host_w.find(n);
        return it == h->host_w.end() ? nullptr : it->second.data.data();
    };
    BlobBuilder bb;
    auto r4 = [](int v) { return (v + 3) & ~3; };
    m.Kcat = m.ksize * m.R + m.cin + m.gin;
    m.KcatPad = r4(m.Kcat);
    m.Spad = r4(m.S);
    // transposed copy of a TensorFlow [K][N] matrix: out[n][k], rows padded with zeros to Kpad
    auto transposed = [](const float *w, int K, int N, int Kpad) {
        std::vector<float> t((size_t)N * Kpad, 0.0f);
        for (int k = 0; k < K; ++k)
            for (int n = 0; n < N; ++n) t[(size_t)n * Kpad + k] = w[(size_t)k * N + n];
        return t;
    };
    size_t o_first_w = bb.add(W("input_convolution/kernel"), (size_t)m.in_ch * m.R);
    size_t o_first_b = bb.add(W("input_convolution/bias"), m.R);
    size_t o_layers = 0, stride = 0;
    for (int l = 0; l < m.L; ++l) {
        std::string p = "ResidualConv1DGLU_" + std::to_string(l) + "/residual_block_";
        // gate pre-activation matrix over the concatenated input [older taps | c | g | x(t)] (DESIGN.md "Arithmetic")
        const float *cw = W(p + "causal_conv/kernel"), *iw = W(p + "cin_conv/kernel"), *gw = W(p + "gin_conv/kernel");
        const float *cb = W(p + "causal_conv/bias"), *ib = W(p + "cin_conv/bias"), *gb = W(p + "gin_conv/bias");
        std::vector<float> cat((size_t)m.G * m.KcatPad, 0.0f), bsum(m.G, 0.0f);
        const int n_old = (m.ksize - 1) * m.R;
        for (int n = 0; n < m.G; ++n) {
            float *row = cat.data() + (size_t)n * m.KcatPad;
            for (int k = 0; k < n_old; ++k) row[k] = cw[(size_t)k * m.G + n];
            for (int k = 0; k < m.cin; ++k) row[n_old + k] = iw[(size_t)k * m.G + n];
            for (int k = 0; k < m.gin; ++k) row[n_old + m.cin + k] = gw[(size_t)k * m.G + n];
            for (int k = 0; k < m.R; ++k) row[n_old + m.cin + m.gin + k] = cw[(size_t)(n_old + k) * m.G + n];
            float bs = cb ? cb[n] : 0.0f;
            if (m.cin > 0) bs = bs + (ib ? ib[n] : 0.0f);
            if (m.gin > 0) bs = bs + (gb ? gb[n] : 0.0f);
            bsum[n] = bs;
        }
        size_t base = bb.add(cat.data(), cat.size());
        if (l == 0) o_layers = base;
        m.off_cat_w = 0;
        m.off_bsum = (int)(bb.add(bsum.data(), m.G) - base);
        m.off_out_w = (int)(bb.add(W(p + "out_conv/kernel"), (size_t)m.H * m.R) - base);
        m.off_out_b = (int)(bb.add(W(p + "out_conv/bias"), m.R) - base);
        m.off_skip_w = (int)(bb.add(W(p + "skip_conv/kernel"), (size_t)m.H * m.S) - base);
        m.off_skip_b = (int)(bb.add(W(p + "skip_conv/bias"), m.S) - base);
        {   // rows of [skip ; out] with k contiguous (16-byte staging copies of wn_loop_generic.cu)
            m.Hpad = r4(m.H);
            const float *sw = W(p + "skip_conv/kernel"), *ow = W(p + "out_conv/kernel");
            std::vector<float> ost((size_t)(m.S + m.R) * m.Hpad, 0.0f);
            for (int k = 0; k < m.H; ++k) {
                for (int n = 0; n < m.S; ++n) ost[(size_t)n * m.Hpad + k] = sw[(size_t)k * m.S + n];
                for (int n = 0; n < m.R; ++n) ost[(size_t)(m.S + n) * m.Hpad + k] = ow[(size_t)k * m.R + n];
            }
            m.off_os_t = (int)(bb.add(ost.data(), ost.size()) - base);
        }
        bb.add(nullptr, 0);
        if (l == 0) stride = bb.v.size() - base;
    }
    std::vector<float> h1t = transposed(W("final_convolution_1/kernel"), m.S, m.S, m.Spad);
    std::vector<float> h2t = transposed(W("final_convolution_2/kernel"), m.S, m.O, m.Spad);
    size_t o_h1w = bb.add(h1t.data(), h1t.size());
    size_t o_h1b = bb.add(W("final_convolution_1/bias"), m.S);
    size_t o_h2w = bb.add(h2t.data(), h2t.size());
    size_t o_h2b = bb.add(W("final_convolution_2/bias"), m.O);
    size_t o_emb = 0;
    const bool has_emb = c.gin_channels > 0 && c.use_speaker_embedding;
    if (has_emb) o_emb = bb.add(W("gc_embedding"), (size_t)c.n_speakers * c.gin_channels);
    size_t o_upw[WN_MAX_UPSAMPLE], o_upb[WN_MAX_UPSAMPLE];
    h->n_up = 0;
    if (c.cin_channels > 0 && c.upsample_type != WN_UP_NEAREST) {
        h->n_up = c.n_upsample_scales;
        for (int i = 0; i < h->n_up; ++i) {
            std::string p = "local_conditioning_upsampling_" + std::to_string(i + 1);
            WnHostTensor &kt = h->host_w[p + "/kernel"];
            std::vector<float> kern = kt.data;
            if (c.upsample_type == WN_UP_SUBPIXEL && !c.nn_init) {   // modules.py:584-592: every filter := filter 0
                const size_t s = (size_t)c.upsample_scales[i], rows = kern.size() / s;
                for (size_t r = 0; r < rows; ++r)
                    for (size_t f = 1; f < s; ++f) kern[r * s + f] = kern[r * s];
            }
            o_upw[i] = bb.add(kern.data(), kern.size());
            WnHostTensor &bt = h->host_w[p + "/bias"];
            o_upb[i] = bb.add(bt.data.data(), bt.data.size());
        }
    }
    h->n_params = 0;
    for (auto &kv : h->host_w) h->n_params += (int64_t)kv.second.data.size();
    h->blob_floats = bb.v.size();
    WN_CUDA_TRY(h, cudaMalloc((void **)&h->d_blob, sizeof(float) * bb.v.size()));
    WN_CUDA_TRY(h, cudaMemcpy(h->d_blob, bb.v.data(), sizeof(float) * bb.v.size(), cudaMemcpyHostToDevice));
    const float *d = h->d_blob;
    m.first_w = d + o_first_w; m.first_b = d + o_first_b;
    m.layers = d + o_layers; m.layer_stride = (long long)stride;
    m.h1_w = d + o_h1w; m.h1_b = d + o_h1b; m.h2_w = d + o_h2w; m.h2_b = d + o_h2b;
    m.emb = has_emb ? d + o_emb : nullptr;
    for (int i = 0; i < h->n_up; ++i) {
        h->up[i].w = d + o_upw[i];
        h->up[i].b = d + o_upb[i];
        h->up[i].scale = c.upsample_scales[i];
    }
    // engine selection: the pipelined persistent kernel when the model fits it, else the generic kernel
    // ksplit (summation order of the out / skip 1x1s): the caller's value, else the pipeline plan's CTAs per layer
    h->engine = WN_ENGINE_GENERIC;
    m.ksplit = c.ksplit > 0 ? c.ksplit : 1;
    if (c.ksplit > m.H) return fail(h, WN_ERR_INVALID, "ksplit exceeds gate_channels / 2");
    if (c.engine != WN_ENGINE_GENERIC) {
        int prc = wn_pipe_plan(h);   // honours cfg.ksplit > 0, else picks its own split and reports it in m.ksplit
        if (prc == 0) h->engine = WN_ENGINE_PIPELINE;
        else if (c.engine == WN_ENGINE_PIPELINE) return fail(h, WN_ERR_INVALID, "pipeline engine unavailable: " + h->err);
        else h->err.clear();
    }
    // the tensor-core loop engine: on request, and AUTO's choice for the block it is built for (parity by tolerance, 1e-3
    // criterion, instead of bit for bit; the fp32 engines stay planned: WaveNet.step's serial path, explicit requests)
    if (c.engine == WN_ENGINE_TC || (c.engine == WN_ENGINE_AUTO && wn_tc_supported(h) && !getenv("WN_AUTO_FP32"))) {
        int trc = wn_tc_plan(h);
        if (trc && c.engine == WN_ENGINE_TC) return fail(h, WN_ERR_INVALID, "tensor-core engine unavailable: " + h->err);
        if (!trc) h->engine = WN_ENGINE_TC; else h->err.clear();
    }
    h->finalized = true;
    return WN_OK;
}

static int check_ready(wn_handle *h, int B, int Tc) {
    if (!h) return WN_ERR_INVALID;
    if (!h->finalized) return fail(h, WN_ERR_STATE, "model not finalized (call wn_load_weights for every tensor, then wn_finalize)");
    if (B <= 0 || Tc <= 0) return fail(h, WN_ERR_INVALID, "batch and frame count must be positive");
    // a handle lives on the device that was current at wn_create: make it current for this call (a caller thread that serves
    // several handles, or another library, may have switched devices in between)
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev != h->device) WN_CUDA_TRY(h, cudaSetDevice(h->device));
    return WN_OK;
}

extern "C" int wn_upsample(wn_handle *h, const float *c, int B, int Tc, float *c_up, void *stream) {
    int rc = check_ready(h, B, Tc);
    if (rc) return rc;
    if (!c || !c_up) return fail(h, WN_ERR_INVALID, "wn_upsample: null buffer");
    return wn_launch_upsample(h, c, B, Tc, c_up, (cudaStream_t)stream);
}

static int prep_conditioning(wn_handle *h, const float *c, const int32_t *g_ids, int B, int Tc, float *c_up_out,
                             const float **c_up, const float **g_emb, cudaStream_t st) {
    const wn_config &cfg = h->cfg;
    *c_up = nullptr;
    *g_emb = nullptr;
    if (cfg.cin_channels > 0) {
        if (!c) return fail(h, WN_ERR_INVALID, "local conditioning enabled but c is NULL");
        float *dst = c_up_out;
        if (!dst) {
            size_t need = (size_t)B * Tc * cfg.hop_size * cfg.cin_channels;
            if (wn_ensure(h, &h->d_cup, &h->cup_floats, need)) return WN_ERR_NOMEM;
            dst = h->d_cup;
        }
        int rc = wn_launch_upsample(h, c, B, Tc, dst, st);
        if (rc) return rc;
        *c_up = dst;
    }
    if (cfg.gin_channels > 0) {
        if (!g_ids) return fail(h, WN_ERR_INVALID, "global conditioning enabled but g_ids is NULL");
        if (!h->m.emb) return fail(h, WN_ERR_INVALID, "gin_channels > 0 needs use_speaker_embedding (ids are the only global feature the Synthesizer feeds)");
        if (wn_ensure(h, &h->d_gemb, &h->gemb_floats, (size_t)B * cfg.gin_channels)) return WN_ERR_NOMEM;
        int n = B * cfg.gin_channels;
        gather_embedding_kernel<<<(n + 127) / 128, 128, 0, st>>>(h->m.emb, g_ids, B, cfg.gin_channels, cfg.n_speakers, h->d_gemb);
        h->launches++;
        WN_CUDA_TRY(h, cudaGetLastError());
        *g_emb = h->d_gemb;
    }
    return WN_OK;
}

extern "C" int wn_synthesize(wn_handle *h, const float *c, const int32_t *g_ids, int B, int Tc, const float *noise,
                             const float *teacher, float *audio, float *raw, float *c_up_out, void *stream) {
    int rc = check_ready(h, B, Tc);
    if (rc) return rc;
    if (!noise || !audio) return fail(h, WN_ERR_INVALID, "wn_synthesize: noise and audio buffers are required");
    cudaStream_t st = (cudaStream_t)stream;
    const int T = Tc * (h->cfg.cin_channels > 0 ? h->cfg.hop_size : 1);
    const float *c_up, *g_emb;
    if (!h->ev[0])
        for (int i = 0; i < 3; ++i) WN_CUDA_TRY(h, cudaEventCreate(&h->ev[i]));
    h->ev_valid = false;
    WN_CUDA_TRY(h, cudaEventRecord(h->ev[0], st));
    rc = prep_conditioning(h, c, g_ids, B, Tc, c_up_out, &c_up, &g_emb, st);
    if (rc) return rc;
    WN_CUDA_TRY(h, cudaEventRecord(h->ev[1], st));
    if (h->engine == WN_ENGINE_TC) rc = wn_launch_loop_tc(h, c_up, noise, teacher, 0, B, T, audio, raw, st);
    else if (h->pipe) rc = wn_launch_loop_pipe(h, c_up, noise, teacher, 0, B, T, audio, raw, st);
    else rc = wn_launch_loop_generic(h, c_up, g_emb, noise, teacher, 0, B, T, audio, raw, st);
    if (rc) return rc;
    WN_CUDA_TRY(h, cudaEventRecord(h->ev[2], st));
    h->ev_valid = true;
    return WN_OK;
}

extern "C" int wn_step_full(wn_handle *h, const float *x, const float *c, const int32_t *g_ids, int B, int Tc,
                            float *raw, void *stream) {
    int rc = check_ready(h, B, Tc);
    if (rc) return rc;
    if (!x || !raw) return fail(h, WN_ERR_INVALID, "wn_step_full: x and raw buffers are required");
    cudaStream_t st = (cudaStream_t)stream;
    const int T = Tc * (h->cfg.cin_channels > 0 ? h->cfg.hop_size : 1);
    const float *c_up, *g_emb;
    if (!h->ev[0])
        for (int i = 0; i < 3; ++i) WN_CUDA_TRY(h, cudaEventCreate(&h->ev[i]));
    h->ev_valid = false;
    WN_CUDA_TRY(h, cudaEventRecord(h->ev[0], st));
    rc = prep_conditioning(h, c, g_ids, B, Tc, nullptr, &c_up, &g_emb, st);
    if (rc) return rc;
    WN_CUDA_TRY(h, cudaEventRecord(h->ev[1], st));
    rc = wn_launch_step_full(h, x, c_up, g_emb, B, T, raw, st);
    if (rc) return rc;
    WN_CUDA_TRY(h, cudaEventRecord(h->ev[2], st));
    h->ev_valid = true;
    return WN_OK;
}

template <typename T>
static int stage(wn_handle *h, T **dp, size_t *cur, size_t n) {
    return wn_ensure(h, (float **)dp, cur, n);
}

extern "C" int wn_synthesize_host(wn_handle *h, const float *c, const int32_t *g_ids, int B, int Tc,
                                  const float *noise, const float *teacher, float *audio, float *raw, void *stream) {
    int rc = check_ready(h, B, Tc);
    if (rc) return rc;
    if (!noise || !audio) return fail(h, WN_ERR_INVALID, "wn_synthesize_host: noise and audio buffers are required");
    cudaStream_t st = (cudaStream_t)stream;
    const wn_config &cfg = h->cfg;
    const bool lc = cfg.cin_channels > 0;
    const size_t T = (size_t)Tc * (lc ? cfg.hop_size : 1);
    const size_t n_c = lc ? (size_t)B * Tc * cfg.cin_channels : 0, n_noise = T * B * h->m.NZ, n_a = (size_t)B * T;
    if (lc && !c) return fail(h, WN_ERR_INVALID, "local conditioning enabled but c is NULL");
    if (lc && stage(h, &h->d_in_c, &h->in_c_n, n_c)) return WN_ERR_NOMEM;
    if (stage(h, &h->d_in_noise, &h->in_noise_n, n_noise)) return WN_ERR_NOMEM;
    if (stage(h, &h->d_out_audio, &h->out_audio_n, n_a)) return WN_ERR_NOMEM;
    if (teacher && stage(h, &h->d_in_teacher, &h->in_teacher_n, n_a)) return WN_ERR_NOMEM;
    if (raw && stage(h, &h->d_out_raw, &h->out_raw_n, n_a * cfg.out_channels)) return WN_ERR_NOMEM;
    if (g_ids && stage(h, &h->d_in_g, &h->in_g_n, (size_t)B)) return WN_ERR_NOMEM;
    if (lc) WN_CUDA_TRY(h, cudaMemcpyAsync(h->d_in_c, c, sizeof(float) * n_c, cudaMemcpyHostToDevice, st));
    WN_CUDA_TRY(h, cudaMemcpyAsync(h->d_in_noise, noise, sizeof(float) * n_noise, cudaMemcpyHostToDevice, st));
    if (teacher) WN_CUDA_TRY(h, cudaMemcpyAsync(h->d_in_teacher, teacher, sizeof(float) * n_a, cudaMemcpyHostToDevice, st));
    if (g_ids) WN_CUDA_TRY(h, cudaMemcpyAsync(h->d_in_g, g_ids, sizeof(int32_t) * B, cudaMemcpyHostToDevice, st));
    rc = wn_synthesize(h, lc ? h->d_in_c : nullptr, g_ids ? h->d_in_g : nullptr, B, Tc, h->d_in_noise,
                       teacher ? h->d_in_teacher : nullptr, h->d_out_audio, raw ? h->d_out_raw : nullptr, nullptr, stream);
    if (rc) return rc;
    WN_CUDA_TRY(h, cudaMemcpyAsync(audio, h->d_out_audio, sizeof(float) * n_a, cudaMemcpyDeviceToHost, st));
    if (raw) WN_CUDA_TRY(h, cudaMemcpyAsync(raw, h->d_out_raw, sizeof(float) * n_a * cfg.out_channels, cudaMemcpyDeviceToHost, st));
    WN_CUDA_TRY(h, cudaStreamSynchronize(st));
    rc = wn_tc_check_abort(h);
    return rc ? rc : wn_pipe_check_abort(h);
}

extern "C" int wn_query(wn_handle *h, wn_info *info) {
    if (!h || !info) return WN_ERR_INVALID;
    const wn_config &c = h->cfg;
    memset(info, 0, sizeof(*info));
    const int64_t R = c.residual_channels, G = c.gate_channels, S = c.skip_out_channels, O = c.out_channels,
                  k = c.kernel_size, H = G / 2, cin = c.cin_channels > 0 ? c.cin_channels : 0,
                  gin = c.gin_channels > 0 ? c.gin_channels : 0, in_ch = scalar_input(c) ? 1 : c.quantize_channels;
    const int per_stack = c.layers / c.stacks;
    int64_t dsum = 0;
    for (int l = 0; l < c.layers; ++l) dsum += 1ll << (l % per_stack);
    info->receptive_field = (int32_t)((k - 1) * dsum + 1);                                   // wavenet.py:54-71
    const int64_t mac_layer = k * R * G + cin * G + gin * G + H * S + H * R;
    const int64_t mac = in_ch * R + c.layers * mac_layer + S * S + S * O;
    info->flops_per_sample = 2 * mac;
    const int64_t nz = (scalar_input(c) && O != 2) ? O / 3 + 1 : 1;
    info->bytes_per_sample = cin * 4 + c.layers * k * R * 4 + nz * 4 + 4;
    if (h->finalized) {
        info->n_params = h->n_params;
        info->weight_bytes = h->n_params * 4;
    }
    info->engine = h->engine;
    info->ksplit = h->finalized ? h->m.ksplit : 0;
    info->sm_count = h->sm_count;
    info->kernel_launches = h->launches;
    info->last_upsample_ms = info->last_loop_ms = -1.0f;
    if (h->ev_valid && cudaEventQuery(h->ev[2]) == cudaSuccess) {
        cudaEventElapsedTime(&info->last_upsample_ms, h->ev[0], h->ev[1]);
        cudaEventElapsedTime(&info->last_loop_ms, h->ev[1], h->ev[2]);
    }
    wn_pipe_fill_info(h, info);
    wn_tc_fill_info(h, info);
    info->aborted = (wn_tc_check_abort(h) != WN_OK || wn_pipe_check_abort(h) != WN_OK) ? 1 : 0;   // meaningful after a stream sync
    return WN_OK;
}

extern "C" const char *wn_last_error(wn_handle *h) { return h ? h->err.c_str() : g_create_err.c_str(); }

extern "C" void wn_destroy(wn_handle *h) {
    if (!h) return;
    wn_pipe_free(h);
    wn_tc_free(h);
    wn_gemm_free(h);
    wn_umma_free(h);
    for (int i = 0; i < 3; ++i)
        if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    float *bufs[] = {h->d_blob, h->d_ring, h->d_cup, h->d_tmp0, h->d_tmp1, h->d_gemb, h->d_in_c, h->d_in_noise,
                     h->d_in_teacher, h->d_out_audio, h->d_out_raw, (float *)h->d_in_g};
    for (float *p : bufs)
        if (p) cudaFree(p);
    delete h;
}

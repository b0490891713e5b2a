// wn_loop_generic.cu - "generic" engine of the autoregressive loop (wavenet.py:805-911): one CTA per utterance,
// the whole time loop inside ONE kernel launch, weights streamed from L2 in their TensorFlow [K][N] layout,
// dilation ring buffers in global memory, sampling fused, noise read from the caller's buffer.
//
// This engine accepts every configuration the reference's hparams allow (any R/G/S/L/kernel_size, scalar or
// one-hot input, Gaussian / mixture-of-logistics / categorical heads, local and global conditioning).  It is the
// correctness anchor and the fall-back for models whose weights do not fit the shared memory of the pipelined
// persistent engine (wn_loop_pipe.cu), which is the fast path.
//
// Arithmetic: every dot product is a sequential fp32 fma chain over k followed by ONE bias add, every
// transcendental comes from include/wn_math.h - the exact order oracle/c/wn_oracle.c uses, so free-running
// audio can be compared bit for bit.
#include "wn_common.cuh"

namespace {

// y[n] = bias[n] + sum_k x[k] * W[k][n]; x in shared memory (broadcast), W rows coalesced over n.
__device__ __forceinline__ float dot_col(const float *__restrict__ xs, int K, const float *__restrict__ W, int N, int n) {
    float acc = 0.0f;
    const float *wp = W + n;
    int k = 0;
    for (; k + 8 <= K; k += 8) {
        float w0 = __ldg(wp + (long long)(k + 0) * N), w1 = __ldg(wp + (long long)(k + 1) * N);
        float w2 = __ldg(wp + (long long)(k + 2) * N), w3 = __ldg(wp + (long long)(k + 3) * N);
        float w4 = __ldg(wp + (long long)(k + 4) * N), w5 = __ldg(wp + (long long)(k + 5) * N);
        float w6 = __ldg(wp + (long long)(k + 6) * N), w7 = __ldg(wp + (long long)(k + 7) * N);
        acc = __fmaf_rn(xs[k + 0], w0, acc);
        acc = __fmaf_rn(xs[k + 1], w1, acc);
        acc = __fmaf_rn(xs[k + 2], w2, acc);
        acc = __fmaf_rn(xs[k + 3], w3, acc);
        acc = __fmaf_rn(xs[k + 4], w4, acc);
        acc = __fmaf_rn(xs[k + 5], w5, acc);
        acc = __fmaf_rn(xs[k + 6], w6, acc);
        acc = __fmaf_rn(xs[k + 7], w7, acc);
    }
    for (; k < K; ++k) acc = __fmaf_rn(xs[k], __ldg(wp + (long long)k * N), acc);
    return acc;
}

__device__ float sample_gaussian(const WnModelDev &m, const float *out, float eps) {  // gaussian.py:39-52
    float mean = out[0];
    float ls = out[1] > m.log_scale_min_gauss ? out[1] : m.log_scale_min_gauss;
    float x = __fadd_rn(mean, __fmul_rn(wn_expf(ls), eps));
    x = x > -1.0f ? x : -1.0f;
    return x < 1.0f ? x : 1.0f;
}

__device__ float sample_mol(const WnModelDev &m, const float *out, const float *u) {  // mixture.py:76-107
    const int K = m.O / 3;
    int best = 0;
    float bestv = 0.0f;
    for (int k = 0; k < K; ++k) {
        float v = __fsub_rn(out[k], wn_logf(-wn_logf(u[k])));
        if (k == 0 || v > bestv) { bestv = v; best = k; }
    }
    float mean = out[K + best];
    float ls = out[2 * K + best] > m.log_scale_min ? out[2 * K + best] : m.log_scale_min;
    float u2 = u[K];
    float lg = __fsub_rn(wn_logf(u2), wn_logf(__fsub_rn(1.0f, u2)));
    float x = __fadd_rn(mean, __fmul_rn(wn_expf(ls), lg));
    x = x > -1.0f ? x : -1.0f;
    return x < 1.0f ? x : 1.0f;
}

__device__ int sample_categorical(const float *logits, int C, float u, float *e) {
    float mx = logits[0];
    for (int k = 1; k < C; ++k) mx = logits[k] > mx ? logits[k] : mx;
    float S = 0.0f;
    for (int k = 0; k < C; ++k) { e[k] = wn_expf(__fsub_rn(logits[k], mx)); S = __fadd_rn(S, e[k]); }
    float thr = __fmul_rn(u, S), acc = 0.0f;
    for (int k = 0; k < C; ++k) {
        acc = __fadd_rn(acc, e[k]);
        if (acc > thr) return k;
    }
    return C - 1;
}

// grid = B utterances, block = NT threads.  Dynamic shared memory carved below.
__global__ void wn_loop_generic_kernel(const WnModelDev m, const float *__restrict__ c_up,
                                       const float *__restrict__ g_emb, const float *__restrict__ noise,
                                       const float *__restrict__ teacher, int inputs_mode, int B, int T,
                                       float *__restrict__ rings,
                                       float *__restrict__ audio, float *__restrict__ raw) {
    extern __shared__ float sm[];
    const int tid = threadIdx.x, NT = blockDim.x, b = blockIdx.x;
    const int R = m.R, G = m.G, S = m.S, O = m.O, H = m.H, ks = m.ksize;
    float *x = sm;                 // [R]   current residual-stream vector
    float *taps = x + R;           // [ks*R]
    float *pre = taps + ks * R;    // [G]
    float *z = pre + G;            // [H]
    float *skips = z + H;          // [S]
    float *h1 = skips + S;         // [S]
    float *outv = h1 + S;          // [O]
    float *ev = outv + O;          // [O]   categorical scratch
    float *cond = ev + O;          // [cin]
    float *gv = cond + (m.cin > 0 ? m.cin : 0);  // [gin]
    float *fb = gv + (m.gin > 0 ? m.gin : 0);    // [2]: fb[0] next scalar input, fb[1] next class
    const bool quant = m.input_type == WN_INPUT_MULAW_QUANTIZE;
    float *ring = rings + (long long)b * m.ring_per_utt;

    if (tid == 0) {  // wavenet.py:433-445: raw/mulaw 0.0; class int((0+1)/2*255)=127
        fb[0] = 0.0f;
        fb[1] = 127.0f;
        // inputs_mode (WaveNet.step semantics, wavenet.py:650-721): teacher[t] IS the network input of step t
        if (inputs_mode) fb[0] = fb[1] = teacher[(long long)b * T];
    }
    for (int i = tid; i < m.gin; i += NT) gv[i] = g_emb ? g_emb[(long long)b * m.gin + i] : 0.0f;
    __syncthreads();

    for (int t = 0; t < T; ++t) {
        // ---- first_conv.incremental_step (wavenet.py:826) ------------------------------------------------
        for (int n = tid; n < R; n += NT) {
            float acc;
            if (quant) acc = __ldg(m.first_w + (long long)((int)fb[1]) * R + n);
            else acc = __fmaf_rn(fb[0], __ldg(m.first_w + n), 0.0f);
            x[n] = __fadd_rn(acc, __ldg(m.first_b + n));
        }
        if (m.cin > 0)
            for (int i = tid; i < m.cin; i += NT) cond[i] = __ldg(c_up + ((long long)b * T + t) * m.cin + i);
        __syncthreads();

        for (int l = 0; l < m.L; ++l) {
            const float *lw = m.layers + (long long)l * m.layer_stride;
            const int d = m.dil[l], ql = m.qlen[l];
            float *rq = ring + m.qoff[l];
            // ---- CausalConv1D incremental: push x_t, gather the k taps (modules.py:273-292) ----------------
            for (int i = tid; i < ks * R; i += NT) {
                const int j = i / R, n = i - j * R;
                const int tt = t - (ks - 1 - j) * d;
                float v;
                if (j == ks - 1) { v = x[n]; rq[(long long)(t % ql) * R + n] = v; }
                else v = tt >= 0 ? rq[(long long)(tt % ql) * R + n] : 0.0f;
                taps[i] = v;
            }
            __syncthreads();
            // ---- dilated conv + conditioning 1x1s (modules.py:294-297, :497-508) --------------------------
            for (int n = tid; n < G; n += NT) {
                float a = __fadd_rn(dot_col(taps, ks * R, lw + m.off_conv_w, G, n), __ldg(lw + m.off_conv_b + n));
                if (m.cin > 0)
                    a = __fadd_rn(a, __fadd_rn(dot_col(cond, m.cin, lw + m.off_cin_w, G, n), __ldg(lw + m.off_cin_b + n)));
                if (m.gin > 0)
                    a = __fadd_rn(a, __fadd_rn(dot_col(gv, m.gin, lw + m.off_gin_w, G, n), __ldg(lw + m.off_gin_b + n)));
                pre[n] = a;
            }
            __syncthreads();
            for (int n = tid; n < H; n += NT) z[n] = __fmul_rn(wn_tanhf(pre[n]), wn_sigmoidf(pre[H + n]));  // :510
            __syncthreads();
            // ---- skip and residual 1x1s (modules.py:512-520, wavenet.py:833-836) ---------------------------
            for (int n = tid; n < S + R; n += NT) {
                if (n < S) {
                    float s = __fadd_rn(dot_col(z, H, lw + m.off_skip_w, S, n), __ldg(lw + m.off_skip_b + n));
                    if (l == 0) skips[n] = s;
                    else {
                        float v = __fadd_rn(skips[n], s);
                        skips[n] = m.legacy ? __fmul_rn(v, WN_SQRT_HALF) : v;
                    }
                } else {
                    const int r = n - S;
                    float o = __fadd_rn(dot_col(z, H, lw + m.off_out_w, R, r), __ldg(lw + m.off_out_b + r));
                    float v = __fadd_rn(o, x[r]);
                    x[r] = m.residual_legacy ? __fmul_rn(v, WN_SQRT_HALF) : v;
                }
            }
            __syncthreads();
        }
        // ---- head (wavenet.py:840-844) ---------------------------------------------------------------------
        for (int n = tid; n < S; n += NT) skips[n] = skips[n] > 0.0f ? skips[n] : 0.0f;
        __syncthreads();
        for (int n = tid; n < S; n += NT) {
            float v = __fadd_rn(dot_col(skips, S, m.h1_w, S, n), __ldg(m.h1_b + n));
            h1[n] = v > 0.0f ? v : 0.0f;
        }
        __syncthreads();
        for (int n = tid; n < O; n += NT) {
            float v = __fadd_rn(dot_col(h1, S, m.h2_w, O, n), __ldg(m.h2_b + n));
            outv[n] = v;
            if (raw) raw[((long long)b * T + t) * O + n] = v;  // wavenet.py:847
        }
        __syncthreads();
        // ---- sampling + feedback (wavenet.py:850-878) ----------------------------------------------------------
        if (tid == 0) {
            const int ti = inputs_mode ? t + 1 : t;  // index of the teacher value that becomes the next input
            if (noise == nullptr) {   // step() semantics: distribution parameters only, nothing is drawn
                if (teacher && ti < T) fb[0] = fb[1] = teacher[(long long)b * T + ti];
            } else if (!quant) {
                const float *nz = noise + ((long long)t * B + b) * m.NZ;
                float smp = (O == 2) ? sample_gaussian(m, outv, nz[0]) : sample_mol(m, outv, nz);
                if (audio) audio[(long long)b * T + t] = m.input_type == WN_INPUT_MULAW ? wn_inv_mulawf(smp) : smp;
                fb[0] = teacher ? (ti < T ? teacher[(long long)b * T + ti] : 0.0f) : smp;
            } else {
                const float *nz = noise + ((long long)t * B + b) * m.NZ;
                int cls = sample_categorical(outv, O, nz[0], ev);
                float yq = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, (float)cls), 255.0f), 1.0f);
                if (audio) audio[(long long)b * T + t] = wn_inv_mulawf(yq);
                fb[1] = teacher ? (ti < T ? teacher[(long long)b * T + ti] : 0.0f) : (float)cls;
            }
        }
        __syncthreads();
    }
}

}  // namespace

int wn_launch_loop_generic(wn_handle *h, const float *c_up, const float *g_emb, const float *noise,
                           const float *teacher, int inputs_mode, int B, int T, float *audio, float *raw,
                           cudaStream_t st) {
    const WnModelDev &m = h->m;
    if (wn_ensure(h, &h->d_ring, &h->ring_floats, (size_t)B * (size_t)m.ring_per_utt)) return WN_ERR_NOMEM;
    WN_CUDA_TRY(h, cudaMemsetAsync(h->d_ring, 0, sizeof(float) * (size_t)B * (size_t)m.ring_per_utt, st));
    int need = m.G > m.R + m.S ? m.G : m.R + m.S;
    int nt = ((need + 31) / 32) * 32;
    if (nt > 1024) nt = 1024;
    if (nt < 64) nt = 64;
    size_t smem = sizeof(float) * (size_t)(m.R + m.ksize * m.R + m.G + m.H + 2 * m.S + 2 * m.O +
                                           (m.cin > 0 ? m.cin : 0) + (m.gin > 0 ? m.gin : 0) + 4);
    if (smem > 48 * 1024)
        WN_CUDA_TRY(h, cudaFuncSetAttribute(wn_loop_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    wn_loop_generic_kernel<<<B, nt, smem, st>>>(m, c_up, g_emb, noise, teacher, inputs_mode, B, T, h->d_ring,
                                                audio, raw);
    h->launches++;
    WN_CUDA_TRY(h, cudaGetLastError());
    return WN_OK;
}

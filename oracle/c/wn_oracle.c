/* TEST INFRASTRUCTURE - see wn_oracle.h.  Build: oracle/build_oracle.py (gcc -O2 -ffp-contract=off -mfma -fopenmp). */
#include "wn_oracle.h"

#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../../include/wn_math.h"

/* ---- the fixed ("canonical") accumulation orders shared with the CUDA kernels ("Arithmetic") -------
 * TensorFlow does not define the reduction order of tf.matmul, so any fixed order is a valid restatement of
 * modules.py:295-297; these are the orders the sm_100a kernels use, which makes kernel and checker bit-comparable.
 *
 * dot16: the K products are dealt to 16 partial sums by groups of four (element k belongs to partial (k/4) % 16),
 * each partial is a sequential fma chain in ascending k, and the partials are added as a balanced binary tree
 * (neighbours first: ((p0+p1)+(p2+p3)) + ...).  On the GPU the 16 partials live in 16 lanes of a warp and the tree
 * is an xor-butterfly of shuffles.  rows[k] points at row k of a [K][N] matrix (TensorFlow layout).
 */
#define WNO_LANES 16
static void dot16_rows(const float *x, int K, const float *const *rows, int N, const float *bias, float *y,
                       float *scratch /* [16][N] */) {
    memset(scratch, 0, sizeof(float) * (size_t)WNO_LANES * N);
    for (int k = 0; k < K; ++k) {
        float *p = scratch + (size_t)((k >> 2) % WNO_LANES) * N;
        const float xk = x[k];
        const float *Wk = rows[k];
        for (int n = 0; n < N; ++n) p[n] = fmaf(xk, Wk[n], p[n]);
    }
    for (int s = 1; s < WNO_LANES; s <<= 1)
        for (int j = 0; j < WNO_LANES; j += 2 * s) {
            float *a = scratch + (size_t)j * N;
            const float *b = scratch + (size_t)(j + s) * N;
            for (int n = 0; n < N; ++n) a[n] = a[n] + b[n];
        }
    for (int n = 0; n < N; ++n) y[n] = scratch[n] + (bias ? bias[n] : 0.0f);
}

static void dot16(const float *x, int K, const float *W, int N, const float *bias, float *y, float *scratch,
                  const float **rows /* [K] scratch */) {
    for (int k = 0; k < K; ++k) rows[k] = W + (size_t)k * N;
    dot16_rows(x, K, rows, N, bias, y, scratch);
}

/* blocked: the K inputs are cut into `nsplit` contiguous blocks [K*p/nsplit, K*(p+1)/nsplit); each block is one
 * sequential fma chain starting from 0, the block sums are added in ascending p, then the bias.  On the GPU every
 * block is computed by a different CTA (the one that owns those gate channels) and the consumer adds them up. */
static void dot_blocked(const float *x, int K, const float *W, int N, const float *bias, int nsplit, float *y,
                        float *scratch /* [N] */) {
    if (nsplit < 1) nsplit = 1;
    for (int p = 0; p < nsplit; ++p) {
        const int k0 = (int)((long long)K * p / nsplit), k1 = (int)((long long)K * (p + 1) / nsplit);
        float *acc = p == 0 ? y : scratch;
        for (int n = 0; n < N; ++n) acc[n] = 0.0f;
        for (int k = k0; k < k1; ++k) {
            const float xk = x[k];
            const float *Wk = W + (size_t)k * N;
            for (int n = 0; n < N; ++n) acc[n] = fmaf(xk, Wk[n], acc[n]);
        }
        if (p > 0)
            for (int n = 0; n < N; ++n) y[n] = y[n] + scratch[n];
    }
    for (int n = 0; n < N; ++n) y[n] = y[n] + (bias ? bias[n] : 0.0f);
}

static inline float act(float y, int kind, float alpha) {
    if (kind == WNO_ACT_RELU) return y > 0.0f ? y : 0.0f;
    if (kind == WNO_ACT_LEAKY) return y > 0.0f ? y : alpha * y;
    return y;
}

/* ---- upsampling net: intermediates are time-major [W][C] per utterance ------------------------------------- */
static void up_layer(const wno_config *cfg, int li, const float *x, int W, const float *K, const float *bias,
                     float *y) {
    const int C = cfg->cin_channels, s = cfg->upsample_scales[li], kh = cfg->freq_axis_kernel_size;
    const int akind = cfg->upsample_activation;
    const float alpha = cfg->leaky_alpha;
    if (cfg->upsample_type == WNO_UP_SUBPIXEL) {
        /* modules.py:596-640: y[h][w*s+f] = act(b[f] + sum_{i,j} x[h+i-pt][w+j-1] K[i][j][0][f]) */
        const int pt = (kh - 1) / 2, kw = 3, pl = 1;
        for (int w = 0; w < W; ++w)
            for (int h = 0; h < C; ++h)
                for (int f = 0; f < s; ++f) {
                    float acc = 0.0f;
                    for (int i = 0; i < kh; ++i) {
                        int hh = h + i - pt;
                        if (hh < 0 || hh >= C) continue;
                        for (int j = 0; j < kw; ++j) {
                            int ww = w + j - pl;
                            if (ww < 0 || ww >= W) continue;
                            acc = fmaf(x[(size_t)ww * C + hh], K[(i * kw + j) * s + f], acc);
                        }
                    }
                    y[((size_t)w * s + f) * C + h] = act(acc + bias[f], akind, alpha);
                }
    } else if (cfg->upsample_type == WNO_UP_2D) {
        /* modules.py:736-759: y[h][w*s+j] = act(b + sum_i x[h+pt-i][w] K[i][j]) */
        const int pt = (kh - 1) / 2;
        for (int w = 0; w < W; ++w)
            for (int h = 0; h < C; ++h)
                for (int j = 0; j < s; ++j) {
                    float acc = 0.0f;
                    for (int i = 0; i < kh; ++i) {
                        int hh = h + pt - i;
                        if (hh < 0 || hh >= C) continue;
                        acc = fmaf(x[(size_t)w * C + hh], K[i * s + j], acc);
                    }
                    y[((size_t)w * s + j) * C + h] = act(acc + bias[0], akind, alpha);
                }
    } else if (cfg->upsample_type == WNO_UP_1D) {
        /* modules.py:697-722: y[co][w*s+j] = act(b[co] + sum_ci x[ci][w] K[0][j][co][ci]) */
        for (int w = 0; w < W; ++w)
            for (int j = 0; j < s; ++j)
                for (int co = 0; co < C; ++co) {
                    float acc = 0.0f;
                    for (int ci = 0; ci < C; ++ci)
                        acc = fmaf(x[(size_t)w * C + ci], K[((size_t)j * C + co) * C + ci], acc);
                    y[((size_t)w * s + j) * C + co] = act(acc + bias[co], akind, alpha);
                }
    } else { /* WNO_UP_RESIZE, modules.py:657-683: nearest resize by s then conv (kh, s) SAME, one filter */
        const int pt = (kh - 1) / 2, pl = (s - 1) / 2, Wo = W * s;
        const float scale = (float)W / (float)Wo;
        for (int u = 0; u < Wo; ++u)
            for (int h = 0; h < C; ++h) {
                float acc = 0.0f;
                for (int i = 0; i < kh; ++i) {
                    int hh = h + i - pt;
                    if (hh < 0 || hh >= C) continue;
                    for (int j = 0; j < s; ++j) {
                        int uu = u + j - pl;
                        if (uu < 0 || uu >= Wo) continue;
                        int src = (int)floorf((float)uu * scale);
                        if (src > W - 1) src = W - 1;
                        acc = fmaf(x[(size_t)src * C + hh], K[i * s + j], acc);
                    }
                }
                y[(size_t)u * C + h] = act(acc + bias[0], akind, alpha);
            }
    }
}

int wno_upsample(const wno_config *cfg, const wno_weights *w, const float *c, int B, int Tc, float *c_up,
                 int nthreads) {
    const int C = cfg->cin_channels, hop = cfg->hop_size;
    if (C <= 0) return -1;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    int err = 0;
#pragma omp parallel for schedule(dynamic)
    for (int b = 0; b < B; ++b) {
        const float *x0 = c + (size_t)b * Tc * C;
        float *out = c_up + (size_t)b * Tc * hop * C;
        if (cfg->upsample_type == WNO_UP_NEAREST) {
            /* modules.py:524-536: src = min(floor(dst * (in/out)), in-1) in float32 */
            const int Wo = Tc * hop;
            const float scale = (float)Tc / (float)Wo;
            for (int u = 0; u < Wo; ++u) {
                int src = (int)floorf((float)u * scale);
                if (src > Tc - 1) src = Tc - 1;
                memcpy(out + (size_t)u * C, x0 + (size_t)src * C, sizeof(float) * C);
            }
            continue;
        }
        int W = Tc;
        float *cur = (float *)malloc(sizeof(float) * (size_t)W * C);
        if (!cur) { err = -2; continue; }
        memcpy(cur, x0, sizeof(float) * (size_t)W * C);
        for (int li = 0; li < cfg->n_upsample; ++li) {
            int s = cfg->upsample_scales[li];
            int last = (li == cfg->n_upsample - 1);
            float *nxt = last ? out : (float *)malloc(sizeof(float) * (size_t)W * s * C);
            if (!nxt) { err = -2; break; }
            up_layer(cfg, li, cur, W, w->up_w[li], w->up_b[li], nxt);
            free(cur);
            cur = last ? NULL : nxt;
            W *= s;
        }
        if (cur) free(cur);
    }
    return err;
}

/* ---- samplers ------------------------------------------------------------------------------------------ */
static float sample_gaussian(const wno_config *cfg, const float *out, float eps) {
    /* gaussian.py:39-52 */
    float mean = out[0];
    float ls = out[1] > cfg->log_scale_min_gauss ? out[1] : cfg->log_scale_min_gauss;
    float x = mean + wn_expf(ls) * eps;
    x = x > -1.0f ? x : -1.0f;
    return x < 1.0f ? x : 1.0f;
}

static float sample_mol(const wno_config *cfg, const float *out, const float *u) {
    /* mixture.py:76-107; u[0..K-1] pick the component (Gumbel-max, lowest index on ties), u[K] drives the logistic */
    const int K = cfg->out_channels / 3;
    int best = 0;
    float bestv = 0.0f;
    for (int k = 0; k < K; ++k) {
        float v = out[k] - wn_logf(-wn_logf(u[k]));
        if (k == 0 || v > bestv) { bestv = v; best = k; }
    }
    float mean = out[K + best];
    float ls = out[2 * K + best] > cfg->log_scale_min ? out[2 * K + best] : cfg->log_scale_min;
    float u2 = u[K];
    float x = mean + wn_expf(ls) * (wn_logf(u2) - wn_logf(1.0f - u2));
    x = x > -1.0f ? x : -1.0f;
    return x < 1.0f ? x : 1.0f;
}

static int sample_categorical(const float *logits, int C, float u, float *e) {
    /* inverse CDF with one uniform (oracle definition) */
    float m = logits[0];
    for (int k = 1; k < C; ++k) m = logits[k] > m ? logits[k] : m;
    float S = 0.0f;
    for (int k = 0; k < C; ++k) { e[k] = wn_expf(logits[k] - m); S = S + e[k]; }
    float thr = u * S, acc = 0.0f;
    for (int k = 0; k < C; ++k) {
        acc = acc + e[k];
        if (acc > thr) return k;
    }
    return C - 1;
}

/* ---- the loop, one utterance ------------------------------------------------------------------------------ */
static int run_utterance(const wno_config *cfg, const wno_weights *w, const float *c_up, const float *g_emb,
                         const float *noise, int noise_stride, int NZ, const float *teacher, int T, float *audio,
                         float *samples, float *raw, int shiftqueue) {
    const int L = cfg->layers, R = cfg->residual_channels, G = cfg->gate_channels, S = cfg->skip_channels;
    const int O = cfg->out_channels, k = cfg->kernel_size, H = G / 2, cin = cfg->cin_channels, gin = cfg->gin_channels;
    const int quant = cfg->input_type == WNO_INPUT_MULAW_QUANTIZE;
    size_t qtot = 0;
    int qlen[WNO_MAX_LAYERS];
    size_t qoff[WNO_MAX_LAYERS];
    for (int l = 0; l < L; ++l) {
        qlen[l] = k + (k - 1) * (cfg->dilations[l] - 1); /* wavenet.py:815 */
        qoff[l] = qtot;
        qtot += (size_t)qlen[l] * R;
    }
    float *queues = (float *)calloc(qtot ? qtot : 1, sizeof(float));
    /* gate pre-activation = ONE dot16 over the concatenation [older taps (k-1)*R | c(t) cin | g gin | x(t) R] against
     * the matching rows of the dilated conv / cin 1x1 / gin 1x1 kernels, plus the sum of their biases
     * ((b_conv + b_cin) + b_gin), i.e. modules.py:294-297 + :497-508 with one fixed reduction order. */
    const int cinu = (cin > 0 && c_up) ? cin : 0, ginu = (gin > 0 && g_emb) ? gin : 0;
    const int Kcat = k * R + cinu + ginu;
    const int Nmax = (G > S ? G : S) > O ? (G > S ? G : S) : O;
    const int Kmax = Kcat > S ? Kcat : S;
    float *x = (float *)malloc(sizeof(float) * R), *xcat = (float *)malloc(sizeof(float) * (size_t)Kcat);
    float *pre = (float *)malloc(sizeof(float) * G), *bsum = (float *)malloc(sizeof(float) * (size_t)L * G);
    float *z = (float *)malloc(sizeof(float) * H), *o = (float *)malloc(sizeof(float) * R);
    float *s = (float *)malloc(sizeof(float) * S), *skips = (float *)malloc(sizeof(float) * S);
    float *h1 = (float *)malloc(sizeof(float) * S), *out = (float *)malloc(sizeof(float) * (O > 0 ? O : 1));
    float *e = (float *)malloc(sizeof(float) * (O > 0 ? O : 1));
    float *scratch = (float *)malloc(sizeof(float) * (size_t)WNO_LANES * Nmax);
    const float **rows = (const float **)malloc(sizeof(float *) * (size_t)L * Kmax);
    if (!queues || !x || !xcat || !pre || !bsum || !z || !o || !s || !skips || !h1 || !out || !e || !scratch || !rows)
        return -2;
    for (int l = 0; l < L; ++l) {
        const float **rw = rows + (size_t)l * Kmax;
        int q = 0;
        for (int i = 0; i < (k - 1) * R; ++i) rw[q++] = w->conv_w[l] + (size_t)i * G;
        for (int i = 0; i < cinu; ++i) rw[q++] = w->cin_w[l] + (size_t)i * G;
        for (int i = 0; i < ginu; ++i) rw[q++] = w->gin_w[l] + (size_t)i * G;
        for (int i = (k - 1) * R; i < k * R; ++i) rw[q++] = w->conv_w[l] + (size_t)i * G;
        for (int n = 0; n < G; ++n) {
            float bs = w->conv_b[l] ? w->conv_b[l][n] : 0.0f;
            if (cinu) bs = bs + (w->cin_b[l] ? w->cin_b[l][n] : 0.0f);
            if (ginu) bs = bs + (w->gin_b[l] ? w->gin_b[l][n] : 0.0f);
            bsum[(size_t)l * G + n] = bs;
        }
    }
    const float **hrows = (const float **)malloc(sizeof(float *) * (size_t)(S > 0 ? S : 1));
    if (!hrows) return -2;
    const int nsplit = cfg->ksplit > 0 ? cfg->ksplit : 1;

    /* initial input (wavenet.py:433-445): raw 0.0, mulaw mulaw(0)=0.0, mulaw-quantize class int((0+1)/2*255)=127 */
    float cur_scalar = 0.0f;
    int cur_class = 127;

    for (int t = 0; t < T; ++t) {
        /* first_conv.incremental_step (wavenet.py:826) */
        for (int n = 0; n < R; ++n) {
            float acc = quant ? w->first_w[(size_t)cur_class * R + n]   /* one-hot row select */
                              : fmaf(cur_scalar, w->first_w[n], 0.0f);
            x[n] = acc + (w->first_b ? w->first_b[n] : 0.0f);
        }
        for (int l = 0; l < L; ++l) {
            const int d = cfg->dilations[l];
            float *q = queues + qoff[l];
            float *older = xcat, *cur = xcat + (size_t)(k - 1) * R + cinu + ginu;
            /* CausalConv1D incremental (modules.py:273-297): taps x(t-(k-1)d) ... x(t-d), x(t) */
            if (shiftqueue) {
                memmove(q, q + R, sizeof(float) * (size_t)(qlen[l] - 1) * R);
                memcpy(q + (size_t)(qlen[l] - 1) * R, x, sizeof(float) * R);
                for (int j = 0; j < k - 1; ++j) memcpy(older + (size_t)j * R, q + (size_t)j * d * R, sizeof(float) * R);
            } else {
                memcpy(q + (size_t)(t % qlen[l]) * R, x, sizeof(float) * R);
                for (int j = 0; j < k - 1; ++j) {
                    int tt = t - (k - 1 - j) * d;
                    if (tt >= 0) memcpy(older + (size_t)j * R, q + (size_t)(tt % qlen[l]) * R, sizeof(float) * R);
                    else memset(older + (size_t)j * R, 0, sizeof(float) * R);
                }
            }
            if (cinu) memcpy(xcat + (size_t)(k - 1) * R, c_up + (size_t)t * cin, sizeof(float) * cin);
            if (ginu) memcpy(xcat + (size_t)(k - 1) * R + cinu, g_emb, sizeof(float) * gin);
            memcpy(cur, x, sizeof(float) * R);
            dot16_rows(xcat, Kcat, rows + (size_t)l * Kmax, G, bsum + (size_t)l * G, pre, scratch);
            for (int n = 0; n < H; ++n) z[n] = wn_tanhf(pre[n]) * wn_sigmoidf(pre[H + n]); /* :494, :510 */
            dot_blocked(z, H, w->skip_w[l], S, w->skip_b[l], nsplit, s, scratch);          /* :512 */
            dot_blocked(z, H, w->out_w[l], R, w->out_b[l], nsplit, o, scratch);            /* :515 */
            for (int n = 0; n < R; ++n) {                                                  /* :517-520 */
                float v = o[n] + x[n];
                x[n] = cfg->residual_legacy ? v * WN_SQRT_HALF : v;
            }
            if (l == 0) {
                memcpy(skips, s, sizeof(float) * S);
            } else { /* wavenet.py:833-836 */
                for (int n = 0; n < S; ++n) {
                    float v = skips[n] + s[n];
                    skips[n] = cfg->legacy ? v * WN_SQRT_HALF : v;
                }
            }
        }
        /* head (wavenet.py:840-844) */
        for (int n = 0; n < S; ++n) skips[n] = skips[n] > 0.0f ? skips[n] : 0.0f;
        dot16(skips, S, w->head1_w, S, w->head1_b, h1, scratch, hrows);
        for (int n = 0; n < S; ++n) h1[n] = h1[n] > 0.0f ? h1[n] : 0.0f;
        dot16(h1, S, w->head2_w, O, w->head2_b, out, scratch, hrows);
        if (raw) memcpy(raw + (size_t)t * O, out, sizeof(float) * O);
        /* sampling (wavenet.py:850-869) */
        const float *nz = noise ? noise + (size_t)t * noise_stride : NULL;
        (void)NZ;
        if (!quant) {
            float smp = (O == 2) ? sample_gaussian(cfg, out, nz ? nz[0] : 0.0f) : sample_mol(cfg, out, nz);
            if (samples) samples[t] = smp;
            audio[t] = cfg->input_type == WNO_INPUT_MULAW ? wn_inv_mulawf(smp) : smp; /* wavenet.py:453-456 */
            cur_scalar = teacher ? teacher[t] : smp;                                  /* :877-878 */
        } else {
            int cls = sample_categorical(out, O, nz ? nz[0] : 0.0f, e);
            if (samples) samples[t] = (float)cls;
            audio[t] = wn_inv_mulawf(2.0f * (float)cls / 255.0f - 1.0f); /* :450-452, util.py:105-129 */
            cur_class = teacher ? (int)teacher[t] : cls;
        }
    }
    free(queues); free(x); free(xcat); free(pre); free(bsum); free(z); free(o); free(s); free(skips); free(h1);
    free(out); free(e); free(scratch); free((void *)rows); free((void *)hrows);
    return 0;
}

static int run_all(const wno_config *cfg, const wno_weights *w, const float *c_up, const float *g_emb,
                   const float *noise, const float *teacher, int B, int T, float *audio, float *samples, float *raw,
                   int nthreads, int shiftqueue) {
    if (cfg->layers > WNO_MAX_LAYERS || cfg->layers <= 0) return -1;
    const int O = cfg->out_channels;
    const int NZ = (cfg->input_type != WNO_INPUT_MULAW_QUANTIZE && O != 2) ? O / 3 + 1 : 1;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    int err = 0;
#pragma omp parallel for schedule(dynamic)
    for (int b = 0; b < B; ++b) {
        int r = run_utterance(cfg, w, c_up ? c_up + (size_t)b * T * cfg->cin_channels : NULL,
                              g_emb ? g_emb + (size_t)b * cfg->gin_channels : NULL,
                              noise ? noise + (size_t)b * NZ : NULL, B * NZ, NZ,
                              teacher ? teacher + (size_t)b * T : NULL, T, audio + (size_t)b * T,
                              samples ? samples + (size_t)b * T : NULL, raw ? raw + (size_t)b * T * O : NULL,
                              shiftqueue);
        if (r) err = r;
    }
    return err;
}

int wno_incremental(const wno_config *cfg, const wno_weights *w, const float *c_up, const float *g_emb,
                    const float *noise, const float *teacher, int B, int T, float *audio, float *samples, float *raw,
                    int nthreads) {
    return run_all(cfg, w, c_up, g_emb, noise, teacher, B, T, audio, samples, raw, nthreads, 0);
}

int wno_incremental_shiftqueue(const wno_config *cfg, const wno_weights *w, const float *c_up, const float *g_emb,
                               const float *noise, const float *teacher, int B, int T, float *audio, float *samples,
                               float *raw, int nthreads) {
    return run_all(cfg, w, c_up, g_emb, noise, teacher, B, T, audio, samples, raw, nthreads, 1);
}

void wno_math(int fn, const float *x, float *y, int64_t n) {
    for (int64_t i = 0; i < n; ++i) {
        switch (fn) {
            case 0: y[i] = wn_expf(x[i]); break;
            case 1: y[i] = wn_logf(x[i]); break;
            case 2: y[i] = wn_tanhf(x[i]); break;
            case 3: y[i] = wn_sigmoidf(x[i]); break;
            default: y[i] = wn_inv_mulawf(x[i]); break;
        }
    }
}

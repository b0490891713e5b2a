// wn_loop_generic.cu - "generic" engine of the autoregressive loop (wavenet.py:805-911): one CTA per utterance,
// the whole time loop inside ONE kernel launch, weights streamed from L2, dilation ring buffers in global memory,
// sampling fused, noise read from the caller's buffer.
//
// This engine accepts every configuration the reference's hparams allow (any R/G/S/L/kernel_size, scalar or
// one-hot input, Gaussian / mixture-of-logistics / categorical heads, local and global conditioning).  It is the
// correctness anchor and the fall-back for models whose weights do not fit the shared memory of the pipelined
// persistent engine (wn_loop_pipe.cu), which is the fast path.
//
// Arithmetic: the fixed fp32 orders of DESIGN.md "Arithmetic" (dot16 for the gate pre-activation and the head,
// blocked sequential chains for the out / skip 1x1s), every transcendental from include/wn_math.h - the exact
// operations oracle/c/wn_oracle.c performs, so free-running audio can be compared bit for bit.
#include "wn_common.cuh"

namespace {

constexpr int GEN_THREADS = 512;
constexpr int LANES = WN_DOT_LANES;   // partial sums per dot reduction

// dot16 of xs[0..K) (shared memory, zero padded to a multiple of 4) with row n of Wt ([N][Kpad], global), for the
// outputs n = nb + group.  Every thread of the CTA calls it with the same trip count (shuffles are warp-wide).
// Lane j owns the float4 chunks c = j, j+16, ... ; the xor butterfly adds neighbours first, which is the tree of
// wn_oracle.c:dot16.  Returns the full sum on every lane of the group.
__device__ __forceinline__ float dot16_lane(const float *xs, int K, int Kpad, const float *__restrict__ Wt, int n,
                                            bool valid, int lane) {
    float acc = 0.0f;
    if (valid) {
        const float4 *w4 = reinterpret_cast<const float4 *>(Wt + (long long)n * Kpad);
        const float4 *x4 = reinterpret_cast<const float4 *>(xs);
        const int nfull = K >> 2;
        // weights stream from L2: issue a batch of independent loads before the dependent fma chain consumes them
        constexpr int UB = 8;
        int c = lane;
        for (; c + (UB - 1) * LANES < nfull; c += UB * LANES) {
            float4 wv[UB];
#pragma unroll
            for (int u = 0; u < UB; ++u) wv[u] = __ldg(w4 + c + u * LANES);
#pragma unroll
            for (int u = 0; u < UB; ++u) {
                const float4 xv = x4[c + u * LANES];
                acc = __fmaf_rn(xv.x, wv[u].x, acc);
                acc = __fmaf_rn(xv.y, wv[u].y, acc);
                acc = __fmaf_rn(xv.z, wv[u].z, acc);
                acc = __fmaf_rn(xv.w, wv[u].w, acc);
            }
        }
        for (; c < nfull; c += LANES) {
            const float4 wv = __ldg(w4 + c);
            const float4 xv = x4[c];
            acc = __fmaf_rn(xv.x, wv.x, acc);
            acc = __fmaf_rn(xv.y, wv.y, acc);
            acc = __fmaf_rn(xv.z, wv.z, acc);
            acc = __fmaf_rn(xv.w, wv.w, acc);
        }
        if ((K & 3) && (nfull % LANES) == lane) {   // ragged tail chunk: only the elements below K
            const float *wr = Wt + (long long)n * Kpad;
            for (int k = nfull * 4; k < K; ++k) acc = __fmaf_rn(xs[k], __ldg(wr + k), acc);
        }
    }
#pragma unroll
    for (int s = 1; s < LANES; s <<= 1) acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, s));
    return acc;
}

// blocked sequential chains (wn_oracle.c:dot_blocked): y = ((o_0 + o_1) + ...) + b, o_p over k in [K*p/P, K*(p+1)/P).
__device__ __forceinline__ float dot_blocked_col(const float *__restrict__ zs, int K, const float *__restrict__ W,
                                                 int N, int n, int nsplit, float bias) {
    float y = 0.0f;
    for (int p = 0; p < nsplit; ++p) {
        const int k0 = (int)((long long)K * p / nsplit), k1 = (int)((long long)K * (p + 1) / nsplit);
        float acc = 0.0f;
        constexpr int UB = 16;   // independent loads in flight per thread; the chain itself stays sequential in k
        int k = k0;
        for (; k + UB <= k1; k += UB) {
            float wv[UB];
#pragma unroll
            for (int u = 0; u < UB; ++u) wv[u] = __ldg(W + (long long)(k + u) * N + n);
#pragma unroll
            for (int u = 0; u < UB; ++u) acc = __fmaf_rn(zs[k + u], wv[u], acc);
        }
        for (; k < k1; ++k) acc = __fmaf_rn(zs[k], __ldg(W + (long long)k * N + n), acc);
        y = p == 0 ? acc : __fadd_rn(y, acc);
    }
    return __fadd_rn(y, bias);
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

__device__ __forceinline__ int round4(int v) { return (v + 3) & ~3; }

// grid = B utterances, block = GEN_THREADS.  Dynamic shared memory carved below.
__global__ void __launch_bounds__(GEN_THREADS) wn_loop_generic_kernel(
    const WnModelDev m, const float *__restrict__ c_up, const float *__restrict__ g_emb,
    const float *__restrict__ noise, const float *__restrict__ teacher, int inputs_mode, int B, int T,
    float *__restrict__ rings, float *__restrict__ audio, float *__restrict__ raw) {
    extern __shared__ __align__(16) float sm[];
    const int tid = threadIdx.x, NT = blockDim.x, b = blockIdx.x;
    const int lane = tid & (LANES - 1), group = tid / LANES, ngroups = NT / LANES;
    const int R = m.R, G = m.G, S = m.S, O = m.O, H = m.H, ks = m.ksize;
    float *x = sm;                            // [round4(R)]  current residual-stream vector
    float *xcat = x + round4(R);              // [KcatPad]    older taps | c(t) | g | x(t)
    float *pre = xcat + m.KcatPad;            // [round4(G)]
    float *z = pre + round4(G);               // [round4(H)]
    float *skips = z + round4(H);             // [Spad]
    float *h1 = skips + m.Spad;               // [Spad]
    float *outv = h1 + m.Spad;                // [round4(O)]
    float *ev = outv + round4(O);             // [round4(O)]  categorical scratch
    float *fb = ev + round4(O);               // [4]: fb[0] next scalar input, fb[1] next class
    const bool quant = m.input_type == WN_INPUT_MULAW_QUANTIZE;
    float *ring = rings + (long long)b * m.ring_per_utt;
    const int o_cond = (ks - 1) * R, o_g = o_cond + m.cin, o_cur = o_g + m.gin;

    if (tid == 0) {  // wavenet.py:433-445: raw/mulaw 0.0; class int((0+1)/2*255)=127
        fb[0] = 0.0f;
        fb[1] = 127.0f;
        // inputs_mode (WaveNet.step semantics, wavenet.py:650-721): teacher[t] IS the network input of step t
        if (inputs_mode) fb[0] = fb[1] = teacher[(long long)b * T];
    }
    for (int i = tid; i < m.KcatPad; i += NT) xcat[i] = 0.0f;
    for (int i = tid; i < 2 * m.Spad; i += NT) skips[i] = 0.0f;   // skips and h1 (padding must stay zero)
    __syncthreads();
    for (int i = tid; i < m.gin; i += NT) xcat[o_g + i] = g_emb ? g_emb[(long long)b * m.gin + i] : 0.0f;
    __syncthreads();

    for (int t = 0; t < T; ++t) {
        // ---- first_conv.incremental_step (wavenet.py:826) ------------------------------------------------
        for (int n = tid; n < R; n += NT) {
            float acc;
            if (quant) acc = __ldg(m.first_w + (long long)((int)fb[1]) * R + n);
            else acc = __fmaf_rn(fb[0], __ldg(m.first_w + n), 0.0f);
            x[n] = __fadd_rn(acc, __ldg(m.first_b + n));
        }
        for (int i = tid; i < m.cin; i += NT) xcat[o_cond + i] = __ldg(c_up + ((long long)b * T + t) * m.cin + i);
        __syncthreads();

        for (int l = 0; l < m.L; ++l) {
            const float *lw = m.layers + (long long)l * m.layer_stride;
            const int d = m.dil[l], ql = m.qlen[l];
            float *rq = ring + m.qoff[l];
            // ---- CausalConv1D incremental: push x_t, gather the k taps (modules.py:273-292) ----------------
            for (int i = tid; i < ks * R; i += NT) {
                const int j = i / R, n = i - j * R;
                const int tt = t - (ks - 1 - j) * d;
                if (j == ks - 1) { const float v = x[n]; rq[(long long)(t % ql) * R + n] = v; xcat[o_cur + n] = v; }
                else xcat[i] = tt >= 0 ? rq[(long long)(tt % ql) * R + n] : 0.0f;
            }
            __syncthreads();
            // ---- gate pre-activation: one dot16 over [taps | c | g | x(t)] (modules.py:294-297, :497-508) ----
            for (int nb = 0; nb < G; nb += ngroups) {
                const int n = nb + group;
                const float a = dot16_lane(xcat, m.Kcat, m.KcatPad, lw + m.off_cat_w, n, n < G, lane);
                if (lane == 0 && n < G) pre[n] = __fadd_rn(a, __ldg(lw + m.off_bsum + n));
            }
            __syncthreads();
            for (int n = tid; n < H; n += NT) z[n] = __fmul_rn(wn_tanhf(pre[n]), wn_sigmoidf(pre[H + n]));  // :510
            __syncthreads();
            // ---- skip and residual 1x1s (modules.py:512-520, wavenet.py:833-836) ---------------------------
            for (int n = tid; n < S + R; n += NT) {
                if (n < S) {
                    float s = dot_blocked_col(z, H, lw + m.off_skip_w, S, n, m.ksplit, __ldg(lw + m.off_skip_b + n));
                    if (l == 0) skips[n] = s;
                    else {
                        float v = __fadd_rn(skips[n], s);
                        skips[n] = m.legacy ? __fmul_rn(v, WN_SQRT_HALF) : v;
                    }
                } else {
                    const int r = n - S;
                    float o = dot_blocked_col(z, H, lw + m.off_out_w, R, r, m.ksplit, __ldg(lw + m.off_out_b + r));
                    float v = __fadd_rn(o, x[r]);
                    x[r] = m.residual_legacy ? __fmul_rn(v, WN_SQRT_HALF) : v;
                }
            }
            __syncthreads();
        }
        // ---- head (wavenet.py:840-844) ---------------------------------------------------------------------
        for (int n = tid; n < S; n += NT) skips[n] = skips[n] > 0.0f ? skips[n] : 0.0f;
        __syncthreads();
        for (int nb = 0; nb < S; nb += ngroups) {
            const int n = nb + group;
            const float a = dot16_lane(skips, S, m.Spad, m.h1_w, n, n < S, lane);
            if (lane == 0 && n < S) { const float v = __fadd_rn(a, __ldg(m.h1_b + n)); h1[n] = v > 0.0f ? v : 0.0f; }
        }
        __syncthreads();
        for (int nb = 0; nb < O; nb += ngroups) {
            const int n = nb + group;
            const float a = dot16_lane(h1, S, m.Spad, m.h2_w, n, n < O, lane);
            if (lane == 0 && n < O) {
                const float v = __fadd_rn(a, __ldg(m.h2_b + n));
                outv[n] = v;
                if (raw) raw[((long long)b * T + t) * O + n] = v;  // wavenet.py:847
            }
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
    auto r4 = [](int v) { return (v + 3) & ~3; };
    size_t smem = sizeof(float) * (size_t)(r4(m.R) + m.KcatPad + r4(m.G) + r4(m.H) + 2 * m.Spad + 2 * r4(m.O) + 4);
    if (smem > 48 * 1024)
        WN_CUDA_TRY(h, cudaFuncSetAttribute(wn_loop_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    wn_loop_generic_kernel<<<B, GEN_THREADS, smem, st>>>(m, c_up, g_emb, noise, teacher, inputs_mode, B, T, h->d_ring,
                                                         audio, raw);
    h->launches++;
    WN_CUDA_TRY(h, cudaGetLastError());
    return WN_OK;
}

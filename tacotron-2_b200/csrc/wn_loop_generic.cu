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
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "wn_common.cuh"

namespace {

constexpr int GEN_THREADS = 512;
#ifndef GEN_UB_LOWREG
#define GEN_UB_LOWREG 4   // independent 16-byte weight loads per round in the 64-register variant (8 spills there)
#endif
constexpr int LANES = WN_DOT_LANES;   // partial sums per dot reduction

// dot16 of xs[0..K) (shared memory, zero padded to a multiple of 4) with row n of Wt ([N][Kpad], global), for the
// outputs n = nb + group.  Every thread of the CTA calls it with the same trip count (shuffles are warp-wide).
// Lane j owns the float4 chunks c = j, j+16, ... ; the xor butterfly adds neighbours first, which is the tree of
// wn_oracle.c:dot16.  Returns the full sum on every lane of the group.
template <int UB = 8>
__device__ __forceinline__ float dot16_lane(const float *xs, int K, int Kpad, const float *__restrict__ Wt, int n,
                                            bool valid, int lane) {
    float acc = 0.0f;
    if (valid) {
        const float4 *w4 = reinterpret_cast<const float4 *>(Wt + (long long)n * Kpad);
        const float4 *x4 = reinterpret_cast<const float4 *>(xs);
        const int nfull = K >> 2;
        // weights stream from L2: a round issues up to UB independent loads (predicated at the end of the row) before the
        // dependent fma chain consumes them, so a row costs ceil(K / (64 UB)) L2 round trips
        for (int c = lane; c < nfull; c += UB * LANES) {
            float4 wv[UB];
#pragma unroll
            for (int u = 0; u < UB; ++u)
                if (c + u * LANES < nfull) wv[u] = __ldg(w4 + c + u * LANES);
#pragma unroll
            for (int u = 0; u < UB; ++u)
                if (c + u * LANES < nfull) {
                    const float4 xv = x4[c + u * LANES];
                    acc = __fmaf_rn(xv.x, wv[u].x, acc);
                    acc = __fmaf_rn(xv.y, wv[u].y, acc);
                    acc = __fmaf_rn(xv.z, wv[u].z, acc);
                    acc = __fmaf_rn(xv.w, wv[u].w, acc);
                }
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

// Two rows (nA, nB) per call: the loads of both rows are issued in the same round, which halves the L2 round trips of a
// CTA that owns more rows than it has lane groups.  Each row's arithmetic is exactly dot16_lane's.
__device__ __forceinline__ void dot16_lane2(const float *xs, int K, int Kpad, const float *__restrict__ Wt, int nA, bool vA,
                                            int nB, bool vB, int lane, float &outA, float &outB) {
    float accA = 0.0f, accB = 0.0f;
    const float4 *wA = reinterpret_cast<const float4 *>(Wt + (long long)(vA ? nA : 0) * Kpad);
    const float4 *wB = reinterpret_cast<const float4 *>(Wt + (long long)(vB ? nB : 0) * Kpad);
    const float4 *x4 = reinterpret_cast<const float4 *>(xs);
    const int nfull = K >> 2;
    constexpr int UB = 8;
    for (int c = lane; c < nfull; c += UB * LANES) {
        float4 a[UB], b[UB];
#pragma unroll
        for (int u = 0; u < UB; ++u)
            if (c + u * LANES < nfull) {
                if (vA) a[u] = __ldg(wA + c + u * LANES);
                if (vB) b[u] = __ldg(wB + c + u * LANES);
            }
#pragma unroll
        for (int u = 0; u < UB; ++u)
            if (c + u * LANES < nfull) {
                const float4 xv = x4[c + u * LANES];
                if (vA) { accA = __fmaf_rn(xv.x, a[u].x, accA); accA = __fmaf_rn(xv.y, a[u].y, accA); accA = __fmaf_rn(xv.z, a[u].z, accA); accA = __fmaf_rn(xv.w, a[u].w, accA); }
                if (vB) { accB = __fmaf_rn(xv.x, b[u].x, accB); accB = __fmaf_rn(xv.y, b[u].y, accB); accB = __fmaf_rn(xv.z, b[u].z, accB); accB = __fmaf_rn(xv.w, b[u].w, accB); }
            }
    }
    if ((K & 3) && (nfull % LANES) == lane) {   // ragged tail chunk: only the elements below K
        for (int k = nfull * 4; k < K; ++k) {
            if (vA) accA = __fmaf_rn(xs[k], __ldg(Wt + (long long)nA * Kpad + k), accA);
            if (vB) accB = __fmaf_rn(xs[k], __ldg(Wt + (long long)nB * Kpad + k), accB);
        }
    }
#pragma unroll
    for (int s = 1; s < LANES; s <<= 1) {
        accA = __fadd_rn(accA, __shfl_xor_sync(0xffffffffu, accA, s));
        accB = __fadd_rn(accB, __shfl_xor_sync(0xffffffffu, accB, s));
    }
    outA = accA; outB = accB;
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

// the same chains with the weights of this CTA's rows staged in shared memory, one row per output: Wrow[k], k contiguous
// (rows are WOS_PAD floats longer than K, so that the 16-byte loads of consecutive threads fall into different banks).
// Four k per load instruction: ptxas, short of registers in this kernel, issues every load right before its fma, so the
// chain pays one shared-memory latency per LOAD - with scalar loads that was 33 cycles per k.
constexpr int WOS_PAD = 4;
__device__ __forceinline__ float dot_blocked_smem(const float *__restrict__ zs, int K, const float *__restrict__ Wrow, int nsplit,
                                                  float bias) {
    float y = 0.0f;
    for (int p = 0; p < nsplit; ++p) {
        const int k0 = (int)((long long)K * p / nsplit), k1 = (int)((long long)K * (p + 1) / nsplit);
        float acc = 0.0f;
        int k = k0;
        for (; k < k1 && (k & 3); ++k) acc = __fmaf_rn(zs[k], Wrow[k], acc);
        for (; k + 4 <= k1; k += 4) {
            const float4 w = *reinterpret_cast<const float4 *>(Wrow + k), z = *reinterpret_cast<const float4 *>(zs + k);
            acc = __fmaf_rn(z.x, w.x, acc);
            acc = __fmaf_rn(z.y, w.y, acc);
            acc = __fmaf_rn(z.z, w.z, acc);
            acc = __fmaf_rn(z.w, w.w, acc);
        }
        for (; k < k1; ++k) acc = __fmaf_rn(zs[k], Wrow[k], acc);
        y = p == 0 ? acc : __fadd_rn(y, acc);
    }
    return __fadd_rn(y, bias);
}
__device__ __forceinline__ void cp_async16(float *dst_smem, const float *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
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

// ---- thread-block cluster: the C CTAs of one utterance split the rows of every inner product and exchange the results
//      through distributed shared memory ------------------------------------------------------------------------------
__device__ __forceinline__ unsigned cluster_rank() { unsigned r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ unsigned cluster_size() { unsigned r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r; }
// v -> the same shared-memory variable of every CTA of the cluster (this one included)
__device__ __forceinline__ void bcast_f32(float *p, float v, int C) {
    if (C == 1) { *p = v; return; }
    const unsigned a = (unsigned)__cvta_generic_to_shared(p);
    for (int q = 0; q < C; ++q) {
        unsigned ra;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(a), "r"(q));
        asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(ra), "f"(v) : "memory");
    }
}
// every thread of every CTA of the cluster; orders the remote stores above before the reads below
__device__ __forceinline__ void cluster_barrier(int C) {
    if (C == 1) { __syncthreads(); return; }
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ int slice_lo(int n, int q, int C) { return (int)((long long)n * q / C); }

// grid = B utterances x C CTAs (one cluster per utterance), block = GEN_THREADS.  Every CTA holds the utterance's small
// state (x, the concatenated operand, z, skips, ...) in its own shared memory; CTA `cr` of the cluster computes the gate
// channels [H cr / C, H (cr+1) / C) (tanh and sigmoid rows), the rows [(S+R) cr / C, ...) of the skip / out 1x1s and the
// rows [S cr / C, ...) of the first head 1x1, and writes its results into all C copies.  The arithmetic of a row does not
// depend on who computes it, so every C gives the same bits.  Two cluster barriers per layer and one in the head.
// Dynamic shared memory carved below.
// MINB = CTAs per SM the register budget allows: 1 -> 128 registers (two gate rows per lane group and round), 2 -> 64
// registers (one row per round), which lets clusters share SMs when one CTA per SM does not keep every utterance resident.
template <int MINB>
__global__ void __launch_bounds__(GEN_THREADS, MINB) wn_loop_generic_kernel(
    const WnModelDev m, const float *__restrict__ c_up, const float *__restrict__ g_emb,
    const float *__restrict__ noise, const float *__restrict__ teacher, int inputs_mode, int B, int T,
    float *__restrict__ rings, float *__restrict__ audio, float *__restrict__ raw, int wos_floats) {
    extern __shared__ __align__(16) float sm[];
    const int C = (int)cluster_size(), cr = (int)cluster_rank();
    const int tid = threadIdx.x, NT = blockDim.x, b = blockIdx.x / C;
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
    float *wos = fb + 4;                      // [rows of this CTA][H + WOS_PAD] skip / out weights of the current layer (if they fit)
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
            // this CTA's rows of [skip ; out], requested now (cp.async) and used after the gate: the blocked chains are
            // sequential in k, so reading their weights from L2 on demand costs one round trip per 16 k
            const int q0 = slice_lo(S + R, cr, C), q1 = slice_lo(S + R, cr + 1, C), nq = q1 - q0;
            const int wld = H + WOS_PAD;                           // row stride of the staged copy
            const bool staged = nq * wld <= wos_floats && !(H & 3);
            if (staged) {   // rows [q0, q1) of the transposed [skip ; out] copy of the blob -> wos[row - q0][k], 16 bytes per copy
                const int H4 = H >> 2;
                const float *src = lw + m.off_os_t + (size_t)q0 * m.Hpad;
                if (NT % H4 == 0) {   // a thread keeps its k and walks the rows with a constant stride: no division per copy
                    const int k4 = tid % H4, jstep = NT / H4;
                    for (int j = tid / H4; j < nq; j += jstep) cp_async16(wos + (size_t)j * wld + 4 * k4, src + (size_t)j * m.Hpad + 4 * k4);
                } else {
                    for (int i = tid; i < nq * H4; i += NT) {
                        const int j = i / H4, k4 = i - j * H4;
                        cp_async16(wos + (size_t)j * wld + 4 * k4, src + (size_t)j * m.Hpad + 4 * k4);
                    }
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
            }
            // ---- CausalConv1D incremental: push x_t, gather the k taps (modules.py:273-292) ----------------
            if (!(R & 3) && !(o_cur & 3)) {   // 16-byte pieces: one round of loads for the usual sizes
                const int R4 = R >> 2;
                float4 *xc4 = reinterpret_cast<float4 *>(xcat), *rq4 = reinterpret_cast<float4 *>(rq);
                for (int i = tid; i < ks * R4; i += NT) {
                    const int j = i / R4, n4 = i - j * R4;
                    const int tt = t - (ks - 1 - j) * d;
                    if (j == ks - 1) {
                        const float4 v = reinterpret_cast<const float4 *>(x)[n4];
                        rq4[(long long)(t % ql) * R4 + n4] = v;
                        xc4[(o_cur >> 2) + n4] = v;
                    } else {
                        xc4[i] = tt >= 0 ? rq4[(long long)(tt % ql) * R4 + n4] : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                }
            } else
            for (int i = tid; i < ks * R; i += NT) {
                const int j = i / R, n = i - j * R;
                const int tt = t - (ks - 1 - j) * d;
                if (j == ks - 1) { const float v = x[n]; rq[(long long)(t % ql) * R + n] = v; xcat[o_cur + n] = v; }
                else xcat[i] = tt >= 0 ? rq[(long long)(tt % ql) * R + n] : 0.0f;
            }
            __syncthreads();
            // ---- gate pre-activation: one dot16 over [taps | c | g | x(t)] (modules.py:294-297, :497-508); this CTA's
            //      channels: rows h0.. (tanh half) and H + h0.. (sigmoid half)
            const int h0 = slice_lo(H, cr, C), hn = slice_lo(H, cr + 1, C) - h0;
            if (MINB == 1) {
                for (int ib = 0; ib < 2 * hn; ib += 2 * ngroups) {
                    const int iA = ib + group, iB = iA + ngroups;
                    const bool vA = iA < 2 * hn, vB = iB < 2 * hn;
                    const int nA = iA < hn ? h0 + iA : H + h0 + (iA - hn), nB = iB < hn ? h0 + iB : H + h0 + (iB - hn);
                    float a, b2;
                    dot16_lane2(xcat, m.Kcat, m.KcatPad, lw + m.off_cat_w, nA, vA, nB, vB, lane, a, b2);
                    if (lane == 0 && vA) pre[nA] = __fadd_rn(a, __ldg(lw + m.off_bsum + nA));
                    if (lane == 0 && vB) pre[nB] = __fadd_rn(b2, __ldg(lw + m.off_bsum + nB));
                }
            } else {
                for (int ib = 0; ib < 2 * hn; ib += ngroups) {
                    const int i = ib + group;
                    const bool valid = i < 2 * hn;
                    const int n = i < hn ? h0 + i : H + h0 + (i - hn);
                    const float a = dot16_lane<GEN_UB_LOWREG>(xcat, m.Kcat, m.KcatPad, lw + m.off_cat_w, n, valid, lane);
                    if (lane == 0 && valid) pre[n] = __fadd_rn(a, __ldg(lw + m.off_bsum + n));
                }
            }
            __syncthreads();
            for (int i = tid; i < hn; i += NT) {   // :510
                const int n = h0 + i;
                bcast_f32(z + n, __fmul_rn(wn_tanhf(pre[n]), wn_sigmoidf(pre[H + n])), C);
            }
            if (staged) asm volatile("cp.async.wait_group 0;" ::: "memory");
            cluster_barrier(C);
            // ---- skip and residual 1x1s (modules.py:512-520, wavenet.py:833-836): this CTA's rows of [skip ; out] ------
            for (int n = q0 + tid; n < q1; n += NT) {
                if (n < S) {
                    float s = staged ? dot_blocked_smem(z, H, wos + (size_t)(n - q0) * wld, m.ksplit, __ldg(lw + m.off_skip_b + n))
                                     : dot_blocked_col(z, H, lw + m.off_skip_w, S, n, m.ksplit, __ldg(lw + m.off_skip_b + n));
                    if (l > 0) {
                        const float v = __fadd_rn(skips[n], s);
                        s = m.legacy ? __fmul_rn(v, WN_SQRT_HALF) : v;
                    }
                    if (l == m.L - 1) bcast_f32(skips + n, s, C);   // the head needs the whole sum
                    else skips[n] = s;                              // the running sum stays with its owner
                } else {
                    const int r = n - S;
                    float o = staged ? dot_blocked_smem(z, H, wos + (size_t)(n - q0) * wld, m.ksplit, __ldg(lw + m.off_out_b + r))
                                     : dot_blocked_col(z, H, lw + m.off_out_w, R, r, m.ksplit, __ldg(lw + m.off_out_b + r));
                    float v = __fadd_rn(o, x[r]);
                    bcast_f32(x + r, m.residual_legacy ? __fmul_rn(v, WN_SQRT_HALF) : v, C);
                }
            }
            cluster_barrier(C);
        }
        // ---- head (wavenet.py:840-844) ---------------------------------------------------------------------
        for (int n = tid; n < S; n += NT) skips[n] = skips[n] > 0.0f ? skips[n] : 0.0f;
        __syncthreads();
        const int s0 = slice_lo(S, cr, C), s1 = slice_lo(S, cr + 1, C);
        for (int nb = s0; nb < s1; nb += ngroups) {
            const int n = nb + group;
            const float a = dot16_lane(skips, S, m.Spad, m.h1_w, n, n < s1, lane);
            if (lane == 0 && n < s1) { const float v = __fadd_rn(a, __ldg(m.h1_b + n)); bcast_f32(h1 + n, v > 0.0f ? v : 0.0f, C); }
        }
        cluster_barrier(C);   // (every CTA evaluates the small output 1x1 and the sampler itself: no feedback broadcast)
        for (int nb = 0; nb < O; nb += ngroups) {
            const int n = nb + group;
            const float a = dot16_lane(h1, S, m.Spad, m.h2_w, n, n < O, lane);
            if (lane == 0 && n < O) {
                const float v = __fadd_rn(a, __ldg(m.h2_b + n));
                outv[n] = v;
                if (raw && cr == 0) raw[((long long)b * T + t) * O + n] = v;  // wavenet.py:847
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
                if (audio && cr == 0) audio[(long long)b * T + t] = m.input_type == WN_INPUT_MULAW ? wn_inv_mulawf(smp) : smp;
                fb[0] = teacher ? (ti < T ? teacher[(long long)b * T + ti] : 0.0f) : smp;
            } else {
                const float *nz = noise + ((long long)t * B + b) * m.NZ;
                int cls = sample_categorical(outv, O, nz[0], ev);
                float yq = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, (float)cls), 255.0f), 1.0f);
                if (audio && cr == 0) audio[(long long)b * T + t] = wn_inv_mulawf(yq);
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
    // CTAs per utterance: the largest power of two (at most 8, the portable cluster size) for which all B clusters are
    // resident at once (cudaOccupancyMaxActiveClusters: clusters live inside one GPC); WN_GEN_CLUSTER overrides.  One
    // CTA per utterance streams every weight from L2 itself (the paper model: 54.7 MB per step and CTA); C CTAs stream
    // 1 / C each.  The CTA's rows of the skip / out 1x1s of one layer are staged in shared memory if that fits next to
    // the state (else they stream from L2 as well).
    const size_t smem_state = smem;
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute at[1];
    int wos_floats = 0;
    const void *kern = nullptr;
    auto configure = [&](int C, int minb) {
        kern = minb == 1 ? (const void *)wn_loop_generic_kernel<1> : (const void *)wn_loop_generic_kernel<2>;
        wos_floats = (int)(((size_t)(m.S + m.R + C - 1) / C + 1) * (size_t)(m.H + WOS_PAD));
        if (smem_state + sizeof(float) * (size_t)wos_floats > (size_t)(minb == 1 ? 160 : 100) * 1024) wos_floats = 0;
        smem = smem_state + sizeof(float) * (size_t)wos_floats;
        cfg.gridDim = dim3((unsigned)(B * C)); cfg.blockDim = dim3(GEN_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = (unsigned)C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(smem, 48 * 1024)) != cudaSuccess) { cudaGetLastError(); return 0; }
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
        return n;
    };
    // Largest cluster first; per size the 128-register variant (one CTA per SM), then the 64-register one that lets
    // clusters share SMs (paper block, 16 utterances: 15 clusters of 8 fit with one CTA per SM, 33 with two; 333 us/step
    // against 348 with clusters of 4).  WN_GEN_CLUSTER / WN_GEN_MINB force the choice.
    int C = 1, minb = 1;
    const char *eC = getenv("WN_GEN_CLUSTER"), *eM = getenv("WN_GEN_MINB");
    if (eC || eM) {
        if (eC) C = atoi(eC);
        if (C != 1 && C != 2 && C != 4 && C != 8) C = 1;
        minb = eM && atoi(eM) == 2 ? 2 : 1;
        if (!eC)
            for (C = 8; C > 1; C >>= 1)
                if (configure(C, minb) >= B) break;
    } else {
        bool found = false;
        for (int c = 8; c > 1 && !found; c >>= 1)
            for (int mb = 1; mb <= 2 && !found; ++mb)
                if (configure(c, mb) >= B) { C = c; minb = mb; found = true; }
    }
    const int resident = configure(C, minb);
    if (getenv("WN_GEN_DEBUG")) fprintf(stderr, "generic engine: B=%d cluster=%d CTAs/SM=%d resident clusters=%d smem=%zu staged=%d\n", B, C, minb, resident, smem, wos_floats != 0);
    if (resident < 1) { h->err = "generic engine: the cluster does not fit the device"; return WN_ERR_CUDA; }
    {
        const WnModelDev mm = m;
        float *ring_p = h->d_ring;
        void *args[] = {(void *)&mm, (void *)&c_up, (void *)&g_emb, (void *)&noise, (void *)&teacher, (void *)&inputs_mode, (void *)&B, (void *)&T,
                        (void *)&ring_p, (void *)&audio, (void *)&raw, (void *)&wos_floats};
        WN_CUDA_TRY(h, cudaLaunchKernelExC(&cfg, kern, args));
    }
    h->launches++;
    WN_CUDA_TRY(h, cudaGetLastError());
    return WN_OK;
}

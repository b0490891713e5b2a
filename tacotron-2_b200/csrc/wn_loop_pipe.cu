// wn_loop_pipe.cu - the pipelined persistent engine of the autoregressive loop (wavenet.py:805-911) for sm_100a.
//
// One cooperative launch runs the whole utterance batch for all T steps.  The network is laid out ACROSS the SMs:
//   * every residual layer l is owned by P CTAs ("layer CTAs", one per SM); CTA (l,p) keeps ITS slice of the layer's
//     weights resident in shared memory for the whole launch: the rows of the dilated conv + cin 1x1 that produce gate
//     channels [h0,h1) (both the tanh and the sigmoid half, so gating is local), and the rows of the residual / skip
//     1x1s that produce output channels [r0,r1) / [s0,s1);
//   * NH "head CTAs" own the two head 1x1s, the first (input) 1x1 and the sampler.
// Utterances travel through that pipeline in groups of NU.  All stages walk the same schedule (t major, group minor),
// so while layer 7 works on group 3 layer 6 works on group 4: the serial chain of one group costs ~(L+1) stage
// latencies per sample, but all SMs stay busy on different groups.  The only feedback edge is sample(t) -> input(t+1).
//
// Hand-off between CTAs never uses a barrier or a separate flag: every float travels as an 8-byte {value, tag} word
// (tag = t+1) written with one 64-bit store and polled with 64-bit loads by the consumer ("LL" protocol), so a value
// is usable the moment it lands in L2.  Three kinds of mailboxes, all in global memory (L2 resident when small):
//   X[l]  the input vector of layer l, [group][slot][u][R].  It has 2*d_l+1 slots and so doubles as the dilation
//         ring buffer: taps x(t-d), x(t-2d) are re-read from older slots (modules.py:285-292 without the queue shift);
//   Z[l]  the gated activations of layer l, exchanged between the P CTAs of the layer (each needs all of z);
//   S[l]  the running skip sum after layer l, slice p -> slice p of layer l+1, and the full vector to the head.
// Everything that does not depend on x_l(t) - the two older taps, the conditioning 1x1 - is prefetched into registers
// one work item ahead and accumulated before x_l(t) arrives, so the per-layer critical path is one K=R fma chain,
// the gate, the z exchange and one K=G/2 chain.
//
// Arithmetic is the same fixed order as wn_loop_generic.cu and oracle/c/wn_oracle.c (sequential fp32 fma chains over
// k, one bias add, include/wn_math.h), so results are bit-identical to both.
//
// Scope: scalar input (raw / mulaw), Gaussian or mixture-of-logistics head, kernel_size 3, local conditioning on,
// global conditioning off, fp32 weights that fit shared memory (the default config uses 165 KB per layer CTA).
// Anything else runs on the generic engine.
#include <cooperative_groups.h>

#include <algorithm>

#include "wn_common.cuh"

namespace {

constexpr int PIPE_THREADS = 256;
constexpr int MAX_PF = 4;            // 16-byte prefetch loads per thread for the two older taps
constexpr unsigned SPIN_LIMIT = 1u << 24;

struct PipeCta {
    int role;                        // 0 layer, 1 head
    int l, p;                        // layer / slice (head: index of the head CTA in l)
    int h0, hn, r0, rn, s0, sn;      // gate / residual / skip channel slices owned
    long long blob;                  // float offset of this CTA's packed weights
};

struct PipeParams {
    // model
    int L, R, G, S, O, H, cin, NZ, input_type, legacy, residual_legacy;
    float log_scale_min, log_scale_min_gauss;
    int dil[WN_MAX_LAYERS], depth[WN_MAX_LAYERS];
    long long xoff[WN_MAX_LAYERS];   // uint64 offset of X[l] inside xbuf
    // plan
    int P, NH, NU, NG;
    int hmax, rmax, smax;            // largest slice sizes (uniform smem layout)
    int KC, KCs, KI, KIs, KH, KHs, KS, KSs;   // chain lengths (padded to 4) and padded row strides
    int layer_blob_floats, head_blob_floats;
    const float *blobs;
    const PipeCta *ctas;
    unsigned long long *xbuf, *zbuf, *sbuf;
    int *abort_flag;
    // call
    const float *c_up, *noise, *teacher;
    float *audio, *raw;
    int B, T;
};

__device__ __forceinline__ unsigned long long pack_ll(float v, unsigned tag) {
    return ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v);
}
__device__ __forceinline__ void st_ll(unsigned long long *p, float v, unsigned tag) {
    *(volatile unsigned long long *)p = pack_ll(v, tag);
}
__device__ __forceinline__ void ld_ll2(const unsigned long long *p, unsigned long long &a, unsigned long long &b) {
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p));
}
__device__ __forceinline__ unsigned long long ld_ll1(const unsigned long long *p) {
    return *(const volatile unsigned long long *)p;
}
__device__ __forceinline__ void ldcg2(const unsigned long long *p, unsigned long long &a, unsigned long long &b) {
    asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p));
}
__device__ __forceinline__ float ll_val(unsigned long long w) { return __uint_as_float((unsigned)(w & 0xffffffffu)); }
__device__ __forceinline__ unsigned ll_tag(unsigned long long w) { return (unsigned)(w >> 32); }

// Poll `npairs` consecutive {value,tag} words starting at src until all carry `tag`.  The words form rows of
// `rowlen` (even) values; value i goes to dst[(i / rowlen) * dstride + i % rowlen] (shared memory).  Each thread owns
// 16-byte chunks.  Returns nonzero if the watchdog fired (a peer CTA never delivered: never expected).
__device__ __forceinline__ int poll_block(const unsigned long long *src, int npairs, unsigned tag, float *dst,
                                          int rowlen, int dstride, int *abort_flag) {
    int bad = 0;
    for (int i = threadIdx.x * 2; i < npairs; i += PIPE_THREADS * 2) {
        unsigned long long a, b;
        unsigned spins = 0;
        for (;;) {
            ld_ll2(src + i, a, b);
            if (ll_tag(a) == tag && ll_tag(b) == tag) break;
            if ((++spins & 0x3ffu) == 0) {
                if (*(volatile int *)abort_flag) { bad = 1; break; }
                if (spins > SPIN_LIMIT) { atomicExch(abort_flag, 1); bad = 1; break; }
            }
        }
        const int row = i / rowlen, col = i - row * rowlen;
        dst[row * dstride + col] = ll_val(a);
        dst[row * dstride + col + 1] = ll_val(b);
    }
    return bad;
}

__device__ __forceinline__ int poll_one(const unsigned long long *src, unsigned tag, float &v, int *abort_flag) {
    unsigned spins = 0;
    for (;;) {
        unsigned long long a = ld_ll1(src);
        if (ll_tag(a) == tag) { v = ll_val(a); return 0; }
        if ((++spins & 0x3ffu) == 0) {
            if (*(volatile int *)abort_flag) return 1;
            if (spins > SPIN_LIMIT) { atomicExch(abort_flag, 1); return 1; }
        }
    }
}

// acc += sum_{k<K} x[k] * w[k], sequential, 4 at a time (K % 4 == 0, 16-byte aligned shared memory)
template <int UT>
__device__ __forceinline__ void chain(float (&acc)[UT], const float *w, const float *x, int xstride, int K) {
    const float4 *w4 = reinterpret_cast<const float4 *>(w);
#pragma unroll 4
    for (int k4 = 0; k4 < K / 4; ++k4) {
        const float4 wv = w4[k4];
#pragma unroll
        for (int j = 0; j < UT; ++j) {
            const float4 xv = reinterpret_cast<const float4 *>(x + j * xstride)[k4];
            acc[j] = __fmaf_rn(xv.x, wv.x, acc[j]);
            acc[j] = __fmaf_rn(xv.y, wv.y, acc[j]);
            acc[j] = __fmaf_rn(xv.z, wv.z, acc[j]);
            acc[j] = __fmaf_rn(xv.w, wv.w, acc[j]);
        }
    }
}

__device__ float pipe_sample_gaussian(const PipeParams &P, const float *out, float eps) {  // gaussian.py:39-52
    float mean = out[0];
    float ls = out[1] > P.log_scale_min_gauss ? out[1] : P.log_scale_min_gauss;
    float x = __fadd_rn(mean, __fmul_rn(wn_expf(ls), eps));
    x = x > -1.0f ? x : -1.0f;
    return x < 1.0f ? x : 1.0f;
}

__device__ float pipe_sample_mol(const PipeParams &P, const float *out, const float *u) {  // mixture.py:76-107
    const int K = P.O / 3;
    int best = 0;
    float bestv = 0.0f;
    for (int k = 0; k < K; ++k) {
        float v = __fsub_rn(out[k], wn_logf(-wn_logf(u[k])));
        if (k == 0 || v > bestv) { bestv = v; best = k; }
    }
    float mean = out[K + best];
    float ls = out[2 * K + best] > P.log_scale_min ? out[2 * K + best] : P.log_scale_min;
    float u2 = u[K];
    float lg = __fsub_rn(wn_logf(u2), wn_logf(__fsub_rn(1.0f, u2)));
    float x = __fadd_rn(mean, __fmul_rn(wn_expf(ls), lg));
    x = x > -1.0f ? x : -1.0f;
    return x < 1.0f ? x : 1.0f;
}

__device__ __forceinline__ int round4(int v) { return (v + 3) & ~3; }


// ---------------------------------------------------------------------------------------------------------------
// layer CTA: owns slice p of residual layer l  (ResidualConv1DGLU.incremental_step, modules.py:468-521)
// ---------------------------------------------------------------------------------------------------------------
template <int UT>
__device__ void layer_role(const PipeParams &P, const PipeCta &me, float *smem) {
    const int tid = threadIdx.x;
    const int R = P.R, H = P.H, NU = P.NU, NG = P.NG, T = P.T, l = me.l;
    const int hn = me.hn, rn = me.rn, sn = me.sn, hmax = P.hmax, rmax = P.rmax;
    const int d = P.dil[l], depth = P.depth[l];
    const int NUT = NU / UT;
    // ---- shared memory carve-up (weights first: identical to the packed blob) ------------------------------
    float *Wc = smem;                                   // [2*hmax][KCs]   dilated conv rows (tanh rows, then sigmoid rows)
    float *Wci = Wc + 2 * hmax * P.KCs;                 // [2*hmax][KIs]   cin 1x1 rows
    float *Wos = Wci + 2 * hmax * P.KIs;                // [rmax+smax][KHs] residual-out rows, then skip rows
    float *bconv = Wos + (rmax + P.smax) * P.KHs;       // [round4(2*hmax)]
    float *bcin = bconv + round4(2 * hmax);
    float *bos = bcin + round4(2 * hmax);               // [round4(rmax+smax)]
    float *xt = smem + P.layer_blob_floats;             // [NU][3R]  tap0 | tap1 | tap2 (= x_l(t))
    float *cond = xt + NU * 3 * R;                      // [NU][KI]
    float *pre = cond + NU * P.KI;                      // [NU][2*hmax]
    float *zf = pre + round4(NU * 2 * hmax);            // [NU][KH]
    {
        const float4 *src = reinterpret_cast<const float4 *>(P.blobs + me.blob);
        float4 *dst = reinterpret_cast<float4 *>(smem);
        for (int i = tid; i < P.layer_blob_floats / 4; i += PIPE_THREADS) dst[i] = __ldg(src + i);
        for (int i = tid; i < NU * (3 * R + P.KI + P.KH) + round4(NU * 2 * hmax); i += PIPE_THREADS) xt[i] = 0.0f;
    }
    __syncthreads();

    unsigned long long *X = P.xbuf + P.xoff[l];
    unsigned long long *Xn = (l + 1 < P.L) ? P.xbuf + P.xoff[l + 1] : nullptr;
    const int depth_n = (l + 1 < P.L) ? P.depth[l + 1] : 1;
    unsigned long long *Z = P.zbuf + (long long)l * NG * NU * H;
    unsigned long long *Sout = P.sbuf + (long long)l * NG * NU * P.S;
    const unsigned long long *Sin = (l > 0) ? P.sbuf + (long long)(l - 1) * NG * NU * P.S : nullptr;

    // my work items (at most one of each per thread, checked by the plan)
    const int n_pre = 2 * hn * NUT, n_os = (rn + sn) * NUT;
    const bool has_pre = tid < n_pre, has_os = tid < n_os;
    const int pj = has_pre ? tid % (2 * hn) : 0, put = has_pre ? tid / (2 * hn) : 0;
    const int prow = pj < hn ? pj : hmax + (pj - hn);
    const int oj = has_os ? tid % (rn + sn) : 0, out_ut = has_os ? tid / (rn + sn) : 0;
    const bool is_out = oj < rn;
    const int orow = is_out ? oj : rmax + (oj - rn);

    // ---- register prefetch of the older taps and the conditioning column of the NEXT work item -------------
    const int n_tap16 = NU * R;          // 16-byte chunks: 2 taps * NU * R/2
    const int n_c16 = NU * P.KI / 4;
    unsigned long long pfa[MAX_PF], pfb[MAX_PF];
    float4 pfc = make_float4(0.f, 0.f, 0.f, 0.f);
    int pf_t = 0;
    int bad = 0;
    auto prefetch = [&](int g, int t) {
        pf_t = t;
#pragma unroll
        for (int k = 0; k < MAX_PF; ++k) {
            const int q = tid + k * PIPE_THREADS;
            pfa[k] = 0ull; pfb[k] = 0ull;
            if (q < n_tap16) {
                const int half = n_tap16 / 2;
                const int tap = q / half, rem = q - tap * half;
                const int u = rem / (R / 2), r = (rem - u * (R / 2)) * 2;
                const int tt = t - (2 - tap) * d;
                if (tt >= 0) ldcg2(X + ((long long)(g * depth + tt % depth) * NU + u) * R + r, pfa[k], pfb[k]);
            }
        }
        pfc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tid < n_c16) {
            const int u = tid / (P.KI / 4), c4 = tid - u * (P.KI / 4);
            const int b = g * NU + u;
            if (b < P.B && c4 * 4 < P.cin)
                pfc = __ldg(reinterpret_cast<const float4 *>(P.c_up + ((long long)b * T + t) * P.cin) + c4);
        }
    };
    auto commit = [&]() {
#pragma unroll
        for (int k = 0; k < MAX_PF; ++k) {
            const int q = tid + k * PIPE_THREADS;
            if (q < n_tap16) {
                const int half = n_tap16 / 2;
                const int tap = q / half, rem = q - tap * half;
                const int u = rem / (R / 2), r = (rem - u * (R / 2)) * 2;
                const int tt = pf_t - (2 - tap) * d;
                // a tap read ahead of time must carry the tag of its own step; anything else is a protocol bug
                if (tt >= 0 && (ll_tag(pfa[k]) != (unsigned)tt + 1u || ll_tag(pfb[k]) != (unsigned)tt + 1u)) {
                    atomicExch(P.abort_flag, 2);
                    bad = 1;
                }
                *reinterpret_cast<float2 *>(xt + u * 3 * R + tap * R + r) = make_float2(ll_val(pfa[k]), ll_val(pfb[k]));
            }
        }
        if (tid < n_c16) reinterpret_cast<float4 *>(cond)[tid] = pfc;
    };

    prefetch(0, 0);
    const long long n_items = (long long)T * NG;
    int t = 0, g = 0;
    for (long long it = 0; it < n_items; ++it) {
        const unsigned tag = (unsigned)t + 1u;
        commit();
        if (__syncthreads_or(bad)) return;                                                           // (A)
        int g2 = g + 1, t2 = t;
        if (g2 == NG) { g2 = 0; t2 = t + 1; }
        // the next item's older taps were delivered in earlier items - except with a single group and d = 1, where
        // tap t+1-d is the x_l(t) this item is about to receive: then the read-ahead is issued after the poll below
        const bool early_pf = NG > 1;
        if (early_pf && t2 < T) prefetch(g2, t2);
        // ---- phase 0: everything that does not need x_l(t): taps t-2d, t-d and the cin 1x1 ----------------
        float acc[UT], cacc[UT];
#pragma unroll
        for (int j = 0; j < UT; ++j) acc[j] = cacc[j] = 0.0f;
        if (has_pre) {
            chain<UT>(acc, Wc + prow * P.KCs, xt + put * UT * 3 * R, 3 * R, 2 * R);
            chain<UT>(cacc, Wci + prow * P.KIs, cond + put * UT * P.KI, P.KI, P.KI);
        }
        // ---- wait for x_l(t) of this group (from layer l-1, or from the head for l = 0) ---------------------
        bad |= poll_block(X + (long long)(g * depth + t % depth) * NU * R, NU * R, tag, xt + 2 * R, R, 3 * R, P.abort_flag);
        if (__syncthreads_or(bad)) return;                                                           // (B)
        if (!early_pf && t2 < T) prefetch(g2, t2);
        // ---- phase 1: current tap, biases, conditioning add (modules.py:294-297, :497-501) -----------------
        if (has_pre) {
            chain<UT>(acc, Wc + prow * P.KCs + 2 * R, xt + put * UT * 3 * R + 2 * R, 3 * R, R);
#pragma unroll
            for (int j = 0; j < UT; ++j)
                pre[(put * UT + j) * 2 * hmax + prow] =
                    __fadd_rn(__fadd_rn(acc[j], bconv[prow]), __fadd_rn(cacc[j], bcin[prow]));
        }
        __syncthreads();                                                                             // (C)
        // ---- gate (modules.py:510) and publish my z channels to the P CTAs of this layer -------------------
        if (tid < hn * NU) {
            const int j = tid % hn, u = tid / hn;
            const float zv = __fmul_rn(wn_tanhf(pre[u * 2 * hmax + j]), wn_sigmoidf(pre[u * 2 * hmax + hmax + j]));
            st_ll(Z + ((long long)g * NU + u) * H + me.h0 + j, zv, tag);
        }
        bad |= poll_block(Z + (long long)g * NU * H, NU * H, tag, zf, H, P.KH, P.abort_flag);
        if (__syncthreads_or(bad)) return;                                                           // (D)
        // ---- phase 2: residual-out rows -> x_{l+1}(t); skip rows -> running skip sum (modules.py:512-520) --
        if (has_os) {
            float a2[UT];
#pragma unroll
            for (int j = 0; j < UT; ++j) a2[j] = 0.0f;
            chain<UT>(a2, Wos + orow * P.KHs, zf + out_ut * UT * P.KH, P.KH, P.KH);
#pragma unroll
            for (int j = 0; j < UT; ++j) {
                const int u = out_ut * UT + j;
                const float v = __fadd_rn(a2[j], bos[orow]);
                if (is_out) {
                    if (Xn) {
                        const int r = me.r0 + oj;
                        const float s2 = __fadd_rn(v, xt[u * 3 * R + 2 * R + r]);
                        const float xn = P.residual_legacy ? __fmul_rn(s2, WN_SQRT_HALF) : s2;
                        st_ll(Xn + ((long long)(g * depth_n + t % depth_n) * NU + u) * R + r, xn, tag);
                    }
                } else {
                    const int sc = me.s0 + (oj - rn);
                    float sk = v;                                          // wavenet.py:833-836
                    if (Sin) {
                        float prev;
                        bad |= poll_one(Sin + ((long long)g * NU + u) * P.S + sc, tag, prev, P.abort_flag);
                        const float s2 = __fadd_rn(prev, v);
                        sk = P.legacy ? __fmul_rn(s2, WN_SQRT_HALF) : s2;
                    }
                    st_ll(Sout + ((long long)g * NU + u) * P.S + sc, sk, tag);
                }
            }
        }
        if (++g == NG) { g = 0; ++t; }
    }
}


// ---------------------------------------------------------------------------------------------------------------
// head CTA: relu -> 1x1 -> relu -> 1x1 (wavenet.py:840-844), sampler (:850-859), feedback + first 1x1 (:826, :877)
// for the groups g = me.l, me.l + NH, ...
// ---------------------------------------------------------------------------------------------------------------
__device__ void head_role(const PipeParams &P, const PipeCta &me, float *smem) {
    const int tid = threadIdx.x;
    const int R = P.R, S = P.S, O = P.O, NU = P.NU, NG = P.NG, T = P.T;
    float *W1 = smem;                          // [S][KSs]
    float *W2 = W1 + S * P.KSs;                // [O][KSs]
    float *b1 = W2 + O * P.KSs;                // [round4(S)]
    float *b2 = b1 + round4(S);                // [round4(O)]
    float *fw = b2 + round4(O);                // [round4(R)]
    float *fb = fw + round4(R);                // [round4(R)]
    float *sk = smem + P.head_blob_floats;     // [NU][KS]
    float *h1 = sk + NU * P.KS;                // [NU][KS]
    float *outv = h1 + NU * P.KS;              // [NU][round4(O)]
    float *nin = outv + NU * round4(O);        // [NU]
    {
        const float4 *src = reinterpret_cast<const float4 *>(P.blobs + me.blob);
        float4 *dst = reinterpret_cast<float4 *>(smem);
        for (int i = tid; i < P.head_blob_floats / 4; i += PIPE_THREADS) dst[i] = __ldg(src + i);
        for (int i = tid; i < NU * (2 * P.KS + round4(O)) + NU; i += PIPE_THREADS) sk[i] = 0.0f;
    }
    __syncthreads();
    unsigned long long *X0 = P.xbuf + P.xoff[0];
    const int depth0 = P.depth[0];
    const unsigned long long *Sl = P.sbuf + (long long)(P.L - 1) * NG * NU * S;

    // initial input (wavenet.py:433-445): raw 0.0, mulaw mulaw(0) = 0.0 -> x_0(0) = first_conv(0)
    for (int g = me.l; g < NG; g += P.NH)
        for (int i = tid; i < R * NU; i += PIPE_THREADS) {
            const int r = i % R, u = i / R;
            const float x0 = __fadd_rn(__fmaf_rn(0.0f, fw[r], 0.0f), fb[r]);
            st_ll(X0 + ((long long)(g * depth0 + 0) * NU + u) * R + r, x0, 1u);
        }

    int bad = 0;
    for (int t = 0; t < T; ++t) {
        const unsigned tag = (unsigned)t + 1u;
        for (int g = me.l; g < NG; g += P.NH) {
            bad |= poll_block(Sl + (long long)g * NU * S, NU * S, tag, sk, S, P.KS, P.abort_flag);
            if (__syncthreads_or(bad)) return;
            for (int i = tid; i < S * NU; i += PIPE_THREADS) {       // relu -> final_convolution_1 -> relu
                const int n = i % S, u = i / S;
                const float4 *w4 = reinterpret_cast<const float4 *>(W1 + n * P.KSs);
                const float4 *x4 = reinterpret_cast<const float4 *>(sk + u * P.KS);
                float acc = 0.0f;
#pragma unroll 4
                for (int k4 = 0; k4 < P.KS / 4; ++k4) {
                    const float4 wv = w4[k4];
                    float4 xv = x4[k4];
                    xv.x = xv.x > 0.f ? xv.x : 0.f; xv.y = xv.y > 0.f ? xv.y : 0.f;
                    xv.z = xv.z > 0.f ? xv.z : 0.f; xv.w = xv.w > 0.f ? xv.w : 0.f;
                    acc = __fmaf_rn(xv.x, wv.x, acc);
                    acc = __fmaf_rn(xv.y, wv.y, acc);
                    acc = __fmaf_rn(xv.z, wv.z, acc);
                    acc = __fmaf_rn(xv.w, wv.w, acc);
                }
                const float v = __fadd_rn(acc, b1[n]);
                h1[u * P.KS + n] = v > 0.0f ? v : 0.0f;
            }
            __syncthreads();
            for (int i = tid; i < O * NU; i += PIPE_THREADS) {       // final_convolution_2
                const int n = i % O, u = i / O;
                float acc1[1] = {0.0f};
                chain<1>(acc1, W2 + n * P.KSs, h1 + u * P.KS, P.KS, P.KS);
                const float v = __fadd_rn(acc1[0], b2[n]);
                outv[u * round4(O) + n] = v;
                const int b = g * NU + u;
                if (P.raw && b < P.B) P.raw[((long long)b * T + t) * O + n] = v;   // wavenet.py:847
            }
            __syncthreads();
            if (tid < NU) {                                           // sampling + feedback
                const int u = tid, b = g * NU + u;
                float nxt = 0.0f;
                if (b < P.B) {
                    const float *nz = P.noise + ((long long)t * P.B + b) * P.NZ;
                    const float *o = outv + u * round4(O);
                    const float smp = (O == 2) ? pipe_sample_gaussian(P, o, nz[0]) : pipe_sample_mol(P, o, nz);
                    P.audio[(long long)b * T + t] = P.input_type == WN_INPUT_MULAW ? wn_inv_mulawf(smp) : smp;
                    nxt = P.teacher ? P.teacher[(long long)b * T + t] : smp;
                }
                nin[u] = nxt;
            }
            __syncthreads();
            if (t + 1 < T) {
                const int slot = (t + 1) % depth0;
                for (int i = tid; i < R * NU; i += PIPE_THREADS) {   // first_conv.incremental_step of step t+1
                    const int r = i % R, u = i / R;
                    const float x0 = __fadd_rn(__fmaf_rn(nin[u], fw[r], 0.0f), fb[r]);
                    st_ll(X0 + ((long long)(g * depth0 + slot) * NU + u) * R + r, x0, tag + 1u);
                }
            }
        }
    }
}

template <int UT>
__global__ void __launch_bounds__(PIPE_THREADS, 1) wn_loop_pipe_kernel(const PipeParams P) {
    extern __shared__ __align__(16) float pipe_smem[];
    const PipeCta me = P.ctas[blockIdx.x];
    if (me.role == 0) layer_role<UT>(P, me, pipe_smem);
    else head_role(P, me, pipe_smem);
}

}  // namespace

// ===============================================================================================================
// host side: plan (once per model), buffers and launch (per call)
// ===============================================================================================================
struct WnPipePlan {
    PipeParams prm;                 // model + plan part filled at plan time
    int max_heads = 16;
    float *d_blobs = nullptr;
    PipeCta *d_ctas = nullptr;
    unsigned long long *d_mail = nullptr;   // xbuf | zbuf | sbuf
    size_t mail_words = 0;
    int *d_abort = nullptr;
    int *h_abort = nullptr;          // pinned mirror, copied after every launch
    size_t smem_bytes = 0;
    int UT = 1;
};

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

static int pad_stride(int K) { return ((K / 4) % 2 == 1) ? K : K + 4; }   // (stride/4) odd: conflict-free LDS.128 across rows
static int r4(int v) { return (v + 3) & ~3; }

void wn_pipe_free(wn_handle *h) {
    WnPipePlan *pl = h->pipe;
    if (!pl) return;
    if (pl->d_blobs) cudaFree(pl->d_blobs);
    if (pl->d_ctas) cudaFree(pl->d_ctas);
    if (pl->d_mail) cudaFree(pl->d_mail);
    if (pl->d_abort) cudaFree(pl->d_abort);
    if (pl->h_abort) cudaFreeHost(pl->h_abort);
    delete pl;
    h->pipe = nullptr;
}

void wn_pipe_fill_info(wn_handle *h, wn_info *info) {
    if (!h->pipe) return;
    info->pipeline_ctas = h->pipe->prm.L * h->pipe->prm.P;
    info->pipeline_ctas_per_layer = h->pipe->prm.P;
    info->pipeline_group_size = h->pipe->prm.NU;
}

int wn_pipe_plan(wn_handle *h) {
    const wn_config &c = h->cfg;
    const WnModelDev &m = h->m;
    auto no = [&](const std::string &why) { h->err = why; return -1; };
    if (c.input_type == WN_INPUT_MULAW_QUANTIZE) return no("one-hot (mulaw-quantize) input runs on the generic engine");
    if (m.ksize != 3) return no("kernel_size != 3 runs on the generic engine");
    if (m.cin <= 0 || m.cin % 4) return no("needs local conditioning with cin_channels % 4 == 0");
    if (m.gin > 0) return no("global conditioning runs on the generic engine");
    if (m.R % 4 || m.S % 4 || m.H % 4) return no("channel counts must be multiples of 4");
    WnPipePlan *pl = new WnPipePlan();
    PipeParams &p = pl->prm;
    memset(&p, 0, sizeof(p));
    p.L = m.L; p.R = m.R; p.G = m.G; p.S = m.S; p.O = m.O; p.H = m.H; p.cin = m.cin; p.NZ = m.NZ;
    p.input_type = m.input_type; p.legacy = m.legacy; p.residual_legacy = m.residual_legacy;
    p.log_scale_min = m.log_scale_min; p.log_scale_min_gauss = m.log_scale_min_gauss;
    for (int l = 0; l < m.L; ++l) { p.dil[l] = m.dil[l]; p.depth[l] = 2 * m.dil[l] + 1; }
    p.P = env_int("WN_PIPE_P", 4);
    p.NU = env_int("WN_PIPE_NU", 4);
    pl->UT = env_int("WN_PIPE_UT", 1);
    auto bail = [&](const std::string &why) { delete pl; h->err = why; return -1; };
    if (p.P < 1 || p.P > m.H || p.P > m.R || p.P > m.S) return bail("bad WN_PIPE_P");
    if (pl->UT != 1 && pl->UT != 2) return bail("WN_PIPE_UT must be 1 or 2");
    if (p.NU < 1 || p.NU % pl->UT) return bail("WN_PIPE_NU must be a multiple of WN_PIPE_UT");
    auto lo = [&](int n, int q) { return (int)((long long)n * q / p.P); };
    p.hmax = p.rmax = p.smax = 0;
    for (int q = 0; q < p.P; ++q) {
        p.hmax = std::max(p.hmax, lo(m.H, q + 1) - lo(m.H, q));
        p.rmax = std::max(p.rmax, lo(m.R, q + 1) - lo(m.R, q));
        p.smax = std::max(p.smax, lo(m.S, q + 1) - lo(m.S, q));
    }
    const int NUT = p.NU / pl->UT;
    if (2 * p.hmax * NUT > PIPE_THREADS || (p.rmax + p.smax) * NUT > PIPE_THREADS || p.hmax * p.NU > PIPE_THREADS)
        return bail("slice x group does not fit 256 threads (raise WN_PIPE_P or lower WN_PIPE_NU)");
    if (p.NU * m.R > MAX_PF * PIPE_THREADS) return bail("tap prefetch does not fit (NU * R > 1024)");
    p.KC = 3 * m.R; p.KCs = pad_stride(p.KC);
    p.KI = r4(m.cin); p.KIs = pad_stride(p.KI);
    p.KH = m.H; p.KHs = pad_stride(p.KH);
    p.KS = m.S; p.KSs = pad_stride(p.KS);
    if (p.NU * p.KI / 4 > PIPE_THREADS) return bail("conditioning prefetch does not fit");
    p.layer_blob_floats = 2 * p.hmax * p.KCs + 2 * p.hmax * p.KIs + (p.rmax + p.smax) * p.KHs + 2 * r4(2 * p.hmax) +
                          r4(p.rmax + p.smax);
    p.head_blob_floats = m.S * p.KSs + m.O * p.KSs + r4(m.S) + r4(m.O) + 2 * r4(m.R);
    const size_t layer_smem = sizeof(float) * ((size_t)p.layer_blob_floats + (size_t)p.NU * (3 * m.R + p.KI + p.KH) + r4(p.NU * 2 * p.hmax));
    const size_t head_smem = sizeof(float) * ((size_t)p.head_blob_floats + (size_t)p.NU * (2 * p.KS + r4(m.O)) + p.NU);
    pl->smem_bytes = std::max(layer_smem, head_smem);
    int max_optin = 0;
    cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device);
    if (pl->smem_bytes > (size_t)max_optin)
        return bail("fp32 weight slice (" + std::to_string(pl->smem_bytes) + " B) exceeds shared memory per CTA (" +
                    std::to_string(max_optin) + " B)");
    if (m.L * p.P + 1 > h->sm_count) return bail("more layer CTAs than SMs");
    pl->max_heads = std::min(16, h->sm_count - m.L * p.P);

    // ---- pack weights: one blob per layer CTA in its shared-memory layout, one head blob ---------------------
    auto Wt = [&](const std::string &n) -> const float * {
        auto it = h->host_w.find(n);
        return it == h->host_w.end() ? nullptr : it->second.data.data();
    };
    std::vector<float> blobs((size_t)m.L * p.P * p.layer_blob_floats + p.head_blob_floats, 0.0f);
    std::vector<PipeCta> ctas;
    for (int l = 0; l < m.L; ++l) {
        std::string pre = "ResidualConv1DGLU_" + std::to_string(l) + "/residual_block_";
        const float *cw = Wt(pre + "causal_conv/kernel"), *cb = Wt(pre + "causal_conv/bias");
        const float *iw = Wt(pre + "cin_conv/kernel"), *ib = Wt(pre + "cin_conv/bias");
        const float *ow = Wt(pre + "out_conv/kernel"), *ob = Wt(pre + "out_conv/bias");
        const float *sw = Wt(pre + "skip_conv/kernel"), *sb = Wt(pre + "skip_conv/bias");
        for (int q = 0; q < p.P; ++q) {
            PipeCta ct;
            ct.role = 0; ct.l = l; ct.p = q;
            ct.h0 = lo(m.H, q); ct.hn = lo(m.H, q + 1) - ct.h0;
            ct.r0 = lo(m.R, q); ct.rn = lo(m.R, q + 1) - ct.r0;
            ct.s0 = lo(m.S, q); ct.sn = lo(m.S, q + 1) - ct.s0;
            ct.blob = (long long)(l * p.P + q) * p.layer_blob_floats;
            float *b = blobs.data() + ct.blob;
            float *Wc = b, *Wci = Wc + 2 * p.hmax * p.KCs, *Wos = Wci + 2 * p.hmax * p.KIs;
            float *bconv = Wos + (p.rmax + p.smax) * p.KHs, *bcin = bconv + r4(2 * p.hmax), *bos = bcin + r4(2 * p.hmax);
            for (int half = 0; half < 2; ++half)
                for (int j = 0; j < ct.hn; ++j) {
                    const int row = half * p.hmax + j, col = half * m.H + ct.h0 + j;   // modules.py:494 split order
                    for (int k = 0; k < 3 * m.R; ++k) Wc[row * p.KCs + k] = cw[(size_t)k * m.G + col];
                    for (int k = 0; k < m.cin; ++k) Wci[row * p.KIs + k] = iw[(size_t)k * m.G + col];
                    bconv[row] = cb ? cb[col] : 0.0f;
                    bcin[row] = ib ? ib[col] : 0.0f;
                }
            for (int j = 0; j < ct.rn; ++j) {
                for (int k = 0; k < m.H; ++k) Wos[j * p.KHs + k] = ow[(size_t)k * m.R + ct.r0 + j];
                bos[j] = ob ? ob[ct.r0 + j] : 0.0f;
            }
            for (int j = 0; j < ct.sn; ++j) {
                for (int k = 0; k < m.H; ++k) Wos[(p.rmax + j) * p.KHs + k] = sw[(size_t)k * m.S + ct.s0 + j];
                bos[p.rmax + j] = sb ? sb[ct.s0 + j] : 0.0f;
            }
            ctas.push_back(ct);
        }
    }
    {
        const long long hb = (long long)m.L * p.P * p.layer_blob_floats;
        float *b = blobs.data() + hb;
        float *W1 = b, *W2 = W1 + m.S * p.KSs, *b1 = W2 + m.O * p.KSs, *b2 = b1 + r4(m.S), *fw = b2 + r4(m.O), *fb = fw + r4(m.R);
        const float *w1 = Wt("final_convolution_1/kernel"), *bb1 = Wt("final_convolution_1/bias");
        const float *w2 = Wt("final_convolution_2/kernel"), *bb2 = Wt("final_convolution_2/bias");
        const float *f = Wt("input_convolution/kernel"), *fbb = Wt("input_convolution/bias");
        for (int n = 0; n < m.S; ++n) {
            for (int k = 0; k < m.S; ++k) W1[n * p.KSs + k] = w1[(size_t)k * m.S + n];
            b1[n] = bb1 ? bb1[n] : 0.0f;
        }
        for (int n = 0; n < m.O; ++n) {
            for (int k = 0; k < m.S; ++k) W2[n * p.KSs + k] = w2[(size_t)k * m.O + n];
            b2[n] = bb2 ? bb2[n] : 0.0f;
        }
        for (int r = 0; r < m.R; ++r) { fw[r] = f[r]; fb[r] = fbb ? fbb[r] : 0.0f; }
        for (int j = 0; j < pl->max_heads; ++j) {
            PipeCta ct;
            memset(&ct, 0, sizeof(ct));
            ct.role = 1; ct.l = j; ct.blob = hb;
            ctas.push_back(ct);
        }
    }
    h->pipe = pl;
    if (cudaMalloc((void **)&pl->d_blobs, blobs.size() * sizeof(float)) != cudaSuccess ||
        cudaMemcpy(pl->d_blobs, blobs.data(), blobs.size() * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMalloc((void **)&pl->d_ctas, ctas.size() * sizeof(PipeCta)) != cudaSuccess ||
        cudaMemcpy(pl->d_ctas, ctas.data(), ctas.size() * sizeof(PipeCta), cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMalloc((void **)&pl->d_abort, sizeof(int)) != cudaSuccess ||
        cudaMallocHost((void **)&pl->h_abort, sizeof(int)) != cudaSuccess) {
        wn_pipe_free(h);
        h->err = std::string("pipeline engine: ") + cudaGetErrorString(cudaGetLastError());
        return -1;
    }
    *pl->h_abort = 0;
    p.blobs = pl->d_blobs;
    p.ctas = pl->d_ctas;
    p.abort_flag = pl->d_abort;
    auto kern = pl->UT == 2 ? wn_loop_pipe_kernel<2> : wn_loop_pipe_kernel<1>;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl->smem_bytes) != cudaSuccess) {
        wn_pipe_free(h);
        h->err = std::string("pipeline engine: ") + cudaGetErrorString(cudaGetLastError());
        return -1;
    }
    return 0;
}

int wn_pipe_check_abort(wn_handle *h) {   // call after the stream has been synchronised
    if (!h->pipe || !h->pipe->h_abort) return WN_OK;
    const int code = *h->pipe->h_abort;
    if (code == 0) return WN_OK;
    h->err = code == 2 ? "pipeline engine: stale dilation tap (protocol error)"
                       : "pipeline engine: watchdog - a CTA waited too long for a peer";
    return WN_ERR_WATCHDOG;
}

int wn_launch_loop_pipe(wn_handle *h, const float *c_up, const float *noise, const float *teacher, int B, int T,
                        float *audio, float *raw, cudaStream_t st) {
    WnPipePlan *pl = h->pipe;
    if (!pl) { h->err = "pipeline engine not planned"; return WN_ERR_STATE; }
    PipeParams p = pl->prm;
    p.NG = (B + p.NU - 1) / p.NU;
    p.NH = std::min(p.NG, pl->max_heads);
    p.B = B; p.T = T;
    p.c_up = c_up; p.noise = noise; p.teacher = teacher; p.audio = audio; p.raw = raw;
    // mailboxes: X (per layer, ring depth 2d+1) | Z | S, all 8-byte {value, tag} words, zeroed (tag 0 = empty)
    size_t words = 0;
    for (int l = 0; l < p.L; ++l) { p.xoff[l] = (long long)words; words += (size_t)p.NG * p.depth[l] * p.NU * p.R; }
    const size_t zoff = words; words += (size_t)p.L * p.NG * p.NU * p.H;
    const size_t soff = words; words += (size_t)p.L * p.NG * p.NU * p.S;
    if (pl->mail_words < words) {
        if (pl->d_mail) cudaFree(pl->d_mail);
        pl->d_mail = nullptr; pl->mail_words = 0;
        cudaError_t e = cudaMalloc((void **)&pl->d_mail, words * 8);
        if (e != cudaSuccess) { h->err = std::string("pipeline mailboxes: ") + cudaGetErrorString(e); return WN_ERR_NOMEM; }
        pl->mail_words = words;
    }
    p.xbuf = pl->d_mail; p.zbuf = pl->d_mail + zoff; p.sbuf = pl->d_mail + soff;
    WN_CUDA_TRY(h, cudaMemsetAsync(pl->d_mail, 0, words * 8, st));
    WN_CUDA_TRY(h, cudaMemsetAsync(pl->d_abort, 0, sizeof(int), st));
    const int grid = p.L * p.P + p.NH;
    void *args[] = {&p};
    const void *kern = pl->UT == 2 ? (const void *)wn_loop_pipe_kernel<2> : (const void *)wn_loop_pipe_kernel<1>;
    WN_CUDA_TRY(h, cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(PIPE_THREADS), args, pl->smem_bytes, st));
    h->launches++;
    WN_CUDA_TRY(h, cudaMemcpyAsync(pl->h_abort, pl->d_abort, sizeof(int), cudaMemcpyDeviceToHost, st));
    return WN_OK;
}

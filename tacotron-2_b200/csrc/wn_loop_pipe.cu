// wn_loop_pipe.cu - the pipelined persistent engine of the autoregressive loop (wavenet.py:805-911) for sm_100a.
//
// One cooperative launch runs the whole utterance batch for all T steps.  The network is laid out ACROSS the SMs:
//   * every residual layer l is owned by P CTAs ("layer CTAs", one per SM).  CTA (l,p) keeps ITS slice of the layer's
//     weights resident in shared memory for the whole launch: the rows of the gate pre-activation matrix (dilated conv +
//     cin 1x1 over the concatenated input [x(t-2d) | x(t-d) | c(t) | x(t)]) that produce gate channels [h0,h1) - the tanh
//     and the sigmoid half, so gating is local - and the COLUMNS [h0,h1) of the residual-out / skip 1x1s;
//   * NH "head CTAs" own the two head 1x1s and the sampler.
// Utterances travel through that pipeline in groups of NU.  All stages walk the same schedule (t major, group minor),
// so while layer 7 works on group 3 layer 6 works on group 4: the serial chain of one group costs ~(L+1) stage
// latencies per sample, but all SMs stay busy on different groups.  The only feedback edge is sample(t) -> input(t+1).
//
// ONE global hand-off per layer.  The P CTAs of a layer form a thread-block cluster.  Each one gates its own channels
// and multiplies them with its column block of the out / skip 1x1s, which yields a PARTIAL sum for every output channel
// (split-K).  The partial sums are reduce-scattered inside the cluster over distributed shared memory: every row goes,
// with one 16-byte st.async, into the inbox of the CTA that owns the row and is counted on that CTA's mbarrier
// (complete_tx); the owner adds the P partials in a fixed order (p = 0, 1, ...), the bias and the residual input, and
// publishes its finished slice of x_{l+1}(t) to all P CTAs of layer l+1.  The running skip sum is carried along the same
// edges (slice p of it by the CTAs (.,p)) and reaches the head finished.
//
// The global hand-off never uses a barrier or a separate flag: every float travels as an 8-byte {value, tag} word
// (tag = t+1) written with 128-bit stores and polled with 128-bit loads by the consumer ("LL" protocol), so a value is
// usable the moment it lands in L2.  Mailboxes are double buffered on t&1 (no CTA can be more than one step ahead of
// another: step t+1 cannot start before the head sampled step t, which needs every CTA's step-t output).  The exchange
// inboxes are double buffered on the item parity: a peer can only be two items ahead of a CTA after that CTA has sent its
// partials for the item in between, i.e. after it finished reading the inbox the peer is about to overwrite.
// The dilation history x_l(t-d), x_l(t-2d) lives in a per-CTA ring in global memory (plain floats, written by the CTA
// itself right after it received x_l(t)); it is read one work item ahead into registers, together with the
// conditioning column c(t), and the part of the gate pre-activation that only needs those is computed BEFORE x_l(t)
// arrives, so the per-layer critical path is: poll -> K=R slice of the dot -> gate -> K=H/P chain -> exchange -> store.
//
// Arithmetic is the fixed order of DESIGN.md "Arithmetic" (dot16 / blocked chains with ksplit = P), identical in
// wn_loop_generic.cu and oracle/c/wn_oracle.c, so results are bit-identical to both.
//
// Scope: scalar input (raw / mulaw), Gaussian or mixture-of-logistics head, kernel_size 3, local conditioning on,
// global conditioning off, fp32 weights that fit shared memory (default config: 125 KB per layer CTA at P = 6).
// Anything else runs on the generic engine.
#include <cooperative_groups.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "wn_common.cuh"

#ifndef WN_CHAIN_ROWS
#define WN_CHAIN_ROWS 1
#endif
#ifndef WN_EARLY_RS
#define WN_EARLY_RS 1
#endif
#ifndef WN_EARLY_X
#define WN_EARLY_X 24          // with at least this many utterance groups (throughput-bound regime: the producer is ahead) the
#endif                         // next item's x words are requested before its x-independent work, which hides one L2 round
                               // trip per item (+7 % at 128 utterances); with fewer groups the extra load only costs (-2 %)

namespace {

constexpr int PIPE_THREADS = 256;     // threads of one TEAM; a CTA runs TEAMS teams that share the resident weights
constexpr int TEAMS = 1;              // and work on different utterance groups (g % TEAMS), out of phase with each other
static_assert(TEAMS == 1, "the layer role runs one team per CTA");
constexpr int LANES = WN_DOT_LANES;            // partial sums per dot16
constexpr int MAX_P = 8;             // CTAs per layer
constexpr int CR = WN_CHAIN_ROWS;    // rows of the out / skip 1x1s per thread of the blocked chains
constexpr int MAX_PF = 2;            // 16-byte prefetch loads per thread for the two older taps
constexpr unsigned long long WATCHDOG_NS = 20ull * 1000 * 1000 * 1000;   // a CTA that waits this long for a peer gives up
// Development aid (tools/ablate.sh): -DWN_ABLATE=<bits> removes one phase of the layer role so that its share of the
// step time can be read off the production kernel (results are then wrong, the data flow is intact).  0 in every build
// that ships.  1 gate -> identity, 2 no current-tap mac, 4 no chains, 8 no phase0, 16 no ring store / prefetch,
// 32 no reduce16, 64 no intra-layer exchange
#ifndef WN_ABLATE
#define WN_ABLATE 0
#endif

struct PipeCta {
    int role;                        // 0 layer, 1 head
    int l, p;                        // layer / slice (head: index of the head CTA in l)
    int h0, hn, r0, rn, s0, sn;      // gate-channel / residual / skip slices owned
    long long blob;                  // float offset of this CTA's packed weights
    long long ring;                  // float offset of this CTA's private dilation ring
};

struct PipeParams {
    // model
    int L, R, H, S, O, cin, NZ, input_type, legacy, residual_legacy;
    float log_scale_min, log_scale_min_gauss;
    int dil[WN_MAX_LAYERS], depth[WN_MAX_LAYERS];
    // plan
    int P, NH, NU, NG;
    int hq, hk;                      // gate channels per CTA padded: hq rows per half (even), hk = K of the partial chains (mult. of 4)
    int NR, NQ;                      // gate rows per CTA (2*hq, multiple of 4) and row quads
    int KC, KC4, C0;                 // concatenated K (3R + cin), in chunks, and the chunks that precede x(t)
    int smax;                        // largest skip slice
    int layer_blob_floats, head_blob_floats;
    const float *blobs;
    const PipeCta *ctas;
    float *rings;
    // mailboxes: one block per consumer stage (layers 0..L-1, head = L): x | rs (see the host side); `in` carries the
    // sample fed back from the head to layer 0
    unsigned long long *mail, *in;
    long long stage_sz, rs_off;      // words
    int rmax, RO;                    // largest residual slice; rows of one source in the exchange inbox (rmax + smax)
    int *abort_flag;
    long long *stats;                // optional [CTA][8] clock64 counters (WN_PIPE_STATS), else nullptr
    // call
    const float *c_up, *noise, *teacher;
    int inputs_mode;                 // WaveNet.step semantics: teacher[t] IS the input of step t, nothing is sampled
    float *audio, *raw;
    int B, T;
};

__device__ __forceinline__ unsigned long long pack_ll(float v, unsigned tag) {
    return ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v);
}
__device__ __forceinline__ void st_ll2(unsigned long long *p, float v0, float v1, unsigned tag) {   // 16-byte aligned
    const unsigned long long a = pack_ll(v0, tag), b = pack_ll(v1, tag);
    asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ void ld_ll2(const unsigned long long *p, unsigned long long &a, unsigned long long &b) {
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ float ll_val(unsigned long long w) { return __uint_as_float((unsigned)(w & 0xffffffffu)); }
__device__ __forceinline__ unsigned ll_tag(unsigned long long w) { return (unsigned)(w >> 32); }

__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Poll one pair of {value,tag} words until both carry `tag`.  Returns nonzero if the watchdog fired.
__device__ __forceinline__ int poll_pair(const unsigned long long *src, unsigned tag, float &v0, float &v1,
                                         int *abort_flag) {
    unsigned spins = 0;
    unsigned long long t_start = 0;
    // One load in flight.  Measured alternatives, both slower: a __nanosleep back-off between polls, and two loads in
    // flight staggered by half a round trip (finer sampling of L2, but twice the poll traffic).
    for (;;) {
        unsigned long long a, b;
        ld_ll2(src, a, b);
        if (ll_tag(a) == tag && ll_tag(b) == tag) { v0 = ll_val(a); v1 = ll_val(b); return 0; }
        if ((++spins & 0x3ffu) == 0) {   // watchdog: wall-clock based, so that a slow run (profiler replay) is not mistaken for a hang
            if (*(volatile int *)abort_flag) return 1;
            const unsigned long long now = global_timer_ns();
            if (t_start == 0) t_start = now;
            else if (now - t_start > WATCHDOG_NS) { atomicExch(abort_flag, 1); return 1; }
        }
    }
}

// ---- thread-block cluster / distributed shared memory (the P CTAs of a layer form one cluster) ------------------
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned mapa_u32(unsigned addr, unsigned rank) {   // my shared address -> the same variable of CTA `rank`
    unsigned r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
// 16-byte store into a peer's shared memory that also counts 16 bytes on the peer's mbarrier (release at cluster scope)
__device__ __forceinline__ void st_async4(unsigned raddr, float4 v, unsigned rmbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.f32 [%0], {%1, %2, %3, %4}, [%5];"
                 ::"r"(raddr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "r"(rmbar) : "memory");
}
__device__ __forceinline__ void mbar_init(unsigned a, unsigned n) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(n) : "memory");
}
__device__ __forceinline__ void mbar_expect(unsigned a, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(a), "r"(bytes) : "memory");
}
__device__ __forceinline__ int mbar_try(unsigned a, unsigned parity) {
    unsigned ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    return (int)ok;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__device__ __forceinline__ int round4(int v) { return (v + 3) & ~3; }

// Barriers of one team (named barrier 1 + team, PIPE_THREADS participants): the teams never wait for each other.
__device__ __forceinline__ void team_sync(int team) {
    asm volatile("bar.sync %0, %1;" ::"r"(1 + team), "r"(PIPE_THREADS) : "memory");
}
__device__ __forceinline__ int team_sync_or(int team, int v) {
    int out;
    asm volatile(
        "{\n\t.reg .pred p, q;\n\tsetp.ne.s32 q, %1, 0;\n\tbar.red.or.pred p, %2, %3, q;\n\tselp.s32 %0, 1, 0, p;\n\t}"
        : "=r"(out)
        : "r"(v), "r"(1 + team), "r"(PIPE_THREADS)
        : "memory");
    return out;
}

// ---- the fp32 inner products ---------------------------------------------------------------------------------
// dot16 operands in shared memory: element k = 4c+e of utterance u sits at xs_pos(k, u) = ((e*NU/4 + u/4)*K4 + c)*4 + u%4,
// i.e. one 16-byte word holds the same k for 4 utterances and, for fixed e, the words of consecutive chunks c are
// contiguous - the 16 lanes of a dot16 (lane = c % 16) then read 256 contiguous bytes, free of bank conflicts.
// Weights are row-major [row][K].
//
// All multiply-adds are issued as FFMA2 (fma.rn.f32x2: two independent IEEE round-to-nearest fp32 fmas per
// instruction, the form in which sm_100 reaches its fp32 peak): the pair is two UTTERANCES of the same output, the
// weight is the scalar broadcast operand.  Each lane of each fma is exactly fmaf(), so the bits equal the oracle's.
template <int NU>
__device__ __forceinline__ int xs_pos(int k, int u, int K4) {
    return (((k & 3) * (NU / 4) + (u >> 2)) * K4 + (k >> 2)) * 4 + (u & 3);
}

__device__ __forceinline__ float2 ffma2(float2 a, float w, float2 c) {
    float2 d;
    const float2 ww = make_float2(w, w);   // ptxas folds the duplicate into FFMA2's scalar-broadcast operand
    asm("fma.rn.f32x2 %0, %1, %2, %3;"
        : "=l"(*reinterpret_cast<unsigned long long *>(&d))
        : "l"(*reinterpret_cast<const unsigned long long *>(&a)),
          "l"(*reinterpret_cast<const unsigned long long *>(&ww)),
          "l"(*reinterpret_cast<const unsigned long long *>(&c)));
    return d;
}

template <int NU, int ROWS>
struct MacOps {
    float4 w[ROWS];
    float4 x[NU];   // [e][u4]
};

template <int NU, int ROWS>
__device__ __forceinline__ void mac_fma(float2 (&acc)[ROWS * NU / 2], const MacOps<NU, ROWS> &o) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
#pragma unroll
        for (int i = 0; i < ROWS; ++i) {
            const float wv = e == 0 ? o.w[i].x : (e == 1 ? o.w[i].y : (e == 2 ? o.w[i].z : o.w[i].w));
#pragma unroll
            for (int u4 = 0; u4 < NU / 4; ++u4) {
                const float4 t = o.x[e * (NU / 4) + u4];
                acc[i * (NU / 2) + 2 * u4] = ffma2(make_float2(t.x, t.y), wv, acc[i * (NU / 2) + 2 * u4]);
                acc[i * (NU / 2) + 2 * u4 + 1] = ffma2(make_float2(t.z, t.w), wv, acc[i * (NU / 2) + 2 * u4 + 1]);
            }
        }
    }
}

// mac_lane: the thread owns a tile of ROWS rows x NU utterances and the float4 chunks c = c_first, c_first+16, ... < c_end
// of K (lane = c % 16): acc[row i][utt u] += sum_e w[i][4c+e] * xs[4c+e][u], e ascending, c ascending - one partial
// of dot16.  Loads of the next chunk are issued before the fmas of the current one (register double buffer).
template <int NU, int ROWS>
__device__ __forceinline__ void mac_lane(float2 (&acc)[ROWS * NU / 2], const float4 *__restrict__ w4 /* first row, chunk 0 */,
                                         int KC4, const float *__restrict__ xs, int c_first, int c_end) {
    const float4 *x4 = reinterpret_cast<const float4 *>(xs);
    auto load = [&](MacOps<NU, ROWS> &o, int c) {
#pragma unroll
        for (int i = 0; i < ROWS; ++i) o.w[i] = w4[i * KC4 + c];
#pragma unroll
        for (int j = 0; j < NU; ++j) o.x[j] = x4[j * KC4 + c];
    };
    int c = c_first;
    if (c >= c_end) return;
    if (ROWS * NU > 16 || TEAMS > 1) {   // large tiles or a 128-register budget: one operand buffer
        for (; c < c_end; c += LANES) {
            MacOps<NU, ROWS> a;
            load(a, c);
            mac_fma<NU, ROWS>(acc, a);
        }
        return;
    }
    MacOps<NU, ROWS> a, b;
    load(a, c);
    for (;;) {
        const int c1 = c + LANES;
        if (c1 < c_end) load(b, c1);
        mac_fma<NU, ROWS>(acc, a);
        if (c1 >= c_end) break;
        c = c1 + LANES;
        if (c < c_end) load(a, c);
        mac_fma<NU, ROWS>(acc, b);
        if (c >= c_end) break;
    }
}

// Same partial as mac_lane over NC chunks c_first, c_first+16, ... with the weights already in registers (the current-tap
// columns of the gate matrix never change: a layer CTA keeps its 4 x NC float4 per thread for the whole launch, which
// takes the weight loads off the critical path between "x_l(t) arrived" and "gate").
template <int NU, int ROWS, int NC>
__device__ __forceinline__ void mac_lane_wreg(float2 (&acc)[ROWS * NU / 2], const float4 (&w)[ROWS][NC],
                                              int KC4, const float *__restrict__ xs, int c_first) {
    const float4 *x4 = reinterpret_cast<const float4 *>(xs);
    MacOps<NU, ROWS> o[NC];
#pragma unroll
    for (int j = 0; j < NC; ++j) {
#pragma unroll
        for (int i = 0; i < ROWS; ++i) o[j].w[i] = w[i][j];
#pragma unroll
        for (int u = 0; u < NU; ++u) o[j].x[u] = x4[u * KC4 + c_first + j * LANES];
    }
#pragma unroll
    for (int j = 0; j < NC; ++j) mac_fma<NU, ROWS>(acc, o[j]);
}

// Butterfly of the 16 lanes' partials: neighbours first (the tree of wn_oracle.c:dot16).  While a lane holds more
// than one value each step also halves what it keeps (the lane with the mask bit set keeps the upper half), so
// the whole reduction of NV values costs NV-1 shuffles instead of 4*NV.  On return the lane holds NV/16 (at least
// 1) finished sums in a[0..]; the original index of a[i] is i + reduce16_base<NV>(lane).
template <int NV>
__device__ __forceinline__ void reduce16(float (&a)[NV], int lane) {
    int cnt = NV;
#pragma unroll
    for (int m = 1; m < LANES; m <<= 1) {
        if (cnt >= 2) {
            const int half = cnt / 2;
            const bool up = (lane & m) != 0;
#pragma unroll
            for (int i = 0; i < NV / 2; ++i) {
                if (i < half) {
                    const float send = up ? a[i] : a[i + half];
                    const float keep = up ? a[i + half] : a[i];
                    a[i] = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, m));
                }
            }
            cnt = half;
        } else {
            a[0] = __fadd_rn(a[0], __shfl_xor_sync(0xffffffffu, a[0], m));
        }
    }
}
template <int NV>
__device__ __forceinline__ int reduce16_base(int lane) {
    int cnt = NV, base = 0;
#pragma unroll
    for (int m = 1; m < LANES; m <<= 1) {
        if (cnt >= 2) { cnt /= 2; if (lane & m) base += cnt; }
    }
    return base;
}
// float2 accumulators (row i, utterance pair) -> flat values v = i*NU + u for reduce16
template <int NV>
__device__ __forceinline__ void unpack2(const float2 (&acc)[NV / 2], float (&a)[NV]) {
#pragma unroll
    for (int i = 0; i < NV / 2; ++i) { a[2 * i] = acc[i].x; a[2 * i + 1] = acc[i].y; }
}

// chain_tile: ROWS rows x NU utterances, K sequential in-thread (the blocked chains of the out / skip 1x1s).
// w4t is chunk-major: w4t[c * nrows + row] holds W[row][4c..4c+3]; the thread's rows are row0 + i*rstride.
// zs is k-major [k][NU]; every lane reads the same words (broadcast).
template <int NU, int ROWS>
__device__ __forceinline__ void chain_tile(float2 (&acc)[ROWS * NU / 2], const float4 *__restrict__ w4t, int nrows, int row0,
                                           int rstride, const float *__restrict__ zs, int K4) {
    const float4 *z4 = reinterpret_cast<const float4 *>(zs);
    auto load = [&](MacOps<NU, ROWS> &o, int c) {
#pragma unroll
        for (int i = 0; i < ROWS; ++i) o.w[i] = w4t[c * nrows + row0 + i * rstride];
#pragma unroll
        for (int j = 0; j < NU; ++j) o.x[j] = z4[c * NU + j];
    };
    if (K4 <= 0) return;
    MacOps<NU, ROWS> a, b;
    load(a, 0);
    int c = 0;
    for (;;) {
        if (c + 1 < K4) load(b, c + 1);
        mac_fma<NU, ROWS>(acc, a);
        if (c + 1 >= K4) break;
        c += 2;
        if (c < K4) load(a, c);
        mac_fma<NU, ROWS>(acc, b);
        if (c >= K4) break;
    }
}

// One instruction stream for both gate activations (modules.py:510): the lanes of a warp hold tanh and sigmoid
// pre-activations side by side, and calling wn_tanhf / wn_sigmoidf under a branch makes the warp run the tanh polynomial,
// the tanh exponential path and the sigmoid one after the other.  Operation for operation this is wn_tanhf(x) for
// sig = false and wn_sigmoidf(x) for sig = true (include/wn_math.h), with selects instead of branches: same bits.
__device__ __forceinline__ float gate_act(float x, bool sig) {
    const float ax = fabsf(x);
    const float e = wn_expf(sig ? -x : __fadd_rn(ax, ax));
    const float qd = __fdiv_rn(sig ? 1.0f : 2.0f, sig ? __fadd_rn(1.0f, e) : __fadd_rn(e, 1.0f));
    float r = sig ? qd : __fsub_rn(1.0f, qd);
    const float z = __fmul_rn(x, x);
    float p = -5.70498872745e-3f;
    p = __fmaf_rn(p, z, 2.06390887954e-2f);
    p = __fmaf_rn(p, z, -5.37397155531e-2f);
    p = __fmaf_rn(p, z, 1.33314422036e-1f);
    p = __fmaf_rn(p, z, -3.33332819422e-1f);
    const float small = __fmaf_rn(__fmul_rn(p, z), x, x);
    if (!sig) {
        if (ax > 9.02f) r = 1.0f;
        if (x < 0.0f) r = -r;
        if (ax < 0.625f) r = small;
        if (!(x == x)) r = x;
    }
    return r;
}

__device__ float pipe_sample_gaussian(const PipeParams &P, const float *out, int ostride, float eps) {  // gaussian.py:39-52
    float mean = out[0];
    float ls = out[ostride] > P.log_scale_min_gauss ? out[ostride] : P.log_scale_min_gauss;
    float x = __fadd_rn(mean, __fmul_rn(wn_expf(ls), eps));
    x = x > -1.0f ? x : -1.0f;
    return x < 1.0f ? x : 1.0f;
}

__device__ float pipe_sample_mol(const PipeParams &P, const float *out, int ostride, const float *u) {  // mixture.py:76-107
    const int K = P.O / 3;
    int best = 0;
    float bestv = 0.0f;
    for (int k = 0; k < K; ++k) {
        float v = __fsub_rn(out[k * ostride], wn_logf(-wn_logf(u[k])));
        if (k == 0 || v > bestv) { bestv = v; best = k; }
    }
    float mean = out[(K + best) * ostride];
    float lsr = out[(2 * K + best) * ostride];
    float ls = lsr > P.log_scale_min ? lsr : P.log_scale_min;
    float u2 = u[K];
    float lg = __fsub_rn(wn_logf(u2), wn_logf(__fsub_rn(1.0f, u2)));
    float x = __fadd_rn(mean, __fmul_rn(wn_expf(ls), lg));
    x = x > -1.0f ? x : -1.0f;
    return x < 1.0f ? x : 1.0f;
}

// ---------------------------------------------------------------------------------------------------------------
// layer CTA: owns gate channels [h0,h1) of residual layer l  (ResidualConv1DGLU.incremental_step, modules.py:468-521)
// ---------------------------------------------------------------------------------------------------------------
// RT / CT: residual (= skip) channels and conditioning channels as compile-time constants (0 = read them from P): with
// constants every index computation below folds into shifts and immediates, which the compiler rematerialises instead
// of keeping ~40 loop-invariant registers alive.
template <int NU, int RT, int CT, bool STATS>
__device__ void layer_role(const PipeParams &P, const PipeCta &me, float *smem) {
    constexpr int NV = 4 * NU, NP = NU / 2;          // accumulators per thread; utterance pairs per group
    constexpr int XT = 4;                            // exchange tasks per thread at most (checked by the plan)
    const int tid = threadIdx.x % PIPE_THREADS, team = threadIdx.x / PIPE_THREADS;
    const int R = RT ? RT : P.R, S = RT ? RT : P.S, cin = CT ? CT : P.cin;
    const int NG = P.NG, T = P.T, l = me.l, PP = P.P, B = P.B, RO = P.RO, rmax = P.rmax;
    const int KC = 3 * R + cin, KC4 = KC / 4, C0 = (2 * R + cin) / 4;
    const int hk = P.hk, NR = P.NR;
    const bool last = (l == P.L - 1);
    // ---- shared memory carve-up (weights first: identical to the packed blob) ------------------------------
    float *Wcat = smem;                                  // [NR][KC]      gate rows: row 2j = tanh half, 2j+1 = sigmoid half of channel h0+j
    float *Wos = Wcat + (size_t)NR * KC;                 // [hk/4][R+S][4] columns [h0,h1) of out (rows 0..R) and skip (rows R..R+S)
    float *bsum = Wos + (size_t)(R + S) * hk;            // [NR]
    float *bpo = bsum + NR;                              // [R]           out-conv bias of this layer
    float *bps = bpo + R;                                // [round4(smax)] skip-conv bias of this layer, my slice
    float *fw = bps + round4(P.smax);                    // [R] first conv (layer 0 only)
    float *fb = fw + R;                                  // [R]
    float *xc = smem + P.layer_blob_floats;              // [KC x NU]     x(t-2d) | x(t-d) | c(t) | x(t) in xs_pos order
    float *pre = xc + (size_t)KC * NU;                   // [NR][NU]      (unused: the gate is fused into the reduction)
    float *zs = pre + (size_t)NR * NU;                   // [hk][NU]
    float *part = zs + (size_t)hk * NU;                  // [2][P][RO][NU] inbox of the intra-layer exchange (written by the
                                                         //               cluster peers with st.async), double buffered on the item parity
    unsigned long long *mbar = reinterpret_cast<unsigned long long *>(part + (size_t)2 * P.P * P.RO * NU);   // [2]
    {
        const float4 *src = reinterpret_cast<const float4 *>(P.blobs + me.blob);
        float4 *dst = reinterpret_cast<float4 *>(smem);
        for (int i = threadIdx.x; i < P.layer_blob_floats / 4; i += PIPE_THREADS * TEAMS) dst[i] = __ldg(src + i);
        for (int i = tid; i < (KC + NR + hk) * NU; i += PIPE_THREADS) xc[i] = 0.0f;
        if (threadIdx.x == 0) {
            mbar_init(smem_u32(&mbar[0]), 1);
            mbar_init(smem_u32(&mbar[1]), 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    __syncthreads();
    cluster_sync_all();   // every peer's inbox barriers exist before the first remote store
    (void)team;

    const int o_cur4 = (2 * R + cin) / 4, R4 = R / 4;
    const int d = P.dil[l], depth = P.depth[l], legacy = P.legacy, residual_legacy = P.residual_legacy;
    float *const ring = P.rings + (size_t)me.ring * NG * R * NU;   // [NG][depth][R x NU]; me.ring counts ring slots
    const float *const c_up = P.c_up;
    int *const abort_flag = P.abort_flag;
    // consumer-side mailboxes of this layer (stage l): x | rs; the producer-side ones (stage l+1) follow at + stage_sz
    unsigned long long *const x_in = l == 0 ? P.in : P.mail + (long long)l * P.stage_sz;
    unsigned long long *const rs_in = P.mail + (long long)l * P.stage_sz + P.rs_off;
    unsigned long long *const x_out = P.mail + (long long)(l + 1) * P.stage_sz, *const rs_out = x_out + P.rs_off;

    // ---- thread roles (everything that only depends on tid is computed once, outside the time loop) -------
    const int lane = tid & (LANES - 1), q = tid / LANES;   // dot16: row quad q, partial `lane`
    const bool has_q = q < P.NQ;
    const bool warp_has_q = (tid / 32) * (32 / LANES) < P.NQ;
    const float4 *wq = reinterpret_cast<const float4 *>(Wcat) + (size_t)(has_q ? 4 * q : 0) * KC4;
    const int c1_first = C0 + ((lane - C0) % LANES + LANES) % LANES;   // first chunk >= C0 owned by this lane
    const int rbase = reduce16_base<NV>(lane);
    const int NT2 = (R + S) / CR;                           // threads of the blocked chains (CR rows each)
    const bool has_t2 = tid < NT2;
    const int NT2r = (NT2 + 31) & ~31;
    const int hbase = (NT2r <= PIPE_THREADS - 64) ? NT2r : 0;   // threads [hbase, 256) finish the exchange
    const int hthreads = PIPE_THREADS - hbase;
    // G: where the two rows of this thread go - the owner CTA of the row (slices lo(n,q) = n*q/P) and the row's place
    // in the owner's inbox; addresses in the cluster window, inbox of parity 0
    unsigned g_dst[CR], g_bar[CR];
    const unsigned part_par_bytes = (unsigned)(PP * RO * NU * sizeof(float));
#pragma unroll
    for (int i = 0; i < CR; ++i) { g_dst[i] = 0u; g_bar[i] = 0u; }
    if (has_t2) {
#pragma unroll
        for (int i = 0; i < CR; ++i) {
            const int row = tid + i * NT2;
            int owner, li;
            if (row < R) { owner = ((row + 1) * PP - 1) / R; li = row - (int)((long long)R * owner / PP); }
            else { const int c = row - R; owner = ((c + 1) * PP - 1) / S; li = rmax + c - (int)((long long)S * owner / PP); }
            g_dst[i] = mapa_u32(smem_u32(part + ((size_t)me.p * RO + li) * NU), (unsigned)owner);
            g_bar[i] = mapa_u32(smem_u32(&mbar[0]), (unsigned)owner);
        }
    }
    // X: the rows this CTA owns - out rows [r0, r0+rn) (not in the last layer) then skip rows [s0, s0+sn), NP tasks each
    const int x_rn = last ? 0 : me.rn, x_tasks = (x_rn + me.sn) * NP;
    const unsigned x_bytes = (unsigned)(PP * (x_rn + me.sn) * NU * sizeof(float));   // what lands in my inbox per item
    // D task of this thread (first loop trip): mailbox row m (e-major: m = (r%4)*R/4 + r/4), utterance pair up
    const int d_m = tid / NP, d_up = tid - d_m * NP, d_e = d_m / R4, d_cr = d_m - d_e * R4;
    // prefetch / commit slots
    const int n_tap16 = 2 * R * NU / 4;                     // float4 loads for both taps
    const int n_c16 = NU * (cin / 4);
    int pf_tap[MAX_PF], pf_rem[MAX_PF], pf_dst[MAX_PF];
#pragma unroll
    for (int k = 0; k < MAX_PF; ++k) {
        const int qf = tid + k * PIPE_THREADS, per = n_tap16 / 2;
        pf_tap[k] = -1; pf_rem[k] = 0; pf_dst[k] = 0;
        if (qf < n_tap16) {   // a ring slot is [e*NU/4 + u/4][R/4][4]: segment `seg` of tap `tap` -> chunks tap*R/4.. of xc
            pf_tap[k] = qf / per;
            pf_rem[k] = qf - pf_tap[k] * per;
            const int seg = pf_rem[k] / R4, cr = pf_rem[k] - seg * R4;
            pf_dst[k] = seg * KC4 + pf_tap[k] * R4 + cr;
        }
    }
    const int pc_u = tid / (cin / 4), pc_c4 = tid - pc_u * (cin / 4);

    // ---- register prefetch of the older taps and the conditioning column of the NEXT work item -------------
    float4 pft[MAX_PF], pfc;
    // `slot` = t % depth of the item being prefetched; depth = 2d+1, so (t-2d) % depth = slot+1 and (t-d) % depth =
    // slot+d+1 (both mod depth): no division in the time loop
    auto prefetch = [&](int g, int t, int slot) {
#pragma unroll
        for (int k = 0; k < MAX_PF; ++k) {
            pft[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (pf_tap[k] >= 0) {
                const int tt = t - (2 - pf_tap[k]) * d;
                int sl = slot + 1 + (pf_tap[k] ? d : 0);
                if (sl >= depth) sl -= depth;
                if (tt >= 0 && !(WN_ABLATE & 16))
                    pft[k] = __ldcg(reinterpret_cast<const float4 *>(ring + (size_t)(g * depth + sl) * R * NU) + pf_rem[k]);
            }
        }
        pfc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tid < n_c16) {
            const int b = g * NU + pc_u;
            if (b < B) pfc = __ldg(reinterpret_cast<const float4 *>(c_up + ((long long)b * T + t) * cin) + pc_c4);
        }
    };
    auto commit = [&]() {
#pragma unroll
        for (int k = 0; k < MAX_PF; ++k)
            if (pf_tap[k] >= 0) reinterpret_cast<float4 *>(xc)[pf_dst[k]] = pft[k];
        if (tid < n_c16) {
            const int k = 2 * R + 4 * pc_c4;
            xc[xs_pos<NU>(k, pc_u, KC4)] = pfc.x; xc[xs_pos<NU>(k + 1, pc_u, KC4)] = pfc.y;
            xc[xs_pos<NU>(k + 2, pc_u, KC4)] = pfc.z; xc[xs_pos<NU>(k + 3, pc_u, KC4)] = pfc.w;
        }
    };

    // current-tap weights of this thread's tile, register resident when the chunk count is a compile-time constant
    constexpr int NC1 = RT ? RT / 4 / LANES : 1;
    constexpr bool W1_REGS = RT != 0 && RT % (4 * LANES) == 0 && NC1 <= 2;
    float4 w1[4][NC1];
    if (W1_REGS) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < NC1; ++j) w1[i][j] = has_q ? wq[i * KC4 + c1_first + j * LANES] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // X: the first exchange task of this thread, if it is a skip row - its running-sum word is requested early
    const int x0_task = tid - hbase;
    const bool x0_skip = tid >= hbase && l >= 1 && x0_task < x_tasks && x0_task / NP >= x_rn;
    const int x0_off = x0_skip ? (me.s0 + x0_task / NP - x_rn) * NU + 2 * (x0_task % NP) : 0;
    unsigned long long rs_ea = 0ull, rs_eb = 0ull, xe_a = 0ull, xe_b = 0ull;
    const bool early_x = WN_EARLY_X > 0 && NG >= WN_EARLY_X;

    float2 acc[NV / 2];
    auto phase0 = [&]() {   // everything that does not need x_l(t): taps t-2d, t-d and the conditioning column
#pragma unroll
        for (int i = 0; i < NV / 2; ++i) acc[i] = make_float2(0.0f, 0.0f);
        if (has_q && !(WN_ABLATE & 8)) mac_lane<NU, 4>(acc, wq, KC4, xc, lane, C0);
    };

    int bad = 0;
    // WN_PIPE_STATS: clock64 spent between consecutive stamps, summed over the items, for thread 0 (dot16 / chain
    // thread) and thread hbase (exchange thread)
    long long st_c[14] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool do_stats = STATS && P.stats != nullptr && (tid == 0 || tid == hbase);
    const long long st_begin = STATS ? clock64() : 0;
    long long st_last = st_begin;
#define STAMP(k) if (do_stats) { const long long now_ = clock64(); st_c[k] += now_ - st_last; st_last = now_; }

    prefetch(0, 0, 0);
    commit();
    team_sync(0);
    phase0();

    const long long n_items = (long long)T * NG;
    int t = 0, g = 0, rslot = 0;   // rslot = t % depth
    for (long long it = 0; it < n_items; ++it) {
        const unsigned tag = (unsigned)t + 1u;
        const int slot_g = (t & 1) * NG + g;
        const int par = (int)(it & 1);
        STAMP(13);
        // ---- D: wait for x_l(t) of group g (layer 0: the sample fed back by the head) --------------------------
        // mailbox rows are stored e-major (position m = (r%4)*R/4 + r/4), so that consecutive threads read
        // consecutive mailbox words AND write consecutive words of xc and of the ring
        for (int task = tid; task < R * NP; task += PIPE_THREADS) {
            int m = d_m, up = d_up, e = d_e, cr = d_cr;
            if (task != tid) { m = task / NP; up = task - m * NP; e = m / R4; cr = m - e * R4; }
            float x0, x1;
            const bool early = early_x && task == tid && ll_tag(xe_a) == tag && ll_tag(xe_b) == tag;
            if (early) { x0 = ll_val(xe_a); x1 = ll_val(xe_b); }
            if (l == 0) {   // first_conv.incremental_step (wavenet.py:826) on the value the head fed back
                const int r = 4 * cr + e;
                float s0 = x0, s1 = x1;
                if (!early) bad |= poll_pair(x_in + slot_g * NU + 2 * up, tag, s0, s1, abort_flag);
                x0 = __fadd_rn(__fmaf_rn(s0, fw[r], 0.0f), fb[r]);
                x1 = __fadd_rn(__fmaf_rn(s1, fw[r], 0.0f), fb[r]);
            } else if (!early) {
                bad |= poll_pair(x_in + (slot_g * R + m) * NU + 2 * up, tag, x0, x1, abort_flag);
            }
            const int seg = e * (NU / 4) + (up >> 1);
            *reinterpret_cast<float2 *>(xc + (size_t)(seg * KC4 + o_cur4 + cr) * 4 + 2 * (up & 1)) = make_float2(x0, x1);
            if (!(WN_ABLATE & 16))
                __stcg(reinterpret_cast<float2 *>(ring + (size_t)(g * depth + rslot) * R * NU +
                                                  (size_t)(seg * R4 + cr) * 4 + 2 * (up & 1)), make_float2(x0, x1));
        }
        STAMP(0);
        if (team_sync_or(0, bad)) break;                                                             // (1)
        STAMP(1);
        int g2 = g + 1, t2 = t, rslot2 = rslot;
        if (g2 >= NG) { g2 = 0; t2 = t + 1; rslot2 = rslot + 1 == depth ? 0 : rslot + 1; }
        if (t2 < T) prefetch(g2, t2, rslot2);   // ring slots of (t2 - d, g2) were written in an earlier item or just above
        STAMP(2);
        // ---- E: current tap, tree reduction, bias (modules.py:294-297, :497-501) ----------------------------
        if (warp_has_q) {   // warp-uniform: the butterfly shuffles need every lane of the warp
            if (has_q && !(WN_ABLATE & 2)) {
                if (W1_REGS) mac_lane_wreg<NU, 4, NC1>(acc, w1, KC4, xc, c1_first);
                else mac_lane<NU, 4>(acc, wq, KC4, xc, c1_first, KC4);
            }
            STAMP(3);
            float a[NV];
            unpack2<NV>(acc, a);
            if (!(WN_ABLATE & 32)) reduce16<NV>(a, lane);
            STAMP(4);
            // gate (modules.py:510).  Rows of a quad are (tanh j, sigmoid j, tanh j+1, sigmoid j+1), so the two halves of
            // one gated value sit in lanes that differ in the bit the butterfly used for tile-row bit 0 (mask 2): every
            // lane evaluates ONE transcendental, one shuffle brings the partner, the tanh lane stores z.
#pragma unroll
            for (int i = 0; i < (NV >= LANES ? NV / LANES : 1); ++i) {
                const int v = rbase + i, ti = v / NU, u = v % NU, row = 4 * q + ti;
                float g_act = 0.0f;
                if (has_q) {
                    const float pv = __fadd_rn(a[i], bsum[row]);
                    g_act = (WN_ABLATE & 1) ? pv : gate_act(pv, (ti & 1) != 0);
                }
                const float o_act = __shfl_xor_sync(0xffffffffu, g_act, 2);
                if (has_q && !(ti & 1)) zs[(size_t)(2 * q + (ti >> 1)) * NU + u] = __fmul_rn(g_act, o_act);
            }
        }
        STAMP(5);
        // the running skip sum left layer l-1 together with x_l(t): ask for it now, look at it after the chains
        if (WN_EARLY_RS && x0_skip) ld_ll2(rs_in + slot_g * S * NU + x0_off, rs_ea, rs_eb);
        team_sync(0);                                                                                // (3)
        STAMP(6);
        // ---- G: my column block of the out / skip 1x1s -> partial sums, each row sent to the CTA of this layer that
        //         owns it (reduce-scatter inside the cluster: st.async into the owner's inbox + complete_tx on its barrier)
        if (has_t2) {
            float2 a2[CR * NP];
#pragma unroll
            for (int i = 0; i < CR * NP; ++i) a2[i] = make_float2(0.0f, 0.0f);
            if (!(WN_ABLATE & 4)) chain_tile<NU, CR>(a2, reinterpret_cast<const float4 *>(Wos), R + S, tid, NT2, zs, hk / 4);
            STAMP(7);
#pragma unroll
            for (int i = 0; i < CR; ++i) {
                if ((last && tid + i * NT2 < R) || (WN_ABLATE & 64)) continue;
                const unsigned dst = g_dst[i] + (par ? part_par_bytes : 0u), bar = g_bar[i] + (par ? 8u : 0u);
#pragma unroll
                for (int u4 = 0; u4 < NU / 4; ++u4)
                    st_async4(dst + 16u * u4, make_float4(a2[i * NP + 2 * u4].x, a2[i * NP + 2 * u4].y,
                                                          a2[i * NP + 2 * u4 + 1].x, a2[i * NP + 2 * u4 + 1].y), bar);
            }
        }
        STAMP(8);
        // ---- X (first half): what the rows I own need besides the partial sums - the residual input x_l(t) (out rows,
        //      modules.py:515-520) or the running skip sum through layer l-1 (skip rows, wavenet.py:833-836)
        float xa0[XT], xa1[XT];
        if (tid >= hbase) {
            if (tid == hbase && !(WN_ABLATE & 64)) mbar_expect(smem_u32(&mbar[par]), x_bytes);
#pragma unroll
            for (int k = 0; k < XT; ++k) {
                const int task = tid - hbase + k * hthreads;
                xa0[k] = 0.0f; xa1[k] = 0.0f;
                if (task < x_tasks) {
                    const int rl = task / NP, up = task - rl * NP;
                    if (rl < x_rn) {
                        const int r = me.r0 + rl, e = r & 3, cr = r >> 2, seg = e * (NU / 4) + (up >> 1);
                        const float2 xv = *reinterpret_cast<const float2 *>(xc + (size_t)(seg * KC4 + o_cur4 + cr) * 4 + 2 * (up & 1));
                        xa0[k] = xv.x; xa1[k] = xv.y;
                    } else if (l >= 1) {
                        const int c = me.s0 + rl - x_rn;
                        if (k == 0 && WN_EARLY_RS && ll_tag(rs_ea) == tag && ll_tag(rs_eb) == tag) {
                            xa0[k] = ll_val(rs_ea); xa1[k] = ll_val(rs_eb);
                        } else {
                            bad |= poll_pair(rs_in + (slot_g * S + c) * NU + 2 * up, tag, xa0[k], xa1[k], abort_flag);
                        }
                    }
                }
            }
        }
        STAMP(9);
        // ---- next item: older taps + conditioning into shared memory ----------------------------------------------
        if (t2 < T) commit();
        STAMP(10);
        if (team_sync_or(0, bad)) break;                                                             // (4)
        STAMP(11);
        // ---- X (second half): partial sums of my rows from all P CTAs of the layer, ascending p, bias, residual /
        //      running skip sum; x_{l+1}(t) goes to every CTA of layer l+1, the skip sum to (l+1, p) or the head
        if (tid >= hbase) {
            unsigned spins = 0;
            unsigned long long t_start = 0;
            while (!(WN_ABLATE & 64) && !mbar_try(smem_u32(&mbar[par]), (unsigned)(it >> 1) & 1u)) {
                if ((++spins & 0xffu) == 0) {
                    if (*(volatile int *)abort_flag) { bad = 1; break; }
                    const unsigned long long now = global_timer_ns();
                    if (t_start == 0) t_start = now;
                    else if (now - t_start > WATCHDOG_NS) { atomicExch(abort_flag, 1); bad = 1; break; }
                }
            }
            const float *pin = part + (size_t)par * PP * RO * NU;
#pragma unroll
            for (int k = 0; k < XT; ++k) {
                const int task = tid - hbase + k * hthreads;
                if (task < x_tasks && !bad) {
                    const int rl = task / NP, up = task - rl * NP;
                    const bool is_out = rl < x_rn;
                    const int li = is_out ? rl : rmax + rl - x_rn;
                    float2 y = *reinterpret_cast<const float2 *>(pin + (size_t)li * NU + 2 * up);
#pragma unroll
                    for (int pp = 1; pp < MAX_P; ++pp)
                        if (pp < PP) {
                            const float2 v = *reinterpret_cast<const float2 *>(pin + ((size_t)pp * RO + li) * NU + 2 * up);
                            y.x = __fadd_rn(y.x, v.x); y.y = __fadd_rn(y.y, v.y);
                        }
                    if (is_out) {
                        const int r = me.r0 + rl;
                        y.x = __fadd_rn(__fadd_rn(y.x, bpo[r]), xa0[k]); y.y = __fadd_rn(__fadd_rn(y.y, bpo[r]), xa1[k]);
                        if (residual_legacy) { y.x = __fmul_rn(y.x, WN_SQRT_HALF); y.y = __fmul_rn(y.y, WN_SQRT_HALF); }
                        st_ll2(x_out + (slot_g * R + (r & 3) * R4 + (r >> 2)) * NU + 2 * up, y.x, y.y, tag);
                    } else {
                        const int cl = rl - x_rn;
                        y.x = __fadd_rn(y.x, bps[cl]); y.y = __fadd_rn(y.y, bps[cl]);
                        if (l >= 1) {
                            y.x = __fadd_rn(xa0[k], y.x); y.y = __fadd_rn(xa1[k], y.y);
                            if (legacy) { y.x = __fmul_rn(y.x, WN_SQRT_HALF); y.y = __fmul_rn(y.y, WN_SQRT_HALF); }
                        }
                        st_ll2(rs_out + (slot_g * S + me.s0 + cl) * NU + 2 * up, y.x, y.y, tag);
                    }
                }
            }
        }
        STAMP(12);
        // ---- the x(t)-independent part of the next item; its x words are requested first ----------------------------------
        if (early_x && t2 < T && tid < R * NP) {
            const int slot_g2 = (t2 & 1) * NG + g2;
            ld_ll2(l == 0 ? x_in + slot_g2 * NU + 2 * d_up : x_in + (slot_g2 * R + d_m) * NU + 2 * d_up, xe_a, xe_b);
        }
        if (t2 < T) phase0();
        g = g2; t = t2; rslot = rslot2;
    }
    if (do_stats) {
        long long *o = P.stats + (long long)blockIdx.x * 32 + (tid == 0 ? 0 : 16);
        for (int i = 0; i < 14; ++i) o[i] = st_c[i];
        o[14] = clock64() - st_begin; o[15] = n_items;
    }
#undef STAMP
    __syncthreads();
    cluster_sync_all();   // no CTA of the cluster leaves while a peer may still store into its inbox
}


// ---------------------------------------------------------------------------------------------------------------
// head CTA: last skip accumulation, relu -> 1x1 -> relu -> 1x1 (wavenet.py:833-844), sampler (:850-859) and the
// feedback value (:877) for the groups g = me.l, me.l + NH, ...
// ---------------------------------------------------------------------------------------------------------------
template <int NU, bool STATS>
__device__ void head_role(const PipeParams &P, const PipeCta &me, float *smem) {
    constexpr int NV = 4 * NU, NP = NU / 2;
    const int tid = threadIdx.x % PIPE_THREADS, team = threadIdx.x / PIPE_THREADS;
    const int S = P.S, O = P.O, NG = P.NG, T = P.T, L = P.L;
    const int S4 = S / 4;
    float *W1 = smem;                          // [S][S]   row n = output channel (transposed TF kernel)
    float *W2 = W1 + (size_t)S * S;            // [O][S]
    float *b1 = W2 + (size_t)O * S;            // [S]
    float *b2 = b1 + S;                        // [round4(O)]
    float *hs = smem + P.head_blob_floats + (size_t)team * (2 * S + round4(O)) * NU;   // per team:
                                               // [S x NU] relu(skips), xs_pos order
    float *h1 = hs + (size_t)S * NU;           // [S x NU] xs_pos order
    float *outv = h1 + (size_t)S * NU;         // [O][NU]
    {
        const float4 *src = reinterpret_cast<const float4 *>(P.blobs + me.blob);
        float4 *dst = reinterpret_cast<float4 *>(smem);
        for (int i = threadIdx.x; i < P.head_blob_floats / 4; i += PIPE_THREADS * TEAMS) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    const int g_first = me.l * TEAMS + team, g_step = P.NH * TEAMS;   // this team's groups
    const unsigned long long *rs_in = P.mail + (long long)L * P.stage_sz + P.rs_off;
    const int lane = tid & (LANES - 1), q16 = tid / LANES;
    const int rbase4 = reduce16_base<NV>(lane), rbase1 = reduce16_base<NU>(lane);

    // initial input (wavenet.py:433-445): raw 0.0, mulaw mulaw(0) = 0.0
    for (int g = g_first; g < NG; g += g_step)
        if (tid < NP) {
            float i0 = 0.0f, i1 = 0.0f;
            if (P.inputs_mode) {   // WaveNet.step (wavenet.py:650-721): the first input is x[0]
                const int b0 = g * NU + 2 * tid;
                if (b0 < P.B) i0 = P.teacher[(long long)b0 * T];
                if (b0 + 1 < P.B) i1 = P.teacher[(long long)(b0 + 1) * T];
            }
            st_ll2(P.in + ((long long)0 * NG + g) * NU + 2 * tid, i0, i1, 1u);
        }

    int bad = 0;
    long long st_c[4] = {0, 0, 0, 0};
    const bool do_stats = STATS && P.stats != nullptr && threadIdx.x == 0;
    const long long st_begin = clock64();
    for (int t = 0; t < T; ++t) {
        const unsigned tag = (unsigned)t + 1u;
        for (int g = g_first; g < NG; g += g_step) {
            const int slot_g = (t & 1) * NG + g;
            const long long c0 = do_stats ? clock64() : 0;
            // the sampler's inputs do not depend on the network: fetch them while the skip sums are still in flight
            float nz0 = 0.0f, tch = 0.0f;
            if (tid < NU && g * NU + tid < P.B) {
                const long long b = g * NU + tid;
                if (O == 2 && P.noise) nz0 = __ldg(P.noise + ((long long)t * P.B + b) * P.NZ);
                if (P.teacher) tch = P.inputs_mode ? (t + 1 < T ? __ldg(P.teacher + b * T + t + 1) : 0.0f) : __ldg(P.teacher + b * T + t);
            }
            // ---- relu of the finished skip sum (wavenet.py:833-840; the last layer's CTAs already added their part) ----
            for (int task = tid; task < S * NP; task += PIPE_THREADS) {
                const int c = task / NP, up = task - c * NP;
                float s0, s1;
                bad |= poll_pair(rs_in + (slot_g * S + c) * NU + 2 * up, tag, s0, s1, P.abort_flag);
                *reinterpret_cast<float2 *>(hs + xs_pos<NU>(c, 2 * up, S4)) =
                    make_float2(s0 > 0.f ? s0 : 0.f, s1 > 0.f ? s1 : 0.f);
            }
            if (team_sync_or(team, bad)) return;
            const long long c1 = do_stats ? clock64() : 0;
            // ---- final_convolution_1 + relu: dot16, 4 rows x NU per thread -------------------------------------
            for (int qb = 0; qb < S4; qb += PIPE_THREADS / LANES) {
                const int q = qb + q16;
                float2 acc[NV / 2];
#pragma unroll
                for (int i = 0; i < NV / 2; ++i) acc[i] = make_float2(0.0f, 0.0f);
                if (q < S4) mac_lane<NU, 4>(acc, reinterpret_cast<const float4 *>(W1) + (size_t)4 * q * S4, S4, hs, lane, S4);
                float a[NV];
                unpack2<NV>(acc, a);
                reduce16<NV>(a, lane);
                if (q < S4) {
#pragma unroll
                    for (int i = 0; i < (NV >= LANES ? NV / LANES : 1); ++i) {
                        const int v = rbase4 + i, row = 4 * q + v / NU, u = v % NU;
                        const float y = __fadd_rn(a[i], b1[row]);
                        h1[xs_pos<NU>(row, u, S4)] = y > 0.0f ? y : 0.0f;
                    }
                }
            }
            team_sync(team);
            const long long c2 = do_stats ? clock64() : 0;
            // ---- final_convolution_2: dot16, 1 row x NU per 16 lanes ----------------------------------------------
            for (int nb = 0; nb < O; nb += PIPE_THREADS / LANES) {
                const int n = nb + q16;
                float2 acc[NP];
#pragma unroll
                for (int u = 0; u < NP; ++u) acc[u] = make_float2(0.0f, 0.0f);
                if (n < O) mac_lane<NU, 1>(acc, reinterpret_cast<const float4 *>(W2) + (size_t)n * S4, S4, h1, lane, S4);
                float a[NU];
                unpack2<NU>(acc, a);
                reduce16<NU>(a, lane);
                if (n < O) {
#pragma unroll
                    for (int i = 0; i < (NU >= LANES ? NU / LANES : 1); ++i) {
                        const int u = rbase1 + i;
                        const float y = __fadd_rn(a[i], b2[n]);
                        outv[(size_t)n * NU + u] = y;
                        const int b = g * NU + u;
                        if (P.raw && b < P.B && (lane >> (NU == 4 ? 2 : 3)) == 0)   // one of the lanes holding this sum
                            P.raw[((long long)b * T + t) * O + n] = y;                     // wavenet.py:847
                    }
                }
            }
            team_sync(team);
            const long long c3 = do_stats ? clock64() : 0;
            // ---- sampling + feedback ------------------------------------------------------------------------------
            if (tid < 32) {
                const int u = tid, b = g * NU + u;
                float nxt = 0.0f;
                if (u < NU && b < P.B && P.inputs_mode) {
                    nxt = tch;
                } else if (u < NU && b < P.B) {
                    const float smp = (O == 2) ? pipe_sample_gaussian(P, outv + u, NU, nz0)
                                               : pipe_sample_mol(P, outv + u, NU, P.noise + ((long long)t * P.B + b) * P.NZ);
                    P.audio[(long long)b * T + t] = P.input_type == WN_INPUT_MULAW ? wn_inv_mulawf(smp) : smp;
                    nxt = P.teacher ? tch : smp;
                }
                const float other = __shfl_down_sync(0xffffffffu, nxt, 1);
                if (u < NU && !(u & 1) && t + 1 < T)
                    st_ll2(P.in + ((long long)((t + 1) & 1) * NG + g) * NU + u, nxt, other, tag + 1u);
            }
            if (do_stats) { st_c[0] += c1 - c0; st_c[1] += c2 - c1; st_c[2] += c3 - c2; st_c[3] += clock64() - c3; }
        }
    }
    if (do_stats) {
        long long *o = P.stats + (long long)blockIdx.x * 32;
        o[0] = st_c[0]; o[1] = st_c[1]; o[2] = st_c[2]; o[3] = st_c[3];
        o[14] = clock64() - st_begin; o[15] = T;
    }
}

template <int NU, int RT, int CT, bool STATS>
__global__ void __launch_bounds__(PIPE_THREADS * TEAMS, 1) wn_loop_pipe_kernel(const PipeParams P) {
    extern __shared__ __align__(16) float pipe_smem[];
    const PipeCta me = P.ctas[blockIdx.x];
    if (me.role == 0) layer_role<NU, RT, CT, STATS>(P, me, pipe_smem);
    else head_role<NU, STATS>(P, me, pipe_smem);
}

}  // namespace

// ===============================================================================================================
// host side: plan (once per model), buffers and launch (per call)
// ===============================================================================================================
struct WnPipePlan {
    PipeParams prm;                 // model + plan part filled at plan time
    int max_heads = 16;
    int max_clusters = 0;           // clusters of P CTAs that can be resident at once
    float *d_blobs = nullptr;
    PipeCta *d_ctas = nullptr;
    unsigned long long *d_mail = nullptr;   // op | xr | sp | rs | in
    size_t mail_words = 0;
    float *d_rings = nullptr;
    size_t ring_floats = 0;
    long long ring_slots = 0;        // sum over layer CTAs of their ring depth
    int *d_abort = nullptr;
    int *h_abort = nullptr;          // pinned mirror, copied after every launch
    long long *d_stats = nullptr;    // WN_PIPE_STATS=<file>: per-CTA clock64 counters, dumped after each synchronised call
    int stats_ctas = 0;
    size_t smem_bytes = 0;
    // groups of 8 utterances (throughput regime): same weight blobs, more shared memory per CTA; 0 = not available
    size_t smem_bytes8 = 0;
    int max_heads8 = 0, nu8_min = 112, last_nu = 4;
};

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}
static int r4(int v) { return (v + 3) & ~3; }

void wn_pipe_free(wn_handle *h) {
    WnPipePlan *pl = h->pipe;
    if (!pl) return;
    if (pl->d_blobs) cudaFree(pl->d_blobs);
    if (pl->d_ctas) cudaFree(pl->d_ctas);
    if (pl->d_mail) cudaFree(pl->d_mail);
    if (pl->d_rings) cudaFree(pl->d_rings);
    if (pl->d_abort) cudaFree(pl->d_abort);
    if (pl->h_abort) cudaFreeHost(pl->h_abort);
    if (pl->d_stats) cudaFree(pl->d_stats);
    delete pl;
    h->pipe = nullptr;
}

void wn_pipe_fill_info(wn_handle *h, wn_info *info) {
    if (!h->pipe) return;
    info->pipeline_ctas = h->pipe->prm.L * h->pipe->prm.P;
    info->pipeline_ctas_per_layer = h->pipe->prm.P;
    info->pipeline_group_size = h->pipe->last_nu;
}

// Specialised for the shipped default block (R = S = 128, 80 mel bands), generic otherwise; the variants with the
// clock64 counters are separate instantiations so that the production kernel carries no timing registers.
static const void *pipe_kernel(const PipeParams &p, bool stats) {
    const bool dflt = p.R == 128 && p.S == 128 && p.cin == 80;
    if (p.NU == 8) {   // groups of 8 utterances: the default block only (throughput regime, see wn_launch_loop_pipe)
        if (!dflt) return nullptr;
        return stats ? (const void *)wn_loop_pipe_kernel<8, 128, 80, true> : (const void *)wn_loop_pipe_kernel<8, 128, 80, false>;
    }
    if (stats) return dflt ? (const void *)wn_loop_pipe_kernel<4, 128, 80, true> : (const void *)wn_loop_pipe_kernel<4, 0, 0, true>;
    return dflt ? (const void *)wn_loop_pipe_kernel<4, 128, 80, false> : (const void *)wn_loop_pipe_kernel<4, 0, 0, false>;
}

// How many clusters of P CTAs with `smem` bytes of dynamic shared memory fit the device at once.
static int pipe_max_clusters(const PipeParams &p, size_t smem) {
    int best = 0;
    for (int stats = 0; stats < 2; ++stats) {
        const void *fn = pipe_kernel(p, stats != 0);
        if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) { cudaGetLastError(); return 0; }
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(p.P * 64); cfg.blockDim = dim3(PIPE_THREADS * TEAMS); cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = p.P; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, fn, &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
        best = stats ? std::min(best, n) : n;
    }
    return best;
}

// Fills the geometry that follows from (model, P, NU); returns the dynamic shared memory the kernel needs, 0 if the
// split does not fit the thread roles.
static size_t pipe_geometry(const WnModelDev &m, PipeParams &p) {
    auto lo = [&](int n, int q) { return (int)((long long)n * q / p.P); };
    int hmax = 0;
    p.smax = 0; p.rmax = 0;
    for (int q = 0; q < p.P; ++q) {
        hmax = std::max(hmax, lo(m.H, q + 1) - lo(m.H, q));
        p.smax = std::max(p.smax, lo(m.S, q + 1) - lo(m.S, q));
        p.rmax = std::max(p.rmax, lo(m.R, q + 1) - lo(m.R, q));
    }
    p.RO = p.rmax + p.smax;
    p.hq = (hmax + 1) & ~1;
    p.hk = r4(hmax);
    p.NR = 2 * p.hq;
    p.NQ = p.NR / 4;
    p.KC = 3 * m.R + m.cin;
    p.KC4 = p.KC / 4;
    p.C0 = (2 * m.R + m.cin) / 4;
    if (p.NQ * LANES > PIPE_THREADS || (m.R + m.S) / CR > PIPE_THREADS || (m.R + m.S) % CR) return 0;
    if (p.NU * (m.cin / 4) > PIPE_THREADS || 2 * m.R * p.NU / 4 > MAX_PF * PIPE_THREADS) return 0;
    {   // the exchange tasks of the rows a CTA owns must fit 4 per thread of the threads that do not run the chains
        const int nt2r = ((m.R + m.S) / CR + 31) & ~31, hthreads = nt2r <= PIPE_THREADS - 64 ? PIPE_THREADS - nt2r : PIPE_THREADS;
        if (p.RO * (p.NU / 2) > 4 * hthreads) return 0;
    }
    p.layer_blob_floats = p.NR * p.KC + (m.R + m.S) * p.hk + p.NR + m.R + r4(p.smax) + 2 * m.R;
    p.head_blob_floats = m.S * m.S + m.O * m.S + m.S + r4(m.O);
    const size_t layer_smem = sizeof(float) * ((size_t)p.layer_blob_floats + (size_t)(p.KC + p.NR + p.hk) * p.NU +
                                               (size_t)2 * p.P * p.RO * p.NU) + 2 * sizeof(unsigned long long);
    const size_t head_smem = sizeof(float) * ((size_t)p.head_blob_floats + (size_t)TEAMS * (2 * m.S + r4(m.O)) * p.NU);
    return std::max(layer_smem, head_smem);
}

int wn_pipe_plan(wn_handle *h) {
    const wn_config &c = h->cfg;
    WnModelDev &m = h->m;
    auto no = [&](const std::string &why) { h->err = why; return -1; };
    if (c.input_type == WN_INPUT_MULAW_QUANTIZE) return no("one-hot (mulaw-quantize) input runs on the generic engine");
    if (m.ksize != 3) return no("kernel_size != 3 runs on the generic engine");
    if (m.cin <= 0 || m.cin % 4) return no("needs local conditioning with cin_channels % 4 == 0");
    if (m.gin > 0) return no("global conditioning runs on the generic engine");
    if (m.R % 4 || m.S % 4) return no("residual / skip channel counts must be multiples of 4");
    WnPipePlan *pl = new WnPipePlan();
    PipeParams &p = pl->prm;
    memset(&p, 0, sizeof(p));
    p.L = m.L; p.R = m.R; p.H = m.H; p.S = m.S; p.O = m.O; p.cin = m.cin; p.NZ = m.NZ;
    p.input_type = m.input_type; p.legacy = m.legacy; p.residual_legacy = m.residual_legacy;
    p.log_scale_min = m.log_scale_min; p.log_scale_min_gauss = m.log_scale_min_gauss;
    for (int l = 0; l < m.L; ++l) { p.dil[l] = m.dil[l]; p.depth[l] = 2 * m.dil[l] + 1; }
    auto bail = [&](const std::string &why) { delete pl; h->err = why; return -1; };
    p.NU = env_int("WN_PIPE_NU", 4);
    if (p.NU != 4 && p.NU != 8) return bail("WN_PIPE_NU must be 4 or 8");
    if (p.NU == 8 && !(m.R == 128 && m.S == 128 && m.cin == 80)) p.NU = 4;
    int max_optin = 0;
    cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device);
    // CTAs per layer: the caller's ksplit (it fixes the summation order), else WN_PIPE_P, else the widest split that
    // leaves SMs for the head CTAs
    // The P CTAs of a layer form one thread-block cluster; all L layer clusters and at least one cluster of head CTAs
    // must be resident at once (clusters are placed inside one GPC, so this is tighter than the SM count).
    auto feasible = [&](int P) {
        if (P < 1 || P > MAX_P || P > m.H || P > m.S || P > m.R) return false;
        if (m.L * P + P > h->sm_count) return false;
        p.P = P;
        const size_t smem = pipe_geometry(m, p);
        if (smem == 0 || smem > (size_t)max_optin) return false;
        pl->max_clusters = pipe_max_clusters(p, smem);
        return pl->max_clusters >= m.L + 1;
    };
    int P = c.ksplit > 0 ? c.ksplit : env_int("WN_PIPE_P", 0);
    if (P > 0) {
        if (!feasible(P)) return bail("split over " + std::to_string(P) + " CTAs per layer does not fit (shared memory " +
                                      std::to_string(max_optin) + " B, " + std::to_string(h->sm_count) + " SMs)");
    } else {
        const int want_heads = 16;   // one head CTA per group up to 64 utterances (the head serves its groups serially)
        for (P = std::min(MAX_P, std::max(1, (h->sm_count - want_heads) / m.L)); P >= 1; --P)
            if (feasible(P)) break;
        if (P < 1)
            for (P = MAX_P; P >= 1; --P)
                if (feasible(P)) break;
        if (P < 1) return bail("fp32 weight slices do not fit shared memory for any split");
    }
    p.P = P;
    pl->smem_bytes = pipe_geometry(m, p);
    pl->max_clusters = pipe_max_clusters(p, pl->smem_bytes);
    pl->max_heads = std::min((32 / p.P > 0 ? 32 / p.P : 1) * p.P, (pl->max_clusters - m.L) * p.P);   // whole clusters of head CTAs
    m.ksplit = p.P;
    // Groups of 8 utterances halve the number of hand-offs per step: +11..14 % once the engine is throughput bound (from
    // ~112 utterances; measured 128: 85.0 -> 76.3 us/step, 256: 169.9 -> 148.6), but a longer stage latency below that
    // (64 utterances: 47.7 -> 73.0 us/step).  The launch picks the group size from the batch; WN_PIPE_NU forces it.
    if (p.NU == 4 && !getenv("WN_PIPE_NU") && m.R == 128 && m.S == 128 && m.cin == 80) {
        PipeParams p8 = p;
        p8.NU = 8;
        const size_t smem8 = pipe_geometry(m, p8);
        if (smem8 != 0 && smem8 <= (size_t)max_optin) {
            const int mc8 = pipe_max_clusters(p8, smem8);
            if (mc8 >= m.L + 1) {
                pl->smem_bytes8 = smem8;
                pl->max_heads8 = std::min(pl->max_heads, (mc8 - m.L) * p.P);
            }
        }
        pl->nu8_min = env_int("WN_PIPE_NU8_MIN", 112);
    }
    pl->last_nu = p.NU;

    // ---- pack weights: one blob per layer CTA in its shared-memory layout, one head blob ---------------------
    auto lo = [&](int n, int q) { return (int)((long long)n * q / p.P); };
    auto Wt = [&](const std::string &n) -> const float * {
        auto it = h->host_w.find(n);
        return it == h->host_w.end() ? nullptr : it->second.data.data();
    };
    const int G = 2 * m.H, R = m.R, S = m.S;
    std::vector<float> blobs((size_t)m.L * p.P * p.layer_blob_floats + p.head_blob_floats, 0.0f);
    std::vector<PipeCta> ctas;
    long long ring_slots = 0;
    for (int l = 0; l < m.L; ++l) {
        std::string pre = "ResidualConv1DGLU_" + std::to_string(l) + "/residual_block_";
        const float *cw = Wt(pre + "causal_conv/kernel"), *cb = Wt(pre + "causal_conv/bias");
        const float *iw = Wt(pre + "cin_conv/kernel"), *ib = Wt(pre + "cin_conv/bias");
        const float *ow = Wt(pre + "out_conv/kernel"), *sw = Wt(pre + "skip_conv/kernel");
        const float *pob = Wt(pre + "out_conv/bias"), *psb = Wt(pre + "skip_conv/bias");
        const float *f = Wt("input_convolution/kernel"), *fbb = Wt("input_convolution/bias");
        for (int q = 0; q < p.P; ++q) {
            PipeCta ct;
            memset(&ct, 0, sizeof(ct));
            ct.role = 0; ct.l = l; ct.p = q;
            ct.h0 = lo(m.H, q); ct.hn = lo(m.H, q + 1) - ct.h0;
            ct.r0 = lo(R, q); ct.rn = lo(R, q + 1) - ct.r0;
            ct.s0 = lo(S, q); ct.sn = lo(S, q + 1) - ct.s0;
            ct.blob = (long long)(l * p.P + q) * p.layer_blob_floats;
            ct.ring = ring_slots;
            ring_slots += p.depth[l];
            float *b = blobs.data() + ct.blob;
            float *Wcat = b, *Wos = Wcat + (size_t)p.NR * p.KC, *bsum = Wos + (size_t)(R + S) * p.hk, *bpo = bsum + p.NR,
                  *bps = bpo + R, *fw = bps + r4(p.smax), *fbv = fw + R;
            for (int half = 0; half < 2; ++half)
                for (int j = 0; j < ct.hn; ++j) {
                    // row 2j = tanh half of gate channel h0+j, row 2j+1 = its sigmoid half (modules.py:494 split order)
                    const int row = 2 * j + half, col = half * m.H + ct.h0 + j;
                    float *w = Wcat + (size_t)row * p.KC;
                    for (int k = 0; k < 2 * R; ++k) w[k] = cw[(size_t)k * G + col];
                    for (int k = 0; k < m.cin; ++k) w[2 * R + k] = iw[(size_t)k * G + col];
                    for (int k = 0; k < R; ++k) w[2 * R + m.cin + k] = cw[(size_t)(2 * R + k) * G + col];
                    bsum[row] = (cb ? cb[col] : 0.0f) + (ib ? ib[col] : 0.0f);
                }
            for (int row = 0; row < R + S; ++row)
                for (int k = 0; k < ct.hn; ++k) {
                    const float v = row < R ? ow[(size_t)(ct.h0 + k) * R + row] : sw[(size_t)(ct.h0 + k) * S + (row - R)];
                    Wos[((size_t)(k / 4) * (R + S) + row) * 4 + (k % 4)] = v;
                }
            for (int r = 0; r < R; ++r) { bpo[r] = pob ? pob[r] : 0.0f; fw[r] = f[r]; fbv[r] = fbb ? fbb[r] : 0.0f; }
            for (int j = 0; j < ct.sn; ++j) bps[j] = psb ? psb[ct.s0 + j] : 0.0f;
            ctas.push_back(ct);
        }
    }
    pl->ring_slots = ring_slots;
    {
        const long long hb = (long long)m.L * p.P * p.layer_blob_floats;
        float *b = blobs.data() + hb;
        float *W1 = b, *W2 = W1 + (size_t)S * S, *b1 = W2 + (size_t)m.O * S, *b2 = b1 + S;
        const float *w1 = Wt("final_convolution_1/kernel"), *bb1 = Wt("final_convolution_1/bias");
        const float *w2 = Wt("final_convolution_2/kernel"), *bb2 = Wt("final_convolution_2/bias");
        for (int n = 0; n < S; ++n) {
            for (int k = 0; k < S; ++k) W1[(size_t)n * S + k] = w1[(size_t)k * S + n];
            b1[n] = bb1 ? bb1[n] : 0.0f;
        }
        for (int n = 0; n < m.O; ++n) {
            for (int k = 0; k < S; ++k) W2[(size_t)n * S + k] = w2[(size_t)k * m.O + n];
            b2[n] = bb2 ? bb2[n] : 0.0f;
        }
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
    if (cudaFuncSetAttribute(pipe_kernel(p, false), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl->smem_bytes) != cudaSuccess ||
        cudaFuncSetAttribute(pipe_kernel(p, true), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl->smem_bytes) != cudaSuccess) {
        wn_pipe_free(h);
        h->err = std::string("pipeline engine: ") + cudaGetErrorString(cudaGetLastError());
        return -1;
    }
    return 0;
}

int wn_pipe_check_abort(wn_handle *h) {   // call after the stream has been synchronised
    if (!h->pipe || !h->pipe->h_abort) return WN_OK;
    const int code = *h->pipe->h_abort;
    if (h->pipe->d_stats && h->pipe->stats_ctas > 0 && getenv("WN_PIPE_STATS")) {   // development aid
        std::vector<long long> st((size_t)h->pipe->stats_ctas * 32);
        if (cudaMemcpy(st.data(), h->pipe->d_stats, st.size() * sizeof(long long), cudaMemcpyDeviceToHost) == cudaSuccess) {
            FILE *f = fopen(getenv("WN_PIPE_STATS"), "a");
            if (f) {
                const PipeParams &q = h->pipe->prm;
                fprintf(f, "# pipe stats: L=%d P=%d NU=%d ctas=%d  cols: cta role(0 layer,1 head) thread(0 first,1 exchange) "
                           "stamp0..stamp13 total items\n", q.L, q.P, q.NU, h->pipe->stats_ctas);
                for (int i = 0; i < h->pipe->stats_ctas; ++i)
                    for (int th = 0; th < 2; ++th) {
                        const long long *o = st.data() + (size_t)i * 32 + th * 16;
                        if (o[15] == 0) continue;
                        fprintf(f, "%d %d %d", i, i >= q.L * q.P ? 1 : 0, th);
                        for (int k = 0; k < 16; ++k) fprintf(f, " %lld", o[k]);
                        fprintf(f, "\n");
                    }
                fclose(f);
            }
        }
        h->pipe->stats_ctas = 0;
    }
    if (code == 0) return WN_OK;
    h->err = "pipeline engine: watchdog - a CTA waited too long for a peer";
    return WN_ERR_WATCHDOG;
}

int wn_launch_loop_pipe(wn_handle *h, const float *c_up, const float *noise, const float *teacher, int inputs_mode,
                        int B, int T, float *audio, float *raw, cudaStream_t st) {
    WnPipePlan *pl = h->pipe;
    if (!pl) { h->err = "pipeline engine not planned"; return WN_ERR_STATE; }
    PipeParams p = pl->prm;
    const bool nu8 = pl->smem_bytes8 != 0 && B >= pl->nu8_min;
    if (nu8) p.NU = 8;
    pl->last_nu = p.NU;
    const size_t smem_bytes = nu8 ? pl->smem_bytes8 : pl->smem_bytes;
    p.NG = (B + p.NU - 1) / p.NU;
    p.NH = std::min((p.NG + p.P - 1) / p.P * p.P, nu8 ? pl->max_heads8 : pl->max_heads);   // whole clusters; surplus head CTAs have no group
    p.B = B; p.T = T;
    p.c_up = c_up; p.noise = noise; p.teacher = teacher; p.audio = audio; p.raw = raw; p.inputs_mode = inputs_mode;
    // mailboxes, 8-byte {value, tag} words, zeroed (tag 0 = empty), double buffered on t&1, one set per consumer stage
    // (layers 0..L-1, head = L):  x [2][NG][R][NU] | rs [2][NG][S][NU]
    const long long slots = 2ll * p.NG * p.NU;
    p.rs_off = slots * p.R; p.stage_sz = slots * (p.R + p.S);
    const size_t stages = (size_t)p.L + 1;
    const size_t words = stages * (size_t)p.stage_sz + (size_t)slots;
    if (pl->mail_words < words) {
        if (pl->d_mail) cudaFree(pl->d_mail);
        pl->d_mail = nullptr; pl->mail_words = 0;
        cudaError_t e = cudaMalloc((void **)&pl->d_mail, words * 8);
        if (e != cudaSuccess) { h->err = std::string("pipeline mailboxes: ") + cudaGetErrorString(e); return WN_ERR_NOMEM; }
        pl->mail_words = words;
    }
    p.mail = pl->d_mail;
    p.in = p.mail + stages * p.stage_sz;
    // private dilation rings of the layer CTAs: [CTA][NG][2d+1][R][NU] floats (never read before written: no memset)
    const size_t ring_floats = (size_t)pl->ring_slots * p.NG * p.R * p.NU;
    if (pl->ring_floats < ring_floats) {
        if (pl->d_rings) cudaFree(pl->d_rings);
        pl->d_rings = nullptr; pl->ring_floats = 0;
        cudaError_t e = cudaMalloc((void **)&pl->d_rings, ring_floats * sizeof(float));
        if (e != cudaSuccess) { h->err = std::string("pipeline rings: ") + cudaGetErrorString(e); return WN_ERR_NOMEM; }
        pl->ring_floats = ring_floats;
    }
    p.rings = pl->d_rings;
    WN_CUDA_TRY(h, cudaMemsetAsync(pl->d_mail, 0, words * 8, st));
    WN_CUDA_TRY(h, cudaMemsetAsync(pl->d_abort, 0, sizeof(int), st));
    const int grid = p.L * p.P + p.NH;
    p.stats = nullptr;
    if (getenv("WN_PIPE_STATS")) {
        if (!pl->d_stats) cudaMalloc((void **)&pl->d_stats, sizeof(long long) * 32 * 512);
        if (pl->d_stats) { cudaMemsetAsync(pl->d_stats, 0, sizeof(long long) * 32 * 512, st); p.stats = pl->d_stats; pl->stats_ctas = grid; }
    }
    // every CTA must be resident for the whole launch (they wait for each other): cooperative launch, clusters of P
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(PIPE_THREADS * TEAMS); cfg.dynamicSmemBytes = smem_bytes; cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = p.P; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeCooperative;
    at[1].val.cooperative = 1;
    // WN_PIPE_NO_COOP=1 (profiling aid): plain cluster launch; residency is then only guaranteed by the plan's occupancy check
    cfg.attrs = at; cfg.numAttrs = env_int("WN_PIPE_NO_COOP", 0) ? 1 : 2;
    void *args[] = {&p};
    WN_CUDA_TRY(h, cudaLaunchKernelExC(&cfg, pipe_kernel(p, p.stats != nullptr), args));
    h->launches++;
    WN_CUDA_TRY(h, cudaMemcpyAsync(pl->h_abort, pl->d_abort, sizeof(int), cudaMemcpyDeviceToHost, st));
    return WN_OK;
}


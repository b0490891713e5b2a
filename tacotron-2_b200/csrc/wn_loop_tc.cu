// wn_loop_tc.cu - the autoregressive loop of WaveNet.incremental (wavenet.py:805-911) on the sm_100a tensor cores.
//
// ONE cooperative launch runs all T steps.  The batch is cut into groups of NA utterances (a multiple of 4, at most NB = 16 or 32 =
// the N of every MMA); a "replica" is a chain of L + 1 CTAs, the head CTA and one per residual layer, each alone on its SM; up to 7
// replicas run side by side and up to 4 groups follow each other through one chain:
//
//   layer CTA l  (ResidualConv1DGLU.incremental_step, modules.py:468-521)
//     * the weights every step needs at once stay ON CHIP for the whole launch, as fp16 (hi, lo) pairs:
//         current-tap rows of the dilated conv (256 x 128) and the residual 1x1 (128 x 128) in TENSOR MEMORY as the A operand
//         of tcgen05.mma (lane = output channel, one 32-bit column = two consecutive k), the skip 1x1 in shared memory;
//     * per step: x_l(t) arrives -> fp16 (hi, lo) image in shared memory (the B operand, K-major) -> 48 MMAs
//       (M = 128, N = NB, K = 16; two gate halves x 8 k steps x three passes hi*hi + hi*lo + lo*hi) on top of the part of the
//       pre-activation that does not depend on x_l(t) -> tcgen05.ld -> tanh * sigmoid -> z image -> 24 MMAs (residual
//       1x1) -> x_{l+1}(t) = (out + b + x_l(t)) * sqrt(0.5) leaves for the next CTA; 24 more MMAs (skip 1x1) -> the skip
//       contribution leaves for the head CTA;
//     * between two steps (the CTA idles ~95 % of a step): the x-independent part of the NEXT step, W[tap0 | tap1 | cin]
//       * [x_l(t+1-2d) ; x_l(t+1-d) ; c(t+1)], is accumulated straight into the gate accumulator in tensor memory; its
//       weights (344 KB per layer) are streamed from L2 through four 16 KB shared-memory stages with cp.async.bulk.
//   head CTA  (wavenet.py:828-878): running skip sum in the reference's order, relu, 1x1, relu, 1x1 (both on the tensor
//     cores), the sampler (gaussian.py:39-52 / mixture.py:76-107) from the caller's noise buffer, audio / raw outputs,
//     and the sample fed back to layer 0.
//
// Hand-offs: consecutive CTAs of a replica share a thread-block cluster where they can (tc_launch_chunk) and pass x - and the head
// the fed-back sample - with st.async into the next CTA's shared memory, counted in bytes on its mbarrier; across cluster
// boundaries and for the skip outputs they are {value, tag} words in L2 (tag = step + 1, double buffered on t & 1): no fence, no
// flag separate from the data.  The dilation rings live in global memory (L2); every ring element is written and later read
// by the same thread.  Groups of <= 8 utterances use two MMA passes instead of three (img_put2 in wn_tc_common.cuh).
//
// Arithmetic: fp32 accumulation everywhere; every fp32 operand is split into two fp16 numbers (22 significant bits) and a
// product is three MMA passes, so results agree with the fp32 engines to ~1e-6 relative - the summation order differs, so not
// bit for bit (tests: tests/test_gpu_tc.py, criterion 1e-3 of BASELINE.json).  Operands must stay inside the fp16 range
// (|v| < 65504): weights and activations of a WaveNet are orders of magnitude below it.
#include <cuda_fp16.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "wn_common.cuh"
#include "wn_umma.cuh"

namespace {

constexpr int CH = 128;                  // residual = gate / 2 = skip channels of the block this engine is built for
constexpr int EPI_THREADS = 256;         // 8 epilogue warps: TMEM lane quarter = warp % 4, utterance half = warp / 4
constexpr int MMA_WARP = 8;              // warp 9 streams the weights
constexpr int TC_THREADS = 320;
constexpr int KB_TOTAL = 11;             // 32-wide k blocks of the streamed weights: tap0 4, tap1 4, cin 3 (80 + 16 zero)
constexpr int NCHUNK = 2 * KB_TOTAL;     // (k block, gate half), 16 KB each: A image [4][128][8] fp16 hi, then lo
constexpr int CHUNK_BYTES = 16384;
constexpr int NSTG = 4;                  // chunks in flight between L2 and the tensor cores: a pass over the stream takes 9.6 k cycles with 2
                                         // stages, 6.5 k with 4 (the L2 -> SM bandwidth bound is 5.6 k) - the MMA warp is busy that long per group
constexpr int CI_K = 96;                 // K of the conditioning image (cin <= 96, zero padded)

// layer blob (bytes)
constexpr int OFF_TMEM = 0;                       // 6 x [128 rows][64 words]: Wcur a hi, a lo, b hi, b lo, Wout hi, lo
constexpr int OFF_WS = 6 * 32768;                 // skip 1x1, A image [16][128][8] hi, lo
constexpr int OFF_STREAM = OFF_WS + 65536;        // NCHUNK chunks
constexpr int OFF_BIAS = OFF_STREAM + NCHUNK * CHUNK_BYTES;   // gate a, gate b, out, skip: 4 x 128 floats
constexpr int LAYER_BLOB_BYTES = OFF_BIAS + 4 * CH * 4;
// head blob: 4 x [128][64] words (W1 hi, lo, W2 hi, lo; W2 rows >= O zero), then b1[128], b2[128]
constexpr int HEAD_BLOB_BYTES = 4 * 32768 + 4 * CH * 4;   // W1, W2 as tensor-memory rows (hi, lo) | b1, b2 | rows 0, 1 of W2 in fp32

// tensor memory columns of a layer CTA
constexpr int COL_WCUR = 0, COL_WOUT = 256, COL_ACC = 384;   // accumulators: gate a | gate b | out | skip, NB columns each
// head CTA
constexpr int COL_W1 = 0, COL_W2 = 128, COL_HACC = 256;      // h1 | out

struct TcParams {
    int L, O, cin, NZ, input_type, legacy, residual_legacy;
    float log_scale_min, log_scale_min_gauss;
    int dil[WN_MAX_LAYERS];
    long long ring_off[WN_MAX_LAYERS];   // first ring slot of layer l inside a replica's ring block
    long long ring_slots;                // slots per replica
    const unsigned char *layer_blobs, *head_blob;
    const float *first_wb;               // input 1x1: w[128], b[128]
    unsigned long long *mail_x, *mail_skip, *mail_in;
    float *ring;
    int *abort_flag;              // [0] set by the first waiter that gives up; [1..2] = the wall-clock limit in ns (lo, hi words)
    long long *stats;
    const float *c_up, *noise, *teacher;
    int inputs_mode;
    float *audio, *raw;
    int B, T, NR;
    int no_w2c;                          // ablation (WN_TC_NOW2C): the single-Gaussian head's second 1x1 on the tensor cores too
    int no_fb;                           // ablation (WN_TC_NOFB): the fed-back sample through the L2 mailbox even inside a cluster
    int stride;                          // CTAs per replica in the grid: L + 1, or padded so that larger clusters divide it (idle CTAs at the end)
    int NG;                              // utterance groups that travel through a replica's chain one after the other (1 .. 4)
    int NA;                              // utterances a group really holds (multiple of 4, <= the MMA's N): the epilogues skip the rest
    int bend;                            // one past the last utterance of this launch
    int test_stall_step;                 // development / tests: the head stops feeding samples back at this step (-1: never)
    int b0;                              // first utterance of this launch in the caller's batch of B (batches above the launch limit run in chunks)
};

constexpr int NG_CAP = 4;                // utterance groups per chain (barrier sets)

template <int NB>
struct Geo {
    static constexpr int NH = NB / 2;            // utterances per epilogue thread
    static constexpr int NU = NH / 2;            // 16-byte units of two {value, tag} words
    static constexpr int LBO = NB * 16 + 16;     // k-chunk stride of the activation images (16 B of padding: conflict-free stores)
    static constexpr int PLANE = 16 * LBO;       // 128 channels
    static constexpr int IMG = 2 * PLANE;        // hi, lo
    static constexpr int CI_PLANE = (CI_K / 8) * LBO;
    // shared memory of a layer CTA: the weights, the images every utterance group rebuilds inside its own turn (z, taps,
    // conditioning), then per group what arrives or lives across turns (x image and the fp32 inbox of x)
    static constexpr int S_WS = 0, S_WST = 65536, S_ZI = S_WST + NSTG * CHUNK_BYTES, S_TAP = S_ZI + IMG, S_CI = S_TAP + 2 * IMG,
                         S_G0 = S_CI + 2 * CI_PLANE, G_XI = 0, G_XF = IMG, GRP = IMG + CH * NB * 4;
    // tensor memory: gate accumulators a | b per group from COL_ACC.  The out | skip accumulators of a group's turn are the SAME columns:
    // the gate epilogue has read them before the out / skip MMAs are issued (z barrier), and the next x-independent part overwrites
    // them only after the out / skip epilogues (p barrier)
    static constexpr int NG_TMEM = (512 - 384) / (2 * NB), NG_SMEM = (226 * 1024 - S_G0) / GRP;
    static constexpr int NG_MAX = NG_TMEM < NG_SMEM ? (NG_TMEM < NG_CAP ? NG_TMEM : NG_CAP) : (NG_SMEM < NG_CAP ? NG_SMEM : NG_CAP);
    static constexpr int S_END = S_G0 + NG_MAX * GRP;
    // head CTA
    static constexpr int S_HI = 0, S_H2 = IMG, S_PRM = 2 * IMG, S_HEND = S_PRM + NB * CH * 4;
    static constexpr int SMEM = S_END > S_HEND ? S_END : S_HEND;
    static constexpr int smem_for(int ng) { return (S_G0 + ng * GRP) > S_HEND ? (S_G0 + ng * GRP) : S_HEND; }
    static constexpr int MAILW = CH * NB;        // words of one mailbox buffer (and floats of one ring slot)
};

#include "wn_tc_common.cuh"

struct GBars {   // per utterance group: every barrier completes once per step of its group
    unsigned long long x, z, gate, out, skip, p, xf;
};
struct Bars {
    GBars g[NG_CAP];
    unsigned long long wfull[NSTG], wempty[NSTG];
    unsigned long long pdone;   // prologue only: the x-independent part of step 0 of a group has been read from the shared images
};

// =========================================================================================================================
// layer CTA
// =========================================================================================================================
template <int NB, int NAH, bool STATS>
__device__ void layer_epilogue(const TcParams &P, const int r, const int l, unsigned char *smem, const uint32_t tmem, Bars &bars) {
    using G = Geo<NB>;
    constexpr int NH = G::NH, NU = G::NU, LBO = G::LBO, PLANE = G::PLANE;
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, q = w & 3, hh = w >> 2;
    constexpr int nah = NAH, nu = NAH / 2, NA = 2 * NAH;   // utterances / two-utterance units this thread really handles (of NH / NU)
    constexpr bool WIDE = NB == 16 && NA <= 8;             // hi | lo side by side in the 16 columns (see img_put)
    const int j = 32 * q + lane, n0 = hh * nah;
    const uint32_t lane_addr = tmem + ((uint32_t)(32 * q) << 16);
    const int L = P.L, T = P.T, cin = P.cin, NG = P.NG;
    const unsigned char *blob = P.layer_blobs + (long long)l * LAYER_BLOB_BYTES;
    const float *bias = reinterpret_cast<const float *>(blob + OFF_BIAS);
    const float ba = __ldg(bias + j), bb = __ldg(bias + CH + j), bo = __ldg(bias + 2 * CH + j), bs = __ldg(bias + 3 * CH + j);
    const float wf = l == 0 ? __ldg(P.first_wb + j) : 0.0f, bf = l == 0 ? __ldg(P.first_wb + CH + j) : 0.0f;
    const float cres = P.residual_legacy ? 0.70710677f : 1.0f;   // np.sqrt(0.5) as float32 (modules.py:517-520)
    const int d = P.dil[l], qmask = 2 * d - 1;
    const int joff = (j >> 3) * LBO + n0 * 16 + (j & 7) * 2;            // element (j, n0) inside an image plane
    const long long unit0 = ((long long)(hh * NU) * CH + j) * 2;      // word offset of this thread's first unit inside a mailbox buffer
    // x arrives / leaves through distributed shared memory when the neighbouring stage shares this CTA's cluster, else
    // through the L2 mailboxes
    const uint32_t crank = cluster_ctarank(), csize = cluster_nctarank();
    const bool in_dsmem = l > 0 && crank > 0, out_dsmem = l + 1 < L && crank + 1 < csize;
    const bool fb_dsmem = l == 0 && crank > 0 && !P.no_fb;   // layer 0: the head is the previous CTA of the cluster
    constexpr int CN = (CI_K * NA + EPI_THREADS - 1) / EPI_THREADS;   // conditioning elements per thread (upper bound)
    long long st_c[6] = {0, 0, 0, 0, 0, 0};
    const bool do_stats = STATS && P.stats != nullptr && tid == 0;

    // ---- the state of utterance group g of this replica (set by setg) -----------------------------------------------------------
    int rg = 0;                                   // index of the group among all groups of the launch
    const unsigned long long *mx_in = nullptr, *min_in = nullptr;
    unsigned long long *mx_out = nullptr, *ms_out = nullptr;
    float *ring = nullptr, *XF = nullptr;
    unsigned char *XI = nullptr, *ZI = nullptr, *TAP = nullptr, *CI = nullptr;
    uint32_t bx = 0, bz = 0, bp = 0, bgate = 0, bout = 0, bskip = 0, bxf = 0, r_xf = 0, r_bxf = 0, acc = 0, acc_os = 0;
    auto setg = [&](int g) {
        rg = r * NG + g;
        mx_in = P.mail_x + (long long)(rg * L + l) * 2 * G::MAILW + unit0;
        mx_out = P.mail_x + (long long)(rg * L + l + 1) * 2 * G::MAILW + unit0;
        ms_out = P.mail_skip + (long long)(rg * L + l) * 2 * G::MAILW + unit0;
        min_in = P.mail_in + (long long)rg * 2 * NB + n0;
        ring = P.ring + ((long long)rg * P.ring_slots + P.ring_off[l]) * G::MAILW + ((long long)(hh * NU) * CH + j) * 2;   // [half][unit][channel][2]
        unsigned char *base = smem + G::S_G0 + g * G::GRP;
        XI = base + G::G_XI; ZI = smem + G::S_ZI; TAP = smem + G::S_TAP; CI = smem + G::S_CI;
        XF = reinterpret_cast<float *>(base + G::G_XF) + ((hh * NU) * CH + j) * 2;        // [half][unit][channel][2 utterances]
        GBars &gb = bars.g[g];
        bx = umma::smem_addr(&gb.x); bz = umma::smem_addr(&gb.z); bp = umma::smem_addr(&gb.p); bgate = umma::smem_addr(&gb.gate);
        bout = umma::smem_addr(&gb.out); bskip = umma::smem_addr(&gb.skip); bxf = umma::smem_addr(&gb.xf);
        r_xf = mapa_u32(umma::smem_addr(XF), out_dsmem ? crank + 1 : crank);
        r_bxf = mapa_u32(bxf, out_dsmem ? crank + 1 : crank);
        acc = lane_addr + COL_ACC + g * 2 * NB;   // gate a | gate b of the group
        acc_os = acc;                             // ... and, later in the turn, its out | skip accumulators
    };

    // operands of the x-independent part of step t: the two older taps from the ring and the conditioning column c(t)
    float tp0[NH], tp1[NH], cnd[CN];
    auto prep_load = [&](int t) {
        const int t0 = t - 2 * d, t1 = t - d;
#pragma unroll
        for (int u = 0; u < NU; ++u) {
            float2 a = make_float2(0.f, 0.f), b = a;
            if (u < nu && t0 >= 0) a = *reinterpret_cast<const float2 *>(ring + (long long)(t0 & qmask) * G::MAILW + (long long)u * CH * 2);
            if (u < nu && t1 >= 0) b = *reinterpret_cast<const float2 *>(ring + (long long)(t1 & qmask) * G::MAILW + (long long)u * CH * 2);
            tp0[2 * u] = a.x; tp0[2 * u + 1] = a.y;
            tp1[2 * u] = b.x; tp1[2 * u + 1] = b.y;
        }
#pragma unroll
        for (int i = 0; i < CN; ++i) {
            const int e = tid + i * EPI_THREADS;
            const int n = e / cin, k = e - n * cin;
            const long long b = (long long)P.b0 + rg * NA + n;
            cnd[i] = (n < NA && b < P.bend) ? __ldg(P.c_up + (b * T + t) * cin + k) : 0.0f;
        }
    };
    auto prep_store = [&]() {
#pragma unroll
        for (int i = 0; i < NH; i += 2) {
            if (i < nah) {
                img_put2<WIDE>(TAP + joff + i * 16, PLANE, tp0[i], tp0[i + 1]);
                img_put2<WIDE>(TAP + G::IMG + joff + i * 16, PLANE, tp1[i], tp1[i + 1]);
            }
        }
#pragma unroll
        for (int i = 0; i < CN; ++i) {
            const int e = tid + i * EPI_THREADS;
            const int n = e / cin, k = e - n * cin;
            if (n < NA) img_put<WIDE>(CI, G::CI_PLANE, LBO, k, n, cnd[i]);
        }
        umma::fence_proxy_async();
        umma::fence_before_sync();
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(bp);
    };
    for (int g = 0; g < NG; ++g) {   // the tap / conditioning images are shared by the groups: one group at a time
        setg(g);
        prep_load(0);
        prep_store();
        if (g + 1 < NG && mbar_wait_wd(umma::smem_addr(&bars.pdone), g & 1, P.abort_flag)) return;
    }

    const long long st_begin = clock64();
    // The groups of a replica follow each other through the chain: group 1 enters a stage when group 0 has left it, so every
    // role simply alternates between them; a group's own step latency is unchanged and the chain is busy twice as often.
    for (int t = 0; t < T; ++t) {
      for (int g = 0; g < NG; ++g) {
        setg(g);
        const int par = t & 1;
        const unsigned tag = (unsigned)t + 1u;
        const long long c0 = do_stats ? clock64() : 0;
        // ---- x_l(t) arrives (layer 0: the fed-back sample through the input 1x1, wavenet.py:826) ---------------------------
        float xin[NH];
        if (l == 0) {
            float s[NH];
            if (fb_dsmem) {   // the group's inbox holds the NA fed-back samples (4 bytes each, counted on the inbox barrier)
                if (tid == 0) umma::mbar_expect_tx(bxf, NA * 4);
                if (mbar_wait_cluster_wd(bxf, par, P.abort_flag)) return;
                const float *fb = reinterpret_cast<const float *>(smem + G::S_G0 + g * G::GRP + G::G_XF);
#pragma unroll
                for (int i = 0; i < NH; ++i) s[i] = i < nah ? fb[n0 + i] : 0.0f;
            } else if (poll_units<NU>(min_in + par * NB, 2, tag, s, nu, P.abort_flag)) return;
#pragma unroll
            for (int i = 0; i < NH; ++i) xin[i] = __fmaf_rn(s[i], wf, bf);
        } else if (in_dsmem) {
            if (tid == 0) umma::mbar_expect_tx(bxf, CH * NA * 4);
            if (mbar_wait_cluster_wd(bxf, par, P.abort_flag)) return;
#pragma unroll
            for (int u = 0; u < NU; ++u) {
                const float2 v = u < nu ? *reinterpret_cast<const float2 *>(XF + u * CH * 2) : make_float2(0.f, 0.f);
                xin[2 * u] = v.x; xin[2 * u + 1] = v.y;
            }
        } else {
            if (poll_units<NU>(mx_in + (long long)par * G::MAILW, (long long)CH * 2, tag, xin, nu, P.abort_flag)) return;
        }
        const long long c1 = do_stats ? clock64() : 0;
        unsigned char *xi = XI;
        {
            unsigned char *px = xi + joff;
#pragma unroll
            for (int i = 0; i < NH; i += 2)
                if (i < nah) img_put2<WIDE>(px + i * 16, PLANE, xin[i], xin[i + 1]);
        }
        umma::fence_proxy_async();
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(bx);
#pragma unroll
        for (int u = 0; u < NU; ++u)
            if (u < nu)
                *reinterpret_cast<float2 *>(ring + (long long)(t & qmask) * G::MAILW + (long long)u * CH * 2) = make_float2(xin[2 * u], xin[2 * u + 1]);
        // ---- gate (modules.py:494-510) ----------------------------------------------------------------------------------------
        if (mbar_wait_wd(bgate, par, P.abort_flag)) return;
        umma::fence_after_sync();
        const long long c2 = do_stats ? clock64() : 0;
        {
            float a[NH], b[NH];
            acc_load<NH, WIDE>(acc + n0, a);
            acc_load<NH, WIDE>(acc + NB + n0, b);
#pragma unroll
            for (int i = 0; i < NH; i += 2)
                if (i < nah)
                    img_put2<WIDE>(ZI + joff + i * 16, PLANE, gate_fast(__fadd_rn(a[i], ba), __fadd_rn(b[i], bb)),
                             gate_fast(__fadd_rn(a[i + 1], ba), __fadd_rn(b[i + 1], bb)));
        }
        umma::fence_proxy_async();
        umma::fence_before_sync();
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(bz);
        const long long c3 = do_stats ? clock64() : 0;
        // ---- residual 1x1 -> x_{l+1}(t) (modules.py:515-520) ----------------------------------------------------------------
        if (mbar_wait_wd(bout, par, P.abort_flag)) return;
        umma::fence_after_sync();
        const long long c4 = do_stats ? clock64() : 0;
        {
            float o[NH];
            acc_load<NH, WIDE>(acc_os + n0, o);
            if (l + 1 < L) {
#pragma unroll
                for (int u = 0; u < NU; ++u) {
                    if (u < nu) {
                        const float v0 = __fmul_rn(__fadd_rn(__fadd_rn(o[2 * u], bo), xin[2 * u]), cres);
                        const float v1 = __fmul_rn(__fadd_rn(__fadd_rn(o[2 * u + 1], bo), xin[2 * u + 1]), cres);
                        if (out_dsmem) st_async2(r_xf + u * CH * 8, v0, v1, r_bxf);
                        else st_ll2(mx_out + (long long)par * G::MAILW + (long long)u * CH * 2, v0, v1, tag);
                    }
                }
            }
        }
        // ---- skip 1x1 -> head (modules.py:512) ------------------------------------------------------------------------------
        if (mbar_wait_wd(bskip, par, P.abort_flag)) return;
        umma::fence_after_sync();
        {
            float s[NH];
            acc_load<NH, WIDE>(acc_os + NB + n0, s);
#pragma unroll
            for (int u = 0; u < NU; ++u)
                if (u < nu) st_ll2(ms_out + (long long)par * G::MAILW + (long long)u * CH * 2, __fadd_rn(s[2 * u], bs), __fadd_rn(s[2 * u + 1], bs), tag);
        }
        // ---- off the critical path: operands of the x-independent part of this group's next step ------------------------------
        if (t + 1 < T) {
            prep_load(t + 1);
            prep_store();
        }
        if (do_stats) {
            const long long c5 = clock64();
            st_c[0] += c1 - c0; st_c[1] += c2 - c1; st_c[2] += c3 - c2; st_c[3] += c4 - c3; st_c[4] += c5 - c4;
        }
      }
    }
    if (do_stats) {
        long long *o = P.stats + (long long)blockIdx.x * 16;
        for (int i = 0; i < 5; ++i) o[i] = st_c[i];
        o[14] = clock64() - st_begin; o[15] = (long long)T * NG;
    }
}

// LAST: the last layer has no residual output - the tensor-memory slot of W_out holds its skip weights (wn_tc_plan), the same
// MMAs go into the skip accumulator and the shared-memory-operand skip MMAs (~60 cycles each, on the step's critical path there) fall away
template <int NB, bool WIDE, bool LAST>
__device__ void layer_mma(const TcParams &P, unsigned char *smem, const uint32_t tmem, Bars &bars) {
    using G = Geo<NB>;
    constexpr int LBO = G::LBO, PLANE = G::PLANE;
    constexpr uint32_t idesc = umma::idesc_f16(CH, NB);
    const int T = P.T, NG = P.NG;
    const uint32_t s0 = umma::smem_addr(smem);
    const uint64_t ws_hi = umma::smem_desc(s0 + G::S_WS, CH * 16, 128), ws_lo = umma::smem_desc(s0 + G::S_WS + 32768, CH * 16, 128);
    const uint64_t st_hi0 = umma::smem_desc(s0 + G::S_WST, CH * 16, 128), st_lo0 = umma::smem_desc(s0 + G::S_WST + 8192, CH * 16, 128);   // stage 0
    const uint32_t wfull0 = umma::smem_addr(&bars.wfull[0]), wempty0 = umma::smem_addr(&bars.wempty[0]);
    constexpr uint64_t KS_B = (uint64_t)((2 * LBO) >> 4);        // descriptor step per k step (16 k) of an activation image
    constexpr uint64_t KS_A = (uint64_t)((2 * CH * 16) >> 4);    // ... of a 128-row weight image

    // the x-independent part of step tp of group g, accumulated from scratch into that group's gate accumulators
    long long pst[3] = {0, 0, 0};
    auto p_phase = [&](int tp, int g) -> bool {
        const long long q0 = P.stats ? clock64() : 0;
        if (mbar_wait_wd(umma::smem_addr(&bars.g[g].p), tp & 1, P.abort_flag)) return true;
        umma::fence_after_sync();
        const long long q1 = P.stats ? clock64() : 0;
        pst[0] += q1 - q0;
        const uint32_t D_A = tmem + COL_ACC + g * 2 * NB, D_B = D_A + NB;
        const uint64_t tap_hi[3] = {umma::smem_desc(s0 + G::S_TAP, LBO, 128), umma::smem_desc(s0 + G::S_TAP + G::IMG, LBO, 128),
                                    umma::smem_desc(s0 + G::S_CI, LBO, 128)};
        const uint64_t tap_lo[3] = {umma::smem_desc(s0 + G::S_TAP + PLANE, LBO, 128), umma::smem_desc(s0 + G::S_TAP + G::IMG + PLANE, LBO, 128),
                                    umma::smem_desc(s0 + G::S_CI + G::CI_PLANE, LBO, 128)};
        const uint32_t pass = (uint32_t)(tp * NG + g);   // passes over the weight stream so far
#pragma unroll
        for (int kb = 0; kb < KB_TOTAL; ++kb) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const uint32_t c = (uint32_t)NCHUNK * pass + 2 * kb + half, stage = c % NSTG, use = c / NSTG;   // chunk number -> stage, turn
                if (mbar_wait_wd(wfull0 + stage * 8, use & 1, P.abort_flag)) return true;
                umma::fence_after_sync();
                if (umma::elect_one()) {
                    const int src = kb / 4, kin = (kb % 4) * 2;   // image (tap0, tap1, cin) and first k step inside it
#pragma unroll
                    for (int kk = 0; kk < 2; ++kk) {
                        const uint64_t ah = st_hi0 + stage * (CHUNK_BYTES >> 4) + kk * KS_A, al = st_lo0 + stage * (CHUNK_BYTES >> 4) + kk * KS_A;
                        const uint64_t bh = tap_hi[src] + (kin + kk) * KS_B, bl = tap_lo[src] + (kin + kk) * KS_B;
                        const uint32_t dd = half ? D_B : D_A;
                        umma::mma_bf16(dd, ah, bh, idesc, (kb | kk) != 0);      // WIDE: plane 0 = hi | lo, plane 1 = hi
                        if (!WIDE) umma::mma_bf16(dd, ah, bl, idesc, 1);
                        umma::mma_bf16(dd, al, WIDE ? bl : bh, idesc, 1);
                    }
                    umma::mma_commit(wempty0 + stage * 8);
                }
                __syncwarp();
            }
        }
        if (P.stats) { pst[1] += clock64() - q1; pst[2] += 1; }
        return false;
    };
    for (int g = 0; g < NG; ++g) {
        if (p_phase(0, g)) return;
        if (umma::elect_one()) umma::mma_commit(umma::smem_addr(&bars.pdone));
        __syncwarp();
    }
    for (int t = 0; t < T; ++t) {
      for (int g = 0; g < NG; ++g) {
        const int par = t & 1;
        GBars &gb = bars.g[g];
        const uint32_t D_A = tmem + COL_ACC + g * 2 * NB, D_B = D_A + NB, D_O = D_A, D_S = D_B;   // out | skip reuse the gate columns
        const uint32_t sg = s0 + G::S_G0 + g * G::GRP;
        const uint64_t xi_hi = umma::smem_desc(sg + G::G_XI, LBO, 128), xi_lo = umma::smem_desc(sg + G::G_XI + PLANE, LBO, 128);
        const uint64_t zi_hi = umma::smem_desc(s0 + G::S_ZI, LBO, 128), zi_lo = umma::smem_desc(s0 + G::S_ZI + PLANE, LBO, 128);
        // current tap of the dilated conv (modules.py:294-297) on top of the x-independent part
        if (mbar_wait_wd(umma::smem_addr(&gb.x), par, P.abort_flag)) return;
        umma::fence_after_sync();
        if (umma::elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                const uint64_t bh = xi_hi + ks * KS_B, bl = xi_lo + ks * KS_B;
                umma::mma_bf16_ts(D_A, tmem + COL_WCUR + ks * 8, bh, idesc, 1);
                umma::mma_bf16_ts(D_B, tmem + COL_WCUR + 128 + ks * 8, bh, idesc, 1);
                if (!WIDE) {
                    umma::mma_bf16_ts(D_A, tmem + COL_WCUR + ks * 8, bl, idesc, 1);
                    umma::mma_bf16_ts(D_B, tmem + COL_WCUR + 128 + ks * 8, bl, idesc, 1);
                }
                umma::mma_bf16_ts(D_A, tmem + COL_WCUR + 64 + ks * 8, WIDE ? bl : bh, idesc, 1);
                umma::mma_bf16_ts(D_B, tmem + COL_WCUR + 192 + ks * 8, WIDE ? bl : bh, idesc, 1);
            }
            umma::mma_commit(umma::smem_addr(&gb.gate));
        }
        __syncwarp();
        // residual 1x1 on z (modules.py:515)
        if (mbar_wait_wd(umma::smem_addr(&gb.z), par, P.abort_flag)) return;
        umma::fence_after_sync();
        if (umma::elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                umma::mma_bf16_ts(LAST ? D_S : D_O, tmem + COL_WOUT + ks * 8, zi_hi + ks * KS_B, idesc, ks != 0);
                if (!WIDE) umma::mma_bf16_ts(LAST ? D_S : D_O, tmem + COL_WOUT + ks * 8, zi_lo + ks * KS_B, idesc, 1);
                umma::mma_bf16_ts(LAST ? D_S : D_O, tmem + COL_WOUT + 64 + ks * 8, (WIDE ? zi_lo : zi_hi) + ks * KS_B, idesc, 1);
            }
            umma::mma_commit(umma::smem_addr(&gb.out));
            if (LAST) umma::mma_commit(umma::smem_addr(&gb.skip));
        }
        __syncwarp();
        // skip 1x1 on z (modules.py:512), weights in shared memory
        if (!LAST && umma::elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                umma::mma_bf16(D_S, ws_hi + ks * KS_A, zi_hi + ks * KS_B, idesc, ks != 0);
                if (!WIDE) umma::mma_bf16(D_S, ws_hi + ks * KS_A, zi_lo + ks * KS_B, idesc, 1);
                umma::mma_bf16(D_S, ws_lo + ks * KS_A, (WIDE ? zi_lo : zi_hi) + ks * KS_B, idesc, 1);
            }
            umma::mma_commit(umma::smem_addr(&gb.skip));
        }
        __syncwarp();
        if (t + 1 < T && p_phase(t + 1, g)) return;
      }
    }
    if (P.stats && (threadIdx.x & 31) == 0) {
        long long *o = P.stats + (long long)blockIdx.x * 16;
        o[8] = pst[0]; o[9] = pst[1]; o[10] = pst[2];
    }
}

// streams the [tap0 | tap1 | cin] weights of the layer, NCHUNK chunks per step and utterance group, through the NSTG stages
__device__ void layer_loader(const TcParams &P, const int l, unsigned char *smem, Bars &bars, int wst_off) {
    if ((threadIdx.x & 31) != 0) return;
    const unsigned char *stream = P.layer_blobs + (long long)l * LAYER_BLOB_BYTES + OFF_STREAM;
    const uint32_t dst0 = umma::smem_addr(smem + wst_off);
    const int passes = P.T * P.NG;
    for (int pass = 0; pass < passes; ++pass)
        for (int kb = 0; kb < KB_TOTAL; ++kb)
            for (int half = 0; half < 2; ++half) {
                const uint32_t c = (uint32_t)(NCHUNK * pass + 2 * kb + half), stage = c % NSTG, use = c / NSTG;
                if (mbar_wait_wd(umma::smem_addr(&bars.wempty[stage]), (use & 1) ^ 1, P.abort_flag)) return;
                const uint32_t full = umma::smem_addr(&bars.wfull[stage]);
                umma::mbar_expect_tx(full, CHUNK_BYTES);
                umma::bulk_g2s(dst0 + stage * CHUNK_BYTES, stream + (size_t)(kb * 2 + half) * CHUNK_BYTES, CHUNK_BYTES, full);
            }
}

// =========================================================================================================================
// head CTA
// =========================================================================================================================
template <int NB, int NAH, bool STATS>
__device__ void head_epilogue(const TcParams &P, const int r, unsigned char *smem, const uint32_t tmem, Bars &bars) {
    using G = Geo<NB>;
    constexpr int NH = G::NH, NU = G::NU, LBO = G::LBO, PLANE = G::PLANE;
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, q = w & 3, hh = w >> 2;
    constexpr int nah = NAH, nu = NAH / 2, NA = 2 * NAH;
    constexpr bool WIDE = NB == 16 && NA <= 8;
    const int j = 32 * q + lane, n0 = hh * nah;
    const uint32_t lane_addr = tmem + ((uint32_t)(32 * q) << 16);
    const int L = P.L, T = P.T, O = P.O, NG = P.NG;
    const float *hb = reinterpret_cast<const float *>(P.head_blob + 4 * 32768);
    const float b1 = __ldg(hb + j), b2 = __ldg(hb + CH + j);
    // single-Gaussian head (O = 2): the second 1x1 has two output rows - fp32 dot products on the CUDA cores (this thread's channel,
    // warp-shuffle tree, the four warps of a utterance half through shared memory) instead of another image -> MMA -> epilogue round trip
    const bool w2_cuda = O == 2 && !P.no_w2c;
    const float w20 = __ldg(hb + 2 * CH + j), w21 = __ldg(hb + 3 * CH + j), b20 = __ldg(hb + CH), b21 = __ldg(hb + CH + 1);
    float *part = reinterpret_cast<float *>(smem + G::S_H2);   // [4 lane quarters][2 rows][NB]   (the image of the MMA path is not used then)
    const uint32_t bpart = umma::smem_addr(&bars.g[1].x);      // 8 warp arrivals per turn
    const float cskip = 0.70710677f;
    const int joff = (j >> 3) * LBO + n0 * 16 + (j & 7) * 2;
    const long long unit0 = ((long long)(hh * NU) * CH + j) * 2;

    unsigned char *HI = smem + G::S_HI, *H2 = smem + G::S_H2;
    float *prm = reinterpret_cast<float *>(smem + G::S_PRM);   // [NB][CH]
    const uint32_t bh = umma::smem_addr(&bars.g[0].x), bh2 = umma::smem_addr(&bars.g[0].z), bh1 = umma::smem_addr(&bars.g[0].gate),
                   bo = umma::smem_addr(&bars.g[0].out);
    // layer 0 is the next CTA of the head's cluster (if there are clusters): the samples go into its per-group inbox with st.async
    const bool fb_dsmem = cluster_nctarank() > 1 && !P.no_fb;
    auto r_fb = [&](int g) { return mapa_u32(umma::smem_addr(smem + G::S_G0 + g * G::GRP + G::G_XF), 1); };
    auto r_bfb = [&](int g) { return mapa_u32(umma::smem_addr(&bars.g[g].xf), 1); };
    // initial input (wavenet.py:433-445): raw 0.0, mulaw mulaw(0) = 0.0; WaveNet.step semantics: the first input is x[0]
    for (int g = 0; g < NG; ++g) {
        const long long b = (long long)P.b0 + (long long)(r * NG + g) * NA + tid;
        const float v0 = (tid < NA && b < P.bend && P.inputs_mode) ? P.teacher[b * T] : 0.0f;
        if (fb_dsmem) { if (tid < NA) st_async_b32(r_fb(g) + tid * 4, __float_as_uint(v0), r_bfb(g)); }
        else if (tid < NB) st_ll1(P.mail_in + (long long)(r * NG + g) * 2 * NB + tid, v0, 1u);
    }
    long long st_c[4] = {0, 0, 0, 0};
    const bool do_stats = STATS && P.stats != nullptr && tid == 0;
    const long long st_begin = clock64();
    for (int t = 0; t < T; ++t) {
      for (int g = 0; g < NG; ++g) {   // the groups of the replica arrive one after the other
        const int rg = r * NG + g;
        const unsigned long long *ms_in = P.mail_skip + (long long)rg * L * 2 * G::MAILW + unit0;
        unsigned long long *min_out = P.mail_in + (long long)rg * 2 * NB;
        // the O <= 32 output rows live in lane quarter 0: warp 0 (first half of the utterances) and warp 4 (second half) hold
        // every parameter of their utterances, so lane i of those warps samples utterance n0 + i without a CTA-wide barrier
        const int ns = n0 + lane;                                              // utterance (inside the group) of a sampler lane
        const long long bu = (long long)P.b0 + (long long)rg * NA + ns;
        const bool slane = q == 0 && lane < nah, sampler = slane && bu < P.bend;
        const int par = t & 1;                      // mailbox buffer of the step
        const int bpar = (t * NG + g) & 1;          // phase of this CTA's barriers (one completion per group and step)
        const unsigned tag = (unsigned)t + 1u;
        const long long c0 = do_stats ? clock64() : 0;
        float nz0 = 0.0f, tch = 0.0f;
        if (sampler) {
            if (O == 2 && P.noise) nz0 = __ldg(P.noise + ((long long)t * P.B + bu) * P.NZ);
            if (P.teacher) tch = P.inputs_mode ? (t + 1 < T ? __ldg(P.teacher + bu * T + t + 1) : 0.0f) : __ldg(P.teacher + bu * T + t);
        }
        // running skip sum in the reference's order (wavenet.py:833-836)
        // One L2 round trip per layer would make the head the slowest stage for the groups that arrive behind the first one of a
        // step (their skip outputs are there already; 20 round trips = 14.7 k cycles per group: a chain could not carry more than
        // three).  Polling many mailboxes at once while the chain is still working, on the other hand, slows the chain (+0.75 us
        // per step).  So: layer by layer while the words have to be waited for; as soon as a mailbox is complete at the first look
        // the head is behind, and it looks at LB mailboxes at a time (all loads in flight together) until it has caught up.
        constexpr int LB = nu <= 1 ? 5 : nu <= 2 ? 4 : 2;
        float S[NH];
        auto add_layer = [&](int l, const float *s) {
#pragma unroll
            for (int i = 0; i < NH; ++i) {
                if (l == 0) S[i] = s[i];
                else S[i] = P.legacy ? __fmul_rn(__fadd_rn(S[i], s[i]), cskip) : __fadd_rn(S[i], s[i]);
            }
        };
        {
            bool behind = false;
            int l = 0;
            while (l < L) {
                const unsigned long long *box = ms_in + ((long long)l * 2 + par) * G::MAILW;
                if (!behind || nu > 4) {
                    float s[NH];
                    bool first;
                    if (poll_units_first<NU>(box, (long long)CH * 2, tag, s, nu, P.abort_flag, first)) return;
                    add_layer(l, s);
                    ++l;
                    behind = first && NG >= 3;   // one or two groups: the head keeps up anyway (and 0.2 us per step is lost otherwise)
                } else {
                    float s[LB][NH];
                    const int nl = L - l < LB ? L - l : LB;
                    const int m = look_boxes<NU, LB>(box, 2LL * G::MAILW, (long long)CH * 2, tag, s, nu, nl);
#pragma unroll
                    for (int b = 0; b < LB; ++b)
                        if (b < m) add_layer(l + b, s[b]);
                    l += m;
                    behind = m == nl;
                }
            }
        }
        const long long c1 = do_stats ? clock64() : 0;
#pragma unroll
        for (int i = 0; i < NH; i += 2)   // relu (wavenet.py:840)
            if (i < nah) img_put2<WIDE>(HI + joff + i * 16, PLANE, S[i] > 0.0f ? S[i] : 0.0f, S[i + 1] > 0.0f ? S[i + 1] : 0.0f);
        umma::fence_proxy_async();
        umma::fence_before_sync();
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(bh);
        if (mbar_wait_wd(bh1, bpar, P.abort_flag)) return;
        umma::fence_after_sync();
        float out2[2] = {0.0f, 0.0f};   // the two parameters of this sampler lane's utterance (w2_cuda)
        if (w2_cuda) {
            float h[NH];
            acc_load<NH, WIDE>(lane_addr + COL_HACC + n0, h);
            float pa[NH], pb[NH];
#pragma unroll
            for (int i = 0; i < NH; ++i) {
                const float y = __fadd_rn(h[i], b1), yr = y > 0.0f ? y : 0.0f;   // relu (wavenet.py:842)
                pa[i] = i < nah ? __fmul_rn(w20, yr) : 0.0f;
                pb[i] = i < nah ? __fmul_rn(w21, yr) : 0.0f;
            }
#pragma unroll
            for (int i = 0; i < NH; ++i) {
                if (i < nah) {
#pragma unroll
                    for (int m = 16; m >= 1; m >>= 1) {
                        pa[i] = __fadd_rn(pa[i], __shfl_xor_sync(0xffffffffu, pa[i], m));
                        pb[i] = __fadd_rn(pb[i], __shfl_xor_sync(0xffffffffu, pb[i], m));
                    }
                    if (lane == 0) {
                        part[(q * 2 + 0) * NB + n0 + i] = pa[i];
                        part[(q * 2 + 1) * NB + n0 + i] = pb[i];
                    }
                }
            }
            umma::fence_before_sync();
            __syncwarp();
            if (lane == 0) umma::mbar_arrive(bpart);
            if (q == 0) {   // the sampler warps
                if (mbar_wait_wd(bpart, bpar, P.abort_flag)) return;
                if (slane) {
                    out2[0] = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(part[0 * NB + ns], part[2 * NB + ns]), part[4 * NB + ns]), part[6 * NB + ns]), b20);
                    out2[1] = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(part[1 * NB + ns], part[3 * NB + ns]), part[5 * NB + ns]), part[7 * NB + ns]), b21);
                    if (P.raw && bu < P.bend) {   // wavenet.py:847
                        P.raw[(bu * T + t) * 2] = out2[0];
                        P.raw[(bu * T + t) * 2 + 1] = out2[1];
                    }
                }
            }
        } else {
        {
            float h[NH];
            acc_load<NH, WIDE>(lane_addr + COL_HACC + n0, h);
#pragma unroll
            for (int i = 0; i < NH; i += 2) {
                const float y0 = __fadd_rn(h[i], b1), y1 = __fadd_rn(h[i + 1], b1);
                if (i < nah) img_put2<WIDE>(H2 + joff + i * 16, PLANE, y0 > 0.0f ? y0 : 0.0f, y1 > 0.0f ? y1 : 0.0f);
            }
        }
        umma::fence_proxy_async();
        umma::fence_before_sync();
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(bh2);
        if (mbar_wait_wd(bo, bpar, P.abort_flag)) return;
        umma::fence_after_sync();
        {
            float o[NH];
            acc_load<NH, WIDE>(lane_addr + COL_HACC + NB + n0, o);
            if (j < O) {
#pragma unroll
                for (int i = 0; i < NH; ++i) {
                    if (i < nah) {
                        const float y = __fadd_rn(o[i], b2);
                        prm[(n0 + i) * CH + j] = y;
                        const long long b = (long long)P.b0 + (long long)rg * NA + n0 + i;
                        if (P.raw && b < P.bend) P.raw[(b * T + t) * O + j] = y;   // wavenet.py:847
                    }
                }
            }
        }
        }
        const long long c2 = do_stats ? clock64() : 0;
        umma::fence_before_sync();
        __syncwarp();
        if (t == P.test_stall_step) return;   // test hook: the chain starves and every waiter must time out (WN_ERR_WATCHDOG)
        // sampling + feedback (wavenet.py:850-878)
        if (slane) {
            float nxt = 0.0f;
            if (sampler && P.inputs_mode) {
                nxt = tch;
            } else if (sampler) {
                const float smp = (O == 2) ? tc_sample_gaussian(P.log_scale_min_gauss, out2, nz0)
                                           : tc_sample_mol(P.log_scale_min, O, prm + ns * CH, P.noise + ((long long)t * P.B + bu) * P.NZ);
                P.audio[bu * T + t] = P.input_type == WN_INPUT_MULAW ? wn_inv_mulawf(smp) : smp;
                nxt = P.teacher ? tch : smp;
            }
            if (t + 1 < T) {
                if (fb_dsmem) st_async_b32(r_fb(g) + ns * 4, __float_as_uint(nxt), r_bfb(g));
                else st_ll1(min_out + ((t + 1) & 1) * NB + ns, nxt, tag + 1u);
            }
        }
        __syncwarp();   // the parameters are overwritten by the same warp in its next turn
        if (do_stats) {
            const long long c3 = clock64();
            st_c[0] += c1 - c0; st_c[1] += c2 - c1; st_c[2] += c3 - c2;
        }
      }
    }
    if (do_stats) {
        long long *o = P.stats + (long long)blockIdx.x * 16;
        for (int i = 0; i < 3; ++i) o[i] = st_c[i];
        o[14] = clock64() - st_begin; o[15] = (long long)T * NG;
    }
}

template <int NB, bool WIDE>
__device__ void head_mma(const TcParams &P, unsigned char *smem, const uint32_t tmem, Bars &bars) {
    using G = Geo<NB>;
    constexpr int LBO = G::LBO, PLANE = G::PLANE;
    constexpr uint32_t idesc = umma::idesc_f16(CH, NB);
    const uint32_t s0 = umma::smem_addr(smem);
    const uint64_t hi_hi = umma::smem_desc(s0 + G::S_HI, LBO, 128), hi_lo = umma::smem_desc(s0 + G::S_HI + PLANE, LBO, 128);
    const uint64_t h2_hi = umma::smem_desc(s0 + G::S_H2, LBO, 128), h2_lo = umma::smem_desc(s0 + G::S_H2 + PLANE, LBO, 128);
    constexpr uint64_t KS_B = (uint64_t)((2 * LBO) >> 4);
    const uint32_t D_H = tmem + COL_HACC, D_O = tmem + COL_HACC + NB;
    const uint32_t bh1 = umma::smem_addr(&bars.g[0].gate), bo = umma::smem_addr(&bars.g[0].out);
    const int steps = P.T * P.NG;   // one pass per step and utterance group
    for (int k = 0; k < steps; ++k) {
        const int par = k & 1;
        if (mbar_wait_wd(umma::smem_addr(&bars.g[0].x), par, P.abort_flag)) return;   // final_convolution_1 (wavenet.py:841)
        umma::fence_after_sync();
        if (umma::elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                umma::mma_bf16_ts(D_H, tmem + COL_W1 + ks * 8, hi_hi + ks * KS_B, idesc, ks != 0);
                if (!WIDE) umma::mma_bf16_ts(D_H, tmem + COL_W1 + ks * 8, hi_lo + ks * KS_B, idesc, 1);
                umma::mma_bf16_ts(D_H, tmem + COL_W1 + 64 + ks * 8, (WIDE ? hi_lo : hi_hi) + ks * KS_B, idesc, 1);
            }
            umma::mma_commit(bh1);
        }
        __syncwarp();
        if (P.O == 2 && !P.no_w2c) continue;   // single Gaussian: the second 1x1 runs on the CUDA cores (head_epilogue)
        if (mbar_wait_wd(umma::smem_addr(&bars.g[0].z), par, P.abort_flag)) return;   // final_convolution_2 (wavenet.py:843), rows >= O zero
        umma::fence_after_sync();
        if (umma::elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                umma::mma_bf16_ts(D_O, tmem + COL_W2 + ks * 8, h2_hi + ks * KS_B, idesc, ks != 0);
                if (!WIDE) umma::mma_bf16_ts(D_O, tmem + COL_W2 + ks * 8, h2_lo + ks * KS_B, idesc, 1);
                umma::mma_bf16_ts(D_O, tmem + COL_W2 + 64 + ks * 8, (WIDE ? h2_lo : h2_hi) + ks * KS_B, idesc, 1);
            }
            umma::mma_commit(bo);
        }
        __syncwarp();
    }
}

// rows [128][64] words of `nmat` matrices -> tensor memory columns [col0, col0 + 64 * nmat): warps 0..3, thread = row
__device__ void load_tmem_weights(const unsigned char *src, int nmat, uint32_t tmem, int col0) {
    const int tid = threadIdx.x;
    if (tid >= 128) return;
    const uint32_t lane_addr = tmem + ((uint32_t)(tid & ~31) << 16);
    for (int m = 0; m < nmat; ++m) {
        const uint4 *row = reinterpret_cast<const uint4 *>(src + ((size_t)m * CH + tid) * 256);
        for (int c0 = 0; c0 < 64; c0 += 16) {
            uint32_t r[16];
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                const uint4 x = __ldg(row + c0 / 4 + v);
                r[4 * v] = x.x; r[4 * v + 1] = x.y; r[4 * v + 2] = x.z; r[4 * v + 3] = x.w;
            }
            umma::tmem_st16(lane_addr + col0 + m * 64 + c0, r);
        }
    }
    umma::tmem_st_wait();
}

template <int NB, int NAH, bool STATS>
__global__ void __launch_bounds__(TC_THREADS, 1) wn_loop_tc_kernel(const TcParams P) {
    using G = Geo<NB>;
    extern __shared__ __align__(128) unsigned char tc_smem[];
    __shared__ Bars bars;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    // slots of a replica in the grid: 0 = the head (so that it shares layer 0's cluster: the fed-back sample travels through
    // distributed shared memory), 1 .. L = the layers, beyond = padding
    const int r = blockIdx.x / P.stride, slot = blockIdx.x - r * P.stride;
    const bool head = slot == 0;
    const int l = head ? P.L : slot - 1;
    if (slot > P.L) {   // padding CTA (the replica's slots are rounded up to a multiple of the cluster size): only the cluster's barriers
        cluster_sync_all();
        cluster_sync_all();
        return;
    }
    if (tid == 0) {
        for (int g = 0; g < NG_CAP; ++g) {
            GBars &gb = bars.g[g];
            umma::mbar_init(umma::smem_addr(&gb.x), EPI_THREADS / 32);
            umma::mbar_init(umma::smem_addr(&gb.z), EPI_THREADS / 32);
            umma::mbar_init(umma::smem_addr(&gb.gate), 1);
            umma::mbar_init(umma::smem_addr(&gb.out), 1);
            umma::mbar_init(umma::smem_addr(&gb.skip), head ? EPI_THREADS / 32 : 1);   // head: "parameters written" barrier
            umma::mbar_init(umma::smem_addr(&gb.p), EPI_THREADS / 32);
            umma::mbar_init(umma::smem_addr(&gb.xf), 1);
        }
        for (int i = 0; i < NSTG; ++i) { umma::mbar_init(umma::smem_addr(&bars.wfull[i]), 1); umma::mbar_init(umma::smem_addr(&bars.wempty[i]), 1); }
        umma::mbar_init(umma::smem_addr(&bars.pdone), 1);
        umma::mbar_fence_init();
    }
    if (warp == 0) umma::tmem_alloc(umma::smem_addr(&tmem_base_s), 512);
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = tmem_base_s;
    // resident weights; zero the activation images (the padding rows / columns of the images are never written again)
    if (head) {
        load_tmem_weights(P.head_blob, 4, tmem, COL_W1);
        for (int i = tid; i < G::S_HEND / 16; i += TC_THREADS) reinterpret_cast<uint4 *>(tc_smem)[i] = make_uint4(0, 0, 0, 0);
    } else {
        const unsigned char *blob = P.layer_blobs + (long long)l * LAYER_BLOB_BYTES;
        load_tmem_weights(blob + OFF_TMEM, 6, tmem, COL_WCUR);
        const uint4 *ws = reinterpret_cast<const uint4 *>(blob + OFF_WS);
        for (int i = tid; i < 65536 / 16; i += TC_THREADS) reinterpret_cast<uint4 *>(tc_smem + G::S_WS)[i] = __ldg(ws + i);
        for (int i = tid; i < (G::S_G0 - G::S_ZI + P.NG * G::GRP) / 16; i += TC_THREADS) reinterpret_cast<uint4 *>(tc_smem + G::S_ZI)[i] = make_uint4(0, 0, 0, 0);
    }
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    cluster_sync_all();   // the peers' mbarriers exist before anything is sent to them
    if (head) {
        if (warp < 8) head_epilogue<NB, NAH, STATS>(P, r, tc_smem, tmem, bars);
        else if (warp == MMA_WARP) head_mma<NB, (NB == 16 && NAH <= 4)>(P, tc_smem, tmem, bars);
    } else {
        if (warp < 8) layer_epilogue<NB, NAH, STATS>(P, r, l, tc_smem, tmem, bars);
        else if (warp == MMA_WARP) {
            if (l + 1 == P.L) layer_mma<NB, (NB == 16 && NAH <= 4), true>(P, tc_smem, tmem, bars);
            else layer_mma<NB, (NB == 16 && NAH <= 4), false>(P, tc_smem, tmem, bars);
        }
        else layer_loader(P, l, tc_smem, bars, G::S_WST);
    }
    // every role returns when the sequence is done (or the watchdog fired): release tensor memory
    umma::fence_before_sync();
    __syncthreads();
    cluster_sync_all();   // nobody leaves while a peer may still write into its shared memory
    if (warp == 0) umma::tmem_dealloc(tmem, 512);
}

}  // namespace

// =========================================================================================================================
// host side
// =========================================================================================================================
struct WnTcPlan {
    TcParams prm;
    unsigned char *d_layer_blobs = nullptr, *d_head_blob = nullptr;
    float *d_first = nullptr;
    unsigned long long *d_mail = nullptr;
    size_t mail_words = 0;
    float *d_ring = nullptr;
    size_t ring_floats = 0;
    int *d_abort = nullptr, *h_abort = nullptr;
    int h_abort_init[4] = {0, 0, 0, 0};
    long long *d_stats = nullptr;
    int stats_ctas = 0;
    int max_replicas = 0;
    int last_nb = 0, last_nr = 0, last_ng = 1, last_cluster = 1, last_stride = 1;
};

static void split_host(float v, uint16_t &hi, uint16_t &lo) {   // fp16 (hi, lo): 22 significant bits of v
    const __half h = __float2half_rn(v);
    const __half l = __float2half_rn(v - __half2float(h));
    memcpy(&hi, &h, 2);
    memcpy(&lo, &l, 2);
}

static bool tc1_supported(const wn_handle *h) {
    const wn_config &c = h->cfg;
    return c.residual_channels == CH && c.gate_channels == 2 * CH && c.skip_out_channels == CH && c.kernel_size == 3 &&
           c.input_type != WN_INPUT_MULAW_QUANTIZE && c.gin_channels <= 0 && c.cin_channels > 0 && c.cin_channels <= CI_K &&
           c.out_channels <= 32 && c.layers >= 1 && c.layers / c.stacks <= 16;   // <= 32 output rows: one TMEM lane quarter (head sampler)
}

bool wn_tc_supported(const wn_handle *h) { return tc1_supported(h) || wn_tc2_supported(h); }

void wn_tc_free(wn_handle *h) {
    wn_tc2_free(h);
    WnTcPlan *pl = h->tc;
    if (!pl) return;
    if (pl->d_layer_blobs) cudaFree(pl->d_layer_blobs);
    if (pl->d_head_blob) cudaFree(pl->d_head_blob);
    if (pl->d_first) cudaFree(pl->d_first);
    if (pl->d_mail) cudaFree(pl->d_mail);
    if (pl->d_ring) cudaFree(pl->d_ring);
    if (pl->d_abort) cudaFree(pl->d_abort);
    if (pl->h_abort) cudaFreeHost(pl->h_abort);
    if (pl->d_stats) cudaFree(pl->d_stats);
    delete pl;
    h->tc = nullptr;
}

// kernel for N = nb and na = 2 * NAH utterances per replica (na a multiple of 4, nb - 12 <= na <= nb)
static const void *tc_kernel(int nb, int na, bool stats) {
#define WN_TC_K(NB_, NAH_) (stats ? (const void *)wn_loop_tc_kernel<NB_, NAH_, true> : (const void *)wn_loop_tc_kernel<NB_, NAH_, false>)
    if (nb == 16) {
        switch (na) { case 4: return WN_TC_K(16, 2); case 8: return WN_TC_K(16, 4); case 12: return WN_TC_K(16, 6); default: return WN_TC_K(16, 8); }
    }
    switch (na) { case 20: return WN_TC_K(32, 10); case 24: return WN_TC_K(32, 12); case 28: return WN_TC_K(32, 14); default: return WN_TC_K(32, 16); }
#undef WN_TC_K
}

// Packs the weights (TF layouts) into the per-CTA blobs described at the top of this file.
int wn_tc_plan(wn_handle *h) {
    if (!tc1_supported(h) && wn_tc2_supported(h)) return wn_tc2_plan(h);   // the 256-channel block: wn_loop_tc2.cu
    if (!tc1_supported(h)) { h->err = "tensor-core loop engine: needs R = G/2 = S = 128, kernel_size 3, scalar input, local conditioning (<= 96 channels), no global conditioning"; return WN_ERR_INVALID; }
    const wn_config &c = h->cfg;
    const int L = c.layers, cin = c.cin_channels, O = c.out_channels, G2 = 2 * CH;
    auto W = [&](const std::string &n) -> const float * {
        auto it = h->host_w.find(n);
        return it == h->host_w.end() ? nullptr : it->second.data.data();
    };
    std::vector<unsigned char> blobs((size_t)L * LAYER_BLOB_BYTES, 0);
    // rows [128][64 words] of a tensor-memory matrix: word c = bf16(k = 2c) | bf16(k = 2c + 1) << 16
    auto put_tmem = [](unsigned char *dst_hi, unsigned char *dst_lo, int row, int k, float v) {
        uint16_t hi, lo;
        split_host(v, hi, lo);
        reinterpret_cast<uint16_t *>(dst_hi)[(size_t)row * 128 + k] = hi;
        reinterpret_cast<uint16_t *>(dst_lo)[(size_t)row * 128 + k] = lo;
    };
    // A image [K/8][128 rows][8] of a shared-memory operand
    auto put_img = [](unsigned char *dst_hi, unsigned char *dst_lo, int row, int k, float v) {
        uint16_t hi, lo;
        split_host(v, hi, lo);
        const size_t e = ((size_t)(k / 8) * CH + row) * 8 + k % 8;
        reinterpret_cast<uint16_t *>(dst_hi)[e] = hi;
        reinterpret_cast<uint16_t *>(dst_lo)[e] = lo;
    };
    for (int l = 0; l < L; ++l) {
        const std::string p = "ResidualConv1DGLU_" + std::to_string(l) + "/residual_block_";
        const float *cw = W(p + "causal_conv/kernel"), *iw = W(p + "cin_conv/kernel"), *ow = W(p + "out_conv/kernel"),
                    *sw = W(p + "skip_conv/kernel");
        const float *cb = W(p + "causal_conv/bias"), *ib = W(p + "cin_conv/bias"), *ob = W(p + "out_conv/bias"),
                    *sb = W(p + "skip_conv/bias");
        unsigned char *b = blobs.data() + (size_t)l * LAYER_BLOB_BYTES;
        for (int j = 0; j < CH; ++j)
            for (int k = 0; k < CH; ++k) {
                // current tap = tap 2 of the [3][R][G] kernel (modules.py:238-251: row = tap * R + ci, tap 0 oldest)
                put_tmem(b + OFF_TMEM, b + OFF_TMEM + 32768, j, k, cw[((size_t)2 * CH + k) * G2 + j]);
                put_tmem(b + OFF_TMEM + 2 * 32768, b + OFF_TMEM + 3 * 32768, j, k, cw[((size_t)2 * CH + k) * G2 + CH + j]);
                // the last layer has no residual output (wavenet.py:833: only its skip is used): the slot holds its skip weights instead
                put_tmem(b + OFF_TMEM + 4 * 32768, b + OFF_TMEM + 5 * 32768, j, k, l + 1 == L ? sw[(size_t)k * CH + j] : ow[(size_t)k * CH + j]);
                put_img(b + OFF_WS, b + OFF_WS + 32768, j, k, sw[(size_t)k * CH + j]);
            }
        for (int kb = 0; kb < KB_TOTAL; ++kb)
            for (int half = 0; half < 2; ++half) {
                unsigned char *ch = b + OFF_STREAM + (size_t)(kb * 2 + half) * CHUNK_BYTES;
                for (int j = 0; j < CH; ++j)
                    for (int kk = 0; kk < 32; ++kk) {
                        const int k = kb * 32 + kk, g = half * CH + j;
                        float v = 0.0f;
                        if (k < 2 * CH) v = cw[(size_t)k * G2 + g];                       // taps 0 and 1: rows 0 .. 2R-1
                        else if (k - 2 * CH < cin) v = iw[(size_t)(k - 2 * CH) * G2 + g];
                        put_img(ch, ch + 8192, j, kk, v);
                    }
            }
        float *bias = reinterpret_cast<float *>(b + OFF_BIAS);
        for (int j = 0; j < CH; ++j) {
            bias[j] = (cb ? cb[j] : 0.0f) + (ib ? ib[j] : 0.0f);
            bias[CH + j] = (cb ? cb[CH + j] : 0.0f) + (ib ? ib[CH + j] : 0.0f);
            bias[2 * CH + j] = ob ? ob[j] : 0.0f;
            bias[3 * CH + j] = sb ? sb[j] : 0.0f;
        }
    }
    std::vector<unsigned char> hb(HEAD_BLOB_BYTES, 0);
    {
        const float *w1 = W("final_convolution_1/kernel"), *w2 = W("final_convolution_2/kernel");
        const float *b1 = W("final_convolution_1/bias"), *b2 = W("final_convolution_2/bias");
        for (int j = 0; j < CH; ++j)
            for (int k = 0; k < CH; ++k) {
                put_tmem(hb.data(), hb.data() + 32768, j, k, w1[(size_t)k * CH + j]);
                put_tmem(hb.data() + 2 * 32768, hb.data() + 3 * 32768, j, k, j < O ? w2[(size_t)k * O + j] : 0.0f);
            }
        float *bias = reinterpret_cast<float *>(hb.data() + 4 * 32768);
        for (int j = 0; j < CH; ++j) {
            bias[j] = b1 ? b1[j] : 0.0f; bias[CH + j] = (b2 && j < O) ? b2[j] : 0.0f;
            bias[2 * CH + j] = w2[(size_t)j * O]; bias[3 * CH + j] = O > 1 ? w2[(size_t)j * O + 1] : 0.0f;
        }
    }
    std::vector<float> first(2 * CH, 0.0f);
    {
        const float *fw = W("input_convolution/kernel"), *fb = W("input_convolution/bias");
        for (int j = 0; j < CH; ++j) { first[j] = fw[j]; first[CH + j] = fb ? fb[j] : 0.0f; }
    }
    WnTcPlan *pl = new WnTcPlan();
    h->tc = pl;
    auto bail = [&](const std::string &msg) { h->err = msg; wn_tc_free(h); return WN_ERR_CUDA; };
    if (cudaMalloc((void **)&pl->d_layer_blobs, blobs.size()) != cudaSuccess) return bail("tensor-core loop engine: cudaMalloc(layer blobs)");
    if (cudaMalloc((void **)&pl->d_head_blob, hb.size()) != cudaSuccess) return bail("tensor-core loop engine: cudaMalloc(head blob)");
    if (cudaMalloc((void **)&pl->d_first, first.size() * 4) != cudaSuccess) return bail("tensor-core loop engine: cudaMalloc(first conv)");
    if (cudaMalloc((void **)&pl->d_abort, 4 * sizeof(int)) != cudaSuccess) return bail("tensor-core loop engine: cudaMalloc(abort flag)");
    if (cudaMallocHost((void **)&pl->h_abort, sizeof(int)) != cudaSuccess) return bail("tensor-core loop engine: cudaMallocHost(abort flag)");
    *pl->h_abort = 0;
    cudaMemcpy(pl->d_layer_blobs, blobs.data(), blobs.size(), cudaMemcpyHostToDevice);
    cudaMemcpy(pl->d_head_blob, hb.data(), hb.size(), cudaMemcpyHostToDevice);
    cudaMemcpy(pl->d_first, first.data(), first.size() * 4, cudaMemcpyHostToDevice);
    TcParams &p = pl->prm;
    memset(&p, 0, sizeof(p));
    p.L = L; p.O = O; p.cin = cin; p.NZ = h->m.NZ; p.input_type = c.input_type; p.legacy = c.legacy; p.residual_legacy = c.residual_legacy;
    p.log_scale_min = c.log_scale_min; p.log_scale_min_gauss = c.log_scale_min_gauss;
    long long slots = 0;
    for (int l = 0; l < L; ++l) { p.dil[l] = h->m.dil[l]; p.ring_off[l] = slots; slots += 2 * p.dil[l]; }
    p.ring_slots = slots;
    p.layer_blobs = pl->d_layer_blobs; p.head_blob = pl->d_head_blob; p.first_wb = pl->d_first; p.abort_flag = pl->d_abort;
    pl->max_replicas = h->sm_count / (L + 1);
    for (int nb : {16, 32})
        for (int na = nb - 12; na <= nb; na += 4)
            for (int st = 0; st < 2; ++st)
                if (cudaFuncSetAttribute(tc_kernel(nb, na, st != 0), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         nb == 16 ? Geo<16>::SMEM : Geo<32>::SMEM) != cudaSuccess)
                    return bail("tensor-core loop engine: shared memory attribute");
    if (pl->max_replicas < 1) return bail("tensor-core loop engine: fewer SMs than layers + 1");
    return WN_OK;
}

// largest batch one launch takes; larger batches run as several launches
int wn_tc_max_batch(const wn_handle *h) { return h->tc ? h->tc->max_replicas * Geo<16>::NG_MAX * 16 : 0; }

void wn_tc_fill_info(wn_handle *h, wn_info *info) {
    wn_tc2_fill_info(h, info);
    if (!h->tc || h->engine != WN_ENGINE_TC) return;
    info->pipeline_ctas = h->tc->last_nr * (h->cfg.layers + 1);
    info->pipeline_ctas_per_layer = h->tc->last_nr;
    info->pipeline_group_size = h->tc->last_nb;
}

int wn_tc_check_abort(wn_handle *h) {   // after the stream has been synchronised
    if (h->tc2) return wn_tc2_check_abort(h);
    WnTcPlan *pl = h->tc;
    if (!pl || !pl->h_abort) return WN_OK;
    if (pl->d_stats && pl->stats_ctas > 0 && getenv("WN_TC_STATS")) {   // development aid
        std::vector<long long> st((size_t)pl->stats_ctas * 16);
        if (cudaMemcpy(st.data(), pl->d_stats, st.size() * sizeof(long long), cudaMemcpyDeviceToHost) == cudaSuccess) {
            FILE *f = fopen(getenv("WN_TC_STATS"), "a");
            if (f) {
                fprintf(f, "# tc stats: L=%d NB=%d NR=%d cluster=%d; cols: cta stage  layer: poll image+arrive->gate gate-epilogue->z z->out out+skip | head: skip-polls 1x1s sampler ; total steps\n",
                        pl->prm.L, pl->last_nb, pl->last_nr, pl->last_cluster);
                for (int i = 0; i < pl->stats_ctas; ++i) {
                    const long long *o = st.data() + (size_t)i * 16;
                    if (o[15] == 0) continue;
                    fprintf(f, "%d %d", i, i % pl->last_stride);
                    for (int k = 0; k < 5; ++k) fprintf(f, " %.0f", (double)o[k] / (double)o[15]);
                    if (o[10]) fprintf(f, " [x-independent part: wait for operands %.0f, stream + MMAs %.0f]", (double)o[8] / (double)o[10], (double)o[9] / (double)o[10]);
                    fprintf(f, " | %.0f %lld\n", (double)o[14] / (double)o[15], o[15]);
                }
                fclose(f);
            }
        }
        pl->stats_ctas = 0;
    }
    if (*pl->h_abort == 0) return WN_OK;
    h->err = "tensor-core loop engine: watchdog - a CTA waited too long for a peer";
    return WN_ERR_WATCHDOG;
}

static int tc_launch_chunk(wn_handle *h, TcParams p, int b0, int nbatch, cudaStream_t st) {
    WnTcPlan *pl = h->tc;
    // How the batch is laid over the chip.  A replica is a chain of L + 1 CTAs; NG utterance groups of NA utterances follow each
    // other through it.  Measured: the step time of a group grows with NA (epilogue work, hand-off bytes:
    // 23.3 / 28.7 / 33.8 / 37.5 us at 4 / 8 / 12 / 16), every further group in the chain costs 0.2 - 0.7 us, more than 5 replicas
    // lose the clusters of 8 (6 replicas: +1.9 us, 7: +4.3 us at one group of 4); idle SMs cost nothing.  The cheapest feasible
    // (NA, NG, NR) wins.
    int nb = 16, na = 16, ng = 1, nr = 0;
    {
        static const float base[4] = {23.3f, 28.7f, 33.8f, 37.5f};
        float best = 1e30f;
        for (int a = 4; a <= 16; a += 4)
            for (int g = 1; g <= Geo<16>::NG_MAX; ++g) {
                const int groups = (nbatch + a - 1) / a, r = (groups + g - 1) / g;
                if (r > pl->max_replicas) continue;
                const float cost = base[a / 4 - 1] + 0.4f * (g - 1) + (r > 6 ? 4.3f : r > 5 ? 1.9f : 0.0f) + 0.01f * r * g;
                if (cost < best) { best = cost; na = a; ng = g; nr = r; }
            }
    }
    if (const char *e = getenv("WN_TC_NB")) { const int v = atoi(e); if (v == 16 || v == 32) nb = v; }
    if (nr == 0) nb = 32;   // more than 14 groups of 16: 32-wide MMAs, one group per chain
    if (nb == 32) {
        ng = 1; na = 32;
        for (int a = 20; a <= 32; a += 4)
            if ((nbatch + a - 1) / a <= pl->max_replicas) { na = a; break; }
    }
    if (const char *e = getenv("WN_TC_NA")) { const int v = atoi(e); if (v >= nb - 12 && v <= nb && v % 4 == 0) na = v; }
    if (const char *e = getenv("WN_TC_NG")) { const int v = atoi(e); if (v >= 1 && v <= (nb == 16 ? Geo<16>::NG_MAX : 1)) ng = v; }
    nr = ((nbatch + na - 1) / na + ng - 1) / ng;
    if (nr > pl->max_replicas) { h->err = "tensor-core loop engine: batch too large for one launch"; return WN_ERR_INVALID; }
    pl->last_nb = na; pl->last_nr = nr; pl->last_ng = ng;
    p.NR = nr; p.NG = ng; p.NA = na; p.b0 = b0; p.bend = b0 + nbatch;
    const int L = p.L;
    // mailboxes, per utterance group: x [groups][L][2][128*NB] | skip [groups][L][2][128*NB] | in [groups][2][NB], 8-byte {value, tag} words, zeroed (tag 0 = empty)
    const size_t mailw = (size_t)CH * nb;
    const size_t ngt = (size_t)nr * ng;   // utterance groups of the launch
    const size_t words = 2 * ngt * L * 2 * mailw + ngt * 2 * nb;
    if (pl->mail_words < words) {
        if (pl->d_mail) cudaFree(pl->d_mail);
        pl->d_mail = nullptr; pl->mail_words = 0;
        if (cudaMalloc((void **)&pl->d_mail, words * 8) != cudaSuccess) { h->err = "tensor-core loop engine: cudaMalloc(mailboxes)"; return WN_ERR_NOMEM; }
        pl->mail_words = words;
    }
    p.mail_x = pl->d_mail;
    p.mail_skip = p.mail_x + ngt * L * 2 * mailw;
    p.mail_in = p.mail_skip + ngt * L * 2 * mailw;
    // dilation rings [NR][sum 2d][128*NB] floats: every slot is written before it is read (steps before 0 read zeros instead)
    const size_t ring_floats = ngt * p.ring_slots * mailw;
    if (pl->ring_floats < ring_floats) {
        if (pl->d_ring) cudaFree(pl->d_ring);
        pl->d_ring = nullptr; pl->ring_floats = 0;
        if (cudaMalloc((void **)&pl->d_ring, ring_floats * sizeof(float)) != cudaSuccess) { h->err = "tensor-core loop engine: cudaMalloc(rings)"; return WN_ERR_NOMEM; }
        pl->ring_floats = ring_floats;
    }
    p.ring = pl->d_ring;
    WN_CUDA_TRY(h, cudaMemsetAsync(pl->d_mail, 0, words * 8, st));
    p.stats = nullptr;
    if (getenv("WN_TC_STATS")) {
        if (!pl->d_stats) cudaMalloc((void **)&pl->d_stats, sizeof(long long) * 16 * 256);
        if (pl->d_stats) { cudaMemsetAsync(pl->d_stats, 0, sizeof(long long) * 16 * 256, st); p.stats = pl->d_stats; pl->stats_ctas = 256; }
    }
    // all CTAs wait for each other: they must be co-resident -> cooperative launch (one CTA per SM, grid <= SM count)
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(TC_THREADS); cfg.stream = st;
    cfg.dynamicSmemBytes = nb == 16 ? Geo<16>::smem_for(ng) : Geo<32>::smem_for(1);
    const void *kern = tc_kernel(nb, na, p.stats != nullptr);
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeCooperative;
    at[0].val.cooperative = 1;
    // Consecutive stages of a replica share a thread-block cluster: inside a cluster x is handed over through distributed shared
    // memory (~0.33 us cheaper per hop than an L2 mailbox).  A replica occupies `stride` consecutive CTAs, L + 1 of them working:
    // padding the 21 stages of the default stack to 24 lets clusters of 12 (one L2 hop inside a replica; one such cluster fits a
    // GPC: up to 4 replicas), 8 (two hops, 5 replicas) and 6 (three hops, 6 replicas) divide it; unpadded, the divisors of L + 1
    // remain (7 replicas: clusters of 3).  The candidate with the fewest L2 hops whose clusters are all co-resident is used.
    struct Cand { int stride, csize, hops; };
    std::vector<Cand> cands;
    const int pad = ((L + 1 + 11) / 12) * 12;
    for (int c : {12, 8, 6, 4}) cands.push_back({pad, c, L / c});   // layer l sits in slot l + 1: its x crosses a cluster boundary when c divides l + 1
    for (int c = 8; c >= 2; --c)
        if ((L + 1) % c == 0) cands.push_back({L + 1, c, L / c});
    std::stable_sort(cands.begin(), cands.end(), [](const Cand &a, const Cand &b) { return a.hops < b.hops; });
    int csize = 1, stride = L + 1;
    const char *ecl = getenv("WN_TC_CLUSTER");
    const int want = ecl ? atoi(ecl) : 0;
    for (const Cand &cd : cands) {
        if (want > 0 && cd.csize != want) continue;
        if (nr * cd.stride > h->sm_count) continue;
        if (cd.csize > 8 && cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) { cudaGetLastError(); continue; }
        cfg.gridDim = dim3(nr * cd.stride);
        at[1].id = cudaLaunchAttributeClusterDimension;
        at[1].val.clusterDim.x = cd.csize; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
        cfg.attrs = at + 1; cfg.numAttrs = 1;
        int ncl = 0;
        if (cudaOccupancyMaxActiveClusters(&ncl, kern, &cfg) == cudaSuccess && ncl * cd.csize >= nr * cd.stride) { csize = cd.csize; stride = cd.stride; break; }
        cudaGetLastError();
    }
    p.stride = stride;
    cfg.gridDim = dim3(nr * stride);
    at[1].id = cudaLaunchAttributeClusterDimension;
    at[1].val.clusterDim.x = csize; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = csize > 1 ? 2 : 1;
    pl->last_cluster = csize; pl->last_stride = stride;
    void *args[] = {&p};
    cudaError_t le = cudaLaunchKernelExC(&cfg, kern, args);
    if (le != cudaSuccess && csize > 1) {   // the occupancy query was optimistic (another context on the GPU, ...): every hand-off through L2
        cudaGetLastError();
        p.stride = L + 1;
        cfg.gridDim = dim3(nr * (L + 1));
        cfg.numAttrs = 1;
        pl->last_cluster = 1; pl->last_stride = L + 1;
        le = cudaLaunchKernelExC(&cfg, kern, args);
    }
    WN_CUDA_TRY(h, le);
    h->launches++;
    return WN_OK;
}

// Batches above the launch limit (7 replicas x 4 groups x 16 utterances on 148 SMs) run as consecutive launches over equal chunks.
int wn_launch_loop_tc(wn_handle *h, const float *c_up, const float *noise, const float *teacher, int inputs_mode, int B, int T,
                      float *audio, float *raw, cudaStream_t st) {
    if (h->tc2) return wn_launch_loop_tc2(h, c_up, noise, teacher, inputs_mode, B, T, audio, raw, st);
    WnTcPlan *pl = h->tc;
    if (!pl) { h->err = "tensor-core loop engine not planned"; return WN_ERR_STATE; }
    TcParams p = pl->prm;
    p.B = B; p.T = T;
    p.c_up = c_up; p.noise = noise; p.teacher = teacher; p.inputs_mode = inputs_mode; p.audio = audio; p.raw = raw;
    {   // abort flag + the watchdog's wall-clock limit
        unsigned long long limit = WATCHDOG_NS_DEFAULT;
        if (const char *e = getenv("WN_WATCHDOG_MS")) { const long long v = atoll(e); if (v > 0) limit = (unsigned long long)v * 1000000ull; }
        pl->h_abort_init[0] = 0; pl->h_abort_init[1] = 0;
        memcpy(&pl->h_abort_init[2], &limit, 8);
        WN_CUDA_TRY(h, cudaMemcpyAsync(pl->d_abort, pl->h_abort_init, 4 * sizeof(int), cudaMemcpyHostToDevice, st));
    }
    p.test_stall_step = -1;
    if (const char *e = getenv("WN_TC_TEST_STALL")) p.test_stall_step = atoi(e);
    p.no_fb = getenv("WN_TC_NOFB") ? 1 : 0;
    p.no_w2c = getenv("WN_TC_NOW2C") ? 1 : 0;
    const int cap = pl->max_replicas * Geo<16>::NG_MAX * 16;   // groups of 16 utterances in every chain
    const int nchunks = (B + cap - 1) / cap, per = (B + nchunks - 1) / nchunks;
    for (int b0 = 0; b0 < B; b0 += per) {
        const int rc = tc_launch_chunk(h, p, b0, std::min(per, B - b0), st);
        if (rc) return rc;
    }
    WN_CUDA_TRY(h, cudaMemcpyAsync(pl->h_abort, pl->d_abort, sizeof(int), cudaMemcpyDeviceToHost, st));
    return WN_OK;
}

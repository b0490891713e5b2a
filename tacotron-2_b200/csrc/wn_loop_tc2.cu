// wn_loop_tc2.cu - the autoregressive loop (wavenet.py:805-911) on the tensor cores for the 256-channel block of
// paper_hparams.py (R = S = G/2 = 256; BASELINE.json configs[3]).  Same construction as wn_loop_tc.cu - a chain of
// layer-stationary CTAs, step-critical weights resident in tensor memory as the A operand of tcgen05.mma, fp16 (hi, lo)
// operand pairs, the x-independent part accumulated between the steps from L2-streamed weights - but a layer's weights are 4x
// those of the default block (cur tap + out + skip: 1 MB as fp16 pairs), so a LAYER IS FOUR CTAs ("quarters"):
//
//   quarter q owns the gate channels [64q, 64q + 64) - one M = 128 tile = their 64 tanh rows + their 64 sigmoid rows over
//   K = 256 - and one M = 128 tile of the second GEMM = 64 residual-out rows + 64 skip rows over K = 256 (the gated channels).
//   Tensor memory: gate tile hi | lo (256 columns), out/skip tile hi (128), one accumulator per utterance group, and behind them the
//   out/skip lo weights of as many k steps as fit; the out/skip lo half (64 KB, used for the remaining k steps) and the two stages
//   of the weight stream sit in shared memory.
//
//   Both GEMMs need ALL 256 channels of their input, produced 64 per quarter: the all-gather is a shared mailbox in L2 -
//   every quarter writes its 64 channels as {value, tag} words, every quarter polls all 256.  Per layer and step:
//     x_l(t) [256] polled -> fp16 image -> gate MMAs (on top of the x-independent part) -> tanh rows / sigmoid rows of a channel
//     sit in different TMEM lanes: the sigmoid warps pass their pre-activations through shared memory -> z [64] -> z mailbox
//     -> z [256] polled -> image -> out/skip MMAs -> x_{l+1} [64] -> next layer's x mailbox, skip [64] -> head's mailbox.
//   Every quarter keeps its own copy of the layer's dilation ring (global memory, written and read by the same thread).
//
//   The head CTA: running skip sum [256] in the reference's order, relu, final 1x1 #1 (two M = 128 tiles on the tensor cores,
//   hi in tensor memory, lo in shared memory), relu, final 1x1 #2 (<= 32 rows: CUDA cores, fp32), sampler, feedback.
//
// Arithmetic and parity: as wn_loop_tc.cu (fp32 accumulation, 22-bit operands; 1e-3 criterion, observed ~1e-5).
#include <cuda_fp16.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "wn_common.cuh"
#include "wn_umma.cuh"

namespace {

#include "wn_tc_common.cuh"

constexpr int C2 = 256;                  // residual = gate / 2 = skip channels
constexpr int QC = 64;                   // channels of a quarter
constexpr int NB = 16;                   // N of every MMA (utterances of a launch, padded)
constexpr int EPI_THREADS = 256;         // thread c <-> channel c for mailboxes / images; TMEM lane = 32 * (warp % 4) + lane
constexpr int MMA_WARP = 8;
constexpr int T2_THREADS = 320;
constexpr int CI_K = 96;
constexpr int KB_TOTAL = (2 * C2 + CI_K) / 32;   // 19 k blocks of the streamed [tap0 | tap1 | cin] rows
constexpr int CHUNK_BYTES = 16384;               // A image [4][128][8] fp16 hi, then lo
constexpr int LBO = NB * 16 + 16, PLANE = (C2 / 8) * LBO, IMG = 2 * PLANE, CI_PLANE = (CI_K / 8) * LBO;

// quarter blob (bytes)
constexpr int OFF_TMEM = 0;                          // 3 x [128 rows][128 words]: gate hi, gate lo, out/skip hi
constexpr int OFF_OSLO = 3 * 65536;                  // out/skip lo: A image [32][128][8]
constexpr int OFF_STREAM = OFF_OSLO + 65536;         // KB_TOTAL chunks
constexpr int OFF_BIAS = OFF_STREAM + KB_TOTAL * CHUNK_BYTES;   // gate rows [128], out/skip rows [128]
constexpr int OFF_OSLO_T = OFF_BIAS + 2 * 128 * 4;              // out/skip lo once more, as tensor-memory rows [128][128 words]
constexpr int QUARTER_BLOB_BYTES = OFF_OSLO_T + 65536;
// head blob: W1 hi 2 x [128][128 words] | W1 lo 2 x A image [32][128][8] | W2 fp32 [32][256] | b1 [256] | b2 [32] | first conv w [256], b [256]
constexpr int HOFF_W1HI = 0, HOFF_W1LO = 2 * 65536, HOFF_W2 = 4 * 65536, HOFF_B1 = HOFF_W2 + 32 * C2 * 4, HOFF_B2 = HOFF_B1 + C2 * 4,
              HOFF_FIRST = HOFF_B2 + 32 * 4, HOFF_W1LO_T = HOFF_FIRST + 2 * C2 * 4 /* W1 lo once more, as tensor-memory rows 2 x [128][128 words] */,
              HEAD_BLOB_BYTES = HOFF_W1LO_T + 2 * 65536;

// tensor memory columns
constexpr int NG_MAX = 3;                                                    // utterance groups that follow each other through the chain
// every accumulator is 2 NB columns: the main terms (hi * hi) | the cross terms (hi * lo + lo * hi), added in the epilogue - the
// tensor core truncates an addend when it aligns it to the accumulator, so the small cross terms accumulate among themselves
constexpr int ACCW = 2 * NB;
constexpr int COL_WG = 0, COL_WOS = 256, COL_AG = 384;                       // layer quarter: weights, then the accumulators of the groups
constexpr int COL_W1 = 0, COL_H0 = 384, COL_H1 = COL_H0 + ACCW;              // head: W1 hi (two tiles) | W1 lo of tile 0 | accumulators | W1 lo of tile 1, k < 128
constexpr int COL_W1LO0 = 256, COL_W1LO1 = COL_H1 + ACCW, HEAD_NKS1 = (512 - COL_W1LO1) / 8;   // 8 of tile 1's 16 lo k steps fit
// Layer quarter, columns from COL_AG on, for a kernel built for NGT groups: one accumulator per group - two-pass kernels (groups of
// <= 8: main | cross inside NB columns) NB columns, three-pass kernels 2 NB - which holds the gate pre-activations and, later in the
// group's turn, the out | skip rows (the gate epilogue has read it before the z barrier; the next x-independent part overwrites it
// only after the p barrier).  Everything behind the last accumulator holds out/skip LO weights, 8 columns per k step: those k
// steps of the lo * hi pass read A from tensor memory (~10 cycles per MMA), the rest from shared memory (~60).
__host__ __device__ constexpr int acc_w(bool wide) { return wide ? NB : 2 * NB; }
__host__ __device__ constexpr int col_acc(bool wide, int g) { return COL_AG + g * acc_w(wide); }
__host__ __device__ constexpr int nks_tmem(bool wide, int ngt) { return (512 - col_acc(wide, ngt)) / 8 < 16 ? (512 - col_acc(wide, ngt)) / 8 : 16; }
__host__ __device__ constexpr int col_oslo(bool wide, int ngt, int ks) { return col_acc(wide, ngt) + 8 * ks; }

// NH = 8 accumulator columns of this thread's lane: main + cross terms (WIDE: both inside the first NB columns, see img_put2)
template <bool WIDE>
__device__ __forceinline__ void acc_load2(uint32_t taddr, float (&v)[8]) {
    if constexpr (WIDE) {
        acc_load<8, true>(taddr, v);
    } else {
        float x[8];
        umma::tmem_ld8(taddr, v);
        umma::tmem_ld8(taddr + NB, x);
        umma::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = __fadd_rn(v[i], x[i]);
    }
}
// shared memory, layer quarter
constexpr int S_OSLO = 0, S_WST = 65536, S_ZI = S_WST + 2 * CHUNK_BYTES, S_TAP = S_ZI + IMG, S_CI = S_TAP + 2 * IMG, S_SG = S_CI + 2 * CI_PLANE,
              S_G0 = S_SG + QC * NB * 4;                                      // shared by the groups: z, taps, conditioning images, sigmoid rows
constexpr int G_XI = 0, G_XR = IMG, GRP = IMG + QC * NB * 4;                  // per group: x image, fp32 x rows of this quarter
// head
constexpr int S_W1LO = 0, S_W2 = 2 * 65536, S_HI = S_W2 + 32 * C2 * 4, S_H1F = S_HI + IMG, S_PRM = S_H1F + C2 * NB * 4, S_HEND = S_PRM + NB * 32 * 4;
constexpr int smem_for(int ng) { return (S_G0 + ng * GRP) > S_HEND ? (S_G0 + ng * GRP) : S_HEND; }
constexpr int MAILW = C2 * NB;           // words of one mailbox buffer / floats of one ring slot

struct Tc2Params {
    int L, O, cin, NZ, input_type, legacy, residual_legacy;
    float log_scale_min, log_scale_min_gauss;
    int dil[WN_MAX_LAYERS];
    long long ring_off[WN_MAX_LAYERS];   // sum of 2 * dilation over the layers before l (every quarter of a layer keeps its own 2d slots)
    const unsigned char *blobs, *head_blob;
    long long ring_slots;                // sum of 2 * dilation over all layers
    unsigned long long *mail_x, *mail_z, *mail_skip, *mail_in;   // [NG][L][2][MAILW] each; in: [NG][2][NB]
    float *ring;                         // [NG][4 * sum 2d][MAILW]
    int NG;                              // utterance groups of the launch (NA utterances each; NA is a template parameter)
    int *abort_flag;
    long long *stats;                    // development aid (WN_TC_STATS): per CTA 16 cycle counters
    const float *c_up, *noise, *teacher;
    int inputs_mode;
    float *audio, *raw;
    int B, T, b0, bend;
    int test_stall_step;                 // test hook (WN_TC_TEST_STALL): the head stops feeding back after this step
};

struct Bars2 {
    unsigned long long x, z, gate, os, p, pdone, sg, h1, part, prm, wfull[2], wempty[2];
};

// the first NA of this channel's 16 words: units of two words, 2 * C2 words apart
template <int NA>
__device__ __forceinline__ bool poll_channel(const unsigned long long *p, unsigned tag, float (&v)[NA], int *abort_flag) {
    return poll_units<NA / 2>(p, 2 * C2, tag, v, NA / 2, abort_flag);
}

// ---------------------------------------------------------------------------------------------------------------------
// layer quarter
// ---------------------------------------------------------------------------------------------------------------------
template <int NA, int NGT, bool STATS>
__device__ void quarter_epilogue(const Tc2Params &P, const int l, const int q, unsigned char *smem, const uint32_t tmem, Bars2 &bars) {
    constexpr bool WIDE = NA <= 8;
    constexpr int nah = NA / 2, NH = 8;
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, q4 = w & 3, hh = w >> 2;
    const int c = tid;                                   // channel of this thread for mailboxes, images, the ring
    const int row = 32 * q4 + lane, n0 = hh * nah;       // TMEM lane and first utterance of this thread in the epilogues
    const uint32_t lane_addr = tmem + ((uint32_t)(32 * q4) << 16);
    const int L = P.L, T = P.T, cin = P.cin, NG = P.NG;
    const unsigned char *blob = P.blobs + (long long)(l * 4 + q) * QUARTER_BLOB_BYTES;
    const float *bias = reinterpret_cast<const float *>(blob + OFF_BIAS);
    const float bg = __ldg(bias + row), bos = __ldg(bias + 128 + row);
    const float *first = reinterpret_cast<const float *>(P.head_blob + HOFF_FIRST);
    const float wf = l == 0 ? __ldg(first + c) : 0.0f, bf = l == 0 ? __ldg(first + C2 + c) : 0.0f;
    const float cres = P.residual_legacy ? 0.70710677f : 1.0f;
    const int d = P.dil[l], qmask = 2 * d - 1;
    const int coff = (c >> 3) * LBO + (c & 7) * 2;       // element (channel c, utterance 0) inside an image plane
    const long long gmail = (long long)L * 2 * MAILW;    // mailbox words of one group
    const unsigned long long *mx_in = P.mail_x + (long long)l * 2 * MAILW + (long long)c * 2;
    const unsigned long long *mz_in = P.mail_z + (long long)l * 2 * MAILW + (long long)c * 2;
    // writers: channel 64q + (row % 64), units of this thread's utterances
    const long long wunit = ((long long)(n0 / 2) * C2 + QC * q + (row & 63)) * 2;
    unsigned long long *mz_out = P.mail_z + (long long)l * 2 * MAILW + wunit;
    unsigned long long *mx_out = P.mail_x + (long long)(l + 1) * 2 * MAILW + wunit;
    unsigned long long *ms_out = P.mail_skip + (long long)l * 2 * MAILW + wunit;
    // this quarter's copy of the layer's ring, per group: [slot][unit][channel][2]
    float *ring0 = P.ring + (4 * P.ring_off[l] + (long long)q * 2 * d) * MAILW + (long long)c * 2;
    const long long gring = 4 * P.ring_slots * MAILW;
    unsigned char *ZI = smem + S_ZI, *TAP = smem + S_TAP, *CI = smem + S_CI;
    float *SG = reinterpret_cast<float *>(smem + S_SG);   // [64][NB]
    const uint32_t bx = umma::smem_addr(&bars.x), bz = umma::smem_addr(&bars.z), bp = umma::smem_addr(&bars.p),
                   bgate = umma::smem_addr(&bars.gate), bosb = umma::smem_addr(&bars.os), bsg = umma::smem_addr(&bars.sg),
                   bpdone = umma::smem_addr(&bars.pdone);
    constexpr int CN = (CI_K * NA + EPI_THREADS - 1) / EPI_THREADS;
    const bool mine = (c >> 6) == q;                      // this thread's channel belongs to this quarter

    // operands of the x-independent part of step t of group g: the taps t - 2d, t - d from the ring, the conditioning frame.
    // The tap / conditioning images are shared by the groups: the MMAs of the previous x-independent part must be done.
    auto prep = [&](int t, int g) -> bool {
        const float *ring = ring0 + g * gring;
        float tp0[NA], tp1[NA], cnd[CN];
        const int t0 = t - 2 * d, t1 = t - d;
#pragma unroll
        for (int u = 0; u < NA / 2; ++u) {
            float2 a = make_float2(0.f, 0.f), b = a;
            if (t0 >= 0) a = *reinterpret_cast<const float2 *>(ring + (long long)(t0 & qmask) * MAILW + (long long)u * C2 * 2);
            if (t1 >= 0) b = *reinterpret_cast<const float2 *>(ring + (long long)(t1 & qmask) * MAILW + (long long)u * C2 * 2);
            tp0[2 * u] = a.x; tp0[2 * u + 1] = a.y;
            tp1[2 * u] = b.x; tp1[2 * u + 1] = b.y;
        }
#pragma unroll
        for (int i = 0; i < CN; ++i) {
            const int e = tid + i * EPI_THREADS;
            const int n = e / cin, k = e - n * cin;
            const long long b = (long long)P.b0 + g * NA + n;
            cnd[i] = (n < NA && b < P.bend) ? __ldg(P.c_up + (b * T + t) * cin + k) : 0.0f;
        }
        const int m = t * NG + g;                         // number of this x-independent part
        if (m > 0 && mbar_wait_wd(bpdone, (m - 1) & 1, P.abort_flag)) return true;
#pragma unroll
        for (int i = 0; i < NA; i += 2) {
            img_put2<WIDE>(TAP + coff + i * 16, PLANE, tp0[i], tp0[i + 1]);
            img_put2<WIDE>(TAP + IMG + coff + i * 16, PLANE, tp1[i], tp1[i + 1]);
        }
#pragma unroll
        for (int i = 0; i < CN; ++i) {
            const int e = tid + i * EPI_THREADS;
            const int n = e / cin, k = e - n * cin;
            if (n < NA) img_put<WIDE>(CI, CI_PLANE, LBO, k, n, cnd[i]);
        }
        umma::fence_proxy_async();
        umma::fence_before_sync();
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(bp);
        return false;
    };
    for (int g = 0; g < NG; ++g)
        if (prep(0, g)) return;
    long long st_c[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, ck[10];
    const bool do_stats = STATS && P.stats != nullptr && tid == 0;
#define WN_CK(i) if (do_stats) ck[i] = clock64()

    for (int t = 0; t < T; ++t) {
      const int par = t & 1;                              // mailbox buffer of the step
      const unsigned tag = (unsigned)t + 1u;
      for (int g = 0; g < NG; ++g) {                      // the groups arrive one after the other
        const int bpar = (t * NG + g) & 1;                // phase of this CTA's barriers (one completion per group and step)
        unsigned char *XI = smem + S_G0 + g * GRP + G_XI;
        float *XR = reinterpret_cast<float *>(smem + S_G0 + g * GRP + G_XR);   // [64][NB]
        float *ring = ring0 + g * gring;
        const long long mo = g * gmail + (long long)par * MAILW;
        WN_CK(0);
        // ---- x_l(t): all 256 channels from the layer's mailbox (layer 0: the fed-back sample through the input 1x1) ---------
        float xin[NA];
        if (l == 0) {
            float s[NA];
            if (poll_units<NA / 2>(P.mail_in + (g * 2 + par) * NB, 2, tag, s, NA / 2, P.abort_flag)) return;
#pragma unroll
            for (int i = 0; i < NA; ++i) xin[i] = __fmaf_rn(s[i], wf, bf);
        } else {
            if (poll_channel<NA>(mx_in + mo, tag, xin, P.abort_flag)) return;
        }
        WN_CK(1);
        {
            unsigned char *px = XI + coff;
#pragma unroll
            for (int i = 0; i < NA; i += 2) img_put2<WIDE>(px + i * 16, PLANE, xin[i], xin[i + 1]);
            if (mine) {
#pragma unroll
                for (int i = 0; i < NA; ++i) XR[(c & 63) * NB + i] = xin[i];   // fp32 copy of this quarter's rows for the residual add
            }
        }
        umma::fence_proxy_async();
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(bx);
        WN_CK(2);
#pragma unroll
        for (int u = 0; u < NA / 2; ++u)
            *reinterpret_cast<float2 *>(ring + (long long)(t & qmask) * MAILW + (long long)u * C2 * 2) = make_float2(xin[2 * u], xin[2 * u + 1]);
        // ---- gate (modules.py:494-510): TMEM lanes 0..63 tanh rows, 64..127 the sigmoid rows of the same channels -------------
        if (mbar_wait_wd(bgate, bpar, P.abort_flag)) return;
        umma::fence_after_sync();
        WN_CK(3);
        {
            float a[NH];
            acc_load2<WIDE>(lane_addr + col_acc(WIDE, g) + n0, a);
            umma::fence_before_sync();
            if (q4 >= 2) {   // sigmoid rows: hand the pre-activations to the warps that hold the tanh rows
#pragma unroll
                for (int i = 0; i < nah; ++i) SG[(row - 64) * NB + n0 + i] = __fadd_rn(a[i], bg);
                __syncwarp();
                if (lane == 0) umma::mbar_arrive(bsg);
            } else {
                __syncwarp();
                if (lane == 0) umma::mbar_arrive(bsg);   // all 8 warps arrive: the XR rows written above are ordered too
                if (mbar_wait_wd(bsg, bpar, P.abort_flag)) return;
#pragma unroll
                for (int u = 0; u < nah / 2; ++u) {
                    const float z0 = gate_fast(__fadd_rn(a[2 * u], bg), SG[row * NB + n0 + 2 * u]);
                    const float z1 = gate_fast(__fadd_rn(a[2 * u + 1], bg), SG[row * NB + n0 + 2 * u + 1]);
                    st_ll2(mz_out + mo + (long long)u * C2 * 2, z0, z1, tag);
                }
            }
        }
        // ---- z(t): all 256 gated channels -> image ------------------------------------------------------------------------------
        WN_CK(4);
        {
            float zin[NA];
            if (poll_channel<NA>(mz_in + mo, tag, zin, P.abort_flag)) return;
            WN_CK(5);
#pragma unroll
            for (int i = 0; i < NA; i += 2) img_put2<WIDE>(ZI + coff + i * 16, PLANE, zin[i], zin[i + 1]);
        }
        umma::fence_proxy_async();
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(bz);
        WN_CK(6);
        // ---- residual-out rows (lanes 0..63) and skip rows (64..127) of this quarter (modules.py:512-520) ---------------------
        if (mbar_wait_wd(bosb, bpar, P.abort_flag)) return;
        umma::fence_after_sync();
        WN_CK(7);
        {
            float o[NH];
            acc_load2<WIDE>(lane_addr + col_acc(WIDE, g) + n0, o);
            umma::fence_before_sync();
            if (q4 < 2) {
                if (l + 1 < L) {
#pragma unroll
                    for (int u = 0; u < nah / 2; ++u) {
                        const float v0 = __fmul_rn(__fadd_rn(__fadd_rn(o[2 * u], bos), XR[row * NB + n0 + 2 * u]), cres);
                        const float v1 = __fmul_rn(__fadd_rn(__fadd_rn(o[2 * u + 1], bos), XR[row * NB + n0 + 2 * u + 1]), cres);
                        st_ll2(mx_out + mo + (long long)u * C2 * 2, v0, v1, tag);
                    }
                }
            } else {
#pragma unroll
                for (int u = 0; u < nah / 2; ++u)
                    st_ll2(ms_out + mo + (long long)u * C2 * 2, __fadd_rn(o[2 * u], bos), __fadd_rn(o[2 * u + 1], bos), tag);
            }
        }
        WN_CK(8);
        if (do_stats) {
            ck[9] = ck[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) st_c[i] += ck[i + 1] - ck[i];
        }
      }
      // ---- off the critical path: operands of the x-independent part of the next step, group by group ------------------------------
      if (t + 1 < T) {
          const long long c0 = do_stats ? clock64() : 0;
          for (int g = 0; g < NG; ++g)
              if (prep(t + 1, g)) return;
          if (do_stats) st_c[8] += clock64() - c0;
      }
    }
#undef WN_CK
    if (do_stats) {
        long long *o = P.stats + (long long)blockIdx.x * 16;
        for (int i = 0; i < 9; ++i) o[i] = st_c[i];
        o[15] = T;
    }
}

template <bool WIDE, int NGT>
__device__ void quarter_mma(const Tc2Params &P, unsigned char *smem, const uint32_t tmem, Bars2 &bars) {
    constexpr uint32_t idesc = umma::idesc_f16(128, NB);
    const int T = P.T, NG = P.NG;
    const uint32_t s0 = umma::smem_addr(smem);
    const uint64_t oslo = umma::smem_desc(s0 + S_OSLO, 128 * 16, 128);
    const uint64_t st_hi0 = umma::smem_desc(s0 + S_WST, 128 * 16, 128), st_hi1 = umma::smem_desc(s0 + S_WST + CHUNK_BYTES, 128 * 16, 128);
    const uint64_t st_lo0 = umma::smem_desc(s0 + S_WST + 8192, 128 * 16, 128), st_lo1 = umma::smem_desc(s0 + S_WST + CHUNK_BYTES + 8192, 128 * 16, 128);
    constexpr uint64_t KS_B = (uint64_t)((2 * LBO) >> 4);        // descriptor step per k step (16 k) of an activation image
    constexpr uint64_t KS_A = (uint64_t)((2 * 128 * 16) >> 4);   // ... of a 128-row weight image
    constexpr int NKS_T = nks_tmem(WIDE, NGT);   // k steps of the out/skip lo weights that are resident in tensor memory
    const uint64_t tap_hi[3] = {umma::smem_desc(s0 + S_TAP, LBO, 128), umma::smem_desc(s0 + S_TAP + IMG, LBO, 128), umma::smem_desc(s0 + S_CI, LBO, 128)};
    const uint64_t tap_lo[3] = {umma::smem_desc(s0 + S_TAP + PLANE, LBO, 128), umma::smem_desc(s0 + S_TAP + IMG + PLANE, LBO, 128),
                                umma::smem_desc(s0 + S_CI + CI_PLANE, LBO, 128)};
    const uint64_t zi_hi = umma::smem_desc(s0 + S_ZI, LBO, 128), zi_lo = umma::smem_desc(s0 + S_ZI + PLANE, LBO, 128);

    // the x-independent part of step tp of group g (taps t - 2d, t - d and the conditioning), accumulated from scratch into the
    // group's gate accumulator
    auto p_phase = [&](int tp, int g) -> bool {
        const uint32_t m = (uint32_t)(tp * NG + g);
        const uint32_t D_G = tmem + col_acc(WIDE, g);
        if (mbar_wait_wd(umma::smem_addr(&bars.p), m & 1u, P.abort_flag)) return true;
        umma::fence_after_sync();
#pragma unroll
        for (int kb = 0; kb < KB_TOTAL; ++kb) {
            const uint32_t use = (uint32_t)KB_TOTAL * m + kb;
            const uint32_t stage = use & 1u;
            if (mbar_wait_wd(umma::smem_addr(&bars.wfull[stage]), (use >> 1) & 1u, P.abort_flag)) return true;
            umma::fence_after_sync();
            if (umma::elect_one()) {
                const int src = kb / 8, kin = (kb % 8) * 2;   // image (tap0, tap1, cin) and first k step inside it
                const uint64_t ah0 = stage ? st_hi1 : st_hi0, al0 = stage ? st_lo1 : st_lo0;
#pragma unroll
                for (int kk = 0; kk < 2; ++kk) {
                    const uint64_t ah = ah0 + kk * KS_A, al = al0 + kk * KS_A;
                    const uint64_t bh = tap_hi[src] + (kin + kk) * KS_B, bl = tap_lo[src] + (kin + kk) * KS_B;
                    umma::mma_bf16(D_G, ah, bh, idesc, (kb | kk) != 0);
                    if (!WIDE) umma::mma_bf16(D_G + NB, ah, bl, idesc, (kb | kk) != 0);
                    umma::mma_bf16(WIDE ? D_G : D_G + NB, al, WIDE ? bl : bh, idesc, 1);
                }
                umma::mma_commit(umma::smem_addr(&bars.wempty[stage]));
            }
            __syncwarp();
        }
        if (umma::elect_one()) umma::mma_commit(umma::smem_addr(&bars.pdone));   // the shared tap / conditioning images are free again
        __syncwarp();
        return false;
    };
    for (int g = 0; g < NG; ++g)
        if (p_phase(0, g)) return;
    for (int t = 0; t < T; ++t) {
      for (int g = 0; g < NG; ++g) {
        const int par = (t * NG + g) & 1;
        const uint32_t D_G = tmem + col_acc(WIDE, g), D_OS = D_G;   // the out | skip accumulator of the turn = the group's gate columns
        const uint32_t xi = s0 + S_G0 + g * GRP + G_XI;
        const uint64_t xi_hi = umma::smem_desc(xi, LBO, 128), xi_lo = umma::smem_desc(xi + PLANE, LBO, 128);
        // current tap of the dilated conv (modules.py:294-297) on top of the x-independent part
        if (mbar_wait_wd(umma::smem_addr(&bars.x), par, P.abort_flag)) return;
        umma::fence_after_sync();
        if (umma::elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 16; ++ks) {
                const uint64_t bh = xi_hi + ks * KS_B, bl = xi_lo + ks * KS_B;
                umma::mma_bf16_ts(D_G, tmem + COL_WG + ks * 8, bh, idesc, 1);
                if (!WIDE) umma::mma_bf16_ts(D_G + NB, tmem + COL_WG + ks * 8, bl, idesc, 1);
                umma::mma_bf16_ts(WIDE ? D_G : D_G + NB, tmem + COL_WG + 128 + ks * 8, WIDE ? bl : bh, idesc, 1);
            }
            umma::mma_commit(umma::smem_addr(&bars.gate));
        }
        __syncwarp();
        // residual 1x1 and skip 1x1 on z (modules.py:512-515): one tile, hi in tensor memory, lo in shared memory
        if (mbar_wait_wd(umma::smem_addr(&bars.z), par, P.abort_flag)) return;
        umma::fence_after_sync();
        if (umma::elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 16; ++ks) {
                umma::mma_bf16_ts(D_OS, tmem + COL_WOS + ks * 8, zi_hi + ks * KS_B, idesc, ks != 0);
                if (!WIDE) umma::mma_bf16_ts(D_OS + NB, tmem + COL_WOS + ks * 8, zi_lo + ks * KS_B, idesc, ks != 0);
                if (ks < NKS_T) umma::mma_bf16_ts(WIDE ? D_OS : D_OS + NB, tmem + col_oslo(WIDE, NGT, ks), (WIDE ? zi_lo : zi_hi) + ks * KS_B, idesc, 1);
                else umma::mma_bf16(WIDE ? D_OS : D_OS + NB, oslo + ks * KS_A, (WIDE ? zi_lo : zi_hi) + ks * KS_B, idesc, 1);
            }
            umma::mma_commit(umma::smem_addr(&bars.os));
        }
        __syncwarp();
      }
      if (t + 1 < T)
          for (int g = 0; g < NG; ++g)
              if (p_phase(t + 1, g)) return;
    }
}

// streams the [tap0 | tap1 | cin] rows of the quarter, KB_TOTAL chunks per step, through the two stages
__device__ void quarter_loader(const Tc2Params &P, const int l, const int q, unsigned char *smem, Bars2 &bars) {
    if ((threadIdx.x & 31) != 0) return;
    const unsigned char *stream = P.blobs + (long long)(l * 4 + q) * QUARTER_BLOB_BYTES + OFF_STREAM;
    const uint32_t dst0 = umma::smem_addr(smem + S_WST);
    const int passes = P.T * P.NG;   // one pass over the stream per step and utterance group
    for (int m = 0; m < passes; ++m)
        for (int kb = 0; kb < KB_TOTAL; ++kb) {
            const uint32_t use = (uint32_t)KB_TOTAL * (uint32_t)m + kb, stage = use & 1u;
            if (mbar_wait_wd(umma::smem_addr(&bars.wempty[stage]), ((use >> 1) & 1u) ^ 1u, P.abort_flag)) return;
            const uint32_t full = umma::smem_addr(&bars.wfull[stage]);
            umma::mbar_expect_tx(full, CHUNK_BYTES);
            umma::bulk_g2s(dst0 + stage * CHUNK_BYTES, stream + (size_t)kb * CHUNK_BYTES, CHUNK_BYTES, full);
        }
}

// ---------------------------------------------------------------------------------------------------------------------
// head CTA
// ---------------------------------------------------------------------------------------------------------------------
template <int NA, bool STATS>
__device__ void head2_epilogue(const Tc2Params &P, unsigned char *smem, const uint32_t tmem, Bars2 &bars) {
    constexpr bool WIDE = NA <= 8;
    constexpr int nah = NA / 2, NH = 8, NPT = NA / 8;    // NPT: utterances per thread in the second 1x1
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, q4 = w & 3, hh = w >> 2;
    const int c = tid, row = 32 * q4 + lane, n0 = hh * nah;
    const uint32_t lane_addr = tmem + ((uint32_t)(32 * q4) << 16);
    const int L = P.L, T = P.T, O = P.O, NG = P.NG;
    const float *b1p = reinterpret_cast<const float *>(P.head_blob + HOFF_B1), *b2p = reinterpret_cast<const float *>(P.head_blob + HOFF_B2);
    const float b1[2] = {__ldg(b1p + row), __ldg(b1p + 128 + row)};
    const float b2 = __ldg(b2p + lane);
    const float cskip = 0.70710677f;
    const int coff = (c >> 3) * LBO + (c & 7) * 2;
    unsigned char *HI = smem + S_HI;
    float *PART = reinterpret_cast<float *>(smem + S_W2);                 // [8 warps][NB][32 rows]: partial sums of the second 1x1
    // final_convolution_2 on the CUDA cores: warp w owns k in [32 w, 32 w + 32), lane = output row; the weights stay in registers
    float4 w2r[8];
#pragma unroll
    for (int k4 = 0; k4 < 8; ++k4) w2r[k4] = __ldg(reinterpret_cast<const float4 *>(P.head_blob + HOFF_W2) + (8 * w + k4) * 32 + lane);
    float *H1F = reinterpret_cast<float *>(smem + S_H1F);                 // [NB][256]
    float *prm = reinterpret_cast<float *>(smem + S_PRM);                 // [NB][32]
    const uint32_t bx = umma::smem_addr(&bars.x), bgate = umma::smem_addr(&bars.gate), bh1 = umma::smem_addr(&bars.h1), bprm = umma::smem_addr(&bars.prm), bpart = umma::smem_addr(&bars.part);
    // initial input (wavenet.py:433-445): raw 0.0, mulaw mulaw(0) = 0.0; WaveNet.step semantics: the first input is x[0]
    if (tid < NB)
        for (int g = 0; g < NG; ++g) {
            const long long b = (long long)P.b0 + g * NA + tid;
            st_ll1(P.mail_in + g * 2 * NB + tid, (tid < NA && b < P.bend && P.inputs_mode) ? P.teacher[b * T] : 0.0f, 1u);
        }
    const bool slane = tid < NA;
    long long st_c[7] = {0, 0, 0, 0, 0, 0, 0}, ck[8];
    const bool do_stats = STATS && P.stats != nullptr && tid == 0;
#define WN_CK(i) if (do_stats) ck[i] = clock64()
    for (int t = 0; t < T; ++t) {
      for (int g = 0; g < NG; ++g) {                       // the groups arrive one after the other
        const int mpar = t & 1;                            // mailbox buffer of the step
        const int par = (t * NG + g) & 1;                  // phase of this CTA's barriers
        const unsigned tag = (unsigned)t + 1u;
        const unsigned long long *ms_in = P.mail_skip + (long long)g * L * 2 * MAILW + (long long)c * 2;
        const long long bu = (long long)P.b0 + g * NA + tid;   // utterance of a sampler thread
        const bool sampler = slane && bu < P.bend;
        WN_CK(0);
        float nz0 = 0.0f, tch = 0.0f;
        if (sampler) {
            if (O == 2 && P.noise) nz0 = __ldg(P.noise + ((long long)t * P.B + bu) * P.NZ);
            if (P.teacher) tch = P.inputs_mode ? (t + 1 < T ? __ldg(P.teacher + bu * T + t + 1) : 0.0f) : __ldg(P.teacher + bu * T + t);
        }
        // mixture of logistics (mixture.py:76-107): the terms that depend on the uniform draws only, while the chain is busy
        float gum[10], lgt = 0.0f;
        if (sampler && O != 2 && !P.inputs_mode) {
            const float *u = P.noise + ((long long)t * P.B + bu) * P.NZ;
            const int K = O / 3;
#pragma unroll
            for (int k = 0; k < 10; ++k) gum[k] = k < K ? wn_logf(-wn_logf(__ldg(u + k))) : 0.0f;
            const float u2 = __ldg(u + K);
            lgt = __fsub_rn(wn_logf(u2), wn_logf(__fsub_rn(1.0f, u2)));
        }
        // running skip sum in the reference's order (wavenet.py:833-836): all utterances of channel c
        float S[NA];
        for (int l = 0; l < L; ++l) {
            float s[NA];
            if (l == L - 1) { WN_CK(1); }
            if (poll_channel<NA>(ms_in + ((long long)l * 2 + mpar) * MAILW, tag, s, P.abort_flag)) return;
#pragma unroll
            for (int i = 0; i < NA; ++i) {
                if (l == 0) S[i] = s[i];
                else S[i] = P.legacy ? __fmul_rn(__fadd_rn(S[i], s[i]), cskip) : __fadd_rn(S[i], s[i]);
            }
        }
        WN_CK(2);
#pragma unroll
        for (int i = 0; i < NA; i += 2)   // relu (wavenet.py:840)
            img_put2<WIDE>(HI + coff + i * 16, PLANE, S[i] > 0.0f ? S[i] : 0.0f, S[i + 1] > 0.0f ? S[i + 1] : 0.0f);
        umma::fence_proxy_async();
        umma::fence_before_sync();
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(bx);
        // final_convolution_1 (wavenet.py:841): two tiles of 128 rows; relu; fp32 copy for the second 1x1
        WN_CK(3);
        if (mbar_wait_wd(bgate, par, P.abort_flag)) return;
        umma::fence_after_sync();
        WN_CK(4);
#pragma unroll
        for (int tile = 0; tile < 2; ++tile) {
            float h[NH];
            acc_load2<WIDE>(lane_addr + (tile ? COL_H1 : COL_H0) + n0, h);
#pragma unroll
            for (int i = 0; i < nah; ++i) {
                const float y = __fadd_rn(h[i], b1[tile]);
                H1F[(n0 + i) * C2 + 128 * tile + row] = y > 0.0f ? y : 0.0f;
            }
        }
        umma::fence_before_sync();
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(bh1);
        if (mbar_wait_wd(bh1, par, P.abort_flag)) return;
        WN_CK(5);
        // final_convolution_2 (wavenet.py:843), fp32 (FFMA2: even and odd k in the two halves): partial sums over this warp's k ...
        for (int n = 0; n < NA; n += 4) {   // four utterances x two chains in flight
            float2 a2[4][2];
#pragma unroll
            for (int j = 0; j < 4; ++j) a2[j][0] = a2[j][1] = make_float2(0.0f, 0.0f);
#pragma unroll
            for (int k4 = 0; k4 < 8; ++k4) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float4 hv = *reinterpret_cast<const float4 *>(H1F + (n + j) * C2 + 32 * w + 4 * k4);
                    a2[j][0] = ffma2v(make_float2(w2r[k4].x, w2r[k4].y), make_float2(hv.x, hv.y), a2[j][0]);
                    a2[j][1] = ffma2v(make_float2(w2r[k4].z, w2r[k4].w), make_float2(hv.z, hv.w), a2[j][1]);
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j)
                PART[(w * NB + n + j) * 32 + lane] = __fadd_rn(__fadd_rn(a2[j][0].x, a2[j][0].y), __fadd_rn(a2[j][1].x, a2[j][1].y));
        }
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(bpart);
        if (mbar_wait_wd(bpart, par, P.abort_flag)) return;
        // ... summed over the warps in a fixed order: warp w -> utterances w, w + 8; lane -> output row
#pragma unroll
        for (int i = 0; i < NPT; ++i) {
            const int n = w + 8 * i;
            float y = PART[n * 32 + lane];
#pragma unroll
            for (int ww = 1; ww < 8; ++ww) y = __fadd_rn(y, PART[(ww * NB + n) * 32 + lane]);
            y = __fadd_rn(y, b2);
            prm[n * 32 + lane] = y;
            const long long b = (long long)P.b0 + g * NA + n;
            if (P.raw && lane < O && b < P.bend) P.raw[(b * T + t) * O + lane] = y;   // wavenet.py:847
        }
        __syncwarp();
        if (lane == 0) umma::mbar_arrive(bprm);
        if (mbar_wait_wd(bprm, par, P.abort_flag)) return;
        WN_CK(6);
        if (t == P.test_stall_step) return;   // test hook: the chain starves and every waiter must time out (WN_ERR_WATCHDOG)
        // sampling + feedback (wavenet.py:850-878)
        if (slane) {
            float nxt = 0.0f;
            if (sampler && P.inputs_mode) {
                nxt = tch;
            } else if (sampler) {
                float smp;
                if (O == 2) {
                    smp = tc_sample_gaussian(P.log_scale_min_gauss, prm + tid * 32, nz0);
                } else {
                    const float *out = prm + tid * 32;
                    const int K = O / 3;
                    int best = 0;
                    float bestv = 0.0f;
#pragma unroll
                    for (int k = 0; k < 10; ++k) {
                        const float v = __fsub_rn(out[k], gum[k]);
                        if (k < K && (k == 0 || v > bestv)) { bestv = v; best = k; }
                    }
                    const float lsr = out[2 * K + best];
                    const float ls = lsr > P.log_scale_min ? lsr : P.log_scale_min;
                    smp = __fadd_rn(out[K + best], __fmul_rn(wn_expf(ls), lgt));
                    smp = smp > -1.0f ? smp : -1.0f;
                    smp = smp < 1.0f ? smp : 1.0f;
                }
                P.audio[bu * T + t] = P.input_type == WN_INPUT_MULAW ? wn_inv_mulawf(smp) : smp;
                nxt = P.teacher ? tch : smp;
            }
            if (t + 1 < T) st_ll1(P.mail_in + (g * 2 + ((t + 1) & 1)) * NB + tid, nxt, tag + 1u);
        }
        if (do_stats) {
            ck[7] = clock64();
#pragma unroll
            for (int i = 0; i < 7; ++i) st_c[i] += ck[i + 1] - ck[i];
        }
        // prm / H1F are rewritten only after the next step's barriers: every thread passes bh1 (t + 1) after the samplers are done?
        // no - the samplers are in warp 0 only; the other warps may reach the second 1x1 of step t + 1 only after bgate(t + 1),
        // which needs x(t + 1), which needs the feedback written above: ordered by the data flow
      }
    }
#undef WN_CK
    if (do_stats) {
        long long *o = P.stats + (long long)blockIdx.x * 16;
        for (int i = 0; i < 7; ++i) o[i] = st_c[i];
        o[15] = T;
    }
}

template <bool WIDE>
__device__ void head2_mma(const Tc2Params &P, unsigned char *smem, const uint32_t tmem, Bars2 &bars) {
    constexpr uint32_t idesc = umma::idesc_f16(128, NB);
    const uint32_t s0 = umma::smem_addr(smem);
    const uint64_t hi_hi = umma::smem_desc(s0 + S_HI, LBO, 128), hi_lo = umma::smem_desc(s0 + S_HI + PLANE, LBO, 128);
    const uint64_t w1lo[2] = {umma::smem_desc(s0 + S_W1LO, 128 * 16, 128), umma::smem_desc(s0 + S_W1LO + 65536, 128 * 16, 128)};
    constexpr uint64_t KS_B = (uint64_t)((2 * LBO) >> 4), KS_A = (uint64_t)((2 * 128 * 16) >> 4);
    const int passes = P.T * P.NG;   // one pass per step and utterance group
    for (int t = 0; t < passes; ++t) {
        if (mbar_wait_wd(umma::smem_addr(&bars.x), t & 1, P.abort_flag)) return;
        umma::fence_after_sync();
        if (umma::elect_one()) {
#pragma unroll
            for (int tile = 0; tile < 2; ++tile) {
                const uint32_t D = tmem + (tile ? COL_H1 : COL_H0);
#pragma unroll
                for (int ks = 0; ks < 16; ++ks) {
                    umma::mma_bf16_ts(D, tmem + COL_W1 + tile * 128 + ks * 8, hi_hi + ks * KS_B, idesc, ks != 0);
                    if (!WIDE) umma::mma_bf16_ts(D + NB, tmem + COL_W1 + tile * 128 + ks * 8, hi_lo + ks * KS_B, idesc, ks != 0);
                    if (tile == 0 || ks < HEAD_NKS1)   // lo weights resident in tensor memory
                        umma::mma_bf16_ts(WIDE ? D : D + NB, tmem + (tile ? COL_W1LO1 : COL_W1LO0) + ks * 8, (WIDE ? hi_lo : hi_hi) + ks * KS_B, idesc, 1);
                    else umma::mma_bf16(WIDE ? D : D + NB, w1lo[tile] + ks * KS_A, (WIDE ? hi_lo : hi_hi) + ks * KS_B, idesc, 1);
                }
            }
            umma::mma_commit(umma::smem_addr(&bars.gate));
        }
        __syncwarp();
    }
}

// rows [128][128 words] of `nmat` matrices -> tensor memory columns [col0, col0 + 128 * nmat): warps 0..3, thread = row
__device__ void load_tmem_weights2(const unsigned char *src, int nmat, uint32_t tmem, int col0) {
    const int tid = threadIdx.x;
    if (tid >= 128) return;
    const uint32_t lane_addr = tmem + ((uint32_t)(tid & ~31) << 16);
    for (int m = 0; m < nmat; ++m) {
        const uint4 *row = reinterpret_cast<const uint4 *>(src + ((size_t)m * 128 + tid) * 512);
        for (int c0 = 0; c0 < 128; c0 += 16) {
            uint32_t r[16];
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                const uint4 x = __ldg(row + c0 / 4 + v);
                r[4 * v] = x.x; r[4 * v + 1] = x.y; r[4 * v + 2] = x.z; r[4 * v + 3] = x.w;
            }
            umma::tmem_st16(lane_addr + col0 + m * 128 + c0, r);
        }
    }
    umma::tmem_st_wait();
}

template <int NA, int NGT, bool STATS>
__global__ void __launch_bounds__(T2_THREADS, 1) wn_loop_tc2_kernel(const Tc2Params P) {
    extern __shared__ __align__(128) unsigned char tc_smem[];
    __shared__ Bars2 bars;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int l = blockIdx.x >> 2, q = blockIdx.x & 3;
    const bool head = (int)blockIdx.x == 4 * P.L;
    if (tid == 0) {
        umma::mbar_init(umma::smem_addr(&bars.x), EPI_THREADS / 32);
        umma::mbar_init(umma::smem_addr(&bars.z), EPI_THREADS / 32);
        umma::mbar_init(umma::smem_addr(&bars.p), EPI_THREADS / 32);
        umma::mbar_init(umma::smem_addr(&bars.sg), EPI_THREADS / 32);
        umma::mbar_init(umma::smem_addr(&bars.h1), EPI_THREADS / 32);
        umma::mbar_init(umma::smem_addr(&bars.prm), EPI_THREADS / 32);
        umma::mbar_init(umma::smem_addr(&bars.part), EPI_THREADS / 32);
        umma::mbar_init(umma::smem_addr(&bars.gate), 1);
        umma::mbar_init(umma::smem_addr(&bars.os), 1);
        umma::mbar_init(umma::smem_addr(&bars.pdone), 1);
        for (int i = 0; i < 2; ++i) { umma::mbar_init(umma::smem_addr(&bars.wfull[i]), 1); umma::mbar_init(umma::smem_addr(&bars.wempty[i]), 1); }
        umma::mbar_fence_init();
    }
    if (warp == 0) umma::tmem_alloc(umma::smem_addr(&tmem_base_s), 512);
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = tmem_base_s;
    // resident weights; zero the activation images (the padding rows / columns of the images are never written again)
    if (head) {
        load_tmem_weights2(P.head_blob + HOFF_W1HI, 2, tmem, COL_W1);
        load_tmem_weights2(P.head_blob + HOFF_W1LO_T, 1, tmem, COL_W1LO0);
        if (tid < 128) {   // tile 1: the leading k steps
            const uint32_t lane_addr = tmem + ((uint32_t)(tid & ~31) << 16);
            const uint4 *row = reinterpret_cast<const uint4 *>(P.head_blob + HOFF_W1LO_T + 65536 + (size_t)tid * 512);
            for (int pc = 0; pc < HEAD_NKS1 / 2; ++pc) {
                uint32_t r[16];
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    const uint4 x = __ldg(row + pc * 4 + v);
                    r[4 * v] = x.x; r[4 * v + 1] = x.y; r[4 * v + 2] = x.z; r[4 * v + 3] = x.w;
                }
                umma::tmem_st16(lane_addr + COL_W1LO1 + 16 * pc, r);
            }
            umma::tmem_st_wait();
        }
        const uint4 *src = reinterpret_cast<const uint4 *>(P.head_blob + HOFF_W1LO);
        for (int i = tid; i < (S_W2 - S_W1LO) / 16; i += T2_THREADS) reinterpret_cast<uint4 *>(tc_smem + S_W1LO)[i] = __ldg(src + i);   // W1 lo
        for (int i = tid; i < (S_HEND - S_HI) / 16; i += T2_THREADS) reinterpret_cast<uint4 *>(tc_smem + S_HI)[i] = make_uint4(0, 0, 0, 0);
    } else {
        const unsigned char *blob = P.blobs + (long long)blockIdx.x * QUARTER_BLOB_BYTES;
        load_tmem_weights2(blob + OFF_TMEM, 3, tmem, COL_WG);
        if (tid < 128) {   // the leading k steps of the out/skip lo weights behind the accumulators
            constexpr bool WIDE = NA <= 8;
            constexpr int NKS_T = nks_tmem(WIDE, NGT);
            const uint32_t lane_addr = tmem + ((uint32_t)(tid & ~31) << 16);
            const uint4 *row = reinterpret_cast<const uint4 *>(blob + OFF_OSLO_T + (size_t)tid * 512);
            for (int pc = 0; pc < NKS_T / 2; ++pc) {   // 16 columns = two k steps at a time
                uint32_t r[16];
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    const uint4 x = __ldg(row + pc * 4 + v);
                    r[4 * v] = x.x; r[4 * v + 1] = x.y; r[4 * v + 2] = x.z; r[4 * v + 3] = x.w;
                }
                umma::tmem_st16(lane_addr + col_oslo(WIDE, NGT, 2 * pc), r);
            }
            umma::tmem_st_wait();
        }
        const uint4 *src = reinterpret_cast<const uint4 *>(blob + OFF_OSLO);
        for (int i = tid; i < 65536 / 16; i += T2_THREADS) reinterpret_cast<uint4 *>(tc_smem + S_OSLO)[i] = __ldg(src + i);
        for (int i = tid; i < (S_G0 + P.NG * GRP - S_ZI) / 16; i += T2_THREADS) reinterpret_cast<uint4 *>(tc_smem + S_ZI)[i] = make_uint4(0, 0, 0, 0);
    }
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    if (head) {
        if (warp < 8) head2_epilogue<NA, STATS>(P, tc_smem, tmem, bars);
        else if (warp == MMA_WARP) head2_mma<(NA <= 8)>(P, tc_smem, tmem, bars);
    } else {
        if (warp < 8) quarter_epilogue<NA, NGT, STATS>(P, l, q, tc_smem, tmem, bars);
        else if (warp == MMA_WARP) quarter_mma<(NA <= 8), NGT>(P, tc_smem, tmem, bars);
        else quarter_loader(P, l, q, tc_smem, bars);
    }
    // every role returns when the sequence is done (or the watchdog fired): release tensor memory
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, 512);
}

}  // namespace

// =========================================================================================================================
// host side
// =========================================================================================================================
struct WnTc2Plan {
    Tc2Params prm;
    unsigned char *d_blobs = nullptr, *d_head_blob = nullptr;
    unsigned long long *d_mail = nullptr;
    size_t mail_words = 0;
    float *d_ring = nullptr;
    int *d_abort = nullptr, *h_abort = nullptr;
    int h_abort_init[4] = {0, 0, 0, 0};
    int last_na = 0, last_ng = 1;
    size_t ring_floats = 0;
    long long *d_stats = nullptr;
    int stats_ctas = 0;
};

static void split_host2(float v, uint16_t &hi, uint16_t &lo) {   // fp16 (hi, lo): 22 significant bits of v
    const __half h = __float2half_rn(v);
    const __half l = __float2half_rn(v - __half2float(h));
    memcpy(&hi, &h, 2);
    memcpy(&lo, &l, 2);
}

bool wn_tc2_supported(const wn_handle *h) {
    const wn_config &c = h->cfg;
    return c.residual_channels == C2 && c.gate_channels == 2 * C2 && c.skip_out_channels == C2 && c.kernel_size == 3 &&
           c.input_type != WN_INPUT_MULAW_QUANTIZE && c.gin_channels <= 0 && c.cin_channels > 0 && c.cin_channels <= CI_K &&
           c.out_channels <= 32 && c.layers >= 1 && 4 * c.layers + 1 <= h->sm_count && c.layers / c.stacks <= 12;
}

void wn_tc2_free(wn_handle *h) {
    WnTc2Plan *pl = h->tc2;
    if (!pl) return;
    if (pl->d_blobs) cudaFree(pl->d_blobs);
    if (pl->d_head_blob) cudaFree(pl->d_head_blob);
    if (pl->d_mail) cudaFree(pl->d_mail);
    if (pl->d_ring) cudaFree(pl->d_ring);
    if (pl->d_abort) cudaFree(pl->d_abort);
    if (pl->h_abort) cudaFreeHost(pl->h_abort);
    if (pl->d_stats) cudaFree(pl->d_stats);
    delete pl;
    h->tc2 = nullptr;
}

// kernel for groups of na (8: two-pass images, 16) utterances and ng groups per chain (the tensor-memory layout depends on both)
static const void *tc2_kernel(int na, int ng, bool stats) {
#define WN_TC2_K(NA_, NG_) (stats ? (const void *)wn_loop_tc2_kernel<NA_, NG_, true> : (const void *)wn_loop_tc2_kernel<NA_, NG_, false>)
    if (na <= 8) return ng <= 1 ? WN_TC2_K(8, 1) : ng == 2 ? WN_TC2_K(8, 2) : WN_TC2_K(8, 3);
    return ng <= 1 ? WN_TC2_K(16, 1) : ng == 2 ? WN_TC2_K(16, 2) : WN_TC2_K(16, 3);
#undef WN_TC2_K
}

// Packs the weights (TF layouts) into the per-CTA blobs described at the top of this file.
int wn_tc2_plan(wn_handle *h) {
    if (!wn_tc2_supported(h)) { h->err = "tensor-core loop engine (256-channel block): needs R = G/2 = S = 256, kernel_size 3, scalar input, local conditioning (<= 96 channels), no global conditioning, 4 * layers + 1 SMs"; return WN_ERR_INVALID; }
    const wn_config &c = h->cfg;
    const int L = c.layers, cin = c.cin_channels, O = c.out_channels, G2 = 2 * C2;
    auto W = [&](const std::string &n) -> const float * {
        auto it = h->host_w.find(n);
        return it == h->host_w.end() ? nullptr : it->second.data.data();
    };
    // rows [128][128 words] of a tensor-memory matrix: word w = fp16(k = 2w) | fp16(k = 2w + 1) << 16
    auto put_tmem = [](unsigned char *dst_hi, unsigned char *dst_lo, int row, int k, float v) {
        uint16_t hi, lo;
        split_host2(v, hi, lo);
        reinterpret_cast<uint16_t *>(dst_hi)[(size_t)row * 256 + k] = hi;
        if (dst_lo) reinterpret_cast<uint16_t *>(dst_lo)[(size_t)row * 256 + k] = lo;
    };
    // A image [K/8][128 rows][8] of a shared-memory operand
    auto put_img = [](unsigned char *dst_hi, unsigned char *dst_lo, int row, int k, float v) {
        uint16_t hi, lo;
        split_host2(v, hi, lo);
        const size_t e = ((size_t)(k / 8) * 128 + row) * 8 + k % 8;
        if (dst_hi) reinterpret_cast<uint16_t *>(dst_hi)[e] = hi;
        if (dst_lo) reinterpret_cast<uint16_t *>(dst_lo)[e] = lo;
    };
    std::vector<unsigned char> blobs((size_t)L * 4 * QUARTER_BLOB_BYTES, 0);
    for (int l = 0; l < L; ++l) {
        const std::string p = "ResidualConv1DGLU_" + std::to_string(l) + "/residual_block_";
        const float *cw = W(p + "causal_conv/kernel"), *iw = W(p + "cin_conv/kernel"), *ow = W(p + "out_conv/kernel"),
                    *sw = W(p + "skip_conv/kernel");
        const float *cb = W(p + "causal_conv/bias"), *ib = W(p + "cin_conv/bias"), *ob = W(p + "out_conv/bias"),
                    *sb = W(p + "skip_conv/bias");
        if (!cw || !iw || !ow || !sw) { h->err = "tensor-core loop engine: missing weights of layer " + std::to_string(l); return WN_ERR_STATE; }
        for (int q = 0; q < 4; ++q) {
            unsigned char *b = blobs.data() + (size_t)(l * 4 + q) * QUARTER_BLOB_BYTES;
            float *bias = reinterpret_cast<float *>(b + OFF_BIAS);
            for (int row = 0; row < 128; ++row) {
                const int j = QC * q + (row & 63);                   // channel of the quarter
                const int g = row < 64 ? j : C2 + j;                 // gate row: tanh half, sigmoid half (modules.py:494-510)
                for (int k = 0; k < C2; ++k) {
                    // current tap = tap 2 of the [3][R][G] kernel (modules.py:238-251: row = tap * R + ci, tap 0 oldest)
                    put_tmem(b + OFF_TMEM, b + OFF_TMEM + 65536, row, k, cw[((size_t)2 * C2 + k) * G2 + g]);
                    const float osv = row < 64 ? ow[(size_t)k * C2 + j] : sw[(size_t)k * C2 + j];
                    put_tmem(b + OFF_TMEM + 2 * 65536, nullptr, row, k, osv);
                    put_img(nullptr, b + OFF_OSLO, row, k, osv);
                    {   // the same lo halves as tensor-memory rows [128][128 words]
                        uint16_t hi16, lo16;
                        split_host2(osv, hi16, lo16);
                        reinterpret_cast<uint16_t *>(b + OFF_OSLO_T)[(size_t)row * 256 + k] = lo16;
                    }
                }
                for (int k = 0; k < KB_TOTAL * 32; ++k) {
                    float v = 0.0f;
                    if (k < 2 * C2) v = cw[(size_t)k * G2 + g];                       // taps 0 and 1: rows 0 .. 2R-1
                    else if (k - 2 * C2 < cin) v = iw[(size_t)(k - 2 * C2) * G2 + g];
                    unsigned char *ch = b + OFF_STREAM + (size_t)(k / 32) * CHUNK_BYTES;
                    put_img(ch, ch + 8192, row, k % 32, v);
                }
                bias[row] = (cb ? cb[g] : 0.0f) + (ib ? ib[g] : 0.0f);
                bias[128 + row] = row < 64 ? (ob ? ob[j] : 0.0f) : (sb ? sb[j] : 0.0f);
            }
        }
    }
    std::vector<unsigned char> hb(HEAD_BLOB_BYTES, 0);
    {
        const float *w1 = W("final_convolution_1/kernel"), *w2 = W("final_convolution_2/kernel");
        const float *b1 = W("final_convolution_1/bias"), *b2 = W("final_convolution_2/bias");
        const float *fw = W("input_convolution/kernel"), *fb = W("input_convolution/bias");
        if (!w1 || !w2 || !fw) { h->err = "tensor-core loop engine: missing head weights"; return WN_ERR_STATE; }
        float *w2s = reinterpret_cast<float *>(hb.data() + HOFF_W2), *b1s = reinterpret_cast<float *>(hb.data() + HOFF_B1),
              *b2s = reinterpret_cast<float *>(hb.data() + HOFF_B2), *first = reinterpret_cast<float *>(hb.data() + HOFF_FIRST);
        for (int j = 0; j < C2; ++j) {
            const int tile = j / 128, row = j % 128;
            for (int k = 0; k < C2; ++k) {
                const float v = w1[(size_t)k * C2 + j];
                put_tmem(hb.data() + HOFF_W1HI + tile * 65536, nullptr, row, k, v);
                put_img(nullptr, hb.data() + HOFF_W1LO + tile * 65536, row, k, v);
                {
                    uint16_t hi16, lo16;
                    split_host2(v, hi16, lo16);
                    reinterpret_cast<uint16_t *>(hb.data() + HOFF_W1LO_T + tile * 65536)[(size_t)row * 256 + k] = lo16;
                }
            }
            b1s[j] = b1 ? b1[j] : 0.0f;
            first[j] = fw[j];
            first[C2 + j] = fb ? fb[j] : 0.0f;
        }
        for (int k = 0; k < C2; ++k)
            for (int o = 0; o < O; ++o) w2s[((size_t)(k / 4) * 32 + o) * 4 + k % 4] = w2[(size_t)k * O + o];
        for (int o = 0; o < O; ++o) b2s[o] = b2 ? b2[o] : 0.0f;
    }
    WnTc2Plan *pl = new WnTc2Plan();
    h->tc2 = pl;
    auto bail = [&](const std::string &msg) { h->err = msg; wn_tc2_free(h); return WN_ERR_CUDA; };
    Tc2Params &p = pl->prm;
    memset(&p, 0, sizeof(p));
    p.L = L; p.O = O; p.cin = cin; p.NZ = h->m.NZ; p.input_type = c.input_type; p.legacy = c.legacy; p.residual_legacy = c.residual_legacy;
    p.log_scale_min = c.log_scale_min; p.log_scale_min_gauss = c.log_scale_min_gauss;
    long long slots = 0;
    for (int l = 0; l < L; ++l) { p.dil[l] = h->m.dil[l]; p.ring_off[l] = slots; slots += 2 * p.dil[l]; }
    p.ring_slots = slots;
    if (cudaMalloc((void **)&pl->d_blobs, blobs.size()) != cudaSuccess) return bail("tensor-core loop engine: cudaMalloc(layer blobs)");
    if (cudaMalloc((void **)&pl->d_head_blob, hb.size()) != cudaSuccess) return bail("tensor-core loop engine: cudaMalloc(head blob)");
    if (cudaMalloc((void **)&pl->d_abort, 4 * sizeof(int)) != cudaSuccess) return bail("tensor-core loop engine: cudaMalloc(abort flag)");
    if (cudaMallocHost((void **)&pl->h_abort, sizeof(int)) != cudaSuccess) return bail("tensor-core loop engine: cudaMallocHost(abort flag)");
    *pl->h_abort = 0;
    cudaMemcpy(pl->d_blobs, blobs.data(), blobs.size(), cudaMemcpyHostToDevice);
    cudaMemcpy(pl->d_head_blob, hb.data(), hb.size(), cudaMemcpyHostToDevice);
    p.blobs = pl->d_blobs; p.head_blob = pl->d_head_blob; p.abort_flag = pl->d_abort;
    for (int na : {8, 16})
        for (int ng = 1; ng <= NG_MAX; ++ng)
            for (int st = 0; st < 2; ++st)
                if (cudaFuncSetAttribute(tc2_kernel(na, ng, st != 0), cudaFuncAttributeMaxDynamicSharedMemorySize, smem_for(NG_MAX)) != cudaSuccess)
                    return bail("tensor-core loop engine: shared memory attribute");
    return WN_OK;
}

void wn_tc2_fill_info(wn_handle *h, wn_info *info) {
    if (!h->tc2 || h->engine != WN_ENGINE_TC) return;
    info->pipeline_ctas = 4 * h->cfg.layers + 1;
    info->pipeline_ctas_per_layer = 4;
    info->pipeline_group_size = h->tc2->last_na;
}

int wn_tc2_check_abort(wn_handle *h) {   // after the stream has been synchronised
    WnTc2Plan *pl = h->tc2;
    if (pl && pl->d_stats && pl->stats_ctas > 0 && getenv("WN_TC_STATS")) {   // development aid
        std::vector<long long> st((size_t)pl->stats_ctas * 16);
        if (cudaMemcpy(st.data(), pl->d_stats, st.size() * sizeof(long long), cudaMemcpyDeviceToHost) == cudaSuccess) {
            FILE *f = fopen(getenv("WN_TC_STATS"), "a");
            if (f) {
                fprintf(f, "# tc2 stats (cycles per step): cta layer quarter | x-poll image gate-wait gate-epilogue z-poll z-image os-wait os-epilogue prep\n");
                for (int i = 0; i < pl->stats_ctas; ++i) {
                    const long long *o = st.data() + (size_t)i * 16;
                    if (o[15] == 0) continue;
                    fprintf(f, "%d %d %d |", i, i / 4, i % 4);   // last row = head: polls of layers 0..L-2, last poll, image, -, 1x1 #1 wait, relu + barrier, 1x1 #2 + barrier, sampler
                    for (int k = 0; k < 9; ++k) fprintf(f, " %.0f", (double)o[k] / (double)o[15]);
                    fprintf(f, "\n");
                }
                fclose(f);
            }
        }
        pl->stats_ctas = 0;
    }
    if (!pl || !pl->h_abort || *pl->h_abort == 0) return WN_OK;
    h->err = "tensor-core loop engine: watchdog - a CTA waited too long for a peer";
    return WN_ERR_WATCHDOG;
}

// One chain of 4 L + 1 CTAs; up to NG_MAX groups of 8 or 16 utterances follow each other through it (a group of 8 steps in ~3/4 of
// the time of a group of 16, and every further group adds only a few us to the step: 16 utterances run as two groups of 8);
// batches above 48 run as consecutive launches.
int wn_launch_loop_tc2(wn_handle *h, const float *c_up, const float *noise, const float *teacher, int inputs_mode, int B, int T,
                       float *audio, float *raw, cudaStream_t st) {
    WnTc2Plan *pl = h->tc2;
    if (!pl) { h->err = "tensor-core loop engine not planned"; return WN_ERR_STATE; }
    Tc2Params p = pl->prm;
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
    const int cap = NG_MAX * NB;
    const int nchunks = (B + cap - 1) / cap, per = (B + nchunks - 1) / nchunks;
    for (int b0 = 0; b0 < B; b0 += per) {
        const int nbatch = std::min(per, B - b0);
        int na = nbatch <= NG_MAX * 8 ? 8 : 16;
        if (const char *e = getenv("WN_TC_NA")) { const int v = atoi(e); if ((v == 8 || v == 16) && v * NG_MAX >= nbatch) na = v; }
        const int ng = (nbatch + na - 1) / na;
        pl->last_na = na; pl->last_ng = ng;
        p.b0 = b0; p.bend = b0 + nbatch; p.NG = ng;
        // mailboxes: x | z | skip [NG][L][2][MAILW] each | in [NG][2][NB], 8-byte {value, tag} words, zeroed (tag 0 = empty)
        const size_t gm = (size_t)p.L * 2 * MAILW, words = 3 * ng * gm + (size_t)ng * 2 * NB;
        if (pl->mail_words < words) {
            if (pl->d_mail) cudaFree(pl->d_mail);
            pl->d_mail = nullptr; pl->mail_words = 0;
            if (cudaMalloc((void **)&pl->d_mail, words * 8) != cudaSuccess) { h->err = "tensor-core loop engine: cudaMalloc(mailboxes)"; return WN_ERR_NOMEM; }
            pl->mail_words = words;
        }
        p.mail_x = pl->d_mail;
        p.mail_z = p.mail_x + ng * gm;
        p.mail_skip = p.mail_z + ng * gm;
        p.mail_in = p.mail_skip + ng * gm;
        // dilation rings [NG][4 quarters' copies of sum 2d slots][MAILW]: every slot is written before it is read
        const size_t ring_floats = (size_t)ng * 4 * p.ring_slots * MAILW;
        if (pl->ring_floats < ring_floats) {
            if (pl->d_ring) cudaFree(pl->d_ring);
            pl->d_ring = nullptr; pl->ring_floats = 0;
            if (cudaMalloc((void **)&pl->d_ring, ring_floats * sizeof(float)) != cudaSuccess) { h->err = "tensor-core loop engine: cudaMalloc(rings)"; return WN_ERR_NOMEM; }
            pl->ring_floats = ring_floats;
        }
        p.ring = pl->d_ring;
        WN_CUDA_TRY(h, cudaMemsetAsync(pl->d_mail, 0, words * 8, st));
        // all CTAs wait for each other: they must be co-resident -> cooperative launch (one CTA per SM, grid <= SM count)
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(4 * p.L + 1); cfg.blockDim = dim3(T2_THREADS); cfg.stream = st;
        cfg.dynamicSmemBytes = smem_for(ng);
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeCooperative;
        at[0].val.cooperative = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        p.stats = nullptr;
        if (getenv("WN_TC_STATS")) {
            if (!pl->d_stats) cudaMalloc((void **)&pl->d_stats, sizeof(long long) * 16 * 160);
            if (pl->d_stats) { cudaMemsetAsync(pl->d_stats, 0, sizeof(long long) * 16 * 160, st); p.stats = pl->d_stats; pl->stats_ctas = 4 * p.L + 1; }
        }
        void *args[] = {&p};
        WN_CUDA_TRY(h, cudaLaunchKernelExC(&cfg, tc2_kernel(na, ng, p.stats != nullptr), args));
        h->launches++;
    }
    WN_CUDA_TRY(h, cudaMemcpyAsync(pl->h_abort, pl->d_abort, sizeof(int), cudaMemcpyDeviceToHost, st));
    return WN_OK;
}

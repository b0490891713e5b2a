// wn_step_gemm.cu - WaveNet.step (wavenet.py:650-721, modules.py:305-325, :471-521 with is_incremental = False) as a
// TIME-PARALLEL computation for sm_100a: teacher forcing makes every position independent, so each residual layer is two
// dense fp32 GEMMs over all N = B*T positions (BASELINE.json config 5):
//
//   gate :  [tanh rows ; sigmoid rows] (G x Kcat)  x  Xcat (Kcat x N),  Xcat(:, n) = [x(t-2d) | x(t-d) | c(t) | x(t)]
//           gathered on the fly from the layer input (zero before the utterance starts: the causal padding of
//           modules.py:309-312), epilogue bias + tanh*sigmoid (modules.py:510)                      -> z  [N][H]
//   out/skip : [W_out ; W_skip] ((R+S) x H) x z, epilogue bias, residual add and scale (modules.py:515-520) -> x' [N][R]
//           and the skip recursion of wavenet.py:833-836                                             -> skips [N][S]
//   head :  relu -> 1x1 -> relu -> 1x1 (wavenet.py:838-844) as two more GEMMs                          -> raw [N][O]
//
// One kernel, four epilogues.  CTA tile 128 (outputs) x 128 (positions), K step 16, 256 threads, 8 x 8 register tile per
// thread; operands staged through shared memory with register prefetch of the next K step; all multiply-adds are FFMA2
// (fma.rn.f32x2: two positions per instruction, the weight is the scalar operand) with K ascending per output, i.e. plain
// fmaf chains - NOT the dot16 order of the sample-by-sample engines: results agree with them to fp32 rounding
// (tests: 1e-3 relative, the criterion of BASELINE.json; observed ~1e-6), not bit for bit.
// Activations are position-major ([N][channels], like the upsampled conditioning), so a K step of one position is 64
// contiguous bytes.  Scope: scalar input, kernel_size 3, local conditioning, no global conditioning, R, S, cin multiples
// of 16 - anything else (and WN_STEP_SERIAL=1) runs through the loop engines in inputs mode (wn_step_full.cu).
#include <cstdlib>
#include <cstring>

#include "wn_common.cuh"

namespace {

constexpr int TM = 128, TN = 128, TK = 16, GT = 256;

enum { EPI_GATE = 0, EPI_OUTSKIP = 1, EPI_RELU = 2, EPI_PLAIN = 3 };

struct GemmArgs {
    const float *W;        // [K][Mpad]: K-major, Mpad multiple of 128 (rows beyond M are zero)
    const float *bias;     // [Mpad]
    int M, Mpad, K;
    long long N;
    int T;                 // positions per utterance
    // operand B
    const float *src;      // EPI_GATE: layer input x [N][R]; else the [N][K] matrix
    const float *cond;     // EPI_GATE: c_up [N][cin]
    int R, cin, d;         // EPI_GATE gather
    int relu_in;
    // outputs
    float *out;            // EPI_GATE: z [N][H]; EPI_RELU: [N][M]; EPI_PLAIN: [N][M]; EPI_OUTSKIP: x' [N][R]
    float *skips;          // EPI_OUTSKIP: [N][S]
    int H, S;
    int first_layer, legacy, residual_legacy;
};

__device__ __forceinline__ float2 ffma2g(float2 a, float w, float2 c) {
    float2 d;
    const float2 ww = make_float2(w, w);
    asm("fma.rn.f32x2 %0, %1, %2, %3;"
        : "=l"(*reinterpret_cast<unsigned long long *>(&d))
        : "l"(*reinterpret_cast<const unsigned long long *>(&a)), "l"(*reinterpret_cast<const unsigned long long *>(&ww)),
          "l"(*reinterpret_cast<const unsigned long long *>(&c)));
    return d;
}

// same operations as wn_tanhf / wn_sigmoidf (include/wn_math.h), one instruction stream (see wn_loop_pipe.cu: gate_act)
__device__ __forceinline__ float gate_act_g(float x, bool sig) {
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

// Thread (ty, tx) of the 16 x 16 thread grid owns rows {4ty..4ty+3} u {64+4ty..64+4ty+3} and positions
// {4tx..4tx+3} u {64+4tx..64+4tx+3} of the tile: both shared-memory reads are then 16-byte words at a 16-byte stride
// (conflict free), and for the gate the two row groups are the tanh and the sigmoid half of the same 4 channels.
template <int EPI>
__global__ void __launch_bounds__(GT, 2) step_gemm_kernel(const GemmArgs a) {
    __shared__ __align__(16) float As[2][TK][TM];
    __shared__ __align__(16) float Bs[2][TK][TN];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const long long n0 = (long long)blockIdx.x * TN;
    const int m0 = blockIdx.y * TM;

    // loader roles: A - 2 float4 per thread (row kk = tid/32 and +8, 4 outputs at (tid%32)*4);
    //               B - position tid%128, k offset 8*(tid/128): 2 float4 along K
    const int a_kk = tid >> 5, a_m = (tid & 31) * 4;
    const int b_n = tid & 127, b_k = (tid >> 7) * 8;
    const long long n_b = n0 + b_n;
    const bool n_ok = n_b < a.N;
    const int t_b = n_ok ? (int)(n_b % a.T) : 0;

    float4 pa[2], pb[2];
    auto load_tiles = [&](int k0) {
        pa[0] = __ldg(reinterpret_cast<const float4 *>(a.W + (size_t)(k0 + a_kk) * a.Mpad + m0 + a_m));
        pa[1] = __ldg(reinterpret_cast<const float4 *>(a.W + (size_t)(k0 + a_kk + 8) * a.Mpad + m0 + a_m));
        const float *p = nullptr;
        if (n_ok) {
            if (EPI == EPI_GATE) {   // which segment of [x(t-2d) | x(t-d) | c(t) | x(t)] this K step lies in (uniform)
                const int R = a.R;
                if (k0 < 2 * R) {
                    const int tap = k0 / R, back = (2 - tap) * a.d;
                    if (t_b >= back) p = a.src + (n_b - back) * R + (k0 - tap * R) + b_k;
                } else if (k0 < 2 * R + a.cin) {
                    p = a.cond + n_b * a.cin + (k0 - 2 * R) + b_k;
                } else {
                    p = a.src + n_b * R + (k0 - 2 * R - a.cin) + b_k;
                }
            } else {
                p = a.src + n_b * a.K + k0 + b_k;
            }
        }
        pb[0] = p ? *reinterpret_cast<const float4 *>(p) : make_float4(0.f, 0.f, 0.f, 0.f);
        pb[1] = p ? *reinterpret_cast<const float4 *>(p + 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        if (a.relu_in) {
            pb[0].x = fmaxf(pb[0].x, 0.f); pb[0].y = fmaxf(pb[0].y, 0.f); pb[0].z = fmaxf(pb[0].z, 0.f); pb[0].w = fmaxf(pb[0].w, 0.f);
            pb[1].x = fmaxf(pb[1].x, 0.f); pb[1].y = fmaxf(pb[1].y, 0.f); pb[1].z = fmaxf(pb[1].z, 0.f); pb[1].w = fmaxf(pb[1].w, 0.f);
        }
    };
    auto store_tiles = [&](int buf) {
        *reinterpret_cast<float4 *>(&As[buf][a_kk][a_m]) = pa[0];
        *reinterpret_cast<float4 *>(&As[buf][a_kk + 8][a_m]) = pa[1];
        Bs[buf][b_k + 0][b_n] = pb[0].x; Bs[buf][b_k + 1][b_n] = pb[0].y; Bs[buf][b_k + 2][b_n] = pb[0].z; Bs[buf][b_k + 3][b_n] = pb[0].w;
        Bs[buf][b_k + 4][b_n] = pb[1].x; Bs[buf][b_k + 5][b_n] = pb[1].y; Bs[buf][b_k + 6][b_n] = pb[1].z; Bs[buf][b_k + 7][b_n] = pb[1].w;
    };

    float2 acc[8][4];   // [row i][position pair j]: positions (4tx, 4tx+1), (4tx+2, 4tx+3), (64+4tx, ..), (64+4tx+2, ..)
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = make_float2(0.f, 0.f);

    load_tiles(0);
    store_tiles(0);
    __syncthreads();
    const int nk = a.K / TK;
    for (int ks = 0; ks < nk; ++ks) {
        const int buf = ks & 1;
        if (ks + 1 < nk) load_tiles((ks + 1) * TK);
#pragma unroll
        for (int kk = 0; kk < TK; ++kk) {
            const float4 a0 = *reinterpret_cast<const float4 *>(&As[buf][kk][4 * ty]);
            const float4 a1 = *reinterpret_cast<const float4 *>(&As[buf][kk][64 + 4 * ty]);
            const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][kk][4 * tx]);
            const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[buf][kk][64 + 4 * tx]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float2 bv[4] = {make_float2(b0.x, b0.y), make_float2(b0.z, b0.w), make_float2(b1.x, b1.y), make_float2(b1.z, b1.w)};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = ffma2g(bv[j], av[i], acc[i][j]);
        }
        if (ks + 1 < nk) store_tiles(buf ^ 1);
        __syncthreads();
    }

    // ---- epilogues ------------------------------------------------------------------------------------------------
    // row of acc index i: m0 + (i < 4 ? 4ty + i : 64 + 4ty + i - 4); position of (j, half): n0 + (j < 2 ? 4tx + 2j : 64 + 4tx + 2(j-2)) + half
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const long long n = n0 + (j < 2 ? 4 * tx + 2 * j : 64 + 4 * tx + 2 * (j - 2)) + half;
            if (n >= a.N) continue;
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = half ? acc[i][j].y : acc[i][j].x;
            if (EPI == EPI_GATE) {
                // rows 4ty..4ty+3 of the tile: tanh half of channels ch0..ch0+3, rows 64+4ty..: their sigmoid half
                const int ch0 = blockIdx.y * 64 + 4 * ty;
                if (ch0 < a.H) {
                    float4 z;
                    float *zp = &z.x;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float ta = gate_act_g(__fadd_rn(v[i], a.bias[m0 + 4 * ty + i]), false);
                        const float sb = gate_act_g(__fadd_rn(v[4 + i], a.bias[m0 + 64 + 4 * ty + i]), true);
                        zp[i] = __fmul_rn(ta, sb);
                    }
                    *reinterpret_cast<float4 *>(a.out + n * a.H + ch0) = z;
                }
            } else {
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    const int m = m0 + (g ? 64 : 0) + 4 * ty;
                    if (m >= a.M) continue;
                    float4 y = make_float4(__fadd_rn(v[4 * g + 0], a.bias[m + 0]), __fadd_rn(v[4 * g + 1], a.bias[m + 1]),
                                           __fadd_rn(v[4 * g + 2], a.bias[m + 2]), __fadd_rn(v[4 * g + 3], a.bias[m + 3]));
                    if (EPI == EPI_OUTSKIP) {
                        if (m < a.R) {          // residual (modules.py:515-520)
                            const float4 xin = *reinterpret_cast<const float4 *>(a.cond + n * a.R + m);   // cond = layer input here
                            y.x = __fadd_rn(y.x, xin.x); y.y = __fadd_rn(y.y, xin.y); y.z = __fadd_rn(y.z, xin.z); y.w = __fadd_rn(y.w, xin.w);
                            if (a.residual_legacy) {
                                y.x = __fmul_rn(y.x, WN_SQRT_HALF); y.y = __fmul_rn(y.y, WN_SQRT_HALF);
                                y.z = __fmul_rn(y.z, WN_SQRT_HALF); y.w = __fmul_rn(y.w, WN_SQRT_HALF);
                            }
                            *reinterpret_cast<float4 *>(a.out + n * a.R + m) = y;
                        } else {                // skips (wavenet.py:833-836)
                            float4 *sp = reinterpret_cast<float4 *>(a.skips + n * a.S + (m - a.R));
                            if (!a.first_layer) {
                                const float4 s = *sp;
                                y.x = __fadd_rn(s.x, y.x); y.y = __fadd_rn(s.y, y.y); y.z = __fadd_rn(s.z, y.z); y.w = __fadd_rn(s.w, y.w);
                                if (a.legacy) {
                                    y.x = __fmul_rn(y.x, WN_SQRT_HALF); y.y = __fmul_rn(y.y, WN_SQRT_HALF);
                                    y.z = __fmul_rn(y.z, WN_SQRT_HALF); y.w = __fmul_rn(y.w, WN_SQRT_HALF);
                                }
                            }
                            *sp = y;
                        }
                    } else {
                        if (EPI == EPI_RELU) { y.x = fmaxf(y.x, 0.f); y.y = fmaxf(y.y, 0.f); y.z = fmaxf(y.z, 0.f); y.w = fmaxf(y.w, 0.f); }
                        float *op = a.out + n * a.M + m;
                        if (m + 3 < a.M && (a.M & 3) == 0) {
                            *reinterpret_cast<float4 *>(op) = y;
                        } else {
                            if (m + 0 < a.M) op[0] = y.x;
                            if (m + 1 < a.M) op[1] = y.y;
                            if (m + 2 < a.M) op[2] = y.z;
                            if (m + 3 < a.M) op[3] = y.w;
                        }
                    }
                }
            }
        }
    }
}

// first_conv on the teacher input (wavenet.py:826 / :702): x0[n][r] = x[n] * w[r] + b[r]
__global__ void step_first_conv_kernel(const float *__restrict__ x, const float *__restrict__ w, const float *__restrict__ b,
                                       long long N, int R, float *__restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N * (R / 4)) return;
    const long long n = i / (R / 4);
    const int r = (int)(i - n * (R / 4)) * 4;
    const float s = x[n];
    const float4 ww = *reinterpret_cast<const float4 *>(w + r), bb = *reinterpret_cast<const float4 *>(b + r);
    float4 o;
    o.x = __fadd_rn(__fmaf_rn(s, ww.x, 0.0f), bb.x); o.y = __fadd_rn(__fmaf_rn(s, ww.y, 0.0f), bb.y);
    o.z = __fadd_rn(__fmaf_rn(s, ww.z, 0.0f), bb.z); o.w = __fadd_rn(__fmaf_rn(s, ww.w, 0.0f), bb.w);
    *reinterpret_cast<float4 *>(out + n * R + r) = o;
}

}  // namespace

// ===============================================================================================================
// host side
// ===============================================================================================================
struct WnGemmPlan {
    float *d_w = nullptr;                 // all GEMM-layout weights
    std::vector<size_t> gate_w, gate_b, os_w, os_b;   // per layer offsets (floats)
    size_t h1_w = 0, h1_b = 0, h2_w = 0, h2_b = 0, first_w = 0, first_b = 0;
    int gate_mpad = 0, os_mpad = 0, h1_mpad = 0, h2_mpad = 0, Kcat = 0;
    float *d_x[2] = {nullptr, nullptr}, *d_z = nullptr, *d_skips = nullptr, *d_h1 = nullptr;
    size_t cap = 0;                        // positions the scratch holds
};

void wn_gemm_free(wn_handle *h) {
    WnGemmPlan *g = h->gemm;
    if (!g) return;
    if (g->d_w) cudaFree(g->d_w);
    for (float *p : {g->d_x[0], g->d_x[1], g->d_z, g->d_skips, g->d_h1})
        if (p) cudaFree(p);
    delete g;
    h->gemm = nullptr;
}

bool wn_gemm_supported(const wn_handle *h) {
    const wn_config &c = h->cfg;
    const WnModelDev &m = h->m;
    if (getenv("WN_STEP_SERIAL") && atoi(getenv("WN_STEP_SERIAL"))) return false;
    if (c.input_type == WN_INPUT_MULAW_QUANTIZE || m.ksize != 3 || m.cin <= 0 || m.gin > 0) return false;
    if (m.R % 16 || m.S % 16 || m.cin % 16 || m.H % 16) return false;
    return true;
}

static int gemm_plan(wn_handle *h) {
    if (h->gemm) return WN_OK;
    const WnModelDev &m = h->m;
    auto Wt = [&](const std::string &n) -> const float * {
        auto it = h->host_w.find(n);
        return it == h->host_w.end() ? nullptr : it->second.data.data();
    };
    auto pad = [](int v) { return (v + TM - 1) / TM * TM; };
    WnGemmPlan *g = new WnGemmPlan();
    const int R = m.R, S = m.S, H = m.H, G = 2 * H, cin = m.cin, O = m.O;
    g->Kcat = 3 * R + cin;
    // gate rows are tiled as 64 tanh rows + the 64 sigmoid rows of the same channels
    const int gate_tiles = (H + 63) / 64;
    g->gate_mpad = gate_tiles * TM; g->os_mpad = pad(R + S); g->h1_mpad = pad(S); g->h2_mpad = pad(O);
    std::vector<float> w;
    auto reserve = [&](size_t n) { size_t o = w.size(); w.resize(o + ((n + 3) & ~(size_t)3), 0.0f); return o; };
    for (int l = 0; l < m.L; ++l) {
        const std::string pre = "ResidualConv1DGLU_" + std::to_string(l) + "/residual_block_";
        const float *cw = Wt(pre + "causal_conv/kernel"), *cb = Wt(pre + "causal_conv/bias");
        const float *iw = Wt(pre + "cin_conv/kernel"), *ib = Wt(pre + "cin_conv/bias");
        const float *ow = Wt(pre + "out_conv/kernel"), *ob = Wt(pre + "out_conv/bias");
        const float *sw = Wt(pre + "skip_conv/kernel"), *sb = Wt(pre + "skip_conv/bias");
        const size_t gw = reserve((size_t)g->Kcat * g->gate_mpad), gb = reserve(g->gate_mpad);
        for (int tile = 0; tile < gate_tiles; ++tile)
            for (int j = 0; j < TM; ++j) {
                const int ch = tile * 64 + (j & 63);
                if (ch >= H) continue;
                const int col = (j < 64 ? 0 : H) + ch, mrow = tile * TM + j;   // modules.py:494 split: first half tanh
                for (int k = 0; k < 2 * R; ++k) w[gw + (size_t)k * g->gate_mpad + mrow] = cw[(size_t)k * G + col];
                for (int k = 0; k < cin; ++k) w[gw + (size_t)(2 * R + k) * g->gate_mpad + mrow] = iw[(size_t)k * G + col];
                for (int k = 0; k < R; ++k) w[gw + (size_t)(2 * R + cin + k) * g->gate_mpad + mrow] = cw[(size_t)(2 * R + k) * G + col];
                w[gb + mrow] = (cb ? cb[col] : 0.0f) + (ib ? ib[col] : 0.0f);
            }
        const size_t ow_o = reserve((size_t)H * g->os_mpad), ob_o = reserve(g->os_mpad);
        for (int k = 0; k < H; ++k) {
            for (int r = 0; r < R; ++r) w[ow_o + (size_t)k * g->os_mpad + r] = ow[(size_t)k * R + r];
            for (int s = 0; s < S; ++s) w[ow_o + (size_t)k * g->os_mpad + R + s] = sw[(size_t)k * S + s];
        }
        for (int r = 0; r < R; ++r) w[ob_o + r] = ob ? ob[r] : 0.0f;
        for (int s = 0; s < S; ++s) w[ob_o + R + s] = sb ? sb[s] : 0.0f;
        g->gate_w.push_back(gw); g->gate_b.push_back(gb); g->os_w.push_back(ow_o); g->os_b.push_back(ob_o);
    }
    {
        const float *w1 = Wt("final_convolution_1/kernel"), *b1 = Wt("final_convolution_1/bias");
        const float *w2 = Wt("final_convolution_2/kernel"), *b2 = Wt("final_convolution_2/bias");
        g->h1_w = reserve((size_t)S * g->h1_mpad); g->h1_b = reserve(g->h1_mpad);
        g->h2_w = reserve((size_t)S * g->h2_mpad); g->h2_b = reserve(g->h2_mpad);
        for (int k = 0; k < S; ++k) {
            for (int n = 0; n < S; ++n) w[g->h1_w + (size_t)k * g->h1_mpad + n] = w1[(size_t)k * S + n];
            for (int n = 0; n < O; ++n) w[g->h2_w + (size_t)k * g->h2_mpad + n] = w2[(size_t)k * O + n];
        }
        for (int n = 0; n < S; ++n) w[g->h1_b + n] = b1 ? b1[n] : 0.0f;
        for (int n = 0; n < O; ++n) w[g->h2_b + n] = b2 ? b2[n] : 0.0f;
        const float *f = Wt("input_convolution/kernel"), *fb = Wt("input_convolution/bias");
        g->first_w = reserve(R); g->first_b = reserve(R);
        for (int r = 0; r < R; ++r) { w[g->first_w + r] = f[r]; w[g->first_b + r] = fb ? fb[r] : 0.0f; }
    }
    if (cudaMalloc((void **)&g->d_w, w.size() * sizeof(float)) != cudaSuccess ||
        cudaMemcpy(g->d_w, w.data(), w.size() * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess) {
        h->err = std::string("step GEMM weights: ") + cudaGetErrorString(cudaGetLastError());
        if (g->d_w) cudaFree(g->d_w);
        delete g;
        return WN_ERR_NOMEM;
    }
    h->gemm = g;
    return WN_OK;
}

template <int EPI>
static cudaError_t launch_gemm(const GemmArgs &a, cudaStream_t st) {
    dim3 grid((unsigned)((a.N + TN - 1) / TN), (unsigned)(a.Mpad / TM));
    step_gemm_kernel<EPI><<<grid, GT, 0, st>>>(a);
    return cudaGetLastError();
}

int wn_launch_step_gemm(wn_handle *h, const float *x, const float *c_up, int B, int T, float *raw, cudaStream_t st) {
    int rc = gemm_plan(h);
    if (rc) return rc;
    WnGemmPlan *g = h->gemm;
    const WnModelDev &m = h->m;
    const long long N = (long long)B * T;
    if (g->cap < (size_t)N) {
        for (float **p : {&g->d_x[0], &g->d_x[1], &g->d_z, &g->d_skips, &g->d_h1}) {
            if (*p) cudaFree(*p);
            *p = nullptr;
        }
        g->cap = 0;
        const size_t sizes[5] = {(size_t)N * m.R, (size_t)N * m.R, (size_t)N * m.H, (size_t)N * m.S, (size_t)N * m.S};
        float **ptrs[5] = {&g->d_x[0], &g->d_x[1], &g->d_z, &g->d_skips, &g->d_h1};
        for (int i = 0; i < 5; ++i)
            if (cudaMalloc((void **)ptrs[i], sizes[i] * sizeof(float)) != cudaSuccess) {
                h->err = std::string("step GEMM scratch: ") + cudaGetErrorString(cudaGetLastError());
                return WN_ERR_NOMEM;
            }
        g->cap = (size_t)N;
    }
    {
        const long long n4 = N * (m.R / 4);
        step_first_conv_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>(x, g->d_w + g->first_w, g->d_w + g->first_b, N, m.R, g->d_x[0]);
        WN_CUDA_TRY(h, cudaGetLastError());
        h->launches++;
    }
    int cur = 0;
    const bool tensor = wn_umma_supported(h);
    for (int l = 0; l < m.L; ++l) {
        if (tensor) {   // tcgen05 path (wn_step_umma.cu)
            rc = wn_umma_layer(h, l, g->d_x[cur], c_up, g->d_z, g->d_x[cur ^ 1], g->d_skips, N, T, st);
            if (rc) return rc;
            cur ^= 1;
            continue;
        }
        GemmArgs a;
        memset(&a, 0, sizeof(a));
        a.N = N; a.T = T; a.R = m.R; a.cin = m.cin; a.d = m.dil[l]; a.H = m.H; a.S = m.S;
        a.legacy = m.legacy; a.residual_legacy = m.residual_legacy; a.first_layer = l == 0;
        // gate
        a.W = g->d_w + g->gate_w[l]; a.bias = g->d_w + g->gate_b[l]; a.M = 2 * m.H; a.Mpad = g->gate_mpad; a.K = g->Kcat;
        a.src = g->d_x[cur]; a.cond = c_up; a.out = g->d_z;
        WN_CUDA_TRY(h, launch_gemm<EPI_GATE>(a, st));
        // out / skip
        a.W = g->d_w + g->os_w[l]; a.bias = g->d_w + g->os_b[l]; a.M = m.R + m.S; a.Mpad = g->os_mpad; a.K = m.H;
        a.src = g->d_z; a.cond = g->d_x[cur]; a.out = g->d_x[cur ^ 1]; a.skips = g->d_skips;
        WN_CUDA_TRY(h, launch_gemm<EPI_OUTSKIP>(a, st));
        h->launches += 2;
        cur ^= 1;
    }
    if (tensor && wn_umma_head_supported(h)) {
        rc = wn_umma_head(h, g->d_skips, g->d_h1, raw, N, T, st);
        if (rc) return rc;
    } else {
        GemmArgs a;
        memset(&a, 0, sizeof(a));
        a.N = N; a.T = T;
        a.W = g->d_w + g->h1_w; a.bias = g->d_w + g->h1_b; a.M = m.S; a.Mpad = g->h1_mpad; a.K = m.S;
        a.src = g->d_skips; a.relu_in = 1; a.out = g->d_h1;
        WN_CUDA_TRY(h, launch_gemm<EPI_RELU>(a, st));
        a.W = g->d_w + g->h2_w; a.bias = g->d_w + g->h2_b; a.M = m.O; a.Mpad = g->h2_mpad; a.K = m.S;
        a.src = g->d_h1; a.relu_in = 0; a.out = raw;
        WN_CUDA_TRY(h, launch_gemm<EPI_PLAIN>(a, st));
        h->launches += 2;
    }
    return WN_OK;
}

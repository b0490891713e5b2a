// wn_upsample.cu - the conditioning upsample network (wavenet.py:781-803 -> modules.py:524-770) for sm_100a.
//
// HBM-bound: the last layer writes B*T*cin*4 bytes (320 B per audio sample for 80 mel bands) and reads 1/hop of
// that.  All tensors are time-major [B][W][C] so that the C (=80) frequency bands of one time column are one
// contiguous 320-byte run: every warp store is a run of full 128-bit vectors, and the loop kernels later read one
// column per step as a single coalesced 320-byte vector.  Each thread owns 4 adjacent bands of one input column,
// keeps its (kh x kw x 6) input window in registers and emits all `scale` output columns that depend on it, so the
// small input (L2 resident, it is hop-times smaller than the output) is read once per thread and the output is
// written exactly once with st.global.v4.  Activation (ReLU / LeakyReLU) and the sub-pixel shuffle
// (out[.., w*s+f] = y[.., w, f], modules.py:604-640) are fused into the store.
//
// Arithmetic order (identical in oracle/c/wn_oracle.c, so results are bit-equal): acc = 0; for i (frequency tap)
// for j (time tap): acc = fma(x, K, acc); y = act(acc + bias).
#include "wn_common.cuh"

namespace {

__device__ __forceinline__ float act_fn(float y, int kind, float alpha) {
    if (kind == WN_ACT_RELU) return y > 0.0f ? y : 0.0f;
    if (kind == WN_ACT_LEAKY) return y > 0.0f ? y : __fmul_rn(alpha, y);
    return y;
}


// SubPixelConvolution (modules.py:539-654): Conv2D SAME (kh x 3), s filters, periodic shuffle.
// Weights K[(i*3+j)*s + f] are staged transposed in shared memory as Ks[f][kh*3] (+bias) for broadcast reads.
template <int KH>
__global__ void __launch_bounds__(256) up_subpixel_kernel(const float *__restrict__ x, float *__restrict__ y,
                                                          const float *__restrict__ K, const float *__restrict__ bias,
                                                          int B, int W, int C, int s, int akind, float alpha) {
    extern __shared__ float sm[];
    constexpr int NT = KH * 3;
    float *Ks = sm;            // [s][NT]
    float *bs = sm + s * NT;   // [s]
    for (int i = threadIdx.x; i < s * NT; i += blockDim.x) {
        int f = i / NT, t = i % NT;
        Ks[i] = K[t * s + f];
    }
    for (int i = threadIdx.x; i < s; i += blockDim.x) bs[i] = bias[i];
    __syncthreads();
    const int C4 = C >> 2;
    const long long total = (long long)B * W * C4;
    constexpr int PT = (KH - 1) / 2;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int h4 = (int)(idx % C4);
        const long long bw = idx / C4;
        const int w = (int)(bw % W);
        const long long b = bw / W;
        const int h0 = h4 * 4;
        // window win[j][r]: time tap j (w-1..w+1), band h0-PT+r for r in [0, 4+KH-1)
        float win[3][4 + KH - 1];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const int ww = w + j - 1;
            const bool okw = (ww >= 0) && (ww < W);
            const float *col = x + ((long long)b * W + (okw ? ww : 0)) * C;
#pragma unroll
            for (int r = 0; r < 4 + KH - 1; ++r) {
                const int hh = h0 - PT + r;
                win[j][r] = (okw && hh >= 0 && hh < C) ? __ldg(col + hh) : 0.0f;
            }
        }
        float *out = y + ((long long)b * W + w) * s * C + h0;
        for (int f = 0; f < s; ++f) {
            const float *kf = Ks + f * NT;
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
            for (int i = 0; i < KH; ++i) {
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    const float kv = kf[i * 3 + j];
                    a0 = __fmaf_rn(win[j][i + 0], kv, a0);
                    a1 = __fmaf_rn(win[j][i + 1], kv, a1);
                    a2 = __fmaf_rn(win[j][i + 2], kv, a2);
                    a3 = __fmaf_rn(win[j][i + 3], kv, a3);
                }
            }
            const float bf = bs[f];
            float4 v;
            v.x = act_fn(__fadd_rn(a0, bf), akind, alpha);
            v.y = act_fn(__fadd_rn(a1, bf), akind, alpha);
            v.z = act_fn(__fadd_rn(a2, bf), akind, alpha);
            v.w = act_fn(__fadd_rn(a3, bf), akind, alpha);
            __stcs(reinterpret_cast<float4 *>(out + (long long)f * C), v);  // streaming store: written once, read much later
        }
    }
}

// ConvTranspose2D (modules.py:736-770): kernel (kh, s), stride (1, s): y[h][w*s+j] = act(b + sum_i x[h+pt-i][w] K[i][j]).
template <int KH>
__global__ void __launch_bounds__(256) up_convt2d_kernel(const float *__restrict__ x, float *__restrict__ y,
                                                         const float *__restrict__ K, const float *__restrict__ bias,
                                                         int B, int W, int C, int s, int akind, float alpha) {
    extern __shared__ float sm[];
    float *Ks = sm;  // [s][KH]
    for (int i = threadIdx.x; i < s * KH; i += blockDim.x) {
        int j = i / KH, t = i % KH;
        Ks[i] = K[t * s + j];
    }
    __syncthreads();
    const float b0 = __ldg(bias);
    const int C4 = C >> 2;
    const long long total = (long long)B * W * C4;
    constexpr int PT = (KH - 1) / 2;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int h4 = (int)(idx % C4);
        const long long bw = idx / C4;
        const int h0 = h4 * 4;
        const float *col = x + bw * C;
        // band h0 + PT - i + q, i in [0,KH), q in [0,4): r = PT - i + q + (KH-1-PT) in [0, 4+KH-1)
        float win[4 + KH - 1];
#pragma unroll
        for (int r = 0; r < 4 + KH - 1; ++r) {
            const int hh = h0 - (KH - 1 - PT) + r;
            win[r] = (hh >= 0 && hh < C) ? __ldg(col + hh) : 0.0f;
        }
        float *out = y + bw * s * C + h0;
        for (int j = 0; j < s; ++j) {
            const float *kj = Ks + j * KH;
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
            for (int i = 0; i < KH; ++i) {
                const float kv = kj[i];
                const int r = PT - i + (KH - 1 - PT);
                a0 = __fmaf_rn(win[r + 0], kv, a0);
                a1 = __fmaf_rn(win[r + 1], kv, a1);
                a2 = __fmaf_rn(win[r + 2], kv, a2);
                a3 = __fmaf_rn(win[r + 3], kv, a3);
            }
            float4 v;
            v.x = act_fn(__fadd_rn(a0, b0), akind, alpha);
            v.y = act_fn(__fadd_rn(a1, b0), akind, alpha);
            v.z = act_fn(__fadd_rn(a2, b0), akind, alpha);
            v.w = act_fn(__fadd_rn(a3, b0), akind, alpha);
            __stcs(reinterpret_cast<float4 *>(out + (long long)j * C), v);
        }
    }
}

// Scalar fall-back for any kh / C not divisible by 4, and for the variants no shipped config uses
// (ConvTranspose1D modules.py:697-733, ResizeConvolution :657-695, NearestNeighborUpsample :524-536).
// One thread per output element, bands fastest so stores stay coalesced.
__global__ void __launch_bounds__(256) up_generic_kernel(const float *__restrict__ x, float *__restrict__ y,
                                                         const float *__restrict__ K, const float *__restrict__ bias,
                                                         int B, int W, int C, int s, int kh, int type, int akind,
                                                         float alpha) {
    const int Wo = W * s;
    const long long total = (long long)B * Wo * C;
    const int pt = (kh - 1) / 2;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int h = (int)(idx % C);
        const long long bu = idx / C;
        const int u = (int)(bu % Wo);
        const long long b = bu / Wo;
        const float *xb = x + b * W * C;
        float acc = 0.0f, bv = 0.0f;
        if (type == WN_UP_SUBPIXEL) {
            const int w = u / s, f = u % s;
            for (int i = 0; i < kh; ++i) {
                const int hh = h + i - pt;
                if (hh < 0 || hh >= C) continue;
                for (int j = 0; j < 3; ++j) {
                    const int ww = w + j - 1;
                    if (ww < 0 || ww >= W) continue;
                    acc = __fmaf_rn(__ldg(xb + (long long)ww * C + hh), __ldg(K + (i * 3 + j) * s + f), acc);
                }
            }
            bv = __ldg(bias + f);
        } else if (type == WN_UP_2D) {
            const int w = u / s, j = u % s;
            for (int i = 0; i < kh; ++i) {
                const int hh = h + pt - i;
                if (hh < 0 || hh >= C) continue;
                acc = __fmaf_rn(__ldg(xb + (long long)w * C + hh), __ldg(K + i * s + j), acc);
            }
            bv = __ldg(bias);
        } else if (type == WN_UP_1D) {
            const int w = u / s, j = u % s;
            const float *kr = K + ((long long)j * C + h) * C;
            for (int ci = 0; ci < C; ++ci) acc = __fmaf_rn(__ldg(xb + (long long)w * C + ci), __ldg(kr + ci), acc);
            bv = __ldg(bias + h);
        } else if (type == WN_UP_RESIZE) {
            const int pl = (s - 1) / 2;
            const float scale = __fdiv_rn((float)W, (float)Wo);
            for (int i = 0; i < kh; ++i) {
                const int hh = h + i - pt;
                if (hh < 0 || hh >= C) continue;
                for (int j = 0; j < s; ++j) {
                    const int uu = u + j - pl;
                    if (uu < 0 || uu >= Wo) continue;
                    int src = (int)floorf(__fmul_rn((float)uu, scale));
                    src = src > W - 1 ? W - 1 : src;
                    acc = __fmaf_rn(__ldg(xb + (long long)src * C + hh), __ldg(K + i * s + j), acc);
                }
            }
            bv = __ldg(bias);
        } else {  // WN_UP_NEAREST: pure gather, no weights, no activation
            const float scale = __fdiv_rn((float)W, (float)Wo);
            int src = (int)floorf(__fmul_rn((float)u, scale));
            src = src > W - 1 ? W - 1 : src;
            y[idx] = __ldg(xb + (long long)src * C + h);
            continue;
        }
        y[idx] = act_fn(__fadd_rn(acc, bv), akind, alpha);
    }
}

int grid_for(long long work_items, int block, int sm_count) {
    // One CTA per 256 consecutive work items, launched in index order: the set of resident CTAs then covers one
    // contiguous window of the output (a few hundred MB at most), which keeps DRAM pages and the 256 MB TLB reach
    // hot.  (A persistent grid-stride loop spread the resident threads over the whole multi-GB output and measured
    // 1.6 TB/s in round 1.)  The grid-stride loops in the kernels remain as a guard for > 2^31-1 CTAs.
    (void)sm_count;
    long long blocks = (work_items + block - 1) / block;
    if (blocks > 0x7fffffffLL) blocks = 0x7fffffffLL;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace

int wn_launch_upsample(wn_handle *h, const float *c, int B, int Tc, float *c_up, cudaStream_t st) {
    const wn_config &cfg = h->cfg;
    const int C = cfg.cin_channels;
    if (C <= 0) { h->err = "wn_upsample: local conditioning is disabled (cin_channels <= 0)"; return WN_ERR_INVALID; }
    const int akind = cfg.upsample_activation;
    const float alpha = cfg.leaky_alpha;
    const int kh = cfg.freq_axis_kernel_size;
    if (cfg.upsample_type == WN_UP_NEAREST) {
        const long long total = (long long)B * Tc * cfg.hop_size * C;
        up_generic_kernel<<<grid_for(total, 256, h->sm_count), 256, 0, st>>>(c, c_up, nullptr, nullptr, B, Tc, C,
                                                                            cfg.hop_size, 1, WN_UP_NEAREST, 0, 0.f);
        h->launches++;
        WN_CUDA_TRY(h, cudaGetLastError());
        return WN_OK;
    }
    // intermediates ping-pong through two scratch buffers; the last layer writes c_up
    size_t max_mid = 0;
    {
        long long W = Tc;
        for (int i = 0; i + 1 < h->n_up; ++i) { W *= h->up[i].scale; max_mid = (size_t)B * W * C; }
    }
    if (max_mid > 0) {
        if (wn_ensure(h, &h->d_tmp0, &h->tmp_floats, max_mid)) return WN_ERR_NOMEM;
        if (h->n_up > 2 && wn_ensure(h, &h->d_tmp1, &h->tmp1_floats, max_mid)) return WN_ERR_NOMEM;   // ping-pong partner
    }
    const float *src = c;
    int W = Tc;
    for (int i = 0; i < h->n_up; ++i) {
        const int s = h->up[i].scale;
        float *dst = (i == h->n_up - 1) ? c_up : ((i & 1) ? h->d_tmp1 : h->d_tmp0);
        const bool vec = (C % 4 == 0) && (kh == 3);
        if (cfg.upsample_type == WN_UP_SUBPIXEL && vec) {
            const long long items = (long long)B * W * (C / 4);
            const size_t smem = sizeof(float) * (size_t)(s * 9 + s);
            up_subpixel_kernel<3><<<grid_for(items, 256, h->sm_count), 256, smem, st>>>(src, dst, h->up[i].w, h->up[i].b,
                                                                                       B, W, C, s, akind, alpha);
        } else if (cfg.upsample_type == WN_UP_2D && vec) {
            const long long items = (long long)B * W * (C / 4);
            const size_t smem = sizeof(float) * (size_t)(s * 3);
            up_convt2d_kernel<3><<<grid_for(items, 256, h->sm_count), 256, smem, st>>>(src, dst, h->up[i].w, h->up[i].b,
                                                                                      B, W, C, s, akind, alpha);
        } else {
            const long long total = (long long)B * W * s * C;
            up_generic_kernel<<<grid_for(total, 256, h->sm_count), 256, 0, st>>>(src, dst, h->up[i].w, h->up[i].b, B, W,
                                                                                C, s, kh, cfg.upsample_type, akind, alpha);
        }
        h->launches++;
        WN_CUDA_TRY(h, cudaGetLastError());
        src = dst;
        W *= s;
    }
    return WN_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// wn_upsample_static: the upsampling network for callers that may not allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same kernels, same arithmetic, same bits as wn_upsample; the only difference is where the intermediate
// layers live: in a file-scope device array of fixed size instead of the handle's grown-on-demand scratch.
//
// * The array is WN_STATIC_SCRATCH_FLOATS floats (16 MiB); every GPU that loads the library holds it until exit.
// * In host code the array's name is only a host-side stand-in, so the device address is looked up once per device
//   with cudaGetSymbolAddress (with that device current) and cached; kernels only ever get that address.
// * Every call shares the array and it keeps the previous call's contents.  Nothing has to start at zero: each layer
//   overwrites every element the next layer reads.  Two calls that overlap on the GPU (different streams, or two
//   handles on one device) race on it: serialise them.
// * This function calls no cudaMalloc / cudaFree and no allocating library routine.
static __device__ __align__(256) float g_up_static_scratch[WN_STATIC_SCRATCH_FLOATS];

namespace {
constexpr int kMaxDevices = 64;
float *g_static_scratch_addr[kMaxDevices] = {};   // device address of g_up_static_scratch, per device, 0 = not looked up

// One layer of the network, identical dispatch to the loop in wn_launch_upsample.
cudaError_t launch_up_layer(const wn_handle *h, int i, const float *src, float *dst, int B, int W, cudaStream_t st) {
    const wn_config &cfg = h->cfg;
    const int C = cfg.cin_channels, kh = cfg.freq_axis_kernel_size, akind = cfg.upsample_activation;
    const float alpha = cfg.leaky_alpha;
    const int s = h->up[i].scale;
    const bool vec = (C % 4 == 0) && (kh == 3);
    if (cfg.upsample_type == WN_UP_SUBPIXEL && vec) {
        const long long items = (long long)B * W * (C / 4);
        const size_t smem = sizeof(float) * (size_t)(s * 9 + s);
        up_subpixel_kernel<3><<<grid_for(items, 256, h->sm_count), 256, smem, st>>>(src, dst, h->up[i].w, h->up[i].b, B, W,
                                                                                   C, s, akind, alpha);
    } else if (cfg.upsample_type == WN_UP_2D && vec) {
        const long long items = (long long)B * W * (C / 4);
        const size_t smem = sizeof(float) * (size_t)(s * 3);
        up_convt2d_kernel<3><<<grid_for(items, 256, h->sm_count), 256, smem, st>>>(src, dst, h->up[i].w, h->up[i].b, B, W,
                                                                                  C, s, akind, alpha);
    } else {
        const long long total = (long long)B * W * s * C;
        up_generic_kernel<<<grid_for(total, 256, h->sm_count), 256, 0, st>>>(src, dst, h->up[i].w, h->up[i].b, B, W, C, s,
                                                                            kh, cfg.upsample_type, akind, alpha);
    }
    return cudaGetLastError();
}
}  // namespace

extern "C" int wn_upsample_static(wn_handle *h, const float *c, int B, int Tc, float *c_up, void *stream) {
    if (!h) return WN_ERR_INVALID;
    if (!h->finalized) { h->err = "wn_upsample_static: model not finalized"; return WN_ERR_STATE; }
    if (!c || !c_up || B <= 0 || Tc <= 0) { h->err = "wn_upsample_static: null buffer or non-positive size"; return WN_ERR_INVALID; }
    const wn_config &cfg = h->cfg;
    const int C = cfg.cin_channels;
    if (C <= 0) { h->err = "wn_upsample_static: local conditioning is disabled (cin_channels <= 0)"; return WN_ERR_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    int dev = -1;
    WN_CUDA_TRY(h, cudaGetDevice(&dev));
    if (dev != h->device) {
        h->err = "wn_upsample_static: the handle was created on device " + std::to_string(h->device) +
                 " but device " + std::to_string(dev) + " is current";
        return WN_ERR_INVALID;
    }
    if (dev < 0 || dev >= kMaxDevices) { h->err = "wn_upsample_static: device ordinal out of range"; return WN_ERR_INVALID; }
    if (cfg.upsample_type == WN_UP_NEAREST) {   // one gather, no intermediate
        const long long total = (long long)B * Tc * cfg.hop_size * C;
        up_generic_kernel<<<grid_for(total, 256, h->sm_count), 256, 0, st>>>(c, c_up, nullptr, nullptr, B, Tc, C,
                                                                            cfg.hop_size, 1, WN_UP_NEAREST, 0, 0.f);
        h->launches++;
        WN_CUDA_TRY(h, cudaGetLastError());
        return WN_OK;
    }
    // size check: the largest intermediate (the input of the last layer); networks of three or more layers ping-pong
    // between the two halves of the array
    size_t max_mid = 0;
    {
        long long W = Tc;
        for (int i = 0; i + 1 < h->n_up; ++i) { W *= h->up[i].scale; max_mid = (size_t)B * W * C; }
    }
    const size_t part = h->n_up > 2 ? WN_STATIC_SCRATCH_FLOATS / 2 : WN_STATIC_SCRATCH_FLOATS;
    if (max_mid > part) {
        h->err = "wn_upsample_static: the intermediate layer needs " + std::to_string(max_mid) +
                 " floats but the static scratch holds " + std::to_string(part) + " per buffer (" +
                 std::to_string(WN_STATIC_SCRATCH_FLOATS) + " floats = 16 MiB in total); use wn_upsample, or a smaller batch";
        return WN_ERR_INVALID;
    }
    if (max_mid > 0 && !g_static_scratch_addr[dev]) {   // set-up, once per device, with that device current
        float *p = nullptr;
        WN_CUDA_TRY(h, cudaGetSymbolAddress((void **)&p, g_up_static_scratch));
        if (!p) { h->err = "wn_upsample_static: cudaGetSymbolAddress returned a null address"; return WN_ERR_CUDA; }
        g_static_scratch_addr[dev] = p;
    }
    float *scratch = g_static_scratch_addr[dev];
    const float *src = c;
    int W = Tc;
    for (int i = 0; i < h->n_up; ++i) {
        float *dst = (i == h->n_up - 1) ? c_up : scratch + ((i & 1) ? part : 0);
        WN_CUDA_TRY(h, launch_up_layer(h, i, src, dst, B, W, st));
        h->launches++;
        src = dst;
        W *= h->up[i].scale;
    }
    return WN_OK;
}

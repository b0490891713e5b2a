// Development aid: clock64 cost of the building blocks of the pipelined layer role, run alone in one CTA of 256 threads
// with operands in shared memory (same code: this file includes the engine's translation unit).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --fmad=false -Iinclude -o build/phase_bench tools/phase_bench.cu \
//        tacotron-2_b200/csrc/wn_api.cu tacotron-2_b200/csrc/wn_upsample.cu tacotron-2_b200/csrc/wn_loop_generic.cu tacotron-2_b200/csrc/wn_step_full.cu
#include "../tacotron-2_b200/csrc/wn_loop_pipe.cu"

namespace {
constexpr int NU = 4, R = 128, S = 128, CIN = 80, KC = 3 * R + CIN, KC4 = KC / 4, C0 = (2 * R + CIN) / 4, NR = 64, HK = 32;

template <int WHAT, int NTHREADS_ACTIVE, int ROWS>
__global__ void __launch_bounds__(256, 1) bench(long long *out, float *sink, int iters) {
    extern __shared__ __align__(16) float sm[];
    float *Wcat = sm, *Wos = Wcat + NR * KC, *xc = Wos + (R + S) * HK, *zs = xc + KC * NU;
    const int tid = threadIdx.x, lane = tid & (LANES - 1), q = tid / LANES;
    for (int i = tid; i < NR * KC + (R + S) * HK + KC * NU + HK * NU; i += 256) sm[i] = 1e-3f * (float)((i * 37) % 101 - 50);
    __syncthreads();
    const float4 *wq = reinterpret_cast<const float4 *>(Wcat) + (size_t)4 * q * KC4;
    const int c1_first = C0 + ((lane - C0) % LANES + LANES) % LANES;
    float keep = 0.f;
    long long total = 0;
    for (int it = 0; it < iters; ++it) {
        __syncthreads();
        const long long t0 = clock64();
        if (WHAT == 0) {   // current-tap mac (weights from shared memory)
            float2 acc[8];
            for (int i = 0; i < 8; ++i) acc[i] = make_float2(keep, 0.f);
            mac_lane<NU, 4>(acc, wq, KC4, xc, c1_first, KC4);
            for (int i = 0; i < 8; ++i) keep += acc[i].x + acc[i].y;
        } else if (WHAT == 1) {   // phase0 mac
            float2 acc[8];
            for (int i = 0; i < 8; ++i) acc[i] = make_float2(keep, 0.f);
            mac_lane<NU, 4>(acc, wq, KC4, xc, lane, C0);
            for (int i = 0; i < 8; ++i) keep += acc[i].x + acc[i].y;
        } else if (WHAT == 2) {   // reduce16 of 16 values
            float a[16];
            for (int i = 0; i < 16; ++i) a[i] = keep + (float)i;
            reduce16<16>(a, lane);
            keep += a[0];
        } else if (WHAT == 3) {   // gate transcendental + partner shuffle
            const float pv = keep * 0.001f + 0.01f * (float)(tid % 97) - 0.4f;
            const float g_act = (tid & 4) ? wn_sigmoidf(pv) : wn_tanhf(pv);
            const float o_act = __shfl_xor_sync(0xffffffffu, g_act, 2);
            keep += g_act * o_act;
        } else if (WHAT == 5) {   // merged gate activation
            const float pv = keep * 0.001f + 0.01f * (float)(tid % 97) - 0.4f;
            const float g_act = gate_act(pv, (tid & 4) != 0);
            const float o_act = __shfl_xor_sync(0xffffffffu, g_act, 2);
            keep += g_act * o_act;
        } else if (WHAT == 6) {   // current-tap mac, weights in registers
            float4 w1[4][2];
            for (int i = 0; i < 4; ++i) for (int j = 0; j < 2; ++j) w1[i][j] = wq[i * KC4 + c1_first + j * LANES];
            float2 acc[8];
            for (int i = 0; i < 8; ++i) acc[i] = make_float2(keep, 0.f);
            mac_lane_wreg<NU, 4, 2>(acc, w1, KC4, xc, c1_first);
            for (int i = 0; i < 8; ++i) keep += acc[i].x + acc[i].y;
        } else if (WHAT == 4) {   // blocked chains, ROWS rows per thread on NTHREADS_ACTIVE threads
            if (tid < NTHREADS_ACTIVE) {
                float2 a2[ROWS * NU / 2];
                for (int i = 0; i < ROWS * NU / 2; ++i) a2[i] = make_float2(keep, 0.f);
                chain_tile<NU, ROWS>(a2, reinterpret_cast<const float4 *>(Wos), R + S, tid, NTHREADS_ACTIVE, zs, HK / 4);
                for (int i = 0; i < ROWS * NU / 2; ++i) keep += a2[i].x + a2[i].y;
            }
        }
        const long long t1 = clock64();
        if (tid == 0) total += t1 - t0;
    }
    if (tid == 0) out[0] = total;
    sink[blockIdx.x * 256 + tid] = keep;
}

template <int WHAT, int NA, int ROWS>
void run(const char *name, long long *out, float *sink) {
    const int iters = 2000;
    const size_t smem = sizeof(float) * (NR * KC + (R + S) * HK + KC * NU + HK * NU);
    cudaFuncSetAttribute(bench<WHAT, NA, ROWS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    long long h = 0;
    for (int rep = 0; rep < 2; ++rep) {
        bench<WHAT, NA, ROWS><<<1, 256, smem>>>(out, sink, iters);
        cudaError_t e = cudaDeviceSynchronize();
        cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
        if (rep) printf("%-56s %7.0f cycles (%s)\n", name, (double)h / iters, cudaGetErrorString(e));
    }
}
}  // namespace

int main() {
    long long *out; float *sink;
    cudaMalloc(&out, 8); cudaMalloc(&sink, 4 * 256);
    run<0, 256, 1>("current-tap mac (2 chunks/lane, smem weights)", out, sink);
    run<1, 256, 1>("phase0 mac (5.25 chunks/lane)", out, sink);
    run<2, 256, 1>("reduce16 of 16 values", out, sink);
    run<3, 256, 1>("gate (tanh | sigmoid, divergent lanes) + shuffle", out, sink);
    run<5, 256, 1>("gate, one instruction stream (gate_act) + shuffle", out, sink);
    run<6, 256, 1>("current-tap mac, weights loaded to registers first", out, sink);
    run<4, 128, 2>("chains: 128 threads x 2 rows", out, sink);
    run<4, 256, 1>("chains: 256 threads x 1 row", out, sink);
    run<4, 64, 4>("chains: 64 threads x 4 rows", out, sink);
    return 0;
}

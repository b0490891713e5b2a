// tc_probe3.cu - development aid: tcgen05.mma issue rate when the issuing code is written the way the loop engine needs
// it: a whole warp runs the (warp-uniform) issue code, one elected lane executes the MMA, every operand address is
// base + compile-time constant (fully unrolled).  NM MMAs (M = 128, N columns) over NACC accumulators.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/tc_probe3 tools/tc_probe3.cu
#include <cstdio>
#include <cstdlib>
#include "../tacotron-2_b200/csrc/wn_umma.cuh"

constexpr int M = 128, K = 128;
constexpr int COL_A = 448;

__device__ __forceinline__ void mma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.b32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred != 0;
}

template <int N, int NM, int NACC, int MODE>
__global__ void __launch_bounds__(192, 1) probe_kernel(long long *clk, int iters) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int LBO_B = N * 16 + 16;
    unsigned char *As = smem, *Bs = smem + M * K * 2;
    __shared__ __align__(8) unsigned long long bars[1];
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid / 32;
    const uint32_t done = umma::smem_addr(&bars[0]);
    if (tid == 0) { umma::mbar_init(done, 1); umma::mbar_fence_init(); }
    if (warp == 0) umma::tmem_alloc(umma::smem_addr(&tmem_base_s), 512);
    for (int i = tid; i < (M * K * 2 + (K / 8) * LBO_B) / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u;
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = tmem_base_s;
    constexpr uint32_t idesc = umma::idesc_bf16(M, N);
    const uint64_t bdesc0 = umma::smem_desc(umma::smem_addr(Bs), LBO_B, 128);
    const uint64_t adesc0 = umma::smem_desc(umma::smem_addr(As), M * 16, 128);
    long long t_issue = 0, t_done = 0, t_wake = 0;
    for (int it = 0; it < iters; ++it) {
        __syncthreads();
        const long long t0 = clock64();
        if (warp == 4) {
            if (elect_one()) {
#pragma unroll
                for (int i = 0; i < NM; ++i) {
                    constexpr int dummy = 0; (void)dummy;
                    const int ks = (i / NACC) % (K / 16), acc = i % NACC;
                    const uint64_t b = bdesc0 + (uint64_t)((ks * 2 * LBO_B) >> 4);
                    if (MODE == 0) mma_bf16_ts(tmem + acc * N, tmem + COL_A + ks * 8, b, idesc, i >= NACC);
                    else umma::mma_bf16(tmem + acc * N, adesc0 + (uint64_t)((ks * 2 * M * 16) >> 4), b, idesc, i >= NACC);
                }
                umma::mma_commit(done);
            }
            __syncwarp();
            t_issue += clock64() - t0;
            umma::mbar_wait(done, it & 1);
            t_done += clock64() - t0;
        } else if (tid == 0) {
            umma::mbar_wait(done, it & 1);
            t_wake += clock64() - t0;
        }
        umma::fence_after_sync();
    }
    __syncthreads();
    if (tid == 128) { clk[0] = t_issue; clk[1] = t_done; }
    if (tid == 0) clk[2] = t_wake;
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, 512);
}

template <int N, int NM, int NACC, int MODE>
void run(long long *dclk) {
    const int iters = 64;
    const int smem = M * K * 2 + (K / 8) * (N * 16 + 16);
    cudaFuncSetAttribute(probe_kernel<N, NM, NACC, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    probe_kernel<N, NM, NACC, MODE><<<1, 192, smem>>>(dclk, iters);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("N %d nm %d nacc %d: %s\n", N, NM, NACC, cudaGetErrorString(e)); exit(1); }
    long long clk[4];
    cudaMemcpy(clk, dclk, 32, cudaMemcpyDeviceToHost);
    printf("A in %s  N %3d  %2d MMAs over %2d accumulators: issue %6.0f  done %6.0f  wake %6.0f cycles\n", MODE ? "smem" : "TMEM", N, NM, NACC,
           (double)clk[0] / iters, (double)clk[1] / iters, (double)clk[2] / iters);
}

int main() {
    long long *dclk;
    cudaMalloc(&dclk, 64);
    run<16, 0, 1, 0>(dclk); run<16, 1, 1, 0>(dclk); run<16, 2, 1, 0>(dclk); run<16, 4, 1, 0>(dclk); run<16, 8, 1, 0>(dclk);
    run<16, 24, 1, 0>(dclk); run<16, 24, 2, 0>(dclk); run<16, 24, 3, 0>(dclk); run<16, 24, 6, 0>(dclk); run<16, 24, 24, 0>(dclk);
    run<16, 48, 1, 0>(dclk); run<16, 48, 6, 0>(dclk); run<16, 72, 6, 0>(dclk);
    run<32, 24, 1, 0>(dclk); run<32, 24, 6, 0>(dclk); run<32, 48, 6, 0>(dclk);
    run<64, 24, 1, 0>(dclk); run<64, 24, 6, 0>(dclk); run<64, 48, 6, 0>(dclk);
    run<128, 24, 1, 0>(dclk); run<128, 24, 3, 0>(dclk); run<256, 8, 1, 0>(dclk);
    run<16, 24, 1, 1>(dclk); run<16, 24, 6, 1>(dclk); run<16, 48, 6, 1>(dclk); run<32, 48, 6, 1>(dclk); run<64, 48, 6, 1>(dclk); run<256, 8, 1, 1>(dclk);
    return 0;
}

// tc_probe2.cu - development aid: how tcgen05.mma (M = 128, small N) pipelines.  Times NM MMAs issued by one thread,
// spread round-robin over NACC independent TMEM accumulators (NACC = 1: one dependent chain), A from tensor memory or
// shared memory, plus the completion latency of a single MMA and of an empty commit.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/tc_probe2 tools/tc_probe2.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../tacotron-2_b200/csrc/wn_umma.cuh"

constexpr int M = 128, K = 128;
constexpr int COL_A = 448;   // 64 columns of A (bf16 pairs); accumulators from column 0

__device__ __forceinline__ void mma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}

// mode: 0 A in TMEM, 1 A in smem.  nm MMAs over nacc accumulators (each N columns).  acc0: first MMA of a chain overwrites.
__global__ void __launch_bounds__(192, 1) probe_kernel(long long *clk, int N, int mode, int nm, int nacc, int iters) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int LBO_B = N * 16 + 16;
    unsigned char *As = smem, *Bs = smem + M * K * 2;
    __shared__ __align__(8) unsigned long long bars[1];
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid / 32;
    const uint32_t done = umma::smem_addr(&bars[0]);
    if (tid == 0) { umma::mbar_init(done, 1); umma::mbar_fence_init(); }
    if (warp == 0) umma::tmem_alloc(umma::smem_addr(&tmem_base_s), 512);
    for (int i = tid; i < (M * K * 2 + (K / 8) * LBO_B) / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u;  // small bf16 values
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = tmem_base_s;
    const uint32_t idesc = umma::idesc_bf16(M, N);
    long long t_issue = 0, t_done = 0, t_wake = 0;
    for (int it = 0; it < iters; ++it) {
        __syncthreads();
        const long long t0 = clock64();
        if (tid == 128) {
            for (int i = 0; i < nm; ++i) {
                const int ks = (i / nacc) % (K / 16), acc = i % nacc;
                const uint64_t b = umma::smem_desc(umma::smem_addr(Bs) + ks * 2 * LBO_B, LBO_B, 128);
                if (mode == 0) mma_bf16_ts(tmem + acc * N, tmem + COL_A + ks * 8, b, idesc, i >= nacc);
                else umma::mma_bf16(tmem + acc * N, umma::smem_desc(umma::smem_addr(As) + ks * 2 * M * 16, M * 16, 128), b, idesc, i >= nacc);
            }
            umma::mma_commit(done);
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

int main() {
    long long *dclk;
    cudaMalloc(&dclk, 64);
    const int iters = 64;
    const int smem = M * K * 2 + (K / 8) * (64 * 16 + 16);
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    struct Cfg { int N, nm, nacc; };
    const Cfg cfgs[] = {{16, 0, 1},  {16, 1, 1},  {16, 2, 1},  {16, 4, 1},  {16, 8, 1},  {16, 24, 1}, {16, 24, 2}, {16, 24, 3},
                        {16, 24, 4}, {16, 24, 6}, {16, 24, 8}, {16, 24, 12}, {16, 24, 24}, {16, 48, 6}, {16, 48, 12}, {16, 72, 6},
                        {32, 24, 1}, {32, 24, 6}, {32, 48, 6}, {32, 48, 12}, {64, 24, 1}, {64, 24, 6}, {64, 48, 6}, {128, 24, 1}, {128, 24, 3}, {256, 24, 1}};
    for (int mode = 0; mode < 2; ++mode)
        for (const Cfg &c : cfgs) {
            if (c.nacc * c.N > COL_A) continue;
            probe_kernel<<<1, 192, smem>>>(dclk, c.N, mode, c.nm, c.nacc, iters);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("N %d nm %d nacc %d: %s\n", c.N, c.nm, c.nacc, cudaGetErrorString(e)); return 1; }
            long long clk[4];
            cudaMemcpy(clk, dclk, 32, cudaMemcpyDeviceToHost);
            printf("A in %s  N %3d  %2d MMAs over %2d accumulators: issue %6.0f  done %6.0f  wake %6.0f cycles\n", mode ? "smem" : "TMEM", c.N,
                   c.nm, c.nacc, (double)clk[0] / iters, (double)clk[1] / iters, (double)clk[2] / iters);
        }
    return 0;
}

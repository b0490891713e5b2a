// tc_probe.cu - development aid for the tensor-core loop engine (csrc/wn_loop_tc.cu): checks, on one CTA,
//   (1) tcgen05.mma with the A operand (a 128 x 128 weight tile, bf16 hi / lo) RESIDENT IN TENSOR MEMORY, written there
//       once with tcgen05.st.32x32b (lane = row, one 32-bit column = two consecutive k, low half = even k),
//   (2) the B operand (activations [K = 128 channels] x [N utterances], N = 16 / 32 / 64) written by one thread per
//       channel with 2-byte stores into the K-major no-swizzle image of wn_umma.cuh (LBO = N * 16 + 16 bytes),
//   (3) three passes hi*hi + hi*lo + lo*hi against a double reference, also with A taken from shared memory,
//   and times the issue of the 24 MMAs, their completion (tcgen05.commit -> mbarrier) and the wake-up of a waiting warp.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/tc_probe tools/tc_probe.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../tacotron-2_b200/csrc/wn_umma.cuh"

constexpr int M = 128, K = 128;
constexpr int COL_D = 0, COL_D2 = 256, COL_AHI = 64, COL_ALO = 128;   // tensor-memory columns

__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
                   "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// D[tmem] (+)= A[tmem] * B[smem]^T
__device__ __forceinline__ void mma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ uint32_t bf16_bits(float f) {   // round to nearest even
    return (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(f));
}

__global__ void __launch_bounds__(192, 1) probe_kernel(const float *A, const float *B, float *D, float *D2, long long *clk, int N, int swap, int iters) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int LBO_B = N * 16 + 16;
    unsigned char *As_hi = smem, *As_lo = smem + M * K * 2, *Bs_hi = smem + 2 * M * K * 2, *Bs_lo = Bs_hi + (K / 8) * LBO_B;
    __shared__ __align__(8) unsigned long long bars[2];
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid / 32, lane = tid % 32;
    const uint32_t done = umma::smem_addr(&bars[0]);
    if (tid == 0) { umma::mbar_init(done, 1); umma::mbar_fence_init(); }
    if (warp == 0) umma::tmem_alloc(umma::smem_addr(&tmem_base_s), 512);
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = tmem_base_s;
    if (warp < 4) {
        const int j = tid;                       // row of A / channel k of B
        const uint32_t lane_base = (uint32_t)(32 * warp) << 16;
        for (int c0 = 0; c0 < K / 2; c0 += 16) {   // 16 columns = 32 k
            uint32_t hi[16], lo[16];
            for (int c = 0; c < 16; ++c) {
                const float v0 = A[j * K + 2 * (c0 + c)], v1 = A[j * K + 2 * (c0 + c) + 1];
                const uint32_t h0 = bf16_bits(v0), h1 = bf16_bits(v1);
                const uint32_t l0 = bf16_bits(v0 - __uint_as_float(h0 << 16)), l1 = bf16_bits(v1 - __uint_as_float(h1 << 16));
                hi[c] = swap ? (h1 | (h0 << 16)) : (h0 | (h1 << 16));
                lo[c] = swap ? (l1 | (l0 << 16)) : (l0 | (l1 << 16));
                // shared-memory copy of A in the K-major image [K/8][128][8]
                const int k0 = 2 * (c0 + c);
                *reinterpret_cast<uint32_t *>(As_hi + ((k0 / 8) * M + j) * 16 + (k0 % 8) * 2) = h0 | (h1 << 16);
                *reinterpret_cast<uint32_t *>(As_lo + ((k0 / 8) * M + j) * 16 + (k0 % 8) * 2) = l0 | (l1 << 16);
            }
            tmem_st16(tmem + lane_base + COL_AHI + c0, hi);
            tmem_st16(tmem + lane_base + COL_ALO + c0, lo);
        }
        tmem_st_wait();
        for (int n = 0; n < N; ++n) {            // B image: thread = channel k, 2-byte stores
            const float v = B[n * K + j];
            const uint32_t h = bf16_bits(v), l = bf16_bits(v - __uint_as_float(h << 16));
            *reinterpret_cast<unsigned short *>(Bs_hi + (j / 8) * LBO_B + n * 16 + (j % 8) * 2) = (unsigned short)h;
            *reinterpret_cast<unsigned short *>(Bs_lo + (j / 8) * LBO_B + n * 16 + (j % 8) * 2) = (unsigned short)l;
        }
        umma::fence_proxy_async();
    }
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t idesc = umma::idesc_bf16(M, N);
    long long t_issue = 0, t_done = 0, t_wake = 0;
    for (int mode = 0; mode < 2; ++mode) {       // 0: A from tensor memory, 1: A from shared memory
        for (int it = 0; it < iters; ++it) {
            __syncthreads();
            const long long t0 = clock64();
            if (tid == 128) {
                for (int ks = 0; ks < K / 16; ++ks) {
                    const uint64_t bh = umma::smem_desc(umma::smem_addr(Bs_hi) + ks * 2 * LBO_B, LBO_B, 128);
                    const uint64_t bl = umma::smem_desc(umma::smem_addr(Bs_lo) + ks * 2 * LBO_B, LBO_B, 128);
                    if (mode == 0) {
                        mma_bf16_ts(tmem + COL_D, tmem + COL_AHI + ks * 8, bh, idesc, ks > 0);
                        mma_bf16_ts(tmem + COL_D, tmem + COL_AHI + ks * 8, bl, idesc, 1);
                        mma_bf16_ts(tmem + COL_D, tmem + COL_ALO + ks * 8, bh, idesc, 1);
                    } else {
                        const uint64_t ah = umma::smem_desc(umma::smem_addr(As_hi) + ks * 2 * M * 16, M * 16, 128);
                        const uint64_t al = umma::smem_desc(umma::smem_addr(As_lo) + ks * 2 * M * 16, M * 16, 128);
                        umma::mma_bf16(tmem + COL_D2, ah, bh, idesc, ks > 0);
                        umma::mma_bf16(tmem + COL_D2, ah, bl, idesc, 1);
                        umma::mma_bf16(tmem + COL_D2, al, bh, idesc, 1);
                    }
                }
                umma::mma_commit(done);
                t_issue += clock64() - t0;
                umma::mbar_wait(done, (mode * iters + it) & 1);
                t_done += clock64() - t0;
            } else if (tid == 0) {
                umma::mbar_wait(done, (mode * iters + it) & 1);
                t_wake += clock64() - t0;
            }
            umma::fence_after_sync();
        }
        __syncthreads();
        if (tid == 128) { clk[mode * 4 + 0] = t_issue; clk[mode * 4 + 1] = t_done; t_issue = t_done = 0; }
        if (tid == 0) { clk[mode * 4 + 2] = t_wake; t_wake = 0; }
    }
    umma::fence_after_sync();
    if (warp < 4) {
        const int row = tid;
        for (int c = 0; c < N; c += 16) {
            float v[16];
            umma::tmem_ld16(tmem + ((uint32_t)(32 * warp) << 16) + COL_D + c, v);
            umma::tmem_ld_wait();
            for (int i = 0; i < 16; ++i) D[row * N + c + i] = v[i];
            umma::tmem_ld16(tmem + ((uint32_t)(32 * warp) << 16) + COL_D2 + c, v);
            umma::tmem_ld_wait();
            for (int i = 0; i < 16; ++i) D2[row * N + c + i] = v[i];
        }
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, 512);
}

int main() {
    std::vector<float> A(M * K);
    srand(1);
    for (auto &v : A) v = (float)rand() / RAND_MAX * 2.f - 1.f;
    float *dA, *dB, *dD, *dD2;
    long long *dclk;
    cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, 64 * K * 4); cudaMalloc(&dD, M * 64 * 4); cudaMalloc(&dD2, M * 64 * 4);
    cudaMalloc(&dclk, 64);
    cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
    const int iters = 64;
    for (int N : {16, 32, 64})
        for (int swap = 0; swap < 2; ++swap) {
            std::vector<float> B(N * K);
            for (auto &v : B) v = (float)rand() / RAND_MAX * 2.f - 1.f;
            cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
            const int smem = 2 * M * K * 2 + 2 * (K / 8) * (N * 16 + 16);
            cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            cudaMemset(dD, 0, M * 64 * 4); cudaMemset(dD2, 0, M * 64 * 4);
            probe_kernel<<<1, 192, smem>>>(dA, dB, dD, dD2, dclk, N, swap, iters);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("N %d swap %d: %s\n", N, swap, cudaGetErrorString(e)); return 1; }
            std::vector<float> D(M * N), D2(M * N);
            long long clk[8];
            cudaMemcpy(D.data(), dD, M * N * 4, cudaMemcpyDeviceToHost);
            cudaMemcpy(D2.data(), dD2, M * N * 4, cudaMemcpyDeviceToHost);
            cudaMemcpy(clk, dclk, 64, cudaMemcpyDeviceToHost);
            double e1 = 0, e2 = 0, mr = 0;
            for (int m = 0; m < M; ++m)
                for (int n = 0; n < N; ++n) {
                    double r = 0;
                    for (int k = 0; k < K; ++k) r += (double)A[m * K + k] * (double)B[n * K + k];
                    e1 = fmax(e1, fabs(r - D[m * N + n]));
                    e2 = fmax(e2, fabs(r - D2[m * N + n]));
                    mr = fmax(mr, fabs(r));
                }
            printf("N %2d swap %d: A in TMEM max|err| %.3e, A in smem max|err| %.3e (max|ref| %.2f) | cycles per 24 MMAs: TMEM-A issue %.0f done %.0f wake %.0f ; smem-A issue %.0f done %.0f wake %.0f\n",
                   N, swap, e1, e2, mr, (double)clk[0] / iters, (double)clk[1] / iters, (double)clk[2] / iters,
                   (double)clk[4] / iters, (double)clk[5] / iters, (double)clk[6] / iters);
        }
    return 0;
}

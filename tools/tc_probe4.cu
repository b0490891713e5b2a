// tc_probe4.cu - development aid: the B operand (activations [K channels] x [N utterances]) in MN-MAJOR no-swizzle layout, so that
// the thread that owns a channel writes its utterances as ONE 16-byte run: element (k, n) at
//   (n % 8) * 2 + (k % 8) * 16 + (n / 8) * NSTRIDE + (k / 8) * KSTRIDE   bytes,
// instruction descriptor with the "B is MN-major" bit (16).  variant 0: descriptor LBO = KSTRIDE, SBO = NSTRIDE; variant 1: swapped.
// A (weights) from tensor memory as in tc_probe.cu, one pass (hi only), compared with a double reference of the rounded operands.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/tc_probe4 tools/tc_probe4.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_fp16.h>
#include "../tacotron-2_b200/csrc/wn_umma.cuh"

constexpr int M = 128, K = 128, N = 16;
constexpr int COL_D = 0, COL_A = 64;
constexpr int KSTRIDE = 128, NSTRIDE = (K / 8) * 128;   // core matrices of consecutive k groups are contiguous

__global__ void __launch_bounds__(192, 1) probe_kernel(const float *A, const float *B, float *D, int variant) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) unsigned long long bars[1];
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid / 32;
    const uint32_t done = umma::smem_addr(&bars[0]);
    if (tid == 0) { umma::mbar_init(done, 1); umma::mbar_fence_init(); }
    if (warp == 0) umma::tmem_alloc(umma::smem_addr(&tmem_base_s), 512);
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = tmem_base_s;
    if (warp < 4) {
        const int j = tid;
        const uint32_t lane_base = (uint32_t)(32 * warp) << 16;
        for (int c0 = 0; c0 < K / 2; c0 += 16) {
            uint32_t hi[16];
            for (int c = 0; c < 16; ++c) {
                const __half h0 = __float2half_rn(A[j * K + 2 * (c0 + c)]), h1 = __float2half_rn(A[j * K + 2 * (c0 + c) + 1]);
                hi[c] = (uint32_t)__half_as_ushort(h0) | ((uint32_t)__half_as_ushort(h1) << 16);
            }
            umma::tmem_st16(tmem + lane_base + COL_A + c0, hi);
        }
        umma::tmem_st_wait();
        for (int ng = 0; ng < N / 8; ++ng) {   // thread = channel k: 8 utterances = one 16-byte store
            uint32_t w[4];
            for (int p = 0; p < 4; ++p) {
                const __half a = __float2half_rn(B[(ng * 8 + 2 * p) * K + j]), b = __float2half_rn(B[(ng * 8 + 2 * p + 1) * K + j]);
                w[p] = (uint32_t)__half_as_ushort(a) | ((uint32_t)__half_as_ushort(b) << 16);
            }
            *reinterpret_cast<uint4 *>(smem + (j % 8) * 16 + ng * NSTRIDE + (j / 8) * KSTRIDE) = make_uint4(w[0], w[1], w[2], w[3]);
        }
        umma::fence_proxy_async();
    }
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t idesc = umma::idesc_f16(M, N) | (1u << 16);   // B MN-major
    if (warp == 4) {
        if (umma::elect_one()) {
            for (int ks = 0; ks < K / 16; ++ks) {
                const uint32_t a = umma::smem_addr(smem) + ks * 2 * KSTRIDE;
                const uint64_t b = variant == 0 ? umma::smem_desc(a, KSTRIDE, NSTRIDE) : umma::smem_desc(a, NSTRIDE, KSTRIDE);
                umma::mma_bf16_ts(tmem + COL_D, tmem + COL_A + ks * 8, b, idesc, ks > 0);
            }
            umma::mma_commit(done);
        }
        __syncwarp();
    }
    umma::mbar_wait(done, 0);
    umma::fence_after_sync();
    if (warp < 4) {
        float v[16];
        umma::tmem_ld16(tmem + ((uint32_t)(32 * warp) << 16) + COL_D, v);
        umma::tmem_ld_wait();
        for (int i = 0; i < 16; ++i) D[tid * N + i] = v[i];
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, 512);
}

int main() {
    std::vector<float> A(M * K), B(N * K);
    srand(2);
    for (auto &v : A) v = (float)rand() / RAND_MAX * 2.f - 1.f;
    for (auto &v : B) v = (float)rand() / RAND_MAX * 2.f - 1.f;
    float *dA, *dB, *dD;
    cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4); cudaMalloc(&dD, M * N * 4);
    cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
    const int smem = 2 * NSTRIDE;
    for (int variant = 0; variant < 2; ++variant) {
        cudaMemset(dD, 0, M * N * 4);
        probe_kernel<<<1, 192, smem>>>(dA, dB, dD, variant);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("variant %d: %s\n", variant, cudaGetErrorString(e)); return 1; }
        std::vector<float> D(M * N);
        cudaMemcpy(D.data(), dD, M * N * 4, cudaMemcpyDeviceToHost);
        double err = 0, mr = 0;
        for (int m = 0; m < M; ++m)
            for (int n = 0; n < N; ++n) {
                double r = 0;
                for (int k = 0; k < K; ++k) r += (double)__half2float(__float2half_rn(A[m * K + k])) * (double)__half2float(__float2half_rn(B[n * K + k]));
                err = fmax(err, fabs(r - D[m * N + n]));
                mr = fmax(mr, fabs(r));
            }
        printf("MN-major B, variant %d (%s): max|err| %.3e (max|ref| %.2f)\n", variant, variant ? "LBO = n-group stride, SBO = k-group stride" : "LBO = k-group stride, SBO = n-group stride", err, mr);
    }
    return 0;
}

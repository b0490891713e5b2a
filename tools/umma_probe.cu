// umma_probe.cu - development aid: checks the tcgen05 operand-descriptor conventions of csrc/wn_umma.cuh on one CTA.
//   D[128 x 256] = A[128 x K] * B[256 x K]^T, K = 64, A and B random fp32 split into bf16 (hi, lo), three passes
//   (hi*hi + hi*lo + lo*hi), B brought in by one bulk copy, D read back with tcgen05.ld; compared with a double
//   reference.  variant 0: LBO = k-chunk stride, SBO = 128 B (the layout wn_umma.cuh documents); variant 1: swapped.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o umma_probe tools/umma_probe.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../tacotron-2_b200/csrc/wn_umma.cuh"

constexpr int M = 128, N = 256, K = 64;

__global__ void __launch_bounds__(256, 1) probe_kernel(const float *A, const unsigned char *Bpk, float *D, int variant, int passes) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char *As_hi = smem, *As_lo = smem + M * K * 2, *Bs = smem + 2 * M * K * 2;   // Bs: hi then lo
    __shared__ __align__(8) unsigned long long bars[2];
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid / 32, lane = tid % 32;
    const uint32_t full = umma::smem_addr(&bars[0]), done = umma::smem_addr(&bars[1]);
    if (tid == 0) { umma::mbar_init(full, 1); umma::mbar_init(done, 1); umma::mbar_fence_init(); }
    if (warp == 0) umma::tmem_alloc(umma::smem_addr(&tmem_base_s), 256);
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = tmem_base_s;
    if (tid == 0) {
        umma::mbar_expect_tx(full, 2 * N * K * 2);
        for (int i = 0; i < 4; ++i) umma::bulk_g2s(umma::smem_addr(Bs) + i * 16384, Bpk + i * 16384, 16384, full);
    }
    {   // A: thread = (row, k half)
        const int row = tid & 127, kh = tid >> 7;
        for (int c = 0; c < K / 16; ++c) {
            const int kc = kh * (K / 16) + c;
            const float4 a = *reinterpret_cast<const float4 *>(A + row * K + kc * 8), b = *reinterpret_cast<const float4 *>(A + row * K + kc * 8 + 4);
            uint4 hi, lo;
            umma::split8(a, b, hi, lo);
            *reinterpret_cast<uint4 *>(As_hi + (kc * M + row) * 16) = hi;
            *reinterpret_cast<uint4 *>(As_lo + (kc * M + row) * 16) = lo;
        }
    }
    umma::fence_proxy_async();
    __syncthreads();
    if (tid == 0) {
        umma::mbar_wait(full, 0);
        umma::fence_after_sync();
        const uint32_t idesc = umma::idesc_bf16(M, N);
        const uint32_t lboA = M * 16, lboB = N * 16;
        for (int ks = 0; ks < K / 16; ++ks) {
            auto dsc = [&](const unsigned char *base, uint32_t lbo) {
                const uint32_t a = umma::smem_addr(base) + ks * 2 * lbo;
                return variant == 0 ? umma::smem_desc(a, lbo, 128) : umma::smem_desc(a, 128, lbo);
            };
            const uint64_t ah = dsc(As_hi, lboA), al = dsc(As_lo, lboA), bh = dsc(Bs, lboB), bl = dsc(Bs + N * K * 2, lboB);
            umma::mma_bf16(tmem, ah, bh, idesc, ks > 0);
            if (passes == 3) { umma::mma_bf16(tmem, ah, bl, idesc, 1); umma::mma_bf16(tmem, al, bh, idesc, 1); }
        }
        umma::mma_commit(done);
    }
    umma::mbar_wait(done, 0);
    umma::fence_after_sync();
    {
        const int row = 32 * (warp & 3) + lane, c0 = (warp >> 2) * 128;
        for (int c = 0; c < 128; c += 16) {
            float v[16];
            umma::tmem_ld16(tmem + ((uint32_t)(32 * (warp & 3)) << 16) + c0 + c, v);
            umma::tmem_ld_wait();
            for (int i = 0; i < 16; ++i) D[row * N + c0 + c + i] = v[i];
        }
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tmem, 256);
}

static unsigned short bf16_rn(float f) {
    unsigned u;
    memcpy(&u, &f, 4);
    const unsigned r = u + 0x7fffu + ((u >> 16) & 1u);
    return (unsigned short)(r >> 16);
}
static float bf16_f(unsigned short h) {
    unsigned u = (unsigned)h << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

int main() {
    std::vector<float> A(M * K), B(N * K);
    srand(1);
    for (auto &v : A) v = (float)rand() / RAND_MAX * 2.f - 1.f;
    for (auto &v : B) v = (float)rand() / RAND_MAX * 2.f - 1.f;
    std::vector<unsigned short> Bpk(2 * N * K);
    for (int n = 0; n < N; ++n)
        for (int k = 0; k < K; ++k) {
            const unsigned short hi = bf16_rn(B[n * K + k]), lo = bf16_rn(B[n * K + k] - bf16_f(hi));
            Bpk[((k / 8) * N + n) * 8 + k % 8] = hi;
            Bpk[N * K + ((k / 8) * N + n) * 8 + k % 8] = lo;
        }
    float *dA, *dD;
    unsigned char *dB;
    cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dD, M * N * 4); cudaMalloc(&dB, Bpk.size() * 2);
    cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, Bpk.data(), Bpk.size() * 2, cudaMemcpyHostToDevice);
    const int smem = 2 * M * K * 2 + 2 * N * K * 2;
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    std::vector<float> D(M * N);
    for (int variant = 0; variant < 2; ++variant)
        for (int passes = 1; passes <= 3; passes += 2) {
            cudaMemset(dD, 0, M * N * 4);
            probe_kernel<<<1, 256, smem>>>(dA, dB, dD, variant, passes);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("variant %d passes %d: %s\n", variant, passes, cudaGetErrorString(e)); return 1; }
            cudaMemcpy(D.data(), dD, M * N * 4, cudaMemcpyDeviceToHost);
            double maxerr = 0, maxref = 0;
            for (int m = 0; m < M; ++m)
                for (int n = 0; n < N; ++n) {
                    double r = 0;
                    for (int k = 0; k < K; ++k) r += (double)A[m * K + k] * (double)B[n * K + k];
                    maxerr = fmax(maxerr, fabs(r - D[m * N + n]));
                    maxref = fmax(maxref, fabs(r));
                }
            printf("variant %d passes %d: max |err| %.3e (max |ref| %.3f)  D[0][0..3] = %f %f %f %f\n", variant, passes, maxerr, maxref,
                   D[0], D[1], D[2], D[3]);
        }
    return 0;
}

// Development aid: one-way latency of the {value,tag} mailbox hand-off between two CTAs on different SMs of a B200.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/pingpong tools/pingpong.cu && build/pingpong
#include <cstdio>
#include <cuda_runtime.h>
#include <cooperative_groups.h>
__device__ __forceinline__ void st_ll2(unsigned long long *p, float v0, float v1, unsigned tag) {
    const unsigned long long a = ((unsigned long long)tag << 32) | __float_as_uint(v0), b = ((unsigned long long)tag << 32) | __float_as_uint(v1);
    asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ void ld_ll2(const unsigned long long *p, unsigned long long &a, unsigned long long &b) {
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
// CTA 0 and CTA `peer` bounce a message of `nthreads` 16-byte words x `nsrc` sources back and forth `iters` times.
__global__ void pingpong(unsigned long long *box, int iters, int nsrc, int peer, long long *out) {
    if (blockIdx.x != 0 && blockIdx.x != peer) return;
    const int me = blockIdx.x == 0 ? 0 : 1, tid = threadIdx.x, n = blockDim.x;
    unsigned long long *mine = box + (size_t)me * 2 * n * nsrc, *theirs = box + (size_t)(1 - me) * 2 * n * nsrc;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 1; i <= iters; ++i) {
        if (me == 0) {
            for (int s = 0; s < nsrc; ++s) st_ll2(theirs + 2 * (s * n + tid), 1.f, 2.f, i);
        }
        // wait for my inbox
        bool ok;
        float acc = 0.f;
        do {
            ok = true;
            for (int s = 0; s < nsrc; ++s) {
                unsigned long long a, b;
                ld_ll2(mine + 2 * (s * n + tid), a, b);
                ok &= (unsigned)(a >> 32) == (unsigned)i && (unsigned)(b >> 32) == (unsigned)i;
                acc += __uint_as_float((unsigned)a);
            }
        } while (!ok);
        __syncthreads();
        if (me == 1) {
            for (int s = 0; s < nsrc; ++s) st_ll2(theirs + 2 * (s * n + tid), acc, 2.f, i);
        }
    }
    if (tid == 0 && me == 0) out[0] = clock64() - t0;
}
int main() {
    unsigned long long *box; long long *out, h;
    cudaMalloc(&box, 1 << 22); cudaMalloc(&out, 8);
    int dev_sms = 0; cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, 0);
    const int iters = 20000;
    for (int peer : {1, 2, 37, 74, 100, 147})
        for (int nt : {32, 256})
            for (int nsrc : {1, 5}) {
                cudaMemset(box, 0, 1 << 22);
                void *args[] = {&box, (void *)&iters, &nsrc, &peer, &out};
                cudaLaunchCooperativeKernel((void *)pingpong, dim3(dev_sms), dim3(nt), args, 0, 0);
                cudaError_t e = cudaDeviceSynchronize();
                cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
                printf("peer CTA %3d threads %3d sources %d: one-way %.0f cycles (%s)\n", peer, nt, nsrc, (double)h / iters / 2, cudaGetErrorString(e));
            }
    return 0;
}

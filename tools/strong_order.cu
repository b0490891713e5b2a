// Development aid: does a strong (relaxed.gpu) load wait for an earlier strong store of the same thread?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/strong_order tools/strong_order.cu && build/strong_order
#include <cstdio>
#include <cuda_runtime.h>
template <int ST, int LD>   // ST: 0 none, 1 weak, 2 relaxed.gpu, 3 weak .cg   LD: 0 none, 1 weak (ca), 2 relaxed.gpu, 3 weak .cg
__global__ void k(unsigned long long *a, const unsigned long long *b, int iters, long long *out, unsigned long long *sink) {
    unsigned long long *pa = a + 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    const unsigned long long *pb = b + 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    unsigned long long acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        unsigned long long v = acc + i, x = 0, y = 0;
        if (ST == 1) asm volatile("st.global.v2.u64 [%0], {%1, %2};" ::"l"(pa), "l"(v), "l"(v) : "memory");
        if (ST == 2) asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(pa), "l"(v), "l"(v) : "memory");
        if (ST == 3) asm volatile("st.global.cg.v2.u64 [%0], {%1, %2};" ::"l"(pa), "l"(v), "l"(v) : "memory");
        if (LD == 1) asm volatile("ld.global.v2.u64 {%0, %1}, [%2];" : "=l"(x), "=l"(y) : "l"(pb) : "memory");
        if (LD == 2) asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(x), "=l"(y) : "l"(pb) : "memory");
        if (LD == 3) asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(x), "=l"(y) : "l"(pb) : "memory");
        acc += x + y;   // the next iteration depends on the load
    }
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = clock64() - t0;
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int ST, int LD>
void run(const char *name, unsigned long long *a, unsigned long long *b, long long *out, unsigned long long *sink) {
    const int iters = 2000;
    long long h;
    k<ST, LD><<<1, 256>>>(a, b, iters, out, sink);
    cudaDeviceSynchronize();
    k<ST, LD><<<1, 256>>>(a, b, iters, out, sink);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
    printf("%-44s %7.0f cycles per iteration (%s)\n", name, (double)h / iters, cudaGetErrorString(e));
}
int main() {
    unsigned long long *a, *b, *sink; long long *out;
    cudaMalloc(&a, 1 << 20); cudaMalloc(&b, 1 << 20); cudaMalloc(&sink, 1 << 20); cudaMalloc(&out, 8);
    cudaMemset(a, 0, 1 << 20); cudaMemset(b, 0, 1 << 20);
    run<0, 1>("load weak (L1 hit)", a, b, out, sink);
    run<0, 2>("load relaxed.gpu", a, b, out, sink);
    run<0, 3>("load .cg", a, b, out, sink);
    run<2, 0>("store relaxed.gpu only", a, b, out, sink);
    run<1, 0>("store weak only", a, b, out, sink);
    run<2, 2>("store relaxed.gpu -> load relaxed.gpu", a, b, out, sink);
    run<1, 2>("store weak        -> load relaxed.gpu", a, b, out, sink);
    run<3, 2>("store .cg         -> load relaxed.gpu", a, b, out, sink);
    run<2, 3>("store relaxed.gpu -> load .cg", a, b, out, sink);
    run<3, 3>("store .cg         -> load .cg", a, b, out, sink);
    run<1, 3>("store weak        -> load .cg", a, b, out, sink);
    run<1, 1>("store weak        -> load weak", a, b, out, sink);
    return 0;
}

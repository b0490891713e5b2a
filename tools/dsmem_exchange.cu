// Development aid: cost of the intra-layer exchange of the pipelined engine - every CTA of a cluster of C sends
// `nthreads` 16-byte words to EVERY CTA of the cluster (st.async + mbarrier complete_tx) and waits for its own inbox.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/dsmem_exchange tools/dsmem_exchange.cu && build/dsmem_exchange
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned mapa(unsigned a, unsigned rank) { unsigned r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank)); return r; }
__device__ __forceinline__ void st_async4(unsigned ra, float4 v, unsigned rmbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.f32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(ra), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "r"(rmbar) : "memory");
}
__device__ __forceinline__ void mbar_init(unsigned a, unsigned n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_expect(unsigned a, unsigned bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(a), "r"(bytes) : "memory"); }
__device__ __forceinline__ int mbar_try(unsigned a, unsigned par) {
    unsigned ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(a), "r"(par) : "memory");
    return ok;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ unsigned cluster_rank() { unsigned r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ unsigned cluster_size() { unsigned r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r; }

__device__ __forceinline__ void exchange_body(int iters, long long *out, float *sink) {
    extern __shared__ __align__(16) float buf[];      // [2][C][nthreads][4]
    __shared__ __align__(8) unsigned long long mb[2];
    const unsigned C = cluster_size(), me = cluster_rank(), n = blockDim.x, tid = threadIdx.x;
    if (tid == 0) { mbar_init(smem_u32(&mb[0]), 1); mbar_init(smem_u32(&mb[1]), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    cluster_sync_all();
    float acc = 0.f;
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        const int par = i & 1;
        if (tid == 0) mbar_expect(smem_u32(&mb[par]), C * n * 16);
        for (unsigned c = 0; c < C; ++c)
            st_async4(mapa(smem_u32(buf + ((par * C + me) * n + tid) * 4), c), make_float4(acc, 1.f, 2.f, (float)i), mapa(smem_u32(&mb[par]), c));
        while (!mbar_try(smem_u32(&mb[par]), (i >> 1) & 1)) {}
        for (unsigned c = 0; c < C; ++c) acc += buf[((par * C + c) * n + tid) * 4 + 3];
        __syncthreads();
    }
    if (tid == 0 && blockIdx.x == 0) out[0] = clock64() - t0;
    sink[blockIdx.x * n + tid] = acc;
    cluster_sync_all();
}
__global__ void exchange(int iters, long long *out, float *sink) { exchange_body(iters, out, sink); }
struct BigParams { int pad[200]; int iters; long long *out; float *sink; };
__global__ void __launch_bounds__(256, 1) exchange_big(const BigParams p) { exchange_body(p.iters, p.out, p.sink); }

int main(int argc, char **argv) {
    long long *out, h; float *sink;
    cudaMalloc(&out, 8); cudaMalloc(&sink, 4 * 148 * 256);
    const int iters = 20000;
    if (argc >= 6) {   // one launch with an explicit configuration: cluster size, clusters, threads, dynamic smem bytes, cooperative
        const int C = atoi(argv[1]), ncl = atoi(argv[2]), nt = atoi(argv[3]), smem = atoi(argv[4]), coop = atoi(argv[5]);
        cudaFuncSetAttribute(exchange, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(C * ncl); cfg.blockDim = dim3(nt); cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute at[2];
        at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        at[1].id = cudaLaunchAttributeCooperative; at[1].val.cooperative = 1;
        cfg.attrs = at; cfg.numAttrs = coop ? 2 : 1;
        cudaError_t e;
        if (argc >= 7) {   // 7th argument: pass the arguments as one 800-byte struct through cudaLaunchKernelExC
            BigParams bp; memset(&bp, 0, sizeof(bp)); bp.iters = iters; bp.out = out; bp.sink = sink;
            cudaFuncSetAttribute(exchange_big, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            void *args[] = {&bp};
            e = cudaLaunchKernelExC(&cfg, (const void *)exchange_big, args);
        } else {
            e = cudaLaunchKernelEx(&cfg, exchange, iters, out, sink);
        }
        cudaError_t e2 = cudaDeviceSynchronize();
        cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
        printf("cluster %d x %d clusters, %d threads, %d B smem, coop %d: %.0f cycles per exchange (%s / %s)\n", C, ncl, nt, smem, coop,
               (double)h / iters, cudaGetErrorString(e), cudaGetErrorString(e2));
        return 0;
    }
    for (int C : {1, 2, 4, 8})
        for (int nt : {32, 128, 256}) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(C * 8); cfg.blockDim = dim3(nt); cfg.dynamicSmemBytes = 2 * C * nt * 16;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            cudaError_t e = cudaLaunchKernelEx(&cfg, exchange, iters, out, sink);
            cudaError_t e2 = cudaDeviceSynchronize();
            cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
            printf("cluster %d threads %3d (%5d B in per CTA): %.0f cycles per exchange (%s / %s)\n", C, nt, C * nt * 16, (double)h / iters,
                   cudaGetErrorString(e), cudaGetErrorString(e2));
        }
    return 0;
}

// Read-bandwidth ceilings for the access flavours the fused kernel can use (diagnosis; numbers go to profiles/).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o stream_read stream_read.cu && ./stream_read
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

// plain loads, V bytes per thread per access, grid-stride
template <typename V>
__global__ void ldgKernel(const V* __restrict__ in, size_t n, float* out) {
    float acc = 0.f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const V v = __ldg(in + i);
        const float* f = reinterpret_cast<const float*>(&v);
#pragma unroll
        for (int k = 0; k < (int)(sizeof(V) / 4); k++) { acc += f[k]; }
    }
    if (acc == 12345.678f) { *out = acc; }
}

// one CTA per tile: cp.async the tile into shared memory (BYTES per copy), wait, touch it once
template <int BYTES, int THREADS, int TILE_BYTES>
__global__ void __launch_bounds__(THREADS) cpAsyncKernel(const char* __restrict__ in, float* out) {
    extern __shared__ __align__(16) char sm[];
    const char* src = in + (size_t)blockIdx.x * TILE_BYTES;
    const unsigned d0 = (unsigned)__cvta_generic_to_shared(sm);
#pragma unroll
    for (int u = 0; u < TILE_BYTES / (BYTES * THREADS); u++) {
        const int o = (threadIdx.x + u * THREADS) * BYTES;
        if (BYTES == 16) { asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d0 + o), "l"(src + o) : "memory"); }
        else { asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(d0 + o), "l"(src + o) : "memory"); }
    }
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
    __syncthreads();
    float acc = reinterpret_cast<float*>(sm)[threadIdx.x];
    if (acc == 12345.678f) { *out = acc; }
}

// one CTA per tile, one elected thread issues a single bulk (TMA) copy of the tile and everyone waits on the mbarrier
template <int THREADS, int TILE_BYTES>
__global__ void __launch_bounds__(THREADS) bulkKernel(const char* __restrict__ in, float* out) {
    extern __shared__ __align__(128) char sm[];
    __shared__ __align__(8) unsigned long long bar;
    const unsigned barA = (unsigned)__cvta_generic_to_shared(&bar);
    const unsigned d0 = (unsigned)__cvta_generic_to_shared(sm);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(barA));
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(barA), "r"(TILE_BYTES) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
                     ::"r"(d0), "l"(in + (size_t)blockIdx.x * TILE_BYTES), "r"(TILE_BYTES), "r"(barA) : "memory");
    }
    unsigned ok = 0;
    while (!ok) {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(barA) : "memory");
    }
    float acc = reinterpret_cast<float*>(sm)[threadIdx.x];
    if (acc == 12345.678f) { *out = acc; }
}

template <typename F>
float timeIt(F f, int reps) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    f(); f();
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    for (int i = 0; i < reps; i++) { f(); }
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, a, b);
    return ms / reps;
}

int main() {
    const size_t bytes = (size_t)512 << 20;
    char* d = nullptr; float* o = nullptr;
    CK(cudaMalloc(&d, bytes)); CK(cudaMalloc(&o, 4));
    CK(cudaMemset(d, 1, bytes));
    const int reps = 20;
    auto report = [&](const char* name, float ms) { printf("%-44s %8.3f ms  %8.1f GB/s\n", name, ms, bytes / 1e6 / ms); };
    report("ldg float2, 148*8 CTAs x 256, grid-stride", timeIt([&] { ldgKernel<float2><<<148 * 8, 256>>>((const float2*)d, bytes / 8, o); }, reps));
    report("ldg float4, 148*8 CTAs x 256, grid-stride", timeIt([&] { ldgKernel<float4><<<148 * 8, 256>>>((const float4*)d, bytes / 16, o); }, reps));
    report("ldg float4, 148*32 CTAs x 256, grid-stride", timeIt([&] { ldgKernel<float4><<<148 * 32, 256>>>((const float4*)d, bytes / 16, o); }, reps));
    {
        constexpr int TB = 16384;
        CK(cudaFuncSetAttribute(cpAsyncKernel<8, 64, TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, TB));
        report("cp.async 8 B, 16 KB tile/CTA of 64 thr", timeIt([&] { cpAsyncKernel<8, 64, TB><<<bytes / TB, 64, TB>>>(d, o); }, reps));
        CK(cudaFuncSetAttribute(cpAsyncKernel<8, 256, TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, TB));
        report("cp.async 8 B, 16 KB tile/CTA of 256 thr", timeIt([&] { cpAsyncKernel<8, 256, TB><<<bytes / TB, 256, TB>>>(d, o); }, reps));
        CK(cudaFuncSetAttribute(cpAsyncKernel<16, 256, TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, TB));
        report("cp.async 16 B, 16 KB tile/CTA of 256 thr", timeIt([&] { cpAsyncKernel<16, 256, TB><<<bytes / TB, 256, TB>>>(d, o); }, reps));
        CK(cudaFuncSetAttribute(bulkKernel<256, TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, TB));
        report("cp.async.bulk, 16 KB tile/CTA of 256 thr", timeIt([&] { bulkKernel<256, TB><<<bytes / TB, 256, TB>>>(d, o); }, reps));
    }
    {
        constexpr int TB = 32768;
        CK(cudaFuncSetAttribute(cpAsyncKernel<8, 128, TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, TB));
        report("cp.async 8 B, 32 KB tile/CTA of 128 thr", timeIt([&] { cpAsyncKernel<8, 128, TB><<<bytes / TB, 128, TB>>>(d, o); }, reps));
        CK(cudaFuncSetAttribute(cpAsyncKernel<16, 128, TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, TB));
        report("cp.async 16 B, 32 KB tile/CTA of 128 thr", timeIt([&] { cpAsyncKernel<16, 128, TB><<<bytes / TB, 128, TB>>>(d, o); }, reps));
        CK(cudaFuncSetAttribute(bulkKernel<128, TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, TB));
        report("cp.async.bulk, 32 KB tile/CTA of 128 thr", timeIt([&] { bulkKernel<128, TB><<<bytes / TB, 128, TB>>>(d, o); }, reps));
    }
    return 0;
}

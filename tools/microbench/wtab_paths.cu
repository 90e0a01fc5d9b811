// Where should the per-VFO rotation table of the channelizer's first stage live?  The inner loop of the fused kernel
// (one sub-block per lane, 32 samples per thread, one pair of VFOs in the two halves of FFMA2, taps as immediates) with
// the table read (a) from shared memory with a broadcast LDS.128 per sample, (b) from the kernel's parameter block in
// the constant bank through uniform loads (LDCU) straight into uniform-register operands.  Prints FMA-lane-ops/clk/SM.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long ra, rb, rc, rd;
    ra = *reinterpret_cast<unsigned long long*>(&a); rb = *reinterpret_cast<unsigned long long*>(&b); rc = *reinterpret_cast<unsigned long long*>(&c);
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2*>(&rd);
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
    unsigned long long ra, rb, rd;
    ra = *reinterpret_cast<unsigned long long*>(&a); rb = *reinterpret_cast<unsigned long long*>(&b);
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    return *reinterpret_cast<float2*>(&rd);
}

constexpr int D = 64, DS = 32, SB = 64, NT = 128, LP = D + 2, Q = 5, NPAIR = 30;
struct Tables { float4 w[NPAIR * D]; };      // 30 KB: (wA.re, wB.re, wA.im, wB.im) per sample of the sub-block and pair

template <int MODE, int RUN>
__device__ __forceinline__ void body(const float2* xrow, const float4* wsm, const Tables& tp, int pass, float2* accX, float2* accY) {
    const float4* x2 = reinterpret_cast<const float4*>(xrow);
#pragma unroll
    for (int r2 = 0; r2 < DS / 2; r2++) {
        const float4 xx = x2[r2];
#pragma unroll
        for (int half = 0; half < 2; half++) {
            const float xr = half ? xx.z : xx.x, xi = half ? xx.w : xx.y;
            const int r = RUN * DS + 2 * r2 + half;
            const float4 w = MODE == 0 ? wsm[r] : tp.w[pass * D + r];
            const float2 wX = make_float2(w.x, w.y), wY = make_float2(w.z, w.w);
            const float2 zX = ffma2(make_float2(-xi, -xi), wY, fmul2(make_float2(xr, xr), wX));
            const float2 zY = ffma2(make_float2(xi, xi), wX, fmul2(make_float2(xr, xr), wY));
#pragma unroll
            for (int q = 0; q < Q - 1; q++) {
                const float h = 0.001f * (q * D + r + 1);
                accX[q] = ffma2(make_float2(h, h), zX, accX[q]);
                accY[q] = ffma2(make_float2(h, h), zY, accY[q]);
            }
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(NT) probe(const __grid_constant__ Tables tp, const float2* in, float2* out, int tiles) {
    extern __shared__ __align__(16) float2 sm[];
    float2* xs = sm;
    float4* wsm = reinterpret_cast<float4*>(xs + SB * LP);
    const int tid = threadIdx.x, sb = tid % SB, run = tid / SB;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        __syncthreads();
        for (int i = tid; i < SB * D; i += NT) { xs[i + 2 * (i / D)] = in[(size_t)tile * SB * D + i]; }
        __syncthreads();
        const float2* xrow = xs + sb * LP + run * DS;
        for (int pass = 0; pass < NPAIR; pass++) {
            if (MODE == 0) {
                __syncthreads();
                if (tid < D) { wsm[tid] = tp.w[pass * D + tid]; }
                __syncthreads();
            }
            float2 accX[Q], accY[Q];
#pragma unroll
            for (int q = 0; q < Q; q++) { accX[q] = make_float2(0.f, 0.f); accY[q] = accX[q]; }
            if (run == 0) { body<MODE, 0>(xrow, wsm, tp, pass, accX, accY); } else { body<MODE, 1>(xrow, wsm, tp, pass, accX, accY); }
            float2 s = make_float2(0.f, 0.f);
#pragma unroll
            for (int q = 0; q < Q - 1; q++) { s.x += accX[q].x + accX[q].y; s.y += accY[q].x + accY[q].y; }
            out[((size_t)tile * NPAIR + pass) * NT + tid] = s;
        }
    }
}

int main() {
    const int tiles = 148 * 5 * 4;
    float2 *in, *out;
    cudaMalloc(&in, (size_t)tiles * SB * D * sizeof(float2));
    cudaMemset(in, 0, (size_t)tiles * SB * D * sizeof(float2));
    cudaMalloc(&out, (size_t)tiles * NPAIR * NT * sizeof(float2));
    static Tables t;
    for (int i = 0; i < NPAIR * D; i++) { t.w[i] = make_float4(0.999f, 0.998f, 0.01f * (i % 7), 0.02f); }
    const size_t smem = (SB * LP) * sizeof(float2) + D * sizeof(float4);
    cudaFuncSetAttribute(probe<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(probe<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    for (int mode = 0; mode < 2; mode++) {
        float ms = 0.f;
        for (int rep = 0; rep < 3; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) { probe<0><<<148 * 5, NT, smem>>>(t, in, out, tiles); } else { probe<1><<<148 * 5, NT, smem>>>(t, in, out, tiles); }
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            cudaEventElapsedTime(&ms, e0, e1);
        }
        cudaError_t e = cudaGetLastError();
        // FMA lane-ops per sample and VFO: 4 (rotation) + 2 * (Q - 1) taps; two VFOs per pair
        const double ops = (double)tiles * SB * D * NPAIR * 2 * (4 + 2 * (Q - 1));
        printf("%-46s %8.3f ms  %6.1f FMA/clk/SM of 128  (%s)\n", mode == 0 ? "table in shared memory (broadcast LDS.128)" : "table in the parameter block (LDCU -> UR operand)",
               ms, ops / (ms * 1e-3) / 148 / (clk * 1e3), cudaGetErrorString(e));
    }
    return 0;
}

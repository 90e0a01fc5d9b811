// The /8192 plan of the channelizer's first stage (D = 128 samples per sub-block, T = 726 taps -> Q = 6 partial sums, 8 runs
// of 16 samples per sub-block): its fully unrolled loop body holds one copy of the tap MACs per run, because every tap is an
// immediate -- 8 x 4.3 KB = 34 KB of code that 8 warps walk at the same time, past the 32 KB instruction-cache level.
// Alternative probed here: the taps as (h, h) pairs in the kernel's parameter block, fetched with uniform loads (LDCU.64 /
// .128) at a warp-uniform run offset into uniform registers that FFMA2 takes as an operand -- one copy of the loop for all
// runs.  Prints FMA lane operations per clock and SM for both forms (table in shared memory in both, as the plan has it).
// Measured (profiles/r02_tap_paths.txt): 77.8 FMA/clk/SM with immediates, 34.7 with uniform loads -- six LDCU.64 per
// sample are more than the uniform datapath sustains -- and 9.5 when the run offset is left in a vector register (LDC).
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

constexpr int D = 128, RS = 8, DS = D / RS, SB = 64, NT = SB * RS, LP = D + 2, Q = 6, NPASS = 16;
struct Taps { float2 hh[Q * D]; };          // 6 KB: every tap twice, so that one uniform load yields the packed operand

template <int MODE, int RUN>
__device__ __forceinline__ void body(const float2* xrow, const float4* wq, const Taps& tp, int base, float2* accX, float2* accY) {
    const float4* x2 = reinterpret_cast<const float4*>(xrow);
#pragma unroll
    for (int r2 = 0; r2 < DS / 2; r2++) {
        const float4 xx = x2[r2];
#pragma unroll
        for (int half = 0; half < 2; half++) {
            const float xr = half ? xx.z : xx.x, xi = half ? xx.w : xx.y;
            const int r = 2 * r2 + half;
            const float4 w = wq[r];
            const float2 wX = make_float2(w.x, w.y), wY = make_float2(w.z, w.w);
            const float2 zX = ffma2(make_float2(-xi, -xi), wY, fmul2(make_float2(xr, xr), wX));
            const float2 zY = ffma2(make_float2(xi, xi), wX, fmul2(make_float2(xr, xr), wY));
#pragma unroll
            for (int q = 0; q < Q; q++) {
                if (MODE == 0) {
                    const float h = 0.001f * (float)(q * D + RUN * DS + r + 1);       // compile time: an immediate
                    accX[q] = ffma2(make_float2(h, h), zX, accX[q]);
                    accY[q] = ffma2(make_float2(h, h), zY, accY[q]);
                }
                else {
                    const float2 hh = tp.hh[q * D + base + r];                         // uniform index: LDCU
                    accX[q] = ffma2(hh, zX, accX[q]);
                    accY[q] = ffma2(hh, zY, accY[q]);
                }
            }
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(NT) probe(const __grid_constant__ Taps tp, const float2* in, const float4* wtab, float2* out, int tiles) {
    extern __shared__ __align__(16) float2 sm[];
    float2* xs = sm;
    float4* wsm = reinterpret_cast<float4*>(xs + SB * LP);
    const int tid = threadIdx.x, sb = tid % SB, run = tid / SB;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        __syncthreads();
        for (int i = tid; i < SB * D; i += NT) { xs[i + 2 * (i / D)] = in[(size_t)tile * SB * D + i]; }
        __syncthreads();
        const float2* xrow = xs + sb * LP + run * DS;
        for (int pass = 0; pass < NPASS; pass++) {
            __syncthreads();
            if (tid < D) { wsm[tid] = wtab[pass * D + tid]; }
            __syncthreads();
            float2 accX[Q], accY[Q];
#pragma unroll
            for (int q = 0; q < Q; q++) { accX[q] = make_float2(0.f, 0.f); accY[q] = accX[q]; }
            const float4* wq = wsm + run * DS;
            if (MODE == 1) { body<1, 0>(xrow, wq, tp, (int)__reduce_max_sync(0xffffffffu, (unsigned)run) * DS, accX, accY); }   // REDUX: the result is a uniform register
            else {
                switch (run) {
                case 0: body<0, 0>(xrow, wq, tp, 0, accX, accY); break;
                case 1: body<0, 1>(xrow, wq, tp, 0, accX, accY); break;
                case 2: body<0, 2>(xrow, wq, tp, 0, accX, accY); break;
                case 3: body<0, 3>(xrow, wq, tp, 0, accX, accY); break;
                case 4: body<0, 4>(xrow, wq, tp, 0, accX, accY); break;
                case 5: body<0, 5>(xrow, wq, tp, 0, accX, accY); break;
                case 6: body<0, 6>(xrow, wq, tp, 0, accX, accY); break;
                default: body<0, 7>(xrow, wq, tp, 0, accX, accY); break;
                }
            }
            float2 s = make_float2(0.f, 0.f);
#pragma unroll
            for (int q = 0; q < Q; q++) { s.x += accX[q].x + accX[q].y; s.y += accY[q].x + accY[q].y; }
            out[((size_t)tile * NPASS + pass) * NT + tid] = s;
        }
    }
}

int main() {
    const int tiles = 148 * 2 * 6;
    float2 *in, *out; float4* wtab;
    cudaMalloc(&in, (size_t)tiles * SB * D * sizeof(float2));
    cudaMemset(in, 0, (size_t)tiles * SB * D * sizeof(float2));
    cudaMalloc(&out, (size_t)tiles * NPASS * NT * sizeof(float2));
    cudaMalloc(&wtab, (size_t)NPASS * D * sizeof(float4));
    cudaMemset(wtab, 0, (size_t)NPASS * D * sizeof(float4));
    static Taps t;
    for (int i = 0; i < Q * D; i++) { t.hh[i] = make_float2(0.001f * (i + 1), 0.001f * (i + 1)); }
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
            if (mode == 0) { probe<0><<<148 * 2, NT, smem>>>(t, in, wtab, out, tiles); } else { probe<1><<<148 * 2, NT, smem>>>(t, in, wtab, out, tiles); }
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            cudaEventElapsedTime(&ms, e0, e1);
        }
        cudaError_t e = cudaGetLastError();
        const double ops = (double)tiles * SB * D * NPASS * 2 * (4 + 2 * Q);
        printf("%-64s %8.3f ms  %6.1f FMA/clk/SM of 128  (%s)\n", mode == 0 ? "taps as immediates, one copy of the loop per run (34 KB)" : "taps as (h, h) pairs from the parameter block (LDCU), one loop",
               ms, ops / (ms * 1e-3) / 148 / (clk * 1e3), cudaGetErrorString(e));
    }
    return 0;
}

// Which form of the first-stage FIR body keeps the FP32 pipes busy when the decimation is large (D = 128, T = 726:
// ratio 8192, the AM channels of cfg4)?  Per thread: DS = 32 samples of a D-sample sub-block (RS = 4 run variants, warp
// uniform), each sample rotated (4 FP32 ops) and accumulated into Q = 6 complex partial outputs (12 FFMA).
//   A  unrolled, taps as constant-bank immediates (what the library does): 4 variants x 8.7 KB of code
//   B  rolled over the samples, taps read from the constant bank through a register index (LDC), layout [d][8]
//   C  rolled, taps from shared memory (LDS.128), layout [d][8]
// Prints FP32 instructions per second against 148 SM x 128 lanes x clock.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ldc_fir ldc_fir.cu && ./ldc_fir
#include <cstdio>
#include <cuda_runtime.h>

constexpr int D = 128, T = 726, Q = 6, DS = 32, RS = 4, SB = 32, NT = SB * RS, LP = D + 2;

__constant__ float c_taps[T + 64];          // natural order (A)
__constant__ float4 c_dq[D * 2];            // [d][q0..7] (B)

__device__ __forceinline__ float2 cmulf(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }

template <int RUN>
__device__ __forceinline__ void bodyA(const float2* x, const float2* w, float2* acc) {
#pragma unroll
    for (int r = 0; r < DS; r++) {
        const float2 z = cmulf(x[r], w[r]);
#pragma unroll
        for (int q = 0; q < Q; q++) {
            const int t = q * D + RUN * DS + r;
            if (t < T) { const float h = c_taps[t]; acc[q].x = fmaf(h, z.x, acc[q].x); acc[q].y = fmaf(h, z.y, acc[q].y); }
        }
    }
}

// E  as A, but the taps are compile-time values: ptxas folds them into the FFMA as 32-bit immediates (no constant bank)
__host__ __device__ constexpr float tapValue(int t) { return t < T ? 1e-3f * (float)((t * 37) % 101 - 50) : 0.f; }
template <int RUN>
__device__ __forceinline__ void bodyE(const float2* x, const float2* w, float2* acc) {
#pragma unroll
    for (int r = 0; r < DS; r++) {
        const float2 z = cmulf(x[r], w[r]);
#pragma unroll
        for (int q = 0; q < Q; q++) {
            const int t = q * D + RUN * DS + r;
            if (t < T) { constexpr int dummy = 0; (void)dummy; const float h = tapValue(t); acc[q].x = fmaf(h, z.x, acc[q].x); acc[q].y = fmaf(h, z.y, acc[q].y); }
        }
    }
}

__device__ __forceinline__ void bodyB(const float2* x, const float2* w, float2* acc, int run) {
    const float4* tq = c_dq + run * DS * 2;
#pragma unroll 4
    for (int r = 0; r < DS; r++) {
        const float2 z = cmulf(x[r], w[r]);
        const float4 a = tq[2 * r], b = tq[2 * r + 1];
        acc[0].x = fmaf(a.x, z.x, acc[0].x); acc[0].y = fmaf(a.x, z.y, acc[0].y);
        acc[1].x = fmaf(a.y, z.x, acc[1].x); acc[1].y = fmaf(a.y, z.y, acc[1].y);
        acc[2].x = fmaf(a.z, z.x, acc[2].x); acc[2].y = fmaf(a.z, z.y, acc[2].y);
        acc[3].x = fmaf(a.w, z.x, acc[3].x); acc[3].y = fmaf(a.w, z.y, acc[3].y);
        acc[4].x = fmaf(b.x, z.x, acc[4].x); acc[4].y = fmaf(b.x, z.y, acc[4].y);
        acc[5].x = fmaf(b.y, z.x, acc[5].x); acc[5].y = fmaf(b.y, z.y, acc[5].y);
    }
}

__device__ __forceinline__ void bodyC(const float2* x, const float2* w, float2* acc, const float4* tq) {
#pragma unroll 4
    for (int r = 0; r < DS; r++) {
        const float2 z = cmulf(x[r], w[r]);
        const float4 a = tq[2 * r], b = tq[2 * r + 1];
        acc[0].x = fmaf(a.x, z.x, acc[0].x); acc[0].y = fmaf(a.x, z.y, acc[0].y);
        acc[1].x = fmaf(a.y, z.x, acc[1].x); acc[1].y = fmaf(a.y, z.y, acc[1].y);
        acc[2].x = fmaf(a.z, z.x, acc[2].x); acc[2].y = fmaf(a.z, z.y, acc[2].y);
        acc[3].x = fmaf(a.w, z.x, acc[3].x); acc[3].y = fmaf(a.w, z.y, acc[3].y);
        acc[4].x = fmaf(b.x, z.x, acc[4].x); acc[4].y = fmaf(b.x, z.y, acc[4].y);
        acc[5].x = fmaf(b.y, z.x, acc[5].x); acc[5].y = fmaf(b.y, z.y, acc[5].y);
    }
}

// D  position-major: a thread owns 8 sample positions of the sub-block for good (their 8 x Q taps and 8 rotation factors
//    live in registers, every FFMA is register-only) and walks over the sub-blocks; partial sums per (thread, sub-block)
__global__ void __launch_bounds__(NT) firPos(const float2* in, float2* out, int iters) {
    __shared__ __align__(16) float2 xs[SB * LP];
    constexpr int PR = D / 8;                 // position runs per sub-block
    constexpr int GR = NT / PR;               // sub-blocks worked on side by side
    const int tid = threadIdx.x, pr = tid % PR, g = tid / PR;
    for (int i = tid; i < SB * D; i += NT) { xs[i + 2 * (i / D)] = in[(size_t)blockIdx.x * SB * D + i]; }
    float h[8][Q]; float2 w[8];
#pragma unroll
    for (int r = 0; r < 8; r++) {
        w[r] = make_float2(0.999f, 0.01f * (pr * 8 + r));
#pragma unroll
        for (int q = 0; q < Q; q++) { const int t = q * D + pr * 8 + r; h[r][q] = t < T ? c_taps[t] : 0.f; }
    }
    __syncthreads();
    float2 tot = make_float2(0.f, 0.f);
    for (int it = 0; it < iters; it++) {
#pragma unroll 1
        for (int sb = g; sb < SB; sb += GR) {
            const float4* x4 = reinterpret_cast<const float4*>(xs + sb * LP + pr * 8);
            float2 acc[Q];
#pragma unroll
            for (int q = 0; q < Q; q++) { acc[q] = make_float2(0.f, 0.f); }
#pragma unroll
            for (int r2 = 0; r2 < 4; r2++) {
                const float4 xx = x4[r2];
                const float2 z0 = cmulf(make_float2(xx.x, xx.y), w[2 * r2]), z1 = cmulf(make_float2(xx.z, xx.w), w[2 * r2 + 1]);
#pragma unroll
                for (int q = 0; q < Q; q++) {
                    acc[q].x = fmaf(h[2 * r2][q], z0.x, acc[q].x); acc[q].y = fmaf(h[2 * r2][q], z0.y, acc[q].y);
                    acc[q].x = fmaf(h[2 * r2 + 1][q], z1.x, acc[q].x); acc[q].y = fmaf(h[2 * r2 + 1][q], z1.y, acc[q].y);
                }
            }
            // the real kernel stores the Q partial sums here; keep them alive at the same cost class (Q shared-memory stores)
#pragma unroll
            for (int q = 0; q < Q; q++) { tot.x += acc[q].x; tot.y += acc[q].y; }
        }
    }
    out[(size_t)blockIdx.x * NT + tid] = tot;
}

template <int MODE>
__global__ void __launch_bounds__(NT) fir(const float2* in, float2* out, int iters) {
    __shared__ __align__(16) float2 xs[SB * LP];
    __shared__ __align__(16) float2 ws[D];
    __shared__ __align__(16) float4 ts[D * 2];
    const int tid = threadIdx.x, sb = tid % SB, run = tid / SB;
    for (int i = tid; i < SB * D; i += NT) { xs[i + 2 * (i / D)] = in[(size_t)blockIdx.x * SB * D + i]; }
    for (int i = tid; i < D; i += NT) { ws[i] = make_float2(0.999f, 0.01f * i); }
    for (int i = tid; i < D * 2; i += NT) { ts[i] = c_dq[i]; }
    __syncthreads();
    float2 acc[Q];
    for (int q = 0; q < Q; q++) { acc[q] = make_float2(0.f, 0.f); }
    const float2* x = xs + sb * LP + run * DS;
    const float2* w = ws + run * DS;
    for (int it = 0; it < iters; it++) {
        if (MODE == 0) {
            switch (run) {
            case 0: bodyA<0>(x, w, acc); break;
            case 1: bodyA<1>(x, w, acc); break;
            case 2: bodyA<2>(x, w, acc); break;
            default: bodyA<3>(x, w, acc); break;
            }
        }
        else if (MODE == 3) {
            switch (run) {
            case 0: bodyE<0>(x, w, acc); break;
            case 1: bodyE<1>(x, w, acc); break;
            case 2: bodyE<2>(x, w, acc); break;
            default: bodyE<3>(x, w, acc); break;
            }
        }
        else if (MODE == 1) { bodyB(x, w, acc, run); }
        else { bodyC(x, w, acc, ts + run * DS * 2); }
    }
    float2 s = make_float2(0.f, 0.f);
    for (int q = 0; q < Q; q++) { s.x += acc[q].x; s.y += acc[q].y; }
    out[(size_t)blockIdx.x * NT + tid] = s;
}

int main() {
    float h[T + 64]; float4 dq[D * 2];
    for (int t = 0; t < T + 64; t++) { h[t] = t < T ? 1e-3f * (float)((t * 37) % 101 - 50) : 0.f; }
    for (int d = 0; d < D; d++) {
        float v[8];
        for (int q = 0; q < 8; q++) { const int t = q * D + d; v[q] = (q < Q && t < T) ? h[t] : 0.f; }
        dq[2 * d] = make_float4(v[0], v[1], v[2], v[3]); dq[2 * d + 1] = make_float4(v[4], v[5], v[6], v[7]);
    }
    cudaMemcpyToSymbol(c_taps, h, sizeof(h)); cudaMemcpyToSymbol(c_dq, dq, sizeof(dq));
    const int ctas = 148 * 16, iters = 200;
    float2 *in, *out;
    cudaMalloc(&in, (size_t)ctas * SB * D * sizeof(float2)); cudaMalloc(&out, (size_t)ctas * NT * sizeof(float2));
    cudaMemset(in, 0, (size_t)ctas * SB * D * sizeof(float2));
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const double peak = (double)p.multiProcessorCount * 128.0 * (double)p.clockRate * 1e3;
    const char* names[] = { "A unrolled, immediate constant taps", "B rolled, constant bank by register index", "C rolled, shared-memory taps", "D position-major, taps in registers", "E unrolled, taps as FFMA immediates" };
    for (int m = 0; m < 5; m++) {
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(a);
            if (m == 0) { fir<0><<<ctas, NT>>>(in, out, iters); } else if (m == 1) { fir<1><<<ctas, NT>>>(in, out, iters); } else if (m == 2) { fir<2><<<ctas, NT>>>(in, out, iters); } else if (m == 3) { firPos<<<ctas, NT>>>(in, out, iters); } else { fir<3><<<ctas, NT>>>(in, out, iters); }
            cudaEventRecord(b); cudaEventSynchronize(b);
        }
        float ms = 0; cudaEventElapsedTime(&ms, a, b);
        const double fp32 = (double)ctas * NT * iters * DS * (4.0 + 2.0 * Q);   // lane-instructions
        printf("%-46s %8.3f ms  %6.2f T lane-instr/s  %.2f of peak (%s)\n", names[m], ms, fp32 / (ms * 1e-3) / 1e12, fp32 / (ms * 1e-3) / peak, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}

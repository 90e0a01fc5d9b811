// Packed FP32 (fma.rn.f32x2 -> SASS FFMA2) against scalar FFMA on B200: issue rate and FMA rate.
//   mode 0: scalar FFMA, three register operands        (64 independent chains / thread-iteration)
//   mode 1: scalar FFMA, one operand a 32-bit immediate (the tap form of the fused channelizer kernel)
//   mode 2: FFMA2, three register-pair operands
//   mode 3: FFMA2 with the multiplier pair built from an immediate (mov.b64 {imm, imm})
//   mode 4: the channelizer's inner step per sample: rotate (FMUL2 + FFMA2) + 4 tap FFMA2 (tap pairs in registers)
//   mode 5: the same step in scalar form: 2 FMUL + 2 FFMA + 8 FFMA-imm
// Prints warp-instructions/clk/SM and FMA/clk/SM at the measured SM clock.
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

template <int MODE>
__global__ void __launch_bounds__(512) probe(float2* out, float s, int iters) {
    float2 a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = make_float2(threadIdx.x + i, threadIdx.x - i); }
    float2 x0 = make_float2(s, s + 1), x1 = make_float2(s + 2, s + 3), w = make_float2(0.999f, 0.01f), wq = make_float2(-0.01f, 0.999f);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
            if (MODE == 0) {
#pragma unroll
                for (int i = 0; i < 8; i++) { a[i].x = fmaf(a[i].x, x0.x, x1.x); a[i].y = fmaf(a[i].y, x0.y, x1.y); }
            }
            else if (MODE == 1) {
#pragma unroll
                for (int i = 0; i < 8; i++) { a[i].x = fmaf(0.123f + 0.01f * (u * 8 + i), x0.x, a[i].x); a[i].y = fmaf(0.123f + 0.01f * (u * 8 + i), x0.y, a[i].y); }
            }
            else if (MODE == 2) {
#pragma unroll
                for (int i = 0; i < 8; i++) { a[i] = ffma2(a[i], x0, x1); }
            }
            else if (MODE == 3) {
#pragma unroll
                for (int i = 0; i < 8; i++) { const float h = 0.123f + 0.01f * (u * 8 + i); a[i] = ffma2(make_float2(h, h), x0, a[i]); }
            }
            else if (MODE == 4) {
                // two samples: z = x.x * (w.x, w.y) + x.y * (-w.y, w.x); 4 accumulators each
#pragma unroll
                for (int k = 0; k < 2; k++) {
                    const float2 x = k ? x1 : x0;
                    float2 z = fmul2(make_float2(x.x, x.x), w);
                    z = ffma2(make_float2(x.y, x.y), wq, z);
#pragma unroll
                    for (int q = 0; q < 4; q++) { const float h = 0.123f + 0.01f * (u * 8 + k * 4 + q); a[k * 4 + q] = ffma2(make_float2(h, h), z, a[k * 4 + q]); }
                    x0.x += z.x * 1e-9f;
                }
            }
            else {
#pragma unroll
                for (int k = 0; k < 2; k++) {
                    const float2 x = k ? x1 : x0;
                    float2 z;
                    z.x = x.x * w.x - x.y * w.y;
                    z.y = x.x * w.y + x.y * w.x;
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        const float h = 0.123f + 0.01f * (u * 8 + k * 4 + q);
                        a[k * 4 + q].x = fmaf(h, z.x, a[k * 4 + q].x); a[k * 4 + q].y = fmaf(h, z.y, a[k * 4 + q].y);
                    }
                    x0.x += z.x * 1e-9f;
                }
            }
        }
    }
    float2 r = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 8; i++) { r.x += a[i].x; r.y += a[i].y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE> void run(float2* out, const char* name, double fmaPerIter, double instrPerIter) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000, grid = 148 * 4, threads = 512;
    float ms = 0.f;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        probe<MODE><<<grid, threads>>>(out, 0.5f, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
    }
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double total = (double)grid * threads * iters;
    const double sec = ms * 1e-3, hz = clk * 1e3;
    printf("%-44s %8.3f ms  %6.1f FMA/clk/SM  %5.2f warp-instr/clk/SM  %6.1f TFMA/s (clock attr %.0f MHz)\n", name, ms, total * fmaPerIter / sec / 148 / hz,
           total * instrPerIter / 32 / sec / 148 / hz, total * fmaPerIter / sec / 1e12, clk / 1e3);
}

int main() {
    float2* out;
    cudaMalloc(&out, 148 * 4 * 512 * sizeof(float2));
    run<0>(out, "FFMA 3-register", 128, 128);
    run<1>(out, "FFMA immediate multiplier", 128, 128);
    run<2>(out, "FFMA2 3-register-pair", 128, 64);
    run<3>(out, "FFMA2 multiplier pair from immediate", 128, 64);
    run<4>(out, "channelizer step, packed (per 16 samples)", 16 * 12, 16 * 6);
    run<5>(out, "channelizer step, scalar (per 16 samples)", 16 * 12, 16 * 12);
    return 0;
}

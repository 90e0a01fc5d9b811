// FP32 FFMA issue throughput on one SM-full of warps: 3-register form vs constant-bank operand form.
#include <cstdio>
#include <cuda_runtime.h>

__constant__ float cTap[64];

template <int MODE>
__global__ void __launch_bounds__(1024) probe(float* out, float s, int iters) {
    float a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    float x0 = s, x1 = s + 1, x2 = s + 2, x3 = s + 3;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
            if (MODE == 0) {          // three register operands
                a0 = fmaf(a0, x0, x1); a1 = fmaf(a1, x1, x2); a2 = fmaf(a2, x2, x3); a3 = fmaf(a3, x3, x0);
                a4 = fmaf(a4, x0, x2); a5 = fmaf(a5, x1, x3); a6 = fmaf(a6, x2, x0); a7 = fmaf(a7, x3, x1);
            }
            else {                    // one operand from the constant bank (tap), like the FIR inner loop
                a0 = fmaf(cTap[u * 8 + 0], x0, a0); a1 = fmaf(cTap[u * 8 + 1], x1, a1); a2 = fmaf(cTap[u * 8 + 2], x2, a2); a3 = fmaf(cTap[u * 8 + 3], x3, a3);
                a4 = fmaf(cTap[u * 8 + 4], x0, a4); a5 = fmaf(cTap[u * 8 + 5], x1, a5); a6 = fmaf(cTap[u * 8 + 6], x2, a6); a7 = fmaf(cTap[u * 8 + 7], x3, a7);
            }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

int main() {
    float* out;
    cudaMalloc(&out, 148 * 2 * 1024 * 4);
    float h[64];
    for (int i = 0; i < 64; i++) h[i] = 1.0f / (i + 2);
    cudaMemcpyToSymbol(cTap, h, sizeof(h));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    for (int mode = 0; mode < 2; mode++) {
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) probe<0><<<148 * 2, 1024>>>(out, 0.5f, iters);
            else probe<1><<<148 * 2, 1024>>>(out, 0.5f, iters);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            double ffma = (double)148 * 2 * 1024 * iters * 64;
            if (rep) printf("%s: %.1f TFFMA/s = %.1f FFMA/clk/SM at 1.965 GHz (%.3f ms)\n", mode ? "const-operand" : "3-register", ffma / (ms * 1e-3) / 1e12,
                            ffma / (ms * 1e-3) / 148 / 1.965e9, ms);
        }
    }
    return 0;
}

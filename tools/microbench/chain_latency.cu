// Dependent-chain latency probes for the AGC recurrence (one lane of one warp active).
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void probe(float* out, long long* cyc, float a0, float s1, float i1, float s2, float i2, int n) {
    if (threadIdx.x != 0) return;
    float am = a0;
    long long t0 = clock64();
#pragma unroll 16
    for (int k = 0; k < n; k++) {
        if (MODE == 0) { am = fmaf(am, s1, i1); }                                              // FFMA chain
        if (MODE == 1) { am = fmaxf(fmaf(am, s1, i1), fmaf(am, s2, i2)); }                     // 2 FFMA + FMNMX
        if (MODE == 2) { float a = fmaf(am, s1, i1), b = fmaf(am, s2, i2); am = (a > b) ? a : b; } // compare/select
        if (MODE == 3) { float a = fmaf(am, s1, i1), b = fmaf(am, s2, i2); am = 0.5f * (a + b + fabsf(a - b)); } // arithmetic max
        if (MODE == 4) { float a = fmaf(am, s1, i1), b = fmaf(am, s2, i2);
                         am = __int_as_float(max(__float_as_int(a), __float_as_int(b))); }     // integer max on positive floats
        if (MODE == 5) { am = fmaxf(am * s1, i1); }                                             // FMUL + FMNMX
    }
    long long t1 = clock64();
    out[0] = am; cyc[0] = t1 - t0;
}

int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 4); cudaMalloc(&cyc, 8);
    const int n = 1 << 16;
    const char* names[] = { "FFMA", "2xFFMA+FMNMX", "2xFFMA+FSETP/FSEL", "2xFFMA+arith max", "2xFFMA+IMNMX", "FMUL+FMNMX" };
    for (int rep = 0; rep < 2; rep++) {
        for (int m = 0; m < 6; m++) {
            switch (m) {
            case 0: probe<0><<<1, 32>>>(out, cyc, 0.01f, 0.998f, 1e-5f, 0.9998f, 1e-6f, n); break;
            case 1: probe<1><<<1, 32>>>(out, cyc, 0.01f, 0.998f, 1e-5f, 0.9998f, 1e-6f, n); break;
            case 2: probe<2><<<1, 32>>>(out, cyc, 0.01f, 0.998f, 1e-5f, 0.9998f, 1e-6f, n); break;
            case 3: probe<3><<<1, 32>>>(out, cyc, 0.01f, 0.998f, 1e-5f, 0.9998f, 1e-6f, n); break;
            case 4: probe<4><<<1, 32>>>(out, cyc, 0.01f, 0.998f, 1e-5f, 0.9998f, 1e-6f, n); break;
            case 5: probe<5><<<1, 32>>>(out, cyc, 0.01f, 0.998f, 1e-5f, 0.9998f, 1e-6f, n); break;
            }
            long long c; float o;
            cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&o, out, 4, cudaMemcpyDeviceToHost);
            if (rep) printf("%-22s %6.2f cycles/step (result %g)\n", names[m], (double)c / n, o);
        }
    }
    return 0;
}

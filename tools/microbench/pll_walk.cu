// Where the pilot PLL kernel's time goes: clock64 around each stage of one warp's chunk loop (phases by atan2f from
// global memory, phase increments, the error-form walk without / with the frequency clamp in its chain, phases out).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pll_walk pll_walk.cu && ./pll_walk
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>
#include <cmath>

constexpr int kChunk = 512;
__device__ __forceinline__ float wrapPi(float v) {
    const float PI = 3.1415926535f;
    return (v > PI) ? v - 2.0f * PI : ((v <= -PI) ? v + 2.0f * PI : v);
}
struct Walk {
    float c1, beta, lo, hi;
    float e, f, eLast, fLo, fHi, rMax;
    template <bool CLAMP, bool WATCH>
    __device__ __forceinline__ void run(const float* __restrict__ DX, float* __restrict__ E, int m, float e0, float f0) {
        e = e0; f = f0; eLast = e0; fLo = f0; fHi = f0; rMax = 0.0f;
        auto step = [&](float dx) {
            const float cur = e;
            const float g = fmaf(c1, e, dx);
            float fn = fmaf(beta, e, f);
            if (WATCH) { fLo = fminf(fLo, fn); fHi = fmaxf(fHi, fn); }
            if (CLAMP) { fn = fminf(fmaxf(fn, lo), hi); }
            f = fn;
            const float r = g - fn;
            if (WATCH) { rMax = fmaxf(rMax, fabsf(r)); }
            if (CLAMP) { const float k = fmaf(r, 0.15915494309189535f, 12582912.0f) - 12582912.0f; e = fmaf(k, -2.0f * 3.1415926535f, r); }   // the kernel's wrap
            else { e = wrapPi(r); }                                                                                                            // compare + select
            return cur;
        };
        int i = 0;
        float4 xa, xb;
        if (m >= 8) { xa = *reinterpret_cast<const float4*>(&DX[0]); xb = *reinterpret_cast<const float4*>(&DX[4]); }
#pragma unroll 1
        for (; i + 8 <= m; i += 8) {
            const float4 ca = xa, cb = xb;
            if (i + 16 <= m) { xa = *reinterpret_cast<const float4*>(&DX[i + 8]); xb = *reinterpret_cast<const float4*>(&DX[i + 12]); }
            float4 oa, ob;
            oa.x = step(ca.x); oa.y = step(ca.y); oa.z = step(ca.z); oa.w = step(ca.w);
            ob.x = step(cb.x); ob.y = step(cb.y); ob.z = step(cb.z); ob.w = step(cb.w);
            eLast = ob.w;
            *reinterpret_cast<float4*>(&E[i]) = oa;
            *reinterpret_cast<float4*>(&E[i + 4]) = ob;
        }
        for (; i < m; i++) { eLast = step(DX[i]); E[i] = eLast; }
    }
};

// the loop in lock: no step leaves (-pi, pi) and the frequency stays inside its limits, so neither the wrap nor the clamp acts;
// with d[i] = dx[i] - f[i-1] formed beside the chain (f runs a step behind e) the chain is one FFMA per sample
struct WalkLinear {
    float k, beta;
    float e, f, eLast, fLo, fHi, rMax;
    __device__ __forceinline__ void run(const float* __restrict__ DX, float* __restrict__ E, int m, float e0, float f0) {
        e = e0; f = f0; eLast = e0; fLo = f0; fHi = f0; rMax = 0.0f;
        auto step = [&](float dx) {
            const float cur = e;
            const float d = dx - f;
            f = fmaf(beta, e, f);
            fLo = fminf(fLo, f); fHi = fmaxf(fHi, f);
            e = fmaf(k, e, d);
            rMax = fmaxf(rMax, fabsf(e));
            return cur;
        };
        int i = 0;
        float4 xa, xb;
        if (m >= 8) { xa = *reinterpret_cast<const float4*>(&DX[0]); xb = *reinterpret_cast<const float4*>(&DX[4]); }
#pragma unroll 1
        for (; i + 8 <= m; i += 8) {
            const float4 ca = xa, cb = xb;
            if (i + 16 <= m) { xa = *reinterpret_cast<const float4*>(&DX[i + 8]); xb = *reinterpret_cast<const float4*>(&DX[i + 12]); }
            float4 oa, ob;
            oa.x = step(ca.x); oa.y = step(ca.y); oa.z = step(ca.z); oa.w = step(ca.w);
            ob.x = step(cb.x); ob.y = step(cb.y); ob.z = step(cb.z); ob.w = step(cb.w);
            eLast = ob.w;
            *reinterpret_cast<float4*>(&E[i]) = oa;
            *reinterpret_cast<float4*>(&E[i + 4]) = ob;
        }
        for (; i < m; i++) { eLast = step(DX[i]); E[i] = eLast; }
    }
};

// the same loop on 32-bit phases (2^32 per turn: wrapping is the integer overflow), gains as 0.32 fixed point
struct WalkFixed {
    int A, Aa, B, lo, hi;          // (alpha + beta), alpha, beta as fractions of 2^32; frequency limits
    int e, f, eLast, fLo, fHi;
    template <bool CLAMP>
    __device__ __forceinline__ void run(const int* __restrict__ DX, int* __restrict__ E, int m, int e0, int f0) {
        e = e0; f = f0; eLast = e0; fLo = f0; fHi = f0;
        auto step = [&](int dx) {
            const int cur = e;
            if (CLAMP) {
                int fn = f + __mulhi(e, B);
                fLo = min(fLo, fn); fHi = max(fHi, fn);
                fn = min(max(fn, lo), hi);
                e = (dx + e) - __mulhi(e, Aa) - fn;
                f = fn;
            }
            else {
                const int d = dx - f;                       // beside the chain: f is one step behind e
                f = f + __mulhi(e, B);
                fLo = min(fLo, f); fHi = max(fHi, f);
                e = (e + d) - __mulhi(e, A);
            }
            return cur;
        };
        int i = 0;
        int4 xa, xb;
        if (m >= 8) { xa = *reinterpret_cast<const int4*>(&DX[0]); xb = *reinterpret_cast<const int4*>(&DX[4]); }
#pragma unroll 1
        for (; i + 8 <= m; i += 8) {
            const int4 ca = xa, cb = xb;
            if (i + 16 <= m) { xa = *reinterpret_cast<const int4*>(&DX[i + 8]); xb = *reinterpret_cast<const int4*>(&DX[i + 12]); }
            int4 oa, ob;
            oa.x = step(ca.x); oa.y = step(ca.y); oa.z = step(ca.z); oa.w = step(ca.w);
            ob.x = step(cb.x); ob.y = step(cb.y); ob.z = step(cb.z); ob.w = step(cb.w);
            eLast = ob.w;
            *reinterpret_cast<int4*>(&E[i]) = oa;
            *reinterpret_cast<int4*>(&E[i + 4]) = ob;
        }
        for (; i < m; i++) { eLast = step(DX[i]); E[i] = eLast; }
    }
};

__global__ void __launch_bounds__(32) probeFixed(const float* __restrict__ pr, const float* __restrict__ pi, float* __restrict__ out, int n, long long* cyc,
                                                 float alpha, float beta, float lo, float hi) {
    __shared__ __align__(16) int sX[kChunk], sE[kChunk], sD[kChunk + 4];
    const int lane = threadIdx.x;
    long long t[4] = {0, 0, 0, 0};
    const double S = 2147483648.0 / 3.14159265358979323846;
    WalkFixed w;
    w.A = (int)llrint(((double)alpha + (double)beta) * 4294967296.0); w.Aa = (int)llrint((double)alpha * 4294967296.0); w.B = (int)llrint((double)beta * 4294967296.0);
    w.lo = __double2int_rn(lo * S); w.hi = __double2int_rn(hi * S);
    int P = 0, F = (w.lo + w.hi) / 2;
    for (int base = 0; base < n; base += kChunk) {
        const int m = min(kChunk, n - base);
        long long a = clock64();
        for (int i = lane; i < m; i += 32) { sX[i] = __double2int_rn((double)atan2f(__ldg(pi + base + i), __ldg(pr + base + i)) * S); }
        __syncwarp();
        for (int i = lane; i < ((m + 3) & ~3); i += 32) { sD[i] = (i + 1 < m) ? sX[i + 1] - sX[i] : 0; }
        __syncwarp();
        long long b = clock64(); t[0] += b - a; a = b;
        w.run<true>(sD, sE, m, sX[0] - P, F);
        const int keep = w.e + w.fLo + w.fHi;
        b = clock64(); t[1] += b - a; a = b;
        w.run<false>(sD, sE, m, sX[0] - P, F);
        b = clock64(); t[2] += b - a; a = b;
        __syncwarp();
        for (int k = lane; k < m; k += 32) { out[base + k] = (float)((double)(sX[k] - sE[k]) * (1.0 / S)) + 1e-30f * (float)keep; }
        P = (sX[m - 1] - w.eLast) + w.f + __mulhi(w.eLast, w.Aa); F = w.f;
        __syncwarp();
        b = clock64(); t[3] += b - a;
    }
    if (lane == 0) { for (int k = 0; k < 4; k++) { cyc[k] = t[k]; } }
}

__global__ void __launch_bounds__(32) probe(const float* __restrict__ pr, const float* __restrict__ pi, float* __restrict__ out, int n, long long* cyc,
                                            float alpha, float beta, float lo, float hi) {
    __shared__ __align__(16) float sIn[kChunk], sOut[kChunk], sDx[kChunk + 4];
    const int lane = threadIdx.x;
    long long t[7] = {0, 0, 0, 0, 0, 0, 0};
    float phase = 0.f, freq = 0.5f * (lo + hi);
    Walk w; w.c1 = 1.0f - alpha; w.beta = beta; w.lo = lo; w.hi = hi;
    for (int base = 0; base < n; base += kChunk) {
        const int m = min(kChunk, n - base);
        long long a = clock64();
        for (int i = lane; i < m; i += 32) { sIn[i] = atan2f(__ldg(pi + base + i), __ldg(pr + base + i)); }
        __syncwarp();
        long long b = clock64(); t[0] += b - a; a = b;
        for (int i = lane; i < ((m + 3) & ~3); i += 32) { sDx[i] = (i + 1 < m) ? wrapPi(sIn[i + 1] - sIn[i]) : 0.0f; }
        __syncwarp();
        b = clock64(); t[1] += b - a; a = b;
        WalkLinear wl; wl.k = 1.0f - alpha - beta; wl.beta = beta;
        wl.run(sDx, sOut, m, wrapPi(sIn[0] - phase), freq);
        const float keep3 = wl.e + wl.fLo + wl.fHi + wl.rMax;
        b = clock64(); t[4] += b - a; a = b;
        w.run<false, true>(sDx, sOut, m, wrapPi(sIn[0] - phase), freq);
        const float keep = w.e + w.fLo + w.fHi + w.rMax;
        b = clock64(); t[2] += b - a; a = b;
        w.run<true, true>(sDx, sOut, m, wrapPi(sIn[0] - phase), freq);
        const float keep2 = w.e + w.fLo + w.fHi + w.rMax;
        b = clock64(); t[3] += b - a; a = b;

        __syncwarp();
        for (int k = lane; k < m; k += 32) { out[base + k] = wrapPi(sIn[k] - sOut[k]) + 1e-30f * (keep + keep2 + keep3); }
        const float p0 = wrapPi(sIn[m - 1] - w.eLast) + (w.f + alpha * w.eLast);
        phase = wrapPi(p0); freq = w.f;
        __syncwarp();
        b = clock64(); t[5] += b - a;
    }
    if (lane == 0) { for (int k = 0; k < 6; k++) { cyc[k] = t[k]; } }
}

int main() {
    const int n = 1 << 18;
    const double sr = 250e3;
    std::vector<float> re(n), im(n);
    for (int i = 0; i < n; i++) { const double p = 2 * M_PI * 19000.0 * i / sr + 0.3; re[i] = (float)(0.1 * cos(p)); im[i] = (float)(0.1 * sin(p)); }
    float *dr, *di, *dout; long long* dc;
    cudaMalloc(&dr, n * 4); cudaMalloc(&di, n * 4); cudaMalloc(&dout, n * 4); cudaMalloc(&dc, 64);
    cudaMemcpy(dr, re.data(), n * 4, cudaMemcpyHostToDevice); cudaMemcpy(di, im.data(), n * 4, cudaMemcpyHostToDevice);
    const float bw = 25000.0f / (float)sr, den = 1.0f + 1.41421356f * bw + bw * bw;
    const float alpha = 2.0f * 1.41421356f * bw / den, beta = 4.0f * bw * bw / den;
    const float lo = (float)(2 * M_PI * 18750.0 / sr), hi = (float)(2 * M_PI * 19250.0 / sr);
    for (int rep = 0; rep < 2; rep++) { probe<<<1, 32>>>(dr, di, dout, n, dc, alpha, beta, lo, hi); cudaDeviceSynchronize(); }
    long long c[6];
    cudaMemcpy(c, dc, 48, cudaMemcpyDeviceToHost);
    const char* names[] = { "atan2f from global -> smem", "phase increments", "error-form walk, select wrap in chain", "error-form walk, clamp + rint wrap in chain",
                            "error-form walk in lock (one FFMA)", "phases out + state" };
    for (int k = 0; k < 6; k++) { printf("%-36s %7.2f cycles/sample\n", names[k], (double)c[k] / n); }
    std::vector<float> a(n), bb(n);
    cudaMemcpy(a.data(), dout, n * 4, cudaMemcpyDeviceToHost);
    for (int rep = 0; rep < 2; rep++) { probeFixed<<<1, 32>>>(dr, di, dout, n, dc, alpha, beta, lo, hi); cudaDeviceSynchronize(); }
    cudaMemcpy(c, dc, 32, cudaMemcpyDeviceToHost);
    cudaMemcpy(bb.data(), dout, n * 4, cudaMemcpyDeviceToHost);
    const char* namesF[] = { "fixed point: phases + increments", "fixed-point walk, clamp in chain", "fixed-point walk, clamp beside chain", "phases out + state" };
    for (int k = 0; k < 4; k++) { printf("%-36s %7.2f cycles/sample\n", namesF[k], (double)c[k] / n); }
    double worst = 0; int at = -1;
    for (int i = 0; i < n; i++) { double d = fabs((double)a[i] - (double)bb[i]); if (d > 3.0) { d = fabs(d - 2 * M_PI); } if (d > worst) { worst = d; at = i; } }
    printf("largest phase difference float walk / fixed-point walk: %.3g rad at sample %d (phases %g %g)\n", worst, at, a[at], bb[at]);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}

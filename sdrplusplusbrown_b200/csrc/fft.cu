// Windowed forward FFT -> log power.  Replaces IQFrontEnd::handler (signal_path/iq_frontend.cpp:267-287):
//   :271  fftIn[i] = data[i] * window[i]   (window carries the (-1)^i fft-shift, :303-311; tail zero, :317)
//   :274  forward c2c FFT of size N (FFTW in the reference, utils/arrays.cpp:533-546)
//   :282  out[k] = 10*log10((re^2+im^2)/N^2) (VOLK power spectrum, log2 clamped at +-127)
//
// Stockham autosort FFT, radix 16/8/4/2 butterflies held in registers, stages exchanged through two
// ping-pong shared-memory buffers (one barrier per stage, no bit reversal), twiddles from float64-designed
// tables.  N <= 512: one CTA per frame.  Larger N: N = N1 x N2 in two kernels (register-resident passes for N1, N2 <= 256,
// see fftColsKernel / fftRowsKernel; the generic shared-memory passes described here beyond that) -- pass 1 transforms C adjacent
// columns (stride N2, so global reads stay 64-128 B contiguous), applies the inter-pass twiddle and leaves the
// N1 x N2 intermediate in a small scratch ring that stays resident in the 126 MB L2; pass 2 transforms R
// adjacent rows and emits dB (or the complex spectrum) with R-wide contiguous stores.  HBM sees the windowed
// samples once and the dB row once.
#include "fft.cuh"

#include <cmath>

namespace b200 {

int fftDirectMaxLog = 13;    // largest log2(N) served by the one-pass kernel (B200_OPT_FFT_DIRECT_MAX_LOG; 9 = never)
int fftDirectPrefetch = 0;   // B200_OPT_FFT_DIRECT_PREFETCH: the one-pass kernel stages its next frame in shared memory while it transforms this one (measured slower: off)
int fftCtasPerSm = 0;       // B200_OPT_FFT_CTAS_PER_SM: persistent CTAs per SM of the two register-resident passes; 0 = as many as fit

namespace {

constexpr int kThreads = 256;

// (Measured: the same helpers on the packed FP32 pair instructions -- FADD2 / FMUL2 / FFMA2, with the multiplication by -i
// folded into the ".LO_HI.NP" operand form -- cut the pass kernels' instruction counts by 11-13 % and made every size
// 3-10 % SLOWER (65536: 0.0875 -> 0.0901 ms per 208 frames, 2^20: 0.113 -> 0.124): the passes wait on memory and
// barriers, not on FP issue slots.  Scalar form kept.)
__device__ __forceinline__ float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 mulNegI(float2 a) { return make_float2(a.y, -a.x); }   // a * (-i)

// shared-memory slot of element i: one pad element per 16 so that the stride-R stores of the first Stockham
// stage (and stride-2^k accesses in general) spread over the banks
__device__ __host__ __forceinline__ int slot(int i) { return i + (i >> 4); }

// exp(-2*pi*i*m/32), m = 0..31 (m is a compile-time constant wherever this is called: the switch folds away)
__device__ __forceinline__ float2 w32(int m) {
    constexpr float C1 = 0.98078528040323044913f, S1 = 0.19509032201612826785f;   // cos, sin of 1 * pi/16
    constexpr float C2 = 0.92387953251128675613f, S2 = 0.38268343236508977173f;   // 2 * pi/16
    constexpr float C3 = 0.83146961230254523708f, S3 = 0.55557023301960222474f;   // 3 * pi/16
    constexpr float C4 = 0.70710678118654752440f;                                  // 4 * pi/16
    float c, sn;
    switch (m & 7) {
    case 0: c = 1.f; sn = 0.f; break;
    case 1: c = C1; sn = S1; break;
    case 2: c = C2; sn = S2; break;
    case 3: c = C3; sn = S3; break;
    case 4: c = C4; sn = C4; break;
    case 5: c = S3; sn = C3; break;
    case 6: c = S2; sn = C2; break;
    default: c = S1; sn = C1; break;
    }
    // e^{-i a} = (c, -sn) for a in the first quadrant; every further quadrant multiplies by -i
    switch ((m >> 3) & 3) {
    case 0: return make_float2(c, -sn);
    case 1: return make_float2(-sn, -c);
    case 2: return make_float2(-c, sn);
    default: return make_float2(sn, c);
    }
}

// ---- R-point forward DFTs in registers (natural order in, natural order out)
template <int R> struct Dft;
template <> struct Dft<2> {
    static __device__ __forceinline__ void run(float2* v) {
        const float2 a = v[0], b = v[1];
        v[0] = cadd(a, b); v[1] = csub(a, b);
    }
};
template <> struct Dft<4> {
    static __device__ __forceinline__ void run(float2* v) {
        const float2 a0 = cadd(v[0], v[2]), a1 = csub(v[0], v[2]), a2 = cadd(v[1], v[3]), a3 = mulNegI(csub(v[1], v[3]));
        v[0] = cadd(a0, a2); v[2] = csub(a0, a2);
        v[1] = cadd(a1, a3); v[3] = csub(a1, a3);
    }
};
// R = R1 * R2 by Cooley-Tukey on registers: n = R2*n1 + n2, k = k1 + R1*k2
template <int R1, int R2>
__device__ __forceinline__ void dftComposite(float2* v) {
    constexpr int R = R1 * R2;
    float2 t[R];
#pragma unroll
    for (int n2 = 0; n2 < R2; n2++) {
        float2 x[R1];
#pragma unroll
        for (int n1 = 0; n1 < R1; n1++) { x[n1] = v[R2 * n1 + n2]; }
        Dft<R1>::run(x);
#pragma unroll
        for (int k1 = 0; k1 < R1; k1++) { t[k1 * R2 + n2] = (n2 * k1 == 0) ? x[k1] : cmul(x[k1], w32(n2 * k1 * (32 / R))); }
    }
#pragma unroll
    for (int k1 = 0; k1 < R1; k1++) {
        float2 y[R2];
#pragma unroll
        for (int n2 = 0; n2 < R2; n2++) { y[n2] = t[k1 * R2 + n2]; }
        Dft<R2>::run(y);
#pragma unroll
        for (int k2 = 0; k2 < R2; k2++) { v[k1 + R1 * k2] = y[k2]; }
    }
}
template <> struct Dft<8> { static __device__ __forceinline__ void run(float2* v) { dftComposite<4, 2>(v); } };
template <> struct Dft<16> { static __device__ __forceinline__ void run(float2* v) { dftComposite<4, 4>(v); } };
template <> struct Dft<32> { static __device__ __forceinline__ void run(float2* v) { dftComposite<8, 4>(v); } };

// One Stockham stage of radix R on nb rows of length L (row pitch LP slots) from src to dst.
// Ns = product of the radices already applied.  tw[m] = exp(-2*pi*i*m/L), m < L (staged in shared memory).
template <int R, int LOGR>
__device__ __forceinline__ void stockhamStage(const float2* __restrict__ src, float2* __restrict__ dst, int logL, int logNs, int LP, int nb,
                                              const float2* __restrict__ tw) {
    const int logPer = logL - LOGR;
    const int per = 1 << logPer;
    const int total = nb << logPer;
    const int Ns = 1 << logNs;
    const int sh = logL - logNs - LOGR;              // twiddle index = (k*r) << sh
    for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
        const int b = idx >> logPer, j = idx & (per - 1);
        const int k = j & (Ns - 1);
        const float2* s = src + b * LP;
        float2 v[R];
#pragma unroll
        for (int r = 0; r < R; r++) { v[r] = s[slot(j + r * per)]; }
        if (logNs > 0) {
#pragma unroll
            for (int r = 1; r < R; r++) { v[r] = cmul(v[r], tw[(k * r) << sh]); }
        }
        Dft<R>::run(v);
        const int j0 = ((j - k) << LOGR) + k;
        float2* d = dst + b * LP;
#pragma unroll
        for (int r = 0; r < R; r++) { d[slot(j0 + r * Ns)] = v[r]; }
    }
}

// all stages of a length-L transform; radix plan packed 4 bits per stage (log2 of the radix), first stage in the low nibble.
// Returns the buffer holding the result.
__device__ float2* fftRows(float2* a, float2* b, int logL, unsigned plan, int LP, int nb, const float2* __restrict__ tw) {
    int logNs = 0;
    float2 *src = a, *dst = b;
    while (plan) {
        const int lr = plan & 15;
        plan >>= 4;
        switch (lr) {
        case 4: stockhamStage<16, 4>(src, dst, logL, logNs, LP, nb, tw); break;
        case 3: stockhamStage<8, 3>(src, dst, logL, logNs, LP, nb, tw); break;
        case 2: stockhamStage<4, 2>(src, dst, logL, logNs, LP, nb, tw); break;
        default: stockhamStage<2, 1>(src, dst, logL, logNs, LP, nb, tw); break;
        }
        __syncthreads();
        logNs += lr;
        float2* t = src; src = dst; dst = t;
    }
    return src;
}

// 8-byte asynchronous global->shared copy; src_size 0 zero-fills without reading
__device__ __forceinline__ void cpAsync8(float2* smemDst, const float2* gsrc, bool valid) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smemDst);
    const int sz = valid ? 8 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(d), "l"(gsrc), "r"(sz) : "memory");
}
__device__ __forceinline__ void cpAsyncWaitAll() {
    asm volatile("cp.async.commit_group;\n" ::: "memory");
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}
// address of frame sample i (the frame may straddle the carry buffer and the current block)
__device__ __forceinline__ const float2* frameAddr(const FrameDesc& f, int i) {
    return (i < f.len0) ? reinterpret_cast<const float2*>(f.src0) + i : reinterpret_cast<const float2*>(f.src1) + (i - f.len0);
}

__device__ __forceinline__ float2 frameSample(const FrameDesc& f, const float* __restrict__ win, int i, int nz) {
    if (i >= nz) { return make_float2(0.f, 0.f); }
    const float2 x = (i < f.len0) ? __ldg(reinterpret_cast<const float2*>(f.src0) + i)
                                  : __ldg(reinterpret_cast<const float2*>(f.src1) + (i - f.len0));
    const float w = win ? __ldg(win + i) : 1.0f;
    return make_float2(x.x * w, x.y * w);
}

__device__ __forceinline__ float powerDb(float2 X, float invN) {
    const float r = X.x * invN, q = X.y * invN;
    // MUFU.LG2: absolute error < 1e-6 in log2, i.e. < 4e-6 dB; an exact zero gives -inf and is clamped like the
    // reference's generic power-spectrum kernel does (log2 limited to +-127)
    float l = __log2f(r * r + q * q);
    l = fminf(fmaxf(l, -127.0f), 127.0f);
    return 3.01029995663981209120f * l;
}

// ---- N <= 8192: one CTA per frame
__global__ void __launch_bounds__(kThreads) fftSingleKernel(const FrameDesc* __restrict__ frames, const float* __restrict__ win, int nz,
                                                            int N, int logN, unsigned plan, const float2* __restrict__ tw,
                                                            float* __restrict__ outDb, cf32* __restrict__ outSpec) {
    extern __shared__ __align__(16) float2 sm[];
    const int LP = slot(N) + 1;
    float2* bufA = sm;
    float2* bufB = sm + LP;
    const FrameDesc f = frames[blockIdx.x];
    float2* sTw = bufB + LP;                     // W_N^m, m < N
    for (int m = threadIdx.x; m < N; m += kThreads) { sTw[m] = __ldg(tw + m); }
#pragma unroll 8
    for (int i = threadIdx.x; i < N; i += kThreads) { bufA[slot(i)] = frameSample(f, win, i, nz); }
    __syncthreads();
    const float2* res = fftRows(bufA, bufB, logN, plan, LP, 1, sTw);
    const float invN = 1.0f / (float)N;
    for (int k = threadIdx.x; k < N; k += kThreads) {
        const float2 X = res[slot(k)];
        if (outDb) { outDb[(size_t)blockIdx.x * N + k] = powerDb(X, invN); }
        if (outSpec) { reinterpret_cast<float2*>(outSpec)[(size_t)blockIdx.x * N + k] = X; }
    }
}

// ---- N = N1 * N2, pass 1: C adjacent columns n2 per CTA, length-N1 transforms over n1 (input stride N2),
// then the inter-pass twiddle W_N^(n2*k1); result T[k1][n2].
__global__ void __launch_bounds__(kThreads, 3) fftPass1Kernel(const FrameDesc* __restrict__ frames, const float* __restrict__ win, int nz,
                                                           int logN1, int logN2, int logC, unsigned plan, const float2* __restrict__ twN1,
                                                           const float2* __restrict__ twLo, const float2* __restrict__ twHi, int logS,
                                                           cf32* __restrict__ scratch, int scratchFrames) {
    extern __shared__ __align__(16) float2 sm[];
    const int N1 = 1 << logN1, C = 1 << logC;
    const FrameDesc f = frames[blockIdx.y];
    const int c0 = blockIdx.x << logC;
    const int LP = slot(N1) + 1;
    float2* bufA = sm;
    float2* bufB = sm + C * LP;
    const int total = N1 << logC;
    float2* sTw = bufB + C * LP;                 // W_N1^m, m < N1
    float2* sLo = sTw + N1;                      // W_N^j, j < S
    float2* sHi = sLo + (1 << logS);             // W_N^(j*S), j < N/S
    // all of this thread's samples are put in flight at once (no register staging), then scaled by the window in place
    constexpr int kBatch = 8;
    for (int base = 0; base < total; base += kBatch * kThreads) {
#pragma unroll
        for (int u = 0; u < kBatch; u++) {
            const int idx = base + u * kThreads + threadIdx.x;
            if (idx < total) {
                const int c = idx & (C - 1), n1 = idx >> logC;
                const int i = (n1 << logN2) + c0 + c;
                const bool ok = i < nz;
                cpAsync8(bufA + c * LP + slot(n1), frameAddr(f, ok ? i : 0), ok);
            }
        }
        float wv[kBatch];
#pragma unroll
        for (int u = 0; u < kBatch; u++) {
            const int idx = base + u * kThreads + threadIdx.x;
            const int i = ((idx >> logC) << logN2) + c0 + (idx & (C - 1));
            wv[u] = (idx < total && i < nz && win) ? __ldg(win + i) : 1.0f;
        }
        if (base == 0) {
            for (int m = threadIdx.x; m < N1; m += kThreads) { sTw[m] = __ldg(twN1 + m); }
            for (int m = threadIdx.x; m < (1 << logS); m += kThreads) { sLo[m] = __ldg(twLo + m); }
            for (int m = threadIdx.x; m < (1 << (logN1 + logN2 - logS)); m += kThreads) { sHi[m] = __ldg(twHi + m); }
        }
        cpAsyncWaitAll();
#pragma unroll
        for (int u = 0; u < kBatch; u++) {
            const int idx = base + u * kThreads + threadIdx.x;
            if (idx < total) {
                float2* p = bufA + (idx & (C - 1)) * LP + slot(idx >> logC);
                const float2 x = *p;
                *p = make_float2(x.x * wv[u], x.y * wv[u]);
            }
        }
    }
    __syncthreads();
    const float2* res = fftRows(bufA, bufB, logN1, plan, LP, C, sTw);
    float2* T = reinterpret_cast<float2*>(scratch) + ((size_t)(blockIdx.y % scratchFrames) << (logN1 + logN2));
    const int maskS = (1 << logS) - 1;
#pragma unroll 4
    for (int idx = threadIdx.x; idx < total; idx += kThreads) {
        const int c = idx & (C - 1), k1 = idx >> logC;
        const int m = (c0 + c) * k1;                     // < N
        const float2 w = cmul(sLo[m & maskS], sHi[m >> logS]);
        T[((size_t)k1 << logN2) + c0 + c] = cmul(res[c * LP + slot(k1)], w);
    }
}

// pass 2: R adjacent rows k1 per CTA, length-N2 transforms over n2; X[k1 + N1*k2].
__global__ void __launch_bounds__(kThreads, 3) fftPass2Kernel(int logN1, int logN2, int logR, unsigned plan, const float2* __restrict__ twN2,
                                                           const cf32* __restrict__ scratch, int scratchFrames, int frameBase,
                                                           float* __restrict__ outDb, cf32* __restrict__ outSpec) {
    extern __shared__ __align__(16) float2 sm[];
    const int N2 = 1 << logN2, R = 1 << logR;
    const int frame = frameBase + blockIdx.y;
    const int r0 = blockIdx.x << logR;
    const int LP = slot(N2) + 1;
    float2* bufA = sm;
    float2* bufB = sm + R * LP;
    const int total = N2 << logR;
    const float2* T = reinterpret_cast<const float2*>(scratch) + ((size_t)(frame % scratchFrames) << (logN1 + logN2));
    float2* sTw = bufB + R * LP;                 // W_N2^m, m < N2
    for (int m = threadIdx.x; m < N2; m += kThreads) { sTw[m] = __ldg(twN2 + m); }
#pragma unroll 16
    for (int idx = threadIdx.x; idx < total; idx += kThreads) {
        const int n2 = idx & (N2 - 1), r = idx >> logN2;
        cpAsync8(bufA + r * LP + slot(n2), T + ((size_t)(r0 + r) << logN2) + n2, true);
    }
    cpAsyncWaitAll();
    __syncthreads();
    const float2* res = fftRows(bufA, bufB, logN2, plan, LP, R, sTw);
    const size_t N = (size_t)1 << (logN1 + logN2);
    const float invN = 1.0f / (float)N;
#pragma unroll 4
    for (int idx = threadIdx.x; idx < total; idx += kThreads) {
        const int r = idx & (R - 1), k2 = idx >> logR;
        const float2 X = res[r * LP + slot(k2)];
        const size_t k = ((size_t)k2 << logN1) + r0 + r;
        if (outDb) { outDb[(size_t)frame * N + k] = powerDb(X, invN); }
        if (outSpec) { reinterpret_cast<float2*>(outSpec)[(size_t)frame * N + k] = X; }
    }
}

// ---- register-resident passes for N1, N2 in {64, 128, 256} (N = 2^14 .. 2^16, the waterfall sizes of cfg2)
// A length-L transform with L = RA * RB is two Stockham stages.  Each thread loads the RA points of its first
// butterfly straight from global memory into registers (window applied on the way), the single exchange between
// the stages goes through shared memory once, and the RB results of the second butterfly leave from registers
// (inter-pass twiddle or |X|^2 -> dB applied on the way).  Shared memory is touched twice per point instead of
// ten times, there is one barrier per CTA, and every global access is a 64-128 byte contiguous run per half-warp.
template <int RA, int RB> struct TwoStage {
    static constexpr int L = RA * RB;
    static constexpr int TPC = RA > RB ? RA : RB;     // threads per column (row): stage 1 uses RB of them, stage 2 RA
    // columns (rows) per tile: 16 where the tile fits twice into shared memory beside a second CTA; 8 for the 512- and
    // 1024-point transforms (radix-32 butterflies hold 64 registers of data per thread, so a CTA is 256 threads)
    static constexpr int LOGC = (L > 256) ? 3 : 4;
    static constexpr int C = 1 << LOGC;
    static constexpr int NT = C * TPC;
    static constexpr int MINB = (L > 256) ? 1 : (512 / NT);
};

// pass 1: a tile = C adjacent columns n2 = c0 .. c0+C-1 of one frame, transforms over n1 (global stride N2).
// Persistent CTAs: CTA b owns column block b % (N2/C) and walks the frames b / (N2/C), += gridDim.x / (N2/C), so
// its window samples and inter-pass twiddles are loaded once; the next tile's samples are in flight (cp.async into the
// other half of a double buffer) while the current tile is transformed.  The exchange between the two stages reuses
// the tile's own buffer.  Tile layout [point p][column] with one spare row of C elements per RA points, so that the
// lanes of a warp (C columns x 32/C consecutive j) hit distinct banks in the stage-1 stores (p = j * RA + r) as well
// as in the contiguous accesses.
template <int RA, int RB>
__global__ void __launch_bounds__(TwoStage<RA, RB>::NT, TwoStage<RA, RB>::MINB)
fftColsKernel(const FrameDesc* __restrict__ frames, int nFrames, const float* __restrict__ win, int nz, int logN2,
              const float2* __restrict__ twN1, const float2* __restrict__ twLo, const float2* __restrict__ twHi, int logS,
              cf32* __restrict__ scratch, int scratchFrames) {
    using TS = TwoStage<RA, RB>;
    constexpr int L = TS::L, C = TS::C, LOGC = TS::LOGC, NT = TS::NT;
    constexpr int TILE = (L + RB) * C;                   // padded tile: slot(p) = (p + p / RA) * C
    extern __shared__ __align__(16) float2 sm[];         // raw[2][TILE]
    const int blocksPerFrame = 1 << (logN2 - LOGC);
    const int cb = blockIdx.x & (blocksPerFrame - 1);
    const int fStep = gridDim.x >> (logN2 - LOGC);
    const int c = threadIdx.x & (C - 1), j = threadIdx.x >> LOGC;
    const int n2 = (cb << LOGC) + c;
    auto slotOf = [](int pnt) { return (pnt + pnt / RA) * C; };

    // this thread's cp.async duties for a tile: elements e = threadIdx.x + u * NT of [n1][c]
    auto issue = [&](int frame, float2* dst) {
        const FrameDesc f = frames[frame];
        // element e = threadIdx.x + u * NT of [n1][c] is frame sample i0 + u * iStep
        const float2* p0 = reinterpret_cast<const float2*>(f.src0);
        const float2* p1 = reinterpret_cast<const float2*>(f.src1) - f.len0;
        int i = ((threadIdx.x >> LOGC) << logN2) + (cb << LOGC) + (threadIdx.x & (C - 1));
        const int iStep = (NT >> LOGC) << logN2;
        if (nz == (L << logN2) && (f.len0 == 0 || f.len0 >= nz)) {
            // the usual frame: as long as the transform and in one piece -- one base pointer, no bounds, constant strides
            const float2* g = (f.len0 == 0 ? reinterpret_cast<const float2*>(f.src1) : p0) + i;
            float2* d = dst + slotOf(threadIdx.x >> LOGC) + c;
            constexpr int PSTEP = NT >> LOGC;                       // points between a thread's consecutive copies
            static_assert(PSTEP % RA == 0 || RA % PSTEP == 0, "slot stride");
#pragma unroll
            for (int u = 0; u < (L * C) / NT; u++) {
                const int pnt = (threadIdx.x >> LOGC) + u * PSTEP;
                cpAsync8(dst + slotOf(pnt) + c, g + (size_t)u * iStep, true);
            }
            (void)d;
        }
        else {
#pragma unroll 4
            for (int u = 0; u < (L * C) / NT; u++) {
                const bool ok = i < nz;
                const int ii = ok ? i : 0;
                const int pnt = (threadIdx.x >> LOGC) + u * (NT >> LOGC);
                cpAsync8(dst + slotOf(pnt) + c, (ii < f.len0 ? p0 : p1) + ii, ok);
                i += iStep;
            }
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };

    // per-thread constants: window of the RA stage-1 inputs, inter-pass twiddle base and step
    float wv[RA];
    float2 g = make_float2(1.f, 0.f), g2 = g, wb = g;
    if (j < RB) {
#pragma unroll
        for (int r = 0; r < RA; r++) {
            const int i = ((j + r * RB) << logN2) + n2;
            wv[r] = (win && i < nz) ? __ldg(win + i) : 1.0f;
        }
    }
    if (j < RA) {
        const int maskS = (1 << logS) - 1;
        const int mb = n2 * j, mg = n2 * RA;             // both < N
        g = cmul(__ldg(twLo + (mg & maskS)), __ldg(twHi + (mg >> logS)));
        g2 = cmul(g, g);
        wb = cmul(__ldg(twLo + (mb & maskS)), __ldg(twHi + (mb >> logS)));
    }

    int frame = blockIdx.x >> (logN2 - LOGC);
    if (frame >= nFrames) { return; }
    issue(frame, sm);
    for (int it = 0; frame < nFrames; frame += fStep, it++) {
        float2* raw = sm + (it & 1) * TILE;
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        __syncthreads();                                 // tile landed; everyone is done with the previous tile
        if (frame + fStep < nFrames) { issue(frame + fStep, sm + ((it & 1) ^ 1) * TILE); }
        float2 v[RA > RB ? RA : RB];
        if (j < RB) {
#pragma unroll
            for (int r = 0; r < RA; r++) {
                const float2 x = raw[slotOf(j + r * RB) + c];
                v[r] = make_float2(x.x * wv[r], x.y * wv[r]);
            }
            Dft<RA>::run(v);
        }
        __syncthreads();                                 // all stage-1 reads done: the buffer becomes the exchange
        if (j < RB) {
#pragma unroll
            for (int r = 0; r < RA; r++) { raw[(j * RA + r + j) * C + c] = v[r]; }     // slotOf(j * RA + r), r < RA
        }
        __syncthreads();
        if (j < RA) {
#pragma unroll
            for (int r = 0; r < RB; r++) { v[r] = raw[(j + r * RA + r) * C + c]; }     // slotOf(j + r * RA), j < RA
#pragma unroll
            for (int r = 1; r < RB; r++) { v[r] = cmul(v[r], __ldg(twN1 + j * r)); }
            Dft<RB>::run(v);
            // inter-pass twiddle W_N^(n2*k1), k1 = j + r*RA: W^(n2*j) * (W^(n2*RA))^r by two interleaved recurrences
            // (even / odd r), at most RB/2 roundings deep
            float2* T = reinterpret_cast<float2*>(scratch) + (size_t)(frame % scratchFrames) * ((size_t)L << logN2) + ((size_t)j << logN2) + n2;
            const size_t tStep = (size_t)RA << logN2;
            float2 we = wb, wo = cmul(wb, g);
#pragma unroll
            for (int r = 0; r < RB; r += 2) {
                T[0] = cmul(v[r], we);
                T[tStep] = cmul(v[r + 1], wo);
                T += 2 * tStep;
                we = cmul(we, g2); wo = cmul(wo, g2);
            }
        }
    }
}

// pass 2: a tile = R adjacent rows k1 = r0 .. r0+R-1 of one frame's intermediate (contiguous in the scratch
// ring), transforms over n2; X[k1 + N1*k2].  Persistent CTAs walk the tiles in order with the next tile's copy
// (16-byte cp.async) in flight behind the current transform; the stage-2 twiddles stay in registers for the CTA's
// lifetime.
template <int RA, int RB>
__global__ void __launch_bounds__(TwoStage<RA, RB>::NT, TwoStage<RA, RB>::MINB)
fftRowsKernel(int logN1, int nTiles, const float2* __restrict__ twN2, const cf32* __restrict__ scratch, int frameBase,
              float* __restrict__ outDb, cf32* __restrict__ outSpec) {
    using TS = TwoStage<RA, RB>;
    constexpr int L = TS::L, R = TS::C, LOGR = TS::LOGC, NT = TS::NT;
    // exchange layout [point p][row] with one pad element per RA points: the stage-1 stores (lanes step p by RA) and the
    // stage-2 loads (lanes step the row) are both conflict-free.  Each buffer first receives the raw tile [row][n2].
    constexpr int BUF = (L * R + RB + 2) & ~1;
    extern __shared__ __align__(16) float2 sm[];
    const size_t N = (size_t)L << logN1;
    const float invN = 1.0f / (float)N;
    const int tilesPerFrame = 1 << (logN1 - LOGR);

    auto issue = [&](int tile, float2* dst) {
        const float4* src = reinterpret_cast<const float4*>(scratch) + (size_t)tile * (L * R / 2);
        const unsigned d0 = (unsigned)__cvta_generic_to_shared(dst);
#pragma unroll 4
        for (int u = 0; u < (L * R / 2) / NT; u++) {
            const int e = threadIdx.x + u * NT;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d0 + e * 16), "l"(src + e) : "memory");
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };

    const int j1 = threadIdx.x % RB, row1 = threadIdx.x / RB;      // stage 1: lanes walk n2
    const int row2 = threadIdx.x & (R - 1), j2 = threadIdx.x >> LOGR;  // stage 2: lanes walk the row (contiguous stores)
    // stage-2 twiddles W_L^(j2 * r): kept in registers where that is cheap, re-read (L1 hits) for the radix-32 forms
    constexpr bool kTwRegs = (RB <= 16);
    float2 tw[kTwRegs ? RB : 1];
    if constexpr (kTwRegs) {
#pragma unroll
        for (int r = 1; r < RB; r++) { tw[r] = (j2 < RA) ? __ldg(twN2 + j2 * r) : make_float2(1.f, 0.f); }
    }

    int tile = blockIdx.x;
    if (tile >= nTiles) { return; }
    issue(tile, sm);
    for (int it = 0; tile < nTiles; tile += gridDim.x, it++) {
        float2* buf = sm + (it & 1) * BUF;
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        __syncthreads();
        if (tile + (int)gridDim.x < nTiles) { issue(tile + gridDim.x, sm + ((it & 1) ^ 1) * BUF); }
        float2 v[RA > RB ? RA : RB];
        if (row1 < R) {
#pragma unroll
            for (int r = 0; r < RA; r++) { v[r] = buf[row1 * L + j1 + r * RB]; }
            Dft<RA>::run(v);
        }
        __syncthreads();
        if (row1 < R) {
#pragma unroll
            for (int r = 0; r < RA; r++) { buf[(j1 * RA + r) * R + j1 + row1] = v[r]; }
        }
        __syncthreads();
        if (j2 < RA) {
#pragma unroll
            for (int r = 0; r < RB; r++) { v[r] = buf[(j2 + r * RA) * R + r + row2]; }
#pragma unroll
            for (int r = 1; r < RB; r++) { v[r] = cmul(v[r], kTwRegs ? tw[kTwRegs ? r : 0] : __ldg(twN2 + j2 * r)); }
            Dft<RB>::run(v);
            const int frame = frameBase + (tile >> (logN1 - LOGR));
            const int r0 = (tile & (tilesPerFrame - 1)) << LOGR;
            const size_t k0 = (size_t)frame * N + ((size_t)j2 << logN1) + r0 + row2;
            const size_t kStep = (size_t)RA << logN1;
            if (outDb) {
                float* o = outDb + k0;
#pragma unroll
                for (int r = 0; r < RB; r++) { *o = powerDb(v[r], invN); o += kStep; }
            }
            if (outSpec) {
                float2* o = reinterpret_cast<float2*>(outSpec) + k0;
#pragma unroll
                for (int r = 0; r < RB; r++) { *o = v[r]; o += kStep; }
            }
        }
    }
}

template <int RA, int RB>
int launchCols(int nf, int N2, cudaStream_t s, const FrameDesc* frames, const float* win, int nz, int logN2, const float2* twN1,
               const float2* twLo, const float2* twHi, int logS, cf32* scratch, int scratchFrames) {
    using TS = TwoStage<RA, RB>;
    const size_t smem = (size_t)2 * (TS::L + RB) * TS::C * sizeof(float2);
    B200_CUDA(cudaFuncSetAttribute(fftColsKernel<RA, RB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // persistent grid: a whole number of column-block sets, as many CTAs per SM as fit
    const int blocksPerFrame = N2 >> TS::LOGC;
    const int ctasPerSm = fftCtasPerSm > 0 ? std::min(fftCtasPerSm, TS::MINB) : TS::MINB;
    int sets = std::max(1, (148 * ctasPerSm) / blocksPerFrame);
    sets = std::min(sets, nf);
    fftColsKernel<RA, RB><<<sets * blocksPerFrame, TS::NT, smem, s>>>(frames, nf, win, nz, logN2, twN1, twLo, twHi, logS, scratch, scratchFrames);
    return B200_OK;
}
template <int RA, int RB>
int launchRows(int nf, int N1, cudaStream_t s, int logN1, const float2* twN2, const cf32* scratch, int frameBase, float* outDb, cf32* outSpec) {
    using TS = TwoStage<RA, RB>;
    constexpr int BUF = (TS::L * TS::C + RB + 2) & ~1;
    const size_t smem = (size_t)2 * BUF * sizeof(float2);
    B200_CUDA(cudaFuncSetAttribute(fftRowsKernel<RA, RB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int nTiles = nf * (N1 >> TS::LOGC);
    const int ctasPerSm = fftCtasPerSm > 0 ? std::min(fftCtasPerSm, TS::MINB) : TS::MINB;
    const int grid = std::min(nTiles, 148 * ctasPerSm);
    fftRowsKernel<RA, RB><<<grid, TS::NT, smem, s>>>(logN1, nTiles, twN2, scratch, frameBase, outDb, outSpec);
    return B200_OK;
}

// ---- N = 2^10 .. 2^14 in one pass: one CTA per frame, three register-resident Stockham stages
// N = R0 * R1 * R2.  Every thread keeps PT = N / T points in registers: the first stage loads them straight from the frame
// (window applied on the way; lanes read consecutive samples), stages hand over through one shared-memory buffer
// (read everything, barrier, write in Stockham order, barrier), and the last stage's results leave from registers as dB
// values (lanes write consecutive bins).  The samples cross HBM once and the rows once; nothing else does.
//   stage with radix R after stages of total size Ns: butterfly j < N/R, k = j mod Ns,
//   in x[j + r N/R] * W_{Ns R}^{k r}, out y[(j - k) R + k + r Ns]
template <int R0, int R1, int R2> struct DirectCfg {
    static constexpr int N = R0 * R1 * R2;
    static constexpr int PT = (R0 == 32 || R1 == 32 || R2 == 32) ? 32 : 16;
    static constexpr int T = N / PT;
    // resident CTAs per SM the register allocation is held to (one frame per CTA and no prefetch: the overlap between
    // a frame's loads and another frame's butterflies comes from the other resident CTAs)
    static constexpr int F = (T >= 256) ? 1 : 256 / T;      // frames a CTA transforms side by side (threadIdx.y): CTAs of 256 threads
    static constexpr int MINB = (PT == 32) ? 1 : 2;
};

template <int R, int PT, int T, int N, int NS, bool FROM_FRAME, bool TO_GLOBAL>
__device__ __forceinline__ void directStage(float2* __restrict__ buf, const FrameDesc& f, const float* __restrict__ wv, int nz,
                                            const float2 (*twr)[R], float* __restrict__ outDb, float2* __restrict__ outSpec, float invN,
                                            const float2* __restrict__ pre = nullptr) {
    constexpr int NB = PT / R;               // butterflies per thread
    constexpr int PER = N / R;               // butterflies per frame = input stride
    float2 v[NB][R];
#pragma unroll
    for (int m = 0; m < NB; m++) {
        const int j = threadIdx.x + m * T;
#pragma unroll
        for (int r = 0; r < R; r++) {
            if constexpr (FROM_FRAME) {
                const int i = j + r * PER;
                float2 x = make_float2(0.f, 0.f);
                if (pre) { x = pre[(m * R + r) * T + threadIdx.x]; }       // staged ahead by this very thread (zero filled past nz)
                else if (i < nz) { x = __ldg(frameAddr(f, i)); }
                v[m][r] = make_float2(x.x * wv[m * R + r], x.y * wv[m * R + r]);
            }
            else { v[m][r] = buf[slot(j + r * PER)]; }
        }
        if constexpr (NS > 1) {
#pragma unroll
            for (int r = 1; r < R; r++) { v[m][r] = cmul(v[m][r], twr[m][r]); }
        }
        Dft<R>::run(v[m]);
    }
    if constexpr (!FROM_FRAME) { __syncthreads(); }       // every thread has read its inputs: the buffer may be overwritten
#pragma unroll
    for (int m = 0; m < NB; m++) {
        const int j = threadIdx.x + m * T;
        const int k = j & (NS - 1);
        const int o = ((j - k) * R) + k;
#pragma unroll
        for (int r = 0; r < R; r++) {
            if constexpr (TO_GLOBAL) {
                if (outDb) { outDb[o + r * NS] = powerDb(v[m][r], invN); }
                if (outSpec) { outSpec[o + r * NS] = v[m][r]; }
            }
            else { buf[slot(o + r * NS)] = v[m][r]; }
        }
    }
    if constexpr (!TO_GLOBAL) { __syncthreads(); }
}

// Persistent CTAs: a thread's butterflies sit at the same places in every frame, so its window samples and the
// twiddles of its second- and third-stage butterflies are loaded once and stay in registers for the CTA's lifetime.
// PRE (B200_OPT_FFT_DIRECT_PREFETCH, off): every thread brings its own PT samples of the CTA's NEXT frame into shared memory
// with cp.async while the current frame's second and third stage run (only the copying thread reads them back: no barrier,
// just its own wait_group), so that a frame's transform does not start with a round trip to HBM that only the other
// resident CTA can hide.  Measured (13 Mi points per batch, us without / with): 2048: 77.8 / 88.4, 4096: 73.0 / 77.4,
// 8192: 88.8 / 99.7 -- slower at every size: the kernel does not wait for those loads, it is short of issue slots and
// shared-memory bandwidth, and the staged copy adds to both (profiles/r02_fft_direct_prefetch.txt).
template <int R0, int R1, int R2, bool PRE>
__global__ void __launch_bounds__(DirectCfg<R0, R1, R2>::T * DirectCfg<R0, R1, R2>::F, DirectCfg<R0, R1, R2>::MINB)
fftDirectKernel(const FrameDesc* __restrict__ frames, int nFrames, const float* __restrict__ win, int nz, const float2* __restrict__ tw,
                float* __restrict__ outDb, cf32* __restrict__ outSpec) {
    using DC = DirectCfg<R0, R1, R2>;
    constexpr int N = DC::N, PT = DC::PT, T = DC::T;
    extern __shared__ __align__(16) float2 sm[];
    const float invN = 1.0f / (float)N;
    float wv[PT];
    float2 tw1[PT / R1][R1], tw2[PT / R2][R2];
#pragma unroll
    for (int m = 0; m < PT / R0; m++) {
#pragma unroll
        for (int r = 0; r < R0; r++) {
            const int i = threadIdx.x + m * T + r * (N / R0);
            wv[m * R0 + r] = (win && i < nz) ? __ldg(win + i) : 1.0f;
        }
    }
#pragma unroll
    for (int m = 0; m < PT / R1; m++) {
        const int k = (threadIdx.x + m * T) & (R0 - 1);
#pragma unroll
        for (int r = 0; r < R1; r++) { tw1[m][r] = __ldg(tw + (k * r) * R2); }              // W_{R0 R1}^(k r)
    }
#pragma unroll
    for (int m = 0; m < PT / R2; m++) {
        const int k = threadIdx.x + m * T;                                                    // < N / R2
#pragma unroll
        for (int r = 0; r < R2; r++) { tw2[m][r] = __ldg(tw + k * r); }                     // W_N^(k r)
    }
    float2* buf = sm + threadIdx.y * (slot(N) + 1);
    float2* pre = PRE ? sm + DC::F * (slot(N) + 1) + threadIdx.y * N : nullptr;     // sample (m, r) of thread t at (m * R0 + r) * T + t
    auto stageAhead = [&](int frame) {
        if (frame < nFrames) {
            const FrameDesc f = frames[frame];
#pragma unroll
            for (int k = 0; k < PT; k++) {
                const int i = threadIdx.x + (k / R0) * T + (k % R0) * (N / R0);
                const bool ok = i < nz;
                cpAsync8(pre + k * T + threadIdx.x, frameAddr(f, ok ? i : 0), ok);
            }
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };
    if (PRE) { stageAhead(blockIdx.x * DC::F + threadIdx.y); }
    for (int base = blockIdx.x * DC::F; base < nFrames; base += gridDim.x * DC::F) {
        const int frame = base + threadIdx.y;
        const bool live = frame < nFrames;                // a CTA's last group may be short: idle slots still meet the barriers
        const FrameDesc f = frames[live ? frame : base];
        float* oDb = (outDb && live) ? outDb + (size_t)frame * N : nullptr;
        float2* oSp = (outSpec && live) ? reinterpret_cast<float2*>(outSpec) + (size_t)frame * N : nullptr;
        if (PRE) { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }
        directStage<R0, PT, T, N, 1, true, false>(buf, f, wv, nz, nullptr, nullptr, nullptr, invN, pre);
        // (behind the first stage's closing barrier: the values read from `pre` have long been consumed)
        if (PRE) { stageAhead(frame + gridDim.x * DC::F); }
        directStage<R1, PT, T, N, R0, false, false>(buf, f, nullptr, nz, tw1, nullptr, nullptr, invN);
        directStage<R2, PT, T, N, R0 * R1, false, true>(buf, f, nullptr, nz, tw2, oDb, oSp, invN);
        __syncthreads();       // the next frame's first stage writes the buffer this frame's last stage has just read
    }
}

template <int R0, int R1, int R2, bool PRE>
int launchDirectAs(int nFrames, cudaStream_t s, const FrameDesc* frames, const float* win, int nz, const float2* tw, float* outDb, cf32* outSpec) {
    using DC = DirectCfg<R0, R1, R2>;
    const size_t smem = (size_t)DC::F * ((slot(DC::N) + 1) + (PRE ? DC::N : 0)) * sizeof(float2);
    static int perSm = 0;
    // (function attributes are per device: set on every use)
    B200_CUDA(cudaFuncSetAttribute(fftDirectKernel<R0, R1, R2, PRE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (perSm == 0) {
        B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, fftDirectKernel<R0, R1, R2, PRE>, DC::T * DC::F, smem));
        if (perSm < 1) { perSm = 1; }
    }
    const int groups = (nFrames + DC::F - 1) / DC::F;
    const int grid = std::min(groups, 148 * perSm);
    fftDirectKernel<R0, R1, R2, PRE><<<grid, dim3(DC::T, DC::F), smem, s>>>(frames, nFrames, win, nz, tw, outDb, outSpec);
    B200_CHECK_LAUNCH();
    countLaunch();
    return B200_OK;
}
template <int R0, int R1, int R2>
int launchDirect(int nFrames, cudaStream_t s, const FrameDesc* frames, const float* win, int nz, const float2* tw, float* outDb, cf32* outSpec) {
    return fftDirectPrefetch ? launchDirectAs<R0, R1, R2, true>(nFrames, s, frames, win, nz, tw, outDb, outSpec)
                             : launchDirectAs<R0, R1, R2, false>(nFrames, s, frames, win, nz, tw, outDb, outSpec);
}

int uploadTwiddles(float2** d, int count, int N, int stride) {
    std::vector<float2> h((size_t)std::max(count, 1));
    for (int j = 0; j < count; j++) {
        const double a = -2.0 * 3.14159265358979323846 * (double)j * (double)stride / (double)N;
        h[j] = make_float2((float)cos(a), (float)sin(a));
    }
    B200_CUDA(cudaMalloc((void**)d, h.size() * sizeof(float2)));
    B200_CUDA(cudaMemcpy(*d, h.data(), h.size() * sizeof(float2), cudaMemcpyHostToDevice));
    return B200_OK;
}

// log2 radices, as even as possible, at most 4 bits (radix 16) per stage; first stage in the low nibble
unsigned radixPlan(int logL) {
    const int stages = (logL + 3) / 4;
    unsigned plan = 0;
    int left = logL;
    for (int s = 0; s < stages; s++) {
        const int lr = (left + (stages - s) - 1) / (stages - s);
        plan |= (unsigned)lr << (4 * s);
        left -= lr;
    }
    return plan;
}

int log2i(int v) { int l = 0; while ((1 << l) < v) { l++; } return l; }

// two ping-pong buffers of nb rows + `extra` twiddle entries
size_t rowsSmem(int L, int nb, int extra = 0) { return ((size_t)2 * nb * (slot(L) + 1) + extra) * sizeof(float2); }

constexpr size_t kSmemSoft = 76 * 1024;    // 3 CTAs per SM
constexpr size_t kSmemHard = 200 * 1024;

// rows (or columns) per CTA: as many as wanted within the soft cap, but never fewer than 8 (32-byte segments)
// unless even that exceeds the hard cap
int pickBatchLog(int L, int wantLog, int extra) {
    int lg = wantLog;
    while (lg > 3 && rowsSmem(L, 1 << lg, extra) > kSmemSoft) { lg--; }
    while (lg > 0 && rowsSmem(L, 1 << lg, extra) > kSmemHard) { lg--; }
    return lg;
}

} // namespace

int FftPlan::create(int N_, int device_) {
    destroy();
    device = device_;
    N = N_;
    logN = log2i(N);
    if ((1 << logN) != N || logN < 5 || logN > 25) { setError("fft size %d is not a power of two in [32, 33554432]", N); return B200_EINVAL; }
    direct = (logN >= 10 && logN <= fftDirectMaxLog);
    if (logN <= 9 || direct) {  // up to 512 points: generic one-CTA kernel; 2^10 .. 2^14: one CTA per frame, three register-resident stages
        N1 = 1; N2 = N;
        B200_TRY(uploadTwiddles(&dTwN2, N, N, 1));
    }
    else {
        const int l2 = (logN + 1) / 2;
        N2 = 1 << l2; N1 = N >> l2;
        B200_TRY(uploadTwiddles(&dTwN1, N1, N1, 1));
        B200_TRY(uploadTwiddles(&dTwN2, N2, N2, 1));
        logS = l2; S = 1 << logS;
        B200_TRY(uploadTwiddles(&dTwLo, S, N, 1));
        B200_TRY(uploadTwiddles(&dTwHi, N >> logS, N, S));
    }
    return B200_OK;
}

void FftPlan::destroy() {
    float2** t[] = { &dTwN1, &dTwN2, &dTwLo, &dTwHi };
    for (auto p : t) { if (*p) { cudaFree(*p); *p = nullptr; } }
    if (dScratch) { cudaFree(dScratch); dScratch = nullptr; }
    scratchFrames = 0;
}

int FftPlan::ensureScratch(int frames) {
    if (N1 == 1) { return B200_OK; }
    // at most 256 MB of intermediates per launch pair (measured: one long persistent launch beats several L2-sized ones)
    int want = (int)std::max<size_t>(1, std::min<size_t>((size_t)frames, ((size_t)256 << 20) / ((size_t)N * sizeof(cf32))));
    if (want <= scratchFrames) { return B200_OK; }
    if (dScratch) { cudaFree(dScratch); dScratch = nullptr; }
    B200_CUDA(cudaMalloc((void**)&dScratch, (size_t)want * N * sizeof(cf32)));
    scratchFrames = want;
    return B200_OK;
}

int FftPlan::run(const FrameDesc* dFrames, int nFrames, const float* dWindow, int nz, float* outDb, cf32* outSpec, cudaStream_t s) {
    if (nFrames <= 0) { return B200_OK; }
    // measurement hook: all passes of this call between one pair of events.  Algorithmic traffic per frame: the nz
    // samples in the window are read once (8 nz), the row is written once (4 N dB values, or 8 N for a complex spectrum)
    size_t slot = 0;
    B200_TRY(profBegin(PROF_FFT, s, (double)nFrames * (8.0 * nz + (outDb ? 4.0 : 0.0) * N + (outSpec ? 8.0 : 0.0) * N),
                       (double)nFrames * 5.0 * (double)N * (double)logN, &slot));
    const int rc = runPasses(dFrames, nFrames, dWindow, nz, outDb, outSpec, s);
    B200_TRY(profEnd(PROF_FFT, slot, s));
    return rc;
}

int FftPlan::runPasses(const FrameDesc* dFrames, int nFrames, const float* dWindow, int nz, float* outDb, cf32* outSpec, cudaStream_t s) {
    if (N1 == 1 && direct) {
        switch (logN) {
        case 10: return launchDirect<16, 8, 8>(nFrames, s, dFrames, dWindow, nz, dTwN2, outDb, outSpec);
        case 11: return launchDirect<16, 16, 8>(nFrames, s, dFrames, dWindow, nz, dTwN2, outDb, outSpec);
        case 12: return launchDirect<16, 16, 16>(nFrames, s, dFrames, dWindow, nz, dTwN2, outDb, outSpec);
        case 13: return launchDirect<32, 16, 16>(nFrames, s, dFrames, dWindow, nz, dTwN2, outDb, outSpec);
        default: return launchDirect<32, 32, 16>(nFrames, s, dFrames, dWindow, nz, dTwN2, outDb, outSpec);
        }
    }
    if (N1 == 1) {
        const size_t smem = rowsSmem(N, 1, N);
        B200_CUDA(cudaFuncSetAttribute(fftSingleKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        fftSingleKernel<<<nFrames, kThreads, smem, s>>>(dFrames, dWindow, nz, N, logN, radixPlan(logN), dTwN2, outDb, outSpec);
        B200_CHECK_LAUNCH();
        countLaunch();
        return B200_OK;
    }
    B200_TRY(ensureScratch(nFrames));
    const int logN1 = log2i(N1), logN2 = log2i(N2);
    if (logN1 >= 4 && logN1 <= 10 && logN2 >= 4 && logN2 <= 10) {
        // register-resident passes (16 = 4x4, 32 = 8x4, 64 = 8x8, 128 = 16x8, 256 = 16x16, 512 = 32x16, 1024 = 32x32)
        for (int f0 = 0; f0 < nFrames; f0 += scratchFrames) {
            const int nf = std::min(scratchFrames, nFrames - f0);
            switch (logN1) {
            case 4: B200_TRY((launchCols<4, 4>(nf, N2, s, dFrames + f0, dWindow, nz, logN2, dTwN1, dTwLo, dTwHi, logS, dScratch, scratchFrames))); break;
            case 5: B200_TRY((launchCols<8, 4>(nf, N2, s, dFrames + f0, dWindow, nz, logN2, dTwN1, dTwLo, dTwHi, logS, dScratch, scratchFrames))); break;
            case 6: B200_TRY((launchCols<8, 8>(nf, N2, s, dFrames + f0, dWindow, nz, logN2, dTwN1, dTwLo, dTwHi, logS, dScratch, scratchFrames))); break;
            case 7: B200_TRY((launchCols<16, 8>(nf, N2, s, dFrames + f0, dWindow, nz, logN2, dTwN1, dTwLo, dTwHi, logS, dScratch, scratchFrames))); break;
            case 8: B200_TRY((launchCols<16, 16>(nf, N2, s, dFrames + f0, dWindow, nz, logN2, dTwN1, dTwLo, dTwHi, logS, dScratch, scratchFrames))); break;
            case 9: B200_TRY((launchCols<32, 16>(nf, N2, s, dFrames + f0, dWindow, nz, logN2, dTwN1, dTwLo, dTwHi, logS, dScratch, scratchFrames))); break;
            default: B200_TRY((launchCols<32, 32>(nf, N2, s, dFrames + f0, dWindow, nz, logN2, dTwN1, dTwLo, dTwHi, logS, dScratch, scratchFrames))); break;
            }
            B200_CHECK_LAUNCH();
            switch (logN2) {
            case 4: B200_TRY((launchRows<4, 4>(nf, N1, s, logN1, dTwN2, dScratch, f0, outDb, outSpec))); break;
            case 5: B200_TRY((launchRows<8, 4>(nf, N1, s, logN1, dTwN2, dScratch, f0, outDb, outSpec))); break;
            case 6: B200_TRY((launchRows<8, 8>(nf, N1, s, logN1, dTwN2, dScratch, f0, outDb, outSpec))); break;
            case 7: B200_TRY((launchRows<16, 8>(nf, N1, s, logN1, dTwN2, dScratch, f0, outDb, outSpec))); break;
            case 8: B200_TRY((launchRows<16, 16>(nf, N1, s, logN1, dTwN2, dScratch, f0, outDb, outSpec))); break;
            case 9: B200_TRY((launchRows<32, 16>(nf, N1, s, logN1, dTwN2, dScratch, f0, outDb, outSpec))); break;
            default: B200_TRY((launchRows<32, 32>(nf, N1, s, logN1, dTwN2, dScratch, f0, outDb, outSpec))); break;
            }
            B200_CHECK_LAUNCH();
            countLaunch(2);
        }
        return B200_OK;
    }
    const int extra1 = N1 + S + (N >> logS), extra2 = N2;      // twiddle tables staged in shared memory
    const int logC = pickBatchLog(N1, 4, extra1);      // up to 16 columns: 128-byte global segments
    const int logR = pickBatchLog(N2, 5, extra2);      // up to 32 rows: 128-byte dB segments
    const size_t smem1 = rowsSmem(N1, 1 << logC, extra1), smem2 = rowsSmem(N2, 1 << logR, extra2);
    B200_CUDA(cudaFuncSetAttribute(fftPass1Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
    B200_CUDA(cudaFuncSetAttribute(fftPass2Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
    const unsigned plan1 = radixPlan(logN1), plan2 = radixPlan(logN2);
    // frames are processed in groups that fit the scratch ring, pass 2 right behind pass 1 (L2 reuse);
    // the ring index is (frame % scratchFrames) and groups start at multiples of scratchFrames
    for (int f0 = 0; f0 < nFrames; f0 += scratchFrames) {
        const int nf = std::min(scratchFrames, nFrames - f0);
        fftPass1Kernel<<<dim3(N2 >> logC, nf), kThreads, smem1, s>>>(dFrames + f0, dWindow, nz, logN1, logN2, logC, plan1, dTwN1, dTwLo, dTwHi,
                                                                     logS, dScratch, scratchFrames);
        B200_CHECK_LAUNCH();
        fftPass2Kernel<<<dim3(N1 >> logR, nf), kThreads, smem2, s>>>(logN1, logN2, logR, plan2, dTwN2, dScratch, scratchFrames, f0, outDb, outSpec);
        B200_CHECK_LAUNCH();
        countLaunch(2);
    }
    return B200_OK;
}

} // namespace b200

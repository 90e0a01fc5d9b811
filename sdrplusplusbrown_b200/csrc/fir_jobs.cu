// Generic streaming-filter kernels: decimating FIR / FIR (filter/decimating_fir.h:45-68,
// filter/fir.h:64-92) and polyphase rational resampler (multirate/polyphase_resampler.h:69-99),
// many independent jobs per launch (grid.y = job).  These serve the low-rate tail of every VFO
// (stages 2.., resampler, channel filter), the front-end decimator and the demodulator filters;
// the full-rate first stage has its own fused kernel (xlate_decim.cu).
//
// Three ways a CTA can work on a job (FirJob::plan, chosen on the host by firJobPlan):
//   blocked   (DECIM_FIR, power-of-two decimation D <= 8): each thread owns 8 consecutive outputs.  A decimating FIR is
//             the sum of D ordinary FIRs over the polyphase components x[D*m + v]; per component the thread slides an
//             8-sample register window along the staged input, so one shared-memory load and one tap feed 16 FFMAs.
//             Long filters are split over `plan` adjacent lanes (partial sums folded with shuffles) so that short
//             blocks still fill the GPU.
//   lanes     (POLYPHASE): 8 lanes per output; a bank row is read as 128-byte runs (one float4 per lane) instead of
//             one row per thread, partial sums folded with three shuffles.
//   plan      (DECIM_FIR whose taps are one of the later-stage tables of the pow-2 decimation plans: (D,T) = (2,69),
//             (2,12), (4,27), (8,54), (8,44), (8,36)): 4 or 8 outputs per thread, the loop over the thread's input
//             window fully unrolled with the taps as constant-bank operands -- each staged sample is loaded once and
//             feeds every output it belongs to (T/D of them), against one load per tap in the simple path.
//   simple    one output per thread; kept for shapes the others do not cover.
#include "fir_jobs.cuh"

#include <algorithm>
#include <vector>

namespace b200 {

namespace {

constexpr int kThreads = 128;
constexpr int kR = 8;                       // outputs per thread, blocked path
constexpr int kPolyLanes = 8;               // lanes per output, polyphase path
constexpr int kPolyOutPerCta = 128;         // outputs per CTA, polyphase path (8 rounds of 16)
constexpr int kSimpleOutPerCta = kThreads;
constexpr size_t kSmemCap = 160 * 1024;

// later-stage coefficient tables of the decimation plans in the constant bank (the same generated file the fused
// first-stage kernel includes)
// (as function-local constant arrays: under the unrolled tap loops of planStage the loads fold and every tap is the
// immediate of its FFMA; see xlate_decim.cu)
#define DECIM_TABLE_U32(name, count) constexpr int n_##name = count; static __device__ __forceinline__ float imm_##name(int i) { const uint32_t t[count] = {
#define DECIM_ROW(...) __VA_ARGS__,
#define DECIM_TABLE_END }; return __uint_as_float(t[i]); }
#define DECIM_PLAN_BEGIN(ratio, nstages)
#define DECIM_STAGE(index, decim, name, ntaps)
#define DECIM_PLAN_END
#include "decim_tables.inc"
#undef DECIM_TABLE_U32
#undef DECIM_ROW
#undef DECIM_TABLE_END
#undef DECIM_PLAN_BEGIN
#undef DECIM_STAGE
#undef DECIM_PLAN_END

template <int D, int T> struct PlanStage;
#define B200_PLAN_STAGE(idx, D_, T_)                                                                         \
    static_assert(n_kDecimTaps_##idx == (T_), "decim_tables.inc changed: stage table size"); \
    template <> struct PlanStage<D_, T_> { static __device__ __forceinline__ float tap(int i) { return imm_kDecimTaps_##idx(i); } };
B200_PLAN_STAGE(0, 2, 69)
B200_PLAN_STAGE(1, 2, 12)
B200_PLAN_STAGE(2, 4, 27)
B200_PLAN_STAGE(3, 8, 54)
B200_PLAN_STAGE(4, 8, 44)
B200_PLAN_STAGE(5, 8, 36)
#undef B200_PLAN_STAGE
constexpr int kPlanBase = 1000;             // FirJob::plan = kPlanBase + D * 100 + T for the plan-table path
__host__ __device__ constexpr int planOutPerThread(int D) { return D >= 8 ? 4 : 8; }

template <int EF> struct Elem;
template <> struct Elem<2> {
    using T = float2;
    static __device__ __forceinline__ T zero() { return make_float2(0.f, 0.f); }
    // complex sample x real tap: one packed FMA (both components round exactly as two scalar fmaf would)
    static __device__ __forceinline__ void fma(T& a, float t, T x) { a = fma2(make_float2(t, t), x, a); }
    static __device__ __forceinline__ void add(T& a, T b) { a.x += b.x; a.y += b.y; }
    static __device__ __forceinline__ T shflXor(T a, int m) { return make_float2(__shfl_xor_sync(0xffffffffu, a.x, m), __shfl_xor_sync(0xffffffffu, a.y, m)); }
    // shared-memory slot of window position p (simple path): one pad element per 16 keeps lanes that step by 1, 2, 4
    // or 8 samples on distinct banks
    static __device__ __forceinline__ int slot(int p) { return p + (p >> 4); }
};
template <> struct Elem<1> {
    using T = float;
    static __device__ __forceinline__ T zero() { return 0.f; }
    static __device__ __forceinline__ void fma(T& a, float t, T x) { a = fmaf(t, x, a); }
    static __device__ __forceinline__ void add(T& a, T b) { a += b; }
    static __device__ __forceinline__ T shflXor(T a, int m) { return __shfl_xor_sync(0xffffffffu, a, m); }
    static __device__ __forceinline__ int slot(int p) { return p + (p >> 5); }
};

// sample at position p of the job's [hist | data] work buffer; zero beyond the end of the block (only ever multiplied
// by padding taps or belonging to outputs that are not stored)
template <int EF>
__device__ __forceinline__ typename Elem<EF>::T loadPos(const FirJob& j, int p) {
    using T = typename Elem<EF>::T;
    if (p < j.H) { return __ldg(reinterpret_cast<const T*>(j.hist) + p); }
    if (p - j.H < j.nIn) { return __ldg(reinterpret_cast<const T*>(j.data) + (p - j.H)); }
    return Elem<EF>::zero();
}

__device__ __forceinline__ int ilog2(int v) { return 31 - __clz(v); }

// stages work-buffer positions [base, base + span) into shared memory, position i in slot i + (i >> padShift).  A window
// that lies wholly in the new block (every CTA but the first few of a job) is copied by a branch-free loop unrolled
// eight deep, so that a thread has eight loads in flight instead of one.
template <int EF>
__device__ __forceinline__ void stageWindow(const FirJob& j, typename Elem<EF>::T* win, int base, int span, int padShift) {
    using T = typename Elem<EF>::T;
    if (base >= j.H && base - j.H + span <= j.nIn) {
        const T* src = reinterpret_cast<const T*>(j.data) + (base - j.H);
        if constexpr (EF == 2) {
            // complex samples: asynchronous 8-byte copies, the whole window in flight at once.  The CTAs of a large bank's
            // later stages are many and short (cfg4's first wave: 262144 CTAs of one 34 KB window each); with the window
            // staged through registers, eight loads per thread at a time, an SM had about 24 KB in flight and the wave read
            // its 2.4 GB at 3.5 TB/s (ncu: warps 34 % active, issue slots 23 %, FMA pipe 15 % busy).  (16-byte loads
            // through registers changed nothing: the width of the loads was not the limit.)
            for (int i = threadIdx.x; i < span; i += kThreads) {
                const unsigned d = (unsigned)__cvta_generic_to_shared(win + i + (i >> padShift));
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(d), "l"(src + i) : "memory");
            }
            asm volatile("cp.async.commit_group;\n" ::: "memory");
            asm volatile("cp.async.wait_group 0;\n" ::: "memory");
            return;
        }
        int i = threadIdx.x;
        for (; i + 7 * kThreads < span; i += 8 * kThreads) {
            T v[8];
#pragma unroll
            for (int u = 0; u < 8; u++) { v[u] = __ldg(src + i + u * kThreads); }
#pragma unroll
            for (int u = 0; u < 8; u++) { const int p = i + u * kThreads; win[p + (p >> padShift)] = v[u]; }
        }
        for (; i < span; i += kThreads) { win[i + (i >> padShift)] = __ldg(src + i); }
    }
    else {
        for (int i = threadIdx.x; i < span; i += kThreads) { win[i + (i >> padShift)] = loadPos<EF>(j, base + i); }
    }
}

// ---- blocked decimating FIR --------------------------------------------------------------------------------------
// CTA tile: (kThreads / S) threads-groups x 8 outputs.  Window position p (relative to the tile's first sample) lives
// in slot p + p / (8*D): one pad element per thread-group stride, so the lanes of a warp (8*D samples apart) hit
// distinct banks.
template <int EF>
__device__ __forceinline__ void blockedFir(const FirJob& j, typename Elem<EF>::T* win, int smemElems) {
    using E = Elem<EF>;
    using T = typename E::T;
    const int S = j.plan;
    const int logS = ilog2(S), logD = ilog2(j.D);
    const int opc = (kThreads >> logS) * kR;
    const int k0 = blockIdx.x * opc;
    if (k0 >= j.nOut) { return; }
    const int nk = min(opc, j.nOut - k0);
    const int base = j.offset + k0 * j.D;
    const int span = (nk - 1) * j.D + j.T + (kR - 1) * j.D;       // the last thread's window may run past its last output
    const int sh = 3 + logD;
    stageWindow<EF>(j, win, base, span, sh);
    __syncthreads();
    const int grp = threadIdx.x >> logS, seg = threadIdx.x & (S - 1);
    const int kt = grp * kR;                                       // first output of this thread within the tile
    T acc[kR];
#pragma unroll
    for (int r = 0; r < kR; r++) { acc[r] = E::zero(); }
    if (kt < nk) {
        // sample (r + u) of polyphase component v sits at window position (kt + r + u) * D + v, i.e. slot
        // grp * (8*D + 1) + v + i * D + (i >> 3) with i = r + u
        const T* wbase = win + grp * ((kR << logD) + 1);
        for (int v = 0; v < j.D; v++) {
            const int nU = (j.T - v + j.D - 1) >> logD;            // taps v, v + D, v + 2D, ... < T
            const int u0 = (seg * nU) >> logS, u1 = ((seg + 1) * nU) >> logS;
            if (u1 <= u0) { continue; }
            const T* wv = wbase + v;
            const float* tp = j.taps + v + ((size_t)u0 << logD);
            T xw[kR];
#pragma unroll
            for (int r = 0; r < kR - 1; r++) { const int i = u0 + r; xw[r] = wv[(i << logD) + (i >> 3)]; }
            int u = u0;
            for (; u + kR <= u1; u += kR) {
#pragma unroll
                for (int q = 0; q < kR; q++) {
                    const int i = u + q + kR - 1;
                    xw[(q + kR - 1) & (kR - 1)] = wv[(i << logD) + (i >> 3)];
                    const float h = __ldg(tp + ((size_t)q << logD));
#pragma unroll
                    for (int r = 0; r < kR; r++) { E::fma(acc[r], h, xw[(q + r) & (kR - 1)]); }
                }
                tp += (size_t)kR << logD;
            }
            // remainder: rotate the window by hand so that xw[r] is again the sample of output r
            for (; u < u1; u++) {
                const int i = u + kR - 1;
                xw[kR - 1] = wv[(i << logD) + (i >> 3)];
                const float h = __ldg(tp);
                tp += j.D;
#pragma unroll
                for (int r = 0; r < kR; r++) { E::fma(acc[r], h, xw[r]); }
#pragma unroll
                for (int r = 0; r < kR - 1; r++) { xw[r] = xw[r + 1]; }
            }
        }
    }
    // fold the tap segments (adjacent lanes)
    for (int m = 1; m < S; m <<= 1) {
#pragma unroll
        for (int r = 0; r < kR; r++) { E::add(acc[r], E::shflXor(acc[r], m)); }
    }
    if (seg == 0 && kt < nk) {
        T* out = reinterpret_cast<T*>(j.out) + k0 + kt;
#pragma unroll
        for (int r = 0; r < kR; r++) { if (kt + r < nk) { out[r] = acc[r]; } }
    }
}

// ---- plan-table decimating FIR -------------------------------------------------------------------------------------
// R consecutive outputs per thread; the thread's window of (R-1)*D + T samples is walked once, every sample feeding
// the outputs r with 0 <= q - r*D < T (all decided at compile time).  Window position p sits in slot p + p/(R*D).
template <int D, int T>
__device__ __forceinline__ void planStage(const FirJob& j, float2* win) {
    constexpr int R = planOutPerThread(D);
    constexpr int RD = R * D;
    constexpr int LOGRD = (RD == 16) ? 4 : 5;
    static_assert(RD == 16 || RD == 32, "pad stride");
    constexpr int OPC = kThreads * R;
    const int k0 = blockIdx.x * OPC;
    if (k0 >= j.nOut) { return; }
    const int nk = min(OPC, j.nOut - k0);
    const int base = j.offset + k0 * D;
    const int span = (((nk + R - 1) / R) * R - 1) * D + T;
    stageWindow<2>(j, win, base, span, LOGRD);
    __syncthreads();
    const int kt = threadIdx.x * R;
    if (kt >= nk) { return; }
    const float2* w = win + threadIdx.x * (RD + 1);
    float2 acc[R];
#pragma unroll
    for (int r = 0; r < R; r++) { acc[r] = make_float2(0.f, 0.f); }
#pragma unroll
    for (int q = 0; q < (R - 1) * D + T; q++) {
        const float2 x = w[q + (q >> LOGRD)];
#pragma unroll
        for (int r = 0; r < R; r++) {
            if (q - r * D >= 0 && q - r * D < T) {
                const float h = PlanStage<D, T>::tap(q - r * D);
                acc[r] = fma2(make_float2(h, h), x, acc[r]);
            }
        }
    }
    float2* out = reinterpret_cast<float2*>(j.out) + k0 + kt;
#pragma unroll
    for (int r = 0; r < R; r++) { if (kt + r < nk) { out[r] = acc[r]; } }
}

// ---- polyphase resampler, 8 lanes per output ----------------------------------------------------------------------
// output k sits at offset + (phase0 + k*M)/L and uses bank row (phase0 + k*M) % L; rows are padded with zeros to a
// multiple of 4 floats (rowStride) and 16-byte aligned.
template <int EF>
__device__ __forceinline__ void lanesPoly(const FirJob& j, typename Elem<EF>::T* win, int smemElems) {
    using E = Elem<EF>;
    using T = typename E::T;
    const int k0 = blockIdx.x * kPolyOutPerCta;
    if (k0 >= j.nOut) { return; }
    const int nk = min(kPolyOutPerCta, j.nOut - k0);
    const long long tot0 = (long long)j.phase0 + (long long)k0 * j.M;
    const long long totL = (long long)j.phase0 + (long long)(k0 + nk - 1) * j.M;
    const int base = j.offset + (int)(tot0 / j.L);
    const unsigned rem0 = (unsigned)(tot0 % j.L);                 // per output: 32-bit arithmetic from here (128 * M < 2^31)
    const int span = (int)(totL / j.L) - (int)(tot0 / j.L) + j.rowStride;
    stageWindow<EF>(j, win, base, span, EF == 2 ? 4 : 5);
    __syncthreads();
    // P lanes per output: 8 (a bank row is read as 128-byte runs) or 1 (plenty of outputs in the launch and a bank small
    // enough to sit in L1: no shuffles, one thread walks the whole row)
    const int P = j.plan, logP = ilog2(P);
    const int q = threadIdx.x & (P - 1), g = threadIdx.x >> logP;
    const int kGroups = kThreads >> logP;
    for (int kk = g; kk < kPolyOutPerCta; kk += kGroups) {          // same trip count for every lane of a warp
        T acc = E::zero();
        if (kk < nk) {
            const unsigned total = rem0 + (unsigned)kk * (unsigned)j.M;
            const unsigned adv = total / (unsigned)j.L;
            const int p0 = (int)adv;
            const float* row = j.taps + (size_t)(total - adv * (unsigned)j.L) * j.rowStride;
            for (int t = 4 * q; t < j.rowStride; t += 4 * P) {
                const float4 h = __ldg(reinterpret_cast<const float4*>(row + t));
                E::fma(acc, h.x, win[E::slot(p0 + t)]);
                E::fma(acc, h.y, win[E::slot(p0 + t + 1)]);
                E::fma(acc, h.z, win[E::slot(p0 + t + 2)]);
                E::fma(acc, h.w, win[E::slot(p0 + t + 3)]);
            }
        }
        for (int m = 1; m < P; m <<= 1) { E::add(acc, E::shflXor(acc, m)); }
        if (q == 0 && kk < nk) { reinterpret_cast<T*>(j.out)[k0 + kk] = acc; }
    }
}

// ---- polyphase resampler with a short bank (L <= 16 rows), one output per thread -----------------------------------
// Outputs k and k + L use the same bank row, and their windows start exactly M samples apart.  A CTA takes 128*L
// consecutive outputs and walks the L residue classes one after the other: within a class every thread reads the
// SAME row (uniform loads, one wavefront) and the threads' windows are M samples apart in shared memory.
constexpr int kClassPlan = 2;
template <int EF>
__device__ __forceinline__ void classPoly(const FirJob& j, typename Elem<EF>::T* win) {
    using E = Elem<EF>;
    using T = typename E::T;
    const int opc = kThreads * j.L;
    const int k0 = blockIdx.x * opc;
    if (k0 >= j.nOut) { return; }
    const int nk = min(opc, j.nOut - k0);
    const long long tot0 = (long long)j.phase0 + (long long)k0 * j.M;
    const int base = j.offset + (int)(tot0 / j.L);
    const unsigned rem0 = (unsigned)(tot0 % j.L);
    const int span = (int)(((long long)rem0 + (long long)(nk - 1) * j.M) / j.L) + j.rowStride + 1;
    stageWindow<EF>(j, win, base, span, 30);       // no padding: lanes M samples apart are conflict-free for odd M
    __syncthreads();
    const int t = threadIdx.x;
    for (int c = 0; c < j.L; c++) {
        const int kk = c + j.L * t;
        if (kk >= nk) { continue; }
        const unsigned total = rem0 + (unsigned)c * (unsigned)j.M;     // + L*t*M, which is a whole number of samples
        const unsigned adv = total / (unsigned)j.L;
        const float* row = j.taps + (size_t)(total - adv * (unsigned)j.L) * j.rowStride;
        const int p0 = (int)adv + t * j.M;
        T acc = E::zero();
        for (int i = 0; i < j.rowStride; i += 4) {
            const float4 h = __ldg(reinterpret_cast<const float4*>(row + i));
            E::fma(acc, h.x, win[p0 + i]);
            E::fma(acc, h.y, win[p0 + i + 1]);
            E::fma(acc, h.z, win[p0 + i + 2]);
            E::fma(acc, h.w, win[p0 + i + 3]);
        }
        reinterpret_cast<T*>(j.out)[k0 + kk] = acc;
    }
}

// ---- polyphase resampler, samples shared between the outputs of a residue class --------------------------------------
// Outputs k and k + L use the same bank row and their windows start M samples apart; with more taps per phase than M
// those windows overlap.  A thread takes R (2 or 3) consecutive members of a class: every sample of their joint window
// is read from shared memory ONCE and feeds up to R accumulators, each with its own tap of the same row (tap index
// i - r * M).  A warp works on one class at a time, so the row is read with uniform 16-byte loads (one wavefront), and
// the lanes' windows, R * M samples apart, are padded to an odd stride.  Against eight lanes per output (lanesPoly) this
// is one shared-memory read per R MACs instead of one per MAC and no shuffles: the 25/48 resampler of the WFM channels
// (146 taps per phase) ran with the L1 data pipe 92 % busy there.
constexpr int kSharePlan = 3;
__host__ __device__ inline int shareR(int rowStride, int M) { const int r = 1 + (rowStride - 1) / M; return r > 3 ? 3 : r; }
// pad shift that makes the lane stride step * (1 + 2^-s) odd (step even), none for an odd step
__host__ __device__ inline int sharePad(int step) {
    if (step & 1) { return 30; }
    int sft = 0;
    while (!((step >> sft) & 1)) { sft++; }
    return sft;
}
template <int EF>
__device__ __forceinline__ void sharedPoly(const FirJob& j, typename Elem<EF>::T* win) {
    using E = Elem<EF>;
    using T = typename E::T;
    const int R = shareR(j.rowStride, j.M);
    const int opc = 32 * R * j.L;
    const int k0 = blockIdx.x * opc;
    if (k0 >= j.nOut) { return; }
    const int nk = min(opc, j.nOut - k0);
    const long long tot0 = (long long)j.phase0 + (long long)k0 * j.M;
    const int base = j.offset + (int)(tot0 / j.L);
    const unsigned rem0 = (unsigned)(tot0 % j.L);
    const int span = (int)(((long long)rem0 + (long long)(nk - 1) * j.M) / j.L) + j.rowStride + 1;
    const int ps = sharePad(R * j.M);
    stageWindow<EF>(j, win, base, span, ps);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int walk = (R - 1) * j.M + j.rowStride;            // joint window of a thread's R members
    for (int c = warp; c < j.L; c += kThreads / 32) {
        const int u0 = lane * R;                             // member u of class c is output c + L * u
        if (c + j.L * u0 >= nk) { continue; }
        const unsigned total = rem0 + (unsigned)c * (unsigned)j.M;
        const unsigned adv = total / (unsigned)j.L;
        const float* row = j.taps + (size_t)(total - adv * (unsigned)j.L) * j.rowStride;
        const int p0 = (int)adv + u0 * j.M;
        T acc[3] = { E::zero(), E::zero(), E::zero() };
        for (int i = 0; i < walk; i += 4) {
            const int p = p0 + i;
            const T x0 = win[p + (p >> ps)], x1 = win[p + 1 + ((p + 1) >> ps)], x2 = win[p + 2 + ((p + 2) >> ps)], x3 = win[p + 3 + ((p + 3) >> ps)];
#pragma unroll
            for (int r = 0; r < 3; r++) {
                const int idx = i - r * j.M;
                if (r < R && idx >= 0 && idx < j.rowStride) {
                    const float4 h = __ldg(reinterpret_cast<const float4*>(row + idx));
                    E::fma(acc[r], h.x, x0); E::fma(acc[r], h.y, x1); E::fma(acc[r], h.z, x2); E::fma(acc[r], h.w, x3);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < 3; r++) {
            const int kk = c + j.L * (u0 + r);
            if (r < R && kk < nk) { reinterpret_cast<T*>(j.out)[k0 + kk] = acc[r]; }
        }
    }
}

// ---- one output per thread ----------------------------------------------------------------------------------------
template <int EF>
__device__ __forceinline__ void simpleJob(const FirJob& j, typename Elem<EF>::T* win, int smemElems) {
    using E = Elem<EF>;
    using T = typename E::T;
    const int k0 = blockIdx.x * kSimpleOutPerCta;
    if (k0 >= j.nOut) { return; }
    const int nk = min(kSimpleOutPerCta, j.nOut - k0);
    const int k = threadIdx.x;
    if (j.kind == JOB_DECIM_FIR) {
        const int base = j.offset + k0 * j.D;                 // first work-buffer position this CTA touches
        const int span = (nk - 1) * j.D + j.T;                // positions [base, base+span)
        const bool staged = E::slot(span) + 1 <= smemElems;
        T acc = E::zero();
        if (staged) {
            stageWindow<EF>(j, win, base, span, EF == 2 ? 4 : 5);
            __syncthreads();
            if (k < nk) {
                const int p0 = k * j.D;
                int t = 0;
                for (; t + 4 <= j.T; t += 4) {
                    const float4 h = __ldg(reinterpret_cast<const float4*>(j.taps) + (t >> 2));
                    E::fma(acc, h.x, win[E::slot(p0 + t)]);
                    E::fma(acc, h.y, win[E::slot(p0 + t + 1)]);
                    E::fma(acc, h.z, win[E::slot(p0 + t + 2)]);
                    E::fma(acc, h.w, win[E::slot(p0 + t + 3)]);
                }
                for (; t < j.T; t++) { E::fma(acc, __ldg(j.taps + t), win[E::slot(p0 + t)]); }
            }
        }
        else if (k < nk) {
            for (int t = 0; t < j.T; t++) { E::fma(acc, __ldg(j.taps + t), loadPos<EF>(j, base + k * j.D + t)); }
        }
        if (k < nk) { reinterpret_cast<T*>(j.out)[k0 + k] = acc; }
    }
    else if (k < nk) {
        // polyphase without staging (a span that does not fit in shared memory)
        const long long total = (long long)j.phase0 + (long long)(k0 + k) * j.M;
        const int pos = j.offset + (int)(total / j.L);
        const float* row = j.taps + (size_t)(total % j.L) * j.rowStride;
        T acc = E::zero();
        for (int t = 0; t < j.T; t++) { E::fma(acc, __ldg(row + t), loadPos<EF>(j, pos + t)); }
        reinterpret_cast<T*>(j.out)[k0 + k] = acc;
    }
}

template <int EF>
__global__ void __launch_bounds__(kThreads) firJobsKernel(const FirJob* __restrict__ jobs, int smemElems) {
    using T = typename Elem<EF>::T;
    extern __shared__ __align__(16) unsigned char smemRaw[];
    T* win = reinterpret_cast<T*>(smemRaw);
    const FirJob j = jobs[blockIdx.y];
    if (j.plan >= kPlanBase) {
        if constexpr (EF == 2) {
            switch (j.plan - kPlanBase) {
            case 2 * 100 + 69: planStage<2, 69>(j, win); break;
            case 2 * 100 + 12: planStage<2, 12>(j, win); break;
            case 4 * 100 + 27: planStage<4, 27>(j, win); break;
            case 8 * 100 + 54: planStage<8, 54>(j, win); break;
            case 8 * 100 + 44: planStage<8, 44>(j, win); break;
            default: planStage<8, 36>(j, win); break;
            }
        }
    }
    else if (j.plan > 0 && j.kind == JOB_DECIM_FIR) { blockedFir<EF>(j, win, smemElems); }
    else if (j.plan == kClassPlan && j.kind == JOB_POLYPHASE) { classPoly<EF>(j, win); }
    else if (j.plan == kSharePlan && j.kind == JOB_POLYPHASE) { sharedPoly<EF>(j, win); }
    else if (j.plan > 0) { lanesPoly<EF>(j, win, smemElems); }
    else { simpleJob<EF>(j, win, smemElems); }
}

// histNext[i] = workbuf[nIn + i], i < H  (the reference's memmove of the last taps-1 samples, fir.h:86-87)
__global__ void __launch_bounds__(256) historyUpdateKernel(const FirJob* __restrict__ jobs) {
    const FirJob j = jobs[blockIdx.y];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= j.H) { return; }
    const int p = j.nIn + i;
    if (j.elemFloats == 2) {
        reinterpret_cast<float2*>(j.histNext)[i] = loadPos<2>(j, p);
    }
    else {
        reinterpret_cast<float*>(j.histNext)[i] = loadPos<1>(j, p);
    }
}

// ---- host-side geometry (must mirror the device code above)
int outPerCta(const FirJob& j) {
    if (j.plan >= kPlanBase) { return kThreads * planOutPerThread(j.D); }
    if (j.plan > 0 && j.kind == JOB_DECIM_FIR) { return (kThreads / j.plan) * kR; }
    if (j.plan == kClassPlan && j.kind == JOB_POLYPHASE) { return kThreads * j.L; }
    if (j.plan == kSharePlan && j.kind == JOB_POLYPHASE) { return 32 * shareR(j.rowStride, j.M) * j.L; }
    if (j.plan > 0) { return kPolyOutPerCta; }
    return kSimpleOutPerCta;
}

// shared-memory elements one CTA of this job wants (0 = reads through L1)
size_t smemElemsFor(const FirJob& j) {
    if (j.plan >= kPlanBase) {
        const size_t span = (size_t)outPerCta(j) * j.D + j.T;
        return span + span / ((size_t)planOutPerThread(j.D) * j.D) + 2;
    }
    if (j.plan > 0 && j.kind == JOB_DECIM_FIR) {
        const size_t span = (size_t)outPerCta(j) * j.D + j.T + (size_t)kR * j.D;
        return span + span / ((size_t)kR * j.D) + 2;
    }
    if (j.plan == kClassPlan && j.kind == JOB_POLYPHASE) {
        const size_t span = (size_t)kThreads * j.M + j.rowStride + 4;
        return span + (span >> 4) + 2;
    }
    if (j.plan == kSharePlan && j.kind == JOB_POLYPHASE) {
        const int R = shareR(j.rowStride, j.M);
        const size_t span = (size_t)32 * R * j.M + j.M + j.rowStride + 8;
        return span + (span >> sharePad(R * j.M)) + 4;
    }
    if (j.plan > 0) {
        const size_t span = (size_t)(((long long)kPolyOutPerCta * j.M) / j.L) + j.rowStride + 2;
        return span + (span >> 4) + 2;
    }
    const size_t span = (j.kind == JOB_DECIM_FIR) ? (size_t)(kSimpleOutPerCta - 1) * j.D + j.T : 0;
    return span + (span >> 4) + 8;
}

} // namespace

void firJobPlan(FirJob& j) {
    const bool fromPlanTable = j.plan == -1;
    j.plan = 0;
    if (fromPlanTable && j.kind == JOB_DECIM_FIR && j.elemFloats == 2) {
        const int shape = j.D * 100 + j.T;
        for (int known : { 2 * 100 + 69, 2 * 100 + 12, 4 * 100 + 27, 8 * 100 + 54, 8 * 100 + 44, 8 * 100 + 36 }) {
            if (shape == known) { j.plan = kPlanBase + shape; return; }
        }
    }
    const size_t elemBytes = (size_t)j.elemFloats * sizeof(float);
    if (j.kind == JOB_DECIM_FIR) {
        if (j.D < 1 || j.D > 8 || (j.D & (j.D - 1)) != 0) { return; }
        // split the taps of each polyphase component over S lanes while that leaves at least 8 taps per lane and the
        // launch has fewer than ~4 warps per scheduler
        const int nU = j.T / j.D;
        if (nU < kR) { return; }        // fewer taps per component than the register window: no reuse to gain
        int S = 1;
        while (S < 16 && (long long)(j.nOut / kR + 1) * S < 148LL * 512 && nU / (2 * S) >= 8) { S *= 2; }
        j.plan = S;
        if (smemElemsFor(j) * elemBytes > kSmemCap) { j.plan = 0; }
    }
    else {
        if (j.rowStride % 4 != 0 || (long long)j.L + (long long)kPolyOutPerCta * j.M >= (1LL << 31)) { return; }
        const int R = shareR(j.rowStride, j.M);
        if (R >= 2 && j.M % 4 == 0 && j.L <= 64 && j.nOut >= 2 * 32 * R * j.L) {
            j.plan = kSharePlan;
            if (smemElemsFor(j) * elemBytes <= kSmemCap / 2) { return; }
        }
        if (j.L <= 16 && j.nOut >= 4 * kThreads * j.L) {
            j.plan = kClassPlan;
            if (smemElemsFor(j) * elemBytes <= kSmemCap / 2) { return; }
        }
        j.plan = kPolyLanes;
        if (smemElemsFor(j) * elemBytes > kSmemCap) { j.plan = 0; }
    }
}

void firJobsBalance(FirJob* jobs, int nJobs) {
    long long ctas = 0;
    bool any = false;
    for (int i = 0; i < nJobs; i++) {
        if (jobs[i].nOut <= 0) { continue; }
        const int opc = outPerCta(jobs[i]);
        ctas += (jobs[i].nOut + opc - 1) / opc;
        any = any || jobs[i].plan >= kPlanBase;
    }
    if (any && ctas < 148 * 8) {
        for (int i = 0; i < nJobs; i++) { if (jobs[i].plan >= kPlanBase) { jobs[i].plan = 0; firJobPlan(jobs[i]); } }   // the generic planner's choice
    }
    // The per-job planner splits long filters over lanes as if the job were alone in its launch.  With many jobs side
    // by side the split only costs shuffles: halve it while the launch still has enough CTAs.
    auto launchCtas = [&]() { long long c = 0; for (int i = 0; i < nJobs; i++) { if (jobs[i].nOut > 0) { const int o = outPerCta(jobs[i]); c += (jobs[i].nOut + o - 1) / o; } } return c; };
    for (;;) {
        std::vector<int> saved((size_t)nJobs);
        bool changed = false;
        for (int i = 0; i < nJobs; i++) {
            FirJob& j = jobs[i];
            saved[i] = j.plan;
            if (j.kind == JOB_DECIM_FIR && j.plan > 1 && j.plan < kPlanBase) { j.plan >>= 1; changed = true; }
        }
        if (!changed) { break; }
        if (launchCtas() < 148 * 8) {
            for (int i = 0; i < nJobs; i++) { jobs[i].plan = saved[i]; }
            break;
        }
    }
    // (polyphase jobs keep their 8 lanes per output: one lane per output was measured slower even with 256 jobs in the
    // launch and a four-row bank, 297 us against 229 us)
}

int firJobTiles(const FirJob& j) {
    if (j.nOut <= 0) { return 0; }
    const int opc = outPerCta(j);
    return (j.nOut + opc - 1) / opc;
}

void firJobsSortByTiles(FirJob* jobs, int nJobs) {
    std::stable_sort(jobs, jobs + nJobs, [](const FirJob& a, const FirJob& b) { return firJobTiles(a) > firJobTiles(b); });
}

static int launchFirJobRange(const FirJob* d_jobs, const FirJob* h_jobs, int nJobs, int elemFloats, cudaStream_t s);

// grid = (tiles of the longest job, jobs): a CTA whose tile lies beyond its job's end still has to be launched and to fetch
// its descriptor before it can leave.  In a mixed bank that is most of them -- cfg4's first later-stage wave: 128 WFM jobs of
// 256 tiles beside 384 jobs of 64 and 512 of 32, 72 % of 262144 CTAs empty -- so a list that comes sorted by tile count
// (firJobsSortByTiles) is launched run by run, a run ending where the tile count falls below half of its first job's.
int launchFirJobs(const FirJob* d_jobs, const FirJob* h_jobs, int nJobs, int elemFloats, cudaStream_t s) {
    if (nJobs <= 0) { return B200_OK; }
    bool sorted = true;
    for (int i = 1; i < nJobs && sorted; i++) { sorted = firJobTiles(h_jobs[i]) <= firJobTiles(h_jobs[i - 1]); }
    if (!sorted) { return launchFirJobRange(d_jobs, h_jobs, nJobs, elemFloats, s); }
    for (int i = 0; i < nJobs;) {
        const int t0 = firJobTiles(h_jobs[i]);
        if (t0 == 0) { break; }                           // only empty jobs are left
        int e = i + 1;
        while (e < nJobs && 2 * firJobTiles(h_jobs[e]) > t0) { e++; }
        if (nJobs - e < 8) { e = nJobs; }                  // not worth a launch of its own
        B200_TRY(launchFirJobRange(d_jobs + i, h_jobs + i, e - i, elemFloats, s));
        i = e;
    }
    return B200_OK;
}

static int launchFirJobRange(const FirJob* d_jobs, const FirJob* h_jobs, int nJobs, int elemFloats, cudaStream_t s) {
    if (nJobs <= 0) { return B200_OK; }
    int tiles = 0;
    size_t elems = 0;
    for (int i = 0; i < nJobs; i++) {
        const FirJob& j = h_jobs[i];
        if (j.nOut <= 0) { continue; }
        const int opc = outPerCta(j);
        tiles = std::max(tiles, (j.nOut + opc - 1) / opc);
        elems = std::max(elems, smemElemsFor(j));
    }
    if (tiles == 0) { return B200_OK; }
    // request only what the widest job of this launch needs (more resident CTAs per SM); simple-path jobs whose window
    // would not fit read through L1 instead
    const size_t elemBytes = (size_t)elemFloats * sizeof(float);
    const size_t smemBytes = std::min<size_t>(std::max<size_t>(elems * elemBytes, 1024), kSmemCap);
    dim3 grid(tiles, nJobs);
    if (elemFloats == 2) {
        B200_CUDA(cudaFuncSetAttribute(firJobsKernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemCap));
        firJobsKernel<2><<<grid, kThreads, smemBytes, s>>>(d_jobs, (int)(smemBytes / elemBytes));
    }
    else {
        B200_CUDA(cudaFuncSetAttribute(firJobsKernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemCap));
        firJobsKernel<1><<<grid, kThreads, smemBytes, s>>>(d_jobs, (int)(smemBytes / elemBytes));
    }
    B200_CHECK_LAUNCH();
    countLaunch();
    return B200_OK;
}

int launchHistoryUpdate(const FirJob* d_jobs, int nJobs, int maxH, cudaStream_t s) {
    if (nJobs <= 0 || maxH <= 0) { return B200_OK; }
    dim3 grid((maxH + 255) / 256, nJobs);
    historyUpdateKernel<<<grid, 256, 0, s>>>(d_jobs);
    B200_CHECK_LAUNCH();
    countLaunch();
    return B200_OK;
}

} // namespace b200

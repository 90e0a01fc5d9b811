// Fused frequency translation + first decimating FIR stage, many VFOs per launch  (K1 + K2 of SURVEY §2.2).
//
// Reference work per VFO and input sample (rx_vfo.h:89-90 -> frequency_xlator.h:47-51 ->
// power_decimator.h:58-65 -> decimating_fir.h:51-61): one complex multiply by the rotator phase, a copy
// into the FIR work buffer, and T/D real-tap MACs; every VFO re-reads its own copy of the block that the
// Splitter made for it (routing/splitter.h:74-83).  Here the block is read from HBM once per group of VFOs:
//
//   y[k] = sum_j h[j] * x[m0 + kD + j] * phase[m0 + kD + j]                      (m0 = offset - (T-1))
//
// is regrouped by input sub-block s of D samples.  With phase[m_s + r] = P_s * w[r], w[r] = phaseDelta^r:
//
//   c_q[s] = sum_r h[qD + r] * (x[m_s + r] * w[r])        q = 0..Q-1, Q = ceil(T/D)
//   y[k]   = sum_q P_{k+q} * c_q[k+q]
//
// so each input sample is rotated once (4 flops-pairs) and feeds Q accumulators (2 FFMA each): ~12-16 FP32
// instructions per sample per VFO, the minimum for this algorithm, instead of T/D rotations per sample.
// One thread owns one sub-block: it walks D samples from shared memory (padded so a warp's lanes hit distinct
// banks), taps come from the constant bank as FFMA immediates (the r loop is fully unrolled), w[r] is a
// warp-uniform shared-memory broadcast.  Partial sums meet in shared memory.  The T-1 samples of rotated
// history that the first outputs need, and the history refresh, are handled by a small edge kernel.
//
// CUDA FP32 pipes only: this is a bank of short real-tap FIRs on a shared stream, not a dense contraction.
#include "xlate_decim.cuh"

#include <algorithm>
#include <cstdlib>

#include "host_design.h"
#include "rotator.cuh"
#include "vfo_bank.cuh"

namespace b200 {

int fusedTileSamples = 0;   // B200_OPT_TILE_SAMPLES: kTileLarge forces the grouped kernel for a one-VFO launch too; 0 / 2048 = default
int fusedVfosPerPass = 0;   // B200_OPT_VFOS_PER_PASS: 0 = by code size (launchPlan), else 2 / 4
bool fusedLeaveRoom = false; // set by the bank for a launch whose block has a deferred tail beside it (streaming mode)
bool fusedParamTables = true;
bool fusedSplitRuns = true;     // B200_OPT_SPLIT_RUNS: the /8192 plan as two launches of half the runs each (13.4 against 14.5 ms per cfg4 block)
int fusedVfosPerCta = 32;   // B200_OPT_VFOS_PER_CTA: VFOs a CTA of the grouped kernel serves from one staged tile
// One VFO in the launch: the variant with 16-byte staging (fusedStreamKernel).  8-byte cp.async.ca staging was what held
// the single-VFO kernel at 0.65 of the copy peak; with 16-byte cp.async.cg and LDS.128 it reaches 0.94.
bool fusedStreamSingle = true;


namespace {

// ---- coefficient tables in the constant bank (all 13 stage filters, 9 KB)
#define DECIM_TABLE_U32(name, count) __constant__ uint32_t c_##name[count] = {
#define DECIM_ROW(...) __VA_ARGS__,
#define DECIM_TABLE_END };
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

// ---- the same tables once more as function-local constant arrays: with the compile-time indices of the unrolled tap
// loops the loads fold away and every tap becomes the 32-bit immediate of its FFMA.  Measured on the D = 128 body
// (tools/microbench/ldc_fir.cu): 0.79 of the FP32 peak with immediates against 0.68 with constant-bank operands.
#define DECIM_TABLE_U32(name, count) static __device__ __forceinline__ float imm_##name(int i) { const uint32_t t[count] = {
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

// ---- FirstStage<ratio>: decimation, tap count and tap accessor of each plan's first stage
template <int RATIO> struct FirstStage;
#define DECIM_TABLE_U32(name, count)
#define DECIM_ROW(...)
#define DECIM_TABLE_END
#define DECIM_PLAN_BEGIN(ratio, nstages) template <> struct FirstStage<ratio> {
#define DECIM_STAGE(index, decim, name, ntaps) DECIM_STAGE_##index(decim, name, ntaps)
#define DECIM_STAGE_0(decim, name, ntaps)                                                      \
    static constexpr int D = decim;                                                            \
    static constexpr int T = ntaps;                                                            \
    static __device__ __forceinline__ float tap(int i) { return __uint_as_float(c_##name[i]); }      \
    static __device__ __forceinline__ float tapImm(int i) { return imm_##name(i); }
#define DECIM_STAGE_1(decim, name, ntaps)
#define DECIM_STAGE_2(decim, name, ntaps)
#define DECIM_STAGE_3(decim, name, ntaps)
#define DECIM_PLAN_END };
#include "decim_tables.inc"
#undef DECIM_TABLE_U32
#undef DECIM_ROW
#undef DECIM_TABLE_END
#undef DECIM_PLAN_BEGIN
#undef DECIM_STAGE
#undef DECIM_STAGE_0
#undef DECIM_STAGE_1
#undef DECIM_STAGE_2
#undef DECIM_STAGE_3
#undef DECIM_PLAN_END

constexpr int kMaxVfosPerCta = 64;     // VFOs a CTA of the grouped kernel serves from one staged tile (B200_OPT_VFOS_PER_CTA)
constexpr int kMinFusedRatio = 16;     // first-stage decimation >= 8 (plans 2/4/8 start with /2 or /4: Q too large to pay off)

// Input samples per CTA tile: 2048 for the single-VFO kernel (fusedStreamKernel, HBM-bound: more resident CTAs), 4096
// for the grouped kernel (FP32-bound from a few VFOs on: fewer overlap sub-blocks recomputed per output; measured
// 0.35 of the FFMA peak against 0.32 with 2048-sample tiles, before pairing and 16-byte staging).
#ifndef B200_D128_SMALL_TILE
#define B200_SPLIT_BODY_BYTES (32 * 1024)   // plans with a larger unrolled body run as two half-launches; 16 KB (the /4096 plan too) measured no faster: 11.30 against 11.23 ms
#define B200_D128_SMALL_TILE 0      // 1 = 4096-sample tiles for the /8192 plan too: measured slower (cfg4 15.3 against 14.6 ms per block)
#endif
constexpr int kTileSmall = 2048, kTileLarge = 4096;

template <int D, int TILE> struct TileCfg {
    // sub-blocks per CTA
    static constexpr int SB = (TILE / D) > 256 ? 256 : ((TILE / D) < 32 ? 32 : (TILE / D));
    // threads per sub-block: a sub-block of D samples is split into RS runs of DS = D/RS <= 32 samples, so that the
    // shared-memory tile (8*D bytes per sub-block) does not cap the resident warps for the large-D plans.  (Measured
    // with one VFO: DS = 8, four times the warps on the same tile, gave the same 0.65 of the copy peak -- 100 M instead
    // of 75 M instructions and the L1TEX data pipe at 76 % -- so the split stays as coarse as the tile allows.)
    // (measured for the many-VFO tile too: DS = 16 -> 0.30 of the FFMA peak, DS = 64 -> 0.31, DS = 32 -> 0.33)
    static constexpr int DSMAX = 32;
    static constexpr int RS = (D / DSMAX) > 8 ? 8 : ((D / DSMAX) < 1 ? 1 : (D / DSMAX));
    static constexpr int DS = D / RS;
    static constexpr int THREADS = SB * RS;
};

// 8-byte asynchronous global->shared copy (LDGSTS); src_size 0 zero-fills without touching memory
__device__ __forceinline__ void cpAsync8(float2* smemDst, const float2* gsrc, bool valid) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smemDst);
    const int sz = valid ? 8 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(d), "l"(gsrc), "r"(sz) : "memory");
}

__device__ __forceinline__ void cpAsync8Full(float2* smemDst, const float2* gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smemDst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(d), "l"(gsrc) : "memory");
}

__device__ __forceinline__ void cpAsync16(float2* smemDst, const float2* gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smemDst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc) : "memory");
}

__device__ __forceinline__ float2 cmulf(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }

// Blackwell's packed FP32 pair arithmetic (SASS FFMA2 / FMUL2): one issue slot for two FMAs, same FMA rate as the scalar
// form (measured: 127 FMA/clk/SM at 2.0 warp-instructions/clk/SM, tools/microbench/ffma2_throughput.cu).  A scalar
// multiplier is given as the pair (h, h): ptxas folds it into the instruction's broadcast operand form (a 32-bit
// immediate for a constant tap, the .F32 register modifier for a loaded value), so no extra moves are issued.
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
// x * w with w given as the quad (w.re, w.im, -w.im, w.re):  x.re * (w.re, w.im) + x.im * (-w.im, w.re)
__device__ __forceinline__ float2 rotate2(float xr, float xi, float4 wq) {
    return ffma2(make_float2(xi, xi), make_float2(wq.z, wq.w), fmul2(make_float2(xr, xr), make_float2(wq.x, wq.y)));
}

// phase at sample m of the block, m may be negative (backward extension from chk[0]; only the m >= 0 part of
// a sub-block that starts in the history region ever multiplies non-zero data)
__device__ __forceinline__ float2 phaseAtAny(const FusedJob& j, int m) {
    if (m >= 0) { return rotPhaseAt(j.chk, j.wtab, m); }
    const float2 c0 = __ldg(j.chk);
    double sd, cd;
    sincos(j.thetaEff * (double)m, &sd, &cd);
    const double r = exp(j.logMag * (double)m);
    const float2 w = make_float2((float)(r * cd), (float)(r * sd));
    return cmulf(c0, w);
}

// One thread's run of DS samples of one sub-block for NP pairs of VFOs at once.  The two VFOs of a pair ride in the
// two lanes of the packed instructions ("structure of arrays" over the pair): zX = (zA.re, zB.re), zY = (zA.im, zB.im),
// and the rotation table holds (wA.re, wB.re, wA.im, wB.im) per sample, so one 16-byte broadcast load serves both VFOs:
//     zX = x.re * wX - x.im * wY          zY = x.re * wY + x.im * wX          (FMUL2 + FFMA2 each)
//     accX[q] += h * zX                   accY[q] += h * zY                   (FFMA2 with the tap as immediate)
// i.e. 2 + T/D packed instructions per sample and VFO.  Each input sample is read from shared memory once for all
// 2 * NP VFOs of the pass.  RUN selects which DS-sample slice of the D-sample sub-block this thread owns (compile time).
// wq: the pass's table, entry [r * NP + p].
template <int RATIO, int RUN, int NP, int DS, int ODD>
__device__ __forceinline__ void accumulateRunPairs(const float2* __restrict__ xrow, const float4* __restrict__ wq, float2 (*accX)[8], float2 (*accY)[8]) {
    using FS = FirstStage<RATIO>;
    constexpr int D = FS::D, T = FS::T + ODD;      // ODD: taps used shifted by one (h'[0] = 0, h'[t] = h[t-1])
    constexpr int Q = (T + D - 1) / D;
    const float4* x2 = reinterpret_cast<const float4*>(xrow);     // even pitch, even run start: 16-byte aligned
#pragma unroll
    for (int r2 = 0; r2 < DS / 2; r2++) {
        const float4 xx = x2[r2];
#pragma unroll
        for (int half = 0; half < 2; half++) {
            const float xr = half ? xx.z : xx.x, xi = half ? xx.w : xx.y;
            const int r = 2 * r2 + half;
#pragma unroll
            for (int p = 0; p < NP; p++) {
                const float4 w = wq[r * NP + p];
                const float2 wX = make_float2(w.x, w.y), wY = make_float2(w.z, w.w);
                const float2 zX = ffma2(make_float2(-xi, -xi), wY, fmul2(make_float2(xr, xr), wX));
                const float2 zY = ffma2(make_float2(xi, xi), wX, fmul2(make_float2(xr, xr), wY));
#pragma unroll
                for (int q = 0; q < Q; q++) {
                    const int t = q * D + RUN * DS + r;
                    if (t < T && t - ODD >= 0) {
                        const float h = FS::tapImm(t - ODD);
                        accX[p][q] = ffma2(make_float2(h, h), zX, accX[p][q]);
                        accY[p][q] = ffma2(make_float2(h, h), zY, accY[p][q]);
                    }
                }
            }
        }
    }
}

// The rotation tables of a launch in the kernel's parameter block (constant bank): a warp-uniform operand is what the
// constant path is for -- `LDCU` loads the entry into uniform registers, which FMUL2 / FFMA2 take directly, and the
// shared-memory pipe is left to the samples.  (A broadcast LDS.128 costs four wavefronts like any other; with the table
// in shared memory the inner loop ran at 83 FMA/clk/SM, from the parameter block at 95: tools/microbench/wtab_paths.cu.)
// 30 KB of the 32 KB parameter limit: D = 64 -> 30 pairs = 60 VFOs per launch, D = 128 -> 30 VFOs, D = 16 -> 240.
constexpr int kParamEntries = 1920;
template <bool CT> struct BankTablesT { float4 w[CT ? kParamEntries : 1]; };      // CT = false: the table path in shared memory, nothing to pass
using BankTables = BankTablesT<true>;

template <int RATIO, int RUN, int NP, int DS, int ODD>
__device__ __forceinline__ void accumulateRunPairsC(const float2* __restrict__ xrow, const BankTablesT<true>& tp, int tbase, float2 (*accX)[8], float2 (*accY)[8]) {
    using FS = FirstStage<RATIO>;
    constexpr int D = FS::D, T = FS::T + ODD;
    constexpr int Q = (T + D - 1) / D;
    const float4* x2 = reinterpret_cast<const float4*>(xrow);
#pragma unroll
    for (int r2 = 0; r2 < DS / 2; r2++) {
        const float4 xx = x2[r2];
#pragma unroll
        for (int half = 0; half < 2; half++) {
            const float xr = half ? xx.z : xx.x, xi = half ? xx.w : xx.y;
            const int r = 2 * r2 + half;
#pragma unroll
            for (int p = 0; p < NP; p++) {
                const float4 w = tp.w[tbase + (RUN * DS + r) * NP + p];      // uniform index: the parameter block is read by LDCU
                const float2 wX = make_float2(w.x, w.y), wY = make_float2(w.z, w.w);
                const float2 zX = ffma2(make_float2(-xi, -xi), wY, fmul2(make_float2(xr, xr), wX));
                const float2 zY = ffma2(make_float2(xi, xi), wX, fmul2(make_float2(xr, xr), wY));
#pragma unroll
                for (int q = 0; q < Q; q++) {
                    const int t = q * D + RUN * DS + r;
                    if (t < T && t - ODD >= 0) {
                        const float h = FS::tapImm(t - ODD);
                        accX[p][q] = ffma2(make_float2(h, h), zX, accX[p][q]);
                        accY[p][q] = ffma2(make_float2(h, h), zY, accY[p][q]);
                    }
                }
            }
        }
    }
}

// one CTA per (output tile, range of GV VFOs); all jobs of a launch share in/count/offset/nOut.
// A CTA stages its input tile once and walks its VFO range in passes of VB = 2 * NP VFOs.  Per pass: the threads
// accumulate (above), multiply by their sub-block's start phase, fold the Q partial sums of neighbouring sub-blocks
// with warp shuffles (output k = sum_q c_q[k + q]; what crosses a warp boundary or comes from another run goes
// through a small shared-memory exchange), one barrier, and the outputs leave.  The next pass's rotation table is
// fetched from global memory at the start of a pass and stored at its end (two table buffers), the start phases of a
// pass are loaded at its start and used at its end, so no global load is waited for inside the pass loop.
// HALF: 0 = a CTA covers all RS runs of its sub-blocks.  1 / 2 = it covers only the lower / upper half of the runs (the
// first / second half of every sub-block's samples) and the launch with HALF = 2 ADDS its sums to the outputs the
// HALF = 1 launch stored: the unrolled loop body of the /8192 plan (D = 128, T = 726: 34 KB for all four runs) does
// not fit the 32 KB instruction cache level, its two halves do.
// CT: the rotation tables of the launch's jobs come in the parameter block `tp` (entry (pair * D + r) * 1, one pair per pass)
// instead of being gathered into shared memory pass by pass.
template <int RATIO, int NP, int TILE, int ODD, int HALF, bool CT>
__global__ void __launch_bounds__(TileCfg<FirstStage<RATIO>::D, TILE>::THREADS / (HALF ? 2 : 1))
fusedBankKernel(const __grid_constant__ BankTablesT<CT> tp, const FusedJob* __restrict__ jobs, int nJobs, int GV, unsigned nItems) {
    using FS = FirstStage<RATIO>;
    using TC = TileCfg<FS::D, TILE>;
    constexpr int D = FS::D, T = FS::T + ODD, H = FS::T - 1;
    constexpr int Q = (T + D - 1) / D;
    static_assert(Q <= 8, "accumulator array is sized for Q <= 8");
    static_assert(HALF == 0 || TC::RS % 2 == 0, "halving needs an even number of runs");
    constexpr int SB = TC::SB, RSA = TC::RS, DS = TC::DS;
    constexpr int RS = HALF ? RSA / 2 : RSA;              // runs this CTA covers
    constexpr int RUN0 = (HALF == 2) ? RSA / 2 : 0;       // ... starting with this one
    constexpr int NT = SB * RS;
    constexpr int NB = SB - (Q - 1);        // outputs per tile
    constexpr int DR = HALF ? D / 2 : D;    // samples of a sub-block this CTA reads (a half-launch stages only its half of every row)
    constexpr int LP = DR + 2;              // even sub-block pitch: rows 16-byte aligned, LDS.128 of a quarter-warp covers all banks
    constexpr int VB = 2 * NP;
    constexpr int WS = SB / 32;             // warps per run
    constexpr int HQ = Q * (Q - 1) / 2;     // halo entries per warp: (q, lane < q)
    static_assert(SB % 32 == 0, "a warp holds 32 consecutive sub-blocks of one run");
    constexpr int EPT = (NP * D + NT - 1) / NT;     // rotation-table entries a thread fetches per pass
    extern __shared__ __align__(16) float2 sm[];
    float2* xs = sm;                                               // SB * LP
    float4* wqBuf = reinterpret_cast<float4*>(xs + SB * LP);       // 2 * NP * D      : rotation tables of two passes
    float2* fbuf = reinterpret_cast<float2*>(wqBuf + 2 * NP * D);  // 2 * NP * 2 * RS * SB : in-warp folded sums [buf][p][X|Y][run][sub-block]
    float2* halo = fbuf + 2 * NP * 2 * RS * SB;                    // 2 * NP * 2 * RS * WS * HQ : what the previous warp's last outputs need
    // per-VFO pointers of the CTA's range, so that nothing inside the pass loop chases a pointer through global memory
    __shared__ const float2* sChk[kMaxVfosPerCta];
    __shared__ const float2* sWtab[kMaxVfosPerCta];
    __shared__ float2* sOutPtr[kMaxVfosPerCta];

    const int tid = threadIdx.x;
    const int sb = tid % SB;                // sub-block of this thread
    const int runLocal = tid / SB;          // which of this CTA's runs (warp-uniform: SB % 32 == 0)
    const int run = runLocal + RUN0;        // which DS-sample slice of the sub-block
    const int lane = tid & 31;
    const int ws = sb >> 5;
    // one-dimensional grid, VFO range fastest: the CTAs that share an input tile run at the same time, so only the first of
    // them brings the tile in from HBM and the others find it in L2
    const int nRanges = (nJobs + GV - 1) / GV;
    // A launch may hold fewer CTAs than work items (streaming mode: a grid that leaves room on every SM for the previous
    // block's later stages); each CTA then walks the items blockIdx.x, + gridDim.x, ...
    for (unsigned item = blockIdx.x; item < nItems; item += gridDim.x) {
    if (item != blockIdx.x) { __syncthreads(); }          // the previous item's shared-memory state is no longer read
    const int g0 = (int)(item % nRanges) * GV;
    const FusedJob j0 = jobs[g0];
    const int k0 = (int)(item / nRanges) * NB;
    if (k0 >= j0.nOut) { continue; }
    const int mStart = j0.offset - H - ODD + k0 * D;     // sample index of this tile's first sub-block (even, 16-byte aligned address)
    const float2* in = reinterpret_cast<const float2*>(j0.in);

    // ---- stage the input tile once for the whole VFO range (zero outside [0, count));
    // all copies of a thread are in flight at once, no register staging
    constexpr bool kImmediateStaging = (HALF == 0) && ((2 * NT) % D == 0);   // a thread's copies land one constant smem stride apart
    if constexpr (HALF != 0) {
        // only the half of every row this CTA reads: 16-byte copies inside the block, guarded 8-byte ones at its ends
        constexpr int DH = D / 2;
        if (mStart >= 0 && mStart + SB * D <= j0.count) {
#pragma unroll 8
            for (int e = tid; e < SB * DH / 2; e += NT) {
                const int rowI = (2 * e) / DH, col = 2 * e - rowI * DH;
                cpAsync16(xs + rowI * LP + col, in + mStart + rowI * D + RUN0 * DS + col);
            }
        }
        else {
#pragma unroll 8
            for (int i = tid; i < SB * DH; i += NT) {
                const int rowI = i / DH, col = i - rowI * DH;
                const int m = mStart + rowI * D + RUN0 * DS + col;
                const bool ok = (m >= 0 && m < j0.count);
                cpAsync8(xs + rowI * LP + col, in + (ok ? m : 0), ok);
            }
        }
    }
    else if (kImmediateStaging && mStart >= 0 && mStart + SB * D <= j0.count) {
        // interior tile: 16-byte copies (two samples, no L1 allocation), no bounds to check, and copy u of a thread is a
        // compile-time offset from copy 0 on both sides (samples 2e, 2e+1 of pair e = tid + u*NT land in slot
        // 2e + 2*(2e/D)), so the loop is LDGSTS with immediates only
        const int i0 = 2 * tid;
        const float2* g = in + mStart + i0;
        float2* d = xs + i0 + 2 * (i0 / D);
#pragma unroll
        for (int u = 0; u < (SB * D) / (2 * NT); u++) { cpAsync16(d + u * (2 * NT + 2 * (2 * NT / D)), g + u * 2 * NT); }
    }
    else {
#pragma unroll 8
        for (int i = tid; i < SB * D; i += NT) {
            const int m = mStart + i;
            const bool ok = (m >= 0 && m < j0.count);
            cpAsync8(xs + i + 2 * (i / D), in + (ok ? m : 0), ok);
        }
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");

    const int nG = min(GV, nJobs - g0);
    const int nPass = (nG + VB - 1) / VB;
    for (int g = tid; g < nG; g += NT) {
        sChk[g] = jobs[g0 + g].chk; sWtab[g] = jobs[g0 + g].wtab; sOutPtr[g] = reinterpret_cast<float2*>(jobs[g0 + g].out);
    }
    __syncthreads();
    // table entry e = r * NP + p of pass `pass`: (wA.re, wB.re, wA.im, wB.im), A = VFO gp + 2p, B = A + 1 (zero beyond the range)
    auto tableEntry = [&](int pass, int e) -> float4 {
        const int eR = e / NP, eP = e % NP;
        const int a = pass * VB + 2 * eP;
        const float2 wA = (a < nG) ? __ldg(sWtab[a] + eR) : make_float2(0.f, 0.f);
        const float2 wB = (a + 1 < nG) ? __ldg(sWtab[a + 1] + eR) : make_float2(0.f, 0.f);
        return make_float4(wA.x, wB.x, wA.y, wB.y);
    };
    if constexpr (!CT) {
#pragma unroll
        for (int u = 0; u < EPT; u++) { const int e = tid + u * NT; if (e < NP * D) { wqBuf[e] = tableEntry(0, e); } }
    }

    const float2* xrow = xs + sb * LP + runLocal * DS;    // (runLocal == run unless this is the second half-launch)
    const int ms = mStart + sb * D;                      // first sample of this thread's sub-block
    const bool msLive = ms < j0.count && ms + D > 0;     // the sub-block holds samples of the block
    const bool msFwd = msLive && ms >= 0;                // ... and starts inside it: phase = chk[ms >> 9] * wtab[ms & 511]
    // start phase of this thread's sub-block for the VFOs of a pass: the two factors are loaded one pass ahead
    float2 pc[VB], pw[VB];
    auto phaseLoads = [&](int pass) {
#pragma unroll
        for (int v = 0; v < VB; v++) {
            const int a = pass * VB + v;
            pc[v] = make_float2(0.f, 0.f); pw[v] = pc[v];
            if (msFwd && a < nG) { pc[v] = __ldg(sChk[a] + (ms >> kRotChunkLog)); pw[v] = __ldg(sWtab[a] + (ms & (kRotChunk - 1))); }
        }
    };
    phaseLoads(0);
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncthreads();

    for (int pass = 0; pass < nPass; pass++) {
        const int gp = pass * VB;
        const int buf = pass & 1;
        // ---- this pass's start phases from the factors loaded during the previous pass; then the loads for the next one
        float2 PX[NP], PY[NP];                           // (PA.re, PB.re), (PA.im, PB.im)
#pragma unroll
        for (int p = 0; p < NP; p++) {
            float2 pa = cmulf(pc[2 * p], pw[2 * p]), pb = cmulf(pc[2 * p + 1], pw[2 * p + 1]);
            if (msLive && !msFwd) {                      // sub-block that starts in the history region (first tile only)
                const int a = gp + 2 * p;
                if (a < nG) { pa = phaseAtAny(jobs[g0 + a], ms); }
                if (a + 1 < nG) { pb = phaseAtAny(jobs[g0 + a + 1], ms); }
            }
            PX[p] = make_float2(pa.x, pb.x); PY[p] = make_float2(pa.y, pb.y);
        }
        float4 nextEntry[EPT];
        const bool haveNext = !CT && (pass + 1 < nPass);
#pragma unroll
        for (int u = 0; u < EPT; u++) {
            const int e = tid + u * NT;
            nextEntry[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (haveNext && e < NP * D) { nextEntry[u] = tableEntry(pass + 1, e); }
        }
        if (pass + 1 < nPass) { phaseLoads(pass + 1); }
        // ---- accumulate
        const float4* wq = wqBuf + buf * (NP * D) + run * DS * NP;
        float2 accX[NP][8], accY[NP][8];
#pragma unroll
        for (int p = 0; p < NP; p++) {
#pragma unroll
            for (int q = 0; q < Q; q++) { accX[p][q] = make_float2(0.f, 0.f); accY[p][q] = make_float2(0.f, 0.f); }
        }
        if constexpr (CT) {
            // this pass's table block in the parameter block: pairs are numbered through the launch
            const int tbase = ((g0 / 2) + pass * NP) * D;
            switch (run) {
            case 0: if constexpr (RUN0 <= 0 && RUN0 + RS > 0) { accumulateRunPairsC<RATIO, 0, NP, DS, ODD>(xrow, tp, tbase, accX, accY); } break;
            case 1: if constexpr (RUN0 <= 1 && RUN0 + RS > 1) { accumulateRunPairsC<RATIO, 1, NP, DS, ODD>(xrow, tp, tbase, accX, accY); } break;
            case 2: if constexpr (RUN0 <= 2 && RUN0 + RS > 2) { accumulateRunPairsC<RATIO, 2, NP, DS, ODD>(xrow, tp, tbase, accX, accY); } break;
            case 3: if constexpr (RUN0 <= 3 && RUN0 + RS > 3) { accumulateRunPairsC<RATIO, 3, NP, DS, ODD>(xrow, tp, tbase, accX, accY); } break;
            case 4: if constexpr (RUN0 <= 4 && RUN0 + RS > 4) { accumulateRunPairsC<RATIO, 4, NP, DS, ODD>(xrow, tp, tbase, accX, accY); } break;
            case 5: if constexpr (RUN0 <= 5 && RUN0 + RS > 5) { accumulateRunPairsC<RATIO, 5, NP, DS, ODD>(xrow, tp, tbase, accX, accY); } break;
            case 6: if constexpr (RUN0 <= 6 && RUN0 + RS > 6) { accumulateRunPairsC<RATIO, 6, NP, DS, ODD>(xrow, tp, tbase, accX, accY); } break;
            default: if constexpr (RUN0 <= 7 && RUN0 + RS > 7) { accumulateRunPairsC<RATIO, 7, NP, DS, ODD>(xrow, tp, tbase, accX, accY); } break;
            }
        }
        else {
            switch (run) {
            case 0: if constexpr (RUN0 <= 0 && RUN0 + RS > 0) { accumulateRunPairs<RATIO, 0, NP, DS, ODD>(xrow, wq, accX, accY); } break;
            case 1: if constexpr (RUN0 <= 1 && RUN0 + RS > 1) { accumulateRunPairs<RATIO, 1, NP, DS, ODD>(xrow, wq, accX, accY); } break;
            case 2: if constexpr (RUN0 <= 2 && RUN0 + RS > 2) { accumulateRunPairs<RATIO, 2, NP, DS, ODD>(xrow, wq, accX, accY); } break;
            case 3: if constexpr (RUN0 <= 3 && RUN0 + RS > 3) { accumulateRunPairs<RATIO, 3, NP, DS, ODD>(xrow, wq, accX, accY); } break;
            case 4: if constexpr (RUN0 <= 4 && RUN0 + RS > 4) { accumulateRunPairs<RATIO, 4, NP, DS, ODD>(xrow, wq, accX, accY); } break;
            case 5: if constexpr (RUN0 <= 5 && RUN0 + RS > 5) { accumulateRunPairs<RATIO, 5, NP, DS, ODD>(xrow, wq, accX, accY); } break;
            case 6: if constexpr (RUN0 <= 6 && RUN0 + RS > 6) { accumulateRunPairs<RATIO, 6, NP, DS, ODD>(xrow, wq, accX, accY); } break;
            default: if constexpr (RUN0 <= 7 && RUN0 + RS > 7) { accumulateRunPairs<RATIO, 7, NP, DS, ODD>(xrow, wq, accX, accY); } break;
            }
        }
        // ---- c_q = acc_q * P, then the in-warp part of y[k] = sum_q c_q[k + q]
#pragma unroll
        for (int p = 0; p < NP; p++) {
            float2 fX = make_float2(0.f, 0.f), fY = fX;
            float2* hl = halo + ((((buf * NP + p) * 2) * RS + runLocal) * WS + ws) * HQ;      // X plane; the Y plane is RS * WS * HQ further
#pragma unroll
            for (int q = 0; q < Q; q++) {
                // (cX, cY) = (accX * PX - accY * PY, accX * PY + accY * PX), lane-wise over the pair
                const float2 nY = make_float2(-accY[p][q].x, -accY[p][q].y);
                const float2 cX = ffma2(nY, PY[p], fmul2(accX[p][q], PX[p]));
                const float2 cY = ffma2(accY[p][q], PX[p], fmul2(accX[p][q], PY[p]));
                if (q == 0) { fX = cX; fY = cY; }
                else {
                    const float tx0 = __shfl_down_sync(0xffffffffu, cX.x, q), tx1 = __shfl_down_sync(0xffffffffu, cX.y, q);
                    const float ty0 = __shfl_down_sync(0xffffffffu, cY.x, q), ty1 = __shfl_down_sync(0xffffffffu, cY.y, q);
                    if (lane + q < 32) { fX.x += tx0; fX.y += tx1; fY.x += ty0; fY.y += ty1; }
                    if (lane < q) {                       // needed by output lane + 32 - q of the previous warp
                        hl[q * (q - 1) / 2 + lane] = cX;
                        hl[RS * WS * HQ + q * (q - 1) / 2 + lane] = cY;
                    }
                }
            }
            float2* fb = fbuf + (((buf * NP + p) * 2) * RS + runLocal) * SB + sb;
            fb[0] = fX;
            fb[RS * SB] = fY;
        }
        if (haveNext) {
#pragma unroll
            for (int u = 0; u < EPT; u++) { const int e = tid + u * NT; if (e < NP * D) { wqBuf[(buf ^ 1) * (NP * D) + e] = nextEntry[u]; } }
        }
        __syncthreads();
        // ---- outputs: sum over the runs, add what came from the next warp, store (coalesced over k per VFO)
        for (int idx = tid; idx < NP * SB; idx += NT) {
            const int p = idx / SB, k = idx - p * SB;
            const int a = gp + 2 * p;
            if (k < NB && k0 + k < j0.nOut && a < nG) {
                const float2* fb = fbuf + ((buf * NP + p) * 2) * RS * SB + k;
                const float2* hl = halo + (((buf * NP + p) * 2) * RS * WS + ((k >> 5) + 1)) * HQ;
                const int lk = k & 31;
                float2 yX = make_float2(0.f, 0.f), yY = yX;
#pragma unroll
                for (int h = 0; h < RS; h++) {
                    const float2 u = fb[h * SB], v = fb[(RS + h) * SB];
                    yX.x += u.x; yX.y += u.y; yY.x += v.x; yY.y += v.y;
#pragma unroll
                    for (int q = 1; q < Q; q++) {
                        if (lk + q >= 32) {
                            const float2 cu = hl[h * WS * HQ + q * (q - 1) / 2 + (lk + q - 32)];
                            const float2 cv = hl[(RS + h) * WS * HQ + q * (q - 1) / 2 + (lk + q - 32)];
                            yX.x += cu.x; yX.y += cu.y; yY.x += cv.x; yY.y += cv.y;
                        }
                    }
                }
                if constexpr (HALF == 2) {
                    // the lower-half launch has stored its sums already (stream order): add this half
                    float2 ya = sOutPtr[a][k0 + k];
                    ya.x += yX.x; ya.y += yY.x;
                    sOutPtr[a][k0 + k] = ya;
                    if (a + 1 < nG) {
                        float2 yb = sOutPtr[a + 1][k0 + k];
                        yb.x += yX.y; yb.y += yY.y;
                        sOutPtr[a + 1][k0 + k] = yb;
                    }
                }
                else {
                    sOutPtr[a][k0 + k] = make_float2(yX.x, yY.x);
                    if (a + 1 < nG) { sOutPtr[a + 1][k0 + k] = make_float2(yX.y, yY.y); }
                }
            }
        }
    }
    }   // work items
}

// ---- single-VFO variant with 16-byte staging --------------------------------------------------------------------
// With one or two VFOs the kernel is bound by how fast an SM can take samples in, not by arithmetic.  This variant
// moves the tile with 16-byte cp.async.cg (two samples per copy, no L1 allocation) and reads it back two samples per
// LDS.128, on an even sub-block pitch (D + 2) where those accesses are aligned and conflict-free.  16-byte copies need
// the tile to start on an even sample: when (offset - H) is odd the tile starts one sample early and the taps are
// used shifted by one (h'[0] = 0, h'[t] = h[t-1]), which changes no result (ODD = 1).
// PERSIST = 1: CTAs stride over the tiles with the next tile in flight (double buffer); 0: one tile per CTA.
#ifndef B200_STREAM_DS
#define B200_STREAM_DS 32
#endif
template <int D> struct StreamCfg {
    static constexpr int SB = (2048 / D) > 256 ? 256 : ((2048 / D) < 32 ? 32 : (2048 / D));
    static constexpr int RS = (D / B200_STREAM_DS) > 8 ? 8 : ((D / B200_STREAM_DS) < 1 ? 1 : (D / B200_STREAM_DS));
    static constexpr int DS = D / RS;
    static constexpr int THREADS = SB * RS;
};

// one thread's run of DS samples, two samples per shared-memory load; taps shifted by ODD
template <int RATIO, int RUN, int DS, int ODD>
__device__ __forceinline__ void accumulateRun2(const float4* __restrict__ x2, const float4* __restrict__ w2, float2* acc) {
    using FS = FirstStage<RATIO>;
    constexpr int D = FS::D, T = FS::T + ODD;
    constexpr int Q = (T + D - 1) / D;
#pragma unroll
    for (int r2 = 0; r2 < DS / 2; r2++) {
        const float4 xx = x2[r2], ww = w2[r2];
        const float2 z0 = cmulf(make_float2(xx.x, xx.y), make_float2(ww.x, ww.y));
        const float2 z1 = cmulf(make_float2(xx.z, xx.w), make_float2(ww.z, ww.w));
#pragma unroll
        for (int q = 0; q < Q; q++) {
            constexpr int dummy = 0; (void)dummy;
            const int t0 = q * D + RUN * DS + 2 * r2, t1 = t0 + 1;
            if (t0 < T && t0 - ODD >= 0) { const float h = FS::tapImm(t0 - ODD); acc[q].x = fmaf(h, z0.x, acc[q].x); acc[q].y = fmaf(h, z0.y, acc[q].y); }
            if (t1 < T && t1 - ODD >= 0) { const float h = FS::tapImm(t1 - ODD); acc[q].x = fmaf(h, z1.x, acc[q].x); acc[q].y = fmaf(h, z1.y, acc[q].y); }
        }
    }
}

template <int RATIO, int ODD, int PERSIST>
__global__ void __launch_bounds__(StreamCfg<FirstStage<RATIO>::D>::THREADS)
fusedStreamKernel(const FusedJob* __restrict__ jobs, int tiles) {
    using FS = FirstStage<RATIO>;
    using SC = StreamCfg<FS::D>;
    constexpr int D = FS::D, T = FS::T + ODD, H = FS::T - 1;
    constexpr int Q = (T + D - 1) / D;
    static_assert(Q <= 8, "accumulator array is sized for Q <= 8");
    constexpr int SB = SC::SB, RS = SC::RS, DS = SC::DS, NT = SC::THREADS;
    constexpr int NB = SB - (Q - 1);
    constexpr int LP = D + 2;                        // even pitch: 16-byte aligned rows
    constexpr int NBUF = PERSIST ? 2 : 1;
    constexpr bool kImmediateStaging = ((2 * NT) % D == 0);
    extern __shared__ __align__(16) float2 sm[];
    float2* xs = sm;                        // NBUF * SB * LP
    float2* ws = xs + NBUF * SB * LP;       // D
    float2* part = ws + D;                  // Q * NT
    float2* pPh = part + Q * NT;            // NBUF * SB
    const int tid = threadIdx.x;
    const int sb = tid % SB, run = tid / SB;
    const FusedJob j = jobs[blockIdx.y];
    const float2* in = reinterpret_cast<const float2*>(j.in);
    float2* out = reinterpret_cast<float2*>(j.out);

    auto issue = [&](int tile, float2* dst) {
        const int mStart = j.offset - H - ODD + tile * NB * D;
        if (kImmediateStaging && mStart >= 0 && mStart + SB * D <= j.count) {
            // pair e = tid + u*NT holds samples 2e, 2e+1 -> slot 2e + 2*(2e/D)
            const int i0 = 2 * tid;
            const float2* g = in + mStart + i0;
            float2* d = dst + i0 + 2 * (i0 / D);
#pragma unroll
            for (int u = 0; u < (SB * D) / (2 * NT); u++) { cpAsync16(d + u * (2 * NT + 2 * (2 * NT / D)), g + u * 2 * NT); }
        }
        else {
            for (int i = tid; i < SB * D; i += NT) {
                const int m = mStart + i;
                const bool ok = (m >= 0 && m < j.count);
                cpAsync8(dst + i + 2 * (i / D), in + (ok ? m : 0), ok);
            }
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };

    int tile = blockIdx.x;
    if (tile >= tiles) { return; }
    issue(tile, xs);
    for (int r = tid; r < D; r += NT) { ws[r] = __ldg(j.wtab + r); }
    for (int it = 0; tile < tiles; tile += gridDim.x, it++) {
        float2* xt = xs + (PERSIST ? (it & 1) : 0) * (SB * LP);
        float2* ph = pPh + (PERSIST ? (it & 1) : 0) * SB;
        const int k0 = tile * NB;
        const int mStart = j.offset - H - ODD + k0 * D;
        if (tid < SB) {                                  // sub-block start phases, while the tile is still in flight
            const int ms = mStart + tid * D;
            float2 p = make_float2(0.f, 0.f);
            if (ms < j.count && ms + D > 0) { p = phaseAtAny(j, ms); }
            ph[tid] = p;
        }
        asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        __syncthreads();                                 // tile and phases visible; everyone is done with the previous tile
        if (PERSIST && tile + (int)gridDim.x < tiles) { issue(tile + gridDim.x, xs + ((it & 1) ^ 1) * (SB * LP)); }
        const float4* x2 = reinterpret_cast<const float4*>(xt + sb * LP + run * DS);
        const float4* w2 = reinterpret_cast<const float4*>(ws + run * DS);
        float2 acc[8];
#pragma unroll
        for (int q = 0; q < Q; q++) { acc[q] = make_float2(0.f, 0.f); }
        switch (run) {
        case 0: accumulateRun2<RATIO, 0, DS, ODD>(x2, w2, acc); break;
        case 1: if constexpr (RS > 1) { accumulateRun2<RATIO, 1, DS, ODD>(x2, w2, acc); } break;
        case 2: if constexpr (RS > 2) { accumulateRun2<RATIO, 2, DS, ODD>(x2, w2, acc); } break;
        case 3: if constexpr (RS > 3) { accumulateRun2<RATIO, 3, DS, ODD>(x2, w2, acc); } break;
        case 4: if constexpr (RS > 4) { accumulateRun2<RATIO, 4, DS, ODD>(x2, w2, acc); } break;
        case 5: if constexpr (RS > 5) { accumulateRun2<RATIO, 5, DS, ODD>(x2, w2, acc); } break;
        case 6: if constexpr (RS > 6) { accumulateRun2<RATIO, 6, DS, ODD>(x2, w2, acc); } break;
        default: if constexpr (RS > 7) { accumulateRun2<RATIO, 7, DS, ODD>(x2, w2, acc); } break;
        }
        const float2 P = ph[sb];
#pragma unroll
        for (int q = 0; q < Q; q++) { part[q * NT + tid] = cmulf(acc[q], P); }
        __syncthreads();
        if (tid < NB && k0 + tid < j.nOut) {
            float2 y = make_float2(0.f, 0.f);
#pragma unroll
            for (int q = 0; q < Q; q++) {
#pragma unroll
                for (int h = 0; h < RS; h++) {
                    const float2 c = part[q * NT + h * SB + tid + q];
                    y.x += c.x; y.y += c.y;
                }
            }
            out[k0 + tid] = y;
        }
        if (!PERSIST) { break; }
    }
}

// One CTA per VFO: (1) add the contribution of the carried, already-rotated history to the first outputs,
// (2) write the rotated history for the next call: histNext[i] = workbuf[count + i] (decimating_fir.h:65).
template <int RATIO>
__global__ void __launch_bounds__(256) fusedEdgeKernel(const FusedJob* __restrict__ jobs) {
    using FS = FirstStage<RATIO>;
    constexpr int D = FS::D, T = FS::T, H = T - 1;
    const FusedJob j = jobs[blockIdx.x];
    const float2* hist = reinterpret_cast<const float2*>(j.hist);
    float2* out = reinterpret_cast<float2*>(j.out);
    // outputs whose window starts inside the history: offset + k*D < H
    const int nHead = (j.offset < H) ? min(j.nOut, (H - j.offset + D - 1) / D) : 0;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int k = warp; k < nHead; k += 8) {
        const int p0 = j.offset + k * D;            // first work-buffer position of output k
        float2 a = make_float2(0.f, 0.f);
        for (int t = lane; p0 + t < H && t < T; t += 32) {
            const float h = FS::tap(t);
            const float2 z = __ldg(hist + p0 + t);
            a.x = fmaf(h, z.x, a.x); a.y = fmaf(h, z.y, a.y);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a.x += __shfl_xor_sync(0xffffffffu, a.x, o);
            a.y += __shfl_xor_sync(0xffffffffu, a.y, o);
        }
        if (lane == 0) { float2 y = out[k]; y.x += a.x; y.y += a.y; out[k] = y; }
    }
    const float2* in = reinterpret_cast<const float2*>(j.in);
    float2* hn = reinterpret_cast<float2*>(j.histNext);
    for (int i = threadIdx.x; i < H; i += 256) {
        const int p = j.count + i;                  // position in [hist | rotated block]
        float2 v;
        if (p < H) { v = __ldg(hist + p); }
        else {
            const int m = p - H;
            v = cmulf(__ldg(in + m), rotPhaseAt(j.chk, j.wtab, m));
        }
        hn[i] = v;
    }
}

template <int RATIO, int NP, int TILE, int ODD, int HALF, bool CT>
int launchBankKernel(const BankTablesT<CT>& tp, const FusedJob* dJobs, int nJobs, int nOut, int GV, cudaStream_t s) {
    using FS = FirstStage<RATIO>;
    constexpr int D = FS::D, T = FS::T + ODD;
    constexpr int Q = (T + D - 1) / D;
    using TC = TileCfg<D, TILE>;
    constexpr int SB = TC::SB, RS = HALF ? TC::RS / 2 : TC::RS;
    constexpr int NB = SB - (Q - 1);
    constexpr size_t smem = ((size_t)SB * ((HALF ? D / 2 : D) + 2) + (size_t)2 * 2 * NP * D + (size_t)2 * NP * 2 * RS * SB
                             + (size_t)2 * NP * 2 * RS * (SB / 32) * (Q * (Q - 1) / 2)) * sizeof(float2);
    static bool attrSet = false;        // per instantiation
    if (!attrSet) {
        B200_CUDA(cudaFuncSetAttribute(fusedBankKernel<RATIO, NP, TILE, ODD, HALF, CT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attrSet = true;
    }
    const int tiles = (nOut + NB - 1) / NB;
    const long long items = (long long)tiles * ((nJobs + GV - 1) / GV);
    if (items > 0x7fffffffLL) { setError("first-stage launch of %lld work items is too large", items); return B200_ETOOBIG; }
    unsigned grid = (unsigned)items;
    if (fusedLeaveRoom) {
        // Streaming mode (B200_OPT_DEFER_TAIL) with B200_OPT_LEAVE_ROOM: the previous block's later stages run on their own
        // low-priority stream at the same time; a persistent grid that leaves them about 40 KB of shared memory and 17 K
        // registers per SM (one CTA of the FIR job kernels) keeps that room free for the whole launch.
        static int perSm = 0, regs = 0;         // per instantiation
        if (perSm == 0) {
            cudaFuncAttributes fa;
            B200_CUDA(cudaFuncGetAttributes(&fa, fusedBankKernel<RATIO, NP, TILE, ODD, HALF, CT>));
            regs = fa.numRegs;
            B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, fusedBankKernel<RATIO, NP, TILE, ODD, HALF, CT>, SB * RS, smem));
            const int bySmem = (int)((227 * 1024 - 40 * 1024) / (smem + 1024));
            const int byRegs = (65536 - 17 * 1024) / std::max(1, regs * SB * RS);
            perSm = std::max(1, std::min(perSm, std::min(bySmem, byRegs)));
        }
        grid = (unsigned)std::min<long long>(items, 148LL * perSm);
    }
    fusedBankKernel<RATIO, NP, TILE, ODD, HALF, CT><<<grid, SB * RS, smem, s>>>(tp, dJobs, nJobs, GV, (unsigned)items);
    B200_CHECK_LAUNCH(); countLaunch();
    return B200_OK;
}

// VFOs per CTA: the more, the better the tile load and the CTA prologue are amortised; but keep the SMs full
static int pickRange(int nJobs, int tiles, int np) {
    int GV = std::min(nJobs, std::min(kMaxVfosPerCta, std::max(2 * np, fusedVfosPerCta)));
    // (ranges start at even VFO indices: the pairs of a launch are numbered through it)
    while (GV > 2 * np && (long long)tiles * ((nJobs + GV - 1) / GV) < 148 * 8) { GV = std::max(2 * np, ((GV / 2) + 1) & ~1); }
    return GV;
}

template <int RATIO, int NP, int TILE, int ODD>
int launchMainV(const FusedJob* dJobs, const FusedJob* hJobs, int nJobs, int nOut, int count, cudaStream_t s) {
    using FS = FirstStage<RATIO>;
    constexpr int D = FS::D, T = FS::T + ODD;
    constexpr int Q = (T + D - 1) / D;
    using TC = TileCfg<D, TILE>;
    const int tiles = (nOut + (TC::SB - (Q - 1)) - 1) / (TC::SB - (Q - 1));
    size_t slot = 0;
    // algorithmic bytes: the block is read once for the whole bank, each VFO writes nOut samples; algorithmic flops
    // (SURVEY 8(d)): 6 for the rotation + 4 per tap and output, i.e. 6 + 4 T/D per input sample and VFO
    B200_TRY(profBegin(PROF_STAGE1, s, 8.0 * (double)count + 8.0 * (double)nOut * (double)nJobs,
                       (6.0 * (double)count + 4.0 * (double)FS::T * (double)nOut) * (double)nJobs, &slot));
    BankTables tp;               // launch arguments are copied at launch time: one host buffer serves every launch of the group
    BankTablesT<false> none{};
    // D = 128 would carry only 30 VFOs per launch: measured slower than the shared-memory tables (cfg4: 14.3 against 12.3 ms)
    bool haveHost = fusedParamTables && NP == 1 && D <= 64;
    for (int i = 0; i < nJobs && haveHost; i++) { haveHost = hJobs[i].hWtab != nullptr; }
    if (haveHost) {
        // the tables ride in the parameter block: as many pairs per launch as fit (the block is re-read per launch, from
        // L2 / HBM at a small fraction of what this FP32-bound kernel leaves unused)
        constexpr int kPairs = kParamEntries / D;
        for (int c0 = 0; c0 < nJobs; c0 += 2 * kPairs) {
            const int nc = std::min(2 * kPairs, nJobs - c0);
            for (int q = 0; 2 * q < nc; q++) {
                const float2* wA = hJobs[c0 + 2 * q].hWtab;
                const float2* wB = (2 * q + 1 < nc) ? hJobs[c0 + 2 * q + 1].hWtab : nullptr;
                float4* dst = tp.w + (size_t)q * D;
                for (int r = 0; r < D; r++) { dst[r] = make_float4(wA[r].x, wB ? wB[r].x : 0.f, wA[r].y, wB ? wB[r].y : 0.f); }
            }
            B200_TRY((launchBankKernel<RATIO, NP, TILE, ODD, 0, true>(tp, dJobs + c0, nc, nOut, pickRange(nc, tiles, NP), s)));
        }
    }
    else {
        const int GV = pickRange(nJobs, tiles, NP);
        // a loop body beyond the 32 KB instruction cache level runs as two launches of half the runs each: every launch
        // stages only its half of each sub-block's samples (half the shared memory, twice the CTAs per SM) and its code
        // (17 KB for the /8192 plan) stays in the instruction cache; the second launch adds to the first one's outputs
        constexpr int kBodyBytes = (int)(D * (5.5 + 2.0 * (double)T / (double)D) * 16.0) * NP;
        if constexpr (kBodyBytes > B200_SPLIT_BODY_BYTES && TC::RS % 2 == 0 && TC::RS >= 2) {
            if (fusedSplitRuns) {
                B200_TRY((launchBankKernel<RATIO, NP, TILE, ODD, 1, false>(none, dJobs, nJobs, nOut, GV, s)));
                B200_TRY((launchBankKernel<RATIO, NP, TILE, ODD, 2, false>(none, dJobs, nJobs, nOut, GV, s)));
            }
            else { B200_TRY((launchBankKernel<RATIO, NP, TILE, ODD, 0, false>(none, dJobs, nJobs, nOut, GV, s))); }
        }
        else { B200_TRY((launchBankKernel<RATIO, NP, TILE, ODD, 0, false>(none, dJobs, nJobs, nOut, GV, s))); }
    }
    B200_TRY(profEnd(PROF_STAGE1, slot, s));
    return B200_OK;
}

template <int RATIO, int NP, int TILE>
int launchMain(const FusedJob* dJobs, const FusedJob* hJobs, int nJobs, int nOut, int count, bool odd, cudaStream_t s) {
    return odd ? launchMainV<RATIO, NP, TILE, 1>(dJobs, hJobs, nJobs, nOut, count, s) : launchMainV<RATIO, NP, TILE, 0>(dJobs, hJobs, nJobs, nOut, count, s);
}

template <int RATIO, int ODD, int PERSIST>
int launchStreamV(const FusedJob* dJobs, int nJobs, int nOut, int count, cudaStream_t s) {
    using FS = FirstStage<RATIO>;
    using SC = StreamCfg<FS::D>;
    constexpr int D = FS::D, T = FS::T + ODD;
    constexpr int Q = (T + D - 1) / D;
    constexpr int NB = SC::SB - (Q - 1);
    constexpr int NBUF = PERSIST ? 2 : 1;
    const size_t smem = ((size_t)NBUF * SC::SB * (D + 2) + D + (size_t)Q * SC::THREADS + NBUF * SC::SB) * sizeof(float2);
    B200_CUDA(cudaFuncSetAttribute(fusedStreamKernel<RATIO, ODD, PERSIST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int tiles = (nOut + NB - 1) / NB;
    int gx = tiles;
    if (PERSIST) {
        static int perSm = 0;           // resident CTAs per SM of this instantiation
        if (perSm == 0) {
            B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, fusedStreamKernel<RATIO, ODD, PERSIST>, SC::THREADS, smem));
            if (perSm < 1) { perSm = 1; }
        }
        const int maxX = std::max(1, (148 * perSm) / nJobs);
        const int rounds = (tiles + maxX - 1) / maxX;          // equal numbers of tiles per CTA
        gx = (tiles + rounds - 1) / rounds;
    }
    size_t slot = 0;
    B200_TRY(profBegin(PROF_STAGE1, s, 8.0 * (double)count + 8.0 * (double)nOut * (double)nJobs,
                       (6.0 * (double)count + 4.0 * (double)FS::T * (double)nOut) * (double)nJobs, &slot));
    fusedStreamKernel<RATIO, ODD, PERSIST><<<dim3(gx, nJobs), SC::THREADS, smem, s>>>(dJobs, tiles);
    B200_CHECK_LAUNCH(); countLaunch();
    B200_TRY(profEnd(PROF_STAGE1, slot, s));
    return B200_OK;
}

// odd: whether (input pointer / 8 + offset - H) is odd, i.e. whether the tile must start one sample early
template <int RATIO>
int launchStream(const FusedJob* dJobs, int nJobs, int nOut, int count, bool odd, cudaStream_t s) {
    // (the persistent, double-buffered form -- PERSIST = 1 -- was measured slower: 0.85 against 0.94 of the copy peak)
    return odd ? launchStreamV<RATIO, 1, 0>(dJobs, nJobs, nOut, count, s) : launchStreamV<RATIO, 0, 0>(dJobs, nJobs, nOut, count, s);
}

template <int RATIO>
int launchPlan(const FusedJob* dJobs, const FusedJob* hJobs, int nJobs, int nOut, int count, const void* in, int offset, cudaStream_t s) {
    using FS = FirstStage<RATIO>;
    constexpr int D = FS::D, T = FS::T;
    constexpr int Q = (T + D - 1) / D;
    if (nOut > 0) {
        // tile start parity: 16-byte copies need an even sample index at a 16-byte aligned address (+ 2^20 keeps the sum non-negative)
        const bool odd = ((((uintptr_t)in >> 3) + (uintptr_t)(long long)(offset - (T - 1) + (1 << 20))) & 1) != 0;
        const bool single = nJobs == 1 && fusedStreamSingle && fusedTileSamples != kTileLarge;
        if (single) {
            // one VFO: HBM-bound, 2048-sample tiles, one tile per CTA
            B200_TRY((launchStream<RATIO>(dJobs, nJobs, nOut, count, odd, s)));
        }
        else {
            // D = 128 (ratio 8192): 64 sub-blocks per tile instead of 32, so that the Q - 1 sub-blocks a tile recomputes for
            // its neighbour are 8 % of it instead of 16 %
            constexpr int kTile = (D >= 128 && !B200_D128_SMALL_TILE) ? 2 * kTileLarge : kTileLarge;
            const int NB = TileCfg<D, kTile>::SB - (Q - 1);
            const int tiles = (nOut + NB - 1) / NB;
            // VFOs per register pass: the loop body is fully unrolled over the D samples of a sub-block (every tap is an
            // immediate) and all RS run variants sit in the same code: about D * (5.5 + 2 T/D) instructions of 16 bytes per
            // PAIR of VFOs.  Two pairs per pass halve the tile reads per VFO, but the body has to stay inside the 32 KB
            // level of the instruction cache (round 1: a 57 KB body stalled on instruction fetch for 39 % of its samples).
            // B200_OPT_VFOS_PER_PASS overrides the choice (2 or 4; results do not depend on it).
            constexpr int kBodyPair = (int)(D * (5.5 + 2.0 * (double)T / (double)D) * 16.0);
            constexpr bool kTwoPairsFit = (2 * kBodyPair <= 34 * 1024) && (2 * D <= TileCfg<D, kTile>::THREADS);
            // measured (cfg3 / cfg4): one pair per pass is faster everywhere -- two pairs need 110-140 registers, and the resident
            // warps that costs outweigh the halved tile reads (/128 plan: 1 CTA per SM, FMA pipe 38 % busy against 84 % for the
            // one-pair /4096 plan) -- so two pairs run only when asked for
            int np = 1;
            if (fusedVfosPerPass == 4 && kTwoPairsFit) { np = 2; }
            if constexpr (kTwoPairsFit) {
                if (np == 2) { B200_TRY((launchMain<RATIO, 2, kTile>(dJobs, hJobs, nJobs, nOut, count, odd, s))); }
                else { B200_TRY((launchMain<RATIO, 1, kTile>(dJobs, hJobs, nJobs, nOut, count, odd, s))); }
            }
            else { B200_TRY((launchMain<RATIO, 1, kTile>(dJobs, hJobs, nJobs, nOut, count, odd, s))); }
        }
    }
    fusedEdgeKernel<RATIO><<<nJobs, 256, 0, s>>>(dJobs);
    B200_CHECK_LAUNCH(); countLaunch();
    return B200_OK;
}

int launchPlanDyn(int ratio, const FusedJob* dJobs, const FusedJob* hJobs, int nJobs, int nOut, int count, const void* in, int offset, cudaStream_t s) {
    switch (ratio) {
    case 16: return launchPlan<16>(dJobs, hJobs, nJobs, nOut, count, in, offset, s);
    case 32: return launchPlan<32>(dJobs, hJobs, nJobs, nOut, count, in, offset, s);
    case 64: return launchPlan<64>(dJobs, hJobs, nJobs, nOut, count, in, offset, s);
    case 128: return launchPlan<128>(dJobs, hJobs, nJobs, nOut, count, in, offset, s);
    case 256: return launchPlan<256>(dJobs, hJobs, nJobs, nOut, count, in, offset, s);
    case 512: return launchPlan<512>(dJobs, hJobs, nJobs, nOut, count, in, offset, s);
    case 1024: return launchPlan<1024>(dJobs, hJobs, nJobs, nOut, count, in, offset, s);
    case 2048: return launchPlan<2048>(dJobs, hJobs, nJobs, nOut, count, in, offset, s);
    case 4096: return launchPlan<4096>(dJobs, hJobs, nJobs, nOut, count, in, offset, s);
    case 8192: return launchPlan<8192>(dJobs, hJobs, nJobs, nOut, count, in, offset, s);
    default: setError("no fused kernel for ratio %d", ratio); return B200_EINVAL;
    }
}

} // namespace

bool fusedEnabled = true;

bool fusedStageSupported(const Vfo& v) {
    return fusedEnabled && (v.mode == 0 || v.mode == 1) && v.ratio >= kMinFusedRatio && !v.decimStages.empty();
}

void fillFusedJob(FusedJob& f, const Stage& st, const Vfo& v, const cf32* dIn, int count) {
    f.in = dIn; f.out = (cf32*)st.dOut;
    f.hist = (const cf32*)st.dHist[st.cur]; f.histNext = (cf32*)st.dHist[st.cur ^ 1];
    f.chk = v.dChk; f.wtab = v.dWtab;
    f.hWtab = (v.hWtab.size() >= (size_t)kRotChunk && !v.tableDirty) ? v.hWtab.data() : nullptr;
    f.thetaEff = atan2((double)v.dIm, (double)v.dRe);
    f.logMag = log(hypot((double)v.dRe, (double)v.dIm));
    f.ratio = v.ratio; f.offset = st.offset; f.count = count; f.nOut = st.outCountFor(count);
}

// hostJobs must already be sorted by (ratio, offset); dJobs is the device copy in the same order
int launchFused(const FusedJob* dJobs, const std::vector<FusedJob>& hostJobs, cudaStream_t s) {
    size_t i = 0;
    while (i < hostJobs.size()) {
        size_t e = i + 1;
        while (e < hostJobs.size() && hostJobs[e].ratio == hostJobs[i].ratio && hostJobs[e].offset == hostJobs[i].offset) { e++; }
        B200_TRY(launchPlanDyn(hostJobs[i].ratio, dJobs + i, hostJobs.data() + i, (int)(e - i), hostJobs[i].nOut, hostJobs[i].count, hostJobs[i].in, hostJobs[i].offset, s));
        i = e;
    }
    return B200_OK;
}

} // namespace b200

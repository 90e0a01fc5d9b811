// Demodulators fed by the VFO bank (stage 2 of the path; decoder_modules/radio/src/radio_module.h:76-107):
//   NFM  dsp::demod::FM<stereo_t>    demod/fm.h:86-103  = Quadrature (quadrature.h:39-46) -> real FIR -> mono->stereo
//   SSB  dsp::demod::SSB<stereo_t>   demod/ssb.h:82-97   = FrequencyXlator(+-bw/2) -> real part -> AGC<float> -> stereo
//   AM   dsp::demod::AM<stereo_t>    demod/am.h:101-133  = [carrier AGC<complex>] -> |x| -> DCBlocker<float> -> [audio AGC<float>] -> FIR -> stereo
//   AGC  dsp::loop::AGC<T>           loop/agc.h:85-139   (data-dependent attack/decay recursion + 4800-sample start ramp)
// The quadrature detector, the rotation, the filters and the interleave are data parallel.  AGC and the
// 1-pole DC blocker are sequential recurrences per VFO: one warp per VFO stages a chunk in shared memory,
// lane 0 walks the recurrence (everything that does not depend on the running state is precomputed by the
// whole warp), then all lanes apply gains and write out.  Many VFOs run side by side in one launch.
#include "demod.cuh"

#include <algorithm>
#include <cmath>

#include "vfo_bank.cuh"

namespace b200 {

struct AgcParams { float setPoint, attack, invAttack, decay, invDecay, maxGain; };
struct AgcDev { float amp; int startEnvelope; };

// device-resident sequential state of one demodulator
struct SeqState {
    AgcDev carrier, audio;
    float dcOffset;
    float quadPhase[2];   // ping-pong: written by one call, read by the next
    float2 rotPhase;      // SSB rotator, EXACT mode: the reference's fp32 phase
    float2 deLast;        // Deemphasis<stereo_t>::lastOut (deephasis.h:92)
    float nbAmp;          // NoiseBlanker::amp (noise_blanker.h:62)
    float pllPhase, pllFreq;   // PhaseControlLoop::phase / freq of the WFM pilot PLL (phase_control_loop.h:69-70)
};

struct DemodState {
    int kind = B200_DEMOD_NONE;
    double p0 = 0, p1 = 0, p2 = 0, p3 = 0;
    double sr = 0, bw = 0;
    // SSB rotation (FrequencyXlator inside SSB, ssb.h:27,84)
    float dRe = 1.f, dIm = 0.f; double theta = 0.0, angle = 0.0;
    // NFM
    float invDeviation = 0.f; bool filtering = false; int quadParity = 0;
    AgcParams agc{};
    int agcMode = 1;
    float dcRate = 0.f;
    Stage fir; bool hasFir = false;          // real-valued FIR (FM audio filter / AM low-pass)
    SeqState* dState = nullptr;
    float2* dWtab = nullptr;                 // SSB: phaseDelta^j, j < 512
    float* dMono = nullptr; float* dStereo = nullptr; int cap = 0;
    float2* dMaps = nullptr; float* dAmps = nullptr;   // SSB fast path scratch
    long long envCount = 0;                            // host mirror of AGC::_startEnvelope (SSB fast path)
    int lastCount = 0;
    // radio AF chain (radio_module.h:101-107): RationalResampler<stereo_t> to the audio rate, then Deemphasis<stereo_t>
    bool afResamp = false; double afOutSR = 0.0;
    std::vector<Stage> afDecim; Stage afPoly; bool afHasPoly = false; int afMode = 3;
    bool deOn = false; float deAlpha = 0.f;
    const float* lastAudio = nullptr; int lastAudioCount = 0;
    // radio IF chain in front of the demodulator (radio_module.h:92-98): NoiseBlanker -> Squelch
    bool nbOn = false, sqOn = false; float nbRate = 0.f, nbLevel = 0.f, sqLevel = 0.f;
    cf32* dIf = nullptr; int ifCap = 0;
    // broadcast FM stereo (demod/broadcast_fm.h:144-190)
    bool wfmStereo = false;
    Stage pilotRe, pilotIm;                  // the complex pilot band-pass on the real MPX = two real FIRs (real / imaginary taps)
    Stage firR;                              // right-channel audio low-pass (`fir` serves the left channel / mono)
    float pllAlpha = 0.f, pllBeta = 0.f, pllMinFreq = 0.f, pllMaxFreq = 0.f;
    int wfmDelay = 0; float* dDelayHist[2] = { nullptr, nullptr }; int delayCur = 0;   // math::Delay of the MPX
    float* dPllPhase = nullptr; float* dL = nullptr; float* dR = nullptr;
    // FM IF noise reduction, the last block of the IF chain (noise_reduction/fm_if.h; radio_module.h:93,98)
    bool fmOn = false; int fmBins = 0;
    float2* dFmTab = nullptr;                // [32][32] window x forward twiddle, then [32] backward twiddles of element bins/2
    cf32* dFmHist[2] = { nullptr, nullptr }; int fmCur = 0;   // the bins-1 samples in front of the next block
    cf32* dFm = nullptr; int fmCap = 0;
    // sink-side volume (audio/volume.h; signal_path/sink.h:73)
    bool volOn = false; float volGain = 1.0f;
    // RDS branch of the broadcast FM demodulator (broadcast_fm.h:52-53,169-175,197-206)
    bool rdsOn = false;
    float rdsDRe = 1.f, rdsDIm = 0.f; double rdsTheta = 0.0, rdsAngle = 0.0;
    float2* dRdsWtab = nullptr; float2* dRdsChk = nullptr; float2* dRdsPhase = nullptr; cf32* dRdsX = nullptr; int rdsCap = 0;
    std::vector<Stage> rdsDecim; Stage rdsPoly; bool rdsHasPoly = false; int rdsMode = 3;
    const float* lastRds = nullptr; int lastRdsCount = 0;
    const cf32* lastIf = nullptr;            // what the demodulator read in the last call (the IF chain's output)
};

// per-call descriptors of the WFM stereo stages
struct PllJob { const float* pr; const float* pi; float* phaseOut; SeqState* st; float alpha, beta, minFreq, maxFreq; int n, pad; };   // pllKernel or pllErrKernel: the bank's rotator mode decides
struct WfmMixJob { const float* mpx; const float* hist; float* histNext; const float* phase; float* l; float* r; int n, delay; };
struct LrJob { const float* l; const float* r; float* stereo; int n, pad; };

// per-call descriptor of the IF chain
struct IfJob { const cf32* in; cf32* out; SeqState* st; float rate, invRate, nbLevel, sqLevel; int n, nbOn, sqOn, pad; };

// per-call descriptors of the FM IF noise reduction, the volume block and the RDS down-conversion
struct FmifJob { const cf32* in; cf32* out; const cf32* hist; cf32* histNext; const float2* tab; int n, bins; };
struct VolJob { float* data; float gain; int nFloats; };
struct RdsXlateJob { const float* mpx; cf32* out; const float2* chk; const float2* wtab; int n, pad; };

// per-call descriptor of the de-emphasis walk
struct DeemphJob { float2* data; SeqState* st; float alpha; int n; };

// per-call descriptor
struct DemodJob {
    const cf32* in; float* mono; float* stereo; const float* stereoSrc; SeqState* st; const float2* wtab;
    float2* maps; float* amps;      // SSB fast path scratch: per-sample AGC intercepts (|x|*attack, |x|*decay; -1 marks a sample the AGC skips), walked amplitudes
    AgcParams agc;
    double angle0, theta; float stepRe, stepIm;
    float invDeviation, dcRate;
    int kind, agcMode, n, quadParity, exact, env0;
};

namespace {

constexpr int kSeqThreads = 128;
constexpr int kChunk = 1024;     // multiple of 512 so rotator renormalisation points fall on chunk-relative multiples of 512
constexpr int kEnvLen = 4800;    // AGC::_totalEnvelopeLength (agc.h:162)

// AGC recurrence over a staged chunk (agc.h:96-112).  The average moves only for non-zero, non-NaN amplitudes;
// a skipped update means gain 1.  The reference's branch
//     amp' = (a > amp) ? amp*(1-att) + a*att : amp*(1-dec) + a*dec
// equals max(A, B) of the two candidates when att >= dec and min(A, B) when att < dec (A - B = (a-amp)(att-dec)),
// and a skipped sample is the identity map (slope 1, intercept 0) in the same form.  The whole CTA precomputes
// per-sample slopes and intercepts, so the single thread that walks the recurrence issues only
// 2 FFMA + 1 FMNMX per sample (~9 cycles of loop-carried latency), and the CTA turns amplitudes into gains
// afterwards.  sAmp: |in| on entry, the gain on exit.
struct AgcScratch { float *sa, *ia, *sd, *id, *r; };

__device__ void agcChunk(float* sAmp, const AgcScratch& w, int n, float* ampState, const AgcParams& p) {
    const int tid = threadIdx.x;
    const int n8 = (n + 7) & ~7;
    for (int i = tid; i < n8; i += kSeqThreads) {
        const float a = (i < n) ? sAmp[i] : 0.0f;
        const bool ok = (a != 0.0f) && (a == a);
        w.sa[i] = ok ? p.invAttack : 1.0f; w.ia[i] = ok ? a * p.attack : 0.0f;
        w.sd[i] = ok ? p.invDecay : 1.0f;  w.id[i] = ok ? a * p.decay : 0.0f;
    }
    __syncthreads();
    if (tid == 0) {
        float am = *ampState;
        const bool useMax = p.attack >= p.decay;
        for (int i = 0; i < n8; i += 8) {
            float sa[8], ia[8], sd[8], id[8], r[8];
#pragma unroll
            for (int u = 0; u < 8; u += 4) {
                *reinterpret_cast<float4*>(&sa[u]) = *reinterpret_cast<const float4*>(&w.sa[i + u]);
                *reinterpret_cast<float4*>(&ia[u]) = *reinterpret_cast<const float4*>(&w.ia[i + u]);
                *reinterpret_cast<float4*>(&sd[u]) = *reinterpret_cast<const float4*>(&w.sd[i + u]);
                *reinterpret_cast<float4*>(&id[u]) = *reinterpret_cast<const float4*>(&w.id[i + u]);
            }
            if (useMax) {
#pragma unroll
                for (int u = 0; u < 8; u++) { am = fmaxf(fmaf(am, sa[u], ia[u]), fmaf(am, sd[u], id[u])); r[u] = am; }
            }
            else {
#pragma unroll
                for (int u = 0; u < 8; u++) { am = fminf(fmaf(am, sa[u], ia[u]), fmaf(am, sd[u], id[u])); r[u] = am; }
            }
#pragma unroll
            for (int u = 0; u < 8; u += 4) { *reinterpret_cast<float4*>(&w.r[i + u]) = *reinterpret_cast<const float4*>(&r[u]); }
        }
        *ampState = am;
    }
    __syncthreads();
    for (int i = tid; i < n; i += kSeqThreads) {
        const float a = sAmp[i];
        const bool ok = (a != 0.0f) && (a == a);
        sAmp[i] = ok ? fminf(__fdiv_rn(p.setPoint, w.r[i]), p.maxGain) : 1.0f;
    }
    __syncthreads();
}

// start ramp (agc.h:131-134): `envelope` is re-initialised to 1 on every process() call and keeps its last
// ramp value for the rest of the call in which the ramp ends.
__device__ __forceinline__ float envelopeAt(int startEnvAtCall, int i) {
    const int se = startEnvAtCall + i;
    if (se < kEnvLen) { return (float)se / (float)kEnvLen; }
    return (startEnvAtCall < kEnvLen) ? (float)(kEnvLen - 1) / (float)kEnvLen : 1.0f;
}

// one CTA per demodulator with a sequential part (SSB, AM)
__global__ void __launch_bounds__(kSeqThreads) seqDemodKernel(const DemodJob* __restrict__ jobs) {
    __shared__ __align__(16) float sA[kChunk + 8];     // amplitudes -> gains
    __shared__ __align__(16) float sSa[kChunk + 8], sIa[kChunk + 8], sSd[kChunk + 8], sId[kChunk + 8], sR[kChunk + 8];
    __shared__ __align__(16) float sV[kChunk];         // real working values
    __shared__ float2 sC[kChunk];                      // complex working values (AM) / rotator phases (SSB)
    __shared__ float2 sW[kRotChunk];                   // SSB: phaseDelta^j
    __shared__ float2 sChk[kChunk / kRotChunk];
    __shared__ SeqState sSt;
    const DemodJob j = jobs[blockIdx.x];
    const int tid = threadIdx.x;
    const AgcScratch agcW{ sSa, sIa, sSd, sId, sR };
    if (tid == 0) { sSt = *j.st; }
    if (j.kind != B200_DEMOD_AM) { for (int i = tid; i < kRotChunk; i += kSeqThreads) { sW[i] = __ldg(j.wtab + i); } }
    __syncthreads();
    const int envCarrier0 = sSt.carrier.startEnvelope, envAudio0 = sSt.audio.startEnvelope;
    const bool agcOn = j.agc.attack > 0.0f;   // agc.h:87-90: attack <= 0 copies the input through

    for (int base = 0; base < j.n; base += kChunk) {
        const int m = min(kChunk, j.n - base);
        if (j.kind == B200_DEMOD_AM) {
            // ---- carrier AGC on the complex IF (am.h:103-106)
            for (int i = tid; i < m; i += kSeqThreads) {
                const float2 x = __ldg(reinterpret_cast<const float2*>(j.in) + base + i);
                sC[i] = x;
                sA[i] = sqrtf(__fadd_rn(__fmul_rn(x.x, x.x), __fmul_rn(x.y, x.y)));   // complex_t::amplitude (types.h:77-79)
            }
            __syncthreads();
            if (j.agcMode == 0 && agcOn) {
                agcChunk(sA, agcW, m, &sSt.carrier.amp, j.agc);
                for (int i = tid; i < m; i += kSeqThreads) {
                    const float g = sA[i], e = envelopeAt(envCarrier0, base + i);
                    float2 x = sC[i];
                    x.x = __fmul_rn(__fmul_rn(x.x, g), e); x.y = __fmul_rn(__fmul_rn(x.y, g), e);
                    sC[i] = x;
                }
                __syncthreads();
            }
            // ---- magnitude (am.h:120) and 1-pole DC blocker (dc_blocker.h:54-60)
            for (int i = tid; i < m; i += kSeqThreads) {
                const float2 x = sC[i];
                sV[i] = sqrtf(__fadd_rn(__fmul_rn(x.x, x.x), __fmul_rn(x.y, x.y)));
            }
            __syncthreads();
            if (tid == 0) {
                float o = sSt.dcOffset;
                for (int i = 0; i < m; i++) {
                    const float y = __fsub_rn(sV[i], o);
                    o = __fadd_rn(o, __fmul_rn(y, j.dcRate));
                    sV[i] = y;
                }
                sSt.dcOffset = o;
            }
            __syncthreads();
            // ---- audio AGC (am.h:122-124)
            if (j.agcMode == 1 && agcOn) {
                for (int i = tid; i < m; i += kSeqThreads) { sA[i] = fabsf(sV[i]); }
                __syncthreads();
                agcChunk(sA, agcW, m, &sSt.audio.amp, j.agc);
                for (int i = tid; i < m; i += kSeqThreads) {
                    sV[i] = __fmul_rn(__fmul_rn(sV[i], sA[i]), envelopeAt(envAudio0, base + i));
                }
                __syncthreads();
            }
            for (int i = tid; i < m; i += kSeqThreads) { j.mono[base + i] = sV[i]; }
            __syncthreads();
        }
        else {
            // ---- SSB: rotate by +-bw/2 (FrequencyXlator inside SSB, ssb.h:84), keep the real part (:86-92).
            // Rotator phases come from per-512 checkpoints x the power table, like every other rotation here:
            //   EXACT  thread 0 replays the reference recursion (renormalising every 512 samples of the call and
            //          after the ragged tail) and records each sample's phase;
            //   FAST   checkpoints in closed form from the calibrated step.
            if (j.exact) {
                if (tid == 0) {
                    float pr = sSt.rotPhase.x, pi = sSt.rotPhase.y;
                    const float dr = j.stepRe, di = j.stepIm;
                    for (int c = 0; c < m; c += kRotChunk) {
                        const int mm = min(kRotChunk, m - c);
                        for (int i = 0; i < mm; i++) {
                            sC[c + i] = make_float2(pr, pi);
                            const float nr = __fsub_rn(__fmul_rn(pr, dr), __fmul_rn(pi, di));
                            const float ni = __fadd_rn(__fmul_rn(pr, di), __fmul_rn(pi, dr));
                            pr = nr; pi = ni;
                        }
                        const float h = __double2float_rn(__dsqrt_rn(__dadd_rn(__dmul_rn((double)pr, (double)pr), __dmul_rn((double)pi, (double)pi))));
                        pr = __fdiv_rn(pr, h); pi = __fdiv_rn(pi, h);
                    }
                    sSt.rotPhase = make_float2(pr, pi);
                }
            }
            else {
                if (tid < kChunk / kRotChunk) {
                    double a = fma(j.theta, (double)(base + tid * kRotChunk), j.angle0);
                    a -= 6.283185307179586476925 * rint(a * 0.15915494309189533577);
                    double sd, cd;
                    sincos(a, &sd, &cd);
                    sChk[tid] = make_float2((float)cd, (float)sd);
                }
                __syncthreads();
                for (int i = tid; i < m; i += kSeqThreads) {
                    const float2 c = sChk[i >> kRotChunkLog], w = sW[i & (kRotChunk - 1)];
                    sC[i] = make_float2(c.x * w.x - c.y * w.y, c.x * w.y + c.y * w.x);
                }
            }
            __syncthreads();
            for (int i = tid; i < m; i += kSeqThreads) {
                const float2 x = __ldg(reinterpret_cast<const float2*>(j.in) + base + i);
                const float2 ph = sC[i];
                const float re = __fsub_rn(__fmul_rn(x.x, ph.x), __fmul_rn(x.y, ph.y));
                sV[i] = re;
                sA[i] = fabsf(re);
            }
            __syncthreads();
            if (agcOn) {
                agcChunk(sA, agcW, m, &sSt.audio.amp, j.agc);
                for (int i = tid; i < m; i += kSeqThreads) {
                    sV[i] = __fmul_rn(__fmul_rn(sV[i], sA[i]), envelopeAt(envAudio0, base + i));
                }
                __syncthreads();
            }
            for (int i = tid; i < m; i += kSeqThreads) { j.mono[base + i] = sV[i]; }
            __syncthreads();
        }
    }
    if (tid == 0) {
        if (j.kind == B200_DEMOD_AM) {
            if (j.agcMode == 0 && agcOn) { sSt.carrier.startEnvelope = envCarrier0 + j.n; }
            if (j.agcMode == 1 && agcOn) { sSt.audio.startEnvelope = envAudio0 + j.n; }
        }
        else if (agcOn) { sSt.audio.startEnvelope = envAudio0 + j.n; }
        j.st->carrier = sSt.carrier; j.st->audio = sSt.audio; j.st->dcOffset = sSt.dcOffset; j.st->rotPhase = sSt.rotPhase;
    }
}

// ---------------------------------------------------------------------------------------------------------
// SSB fast path: the same computation as the SSB branch of seqDemodKernel, split so that only the AGC
// recurrence itself is sequential:
//   ssbPrepKernel  (parallel)   rotate, take the real part, turn |re| into the AGC's per-sample slope/intercept pairs
//   agcWalkKernel  (one warp per VFO)  lane 0 walks amp' = max/min(amp*sa+ia, amp*sd+id); the warp streams the
//                                     operands through double-buffered shared memory with cp.async
//   ssbFinishKernel (parallel)  gain = min(setPoint/amp, maxGain), start ramp, mono -> stereo
// Used in FAST rotator mode; EXACT mode keeps seqDemodKernel, whose thread 0 also replays the rotator recursion.
__global__ void __launch_bounds__(256) ssbPrepKernel(const DemodJob* __restrict__ jobs) {
    const DemodJob j = jobs[blockIdx.y];
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= j.n) { return; }
    const int c = i >> kRotChunkLog;
    double a = fma(j.theta, (double)(c << kRotChunkLog), j.angle0);
    a -= 6.283185307179586476925 * rint(a * 0.15915494309189533577);
    // one checkpoint per 512 samples, evaluated by the first lane of each warp and broadcast
    float2 chk;
    {
        double sd = 0.0, cd = 0.0;
        if ((threadIdx.x & 31) == 0) { sincos(a, &sd, &cd); }
        chk.x = __shfl_sync(0xffffffffu, (float)cd, 0);
        chk.y = __shfl_sync(0xffffffffu, (float)sd, 0);
    }
    const float2 w = __ldg(j.wtab + (i & (kRotChunk - 1)));
    const float2 ph = make_float2(chk.x * w.x - chk.y * w.y, chk.x * w.y + chk.y * w.x);
    const float2 x = __ldg(reinterpret_cast<const float2*>(j.in) + i);
    const float re = __fsub_rn(__fmul_rn(x.x, ph.x), __fmul_rn(x.y, ph.y));
    j.mono[i] = re;
    const float am = fabsf(re);
    const bool ok = (am != 0.0f) && (am == am);
    // the two candidates of the next amplitude are amp*(1-attack) + |x|*attack and amp*(1-decay) + |x|*decay: only the
    // intercepts depend on the sample.  A sample the AGC skips (zero or NaN amplitude, agc.h:99) is marked by -1.
    j.maps[i] = ok ? make_float2(am * j.agc.attack, am * j.agc.decay) : make_float2(-1.0f, -1.0f);
}

constexpr int kWalkChunk = 512;

__device__ __forceinline__ void cpAsync16(void* smemDst, const void* gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smemDst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc) : "memory");
}

// amp' = max/min(amp * (1-attack) + ia, amp * (1-decay) + id) over one staged chunk of intercept pairs, two samples per
// 16-byte shared-memory read; 16 samples per trip in two register sets: while one set's chain runs, the other set's
// operands are already in flight.  SKIPS: the chunk holds samples the AGC leaves the amplitude alone for (marked by a
// negative intercept); the select then sits in the chain, which only matters for silence.
template <bool MAX, bool SKIPS>
__device__ __forceinline__ float walkChunk(const float2* __restrict__ M, float* __restrict__ R, int m, float am, float sa, float sd) {
    auto one = [&](float ia, float id) {
        const float a = fmaf(am, sa, ia), d = fmaf(am, sd, id);
        const float nx = MAX ? fmaxf(a, d) : fminf(a, d);
        am = (SKIPS && ia < 0.0f) ? am : nx;
        return am;
    };
    auto walk8 = [&](const float4* q, int at) {
        float r[8];
#pragma unroll
        for (int u = 0; u < 4; u++) { r[2 * u] = one(q[u].x, q[u].y); r[2 * u + 1] = one(q[u].z, q[u].w); }
        *reinterpret_cast<float4*>(&R[at]) = make_float4(r[0], r[1], r[2], r[3]);
        *reinterpret_cast<float4*>(&R[at + 4]) = make_float4(r[4], r[5], r[6], r[7]);
    };
    const float4* M4 = reinterpret_cast<const float4*>(M);
    int i = 0;
    float4 qa[4], qb[4];
    if (m >= 16) {
#pragma unroll
        for (int u = 0; u < 4; u++) { qa[u] = M4[u]; }
    }
#pragma unroll 1
    for (; i + 16 <= m; i += 16) {
#pragma unroll
        for (int u = 0; u < 4; u++) { qb[u] = M4[(i >> 1) + 4 + u]; }
        walk8(qa, i);
        if (i + 32 <= m) {
#pragma unroll
            for (int u = 0; u < 4; u++) { qa[u] = M4[(i >> 1) + 8 + u]; }
        }
        walk8(qb, i + 8);
    }
    for (; i < m; i++) { R[i] = one(M[i].x, M[i].y); }
    return am;
}

__global__ void __launch_bounds__(32) agcWalkKernel(const DemodJob* __restrict__ jobs) {
    __shared__ __align__(16) float2 sM[2][kWalkChunk];
    __shared__ __align__(16) float sR[kWalkChunk];
    const DemodJob j = jobs[blockIdx.x];
    const int lane = threadIdx.x;
    const int nChunks = (j.n + kWalkChunk - 1) / kWalkChunk;
    float am = j.st->audio.amp;
    const bool useMax = j.agc.attack >= j.agc.decay;
    auto prefetch = [&](int c) {
        const int base = c * kWalkChunk;
        const int m = min(kWalkChunk, j.n - base);
        // 16 bytes = two samples per copy (the buffer is allocated with slack, an odd tail reads one sample past the block)
        for (int i = lane; 2 * i < m; i += 32) { cpAsync16(&sM[c & 1][2 * i], j.maps + base + 2 * i); }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };
    if (nChunks > 0) { prefetch(0); }
    for (int c = 0; c < nChunks; c++) {
        if (c + 1 < nChunks) {
            prefetch(c + 1);
            asm volatile("cp.async.wait_group 1;\n" ::: "memory");
        }
        else {
            asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        }
        __syncwarp();
        const int base = c * kWalkChunk;
        const int m = min(kWalkChunk, j.n - base);
        const float2* M = sM[c & 1];
        bool skip = false;
        for (int i = lane; i < m; i += 32) { skip = skip || (M[i].x < 0.0f); }
        const bool skips = __any_sync(0xffffffffu, skip);
        // Every lane walks the same chain: uniform control flow (no divergence bookkeeping around the walk), operands read
        // as shared-memory broadcasts, identical stores.
        if (useMax) { am = skips ? walkChunk<true, true>(M, sR, m, am, j.agc.invAttack, j.agc.invDecay) : walkChunk<true, false>(M, sR, m, am, j.agc.invAttack, j.agc.invDecay); }
        else { am = skips ? walkChunk<false, true>(M, sR, m, am, j.agc.invAttack, j.agc.invDecay) : walkChunk<false, false>(M, sR, m, am, j.agc.invAttack, j.agc.invDecay); }
        __syncwarp();
        for (int i = lane; i < m; i += 32) { j.amps[base + i] = sR[i]; }
        __syncwarp();
    }
    if (lane == 0) { j.st->audio.amp = am; }
}

__global__ void __launch_bounds__(256) ssbFinishKernel(const DemodJob* __restrict__ jobs) {
    const DemodJob j = jobs[blockIdx.y];
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= j.n) { return; }
    const float re = j.mono[i];
    const float a = fabsf(re);
    const bool ok = (a != 0.0f) && (a == a);
    const float gain = ok ? fminf(__fdiv_rn(j.agc.setPoint, __ldg(j.amps + i)), j.agc.maxGain) : 1.0f;
    const float v = __fmul_rn(__fmul_rn(re, gain), envelopeAt(j.env0, i));
    reinterpret_cast<float2*>(j.stereo)[i] = make_float2(v, v);
}

// Quadrature detector (quadrature.h:39-46): out[i] = normalizePhase(phase(x[i]) - phase(x[i-1])) / deviation
__global__ void __launch_bounds__(256) quadratureKernel(const DemodJob* __restrict__ jobs) {
    const DemodJob j = jobs[blockIdx.y];
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= j.n) { return; }
    const float2 x = __ldg(reinterpret_cast<const float2*>(j.in) + i);
    const float ph = atan2f(x.y, x.x);
    float prev;
    if (i == 0) { prev = j.st->quadPhase[j.quadParity]; }
    else { const float2 xp = __ldg(reinterpret_cast<const float2*>(j.in) + i - 1); prev = atan2f(xp.y, xp.x); }
    float d = __fsub_rn(ph, prev);
    const float PI = 3.1415926535f;   // FL_M_PI (math/constants.h:4)
    if (d > PI) { d = __fsub_rn(d, 2.0f * PI); }
    else if (d <= -PI) { d = __fadd_rn(d, 2.0f * PI); }
    j.mono[i] = __fmul_rn(d, j.invDeviation);
    if (i == j.n - 1) { j.st->quadPhase[j.quadParity ^ 1] = ph; }
}

// MonoToStereo (convert/mono_to_stereo.h:12-15)
__global__ void __launch_bounds__(256) stereoKernel(const DemodJob* __restrict__ jobs) {
    const DemodJob j = jobs[blockIdx.y];
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= j.n) { return; }
    const float v = __ldg(j.stereoSrc + i);
    reinterpret_cast<float2*>(j.stereo)[i] = make_float2(v, v);
}

// ---- broadcast FM stereo ------------------------------------------------------------------------------------
// Pilot PLL (loop/pll.h:64-70 over phase_control_loop.h:58-66), one warp per VFO.  The warp turns a chunk of pilot
// samples into phases (atan2f) in shared memory, then walks the second-order loop -- frequency and phase both depend
// on the previous sample's error, so the walk is sequential -- and writes the VCO phase of every sample out.
constexpr int kPllChunk = 512;
__global__ void __launch_bounds__(32) pllKernel(const PllJob* __restrict__ jobs) {
    const PllJob j = jobs[blockIdx.x];
    __shared__ __align__(16) float sIn[kPllChunk], sOut[kPllChunk];
    const int lane = threadIdx.x;
    const float PI = 3.1415926535f;      // FL_M_PI (math/constants.h:4); phase limits -PI..PI, phaseDelta = 2*PI
    float phase = j.st->pllPhase, freq = j.st->pllFreq;
    for (int base = 0; base < j.n; base += kPllChunk) {
        const int m = min(kPllChunk, j.n - base);
        for (int i = lane; i < m; i += 32) { sIn[i] = atan2f(__ldg(j.pi + base + i), __ldg(j.pr + base + i)); }
        __syncwarp();
        // every lane walks the same loop (uniform control flow, broadcast reads, identical writes): no divergence
        // bookkeeping inside the dependent chain
        // The wraps are written as selects so that the dependent chain holds no branches: the reference's while loops
        // run at most once as long as a step stays under 2*pi.  Whether that held is collected in a flag; a chunk where
        // it did not (never seen with the loop's gains) is walked again from its start with the loops as written.
        const float TWO_PI = __fsub_rn(PI, -PI);
        const float phase0 = phase, freq0 = freq;
        bool bad = false;
        auto step = [&](float in) {
            const float out = phase;
            float err = __fsub_rn(in, phase);
            const float eHi = __fsub_rn(err, 2.0f * PI), eLo = __fadd_rn(err, 2.0f * PI);
            err = (err > PI) ? eHi : ((err <= -PI) ? eLo : err);
            freq = __fadd_rn(freq, __fmul_rn(j.beta, err));
            freq = fminf(fmaxf(freq, j.minFreq), j.maxFreq);
            const float p0 = __fadd_rn(phase, __fadd_rn(freq, __fmul_rn(j.alpha, err)));
            float p1 = (p0 > PI) ? __fsub_rn(p0, TWO_PI) : p0;
            p1 = (p1 < -PI) ? __fadd_rn(p1, TWO_PI) : p1;
            bad = bad || (fabsf(p1) > PI);
            phase = p1;
            return out;
        };
        int i = 0;
        // eight samples per trip: inputs read ahead of the chain (two 16-byte reads), phases written behind it
        float4 xa, xb;
        if (m >= 8) { xa = *reinterpret_cast<const float4*>(&sIn[0]); xb = *reinterpret_cast<const float4*>(&sIn[4]); }
#pragma unroll 1
        for (; i + 8 <= m; i += 8) {
            const float4 ca = xa, cb = xb;
            if (i + 16 <= m) { xa = *reinterpret_cast<const float4*>(&sIn[i + 8]); xb = *reinterpret_cast<const float4*>(&sIn[i + 12]); }
            float4 oa, ob;
            oa.x = step(ca.x); oa.y = step(ca.y); oa.z = step(ca.z); oa.w = step(ca.w);
            ob.x = step(cb.x); ob.y = step(cb.y); ob.z = step(cb.z); ob.w = step(cb.w);
            *reinterpret_cast<float4*>(&sOut[i]) = oa;
            *reinterpret_cast<float4*>(&sOut[i + 4]) = ob;
        }
        for (; i < m; i++) { sOut[i] = step(sIn[i]); }
        if (bad) {
            phase = phase0; freq = freq0;
            for (int i = 0; i < m; i++) {
                sOut[i] = phase;
                float err = __fsub_rn(sIn[i], phase);
                if (err > PI) { err = __fsub_rn(err, 2.0f * PI); }
                else if (err <= -PI) { err = __fadd_rn(err, 2.0f * PI); }
                freq = __fadd_rn(freq, __fmul_rn(j.beta, err));
                freq = fminf(fmaxf(freq, j.minFreq), j.maxFreq);
                phase = __fadd_rn(phase, __fadd_rn(freq, __fmul_rn(j.alpha, err)));
                while (phase > PI) { phase = __fsub_rn(phase, TWO_PI); }
                while (phase < -PI) { phase = __fadd_rn(phase, TWO_PI); }
            }
        }
        __syncwarp();
        for (int i = lane; i < m; i += 32) { j.phaseOut[base + i] = sOut[i]; }
        __syncwarp();
    }
    if (lane == 0) { j.st->pllPhase = phase; j.st->pllFreq = freq; }
}

// ---- the pilot PLL of the default (B200_ROTATOR_FAST) banks: the same loop walked in its ERROR form
// With x[i] the pilot's phase, e[i] = wrap(x[i] - phase[i]) and dx[i] = wrap(x[i+1] - x[i]) (no recurrence: all lanes), the
// reference's two updates  freq += beta * e;  phase += freq + alpha * e  become
//     freq[i] = clamp(freq[i-1] + beta * e[i])
//     e[i+1]  = wrap(dx[i] + (1 - alpha) * e[i] - freq[i])
// and phase[i] = wrap(x[i] - e[i]) is recovered afterwards, in parallel.  In exact arithmetic that IS the reference's
// recurrence; in fp32 a step rounds differently by about an ulp of the error, and the loop contracts such differences by
// (1 - alpha) per sample instead of accumulating them: the VCO phases agree with the reference's to ~1e-6 rad (tests: the
// decoded channels to -80 dB, the bar of the exact walk).  What it buys is the length of the dependent chain, which is all
// a sequential walk costs (fourteen dependent instructions per sample in the reference's order, 135 cycles):
//   * LINEAR walk, for a loop in lock: while no step leaves (-pi, pi] and the frequency keeps clear of its limits, neither
//     the wrap nor the clamp does anything, and with d[i] = dx[i] - freq[i-1] formed beside the chain (freq runs one step
//     behind e) the chain is ONE instruction:  e[i+1] = fma(1 - alpha - beta, e[i], d[i]).  The largest step and the range
//     of the frequency are watched beside the chain and judged after the chunk;
//   * FULL walk, clamp and wrap in the chain (seven dependent plain FP32 instructions), for the chunks the linear walk was not valid
//     for (they are walked again from the chunk's start) and the chunks after them until one would have been.
// A chunk that defeats even the full walk (a step beyond 3 pi: not with the loop's gains) is walked in the reference's
// order.  Every chunk starts from (phase, freq) and ends by advancing them as the reference does, so the carried state
// means the same in both kernels.  NaN pilot samples: the reference's phase is NaN from the sample after the first one on,
// for good; so it is here.
// One CTA per VFO, four warps: warp 0 walks chunk c while warps 1-3 turn chunk c + 1 into phases and increments
// (atan2f: 18 cycles per sample when the walking warp does it itself) and write chunk c - 1's phases out.
constexpr int kPeThreads = 128;
__device__ __forceinline__ float pllWrap(float v) {
    const float PI = 3.1415926535f;
    return (v > PI) ? v - 2.0f * PI : ((v <= -PI) ? v + 2.0f * PI : v);
}
// the loop's own phase clamp (phase_control_loop.h clampPhase: limits -PI .. PI, both inclusive)
__device__ __forceinline__ float pllClampPhase(float p) {
    const float PI = 3.1415926535f;
    return (p > PI) ? p - 2.0f * PI : ((p < -PI) ? p + 2.0f * PI : p);
}
struct PllErrWalk {
    float c1, k, beta, lo, hi;             // 1 - alpha, 1 - alpha - beta
    float e, f, eLast, fLo, fHi, rMax;     // e[m], frequency after the chunk, e[m-1]; extremes of the unclamped frequency and of |step|
    template <bool LINEAR>
    __device__ __forceinline__ void run(const float* __restrict__ DX, float* __restrict__ E, int m, float e0, float f0) {
        e = e0; f = f0; eLast = e0; fLo = f0; fHi = f0; rMax = 0.0f;
        auto step = [&](float dx) {
            const float cur = e;
            if (LINEAR) {
                const float d = dx - f;                 // beside the chain: f is a step behind e
                f = fmaf(beta, e, f);
                fLo = fminf(fLo, f); fHi = fmaxf(fHi, f);
                e = fmaf(k, e, d);
                rMax = fmaxf(rMax, fabsf(e));
            }
            else {
                const float g = fmaf(c1, e, dx);
                float fn = fmaf(beta, e, f);
                fLo = fminf(fLo, fn); fHi = fmaxf(fHi, fn);
                fn = fminf(fmaxf(fn, lo), hi);
                f = fn;
                const float r = g - fn;
                rMax = fmaxf(rMax, fabsf(r));
                // wrap = r - 2 pi rint(r / 2 pi), the rounding by the magic-number addition: three plain FP32 instructions in the
                // chain where compare + select cost a predicate round trip (41 -> 30 cycles per sample for the walk alone).  At
                // the very edge (|r| within an ulp of pi) e may stay a hair outside (-pi, pi]; the loop does not care.
                const float k = fmaf(r, 0.15915494309189535f, 12582912.0f) - 12582912.0f;
                e = fmaf(k, -2.0f * 3.1415926535f, r);
            }
            return cur;
        };
        int i = 0;
        // eight samples per trip: increments read ahead of the chain (two 16-byte reads), errors written behind it
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

__global__ void __launch_bounds__(kPeThreads) pllErrKernel(const PllJob* __restrict__ jobs) {
    const PllJob j = jobs[blockIdx.x];
    __shared__ __align__(16) float sX[2][kPllChunk], sD[2][kPllChunk + 4], sE[2][kPllChunk];
    __shared__ int sNanAt[2];        // first NaN pilot sample of the chunk in this buffer (helpers), INT_MAX if none
    __shared__ int sFrom[2];         // what the helpers write for a walked chunk: phases up to this sample, NaN behind it; -1 all NaN; -2 nothing (already written)
    const int tid = threadIdx.x, lane = tid & 31;
    const bool walker = tid < 32;
    const int ht = tid - 32, HT = kPeThreads - 32;
    const float PI = 3.1415926535f;
    const int nChunks = (j.n + kPllChunk - 1) / kPllChunk;
    auto helperBarrier = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(kPeThreads - 32) : "memory"); };
    // phases and increments of chunk c into buffer c & 1 (helpers)
    auto prepare = [&](int c) {
        const int base = c * kPllChunk, m = min(kPllChunk, j.n - base), b = c & 1;
        for (int i = ht; i < m; i += HT) {
            float x = atan2f(__ldg(j.pi + base + i), __ldg(j.pr + base + i));
            if (!(x == x)) { atomicMin(&sNanAt[b], i); x = 0.0f; }
            sX[b][i] = x;
        }
        helperBarrier();
        for (int i = ht; i < ((m + 3) & ~3); i += HT) { sD[b][i] = (i + 1 < m) ? pllWrap(sX[b][i + 1] - sX[b][i]) : 0.0f; }
    };
    if (tid < 2) { sNanAt[tid] = 0x7fffffff; }
    __syncthreads();
    if (!walker && nChunks > 0) { prepare(0); }
    __syncthreads();
    float phase = 0.f, freq = 0.f;
    bool dead = false, full = false;
    PllErrWalk w;
    if (walker) {
        phase = j.st->pllPhase; freq = j.st->pllFreq;
        dead = !(phase == phase && freq == freq);
        w.c1 = 1.0f - j.alpha; w.k = 1.0f - j.alpha - j.beta; w.beta = j.beta; w.lo = j.minFreq; w.hi = j.maxFreq;
        full = !(freq > j.minFreq && freq < j.maxFreq);
    }
    for (int c = 0; c <= nChunks; c++) {
        const int b = c & 1;
        if (walker) {
            if (c < nChunks) {
                const int base = c * kPllChunk, m = min(kPllChunk, j.n - base);
                const int nanAt = sNanAt[b];
                __syncwarp();
                if (lane == 0) { sNanAt[b] = 0x7fffffff; }      // for the chunk after next, which the helpers prepare into this buffer
                int from = dead ? -1 : min(nanAt, m);
                if (!dead) {
                    const float* X = sX[b];
                    const float e0 = pllWrap(X[0] - phase);
                    bool ok = false;
                    if (!full) {
                        w.run<true>(sD[b], sE[b], m, e0, freq);
                        // no step left (-pi, pi) and the frequency stayed inside its limits: wrap and clamp would have done nothing
                        ok = w.rMax < PI && fabsf(e0) < PI && w.fLo >= j.minFreq && w.fHi <= j.maxFreq && w.e == w.e && w.f == w.f;
                    }
                    if (!ok) {
                        w.run<false>(sD[b], sE[b], m, e0, freq);
                        ok = w.rMax < 3.0f * PI && w.e == w.e && w.f == w.f;
                        // back to the linear walk once a chunk would have been valid for it
                        full = !(w.rMax < PI && w.fLo > j.minFreq && w.fHi < j.maxFreq);
                    }
                    if (ok) {
                        const float pLast = pllClampPhase(X[m - 1] - w.eLast);
                        phase = pllClampPhase(pLast + (w.f + j.alpha * w.eLast));
                        freq = w.f;
                    }
                    else {
                        // the loop as the reference writes it (loop/pll.h:64-70, phase_control_loop.h:58-66), phases straight out
                        const float TWO_PI = __fsub_rn(PI, -PI);
                        for (int i = 0; i < m; i++) {
                            if (lane == 0) { j.phaseOut[base + i] = (i > nanAt) ? __int_as_float(0x7fc00000) : phase; }
                            float err = __fsub_rn(X[i], phase);
                            if (err > PI) { err = __fsub_rn(err, 2.0f * PI); }
                            else if (err <= -PI) { err = __fadd_rn(err, 2.0f * PI); }
                            freq = __fadd_rn(freq, __fmul_rn(j.beta, err));
                            freq = fminf(fmaxf(freq, j.minFreq), j.maxFreq);
                            phase = __fadd_rn(phase, __fadd_rn(freq, __fmul_rn(j.alpha, err)));
                            while (phase > PI) { phase = __fsub_rn(phase, TWO_PI); }
                            while (phase < -PI) { phase = __fadd_rn(phase, TWO_PI); }
                        }
                        from = -2;
                        full = true;
                    }
                    if (nanAt < m) { dead = true; }
                }
                if (lane == 0) { sFrom[b] = from; }
            }
        }
        else {
            if (c >= 1) {
                // phases of chunk c - 1: phase[i] = x[i] - e[i], brought into the loop's phase limits
                const int pb = b ^ 1, base = (c - 1) * kPllChunk, m = min(kPllChunk, j.n - base);
                const int from = sFrom[pb];
                if (from != -2) {
                    for (int i = ht; i < m; i += HT) {
                        const float p = pllClampPhase(sX[pb][i] - sE[pb][i]);
                        j.phaseOut[base + i] = (from == -1 || i > from) ? __int_as_float(0x7fc00000) : p;
                    }
                }
            }
            helperBarrier();
            if (c + 1 < nChunks) { prepare(c + 1); }
            else { helperBarrier(); }
        }
        __syncthreads();
    }
    if (tid == 0) {
        j.st->pllPhase = dead ? __int_as_float(0x7fc00000) : phase;
        j.st->pllFreq = dead ? __int_as_float(0x7fc00000) : freq;
    }
}

// broadcast_fm.h:157-179: the MPX delayed by the pilot filter's group delay is mixed down by the squared conjugate VCO
// (the 19 kHz pilot doubled to the 38 kHz subcarrier), L-R = 2 * Re(.), L = (L+R) + (L-R), R = (L+R) - (L-R).
// Also refreshes the delay line: histNext[i] = [hist | mpx][n + i].
__global__ void __launch_bounds__(256) wfmMixKernel(const WfmMixJob* __restrict__ jobs) {
    const WfmMixJob j = jobs[blockIdx.y];
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < j.delay) {
        const int p = j.n + i;                        // position in [hist (delay) | mpx (n)]
        j.histNext[i] = (p < j.delay) ? j.hist[p] : j.mpx[p - j.delay];
    }
    if (i >= j.n) { return; }
    const float x = (i < j.delay) ? j.hist[i] : j.mpx[i - j.delay];
    float sn, cs;
    sincosf(j.phase[i], &sn, &cs);
    const float cr = cs, ci = -sn;                    // conjugate of the VCO phasor
    // (x, 0) * c * c with the reference's complex product (re*re - im*im, re*im + im*re)
    const float a1r = __fsub_rn(__fmul_rn(x, cr), __fmul_rn(0.0f, ci)), a1i = __fadd_rn(__fmul_rn(x, ci), __fmul_rn(0.0f, cr));
    const float a2r = __fsub_rn(__fmul_rn(a1r, cr), __fmul_rn(a1i, ci));
    const float lmr = __fmul_rn(a2r, 2.0f);
    j.l[i] = __fadd_rn(x, lmr);
    j.r[i] = __fsub_rn(x, lmr);
}

// LRToStereo (convert/l_r_to_stereo.h)
__global__ void __launch_bounds__(256) lrKernel(const LrJob* __restrict__ jobs) {
    const LrJob j = jobs[blockIdx.y];
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= j.n) { return; }
    reinterpret_cast<float2*>(j.stereo)[i] = make_float2(__ldg(j.l + i), __ldg(j.r + i));
}

// radio IF chain, one CTA per VFO.
// NoiseBlanker (noise_blanker.h:38-56): amp follows the input amplitude as a first-order recurrence that only advances
// on non-zero samples; a sample whose amplitude exceeds level x amp is scaled down to amp.  The recurrence is linear,
// so every thread first composes the affine map of its own contiguous segment, the CTA scans the 256 maps to find
// each segment's starting amp, and the segments are then walked in the reference's own arithmetic.
// Squelch (squelch.h:55-69): the whole block passes when 10*log10(mean |x|) >= level, else it is zeroed.
__global__ void __launch_bounds__(256) ifChainKernel(const IfJob* __restrict__ jobs) {
    const IfJob j = jobs[blockIdx.x];
    __shared__ float sA[256], sB[256];
    __shared__ double sSum[256];
    const int tid = threadIdx.x;
    const int seg = (j.n + 255) / 256;
    const int i0 = min(j.n, tid * seg), i1 = min(j.n, i0 + seg);
    const float2* in = reinterpret_cast<const float2*>(j.in);
    float2* out = reinterpret_cast<float2*>(j.out);
    if (j.nbOn) {
        float A = 1.0f, B = 0.0f;
        for (int i = i0; i < i1; i++) {
            const float2 x = __ldg(in + i);
            const float a = sqrtf(__fadd_rn(__fmul_rn(x.x, x.x), __fmul_rn(x.y, x.y)));
            if (a != 0.0f) { A *= j.invRate; B = fmaf(B, j.invRate, a * j.rate); }
        }
        sA[tid] = A; sB[tid] = B;
        __syncthreads();
        // inclusive scan of the maps (apply the earlier map first): (A2,B2) o (A1,B1) = (A1*A2, B1*A2 + B2)
        for (int o = 1; o < 256; o <<= 1) {
            float pa = 1.0f, pb = 0.0f;
            if (tid >= o) { pa = sA[tid - o]; pb = sB[tid - o]; }
            __syncthreads();
            if (tid >= o) { const float a2 = sA[tid], b2 = sB[tid]; sA[tid] = pa * a2; sB[tid] = fmaf(pb, a2, b2); }
            __syncthreads();
        }
        const float amp0 = j.st->nbAmp;
        float amp = tid == 0 ? amp0 : fmaf(amp0, sA[tid - 1], sB[tid - 1]);
        for (int i = i0; i < i1; i++) {
            const float2 x = __ldg(in + i);
            const float a = sqrtf(__fadd_rn(__fmul_rn(x.x, x.x), __fmul_rn(x.y, x.y)));
            float gain = 1.0f;
            if (a != 0.0f) {
                amp = __fadd_rn(__fmul_rn(amp, j.invRate), __fmul_rn(a, j.rate));
                const float excess = __fdiv_rn(a, amp);
                if (excess > j.nbLevel) { gain = __fdiv_rn(1.0f, excess); }
            }
            out[i] = make_float2(__fmul_rn(x.x, gain), __fmul_rn(x.y, gain));
        }
        __syncthreads();
        if (tid == 255 && j.n > 0) { j.st->nbAmp = fmaf(amp0, sA[255], sB[255]); }
        in = out;                                    // the squelch looks at the blanked block
        __syncthreads();
    }
    if (j.sqOn) {
        double acc = 0.0;
        for (int i = i0; i < i1; i++) {
            const float2 x = in[i];
            acc += (double)sqrtf(__fadd_rn(__fmul_rn(x.x, x.x), __fmul_rn(x.y, x.y)));
        }
        sSum[tid] = acc;
        __syncthreads();
        for (int o = 128; o > 0; o >>= 1) { if (tid < o) { sSum[tid] += sSum[tid + o]; } __syncthreads(); }
        const float mean = (float)(sSum[0] / (double)max(j.n, 1));
        const bool open = (10.0f * log10f(mean)) >= j.sqLevel;
        for (int i = i0; i < i1; i++) { out[i] = open ? in[i] : make_float2(0.f, 0.f); }
    }
    else if (!j.nbOn) {
        for (int i = i0; i < i1; i++) { out[i] = in[i]; }
    }
}

// Deemphasis<stereo_t> (filter/deephasis.h:58-76): out[i] = alpha*in[i] + (1-alpha)*out[i-1], in place, state carried
// between calls.  A first-order recurrence whose fp32 rounding depends on the order, so it is walked in order: the CTA
// stages a chunk and forms alpha*in[i] for every sample, two threads (left, right) walk FMUL -> FADD, the CTA stores.
constexpr int kDeChunk = 1024;
__global__ void __launch_bounds__(64) deemphKernel(const DeemphJob* __restrict__ jobs) {
    const DeemphJob j = jobs[blockIdx.x];
    __shared__ float sv[2][kDeChunk + 1];
    const float beta = __fsub_rn(1.0f, j.alpha);
    float prev = 0.f;
    if (threadIdx.x < 2) { prev = threadIdx.x == 0 ? j.st->deLast.x : j.st->deLast.y; }
    for (int base = 0; base < j.n; base += kDeChunk) {
        const int m = min(kDeChunk, j.n - base);
        for (int i = threadIdx.x; i < m; i += 64) {
            const float2 v = j.data[base + i];
            sv[0][i] = __fmul_rn(j.alpha, v.x);
            sv[1][i] = __fmul_rn(j.alpha, v.y);
        }
        __syncthreads();
        if (threadIdx.x < 2) {
            float* row = sv[threadIdx.x];
#pragma unroll 8
            for (int i = 0; i < m; i++) {
                prev = __fadd_rn(row[i], __fmul_rn(beta, prev));
                row[i] = prev;
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < m; i += 64) { j.data[base + i] = make_float2(sv[0][i], sv[1][i]); }
        __syncthreads();
    }
    if (threadIdx.x == 0 && j.n > 0) { j.st->deLast.x = prev; }
    if (threadIdx.x == 1 && j.n > 0) { j.st->deLast.y = prev; }
}

// FM IF noise reduction (noise_reduction/fm_if.h:45-80).  For every sample the reference windows the newest `bins`
// samples, runs a forward FFT, keeps the strongest bin alone, runs the backward FFT and takes element bins/2: that is
//     out[i] = X_i[k*] * exp(+2 pi i k* (bins/2) / bins),   X_i[k] = sum_n win[n] x[i+n] exp(-2 pi i k n / bins),
// k* the first of the strongest bins.  Windows of neighbouring samples are independent, so one warp takes a sample at a
// time with lane k holding bin k: the lane's column of the (window x twiddle) matrix stays in registers, the samples
// come out of shared memory as broadcasts, the strongest bin is found with a warp reduction (ties to the lower index,
// NaN handled like VOLK's generic index_max: only a NaN in bin 0 wins).  bins <= 32 (the reference's presets are 9,
// 15, 31 and 32; radio/src/main.cpp:11-16).
constexpr int kFmTile = 512, kFmThreads = 128, kFmMaxBins = 32;
__global__ void __launch_bounds__(kFmThreads) fmifKernel(const FmifJob* __restrict__ jobs) {
    const FmifJob j = jobs[blockIdx.y];
    const int base = blockIdx.x * kFmTile;
    if (base >= j.n) { return; }
    __shared__ float2 sx[kFmTile + kFmMaxBins];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int bins = j.bins, H = bins - 1;
    const int m = min(kFmTile, j.n - base);
    const float2* in = reinterpret_cast<const float2*>(j.in);
    const float2* hist = reinterpret_cast<const float2*>(j.hist);
    for (int i = tid; i < m + H; i += kFmThreads) {
        const int g = base + i - H;                        // position in this block's input; negative = carried history
        sx[i] = g >= 0 ? __ldg(in + g) : __ldg(hist + H + g);
    }
    if (blockIdx.x == 0) {
        // the samples in front of the next block (fm_if.h:72): the last bins-1 of [history | block]
        float2* next = reinterpret_cast<float2*>(j.histNext);
        for (int t = tid; t < H; t += kFmThreads) {
            const int g = j.n + t - H;
            next[t] = g >= 0 ? __ldg(in + g) : __ldg(hist + H + g);
        }
    }
    float2 t[kFmMaxBins];
#pragma unroll
    for (int n = 0; n < kFmMaxBins; n++) { t[n] = (n < bins && lane < bins) ? __ldg(j.tab + n * kFmMaxBins + lane) : make_float2(0.f, 0.f); }
    const float2 back = lane < bins ? __ldg(j.tab + kFmMaxBins * kFmMaxBins + lane) : make_float2(0.f, 0.f);
    __syncthreads();
    float2* out = reinterpret_cast<float2*>(j.out);
    for (int g0 = warp * 32; g0 < m; g0 += (kFmThreads / 32) * 32) {
        const int cnt = min(32, m - g0);
        float2 res = make_float2(0.f, 0.f);
        for (int q = 0; q < cnt; q++) {
            const float2* x = sx + g0 + q;
            float re0 = 0.f, im0 = 0.f, re1 = 0.f, im1 = 0.f;
#pragma unroll
            for (int n = 0; n < kFmMaxBins; n++) {
                if (n < bins) {
                    const float2 v = x[n];
                    re0 = fmaf(v.x, t[n].x, re0); re1 = fmaf(-v.y, t[n].y, re1);
                    im0 = fmaf(v.x, t[n].y, im0); im1 = fmaf(v.y, t[n].x, im1);
                }
            }
            const float re = re0 + re1, im = im0 + im1;
            float mag = sqrtf(__fadd_rn(__fmul_rn(re, re), __fmul_rn(im, im)));
            const bool nan0 = __shfl_sync(0xffffffffu, mag != mag ? 1 : 0, 0) != 0;
            if (lane >= bins || mag != mag) { mag = -1.0f; }
            int idx = lane;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float om = __shfl_xor_sync(0xffffffffu, mag, o);
                const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
                if (om > mag || (om == mag && oi < idx)) { mag = om; idx = oi; }
            }
            if (nan0) { idx = 0; }
            const float xr = __shfl_sync(0xffffffffu, re, idx), xi = __shfl_sync(0xffffffffu, im, idx);
            const float br = __shfl_sync(0xffffffffu, back.x, idx), bi = __shfl_sync(0xffffffffu, back.y, idx);
            if (lane == q) { res = make_float2(xr * br - xi * bi, xr * bi + xi * br); }
        }
        if (lane < cnt) { out[base + g0 + lane] = res; }
    }
}

// audio::Volume (audio/volume.h:47-50): every float of the stereo block times volume^2 (0 when muted), in place
__global__ void __launch_bounds__(256) volumeKernel(const VolJob* __restrict__ jobs) {
    const VolJob j = jobs[blockIdx.y];
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < j.nFloats) { j.data[i] = __fmul_rn(j.data[i], j.gain); }
}

// RDS branch, first half (broadcast_fm.h:171,200-203): RealToComplex of the MPX, then FrequencyXlator by -57 kHz.  With a
// zero imaginary part the rotator's complex product reduces to (x * phase.re, x * phase.im).
__global__ void __launch_bounds__(256) rdsXlateKernel(const RdsXlateJob* __restrict__ jobs) {
    const RdsXlateJob j = jobs[blockIdx.y];
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= j.n) { return; }
    const float2 ph = rotPhaseAt(j.chk, j.wtab, i);
    const float x = __ldg(j.mpx + i);
    reinterpret_cast<float2*>(j.out)[i] = make_float2(__fmul_rn(x, ph.x), __fmul_rn(x, ph.y));
}

} // namespace

void demodDestroy(DemodState* d) {
    if (!d) { return; }
    d->fir.destroy();
    for (auto& st : d->afDecim) { st.destroy(); }
    d->afPoly.destroy();
    if (d->dState) { cudaFree(d->dState); }
    if (d->dWtab) { cudaFree(d->dWtab); }
    if (d->dMono) { cudaFree(d->dMono); }
    if (d->dMaps) { cudaFree(d->dMaps); }
    if (d->dAmps) { cudaFree(d->dAmps); }
    if (d->dStereo) { cudaFree(d->dStereo); }
    if (d->dIf) { cudaFree(d->dIf); }
    d->pilotRe.destroy(); d->pilotIm.destroy(); d->firR.destroy();
    for (int i = 0; i < 2; i++) { if (d->dDelayHist[i]) { cudaFree(d->dDelayHist[i]); } }
    if (d->dPllPhase) { cudaFree(d->dPllPhase); }
    if (d->dL) { cudaFree(d->dL); }
    if (d->dR) { cudaFree(d->dR); }
    if (d->dFmTab) { cudaFree(d->dFmTab); }
    for (int i = 0; i < 2; i++) { if (d->dFmHist[i]) { cudaFree(d->dFmHist[i]); } }
    if (d->dFm) { cudaFree(d->dFm); }
    for (auto& st : d->rdsDecim) { st.destroy(); }
    d->rdsPoly.destroy();
    for (void* q : { (void*)d->dRdsWtab, (void*)d->dRdsChk, (void*)d->dRdsPhase, (void*)d->dRdsX }) { if (q) { cudaFree(q); } }
    delete d;
}

int demodSwitchRotatorMode(DemodState* d, int newMode) {
    if (d && d->rdsOn && d->dRdsPhase) {
        // the RDS branch's -57 kHz rotator: fp32 phase on the device in EXACT mode, a float64 angle on the host in FAST mode
        if (newMode == B200_ROTATOR_EXACT) {
            const float2 p = make_float2((float)cos(d->rdsAngle), (float)sin(d->rdsAngle));
            B200_CUDA(cudaMemcpy(d->dRdsPhase, &p, sizeof(float2), cudaMemcpyHostToDevice));
        }
        else {
            float2 p;
            B200_CUDA(cudaMemcpy(&p, d->dRdsPhase, sizeof(float2), cudaMemcpyDeviceToHost));
            d->rdsAngle = atan2((double)p.y, (double)p.x);
        }
    }
    if (!d || !d->dWtab) { return B200_OK; }
    float2* dPhase = &d->dState->rotPhase;
    int* dEnv = &d->dState->audio.startEnvelope;
    if (newMode == B200_ROTATOR_EXACT) {
        const float2 p = make_float2((float)cos(d->angle), (float)sin(d->angle));
        B200_CUDA(cudaMemcpy(dPhase, &p, sizeof(float2), cudaMemcpyHostToDevice));
        const int env = (int)std::min<long long>(d->envCount, 1 << 30);
        B200_CUDA(cudaMemcpy(dEnv, &env, sizeof(int), cudaMemcpyHostToDevice));
    }
    else {
        float2 p; int env = 0;
        B200_CUDA(cudaMemcpy(&p, dPhase, sizeof(float2), cudaMemcpyDeviceToHost));
        B200_CUDA(cudaMemcpy(&env, dEnv, sizeof(int), cudaMemcpyDeviceToHost));
        d->angle = atan2((double)p.y, (double)p.x);
        d->envCount = env;
    }
    return B200_OK;
}

static int demodEnsureCap(DemodState& d, int n) {
    if (n <= d.cap) { return B200_OK; }
    if (d.dMono) { cudaFree(d.dMono); d.dMono = nullptr; }
    if (d.dStereo) { cudaFree(d.dStereo); d.dStereo = nullptr; }
    if (d.dMaps) { cudaFree(d.dMaps); d.dMaps = nullptr; }
    if (d.dAmps) { cudaFree(d.dAmps); d.dAmps = nullptr; }
    d.cap = n + 64;
    if (d.dWtab) {
        B200_CUDA(cudaMalloc((void**)&d.dMaps, (size_t)d.cap * sizeof(float2)));
        B200_CUDA(cudaMalloc((void**)&d.dAmps, (size_t)d.cap * sizeof(float)));
    }
    B200_CUDA(cudaMalloc((void**)&d.dMono, (size_t)d.cap * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&d.dStereo, (size_t)d.cap * 2 * sizeof(float)));
    if (d.hasFir) { B200_TRY(d.fir.ensureOutCap(d.cap)); }
    if (d.wfmStereo) {
        for (float** p : { &d.dPllPhase, &d.dL, &d.dR }) {
            if (*p) { cudaFree(*p); *p = nullptr; }
            B200_CUDA(cudaMalloc((void**)p, (size_t)d.cap * sizeof(float)));
        }
        B200_TRY(d.pilotRe.ensureOutCap(d.cap));
        B200_TRY(d.pilotIm.ensureOutCap(d.cap));
        if (d.hasFir) { B200_TRY(d.firR.ensureOutCap(d.cap)); }
    }
    if (d.rdsOn && d.rdsCap < d.cap) {
        if (d.dRdsX) { cudaFree(d.dRdsX); d.dRdsX = nullptr; }
        if (d.dRdsChk) { cudaFree(d.dRdsChk); d.dRdsChk = nullptr; }
        d.rdsCap = d.cap;
        B200_CUDA(cudaMalloc((void**)&d.dRdsX, (size_t)d.rdsCap * sizeof(cf32)));
        B200_CUDA(cudaMalloc((void**)&d.dRdsChk, (size_t)(d.rdsCap / kRotChunk + 2) * sizeof(float2)));
    }
    return B200_OK;
}

// dsp::multirate::RationalResampler<T> for a two-float element (complex_t or stereo_t; rational_resampler.h:128-173) as
// a cascade of stages: the power-of-two decimator's plan, then the polyphase resampler
static int buildResampler(double inSR, double outSR, int cap, std::vector<Stage>& decim, Stage& poly, bool& hasPoly, int& mode) {
    for (auto& st : decim) { st.destroy(); }
    decim.clear();
    poly.destroy(); hasPoly = false;
    design::ResampPlan p = design::planResampler(inSR, outSR);
    mode = p.mode;
    if (p.mode == 0 || p.mode == 1) {
        const design::PlanDef* plan = design::findPlan(p.ratio);
        if (!plan) { setError("no decimation plan for ratio %d", p.ratio); return B200_EINVAL; }
        decim.resize(plan->nstages);
        for (int i = 0; i < plan->nstages; i++) {
            std::vector<float> t(plan->stages[i].ntaps);
            memcpy(t.data(), plan->stages[i].taps, sizeof(float) * t.size());   // the table holds IEEE bit patterns
            cap = cap / plan->stages[i].decim + 2;
            B200_TRY(decim[i].create(JOB_DECIM_FIR, t, plan->stages[i].decim, 1, 1, 2, cap, true));
        }
    }
    if (p.mode == 0 || p.mode == 2) {
        const int pcap = (int)(((int64_t)cap * p.interp) / p.decim) + 4;
        B200_TRY(poly.create(JOB_POLYPHASE, p.taps, 1, p.interp, p.decim, 2, pcap));
        hasPoly = true;
    }
    return B200_OK;
}

// builds the demodulator exactly as the radio module's demodulators/{nfm,am,usb,lsb,dsb}.h init() do
static int demodConfigure(VfoBank& bank, Vfo& v, int kind, double p0, double p1, double p2, double p3) {
    demodDestroy(v.demod);
    v.demod = nullptr;
    if (kind == B200_DEMOD_NONE) { return B200_OK; }
    auto* d = new (std::nothrow) DemodState;
    if (!d) { return B200_ENOMEM; }
    v.demod = d;
    d->kind = kind; d->p0 = p0; d->p1 = p1; d->p2 = p2; d->p3 = p3;
    d->sr = v.outSR; d->bw = v.bw;
    SeqState init{};
    init.rotPhase = make_float2(1.0f, 0.0f);
    init.nbAmp = 1.0f;
    auto agcInit = [&](double attack, double decay) {
        // agc.init(NULL, 1.0, attack, decay, 10e6, 10.0, INFINITY) (ssb.h:29, am.h:34-35): amp = setPoint / initGain = 0
        d->agc.setPoint = 1.0f; d->agc.attack = (float)attack; d->agc.invAttack = 1.0f - d->agc.attack;
        d->agc.decay = (float)decay; d->agc.invDecay = 1.0f - d->agc.decay; d->agc.maxGain = (float)10e6;
        init.carrier.amp = init.audio.amp = 1.0f / INFINITY;
        init.carrier.startEnvelope = init.audio.startEnvelope = 0;
    };
    std::vector<float> firTaps;
    if (kind == B200_DEMOD_NFM) {
        d->invDeviation = (float)(1.0 / design::hzToRads(d->bw / 2.0, d->sr));   // quadrature.h:19-26 via fm.h:31
        const bool lp = p0 != 0.0, hp = p1 != 0.0;
        d->filtering = lp || hp;
        if (lp && hp) { firTaps = design::bandPass(300.0, d->bw / 2.0, 100.0, d->sr); }       // fm.h:127-129
        else if (hp) { firTaps = design::highPass(300.0, 100.0, d->sr); }                       // fm.h:130-132
        else if (lp) { firTaps = design::lowPass(d->bw / 2.0, (d->bw / 2.0) * 0.1, d->sr); }   // fm.h:133-135
    }
    else if (kind == B200_DEMOD_WFM) {
        // demod.init(input, bandwidth / 2, 500 kS/s, stereo, lowPass, rds) (radio/src/demodulators/wfm.h:76)
        d->rdsOn = p2 != 0.0;
        d->invDeviation = (float)(1.0 / design::hzToRads(d->bw / 2.0, d->sr));
        d->wfmStereo = p0 != 0.0;
        d->filtering = p1 != 0.0;
        if (d->filtering) { firTaps = design::lowPass(15000.0, 4000.0, d->sr); }              // broadcast_fm.h:49
        if (d->wfmStereo) {
            std::vector<float> tr, ti;
            design::bandPassComplex(18750.0, 19250.0, 3000.0, d->sr, true, tr, ti);             // broadcast_fm.h:43
            if (tr.empty()) { setError("pilot filter has no taps at %g S/s", d->sr); v.demod = nullptr; delete d; return B200_EINVAL; }
            B200_TRY(d->pilotRe.create(JOB_DECIM_FIR, tr, 1, 1, 1, 1, 64));
            B200_TRY(d->pilotIm.create(JOB_DECIM_FIR, ti, 1, 1, 1, 1, 64));
            d->wfmDelay = (((int)tr.size() - 1) / 2) + 1;                                        // broadcast_fm.h:47-48
            for (int i = 0; i < 2; i++) {
                B200_CUDA(cudaMalloc((void**)&d->dDelayHist[i], (size_t)d->wfmDelay * sizeof(float)));
                B200_CUDA(cudaMemset(d->dDelayHist[i], 0, (size_t)d->wfmDelay * sizeof(float)));
            }
            // PLL::init(bandwidth = 25000 / fs, phase 0, freq 19 kHz, limits 18.75 .. 19.25 kHz) (broadcast_fm.h:46),
            // PhaseControlLoop<float>::criticallyDamped evaluated in float (phase_control_loop.h:30-35)
            const float bandwidth = (float)(25000.0 / d->sr);
            const float damp = sqrt(2.0) / 2.0;
            const float den = (1.0 + 2.0 * damp * bandwidth + bandwidth * bandwidth);
            d->pllAlpha = (4 * damp * bandwidth) / den;
            d->pllBeta = (4 * bandwidth * bandwidth) / den;
            d->pllMinFreq = (float)design::hzToRads(18750.0, d->sr);
            d->pllMaxFreq = (float)design::hzToRads(19250.0, d->sr);
            init.pllPhase = 0.0f;
            init.pllFreq = (float)design::hzToRads(19000.0, d->sr);
        }
    }
    else if (kind == B200_DEMOD_AM) {
        d->agcMode = (int)p0;
        agcInit(p1, p2);
        d->dcRate = (float)p3;
        firTaps = design::lowPass(d->bw / 2.0, (d->bw / 2.0) * 0.1, d->sr);   // am.h:37
        d->filtering = true;
    }
    else if (kind == B200_DEMOD_USB || kind == B200_DEMOD_LSB || kind == B200_DEMOD_DSB) {
        const double tr = (kind == B200_DEMOD_USB) ? d->bw / 2.0 : (kind == B200_DEMOD_LSB) ? -d->bw / 2.0 : 0.0;   // ssb.h:111-123
        const double w = design::hzToRads(tr, d->sr);
        d->dRe = (float)cos(w); d->dIm = (float)sin(w);
        d->theta = design::calibrateRotatorStep(d->dRe, d->dIm, bank.calSteps > 0 ? std::min(bank.calSteps, 1 << 18) : 0);
        if (bank.calSteps <= 0) { d->theta = atan2((double)d->dIm, (double)d->dRe); }
        d->angle = 0.0;
        agcInit(p1, p2);
        float2 tab[kRotChunk];
        makeRotatorTable(d->dRe, d->dIm, tab);
        B200_CUDA(cudaMalloc((void**)&d->dWtab, sizeof(tab)));
        B200_CUDA(cudaMemcpy(d->dWtab, tab, sizeof(tab), cudaMemcpyHostToDevice));
    }
    else {
        setError("unknown demodulator kind %d", kind);
        return B200_EINVAL;
    }
    if (!firTaps.empty()) {
        B200_TRY(d->fir.create(JOB_DECIM_FIR, firTaps, 1, 1, 1, 1, 64));
        if (d->wfmStereo) { B200_TRY(d->firR.create(JOB_DECIM_FIR, firTaps, 1, 1, 1, 1, 64)); }
        d->hasFir = true;
    }
    if (d->rdsOn) {
        // xlator.init(NULL, -57000.0, samplerate); rdsResamp.init(NULL, samplerate, 5000.0) (broadcast_fm.h:52-53)
        const double w = design::hzToRads(-57000.0, d->sr);
        d->rdsDRe = (float)cos(w); d->rdsDIm = (float)sin(w);
        d->rdsTheta = (bank.calSteps > 0) ? design::calibrateRotatorStep(d->rdsDRe, d->rdsDIm, std::min(bank.calSteps, 1 << 18))
                                          : atan2((double)d->rdsDIm, (double)d->rdsDRe);
        d->rdsAngle = 0.0;
        float2 tab[kRotChunk];
        makeRotatorTable(d->rdsDRe, d->rdsDIm, tab);
        B200_CUDA(cudaMalloc((void**)&d->dRdsWtab, sizeof(tab)));
        B200_CUDA(cudaMemcpy(d->dRdsWtab, tab, sizeof(tab), cudaMemcpyHostToDevice));
        const float2 one = make_float2(1.0f, 0.0f);
        B200_CUDA(cudaMalloc((void**)&d->dRdsPhase, sizeof(float2)));
        B200_CUDA(cudaMemcpy(d->dRdsPhase, &one, sizeof(float2), cudaMemcpyHostToDevice));
        const int cap0 = bank.maxOut((int)(&v - bank.vfos.data()), bank.maxBlock) + 8 + 64;
        B200_TRY(buildResampler(d->sr, 5000.0, cap0, d->rdsDecim, d->rdsPoly, d->rdsHasPoly, d->rdsMode));
    }
    B200_CUDA(cudaMalloc((void**)&d->dState, sizeof(SeqState)));
    B200_CUDA(cudaMemcpy(d->dState, &init, sizeof(SeqState), cudaMemcpyHostToDevice));
    return demodEnsureCap(*d, bank.maxOut((int)(&v - bank.vfos.data()), bank.maxBlock) + 8);
}

static int afMaxOut(const DemodState& d, int count) {
    int64_t n = count;
    if (!d.afResamp) { return (int)n; }
    if (d.afMode == 0 || d.afMode == 1) { for (auto& st : d.afDecim) { n = n / st.D + 2; } }
    if (d.afMode == 0 || d.afMode == 2) { n = (n * d.afPoly.L) / d.afPoly.M + 4; }
    return (int)n;
}

// afChain of the radio module: resamp.init(NULL, afRate, audioRate) (radio_module.h:101,757-758) and
// deemp.init(NULL, tau, audioRate) (:102).  audioSR <= 0 leaves the audio at the demodulator's rate, tau <= 0
// leaves the de-emphasis out (its default state in the reference).
static int afConfigure(VfoBank& bank, Vfo& v, double audioSR, double tau) {
    if (!v.demod) { setError("audio post-processing needs a demodulator on the VFO first"); return B200_EINVAL; }
    DemodState& d = *v.demod;
    for (auto& st : d.afDecim) { st.destroy(); }
    d.afDecim.clear();
    d.afPoly.destroy(); d.afHasPoly = false;
    d.afResamp = audioSR > 0.0;
    d.afOutSR = d.afResamp ? audioSR : d.sr;
    d.afMode = 3;
    if (d.afResamp) { B200_TRY(buildResampler(d.sr, audioSR, std::max(d.cap, 64), d.afDecim, d.afPoly, d.afHasPoly, d.afMode)); }
    d.deOn = tau > 0.0;
    if (d.deOn) {
        const float dt = 1.0f / d.afOutSR;          // Deemphasis::updateAlpha (deephasis.h:82-85)
        d.deAlpha = dt / (tau + dt);
    }
    const float2 zero = make_float2(0.f, 0.f);
    B200_CUDA(cudaMemcpy(&d.dState->deLast, &zero, sizeof(zero), cudaMemcpyHostToDevice));
    d.lastAudio = nullptr; d.lastAudioCount = 0;
    return B200_OK;
}

// The sequential walks (AGC, pilot PLL) run one warp per VFO and are bound by the latency of a dependent chain; a warp
// that shares its SM with the wide kernels of the next block waits for issue slots at every step of that chain (measured:
// the AGC walk of the bench's one VFO takes 0.58 ms alone and 1.09 ms beside the next block's kernels).  While there are
// few walks, each asks for so much dynamic shared memory that it gets an SM to itself; with many VFOs the walks fill the
// machine themselves and take what fits.
constexpr size_t kSoloSmemBytes = 180 * 1024;
constexpr size_t kSoloMaxJobs = 24;
static size_t soloSmem(size_t nJobs) {
    if (nJobs > kSoloMaxJobs) { return 0; }
    // (function attributes are per device: set on every use, as the FIR launches do)
    cudaFuncSetAttribute(agcWalkKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSoloSmemBytes);
    cudaFuncSetAttribute(pllKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSoloSmemBytes);
    cudaFuncSetAttribute(pllErrKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSoloSmemBytes);
    return kSoloSmemBytes;
}

int processDemods(VfoBank& bank, cudaStream_t s) {
    std::vector<DemodJob> seq, quad, all, ssb;
    std::vector<FirJob> firs;
    std::vector<std::vector<FirJob>> afWaves;   // wave k = stage k of every VFO's audio resampler
    std::vector<DeemphJob> deJobs;
    std::vector<IfJob> ifJobs;
    std::vector<FirJob> wfmA, wfmB;             // WFM stereo: pilot band-pass (two real FIRs), then the two audio low-passes
    std::vector<PllJob> pllJobs;
    std::vector<WfmMixJob> mixJobs;
    std::vector<LrJob> lrJobs;
    std::vector<FmifJob> fmJobs;
    std::vector<VolJob> volJobs;
    std::vector<RotJob> rdsRot;
    std::vector<RdsXlateJob> rdsX;
    std::vector<std::vector<FirJob>> rdsWaves;  // wave k = stage k of every VFO's RDS resampler
    const bool exactRot = bank.rotatorMode == B200_ROTATOR_EXACT;
    int maxN = 0, maxMix = 0, maxFm = 0, maxVol = 0, maxRds = 0;
    for (size_t vi = 0; vi < bank.vfos.size(); vi++) {
        Vfo& v = bank.vfos[vi];
        if (!v.live || !v.demod) { continue; }
        DemodState& d = *v.demod;
        const int n = v.lastOut;
        d.lastCount = n;
        if (n > d.cap) { B200_CUDA(cudaStreamSynchronize(s)); B200_TRY(demodEnsureCap(d, n)); }
        DemodJob j{};
        const cf32* demodIn = v.lastOutPtr;
        if ((d.nbOn || d.sqOn) && n > 0) {
            if (n > d.ifCap) {
                B200_CUDA(cudaStreamSynchronize(s));
                if (d.dIf) { cudaFree(d.dIf); d.dIf = nullptr; }
                d.ifCap = std::max(n + 64, d.cap);
                B200_CUDA(cudaMalloc((void**)&d.dIf, (size_t)d.ifCap * sizeof(cf32)));
            }
            IfJob f{};
            f.in = v.lastOutPtr; f.out = d.dIf; f.st = d.dState; f.rate = d.nbRate; f.invRate = 1.0f - d.nbRate; f.nbLevel = d.nbLevel;
            f.sqLevel = d.sqLevel; f.n = n; f.nbOn = d.nbOn ? 1 : 0; f.sqOn = d.sqOn ? 1 : 0;
            ifJobs.push_back(f);
            demodIn = d.dIf;
        }
        if (d.fmOn && n > 0) {
            if (n > d.fmCap) {
                B200_CUDA(cudaStreamSynchronize(s));
                if (d.dFm) { cudaFree(d.dFm); d.dFm = nullptr; }
                d.fmCap = std::max(n + 64, d.cap);
                B200_CUDA(cudaMalloc((void**)&d.dFm, (size_t)d.fmCap * sizeof(cf32)));
            }
            fmJobs.push_back(FmifJob{ demodIn, d.dFm, d.dFmHist[d.fmCur], d.dFmHist[d.fmCur ^ 1], d.dFmTab, n, d.fmBins });
            d.fmCur ^= 1;
            demodIn = d.dFm;
            maxFm = std::max(maxFm, n);
        }
        d.lastIf = demodIn;
        j.in = demodIn; j.mono = d.dMono; j.stereo = d.dStereo; j.st = d.dState; j.wtab = d.dWtab;
        j.exact = (bank.rotatorMode == B200_ROTATOR_EXACT) ? 1 : 0;
        j.maps = d.dMaps; j.amps = d.dAmps;
        j.env0 = (int)std::min<long long>(d.envCount, 1 << 30);
        j.agc = d.agc; j.angle0 = d.angle; j.theta = d.theta; j.stepRe = d.dRe; j.stepIm = d.dIm;
        j.invDeviation = d.invDeviation; j.dcRate = d.dcRate; j.kind = d.kind; j.agcMode = d.agcMode; j.n = n; j.quadParity = d.quadParity;
        const float* src = d.dMono;
        const bool wfmStereo = d.kind == B200_DEMOD_WFM && d.wfmStereo;
        if (wfmStereo) {
            FirJob f;
            d.pilotRe.fillJob(f, d.dMono, n); wfmA.push_back(f); d.pilotRe.advance(n);
            d.pilotIm.fillJob(f, d.dMono, n); wfmA.push_back(f); d.pilotIm.advance(n);
            PllJob pj{};
            pj.pr = (const float*)d.pilotRe.dOut; pj.pi = (const float*)d.pilotIm.dOut; pj.phaseOut = d.dPllPhase; pj.st = d.dState;
            pj.alpha = d.pllAlpha; pj.beta = d.pllBeta; pj.minFreq = d.pllMinFreq; pj.maxFreq = d.pllMaxFreq; pj.n = n;
            if (n > 0) { pllJobs.push_back(pj); }
            WfmMixJob mj{};
            mj.mpx = d.dMono; mj.hist = d.dDelayHist[d.delayCur]; mj.histNext = d.dDelayHist[d.delayCur ^ 1]; mj.phase = d.dPllPhase;
            mj.l = d.dL; mj.r = d.dR; mj.n = n; mj.delay = d.wfmDelay;
            mixJobs.push_back(mj);
            d.delayCur ^= 1;
            maxMix = std::max(maxMix, std::max(n, d.wfmDelay));
            const float* lsrc = d.dL; const float* rsrc = d.dR;
            if (d.hasFir && d.filtering) {
                d.fir.fillJob(f, d.dL, n); wfmB.push_back(f); d.fir.advance(n);
                d.firR.fillJob(f, d.dR, n); wfmB.push_back(f); d.firR.advance(n);
                lsrc = (const float*)d.fir.dOut; rsrc = (const float*)d.firR.dOut;
            }
            if (n > 0) { lrJobs.push_back(LrJob{ lsrc, rsrc, d.dStereo, n, 0 }); }
        }
        else if (d.hasFir && d.filtering) {
            FirJob f; d.fir.fillJob(f, d.dMono, n);
            firs.push_back(f);
            d.fir.advance(n);
            src = (const float*)d.fir.dOut;
        }
        j.stereoSrc = src;
        {   // audio chain bookkeeping: what leaves the demodulator goes through resampler stages, then the de-emphasis
            const void* data = d.dStereo; int m = n; size_t wave = 0;
            if (d.afResamp) {
                auto pushStage = [&](Stage& st) -> int {
                    const int outN = st.outCountFor(m);
                    if (outN > st.outCap) { B200_CUDA(cudaStreamSynchronize(s)); B200_TRY(st.ensureOutCap(outN + 64)); }
                    FirJob f; st.fillJob(f, data, m);
                    if (afWaves.size() <= wave) { afWaves.resize(wave + 1); }
                    afWaves[wave++].push_back(f);
                    st.advance(m);
                    data = st.dOut; m = f.nOut;
                    return B200_OK;
                };
                if (d.afMode == 0 || d.afMode == 1) { for (auto& st : d.afDecim) { B200_TRY(pushStage(st)); } }
                if (d.afHasPoly) { B200_TRY(pushStage(d.afPoly)); }
            }
            if (d.deOn && m > 0) { deJobs.push_back(DeemphJob{ (float2*)data, d.dState, d.deAlpha, m }); }
            if (d.volOn && m > 0) { volJobs.push_back(VolJob{ (float*)data, d.volGain, 2 * m }); maxVol = std::max(maxVol, 2 * m); }
            d.lastAudio = (const float*)data; d.lastAudioCount = m;
        }
        if (d.rdsOn) {
            // RDS branch: MPX -> complex, down by 57 kHz, RationalResampler<complex_t> to 5 kS/s (broadcast_fm.h:169-175,197-206)
            const void* data = d.dRdsX; int m = n; size_t wave = 0;
            if (n > 0) {
                RotJob r{};
                r.chk = d.dRdsChk; r.state = d.dRdsPhase; r.angle0 = d.rdsAngle; r.theta = d.rdsTheta;
                r.dRe = d.rdsDRe; r.dIm = d.rdsDIm; r.count = n; r.exact = exactRot ? 1 : 0;
                rdsRot.push_back(r);
                rdsX.push_back(RdsXlateJob{ d.dMono, d.dRdsX, d.dRdsChk, d.dRdsWtab, n, 0 });
                if (!exactRot) { d.rdsAngle = fmod(d.rdsAngle + d.rdsTheta * (double)n, 2.0 * design::kPi); }
                maxRds = std::max(maxRds, n);
            }
            auto pushStage = [&](Stage& st) -> int {
                const int outN = st.outCountFor(m);
                if (outN > st.outCap) { B200_CUDA(cudaStreamSynchronize(s)); B200_TRY(st.ensureOutCap(outN + 64)); }
                FirJob f; st.fillJob(f, data, m);
                if (rdsWaves.size() <= wave) { rdsWaves.resize(wave + 1); }
                rdsWaves[wave++].push_back(f);
                st.advance(m);
                data = st.dOut; m = f.nOut;
                return B200_OK;
            };
            if (d.rdsMode == 0 || d.rdsMode == 1) { for (auto& st : d.rdsDecim) { B200_TRY(pushStage(st)); } }
            if (d.rdsHasPoly) { B200_TRY(pushStage(d.rdsPoly)); }
            d.lastRds = (const float*)data; d.lastRdsCount = m;
        }
        if (n > 0) {
            if (d.kind == B200_DEMOD_NFM || d.kind == B200_DEMOD_WFM) {
                quad.push_back(j); d.quadParity ^= 1;
                if (wfmStereo) { maxN = std::max(maxN, n); continue; }      // its stereo rows come from lrKernel, not stereoKernel
            }
            else {
                const bool isSsb = (d.kind != B200_DEMOD_AM);
                const bool fastSsb = isSsb && !j.exact && d.agc.attack > 0.0f && d.dMaps;
                if (fastSsb) { ssb.push_back(j); } else { seq.push_back(j); }
                if (isSsb && !j.exact) { d.angle = fmod(d.angle + d.theta * (double)n, 2.0 * design::kPi); }
                if (isSsb && d.agc.attack > 0.0f) { d.envCount += n; }
                if (!fastSsb) { all.push_back(j); }
                maxN = std::max(maxN, n);
                continue;
            }
            all.push_back(j);
            maxN = std::max(maxN, n);
        }
    }
    if (all.empty() && firs.empty() && ssb.empty() && afWaves.empty() && wfmA.empty() && rdsWaves.empty()) { return B200_OK; }
    for (auto& w : afWaves) { firJobsBalance(w.data(), (int)w.size()); }
    for (auto& w : rdsWaves) { firJobsBalance(w.data(), (int)w.size()); }
    firJobsBalance(firs.data(), (int)firs.size());
    firJobsBalance(wfmA.data(), (int)wfmA.size());
    firJobsBalance(wfmB.data(), (int)wfmB.size());
    // one upload: [seq][quad][all][firs][ssb][af waves][deemph]
    size_t afJobs = 0;
    for (auto& w : afWaves) { afJobs += w.size(); }
    for (auto& w : rdsWaves) { afJobs += w.size(); }
    const size_t bytes = (seq.size() + quad.size() + all.size() + ssb.size()) * sizeof(DemodJob) + (firs.size() + afJobs) * sizeof(FirJob) +
                         deJobs.size() * sizeof(DeemphJob) + ifJobs.size() * sizeof(IfJob) + (wfmA.size() + wfmB.size()) * sizeof(FirJob) +
                         pllJobs.size() * sizeof(PllJob) + mixJobs.size() * sizeof(WfmMixJob) + lrJobs.size() * sizeof(LrJob) +
                         fmJobs.size() * sizeof(FmifJob) + volJobs.size() * sizeof(VolJob) + rdsRot.size() * sizeof(RotJob) + rdsX.size() * sizeof(RdsXlateJob) +
                         64 * (afWaves.size() + rdsWaves.size() + 24);
    void* hp = nullptr; int slot = 0;
    B200_TRY(bank.ring.acquire(bytes, &hp, &slot));
    char* h = (char*)hp; size_t off = 0;
    auto put = [&](const void* p, size_t n) { size_t o = off; if (n) { memcpy(h + off, p, n); } off = (off + n + 63) & ~(size_t)63; return o; };
    const size_t oS = put(seq.data(), seq.size() * sizeof(DemodJob));
    const size_t oQ = put(quad.data(), quad.size() * sizeof(DemodJob));
    const size_t oA = put(all.data(), all.size() * sizeof(DemodJob));
    const size_t oF = put(firs.data(), firs.size() * sizeof(FirJob));
    const size_t oB = put(ssb.data(), ssb.size() * sizeof(DemodJob));
    std::vector<size_t> oW(afWaves.size());
    for (size_t w = 0; w < afWaves.size(); w++) { oW[w] = put(afWaves[w].data(), afWaves[w].size() * sizeof(FirJob)); }
    const size_t oD = put(deJobs.data(), deJobs.size() * sizeof(DeemphJob));
    const size_t oI = put(ifJobs.data(), ifJobs.size() * sizeof(IfJob));
    const size_t oWA = put(wfmA.data(), wfmA.size() * sizeof(FirJob));
    const size_t oWB = put(wfmB.data(), wfmB.size() * sizeof(FirJob));
    const size_t oP = put(pllJobs.data(), pllJobs.size() * sizeof(PllJob));
    const size_t oM = put(mixJobs.data(), mixJobs.size() * sizeof(WfmMixJob));
    const size_t oL = put(lrJobs.data(), lrJobs.size() * sizeof(LrJob));
    const size_t oFm = put(fmJobs.data(), fmJobs.size() * sizeof(FmifJob));
    const size_t oV = put(volJobs.data(), volJobs.size() * sizeof(VolJob));
    const size_t oRr = put(rdsRot.data(), rdsRot.size() * sizeof(RotJob));
    const size_t oRx = put(rdsX.data(), rdsX.size() * sizeof(RdsXlateJob));
    std::vector<size_t> oRw(rdsWaves.size());
    for (size_t w = 0; w < rdsWaves.size(); w++) { oRw[w] = put(rdsWaves[w].data(), rdsWaves[w].size() * sizeof(FirJob)); }
    if (off > bank.dDemodDescBytes) {
        B200_CUDA(cudaStreamSynchronize(s));
        if (bank.dDemodDesc) { cudaFree(bank.dDemodDesc); }
        bank.dDemodDescBytes = off * 2;
        B200_CUDA(cudaMalloc(&bank.dDemodDesc, bank.dDemodDescBytes));
    }
    B200_TRY(uploadDesc(bank.dDemodDesc, h, off, s, bank.descViaParams));
    char* dd = (char*)bank.dDemodDesc;
    if (!ifJobs.empty()) {
        ifChainKernel<<<(unsigned)ifJobs.size(), 256, 0, s>>>((const IfJob*)(dd + oI));
        B200_CHECK_LAUNCH(); countLaunch();
    }
    if (!fmJobs.empty()) {
        fmifKernel<<<dim3((maxFm + kFmTile - 1) / kFmTile, (unsigned)fmJobs.size()), kFmThreads, 0, s>>>((const FmifJob*)(dd + oFm));
        B200_CHECK_LAUNCH(); countLaunch();
    }
    if (!quad.empty()) {
        quadratureKernel<<<dim3((maxN + 255) / 256, (unsigned)quad.size()), 256, 0, s>>>((const DemodJob*)(dd + oQ));
        B200_CHECK_LAUNCH(); countLaunch();
    }
    if (!rdsX.empty()) {
        B200_TRY(launchCheckpoints((const RotJob*)(dd + oRr), (int)rdsRot.size(), exactRot ? (int)rdsRot.size() : 0, maxRds, s));
        rdsXlateKernel<<<dim3((maxRds + 255) / 256, (unsigned)rdsX.size()), 256, 0, s>>>((const RdsXlateJob*)(dd + oRx));
        B200_CHECK_LAUNCH(); countLaunch();
    }
    for (size_t w = 0; w < rdsWaves.size(); w++) {
        int maxH = 0;
        for (auto& f : rdsWaves[w]) { maxH = std::max(maxH, f.H); }
        B200_TRY(launchFirJobs((const FirJob*)(dd + oRw[w]), rdsWaves[w].data(), (int)rdsWaves[w].size(), 2, s));
        B200_TRY(launchHistoryUpdate((const FirJob*)(dd + oRw[w]), (int)rdsWaves[w].size(), maxH, s));
    }
    if (!wfmA.empty()) {
        // broadcast FM stereo: pilot band-pass -> PLL -> mix-down + matrix -> audio low-passes -> interleave
        int maxH = 0;
        for (auto& f : wfmA) { maxH = std::max(maxH, f.H); }
        B200_TRY(launchFirJobs((const FirJob*)(dd + oWA), wfmA.data(), (int)wfmA.size(), 1, s));
        B200_TRY(launchHistoryUpdate((const FirJob*)(dd + oWA), (int)wfmA.size(), maxH, s));
        if (!pllJobs.empty()) {
            // exact banks: the loop in the reference's operation order; default: its error form (see pllErrKernel)
            if (bank.rotatorMode == B200_ROTATOR_EXACT) { pllKernel<<<(unsigned)pllJobs.size(), 32, soloSmem(pllJobs.size()), s>>>((const PllJob*)(dd + oP)); }
            else { pllErrKernel<<<(unsigned)pllJobs.size(), kPeThreads, soloSmem(pllJobs.size()), s>>>((const PllJob*)(dd + oP)); }
            B200_CHECK_LAUNCH(); countLaunch();
        }
        wfmMixKernel<<<dim3((maxMix + 255) / 256, (unsigned)mixJobs.size()), 256, 0, s>>>((const WfmMixJob*)(dd + oM));
        B200_CHECK_LAUNCH(); countLaunch();
        if (!wfmB.empty()) {
            maxH = 0;
            for (auto& f : wfmB) { maxH = std::max(maxH, f.H); }
            B200_TRY(launchFirJobs((const FirJob*)(dd + oWB), wfmB.data(), (int)wfmB.size(), 1, s));
            B200_TRY(launchHistoryUpdate((const FirJob*)(dd + oWB), (int)wfmB.size(), maxH, s));
        }
        if (!lrJobs.empty()) {
            lrKernel<<<dim3((maxN + 255) / 256, (unsigned)lrJobs.size()), 256, 0, s>>>((const LrJob*)(dd + oL));
            B200_CHECK_LAUNCH(); countLaunch();
        }
    }
    if (!ssb.empty()) {
        int maxS = 0;
        for (auto& j : ssb) { maxS = std::max(maxS, j.n); }
        const dim3 grid((maxS + 255) / 256, (unsigned)ssb.size());
        ssbPrepKernel<<<grid, 256, 0, s>>>((const DemodJob*)(dd + oB));
        B200_CHECK_LAUNCH();
        agcWalkKernel<<<(unsigned)ssb.size(), 32, soloSmem(ssb.size()), s>>>((const DemodJob*)(dd + oB));
        B200_CHECK_LAUNCH();
        ssbFinishKernel<<<grid, 256, 0, s>>>((const DemodJob*)(dd + oB));
        B200_CHECK_LAUNCH();
        countLaunch(3);
    }
    if (!seq.empty()) {
        seqDemodKernel<<<(unsigned)seq.size(), kSeqThreads, 0, s>>>((const DemodJob*)(dd + oS));
        B200_CHECK_LAUNCH(); countLaunch();
    }
    if (!firs.empty()) {
        int maxH = 0;
        for (auto& f : firs) { maxH = std::max(maxH, f.H); }
        B200_TRY(launchFirJobs((const FirJob*)(dd + oF), firs.data(), (int)firs.size(), 1, s));
        B200_TRY(launchHistoryUpdate((const FirJob*)(dd + oF), (int)firs.size(), maxH, s));
    }
    if (!all.empty()) {
        stereoKernel<<<dim3((maxN + 255) / 256, (unsigned)all.size()), 256, 0, s>>>((const DemodJob*)(dd + oA));
        B200_CHECK_LAUNCH(); countLaunch();
    }
    for (size_t w = 0; w < afWaves.size(); w++) {
        int maxH = 0;
        for (auto& f : afWaves[w]) { maxH = std::max(maxH, f.H); }
        B200_TRY(launchFirJobs((const FirJob*)(dd + oW[w]), afWaves[w].data(), (int)afWaves[w].size(), 2, s));
        B200_TRY(launchHistoryUpdate((const FirJob*)(dd + oW[w]), (int)afWaves[w].size(), maxH, s));
    }
    if (!deJobs.empty()) {
        deemphKernel<<<(unsigned)deJobs.size(), 64, 0, s>>>((const DeemphJob*)(dd + oD));
        B200_CHECK_LAUNCH(); countLaunch();
    }
    if (!volJobs.empty()) {
        volumeKernel<<<dim3((maxVol + 255) / 256, (unsigned)volJobs.size()), 256, 0, s>>>((const VolJob*)(dd + oV));
        B200_CHECK_LAUNCH(); countLaunch();
    }
    B200_TRY(bank.ring.release(slot, s));   // after the last reader of the slot was enqueued
    return B200_OK;
}

} // namespace b200

using namespace b200;

extern "C" {

int b200_vfo_set_demod(b200_vfo_bank* bank, int vfo, int kind, double p0, double p1, double p2, double p3) {
    if (!bank) { setError("NULL bank handle"); return B200_EINVAL; }
    B200_TRY(bank->impl.check(vfo));
    DeviceGuard g(bank->impl.device);
    B200_TRY(bank->impl.syncLast());
    return demodConfigure(bank->impl, bank->impl.vfos[vfo], kind, p0, p1, p2, p3);
}

int b200_vfo_set_audio_post(b200_vfo_bank* bank, int vfo, double audio_samplerate, double deemph_tau) {
    if (!bank) { setError("NULL bank handle"); return B200_EINVAL; }
    B200_TRY(bank->impl.check(vfo));
    DeviceGuard g(bank->impl.device);
    B200_TRY(bank->impl.syncLast());
    return afConfigure(bank->impl, bank->impl.vfos[vfo], audio_samplerate, deemph_tau);
}

int b200_vfo_set_if_chain(b200_vfo_bank* bank, int vfo, int nb_enabled, double nb_rate, double nb_level, int squelch_enabled, double squelch_level) {
    if (!bank) { setError("NULL bank handle"); return B200_EINVAL; }
    B200_TRY(bank->impl.check(vfo));
    Vfo& v = bank->impl.vfos[vfo];
    if (!v.demod) { setError("the IF chain sits in front of a demodulator: attach one to the VFO first"); return B200_EINVAL; }
    DeviceGuard g(bank->impl.device);
    B200_TRY(bank->impl.syncLast());
    DemodState& d = *v.demod;
    const bool wasNb = d.nbOn;
    d.nbOn = nb_enabled != 0; d.nbRate = (float)nb_rate; d.nbLevel = (float)nb_level;
    d.sqOn = squelch_enabled != 0; d.sqLevel = (float)squelch_level;
    if (d.nbOn && !wasNb) {
        // a block that joins the chain starts from its initial state (NoiseBlanker::amp = 1)
        const float one = 1.0f;
        B200_CUDA(cudaMemcpy(&d.dState->nbAmp, &one, sizeof(float), cudaMemcpyHostToDevice));
    }
    return B200_OK;
}

int b200_vfo_set_if_nr(b200_vfo_bank* bank, int vfo, int enabled, int bins) {
    if (!bank) { setError("NULL bank handle"); return B200_EINVAL; }
    B200_TRY(bank->impl.check(vfo));
    Vfo& v = bank->impl.vfos[vfo];
    if (!v.demod) { setError("the IF chain sits in front of a demodulator: attach one to the VFO first"); return B200_EINVAL; }
    if (enabled && (bins < 2 || bins > kFmMaxBins)) { setError("IF noise reduction takes 2..%d bins, got %d", kFmMaxBins, bins); return B200_EINVAL; }
    DeviceGuard g(bank->impl.device);
    B200_TRY(bank->impl.syncLast());
    DemodState& d = *v.demod;
    if (!enabled) { d.fmOn = false; return B200_OK; }
    const bool fresh = !d.fmOn || d.fmBins != bins;
    if (!d.dFmTab) {
        B200_CUDA(cudaMalloc((void**)&d.dFmTab, (size_t)(kFmMaxBins * kFmMaxBins + kFmMaxBins) * sizeof(float2)));
        for (int i = 0; i < 2; i++) { B200_CUDA(cudaMalloc((void**)&d.dFmHist[i], kFmMaxBins * sizeof(cf32))); }
    }
    if (fresh) {
        // FMIF::initBuffers (fm_if.h:106-124): Nuttall window over bins-1, cleared delay line
        std::vector<float2> tab((size_t)kFmMaxBins * kFmMaxBins + kFmMaxBins, make_float2(0.f, 0.f));
        for (int n = 0; n < bins; n++) {
            const float w = (float)design::nuttall(n, bins - 1);
            for (int k = 0; k < bins; k++) {
                const double a = -2.0 * design::kPi * (double)((k * n) % bins) / (double)bins;
                tab[(size_t)n * kFmMaxBins + k] = make_float2((float)((double)w * cos(a)), (float)((double)w * sin(a)));
            }
        }
        for (int k = 0; k < bins; k++) {
            const double a = 2.0 * design::kPi * (double)((k * (bins / 2)) % bins) / (double)bins;
            tab[(size_t)kFmMaxBins * kFmMaxBins + k] = make_float2((float)cos(a), (float)sin(a));
        }
        B200_CUDA(cudaMemcpy(d.dFmTab, tab.data(), tab.size() * sizeof(float2), cudaMemcpyHostToDevice));
        for (int i = 0; i < 2; i++) { B200_CUDA(cudaMemset(d.dFmHist[i], 0, kFmMaxBins * sizeof(cf32))); }
        d.fmCur = 0;
    }
    d.fmOn = true; d.fmBins = bins;
    return B200_OK;
}

int b200_vfo_set_volume(b200_vfo_bank* bank, int vfo, double volume, int muted) {
    if (!bank) { setError("NULL bank handle"); return B200_EINVAL; }
    B200_TRY(bank->impl.check(vfo));
    Vfo& v = bank->impl.vfos[vfo];
    if (!v.demod) { setError("the volume block follows a demodulator: attach one to the VFO first"); return B200_EINVAL; }
    B200_TRY(bank->impl.syncLast());
    v.demod->volOn = true;
    v.demod->volGain = muted ? 0.0f : powf((float)volume, 2.0f);   // Volume::init / setVolume (audio/volume.h:13-16,22-26)
    return B200_OK;
}

const float* b200_vfo_if_ptr(b200_vfo_bank* bank, int vfo) {
    if (!bank || bank->impl.check(vfo) != B200_OK) { return nullptr; }
    if (bank->impl.tailUnjoined) { DeviceGuard g(bank->impl.device); if (bank->impl.joinTail() != B200_OK) { return nullptr; } }
    const Vfo& v = bank->impl.vfos[vfo];
    if (!v.demod) { return nullptr; }
    return (const float*)v.demod->lastIf;
}

const float* b200_vfo_rds_ptr(b200_vfo_bank* bank, int vfo) {
    if (!bank || bank->impl.check(vfo) != B200_OK) { return nullptr; }
    if (bank->impl.tailUnjoined) { DeviceGuard g(bank->impl.device); if (bank->impl.joinTail() != B200_OK) { return nullptr; } }
    const Vfo& v = bank->impl.vfos[vfo];
    if (!v.demod || !v.demod->rdsOn) { return nullptr; }
    return v.demod->lastRds;
}

int b200_vfo_rds_count(b200_vfo_bank* bank, int vfo) {
    if (!bank) { setError("NULL bank handle"); return B200_EINVAL; }
    B200_TRY(bank->impl.check(vfo));
    const Vfo& v = bank->impl.vfos[vfo];
    if (!v.demod || !v.demod->rdsOn) { setError("VFO %d has no RDS branch (broadcast FM demodulator with p2 != 0)", vfo); return B200_EINVAL; }
    return v.demod->lastRdsCount;
}

const float* b200_vfo_audio_ptr(b200_vfo_bank* bank, int vfo) {
    if (!bank || bank->impl.check(vfo) != B200_OK) { return nullptr; }
    if (bank->impl.tailUnjoined) { DeviceGuard g(bank->impl.device); if (bank->impl.joinTail() != B200_OK) { return nullptr; } }
    const Vfo& v = bank->impl.vfos[vfo];
    if (!v.demod) { return nullptr; }
    return v.demod->lastAudio ? v.demod->lastAudio : v.demod->dStereo;
}

int b200_vfo_audio_count(b200_vfo_bank* bank, int vfo) {
    if (!bank) { setError("NULL bank handle"); return B200_EINVAL; }
    B200_TRY(bank->impl.check(vfo));
    const Vfo& v = bank->impl.vfos[vfo];
    if (!v.demod) { setError("VFO %d has no demodulator", vfo); return B200_EINVAL; }
    return v.demod->lastAudioCount;
}

int b200_vfo_audio_max_out(b200_vfo_bank* bank, int vfo, int count) {
    if (!bank) { setError("NULL bank handle"); return B200_EINVAL; }
    B200_TRY(bank->impl.check(vfo));
    const Vfo& v = bank->impl.vfos[vfo];
    if (!v.demod) { setError("VFO %d has no demodulator", vfo); return B200_EINVAL; }
    return afMaxOut(*v.demod, bank->impl.maxOut(vfo, count));
}

} // extern "C"

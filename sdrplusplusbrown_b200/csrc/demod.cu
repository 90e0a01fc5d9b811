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
    float* dMono = nullptr; float* dStereo = nullptr; int cap = 0;
    int lastCount = 0;
};

// per-call descriptor
struct DemodJob {
    const cf32* in; float* mono; float* stereo; const float* stereoSrc; SeqState* st;
    AgcParams agc;
    double angle0, theta; float stepRe, stepIm;
    float invDeviation, dcRate;
    int kind, agcMode, n, quadParity;
};

namespace {

constexpr int kChunk = 1024;
constexpr int kEnvLen = 4800;    // AGC::_totalEnvelopeLength (agc.h:162)

// AGC recurrence over a staged chunk (agc.h:96-112).  The average moves only for non-zero, non-NaN amplitudes;
// a skipped update means gain 1.  Because attack > decay, the reference's branch
//     amp' = (a > amp) ? amp*(1-att) + a*att : amp*(1-dec) + a*dec
// equals max(amp*(1-att) + a*att, amp*(1-dec) + a*dec): both candidates are evaluated as one FFMA each and
// joined by FMNMX, so the loop-carried dependency is ~9 cycles per sample instead of a compare/select chain.
// Products a*att / a*dec and the gains setPoint/amp do not depend on the running state and are computed by
// the whole warp before / after the walk.  sAmp: |in| on entry; on exit the gain.  sPa/sPd: scratch.
__device__ void agcChunk(float* sAmp, float* sPa, float* sPd, int n, float& amp, const AgcParams& p, int lane) {
    const bool ordered = p.attack >= p.decay;   // the max form needs attack >= decay (always true for the radio's ranges)
    for (int i = lane; i < n; i += 32) {
        const float a = sAmp[i];
        sPa[i] = a * p.attack;
        sPd[i] = a * p.decay;
    }
    __syncwarp();
    if (lane == 0) {
        float am = amp;
        if (ordered) {
            for (int i = 0; i < n; i++) {
                const float a = sAmp[i];
                const float cand = fmaxf(fmaf(am, p.invAttack, sPa[i]), fmaf(am, p.invDecay, sPd[i]));
                const bool upd = (a != 0.0f) && !isnan(cand);
                am = upd ? cand : am;
                sAmp[i] = upd ? am : -1.0f;          // -1 marks "no update": gain 1
            }
        }
        else {
            for (int i = 0; i < n; i++) {
                const float a = sAmp[i];
                const float cand = (a > am) ? fmaf(am, p.invAttack, sPa[i]) : fmaf(am, p.invDecay, sPd[i]);
                const bool upd = (a != 0.0f) && !isnan(cand);
                am = upd ? cand : am;
                sAmp[i] = upd ? am : -1.0f;
            }
        }
        amp = am;
    }
    __syncwarp();
    for (int i = lane; i < n; i += 32) {
        const float am = sAmp[i];
        sAmp[i] = (am < 0.0f) ? 1.0f : fminf(__fdiv_rn(p.setPoint, am), p.maxGain);
    }
    __syncwarp();
}

// start ramp (agc.h:131-134): `envelope` is re-initialised to 1 on every process() call and keeps its last
// ramp value for the rest of the call in which the ramp ends.
__device__ __forceinline__ float envelopeAt(int startEnvAtCall, int i) {
    const int se = startEnvAtCall + i;
    if (se < kEnvLen) { return (float)se / (float)kEnvLen; }
    return (startEnvAtCall < kEnvLen) ? (float)(kEnvLen - 1) / (float)kEnvLen : 1.0f;
}

// one warp per demodulator with a sequential part (SSB, AM)
__global__ void __launch_bounds__(32) seqDemodKernel(const DemodJob* __restrict__ jobs) {
    __shared__ float sA[kChunk];     // amplitudes -> gains
    __shared__ float sPa[kChunk], sPd[kChunk];
    __shared__ float sV[kChunk];     // real working values
    __shared__ float2 sC[kChunk];    // complex working values (AM carrier AGC)
    const DemodJob j = jobs[blockIdx.x];
    const int lane = threadIdx.x;
    SeqState st = *j.st;
    const int envCarrier0 = st.carrier.startEnvelope, envAudio0 = st.audio.startEnvelope;
    const bool agcOn = j.agc.attack > 0.0f;   // agc.h:87-90: attack <= 0 copies the input through

    for (int base = 0; base < j.n; base += kChunk) {
        const int m = min(kChunk, j.n - base);
        if (j.kind == B200_DEMOD_AM) {
            // ---- carrier AGC on the complex IF (am.h:103-106)
            for (int i = lane; i < m; i += 32) {
                const float2 x = __ldg(reinterpret_cast<const float2*>(j.in) + base + i);
                sC[i] = x;
                sA[i] = sqrtf(__fadd_rn(__fmul_rn(x.x, x.x), __fmul_rn(x.y, x.y)));   // complex_t::amplitude (types.h:77-79)
            }
            __syncwarp();
            if (j.agcMode == 0 && agcOn) {
                agcChunk(sA, sPa, sPd, m, st.carrier.amp, j.agc, lane);
                for (int i = lane; i < m; i += 32) {
                    const float g = sA[i], e = envelopeAt(envCarrier0, base + i);
                    float2 x = sC[i];
                    x.x = __fmul_rn(__fmul_rn(x.x, g), e); x.y = __fmul_rn(__fmul_rn(x.y, g), e);
                    sC[i] = x;
                }
                __syncwarp();
            }
            // ---- magnitude (am.h:120) and 1-pole DC blocker (dc_blocker.h:54-60)
            for (int i = lane; i < m; i += 32) {
                const float2 x = sC[i];
                sV[i] = sqrtf(__fadd_rn(__fmul_rn(x.x, x.x), __fmul_rn(x.y, x.y)));
            }
            __syncwarp();
            if (lane == 0) {
                float o = st.dcOffset;
                for (int i = 0; i < m; i++) {
                    const float y = __fsub_rn(sV[i], o);
                    o = __fadd_rn(o, __fmul_rn(y, j.dcRate));
                    sV[i] = y;
                }
                st.dcOffset = o;
            }
            __syncwarp();
            // ---- audio AGC (am.h:122-124)
            if (j.agcMode == 1 && agcOn) {
                for (int i = lane; i < m; i += 32) { sA[i] = fabsf(sV[i]); }
                __syncwarp();
                agcChunk(sA, sPa, sPd, m, st.audio.amp, j.agc, lane);
                for (int i = lane; i < m; i += 32) {
                    sV[i] = __fmul_rn(__fmul_rn(sV[i], sA[i]), envelopeAt(envAudio0, base + i));
                }
                __syncwarp();
            }
            for (int i = lane; i < m; i += 32) { j.mono[base + i] = sV[i]; }
            __syncwarp();
        }
        else {
            // ---- SSB: rotate by +-bw/2, keep the real part (ssb.h:84-92)
            for (int i = lane; i < m; i += 32) {
                const float2 x = __ldg(reinterpret_cast<const float2*>(j.in) + base + i);
                double a = fma(j.theta, (double)(base + i), j.angle0);
                a -= 6.283185307179586476925 * rint(a * 0.15915494309189533577);
                double sd, cd;
                sincos(a, &sd, &cd);
                const float re = __fsub_rn(__fmul_rn(x.x, (float)cd), __fmul_rn(x.y, (float)sd));
                sV[i] = re;
                sA[i] = fabsf(re);
            }
            __syncwarp();
            if (agcOn) {
                agcChunk(sA, sPa, sPd, m, st.audio.amp, j.agc, lane);
                for (int i = lane; i < m; i += 32) {
                    sV[i] = __fmul_rn(__fmul_rn(sV[i], sA[i]), envelopeAt(envAudio0, base + i));
                }
                __syncwarp();
            }
            for (int i = lane; i < m; i += 32) { j.mono[base + i] = sV[i]; }
            __syncwarp();
        }
    }
    if (lane == 0) {
        if (j.kind == B200_DEMOD_AM) {
            if (j.agcMode == 0 && agcOn) { st.carrier.startEnvelope = envCarrier0 + j.n; }
            if (j.agcMode == 1 && agcOn) { st.audio.startEnvelope = envAudio0 + j.n; }
        }
        else if (agcOn) { st.audio.startEnvelope = envAudio0 + j.n; }
        j.st->carrier = st.carrier; j.st->audio = st.audio; j.st->dcOffset = st.dcOffset;
    }
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

} // namespace

void demodDestroy(DemodState* d) {
    if (!d) { return; }
    d->fir.destroy();
    if (d->dState) { cudaFree(d->dState); }
    if (d->dMono) { cudaFree(d->dMono); }
    if (d->dStereo) { cudaFree(d->dStereo); }
    delete d;
}

static int demodEnsureCap(DemodState& d, int n) {
    if (n <= d.cap) { return B200_OK; }
    if (d.dMono) { cudaFree(d.dMono); d.dMono = nullptr; }
    if (d.dStereo) { cudaFree(d.dStereo); d.dStereo = nullptr; }
    d.cap = n + 64;
    B200_CUDA(cudaMalloc((void**)&d.dMono, (size_t)d.cap * sizeof(float)));
    B200_CUDA(cudaMalloc((void**)&d.dStereo, (size_t)d.cap * 2 * sizeof(float)));
    if (d.hasFir) { B200_TRY(d.fir.ensureOutCap(d.cap)); }
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
    }
    else {
        setError("unknown demodulator kind %d", kind);
        return B200_EINVAL;
    }
    if (!firTaps.empty()) {
        B200_TRY(d->fir.create(JOB_DECIM_FIR, firTaps, 1, 1, 1, 1, 64));
        d->hasFir = true;
    }
    B200_CUDA(cudaMalloc((void**)&d->dState, sizeof(SeqState)));
    B200_CUDA(cudaMemcpy(d->dState, &init, sizeof(SeqState), cudaMemcpyHostToDevice));
    return demodEnsureCap(*d, bank.maxOut((int)(&v - bank.vfos.data()), bank.maxBlock) + 8);
}

int processDemods(VfoBank& bank, cudaStream_t s) {
    std::vector<DemodJob> seq, quad, all;
    std::vector<FirJob> firs;
    int maxN = 0;
    for (size_t vi = 0; vi < bank.vfos.size(); vi++) {
        Vfo& v = bank.vfos[vi];
        if (!v.live || !v.demod) { continue; }
        DemodState& d = *v.demod;
        const int n = v.lastOut;
        d.lastCount = n;
        if (n > d.cap) { B200_CUDA(cudaStreamSynchronize(s)); B200_TRY(demodEnsureCap(d, n)); }
        DemodJob j{};
        j.in = v.lastOutPtr; j.mono = d.dMono; j.stereo = d.dStereo; j.st = d.dState;
        j.agc = d.agc; j.angle0 = d.angle; j.theta = d.theta; j.stepRe = d.dRe; j.stepIm = d.dIm;
        j.invDeviation = d.invDeviation; j.dcRate = d.dcRate; j.kind = d.kind; j.agcMode = d.agcMode; j.n = n; j.quadParity = d.quadParity;
        const float* src = d.dMono;
        if (d.hasFir && d.filtering) {
            FirJob f; d.fir.fillJob(f, d.dMono, n);
            firs.push_back(f);
            d.fir.advance(n);
            src = (const float*)d.fir.dOut;
        }
        j.stereoSrc = src;
        if (n > 0) {
            if (d.kind == B200_DEMOD_NFM) { quad.push_back(j); d.quadParity ^= 1; }
            else { seq.push_back(j); d.angle = fmod(d.angle + d.theta * (double)n, 2.0 * design::kPi); }
            all.push_back(j);
            maxN = std::max(maxN, n);
        }
    }
    if (all.empty() && firs.empty()) { return B200_OK; }
    // one upload: [seq][quad][all][firs]
    const size_t bytes = (seq.size() + quad.size() + all.size()) * sizeof(DemodJob) + firs.size() * sizeof(FirJob) + 256;
    static thread_local std::vector<char> scratch;
    (void)scratch;
    void* hp = nullptr; int slot = 0;
    B200_TRY(bank.ring.acquire(bytes, &hp, &slot));
    char* h = (char*)hp; size_t off = 0;
    auto put = [&](const void* p, size_t n) { size_t o = off; if (n) { memcpy(h + off, p, n); } off = (off + n + 63) & ~(size_t)63; return o; };
    const size_t oS = put(seq.data(), seq.size() * sizeof(DemodJob));
    const size_t oQ = put(quad.data(), quad.size() * sizeof(DemodJob));
    const size_t oA = put(all.data(), all.size() * sizeof(DemodJob));
    const size_t oF = put(firs.data(), firs.size() * sizeof(FirJob));
    if (off > bank.dDemodDescBytes) {
        B200_CUDA(cudaStreamSynchronize(s));
        if (bank.dDemodDesc) { cudaFree(bank.dDemodDesc); }
        bank.dDemodDescBytes = off * 2;
        B200_CUDA(cudaMalloc(&bank.dDemodDesc, bank.dDemodDescBytes));
    }
    B200_CUDA(cudaMemcpyAsync(bank.dDemodDesc, h, off, cudaMemcpyHostToDevice, s));
    B200_TRY(bank.ring.release(slot, s));
    char* dd = (char*)bank.dDemodDesc;
    if (!quad.empty()) {
        quadratureKernel<<<dim3((maxN + 255) / 256, (unsigned)quad.size()), 256, 0, s>>>((const DemodJob*)(dd + oQ));
        B200_CHECK_LAUNCH(); countLaunch();
    }
    if (!seq.empty()) {
        seqDemodKernel<<<(unsigned)seq.size(), 32, 0, s>>>((const DemodJob*)(dd + oS));
        B200_CHECK_LAUNCH(); countLaunch();
    }
    if (!firs.empty()) {
        int maxOut = 0, maxH = 0;
        for (auto& f : firs) { maxOut = std::max(maxOut, f.nOut); maxH = std::max(maxH, f.H); }
        B200_TRY(launchFirJobs((const FirJob*)(dd + oF), (int)firs.size(), maxOut, 1, s));
        B200_TRY(launchHistoryUpdate((const FirJob*)(dd + oF), (int)firs.size(), maxH, s));
    }
    if (!all.empty()) {
        stereoKernel<<<dim3((maxN + 255) / 256, (unsigned)all.size()), 256, 0, s>>>((const DemodJob*)(dd + oA));
        B200_CHECK_LAUNCH(); countLaunch();
    }
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

const float* b200_vfo_audio_ptr(b200_vfo_bank* bank, int vfo) {
    if (!bank || bank->impl.check(vfo) != B200_OK) { return nullptr; }
    const Vfo& v = bank->impl.vfos[vfo];
    return v.demod ? v.demod->dStereo : nullptr;
}

} // extern "C"

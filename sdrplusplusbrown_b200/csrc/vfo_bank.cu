// RxVFO bank — host orchestration.  Reference: dsp::channel::RxVFO (channel/rx_vfo.h:6-135) =
// FrequencyXlator -> RationalResampler{PowerDecimator -> PolyphaseResampler} -> FIR, one instance
// per VFO behind a Splitter.  Here all VFOs of a bank advance together, one launch per pipeline
// level, with the integer state machines (output counts, decimation offsets, resampler phase)
// mirrored on the host so counts are known without touching the device.
#include "vfo_bank.cuh"

#include <algorithm>
#include <cmath>
#include <thread>

#include "demod.cuh"
#include "xlate_decim.cuh"

namespace b200 {

bool deferTail = false;               // B200_OPT_DEFER_TAIL
bool fusedLeaveRoomEnabled = false;  // B200_OPT_LEAVE_ROOM (measured slower: see include/b200dsp.h)

// ------------------------------------------------------------------ Stage
int Stage::create(int kind_, const std::vector<float>& taps, int D_, int L_, int M_, int elemFloats_, int outCap_, bool planTaps_) {
    destroy();
    planTaps = planTaps_;
    kind = kind_; D = D_; L = L_; M = M_; elemFloats = elemFloats_;
    hostTaps = taps;
    offset = 0; phase = 0; cur = 0;
    std::vector<float> dev;
    if (kind == JOB_POLYPHASE) {
        int tpp = 0, rs = 0;
        dev = design::polyphaseBank(taps, L, tpp, rs);
        T = tpp; rowStride = rs;
    }
    else {
        dev = taps; T = (int)taps.size(); rowStride = 0;
    }
    H = T - 1;
    B200_CUDA(cudaMalloc(&dTaps, std::max<size_t>(dev.size(), 1) * sizeof(float)));
    B200_CUDA(cudaMemcpy(dTaps, dev.data(), dev.size() * sizeof(float), cudaMemcpyHostToDevice));
    histCapBytes = std::max<size_t>((size_t)H, 1) * elemFloats * sizeof(float);
    for (int i = 0; i < 2; i++) {
        B200_CUDA(cudaMalloc(&dHist[i], histCapBytes));
        B200_CUDA(cudaMemset(dHist[i], 0, histCapBytes));
    }
    return ensureOutCap(outCap_);
}

void Stage::destroy() {
    if (dTaps) { cudaFree(dTaps); dTaps = nullptr; }
    for (int i = 0; i < 2; i++) { if (dHist[i]) { cudaFree(dHist[i]); dHist[i] = nullptr; } }
    if (dOut) { cudaFree(dOut); dOut = nullptr; }
    if (dOutAlt) { cudaFree(dOutAlt); dOutAlt = nullptr; }
    outCap = 0; T = 0; H = 0;
}

int Stage::ensureOutCap(int cap) {
    if (cap <= outCap && dOut) { return B200_OK; }
    if (dOut) { cudaFree(dOut); dOut = nullptr; }
    if (dOutAlt) { cudaFree(dOutAlt); dOutAlt = nullptr; }
    outCap = std::max(cap, 16);
    B200_CUDA(cudaMalloc(&dOut, (size_t)outCap * elemFloats * sizeof(float)));
    return B200_OK;
}

int Stage::flipOut() {
    if (!dOutAlt) { B200_CUDA(cudaMalloc(&dOutAlt, (size_t)outCap * elemFloats * sizeof(float))); }
    std::swap(dOut, dOutAlt);
    return B200_OK;
}

int Stage::resetState() {
    offset = 0; phase = 0;
    if (dHist[cur] && H > 0) { B200_CUDA(cudaMemset(dHist[cur], 0, (size_t)H * elemFloats * sizeof(float))); }
    return B200_OK;
}

// FIR::setTaps (filter/fir.h:33-54): the newest history samples stay aligned with the end of the buffer;
// a longer filter sees zeros in front of them.
int Stage::setTapsKeepHistory(const std::vector<float>& taps) {
    const int oldT = T, newT = (int)taps.size();
    const size_t es = (size_t)elemFloats * sizeof(float);
    float* nt = nullptr;
    B200_CUDA(cudaMalloc(&nt, std::max<size_t>(taps.size(), 1) * sizeof(float)));
    B200_CUDA(cudaMemcpy(nt, taps.data(), taps.size() * sizeof(float), cudaMemcpyHostToDevice));
    void* nh[2] = { nullptr, nullptr };
    const size_t newBytes = std::max<size_t>((size_t)(newT - 1), 1) * es;
    for (int i = 0; i < 2; i++) {
        B200_CUDA(cudaMalloc(&nh[i], newBytes));
        B200_CUDA(cudaMemset(nh[i], 0, newBytes));
    }
    const int keep = std::min(oldT, newT) - 1;     // samples that survive
    if (keep > 0) {
        const char* src = (const char*)dHist[cur] + (size_t)((oldT - 1) - keep) * es;
        char* dst = (char*)nh[0] + (size_t)((newT - 1) - keep) * es;
        B200_CUDA(cudaMemcpy(dst, src, (size_t)keep * es, cudaMemcpyDeviceToDevice));
    }
    cudaFree(dTaps);
    for (int i = 0; i < 2; i++) { cudaFree(dHist[i]); dHist[i] = nh[i]; }
    dTaps = nt; cur = 0; histCapBytes = newBytes;
    hostTaps = taps; T = newT; H = newT - 1;
    return B200_OK;
}

int Stage::outCountFor(int nIn) const {
    if (kind == JOB_POLYPHASE) { return design::polyOutCount(phase, offset, L, M, nIn); }
    return design::decimOutCount(offset, D, nIn);
}

void Stage::advance(int nIn) {
    if (kind == JOB_POLYPHASE) { design::polyNextState(phase, offset, L, M, nIn); }
    else { offset = design::decimNextOffset(offset, D, nIn); }
    cur ^= 1;
}

void Stage::fillJob(FirJob& j, const void* data, int nIn) const {
    j.hist = dHist[cur]; j.histNext = dHist[cur ^ 1];
    j.data = data; j.out = dOut; j.taps = dTaps;
    j.H = H; j.T = T; j.D = D; j.offset = offset; j.nIn = nIn; j.nOut = outCountFor(nIn);
    j.kind = kind; j.L = L; j.M = M; j.phase0 = phase; j.rowStride = rowStride;
    j.elemFloats = elemFloats; j.plan = planTaps ? -1 : 0;     // -1: firJobPlan may pick the plan-table kernel
    firJobPlan(j);
}

// ------------------------------------------------------------------ bank
int VfoBank::init(double inSR_, int maxBlock_, int device_) {
    device = device_; inSR = inSR_; maxBlock = maxBlock_;
    DeviceGuard g(device);
    if (!g.ok) { setError("cannot select CUDA device %d", device); return B200_ECUDA; }
    B200_TRY(ring.init(64 * 1024));
    B200_CUDA(cudaEventCreateWithFlags(&evStage1, cudaEventDisableTiming));
    return B200_OK;
}

void VfoBank::destroy() {
    DeviceGuard g(device);
    if (everRan) { cudaStreamSynchronize(lastStream); }
    if (tailStream) { cudaStreamSynchronize(tailStream); cudaStreamDestroy(tailStream); tailStream = nullptr; }
    for (int i = 0; i < 2; i++) { if (evTail[i]) { cudaEventDestroy(evTail[i]); evTail[i] = nullptr; } }
    if (dTailDesc) { cudaFree(dTailDesc); dTailDesc = nullptr; }
    if (evStage1) { cudaEventDestroy(evStage1); evStage1 = nullptr; }
    for (auto& v : vfos) {
        if (!v.live) { continue; }
        for (auto& st : v.decimStages) { st.destroy(); }
        v.poly.destroy(); v.chan.destroy();
        if (v.dZ) { cudaFree(v.dZ); }
        freeRotator(v);
        demodDestroy(v.demod); v.demod = nullptr;
    }
    vfos.clear();
    if (dDemodDesc) { cudaFree(dDemodDesc); dDemodDesc = nullptr; }
    ring.destroy();
    if (dDesc) { cudaFree(dDesc); dDesc = nullptr; }
    if (hPinnedIn) { cudaFreeHost(hPinnedIn); hPinnedIn = nullptr; }
    if (dStageIn) { cudaFree(dStageIn); dStageIn = nullptr; }
    if (hPinnedOut) { cudaFreeHost(hPinnedOut); hPinnedOut = nullptr; }
}

int VfoBank::check(int v) const {
    if (v < 0 || v >= (int)vfos.size() || !vfos[v].live) { setError("bad VFO index %d", v); return B200_EINVAL; }
    return B200_OK;
}

int VfoBank::liveCount() const {
    int n = 0;
    for (auto& v : vfos) { n += v.live ? 1 : 0; }
    return n;
}

int VfoBank::syncLast() {
    if (everRan) { B200_CUDA(cudaStreamSynchronize(lastStream)); }
    if (tailStream) { B200_CUDA(cudaStreamSynchronize(tailStream)); }
    tailUnjoined = false; tailPending[0] = tailPending[1] = false;
    return B200_OK;
}

int VfoBank::joinTail() {
    if (tailUnjoined) {
        B200_CUDA(cudaStreamWaitEvent(lastStream, evTail[lastTailPar], 0));
        tailUnjoined = false;
    }
    return B200_OK;
}

// FrequencyXlator::init/setOffset (frequency_xlator.h:15-29): phaseDelta = (float)cos/sin of 2*pi*(-offset)/inSR
int VfoBank::configureXlator(Vfo& v) {
    const double w = design::hzToRads(-v.offset, inSR);
    v.dRe = (float)cos(w); v.dIm = (float)sin(w);
    v.calDirty = true;
    v.tableDirty = true;
    return B200_OK;
}

// RationalResampler::reconfigure (rational_resampler.h:128-173).  decim.setRatio() rebuilds the stage FIRs
// with empty history (power_decimator.h:92-110); resamp.setRatio() resets the resampler (polyphase_resampler.h:37-57).
int VfoBank::reconfigureResampler(Vfo& v) {
    design::ResampPlan p = design::planResampler(inSR, v.outSR);
    v.mode = p.mode; v.interp = p.interp; v.decim = p.decim;
    int cap = maxBlock;
    if (p.useDecim) {
        for (auto& st : v.decimStages) { st.destroy(); }
        v.decimStages.clear();
        const design::PlanDef* plan = design::findPlan(p.ratio);
        if (!plan) { setError("no decimation plan for ratio %d", p.ratio); return B200_EINVAL; }
        v.ratio = p.ratio;
        v.decimStages.resize(plan->nstages);
        for (int i = 0; i < plan->nstages; i++) {
            std::vector<float> t(plan->stages[i].ntaps);
            memcpy(t.data(), plan->stages[i].taps, sizeof(float) * t.size());
            cap = cap / plan->stages[i].decim + 2;
            B200_TRY(v.decimStages[i].create(JOB_DECIM_FIR, t, plan->stages[i].decim, 1, 1, 2, cap, true));
        }
    }
    const bool usesDecim = (p.mode == 0 || p.mode == 1);
    if (!usesDecim) { cap = maxBlock; v.ratio = 1; }
    if (p.mode == 0 || p.mode == 2) {
        v.resampTaps = p.taps;
        v.resampTapCount = (int)p.taps.size();
        const int pcap = (int)(((int64_t)cap * p.interp) / p.decim) + 4;
        B200_TRY(v.poly.create(JOB_POLYPHASE, p.taps, 1, p.interp, p.decim, 2, pcap));
        v.hasPoly = true;
        cap = pcap;
    }
    else {
        v.hasPoly = false;
    }
    if (v.chanCreated) { B200_TRY(v.chan.ensureOutCap(cap)); }
    return B200_OK;
}

// RxVFO::generateTaps (rx_vfo.h:117-121) + filter.init / filter.setTaps
int VfoBank::regenChannelTaps(Vfo& v, bool create) {
    const double w = v.bw / 2.0;
    std::vector<float> t = design::lowPass(w, w * 0.1, v.outSR);
    if (t.empty()) { setError("channel filter has no taps (bandwidth %g at %g S/s)", v.bw, v.outSR); return B200_EINVAL; }
    int cap = maxOut((int)(&v - vfos.data()), maxBlock) + 4;
    if (create) {
        B200_TRY(v.chan.create(JOB_DECIM_FIR, t, 1, 1, 1, 2, cap));
        v.chanCreated = true;
    }
    else {
        B200_TRY(v.chan.setTapsKeepHistory(t));
        B200_TRY(v.chan.ensureOutCap(cap));
    }
    return B200_OK;
}

int VfoBank::maxOut(int vi, int count) const {
    const Vfo& v = vfos[vi];
    int64_t n = count;
    if (v.mode == 0 || v.mode == 1) { for (auto& st : v.decimStages) { n = n / st.D + 2; } }
    if (v.mode == 0 || v.mode == 2) { n = (n * v.interp) / v.decim + 4; }
    return (int)n;
}

int VfoBank::addVfo(double outSR, double bw, double offset) {
    DeviceGuard g(device);
    B200_TRY(syncLast());
    int idx = -1;
    for (size_t i = 0; i < vfos.size(); i++) { if (!vfos[i].live) { idx = (int)i; break; } }
    if (idx < 0) {
        // keep addresses stable: Vfo objects hold no self-pointers, vector growth is fine
        vfos.emplace_back();
        idx = (int)vfos.size() - 1;
    }
    Vfo& v = vfos[idx];
    v = Vfo();
    v.live = true; v.outSR = outSR; v.bw = bw; v.offset = offset;
    v.filterNeeded = (bw != outSR);                 // rx_vfo.h:24
    v.angle = 0.0;                                  // phase = 1 + 0i
    configureXlator(v);
    int rc = allocRotator(v);
    if (rc == B200_OK) { rc = reconfigureResampler(v); }
    if (rc == B200_OK) { rc = regenChannelTaps(v, true); }
    if (rc != B200_OK) { removeVfo(idx); return rc; }
    return idx;
}

int VfoBank::removeVfo(int vi) {
    B200_TRY(check(vi));
    DeviceGuard g(device);
    B200_TRY(syncLast());
    Vfo& v = vfos[vi];
    for (auto& st : v.decimStages) { st.destroy(); }
    v.decimStages.clear();
    v.poly.destroy(); v.chan.destroy();
    if (v.dZ) { cudaFree(v.dZ); v.dZ = nullptr; }
    freeRotator(v);
    demodDestroy(v.demod); v.demod = nullptr;
    v.live = false;
    return B200_OK;
}

int VfoBank::resetVfo(Vfo& v) {
    B200_TRY(setPhase(v, 1.0, 0.0));                // FrequencyXlator::reset (frequency_xlator.h:35-41)
    for (auto& st : v.decimStages) { B200_TRY(st.resetState()); }
    if (v.hasPoly) { B200_TRY(v.poly.resetState()); }
    B200_TRY(v.chan.resetState());                  // FIR::reset (fir.h:56-62)
    return B200_OK;
}

int VfoBank::calibrateDirty() {
    std::vector<Vfo*> dirty;
    for (auto& v : vfos) { if (v.live && v.calDirty) { dirty.push_back(&v); } }
    if (dirty.empty()) { return B200_OK; }
    const int steps = calSteps;
    auto work = [&](size_t lo, size_t hi) {
        for (size_t i = lo; i < hi; i++) {
            Vfo* v = dirty[i];
            v->theta = (steps > 0) ? design::calibrateRotatorStep(v->dRe, v->dIm, steps) : atan2((double)v->dIm, (double)v->dRe);
            v->calDirty = false;
        }
    };
    unsigned nt = std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), (unsigned)dirty.size());
    if (nt <= 1) { work(0, dirty.size()); return B200_OK; }
    std::vector<std::thread> th;
    const size_t per = (dirty.size() + nt - 1) / nt;
    for (unsigned t = 0; t < nt; t++) {
        const size_t lo = t * per, hi = std::min(dirty.size(), lo + per);
        if (lo < hi) { th.emplace_back(work, lo, hi); }
    }
    for (auto& t : th) { t.join(); }
    return B200_OK;
}

int VfoBank::allocRotator(Vfo& v) {
    v.chkCap = maxBlock / kRotChunk + 2;
    B200_CUDA(cudaMalloc((void**)&v.dChk, (size_t)v.chkCap * sizeof(float2)));
    B200_CUDA(cudaMalloc((void**)&v.dWtab, kRotChunk * sizeof(float2)));
    B200_CUDA(cudaMalloc((void**)&v.dRotState, sizeof(float2)));
    const float2 one = make_float2(1.0f, 0.0f);     // phase = 1 + 0i (frequency_xlator.h:16)
    B200_CUDA(cudaMemcpy(v.dRotState, &one, sizeof(float2), cudaMemcpyHostToDevice));
    v.tableDirty = true;
    return B200_OK;
}

void VfoBank::freeRotator(Vfo& v) {
    if (v.dChk) { cudaFree(v.dChk); v.dChk = nullptr; }
    if (v.dWtab) { cudaFree(v.dWtab); v.dWtab = nullptr; }
    if (v.dRotState) { cudaFree(v.dRotState); v.dRotState = nullptr; }
}

int VfoBank::getPhase(Vfo& v, double* re, double* im) {
    if (rotatorMode == B200_ROTATOR_EXACT) {
        DeviceGuard g(device);
        B200_TRY(syncLast());
        float2 p;
        B200_CUDA(cudaMemcpy(&p, v.dRotState, sizeof(float2), cudaMemcpyDeviceToHost));
        *re = p.x; *im = p.y;
    }
    else { *re = cos(v.angle); *im = sin(v.angle); }
    return B200_OK;
}

int VfoBank::setPhase(Vfo& v, double re, double im) {
    if (rotatorMode == B200_ROTATOR_EXACT) {
        DeviceGuard g(device);
        B200_TRY(syncLast());
        const float2 p = make_float2((float)re, (float)im);
        B200_CUDA(cudaMemcpy(v.dRotState, &p, sizeof(float2), cudaMemcpyHostToDevice));
    }
    v.angle = atan2(im, re);
    return B200_OK;
}

int VfoBank::setRotatorMode(int mode) {
    if (mode != B200_ROTATOR_FAST && mode != B200_ROTATOR_EXACT) { setError("bad rotator mode %d", mode); return B200_EINVAL; }
    if (mode == rotatorMode) { return B200_OK; }
    // carry every VFO's phase across the switch
    std::vector<std::pair<double, double>> ph(vfos.size());
    for (size_t i = 0; i < vfos.size(); i++) { if (vfos[i].live) { B200_TRY(getPhase(vfos[i], &ph[i].first, &ph[i].second)); } }
    rotatorMode = mode;
    for (size_t i = 0; i < vfos.size(); i++) {
        if (!vfos[i].live) { continue; }
        B200_TRY(setPhase(vfos[i], ph[i].first, ph[i].second));
        B200_TRY(demodSwitchRotatorMode(vfos[i].demod, mode));
    }
    return B200_OK;
}

// RxVFO::process for every live VFO (rx_vfo.h:89-100)
int VfoBank::process(const cf32* dIn, int count, cudaStream_t s, const std::function<int()>* afterStage1) {
    if (count < 0 || count > maxBlock) { setError("block of %d samples exceeds max_block %d", count, maxBlock); return B200_ETOOBIG; }
    DeviceGuard g(device);
    if (everRan && s != lastStream) { B200_TRY(syncLast()); }
    lastStream = s; everRan = true;
    const bool exact = (rotatorMode == B200_ROTATOR_EXACT);
    if (!exact) { B200_TRY(calibrateDirty()); }

    // Deferred tail (B200_OPT_DEFER_TAIL): possible when every VFO's first stage is the fused full-rate kernel, whose
    // output buffer then alternates between two allocations so that block k's tail can still read one while block k+1's
    // first stage fills the other.
    bool defer = deferTail && count > 0;
    if (defer) {
        int live = 0;
        for (auto& v : vfos) {
            if (!v.live) { continue; }
            live++;
            const bool hasChain = ((v.mode == 0 || v.mode == 1) && !v.decimStages.empty());
            if (!hasChain || !fusedStageSupported(v)) { defer = false; break; }
        }
        if (live == 0) { defer = false; }
    }
    const int par = tailPar;
    if (defer) {
        if (!tailStream) {
            int lo = 0, hi = 0;
            B200_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
            // LOWEST priority: the tail is bandwidth-bound filler for what the next block's full-rate kernel leaves free.  With the
            // highest priority (round 1) the tail's kernels went in front of the next block's rotator-checkpoint kernel, which then
            // waited 2.8 ms for an SM (cfg4, 1024 VFOs) and held the whole next block back.  A caller that wants the overlap gives
            // its own stream a higher priority than the default.
            (void)hi;
            B200_CUDA(cudaStreamCreateWithPriority(&tailStream, cudaStreamNonBlocking, lo));
            for (int i = 0; i < 2; i++) { B200_CUDA(cudaEventCreateWithFlags(&evTail[i], cudaEventDisableTiming)); }
        }
        // at most two blocks in flight: this block reuses the first-stage output buffers of block k-2
        if (tailPending[par]) { B200_CUDA(cudaStreamWaitEvent(s, evTail[par], 0)); tailPending[par] = false; }
    }
    else {
        B200_TRY(joinTail());      // an inline tail follows every deferred one
    }

    // ---- build this call's descriptors
    std::vector<RotJob> rj;
    std::vector<XlateJob> xj;
    std::vector<FusedJob> fj;
    std::vector<std::vector<FirJob>> levels;
    std::vector<FirJob> histJobs;
    std::vector<Vfo*> dirtyTables;
    for (auto& v : vfos) {
        if (!v.live) { continue; }
        if (v.tableDirty) { dirtyTables.push_back(&v); }
        RotJob r;
        r.chk = v.dChk; r.state = v.dRotState; r.angle0 = v.angle; r.theta = v.theta;
        r.dRe = v.dRe; r.dIm = v.dIm; r.count = count; r.exact = exact ? 1 : 0;
        rj.push_back(r);

        std::vector<Stage*> chain;
        if (v.mode == 0 || v.mode == 1) { for (auto& st : v.decimStages) { chain.push_back(&st); } }
        if (v.hasPoly && (v.mode == 0 || v.mode == 2)) { chain.push_back(&v.poly); }
        if (v.filterNeeded) { chain.push_back(&v.chan); }

        const void* data = v.dZ;
        int n = count;
        size_t first = 0;
        if (!chain.empty() && fusedStageSupported(v)) {
            // stage 1 consumes the shared input directly, rotation folded in; its history refresh is done
            // by the fused edge kernel, so it is not added to histJobs
            FusedJob f;
            if (defer) { B200_TRY(chain[0]->flipOut()); }
            fillFusedJob(f, *chain[0], v, dIn, count);
            fj.push_back(f);
            n = f.nOut;
            data = chain[0]->dOut;
            chain[0]->advance(count);
            first = 1;
        }
        else {
            if ((!v.dZ || v.zCap < count) && count > 0) {
                if (v.dZ) { B200_CUDA(cudaStreamSynchronize(s)); cudaFree(v.dZ); v.dZ = nullptr; }
                v.zCap = maxBlock;
                B200_CUDA(cudaMalloc((void**)&v.dZ, (size_t)v.zCap * sizeof(cf32)));
                data = v.dZ;
            }
            XlateJob x;
            x.in = dIn; x.out = v.dZ; x.chk = v.dChk; x.wtab = v.dWtab; x.n = count; x.pad = 0;
            xj.push_back(x);
        }
        for (size_t l = first; l < chain.size(); l++) {
            FirJob j; chain[l]->fillJob(j, data, n);
            if (levels.size() <= l) { levels.resize(l + 1); }
            levels[l].push_back(j);
            histJobs.push_back(j);
            chain[l]->advance(n);
            data = chain[l]->dOut; n = j.nOut;
        }
        v.lastOut = n;
        v.lastOutPtr = (const cf32*)data;
        // FAST mode: the angle the reference's phase reaches after `count` recursion steps
        if (!exact) { v.angle = fmod(v.angle + v.theta * (double)count, 2.0 * design::kPi); }
    }
    std::stable_sort(fj.begin(), fj.end(), [](const FusedJob& a, const FusedJob& b) {
        return a.ratio != b.ratio ? a.ratio < b.ratio : a.offset < b.offset;
    });

    // ---- one upload (descriptors + any retuned rotator tables), then one launch per level.  With a deferred tail the
    // descriptors of the later stages go to their own device buffer, uploaded on the tail's stream.
    // (Measured and dropped: sorting a level's jobs by tile count and launching them run by run -- cfg4's first later-stage
    // wave holds 72 % empty CTAs, whose tiles lie beyond their job's end -- 13.32 -> 13.23 ms per device-resident step, but
    // 1066 -> 1020 MS/s end to end, where the extra launches sit between the next chunk's kernels.  firJobsSortByTiles /
    // launchFirJobs still do it for a caller that sorts.)
    for (auto& l : levels) { firJobsBalance(l.data(), (int)l.size()); }
    size_t bytes = rj.size() * sizeof(RotJob) + xj.size() * sizeof(XlateJob) + fj.size() * sizeof(FusedJob) + histJobs.size() * sizeof(FirJob);
    for (auto& l : levels) { bytes += l.size() * sizeof(FirJob); }
    bytes += dirtyTables.size() * (kRotChunk * sizeof(float2)) + 64 * (16 + dirtyTables.size() + levels.size());
    void* hp = nullptr; int slot = 0;
    B200_TRY(ring.acquire(bytes, &hp, &slot));
    char* h = (char*)hp; size_t off = 0;
    auto put = [&](const void* src, size_t n) { size_t o = off; if (n) { memcpy(h + off, src, n); } off = (off + n + 63) & ~(size_t)63; return o; };
    const size_t oR = put(rj.data(), rj.size() * sizeof(RotJob));
    const size_t oX = put(xj.data(), xj.size() * sizeof(XlateJob));
    const size_t oF = put(fj.data(), fj.size() * sizeof(FusedJob));
    const size_t headBytes = off;
    std::vector<size_t> oL;
    for (auto& l : levels) { oL.push_back(put(l.data(), l.size() * sizeof(FirJob))); }
    const size_t oH = put(histJobs.data(), histJobs.size() * sizeof(FirJob));
    const size_t descBytes = off;
    const size_t firstBytes = defer ? headBytes : descBytes;
    if (firstBytes > dDescBytes) {
        B200_TRY(syncLast());
        if (dDesc) { cudaFree(dDesc); }
        dDescBytes = firstBytes * 2;
        B200_CUDA(cudaMalloc(&dDesc, dDescBytes));
    }
    if (defer && descBytes - headBytes > dTailDescBytes) {
        B200_TRY(syncLast());
        if (dTailDesc) { cudaFree(dTailDesc); }
        dTailDescBytes = (descBytes - headBytes) * 2;
        B200_CUDA(cudaMalloc(&dTailDesc, dTailDescBytes));
    }
    B200_TRY(uploadDesc(dDesc, h, firstBytes, s, descViaParams));
    for (Vfo* v : dirtyTables) {
        float2* tab = (float2*)(h + off);
        makeRotatorTable(v->dRe, v->dIm, tab);
        v->hWtab.assign(tab, tab + kRotChunk);
        off += kRotChunk * sizeof(float2);
        B200_CUDA(cudaMemcpyAsync(v->dWtab, tab, kRotChunk * sizeof(float2), cudaMemcpyHostToDevice, s));
        v->tableDirty = false;
    }
    char* d = (char*)dDesc;
    // where the later stages find their descriptors, and the stream they run on
    char* dt = defer ? (char*)dTailDesc : d;
    const size_t tailOff = defer ? headBytes : 0;
    cudaStream_t ts = defer ? tailStream : s;

    int maxH = 0;
    for (auto& j : histJobs) { maxH = std::max(maxH, j.H); }
    if (count > 0) {
        B200_TRY(launchCheckpoints((const RotJob*)(d + oR), (int)rj.size(), exact ? (int)rj.size() : 0, count, s));
        B200_TRY(launchXlate((const XlateJob*)(d + oX), (int)xj.size(), count, s));
        fusedLeaveRoom = defer && fusedLeaveRoomEnabled;      // streaming mode: the previous block's tail runs beside this launch
        const int rcFused = launchFused((const FusedJob*)(d + oF), fj, s);
        fusedLeaveRoom = false;
        B200_TRY(rcFused);
        if (afterStage1) { B200_TRY((*afterStage1)()); afterStage1 = nullptr; }
        B200_CUDA(cudaEventRecord(evStage1, s));
        if (defer) {
            B200_CUDA(cudaStreamWaitEvent(ts, evStage1, 0));
            B200_TRY(uploadDesc(dTailDesc, h + headBytes, descBytes - headBytes, ts, true));
        }
        for (size_t l = 0; l < levels.size(); l++) {
            B200_TRY(launchFirJobs((const FirJob*)(dt + oL[l] - tailOff), levels[l].data(), (int)levels[l].size(), 2, ts));
        }
    }
    else if (!fj.empty()) {
        // an empty block: the fused stages still flip their history ping-pong; copy the history across
        fusedLeaveRoom = defer && fusedLeaveRoomEnabled;      // streaming mode: the previous block's tail runs beside this launch
        const int rcFused = launchFused((const FusedJob*)(d + oF), fj, s);
        fusedLeaveRoom = false;
        B200_TRY(rcFused);
    }
    if (afterStage1) { B200_TRY((*afterStage1)()); }
    B200_TRY(launchHistoryUpdate((const FirJob*)(dt + oH - tailOff), (int)histJobs.size(), maxH, ts));
    B200_TRY(ring.release(slot, ts));      // after the last reader of the slot was enqueued
    B200_TRY(processDemods(*this, ts));
    if (defer) {
        B200_CUDA(cudaEventRecord(evTail[par], ts));
        tailPending[par] = true; lastTailPar = par; tailPar = par ^ 1; tailUnjoined = true;
    }
    return B200_OK;
}

} // namespace b200

// extern "C" surface of the RxVFO bank (include/b200dsp.h).
#include <cmath>
#include <cstring>

#include <mutex>

#include "vfo_bank.cuh"
#include "fft.cuh"
#include "xlate_decim.cuh"

using namespace b200;

#define BANK_OR_FAIL(b) \
    if (!(b)) { setError("NULL bank handle"); return B200_EINVAL; }

extern "C" {

int b200_vfo_bank_create(b200_vfo_bank** out, double in_samplerate, int max_block, int device) {
    if (!out || in_samplerate <= 0 || max_block <= 0) { setError("bad arguments to b200_vfo_bank_create"); return B200_EINVAL; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        setError("no CUDA device available: libb200dsp has no CPU fallback");
        return B200_ECUDA;
    }
    if (device < 0 || device >= ndev) { setError("device %d out of range (%d devices)", device, ndev); return B200_EINVAL; }
    auto* b = new (std::nothrow) b200_vfo_bank;
    if (!b) { return B200_ENOMEM; }
    int rc = b->impl.init(in_samplerate, max_block, device);
    if (rc != B200_OK) { delete b; return rc; }
    *out = b;
    return B200_OK;
}

int b200_vfo_bank_destroy(b200_vfo_bank* bank) {
    BANK_OR_FAIL(bank);
    if (bank->ownedByFrontend) { setError("bank is owned by a front end"); return B200_ESTATE; }
    bank->impl.destroy();
    delete bank;
    return B200_OK;
}

int b200_vfo_bank_add(b200_vfo_bank* bank, double out_samplerate, double bandwidth, double offset) {
    BANK_OR_FAIL(bank);
    if (out_samplerate <= 0 || bandwidth <= 0) { setError("bad VFO parameters"); return B200_EINVAL; }
    return bank->impl.addVfo(out_samplerate, bandwidth, offset);
}

int b200_vfo_bank_remove(b200_vfo_bank* bank, int vfo) { BANK_OR_FAIL(bank); return bank->impl.removeVfo(vfo); }
int b200_vfo_bank_count(b200_vfo_bank* bank) { BANK_OR_FAIL(bank); return bank->impl.liveCount(); }

int b200_vfo_set_offset(b200_vfo_bank* bank, int vfo, double offset) {
    BANK_OR_FAIL(bank);
    B200_TRY(bank->impl.check(vfo));
    Vfo& v = bank->impl.vfos[vfo];
    v.offset = offset;
    return bank->impl.configureXlator(v);   // phase stays continuous, only the step changes (frequency_xlator.h:25-29)
}

int b200_vfo_set_bandwidth(b200_vfo_bank* bank, int vfo, double bandwidth) {
    BANK_OR_FAIL(bank);
    B200_TRY(bank->impl.check(vfo));
    if (bandwidth <= 0) { setError("bad bandwidth"); return B200_EINVAL; }
    DeviceGuard g(bank->impl.device);
    B200_TRY(bank->impl.syncLast());
    Vfo& v = bank->impl.vfos[vfo];
    v.bw = bandwidth;
    v.filterNeeded = (v.bw != v.outSR);
    if (v.filterNeeded) { B200_TRY(bank->impl.regenChannelTaps(v, false)); }
    return B200_OK;
}

int b200_vfo_set_out_samplerate(b200_vfo_bank* bank, int vfo, double out_sr, double bandwidth) {
    BANK_OR_FAIL(bank);
    B200_TRY(bank->impl.check(vfo));
    if (out_sr <= 0 || bandwidth <= 0) { setError("bad rate/bandwidth"); return B200_EINVAL; }
    DeviceGuard g(bank->impl.device);
    B200_TRY(bank->impl.syncLast());
    Vfo& v = bank->impl.vfos[vfo];
    v.outSR = out_sr; v.bw = bandwidth;
    v.filterNeeded = (v.bw != v.outSR);
    B200_TRY(bank->impl.reconfigureResampler(v));
    if (v.filterNeeded) { B200_TRY(bank->impl.regenChannelTaps(v, false)); }
    return B200_OK;
}

int b200_vfo_bank_set_in_samplerate(b200_vfo_bank* bank, double in_samplerate) {
    BANK_OR_FAIL(bank);
    if (in_samplerate <= 0) { setError("bad sample rate"); return B200_EINVAL; }
    DeviceGuard g(bank->impl.device);
    B200_TRY(bank->impl.syncLast());
    bank->impl.inSR = in_samplerate;
    for (auto& v : bank->impl.vfos) {
        if (!v.live) { continue; }
        bank->impl.configureXlator(v);
        B200_TRY(bank->impl.reconfigureResampler(v));
    }
    return B200_OK;
}

int b200_vfo_reset(b200_vfo_bank* bank, int vfo) {
    BANK_OR_FAIL(bank);
    B200_TRY(bank->impl.check(vfo));
    DeviceGuard g(bank->impl.device);
    B200_TRY(bank->impl.syncLast());
    return bank->impl.resetVfo(bank->impl.vfos[vfo]);
}

int b200_vfo_get_phase(b200_vfo_bank* bank, int vfo, double* re, double* im) {
    BANK_OR_FAIL(bank);
    B200_TRY(bank->impl.check(vfo));
    if (!re || !im) { return B200_EINVAL; }
    return bank->impl.getPhase(bank->impl.vfos[vfo], re, im);
}

int b200_vfo_set_phase(b200_vfo_bank* bank, int vfo, double re, double im) {
    BANK_OR_FAIL(bank);
    B200_TRY(bank->impl.check(vfo));
    if (re == 0.0 && im == 0.0) { setError("zero phasor"); return B200_EINVAL; }
    return bank->impl.setPhase(bank->impl.vfos[vfo], re, im);
}

int b200_vfo_bank_set_rotator_mode(b200_vfo_bank* bank, int mode) {
    BANK_OR_FAIL(bank);
    return bank->impl.setRotatorMode(mode);
}

// Process-wide tuning switches (they select kernels / schedules, never results).  One mutex orders concurrent callers; a
// switch that changes how the next block is scheduled takes effect at each handle's next process call.
static std::mutex g_optMtx;

int b200_set_global_option(int option, int value) {
    std::lock_guard<std::mutex> lk(g_optMtx);
    switch (option) {
    case B200_OPT_FUSED_STAGE1: fusedEnabled = value != 0; return B200_OK;
    case B200_OPT_TILE_SAMPLES:
        if (value != 0 && value != 2048 && value != 4096) { setError("tile must be 0 (automatic), 2048 or 4096"); return B200_EINVAL; }
        fusedTileSamples = value;
        return B200_OK;
    case B200_OPT_VFOS_PER_PASS:
        if (value != 0 && value != 2 && value != 4) { setError("VFOs per register pass must be 0 (automatic), 2 or 4"); return B200_EINVAL; }
        fusedVfosPerPass = value;
        return B200_OK;
    case B200_OPT_SPLIT_RUNS: fusedSplitRuns = value != 0; return B200_OK;
    case B200_OPT_LEAVE_ROOM: fusedLeaveRoomEnabled = value != 0; return B200_OK;
    case B200_OPT_PARAM_TABLES: fusedParamTables = value != 0; return B200_OK;
    case B200_OPT_VFOS_PER_CTA:
        if (value < 2 || value > 4096) { setError("VFOs per CTA must be 2 .. 4096"); return B200_EINVAL; }
        fusedVfosPerCta = value;
        return B200_OK;
    case B200_OPT_FFT_DIRECT_MAX_LOG:
        if (value < 9 || value > 14) { setError("one-pass FFT limit must be 9 (off) .. 14"); return B200_EINVAL; }
        fftDirectMaxLog = value;
        return B200_OK;
    case B200_OPT_FFT_CTAS_PER_SM:
        if (value < 0 || value > 4) { setError("FFT CTAs per SM must be 0 (automatic) .. 4"); return B200_EINVAL; }
        fftCtasPerSm = value;
        return B200_OK;
    case B200_OPT_BRANCH_ORDER:
        if (value != 0 && value != 1) { setError("branch order must be 0 or 1"); return B200_EINVAL; }
        branchOrder = value;
        return B200_OK;
    case B200_OPT_FFT_DIRECT_PREFETCH: fftDirectPrefetch = value != 0; return B200_OK;
    case B200_OPT_HOST_CHUNK_MIB:
        if (value < 0 || value > 1024) { setError("host chunk must be 0 (automatic) .. 1024 MiB"); return B200_EINVAL; }
        hostChunkMiB = value;
        return B200_OK;
    default:
        setError("option %d is not process-wide (or unknown)", option);
        return B200_EINVAL;
    }
}

int b200_vfo_bank_set_option(b200_vfo_bank* bank, int option, int value) {
    BANK_OR_FAIL(bank);
    // per-handle switches
    if (option == B200_OPT_CALIBRATION_STEPS) {
        if (value < 0) { setError("bad calibration length"); return B200_EINVAL; }
        bank->impl.calSteps = value;
        for (auto& v : bank->impl.vfos) { v.calDirty = true; }
        return B200_OK;
    }
    if (option == B200_OPT_DEFER_TAIL) {
        // the mode itself is process-wide, but switching it needs THIS bank's deferred work drained first; other banks
        // pick the new mode up at their next process call (each drains its own tail when it finds the mode changed)
        B200_TRY(bank->impl.syncLast());
        std::lock_guard<std::mutex> lk(g_optMtx);
        deferTail = value != 0;
        return B200_OK;
    }
    // everything else is process-wide: same effect as b200_set_global_option, whichever handle is named
    return b200_set_global_option(option, value);
}

int b200_vfo_describe(b200_vfo_bank* bank, int vfo, int* info) {
    BANK_OR_FAIL(bank);
    B200_TRY(bank->impl.check(vfo));
    if (!info) { return B200_EINVAL; }
    const Vfo& v = bank->impl.vfos[vfo];
    const bool hasDecim = (v.mode == 0 || v.mode == 1), hasRes = (v.mode == 0 || v.mode == 2);
    info[0] = v.mode;
    info[1] = hasDecim ? v.ratio : 1;
    info[2] = hasRes ? v.interp : 1;
    info[3] = hasRes ? v.decim : 1;
    info[4] = hasRes ? v.resampTapCount : 0;
    info[5] = hasRes ? v.poly.T : 0;
    info[6] = v.filterNeeded ? v.chan.T : 0;
    info[7] = (hasDecim && !v.decimStages.empty() && fusedStageSupported(v)) ? 1 : 0;
    return B200_OK;
}

int b200_vfo_get_taps(b200_vfo_bank* bank, int vfo, int which, float* out, int cap) {
    BANK_OR_FAIL(bank);
    B200_TRY(bank->impl.check(vfo));
    const Vfo& v = bank->impl.vfos[vfo];
    const std::vector<float>* t = nullptr;
    if (which == 0) { t = &v.resampTaps; }
    else if (which == 1) { t = &v.chan.hostTaps; }
    else if (which >= 10 && which - 10 < (int)v.decimStages.size()) { t = &v.decimStages[which - 10].hostTaps; }
    if (!t) { setError("bad tap table selector %d", which); return B200_EINVAL; }
    if (out && cap > 0) { memcpy(out, t->data(), sizeof(float) * std::min<size_t>(t->size(), (size_t)cap)); }
    return (int)t->size();
}

int b200_vfo_bank_process(b200_vfo_bank* bank, const float* d_in, int count, void* stream) {
    BANK_OR_FAIL(bank);
    if (!d_in && count > 0) { setError("NULL input"); return B200_EINVAL; }
    return bank->impl.process((const cf32*)d_in, count, (cudaStream_t)stream);
}

int b200_vfo_out_count(b200_vfo_bank* bank, int vfo) {
    BANK_OR_FAIL(bank);
    B200_TRY(bank->impl.check(vfo));
    return bank->impl.vfos[vfo].lastOut;
}

const float* b200_vfo_out_ptr(b200_vfo_bank* bank, int vfo) {
    if (!bank || bank->impl.check(vfo) != B200_OK) { return nullptr; }
    if (bank->impl.tailUnjoined) { DeviceGuard g(bank->impl.device); if (bank->impl.joinTail() != B200_OK) { return nullptr; } }
    return (const float*)bank->impl.vfos[vfo].lastOutPtr;
}

int b200_vfo_max_out(b200_vfo_bank* bank, int vfo, int count) {
    BANK_OR_FAIL(bank);
    B200_TRY(bank->impl.check(vfo));
    return bank->impl.maxOut(vfo, count) + 4;
}

int b200_vfo_bank_process_host(b200_vfo_bank* bank, const float* h_in, int count, float* const* h_out, int* counts, void* stream) {
    BANK_OR_FAIL(bank);
    VfoBank& b = bank->impl;
    if (count < 0 || count > b.maxBlock) { setError("block of %d samples exceeds max_block %d", count, b.maxBlock); return B200_ETOOBIG; }
    if (!h_in && count > 0) { setError("NULL input"); return B200_EINVAL; }
    DeviceGuard g(b.device);
    cudaStream_t s = (cudaStream_t)stream;
    if (!b.dStageIn) {
        b.dStageInCap = (size_t)b.maxBlock;
        B200_CUDA(cudaMalloc((void**)&b.dStageIn, b.dStageInCap * sizeof(cf32)));
    }
    // the caller's buffer may be pageable; cudaMemcpyAsync then stages internally (dsp::stream buffers of
    // include/dsp/stream.h are pinned, so the drop-in path copies at full PCIe rate)
    B200_CUDA(cudaMemcpyAsync(b.dStageIn, h_in, (size_t)count * sizeof(cf32), cudaMemcpyHostToDevice, s));
    B200_TRY(b.process(b.dStageIn, count, s));
    B200_TRY(b.joinTail());
    for (size_t i = 0; i < b.vfos.size(); i++) {
        const Vfo& v = b.vfos[i];
        if (!v.live) { continue; }
        if (counts) { counts[i] = v.lastOut; }
        if (h_out && h_out[i] && v.lastOut > 0) {
            B200_CUDA(cudaMemcpyAsync(h_out[i], v.lastOutPtr, (size_t)v.lastOut * sizeof(cf32), cudaMemcpyDeviceToHost, s));
        }
    }
    B200_CUDA(cudaStreamSynchronize(s));
    return B200_OK;
}

} // extern "C"

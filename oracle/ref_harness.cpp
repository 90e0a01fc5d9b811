// TEST INFRASTRUCTURE ONLY (oracle/_ref).  Never linked into, or loaded by, the product.
//
// Instantiates the reference's OWN header-only DSP blocks, compiled in place from
// $REF_ROOT/core/src (default /root/reference), and exposes them through
// oracle_api.h.  The only stand-ins are oracle/shim/{volk/volk.h,core.h,
// sdrpp_export.h,utils/flog.h}; no reference source is copied into this repo.
// iq_frontend.cpp / arrays.cpp / waterfall.cpp cannot be compiled here (GL, FFTW,
// ImGui); those few functions come from frontend_restated.h.
//
// Build: see oracle/Makefile (g++ -O2 -ffp-contract=off -std=c++17).

// Standard headers first, so that the access-specifier override below only
// affects the reference's headers.
#include <algorithm>
#include <atomic>
#include <cassert>
#include <cmath>
#include <complex>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <functional>
#include <map>
#include <mutex>
#include <numeric>
#include <string>
#include <thread>
#include <vector>
#include <typeinfo>

// The harness needs FrequencyXlator::phase, RationalResampler::mode, RxVFO::ftaps ...
// which the reference keeps protected.  Test-only access, no behavioural change.
#define protected public
#define private public
#include <dsp/channel/rx_vfo.h>
#include <dsp/multirate/power_decimator.h>
#include <dsp/correction/dc_blocker.h>
#include <dsp/math/conjugate.h>
#include <dsp/window/blackman.h>
#include <dsp/window/nuttall.h>
#include <dsp/taps/low_pass.h>
#include <dsp/demod/fm.h>
#include <dsp/demod/ssb.h>
#include <dsp/demod/am.h>
#include <dsp/loop/agc.h>
#include <dsp/filter/deephasis.h>
#include <dsp/demod/broadcast_fm.h>
#include <dsp/noise_reduction/noise_blanker.h>
#include <dsp/noise_reduction/squelch.h>
#include <dsp/noise_reduction/fm_if.h>
#include <dsp/audio/volume.h>
#include <utils/wav.h>
#include <dsp/compression/sample_stream_compressor.h>
#include <dsp/compression/sample_stream_decompressor.h>
#include <dsp/routing/splitter.h>
#include <dsp/sink/null_sink.h>
#include <chrono>
#include <memory>
#undef protected
#undef private

#include "frontend_restated.h"
#include "oracle_api.h"

using dsp::complex_t;
using dsp::stereo_t;

// dsp/stream.h:16 declares it; defined in the reference's core (not on the path)
void logDebugMessage(const char*) {}

namespace {
struct Preproc {
    dsp::multirate::PowerDecimator<complex_t> decim;
    dsp::correction::DCBlocker<complex_t> dcBlock;
    dsp::math::Conjugate conjugate;
    bool useDecim, useDc, useConj;
    std::vector<complex_t> tmp;
};
}

extern "C" {

const char* orc_kind(void) { return "reference"; }

// The graph the reference runs for a bank of VFOs: Splitter -> per-VFO stream -> RxVFO thread -> consumer
// (iq_frontend.cpp:159-180; the consumer here is a Null sink).  Timing only.
double orc_graph_run(double inSR, int nVfos, const double* outSR, const double* bandwidth, const double* offset,
                     int blockLen, int nBlocks, const float* block) {
    if (nVfos <= 0 || blockLen <= 0 || blockLen > STREAM_BUFFER_SIZE || nBlocks <= 0) { return -1.0; }
    dsp::stream<dsp::complex_t> src;
    dsp::routing::Splitter<dsp::complex_t> split;
    split.init(&src);
    std::vector<std::unique_ptr<dsp::stream<dsp::complex_t>>> taps;
    std::vector<std::unique_ptr<dsp::channel::RxVFO>> vfos;
    std::vector<std::unique_ptr<dsp::sink::Null<dsp::complex_t>>> sinks;
    for (int i = 0; i < nVfos; i++) {
        taps.emplace_back(new dsp::stream<dsp::complex_t>());
        split.bindStream(taps.back().get());
        vfos.emplace_back(new dsp::channel::RxVFO(taps.back().get(), inSR, outSR[i], bandwidth[i], offset[i]));
        sinks.emplace_back(new dsp::sink::Null<dsp::complex_t>());
        sinks.back()->init(&vfos.back()->out);
    }
    for (auto& s : sinks) { s->start(); }
    for (auto& v : vfos) { v->start(); }
    split.start();
    auto feed = [&](int n) {
        memcpy(src.writeBuf, block, (size_t)n * sizeof(dsp::complex_t));
        return src.swap(n);
    };
    feed(blockLen);                      // one block untimed: threads up, buffers touched
    const auto t0 = std::chrono::steady_clock::now();
    for (int b = 0; b < nBlocks; b++) { feed(blockLen); }
    // a swap returns once the reader has flushed the block before: after two more (small) blocks have gone through the
    // Splitter, every VFO has flushed -- i.e. processed -- the last timed block
    feed(16);
    feed(16);
    const auto t1 = std::chrono::steady_clock::now();
    split.stop();
    for (auto& v : vfos) { v->stop(); }
    for (auto& s : sinks) { s->stop(); }
    return std::chrono::duration<double>(t1 - t0).count();
}

// ---------------------------------------------------------------- RxVFO
void* orc_vfo_create(double inSR, double outSR, double bandwidth, double offset) {
    return new dsp::channel::RxVFO(NULL, inSR, outSR, bandwidth, offset);
}
void orc_vfo_destroy(void* h) { delete (dsp::channel::RxVFO*)h; }
int orc_vfo_process(void* h, int count, const float* in, float* out) {
    return ((dsp::channel::RxVFO*)h)->process(count, (const complex_t*)in, (complex_t*)out);
}
void orc_vfo_set_offset(void* h, double offset) { ((dsp::channel::RxVFO*)h)->setOffset(offset); }
void orc_vfo_set_bandwidth(void* h, double bw) { ((dsp::channel::RxVFO*)h)->setBandwidth(bw); }
void orc_vfo_set_in_samplerate(void* h, double sr) { ((dsp::channel::RxVFO*)h)->setInSamplerate(sr); }
void orc_vfo_set_out_samplerate(void* h, double sr, double bw) { ((dsp::channel::RxVFO*)h)->setOutSamplerate(sr, bw); }
void orc_vfo_reset(void* h) { ((dsp::channel::RxVFO*)h)->reset(); }
void orc_vfo_get_phase(void* h, float* reim) {
    auto* v = (dsp::channel::RxVFO*)h;
    reim[0] = v->xlator.phase.real(); reim[1] = v->xlator.phase.imag();
}
void orc_vfo_set_phase(void* h, float re, float im) { ((dsp::channel::RxVFO*)h)->xlator.phase = lv_cmake(re, im); }
void orc_vfo_get_phase_delta(void* h, float* reim) {
    auto* v = (dsp::channel::RxVFO*)h;
    reim[0] = v->xlator.phaseDelta.real(); reim[1] = v->xlator.phaseDelta.imag();
}
void orc_vfo_describe(void* h, int* info) {
    auto* v = (dsp::channel::RxVFO*)h;
    auto& r = v->resamp;
    const bool hasDecim = (r.mode == r.BOTH || r.mode == r.DECIM_ONLY);
    const bool hasResamp = (r.mode == r.BOTH || r.mode == r.RESAMP_ONLY);
    info[0] = (int)r.mode;
    info[1] = hasDecim ? (int)r.decim._ratio : 1;
    info[2] = hasResamp ? r.resamp._interp : 1;
    info[3] = hasResamp ? r.resamp._decim : 1;
    info[4] = hasResamp ? (int)r.rtaps.size : 0;
    info[5] = hasResamp ? r.resamp.phases.tapsPerPhase : 0;
    info[6] = v->filterNeeded ? (int)v->ftaps.size : 0;
}
int orc_vfo_get_resamp_taps(void* h, float* out, int cap) {
    auto* v = (dsp::channel::RxVFO*)h;
    int n = (int)v->resamp.rtaps.size;
    if (out) { memcpy(out, v->resamp.rtaps.taps, sizeof(float) * std::min(n, cap)); }
    return n;
}
int orc_vfo_get_chan_taps(void* h, float* out, int cap) {
    auto* v = (dsp::channel::RxVFO*)h;
    int n = (int)v->ftaps.size;
    if (out) { memcpy(out, v->ftaps.taps, sizeof(float) * std::min(n, cap)); }
    return n;
}

// ---------------------------------------------------------------- PowerDecimator
void* orc_decim_create(int ratio) {
    auto* d = new dsp::multirate::PowerDecimator<complex_t>(NULL, (unsigned)ratio);
    return d;
}
void orc_decim_destroy(void* h) { delete (dsp::multirate::PowerDecimator<complex_t>*)h; }
int orc_decim_process(void* h, int count, const float* in, float* out) {
    return ((dsp::multirate::PowerDecimator<complex_t>*)h)->process(count, (const complex_t*)in, (complex_t*)out);
}
int orc_decim_plan(int ratio, int* decims, int* tapCounts, int cap) {
    if (ratio < 2) { return 0; }
    int planId = (int)log2(ratio) - 1;   // multirate/power_decimator.h:97
    const auto& plan = dsp::multirate::decim::plans[planId];
    for (unsigned i = 0; i < plan.stageCount && (int)i < cap; i++) {
        decims[i] = plan.stages[i].decimation;
        tapCounts[i] = plan.stages[i].tapcount;
    }
    return plan.stageCount;
}
int orc_decim_stage_taps(int ratio, int stage, float* out, int cap) {
    int planId = (int)log2(ratio) - 1;
    const auto& st = dsp::multirate::decim::plans[planId].stages[stage];
    if (out) { memcpy(out, st.taps, sizeof(float) * std::min<int>(st.tapcount, cap)); }
    return st.tapcount;
}

// ---------------------------------------------------------------- DCBlocker
void* orc_dcblock_create(double rate) { return new dsp::correction::DCBlocker<complex_t>(NULL, rate); }
void orc_dcblock_destroy(void* h) { delete (dsp::correction::DCBlocker<complex_t>*)h; }
int orc_dcblock_process(void* h, int count, const float* in, float* out) {
    return ((dsp::correction::DCBlocker<complex_t>*)h)->process(count, (complex_t*)in, (complex_t*)out);
}

// ---------------------------------------------------------------- pre-processing chain
void* orc_preproc_create(double sampleRate, int decimRatio, int dcBlocking, int invertIQ) {
    auto* p = new Preproc;
    // same order and parameters as IQFrontEnd::init, iq_frontend.cpp:29-44
    double effectiveSr = sampleRate / decimRatio;
    p->decim.init(NULL, decimRatio);
    p->dcBlock.init(NULL, 50.0 / effectiveSr);   // genDCBlockRate, iq_frontend.h:84-86
    p->conjugate.init(NULL);
    p->useDecim = decimRatio > 1;
    p->useDc = dcBlocking != 0;
    p->useConj = invertIQ != 0;
    p->tmp.resize(STREAM_BUFFER_SIZE);
    return p;
}
void orc_preproc_destroy(void* h) { delete (Preproc*)h; }
int orc_preproc_process(void* h, int count, const float* in, float* out) {
    auto* p = (Preproc*)h;
    complex_t* o = (complex_t*)out;
    memcpy(o, in, sizeof(complex_t) * count);
    if (p->useDecim) {
        count = p->decim.process(count, o, p->tmp.data());
        memcpy(o, p->tmp.data(), sizeof(complex_t) * count);
    }
    if (p->useDc) { p->dcBlock.process(count, o, o); }
    if (p->useConj) { p->conjugate.process(count, o, o); }
    return count;
}

// ---------------------------------------------------------------- taps & windows
int orc_lowpass_taps(double cutoff, double transWidth, double sampleRate, float* out, int cap) {
    dsp::tap<float> t = dsp::taps::lowPass(cutoff, transWidth, sampleRate);
    int n = (int)t.size;
    if (out) { memcpy(out, t.taps, sizeof(float) * std::min(n, cap)); }
    dsp::taps::free(t);
    return n;
}
void orc_fft_window(int type, int nz, float* out) {
    // IQFrontEnd::updateFFTPath, iq_frontend.cpp:303-311 (same expressions, same types)
    if (type == 0) {
        for (int i = 0; i < nz; i++) { out[i] = 1.0f * ((i % 2) ? -1.0f : 1.0f); }
    }
    else if (type == 1) {
        for (int i = 0; i < nz; i++) { out[i] = dsp::window::blackman(i, nz) * ((i % 2) ? -1.0f : 1.0f); }
    }
    else {
        for (int i = 0; i < nz; i++) { out[i] = dsp::window::nuttall(i, nz) * ((i % 2) ? -1.0f : 1.0f); }
    }
}
void orc_reshape_params(double sampleRate, int fftSize, double fftRate, int* skip, int* nz) {
    orc_restated::reshapeParams(sampleRate, fftSize, fftRate, skip, nz);
}

// ---------------------------------------------------------------- FFT frame + waterfall math
void orc_fft_frame(const float* iq, const float* window, int nz, int N, float* outDb, float* spectrum) {
    orc_restated::fftFrame(iq, window, nz, N, outDb, spectrum,
        [](float* o, const float* X, float norm, int n) { volk_32fc_s32f_power_spectrum_32f(o, (const lv_32fc_t*)X, norm, n); });
}
void orc_do_zoom(int offset, int width, int inSize, int outSize, const float* in, float* out) {
    orc_restated::doZoom(offset, width, inSize, outSize, in, out);
}
void orc_zoom_indices(int offset, int width, int inSize, int outSize, int* sId, int* uFactor) {
    orc_restated::zoomIndices(offset, width, inSize, outSize, sId, uFactor);
}
void orc_palette_indices(const float* row, int n, float wfMin, float wfMax, int* idx) {
    orc_restated::paletteIndices(row, n, wfMin, wfMax, idx);
}
int orc_vfo_signal_info(const float* row, int rawFFTSize, double wholeBandwidth, double centerOffset, double bandwidth, float* strength, float* snr) {
    return orc_restated::vfoSignalInfo(row, rawFFTSize, wholeBandwidth, centerOffset, bandwidth, strength, snr) ? 1 : 0;
}
void orc_smooth_hold(float* latest, float* smoothingBuf, float* hold, int width, int smoothing, float alpha, int doHold, float holdSpeed) {
    orc_restated::smoothAndHold(latest, smoothingBuf, hold, width, smoothing != 0, alpha, doHold != 0, holdSpeed);
}
void orc_make_palette(const float* colors, int colorCount, unsigned int* lut) {
    orc_restated::makePalette(colors, colorCount, lut);
}

// ---------------------------------------------------------------- demodulators
void* orc_fm_create(double samplerate, double bandwidth, int lowPass, int highPass) {
    auto* d = new dsp::demod::FM<stereo_t>();
    d->init(NULL, samplerate, bandwidth, lowPass != 0, highPass != 0);
    return d;
}
void orc_fm_destroy(void* h) { delete (dsp::demod::FM<stereo_t>*)h; }
int orc_fm_process(void* h, int count, const float* in, float* out) {
    return ((dsp::demod::FM<stereo_t>*)h)->process(count, (complex_t*)in, (stereo_t*)out);
}
void* orc_ssb_create(int mode, double bandwidth, double samplerate, double agcAttack, double agcDecay) {
    auto* d = new dsp::demod::SSB<stereo_t>();
    d->init(NULL, (dsp::demod::SSB<stereo_t>::Mode)mode, bandwidth, samplerate, agcAttack, agcDecay);
    return d;
}
void orc_ssb_destroy(void* h) { delete (dsp::demod::SSB<stereo_t>*)h; }
int orc_ssb_process(void* h, int count, const float* in, float* out) {
    return ((dsp::demod::SSB<stereo_t>*)h)->process(count, (const complex_t*)in, (stereo_t*)out);
}
void* orc_am_create(int agcMode, double bandwidth, double agcAttack, double agcDecay, double dcBlockRate, double samplerate) {
    auto* d = new dsp::demod::AM<stereo_t>();
    d->init(NULL, (dsp::demod::AM<stereo_t>::AGCMode)agcMode, bandwidth, agcAttack, agcDecay, dcBlockRate, samplerate);
    return d;
}
void orc_am_destroy(void* h) { delete (dsp::demod::AM<stereo_t>*)h; }
int orc_am_process(void* h, int count, const float* in, float* out) {
    return ((dsp::demod::AM<stereo_t>*)h)->process(count, (complex_t*)in, (stereo_t*)out);
}
void* orc_wfm_create(double deviation, double samplerate, int stereo, int lowPass) {
    auto* d = new dsp::demod::BroadcastFM;
    d->init(NULL, deviation, samplerate, stereo != 0, lowPass != 0, false);
    return d;
}
void orc_wfm_destroy(void* h) { delete (dsp::demod::BroadcastFM*)h; }
int orc_wfm_process(void* h, int count, const float* in, float* out) {
    int rdsCount = 0;
    return ((dsp::demod::BroadcastFM*)h)->process(count, (complex_t*)in, (stereo_t*)out, rdsCount, NULL);
}
void* orc_wfm_create_rds(double deviation, double samplerate, int stereo, int lowPass) {
    auto* d = new dsp::demod::BroadcastFM;
    d->init(NULL, deviation, samplerate, stereo != 0, lowPass != 0, true);
    return d;
}
int orc_wfm_process_rds(void* h, int count, const float* in, float* out, float* rdsOut) {
    int rdsCount = 0;
    ((dsp::demod::BroadcastFM*)h)->process(count, (complex_t*)in, (stereo_t*)out, rdsCount, (complex_t*)rdsOut);
    return rdsCount;
}
void* orc_fmif_create(int bins) {
    auto* f = new dsp::noise_reduction::FMIF;
    f->init(NULL, bins);
    return f;
}
void orc_fmif_destroy(void* h) { delete (dsp::noise_reduction::FMIF*)h; }
int orc_fmif_process(void* h, int count, const float* in, float* out) {
    return ((dsp::noise_reduction::FMIF*)h)->process(count, (const complex_t*)in, (complex_t*)out);
}
int orc_volume_process(double volume, int muted, int count, const float* in, float* out) {
    dsp::audio::Volume v;
    v.init(NULL, volume, muted != 0);
    return v.process(count, (const stereo_t*)in, (stereo_t*)out);
}
int orc_stream_compress(int count, int pcmType, const float* in, unsigned char* out) {
    return dsp::compression::SampleStreamCompressor::process(count, (dsp::compression::PCMType)pcmType, (const complex_t*)in, out);
}
int orc_stream_decompress(int nbytes, const unsigned char* in, float* out) {
    return dsp::compression::SampleStreamDecompressor::process(nbytes, in, (complex_t*)out);
}
int orc_wav_write(const char* path, int channels, double samplerate, int sampleType, const float* samples, int count, int blockLen) {
    wav::Writer w(channels, (uint64_t)samplerate, wav::FORMAT_WAV, (wav::SampleType)sampleType);
    if (!w.open(path)) { return -1; }
    for (int i = 0; i < count; i += blockLen) {
        w.write((float*)samples + (size_t)i * channels, std::min(blockLen, count - i));
    }
    w.close();
    return 0;
}
namespace { struct IfChain { dsp::noise_reduction::NoiseBlanker nb; dsp::noise_reduction::Squelch sq; bool nbOn, sqOn; }; }
void* orc_ifchain_create(int nbOn, double nbRate, double nbLevel, int squelchOn, double squelchLevel) {
    auto* c = new IfChain;
    c->nb.init(NULL, nbRate, nbLevel);
    c->sq.init(NULL, squelchLevel);
    c->nbOn = nbOn != 0; c->sqOn = squelchOn != 0;
    return c;
}
void orc_ifchain_destroy(void* h) { delete (IfChain*)h; }
int orc_ifchain_process(void* h, int count, const float* in, float* out) {
    auto* c = (IfChain*)h;
    memcpy(out, in, (size_t)count * sizeof(complex_t));
    if (c->nbOn) { c->nb.process(count, (complex_t*)out, (complex_t*)out); }
    if (c->sqOn && count > 0) { c->sq.process(count, (const complex_t*)out, (complex_t*)out); }
    return count;
}
namespace { struct AfPost { dsp::multirate::RationalResampler<stereo_t> resamp; dsp::filter::Deemphasis<stereo_t> deemp; bool deOn; }; }
void* orc_afpost_create(double inSamplerate, double audioSamplerate, double tau) {
    auto* a = new AfPost;
    a->resamp.init(NULL, inSamplerate, audioSamplerate);
    a->deOn = tau > 0;
    a->deemp.init(NULL, tau > 0 ? tau : 50e-6, audioSamplerate);
    return a;
}
void orc_afpost_destroy(void* h) { delete (AfPost*)h; }
int orc_afpost_process(void* h, int count, const float* in, float* out) {
    auto* a = (AfPost*)h;
    int n = a->resamp.process(count, (const stereo_t*)in, (stereo_t*)out);
    if (a->deOn && n > 0) { a->deemp.process(n, (stereo_t*)out, (stereo_t*)out); }
    return n;
}
void* orc_agc_create(double setPoint, double attack, double decay, double maxGain, double maxOutputAmp, double initGain) {
    auto* a = new dsp::loop::AGC<float>();
    a->init(NULL, setPoint, attack, decay, maxGain, maxOutputAmp, initGain);
    return a;
}
void orc_agc_destroy(void* h) { delete (dsp::loop::AGC<float>*)h; }
int orc_agc_process(void* h, int count, const float* in, float* out) {
    return ((dsp::loop::AGC<float>*)h)->process(count, (float*)in, out);
}

} // extern "C"

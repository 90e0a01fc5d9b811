// TEST INFRASTRUCTURE ONLY (oracle, reference side).  Part of oracle/_ref/libref_oracle.so.
//
// The reference's own logmmse IF noise reduction, compiled from the sources where they lie:
//   $REF_ROOT/misc_modules/noise_reduction_logmmse/src/if_nr.h     dsp::IFNRLogMMSE, the processor the module adds to the
//                                                                  front end's pre-processing chain (main.cpp:165)
//   $REF_ROOT/misc_modules/noise_reduction_logmmse/src/logmmse.h   LogMMSE::logmmse_sample / logmmse_all / SavedParamsC
//   $REF_ROOT/misc_modules/noise_reduction_logmmse/src/bgnoise.h   BackgroundNoiseCaltulator
//   $REF_ROOT/core/src/utils/arrays.cpp                            array helpers, FFT plans, the E1 table (linked by oracle/Makefile)
// if_nr.h reaches for two globals of the application (sigpath::iqFrontEnd.getSampleRate() for the rate, sigpath::txState for
// the transmit hold): both are stood in for here; the rate is forced through the processor's own params.forceSampleRate.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <iostream>
#include <list>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include <utils/event.h>
#include <utils/flog.h>

namespace sigpath {
    struct StandInFrontEnd { double getSampleRate() { return 0.0; } };
    static StandInFrontEnd iqFrontEnd;
    static Event<bool> txState;
}

#include <if_nr.h>

#include "oracle_api.h"

extern "C" {

void* orc_logmmse_create(int sampleRate) {
    // logmmse.h reports its progress on std::cout; keep the test output clean (and keep number formatting out of a
    // process whose host program may have loaded another libstdc++): a stream in a failed state formats nothing
    std::cout.setstate(std::ios_base::failbit);
    auto* p = new dsp::IFNRLogMMSE;
    p->params.forceSampleRate = sampleRate;
    return p;
}
void orc_logmmse_destroy(void* h) { delete (dsp::IFNRLogMMSE*)h; }
void orc_logmmse_reset(void* h) { ((dsp::IFNRLogMMSE*)h)->reset(); }
void orc_logmmse_set_hold(void* h, int hold) { ((dsp::IFNRLogMMSE*)h)->setHold(hold != 0); }
int orc_logmmse_process(void* h, int count, const float* in, float* out) {
    int outCount = 0;
    ((dsp::IFNRLogMMSE*)h)->process((dsp::complex_t*)in, count, (dsp::complex_t*)out, outCount);
    return outCount;
}
float orc_expn(float x) { return dsp::math::expn(x); }
int orc_logmmse_noise(void* h, float* out, int cap) {
    auto& p = ((dsp::IFNRLogMMSE*)h)->params;
    if (!p.noise_mu2) { return 0; }
    const int n = (int)p.noise_mu2->size();
    if (out) { memcpy(out, p.noise_mu2->data(), (size_t)std::min(n, cap) * sizeof(float)); }
    return n;
}

}

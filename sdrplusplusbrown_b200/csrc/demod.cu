#include "demod.cuh"

#include "vfo_bank.cuh"

namespace b200 {
struct DemodState { int kind = 0; };
void demodDestroy(DemodState* d) { delete d; }
int processDemods(VfoBank&, cudaStream_t) { return B200_OK; }
} // namespace b200

extern "C" {
int b200_vfo_set_demod(b200_vfo_bank*, int, int, double, double, double, double) { b200::setError("demodulators not built yet"); return B200_ESTATE; }
const float* b200_vfo_audio_ptr(b200_vfo_bank*, int) { return nullptr; }
}

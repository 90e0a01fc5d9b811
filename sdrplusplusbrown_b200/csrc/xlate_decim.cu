#include "xlate_decim.cuh"

#include "vfo_bank.cuh"

namespace b200 {

bool fusedStageSupported(const Stage&) { return false; }
void fillFusedJob(FusedJob&, const Stage&, const Vfo&, const cf32*, int) {}
int launchFused(const FusedJob*, const std::vector<FusedJob>& hostJobs, cudaStream_t) {
    return hostJobs.empty() ? B200_OK : B200_ESTATE;
}

} // namespace b200

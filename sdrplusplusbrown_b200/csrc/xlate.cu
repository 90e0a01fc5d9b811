// Stand-alone frequency translation (FrequencyXlator::process, channel/frequency_xlator.h:43-56):
// out[n] = in[n] * phase[n], phase from the shared checkpoint source (rotator.cuh).  Used for VFOs whose
// first filter stage is not fused with the rotation: no decimation at all, or pow-2 ratios < 16 whose first
// stage decimates by only 2 or 4.
#include "rotator.cuh"
#include "vfo_bank.cuh"

namespace b200 {

namespace {
constexpr int kPer = 4;
constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads) xlateKernel(const XlateJob* __restrict__ jobs) {
    const XlateJob j = jobs[blockIdx.y];
    const int m0 = (blockIdx.x * kThreads + threadIdx.x) * kPer;
    if (m0 >= j.n) { return; }
    const float2* in = reinterpret_cast<const float2*>(j.in);
    float2* out = reinterpret_cast<float2*>(j.out);
#pragma unroll
    for (int i = 0; i < kPer; i++) {
        const int m = m0 + i;
        if (m < j.n) {
            const float2 x = __ldg(in + m);
            const float2 p = rotPhaseAt(j.chk, j.wtab, m);
            out[m] = make_float2(x.x * p.x - x.y * p.y, x.x * p.y + x.y * p.x);
        }
    }
}
} // namespace

int launchXlate(const XlateJob* dJobs, int nJobs, int maxN, cudaStream_t s) {
    if (nJobs <= 0 || maxN <= 0) { return B200_OK; }
    dim3 grid((maxN + kThreads * kPer - 1) / (kThreads * kPer), nJobs);
    xlateKernel<<<grid, kThreads, 0, s>>>(dJobs);
    B200_CHECK_LAUNCH();
    countLaunch();
    return B200_OK;
}

} // namespace b200

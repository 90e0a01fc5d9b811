// Stand-alone frequency translation (FrequencyXlator::process, channel/frequency_xlator.h:43-56):
// out[n] = in[n] * phase[n].  Used for VFOs whose first filter stage is not fused with the
// rotation (no decimation, or pow-2 ratios < 16 whose first stage decimates by 2 or 4) and by the SSB
// demodulator.  The phasor is evaluated in closed form: angle0 + theta*n in float64 at the start of
// each 8-sample run, then an fp32 recurrence inside the run (DESIGN.md "rotator").
#include "vfo_bank.cuh"

namespace b200 {

namespace {
constexpr int kRun = 8;
constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads) xlateKernel(const XlateJob* __restrict__ jobs) {
    const XlateJob j = jobs[blockIdx.y];
    const long long m0 = ((long long)blockIdx.x * kThreads + threadIdx.x) * kRun;
    if (m0 >= j.n) { return; }
    double a = fma(j.theta, (double)m0, j.angle0);
    a -= 6.283185307179586476925 * rint(a * 0.15915494309189533577);
    double sd, cd;
    sincos(a, &sd, &cd);
    float pr = (float)cd, pi = (float)sd;
    const float2* in = reinterpret_cast<const float2*>(j.in) + m0;
    float2* out = reinterpret_cast<float2*>(j.out) + m0;
    const int n = (int)min((long long)kRun, (long long)j.n - m0);
#pragma unroll
    for (int i = 0; i < kRun; i++) {
        if (i < n) {
            const float2 x = __ldg(in + i);
            out[i] = make_float2(x.x * pr - x.y * pi, x.x * pi + x.y * pr);
            const float nr = pr * j.stepRe - pi * j.stepIm;
            const float ni = pr * j.stepIm + pi * j.stepRe;
            pr = nr; pi = ni;
        }
    }
}
} // namespace

int launchXlate(const XlateJob* dJobs, int nJobs, int maxN, cudaStream_t s) {
    if (nJobs <= 0 || maxN <= 0) { return B200_OK; }
    dim3 grid((maxN + kThreads * kRun - 1) / (kThreads * kRun), nJobs);
    xlateKernel<<<grid, kThreads, 0, s>>>(dJobs);
    B200_CHECK_LAUNCH();
    countLaunch();
    return B200_OK;
}

} // namespace b200

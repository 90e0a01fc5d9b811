#include "rotator.cuh"

#include <cmath>
#include <complex>

namespace b200 {

namespace {

// EXACT: bit-for-bit replay of volk_32fc_s32fc_x2_rotator_32fc_generic's phase recursion.
// glibc's hypotf equals (float)sqrt((double)x*x + (double)y*y) (checked over 2e8 unit-circle inputs);
// __fmul_rn/__fadd_rn/__fsub_rn/__fdiv_rn forbid FMA contraction and are IEEE round-to-nearest.
__global__ void __launch_bounds__(32) rotChainKernel(const RotJob* __restrict__ jobs, int nJobs) {
    const int v = blockIdx.x * 32 + threadIdx.x;
    if (v >= nJobs) { return; }
    const RotJob j = jobs[v];
    float2 st = *j.state;
    float pr = st.x, pi = st.y;
    const float dr = j.dRe, di = j.dIm;
    for (int n = 0; n < j.count; n += kRotChunk) {
        j.chk[n >> kRotChunkLog] = make_float2(pr, pi);
        const int m = min(kRotChunk, j.count - n);
        for (int i = 0; i < m; i++) {
            const float nr = __fsub_rn(__fmul_rn(pr, dr), __fmul_rn(pi, di));
            const float ni = __fadd_rn(__fmul_rn(pr, di), __fmul_rn(pi, dr));
            pr = nr; pi = ni;
        }
        const float h = __double2float_rn(__dsqrt_rn(__dadd_rn(__dmul_rn((double)pr, (double)pr), __dmul_rn((double)pi, (double)pi))));
        pr = __fdiv_rn(pr, h); pi = __fdiv_rn(pi, h);
    }
    *j.state = make_float2(pr, pi);
}

// FAST: closed-form checkpoints
__global__ void __launch_bounds__(256) rotClosedFormKernel(const RotJob* __restrict__ jobs) {
    const RotJob j = jobs[blockIdx.y];
    const int c = blockIdx.x * 256 + threadIdx.x;
    if ((long long)c * kRotChunk >= j.count && c > 0) { return; }
    double a = fma(j.theta, (double)c * (double)kRotChunk, j.angle0);
    a -= 6.283185307179586476925 * rint(a * 0.15915494309189533577);
    double sd, cd;
    sincos(a, &sd, &cd);
    j.chk[c] = make_float2((float)cd, (float)sd);
}

} // namespace

// jobs [0, nExact) are exact-mode, [nExact, nJobs) fast-mode
int launchCheckpoints(const RotJob* dJobs, int nJobs, int nExact, int maxCount, cudaStream_t s) {
    if (nJobs <= 0 || maxCount <= 0) { return B200_OK; }
    if (nExact > 0) {
        rotChainKernel<<<(nExact + 31) / 32, 32, 0, s>>>(dJobs, nExact);
        B200_CHECK_LAUNCH(); countLaunch();
    }
    if (nJobs > nExact) {
        const int nChk = (maxCount + kRotChunk - 1) / kRotChunk;
        rotClosedFormKernel<<<dim3((nChk + 255) / 256, nJobs - nExact), 256, 0, s>>>(dJobs + nExact);
        B200_CHECK_LAUNCH(); countLaunch();
    }
    return B200_OK;
}

void makeRotatorTable(float dRe, float dIm, float2* out) {
    // complex powers of the fp32-rounded step, in float64 polar form
    const double mag = std::hypot((double)dRe, (double)dIm), ang = std::atan2((double)dIm, (double)dRe);
    const double lmag = std::log(mag);
    for (int j = 0; j < kRotChunk; j++) {
        const double r = std::exp(lmag * j), a = ang * j;
        out[j] = make_float2((float)(r * std::cos(a)), (float)(r * std::sin(a)));
    }
}

} // namespace b200

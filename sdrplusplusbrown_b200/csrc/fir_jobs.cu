// Generic streaming-filter kernels: decimating FIR / FIR (filter/decimating_fir.h:45-68,
// filter/fir.h:64-92) and polyphase rational resampler (multirate/polyphase_resampler.h:69-99),
// many independent jobs per launch (grid.y = job).  These serve the low-rate tail of every VFO
// (stages 2.., resampler, channel filter), the front-end decimator and the demodulator filters;
// the full-rate first stage has its own fused kernel (xlate_decim.cu).
#include "fir_jobs.cuh"

namespace b200 {

namespace {

constexpr int kThreads = 256;
constexpr int kOutPerThread = 4;             // consecutive outputs per thread (register blocking)
constexpr int kOutPerCta = kThreads * kOutPerThread;

template <int EF> struct Elem;
template <> struct Elem<2> {
    using T = float2;
    static __device__ __forceinline__ T zero() { return make_float2(0.f, 0.f); }
    static __device__ __forceinline__ void fma(T& a, float t, T x) { a.x = fmaf(t, x.x, a.x); a.y = fmaf(t, x.y, a.y); }
};
template <> struct Elem<1> {
    using T = float;
    static __device__ __forceinline__ T zero() { return 0.f; }
    static __device__ __forceinline__ void fma(T& a, float t, T x) { a = fmaf(t, x, a); }
};

// sample at position p of the job's [hist | data] work buffer
template <int EF>
__device__ __forceinline__ typename Elem<EF>::T loadPos(const FirJob& j, int p) {
    using T = typename Elem<EF>::T;
    return (p < j.H) ? __ldg(reinterpret_cast<const T*>(j.hist) + p) : __ldg(reinterpret_cast<const T*>(j.data) + (p - j.H));
}

// One CTA = up to kOutPerCta consecutive outputs of one job.  The input window of the CTA is staged
// in shared memory when it fits (coalesced global reads, reuse T/D times), taps are read through
// the read-only path (warp-uniform address -> broadcast).
template <int EF>
__global__ void __launch_bounds__(kThreads) firJobsKernel(const FirJob* __restrict__ jobs, int smemElems) {
    using E = Elem<EF>;
    using T = typename E::T;
    extern __shared__ __align__(16) unsigned char smemRaw[];
    T* win = reinterpret_cast<T*>(smemRaw);

    const FirJob j = jobs[blockIdx.y];
    const int k0 = blockIdx.x * kOutPerCta;
    if (k0 >= j.nOut) { return; }
    const int nk = min(kOutPerCta, j.nOut - k0);

    if (j.kind == JOB_DECIM_FIR) {
        const int base = j.offset + k0 * j.D;                 // first work-buffer position this CTA touches
        const int span = (nk - 1) * j.D + j.T;                // positions [base, base+span)
        const bool staged = span <= smemElems;
        if (staged) {
            for (int i = threadIdx.x; i < span; i += kThreads) { win[i] = loadPos<EF>(j, base + i); }
            __syncthreads();
        }
        // thread t owns outputs k0 + t + i*kThreads (i < kOutPerThread): lanes step by D samples in the window
        T acc[kOutPerThread];
#pragma unroll
        for (int i = 0; i < kOutPerThread; i++) { acc[i] = E::zero(); }
        if (staged) {
            for (int t = 0; t < j.T; t++) {
                const float h = __ldg(j.taps + t);
#pragma unroll
                for (int i = 0; i < kOutPerThread; i++) {
                    const int k = threadIdx.x + i * kThreads;
                    if (k < nk) { E::fma(acc[i], h, win[k * j.D + t]); }
                }
            }
        }
        else {
            for (int t = 0; t < j.T; t++) {
                const float h = __ldg(j.taps + t);
#pragma unroll
                for (int i = 0; i < kOutPerThread; i++) {
                    const int k = threadIdx.x + i * kThreads;
                    if (k < nk) { E::fma(acc[i], h, loadPos<EF>(j, base + k * j.D + t)); }
                }
            }
        }
#pragma unroll
        for (int i = 0; i < kOutPerThread; i++) {
            const int k = threadIdx.x + i * kThreads;
            if (k < nk) { reinterpret_cast<T*>(j.out)[k0 + k] = acc[i]; }
        }
    }
    else {
        // polyphase: output k sits at offset + (phase0 + k*M)/L and uses bank row (phase0 + k*M) % L
#pragma unroll
        for (int i = 0; i < kOutPerThread; i++) {
            const int k = threadIdx.x + i * kThreads;
            if (k >= nk) { continue; }
            const long long total = (long long)j.phase0 + (long long)(k0 + k) * j.M;
            const int pos = j.offset + (int)(total / j.L);
            const float* row = j.taps + (size_t)(total % j.L) * j.rowStride;
            T acc = E::zero();
            for (int t = 0; t < j.T; t++) { E::fma(acc, __ldg(row + t), loadPos<EF>(j, pos + t)); }
            reinterpret_cast<T*>(j.out)[k0 + k] = acc;
        }
    }
}

// histNext[i] = workbuf[nIn + i], i < H  (the reference's memmove of the last taps-1 samples, fir.h:86-87)
__global__ void __launch_bounds__(256) historyUpdateKernel(const FirJob* __restrict__ jobs) {
    const FirJob j = jobs[blockIdx.y];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= j.H) { return; }
    const int p = j.nIn + i;
    if (j.elemFloats == 2) {
        reinterpret_cast<float2*>(j.histNext)[i] = loadPos<2>(j, p);
    }
    else {
        reinterpret_cast<float*>(j.histNext)[i] = loadPos<1>(j, p);
    }
}

} // namespace

int launchFirJobs(const FirJob* d_jobs, int nJobs, int maxOut, int elemFloats, cudaStream_t s) {
    if (nJobs <= 0 || maxOut <= 0) { return B200_OK; }
    dim3 grid((maxOut + kOutPerCta - 1) / kOutPerCta, nJobs);
    const int smemBytes = 96 * 1024;
    if (elemFloats == 2) {
        static bool attr2 = false;
        if (!attr2) { B200_CUDA(cudaFuncSetAttribute(firJobsKernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemBytes)); attr2 = true; }
        firJobsKernel<2><<<grid, kThreads, smemBytes, s>>>(d_jobs, smemBytes / 8);
    }
    else {
        static bool attr1 = false;
        if (!attr1) { B200_CUDA(cudaFuncSetAttribute(firJobsKernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemBytes)); attr1 = true; }
        firJobsKernel<1><<<grid, kThreads, smemBytes, s>>>(d_jobs, smemBytes / 4);
    }
    B200_CHECK_LAUNCH();
    countLaunch();
    return B200_OK;
}

int launchHistoryUpdate(const FirJob* d_jobs, int nJobs, int maxH, cudaStream_t s) {
    if (nJobs <= 0 || maxH <= 0) { return B200_OK; }
    dim3 grid((maxH + 255) / 256, nJobs);
    historyUpdateKernel<<<grid, 256, 0, s>>>(d_jobs);
    B200_CHECK_LAUNCH();
    countLaunch();
    return B200_OK;
}

} // namespace b200

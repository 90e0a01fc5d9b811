// Generic streaming-filter kernels: decimating FIR / FIR (filter/decimating_fir.h:45-68,
// filter/fir.h:64-92) and polyphase rational resampler (multirate/polyphase_resampler.h:69-99),
// many independent jobs per launch (grid.y = job).  These serve the low-rate tail of every VFO
// (stages 2.., resampler, channel filter), the front-end decimator and the demodulator filters;
// the full-rate first stage has its own fused kernel (xlate_decim.cu).
#include "fir_jobs.cuh"

#include <algorithm>

namespace b200 {

namespace {

constexpr int kThreads = 128;               // one output per thread: even a 20 k-sample tail stage fills the 148 SMs
constexpr int kOutPerCta = kThreads;

template <int EF> struct Elem;
template <> struct Elem<2> {
    using T = float2;
    static __device__ __forceinline__ T zero() { return make_float2(0.f, 0.f); }
    static __device__ __forceinline__ void fma(T& a, float t, T x) { a.x = fmaf(t, x.x, a.x); a.y = fmaf(t, x.y, a.y); }
    // shared-memory slot of window position p: one pad element per 16 keeps lanes that step by 1, 2, 4 or 8
    // samples (the decimations of every non-fused stage) on distinct banks
    static __device__ __forceinline__ int slot(int p) { return p + (p >> 4); }
};
template <> struct Elem<1> {
    using T = float;
    static __device__ __forceinline__ T zero() { return 0.f; }
    static __device__ __forceinline__ void fma(T& a, float t, T x) { a = fmaf(t, x, a); }
    static __device__ __forceinline__ int slot(int p) { return p + (p >> 5); }
};

// sample at position p of the job's [hist | data] work buffer
template <int EF>
__device__ __forceinline__ typename Elem<EF>::T loadPos(const FirJob& j, int p) {
    using T = typename Elem<EF>::T;
    return (p < j.H) ? __ldg(reinterpret_cast<const T*>(j.hist) + p) : __ldg(reinterpret_cast<const T*>(j.data) + (p - j.H));
}

// One CTA = up to kOutPerCta consecutive outputs of one job.  The CTA's input window is staged in shared
// memory (coalesced global reads, each sample reused T/D times); taps are read through the read-only path
// (warp-uniform address -> broadcast) for FIR jobs and per-thread bank rows for the polyphase resampler.
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
    const int k = threadIdx.x;

    if (j.kind == JOB_DECIM_FIR) {
        const int base = j.offset + k0 * j.D;                 // first work-buffer position this CTA touches
        const int span = (nk - 1) * j.D + j.T;                // positions [base, base+span)
        const bool staged = E::slot(span) + 1 <= smemElems;
        T acc = E::zero();
        if (staged) {
            for (int i = threadIdx.x; i < span; i += kThreads) { win[E::slot(i)] = loadPos<EF>(j, base + i); }
            __syncthreads();
            if (k < nk) {
                const int p0 = k * j.D;
                int t = 0;
                for (; t + 4 <= j.T; t += 4) {
                    const float4 h = __ldg(reinterpret_cast<const float4*>(j.taps) + (t >> 2));
                    E::fma(acc, h.x, win[E::slot(p0 + t)]);
                    E::fma(acc, h.y, win[E::slot(p0 + t + 1)]);
                    E::fma(acc, h.z, win[E::slot(p0 + t + 2)]);
                    E::fma(acc, h.w, win[E::slot(p0 + t + 3)]);
                }
                for (; t < j.T; t++) { E::fma(acc, __ldg(j.taps + t), win[E::slot(p0 + t)]); }
            }
        }
        else if (k < nk) {
            for (int t = 0; t < j.T; t++) { E::fma(acc, __ldg(j.taps + t), loadPos<EF>(j, base + k * j.D + t)); }
        }
        if (k < nk) { reinterpret_cast<T*>(j.out)[k0 + k] = acc; }
    }
    else {
        // polyphase: output k sits at offset + (phase0 + k*M)/L and uses bank row (phase0 + k*M) % L
        const long long tot0 = (long long)j.phase0 + (long long)k0 * j.M;
        const long long totL = (long long)j.phase0 + (long long)(k0 + nk - 1) * j.M;
        const int base = j.offset + (int)(tot0 / j.L);
        const int span = (int)(totL / j.L) - (int)(tot0 / j.L) + j.T;
        const bool staged = E::slot(span) + 1 <= smemElems;
        if (staged) {
            for (int i = threadIdx.x; i < span; i += kThreads) { win[E::slot(i)] = loadPos<EF>(j, base + i); }
            __syncthreads();
        }
        if (k < nk) {
            const long long total = (long long)j.phase0 + (long long)(k0 + k) * j.M;
            const int pos = j.offset + (int)(total / j.L);
            const float* row = j.taps + (size_t)(total % j.L) * j.rowStride;
            T acc = E::zero();
            if (staged) {
                const int p0 = pos - base;
                int t = 0;
                for (; t + 4 <= j.T; t += 4) {          // rows are padded to a multiple of 4 floats and 16-byte aligned
                    const float4 h = __ldg(reinterpret_cast<const float4*>(row) + (t >> 2));
                    E::fma(acc, h.x, win[E::slot(p0 + t)]);
                    E::fma(acc, h.y, win[E::slot(p0 + t + 1)]);
                    E::fma(acc, h.z, win[E::slot(p0 + t + 2)]);
                    E::fma(acc, h.w, win[E::slot(p0 + t + 3)]);
                }
                for (; t < j.T; t++) { E::fma(acc, __ldg(row + t), win[E::slot(p0 + t)]); }
            }
            else {
                for (int t = 0; t < j.T; t++) { E::fma(acc, __ldg(row + t), loadPos<EF>(j, pos + t)); }
            }
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

// input window (in elements) a CTA of this job needs to stage
int firJobSpan(const FirJob& j) {
    if (j.kind == JOB_DECIM_FIR) { return (kOutPerCta - 1) * j.D + j.T; }
    return (int)(((long long)kOutPerCta * j.M) / j.L) + j.T + 2;
}

int launchFirJobs(const FirJob* d_jobs, int nJobs, int maxOut, int elemFloats, int maxSpanElems, cudaStream_t s) {
    if (nJobs <= 0 || maxOut <= 0) { return B200_OK; }
    dim3 grid((maxOut + kOutPerCta - 1) / kOutPerCta, nJobs);
    // request only what the widest job of this launch needs (more resident CTAs per SM); jobs whose window
    // would not fit in 160 KB read through L1 instead
    const size_t elemBytes = (size_t)elemFloats * sizeof(float);
    size_t want = ((size_t)maxSpanElems + (maxSpanElems >> 4) + 8) * elemBytes;
    const size_t smemBytes = std::min<size_t>(std::max<size_t>(want, 1024), 160 * 1024);
    if (elemFloats == 2) {
        B200_CUDA(cudaFuncSetAttribute(firJobsKernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(160 * 1024)));
        firJobsKernel<2><<<grid, kThreads, smemBytes, s>>>(d_jobs, (int)(smemBytes / elemBytes));
    }
    else {
        B200_CUDA(cudaFuncSetAttribute(firJobsKernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(160 * 1024)));
        firJobsKernel<1><<<grid, kThreads, smemBytes, s>>>(d_jobs, (int)(smemBytes / elemBytes));
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

#include "common.cuh"

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_set>

namespace b200 {

static thread_local char t_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void setError(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof(t_err), fmt, ap);
    va_end(ap);
}

int PinnedRing::init(size_t nbytes) {
    bytes = nbytes;
    for (int i = 0; i < kSlots; i++) {
        B200_CUDA(cudaHostAlloc(&host[i], bytes, cudaHostAllocDefault));
        B200_CUDA(cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming));
        used[i] = false;
    }
    return B200_OK;
}

void PinnedRing::destroy() {
    for (int i = 0; i < kSlots; i++) {
        if (host[i]) { cudaFreeHost(host[i]); host[i] = nullptr; }
        if (done[i]) { cudaEventDestroy(done[i]); done[i] = nullptr; }
    }
}

int PinnedRing::acquire(size_t nbytes, void** hostPtr, int* slot) {
    if (nbytes > bytes) {
        // grow every slot (rare: the VFO count grew); wait for in-flight copies first
        for (int i = 0; i < kSlots; i++) {
            if (used[i]) { B200_CUDA(cudaEventSynchronize(done[i])); used[i] = false; }
            B200_CUDA(cudaFreeHost(host[i]));
            host[i] = nullptr;
        }
        bytes = nbytes * 2;
        for (int i = 0; i < kSlots; i++) { B200_CUDA(cudaHostAlloc(&host[i], bytes, cudaHostAllocDefault)); }
    }
    const int sidx = next;
    next = (next + 1) % kSlots;
    if (used[sidx]) { B200_CUDA(cudaEventSynchronize(done[sidx])); used[sidx] = false; }
    *hostPtr = host[sidx];
    *slot = sidx;
    return B200_OK;
}

namespace {
struct DescBlob { uint4 q[248]; };     // 3968 bytes: with the two scalars the launch stays under the 4 KB parameter limit
__global__ void __launch_bounds__(256) putDescKernel(const __grid_constant__ DescBlob blob, uint4* __restrict__ dst, int n16) {
    if ((int)threadIdx.x < n16) { dst[threadIdx.x] = blob.q[threadIdx.x]; }
}
// descriptors too long for a few parameter blocks (a bank of a thousand VFOs): the SMs read the page-locked ring slot over PCIe
__global__ void __launch_bounds__(256) pullDescKernel(const uint4* __restrict__ hostSrc, uint4* __restrict__ dst, int n16) {
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n16; i += gridDim.x * 256) { dst[i] = hostSrc[i]; }
}
} // namespace

int uploadDesc(void* dDst, const void* hSrc, size_t bytes, cudaStream_t s, bool viaParams) {
    if (bytes == 0) { return B200_OK; }
    if (viaParams && bytes > 16 * sizeof(DescBlob)) {
        // not through the copy engine either (the point of viaParams: the upload must not queue behind the block copies)
        void* alias = nullptr;
        if (cudaHostGetDevicePointer(&alias, const_cast<void*>(hSrc), 0) == cudaSuccess && alias) {
            const int n16 = (int)((bytes + 15) / 16);
            pullDescKernel<<<std::min(64, (n16 + 255) / 256), 256, 0, s>>>((const uint4*)alias, (uint4*)dDst, n16);
            B200_CHECK_LAUNCH();
            return B200_OK;
        }
        cudaGetLastError();
    }
    if (!viaParams || bytes > 16 * sizeof(DescBlob)) {
        B200_CUDA(cudaMemcpyAsync(dDst, hSrc, bytes, cudaMemcpyHostToDevice, s));
        return B200_OK;
    }
    // descriptor buffers are allocated with slack and the ring slots are larger than any upload, so rounding the
    // length up to 16 bytes stays inside both
    const unsigned char* src = (const unsigned char*)hSrc;
    unsigned char* dst = (unsigned char*)dDst;
    for (size_t off = 0; off < bytes; off += sizeof(DescBlob)) {
        const size_t n = std::min(sizeof(DescBlob), bytes - off);
        DescBlob blob;
        memcpy(&blob, src + off, n);
        putDescKernel<<<1, 256, 0, s>>>(blob, (uint4*)(dst + off), (int)((n + 15) / 16));
        B200_CHECK_LAUNCH();
    }
    return B200_OK;
}

// ---- optional per-launch timing (CUDA events on the launch stream), one accumulator per kernel family
namespace {
struct ProfFamily {
    std::vector<cudaEvent_t> start, stop;
    std::vector<double> bytes, flops;
    size_t used = 0;
    double msTotal = 0.0, bytesTotal = 0.0, flopsTotal = 0.0;
    long long launches = 0;
};
struct Profiler {
    bool enabled = false;
    std::mutex mtx;
    ProfFamily fam[PROF_KINDS];
} g_prof;

int profDrainLocked(ProfFamily& f) {
    for (size_t i = 0; i < f.used; i++) {
        float ms = 0.f;
        B200_CUDA(cudaEventSynchronize(f.stop[i]));
        B200_CUDA(cudaEventElapsedTime(&ms, f.start[i], f.stop[i]));
        f.msTotal += ms; f.bytesTotal += f.bytes[i]; f.flopsTotal += f.flops[i]; f.launches++;
    }
    f.used = 0;
    return B200_OK;
}
} // namespace

bool profEnabled() { return g_prof.enabled; }

int profBegin(int kind, cudaStream_t s, double bytes, double flops, size_t* slot) {
    if (!g_prof.enabled) { return B200_OK; }
    std::lock_guard<std::mutex> lk(g_prof.mtx);
    ProfFamily& f = g_prof.fam[kind];
    if (f.used == f.start.size()) {
        if (f.used >= 4096) { B200_TRY(profDrainLocked(f)); }
        else {
            cudaEvent_t a, b;
            B200_CUDA(cudaEventCreate(&a)); B200_CUDA(cudaEventCreate(&b));
            f.start.push_back(a); f.stop.push_back(b); f.bytes.push_back(0.0); f.flops.push_back(0.0);
        }
    }
    *slot = f.used++;
    f.bytes[*slot] = bytes; f.flops[*slot] = flops;
    B200_CUDA(cudaEventRecord(f.start[*slot], s));
    return B200_OK;
}

int profEnd(int kind, size_t slot, cudaStream_t s) {
    if (!g_prof.enabled) { return B200_OK; }
    std::lock_guard<std::mutex> lk(g_prof.mtx);
    B200_CUDA(cudaEventRecord(g_prof.fam[kind].stop[slot], s));
    return B200_OK;
}

// ---- FP32 roof, measured: independent chains of packed FMAs (fma.rn.f32x2 -> FFMA2), every SM full of warps.  This is
// the instruction form that reaches the CUDA cores' full rate of 128 FMA per clock and SM on B200 (measured 127.5;
// three-register scalar FFMA chains stop at 77-85 because of operand bandwidth, tools/microbench/ffma2_throughput.cu).
namespace {
__device__ __forceinline__ float2 peakFma2(float2 a, float2 b, float2 c) {
    unsigned long long ra, rb, rc, rd;
    ra = *reinterpret_cast<unsigned long long*>(&a); rb = *reinterpret_cast<unsigned long long*>(&b); rc = *reinterpret_cast<unsigned long long*>(&c);
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2*>(&rd);
}
__global__ void __launch_bounds__(512) ffmaPeakKernel(float2* out, float s, int iters) {
    float2 a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = make_float2((float)threadIdx.x + i, (float)threadIdx.x - i); }
    const float2 x0 = make_float2(s, s + 1.f), x1 = make_float2(s + 2.f, s + 3.f);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
#pragma unroll
            for (int i = 0; i < 8; i++) { a[i] = peakFma2(a[i], x0, x1); }
        }
    }
    float2 r = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 8; i++) { r.x += a[i].x; r.y += a[i].y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
} // namespace

int PinnedRing::release(int slot, cudaStream_t s) {
    B200_CUDA(cudaEventRecord(done[slot], s));
    used[slot] = true;
    return B200_OK;
}

} // namespace b200

// page-locked allocations are remembered so b200_host_free knows how to release them
static std::mutex g_hostMtx;
static std::unordered_set<void*> g_pinned;

extern "C" {

void* b200_host_alloc(size_t bytes) {
    if (bytes == 0) { bytes = 64; }
    void* p = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) == cudaSuccess && ndev > 0 && cudaHostAlloc(&p, bytes, cudaHostAllocPortable) == cudaSuccess && p) {
        std::lock_guard<std::mutex> lk(g_hostMtx);
        g_pinned.insert(p);
        return p;
    }
    cudaGetLastError();   // clear the sticky "no device" error
    if (posix_memalign(&p, 64, bytes) != 0) { return nullptr; }
    return p;
}

void b200_host_free(void* p) {
    if (!p) { return; }
    bool pinned = false;
    {
        std::lock_guard<std::mutex> lk(g_hostMtx);
        auto it = g_pinned.find(p);
        if (it != g_pinned.end()) { g_pinned.erase(it); pinned = true; }
    }
    if (pinned) { cudaFreeHost(p); } else { free(p); }
}

int b200_device_to_host(void* h_dst, const void* d_src, size_t bytes, void* stream) {
    if (bytes == 0) { return B200_OK; }
    if (!h_dst || !d_src) { b200::setError("b200_device_to_host: NULL pointer"); return B200_EINVAL; }
    B200_CUDA(cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    B200_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return B200_OK;
}

int b200_profile_enable(int enabled) {
    using namespace b200;
    if (!enabled && g_prof.enabled) {
        std::lock_guard<std::mutex> lk(g_prof.mtx);
        for (auto& f : g_prof.fam) { B200_TRY(profDrainLocked(f)); }
    }
    g_prof.enabled = enabled != 0;
    return B200_OK;
}

int b200_profile_read_kind(int kind, double* ms_total, long long* launches, double* algorithmic_bytes, double* algorithmic_flops) {
    using namespace b200;
    if (kind < 0 || kind >= PROF_KINDS) { setError("b200_profile_read_kind: unknown kernel family %d", kind); return B200_EINVAL; }
    std::lock_guard<std::mutex> lk(g_prof.mtx);
    ProfFamily& f = g_prof.fam[kind];
    B200_TRY(profDrainLocked(f));
    if (ms_total) { *ms_total = f.msTotal; }
    if (launches) { *launches = f.launches; }
    if (algorithmic_bytes) { *algorithmic_bytes = f.bytesTotal; }
    if (algorithmic_flops) { *algorithmic_flops = f.flopsTotal; }
    f.msTotal = 0.0; f.bytesTotal = 0.0; f.flopsTotal = 0.0; f.launches = 0;
    return B200_OK;
}

int b200_profile_read(double* ms_total, long long* launches, double* algorithmic_bytes) {
    return b200_profile_read_kind(b200::PROF_STAGE1, ms_total, launches, algorithmic_bytes, nullptr);
}

int b200_measure_fp32_peak(int device, double* ffma_per_second) {
    using namespace b200;
    if (!ffma_per_second) { setError("b200_measure_fp32_peak: NULL result pointer"); return B200_EINVAL; }
    DeviceGuard g(device);
    if (!g.ok) { setError("cannot select CUDA device %d", device); return B200_ECUDA; }
    int sms = 0;
    B200_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    const int grid = sms * 4, threads = 512, iters = 8000;
    float2* out = nullptr;
    B200_CUDA(cudaMalloc((void**)&out, (size_t)grid * threads * sizeof(float2)));
    cudaEvent_t e0, e1;
    B200_CUDA(cudaEventCreate(&e0)); B200_CUDA(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 4; rep++) {
        B200_CUDA(cudaEventRecord(e0, 0));
        ffmaPeakKernel<<<grid, threads>>>(out, 0.5f, iters);
        B200_CHECK_LAUNCH();
        B200_CUDA(cudaEventRecord(e1, 0));
        B200_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        B200_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double rate = (double)grid * threads * iters * 128.0 / (ms * 1e-3);      // 64 FFMA2 = 128 FMAs per thread and iteration
        if (rep > 0 && rate > best) { best = rate; }
    }
    countLaunch(4);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(out);
    *ffma_per_second = best;
    return B200_OK;
}

const char* b200_last_error(void) { return b200::t_err; }
const char* b200_version(void) { return "b200dsp 0.1 (sm_100a)"; }
uint64_t b200_launch_count(void) { return b200::g_launches.load(); }
}

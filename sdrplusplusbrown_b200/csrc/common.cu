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
} // namespace

int uploadDesc(void* dDst, const void* hSrc, size_t bytes, cudaStream_t s, bool viaParams) {
    if (bytes == 0) { return B200_OK; }
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

const char* b200_last_error(void) { return b200::t_err; }
const char* b200_version(void) { return "b200dsp 0.1 (sm_100a)"; }
uint64_t b200_launch_count(void) { return b200::g_launches.load(); }
}

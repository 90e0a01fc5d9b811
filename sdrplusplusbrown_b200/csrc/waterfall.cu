// Waterfall row reduction: doZoom (gui/widgets/waterfall.cpp:98-154) + dB -> palette index -> RGBA
// (pushFFT, :1229-1235).  The fp32 running start index of doZoom is reproduced on the host with the
// identical loop (host_design.h zoomTable) and uploaded, so the bins each pixel covers are the
// reference's, bit for bit.  Headless: rows land in device buffers (no GL texture).
#include "common.cuh"
#include "host_design.h"

namespace b200 {
namespace {

// G = lanes per output pixel (power of two <= 32, chosen from the widest pixel): the lanes of a group walk the pixel's
// bins side by side, so a warp reads contiguous runs of the dB row, and fold their maxima with shuffles.
__global__ void __launch_bounds__(256) zoomPaletteKernel(const float* __restrict__ rows, int rawSize, int width, const int* __restrict__ sid,
                                                         const int* __restrict__ uf, float wfMin, float wfMax, const uint32_t* __restrict__ lut,
                                                         float* __restrict__ zoom, uint32_t* __restrict__ rgba, int logG,
                                                         int ringHeight, int ringBase, int liveRows) {
    const int G = 1 << logG;
    const int t = blockIdx.x * 256 + threadIdx.x;
    const int pix = t >> logG, sub = t & (G - 1);
    const bool live = pix < width;
    const int i = live ? pix : width - 1;
    if ((int)blockIdx.y >= liveRows) {
        // regeneration of a history that is not full yet: opaque black below the last stored line (waterfall.cpp:846-855)
        if (live && sub == 0 && rgba) { rgba[(size_t)blockIdx.y * width + i] = 0xFF000000u; }
        return;
    }
    // ringHeight > 0: output row y shows history line (ringBase + y) % ringHeight, newest first (waterfall.cpp:826)
    const int srcRow = ringHeight > 0 ? (int)((ringBase + blockIdx.y) % ringHeight) : (int)blockIdx.y;
    const float* row = rows + (size_t)srcRow * rawSize;
    const int s = __ldg(sid + i), n = __ldg(uf + i);
    float m = -INFINITY;
    // eight loads in flight per lane: the kernel is bound by memory latency, not by the maxima
    for (int j = sub; j < n; j += 8 * G) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; u++) { v[u] = (j + u * G < n) ? __ldg(row + s + j + u * G) : -INFINITY; }
#pragma unroll
        for (int u = 0; u < 8; u++) { m = fmaxf(m, v[u]); }
    }
    for (int o = G >> 1; o > 0; o >>= 1) { m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o)); }
    if (!live || sub != 0) { return; }
    if (zoom) { zoom[(size_t)blockIdx.y * width + i] = m; }
    if (rgba) {
        const float range = wfMax - wfMin;
        const float px = (fminf(fmaxf(m, wfMin), wfMax) - wfMin) / range;
        const int id = (int)(px * 999.0f);
        rgba[(size_t)blockIdx.y * width + i] = __ldg(lut + id);
    }
}

// The same reduction for views whose pixels cover a few bins to a few hundred: a CTA takes kZoomPx adjacent pixels of one
// row, brings the stretch of the dB row they cover into shared memory with coalesced 16-byte loads (a quarter of the load
// instructions of the per-pixel walk above, every sector used once), and four lanes per pixel fold their bins from there.
// tileLo / tileHi: first (rounded down to 4) and one-past-last bin of every pixel tile, from the host's zoom table.
constexpr int kZoomPx = 64;
constexpr int kZoomSpanCap = 12288;      // floats of shared memory per CTA (48 KB)
__global__ void __launch_bounds__(256) zoomTileKernel(const float* __restrict__ rows, int rawSize, int width, const int* __restrict__ sid,
                                                      const int* __restrict__ uf, const int* __restrict__ tileLo, const int* __restrict__ tileHi,
                                                      float wfMin, float wfMax, const uint32_t* __restrict__ lut,
                                                      float* __restrict__ zoom, uint32_t* __restrict__ rgba,
                                                      int ringHeight, int ringBase, int liveRows) {
    extern __shared__ __align__(16) float sRow[];
    const int p0 = blockIdx.x * kZoomPx;
    const int pix = p0 + (threadIdx.x >> 2), sub = threadIdx.x & 3;
    const bool live = pix < width;
    if ((int)blockIdx.y >= liveRows) {
        if (live && sub == 0 && rgba) { rgba[(size_t)blockIdx.y * width + pix] = 0xFF000000u; }
        return;
    }
    const int srcRow = ringHeight > 0 ? (int)((ringBase + blockIdx.y) % ringHeight) : (int)blockIdx.y;
    const float* row = rows + (size_t)srcRow * rawSize;
    const int lo = __ldg(tileLo + blockIdx.x), hi = __ldg(tileHi + blockIdx.x);
    const int n4 = (hi - lo + 3) >> 2;
    const float4* src = reinterpret_cast<const float4*>(row + lo);
    for (int k = threadIdx.x; k < n4; k += 256) {
        // rawSize is a multiple of 4 and lo is too: a float4 lies inside the row or wholly past its end
        reinterpret_cast<float4*>(sRow)[k] = (lo + 4 * k < rawSize) ? __ldg(src + k) : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    }
    __syncthreads();
    const int i = live ? pix : width - 1;
    const int s = __ldg(sid + i) - lo, n = live ? __ldg(uf + i) : 0;
    float m = -INFINITY;
    for (int j = sub; j < n; j += 4) { m = fmaxf(m, sRow[s + j]); }
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
    m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
    if (!live || sub != 0) { return; }
    if (zoom) { zoom[(size_t)blockIdx.y * width + i] = m; }
    if (rgba) {
        const float range = wfMax - wfMin;
        const float px = (fminf(fmaxf(m, wfMin), wfMax) - wfMin) / range;
        const int id = (int)(px * 999.0f);
        rgba[(size_t)blockIdx.y * width + i] = __ldg(lut + id);
    }
}

// WaterFall::getFFTBuffer (waterfall.cpp:1188-1199) for a batch: pushed row r (oldest first) becomes history line
// (cur0 - 1 - r) mod height -- the line index steps DOWN with every new row, so ascending lines run newest to oldest.
__global__ void __launch_bounds__(256) storeRawKernel(const float* __restrict__ rows, int firstRow, int rawSize, float* __restrict__ ring, int height, int cur0) {
    const int r = firstRow + blockIdx.y;
    int line = (cur0 - 1 - r) % height;
    if (line < 0) { line += height; }
    const float4* src = reinterpret_cast<const float4*>(rows + (size_t)r * rawSize);
    float4* dst = reinterpret_cast<float4*>(ring + (size_t)line * rawSize);
    for (int i = blockIdx.x * 256 + threadIdx.x; i < rawSize / 4; i += gridDim.x * 256) { dst[i] = __ldg(src + i); }
    if (blockIdx.x == 0 && threadIdx.x < (rawSize & 3)) { ring[(size_t)line * rawSize + (rawSize & ~3) + threadIdx.x] = rows[(size_t)r * rawSize + (rawSize & ~3) + threadIdx.x]; }
}

// WaterFall::calculateVFOSignalInfo (waterfall.cpp:688-759), one CTA per dB row.  The side-band bins (one bandwidth on
// each side of the VFO) are staged in shared memory; their mean, the mean of the bins at or below the lower-quartile
// order statistic (found by rank counting, no sort needed) and the in-band peak give strength and SNR.
__global__ void __launch_bounds__(256) vfoSignalKernel(const float* __restrict__ rows, int rawSize, int minSide, int minOff, int maxOff, int maxSide,
                                                       float* __restrict__ out) {
    extern __shared__ float side[];
    __shared__ double red[256];
    __shared__ float sKth;
    const float* row = rows + (size_t)blockIdx.x * rawSize;
    const int nL = minOff - minSide, nR = max(0, maxSide - (maxOff + 1));
    const int n = nL + nR;
    const int tid = threadIdx.x;
    double acc = 0.0;
    for (int i = tid; i < n; i += 256) {
        const float v = (i < nL) ? __ldg(row + minSide + i) : __ldg(row + maxOff + 1 + (i - nL));
        side[i] = v;
        acc += v;
    }
    red[tid] = acc;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) { if (tid < o) { red[tid] += red[tid + o]; } __syncthreads(); }
    const double avg = red[0] / (double)n;
    __syncthreads();
    const int lower = n / 4;
    for (int i = tid; i < n; i += 256) {
        const float v = side[i];
        int rank = 0;
        for (int j = 0; j < n; j++) { const float u = side[j]; rank += (u < v || (u == v && j < i)) ? 1 : 0; }
        if (rank == lower) { sKth = v; }
    }
    __syncthreads();
    const float kth = sKth;
    acc = 0.0;
    for (int i = tid; i < n; i += 256) { if (side[i] <= kth) { acc += side[i]; } }
    red[tid] = acc;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) { if (tid < o) { red[tid] += red[tid + o]; } __syncthreads(); }
    const double qavg = red[0] / (double)lower;
    __syncthreads();
    float mx = -INFINITY;
    for (int i = minOff + tid; i <= maxOff && i < rawSize; i += 256) { mx = fmaxf(mx, __ldg(row + i)); }
    red[tid] = mx;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) { if (tid < o) { red[tid] = fmax(red[tid], red[tid + o]); } __syncthreads(); }
    if (tid == 0) {
        const double avgdiff = avg - qavg;
        const float peak = (float)red[0];
        out[2 * blockIdx.x] = (float)((double)peak - avgdiff);
        out[2 * blockIdx.x + 1] = (float)((double)peak - avg - avgdiff);
    }
}

// pushFFT smoothing (waterfall.cpp:1247-1252) and peak hold (:1269-1272): recurrences over successive rows, so one
// thread owns one pixel column and walks the rows of the call in order; state persists in the handle.
__global__ void __launch_bounds__(256) smoothHoldKernel(float* __restrict__ rows, int nrows, int width, float* __restrict__ smoothBuf,
                                                        float* __restrict__ holdBuf, int smoothing, float alpha, int doHold, float holdSpeed,
                                                        float* __restrict__ holdOut) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= width) { return; }
    float s = smoothBuf[i], h = holdBuf[i];
    const float beta = 1.0f - alpha;
    for (int r = 0; r < nrows; r++) {
        float v = rows[(size_t)r * width + i];
        if (smoothing) {
            s = __fadd_rn(__fmul_rn(v, alpha), __fmul_rn(s, beta));
            v = s;
            rows[(size_t)r * width + i] = v;
        }
        if (doHold && i >= 1) { h = fmaxf(v, __fsub_rn(h, holdSpeed)); }
        if (holdOut) { holdOut[(size_t)r * width + i] = h; }
    }
    smoothBuf[i] = s; holdBuf[i] = h;
}

} // namespace
} // namespace b200

using namespace b200;

struct b200_waterfall {
    int device = 0, rawSize = 0, width = 0;
    int viewOffset = 0, viewWidth = 0;
    float wfMin = -70.0f, wfMax = 0.0f;     // WaterFall::WaterFall defaults (waterfall.cpp:159-166)
    std::vector<int> sid, uf;
    int* dSid = nullptr; int* dUf = nullptr; uint32_t* dLut = nullptr;
    float* dSmooth = nullptr; float* dHold = nullptr;     // smoothingBuf / latestFFTHold, both start at -1000 (waterfall.cpp:1073,1090)
    bool tableDirty = true;
    int maxUf = 1;            // widest pixel (bins), picks the lanes-per-pixel of the zoom kernel
    std::vector<int> tileLo, tileHi; int* dTileLo = nullptr; int* dTileHi = nullptr; int tileSpan = 0;   // zoomTileKernel geometry (tileSpan 0 = not usable)
    // raw dB history (WaterFall::rawFFTs / currentFFTLine / fftLines, waterfall.cpp:1188-1199)
    float* dRaw = nullptr; int height = 0, currentLine = 0, lines = 0;
};

extern "C" {

int b200_waterfall_create(b200_waterfall** out, int raw_fft_size, int data_width, int device) {
    if (!out || raw_fft_size <= 0 || data_width <= 0) { setError("bad arguments to b200_waterfall_create"); return B200_EINVAL; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { setError("no CUDA device available: libb200dsp has no CPU fallback"); return B200_ECUDA; }
    auto* w = new (std::nothrow) b200_waterfall;
    if (!w) { return B200_ENOMEM; }
    w->device = device; w->rawSize = raw_fft_size; w->width = data_width;
    w->viewOffset = 0; w->viewWidth = raw_fft_size;
    DeviceGuard g(device);
    std::vector<uint32_t> grey(1000);
    for (int i = 0; i < 1000; i++) { const uint32_t v = (uint32_t)(i * 255 / 999); grey[i] = 0xFF000000u | (v << 16) | (v << 8) | v; }
    if (cudaMalloc((void**)&w->dSid, data_width * sizeof(int)) != cudaSuccess || cudaMalloc((void**)&w->dUf, data_width * sizeof(int)) != cudaSuccess ||
        cudaMalloc((void**)&w->dLut, 1000 * sizeof(uint32_t)) != cudaSuccess ||
        cudaMalloc((void**)&w->dSmooth, data_width * sizeof(float)) != cudaSuccess || cudaMalloc((void**)&w->dHold, data_width * sizeof(float)) != cudaSuccess ||
        cudaMemcpy(w->dLut, grey.data(), 1000 * sizeof(uint32_t), cudaMemcpyHostToDevice) != cudaSuccess) {
        setError("device allocation failed in b200_waterfall_create");
        delete w;
        return B200_ENOMEM;
    }
    {
        std::vector<float> init((size_t)data_width, -1000.0f);
        cudaMemcpy(w->dSmooth, init.data(), init.size() * sizeof(float), cudaMemcpyHostToDevice);
        cudaMemcpy(w->dHold, init.data(), init.size() * sizeof(float), cudaMemcpyHostToDevice);
    }
    *out = w;
    return B200_OK;
}

int b200_waterfall_destroy(b200_waterfall* wf) {
    if (!wf) { setError("NULL waterfall handle"); return B200_EINVAL; }
    DeviceGuard g(wf->device);
    cudaFree(wf->dSid); cudaFree(wf->dUf); cudaFree(wf->dTileLo); cudaFree(wf->dTileHi); cudaFree(wf->dLut); cudaFree(wf->dSmooth); cudaFree(wf->dHold);
    if (wf->dRaw) { cudaFree(wf->dRaw); }
    delete wf;
    return B200_OK;
}

int b200_waterfall_set_view(b200_waterfall* wf, int offset, int width) {
    if (!wf || width <= 0) { setError("bad arguments to b200_waterfall_set_view"); return B200_EINVAL; }
    wf->viewOffset = offset; wf->viewWidth = width; wf->tableDirty = true;
    return B200_OK;
}

int b200_waterfall_set_range(b200_waterfall* wf, float wf_min, float wf_max) {
    if (!wf) { setError("NULL waterfall handle"); return B200_EINVAL; }
    wf->wfMin = wf_min; wf->wfMax = wf_max;
    return B200_OK;
}

int b200_waterfall_set_palette(b200_waterfall* wf, const uint32_t* lut1000) {
    if (!wf || !lut1000) { setError("bad arguments to b200_waterfall_set_palette"); return B200_EINVAL; }
    DeviceGuard g(wf->device);
    B200_CUDA(cudaMemcpy(wf->dLut, lut1000, 1000 * sizeof(uint32_t), cudaMemcpyHostToDevice));
    return B200_OK;
}

// WaterFall::updatePallette (waterfall.cpp:1282-1295): linear blend between neighbouring colours,
// channel values truncated to integers, packed A B G R.
int b200_waterfall_make_palette(const float* colors_rgb, int color_count, uint32_t* lut1000) {
    if (!colors_rgb || color_count <= 0 || !lut1000) { setError("bad arguments to b200_waterfall_make_palette"); return B200_EINVAL; }
    for (int i = 0; i < 1000; i++) {
        const float pos = ((float)i / 1000.0f) * color_count;
        const int lo = std::min(std::max((int)floorf(pos), 0), color_count - 1);
        const int hi = std::min(std::max((int)ceilf(pos), 0), color_count - 1);
        const float ratio = pos - lo;
        uint32_t ch[3];
        for (int c = 0; c < 3; c++) {
            const float v = (colors_rgb[lo * 3 + c] * (1.0 - ratio)) + (colors_rgb[hi * 3 + c] * (ratio));
            ch[c] = (uint32_t)v;
        }
        lut1000[i] = (255u << 24) | (ch[2] << 16) | (ch[1] << 8) | ch[0];
    }
    return B200_OK;
}

static int launchZoomTiles(b200_waterfall* wf, const float* rows, float* d_zoom, uint32_t* d_rgba, int outRows, int ringHeight, int ringBase, int liveRows, cudaStream_t s) {
    const size_t smem = (size_t)wf->tileSpan * sizeof(float);
    B200_CUDA(cudaFuncSetAttribute(zoomTileKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kZoomSpanCap * sizeof(float) + 64)));   // per device
    dim3 grid((unsigned)((wf->width + kZoomPx - 1) / kZoomPx), (unsigned)outRows);
    zoomTileKernel<<<grid, 256, smem, s>>>(rows, wf->rawSize, wf->width, wf->dSid, wf->dUf, wf->dTileLo, wf->dTileHi, wf->wfMin, wf->wfMax, wf->dLut,
                                            d_zoom, d_rgba, ringHeight, ringBase, liveRows);
    B200_CHECK_LAUNCH();
    return B200_OK;
}

static int refreshTable(b200_waterfall* wf) {
    if (!wf->tableDirty) { return B200_OK; }
    design::zoomTable(wf->viewOffset, wf->viewWidth, wf->rawSize, wf->width, wf->sid, wf->uf);
    // the reference indexes the raw row unchecked; keep every read inside the row
    for (int i = 0; i < wf->width; i++) {
        if (wf->sid[i] < 0) { wf->sid[i] = 0; }
        if (wf->sid[i] >= wf->rawSize) { wf->sid[i] = wf->rawSize - 1; wf->uf[i] = 0; }
        if (wf->uf[i] < 0) { wf->uf[i] = 0; }
    }
    wf->maxUf = 1;
    for (int i = 0; i < wf->width; i++) { wf->maxUf = std::max(wf->maxUf, wf->uf[i]); }
    B200_CUDA(cudaMemcpy(wf->dSid, wf->sid.data(), wf->width * sizeof(int), cudaMemcpyHostToDevice));
    B200_CUDA(cudaMemcpy(wf->dUf, wf->uf.data(), wf->width * sizeof(int), cudaMemcpyHostToDevice));
    // pixel tiles of the shared-memory kernel: usable when every tile's stretch of the row fits its buffer
    const int nTiles = (wf->width + kZoomPx - 1) / kZoomPx;
    wf->tileLo.assign(nTiles, 0); wf->tileHi.assign(nTiles, 0);
    wf->tileSpan = 0;
    bool usable = (wf->rawSize & 3) == 0;
    for (int t = 0; t < nTiles && usable; t++) {
        int lo = wf->rawSize, hi = 0;
        for (int i = t * kZoomPx; i < std::min(wf->width, (t + 1) * kZoomPx); i++) {
            if (wf->uf[i] <= 0) { continue; }
            lo = std::min(lo, wf->sid[i]); hi = std::max(hi, wf->sid[i] + wf->uf[i]);
        }
        if (hi <= lo) { lo = 0; hi = 0; }
        lo &= ~3;
        wf->tileLo[t] = lo; wf->tileHi[t] = hi;
        wf->tileSpan = std::max(wf->tileSpan, ((hi - lo + 3) & ~3) + 4);
        if (wf->tileSpan > kZoomSpanCap) { usable = false; }
    }
    if (!usable) { wf->tileSpan = 0; }
    if (!wf->dTileLo) {
        const int cap = (wf->width + kZoomPx - 1) / kZoomPx;
        B200_CUDA(cudaMalloc((void**)&wf->dTileLo, cap * sizeof(int)));
        B200_CUDA(cudaMalloc((void**)&wf->dTileHi, cap * sizeof(int)));
    }
    B200_CUDA(cudaMemcpy(wf->dTileLo, wf->tileLo.data(), nTiles * sizeof(int), cudaMemcpyHostToDevice));
    B200_CUDA(cudaMemcpy(wf->dTileHi, wf->tileHi.data(), nTiles * sizeof(int), cudaMemcpyHostToDevice));
    wf->tableDirty = false;
    return B200_OK;
}

int b200_waterfall_push(b200_waterfall* wf, const float* d_rows, int nrows, float* d_zoom, uint32_t* d_rgba, void* stream) {
    if (!wf || !d_rows || nrows < 0) { setError("bad arguments to b200_waterfall_push"); return B200_EINVAL; }
    if (nrows == 0) { return B200_OK; }
    DeviceGuard g(wf->device);
    B200_TRY(refreshTable(wf));
    int logG = 0;
    while (logG < 5 && (8 << logG) < wf->maxUf) { logG++; }          // up to eight bins per lane before adding lanes
    dim3 grid((unsigned)((((size_t)wf->width << logG) + 255) / 256), nrows);
    if (d_zoom || d_rgba) {
        if (wf->tileSpan > 0 && (((uintptr_t)d_rows) & 15) == 0) {
            B200_TRY(launchZoomTiles(wf, d_rows, d_zoom, d_rgba, nrows, 0, 0, nrows, (cudaStream_t)stream));
        }
        else {
            zoomPaletteKernel<<<grid, 256, 0, (cudaStream_t)stream>>>(d_rows, wf->rawSize, wf->width, wf->dSid, wf->dUf, wf->wfMin, wf->wfMax, wf->dLut, d_zoom, d_rgba,
                                                                      logG, 0, 0, nrows);
            B200_CHECK_LAUNCH();
        }
        countLaunch();
    }
    if (wf->dRaw) {
        // keep the raw rows for a later full regeneration; of a batch longer than the history only the newest rows stay
        const int first = std::max(0, nrows - wf->height);
        if ((((uintptr_t)d_rows) & 15) != 0 || (wf->rawSize & 3) != 0) { setError("history rows must be 16-byte aligned with a multiple-of-4 size"); return B200_EINVAL; }
        storeRawKernel<<<dim3(std::max(1, std::min(64, wf->rawSize / 1024)), nrows - first), 256, 0, (cudaStream_t)stream>>>(d_rows, first, wf->rawSize, wf->dRaw, wf->height,
                                                                                                                  wf->currentLine);
        B200_CHECK_LAUNCH();
        countLaunch();
        wf->currentLine = (int)((((long long)wf->currentLine - nrows) % wf->height + wf->height) % wf->height);
        wf->lines = std::min(wf->height, wf->lines + nrows);
    }
    return B200_OK;
}

// WaterFall::rawFFTs: keep the last `height` raw dB rows on the device (0 = none)
int b200_waterfall_set_history(b200_waterfall* wf, int height) {
    if (!wf || height < 0) { setError("bad arguments to b200_waterfall_set_history"); return B200_EINVAL; }
    DeviceGuard g(wf->device);
    B200_CUDA(cudaDeviceSynchronize());
    if (wf->dRaw) { cudaFree(wf->dRaw); wf->dRaw = nullptr; }
    wf->height = height; wf->currentLine = 0; wf->lines = 0;
    if (height > 0) { B200_CUDA(cudaMalloc((void**)&wf->dRaw, (size_t)height * wf->rawSize * sizeof(float))); }
    return B200_OK;
}

// WaterFall::updateWaterfallFb (waterfall.cpp:771-866): every stored line through the CURRENT view, range and palette
// in one launch (the reference splits the lines over four CPU threads).  Output row 0 is the newest line; rows past
// the number of stored lines are opaque black.  Returns the number of stored lines.
int b200_waterfall_regenerate(b200_waterfall* wf, float* d_zoom, uint32_t* d_rgba, void* stream) {
    if (!wf || (!d_zoom && !d_rgba)) { setError("bad arguments to b200_waterfall_regenerate"); return B200_EINVAL; }
    if (!wf->dRaw) { setError("b200_waterfall_regenerate needs b200_waterfall_set_history first"); return B200_ESTATE; }
    DeviceGuard g(wf->device);
    B200_TRY(refreshTable(wf));
    int logG = 0;
    while (logG < 5 && (8 << logG) < wf->maxUf) { logG++; }
    dim3 grid((unsigned)((((size_t)wf->width << logG) + 255) / 256), wf->height);
    if (wf->tileSpan > 0) {
        B200_TRY(launchZoomTiles(wf, wf->dRaw, d_zoom, d_rgba, wf->height, wf->height, wf->currentLine, wf->lines, (cudaStream_t)stream));
    }
    else {
        zoomPaletteKernel<<<grid, 256, 0, (cudaStream_t)stream>>>(wf->dRaw, wf->rawSize, wf->width, wf->dSid, wf->dUf, wf->wfMin, wf->wfMax, wf->dLut, d_zoom, d_rgba,
                                                                  logG, wf->height, wf->currentLine, wf->lines);
        B200_CHECK_LAUNCH();
    }
    countLaunch();
    return wf->lines;
}

int b200_waterfall_zoom_table(b200_waterfall* wf, int* sid, int* ufactor) {
    if (!wf || !sid || !ufactor) { setError("bad arguments to b200_waterfall_zoom_table"); return B200_EINVAL; }
    DeviceGuard g(wf->device);
    B200_TRY(refreshTable(wf));
    memcpy(sid, wf->sid.data(), wf->width * sizeof(int));
    memcpy(ufactor, wf->uf.data(), wf->width * sizeof(int));
    return B200_OK;
}

// WaterFall::rawFFTIndex (waterfall.cpp:761-763)
static int rawFftIndex(double f, double wholeBw, int n) {
    const int v = (int)(((f / (wholeBw / 2.0)) * (double)(n / 2)) + (n / 2));
    return std::min(std::max(v, 0), n);
}

int b200_waterfall_vfo_signal(b200_waterfall* wf, const float* d_rows, int nrows, double whole_bandwidth, double center_offset, double bandwidth,
                              float* d_out, void* stream) {
    if (!wf || !d_rows || !d_out || nrows < 0 || whole_bandwidth <= 0 || bandwidth <= 0) { setError("bad arguments to b200_waterfall_vfo_signal"); return B200_EINVAL; }
    const int n = wf->rawSize;
    const int minSide = rawFftIndex(center_offset - bandwidth, whole_bandwidth, n), minOff = rawFftIndex(center_offset - bandwidth / 2.0, whole_bandwidth, n);
    const int maxOff = rawFftIndex(center_offset + bandwidth / 2.0, whole_bandwidth, n), maxSide = rawFftIndex(center_offset + bandwidth, whole_bandwidth, n);
    const int side = std::max(0, minOff - minSide) + std::max(0, maxSide - (maxOff + 1));
    if (side <= 0 || side / 4 <= 0) { return 1; }          // the reference returns false: nothing to measure against
    if (side > 10000) { setError("VFO side bands span %d bins; at most 10000 are supported", side); return B200_ETOOBIG; }
    if (nrows == 0) { return B200_OK; }
    DeviceGuard g(wf->device);
    vfoSignalKernel<<<nrows, 256, (size_t)side * sizeof(float), (cudaStream_t)stream>>>(d_rows, n, minSide, minOff, maxOff, maxSide, d_out);
    B200_CHECK_LAUNCH();
    countLaunch();
    return B200_OK;
}

int b200_waterfall_smooth_hold(b200_waterfall* wf, float* d_zoom_rows, int nrows, int smoothing, float alpha, int hold, float hold_speed,
                               float* d_hold_rows, void* stream) {
    if (!wf || !d_zoom_rows || nrows < 0) { setError("bad arguments to b200_waterfall_smooth_hold"); return B200_EINVAL; }
    if (nrows == 0 || (!smoothing && !hold)) { return B200_OK; }
    DeviceGuard g(wf->device);
    smoothHoldKernel<<<(wf->width + 255) / 256, 256, 0, (cudaStream_t)stream>>>(d_zoom_rows, nrows, wf->width, wf->dSmooth, wf->dHold, smoothing, alpha, hold,
                                                                              hold_speed, d_hold_rows);
    B200_CHECK_LAUNCH();
    countLaunch();
    return B200_OK;
}

} // extern "C"

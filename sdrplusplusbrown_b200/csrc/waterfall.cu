// Waterfall row reduction: doZoom (gui/widgets/waterfall.cpp:98-154) + dB -> palette index -> RGBA
// (pushFFT, :1229-1235).  The fp32 running start index of doZoom is reproduced on the host with the
// identical loop (host_design.h zoomTable) and uploaded, so the bins each pixel covers are the
// reference's, bit for bit.  Headless: rows land in device buffers (no GL texture).
#include "common.cuh"
#include "host_design.h"

namespace b200 {
namespace {

__global__ void __launch_bounds__(256) zoomPaletteKernel(const float* __restrict__ rows, int rawSize, int width, const int* __restrict__ sid,
                                                         const int* __restrict__ uf, float wfMin, float wfMax, const uint32_t* __restrict__ lut,
                                                         float* __restrict__ zoom, uint32_t* __restrict__ rgba) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= width) { return; }
    const float* row = rows + (size_t)blockIdx.y * rawSize;
    const int s = sid[i], n = uf[i];
    float m = -INFINITY;
    for (int j = 0; j < n; j++) { m = fmaxf(m, __ldg(row + s + j)); }
    if (zoom) { zoom[(size_t)blockIdx.y * width + i] = m; }
    if (rgba) {
        const float range = wfMax - wfMin;
        const float px = (fminf(fmaxf(m, wfMin), wfMax) - wfMin) / range;
        const int id = (int)(px * 999.0f);
        rgba[(size_t)blockIdx.y * width + i] = lut[id];
    }
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
    bool tableDirty = true;
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
        cudaMemcpy(w->dLut, grey.data(), 1000 * sizeof(uint32_t), cudaMemcpyHostToDevice) != cudaSuccess) {
        setError("device allocation failed in b200_waterfall_create");
        delete w;
        return B200_ENOMEM;
    }
    *out = w;
    return B200_OK;
}

int b200_waterfall_destroy(b200_waterfall* wf) {
    if (!wf) { setError("NULL waterfall handle"); return B200_EINVAL; }
    DeviceGuard g(wf->device);
    cudaFree(wf->dSid); cudaFree(wf->dUf); cudaFree(wf->dLut);
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

static int refreshTable(b200_waterfall* wf) {
    if (!wf->tableDirty) { return B200_OK; }
    design::zoomTable(wf->viewOffset, wf->viewWidth, wf->rawSize, wf->width, wf->sid, wf->uf);
    // the reference indexes the raw row unchecked; keep every read inside the row
    for (int i = 0; i < wf->width; i++) {
        if (wf->sid[i] < 0) { wf->sid[i] = 0; }
        if (wf->sid[i] >= wf->rawSize) { wf->sid[i] = wf->rawSize - 1; wf->uf[i] = 0; }
        if (wf->uf[i] < 0) { wf->uf[i] = 0; }
    }
    B200_CUDA(cudaMemcpy(wf->dSid, wf->sid.data(), wf->width * sizeof(int), cudaMemcpyHostToDevice));
    B200_CUDA(cudaMemcpy(wf->dUf, wf->uf.data(), wf->width * sizeof(int), cudaMemcpyHostToDevice));
    wf->tableDirty = false;
    return B200_OK;
}

int b200_waterfall_push(b200_waterfall* wf, const float* d_rows, int nrows, float* d_zoom, uint32_t* d_rgba, void* stream) {
    if (!wf || !d_rows || nrows < 0) { setError("bad arguments to b200_waterfall_push"); return B200_EINVAL; }
    if (nrows == 0) { return B200_OK; }
    DeviceGuard g(wf->device);
    B200_TRY(refreshTable(wf));
    dim3 grid((wf->width + 255) / 256, nrows);
    zoomPaletteKernel<<<grid, 256, 0, (cudaStream_t)stream>>>(d_rows, wf->rawSize, wf->width, wf->dSid, wf->dUf, wf->wfMin, wf->wfMax, wf->dLut, d_zoom, d_rgba);
    B200_CHECK_LAUNCH();
    countLaunch();
    return B200_OK;
}

int b200_waterfall_zoom_table(b200_waterfall* wf, int* sid, int* ufactor) {
    if (!wf || !sid || !ufactor) { setError("bad arguments to b200_waterfall_zoom_table"); return B200_EINVAL; }
    DeviceGuard g(wf->device);
    B200_TRY(refreshTable(wf));
    memcpy(sid, wf->sid.data(), wf->width * sizeof(int));
    memcpy(ufactor, wf->uf.data(), wf->width * sizeof(int));
    return B200_OK;
}

} // extern "C"

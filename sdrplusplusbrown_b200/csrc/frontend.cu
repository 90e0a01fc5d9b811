// IQ front end: pre-processing chain, FFT framing and the VFO bank it feeds.
// Reference: IQFrontEnd (signal_path/iq_frontend.{h,cpp}); block order of iq_frontend.cpp:40-44 is
// PowerDecimator -> DCBlocker -> Conjugate, then the Splitter (:46) fans the result out to the FFT
// branch (Reshaper + handler, :50-52,267-287) and to every VFO (:159-179).  On the GPU the fan-out
// costs nothing: every consumer reads the same device buffer.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>

#include "fft.cuh"
#include "detector.cuh"
#include "vfo_bank.cuh"

namespace b200 {

int branchOrder = 0;        // B200_OPT_BRANCH_ORDER
int hostChunkMiB = 0;       // B200_OPT_HOST_CHUNK_MIB: chunk of the *_process_host entry points in MiB of host data; 0 = 16 (64 for a bank of 64 VFOs or more)

namespace {

// ------------------------------------------------------------------ DC blocker (correction/dc_blocker.h:54-60)
//   out[i] = in[i] - o;  o += out[i] * r      <=>      o' = (1-r) * o + r * in[i]
// A first-order linear recurrence: each CTA folds its chunk into an affine map o -> A*o + B, a
// single small CTA chains the chunk maps (float64), and the apply pass re-walks each chunk from
// its exact starting offset.  The optional conjugate (math/conjugate.h:12-15) rides on the apply pass.
constexpr int kDcThreads = 256;
constexpr int kDcPerThread = 8;
constexpr int kDcChunk = kDcThreads * kDcPerThread;

struct DcParams {
    float rate;        // (float)(50/sr), iq_frontend.h:84-86, DCBlocker::_rate is float
    double a;          // 1 - rate
    double aRun;       // a^kDcPerThread
    double aChunk;     // a^kDcChunk
};

// B for a run of samples starting with offset 0: returns o after the run
__device__ __forceinline__ float2 dcRunFold(const float2* x, int n, float r) {
    float2 o = make_float2(0.f, 0.f);
    for (int i = 0; i < n; i++) {
        o.x += (x[i].x - o.x) * r;
        o.y += (x[i].y - o.y) * r;
    }
    return o;
}

__global__ void __launch_bounds__(kDcThreads) dcReduceKernel(const cf32* __restrict__ in, int n, DcParams p, double2* __restrict__ chunkB) {
    __shared__ double2 sB[kDcThreads];
    const int base = blockIdx.x * kDcChunk + threadIdx.x * kDcPerThread;
    float2 x[kDcPerThread];
    int cnt = 0;
#pragma unroll
    for (int i = 0; i < kDcPerThread; i++) {
        if (base + i < n) { x[i] = __ldg(reinterpret_cast<const float2*>(in) + base + i); cnt = i + 1; }
        else { x[i] = make_float2(0.f, 0.f); }
    }
    // a zero sample leaves o' = a*o, so padding the tail with zeros keeps A = aRun for every thread;
    // the true end-of-block state is recovered by the scan kernel from the exact tail length.
    (void)cnt;
    const float2 b = dcRunFold(x, kDcPerThread, p.rate);
    sB[threadIdx.x] = make_double2((double)b.x, (double)b.y);
    __syncthreads();
    // inclusive scan of affine maps with equal slopes: S_t = aRun * S_{t-1} + B_t
    double mul = p.aRun;
    for (int d = 1; d < kDcThreads; d <<= 1) {
        double2 add = make_double2(0.0, 0.0);
        if ((int)threadIdx.x >= d) { add = sB[threadIdx.x - d]; }
        __syncthreads();
        if ((int)threadIdx.x >= d) { sB[threadIdx.x].x += mul * add.x; sB[threadIdx.x].y += mul * add.y; }
        __syncthreads();
        mul *= mul;
    }
    if (threadIdx.x == kDcThreads - 1) { chunkB[blockIdx.x] = sB[threadIdx.x]; }
}

// chunkStart[c] = offset entering chunk c; state[0] = offset after the whole block
__global__ void __launch_bounds__(1024) dcScanKernel(const double2* __restrict__ chunkB, int nChunks, int n, DcParams p,
                                                     double2* __restrict__ chunkStart, double2* __restrict__ state) {
    __shared__ double2 sB[1024];
    __shared__ double sA[1024];
    const int per = (nChunks + 1023) / 1024;
    const int c0 = threadIdx.x * per;
    // fold this thread's run of chunks
    double A = 1.0; double2 B = make_double2(0.0, 0.0);
    for (int c = c0; c < min(c0 + per, nChunks); c++) {
        const double2 b = chunkB[c];
        B.x = p.aChunk * B.x + b.x; B.y = p.aChunk * B.y + b.y;
        A *= p.aChunk;
    }
    sA[threadIdx.x] = A; sB[threadIdx.x] = B;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        double a2 = 1.0; double2 b2 = make_double2(0.0, 0.0);
        const bool act = (int)threadIdx.x >= d;
        if (act) { a2 = sA[threadIdx.x - d]; b2 = sB[threadIdx.x - d]; }
        __syncthreads();
        if (act) {
            const double myA = sA[threadIdx.x];
            sB[threadIdx.x].x += myA * b2.x; sB[threadIdx.x].y += myA * b2.y;
            sA[threadIdx.x] = myA * a2;
        }
        __syncthreads();
    }
    // exclusive prefix for this thread = inclusive of thread-1, applied to the carried-in state
    const double2 o0 = state[0];
    double2 o;
    if (threadIdx.x == 0) { o = o0; }
    else {
        const double Ap = sA[threadIdx.x - 1]; const double2 Bp = sB[threadIdx.x - 1];
        o = make_double2(Ap * o0.x + Bp.x, Ap * o0.y + Bp.y);
    }
    for (int c = c0; c < min(c0 + per, nChunks); c++) {
        chunkStart[c] = o;
        const double2 b = chunkB[c];
        o.x = p.aChunk * o.x + b.x; o.y = p.aChunk * o.y + b.y;
    }
    __syncthreads();
    if (threadIdx.x == 1023) {
        // the last chunk was zero padded to a full chunk: undo the surplus decay a^(pad)
        const int pad = nChunks * kDcChunk - n;
        const double Aall = sA[1023]; const double2 Ball = sB[1023];
        double2 e = make_double2(Aall * o0.x + Ball.x, Aall * o0.y + Ball.y);
        const double undo = pow(p.a, -(double)pad);
        state[1] = make_double2(e.x * undo, e.y * undo);
    }
}

// `in` and `out` may be the same buffer (b200_dc_block_noalloc allows it): neither is __restrict__, the loads are
// ordinary loads (not the read-only path), and a thread writes only the eight samples it has itself read before.
__global__ void __launch_bounds__(kDcThreads) dcApplyKernel(const cf32* in, cf32* out, int n, DcParams p,
                                                            const double2* __restrict__ chunkStart, int conj) {
    __shared__ double2 sB[kDcThreads];
    const int base = blockIdx.x * kDcChunk + threadIdx.x * kDcPerThread;
    float2 x[kDcPerThread];
#pragma unroll
    for (int i = 0; i < kDcPerThread; i++) {
        x[i] = (base + i < n) ? reinterpret_cast<const float2*>(in)[base + i] : make_float2(0.f, 0.f);
    }
    const float2 b = dcRunFold(x, kDcPerThread, p.rate);
    sB[threadIdx.x] = make_double2((double)b.x, (double)b.y);
    __syncthreads();
    double mul = p.aRun;
    for (int d = 1; d < kDcThreads; d <<= 1) {
        double2 add = make_double2(0.0, 0.0);
        if ((int)threadIdx.x >= d) { add = sB[threadIdx.x - d]; }
        __syncthreads();
        if ((int)threadIdx.x >= d) { sB[threadIdx.x].x += mul * add.x; sB[threadIdx.x].y += mul * add.y; }
        __syncthreads();
        mul *= mul;
    }
    // offset entering this thread's run: a^(run*t) * chunkStart + S_{t-1}
    const double2 cs = chunkStart[blockIdx.x];
    const double at = pow(p.aRun, (double)threadIdx.x);
    double2 o = make_double2(at * cs.x, at * cs.y);
    if (threadIdx.x > 0) { o.x += sB[threadIdx.x - 1].x; o.y += sB[threadIdx.x - 1].y; }
    float ox = (float)o.x, oy = (float)o.y;
    const float sgn = conj ? -1.0f : 1.0f;
#pragma unroll
    for (int i = 0; i < kDcPerThread; i++) {
        if (base + i < n) {
            const float yx = x[i].x - ox, yy = x[i].y - oy;
            ox += yx * p.rate; oy += yy * p.rate;
            reinterpret_cast<float2*>(out)[base + i] = make_float2(yx, sgn * yy);
        }
    }
}

// ---- the same recurrence in ONE pass over the block (16 bytes of HBM traffic per sample instead of 24): chained scan with
// decoupled look-back.  CTAs take chunk numbers from a counter in issue order, so a chunk only ever waits for chunks
// that are already running or done.  A chunk (2048 samples) is staged in shared memory with coalesced 16-byte copies,
// every thread folds its 8 samples (o -> aRun * o + B_t), the CTA scans the thread folds, publishes the chunk's fold
// (flag 1), looks back over its predecessors -- a warp at a time: 32 chunks per step, each either a bare fold or an
// inclusive prefix, combined with the powers of aChunk -- until it meets an inclusive prefix, publishes its own
// inclusive prefix (flag 2), and re-walks its samples from their exact entering offsets.  Same arithmetic per sample
// as dcReduce/dcScan/dcApply above (those stay in use for b200_dc_block_noalloc).
struct DcLookback {
    int counter;             // next chunk to hand out
    int pad[3];
    // followed by: int flag[nChunks] (0 nothing, 1 fold, 2 inclusive prefix), double2 fold[nChunks], double2 prefix[nChunks]
};

__global__ void __launch_bounds__(kDcThreads, 4) dcSinglePassKernel(const cf32* __restrict__ in, cf32* __restrict__ out, int n, int nChunks, DcParams p,
                                                                 DcLookback* __restrict__ lb, int* __restrict__ flag, double2* __restrict__ foldOf,
                                                                 double2* __restrict__ prefixOf, double2* __restrict__ state, int conj) {
    const bool aligned16 = (reinterpret_cast<uintptr_t>(in) & 15) == 0;      // a view that starts on an odd sample takes 8-byte loads
    constexpr int PITCH = kDcPerThread + 1;                  // 9 float2 per thread: 72-byte stride, conflict-free 64-bit accesses
    constexpr int NW = kDcThreads / 32;
    __shared__ __align__(16) float2 tile[kDcThreads * PITCH];
    __shared__ double2 sWarp[NW];                            // inclusive fold of each warp's 32 runs
    __shared__ int sChunk;
    __shared__ double2 sEnter;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const double2 o0 = state[0];                            // offset entering the block (only chunk 0's CTA uses it)
    // powers of the slopes, once per CTA: aRun^(2^k) for the in-warp scan, aRun^32 per warp, aRun^tid for the re-walk,
    // aChunk^lane and aChunk^32 for the look-back
    double aPow[5];
    aPow[0] = p.aRun;
#pragma unroll
    for (int k = 1; k < 5; k++) { aPow[k] = aPow[k - 1] * aPow[k - 1]; }
    const double aWarp = aPow[4] * aPow[4];                 // aRun^32
    const double aLane1 = pow(p.aRun, (double)(lane + 1));  // slope of this lane's inclusive in-warp fold
    const double aTid = pow(p.aRun, (double)tid);
    const double wLane = pow(p.aChunk, (double)lane);
    const double w32 = pow(p.aChunk, 32.0);
    while (true) {
        __syncthreads();                                     // previous chunk's tile / sWarp / sChunk no longer in use
        if (tid == 0) { sChunk = atomicAdd(&lb->counter, 1); }
        __syncthreads();
        const int c = sChunk;
        if (c >= nChunks) { break; }
        const int base = c * kDcChunk;
        // ---- coalesced load: thread t brings samples 2e, 2e+1 (e = t + u * 256) into the slot of the thread that owns them
#pragma unroll
        for (int u = 0; u < kDcChunk / 2 / kDcThreads; u++) {
            const int e = tid + u * kDcThreads, i = 2 * e;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (aligned16 && base + i + 1 < n) { v = __ldcs(reinterpret_cast<const float4*>(in + base) + e); }
            else {
                if (base + i < n) { const float2 a = reinterpret_cast<const float2*>(in)[base + i]; v.x = a.x; v.y = a.y; }
                if (base + i + 1 < n) { const float2 a = reinterpret_cast<const float2*>(in)[base + i + 1]; v.z = a.x; v.w = a.y; }
            }
            float2* d = tile + (i / kDcPerThread) * PITCH + (i % kDcPerThread);
            d[0] = make_float2(v.x, v.y); d[1] = make_float2(v.z, v.w);
        }
        __syncthreads();
        float2 x[kDcPerThread];
#pragma unroll
        for (int i = 0; i < kDcPerThread; i++) { x[i] = tile[tid * PITCH + i]; }
        // ---- fold of this thread's run (zero padding keeps every slope at aRun, see dcReduceKernel), inclusive scan over the
        // warp by shuffles: S_t = aRun^d * S_{t-d} + S_t
        const float2 b = dcRunFold(x, kDcPerThread, p.rate);
        double sx = (double)b.x, sy = (double)b.y;
#pragma unroll
        for (int k = 0; k < 5; k++) {
            const int d = 1 << k;
            const double ux = __shfl_up_sync(0xffffffffu, sx, d), uy = __shfl_up_sync(0xffffffffu, sy, d);
            if (lane >= d) { sx += aPow[k] * ux; sy += aPow[k] * uy; }
        }
        if (lane == 31) { sWarp[warp] = make_double2(sx, sy); }
        __syncthreads();
        // fold of everything in front of this warp inside the chunk (serial over at most 7 warp folds)
        double px = 0.0, py = 0.0;
        for (int w = 0; w < warp; w++) { const double2 t = sWarp[w]; px = aWarp * px + t.x; py = aWarp * py + t.y; }
        // ---- publish the chunk's fold, look back for the offset entering the chunk, publish the inclusive prefix
        if (warp == NW - 1) {
            // the last warp holds the whole chunk's fold in its lane 31: B = aWarp * P + S_31
            const double Bx = __shfl_sync(0xffffffffu, aWarp * px + sx, 31), By = __shfl_sync(0xffffffffu, aWarp * py + sy, 31);
            double2 enter;
            if (c == 0) { enter = o0; }
            else {
                if (lane == 0) {
                    foldOf[c] = make_double2(Bx, By);
                    __threadfence();
                    atomicExch(&flag[c], 1);
                }
                // lane l looks at chunk q = c - 1 - l - 32 * step; weight of its value: aChunk^(l + 32 * step)
                double w = wLane;
                double2 acc = make_double2(0.0, 0.0);
                for (int q0 = c - 1; ; q0 -= 32) {
                    const int q = q0 - lane;
                    int fl = 2;                                      // lanes before the block's start read as "prefix 0": stops the walk
                    if (q >= 0) { do { fl = atomicAdd(&flag[q], 0); } while (fl == 0); }
                    __threadfence();
                    const unsigned full = __ballot_sync(0xffffffffu, fl == 2);
                    const int stop = full ? (__ffs(full) - 1) : 32;    // first lane (nearest chunk) that holds an inclusive prefix
                    double2 v = make_double2(0.0, 0.0);
                    if (q >= 0 && lane <= stop) { v = (lane == stop) ? __ldcg(prefixOf + q) : __ldcg(foldOf + q); }
                    double tx = w * v.x, ty = w * v.y;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) { tx += __shfl_xor_sync(0xffffffffu, tx, o); ty += __shfl_xor_sync(0xffffffffu, ty, o); }
                    acc.x += tx; acc.y += ty;
                    if (full) { break; }
                    w *= w32;
                }
                enter = acc;
            }
            if (lane == 0) {
                prefixOf[c] = make_double2(p.aChunk * enter.x + Bx, p.aChunk * enter.y + By);
                __threadfence();
                atomicExch(&flag[c], 2);
                sEnter = enter;
                if (c == nChunks - 1) {
                    // the last chunk was zero padded to a full chunk: undo the surplus decay a^(pad)
                    const int padN = nChunks * kDcChunk - n;
                    const double undo = pow(p.a, -(double)padN);
                    state[1] = make_double2((p.aChunk * enter.x + Bx) * undo, (p.aChunk * enter.y + By) * undo);
                }
            }
        }
        __syncthreads();
        // ---- offset entering this thread's run: aRun^t * enter + (fold of the runs in front of it); then the reference's own
        // recurrence per sample.  Fold in front = aRun^lane * (warps in front) + (this warp's lanes in front).
        const double2 cs = sEnter;
        double ex = __shfl_up_sync(0xffffffffu, sx, 1), ey = __shfl_up_sync(0xffffffffu, sy, 1);
        if (lane == 0) { ex = 0.0; ey = 0.0; }
        const double aLane = (lane == 0) ? 1.0 : aLane1 / p.aRun;      // aRun^lane
        const double ox0 = aTid * cs.x + aLane * px + ex, oy0 = aTid * cs.y + aLane * py + ey;
        float ox = (float)ox0, oy = (float)oy0;
        const float sgn = conj ? -1.0f : 1.0f;
#pragma unroll
        for (int i = 0; i < kDcPerThread; i++) {
            const float yx = x[i].x - ox, yy = x[i].y - oy;
            ox += yx * p.rate; oy += yy * p.rate;
            tile[tid * PITCH + i] = make_float2(yx, sgn * yy);
        }
        __syncthreads();
#pragma unroll
        for (int u = 0; u < kDcChunk / 2 / kDcThreads; u++) {
            const int e = tid + u * kDcThreads, i = 2 * e;
            const float2* sp = tile + (i / kDcPerThread) * PITCH + (i % kDcPerThread);
            if (base + i + 1 < n) { __stcs(reinterpret_cast<float4*>(out + base) + e, make_float4(sp[0].x, sp[0].y, sp[1].x, sp[1].y)); }
            else if (base + i < n) { reinterpret_cast<float2*>(out)[base + i] = sp[0]; }
        }
    }
}

__global__ void __launch_bounds__(256) conjKernel(const cf32* __restrict__ in, cf32* __restrict__ out, int n) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < n) {
        const float2 x = __ldg(reinterpret_cast<const float2*>(in) + i);
        reinterpret_cast<float2*>(out)[i] = make_float2(x.x, -x.y);
    }
}

} // namespace

// ------------------------------------------------------------------ front end object
// int16 I/Q -> cf32: out = (float)in / scalar, the arithmetic of VOLK's generic 16i_s32f_convert_32f
// (reference call site: source_modules/file_source/src/main.cpp:409, scalar 32768)
// ---- the SDR++ server's sample packet (dsp/compression/sample_stream_compressor.h:30-62): [u16 0][u16 pcmType][f32 scaler][data]
__global__ void streamHeaderKernel(unsigned char* out, int pcmType, float scaler) {
    reinterpret_cast<unsigned short*>(out)[0] = 0;
    reinterpret_cast<unsigned short*>(out)[1] = (unsigned short)pcmType;
    reinterpret_cast<float*>(out)[1] = scaler;
}
// largest float of the block (signed), into the packet's scaler slot.  The slot starts at -inf; a non-negative candidate
// goes in by a signed-integer max of its bit pattern, a negative one by an unsigned min (IEEE order within each sign).
__global__ void __launch_bounds__(256) streamMaxKernel(const float2* __restrict__ in, int count, float* slot) {
    float m = -INFINITY;
    for (int i = blockIdx.x * 256 + threadIdx.x; i < count; i += gridDim.x * 256) {
        const float2 v = __ldg(in + i);
        m = fmaxf(m, fmaxf(v.x, v.y));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o)); }
    __shared__ float sm[8];
    if ((threadIdx.x & 31) == 0) { sm[threadIdx.x >> 5] = m; }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int w = 1; w < 8; w++) { m = fmaxf(m, sm[w]); }
        if (m >= 0.0f) { atomicMax(reinterpret_cast<int*>(slot), __float_as_int(m)); }
        else { atomicMin(reinterpret_cast<unsigned int*>(slot), __float_as_uint(m)); }
    }
}
// volk_32f_s32f_convert_8i / _16i (generic): scale, clamp, round to nearest even
template <int TYPE>
__global__ void __launch_bounds__(256) streamConvertKernel(const float2* __restrict__ in, int count, const float* __restrict__ slot, void* __restrict__ data) {
    const float scale = __fdiv_rn(TYPE == 0 ? 128.0f : 32768.0f, *slot);
    const float hi = TYPE == 0 ? 127.0f : 32767.0f, lo = TYPE == 0 ? -128.0f : -32768.0f;
    for (int i = blockIdx.x * 256 + threadIdx.x; i < count; i += gridDim.x * 256) {
        const float2 v = __ldg(in + i);
        float a = __fmul_rn(v.x, scale), b = __fmul_rn(v.y, scale);
        a = (a > hi) ? hi : ((a < lo) ? lo : a);
        b = (b > hi) ? hi : ((b < lo) ? lo : b);
        if (TYPE == 0) { reinterpret_cast<char2*>(data)[i] = make_char2((signed char)__float2int_rn(a), (signed char)__float2int_rn(b)); }
        else { reinterpret_cast<short2*>(data)[i] = make_short2((short)__float2int_rn(a), (short)__float2int_rn(b)); }
    }
}

template <int TYPE>
__global__ void __launch_bounds__(256) toPcmKernel(const float* __restrict__ in, void* __restrict__ out, int n) {
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
        const float x = __ldg(in + i);
        if (TYPE == 0) {
            // the C conversion float -> uint8_t as x86 compilers emit it: truncate to int, keep the low byte
            reinterpret_cast<unsigned char*>(out)[i] = (unsigned char)(__float2int_rz(__fadd_rn(__fmul_rn(x, 127.0f), 128.0f)) & 0xff);
        }
        else if (TYPE == 1) {
            float r = __fmul_rn(x, 32767.0f);
            r = (r > 32767.0f) ? 32767.0f : ((r < -32768.0f) ? -32768.0f : r);
            reinterpret_cast<short*>(out)[i] = (short)__float2int_rn(r);
        }
        else {
            const float r = __fmul_rn(x, 2147483647.0f);
            reinterpret_cast<int*>(out)[i] = (r >= 2147483648.0f) ? 2147483647 : ((r < -2147483648.0f) ? (-2147483647 - 1) : __float2int_rn(r));
        }
    }
}

__global__ void __launch_bounds__(256) convertI16Kernel(const short2* __restrict__ in, float2* __restrict__ out, int n, float scalar) {
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
        const short2 v = __ldg(in + i);
        out[i] = make_float2(__fdiv_rn((float)v.x, scalar), __fdiv_rn((float)v.y, scalar));
    }
}

// int8 I/Q -> cf32 (VOLK generic 8i_s32f_convert_32f; the server source's PCM_TYPE_I8 path,
// core/src/dsp/compression/sample_stream_decompressor.h, scalar 128 / scaler)
__global__ void __launch_bounds__(256) convertI8Kernel(const char2* __restrict__ in, float2* __restrict__ out, int n, float scalar) {
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
        const char2 v = in[i];
        out[i] = make_float2(__fdiv_rn((float)v.x, scalar), __fdiv_rn((float)v.y, scalar));
    }
}

struct FrontEnd {
    int device = 0;
    double sampleRate = 0; int decimRatio = 1; bool dcBlocking = false, invertIQ = false;
    int fftSize = 0; double fftRate = 0; int fftWindow = 0;
    int maxBlock = 0;
    double effectiveSr = 0;
    b200_vfo_bank* bank = nullptr;
    cudaStream_t lastStream = nullptr; bool everRan = false;
    // the FFT branch and the VFO branch are independent consumers of the pre-processed block (the reference
    // runs them on separate threads behind the Splitter): fork the FFT branch onto a side stream and join
    cudaStream_t fftStream = nullptr;
    cudaEvent_t evFork = nullptr, evJoin = nullptr;

    // pre-processing
    std::vector<Stage> decimStages;
    cf32* dDcOut = nullptr; size_t dcCap = 0;
    double2* dDcState = nullptr; int dcChunksCap = 0;
    void* dDcLookback = nullptr; size_t dcFlagBytes = 0;      // chunk counter, flags, folds and prefixes of the single-pass DC blocker
    void* dJobDesc = nullptr; size_t jobDescBytes = 0;
    PinnedRing ring;
    const cf32* preOut = nullptr; int preCount = 0;
    // the last block of the pre-processing chain: the signal detector (iq_frontend.cpp:37-44), created when first asked for
    b200_sigdet* detector = nullptr;
    int ensureDetector();

    // FFT branch
    FftPlan plan;
    int nz = 0, skip = 0, interval = 0;
    float* dWindow = nullptr;
    int64_t streamPos = 0;       // samples of the pre-processed stream seen since the last FFT reconfiguration
    int64_t nextFrame = 0;
    cf32* dCarry = nullptr; int carryLen = 0;
    float* dRows = nullptr; int rowsCap = 0;
    FrameDesc* dFrameDesc = nullptr; int frameDescCap = 0;
    std::vector<int64_t> frameIdx;   // frames produced by the last call
    int shardRank = 0, shardWorld = 1;
    // waterfall reduction attached to the FFT branch (the reference's FFT sink thread calls WaterFall::pushFFT itself)
    b200_waterfall* wf = nullptr; float* wfZoom = nullptr; uint32_t* wfRgba = nullptr; int wfMaxRows = 0, wfWidth = 0;
    // host staging
    cf32* dStageIn = nullptr;
    short2* dStageI16 = nullptr;                // int16 I/Q as it arrives from the host (process_host_i16)
    cudaStream_t hostCopyStream = nullptr;      // H2D of the next chunks while the current one is processed
    cudaEvent_t evRowsReady = nullptr;          // a chunk's dB rows are complete in the caller's stream (host path in streaming mode)
    bool rowsOnCopyStream = false;              // rows of this call went back on hostCopyStream: the call ends by draining it
    std::vector<cudaEvent_t> hostCopyDone;
    cudaEvent_t evHostStart = nullptr;
    void* dCopyJobs = nullptr; size_t copyJobBytes = 0;                  // scatterToHostKernel's job table
    std::vector<std::pair<const void*, float2*>> hostAlias;               // per VFO: last host output pointer seen and its device alias

    int init();
    void destroy();
    int syncLast() { if (everRan) { B200_CUDA(cudaStreamSynchronize(lastStream)); } if (bank) { B200_TRY(bank->impl.syncLast()); } return B200_OK; }
    int rebuildDecim();
    int updateFftPath();
    int preprocess(const cf32* dIn, int count, cudaStream_t s);
    int fftBranch(cudaStream_t s, bool append);
    int process(const cf32* dIn, int count, cudaStream_t s, bool append = false);   // append: keep the rows of the previous call(s)
    int branches(cudaStream_t s, bool append);   // FFT branch + VFO bank on the pre-processed block
};

int FrontEnd::rebuildDecim() {
    for (auto& st : decimStages) { st.destroy(); }
    decimStages.clear();
    if (decimRatio > 1) {
        const design::PlanDef* plan = design::findPlan(decimRatio);
        if (!plan) { setError("decimation ratio %d is not a supported power of two", decimRatio); return B200_EINVAL; }
        decimStages.resize(plan->nstages);
        int cap = maxBlock;
        for (int i = 0; i < plan->nstages; i++) {
            std::vector<float> t(plan->stages[i].ntaps);
            memcpy(t.data(), plan->stages[i].taps, sizeof(float) * t.size());
            cap = cap / plan->stages[i].decim + 2;
            B200_TRY(decimStages[i].create(JOB_DECIM_FIR, t, plan->stages[i].decim, 1, 1, 2, cap, true));
        }
    }
    return B200_OK;
}

// IQFrontEnd::updateFFTPath (iq_frontend.cpp:289-325)
int FrontEnd::updateFftPath() {
    design::reshapeParams(effectiveSr, fftSize, fftRate, skip, nz);
    interval = nz + skip;
    if (nz <= 0) { setError("fft rate %g too high for sample rate %g", fftRate, effectiveSr); return B200_EINVAL; }
    if (plan.N != fftSize) { B200_TRY(plan.create(fftSize, device)); }
    std::vector<float> w = design::fftWindow(fftWindow, nz);
    if (dWindow) { cudaFree(dWindow); dWindow = nullptr; }
    B200_CUDA(cudaMalloc((void**)&dWindow, (size_t)nz * sizeof(float)));
    B200_CUDA(cudaMemcpy(dWindow, w.data(), (size_t)nz * sizeof(float), cudaMemcpyHostToDevice));
    if (dCarry) { cudaFree(dCarry); dCarry = nullptr; }
    B200_CUDA(cudaMalloc((void**)&dCarry, (size_t)nz * sizeof(cf32)));
    carryLen = 0; streamPos = 0; nextFrame = 0;
    const int maxFrames = maxBlock / std::max(interval, 1) + 2;
    if (dRows) { cudaFree(dRows); dRows = nullptr; }
    rowsCap = maxFrames;
    B200_CUDA(cudaMalloc((void**)&dRows, (size_t)rowsCap * fftSize * sizeof(float)));
    if (dFrameDesc) { cudaFree(dFrameDesc); dFrameDesc = nullptr; }
    frameDescCap = maxFrames;
    B200_CUDA(cudaMalloc((void**)&dFrameDesc, (size_t)frameDescCap * sizeof(FrameDesc) + 16));
    frameIdx.clear();
    return B200_OK;
}

int FrontEnd::init() {
    DeviceGuard g(device);
    if (!g.ok) { setError("cannot select CUDA device %d", device); return B200_ECUDA; }
    effectiveSr = sampleRate / decimRatio;
    B200_TRY(ring.init(64 * 1024));
    B200_CUDA(cudaStreamCreateWithFlags(&fftStream, cudaStreamNonBlocking));
    B200_CUDA(cudaEventCreateWithFlags(&evFork, cudaEventDisableTiming));
    B200_CUDA(cudaEventCreateWithFlags(&evJoin, cudaEventDisableTiming));
    B200_TRY(rebuildDecim());
    B200_CUDA(cudaMalloc((void**)&dDcState, 2 * sizeof(double2)));
    B200_CUDA(cudaMemset(dDcState, 0, 2 * sizeof(double2)));
    B200_TRY(updateFftPath());
    bank = new b200_vfo_bank;
    bank->ownedByFrontend = true;
    return bank->impl.init(effectiveSr, maxBlock, device);
}

void FrontEnd::destroy() {
    DeviceGuard g(device);
    if (everRan) { cudaStreamSynchronize(lastStream); }
    if (fftStream) { cudaStreamSynchronize(fftStream); cudaStreamDestroy(fftStream); fftStream = nullptr; }
    if (evFork) { cudaEventDestroy(evFork); evFork = nullptr; }
    if (evJoin) { cudaEventDestroy(evJoin); evJoin = nullptr; }
    if (hostCopyStream) { cudaStreamSynchronize(hostCopyStream); cudaStreamDestroy(hostCopyStream); hostCopyStream = nullptr; }
    for (auto e : hostCopyDone) { cudaEventDestroy(e); }
    hostCopyDone.clear();
    if (evHostStart) { cudaEventDestroy(evHostStart); evHostStart = nullptr; }
    if (evRowsReady) { cudaEventDestroy(evRowsReady); evRowsReady = nullptr; }
    if (bank) { bank->impl.destroy(); delete bank; bank = nullptr; }
    if (detector) { detector->impl.destroy(); delete detector; detector = nullptr; }
    for (auto& st : decimStages) { st.destroy(); }
    decimStages.clear();
    plan.destroy();
    ring.destroy();
    void* ptrs[] = { dDcOut, dDcLookback, dDcState, dJobDesc, dWindow, dCarry, dRows, dFrameDesc, dStageIn, dStageI16, dCopyJobs };
    for (void* p : ptrs) { if (p) { cudaFree(p); } }
}

int FrontEnd::preprocess(const cf32* dIn, int count, cudaStream_t s) {
    const cf32* cur = dIn;
    int n = count;
    // ---- PowerDecimator (multirate/power_decimator.h:51-67)
    if (decimRatio > 1 && !decimStages.empty()) {
        std::vector<FirJob> jobs;
        const void* data = cur;
        for (auto& st : decimStages) {
            FirJob j; st.fillJob(j, data, n);
            jobs.push_back(j);
            st.advance(n);
            data = st.dOut; n = j.nOut;
        }
        for (auto& fjob : jobs) { firJobsBalance(&fjob, 1); }     // each stage is its own launch here
        const size_t bytes = jobs.size() * sizeof(FirJob);
        if (!dJobDesc) { jobDescBytes = 16 * sizeof(FirJob); B200_CUDA(cudaMalloc(&dJobDesc, jobDescBytes)); }
        void* hp = nullptr; int slot = 0;
        B200_TRY(ring.acquire(bytes, &hp, &slot));
        memcpy(hp, jobs.data(), bytes);
        B200_CUDA(cudaMemcpyAsync(dJobDesc, hp, bytes, cudaMemcpyHostToDevice, s));
        B200_TRY(ring.release(slot, s));
        int maxH = 0;
        for (size_t i = 0; i < jobs.size(); i++) {
            B200_TRY(launchFirJobs((const FirJob*)dJobDesc + i, &jobs[i], 1, 2, s));
            maxH = std::max(maxH, jobs[i].H);
        }
        B200_TRY(launchHistoryUpdate((const FirJob*)dJobDesc, (int)jobs.size(), maxH, s));
        cur = (const cf32*)data;
    }
    // ---- DCBlocker (+ Conjugate folded into the same pass)
    bool conjDone = false;
    if (dcBlocking && n > 0) {
        if ((size_t)n > dcCap) {
            if (dDcOut) { B200_CUDA(cudaStreamSynchronize(s)); cudaFree(dDcOut); }
            dcCap = (size_t)std::max(n, maxBlock / std::max(decimRatio, 1) + 2);
            B200_CUDA(cudaMalloc((void**)&dDcOut, dcCap * sizeof(cf32)));
        }
        const int nChunks = (n + kDcChunk - 1) / kDcChunk;
        if (nChunks > dcChunksCap) {
            if (dDcLookback) { B200_CUDA(cudaStreamSynchronize(s)); cudaFree(dDcLookback); dDcLookback = nullptr; }
            dcChunksCap = std::max(nChunks, (int)(dcCap / kDcChunk) + 2);
            // [DcLookback][flag: int x cap, padded to 16 bytes][fold: double2 x cap][prefix: double2 x cap]
            dcFlagBytes = ((size_t)dcChunksCap * sizeof(int) + 15) & ~(size_t)15;
            B200_CUDA(cudaMalloc(&dDcLookback, sizeof(DcLookback) + dcFlagBytes + 2 * (size_t)dcChunksCap * sizeof(double2)));
        }
        DcParams p;
        p.rate = (float)(50.0 / effectiveSr);
        p.a = 1.0 - (double)p.rate;
        p.aRun = pow(p.a, (double)kDcPerThread);
        p.aChunk = pow(p.a, (double)kDcChunk);
        size_t slot = 0;
        B200_TRY(profBegin(PROF_DC, s, 16.0 * n, 6.0 * n, &slot));      // algorithmic: every sample read once, written once
        // counter and flags start at zero for every block; folds and prefixes are written before they are read
        B200_CUDA(cudaMemsetAsync(dDcLookback, 0, sizeof(DcLookback) + dcFlagBytes, s));
        {
            DcLookback* lb = (DcLookback*)dDcLookback;
            int* flag = (int*)((char*)dDcLookback + sizeof(DcLookback));
            double2* foldOf = (double2*)((char*)flag + dcFlagBytes);
            double2* prefixOf = foldOf + dcChunksCap;
            static int perSm = 0;
            if (perSm == 0) {
                B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, dcSinglePassKernel, kDcThreads, 0));
                if (perSm < 1) { perSm = 1; }
            }
            // never more CTAs than can be resident at once: a waiting CTA must not keep the chunk it waits for off the GPU
            const int grid = std::min(nChunks, 148 * perSm);
            dcSinglePassKernel<<<grid, kDcThreads, 0, s>>>(cur, dDcOut, n, nChunks, p, lb, flag, foldOf, prefixOf, dDcState, invertIQ ? 1 : 0);
            B200_CHECK_LAUNCH();
        }
        B200_TRY(profEnd(PROF_DC, slot, s));
        // state[1] -> state[0] for the next block
        B200_CUDA(cudaMemcpyAsync(dDcState, dDcState + 1, sizeof(double2), cudaMemcpyDeviceToDevice, s));
        countLaunch(1);
        cur = dDcOut;
        conjDone = invertIQ;
    }
    if (invertIQ && !conjDone && n > 0) {
        if ((size_t)n > dcCap) {
            if (dDcOut) { B200_CUDA(cudaStreamSynchronize(s)); cudaFree(dDcOut); }
            dcCap = (size_t)std::max(n, maxBlock / std::max(decimRatio, 1) + 2);
            B200_CUDA(cudaMalloc((void**)&dDcOut, dcCap * sizeof(cf32)));
        }
        conjKernel<<<(n + 255) / 256, 256, 0, s>>>(cur, dDcOut, n);
        B200_CHECK_LAUNCH();
        countLaunch();
        cur = dDcOut;
    }
    preOut = cur; preCount = n;
    return B200_OK;
}

// Reshaper framing (buffer/reshaper.h:101-128 with skip >= 0): frame f = samples [f*I, f*I + nz)
int FrontEnd::fftBranch(cudaStream_t s, bool append) {
    if (!append) { frameIdx.clear(); }
    const size_t rowBase = frameIdx.size();
    const int64_t A = streamPos, E = streamPos + preCount;
    std::vector<FrameDesc> descs;
    // a frame that ends beyond this block is stashed in the carry buffer, but only after this call's FFT
    // kernels (which may still read the carry) have been enqueued
    const cf32* stashSrc = nullptr; int stashLen = 0, stashAt = 0;
    while (true) {
        const int64_t start = nextFrame * (int64_t)interval, end = start + nz;
        const bool mine = (nextFrame % shardWorld) == shardRank;
        if (end > E) {
            if (start < E && mine) {
                const int64_t from = std::max(start, A);
                stashSrc = preOut + (from - A); stashLen = (int)(E - from); stashAt = carryLen;
                carryLen += stashLen;
            }
            break;
        }
        if (mine) {
            FrameDesc d;
            const int64_t from = std::max(start, A);
            d.src0 = dCarry; d.len0 = carryLen; d.src1 = preOut + (from - A); d.pad = 0;
            if (carryLen == 0) { d.src0 = d.src1; }
            descs.push_back(d);
            frameIdx.push_back(nextFrame);
            carryLen = 0;   // only the first frame of a call can start in the carry buffer
        }
        nextFrame++;
    }
    streamPos = E;
    int rc = B200_OK;
    if (!descs.empty()) {
        if ((int)(rowBase + descs.size()) > rowsCap) { setError("internal: %d frames exceed row capacity %d", (int)(rowBase + descs.size()), rowsCap); return B200_ESTATE; }
        void* hp = nullptr; int slot = 0;
        const size_t bytes = descs.size() * sizeof(FrameDesc);
        B200_TRY(ring.acquire(bytes, &hp, &slot));
        memcpy(hp, descs.data(), bytes);
        B200_TRY(uploadDesc(dFrameDesc, hp, bytes, s, bank->impl.descViaParams));
        rc = plan.run(dFrameDesc, (int)descs.size(), dWindow, nz, dRows + rowBase * (size_t)fftSize, nullptr, s);
        B200_TRY(ring.release(slot, s));
        if (rc == B200_OK && wf) {
            // rows of this call go to output rows rowBase.. of the attached buffers, on the FFT branch's own stream
            const int n = std::min((int)descs.size(), wfMaxRows - (int)rowBase);
            if (n > 0) {
                rc = b200_waterfall_push(wf, dRows + rowBase * (size_t)fftSize, n, wfZoom ? wfZoom + rowBase * (size_t)wfWidth : nullptr,
                                         wfRgba ? wfRgba + rowBase * (size_t)wfWidth : nullptr, s);
            }
        }
    }
    if (rc == B200_OK && stashLen > 0) {
        B200_CUDA(cudaMemcpyAsync(dCarry + stashAt, stashSrc, (size_t)stashLen * sizeof(cf32), cudaMemcpyDeviceToDevice, s));
    }
    return rc;
}

int FrontEnd::process(const cf32* dIn, int count, cudaStream_t s, bool append) {
    if (count < 0 || count > maxBlock) { setError("block of %d samples exceeds max_block %d", count, maxBlock); return B200_ETOOBIG; }
    DeviceGuard g(device);
    if (everRan && s != lastStream) { B200_TRY(syncLast()); }
    lastStream = s; everRan = true;
    B200_TRY(preprocess(dIn, count, s));
    return branches(s, append);
}

// the two consumers of the pre-processed block (preOut / preCount): FFT branch and VFO bank
int FrontEnd::ensureDetector() {
    if (detector) { return B200_OK; }
    auto* d = new (std::nothrow) b200_sigdet;
    if (!d) { return B200_ENOMEM; }
    d->ownedByFrontend = true;
    d->impl.enabled = false;                    // preproc.addBlock(&detectorPreprocessor, false)
    int rc = d->impl.init(device);
    if (rc == B200_OK) { rc = d->impl.setSampleRate(effectiveSr); }
    if (rc != B200_OK) { d->impl.destroy(); delete d; return rc; }
    detector = d;
    return B200_OK;
}

int FrontEnd::branches(cudaStream_t s, bool append) {
    // the detector is the chain's last block and passes the samples through: it reads the block the two branches read
    if (detector && detector->impl.enabled) { B200_TRY(detector->impl.process(preOut, preCount, s)); }
    const bool fork = bank->impl.liveCount() > 0;
    if (fork) {
        B200_CUDA(cudaEventRecord(evFork, s));
        // The FFT branch is enqueued first and runs beside the channelizer.  Measured alternative: starting it behind the
        // channelizer's full-rate kernel (VfoBank::evStage1) lets that HBM-bound kernel run alone at 0.94 of the copy peak
        // instead of 0.78, but the step gets 4 % slower (0.266 against 0.256 ms): the FFT passes then collide with the
        // latency-bound tail stages instead.
        if (branchOrder == 0) {
            B200_CUDA(cudaStreamWaitEvent(fftStream, evFork, 0));
            B200_TRY(fftBranch(fftStream, append));
            B200_CUDA(cudaEventRecord(evJoin, fftStream));
            B200_TRY(bank->impl.process(preOut, preCount, s));
        }
        else {
            // B200_OPT_BRANCH_ORDER 1: the FFT branch is enqueued right behind the channelizer's full-rate kernel, without
            // waiting for it, so its CTAs take over the SMs as that kernel drains and run beside the latency-bound tail
            const std::function<int()> hook = [&]() -> int {
                B200_CUDA(cudaStreamWaitEvent(fftStream, evFork, 0));
                B200_TRY(fftBranch(fftStream, append));
                B200_CUDA(cudaEventRecord(evJoin, fftStream));
                return B200_OK;
            };
            B200_TRY(bank->impl.process(preOut, preCount, s, &hook));
        }
        B200_CUDA(cudaStreamWaitEvent(s, evJoin, 0));
    }
    else {
        B200_TRY(fftBranch(s, append));
    }
    return B200_OK;
}

} // namespace b200

using namespace b200;

struct b200_frontend { FrontEnd impl; };

#define FE_OR_FAIL(f) \
    if (!(f)) { setError("NULL front-end handle"); return B200_EINVAL; }

extern "C" {

int b200_frontend_create(b200_frontend** out, double sample_rate, int decim_ratio, int dc_blocking, int invert_iq,
                         int fft_size, double fft_rate, int fft_window, int max_block, int device) {
    if (!out || sample_rate <= 0 || max_block <= 0 || fft_rate <= 0 || fft_window < 0 || fft_window > 2) {
        setError("bad arguments to b200_frontend_create"); return B200_EINVAL;
    }
    if (!design::validDecimRatio(decim_ratio)) { setError("decimation ratio %d is not a power of two <= %d", decim_ratio, design::maxDecimRatio()); return B200_EINVAL; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { setError("no CUDA device available: libb200dsp has no CPU fallback"); return B200_ECUDA; }
    if (device < 0 || device >= ndev) { setError("device %d out of range (%d devices)", device, ndev); return B200_EINVAL; }
    auto* f = new (std::nothrow) b200_frontend;
    if (!f) { return B200_ENOMEM; }
    f->impl.device = device; f->impl.sampleRate = sample_rate; f->impl.decimRatio = decim_ratio;
    f->impl.dcBlocking = dc_blocking != 0; f->impl.invertIQ = invert_iq != 0;
    f->impl.fftSize = fft_size; f->impl.fftRate = fft_rate; f->impl.fftWindow = fft_window; f->impl.maxBlock = max_block;
    int rc = f->impl.init();
    if (rc != B200_OK) { f->impl.destroy(); delete f; return rc; }
    *out = f;
    return B200_OK;
}

int b200_frontend_destroy(b200_frontend* fe) {
    FE_OR_FAIL(fe);
    fe->impl.destroy();
    delete fe;
    return B200_OK;
}

// IQFrontEnd::setSampleRate (iq_frontend.cpp:90-118): new effective rate -> DC-block rate, every VFO's
// input rate, FFT path.
int b200_frontend_set_sample_rate(b200_frontend* fe, double sample_rate) {
    FE_OR_FAIL(fe);
    if (sample_rate <= 0) { setError("bad sample rate"); return B200_EINVAL; }
    FrontEnd& f = fe->impl;
    DeviceGuard g(f.device);
    B200_TRY(f.syncLast());
    f.sampleRate = sample_rate;
    f.effectiveSr = f.sampleRate / f.decimRatio;
    B200_TRY(b200_vfo_bank_set_in_samplerate(f.bank, f.effectiveSr));
    if (f.detector) { B200_TRY(f.detector->impl.setSampleRate(f.effectiveSr)); }     // iq_frontend.cpp:102
    return f.updateFftPath();
}

// IQFrontEnd::detectorPreprocessor (signal_path/iq_frontend.h:77): part of the pre-processing chain, disabled until toggled
b200_sigdet* b200_frontend_detector(b200_frontend* fe) {
    if (!fe) { setError("NULL front-end handle"); return nullptr; }
    DeviceGuard g(fe->impl.device);
    if (fe->impl.ensureDetector() != B200_OK) { return nullptr; }
    return fe->impl.detector;
}

int b200_frontend_set_detector(b200_frontend* fe, int enabled) {
    FE_OR_FAIL(fe);
    DeviceGuard g(fe->impl.device);
    if (!enabled && !fe->impl.detector) { return B200_OK; }
    B200_TRY(fe->impl.ensureDetector());
    fe->impl.detector->impl.enabled = enabled != 0;
    return B200_OK;
}

// IQFrontEnd::setDecimation (iq_frontend.cpp:124-141)
int b200_frontend_set_decimation(b200_frontend* fe, int ratio) {
    FE_OR_FAIL(fe);
    if (!design::validDecimRatio(ratio)) { setError("decimation ratio %d is not a power of two <= %d", ratio, design::maxDecimRatio()); return B200_EINVAL; }
    FrontEnd& f = fe->impl;
    DeviceGuard g(f.device);
    B200_TRY(f.syncLast());
    f.decimRatio = ratio;
    if (ratio > 1) { B200_TRY(f.rebuildDecim()); }   // decim.setRatio() only when > 1 (:130)
    return b200_frontend_set_sample_rate(fe, f.sampleRate);
}

int b200_frontend_set_dc_blocking(b200_frontend* fe, int enabled) { FE_OR_FAIL(fe); fe->impl.dcBlocking = enabled != 0; return B200_OK; }
int b200_frontend_set_invert_iq(b200_frontend* fe, int enabled) { FE_OR_FAIL(fe); fe->impl.invertIQ = enabled != 0; return B200_OK; }

int b200_frontend_set_fft_size(b200_frontend* fe, int size) {
    FE_OR_FAIL(fe);
    FrontEnd& f = fe->impl;
    DeviceGuard g(f.device);
    B200_TRY(f.syncLast());
    const int old = f.fftSize;
    f.fftSize = size;
    int rc = f.updateFftPath();
    if (rc != B200_OK) { f.fftSize = old; f.updateFftPath(); }
    return rc;
}

int b200_frontend_set_fft_rate(b200_frontend* fe, double rate) {
    FE_OR_FAIL(fe);
    if (rate <= 0) { setError("bad fft rate"); return B200_EINVAL; }
    FrontEnd& f = fe->impl;
    DeviceGuard g(f.device);
    B200_TRY(f.syncLast());
    f.fftRate = rate;
    return f.updateFftPath();
}

int b200_frontend_set_fft_window(b200_frontend* fe, int window) {
    FE_OR_FAIL(fe);
    if (window < 0 || window > 2) { setError("bad window type"); return B200_EINVAL; }
    FrontEnd& f = fe->impl;
    DeviceGuard g(f.device);
    B200_TRY(f.syncLast());
    f.fftWindow = window;
    return f.updateFftPath();
}

double b200_frontend_effective_samplerate(b200_frontend* fe) { return fe ? fe->impl.effectiveSr : 0.0; }

int b200_frontend_set_frame_shard(b200_frontend* fe, int rank, int world) {
    FE_OR_FAIL(fe);
    if (world < 1 || rank < 0 || rank >= world) { setError("bad shard %d of %d", rank, world); return B200_EINVAL; }
    fe->impl.shardRank = rank; fe->impl.shardWorld = world;
    return B200_OK;
}

b200_vfo_bank* b200_frontend_vfo_bank(b200_frontend* fe) { return fe ? fe->impl.bank : nullptr; }

int b200_frontend_attach_waterfall(b200_frontend* fe, b200_waterfall* wf, int data_width, float* d_zoom, uint32_t* d_rgba, int max_rows) {
    FE_OR_FAIL(fe);
    if (wf && (data_width <= 0 || max_rows <= 0 || (!d_zoom && !d_rgba))) { setError("bad arguments to b200_frontend_attach_waterfall"); return B200_EINVAL; }
    FrontEnd& f = fe->impl;
    DeviceGuard g(f.device);
    B200_TRY(f.syncLast());
    f.wf = wf; f.wfZoom = d_zoom; f.wfRgba = d_rgba; f.wfMaxRows = wf ? max_rows : 0; f.wfWidth = data_width;
    return B200_OK;
}

int b200_frontend_process(b200_frontend* fe, const float* d_in, int count, void* stream) {
    FE_OR_FAIL(fe);
    if (!d_in && count > 0) { setError("NULL input"); return B200_EINVAL; }
    return fe->impl.process((const cf32*)d_in, count, (cudaStream_t)stream);
}

int b200_frontend_preproc_out(b200_frontend* fe, const float** d_out, int* count) {
    FE_OR_FAIL(fe);
    if (d_out) { *d_out = (const float*)fe->impl.preOut; }
    if (count) { *count = fe->impl.preCount; }
    return B200_OK;
}

int b200_frontend_frame_count(b200_frontend* fe) { FE_OR_FAIL(fe); return (int)fe->impl.frameIdx.size(); }

const float* b200_frontend_frame_ptr(b200_frontend* fe, int i) {
    if (!fe || i < 0 || i >= (int)fe->impl.frameIdx.size()) { return nullptr; }
    return fe->impl.dRows + (size_t)i * fe->impl.fftSize;
}

int64_t b200_frontend_frame_index(b200_frontend* fe, int i) {
    if (!fe || i < 0 || i >= (int)fe->impl.frameIdx.size()) { return -1; }
    return fe->impl.frameIdx[i];
}

// Many small results back to the host in one go: when every destination is page-locked (mapped) host memory, one
// kernel stores all VFO outputs of a chunk straight into the callers' buffers over PCIe -- a bank of a thousand VFOs
// would otherwise pay a thousand small cudaMemcpyAsync calls per chunk, each with its own engine latency.
struct HostCopyJob { const float2* src; float2* dst; int n; int pad; };
static __global__ void __launch_bounds__(256) scatterToHostKernel(const HostCopyJob* __restrict__ jobs) {
    const HostCopyJob j = jobs[blockIdx.x];
    for (int i = threadIdx.x; i < j.n; i += 256) { j.dst[i] = j.src[i]; }
}
// device-visible alias of a host pointer, or NULL when the memory is pageable (plain malloc)
static float2* mappedAlias(const void* h) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, h) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (a.type != cudaMemoryTypeHost || !a.devicePointer) { return nullptr; }
    return reinterpret_cast<float2*>(a.devicePointer);
}

// D2H of what the last f.process() call produced; host buffers are filled from `framesDone` / `done[v]` on
static int copyOutputs(FrontEnd& f, float* h_frames, int max_frames, int& framesDone, float* const* h_vfo_out, float* const* h_audio,
                       std::vector<int>& done, std::vector<int>& audioDone, cudaStream_t s) {
    const int total = std::min((int)f.frameIdx.size(), max_frames);
    VfoBank& b = f.bank->impl;
    // the chunk's dB rows; `rowsFrom` = the stream they leave on
    bool rowsPending = h_frames && total > framesDone;
    auto copyRows = [&](cudaStream_t rowsFrom) -> int {
        if (!rowsPending) { return B200_OK; }
        rowsPending = false;
        if (rowsFrom != s) {
            // the rows were produced in the caller's stream
            if (!f.evRowsReady) { B200_CUDA(cudaEventCreateWithFlags(&f.evRowsReady, cudaEventDisableTiming)); }
            B200_CUDA(cudaEventRecord(f.evRowsReady, s));
            B200_CUDA(cudaStreamWaitEvent(rowsFrom, f.evRowsReady, 0));
        }
        B200_CUDA(cudaMemcpyAsync(h_frames + (size_t)framesDone * f.fftSize, f.dRows + (size_t)framesDone * f.fftSize,
                                  (size_t)(total - framesDone) * f.fftSize * sizeof(float), cudaMemcpyDeviceToHost, rowsFrom));
        if (rowsFrom != s && rowsFrom != b.tailStream) { f.rowsOnCopyStream = true; }
        return B200_OK;
    };
    // Streaming mode (B200_OPT_DEFER_TAIL) with the outputs of a large bank going back through the scatter kernel: the
    // kernel goes on the bank's tail stream, right behind the later stages that produce what it copies and in front of the
    // next chunk's (which overwrite it), so the caller's stream -- the next chunk's full-rate kernels -- never waits for a
    // chunk's results to cross PCIe.  The tail event is recorded again behind the kernel: whoever joins the tail joins it.
    // The rows go the same way, so that the caller's stream holds nothing but the chunks' kernels.
    const bool onTail = b.tailUnjoined && b.tailStream && h_vfo_out && !h_audio && b.liveCount() >= 16;
    cudaStream_t cs = onTail ? b.tailStream : s;
    if (!onTail) { B200_TRY(copyRows(s)); B200_TRY(b.joinTail()); }
    // a large bank: gather the (device pointer, mapped host pointer, count) triples and let one kernel do the copies
    bool scattered = false;
    if (h_vfo_out && b.liveCount() >= 16) {
        std::vector<HostCopyJob> jobs;
        jobs.reserve(b.vfos.size());
        bool allMapped = true;
        if (f.hostAlias.size() != b.vfos.size()) { f.hostAlias.assign(b.vfos.size(), std::make_pair((const void*)nullptr, (float2*)nullptr)); }
        for (size_t i = 0; i < b.vfos.size() && allMapped; i++) {
            const Vfo& v = b.vfos[i];
            if (!v.live || v.lastOut <= 0 || !h_vfo_out[i]) { continue; }
            if (f.hostAlias[i].first != (const void*)h_vfo_out[i]) { f.hostAlias[i] = std::make_pair((const void*)h_vfo_out[i], mappedAlias(h_vfo_out[i])); }
            float2* alias = f.hostAlias[i].second;
            if (!alias) { allMapped = false; break; }
            HostCopyJob j; j.src = reinterpret_cast<const float2*>(v.lastOutPtr); j.dst = alias + done[i]; j.n = v.lastOut; j.pad = 0;
            jobs.push_back(j);
        }
        if (allMapped && !jobs.empty()) {
            const size_t bytes = jobs.size() * sizeof(HostCopyJob);
            if (bytes > f.copyJobBytes) {
                if (f.dCopyJobs) { B200_CUDA(cudaStreamSynchronize(s)); B200_TRY(b.syncLast()); cudaFree(f.dCopyJobs); f.dCopyJobs = nullptr; }
                f.copyJobBytes = std::max(bytes * 2, (size_t)64 * 1024);
                B200_CUDA(cudaMalloc(&f.dCopyJobs, f.copyJobBytes + 16));
            }
            void* hp = nullptr; int slot = 0;
            B200_TRY(f.ring.acquire(bytes, &hp, &slot));
            memcpy(hp, jobs.data(), bytes);
            // (in streaming mode the rows leave on the stream that brought the block in: its copies are long done, and the
            // rows then cross beside the scatter kernel's stores instead of in front of them)
            B200_TRY(copyRows(onTail && f.hostCopyStream ? f.hostCopyStream : cs));
            B200_TRY(uploadDesc(f.dCopyJobs, hp, bytes, cs, onTail ? true : b.descViaParams));
            B200_TRY(f.ring.release(slot, cs));
            scatterToHostKernel<<<(int)jobs.size(), 256, 0, cs>>>((const HostCopyJob*)f.dCopyJobs);
            B200_CHECK_LAUNCH(); countLaunch();
            if (onTail) { B200_CUDA(cudaEventRecord(b.evTail[b.lastTailPar], cs)); }
            scattered = true;
        }
    }
    if (onTail && !scattered) { B200_TRY(b.joinTail()); }      // (pageable destinations: plain copies on the caller's stream)
    B200_TRY(copyRows(s));
    framesDone = std::max(framesDone, total);
    for (size_t i = 0; i < b.vfos.size(); i++) {
        const Vfo& v = b.vfos[i];
        if (!v.live) { continue; }
        if (!scattered && v.lastOut > 0 && h_vfo_out && h_vfo_out[i]) {
            B200_CUDA(cudaMemcpyAsync(h_vfo_out[i] + 2 * (size_t)done[i], v.lastOutPtr, (size_t)v.lastOut * sizeof(cf32), cudaMemcpyDeviceToHost, s));
        }
        done[i] += v.lastOut;
        if (h_audio && h_audio[i]) {
            const float* ap = b200_vfo_audio_ptr(f.bank, (int)i);
            const int an = ap ? b200_vfo_audio_count(f.bank, (int)i) : 0;   // differs from lastOut behind an audio resampler
            if (an > 0) { B200_CUDA(cudaMemcpyAsync(h_audio[i] + 2 * (size_t)audioDone[i], ap, (size_t)an * 2 * sizeof(float), cudaMemcpyDeviceToHost, s)); }
            audioDone[i] += an;
        }
    }
    return B200_OK;
}

// sampleBytes = 8: h_in holds cf32; sampleBytes = 4: h_in holds interleaved int16 I/Q, converted on the device
// (out = (float)in / scalar, the file source's volk_16i_s32f_convert_32f, file_source/src/main.cpp:409)
static int processHostImpl(b200_frontend* fe, const void* h_in, int sampleBytes, float scalar, int count, float* h_frames, int max_frames,
                           float* const* h_vfo_out, float* const* h_audio, int* vfo_counts, void* stream) {
    FE_OR_FAIL(fe);
    FrontEnd& f = fe->impl;
    if (count < 0 || count > f.maxBlock) { setError("block of %d samples exceeds max_block %d", count, f.maxBlock); return B200_ETOOBIG; }
    if (!h_in && count > 0) { setError("NULL input"); return B200_EINVAL; }
    DeviceGuard g(f.device);
    cudaStream_t s = (cudaStream_t)stream;
    if (!f.dStageIn) { B200_CUDA(cudaMalloc((void**)&f.dStageIn, (size_t)f.maxBlock * sizeof(cf32))); }
    const bool i16 = sampleBytes != 8;          // integer I/Q (int16: 4 bytes per sample, int8: 2), staged raw and converted on the device
    if (i16 && !f.dStageI16) { B200_CUDA(cudaMalloc((void**)&f.dStageI16, (size_t)f.maxBlock * sizeof(short2))); }
    if (i16 && !(scalar != 0.0f)) { setError("integer sample scale must not be zero"); return B200_EINVAL; }
    const char* hIn = (const char*)h_in;
    // brings chunk [off, off+n) to dStageIn: the copy goes on `cs`, the int16 -> float conversion (if any) on `ks`
    auto stageCopy = [&](size_t off, int n, cudaStream_t cs) -> int {
        void* dst = i16 ? (void*)((char*)f.dStageI16 + off * sampleBytes) : (void*)(f.dStageIn + off);
        B200_CUDA(cudaMemcpyAsync(dst, hIn + off * sampleBytes, (size_t)n * sampleBytes, cudaMemcpyHostToDevice, cs));
        return B200_OK;
    };
    auto stageConvert = [&](size_t off, int n, cudaStream_t ks) -> int {
        if (!i16 || n <= 0) { return B200_OK; }
        if (sampleBytes == 4) {
            convertI16Kernel<<<std::min(148 * 8, (n + 255) / 256), 256, 0, ks>>>(f.dStageI16 + off, reinterpret_cast<float2*>(f.dStageIn + off), n, scalar);
        }
        else {
            convertI8Kernel<<<std::min(148 * 8, (n + 255) / 256), 256, 0, ks>>>(reinterpret_cast<const char2*>(f.dStageI16) + off, reinterpret_cast<float2*>(f.dStageIn + off), n, scalar);
        }
        B200_CHECK_LAUNCH(); countLaunch();
        return B200_OK;
    };
    VfoBank& b = f.bank->impl;
    std::vector<int> done(b.vfos.size(), 0), audioDone(b.vfos.size(), 0);
    int framesDone = 0;
    // A long block is cut into chunks: the copy engine brings chunk k+1 (and the ones behind it) in while chunk k is
    // processed and its results go back over the other PCIe direction.  The stages are streaming (results do not
    // depend on how a block is cut), and the dB rows of all chunks stay in the row buffer.  Only without
    // pre-processing, so that b200_frontend_preproc_out still names the whole block afterwards.
    // 16 MiB of host data per chunk (2^21 cf32 or 2^22 int16 samples) when the block's kernels are short beside its copy;
    // a heavy bank (its kernels take several times the copy) runs 64 MiB chunks, so that the per-call fixed costs of
    // its later stages are paid four times less often
    const int kChunk = ((hostChunkMiB > 0 ? hostChunkMiB : (f.bank->impl.liveCount() >= 64 ? 64 : 16)) << 20) / sampleBytes;
    const bool plain = f.decimRatio == 1 && !f.dcBlocking && !f.invertIQ;
    if (plain && count >= 2 * kChunk) {
        // (chunks are multiples of 512 samples: an exact-mode rotator renormalises every 512 samples of a call, as the reference's
        // does, so only such cuts leave its phases what one call over the whole block gives)
        const int nChunks = (count + kChunk - 1) / kChunk;
        if (!f.hostCopyStream) {
            B200_CUDA(cudaStreamCreateWithFlags(&f.hostCopyStream, cudaStreamNonBlocking));
            B200_CUDA(cudaEventCreateWithFlags(&f.evHostStart, cudaEventDisableTiming));
        }
        while ((int)f.hostCopyDone.size() < nChunks) {
            cudaEvent_t e;
            B200_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            f.hostCopyDone.push_back(e);
        }
        // the staging buffer may still be read by work already queued on s
        B200_CUDA(cudaEventRecord(f.evHostStart, s));
        B200_CUDA(cudaStreamWaitEvent(f.hostCopyStream, f.evHostStart, 0));
        for (int c = 0; c < nChunks; c++) {
            const size_t off = (size_t)c * kChunk;
            const int n = (int)std::min<size_t>(kChunk, (size_t)count - off);
            B200_TRY(stageCopy(off, n, f.hostCopyStream));
            B200_CUDA(cudaEventRecord(f.hostCopyDone[c], f.hostCopyStream));
        }
        b.descViaParams = true;
        int rc = B200_OK;
        for (int c = 0; c < nChunks && rc == B200_OK; c++) {
            const size_t off = (size_t)c * kChunk;
            const int n = (int)std::min<size_t>(kChunk, (size_t)count - off);
            if (cudaStreamWaitEvent(s, f.hostCopyDone[c], 0) != cudaSuccess) { setError("cudaStreamWaitEvent failed"); rc = B200_ECUDA; break; }
            rc = stageConvert(off, n, s);
            if (rc == B200_OK) { rc = f.process(f.dStageIn + off, n, s, c > 0); }
            if (rc == B200_OK) { rc = copyOutputs(f, h_frames, max_frames, framesDone, h_vfo_out, h_audio, done, audioDone, s); }
        }
        b.descViaParams = false;
        if (rc != B200_OK) { cudaStreamSynchronize(f.hostCopyStream); return rc; }
        f.preOut = f.dStageIn; f.preCount = count;
    }
    else {
        B200_TRY(stageCopy(0, count, s));
        B200_TRY(stageConvert(0, count, s));
        B200_TRY(f.process(f.dStageIn, count, s));
        B200_TRY(copyOutputs(f, h_frames, max_frames, framesDone, h_vfo_out, h_audio, done, audioDone, s));
    }
    if (vfo_counts) { for (size_t i = 0; i < b.vfos.size(); i++) { if (b.vfos[i].live) { vfo_counts[i] = done[i]; } } }
    B200_TRY(b.joinTail());        // streaming mode: the last chunk's later stages and its copies back
    if (f.rowsOnCopyStream) { f.rowsOnCopyStream = false; B200_CUDA(cudaStreamSynchronize(f.hostCopyStream)); }
    B200_CUDA(cudaStreamSynchronize(s));
    return B200_OK;
}

int b200_frontend_process_host(b200_frontend* fe, const float* h_in, int count, float* h_frames, int max_frames,
                               float* const* h_vfo_out, float* const* h_audio, int* vfo_counts, void* stream) {
    return processHostImpl(fe, h_in, 8, 1.0f, count, h_frames, max_frames, h_vfo_out, h_audio, vfo_counts, stream);
}

int b200_frontend_process_host_i16(b200_frontend* fe, const int16_t* h_in, int count, float scalar, float* h_frames, int max_frames,
                                   float* const* h_vfo_out, float* const* h_audio, int* vfo_counts, void* stream) {
    return processHostImpl(fe, h_in, 4, scalar, count, h_frames, max_frames, h_vfo_out, h_audio, vfo_counts, stream);
}

int b200_frontend_process_host_i8(b200_frontend* fe, const int8_t* h_in, int count, float scalar, float* h_frames, int max_frames,
                                  float* const* h_vfo_out, float* const* h_audio, int* vfo_counts, void* stream) {
    return processHostImpl(fe, h_in, 2, scalar, count, h_frames, max_frames, h_vfo_out, h_audio, vfo_counts, stream);
}

int b200_convert_i16(const int16_t* d_in, int count, float scalar, float* d_out, void* stream) {
    if (count < 0 || (count > 0 && (!d_in || !d_out)) || !(scalar != 0.0f)) { setError("bad arguments to b200_convert_i16"); return B200_EINVAL; }
    if (count == 0) { return B200_OK; }
    convertI16Kernel<<<std::min(148 * 8, (count + 255) / 256), 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const short2*>(d_in), reinterpret_cast<float2*>(d_out), count, scalar);
    B200_CHECK_LAUNCH(); countLaunch();
    return B200_OK;
}

// wav::Writer::write (utils/wav.cpp:148-178), the recorder's float -> PCM conversion, on the device so that recorded audio or
// baseband crosses PCIe at its file size: uint8 (x*127)+128 truncated, int16 / int32 as VOLK's generic convert (scale by
// 32767 / 2147483647, clamp, round to nearest even).
int b200_convert_pcm(const float* d_in, void* d_out, int n_floats, int sample_type, void* stream) {
    if (n_floats < 0 || (n_floats > 0 && (!d_in || !d_out))) { setError("bad arguments to b200_convert_pcm"); return B200_EINVAL; }
    if (sample_type != B200_PCM_UINT8 && sample_type != B200_PCM_INT16 && sample_type != B200_PCM_INT32) {
        setError("b200_convert_pcm converts to uint8, int16 or int32 (float32 files take the samples as they are)"); return B200_EINVAL;
    }
    if (n_floats == 0) { return B200_OK; }
    const int grid = std::min(148 * 8, (n_floats + 255) / 256);
    cudaStream_t s = (cudaStream_t)stream;
    if (sample_type == B200_PCM_UINT8) { toPcmKernel<0><<<grid, 256, 0, s>>>(d_in, d_out, n_floats); }
    else if (sample_type == B200_PCM_INT16) { toPcmKernel<1><<<grid, 256, 0, s>>>(d_in, d_out, n_floats); }
    else { toPcmKernel<2><<<grid, 256, 0, s>>>(d_in, d_out, n_floats); }
    B200_CHECK_LAUNCH(); countLaunch();
    return B200_OK;
}

// SampleStreamCompressor::process (compression/sample_stream_compressor.h:30-62) on the device: the server's sample packet is
// built in d_out (8 + count * {2, 4, 8} bytes for int8 / int16 / float32), ready for one D2H copy at its wire size.
int b200_stream_compress(const float* d_in, int count, int pcm_type, void* d_out, void* stream) {
    if (count < 0 || !d_out || (count > 0 && !d_in) || pcm_type < B200_STREAM_PCM_I8 || pcm_type > B200_STREAM_PCM_F32) {
        setError("bad arguments to b200_stream_compress"); return B200_EINVAL;
    }
    if (count > (1 << 28) - 1) { setError("block of %d samples is too long for one packet", count); return B200_ETOOBIG; }
    cudaStream_t s = (cudaStream_t)stream;
    unsigned char* out = (unsigned char*)d_out;
    if (pcm_type == B200_STREAM_PCM_F32) {
        streamHeaderKernel<<<1, 1, 0, s>>>(out, pcm_type, 0.0f);
        B200_CHECK_LAUNCH(); countLaunch();
        if (count > 0) { B200_CUDA(cudaMemcpyAsync(out + 8, d_in, (size_t)count * 8, cudaMemcpyDeviceToDevice, s)); }
        return 8 + count * 8;
    }
    // `maxVal = count == 0 ? 1.0 : in[maxIdx]` (:48-50)
    streamHeaderKernel<<<1, 1, 0, s>>>(out, pcm_type, count == 0 ? 1.0f : -INFINITY);
    B200_CHECK_LAUNCH(); countLaunch();
    if (count == 0) { return 8; }
    const int grid = std::min(148 * 8, (count + 255) / 256);
    streamMaxKernel<<<grid, 256, 0, s>>>(reinterpret_cast<const float2*>(d_in), count, reinterpret_cast<float*>(out) + 1);
    B200_CHECK_LAUNCH();
    if (pcm_type == B200_STREAM_PCM_I8) { streamConvertKernel<0><<<grid, 256, 0, s>>>(reinterpret_cast<const float2*>(d_in), count, reinterpret_cast<const float*>(out) + 1, out + 8); }
    else { streamConvertKernel<1><<<grid, 256, 0, s>>>(reinterpret_cast<const float2*>(d_in), count, reinterpret_cast<const float*>(out) + 1, out + 8); }
    B200_CHECK_LAUNCH(); countLaunch(2);
    return 8 + count * (pcm_type == B200_STREAM_PCM_I8 ? 2 : 4);
}

// complex spectrum of one already-windowed frame (tests of the FFT alone)
int b200_fft_forward(const float* d_in, int nz, int fft_size, float* d_out, void* stream) {
    if (!d_in || !d_out || nz <= 0 || nz > fft_size) { setError("bad arguments to b200_fft_forward"); return B200_EINVAL; }
    int dev = 0;
    B200_CUDA(cudaGetDevice(&dev));
    FftPlan plan;
    int rc = plan.create(fft_size, dev);
    if (rc != B200_OK) { return rc; }
    FrameDesc d; d.src0 = (const cf32*)d_in; d.src1 = (const cf32*)d_in; d.len0 = nz; d.pad = 0;
    FrameDesc* dd = nullptr;
    cudaStream_t s = (cudaStream_t)stream;
    if (cudaMalloc((void**)&dd, sizeof(FrameDesc)) != cudaSuccess) { plan.destroy(); setError("cudaMalloc failed"); return B200_ENOMEM; }
    cudaMemcpyAsync(dd, &d, sizeof(FrameDesc), cudaMemcpyHostToDevice, s);
    rc = plan.run(dd, 1, nullptr, nz, nullptr, (cf32*)d_out, s);
    cudaStreamSynchronize(s);
    cudaFree(dd);
    plan.destroy();
    return rc;
}

} // extern "C"

// ------------------------------------------------------------------------------------------------------
// DC blocker for callers that must not allocate.
//
// Some hosts of this library are not allowed to call cudaMalloc once they are running: an audio / DSP
// callback thread with a hard deadline (cudaMalloc may synchronise the whole device and has no latency
// bound), a plug-in sandbox that hands each module a fixed memory budget at load time, or a process whose
// device memory is already owned in full by another allocator's pool.  b200_frontend_process grows its
// scan scratch with cudaMalloc when a larger block arrives; this entry point runs the same three kernels
// (dcReduceKernel -> dcScanKernel -> dcApplyKernel) with the scratch in a fixed array that is part of the
// library image, so it never allocates or frees anything, on the host or on the device.
//
// The array is per GPU: every device that loads the library carries its ~256 KB until the process exits.
// Inside host code the array's name is only a host-side handle, not a device address, so its address is
// fetched once per device with cudaGetSymbolAddress (return code checked) and cached; that happens on the
// first call made while that device is current.  Make that first call before starting any stream capture.
#define B200_DC_NOALLOC_MAX_CHUNKS (B200_DC_NOALLOC_MAX_SAMPLES / 2048)
static_assert(B200_DC_NOALLOC_MAX_SAMPLES % 2048 == 0, "limit must be a whole number of scan chunks");

namespace {
struct DcNoAllocScratch {
    double2 state[2];                                   // [0] offset entering the block, [1] offset leaving it
    double2 pad[14];                                    // keep the arrays below 256-byte aligned
    double2 chunkB[B200_DC_NOALLOC_MAX_CHUNKS];         // per-chunk fold, written by dcReduceKernel
    double2 chunkStart[B200_DC_NOALLOC_MAX_CHUNKS];     // per-chunk entry offset, written by dcScanKernel
};
}
static __device__ __align__(256) unsigned char g_dcNoAllocScratch[sizeof(DcNoAllocScratch)];

static std::mutex g_dcNoAllocMtx;
static DcNoAllocScratch* g_dcNoAllocPtr[64] = {};

// device address of the scratch on the current device (looked up once per device)
static int dcNoAllocScratch(DcNoAllocScratch** out) {
    int dev = -1;
    B200_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { setError("b200_dc_block_noalloc: unsupported device ordinal %d", dev); return B200_EINVAL; }
    std::lock_guard<std::mutex> lk(g_dcNoAllocMtx);
    if (!g_dcNoAllocPtr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_dcNoAllocScratch);
        if (e != cudaSuccess || !p) {
            setError("b200_dc_block_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            return B200_ECUDA;
        }
        g_dcNoAllocPtr[dev] = static_cast<DcNoAllocScratch*>(p);
    }
    *out = g_dcNoAllocPtr[dev];
    return B200_OK;
}

extern "C" int b200_dc_block_noalloc(const float* d_in, float* d_out, int count, double rate, double* d_offset, void* stream) {
    if (count < 0 || (!d_in && count > 0) || (!d_out && count > 0)) { setError("b200_dc_block_noalloc: bad arguments"); return B200_EINVAL; }
    if (count > B200_DC_NOALLOC_MAX_SAMPLES) {
        setError("b200_dc_block_noalloc: block of %d samples exceeds the fixed scratch limit of %d samples; split the block "
                 "(pass d_offset to carry the state) or use b200_frontend_process", count, B200_DC_NOALLOC_MAX_SAMPLES);
        return B200_ETOOBIG;
    }
    // the scan undoes the zero padding of the last chunk with a^-pad (pad < 2048): a must stay well away from 0
    if (!(rate >= 0.0 && rate <= 0.25)) {
        setError("b200_dc_block_noalloc: rate %g is outside [0, 0.25] (IQFrontEnd uses 50 / sampleRate)", rate);
        return B200_EINVAL;
    }
    DcNoAllocScratch* sc = nullptr;
    B200_TRY(dcNoAllocScratch(&sc));
    if (count == 0) { return B200_OK; }
    cudaStream_t s = (cudaStream_t)stream;
    // The array keeps whatever the previous call left in it.  Only state[0] is read before it is written, so
    // it is the one thing that has to be set: the caller's carried offset, or zero for a fresh blocker.
    // chunkB / chunkStart entries are always written by this call's own kernels before they are read.
    if (d_offset) { B200_CUDA(cudaMemcpyAsync(&sc->state[0], d_offset, sizeof(double2), cudaMemcpyDeviceToDevice, s)); }
    else { B200_CUDA(cudaMemsetAsync(&sc->state[0], 0, sizeof(double2), s)); }
    DcParams p;
    p.rate = (float)rate;                                // DCBlocker::_rate is float (dc_blocker.h:74)
    p.a = 1.0 - (double)p.rate;
    p.aRun = pow(p.a, (double)kDcPerThread);
    p.aChunk = pow(p.a, (double)kDcChunk);
    const int nChunks = (count + kDcChunk - 1) / kDcChunk;
    const cf32* in = (const cf32*)d_in;
    dcReduceKernel<<<nChunks, kDcThreads, 0, s>>>(in, count, p, sc->chunkB);
    B200_CHECK_LAUNCH();
    dcScanKernel<<<1, 1024, 0, s>>>(sc->chunkB, nChunks, count, p, sc->chunkStart, sc->state);
    B200_CHECK_LAUNCH();
    dcApplyKernel<<<nChunks, kDcThreads, 0, s>>>(in, (cf32*)d_out, count, p, sc->chunkStart, 0);
    B200_CHECK_LAUNCH();
    countLaunch(3);
    if (d_offset) { B200_CUDA(cudaMemcpyAsync(d_offset, &sc->state[1], sizeof(double2), cudaMemcpyDeviceToDevice, s)); }
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------------
// Front-end group: one IQ front end spread over several GPUs of one node, inside ONE process.
//
// The reference fans every pre-processed block out to its consumers with Splitter::run (routing/splitter.h:74-83,
// bound in IQFrontEnd: iq_frontend.cpp:46,79,159-179).  Here member 0 (the ingest GPU) runs the pre-processing
// chain, the block is broadcast to the other members over NVLink, every member channelizes its own share of the VFO
// bank (VFOs are dealt out by cost when they are added) and computes the FFT frames f with f % N == member.  The dB rows
// travel back to the ingest GPU, which is the display owner: there they sit in frame order, so the waterfall
// reduction and the frame-order recurrences behind it (smoothing, peak hold: waterfall.cpp:1247-1273) see every frame.
//
// Transport: ncclBroadcast from libnccl (loaded with dlopen; one communicator per member, one call group per block --
// NCCL sends every byte out of the ingest GPU once, through NVSwitch multicast where it is available); when libnccl
// cannot be loaded, cudaMemcpyPeerAsync from the ingest GPU to every member.  Receive buffers are double buffered,
// so the broadcast of block k + 1 runs beside the kernels of block k.  Integer sources are broadcast as they arrive
// (4 or 2 bytes per sample) and converted on every member.
#include <dlfcn.h>
#include <nccl.h>

namespace b200 {
namespace {

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*commInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*commDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*groupStart)() = nullptr;
    ncclResult_t (*groupEnd)() = nullptr;
    ncclResult_t (*broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*errorString)(ncclResult_t) = nullptr;
    bool ok = false;
};

NcclApi& ncclApi() {
    static NcclApi a;
    static std::once_flag once;
    std::call_once(once, [] {
        if (getenv("B200_GROUP_NO_NCCL")) { return; }
        a.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
        if (!a.lib) { return; }
        a.commInitAll = (decltype(a.commInitAll))dlsym(a.lib, "ncclCommInitAll");
        a.commDestroy = (decltype(a.commDestroy))dlsym(a.lib, "ncclCommDestroy");
        a.groupStart = (decltype(a.groupStart))dlsym(a.lib, "ncclGroupStart");
        a.groupEnd = (decltype(a.groupEnd))dlsym(a.lib, "ncclGroupEnd");
        a.broadcast = (decltype(a.broadcast))dlsym(a.lib, "ncclBroadcast");
        a.errorString = (decltype(a.errorString))dlsym(a.lib, "ncclGetErrorString");
        a.ok = a.commInitAll && a.commDestroy && a.groupStart && a.groupEnd && a.broadcast && a.errorString;
    });
    return a;
}

#define B200_NCCL(call)                                                                                          \
    do {                                                                                                         \
        ncclResult_t r__ = (call);                                                                               \
        if (r__ != ncclSuccess) { setError("%s failed: %s", #call, ncclApi().errorString(r__)); return B200_ECUDA; } \
    } while (0)

// FP32 operations per input sample of one VFO (what the dealing-out balances; same model as sharding.vfo_cost)
double vfoCost(double inSr, double outSr, double bw) {
    const double ratio = std::max(1.0, inSr / outSr);
    const double stage1 = (ratio >= 16.0) ? 12.0 : 4.0 + 2.0 * 35.0;
    const double chanTaps = (bw != outSr) ? 3.8 * outSr / (bw / 20.0) : 0.0;
    return stage1 + (2.0 * chanTaps + 2.0 * 130.0) * outSr / inSr;
}

} // namespace

struct FrontEndGroup {
    int n = 0;
    std::vector<int> devices;
    std::vector<b200_frontend*> fe;                 // fe[0] = ingest GPU / display owner
    bool useNccl = false;
    std::vector<ncclComm_t> comms;
    std::vector<cudaStream_t> commStream, procStream;
    std::vector<void*> rx[2];                       // receive buffers of the members (index 0 unused)
    size_t rxBytes = 0;
    std::vector<cudaEvent_t> evReady[2], evFree[2], evRows;
    bool freeArmed[2] = { false, false };
    cudaEvent_t evPre = nullptr, evB0 = nullptr, evB1 = nullptr, evOwner = nullptr;
    bool bTimed = false;
    int par = 0;
    int maxBlock = 0, fftSize = 0;
    double effectiveSr = 0;
    // VFOs: group id -> (member, index in the member's bank)
    struct Loc { int member = -1, local = -1; double cost = 0; };
    std::vector<Loc> vfos;
    std::vector<double> load;
    // display owner: rows of the last call in frame order
    float* dRowsAll = nullptr; int rowsCap = 0;
    std::vector<int64_t> frameIdx;
    b200_waterfall* wf = nullptr; float* wfZoom = nullptr; uint32_t* wfRgba = nullptr; int wfMaxRows = 0, wfWidth = 0;
    // host staging on the ingest GPU
    void* dStage = nullptr; size_t stageBytes = 0;
    cf32* dConv = nullptr; size_t convSamples = 0;    // converted block of an integer source on the ingest GPU

    FrontEnd& member(int m) { return fe[m]->impl; }
    int create(const int* devs, int nDev, double sampleRate, int decimRatio, int dcBlocking, int invertIq, int fftSize, double fftRate,
               int fftWindow, int maxBlock);
    void destroy();
    int ensureRx(size_t bytes);
    int broadcast(const void* src, size_t bytes, cudaStream_t s);       // ingest GPU's `src` -> rx[par][m] of every member
    int run(const cf32* rootBlock, int rawKind, float scalar, int count, cudaStream_t s);
};

int FrontEndGroup::create(const int* devs, int nDev, double sampleRate, int decimRatio, int dcBlocking, int invertIq, int fftSize_, double fftRate,
                          int fftWindow, int maxBlock_) {
    n = nDev; maxBlock = maxBlock_; fftSize = fftSize_;
    devices.assign(devs, devs + nDev);
    fe.assign(n, nullptr);
    for (int m = 0; m < n; m++) {
        // the ingest GPU owns the pre-processing chain; the others receive its output
        if (m == 0) { B200_TRY(b200_frontend_create(&fe[m], sampleRate, decimRatio, dcBlocking, invertIq, fftSize, fftRate, fftWindow, maxBlock, devices[m])); }
        else { B200_TRY(b200_frontend_create(&fe[m], sampleRate / decimRatio, 1, 0, 0, fftSize, fftRate, fftWindow, maxBlock / decimRatio + 2, devices[m])); }
        B200_TRY(b200_frontend_set_frame_shard(fe[m], m, n));
    }
    effectiveSr = member(0).effectiveSr;
    load.assign(n, 0.0);
    commStream.assign(n, nullptr); procStream.assign(n, nullptr);
    for (int k = 0; k < 2; k++) { rx[k].assign(n, nullptr); evReady[k].assign(n, nullptr); evFree[k].assign(n, nullptr); }
    evRows.assign(n, nullptr);
    for (int m = 0; m < n; m++) {
        DeviceGuard g(devices[m]);
        B200_CUDA(cudaStreamCreateWithFlags(&commStream[m], cudaStreamNonBlocking));
        B200_CUDA(cudaStreamCreateWithFlags(&procStream[m], cudaStreamNonBlocking));
        for (int k = 0; k < 2; k++) {
            B200_CUDA(cudaEventCreateWithFlags(&evReady[k][m], cudaEventDisableTiming));
            B200_CUDA(cudaEventCreateWithFlags(&evFree[k][m], cudaEventDisableTiming));
        }
        B200_CUDA(cudaEventCreateWithFlags(&evRows[m], cudaEventDisableTiming));
        if (m > 0) {
            // peer access both ways: block copies out of the ingest GPU, dB rows back into it
            int can = 0;
            B200_CUDA(cudaDeviceCanAccessPeer(&can, devices[m], devices[0]));
            if (can) { cudaError_t e = cudaDeviceEnablePeerAccess(devices[0], 0); if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { B200_CUDA(e); } cudaGetLastError(); }
        }
    }
    {
        DeviceGuard g(devices[0]);
        for (int m = 1; m < n; m++) {
            int can = 0;
            B200_CUDA(cudaDeviceCanAccessPeer(&can, devices[0], devices[m]));
            if (can) { cudaError_t e = cudaDeviceEnablePeerAccess(devices[m], 0); if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { B200_CUDA(e); } cudaGetLastError(); }
        }
        B200_CUDA(cudaEventCreateWithFlags(&evPre, cudaEventDisableTiming));
        B200_CUDA(cudaEventCreateWithFlags(&evOwner, cudaEventDisableTiming));
        B200_CUDA(cudaEventCreate(&evB0)); B200_CUDA(cudaEventCreate(&evB1));
        rowsCap = member(0).rowsCap * n + 2;
        B200_CUDA(cudaMalloc((void**)&dRowsAll, (size_t)rowsCap * fftSize * sizeof(float)));
    }
    useNccl = false;
    if (n > 1 && ncclApi().ok) {
        comms.assign(n, nullptr);
        B200_NCCL(ncclApi().commInitAll(comms.data(), n, devices.data()));
        useNccl = true;
    }
    return B200_OK;
}

void FrontEndGroup::destroy() {
    for (int m = 0; m < n; m++) {
        DeviceGuard g(devices[m]);
        if (procStream[m]) { cudaStreamSynchronize(procStream[m]); }
        if (commStream[m]) { cudaStreamSynchronize(commStream[m]); }
    }
    if (useNccl) { for (auto c : comms) { if (c) { ncclApi().commDestroy(c); } } }
    for (int m = 0; m < n; m++) {
        DeviceGuard g(devices[m]);
        if (fe[m]) { b200_frontend_destroy(fe[m]); }
        for (int k = 0; k < 2; k++) {
            if (rx[k][m]) { cudaFree(rx[k][m]); }
            if (evReady[k][m]) { cudaEventDestroy(evReady[k][m]); }
            if (evFree[k][m]) { cudaEventDestroy(evFree[k][m]); }
        }
        if (evRows[m]) { cudaEventDestroy(evRows[m]); }
        if (procStream[m]) { cudaStreamDestroy(procStream[m]); }
        if (commStream[m]) { cudaStreamDestroy(commStream[m]); }
    }
    DeviceGuard g(devices.empty() ? 0 : devices[0]);
    cudaEvent_t evs[] = { evPre, evOwner, evB0, evB1 };
    for (auto e : evs) { if (e) { cudaEventDestroy(e); } }
    if (dRowsAll) { cudaFree(dRowsAll); }
    if (dStage) { cudaFree(dStage); }
    if (dConv) { cudaFree(dConv); }
}

int FrontEndGroup::ensureRx(size_t bytes) {
    if (bytes <= rxBytes) { return B200_OK; }
    const size_t want = std::max(bytes, (size_t)maxBlock * sizeof(cf32));
    for (int m = 1; m < n; m++) {
        DeviceGuard g(devices[m]);
        B200_CUDA(cudaStreamSynchronize(procStream[m]));
        B200_CUDA(cudaStreamSynchronize(commStream[m]));
        for (int k = 0; k < 2; k++) {
            if (rx[k][m]) { cudaFree(rx[k][m]); rx[k][m] = nullptr; }
            B200_CUDA(cudaMalloc(&rx[k][m], want));
        }
    }
    rxBytes = want;
    freeArmed[0] = freeArmed[1] = false;
    return B200_OK;
}

// `src` lives on the ingest GPU and is complete once everything enqueued in `s` so far has run
int FrontEndGroup::broadcast(const void* src, size_t bytes, cudaStream_t s) {
    if (n == 1 || bytes == 0) { return B200_OK; }
    B200_TRY(ensureRx(bytes));
    {
        DeviceGuard g(devices[0]);
        B200_CUDA(cudaEventRecord(evPre, s));
        B200_CUDA(cudaStreamWaitEvent(commStream[0], evPre, 0));
        // the ingest GPU sends out of `src` on its communication stream; it must also wait until every member has let go
        // of the receive buffer this block goes into (the block two calls ago)
        if (freeArmed[par]) { for (int m = 1; m < n; m++) { B200_CUDA(cudaStreamWaitEvent(commStream[0], evFree[par][m], 0)); } }
        B200_CUDA(cudaEventRecord(evB0, commStream[0]));
    }
    if (useNccl) {
        for (int m = 1; m < n; m++) {
            DeviceGuard g(devices[m]);
            if (freeArmed[par]) { B200_CUDA(cudaStreamWaitEvent(commStream[m], evFree[par][m], 0)); }
        }
        B200_NCCL(ncclApi().groupStart());
        for (int m = 0; m < n; m++) {
            DeviceGuard g(devices[m]);
            void* dst = (m == 0) ? const_cast<void*>(src) : rx[par][m];     // in place on the root: nothing is written there
            ncclResult_t r = ncclApi().broadcast(m == 0 ? src : dst, dst, bytes, ncclChar, 0, comms[m], commStream[m]);
            if (r != ncclSuccess) { ncclApi().groupEnd(); setError("ncclBroadcast failed: %s", ncclApi().errorString(r)); return B200_ECUDA; }
        }
        B200_NCCL(ncclApi().groupEnd());
        for (int m = 1; m < n; m++) {
            DeviceGuard g(devices[m]);
            B200_CUDA(cudaEventRecord(evReady[par][m], commStream[m]));
        }
    }
    else {
        // peer copies, one per member, each on the member's own communication stream (the copies share the ingest GPU's
        // NVLink egress: N - 1 times the block leaves it, where NCCL sends it once)
        for (int m = 1; m < n; m++) {
            DeviceGuard g(devices[m]);
            B200_CUDA(cudaStreamWaitEvent(commStream[m], evPre, 0));
            if (freeArmed[par]) { B200_CUDA(cudaStreamWaitEvent(commStream[m], evFree[par][m], 0)); }
            B200_CUDA(cudaMemcpyPeerAsync(rx[par][m], devices[m], src, devices[0], bytes, commStream[m]));
            B200_CUDA(cudaEventRecord(evReady[par][m], commStream[m]));
        }
        DeviceGuard g(devices[0]);
        for (int m = 1; m < n; m++) { B200_CUDA(cudaStreamWaitEvent(commStream[0], evReady[par][m], 0)); }
    }
    {
        DeviceGuard g(devices[0]);
        B200_CUDA(cudaEventRecord(evB1, commStream[0]));
        bTimed = true;
        // the caller's stream may reuse `src` (its own input buffer) once the block has left the ingest GPU
        B200_CUDA(cudaStreamWaitEvent(s, evB1, 0));
    }
    return B200_OK;
}

// rawKind 0: rootBlock is cf32 on the ingest GPU, goes through the pre-processing chain, the result is broadcast.
// rawKind 2 / 1: dStage holds int16 / int8 I/Q (no pre-processing configured): the integers are broadcast and every
// member converts its own copy with the file source's arithmetic (x / scalar).
int FrontEndGroup::run(const cf32* rootBlock, int rawKind, float scalar, int count, cudaStream_t s) {
    if (count < 0 || count > maxBlock) { setError("block of %d samples exceeds max_block %d", count, maxBlock); return B200_ETOOBIG; }
    FrontEnd& root = member(0);
    int nPre = count;
    const int64_t frameBefore = root.nextFrame;
    {
        DeviceGuard g(devices[0]);
        if (root.everRan && s != root.lastStream) { B200_TRY(root.syncLast()); }
        root.lastStream = s; root.everRan = true;
        if (rawKind == 0) {
            B200_TRY(root.preprocess(rootBlock, count, s));
            nPre = root.preCount;
            B200_TRY(broadcast(root.preOut, (size_t)nPre * sizeof(cf32), s));
        }
        else {
            const size_t bytes = (size_t)count * (rawKind == 2 ? sizeof(short2) : sizeof(char2));
            B200_TRY(broadcast(dStage, bytes, s));
            if ((size_t)count > convSamples) {
                if (dConv) { B200_CUDA(cudaStreamSynchronize(s)); cudaFree(dConv); dConv = nullptr; }
                convSamples = (size_t)std::max(count, maxBlock);
                B200_CUDA(cudaMalloc((void**)&dConv, convSamples * sizeof(cf32)));
            }
            if (count > 0) {
                if (rawKind == 2) { convertI16Kernel<<<std::min(148 * 8, (count + 255) / 256), 256, 0, s>>>((const short2*)dStage, (float2*)dConv, count, scalar); }
                else { convertI8Kernel<<<std::min(148 * 8, (count + 255) / 256), 256, 0, s>>>((const char2*)dStage, (float2*)dConv, count, scalar); }
                B200_CHECK_LAUNCH(); countLaunch();
            }
            root.preOut = dConv; root.preCount = count;
        }
        B200_TRY(root.branches(s, false));
    }
    // ---- the other members: their share of the bank and of the frames, on their own streams
    for (int m = 1; m < n; m++) {
        FrontEnd& f = member(m);
        DeviceGuard g(devices[m]);
        cudaStream_t ps = procStream[m];
        B200_CUDA(cudaStreamWaitEvent(ps, evReady[par][m], 0));
        const cf32* blk = (const cf32*)rx[par][m];
        if (rawKind != 0) {
            // converted in place is not possible (the integers are narrower): the member's staging buffer takes the floats
            if (!f.dStageIn) { B200_CUDA(cudaMalloc((void**)&f.dStageIn, (size_t)f.maxBlock * sizeof(cf32))); }
            if (count > 0) {
                if (rawKind == 2) { convertI16Kernel<<<std::min(148 * 8, (count + 255) / 256), 256, 0, ps>>>((const short2*)rx[par][m], (float2*)f.dStageIn, count, scalar); }
                else { convertI8Kernel<<<std::min(148 * 8, (count + 255) / 256), 256, 0, ps>>>((const char2*)rx[par][m], (float2*)f.dStageIn, count, scalar); }
                B200_CHECK_LAUNCH(); countLaunch();
            }
            blk = f.dStageIn;
        }
        B200_TRY(f.process(blk, nPre, ps));
        B200_CUDA(cudaEventRecord(evFree[par][m], ps));
    }
    if (n > 1) { freeArmed[par] = true; }
    // ---- dB rows to the display owner, in frame order
    const int64_t frameAfter = root.nextFrame;
    const int nFrames = (int)(frameAfter - frameBefore);
    if (nFrames > rowsCap) { setError("internal: %d frames exceed the owner's row capacity %d", nFrames, rowsCap); return B200_ESTATE; }
    frameIdx.resize(nFrames);
    for (int i = 0; i < nFrames; i++) { frameIdx[i] = frameBefore + i; }
    const size_t rowBytes = (size_t)fftSize * sizeof(float);
    for (int m = 0; m < n; m++) {
        FrontEnd& f = member(m);
        DeviceGuard g(devices[m]);
        cudaStream_t ps = (m == 0) ? s : procStream[m];
        for (size_t i = 0; i < f.frameIdx.size(); i++) {
            const int64_t slot = f.frameIdx[i] - frameBefore;
            if (slot < 0 || slot >= nFrames) { setError("internal: member %d produced frame %lld outside the block", m, (long long)f.frameIdx[i]); return B200_ESTATE; }
            float* dst = dRowsAll + (size_t)slot * fftSize;
            const float* src = f.dRows + i * (size_t)fftSize;
            if (m == 0) { B200_CUDA(cudaMemcpyAsync(dst, src, rowBytes, cudaMemcpyDeviceToDevice, ps)); }
            else { B200_CUDA(cudaMemcpyPeerAsync(dst, devices[0], src, devices[m], rowBytes, ps)); }
        }
        if (m > 0) { B200_CUDA(cudaEventRecord(evRows[m], ps)); }
    }
    {
        DeviceGuard g(devices[0]);
        for (int m = 1; m < n; m++) { B200_CUDA(cudaStreamWaitEvent(s, evRows[m], 0)); }
        if (wf && nFrames > 0) {
            const int k = std::min(nFrames, wfMaxRows);
            B200_TRY(b200_waterfall_push(wf, dRowsAll, k, wfZoom, wfRgba, s));
        }
    }
    par ^= 1;
    return B200_OK;
}

} // namespace b200

struct b200_frontend_group { b200::FrontEndGroup impl; };

#define GROUP_OR_FAIL(g) \
    if (!(g)) { setError("NULL front-end group handle"); return B200_EINVAL; }

extern "C" {

int b200_frontend_group_create(b200_frontend_group** out, const int* devices, int n_devices, double sample_rate, int decim_ratio, int dc_blocking,
                               int invert_iq, int fft_size, double fft_rate, int fft_window, int max_block) {
    if (!out || !devices || n_devices < 1 || n_devices > 64) { setError("bad arguments to b200_frontend_group_create"); return B200_EINVAL; }
    for (int i = 0; i < n_devices; i++) {
        for (int j = 0; j < i; j++) { if (devices[i] == devices[j]) { setError("device %d is listed twice", devices[i]); return B200_EINVAL; } }
    }
    auto* g = new (std::nothrow) b200_frontend_group;
    if (!g) { return B200_ENOMEM; }
    const int rc = g->impl.create(devices, n_devices, sample_rate, decim_ratio < 1 ? 1 : decim_ratio, dc_blocking, invert_iq, fft_size, fft_rate, fft_window, max_block);
    if (rc != B200_OK) { g->impl.destroy(); delete g; return rc; }
    *out = g;
    return B200_OK;
}

int b200_frontend_group_destroy(b200_frontend_group* g) {
    GROUP_OR_FAIL(g);
    g->impl.destroy();
    delete g;
    return B200_OK;
}

int b200_frontend_group_size(b200_frontend_group* g) { GROUP_OR_FAIL(g); return g->impl.n; }
int b200_frontend_group_uses_nccl(b200_frontend_group* g) { GROUP_OR_FAIL(g); return g->impl.useNccl ? 1 : 0; }

b200_frontend* b200_frontend_group_member(b200_frontend_group* g, int member) {
    if (!g || member < 0 || member >= g->impl.n) { return nullptr; }
    return g->impl.fe[member];
}

int b200_frontend_group_add_vfo(b200_frontend_group* g, double out_samplerate, double bandwidth, double offset) {
    GROUP_OR_FAIL(g);
    FrontEndGroup& G = g->impl;
    const double cost = vfoCost(G.effectiveSr, out_samplerate, bandwidth);
    int best = 0;
    for (int m = 1; m < G.n; m++) { if (G.load[m] < G.load[best]) { best = m; } }
    const int local = b200_vfo_bank_add(b200_frontend_vfo_bank(G.fe[best]), out_samplerate, bandwidth, offset);
    if (local < 0) { return local; }
    FrontEndGroup::Loc loc; loc.member = best; loc.local = local; loc.cost = cost;
    G.load[best] += cost;
    for (size_t i = 0; i < G.vfos.size(); i++) { if (G.vfos[i].member < 0) { G.vfos[i] = loc; return (int)i; } }
    G.vfos.push_back(loc);
    return (int)G.vfos.size() - 1;
}

int b200_frontend_group_remove_vfo(b200_frontend_group* g, int vfo) {
    GROUP_OR_FAIL(g);
    FrontEndGroup& G = g->impl;
    if (vfo < 0 || vfo >= (int)G.vfos.size() || G.vfos[vfo].member < 0) { setError("no VFO %d in the group", vfo); return B200_EINVAL; }
    const FrontEndGroup::Loc loc = G.vfos[vfo];
    B200_TRY(b200_vfo_bank_remove(b200_frontend_vfo_bank(G.fe[loc.member]), loc.local));
    G.load[loc.member] -= loc.cost;
    G.vfos[vfo] = FrontEndGroup::Loc();
    return B200_OK;
}

int b200_frontend_group_vfo_location(b200_frontend_group* g, int vfo, int* member, int* local_index) {
    GROUP_OR_FAIL(g);
    FrontEndGroup& G = g->impl;
    if (vfo < 0 || vfo >= (int)G.vfos.size() || G.vfos[vfo].member < 0) { setError("no VFO %d in the group", vfo); return B200_EINVAL; }
    if (member) { *member = G.vfos[vfo].member; }
    if (local_index) { *local_index = G.vfos[vfo].local; }
    return B200_OK;
}

int b200_frontend_group_attach_waterfall(b200_frontend_group* g, b200_waterfall* wf, int data_width, float* d_zoom, uint32_t* d_rgba, int max_rows) {
    GROUP_OR_FAIL(g);
    if (wf && (data_width <= 0 || max_rows <= 0 || (!d_zoom && !d_rgba))) { setError("bad arguments to b200_frontend_group_attach_waterfall"); return B200_EINVAL; }
    FrontEndGroup& G = g->impl;
    B200_TRY(G.member(0).syncLast());
    G.wf = wf; G.wfZoom = d_zoom; G.wfRgba = d_rgba; G.wfMaxRows = wf ? max_rows : 0; G.wfWidth = data_width;
    return B200_OK;
}

int b200_frontend_group_process(b200_frontend_group* g, const float* d_in, int count, void* stream) {
    GROUP_OR_FAIL(g);
    if (!d_in && count > 0) { setError("NULL input"); return B200_EINVAL; }
    return g->impl.run((const cf32*)d_in, 0, 1.0f, count, (cudaStream_t)stream);
}

static int groupStage(FrontEndGroup& G, const void* h_in, size_t bytes, cudaStream_t s) {
    DeviceGuard dg(G.devices[0]);
    if (bytes > G.stageBytes) {
        if (G.dStage) { B200_CUDA(cudaStreamSynchronize(s)); cudaFree(G.dStage); G.dStage = nullptr; }
        G.stageBytes = std::max(bytes, (size_t)G.maxBlock * sizeof(cf32));
        B200_CUDA(cudaMalloc(&G.dStage, G.stageBytes));
    }
    // the staging buffer of the previous call may still be read by its broadcast: the broadcast's end is ordered into `s`
    if (bytes) { B200_CUDA(cudaMemcpyAsync(G.dStage, h_in, bytes, cudaMemcpyHostToDevice, s)); }
    return B200_OK;
}

int b200_frontend_group_process_host(b200_frontend_group* g, const float* h_in, int count, void* stream) {
    GROUP_OR_FAIL(g);
    if ((!h_in && count > 0) || count < 0 || count > g->impl.maxBlock) { setError("bad block"); return B200_EINVAL; }
    B200_TRY(groupStage(g->impl, h_in, (size_t)count * sizeof(cf32), (cudaStream_t)stream));
    return g->impl.run((const cf32*)g->impl.dStage, 0, 1.0f, count, (cudaStream_t)stream);
}

int b200_frontend_group_process_host_i16(b200_frontend_group* g, const int16_t* h_in, int count, float scalar, void* stream) {
    GROUP_OR_FAIL(g);
    FrontEndGroup& G = g->impl;
    if ((!h_in && count > 0) || count < 0 || count > G.maxBlock || !(scalar != 0.0f)) { setError("bad block"); return B200_EINVAL; }
    cudaStream_t s = (cudaStream_t)stream;
    B200_TRY(groupStage(G, h_in, (size_t)count * sizeof(short2), s));
    FrontEnd& root = G.member(0);
    if (root.decimRatio == 1 && !root.dcBlocking && !root.invertIQ) { return G.run(nullptr, 2, scalar, count, s); }
    // a pre-processing chain is configured: it runs on the ingest GPU on the converted block, and its output is broadcast
    DeviceGuard dg(G.devices[0]);
    if ((size_t)count > G.convSamples) {
        if (G.dConv) { B200_CUDA(cudaStreamSynchronize(s)); cudaFree(G.dConv); G.dConv = nullptr; }
        G.convSamples = (size_t)std::max(count, G.maxBlock);
        B200_CUDA(cudaMalloc((void**)&G.dConv, G.convSamples * sizeof(cf32)));
    }
    if (count > 0) {
        convertI16Kernel<<<std::min(148 * 8, (count + 255) / 256), 256, 0, s>>>((const short2*)G.dStage, (float2*)G.dConv, count, scalar);
        B200_CHECK_LAUNCH(); countLaunch();
    }
    return G.run(G.dConv, 0, 1.0f, count, s);
}

int b200_frontend_group_frame_count(b200_frontend_group* g) { GROUP_OR_FAIL(g); return (int)g->impl.frameIdx.size(); }

const float* b200_frontend_group_frame_ptr(b200_frontend_group* g, int i) {
    if (!g || i < 0 || i >= (int)g->impl.frameIdx.size()) { return nullptr; }
    return g->impl.dRowsAll + (size_t)i * g->impl.fftSize;
}

int64_t b200_frontend_group_frame_index(b200_frontend_group* g, int i) {
    if (!g || i < 0 || i >= (int)g->impl.frameIdx.size()) { return -1; }
    return g->impl.frameIdx[i];
}

int b200_frontend_group_sync(b200_frontend_group* g) {
    GROUP_OR_FAIL(g);
    FrontEndGroup& G = g->impl;
    for (int m = 0; m < G.n; m++) {
        DeviceGuard dg(G.devices[m]);
        B200_CUDA(cudaStreamSynchronize(G.procStream[m]));
        B200_CUDA(cudaStreamSynchronize(G.commStream[m]));
        B200_TRY(G.member(m).syncLast());
    }
    return B200_OK;
}

int b200_frontend_group_broadcast_ms(b200_frontend_group* g, double* ms) {
    GROUP_OR_FAIL(g);
    FrontEndGroup& G = g->impl;
    if (!ms) { return B200_EINVAL; }
    *ms = 0.0;
    if (!G.bTimed) { return B200_OK; }
    DeviceGuard dg(G.devices[0]);
    B200_CUDA(cudaEventSynchronize(G.evB1));
    float t = 0.f;
    B200_CUDA(cudaEventElapsedTime(&t, G.evB0, G.evB1));
    *ms = t;
    return B200_OK;
}

} // extern "C"

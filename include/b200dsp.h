/*
 * b200dsp.h — C ABI of libb200dsp.so: SDR++Brown's IQ-front-end hot path on NVIDIA B200 (sm_100a).
 *
 * The reference has no C ABI for this path; its boundary is the C++ classes
 * IQFrontEnd (core/src/signal_path/iq_frontend.h:18-135) and dsp::channel::RxVFO
 * (core/src/dsp/channel/rx_vfo.h:6-135) plus the radio demodulators they feed.  Each entry
 * point below names the reference member it replaces.  include/dsp/ holds header-compatible
 * C++ classes (same names, same signatures) that sit on top of this ABI; INTEGRATION.md shows
 * the binding a maintainer would add.
 *
 * Conventions
 *   - Complex samples are interleaved float {re, im} (dsp::complex_t, dsp/types.h:93-94);
 *     stereo samples are interleaved {l, r}.  Counts are in samples.
 *   - Every function returns 0 (B200_OK) or a negative B200_E* code; no exception crosses the ABI.
 *     b200_last_error() returns a thread-local message for the last failure.
 *   - "device" pointers are CUDA device pointers on the handle's device.  `stream` is a
 *     cudaStream_t passed as void* (NULL = the legacy default stream).  Work is stream-ordered;
 *     one submitting thread per handle.  Setters may be called between process calls and take
 *     effect at the next block boundary, in call order (the reference applies them under
 *     ctrlMtx between blocks: rx_vfo.h:35-87).
 *   - All DSP state (FIR histories, decimation offsets, resampler phase, rotator phase, framing
 *     position, AGC state) lives in the handle, so any partition of a stream into blocks yields
 *     identical sample counts and decimation phase as the reference, bit for bit.
 *   - There is no CPU fallback: every process call runs CUDA kernels or fails.
 */
#ifndef B200DSP_H
#define B200DSP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_EINVAL (-1)   /* bad argument (NULL handle, bad index, unsupported size ...) */
#define B200_ECUDA (-2)    /* a CUDA runtime call failed; see b200_last_error() */
#define B200_ENOMEM (-3)   /* host or device allocation failed */
#define B200_ESTATE (-4)   /* call not valid in the handle's current state */
#define B200_ETOOBIG (-5)  /* block larger than the handle's max_block / a static limit */

const char* b200_last_error(void);
const char* b200_version(void);
/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
uint64_t b200_launch_count(void);

/* Page-locked host memory for sample buffers (dsp::stream's buffers in include/dsp/stream.h; the reference uses
 * volk_malloc, dsp/buffer/buffer.h:20-24) and a blocking device->host copy, so that the C++ drop-in headers need no
 * CUDA headers.  b200_host_alloc falls back to ordinary aligned memory when no CUDA device is present (streams
 * still work; every process call still fails without a GPU). */
void* b200_host_alloc(size_t bytes);
void b200_host_free(void* p);
int b200_device_to_host(void* h_dst, const void* d_src, size_t bytes, void* stream);

/* Measurement hooks (bench.py's roofline line): when enabled, every launch of the fused
 * xlate+decimate kernel -- the kernel that streams the full-rate input -- is bracketed by CUDA events
 * on its launch stream.  b200_profile_read synchronises those events and returns, since the last read,
 * the summed kernel time (ms), the number of launches and the algorithmic bytes they moved
 * (8 B per input sample read once per bank + 8 B per stage-1 output sample per VFO). */
int b200_profile_enable(int enabled);
int b200_profile_read(double* ms_total, long long* launches, double* algorithmic_bytes);
/* The same per kernel family: B200_PROF_STAGE1 (what b200_profile_read reports), B200_PROF_FFT (all passes of one FFT-branch
 * call -- window, FFT, log power -- between one pair of events; bytes = 8 per windowed sample + 4 per dB value written),
 * B200_PROF_DC (the DC blocker of a block; 16 bytes per sample).  algorithmic_flops (may be NULL): SURVEY 8(d)'s count,
 * 6 + 4 T/D per input sample and VFO for the fused stage, 5 N log2 N per FFT frame. */
#define B200_PROF_STAGE1 0
#define B200_PROF_FFT 1
#define B200_PROF_DC 2
int b200_profile_read_kind(int kind, double* ms_total, long long* launches, double* algorithmic_bytes, double* algorithmic_flops);
/* The FP32 roof bench.py quotes the channelizer against, measured on `device` in the same run: independent chains of
 * packed FMAs (FFMA2, the form that reaches the CUDA cores' 128 FMA per clock and SM), every SM full of warps, best of
 * three launches.  Result in FMAs per second (x2 = flop/s). */
int b200_measure_fp32_peak(int device, double* ffma_per_second);

/* ------------------------------------------------------------------------------------------
 * RxVFO bank — replaces N instances of dsp::channel::RxVFO (channel/rx_vfo.h) fed by the same
 * Splitter output (routing/splitter.h:59-92).  One launch sequence serves every VFO of the bank.
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_vfo_bank b200_vfo_bank;

/* max_block: largest `count` a process call may carry (the reference's streams carry at most
 * STREAM_BUFFER_SIZE = 1 000 000, dsp/stream.h:14; larger blocks are allowed here). */
int b200_vfo_bank_create(b200_vfo_bank** out, double in_samplerate, int max_block, int device);
int b200_vfo_bank_destroy(b200_vfo_bank* bank);

/* RxVFO::RxVFO(in, inSamplerate, outSamplerate, bandwidth, offset) — rx_vfo.h:11,19-33.
 * Returns the VFO's index (>= 0) or a negative error. */
int b200_vfo_bank_add(b200_vfo_bank* bank, double out_samplerate, double bandwidth, double offset);
/* IQFrontEnd::removeVFO — iq_frontend.cpp:181-202.  Indices of other VFOs do not change. */
int b200_vfo_bank_remove(b200_vfo_bank* bank, int vfo);
int b200_vfo_bank_count(b200_vfo_bank* bank);   /* number of live VFOs */

int b200_vfo_set_offset(b200_vfo_bank* bank, int vfo, double offset);                          /* RxVFO::setOffset        rx_vfo.h:72-77 */
int b200_vfo_set_bandwidth(b200_vfo_bank* bank, int vfo, double bandwidth);                    /* RxVFO::setBandwidth     rx_vfo.h:60-70 */
int b200_vfo_set_out_samplerate(b200_vfo_bank* bank, int vfo, double out_sr, double bandwidth);/* RxVFO::setOutSamplerate rx_vfo.h:45-58 */
int b200_vfo_bank_set_in_samplerate(b200_vfo_bank* bank, double in_samplerate);                /* RxVFO::setInSamplerate  rx_vfo.h:35-43, all VFOs */
int b200_vfo_reset(b200_vfo_bank* bank, int vfo);                                              /* RxVFO::reset            rx_vfo.h:79-87 */

/* Rotator phase (FrequencyXlator::phase, frequency_xlator.h:70).  The reference carries it as a
 * recursive fp32 product; this library carries the angle in float64 and advances it by a step
 * calibrated against that recursion (DESIGN.md "rotator").  get/set exist so tests can compare
 * block by block from the same starting phase; set also serves checkpoint/restore. */
int b200_vfo_get_phase(b200_vfo_bank* bank, int vfo, double* re, double* im);
int b200_vfo_set_phase(b200_vfo_bank* bank, int vfo, double re, double im);

/* How the per-sample rotator phase is produced (DESIGN.md "rotator"):
 *   B200_ROTATOR_FAST  (default) closed form in float64 from a step calibrated against the reference's
 *       fp32 recursion at every retune.  Fully parallel; matches the reference to better than -80 dB per
 *       block of <= 1 M samples when the block starts from the reference's phase.
 *   B200_ROTATOR_EXACT one GPU thread per VFO replays VOLK's generic recursion bit for bit (phase *= delta,
 *       renormalise every 512 samples of a call) and the kernels interpolate between its checkpoints.
 *       Tracks the reference free-running for any offset and any stream length; sequential, so a stream
 *       advances at most ~230 MS/s (still above every configuration's real-time rate). */
#define B200_ROTATOR_FAST 0
#define B200_ROTATOR_EXACT 1
int b200_vfo_bank_set_rotator_mode(b200_vfo_bank* bank, int mode);
/* Tuning/test switches: B200_OPT_FUSED_STAGE1 (1 = fused xlate+decimate kernel where available, default;
 * 0 = separate rotation and generic FIR kernels, process-wide), B200_OPT_CALIBRATION_STEPS (length of the
 * FAST-mode calibration run, default 1048576; 0 = use atan2 of the fp32 step), B200_OPT_TILE_SAMPLES (process-wide
 * choice of the first-stage kernel for a launch that holds ONE VFO: 0 or 2048 = the single-VFO kernel with
 * 2048-sample tiles (default), 4096 = the grouped kernel with 4096-sample tiles, which launches of two or more
 * VFOs always use; results do not depend on it).
 * Scheduling switches (process-wide; results do not depend on them, tools/order_sweep.py measures them):
 * B200_OPT_DEFER_TAIL (default 0).  1 = streaming mode: the stages behind the full-rate first stage of every VFO, and the
 *   demodulators, of block k run on an internal stream beside the full-rate stage of block k+1 (the first stage's output
 *   is double-buffered; at most two blocks are in flight).  b200_vfo_bank_process / b200_frontend_process then return
 *   with that work still running and NOT ordered into `stream`; it is ordered into `stream` by the next call of
 *   b200_vfo_out_ptr / b200_vfo_audio_ptr / b200_vfo_if_ptr / b200_vfo_rds_ptr (call them after every process call whose
 *   results are wanted -- the pointers may alternate between calls), by the *_process_host entry points, by any setter,
 *   and by a device-wide synchronisation.  Used when every live VFO has a power-of-two decimator in front (the fused
 *   first stage); otherwise the call runs inline as with 0.
 * B200_OPT_BRANCH_ORDER (default 0): 0 = a front end enqueues its FFT branch before the channelizer, 1 = right behind
 *   the channelizer's full-rate kernel.  B200_OPT_FFT_CTAS_PER_SM (default 0 = as many as fit): persistent CTAs per SM
 *   of the two-pass FFT kernels. */
#define B200_OPT_FUSED_STAGE1 1
#define B200_OPT_CALIBRATION_STEPS 2
#define B200_OPT_TILE_SAMPLES 3
#define B200_OPT_FFT_CTAS_PER_SM 4
#define B200_OPT_BRANCH_ORDER 5
#define B200_OPT_DEFER_TAIL 6
/* B200_OPT_VFOS_PER_PASS (process-wide, default 0 = chosen by the size of the unrolled loop body): how many VFOs the
 * grouped first-stage kernel accumulates in registers per pass over the staged input tile, 2 or 4.
 * B200_OPT_VFOS_PER_CTA (process-wide, default 32): how many VFOs a CTA of that kernel serves from one staged tile
 * (lowered automatically when the launch would not fill the GPU).  Results depend on neither; tools/bench_bank.py
 * sweeps them. */
#define B200_OPT_VFOS_PER_PASS 7
#define B200_OPT_VFOS_PER_CTA 8
/* B200_OPT_FFT_DIRECT_MAX_LOG (process-wide, default 13; 9 = off): FFT plans created afterwards use the one-pass kernel (one CTA
 * per frame) for sizes up to 2^value and the two-pass kernels above.  Results agree to rounding. */
#define B200_OPT_FFT_DIRECT_MAX_LOG 9
/* B200_OPT_SPLIT_RUNS (process-wide, default 1): first-stage plans whose unrolled loop body exceeds the 32 KB instruction
 * cache level (the /8192 plan: 34 KB, 28 % of its stall samples were instruction fetches) run as two launches, each over
 * half of every sub-block's samples -- it stages only that half, so two such CTAs take the shared memory of one -- the second
 * adding to the first's outputs.  Measured on cfg4's bank: 11.06 against 12.17 ms per 16 Mi-sample block.  (A first form that
 * staged whole tiles in both launches was slower, 13.8 ms.)  Results agree to rounding. */
#define B200_OPT_SPLIT_RUNS 10
/* B200_OPT_LEAVE_ROOM (process-wide, default 0): 1 = in streaming mode (B200_OPT_DEFER_TAIL) the first-stage bank kernel is
 * launched as a persistent grid that leaves about 40 KB of shared memory and 17 K registers free on every SM, so that the
 * previous block's later stages (bandwidth-bound, on their own low-priority stream) run beside it.  Measured on cfg4: the
 * later stages then do overlap (profiles/r02_timeline_cfg4_room.txt), but the bank kernels lose more (12.2 -> 16.3 ms per
 * block with one resident CTA fewer per SM and the shared L1 / issue slots) than the 2.7 ms the overlap hides: 16.8 against
 * 15.2 ms per step.  Off. */
#define B200_OPT_LEAVE_ROOM 11
/* B200_OPT_PARAM_TABLES (process-wide, default 1): the first-stage bank kernel takes its per-VFO rotation tables from the
 * kernel parameter block (constant bank, uniform loads) instead of gathering them into shared memory pass by pass; a
 * launch then carries at most 1920 / D VFO pairs and a bank larger than that runs as several launches.  0 = the
 * shared-memory table path for every launch.  Same table values either way: results are identical. */
#define B200_OPT_PARAM_TABLES 12
/* B200_OPT_HOST_CHUNK_MIB (process-wide, default 0 = automatic): the *_process_host entry points cut a long block into chunks
 * of this many MiB of host data, so that the copy of chunk k+1 runs beside the kernels of chunk k (automatic: 16 MiB, 64 MiB
 * for a bank of 64 VFOs or more, whose per-chunk fixed costs weigh more).  Results do not depend on it. */
#define B200_OPT_HOST_CHUNK_MIB 13
/* B200_OPT_FFT_DIRECT_PREFETCH (process-wide, default 0): 1 = the one-pass FFT kernel (sizes up to
 * 2^B200_OPT_FFT_DIRECT_MAX_LOG) stages the samples of its next frame in shared memory with asynchronous copies while it
 * transforms the current one; 0 = every frame starts by loading its samples from global memory.  Measured 6-13 % slower
 * at 2048 ... 8192 points (the kernel is short of issue slots, not waiting for loads): off.  Same arithmetic either way. */
#define B200_OPT_FFT_DIRECT_PREFETCH 14
int b200_vfo_bank_set_option(b200_vfo_bank* bank, int option, int value);
/* The process-wide switches above without a handle (B200_OPT_FUSED_STAGE1, _TILE_SAMPLES, _FFT_CTAS_PER_SM, _BRANCH_ORDER,
 * _VFOS_PER_PASS, _VFOS_PER_CTA, _FFT_DIRECT_MAX_LOG, _SPLIT_RUNS, _LEAVE_ROOM, _PARAM_TABLES, _HOST_CHUNK_MIB, _FFT_DIRECT_PREFETCH).  b200_vfo_bank_set_option forwards these here: naming a
 * bank does NOT confine them to that bank -- every bank and front end of the process sees the change.  Only
 * B200_OPT_CALIBRATION_STEPS is per bank; B200_OPT_DEFER_TAIL is process-wide too but is switched through a bank handle,
 * because that bank's deferred work is drained first.  Calls are serialised by the library. */
int b200_set_global_option(int option, int value);

/* info[0]=mode (0 BOTH, 1 DECIM_ONLY, 2 RESAMP_ONLY, 3 NONE; rational_resampler.h:120-125)
 * [1]=power-of-two ratio [2]=interp [3]=decim [4]=resampler tap count [5]=taps per phase
 * [6]=channel-filter tap count (0 when bypassed, rx_vfo.h:24,91-93) [7]=1 if stage 1 runs fused */
int b200_vfo_describe(b200_vfo_bank* bank, int vfo, int* info8);
/* Copies tap tables to host (which: 0 = resampler prototype after the *interp gain,
 * 1 = channel filter, 10+s = decimator stage s).  Returns the tap count. */
int b200_vfo_get_taps(b200_vfo_bank* bank, int vfo, int which, float* out, int cap);

/* RxVFO::process(count, in, out) for every VFO — rx_vfo.h:89-100.  d_in: device cf32[count].
 * Results stay on the device until fetched; out counts are exact and available on return
 * (they depend only on counts and integer state, never on sample values). */
int b200_vfo_bank_process(b200_vfo_bank* bank, const float* d_in, int count, void* stream);
int b200_vfo_out_count(b200_vfo_bank* bank, int vfo);             /* outCount of the last process call */
const float* b200_vfo_out_ptr(b200_vfo_bank* bank, int vfo);      /* device cf32[out_count] */
/* Same, host buffers in and out (pinned staging + H2D/D2H inside; blocks until done).
 * h_out[v] must hold b200_vfo_max_out(bank, v, count) samples; counts[v] receives outCount. */
int b200_vfo_bank_process_host(b200_vfo_bank* bank, const float* h_in, int count, float* const* h_out, int* counts, void* stream);
int b200_vfo_max_out(b200_vfo_bank* bank, int vfo, int count);

/* Demodulator attached to a VFO's output (decoder_modules/radio/src/radio_module.h:76-107):
 *   B200_DEMOD_NFM  dsp::demod::FM<stereo_t>   (demod/fm.h:86-103, quadrature.h:39-46); p0=lowPass p1=highPass (flags)
 *   B200_DEMOD_AM   dsp::demod::AM<stereo_t>   (demod/am.h:101-133); p0=agcMode (0 CARRIER,1 AUDIO) p1=attack p2=decay p3=dcBlockRate
 *   B200_DEMOD_USB/LSB/DSB dsp::demod::SSB<stereo_t> (demod/ssb.h:82-97); p1=attack p2=decay
 *   B200_DEMOD_WFM  dsp::demod::BroadcastFM      (demod/broadcast_fm.h:144-190; deviation = bandwidth / 2 as in
 *                   radio/src/demodulators/wfm.h:76); p0=stereo p1=lowPass p2=rdsOut (flags).  With rdsOut the RDS branch
 *                   (broadcast_fm.h:169-175,197-206: MPX as complex, FrequencyXlator by -57 kHz, RationalResampler<complex_t>
 *                   to 5 kS/s) runs too; its output is read with b200_vfo_rds_ptr / b200_vfo_rds_count and is what the
 *                   reference hands to its host-side RDS bit decoder (radio/src/demodulators/wfm.h rdsHandler).
 * attack/decay/dcBlockRate are per-sample coefficients as the radio module passes them
 * (e.g. 50.0/IF rate, usb.h:34).  Demodulation runs at the VFO's out_samplerate with
 * `bandwidth` = the VFO bandwidth.  Audio (stereo {l,r}, same count as the VFO output) is
 * produced by b200_vfo_bank_process right after the channelizer. */
#define B200_DEMOD_NONE 0
#define B200_DEMOD_NFM 1
#define B200_DEMOD_AM 2
#define B200_DEMOD_USB 3
#define B200_DEMOD_LSB 4
#define B200_DEMOD_DSB 5
#define B200_DEMOD_WFM 6
int b200_vfo_set_demod(b200_vfo_bank* bank, int vfo, int kind, double p0, double p1, double p2, double p3);
/* The radio module's IF chain between the VFO and the demodulator (decoder_modules/radio/src/radio_module.h:92-98):
 * dsp::noise_reduction::NoiseBlanker (core/src/dsp/noise_reduction/noise_blanker.h:38-56; rate and level as given to
 * its setRate / setLevel, the radio uses rate = 500 / IF samplerate) followed by dsp::noise_reduction::Squelch
 * (squelch.h:55-69; level in dB against 10*log10 of the block's mean magnitude -- per processed block, as in the
 * reference).  Both are off by default, as in the reference.  b200_vfo_out_ptr keeps naming the VFO output in front
 * of the chain.  Needs a demodulator on the VFO; b200_vfo_set_demod() drops the setting again. */
int b200_vfo_set_if_chain(b200_vfo_bank* bank, int vfo, int nb_enabled, double nb_rate, double nb_level, int squelch_enabled, double squelch_level);
/* The last block of that chain: dsp::noise_reduction::FMIF (core/src/dsp/noise_reduction/fm_if.h:45-80; radio_module.h:93,98,
 * setFMIFNREnabled :856-860, setIFNRPreset / FMIF::setBins :880-891).  Per sample: Nuttall-windowed `bins`-point DFT of the
 * newest `bins` samples, strongest bin kept alone, backward DFT, element bins/2.  bins in 2..32 (the reference's
 * presets: 9 NOAA APT, 15 voice, 31 narrow band, 32 broadcast; radio/src/main.cpp:11-16).  Enabling it, or changing bins,
 * starts from a cleared delay line as FMIF::setBins does. */
int b200_vfo_set_if_nr(b200_vfo_bank* bank, int vfo, int enabled, int bins);
/* What the demodulator read in the last process call: the IF chain's output (ifChain.out, radio_module.h:92-98,
 * selectedDemod->setInput at :859), device cf32[out_count]; the VFO output itself when no chain block is on.  NULL
 * without a demodulator. */
const float* b200_vfo_if_ptr(b200_vfo_bank* bank, int vfo);
/* The radio module's AF chain behind the demodulator (decoder_modules/radio/src/radio_module.h:101-107,756-765):
 * dsp::multirate::RationalResampler<stereo_t> from the demodulator's rate to the audio device rate, then
 * dsp::filter::Deemphasis<stereo_t> (core/src/dsp/filter/deephasis.h:58-92).  audio_samplerate <= 0 keeps the audio at
 * the demodulator's rate; deemph_tau <= 0 leaves the de-emphasis out (the reference's default).  Needs a demodulator
 * on the VFO; b200_vfo_set_demod() drops the setting again (call this after it). */
int b200_vfo_set_audio_post(b200_vfo_bank* bank, int vfo, double audio_samplerate, double deemph_tau);
const float* b200_vfo_audio_ptr(b200_vfo_bank* bank, int vfo);    /* device stereo[audio_count] or NULL */
int b200_vfo_audio_count(b200_vfo_bank* bank, int vfo);           /* stereo samples of the last call (== out_count without a resampler) */
int b200_vfo_audio_max_out(b200_vfo_bank* bank, int vfo, int count);   /* upper bound of audio_count for `count` input samples */
/* The sink stream's volume block behind the AF chain: dsp::audio::Volume (core/src/dsp/audio/volume.h:13-50;
 * SinkManager::Stream::volumeAjust, signal_path/sink.h:73, setVolume :44).  Audio is scaled by volume^2 (0 when muted),
 * in place in the buffer b200_vfo_audio_ptr names.  Off until first called; b200_vfo_set_demod() drops it again. */
int b200_vfo_set_volume(b200_vfo_bank* bank, int vfo, double volume, int muted);
/* RDS baseband of a B200_DEMOD_WFM demodulator created with p2 != 0: device cf32[rds_count] at 5 kS/s
 * (BroadcastFM::rdsOut, broadcast_fm.h:233). */
const float* b200_vfo_rds_ptr(b200_vfo_bank* bank, int vfo);
int b200_vfo_rds_count(b200_vfo_bank* bank, int vfo);

/* ------------------------------------------------------------------------------------------
 * IQ front end — replaces IQFrontEnd (signal_path/iq_frontend.{h,cpp}): pre-processing chain
 * (PowerDecimator -> DCBlocker -> Conjugate, iq_frontend.cpp:40-44), FFT branch (Reshaper framing
 * iq_frontend.h:88-92 + handler iq_frontend.cpp:267-287) and the VFO bank it owns (addVFO :159-179).
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_frontend b200_frontend;

#define B200_WINDOW_RECTANGULAR 0
#define B200_WINDOW_BLACKMAN 1
#define B200_WINDOW_NUTTALL 2

/* IQFrontEnd::init(in, sampleRate, buffering, decimRatio, dcBlocking, fftSize, fftRate, fftWindow, ...)
 * — iq_frontend.cpp:19-83 followed by the updateFFTPath() every setter triggers (:289-325), i.e. the
 * sign-alternating window is in force from the first frame.  fft_size must be a power of two
 * in [32, 33554432]. */
int b200_frontend_create(b200_frontend** out, double sample_rate, int decim_ratio, int dc_blocking, int invert_iq,
                         int fft_size, double fft_rate, int fft_window, int max_block, int device);
int b200_frontend_destroy(b200_frontend* fe);

int b200_frontend_set_sample_rate(b200_frontend* fe, double sample_rate);   /* IQFrontEnd::setSampleRate  iq_frontend.cpp:90-118 */
int b200_frontend_set_decimation(b200_frontend* fe, int ratio);             /* IQFrontEnd::setDecimation  :124-141 */
int b200_frontend_set_dc_blocking(b200_frontend* fe, int enabled);          /* IQFrontEnd::setDCBlocking  :143-145 */
int b200_frontend_set_invert_iq(b200_frontend* fe, int enabled);            /* IQFrontEnd::setInvertIQ    :147-149 */
int b200_frontend_set_fft_size(b200_frontend* fe, int size);                /* IQFrontEnd::setFFTSize     :204-207 */
int b200_frontend_set_fft_rate(b200_frontend* fe, double rate);             /* IQFrontEnd::setFFTRate     :209-212 */
int b200_frontend_set_fft_window(b200_frontend* fe, int window);            /* IQFrontEnd::setFFTWindow   :214-217 */
double b200_frontend_effective_samplerate(b200_frontend* fe);               /* IQFrontEnd::getEffectiveSamplerate :263-265 */
/* Multi-GPU time sharding of FFT frames: this handle computes only frames with
 * (frame_index % world) == rank; every rank still tracks the framing position. Default 0 of 1. */
int b200_frontend_set_frame_shard(b200_frontend* fe, int rank, int world);

/* The bank fed by this front end's pre-processed stream (IQFrontEnd::addVFO/removeVFO go
 * through b200_vfo_bank_add/remove on it).  Owned by the front end. */
b200_vfo_bank* b200_frontend_vfo_bank(b200_frontend* fe);
/* Runs a waterfall's row reduction (b200_waterfall_push) inside the front end's FFT branch, as the reference's FFT sink
 * thread does when releaseFFTBuffer ends in WaterFall::pushFFT (gui/widgets/waterfall.cpp:1204-1276): frame i of a
 * process call lands in row i of d_zoom (float[max_rows][data_width]) and d_rgba (uint32[max_rows][data_width]); either
 * may be NULL.  data_width must be the waterfall's.  On the device the reduction then overlaps the channelizer's tail
 * stages instead of following them.  wf = NULL detaches. */
struct b200_waterfall;
int b200_frontend_attach_waterfall(b200_frontend* fe, struct b200_waterfall* wf, int data_width, float* d_zoom, uint32_t* d_rgba, int max_rows);

/* One block through the whole front end: pre-processing -> {FFT frames, every VFO (+demod)}.
 * d_in: device cf32[count].  After return:
 *   b200_frontend_preproc_out  -> the stream the Splitter would fan out (device cf32[n])
 *   b200_frontend_frame_count  -> FFT frames completed by this block (on this rank's shard)
 *   b200_frontend_frame_ptr(i) -> device float[fft_size], 10*log10(|X|^2/N^2), DC at N/2 */
int b200_frontend_process(b200_frontend* fe, const float* d_in, int count, void* stream);
int b200_frontend_preproc_out(b200_frontend* fe, const float** d_out, int* count);
int b200_frontend_frame_count(b200_frontend* fe);
const float* b200_frontend_frame_ptr(b200_frontend* fe, int i);
int64_t b200_frontend_frame_index(b200_frontend* fe, int i);   /* global index of frame i (for sharded reassembly) */
/* Host-buffer variant (the drop-in path of include/dsp/iq_frontend.h): H2D of the block, then D2H
 * of `h_frames` (frame_count x fft_size floats, may be NULL) and each VFO's output (h_vfo_out[v],
 * may be NULL) / audio (h_audio[v], may be NULL; b200_vfo_audio_count(v) stereo samples, size it with
 * b200_vfo_audio_max_out).  Blocks until done.
 * Blocks of 32 MiB or more (without decimation / DC blocking / IQ inversion) are cut into 16 MiB chunks (2^21 cf32 samples;
 * 64 MiB chunks, from 128 MiB on, when the bank holds 64 VFOs or more):
 * the copy of the following chunks runs behind the kernels of the current one and the results travel back over the
 * other PCIe direction meanwhile (give pinned host buffers, b200_host_alloc, for that to overlap).  All dB rows of
 * the block stay resident (frame_count / frame_ptr / frame_index cover the whole block); the per-VFO device
 * accessors (b200_vfo_out_ptr / out_count / audio_ptr) then describe the last chunk only -- use the host buffers and
 * vfo_counts, which always cover the whole block. */
int b200_frontend_process_host(b200_frontend* fe, const float* h_in, int count, float* h_frames, int max_frames,
                               float* const* h_vfo_out, float* const* h_audio, int* vfo_counts, void* stream);
/* The same call for sources that deliver interleaved int16 I/Q (the file source's WAV path, most SDR hardware): the
 * block crosses PCIe as 4 bytes per sample and becomes cf32 on the device, out = (float)in / scalar -- the arithmetic
 * of volk_16i_s32f_convert_32f(stream.writeBuf, inBuf, 32768.0f, ...) at source_modules/file_source/src/main.cpp:409.
 * Results are bit-identical to b200_frontend_process_host on the converted block. */
int b200_frontend_process_host_i16(b200_frontend* fe, const int16_t* h_in, int count, float scalar, float* h_frames, int max_frames,
                                   float* const* h_vfo_out, float* const* h_audio, int* vfo_counts, void* stream);
/* And for interleaved int8 I/Q (the SDR++ server stream's PCM_TYPE_I8, converted by volk_8i_s32f_convert_32f with
 * scalar 128 / scaler in core/src/dsp/compression/sample_stream_decompressor.h): 2 bytes per sample over PCIe. */
int b200_frontend_process_host_i8(b200_frontend* fe, const int8_t* h_in, int count, float scalar, float* h_frames, int max_frames,
                                  float* const* h_vfo_out, float* const* h_audio, int* vfo_counts, void* stream);
/* Device-resident form of that conversion: count complex samples, d_in int16[2*count] -> d_out float[2*count]. */
int b200_convert_i16(const int16_t* d_in, int count, float scalar, float* d_out, void* stream);
/* The other direction, for the recorder (misc_modules/recorder/src/main.cpp:528-558 -> wav::Writer::write,
 * core/src/utils/wav.cpp:148-178): n_floats device floats (stereo audio or complex baseband, interleaved) -> PCM of the
 * WAV file's sample type, bit for bit as the reference's writer converts them (uint8: (x*127)+128 truncated; int16 / int32:
 * volk_32f_s32f_convert_16i / _32i with 32767 / 2147483647).  Recorded data then crosses PCIe at its file size; the RIFF
 * container is host code (wav::Writer in include/utils/wav.h takes the converted bytes). */
/* The SDR++ server's sample packet, built on the device (dsp::compression::SampleStreamCompressor::process,
 * core/src/dsp/compression/sample_stream_compressor.h:30-62; sent by core/src/server.cpp:444-459): d_out receives
 * [u16 0][u16 pcm_type][f32 scaler][data], scaler = the largest float of the block (signed, as volk_32f_index_max_32u picks
 * it; NaN samples are passed over), data = int8 (x * 128/scaler) or int16 (x * 32768/scaler) clamped and rounded to nearest
 * even, or the floats themselves.  d_out must hold 8 + count * {2, 4, 8} bytes and be 8-byte aligned; the return value is
 * the packet's byte count (negative = error).  The receiving side is b200_frontend_process_host_i8 / _i16 with
 * scalar = 128 / scaler resp. 32768 / scaler (sample_stream_decompressor.h:24-32). */
#define B200_STREAM_PCM_I8 0
#define B200_STREAM_PCM_I16 1
#define B200_STREAM_PCM_F32 2
int b200_stream_compress(const float* d_in, int count, int pcm_type, void* d_out, void* stream);
#define B200_PCM_UINT8 0
#define B200_PCM_INT16 1
#define B200_PCM_INT32 2
int b200_convert_pcm(const float* d_in, void* d_out, int n_floats, int sample_type, void* stream);

/* DC blocker (dsp::correction::DCBlocker<complex_t>::process, correction/dc_blocker.h:54-60) for callers that
 * are not allowed to allocate.  This entry point has no counterpart in the reference; it exists for hosts that
 * cannot call cudaMalloc while running -- a real-time DSP/audio callback thread (cudaMalloc can stall the whole
 * device and has no latency bound), a plug-in sandbox with a memory budget fixed at load time, or a process whose
 * device memory already belongs to another allocator's pool.  b200_frontend_process may cudaMalloc scan scratch
 * when a larger block arrives; this call runs the same kernels with the scratch in a fixed ~256 KB array inside
 * the library image (held by every GPU that loads the library until the process exits) and performs no
 * allocation of any kind.
 *   d_in, d_out   device cf32[count] on the current device; d_out == d_in (in place) is allowed, partial overlap is not
 *   count         at most B200_DC_NOALLOC_MAX_SAMPLES; larger blocks are refused with B200_ETOOBIG
 *   rate          DCBlocker rate (IQFrontEnd uses 50 / sampleRate, iq_frontend.h:84-86); must lie in [0, 0.25],
 *                 anything else is refused with B200_EINVAL
 *   d_offset      device double[2] {re, im}: the blocker's offset, read before and written after the block, so
 *                 consecutive calls continue one stream; NULL = start from zero and discard the final offset
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress); make that call
 * before any stream capture begins.  All calls on a device share the one array: calls whose GPU work can
 * overlap (different streams without ordering, or several host threads) race on it -- serialise them. */
#define B200_DC_NOALLOC_MAX_SAMPLES (1 << 24)
int b200_dc_block_noalloc(const float* d_in, float* d_out, int count, double rate, double* d_offset, void* stream);

/* ------------------------------------------------------------------------------------------
 * Front-end group — one IQ front end over several GPUs of one node, in one process.  It replaces the
 * same IQFrontEnd, with its Splitter fan-out (routing/splitter.h:74-83; bound in iq_frontend.cpp:46,79,159-179)
 * carried over NVLink: member 0 (devices[0], the ingest GPU and display owner) runs the pre-processing chain, every
 * block is broadcast to the other members (ncclBroadcast when libnccl.so.2 can be loaded -- the block leaves the
 * ingest GPU once; cudaMemcpyPeerAsync to every member otherwise), each member channelizes the VFOs it was dealt
 * (b200_frontend_group_add_vfo places a VFO on the member with the least work) and computes the FFT frames f with
 * f % n == member, and the dB rows come back to the ingest GPU in frame order, so the waterfall reduction and the
 * frame-order recurrences behind it (b200_waterfall_smooth_hold; waterfall.cpp:1247-1273) see every frame.
 * Receive buffers are double buffered: the broadcast of block k + 1 overlaps the kernels of block k.
 * One submitting thread per group.  `stream` belongs to devices[0]; the other members work on internal streams --
 * b200_frontend_group_sync() (or a synchronisation of every device) before reading results from another member.
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_frontend_group b200_frontend_group;
int b200_frontend_group_create(b200_frontend_group** out, const int* devices, int n_devices, double sample_rate, int decim_ratio, int dc_blocking,
                               int invert_iq, int fft_size, double fft_rate, int fft_window, int max_block);
int b200_frontend_group_destroy(b200_frontend_group* g);
int b200_frontend_group_size(b200_frontend_group* g);
int b200_frontend_group_uses_nccl(b200_frontend_group* g);      /* 1: ncclBroadcast, 0: peer copies */
/* Member m's own front end (on devices[m]): its setters, its VFO bank (b200_frontend_vfo_bank) and the per-VFO calls
 * (b200_vfo_set_offset, b200_vfo_set_demod, b200_vfo_out_ptr ...) with the local index from
 * b200_frontend_group_vfo_location.  FFT / sample-rate setters must be applied to every member alike. */
b200_frontend* b200_frontend_group_member(b200_frontend_group* g, int member);
/* IQFrontEnd::addVFO / removeVFO over the group; returns the VFO's group id (>= 0). */
int b200_frontend_group_add_vfo(b200_frontend_group* g, double out_samplerate, double bandwidth, double offset);
int b200_frontend_group_remove_vfo(b200_frontend_group* g, int vfo);
int b200_frontend_group_vfo_location(b200_frontend_group* g, int vfo, int* member, int* local_index);
/* zoom + palette rows of every frame, in frame order, on the ingest GPU (cf. b200_frontend_attach_waterfall) */
int b200_frontend_group_attach_waterfall(b200_frontend_group* g, struct b200_waterfall* wf, int data_width, float* d_zoom, uint32_t* d_rgba, int max_rows);
/* One block through the group.  d_in: cf32[count] on devices[0].  The host variants copy the block to the ingest GPU
 * first; the int16 one broadcasts the integers (4 bytes per sample over NVLink instead of 8) and converts on every
 * member when no pre-processing is configured (volk_16i_s32f_convert_32f's arithmetic, file_source/src/main.cpp:409). */
int b200_frontend_group_process(b200_frontend_group* g, const float* d_in, int count, void* stream);
int b200_frontend_group_process_host(b200_frontend_group* g, const float* h_in, int count, void* stream);
int b200_frontend_group_process_host_i16(b200_frontend_group* g, const int16_t* h_in, int count, float scalar, void* stream);
/* dB rows of the last call on the ingest GPU, ALL frames of the block in frame order (whichever member computed them) */
int b200_frontend_group_frame_count(b200_frontend_group* g);
const float* b200_frontend_group_frame_ptr(b200_frontend_group* g, int i);
int64_t b200_frontend_group_frame_index(b200_frontend_group* g, int i);
int b200_frontend_group_sync(b200_frontend_group* g);
/* duration of the last block's broadcast on the ingest GPU's communication stream (waits for it to finish) */
int b200_frontend_group_broadcast_ms(b200_frontend_group* g, double* ms);

/* ---- dsp::detector::SignalDetector, the front end's detector pre-processor.
 * Replaces core/src/dsp/detector/signal_detector.{h,cpp} (class :78-129; run :359-414, getLineCandidates :151-267,
 * addSingleFFTRow :454-481, aggregateAndDetect :416-452), which IQFrontEnd owns as `detectorPreprocessor`
 * (signal_path/iq_frontend.h:77, iq_frontend.cpp:37-44,102) and the display menu toggles (gui/menus/display.cpp:144-151).
 * The detector collects (int)(sample_rate / 10) samples, turns them into a dB row (Blackman window, transform of that --
 * arbitrary -- length, halves swapped, 20 log10(|X| + 1e-10)), the row into "line candidates", keeps 100 candidate rows and
 * with every 10th row (once more than 20 are kept) publishes their time average smoothed over three bins: sigs_smoothed.
 * The samples pass through unchanged, so process() only reads them.  Rows of up to 2^24 points (167 MS/s) are served.
 * For the same dB row the candidates, history and result equal the reference's bit for bit; the dB row agrees to FFT
 * rounding.  Difference: a sample-rate change that changes the row length restarts the history. */
typedef struct b200_sigdet b200_sigdet;
int b200_sigdet_create(b200_sigdet** out, double sample_rate, int device);       /* SignalDetector() + setSampleRate */
void b200_sigdet_destroy(b200_sigdet* d);                                           /* no-op for a front end's own detector */
int b200_sigdet_set_sample_rate(b200_sigdet* d, double sample_rate);              /* :297-307 */
int b200_sigdet_set_center_frequency(b200_sigdet* d, double center_frequency);    /* :309-321: drops the partly collected row */
int b200_sigdet_set_enabled(b200_sigdet* d, int enabled);                         /* signal_detector.h:96 */
int b200_sigdet_is_enabled(b200_sigdet* d);
int b200_sigdet_fft_size(b200_sigdet* d);
/* one block of DEVICE samples (interleaved cf32), stream-ordered; rows complete as the samples add up */
int b200_sigdet_process(b200_sigdet* d, const float* d_in, int count, void* stream);
int b200_sigdet_process_host(b200_sigdet* d, const float* h_in, int count, void* stream);   /* the same for HOST samples */
/* addSingleFFTRow with a ready-made dB row of fft_size floats on the device (tests; a caller with its own spectrum) */
int b200_sigdet_add_row(b200_sigdet* d, const float* d_row_db, void* stream);
int b200_sigdet_sync(b200_sigdet* d);
int64_t b200_sigdet_rows(b200_sigdet* d);                                          /* rows completed since creation */
/* sigs_smoothed to the host (waits for the detector's stream): returns its length, 0 before the first detection, < 0 on error */
int b200_sigdet_result(b200_sigdet* d, float* h_out, int cap);
const float* b200_sigdet_result_ptr(b200_sigdet* d);                               /* device pointer, NULL before the first detection */
const float* b200_sigdet_last_row_ptr(b200_sigdet* d);                             /* device pointer: the last dB row */
/* candidate row `row` of the kept history (0 = oldest) to the host; returns the number of kept rows */
int b200_sigdet_history(b200_sigdet* d, int row, float* h_out, int cap);
/* The front end's own detector (IQFrontEnd::detectorPreprocessor): created disabled, follows the effective sample rate,
 * fed with every pre-processed block while enabled (togglePreprocessor(&detectorPreprocessor, on), iq_frontend.cpp:44). */
b200_sigdet* b200_frontend_detector(b200_frontend* fe);
int b200_frontend_set_detector(b200_frontend* fe, int enabled);

/* ---- dsp::IFNRLogMMSE, the logmmse module's front-end pre-processor: wide-band log-MMSE noise reduction.
 * Replaces misc_modules/noise_reduction_logmmse/src/if_nr.h:15-113 (process :63-113, reset :58-61, setHold :46-48) over
 * logmmse.h (logmmse_sample :244-296, logmmse_all :298-405, add_noise_history :117-143, update_noise_mu2 :150-242) and
 * bgnoise.h:27-59; the module adds it to the front end's chain with addPreprocessor (main.cpp:165).
 * Samples are collected until 28 hops (of sample_rate / 100) are there; the first 12 frames give the noise estimate; from
 * then on every call returns all complete 20 ms frames, de-noised and overlap-added (hop = half a frame), times 4.  The
 * output count of a call is therefore not the input count (0 at first, then a burst, then multiples of the hop).
* sample_rate is the integer rate the reference derives its frame length from (if_nr.h:72); up to 419 MS/s
 * (transform length 2 * even(floor(0.02 * sample_rate)) <= 2^24).  max_block bounds `count` of one call. */
typedef struct b200_logmmse b200_logmmse;
int b200_logmmse_create(b200_logmmse** out, int sample_rate, int max_block, int device);
void b200_logmmse_destroy(b200_logmmse* d);
int b200_logmmse_reset(b200_logmmse* d);                  /* if_nr.h:58-61: forget the noise profile and the buffered samples (at the next call) */
int b200_logmmse_set_hold(b200_logmmse* d, int hold);     /* if_nr.h:46-48: freeze the noise histories (while transmitting) */
int b200_logmmse_fft_size(b200_logmmse* d);
int b200_logmmse_max_out(b200_logmmse* d);                /* the most samples one call can return */
/* DEVICE samples in; *d_out receives a device pointer to *out_count de-noised samples, valid until the next call (stream-ordered) */
int b200_logmmse_process(b200_logmmse* d, const float* d_in, int count, const float** d_out, int* out_count, void* stream);
/* HOST samples in and out (the stand-alone block on dsp::stream buffers); returns when h_out is filled */
int b200_logmmse_process_host(b200_logmmse* d, const float* h_in, int count, float* h_out, int out_cap, int* out_count, void* stream);
int b200_logmmse_sync(b200_logmmse* d);
int b200_logmmse_noise(b200_logmmse* d, float* h_out, int cap);   /* the noise power estimate per bin: returns its length, 0 before sampling */

/* ---- dsp::compression::ExperimentalFFTCompressor: the spectral compressor of the SDR++ server's sample
 * stream.  Replaces core/src/dsp/compression/experimental_fft_compressor.h (setSampleRate :66-88, process :198-427,
 * filterSignal :116-177, estimateNoise :181-189, setMaskedFrequencies :99-103), which core/src/server.cpp:101-110 puts
 * between its resampler and the PCM sample-packet compressor (here: b200_stream_compress on the frame this returns).
 * Every 2^floor(log2(sample_rate / 20)) input samples become one frame of as many frequency-domain samples (DC in the
 * middle), ten frames late: bins that do not rise above the tracked noise floor by the allowance (loss_rate scales it;
 * 0 = keep every bin) are zeroed, the others carry |X|^(1/4) at X's phase.  As in the reference, a call takes the whole
 * block into its input buffer but lets at most ONE frame out: feed blocks of at most a frame.  For the same dB rows
 * every bin is kept or zeroed exactly as by the reference (the noise-floor chain runs in its summation order); the rows
 * themselves agree to FFT rounding.  Transmit mode (setTxMode) is not served. */
typedef struct b200_fftcomp b200_fftcomp;
int b200_fftcomp_create(b200_fftcomp** out, int sample_rate, double loss_rate, int max_block, int device);
void b200_fftcomp_destroy(b200_fftcomp* c);
int b200_fftcomp_fft_size(b200_fftcomp* c);
int b200_fftcomp_set_loss_rate(b200_fftcomp* c, double loss_rate);                       /* the public lossRate member (server.cpp:544) */
int b200_fftcomp_set_masked(b200_fftcomp* c, const int32_t* from_to_hz, int n);         /* setMaskedFrequencies: n values = n / 2 ranges always kept */
/* DEVICE samples in; *d_out receives a device pointer to *out_count (0 or fft_size) samples, valid until the next call */
int b200_fftcomp_process(b200_fftcomp* c, const float* d_in, int count, const float** d_out, int* out_count, void* stream);
int b200_fftcomp_process_host(b200_fftcomp* c, const float* h_in, int count, float* h_out, int* out_count, void* stream);
int b200_fftcomp_sync(b200_fftcomp* c);
int b200_fftcomp_noise_figure(b200_fftcomp* c, float* h_out, int cap);                   /* 30 values (server.cpp:470-474); 0 before the first filtered frame */

/* One complex spectrum for tests: FFT of an already-windowed frame, no log (checks K9 alone). */
int b200_fft_forward(const float* d_in, int nz, int fft_size, float* d_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * Waterfall math — WaterFall::pushFFT (gui/widgets/waterfall.cpp:1204-1276): doZoom (:98-154) and the
 * dB -> palette mapping (:1229-1235).  Headless: results go to device buffers, no GL.
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_waterfall b200_waterfall;
int b200_waterfall_create(b200_waterfall** out, int raw_fft_size, int data_width, int device);
int b200_waterfall_destroy(b200_waterfall* wf);
/* view: drawDataStart / drawDataSize of pushFFT (:1208-1210) */
int b200_waterfall_set_view(b200_waterfall* wf, int offset, int width);
int b200_waterfall_set_range(b200_waterfall* wf, float wf_min, float wf_max);
/* 1000-entry RGBA LUT (WATERFALL_RESOLUTION); b200_waterfall_make_palette restates updatePallette (:1282-1295) */
int b200_waterfall_set_palette(b200_waterfall* wf, const uint32_t* lut1000);
int b200_waterfall_make_palette(const float* colors_rgb, int color_count, uint32_t* lut1000);
/* nrows dB rows (device, row stride raw_fft_size) -> zoomed rows (device float[nrows][data_width], may be
 * NULL) and RGBA rows (device uint32[nrows][data_width], may be NULL). */
int b200_waterfall_push(b200_waterfall* wf, const float* d_rows, int nrows, float* d_zoom, uint32_t* d_rgba, void* stream);
/* WaterFall::rawFFTs / getFFTBuffer (waterfall.cpp:1188-1199): keep the newest `height` raw dB rows of every push on
 * the device (0 = keep none, the default).  Rows handed to b200_waterfall_push must then be 16-byte aligned. */
int b200_waterfall_set_history(b200_waterfall* wf, int height);
/* WaterFall::updateWaterfallFb (waterfall.cpp:771-866): all stored rows through the current view / range / palette in
 * one launch -- what the reference does on four CPU threads after a zoom, pan or palette change.  d_zoom
 * (float[height][data_width]) and d_rgba (uint32[height][data_width]) may each be NULL; output row 0 is the newest
 * stored row, rows past the number stored are opaque black (0xFF000000).  Returns the number of stored rows. */
int b200_waterfall_regenerate(b200_waterfall* wf, float* d_zoom, uint32_t* d_rgba, void* stream);
/* WaterFall::calculateVFOSignalInfo (waterfall.cpp:688-759): strength and SNR of one VFO (centre offset and bandwidth
 * in Hz, relative to a display `whole_bandwidth` wide) from each raw dB row: d_out[row] = {strength, snr}.
 * Returns 1 (and launches nothing) where the reference returns false (empty side bands). */
int b200_waterfall_vfo_signal(b200_waterfall* wf, const float* d_rows, int nrows, double whole_bandwidth, double center_offset,
                              double bandwidth, float* d_out, void* stream);
/* pushFFT's FFT smoothing (waterfall.cpp:1247-1252) and peak hold (:1269-1272) over nrows zoomed rows in frame order.
 * d_zoom_rows is smoothed in place; d_hold_rows (may be NULL) receives the hold trace after each row.  The running
 * buffers live in the handle and start at -1000 dB like the reference's. */
int b200_waterfall_smooth_hold(b200_waterfall* wf, float* d_zoom_rows, int nrows, int smoothing, float alpha, int hold, float hold_speed,
                               float* d_hold_rows, void* stream);
/* the start indices / spans the zoom uses (host copies, data_width each) */
int b200_waterfall_zoom_table(b200_waterfall* wf, int* sid, int* ufactor);

#ifdef __cplusplus
}
#endif
#endif /* B200DSP_H */

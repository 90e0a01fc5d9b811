// TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product; only
// tests/, __graft_entry__.smoke() and bench.py's CPU-baseline / --impl reference
// legs may load these libraries.
//
// One C API, two implementations:
//   oracle/_ref/libref_oracle.so   ref_harness.cpp  — the reference's OWN header-only
//                                  DSP blocks, compiled in place from $REF_ROOT/core/src
//                                  against oracle/shim (generic-C VOLK stand-in).
//   oracle/libport_oracle.so       oracle_port.cpp — a self-contained restatement of the
//                                  same algorithms (no reference includes) that travels
//                                  to machines where /root/reference does not exist.
// tests/test_oracle_*.py pin the port against _ref bit-for-bit and both against
// tests/golden/*.npz.
//
// All complex buffers are interleaved float {re, im} (dsp::complex_t, dsp/types.h:93-94);
// stereo buffers are interleaved {l, r}.  `count` is in samples, not floats.
#pragma once
#ifdef __cplusplus
extern "C" {
#endif

const char* orc_kind(void); // "reference" or "port"

// ---- dsp::channel::RxVFO (channel/rx_vfo.h:6-135)
void* orc_vfo_create(double inSR, double outSR, double bandwidth, double offset);
void orc_vfo_destroy(void* h);
int orc_vfo_process(void* h, int count, const float* in, float* out); // returns out sample count
void orc_vfo_set_offset(void* h, double offset);
void orc_vfo_set_bandwidth(void* h, double bandwidth);
void orc_vfo_set_in_samplerate(void* h, double inSR);
void orc_vfo_set_out_samplerate(void* h, double outSR, double bandwidth);
void orc_vfo_reset(void* h);
void orc_vfo_get_phase(void* h, float* reim);               // FrequencyXlator::phase (frequency_xlator.h:70)
void orc_vfo_set_phase(void* h, float re, float im);
void orc_vfo_get_phase_delta(void* h, float* reim);         // FrequencyXlator::phaseDelta (:71)
// info[0]=mode (0 BOTH,1 DECIM_ONLY,2 RESAMP_ONLY,3 NONE) [1]=pow2 ratio [2]=interp [3]=decim
// [4]=resampler tap count [5]=tapsPerPhase [6]=channel-filter tap count (0 when bypassed)
void orc_vfo_describe(void* h, int* info);
int orc_vfo_get_resamp_taps(void* h, float* out, int cap);  // RationalResampler::rtaps after *interp
int orc_vfo_get_chan_taps(void* h, float* out, int cap);    // RxVFO::ftaps

// ---- dsp::multirate::PowerDecimator<complex_t> (multirate/power_decimator.h)
void* orc_decim_create(int ratio);
void orc_decim_destroy(void* h);
int orc_decim_process(void* h, int count, const float* in, float* out);
// plan introspection: returns stage count; stage s decimation / tap count / taps
int orc_decim_plan(int ratio, int* decims, int* tapCounts, int cap);
int orc_decim_stage_taps(int ratio, int stage, float* out, int cap);

// ---- dsp::correction::DCBlocker<complex_t> (correction/dc_blocker.h:54-60), rate as double like IQFrontEnd::genDCBlockRate
void* orc_dcblock_create(double rate);
void orc_dcblock_destroy(void* h);
int orc_dcblock_process(void* h, int count, const float* in, float* out);

// ---- IQFrontEnd pre-processing chain order (signal_path/iq_frontend.cpp:40-44):
// PowerDecimator -> DCBlocker -> Conjugate, each optional.
void* orc_preproc_create(double sampleRate, int decimRatio, int dcBlocking, int invertIQ);
void orc_preproc_destroy(void* h);
int orc_preproc_process(void* h, int count, const float* in, float* out);

// ---- taps & windows
int orc_lowpass_taps(double cutoff, double transWidth, double sampleRate, float* out, int cap); // taps/low_pass.h:8-15
// window type 0 RECT, 1 BLACKMAN, 2 NUTTALL; includes the (-1)^i alternation of updateFFTPath (iq_frontend.cpp:303-311)
void orc_fft_window(int type, int nz, float* out);
void orc_reshape_params(double sampleRate, int fftSize, double fftRate, int* skip, int* nz); // iq_frontend.h:88-92

// ---- IQFrontEnd::handler (iq_frontend.cpp:267-287): window -> zero-pad -> forward FFT -> 10log10(|X|^2/N^2)
// spectrum (optional, may be NULL) receives the N complex FFT outputs (before log).
void orc_fft_frame(const float* iq, const float* window, int nz, int N, float* outDb, float* spectrum);

// ---- waterfall math (gui/widgets/waterfall.cpp)
void orc_do_zoom(int offset, int width, int inSize, int outSize, const float* in, float* out);   // :98-154
void orc_zoom_indices(int offset, int width, int inSize, int outSize, int* sId, int* uFactor);   // the fp32 index walk of :108-152
void orc_palette_indices(const float* row, int n, float wfMin, float wfMax, int* idx);          // :1229-1235 (index into the 1000-entry LUT)
void orc_make_palette(const float* colors /*[n][3]*/, int colorCount, unsigned int* lut /*[1000]*/); // updatePallette :1282-1295

int orc_vfo_signal_info(const float* row, int rawFFTSize, double wholeBandwidth, double centerOffset, double bandwidth,
                         float* strength, float* snr);                                                    // :688-759, returns 1 when computed
void orc_smooth_hold(float* latest, float* smoothingBuf, float* hold, int width, int smoothing, float alpha, int doHold, float holdSpeed); // :1247-1272

// ---- demodulators (stereo_t output, as the radio module instantiates them)
void* orc_fm_create(double samplerate, double bandwidth, int lowPass, int highPass); // demod/fm.h
void orc_fm_destroy(void* h);
int orc_fm_process(void* h, int count, const float* in, float* outStereo);
void* orc_ssb_create(int mode /*0 USB 1 LSB 2 DSB*/, double bandwidth, double samplerate, double agcAttack, double agcDecay); // demod/ssb.h
void orc_ssb_destroy(void* h);
int orc_ssb_process(void* h, int count, const float* in, float* outStereo);
void* orc_am_create(int agcMode /*0 CARRIER 1 AUDIO*/, double bandwidth, double agcAttack, double agcDecay, double dcBlockRate, double samplerate); // demod/am.h
void orc_am_destroy(void* h);
int orc_am_process(void* h, int count, const float* in, float* outStereo);
// demod::BroadcastFM (demod/broadcast_fm.h:34-215; radio wiring demodulators/wfm.h:76,261-263): quadrature detector, then mono
// (MPX through the 15 kHz audio low-pass into both channels) or stereo (complex band-pass on the 19 kHz pilot, PLL,
// delayed MPX mixed down by the squared conjugate VCO, L = (L+R)+(L-R), R = (L+R)-(L-R), audio low-pass on both).  No RDS branch.
void* orc_wfm_create(double deviation, double samplerate, int stereo, int lowPass);
void orc_wfm_destroy(void* h);
int orc_wfm_process(void* h, int count, const float* in, float* outStereo);
// The same demodulator with its RDS branch on (broadcast_fm.h:169-175,197-206): the real MPX as complex, FrequencyXlator by
// -57 kHz, RationalResampler<complex_t> to 5 kS/s.  Writes the stereo audio (count samples) and returns the RDS sample count.
void* orc_wfm_create_rds(double deviation, double samplerate, int stereo, int lowPass);
int orc_wfm_process_rds(void* h, int count, const float* in, float* outStereo, float* rdsOut);
// radio IF chain between the VFO and the demodulator (radio_module.h:92-98): NoiseBlanker (noise_reduction/noise_blanker.h:38-56)
// then Squelch (noise_reduction/squelch.h:55-69, the complex_t overload); either may be off.
void* orc_ifchain_create(int nbOn, double nbRate, double nbLevel, int squelchOn, double squelchLevel);
void orc_ifchain_destroy(void* h);
int orc_ifchain_process(void* h, int count, const float* in, float* out);
// dsp::noise_reduction::FMIF (noise_reduction/fm_if.h:45-80), the last block of the IF chain (radio_module.h:93,98): per
// sample, Nuttall-windowed `bins`-point forward DFT of the newest `bins` samples, all bins but the strongest zeroed,
// backward DFT, element bins/2 of the result.  FFTW is served by oracle/shim/fftw3.h.
void* orc_fmif_create(int bins);
void orc_fmif_destroy(void* h);
int orc_fmif_process(void* h, int count, const float* in, float* out);
// dsp::audio::Volume (audio/volume.h:13-50), the sink stream's volume block (signal_path/sink.h:73): gain = volume^2 in float, 0 when muted
int orc_volume_process(double volume, int muted, int count, const float* inStereo, float* outStereo);
// dsp::compression::SampleStreamCompressor::process (compression/sample_stream_compressor.h:30-62), the SDR++ server's
// sample packet: [u16 0][u16 pcmType][f32 scaler][data]; pcmType 0 int8, 1 int16, 2 float32; scaler = the largest float of
// the block (signed, first of equals).  Returns the byte count.  And SampleStreamDecompressor::process (:14-35).
int orc_stream_compress(int count, int pcmType, const float* in, unsigned char* out);
int orc_stream_decompress(int nbytes, const unsigned char* in, float* out);
// wav::Writer (utils/wav.cpp:20-182 over riff::Writer, utils/riff.cpp): the recorder's file writer (misc_modules/recorder/src/main.cpp:168-177,528-558).
// Writes `count` samples of `channels` floats each in blocks of blockLen samples; sampleType 0 UINT8, 1 INT16, 2 INT32, 3 FLOAT32.  Returns 0 when written.
int orc_wav_write(const char* path, int channels, double samplerate, int sampleType, const float* samples, int count, int blockLen);
// radio AF chain (decoder_modules/radio/src/radio_module.h:101-107,756-765): RationalResampler<stereo_t> to the audio
// device rate, then optional Deemphasis<stereo_t> (filter/deephasis.h:58-76).  tau <= 0 disables the de-emphasis.
void* orc_afpost_create(double inSamplerate, double audioSamplerate, double tau);
void orc_afpost_destroy(void* h);
int orc_afpost_process(void* h, int count, const float* inStereo, float* outStereo);   // returns out sample count
// loop::AGC<float> stand-alone (loop/agc.h:85-139)
void* orc_agc_create(double setPoint, double attack, double decay, double maxGain, double maxOutputAmp, double initGain);
void orc_agc_destroy(void* h);
int orc_agc_process(void* h, int count, const float* in, float* out);

// ---- dsp::detector::SignalDetector (dsp/detector/signal_detector.cpp; the front end's detectorPreprocessor,
// signal_path/iq_frontend.cpp:37-44).  FFT size = (int)(sampleRate / 10) -- any length, not a power of two.
void* orc_sigdet_create(double sampleRate);
void orc_sigdet_destroy(void* h);
int orc_sigdet_fft_size(void* h);
void orc_sigdet_set_center_frequency(void* h, double f);                 // :309-321 (clears the sample buffer, keeps the row history)
// run() (:359-414) on one block; returns the number of rows completed since creation
int orc_sigdet_process(void* h, int count, const float* in);
int orc_sigdet_result(void* h, float* out, int cap);                     // sigs_smoothed (:436-450); returns its length, 0 before the first detection
int orc_sigdet_history(void* h, int row, float* out, int cap);           // candidate row `row` of the kept history (0 = oldest); returns the number of kept rows
// one row of run(): Blackman window (no alternation), forward DFT, swapfft, |X|, 20 log10(|X| + 1e-10)  (:373-397)
void orc_sigdet_row(const float* iq, int N, float* rowDb);
// getLineCandidates (:151-267) of one dB row: -20 everywhere except the marked line positions
void orc_sigdet_candidates(const float* rowDb, int N, float* out);
// addSingleFFTRow (:454-481) with a ready-made dB row (history, detection interval, aggregateAndDetect :416-452)
void orc_sigdet_add_row(void* h, const float* rowDb, int N);

// ---- dsp::IFNRLogMMSE (misc_modules/noise_reduction_logmmse/src/if_nr.h:15-113 over logmmse.h), the pre-processor the
// logmmse module adds to the front end (main.cpp:165): wide-band log-MMSE spectral noise reduction, 20 ms Hann frames with
// 50 % overlap, transform length 2 * even(floor(0.02 * sampleRate)).
void* orc_logmmse_create(int sampleRate);
void orc_logmmse_destroy(void* h);
void orc_logmmse_reset(void* h);                                          // if_nr.h:58-61 (the module calls it when the tuning changes)
void orc_logmmse_set_hold(void* h, int hold);                             // :46-48 (noise profile frozen while transmitting)
int orc_logmmse_process(void* h, int count, const float* in, float* out); // process() :63-113; returns the output sample count
float orc_expn(float x);                                                  // dsp::math::expn (utils/arrays.cpp:60-431): E1 from its table
int orc_logmmse_noise(void* h, float* out, int cap);                      // SavedParamsC::noise_mu2; returns its length (0 before sampling)

// ---- the reference's threaded block graph for the many-VFO case (SURVEY 8(d) CPU baseline mode (ii); timing only):
// one dsp::routing::Splitter (routing/splitter.h:59-92) fanning the input stream out to nVfos dsp::channel::RxVFO blocks,
// each on its own thread (dsp/block.h:82-84) behind its own stream, wired as IQFrontEnd::addVFO does
// (signal_path/iq_frontend.cpp:159-180), every VFO output drained by a dsp::sink::Null.  Feeds `block` (blockLen samples)
// nBlocks times and returns the seconds until every VFO has processed all of them.  The port has no threads or streams:
// it returns -1.
double orc_graph_run(double inSR, int nVfos, const double* outSR, const double* bandwidth, const double* offset,
                     int blockLen, int nBlocks, const float* block);

// ---- dsp::compression::ExperimentalFFTCompressor (dsp/compression/experimental_fft_compressor.h), the block in front of the
// server's sample-packet compressor (core/src/server.cpp:101-110): frames of fftSize = 2^floor(log2(sampleRate / 20))
// samples -> frequency domain, bins under the tracked noise floor zeroed, the rest amplitude-companded; ten frames of delay.
void* orc_fftcomp_create(int sampleRate, double lossRate);
void orc_fftcomp_destroy(void* h);
int orc_fftcomp_fft_size(void* h);
void orc_fftcomp_set_masked(void* h, const int* fromToHz, int n);            // setMaskedFrequencies :99-103 (pairs from, to)
// appends the block to the input buffer and runs process() once (:198-427, as run() :431-452 does): returns 0 or fftSize
int orc_fftcomp_process(void* h, int count, const float* in, float* out);
int orc_fftcomp_noise_figure(void* h, float* out, int cap);                  // noiseFigure (30 values; server.cpp:470-474 sends it on)

#ifdef __cplusplus
}
#endif

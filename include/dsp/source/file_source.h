// WAV file source (reference: source_modules/file_source/src/main.cpp:396-448, `worker` / `floatWorker`).
// The reference's worker reads blockSize = min(sampleRate / 200, STREAM_BUFFER_SIZE) samples per turn through
// wav::Reader::readSamples (the file loops at its end), converts int16 I/Q with
// volk_16i_s32f_convert_32f(stream.writeBuf, inBuf, 32768.0f, 2 * blockSize) and swaps its stream, which is the input
// of IQFrontEnd.  Two ways to run the same loop here:
//   * start() / stop(): a worker thread fills `stream` exactly like that (conversion on the host), for any consumer of a
//     dsp::stream<complex_t> -- IQFrontEnd::setInput(&source.stream) as in the reference;
//   * pump(fe, ...): the blocks go to a b200_frontend as they sit in the file: int16 I/Q crosses PCIe at 4 bytes per
//     sample and is converted on the device with the same arithmetic (b200_frontend_process_host_i16), float files go
//     through b200_frontend_process_host.
#pragma once
#include <algorithm>
#include <atomic>
#include <memory>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../utils/wav.h"
#include "../stream.h"
#include "../types.h"

namespace dsp::source {
    class WavFileSource {
    public:
        dsp::stream<complex_t> stream;
        std::string error;

        ~WavFileSource() { stop(); }

        bool open(const std::string& path) {
            stop();
            reader.reset(new wav::Reader(path));
            if (!reader->isValid()) { error = reader->error; reader.reset(); return false; }
            if (reader->getChannelCount() != 2) { error = "not an I/Q file (two channels expected)"; reader.reset(); return false; }
            const auto t = reader->getSampleType();
            if (t != wav::SAMP_TYPE_INT16 && t != wav::SAMP_TYPE_FLOAT32) { error = "only int16 and float32 I/Q files are read (as in the reference)"; reader.reset(); return false; }
            isFloat = (t == wav::SAMP_TYPE_FLOAT32);
            return true;
        }
        bool isOpen() const { return (bool)reader; }
        double getSampleRate() const { return reader ? std::max<uint32_t>(reader->getSampleRate(), 1u) : 0.0; }
        bool floatSamples() const { return isFloat; }
        // the reference's block length: sampleRate / 200 samples, at most one stream buffer (main.cpp:399)
        int blockSize() const { return std::min((int)((float)getSampleRate() / 200.0f), (int)STREAM_BUFFER_SIZE); }
        void rewind() { if (reader) { reader->rewind(); } }

        // ---- the reference's worker thread
        void start() {
            if (running || !reader) { return; }
            running = true;
            stream.clearWriteStop();
            worker = std::thread([this] { isFloat ? floatLoop() : int16Loop(); });
        }
        void stop() {
            if (!running) { return; }
            stream.stopWriter();
            if (worker.joinable()) { worker.join(); }
            stream.clearWriteStop();
            running = false;
        }

        // ---- blocks straight into a GPU front end: `blocks` turns of the worker loop (the file loops at its end like the
        // reference's reader), results through the same arguments as b200_frontend_process_host; onBlock(count) after each.
        template <typename F>
        int pump(b200_frontend* fe, int blocks, float* h_frames, int max_frames, float* const* h_vfo_out, float* const* h_audio, int* vfo_counts, F onBlock) {
            if (!reader) { return B200_ESTATE; }
            const int n = blockSize();
            if (isFloat) { fbuf.resize((size_t)n * 2); } else { ibuf.resize((size_t)n * 2); }
            for (int b = 0; b < blocks; b++) {
                int rc;
                if (isFloat) {
                    reader->readSamples(fbuf.data(), (size_t)n * sizeof(complex_t));
                    rc = b200_frontend_process_host(fe, fbuf.data(), n, h_frames, max_frames, h_vfo_out, h_audio, vfo_counts, nullptr);
                }
                else {
                    reader->readSamples(ibuf.data(), (size_t)n * 2 * sizeof(int16_t));
                    rc = b200_frontend_process_host_i16(fe, ibuf.data(), n, 32768.0f, h_frames, max_frames, h_vfo_out, h_audio, vfo_counts, nullptr);
                }
                if (rc != B200_OK) { return rc; }
                onBlock(n);
            }
            return B200_OK;
        }

    private:
        void int16Loop() {
            const int n = blockSize();
            std::vector<int16_t> in((size_t)n * 2);
            while (true) {
                reader->readSamples(in.data(), (size_t)n * 2 * sizeof(int16_t));
                float* out = reinterpret_cast<float*>(stream.writeBuf);
                for (int i = 0; i < 2 * n; i++) { out[i] = (float)in[i] / 32768.0f; }     // volk_16i_s32f_convert_32f
                if (!stream.swap(n)) { break; }
            }
        }
        void floatLoop() {
            const int n = blockSize();
            while (true) {
                reader->readSamples(stream.writeBuf, (size_t)n * sizeof(complex_t));
                if (!stream.swap(n)) { break; }
            }
        }

        std::unique_ptr<wav::Reader> reader;
        bool isFloat = false;
        std::thread worker;
        bool running = false;
        std::vector<int16_t> ibuf;
        std::vector<float> fbuf;
    };
}

// wav::Reader — the baseband file reader the file source uses (reference: core/src/utils/wav.h:119-215 and
// source_modules/file_source/src/main.cpp:396-448).  Same members: isValid / error / getBitDepth / getChannelCount /
// getSampleRate / readSamples (wraps to the first sample at end of file, so a capture loops for ever) /
// readSamples2 (plain read, returns the byte count) / rewind / close.  One addition for the GPU path:
// getSampleType() tells int16 PCM from float32, so a caller can hand int16 blocks to
// b200_frontend_process_host_i16 (4 bytes per sample over PCIe, converted on the device) instead of converting on
// the host as the file source's worker does (main.cpp:409).
#pragma once
#include <cstdint>
#include <cstring>
#include <fstream>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

namespace wav {
    enum SampleType { SAMP_TYPE_UINT8, SAMP_TYPE_INT16, SAMP_TYPE_INT32, SAMP_TYPE_FLOAT32 };
    enum Codec { CODEC_PCM = 1, CODEC_FLOAT = 3 };

    class Reader {
    public:
        std::string error;

        explicit Reader(const std::string& path) : file(path.c_str(), std::ios::binary) {
            if (!file.is_open()) { error = "cannot open file"; return; }
            file.read(reinterpret_cast<char*>(&hdr), sizeof(hdr));
            if (file.gcount() != (std::streamsize)sizeof(hdr) || memcmp(hdr.signature, "RIFF", 4) != 0 || memcmp(hdr.fileType, "WAVE", 4) != 0) {
                error = "signature mismatch";
                return;
            }
            valid = true;
        }

        uint16_t getBitDepth() const { return hdr.bitDepth; }
        uint16_t getChannelCount() const { return hdr.channelCount; }
        uint32_t getSampleRate() const { return hdr.sampleRate; }
        bool isValid() const { return valid; }
        SampleType getSampleType() const {
            if (hdr.sampleType == CODEC_FLOAT) { return SAMP_TYPE_FLOAT32; }
            return hdr.bitDepth == 8 ? SAMP_TYPE_UINT8 : hdr.bitDepth == 32 ? SAMP_TYPE_INT32 : SAMP_TYPE_INT16;
        }

        // fills all `size` bytes; at end of file the read continues from the first sample
        void readSamples(void* data, size_t size) {
            char* d = static_cast<char*>(data);
            size_t got = 0;
            while (got < size) {
                file.read(d + got, (std::streamsize)(size - got));
                const size_t r = (size_t)file.gcount();
                got += r;
                if (got < size) {
                    file.clear();
                    file.seekg(sizeof(hdr));
                    if (r == 0 && looped) { memset(d + got, 0, size - got); break; }   // a file with no samples at all
                    looped = true;
                }
            }
            looped = false;
            bytesRead += size;
        }

        // plain read; returns the number of bytes delivered (less than `size` at end of file)
        size_t readSamples2(void* data, size_t size) {
            file.read(static_cast<char*>(data), (std::streamsize)size);
            return (size_t)file.gcount();
        }

        void rewind() { file.clear(); file.seekg(sizeof(hdr)); }
        void close() { file.close(); }

    private:
#pragma pack(push, 1)
        struct Header {
            char signature[4];             // "RIFF"
            uint32_t fileSize;
            char fileType[4];              // "WAVE"
            char formatMarker[4];          // "fmt "
            uint32_t formatHeaderLength;   // 16
            uint16_t sampleType;           // 1 PCM, 3 float
            uint16_t channelCount;
            uint32_t sampleRate;
            uint32_t bytesPerSecond;
            uint16_t bytesPerSample;
            uint16_t bitDepth;
            char dataMarker[4];            // "data"
            uint32_t dataSize;
        } hdr{};
#pragma pack(pop)
        bool valid = false, looped = false;
        std::ifstream file;
        size_t bytesRead = 0;
    };

    // wav::Writer — the recorder's file writer (reference: core/src/utils/wav.h:41-76, wav.cpp:20-182 over riff::Writer,
    // utils/riff.cpp; used by misc_modules/recorder/src/main.cpp:168-177,225,528-558).  Same members: open / isOpen / close /
    // setChannels / setSamplerate / setFormat / setSampleType / getSamplesWritten / write; the setters throw
    // std::runtime_error while a file is open, as the reference's do.  The container is host code.  The float -> PCM
    // conversion of write() is not: for the uint8 / int16 / int32 sample types this writer takes samples that
    // b200_convert_pcm already converted on the device (writePcm), so recorded data crosses PCIe at its file size;
    // write(float*) serves SAMP_TYPE_FLOAT32, where the reference writes the floats as they are.
    enum Format { FORMAT_WAV, FORMAT_RF64 };

    class Writer {
    public:
        Writer(int channels = 2, uint64_t samplerate = 48000, Format format = FORMAT_WAV, SampleType type = SAMP_TYPE_INT16) {
            if (channels < 1) { throw std::runtime_error("Channel count must be greater or equal to 1"); }
            if (!samplerate) { throw std::runtime_error("Samplerate must be non-zero"); }
            _channels = channels; _samplerate = samplerate; _format = format; _type = type;
        }
        ~Writer() { close(); }

        bool open(std::string path) {
            std::lock_guard<std::recursive_mutex> lck(mtx);
            if (file.is_open()) { close(); }
            samplesWritten = 0; dataBytes = 0;
            const int bits = (_type == SAMP_TYPE_UINT8) ? 8 : (_type == SAMP_TYPE_INT16) ? 16 : 32;
            bytesPerSamp = (size_t)(bits / 8) * (size_t)_channels;
            file.open(path.c_str(), std::ios::out | std::ios::binary);
            if (!file.is_open()) { return false; }
            // "RIFF" <size> "WAVE" | "fmt " 16 <format header> | "data" <size>; the sizes are patched in by close()
            put("RIFF", 4); put32(0); put("WAVE", 4);
            put("fmt ", 4); put32(16);
            put16((uint16_t)((_type == SAMP_TYPE_FLOAT32) ? CODEC_FLOAT : CODEC_PCM)); put16((uint16_t)_channels);
            put32((uint32_t)_samplerate); put32((uint32_t)(bytesPerSamp * _samplerate));
            put16((uint16_t)bytesPerSamp); put16((uint16_t)bits);
            put("data", 4); put32(0);
            return true;
        }
        bool isOpen() { std::lock_guard<std::recursive_mutex> lck(mtx); return file.is_open(); }
        void close() {
            std::lock_guard<std::recursive_mutex> lck(mtx);
            if (!file.is_open()) { return; }
            file.seekp(40); put32((uint32_t)dataBytes);
            file.seekp(4); put32((uint32_t)(4 + (8 + 16) + (8 + dataBytes)));
            file.close();
        }

        void setChannels(int channels) {
            std::lock_guard<std::recursive_mutex> lck(mtx);
            guard();
            if (channels < 1) { throw std::runtime_error("Channel count must be greater or equal to 1"); }
            _channels = channels;
        }
        void setSamplerate(uint64_t samplerate) {
            std::lock_guard<std::recursive_mutex> lck(mtx);
            guard();
            if (!samplerate) { throw std::runtime_error("Samplerate must be non-zero"); }
            _samplerate = samplerate;
        }
        void setFormat(Format format) { std::lock_guard<std::recursive_mutex> lck(mtx); guard(); _format = format; }
        void setSampleType(SampleType type) { std::lock_guard<std::recursive_mutex> lck(mtx); guard(); _type = type; }
        SampleType getSampleType() const { return _type; }
        int getChannels() const { return _channels; }
        size_t getSamplesWritten() { return samplesWritten; }

        // `count` samples (frames of `channels` values) already in the file's sample type: the output of b200_convert_pcm
        // copied to the host, or floats for SAMP_TYPE_FLOAT32
        void writePcm(const void* data, int count) {
            std::lock_guard<std::recursive_mutex> lck(mtx);
            if (!file.is_open()) { return; }
            const size_t bytes = (size_t)count * bytesPerSamp;
            file.write(static_cast<const char*>(data), (std::streamsize)bytes);
            dataBytes += bytes;
            samplesWritten += (size_t)count;
        }
        // wav::Writer::write for float files (wav.cpp:168-170)
        void write(float* samples, int count) {
            if (_type != SAMP_TYPE_FLOAT32) { throw std::runtime_error("PCM sample types are converted on the device: b200_convert_pcm + writePcm"); }
            writePcm(samples, count);
        }

    private:
        void guard() { if (file.is_open()) { throw std::runtime_error("Cannot change parameters while file is open"); } }
        void put(const char* p, size_t n) { file.write(p, (std::streamsize)n); }
        void put32(uint32_t v) { file.write(reinterpret_cast<const char*>(&v), 4); }
        void put16(uint16_t v) { file.write(reinterpret_cast<const char*>(&v), 2); }
        std::recursive_mutex mtx;
        std::ofstream file;
        int _channels; uint64_t _samplerate; Format _format; SampleType _type;
        size_t bytesPerSamp = 0, samplesWritten = 0, dataBytes = 0;
    };
}

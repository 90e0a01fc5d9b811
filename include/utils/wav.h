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
}

// A module in the reference's plug-in format (module.h ABI) whose instances each own one GPU channelizer:
// dsp::channel::RxVFO over libb200dsp feeding an NFM / AM / USB demodulator with the radio module's IF rates
// (decoder_modules/radio/src/demodulators/{nfm,am,usb}.h).  It shows the shape of a decoder module after the switch:
// the instance creates its VFO in enable(), hands consumers the VFO's output stream through getInterface(), and
// answers the debug server's commands; everything numerical happens behind the C ABI.
//
//   g++ -std=c++17 -shared -fPIC -Iinclude examples/b200_channel_module.cpp -Lsdrplusplusbrown_b200 -lb200dsp -o b200_channel.so
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>

#include <dsp/channel/rx_vfo.h>
#include <module.h>

SDRPP_MOD_INFO{
    /* Name:            */ "b200_channel",
    /* Description:     */ "RxVFO channelizer + demodulator on an NVIDIA B200 (libb200dsp)",
    /* Author:          */ "sdrplusplusbrown_b200",
    /* Version:         */ 0, 1, 0,
    /* Max instances    */ -1
};

namespace {
int g_initCalls = 0;

struct Mode { const char* name; double ifRate, bandwidth; };
const Mode kModes[] = { { "NFM", 50000.0, 12500.0 }, { "AM", 15000.0, 10000.0 }, { "USB", 24000.0, 2800.0 } };

class ChannelModule : public ModuleManager::Instance {
public:
    explicit ChannelModule(std::string n) : name(std::move(n)) {}
    ~ChannelModule() override { disable(); }

    void postInit() override {}

    // the input stream and rate come from whoever wires the module (the reference binds to sigpath::iqFrontEnd)
    void enable() override {
        if (enabled || !input) { return; }
        vfo.reset(new dsp::channel::RxVFO(input, inputRate, kModes[mode].ifRate, kModes[mode].bandwidth, offset));
        vfo->start();
        enabled = true;
    }

    void disable() override {
        if (!enabled) { return; }
        vfo->stop();
        vfo.reset();
        enabled = false;
    }

    bool isEnabled() override { return enabled; }

    void* getInterface(const char* what) override {
        if (!strcmp(what, "output")) { return vfo ? (void*)&vfo->out : nullptr; }      // dsp::stream<complex_t>*
        if (!strcmp(what, "input")) { return (void*)&input; }                          // dsp::stream<complex_t>** to set
        if (!strcmp(what, "input_rate")) { return (void*)&inputRate; }                 // double*
        return nullptr;
    }

    std::string handleDebugCommand(const std::string& cmd, const std::string& args) override {
        if (cmd == "set_demod") {
            for (int i = 0; i < 3; i++) {
                if (args == kModes[i].name) {
                    mode = i;
                    if (enabled) { vfo->setOutSamplerate(kModes[i].ifRate, kModes[i].bandwidth); }
                    return std::string("{\"demod\":\"") + kModes[i].name + "\"}";
                }
            }
            return "{\"error\":\"unknown demodulator\"}";
        }
        if (cmd == "set_offset") {
            offset = atof(args.c_str());
            if (enabled) { vfo->setOffset(offset); }
            return "{\"ok\":true}";
        }
        if (cmd == "status") {
            char b[160];
            snprintf(b, sizeof(b), "{\"name\":\"%s\",\"enabled\":%s,\"demod\":\"%s\",\"offset\":%.1f}", name.c_str(), enabled ? "true" : "false", kModes[mode].name, offset);
            return b;
        }
        return "{}";
    }

private:
    std::string name;
    dsp::stream<dsp::complex_t>* input = nullptr;
    double inputRate = 2.4e6, offset = 0.0;
    int mode = 0;
    bool enabled = false;
    std::unique_ptr<dsp::channel::RxVFO> vfo;
};
} // namespace

MOD_EXPORT void _INIT_() { g_initCalls++; }

MOD_EXPORT ModuleManager::Instance* _CREATE_INSTANCE_(std::string name) { return new ChannelModule(name); }

MOD_EXPORT void _DELETE_INSTANCE_(void* instance) { delete (ChannelModule*)instance; }

MOD_EXPORT void _END_() { g_initCalls = 0; }

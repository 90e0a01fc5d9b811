// Loads examples/b200_channel_module.cpp the way the reference's ModuleManager does (core/src/module.cpp:79-98:
// dlopen RTLD_NOW | RTLD_LOCAL, then the five symbols) and drives an instance through ModuleManager::Instance.
//   usage: module_loader <module.so> [gpu]
// Without "gpu" only the parts that need no device run (symbols, info, create / debug commands / delete).
#include <dlfcn.h>

#include <cstdio>
#include <cstring>
#include <string>

#include <dsp/stream.h>
#include <dsp/types.h>
#include <module.h>

#define CHECK(c) do { if (!(c)) { fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

int main(int argc, char** argv) {
    if (argc < 2) { return 2; }
    const bool gpu = argc > 2 && !strcmp(argv[2], "gpu");
    void* h = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
    if (!h) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 1; }
    auto* info = (ModuleManager::ModuleInfo_t*)dlsym(h, "_INFO_");
    auto init = (void (*)())dlsym(h, "_INIT_");
    auto create = (ModuleManager::Instance * (*)(std::string)) dlsym(h, "_CREATE_INSTANCE_");
    auto destroy = (void (*)(void*))dlsym(h, "_DELETE_INSTANCE_");
    auto end = (void (*)())dlsym(h, "_END_");
    CHECK(info && init && create && destroy && end);
    CHECK(!strcmp(info->name, "b200_channel") && info->maxInstances == -1 && info->versionMinor == 1);
    init();
    ModuleManager::Instance* inst = create("Channel A");
    CHECK(inst != nullptr);
    inst->postInit();
    CHECK(!inst->isEnabled());
    CHECK(inst->getInterface("output") == nullptr);
    CHECK(inst->getInterface("no such interface") == nullptr);
    CHECK(inst->handleDebugCommand("set_demod", "USB") == "{\"demod\":\"USB\"}");
    CHECK(inst->handleDebugCommand("set_demod", "XYZ").find("error") != std::string::npos);
    CHECK(inst->handleDebugCommand("status", "").find("\"enabled\":false") != std::string::npos);
    CHECK(inst->handleDebugCommand("anything else", "") == "{}");
    inst->enable();                                   // no input wired yet: stays disabled
    CHECK(!inst->isEnabled());
    if (gpu) {
        dsp::stream<dsp::complex_t> src;
        *(dsp::stream<dsp::complex_t>**)inst->getInterface("input") = &src;
        *(double*)inst->getInterface("input_rate") = 2.4e6;
        CHECK(inst->handleDebugCommand("set_demod", "NFM") == "{\"demod\":\"NFM\"}");
        inst->enable();
        CHECK(inst->isEnabled());
        auto* out = (dsp::stream<dsp::complex_t>*)inst->getInterface("output");
        CHECK(out != nullptr);
        int total = 0;
        for (int b = 0; b < 4; b++) {
            for (int i = 0; i < 12000; i++) { src.writeBuf[i] = dsp::complex_t{ 0.25f, -0.125f }; }
            CHECK(src.swap(12000));
            const int n = out->read();
            CHECK(n == 250);                          // 2.4 MS/s -> 50 kS/s
            total += n;
            out->flush();
        }
        CHECK(total == 1000);
        inst->disable();
        CHECK(!inst->isEnabled());
    }
    destroy(inst);
    end();
    dlclose(h);
    printf("module_loader ok%s\n", gpu ? " (gpu)" : "");
    return 0;
}

// SDR++ module ABI, as the reference's loader expects it (core/src/module.h:15-52,152; loader core/src/module.cpp:79-98):
// a module is a shared object exporting five C symbols
//     _INFO_             const ModuleManager::ModuleInfo_t
//     _INIT_()           once after dlopen
//     _CREATE_INSTANCE_(std::string name) -> ModuleManager::Instance*
//     _DELETE_INSTANCE_(void* instance)
//     _END_()            before unload
// and instances are driven through the ModuleManager::Instance virtual interface.  The declarations below keep the
// reference's member order (the vtable layout is part of the ABI: the loader calls through Instance*), so a module
// built against this header loads into the reference and the other way round, given the same libstdc++
// (std::string crosses the boundary by value).
#pragma once
#include <string>

#define MOD_EXPORT extern "C"
#define SDRPP_MOD_EXTENTSION ".so"

class ModuleManager {
public:
    struct ModuleInfo_t {
        const char* name;
        const char* description;
        const char* author;
        const int versionMajor;
        const int versionMinor;
        const int versionBuild;
        const int maxInstances;
    };

    class Instance {
    public:
        virtual ~Instance() {}
        virtual void postInit() = 0;
        virtual void enable() = 0;
        virtual void disable() = 0;
        virtual bool isEnabled() = 0;
        virtual void* getInterface(const char* name) { (void)name; return nullptr; }
        // automation channel of the fork's debug server: command name + module-defined arguments -> JSON text
        virtual std::string handleDebugCommand(const std::string& cmd, const std::string& args) { (void)cmd; (void)args; return "{}"; }
    };
};

#define SDRPP_MOD_INFO MOD_EXPORT const ModuleManager::ModuleInfo_t _INFO_

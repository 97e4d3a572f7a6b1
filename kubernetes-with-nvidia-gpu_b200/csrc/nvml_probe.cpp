// nvml_probe.cpp — see nvml_probe.h.  No link-time dependency on NVML: everything goes through dlsym.
#include "nvml_probe.h"

#include <dlfcn.h>

#include <cstring>
#include <map>

namespace nvmlprobe {

namespace {

using Dev = void*;
using FnInit = int (*)();
using FnByBus = int (*)(const char*, Dev*);
using FnClock = int (*)(Dev, int, unsigned*);
using FnTemp = int (*)(Dev, int, unsigned*);
using FnU32 = int (*)(Dev, unsigned*);
using FnU64 = int (*)(Dev, unsigned long long*);

struct Api {
    bool tried = false, ok = false;
    FnByBus by_bus = nullptr;
    FnClock clock = nullptr, max_clock = nullptr;
    FnTemp temp = nullptr;
    FnU32 power = nullptr, power_limit = nullptr;
    FnU64 reasons = nullptr;
    std::map<std::string, Dev> handles;
};

Api& api()
{
    static Api a;
    if (a.tried) return a;
    a.tried = true;
    void* lib = dlopen("libnvidia-ml.so.1", RTLD_NOW | RTLD_LOCAL);
    if (!lib) return a;
    auto sym = [&](const char* n) { return dlsym(lib, n); };
    FnInit init = reinterpret_cast<FnInit>(sym("nvmlInit_v2"));
    a.by_bus = reinterpret_cast<FnByBus>(sym("nvmlDeviceGetHandleByPciBusId_v2"));
    a.clock = reinterpret_cast<FnClock>(sym("nvmlDeviceGetClockInfo"));
    a.max_clock = reinterpret_cast<FnClock>(sym("nvmlDeviceGetMaxClockInfo"));
    a.temp = reinterpret_cast<FnTemp>(sym("nvmlDeviceGetTemperature"));
    a.power = reinterpret_cast<FnU32>(sym("nvmlDeviceGetPowerUsage"));
    a.power_limit = reinterpret_cast<FnU32>(sym("nvmlDeviceGetEnforcedPowerLimit"));
    a.reasons = reinterpret_cast<FnU64>(sym("nvmlDeviceGetCurrentClocksEventReasons"));
    if (!a.reasons) a.reasons = reinterpret_cast<FnU64>(sym("nvmlDeviceGetCurrentClocksThrottleReasons"));
    if (!init || !a.by_bus || !a.clock) return a;
    a.ok = init() == 0;
    return a;
}

}  // namespace

bool available() { return api().ok; }

bool sample(const char* pci_bus_id, Sample* out)
{
    Api& a = api();
    if (!a.ok || !pci_bus_id || !out) return false;
    Dev dev = nullptr;
    auto it = a.handles.find(pci_bus_id);
    if (it != a.handles.end()) {
        dev = it->second;
    } else {
        if (a.by_bus(pci_bus_id, &dev) != 0) return false;
        a.handles[pci_bus_id] = dev;
    }
    Sample s;
    if (a.clock(dev, 1 /* NVML_CLOCK_SM */, &s.sm_mhz) != 0) return false;
    a.clock(dev, 2 /* NVML_CLOCK_MEM */, &s.mem_mhz);
    if (a.max_clock) a.max_clock(dev, 1, &s.sm_max_mhz);
    if (a.temp) a.temp(dev, 0 /* NVML_TEMPERATURE_GPU */, &s.temp_c);
    unsigned mw = 0;
    if (a.power && a.power(dev, &mw) == 0) s.power_w = mw / 1000.0;
    if (a.power_limit && a.power_limit(dev, &mw) == 0) s.power_limit_w = mw / 1000.0;
    if (a.reasons) a.reasons(dev, &s.reasons);
    *out = s;
    return true;
}

std::string reasons_text(unsigned long long mask)
{
    static const struct { unsigned long long bit; const char* name; } names[] = {
        {kSwPowerCap, "sw_power_cap"}, {kHwSlowdown, "hw_slowdown"}, {kSwThermal, "sw_thermal_slowdown"},
        {kHwThermal, "hw_thermal_slowdown"}, {kHwPowerBrake, "hw_power_brake"},
    };
    std::string t;
    for (auto& n : names)
        if (mask & n.bit) t += (t.empty() ? "" : ",") + std::string(n.name);
    return t.empty() ? "none" : t;
}

}  // namespace nvmlprobe

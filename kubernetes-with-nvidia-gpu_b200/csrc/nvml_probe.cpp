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
using FnEcc = int (*)(Dev, int, int, unsigned long long*);
using FnLink = int (*)(Dev, unsigned, int*);

struct Api {
    bool tried = false, ok = false;
    FnByBus by_bus = nullptr;
    FnClock clock = nullptr, max_clock = nullptr;
    FnTemp temp = nullptr;
    FnU32 power = nullptr, power_limit = nullptr;
    FnU64 reasons = nullptr;
    FnU32 pcie_gen = nullptr, pcie_width = nullptr, pcie_gen_max = nullptr, pcie_width_max = nullptr;
    FnEcc ecc = nullptr;
    FnLink nvlink_state = nullptr;
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
    a.pcie_gen = reinterpret_cast<FnU32>(sym("nvmlDeviceGetCurrPcieLinkGeneration"));
    a.pcie_width = reinterpret_cast<FnU32>(sym("nvmlDeviceGetCurrPcieLinkWidth"));
    a.pcie_gen_max = reinterpret_cast<FnU32>(sym("nvmlDeviceGetMaxPcieLinkGeneration"));
    a.pcie_width_max = reinterpret_cast<FnU32>(sym("nvmlDeviceGetMaxPcieLinkWidth"));
    a.ecc = reinterpret_cast<FnEcc>(sym("nvmlDeviceGetTotalEccErrors"));
    a.nvlink_state = reinterpret_cast<FnLink>(sym("nvmlDeviceGetNvLinkState"));
    if (!init || !a.by_bus || !a.clock) return a;
    a.ok = init() == 0;
    return a;
}

}  // namespace

bool available() { return api().ok; }

static bool handle_for(Api& a, const char* pci_bus_id, Dev* dev)
{
    auto it = a.handles.find(pci_bus_id);
    if (it != a.handles.end()) {
        *dev = it->second;
        return true;
    }
    if (a.by_bus(pci_bus_id, dev) != 0) return false;
    a.handles[pci_bus_id] = *dev;
    return true;
}

bool health(const char* pci_bus_id, Health* out)
{
    Api& a = api();
    if (!a.ok || !pci_bus_id || !out) return false;
    Dev dev = nullptr;
    if (!handle_for(a, pci_bus_id, &dev)) return false;
    Health h;
    unsigned v = 0;
    if (a.pcie_gen && a.pcie_gen(dev, &v) == 0) h.pcie_gen = (int)v;
    if (a.pcie_width && a.pcie_width(dev, &v) == 0) h.pcie_width = (int)v;
    if (a.pcie_gen_max && a.pcie_gen_max(dev, &v) == 0) h.pcie_gen_max = (int)v;
    if (a.pcie_width_max && a.pcie_width_max(dev, &v) == 0) h.pcie_width_max = (int)v;
    if (a.nvlink_state) {
        int total = 0, active = 0;
        for (unsigned link = 0; link < 18; ++link) {      // NVLink 5: 18 links per GPU
            int st = 0;
            if (a.nvlink_state(dev, link, &st) != 0) break;
            ++total;
            active += st == 1;
        }
        if (total) {
            h.nvlinks_total = total;
            h.nvlinks_active = active;
        }
    }
    if (a.ecc) {
        unsigned long long c = 0;
        // (errorType: 0 corrected, 1 uncorrected; counterType: 0 volatile)
        if (a.ecc(dev, 0, 0, &c) == 0) h.ecc_corrected = (long long)c;
        if (a.ecc(dev, 1, 0, &c) == 0) h.ecc_uncorrected = (long long)c;
    }
    *out = h;
    return true;
}

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

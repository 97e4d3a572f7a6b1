// nvml_probe.h — clocks, temperature, power and throttle reasons for the discovery report and the
// burn-in summary, read through NVML if libnvidia-ml.so.1 can be dlopen'ed (the NVIDIA container
// runtime injects it with the "utility" capability).  Optional by design: when the library is not
// there the launcher says so and carries on — these readings never decide a verdict unless
// --fail-on-throttle is given.  This is the part of `nvidia-smi`'s table (reference README.md:335)
// a validation run actually needs.
#pragma once

#include <string>

namespace nvmlprobe {

struct Sample {
    unsigned sm_mhz = 0, sm_max_mhz = 0, mem_mhz = 0, temp_c = 0;
    double power_w = 0, power_limit_w = 0;
    unsigned long long reasons = 0;   // nvmlClocksEventReasons bit mask
};

// reasons that mean the GPU is being held back by heat or a hardware brake (not the power cap)
constexpr unsigned long long kHwSlowdown = 0x8, kSwThermal = 0x20, kHwThermal = 0x40, kHwPowerBrake = 0x80;
constexpr unsigned long long kSwPowerCap = 0x4;
constexpr unsigned long long kBadReasons = kHwSlowdown | kSwThermal | kHwThermal | kHwPowerBrake;

// link and memory-error state: what a node check wants to know besides clocks
struct Health {
    int pcie_gen = -1, pcie_width = -1, pcie_gen_max = -1, pcie_width_max = -1;
    int nvlinks_active = -1, nvlinks_total = -1;
    long long ecc_corrected = -1, ecc_uncorrected = -1;    // volatile (since driver load) totals; -1 = not reported
};

bool available();                                         // dlopen + nvmlInit succeeded (cached)
bool health(const char* pci_bus_id, Health* out);
bool sample(const char* pci_bus_id, Sample* out);         // false if the device cannot be read
std::string reasons_text(unsigned long long mask);        // "sw_power_cap,hw_thermal_slowdown" or "none"

}  // namespace nvmlprobe

// launcher_common.h — what the translation units of `gpu-validate` share: the argument block, the
// exit codes (the pod's machine-readable verdict, reference README.md:309), the verdict helpers and
// the mode entry points.  launcher.cpp says what the binary is and which reference lines it replaces.
#pragma once

#include <algorithm>
#include <cerrno>
#include <chrono>
#include <cinttypes>
#include <csignal>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <string>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <poll.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

#include "b200v.h"
#include "hostcheck.h"
#include "nvml_probe.h"

namespace gpuvalidate {

// exit codes: the pod's machine-readable verdict (README:309)
enum Exit { EX_PASS = 0, EX_MISMATCH = 1, EX_NO_DEVICE = 2, EX_ARCH = 3, EX_CUDA = 4, EX_PEER = 5, EX_USAGE_ERR = 64 };

constexpr double kNominalGBs = 8000.0;   // BASELINE.json: "~8 TB/s per-GPU peak"
constexpr double kNvlinkGBs = 900.0;     // per direction per GPU

struct Args {
    std::string mode = "smoke";
    int device = -1;             // smoke: -1 = every visible GPU; shard children get their own
    uint64_t n = 50000;          // smoke: cuda-samples vectorAdd size
    bool n_set = false;
    int log2n = -1;
    std::string dtype = "f32";
    std::string op = "";
    int iters = 20;
    int warmup = 3;
    double seconds = 10.0;
    int gpus = 0;                // 0 = all visible
    uint32_t seed = 20261106u;
    bool inject_fault = false;
    bool fail_on_throttle = false;   // burn-in: thermal / hardware slowdown reasons fail the run
    double hbm_fraction = 0.0;       // burn-in: size the three arrays to this share of each GPU's free HBM
    uint64_t p2p_bytes = 1ull << 30;
    int min_log2_bytes = 20, max_log2_bytes = 34;
    double measured_copy_gbs = 0.0;   // optional second denominator (--measured-copy-gbs)
    std::string json_path;
    // shard child plumbing
    int rank = 0, world = 1, go_fd = -1, out_fd = -1;
    bool quiet = false;
};

int exit_code_for(int status);
// Prints the reason and the verdict line `Test FAILED`; returns the process exit code.
int fail(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
int fail_status(int status, const char* what);
// Prints `Test PASSED`; returns 0.
int pass();

#define TRY(call)                                                             \
    do {                                                                      \
        int st__ = (call);                                                    \
        if (st__ != B200V_OK) return ::gpuvalidate::fail_status(st__, #call); \
    } while (0)

inline int dtype_id(const std::string& s) { return s == "bf16" ? B200V_BF16 : B200V_F32; }
inline size_t esize(int dtype) { return dtype == B200V_BF16 ? 2 : 4; }
const char* op_name(int op);
int op_id(const std::string& s);

// discovery report: what `kubectl logs` shows in place of the nvidia-smi table (README:335)
int print_devices(int* count_out, bool all_details);

// device arrays and checksum scratch used by bench / node / burnin / p2p
struct DevArrays {
    void *x = nullptr, *y = nullptr, *d = nullptr, *ws = nullptr, *out = nullptr;
    void release()
    {
        b200v_free(x); b200v_free(y); b200v_free(d); b200v_free(ws); b200v_free(out);
        x = y = d = ws = out = nullptr;
    }
};
int alloc_arrays(DevArrays& v, size_t bytes);
int device_checksum(const DevArrays& v, const void* buf, size_t nbytes, uint64_t word_offset, uint64_t out[2]);
double now_s();

// modes (launcher.cpp: smoke, list, host-reference; mode_bench.cpp; mode_node.cpp; mode_p2p.cpp)
int mode_smoke(const Args& a);
int mode_list(const Args& a);
int mode_exit_codes(const Args& a);
int mode_host_reference(const Args& a);
int mode_bench(const Args& a);
int mode_node(const Args& a, bool burnin, bool print_verdict);
int mode_shard_child(const Args& a);
int mode_p2p(const Args& a, bool print_verdict);

}  // namespace gpuvalidate

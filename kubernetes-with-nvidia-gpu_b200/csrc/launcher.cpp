// launcher.cpp — `gpu-validate`, the pod entrypoint.
//
// Drop-in for the payload of the reference's Step 9 test pod (/root/reference/README.md:302-319):
// where the reference runs `/bin/sh -c nvidia-smi` (README:313-314) and a human looks for a GPU in
// `kubectl logs` (README:332-335), this binary lists the GPUs the container was granted
// (README:315-317), runs a vectorAdd on the GPU, checks it bit for bit against a host loop, prints
// `Test PASSED` or `Test FAILED` as the last verdict line and exits 0 / non-zero, which is what
// `restartPolicy: OnFailure` (README:309) acts on.  With no arguments it does exactly that, on every
// GPU the pod was granted (one, under the reference's limit), well inside the 15 s the guide waits (README:326).
//
// All GPU work goes through the C ABI in include/b200v.h.  The host loops in hostcheck.cpp are the
// verification half of the workload and the OpenMP baseline; there is no CPU fallback for the
// GPU path: without a usable sm_100 device the verdict is FAILED.
//
//   gpu-validate                       vectorAdd N=50,000 on every granted GPU (BASELINE.json configs[0/1] contract)
//   gpu-validate --mode list           device discovery only (the nvidia-smi-style listing)
//   gpu-validate --mode bench          STREAM copy/scale/add/triad sweep, fp32+bf16, + OpenMP baseline (configs[2])
//   gpu-validate --mode node           one process per GPU, sharded triad, max over ranks       (configs[3])
//   gpu-validate --mode p2p            all ordered GPU pairs: peer copy, peer-load and peer-store kernels, verified (configs[4])
//   gpu-validate --mode burnin         per-GPU triad burn-in, every pass verified               (configs[4])
//   gpu-validate --mode full           burnin + p2p
//   gpu-validate --mode host-reference the host loop alone (no GPU is touched or validated)
//   gpu-validate --mode exit-codes     the status -> exit code table (no GPU is touched)
#include "launcher_common.h"

namespace gpuvalidate {

int exit_code_for(int status)
{
    switch (status) {
    case B200V_OK: return EX_PASS;
    case B200V_ERR_NO_DEVICE: return EX_NO_DEVICE;
    case B200V_ERR_ARCH: return EX_ARCH;
    case B200V_ERR_PEER: return EX_PEER;
    case B200V_ERR_MISMATCH: return EX_MISMATCH;
    default: return EX_CUDA;
    }
}

// Prints the reason and the verdict line; returns the process exit code.
int fail(int code, const char* fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    fprintf(stderr, "gpu-validate: %s\n", buf);
    printf("reason: %s\n", buf);
    printf("Test FAILED\n");
    fflush(stdout);
    return code == EX_PASS ? EX_MISMATCH : code;
}

int fail_status(int status, const char* what)
{
    return fail(exit_code_for(status), "%s: %s: %s", what, b200v_status_name(status), b200v_last_error());
}

int pass()
{
    printf("Test PASSED\n");
    fflush(stdout);
    return EX_PASS;
}

const char* op_name(int op)
{
    static const char* n[] = {"add", "copy", "scale", "triad"};
    return n[op & 3];
}
int op_id(const std::string& s)
{
    if (s == "copy") return B200V_OP_COPY;
    if (s == "scale") return B200V_OP_SCALE;
    if (s == "triad") return B200V_OP_TRIAD;
    return B200V_OP_ADD;
}

// ----------------------------------------------------------------------------------------------
// discovery report: what `kubectl logs` shows in place of the nvidia-smi table (README:335)

int print_devices(int* count_out, bool all_details)
{
    int count = 0;
    int st = b200v_device_count(&count);
    if (st != B200V_OK) {
        printf("[discovery] no usable CUDA device: %s\n", b200v_last_error());
        printf("            check: nvidia.com/gpu in the pod's resources.limits, the NVIDIA container runtime,\n"
               "            and that the host driver is new enough for B200 (R570 or later)\n");
        return st;
    }
    b200v_device_info info;
    for (int d = 0; d < count; ++d) {
        memset(&info, 0, sizeof info);
        info.struct_size = sizeof info;
        // the detailed listing visits every device (and so creates a context on each) to read its
        // free memory; the short one leaves the other GPUs untouched
        if (all_details && (st = b200v_set_device(d)) != B200V_OK) return st;
        st = b200v_device_info_get(d, &info);
        if (st != B200V_OK) return st;
        if (d == 0)
            printf("[discovery] %d CUDA device%s visible; driver API %d.%d, runtime %d.%d\n", count, count == 1 ? "" : "s",
                   info.driver_version / 1000, (info.driver_version % 1000) / 10, info.runtime_version / 1000,
                   (info.runtime_version % 1000) / 10);
        char free_txt[48] = "";
        if (info.hbm_free_bytes) snprintf(free_txt, sizeof free_txt, " (%.0f MiB free)", info.hbm_free_bytes / 1048576.0);
        printf("  GPU %d: %s  sm_%d%d  %d SMs  %.0f MiB HBM%s  L2 %.0f MiB  %s  %s\n", d, info.name, info.cc_major,
               info.cc_minor, info.sm_count, info.hbm_bytes / 1048576.0, free_txt, info.l2_bytes / 1048576.0, info.pci_bus_id,
               info.uuid);
        if (all_details)
            printf("         SM clock %d MHz, memory clock %d MHz, bus %d bit, ECC %s\n", info.clock_khz / 1000,
                   info.mem_clock_khz / 1000, info.mem_bus_bits, info.ecc_enabled ? "on" : "off");
        nvmlprobe::Sample nv;
        if (nvmlprobe::sample(info.pci_bus_id, &nv))
            printf("         now: SM %u MHz (max %u), memory %u MHz, %u C, %.0f W of %.0f W, clock limits: %s\n", nv.sm_mhz,
                   nv.sm_max_mhz, nv.mem_mhz, nv.temp_c, nv.power_w, nv.power_limit_w, nvmlprobe::reasons_text(nv.reasons).c_str());
        else if (d == 0 && all_details)
            printf("         (NVML not loadable in this container: no live clocks / temperature / power)\n");
        nvmlprobe::Health hl;
        if (all_details && nvmlprobe::health(info.pci_bus_id, &hl)) {
            printf("         PCIe gen %d x%d (max gen %d x%d)%s", hl.pcie_gen, hl.pcie_width, hl.pcie_gen_max, hl.pcie_width_max,
                   hl.pcie_width >= 0 && hl.pcie_width < hl.pcie_width_max ? "  <-- WARNING: link is narrower than the slot allows" : "");
            if (hl.nvlinks_total >= 0)
                printf(", NVLink %d of %d links up%s", hl.nvlinks_active, hl.nvlinks_total,
                       hl.nvlinks_active < hl.nvlinks_total ? "  <-- WARNING: links down" : "");
            if (hl.ecc_uncorrected >= 0)
                printf(", ECC since driver load: %lld corrected, %lld uncorrected%s", hl.ecc_corrected, hl.ecc_uncorrected,
                       hl.ecc_uncorrected > 0 ? "  <-- WARNING" : "");
            printf("\n");
        }
    }
    if (all_details) b200v_set_device(0);
    if (all_details && count > 1) {
        std::vector<int> m((size_t)count * count);
        if (b200v_peer_matrix(count, m.data()) == B200V_OK) {
            printf("  peer access (row can map column):\n");
            for (int i = 0; i < count; ++i) {
                printf("    GPU %d:", i);
                for (int j = 0; j < count; ++j) printf(" %c", i == j ? '-' : (m[(size_t)i * count + j] ? 'Y' : 'n'));
                printf("\n");
            }
        }
    }
    *count_out = count;
    return B200V_OK;
}

// ----------------------------------------------------------------------------------------------
// smoke: the cuda-samples vectorAdd flow — host arrays, H2D, kernel, D2H, host re-computation.

// One GPU's share of the smoke test: H2D, kernel, D2H, bit-compare with the host loop's result.  Runs on its own
// host thread (one thread per device, as include/b200v.h asks), so it prints nothing: the outcome goes into `out`.
struct SmokeOutcome {
    int exit_code = EX_PASS;
    std::string line;      // the per-GPU log line, or the failure reason
};

static void smoke_one_gpu(int device, const std::vector<float>& hA, const std::vector<float>& hB, const std::vector<float>& want,
                          bool inject, SmokeOutcome& out)
{
    const size_t n = hA.size();
    std::vector<float> hC(n);
    void *dA = nullptr, *dB = nullptr, *dC = nullptr;
    char buf[512];
    int st = B200V_OK;
    const char* what = "";
#define STEP(call)                         \
    if (st == B200V_OK) {                  \
        st = (call);                       \
        if (st != B200V_OK) what = #call;  \
    }
    STEP(b200v_set_device(device));
    STEP(b200v_require_sm100(device));
    STEP(b200v_malloc(&dA, n * 4));
    STEP(b200v_malloc(&dB, n * 4));
    STEP(b200v_malloc(&dC, n * 4));
    STEP(b200v_memcpy_h2d(dA, hA.data(), n * 4, nullptr));
    STEP(b200v_memcpy_h2d(dB, hB.data(), n * 4, nullptr));
    STEP(b200v_add_f32((float*)dC, (const float*)dA, (const float*)dB, n, nullptr));
    STEP(b200v_memcpy_d2h(hC.data(), dC, n * 4, nullptr));
    STEP(b200v_device_sync());
#undef STEP
    if (st != B200V_OK) {       // b200v_last_error() is per thread: read it here
        snprintf(buf, sizeof buf, "GPU %d: %s: %s: %s", device, what, b200v_status_name(st), b200v_last_error());
        out.exit_code = exit_code_for(st);
        out.line = buf;
    }
    b200v_free(dA);
    b200v_free(dB);
    b200v_free(dC);
    if (st != B200V_OK) return;

    if (inject && n) {
        uint32_t u;
        memcpy(&u, &hC[n / 2], 4);
        u ^= 1u;   // one flipped bit in one element must turn the verdict
        memcpy(&hC[n / 2], &u, 4);
    }
    const size_t bad = hostcheck::first_mismatch(hostcheck::F32, hC.data(), want.data(), n);
    if (bad != n) {
        uint32_t g, w;
        memcpy(&g, &hC[bad], 4);
        memcpy(&w, &want[bad], 4);
        snprintf(buf, sizeof buf, "result verification failed at element %zu on GPU %d: device 0x%08x, host loop 0x%08x", bad, device, g, w);
        out.exit_code = EX_MISMATCH;
    } else {
        snprintf(buf, sizeof buf, "[vectorAdd] GPU %d: %zu of %zu elements bit-identical to the host loop", device, n, n);
    }
    out.line = buf;
}

int mode_smoke(const Args& a)
{
    printf("[gpu-validate] B200 validation workload (vectorAdd, fp32, %" PRIu64 " elements)\n", a.n);
    int count = 0;
    int st = print_devices(&count, false);
    if (st != B200V_OK) return fail_status(st, "device discovery");
    if (a.device >= count) return fail(EX_NO_DEVICE, "device %d requested, %d visible", a.device, count);
    // Every GPU the container was granted (resources.limits nvidia.com/gpu: N, README:315-317) is checked; --device D
    // restricts the run to one.  With the reference's limit of 1 that is GPU 0, as before.  The GPUs are driven by one
    // host thread each, at the same time: creating eight CUDA contexts one after the other would take most of the
    // 15 s the guide waits before it reads the log (README:326).
    const int first = a.device >= 0 ? a.device : 0;
    const int last = a.device >= 0 ? a.device : count - 1;

    const size_t n = (size_t)a.n;
    std::vector<float> hA(n), hB(n), want(n);
    hostcheck::fill(hostcheck::F32, hA.data(), 0, n, a.seed, 0, hostcheck::UNIT);
    hostcheck::fill(hostcheck::F32, hB.data(), 0, n, a.seed, 1, hostcheck::UNIT);
    hostcheck::apply(hostcheck::ADD, hostcheck::F32, want.data(), hA.data(), hB.data(), 0.0f, n, false);

    printf("[vectorAdd] host -> device copy of A and B, kernel, device -> host copy of C, on %s\n",
           first == last ? "one GPU" : "every visible GPU");
    if (a.inject_fault && n) printf("[vectorAdd] --inject-fault: flipping the low bit of element %zu (GPU %d)\n", n / 2, last);
    fflush(stdout);
    std::vector<SmokeOutcome> outcome((size_t)(last - first + 1));
    std::vector<std::thread> workers;
    for (int d = first; d <= last; ++d)
        workers.emplace_back(smoke_one_gpu, d, std::cref(hA), std::cref(hB), std::cref(want), a.inject_fault && d == last,
                             std::ref(outcome[(size_t)(d - first)]));
    for (auto& w : workers) w.join();
    for (const auto& o : outcome)      // the log lists every GPU's result before the verdict
        if (o.exit_code == EX_PASS) printf("%s\n", o.line.c_str());
    for (const auto& o : outcome)
        if (o.exit_code != EX_PASS) return fail(o.exit_code, "%s", o.line.c_str());
    return pass();
}

// Prints how every library status maps to the pod's exit code and to the reference's troubleshooting
// checklists (README:339-357).  No GPU is touched: this is how the mapping for hardware this pool does
// not have (a non-sm_100 GPU -> 3, GPUs without peer access -> 5) is kept under test.
int mode_exit_codes(const Args&)
{
    static const struct { int status; const char* hint; } rows[] = {
        {B200V_OK, "pass"},
        {B200V_ERR_INVALID, "internal/usage error"},
        {B200V_ERR_NO_DEVICE, "README:353-357 'Pod Cannot Access GPU': check resources.limits nvidia.com/gpu, the container runtime, the driver"},
        {B200V_ERR_CUDA, "CUDA runtime error: see the reason line"},
        {B200V_ERR_ARCH, "the GPU is not a B200 (sm_100): this image carries sm_100a code only"},
        {B200V_ERR_ALIGN, "internal/usage error"},
        {B200V_ERR_PEER, "peer access between two GPUs is unavailable (NVLink/NVSwitch fabric, ACS, or the container's device set)"},
        {B200V_ERR_MISMATCH, "GPU result differs from the host loop"},
        {B200V_ERR_NOMEM, "device or pinned-host allocation failed"},
    };
    for (const auto& r : rows)
        printf("status %d %-19s -> exit %d  (%s)\n", r.status, b200v_status_name(r.status), exit_code_for(r.status), r.hint);
    printf("usage error             -> exit %d\n", EX_USAGE_ERR);
    return EX_PASS;
}

int mode_list(const Args&)
{
    int count = 0;
    int st = print_devices(&count, true);
    if (st != B200V_OK) return fail_status(st, "device discovery");
    return EX_PASS;
}

int mode_host_reference(const Args& a)
{
    const size_t n = (size_t)a.n;
    printf("[host-reference] host loop only: no GPU is used or validated by this mode\n");
    std::vector<float> hA(n), hB(n), serial(n), par(n);
    hostcheck::fill(hostcheck::F32, hA.data(), 0, n, a.seed, 0, hostcheck::UNIT);
    hostcheck::fill(hostcheck::F32, hB.data(), 0, n, a.seed, 1, hostcheck::UNIT);
    hostcheck::apply(hostcheck::ADD, hostcheck::F32, serial.data(), hA.data(), hB.data(), 0.0f, n, false);
    hostcheck::apply(hostcheck::ADD, hostcheck::F32, par.data(), hA.data(), hB.data(), 0.0f, n, true);
    if (a.inject_fault && n) {
        uint32_t u;
        memcpy(&u, &par[n / 2], 4);
        u ^= 1u;
        memcpy(&par[n / 2], &u, 4);
    }
    uint64_t ck[2], ckt[2];
    hostcheck::checksum(serial.data(), n * 4, 0, ck);
    hostcheck::expected_checksum(hostcheck::TRIAD, hostcheck::F32, 0, n, a.seed, hostcheck::UNIT, 3.0f, 0, ckt);
    printf("[host-reference] N=%zu seed=%u threads=%d add_checksum=%016" PRIx64 ":%016" PRIx64 " triad_checksum=%016" PRIx64
           ":%016" PRIx64 "\n",
           n, a.seed, hostcheck::max_threads(), ck[0], ck[1], ckt[0], ckt[1]);
    if (hostcheck::first_mismatch(hostcheck::F32, par.data(), serial.data(), n) != n)
        return fail(EX_MISMATCH, "serial and OpenMP host loops differ");
    return pass();
}

// ----------------------------------------------------------------------------------------------
// helpers shared by bench / node / burnin
int alloc_arrays(DevArrays& v, size_t bytes)
{
    int st;
    if ((st = b200v_malloc(&v.x, bytes))) return st;
    if ((st = b200v_malloc(&v.y, bytes))) return st;
    if ((st = b200v_malloc(&v.d, bytes))) return st;
    if ((st = b200v_malloc(&v.ws, b200v_checksum_workspace_bytes()))) return st;
    if ((st = b200v_malloc(&v.out, 16))) return st;
    return B200V_OK;
}

int device_checksum(const DevArrays& v, const void* buf, size_t nbytes, uint64_t word_offset, uint64_t out[2])
{
    int st = b200v_checksum(buf, nbytes, word_offset, v.ws, b200v_checksum_workspace_bytes(), (uint64_t*)v.out, nullptr);
    if (st) return st;
    if ((st = b200v_memcpy_d2h(out, v.out, 16, nullptr))) return st;
    return b200v_device_sync();
}

double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ----------------------------------------------------------------------------------------------

void usage()
{
    fprintf(stderr,
            "usage: gpu-validate [--mode smoke|list|bench|node|p2p|burnin|full|host-reference|exit-codes]\n"
            "                    [--device D] [--gpus G] [--n N | --log2n K] [--dtype f32|bf16] [--op add|copy|scale|triad]\n"
            "                    [--iters I] [--warmup W] [--seconds S] [--seed X] [--p2p-bytes B]\n"
            "                    [--min-log2-bytes a --max-log2-bytes b] [--measured-copy-gbs G] [--json FILE] [--inject-fault] [--fail-on-throttle]\n"
            "                    [--hbm-fraction F]   (burnin/full: arrays cover F of each GPU's free HBM instead of --log2n)\n"
            "no arguments: vectorAdd N=50000 on every visible GPU (--device D: one), prints Test PASSED / Test FAILED, exit code 0 / non-zero\n");
}

bool parse(int argc, char** argv, Args& a)
{
    for (int i = 1; i < argc; ++i) {
        std::string k = argv[i];
        auto val = [&](const char** out) {
            if (i + 1 >= argc) return false;
            *out = argv[++i];
            return true;
        };
        const char* v = nullptr;
        if (k == "--inject-fault") a.inject_fault = true;
        else if (k == "--fail-on-throttle") a.fail_on_throttle = true;
        else if (k == "--quiet") a.quiet = true;
        else if (k == "--help" || k == "-h") return false;
        else if (!val(&v)) return false;
        else if (k == "--mode") a.mode = v;
        else if (k == "--device") a.device = atoi(v);
        else if (k == "--gpus") a.gpus = atoi(v);
        else if (k == "--n") { a.n = strtoull(v, nullptr, 10); a.n_set = true; }
        else if (k == "--log2n") { a.log2n = atoi(v); if (a.log2n >= 0) { a.n = 1ull << a.log2n; a.n_set = true; } }
        else if (k == "--dtype") a.dtype = v;
        else if (k == "--op") a.op = v;
        else if (k == "--iters") a.iters = atoi(v);
        else if (k == "--warmup") a.warmup = atoi(v);
        else if (k == "--seconds") a.seconds = atof(v);
        else if (k == "--hbm-fraction") a.hbm_fraction = atof(v);
        else if (k == "--seed") a.seed = (uint32_t)strtoul(v, nullptr, 10);
        else if (k == "--p2p-bytes") a.p2p_bytes = strtoull(v, nullptr, 10);
        else if (k == "--min-log2-bytes") a.min_log2_bytes = atoi(v);
        else if (k == "--max-log2-bytes") a.max_log2_bytes = atoi(v);
        else if (k == "--measured-copy-gbs") a.measured_copy_gbs = atof(v);
        else if (k == "--json") a.json_path = v;
        else if (k == "--rank") a.rank = atoi(v);
        else if (k == "--world") a.world = atoi(v);
        else if (k == "--go-fd") a.go_fd = atoi(v);
        else if (k == "--out-fd") a.out_fd = atoi(v);
        else return false;
    }
    if (a.dtype != "f32" && a.dtype != "bf16") return false;
    if (!a.op.empty() && a.op != "add" && a.op != "copy" && a.op != "scale" && a.op != "triad") return false;
    if (a.hbm_fraction < 0 || a.hbm_fraction > 0.97) return false;
    if (a.iters < 1 || a.warmup < 0 || a.log2n > 40 || a.min_log2_bytes < 4 || a.max_log2_bytes > 40 ||
        a.min_log2_bytes > a.max_log2_bytes || a.world < 1 || a.rank < 0 || a.rank >= a.world)
        return false;
    return true;
}
}  // namespace gpuvalidate

using namespace gpuvalidate;

int main(int argc, char** argv)
{
    setvbuf(stdout, nullptr, _IOLBF, 0);
    signal(SIGPIPE, SIG_IGN);   // a shard child that died must not take the parent down through its go pipe
    Args a;
    if (!parse(argc, argv, a)) {
        usage();
        printf("Test FAILED\n");
        return EX_USAGE_ERR;
    }
    if (a.mode == "smoke") return mode_smoke(a);
    if (a.mode == "list") return mode_list(a);
    if (a.mode == "exit-codes") return mode_exit_codes(a);
    if (a.mode == "host-reference") return mode_host_reference(a);
    if (a.mode == "bench") return mode_bench(a);
    if (a.mode == "node") return mode_node(a, false, true);
    if (a.mode == "burnin") return mode_node(a, true, true);
    if (a.mode == "p2p") return mode_p2p(a, true);
    if (a.mode == "shard-child" || a.mode == "burnin-child") return mode_shard_child(a);
    if (a.mode == "full") {
        // BASELINE.json configs[4]: per-GPU triad burn-in, then the all-pairs NVLink check
        int rc = mode_node(a, true, false);
        if (rc != EX_PASS) return rc;
        rc = mode_p2p(a, false);
        if (rc != EX_PASS) return rc;
        return pass();
    }
    usage();
    printf("Test FAILED\n");
    return EX_USAGE_ERR;
}

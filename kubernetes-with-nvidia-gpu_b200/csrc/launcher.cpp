// launcher.cpp — `gpu-validate`, the pod entrypoint.
//
// Drop-in for the payload of the reference's Step 9 test pod (/root/reference/README.md:302-319):
// where the reference runs `/bin/sh -c nvidia-smi` (README:313-314) and a human looks for a GPU in
// `kubectl logs` (README:332-335), this binary lists the GPUs the container was granted
// (README:315-317), runs a vectorAdd on the GPU, checks it bit for bit against a host loop, prints
// `Test PASSED` or `Test FAILED` as the last verdict line and exits 0 / non-zero, which is what
// `restartPolicy: OnFailure` (README:309) acts on.  With no arguments it does exactly that, on
// device 0 only, well inside the 15 s the guide waits (README:326).
//
// All GPU work goes through the C ABI in include/b200v.h.  The host loops in hostcheck.cpp are the
// verification half of the workload and the OpenMP baseline; there is no CPU fallback for the
// GPU path: without a usable sm_100 device the verdict is FAILED.
//
//   gpu-validate                       vectorAdd N=50,000 on GPU 0          (BASELINE.json configs[0/1] contract)
//   gpu-validate --mode list           device discovery only (the nvidia-smi-style listing)
//   gpu-validate --mode bench          STREAM copy/scale/add/triad sweep, fp32+bf16, + OpenMP baseline (configs[2])
//   gpu-validate --mode node           one process per GPU, sharded triad, max over ranks       (configs[3])
//   gpu-validate --mode p2p            all ordered GPU pairs: peer copy + peer-load kernel, verified (configs[4])
//   gpu-validate --mode burnin         per-GPU triad burn-in, every pass verified               (configs[4])
//   gpu-validate --mode full           burnin + p2p
//   gpu-validate --mode host-reference the host loop alone (no GPU is touched or validated)
#include <algorithm>
#include <cerrno>
#include <chrono>
#include <cinttypes>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <fcntl.h>
#include <poll.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

#include "b200v.h"
#include "hostcheck.h"
#include "nvml_probe.h"

namespace {

// exit codes: the pod's machine-readable verdict (README:309)
enum Exit { EX_PASS = 0, EX_MISMATCH = 1, EX_NO_DEVICE = 2, EX_ARCH = 3, EX_CUDA = 4, EX_PEER = 5, EX_USAGE_ERR = 64 };

constexpr double kNominalGBs = 8000.0;   // BASELINE.json: "~8 TB/s per-GPU peak"
constexpr double kNvlinkGBs = 900.0;     // per direction per GPU

struct Args {
    std::string mode = "smoke";
    int device = 0;
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
    uint64_t p2p_bytes = 1ull << 30;
    int min_log2_bytes = 20, max_log2_bytes = 34;
    double measured_copy_gbs = 0.0;   // optional second denominator (--measured-copy-gbs)
    std::string json_path;
    // shard child plumbing
    int rank = 0, world = 1, go_fd = -1, out_fd = -1;
    bool quiet = false;
};

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

#define TRY(call)                                          \
    do {                                                   \
        int st__ = (call);                                 \
        if (st__ != B200V_OK) return fail_status(st__, #call); \
    } while (0)

int pass()
{
    printf("Test PASSED\n");
    fflush(stdout);
    return EX_PASS;
}

int dtype_id(const std::string& s) { return s == "bf16" ? B200V_BF16 : B200V_F32; }
size_t esize(int dtype) { return dtype == B200V_BF16 ? 2 : 4; }
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

int mode_smoke(const Args& a)
{
    printf("[gpu-validate] B200 validation workload (vectorAdd, fp32, %" PRIu64 " elements)\n", a.n);
    int count = 0;
    int st = print_devices(&count, false);
    if (st != B200V_OK) return fail_status(st, "device discovery");
    if (a.device >= count) return fail(EX_NO_DEVICE, "device %d requested, %d visible", a.device, count);
    TRY(b200v_set_device(a.device));
    TRY(b200v_require_sm100(a.device));

    const size_t n = (size_t)a.n;
    std::vector<float> hA(n), hB(n), hC(n), want(n);
    hostcheck::fill(hostcheck::F32, hA.data(), 0, n, a.seed, 0, hostcheck::UNIT);
    hostcheck::fill(hostcheck::F32, hB.data(), 0, n, a.seed, 1, hostcheck::UNIT);

    printf("[vectorAdd] host -> device copy of A and B, kernel, device -> host copy of C (GPU %d)\n", a.device);
    void *dA = nullptr, *dB = nullptr, *dC = nullptr;
    TRY(b200v_malloc(&dA, n * 4));
    TRY(b200v_malloc(&dB, n * 4));
    TRY(b200v_malloc(&dC, n * 4));
    TRY(b200v_memcpy_h2d(dA, hA.data(), n * 4, nullptr));
    TRY(b200v_memcpy_h2d(dB, hB.data(), n * 4, nullptr));
    TRY(b200v_add_f32((float*)dC, (const float*)dA, (const float*)dB, n, nullptr));
    TRY(b200v_memcpy_d2h(hC.data(), dC, n * 4, nullptr));
    TRY(b200v_device_sync());
    TRY(b200v_free(dA));
    TRY(b200v_free(dB));
    TRY(b200v_free(dC));

    if (a.inject_fault && n) {
        uint32_t u;
        memcpy(&u, &hC[n / 2], 4);
        u ^= 1u;   // one flipped bit in one element must turn the verdict
        memcpy(&hC[n / 2], &u, 4);
        printf("[vectorAdd] --inject-fault: flipped the low bit of element %zu\n", n / 2);
    }
    hostcheck::apply(hostcheck::ADD, hostcheck::F32, want.data(), hA.data(), hB.data(), 0.0f, n, false);
    const size_t bad = hostcheck::first_mismatch(hostcheck::F32, hC.data(), want.data(), n);
    if (bad != n) {
        uint32_t g, w;
        memcpy(&g, &hC[bad], 4);
        memcpy(&w, &want[bad], 4);
        return fail(EX_MISMATCH, "result verification failed at element %zu: device 0x%08x, host loop 0x%08x", bad, g, w);
    }
    printf("[vectorAdd] %zu of %zu elements bit-identical to the host loop\n", n, n);
    return pass();
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

struct DevArrays {
    void *x = nullptr, *y = nullptr, *d = nullptr, *ws = nullptr, *out = nullptr;
    void release()
    {
        b200v_free(x); b200v_free(y); b200v_free(d); b200v_free(ws); b200v_free(out);
        x = y = d = ws = out = nullptr;
    }
};

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
// bench: STREAM sweep (BASELINE.json configs[2]) + the OpenMP baseline in the same run

int mode_bench(const Args& a)
{
    printf("[gpu-validate] STREAM sweep: copy/scale/add/triad, fp32 + bf16, per-array 2^%d..2^%d bytes\n", a.min_log2_bytes,
           a.max_log2_bytes);
    int count = 0;
    int st = print_devices(&count, false);
    if (st != B200V_OK) return fail_status(st, "device discovery");
    if (a.device >= count) return fail(EX_NO_DEVICE, "device %d requested, %d visible", a.device, count);
    TRY(b200v_set_device(a.device));
    TRY(b200v_require_sm100(a.device));
    b200v_device_info info;   // queried again now that a.device is current: free memory is filled in
    memset(&info, 0, sizeof info);
    info.struct_size = sizeof info;
    TRY(b200v_device_info_get(a.device, &info));

    int max_log2 = a.max_log2_bytes;
    while (max_log2 > a.min_log2_bytes && 3.2 * (double)(1ull << max_log2) > (double)info.hbm_free_bytes) --max_log2;
    if (max_log2 != a.max_log2_bytes)
        printf("[bench] largest array limited to 2^%d bytes by free HBM\n", max_log2);

    FILE* js = a.json_path.empty() ? nullptr : fopen(a.json_path.c_str(), "w");
    if (js) fprintf(js, "{\"gpu\": \"%s\", ", info.name);
    bool first_row = true, all_ok = true;

    DevArrays v;
    TRY(alloc_arrays(v, (size_t)1 << max_log2));

    // same-box reference points: the driver's device-to-device copy and memset at 1 GiB (or the largest size run)
    double copy_gbs = a.measured_copy_gbs, memset_gbs = 0;
    {
        const size_t rb = (size_t)1 << std::min(max_log2, 30);
        void *e0 = nullptr, *e1 = nullptr;
        TRY(b200v_event_create(&e0));
        TRY(b200v_event_create(&e1));
        double best[2] = {1e30, 1e30};
        for (int which = 0; which < 2; ++which)
            for (int it = 0; it < a.warmup + 10; ++it) {
                TRY(b200v_event_record(e0, nullptr));
                if (which == 0) TRY(b200v_memcpy_d2d(v.d, v.x, rb, nullptr));
                else TRY(b200v_memset(v.d, 1, rb, nullptr));
                TRY(b200v_event_record(e1, nullptr));
                TRY(b200v_event_sync(e1));
                float ms = 0;
                TRY(b200v_event_elapsed_ms(e0, e1, &ms));
                if (it >= a.warmup) best[which] = std::min(best[which], (double)ms);
            }
        TRY(b200v_event_destroy(e0));
        TRY(b200v_event_destroy(e1));
        const double d2d = 2.0 * rb / best[0] * 1e-6;
        memset_gbs = (double)rb / best[1] * 1e-6;
        printf("[bench] same-box reference points at %zu MiB: cudaMemcpy D2D %.1f GB/s (read+write), cudaMemset %.1f GB/s\n",
               rb >> 20, d2d, memset_gbs);
        if (copy_gbs <= 0) copy_gbs = d2d;   // %copy column: against this run's own D2D copy unless a figure was passed in
        if (js) fprintf(js, "\"memcpy_d2d_gbs\": %.1f, \"memset_gbs\": %.1f, \"copy_denominator_gbs\": %.1f, \"rows_follow\": true, \"rows\": [\n", d2d, memset_gbs, copy_gbs);
    }
    printf("%-6s %-5s %12s %10s %10s %8s %10s", "op", "dtype", "bytes/array", "best ms", "GB/s", "%8TB/s", "b2b GB/s");
    printf(" %8s", "%copy");
    printf("  verify  note\n");

    const int ops[] = {B200V_OP_COPY, B200V_OP_SCALE, B200V_OP_ADD, B200V_OP_TRIAD};
    for (int dtype : {B200V_F32, B200V_BF16}) {
        for (int lg = a.min_log2_bytes; lg <= max_log2; ++lg) {
            const size_t bytes = (size_t)1 << lg;
            const size_t n = bytes / esize(dtype);
            TRY(b200v_fill(dtype, v.x, 0, n, a.seed, 0, B200V_GEN_UNIT, nullptr));
            TRY(b200v_fill(dtype, v.y, 0, n, a.seed, 1, B200V_GEN_UNIT, nullptr));
            for (int op : ops) {
                const bool two = op == B200V_OP_ADD || op == B200V_OP_TRIAD;
                b200v_timing t;
                memset(&t, 0, sizeof t);
                t.struct_size = sizeof t;
                TRY(b200v_time_op(op, dtype, v.d, v.x, two ? v.y : nullptr, 3.0f, n, a.warmup, a.iters, nullptr, &t));
                // verify the last launch's output: full checksum up to 1 GiB arrays, head/tail windows above
                bool ok = true;
                const char* how = "full";
                if (lg <= 30) {
                    uint64_t got[2], want[2];
                    TRY(device_checksum(v, v.d, bytes, 0, got));
                    hostcheck::expected_checksum(op, dtype, 0, n, a.seed, hostcheck::UNIT, 3.0f, 0, want);
                    ok = got[0] == want[0] && got[1] == want[1];
                } else {
                    how = "windows";
                    const size_t wn = (size_t)1 << 22;
                    for (size_t off : {(size_t)0, n / 2 + 1024, n - wn}) {
                        off -= off % 2;
                        uint64_t got[2], want[2];
                        TRY(device_checksum(v, (const char*)v.d + off * esize(dtype), wn * esize(dtype), 0, got));
                        hostcheck::expected_checksum(op, dtype, off, wn, a.seed, hostcheck::UNIT, 3.0f, 0, want);
                        ok = ok && got[0] == want[0] && got[1] == want[1];
                    }
                }
                all_ok = all_ok && ok;
                const double gbs = (double)t.bytes_per_launch / t.best_ms * 1e-6;
                const size_t footprint = bytes * (two ? 3 : 2);
                const bool resident = footprint <= info.l2_bytes;
                // "GB/s" = fastest single launch (STREAM convention); "b2b" = the same launches back to back under
                // one event pair, which hides the per-launch latency that dominates the small rows
                printf("%-6s %-5s %12zu %10.4f %10.1f %7.1f%% %10.1f", op_name(op), dtype == B200V_F32 ? "f32" : "bf16", bytes,
                       t.best_ms, gbs, 100.0 * gbs / kNominalGBs, (double)t.bytes_per_launch / t.mean_ms * 1e-6);
                printf(" %7.1f%%", 100.0 * gbs / copy_gbs);
                printf("  %-7s %s%s\n", ok ? how : "FAILED", resident ? "cache-resident (not an HBM number)" : "",
                       !resident && footprint <= 2 * info.l2_bytes ? "partly L2-resident" : "");
                if (js) {
                    fprintf(js,
                            "%s{\"op\": \"%s\", \"dtype\": \"%s\", \"bytes_per_array\": %zu, \"best_ms\": %.5f, \"median_ms\": %.5f, "
                            "\"mean_ms\": %.5f, \"gbs\": %.1f, \"frac_nominal\": %.4f, \"cache_resident\": %s, \"verified\": %s}",
                            first_row ? "" : ",\n", op_name(op), dtype == B200V_F32 ? "f32" : "bf16", bytes, t.best_ms, t.median_ms,
                            t.mean_ms, gbs, gbs / kNominalGBs, resident ? "true" : "false", ok ? "true" : "false");
                    first_row = false;
                }
                fflush(stdout);
            }
        }
    }
    v.release();

    // OpenMP baseline on this box's host cores, same loops, same run (a reported baseline, not the target)
    {
        const size_t n = (size_t)1 << 26;
        std::vector<float> hx(n), hy(n), hd(n);
        hostcheck::fill(hostcheck::F32, hx.data(), 0, n, a.seed, 0, hostcheck::UNIT);
        hostcheck::fill(hostcheck::F32, hy.data(), 0, n, a.seed, 1, hostcheck::UNIT);
        printf("[cpu baseline] OpenMP host loops, %d threads, N=2^26 fp32 per array\n", hostcheck::max_threads());
        if (js) fprintf(js, "\n], \"cpu_threads\": %d, \"cpu_rows\": [\n", hostcheck::max_threads());
        bool firstc = true;
        for (int op : ops) {
            hostcheck::apply(op, hostcheck::F32, hd.data(), hx.data(), hy.data(), 3.0f, n, true);
            double best = 1e30;
            for (int i = 0; i < 5; ++i) {
                const double t0 = now_s();
                hostcheck::apply(op, hostcheck::F32, hd.data(), hx.data(), hy.data(), 3.0f, n, true);
                best = std::min(best, now_s() - t0);
            }
            const double gbs = (double)b200v_op_bytes(op, B200V_F32, n) / best * 1e-9;
            printf("  %-6s f32 %10.3f ms %8.1f GB/s\n", op_name(op), best * 1e3, gbs);
            if (js) {
                fprintf(js, "%s{\"op\": \"%s\", \"gbs\": %.2f}", firstc ? "" : ",\n", op_name(op), gbs);
                firstc = false;
            }
        }
    }
    if (js) {
        fprintf(js, "\n]}\n");
        fclose(js);
    }
    if (!all_ok) return fail(EX_MISMATCH, "at least one sweep row failed verification");
    return pass();
}

// ----------------------------------------------------------------------------------------------
// shard child: one process, one GPU, one contiguous shard of the global vector

struct ShardResult {
    int rank = -1;
    int ok = 0;
    double best_ms = 0, mean_ms = 0;
    uint64_t count = 0;
    long passes = 0;
    int exit_code = EX_CUDA;
    std::string text;
    // NVML readings over the run (burn-in): -1 = not available
    int sm_min_mhz = -1, temp_max_c = -1;
    double power_max_w = -1;
    unsigned long long reasons = 0;
};

void child_say(const Args& a, const char* fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    int len = vsnprintf(buf, sizeof buf - 1, fmt, ap);
    va_end(ap);
    if (len < 0) return;
    len = std::min(len, (int)sizeof buf - 2);
    buf[len++] = '\n';
    if (a.out_fd >= 0) {
        ssize_t r = write(a.out_fd, buf, (size_t)len);
        (void)r;
    } else {
        fwrite(buf, 1, (size_t)len, stdout);
        fflush(stdout);
    }
}

int child_fail(const Args& a, int status, const char* what)
{
    child_say(a, "ERROR %d %d %s: %s: %s", a.rank, exit_code_for(status), what, b200v_status_name(status), b200v_last_error());
    return exit_code_for(status);
}

#define CTRY(call)                                            \
    do {                                                      \
        int st__ = (call);                                    \
        if (st__ != B200V_OK) return child_fail(a, st__, #call); \
    } while (0)

// scalars rotated between burn-in passes so a stale output can never pass verification
const float kScalars[4] = {3.0f, -0.75f, 1.0009765625f, 2.5f};

int mode_shard_child(const Args& a)
{
    const bool burnin = a.mode == "burnin-child";
    const int dtype = dtype_id(a.dtype);
    const int op = a.op.empty() ? B200V_OP_TRIAD : op_id(a.op);
    CTRY(b200v_set_device(a.device));
    CTRY(b200v_require_sm100(a.device));
    const uint64_t total = a.log2n >= 0 ? (1ull << a.log2n) : a.n;
    uint64_t begin = 0, count = 0;
    CTRY(b200v_shard_range(total, a.rank, a.world, &begin, &count));
    if (dtype == B200V_BF16 && (begin & 1)) {   // keep bf16 shards on 32-bit word boundaries
        begin += 1;
        count -= count ? 1 : 0;
    }
    const size_t bytes = (size_t)count * esize(dtype);
    DevArrays v;
    CTRY(alloc_arrays(v, bytes));
    CTRY(b200v_fill(dtype, v.x, begin, count, a.seed, 0, B200V_GEN_UNIT, nullptr));
    CTRY(b200v_fill(dtype, v.y, begin, count, a.seed, 1, B200V_GEN_UNIT, nullptr));
    const bool two = op == B200V_OP_ADD || op == B200V_OP_TRIAD;
    for (int i = 0; i < a.warmup; ++i) CTRY(b200v_stream_op(op, dtype, v.d, v.x, two ? v.y : nullptr, kScalars[0], count, nullptr));
    CTRY(b200v_device_sync());

    // expected checksums from the host loop (word indices are shard-local)
    const int nscal = burnin ? 4 : 1;
    uint64_t want[4][2];
    for (int k = 0; k < nscal; ++k)
        hostcheck::expected_checksum(op, dtype, begin, count, a.seed, hostcheck::UNIT, kScalars[k], 0, want[k]);

    // start barrier: tell the parent we are ready, wait for its go byte
    child_say(a, "READY %d", a.rank);
    if (a.go_fd >= 0) {
        char c;
        ssize_t r;
        do r = read(a.go_fd, &c, 1); while (r < 0 && errno == EINTR);
        if (r != 1) return child_fail(a, B200V_ERR_INVALID, "start barrier closed");
    }

    int ok = 1;
    double best_ms = 0, mean_ms = 0;
    long passes = 0;
    int nv_min_mhz = -1, nv_max_temp = -1;
    double nv_max_power = -1;
    unsigned long long nv_reasons = 0;
    if (!burnin) {
        b200v_timing t;
        memset(&t, 0, sizeof t);
        t.struct_size = sizeof t;
        CTRY(b200v_time_op(op, dtype, v.d, v.x, two ? v.y : nullptr, kScalars[0], count, 0, a.iters, nullptr, &t));
        best_ms = t.best_ms;
        mean_ms = t.mean_ms;
        passes = 2L * a.iters;
        uint64_t got[2];
        CTRY(device_checksum(v, v.d, bytes, 0, got));
        ok = got[0] == want[0][0] && got[1] == want[0][1];
    } else {
        void *e0 = nullptr, *e1 = nullptr;
        CTRY(b200v_event_create(&e0));
        CTRY(b200v_event_create(&e1));
        const double t_end = now_s() + a.seconds;
        double sum_ms = 0;
        best_ms = 1e30;
        b200v_device_info dinfo;
        memset(&dinfo, 0, sizeof dinfo);
        dinfo.struct_size = sizeof dinfo;
        const bool have_nvml = b200v_device_info_get(a.device, &dinfo) == B200V_OK && nvmlprobe::available();
        double next_probe = 0;
        while (now_s() < t_end) {
            if (have_nvml && now_s() >= next_probe) {   // a few readings per second while the GPU is loaded
                nvmlprobe::Sample nv;
                if (nvmlprobe::sample(dinfo.pci_bus_id, &nv)) {
                    if (passes > 0) {                   // skip the idle reading before the first pass
                        nv_min_mhz = nv_min_mhz < 0 ? (int)nv.sm_mhz : std::min(nv_min_mhz, (int)nv.sm_mhz);
                        nv_reasons |= nv.reasons;
                    }
                    nv_max_temp = std::max(nv_max_temp, (int)nv.temp_c);
                    nv_max_power = std::max(nv_max_power, nv.power_w);
                }
                next_probe = now_s() + 0.2;
            }
            const int k = (int)(passes % 4);
            CTRY(b200v_event_record(e0, nullptr));
            CTRY(b200v_stream_op(op, dtype, v.d, v.x, two ? v.y : nullptr, kScalars[k], count, nullptr));
            CTRY(b200v_event_record(e1, nullptr));
            uint64_t got[2];
            CTRY(device_checksum(v, v.d, bytes, 0, got));
            float ms = 0;
            CTRY(b200v_event_elapsed_ms(e0, e1, &ms));
            best_ms = std::min(best_ms, (double)ms);
            sum_ms += ms;
            ++passes;
            if (got[0] != want[k][0] || got[1] != want[k][1]) {
                ok = 0;
                child_say(a, "NOTE %d pass %ld failed verification", a.rank, passes);
                break;
            }
        }
        mean_ms = passes ? sum_ms / passes : 0;
        b200v_event_destroy(e0);
        b200v_event_destroy(e1);
    }
    if (a.inject_fault && a.rank == a.world - 1) ok = 0;
    v.release();
    if (nv_max_temp >= 0) child_say(a, "NVML %d %d %d %.1f %llu", a.rank, nv_min_mhz, nv_max_temp, nv_max_power, nv_reasons);
    child_say(a, "RESULT %d %d %.6f %.6f %" PRIu64 " %ld", a.rank, ok, best_ms, mean_ms, count, passes);
    return ok ? EX_PASS : EX_MISMATCH;
}

// ----------------------------------------------------------------------------------------------
// node / burnin parent: fork+exec one child per GPU *before* touching CUDA in this process.

int probe_device_count()
{
    // a throw-away child asks the runtime, so this process never holds a CUDA context across fork()
    int fds[2];
    if (pipe(fds) != 0) return -1;
    pid_t pid = fork();
    if (pid < 0) return -1;
    if (pid == 0) {
        close(fds[0]);
        int n = 0;
        if (b200v_device_count(&n) != B200V_OK) n = 0;
        ssize_t r = write(fds[1], &n, sizeof n);
        (void)r;
        _exit(0);
    }
    close(fds[1]);
    int n = 0;
    ssize_t r = read(fds[0], &n, sizeof n);
    close(fds[0]);
    int wst = 0;
    waitpid(pid, &wst, 0);
    return r == (ssize_t)sizeof n ? n : -1;
}

struct Child {
    pid_t pid = -1;
    int go_w = -1, out_r = -1;
    std::string pending;
    bool ready = false, done = false;
    ShardResult res;
};

int run_children(const Args& a, const char* child_mode, int world, std::vector<ShardResult>& results, double timeout_s)
{
    char self[4096];
    ssize_t len = readlink("/proc/self/exe", self, sizeof self - 1);
    if (len <= 0) return fail(EX_CUDA, "cannot resolve /proc/self/exe");
    self[len] = 0;

    std::vector<Child> kids((size_t)world);
    const int threads = std::max(1, hostcheck::max_threads() / world);
    for (int r = 0; r < world; ++r) {
        int go[2], out[2];
        if (pipe(go) != 0 || pipe(out) != 0) return fail(EX_CUDA, "pipe: %s", strerror(errno));
        pid_t pid = fork();
        if (pid < 0) return fail(EX_CUDA, "fork: %s", strerror(errno));
        if (pid == 0) {
            close(go[1]);
            close(out[0]);
            for (int q = 0; q < r; ++q) {
                close(kids[q].go_w);
                close(kids[q].out_r);
            }
            char b[16][64];
            snprintf(b[0], 64, "%d", r);
            snprintf(b[1], 64, "%d", world);
            snprintf(b[2], 64, "%d", r);   // device = rank among the GPUs this container was granted
            snprintf(b[3], 64, "%d", go[0]);
            snprintf(b[4], 64, "%d", out[1]);
            snprintf(b[5], 64, "%d", a.log2n);
            snprintf(b[6], 64, "%" PRIu64, a.n);
            snprintf(b[7], 64, "%d", a.iters);
            snprintf(b[8], 64, "%d", a.warmup);
            snprintf(b[9], 64, "%.3f", a.seconds);
            snprintf(b[10], 64, "%u", a.seed);
            snprintf(b[11], 64, "%d", threads);
            setenv("OMP_NUM_THREADS", b[11], 1);
            std::vector<const char*> av = {self, "--mode", child_mode, "--rank", b[0], "--world", b[1], "--device", b[2],
                                           "--go-fd", b[3], "--out-fd", b[4], "--log2n", b[5], "--n", b[6], "--iters", b[7],
                                           "--warmup", b[8], "--seconds", b[9], "--seed", b[10], "--dtype", a.dtype.c_str()};
            if (!a.op.empty()) { av.push_back("--op"); av.push_back(a.op.c_str()); }
            if (a.inject_fault) av.push_back("--inject-fault");
            av.push_back(nullptr);
            execv(self, const_cast<char* const*>(av.data()));
            _exit(127);
        }
        close(go[0]);
        close(out[1]);
        kids[r].pid = pid;
        kids[r].go_w = go[1];
        kids[r].out_r = out[0];
        kids[r].res.rank = r;
    }

    // read READY / RESULT / ERROR lines; release every child at once when all are ready
    bool released = false;
    const double deadline = now_s() + timeout_s;
    int open_pipes = world;
    while (open_pipes > 0) {
        std::vector<pollfd> pf;
        for (auto& k : kids)
            if (k.out_r >= 0) pf.push_back({k.out_r, POLLIN, 0});
        const double left = deadline - now_s();
        if (left <= 0) break;
        int pr = poll(pf.data(), (nfds_t)pf.size(), (int)std::min(left * 1e3, 1000.0));
        if (pr < 0 && errno != EINTR) break;
        for (auto& p : pf) {
            if (!(p.revents & (POLLIN | POLLHUP))) continue;
            Child* k = nullptr;
            for (auto& c : kids)
                if (c.out_r == p.fd) k = &c;
            char buf[1024];
            ssize_t r = read(p.fd, buf, sizeof buf);
            if (r <= 0) {
                close(k->out_r);
                k->out_r = -1;
                --open_pipes;
                continue;
            }
            k->pending.append(buf, (size_t)r);
            size_t nl;
            while ((nl = k->pending.find('\n')) != std::string::npos) {
                std::string line = k->pending.substr(0, nl);
                k->pending.erase(0, nl + 1);
                if (line.rfind("READY", 0) == 0) {
                    k->ready = true;
                } else if (line.rfind("NVML", 0) == 0) {
                    ShardResult& q = k->res;
                    int rk = 0;
                    sscanf(line.c_str(), "NVML %d %d %d %lf %llu", &rk, &q.sm_min_mhz, &q.temp_max_c, &q.power_max_w, &q.reasons);
                } else if (line.rfind("RESULT", 0) == 0) {
                    ShardResult& s = k->res;
                    sscanf(line.c_str(), "RESULT %d %d %lf %lf %" SCNu64 " %ld", &s.rank, &s.ok, &s.best_ms, &s.mean_ms, &s.count,
                           &s.passes);
                    k->done = true;
                } else {
                    k->res.text += line + "\n";
                }
            }
        }
        if (!released) {
            bool all = true, any_dead = false;
            for (auto& k : kids) {
                all = all && k.ready;
                any_dead = any_dead || (k.out_r < 0 && !k.ready);
            }
            if (all || any_dead) {
                for (auto& k : kids) {
                    if (all) {
                        ssize_t w = write(k.go_w, "g", 1);
                        (void)w;
                    }
                    close(k.go_w);   // closing without a byte tells a waiting child to give up
                    k.go_w = -1;
                }
                released = true;
            }
        }
    }
    for (auto& k : kids) {
        if (k.go_w >= 0) close(k.go_w);
        if (k.out_r >= 0) close(k.out_r);
        int wst = 0;
        if (now_s() >= deadline) kill(k.pid, SIGKILL);
        waitpid(k.pid, &wst, 0);
        k.res.exit_code = WIFEXITED(wst) ? WEXITSTATUS(wst) : EX_CUDA;
        results.push_back(k.res);
    }
    return EX_PASS;
}

int mode_node(const Args& a, bool burnin, bool print_verdict)
{
    const int visible = probe_device_count();
    printf("[gpu-validate] %s: one process per GPU, %s\n", burnin ? "burn-in" : "sharded run",
           burnin ? "every pass verified" : "device-timed, max over ranks");
    if (visible <= 0) return fail(EX_NO_DEVICE, "no CUDA device is visible to this container");
    const int world = a.gpus > 0 ? a.gpus : visible;
    if (world > visible) return fail(EX_NO_DEVICE, "%d GPUs requested, %d visible", world, visible);
    const int dtype = dtype_id(a.dtype);
    const int op = a.op.empty() ? B200V_OP_TRIAD : op_id(a.op);
    Args b = a;
    if (b.log2n < 0 && !b.n_set) b.log2n = burnin ? 30 : 32;   // configs[3]: 2^32 fp32 elements in total
    const uint64_t total = b.log2n >= 0 ? (1ull << b.log2n) : b.n;
    printf("[node] %s %s, %" PRIu64 " elements in total over %d GPU%s (%d visible)\n", op_name(op), a.dtype.c_str(), total, world,
           world == 1 ? "" : "s", visible);
    fflush(stdout);

    std::vector<ShardResult> res;
    int rc = run_children(b, burnin ? "burnin-child" : "shard-child", world, res, 600.0 + 4 * a.seconds);
    if (rc != EX_PASS) return rc;

    bool ok = true, throttled = false;
    double max_best = 0, max_mean = 0;
    uint64_t elems = 0;
    int worst_exit = EX_PASS;
    for (auto& s : res) {
        const bool good = s.exit_code == EX_PASS && s.ok == 1;
        ok = ok && good;
        if (s.exit_code != EX_PASS && worst_exit == EX_PASS) worst_exit = s.exit_code;
        const double bytes = (double)b200v_op_bytes(op, dtype, (size_t)s.count);
        printf("  rank %d (GPU %d): %" PRIu64 " elements, best %.4f ms (%.1f GB/s), mean %.4f ms, %ld passes, %s\n", s.rank, s.rank,
               s.count, s.best_ms, s.best_ms > 0 ? bytes / s.best_ms * 1e-6 : 0.0, s.mean_ms, s.passes,
               good ? "verified" : "FAILED");
        if (s.temp_max_c >= 0) {
            const bool bad = (s.reasons & nvmlprobe::kBadReasons) != 0;
            printf("          under load: SM clock down to %d MHz, up to %d C, up to %.0f W, clock limits seen: %s%s\n", s.sm_min_mhz,
                   s.temp_max_c, s.power_max_w, nvmlprobe::reasons_text(s.reasons).c_str(),
                   bad ? "  <-- WARNING: thermal or hardware slowdown" : "");
            throttled = throttled || bad;
        }
        if (!s.text.empty()) printf("%s", s.text.c_str());
        max_best = std::max(max_best, s.best_ms);
        max_mean = std::max(max_mean, s.mean_ms);
        elems += s.count;
    }
    double node_gbs = 0;
    if (max_best > 0) {
        const double bytes = (double)b200v_op_bytes(op, dtype, (size_t)elems);
        node_gbs = bytes / max_best * 1e-6;
        printf("[node] %d GPU%s: %.1f GB/s (bytes / max over ranks of best ms = %.4f), %.1f GB/s sustained (max mean ms = %.4f)\n",
               world, world == 1 ? "" : "s", node_gbs, max_best, bytes / max_mean * 1e-6, max_mean);
        printf("[node] per GPU %.1f GB/s = %.1f%% of the %.0f GB/s nominal HBM peak\n", node_gbs / world,
               100.0 * node_gbs / world / kNominalGBs, kNominalGBs);
    }
    if (!a.json_path.empty()) {
        if (FILE* js = fopen(a.json_path.c_str(), "w")) {
            fprintf(js, "{\"mode\": \"%s\", \"op\": \"%s\", \"dtype\": \"%s\", \"gpus\": %d, \"elements\": %" PRIu64
                        ", \"max_best_ms\": %.6f, \"max_mean_ms\": %.6f, \"node_gbs\": %.1f, \"ok\": %s, \"ranks\": [",
                    burnin ? "burnin" : "node", op_name(op), a.dtype.c_str(), world, elems, max_best, max_mean, node_gbs,
                    ok ? "true" : "false");
            for (size_t i = 0; i < res.size(); ++i)
                fprintf(js, "%s{\"rank\": %d, \"ok\": %d, \"best_ms\": %.6f, \"mean_ms\": %.6f, \"elements\": %" PRIu64 ", \"passes\": %ld, "
                            "\"sm_min_mhz\": %d, \"temp_max_c\": %d, \"power_max_w\": %.1f, \"clock_limit_reasons\": \"%s\"}",
                        i ? ", " : "", res[i].rank, res[i].ok, res[i].best_ms, res[i].mean_ms, res[i].count, res[i].passes,
                        res[i].sm_min_mhz, res[i].temp_max_c, res[i].power_max_w,
                        res[i].temp_max_c >= 0 ? nvmlprobe::reasons_text(res[i].reasons).c_str() : "unavailable");
            fprintf(js, "]}\n");
            fclose(js);
        }
    }
    if (!ok) return fail(worst_exit == EX_PASS ? EX_MISMATCH : worst_exit, "at least one rank failed");
    if (throttled && a.fail_on_throttle)
        return fail(EX_MISMATCH, "a GPU reported thermal or hardware slowdown under load (--fail-on-throttle)");
    return print_verdict ? pass() : EX_PASS;
}

// ----------------------------------------------------------------------------------------------
// p2p: all ordered pairs, copy engine and direct peer loads, both verified on the destination

int mode_p2p(const Args& a, bool print_verdict)
{
    printf("[gpu-validate] NVLink peer-to-peer check: all ordered pairs, %.0f MiB per transfer\n", a.p2p_bytes / 1048576.0);
    int count = 0;
    int st = print_devices(&count, true);
    if (st != B200V_OK) return fail_status(st, "device discovery");
    const int g = a.gpus > 0 ? std::min(a.gpus, count) : count;
    if (g < 2) {
        printf("[p2p] only %d GPU visible: nothing to check\n", g);
        return print_verdict ? pass() : EX_PASS;
    }
    const size_t bytes = (size_t)a.p2p_bytes & ~(size_t)31;
    const size_t n = bytes / 4;
    std::vector<void*> src((size_t)g), dst((size_t)g);
    std::vector<DevArrays> scratch((size_t)g);
    std::vector<uint64_t> want((size_t)g * 2);
    for (int d = 0; d < g; ++d) {
        TRY(b200v_set_device(d));
        TRY(b200v_require_sm100(d));
        TRY(b200v_malloc(&src[d], bytes));
        TRY(b200v_malloc(&dst[d], bytes));
        TRY(b200v_malloc(&scratch[d].ws, b200v_checksum_workspace_bytes()));
        TRY(b200v_malloc(&scratch[d].out, 16));
        TRY(b200v_fill(B200V_F32, src[d], 0, n, a.seed, 100 + d, B200V_GEN_BITS, nullptr));   // each source has its own stream id
        TRY(b200v_device_sync());
        hostcheck::generator_checksum(hostcheck::F32, 0, n, a.seed, 100 + d, hostcheck::BITS, 0, &want[2 * d]);
    }
    for (int i = 0; i < g; ++i)
        for (int j = 0; j < g; ++j)
            if (i != j) TRY(b200v_peer_enable(i, j));

    std::vector<double> copy_gbs, pull_gbs;
    bool ok = true;
    printf("  %-8s %14s %14s  verify\n", "src->dst", "memcpyPeer GB/s", "peer-load GB/s");
    for (int i = 0; i < g; ++i) {
        for (int j = 0; j < g; ++j) {
            if (i == j) continue;
            double best[2] = {1e30, 1e30};
            bool pair_ok = true;
            for (int method = 0; method < 2; ++method) {
                // method 0: copy engine, issued on the source device; method 1: kernel on the destination pulls
                TRY(b200v_set_device(method == 0 ? i : j));
                void *e0 = nullptr, *e1 = nullptr;
                TRY(b200v_event_create(&e0));
                TRY(b200v_event_create(&e1));
                for (int it = 0; it < a.warmup + std::max(a.iters / 4, 3); ++it) {
                    TRY(b200v_event_record(e0, nullptr));
                    if (method == 0) TRY(b200v_peer_copy(dst[j], j, src[i], i, bytes, nullptr));
                    else TRY(b200v_peer_pull(dst[j], src[i], bytes, nullptr));
                    TRY(b200v_event_record(e1, nullptr));
                    TRY(b200v_event_sync(e1));
                    float ms = 0;
                    TRY(b200v_event_elapsed_ms(e0, e1, &ms));
                    if (it >= a.warmup) best[method] = std::min(best[method], (double)ms);
                }
                TRY(b200v_event_destroy(e0));
                TRY(b200v_event_destroy(e1));
                // verify on the destination device, then scrub the destination for the next method
                TRY(b200v_set_device(j));
                uint64_t got[2];
                TRY(device_checksum(scratch[j], dst[j], bytes, 0, got));
                bool good = got[0] == want[2 * i] && got[1] == want[2 * i + 1];
                if (a.inject_fault && i == g - 1 && j == 0 && method == 1) good = false;
                pair_ok = pair_ok && good;
                TRY(b200v_memset(dst[j], 0xEE, bytes, nullptr));
                TRY(b200v_device_sync());
            }
            const double c = bytes / best[0] * 1e-6, p = bytes / best[1] * 1e-6;
            copy_gbs.push_back(c);
            pull_gbs.push_back(p);
            ok = ok && pair_ok;
            printf("  %d -> %d   %14.1f %14.1f  %s\n", i, j, c, p, pair_ok ? "ok" : "FAILED");
            fflush(stdout);
        }
    }
    // All GPUs at once: in round k every GPU j pulls from GPU (j+k) mod g, so every NVLink port is
    // busy in both directions and the switch carries g concurrent flows.  Kernels on different GPUs
    // are independent (nobody waits on anybody); time = max over GPUs of the device-timed pull.
    std::vector<double> ring_gbs;
    {
        std::vector<void*> e0((size_t)g), e1((size_t)g);
        for (int j = 0; j < g; ++j) {
            TRY(b200v_set_device(j));
            TRY(b200v_event_create(&e0[j]));
            TRY(b200v_event_create(&e1[j]));
        }
        printf("  concurrent rounds (every GPU pulls from its k-th neighbour at the same time):\n");
        for (int k = 1; k < g; ++k) {
            double worst = 0;
            bool round_ok = true;
            for (int rep = 0; rep < 3; ++rep) {
                for (int j = 0; j < g; ++j) {
                    TRY(b200v_set_device(j));
                    TRY(b200v_event_record(e0[j], nullptr));
                    TRY(b200v_peer_pull(dst[j], src[(j + k) % g], bytes, nullptr));
                    TRY(b200v_event_record(e1[j], nullptr));
                }
                double slowest = 0;
                for (int j = 0; j < g; ++j) {
                    TRY(b200v_set_device(j));
                    TRY(b200v_event_sync(e1[j]));
                    float ms = 0;
                    TRY(b200v_event_elapsed_ms(e0[j], e1[j], &ms));
                    slowest = std::max(slowest, (double)ms);
                }
                if (rep == 0 || slowest < worst) worst = slowest;   // best of the repetitions, slowest GPU in each
            }
            for (int j = 0; j < g; ++j) {
                TRY(b200v_set_device(j));
                uint64_t got[2];
                TRY(device_checksum(scratch[j], dst[j], bytes, 0, got));
                const int i = (j + k) % g;
                round_ok = round_ok && got[0] == want[2 * i] && got[1] == want[2 * i + 1];
                TRY(b200v_memset(dst[j], 0xEE, bytes, nullptr));
                TRY(b200v_device_sync());
            }
            const double per_gpu = bytes / worst * 1e-6;
            ring_gbs.push_back(per_gpu);
            ok = ok && round_ok;
            printf("    shift %d: %.1f GB/s into every GPU at once (%.1f GB/s through the switch)  %s\n", k, per_gpu, per_gpu * g,
                   round_ok ? "ok" : "FAILED");
            fflush(stdout);
        }
        for (int j = 0; j < g; ++j) {
            b200v_set_device(j);
            b200v_event_destroy(e0[j]);
            b200v_event_destroy(e1[j]);
        }
    }
    auto summary = [&](const char* name, std::vector<double> v) {
        std::sort(v.begin(), v.end());
        printf("[p2p] %-10s min %.1f  median %.1f  max %.1f GB/s per direction (%.1f%% of the %.0f GB/s NVLink 5 figure at the median)\n",
               name, v.front(), v[v.size() / 2], v.back(), 100.0 * v[v.size() / 2] / kNvlinkGBs, kNvlinkGBs);
    };
    summary("memcpyPeer", copy_gbs);
    summary("peer-load", pull_gbs);
    if (!ring_gbs.empty()) summary("concurrent", ring_gbs);
    if (!a.json_path.empty()) {
        if (FILE* js = fopen((a.json_path + (a.mode == "full" ? ".p2p" : "")).c_str(), "w")) {
            fprintf(js, "{\"mode\": \"p2p\", \"gpus\": %d, \"bytes\": %zu, \"ok\": %s, \"pairs\": [", g, bytes, ok ? "true" : "false");
            size_t k = 0;
            for (int i = 0; i < g; ++i)
                for (int j = 0; j < g; ++j)
                    if (i != j) {
                        fprintf(js, "%s{\"src\": %d, \"dst\": %d, \"memcpy_peer_gbs\": %.1f, \"peer_load_gbs\": %.1f}", k ? ", " : "", i, j,
                                copy_gbs[k], pull_gbs[k]);
                        ++k;
                    }
            fprintf(js, "], \"concurrent_shift_gbs_per_gpu\": [");
            for (size_t q = 0; q < ring_gbs.size(); ++q) fprintf(js, "%s%.1f", q ? ", " : "", ring_gbs[q]);
            fprintf(js, "]}\n");
            fclose(js);
        }
    }
    for (int d = 0; d < g; ++d) {
        b200v_set_device(d);
        b200v_free(src[d]);
        b200v_free(dst[d]);
        scratch[d].release();
    }
    if (!ok) return fail(EX_MISMATCH, "peer-to-peer data verification failed");
    return print_verdict ? pass() : EX_PASS;
}

// ----------------------------------------------------------------------------------------------

void usage()
{
    fprintf(stderr,
            "usage: gpu-validate [--mode smoke|list|bench|node|p2p|burnin|full|host-reference]\n"
            "                    [--device D] [--gpus G] [--n N | --log2n K] [--dtype f32|bf16] [--op add|copy|scale|triad]\n"
            "                    [--iters I] [--warmup W] [--seconds S] [--seed X] [--p2p-bytes B]\n"
            "                    [--min-log2-bytes a --max-log2-bytes b] [--measured-copy-gbs G] [--json FILE] [--inject-fault] [--fail-on-throttle]\n"
            "no arguments: vectorAdd N=50000 on GPU 0, prints Test PASSED / Test FAILED, exit code 0 / non-zero\n");
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
    if (a.iters < 1 || a.warmup < 0 || a.log2n > 40 || a.min_log2_bytes < 4 || a.max_log2_bytes > 40 ||
        a.min_log2_bytes > a.max_log2_bytes || a.world < 1 || a.rank < 0 || a.rank >= a.world)
        return false;
    return true;
}

}  // namespace

int main(int argc, char** argv)
{
    setvbuf(stdout, nullptr, _IOLBF, 0);
    Args a;
    if (!parse(argc, argv, a)) {
        usage();
        printf("Test FAILED\n");
        return EX_USAGE_ERR;
    }
    if (a.mode == "smoke") return mode_smoke(a);
    if (a.mode == "list") return mode_list(a);
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

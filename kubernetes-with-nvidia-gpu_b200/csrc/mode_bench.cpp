// mode_bench.cpp — `gpu-validate --mode bench`: the STREAM sweep (BASELINE.json configs[2]) with the
// OpenMP host baseline timed in the same run.
#include "launcher_common.h"

namespace gpuvalidate {

// ----------------------------------------------------------------------------------------------
// bench: STREAM sweep (BASELINE.json configs[2]) + the OpenMP baseline in the same run

int mode_bench(const Args& a)
{
    printf("[gpu-validate] STREAM sweep: copy/scale/add/triad, fp32 + bf16, per-array 2^%d..2^%d bytes, s = 3.0, seed %u\n",
           a.min_log2_bytes, a.max_log2_bytes, a.seed);
    int count = 0;
    int st = print_devices(&count, false);
    if (st != B200V_OK) return fail_status(st, "device discovery");
    if (a.device >= count) return fail(EX_NO_DEVICE, "device %d requested, %d visible", a.device, count);
    const int device = a.device >= 0 ? a.device : 0;      // the sweep runs on one GPU: --device D, else GPU 0
    TRY(b200v_set_device(device));
    TRY(b200v_require_sm100(device));
    b200v_device_info info;   // queried again now that the device is current: free memory is filled in
    memset(&info, 0, sizeof info);
    info.struct_size = sizeof info;
    TRY(b200v_device_info_get(device, &info));

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

}  // namespace gpuvalidate

// mode_p2p.cpp — `gpu-validate --mode p2p`: the NVLink peer-to-peer check (BASELINE.json configs[4]).
#include "launcher_common.h"

namespace gpuvalidate {

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
    // All GPUs at once: in round k every GPU j receives from GPU (j+k) mod g, so every NVLink port is
    // busy in both directions and the switch carries g concurrent flows.  Two ways of moving the data:
    // "pull" = a kernel on the destination loads from the peer; "push" = cudaMemcpyPeerAsync issued on
    // the source's copy engine.  Kernels and copies on different GPUs are independent (nobody waits on
    // anybody); time = max over GPUs of the device-timed transfer.
    std::vector<double> ring_gbs, ring_push_gbs;
    {
        // one non-blocking stream per GPU: the legacy default stream would couple the devices' queues
        std::vector<void*> e0((size_t)g), e1((size_t)g), strm((size_t)g);
        for (int j = 0; j < g; ++j) {
            TRY(b200v_set_device(j));
            TRY(b200v_event_create(&e0[j]));
            TRY(b200v_event_create(&e1[j]));
            TRY(b200v_stream_create(&strm[j]));
        }
        printf("  concurrent rounds (every GPU receives from its k-th neighbour at the same time):\n");
        for (int k = 1; k < g; ++k) {
            double per_gpu[2] = {0, 0};
            bool round_ok = true;
            for (int method = 0; method < 2; ++method) {
                double best = 0;
                for (int rep = 0; rep < 3; ++rep) {
                    // owner = the GPU that issues and times the transfer (its events live there):
                    // the destination for a pull, the source for a push
                    for (int owner = 0; owner < g; ++owner) {
                        const int j = method == 0 ? owner : (owner - k + g) % g;   // destination
                        const int i = (j + k) % g;                                 // source
                        TRY(b200v_set_device(owner));
                        TRY(b200v_event_record(e0[owner], strm[owner]));
                        if (method == 0) TRY(b200v_peer_pull(dst[j], src[i], bytes, strm[owner]));
                        else TRY(b200v_peer_copy(dst[j], j, src[i], i, bytes, strm[owner]));
                        TRY(b200v_event_record(e1[owner], strm[owner]));
                    }
                    double slowest = 0;
                    for (int owner = 0; owner < g; ++owner) {
                        TRY(b200v_set_device(owner));
                        TRY(b200v_event_sync(e1[owner]));
                        float ms = 0;
                        TRY(b200v_event_elapsed_ms(e0[owner], e1[owner], &ms));
                        slowest = std::max(slowest, (double)ms);
                    }
                    if (rep == 0 || slowest < best) best = slowest;   // best repetition, slowest GPU within it
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
                per_gpu[method] = bytes / best * 1e-6;
            }
            ring_gbs.push_back(per_gpu[0]);
            ring_push_gbs.push_back(per_gpu[1]);
            ok = ok && round_ok;
            printf("    shift %d: pull %.1f, push %.1f GB/s into every GPU at once (%.1f / %.1f GB/s through the switch)  %s\n", k,
                   per_gpu[0], per_gpu[1], per_gpu[0] * g, per_gpu[1] * g, round_ok ? "ok" : "FAILED");
            fflush(stdout);
        }
        for (int j = 0; j < g; ++j) {
            b200v_set_device(j);
            b200v_event_destroy(e0[j]);
            b200v_event_destroy(e1[j]);
            b200v_stream_destroy(strm[j]);
        }
    }
    auto summary = [&](const char* name, std::vector<double> v) {
        std::sort(v.begin(), v.end());
        printf("[p2p] %-10s min %.1f  median %.1f  max %.1f GB/s per direction (%.1f%% of the %.0f GB/s NVLink 5 figure at the median)\n",
               name, v.front(), v[v.size() / 2], v.back(), 100.0 * v[v.size() / 2] / kNvlinkGBs, kNvlinkGBs);
    };
    summary("memcpyPeer", copy_gbs);
    summary("peer-load", pull_gbs);
    if (!ring_gbs.empty()) summary("all pull", ring_gbs);
    if (!ring_push_gbs.empty()) summary("all push", ring_push_gbs);
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
            fprintf(js, "], \"concurrent_shift_push_gbs_per_gpu\": [");
            for (size_t q = 0; q < ring_push_gbs.size(); ++q) fprintf(js, "%s%.1f", q ? ", " : "", ring_push_gbs[q]);
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

}  // namespace gpuvalidate

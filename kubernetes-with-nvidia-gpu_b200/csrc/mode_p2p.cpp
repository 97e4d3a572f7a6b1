// mode_p2p.cpp — `gpu-validate --mode p2p`: the NVLink peer-to-peer check (BASELINE.json configs[4]).
#include "launcher_common.h"

namespace gpuvalidate {

namespace {

// Three ways of moving a buffer from GPU i to GPU j, each verified on the destination:
//   COPY  cudaMemcpyPeerAsync, issued on the source (the copy engine pushes)
//   PULL  a kernel on the destination loading from the source's memory (b200v_peer_pull)
//   PUSH  a kernel on the source storing into the destination's memory (b200v_peer_push)
enum Method { COPY = 0, PULL = 1, PUSH = 2, kMethods = 3 };
const char* const kMethodKey[kMethods] = {"memcpy_peer", "peer_load", "peer_store"};

// the GPU that issues (and times) the transfer i -> j
int owner_of(int method, int i, int j) { return method == PULL ? j : i; }

int issue(int method, void* dst, int j, const void* src, int i, size_t bytes, void* stream)
{
    switch (method) {
    case COPY: return b200v_peer_copy(dst, j, src, i, bytes, stream);
    case PULL: return b200v_peer_pull(dst, src, bytes, stream);
    default: return b200v_peer_push(dst, src, bytes, stream);
    }
}

struct Stats {
    double min = 0, median = 0, max = 0;
    bool any = false;
};

Stats stats_of(std::vector<double> v)
{
    Stats s;
    if (v.empty()) return s;
    std::sort(v.begin(), v.end());
    s.min = v.front();
    s.median = v[v.size() / 2];
    s.max = v.back();
    s.any = true;
    return s;
}

}  // namespace

// ----------------------------------------------------------------------------------------------
// p2p: all ordered pairs by copy engine, peer loads and peer stores; then every GPU at once

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

    std::vector<double> pair_gbs[kMethods];
    bool ok = true;
    printf("  %-8s %15s %15s %15s  verify\n", "src->dst", "memcpyPeer GB/s", "peer-load GB/s", "peer-store GB/s");
    for (int i = 0; i < g; ++i) {
        for (int j = 0; j < g; ++j) {
            if (i == j) continue;
            double best[kMethods] = {1e30, 1e30, 1e30};
            bool pair_ok = true;
            for (int method = 0; method < kMethods; ++method) {
                TRY(b200v_set_device(owner_of(method, i, j)));
                void *e0 = nullptr, *e1 = nullptr;
                TRY(b200v_event_create(&e0));
                TRY(b200v_event_create(&e1));
                for (int it = 0; it < a.warmup + std::max(a.iters / 4, 3); ++it) {
                    TRY(b200v_event_record(e0, nullptr));
                    TRY(issue(method, dst[j], j, src[i], i, bytes, nullptr));
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
                if (a.inject_fault && i == g - 1 && j == 0 && method == PULL) good = false;
                pair_ok = pair_ok && good;
                TRY(b200v_memset(dst[j], 0xEE, bytes, nullptr));
                TRY(b200v_device_sync());
            }
            double gbs[kMethods];
            for (int m = 0; m < kMethods; ++m) {
                gbs[m] = bytes / best[m] * 1e-6;
                pair_gbs[m].push_back(gbs[m]);
            }
            ok = ok && pair_ok;
            printf("  %d -> %d   %15.1f %15.1f %15.1f  %s\n", i, j, gbs[COPY], gbs[PULL], gbs[PUSH], pair_ok ? "ok" : "FAILED");
            fflush(stdout);
        }
    }
    // All GPUs at once: in round k every GPU j receives from GPU (j+k) mod g, so every NVLink port is
    // busy in both directions and the switch carries g concurrent flows.  The transfers on different GPUs
    // are independent (nobody waits on anybody); time = max over GPUs of the device-timed transfer.
    std::vector<double> ring_gbs[kMethods];
    {
        // one non-blocking stream per GPU: the legacy default stream would couple the devices' queues
        std::vector<void*> e0((size_t)g), e1((size_t)g), strm((size_t)g);
        for (int j = 0; j < g; ++j) {
            TRY(b200v_set_device(j));
            TRY(b200v_event_create(&e0[j]));
            TRY(b200v_event_create(&e1[j]));
            TRY(b200v_stream_create(&strm[j]));
        }
        printf("  concurrent rounds (every GPU receives from its k-th neighbour at the same time), GB/s into every GPU:\n");
        for (int k = 1; k < g; ++k) {
            double per_gpu[kMethods] = {0, 0, 0};
            bool round_ok = true;
            for (int method = 0; method < kMethods; ++method) {
                double best = 0;
                for (int rep = 0; rep < 3; ++rep) {
                    for (int j = 0; j < g; ++j) {          // destination j, source i = (j+k) mod g
                        const int i = (j + k) % g;
                        const int owner = owner_of(method, i, j);
                        TRY(b200v_set_device(owner));
                        TRY(b200v_event_record(e0[owner], strm[owner]));
                        TRY(issue(method, dst[j], j, src[i], i, bytes, strm[owner]));
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
                ring_gbs[method].push_back(per_gpu[method]);
            }
            ok = ok && round_ok;
            printf("    shift %d: memcpyPeer %.1f, peer-load kernel %.1f, peer-store kernel %.1f  (%.1f / %.1f / %.1f GB/s through the switch)  %s\n",
                   k, per_gpu[COPY], per_gpu[PULL], per_gpu[PUSH], per_gpu[COPY] * g, per_gpu[PULL] * g, per_gpu[PUSH] * g,
                   round_ok ? "ok" : "FAILED");
            fflush(stdout);
        }
        for (int j = 0; j < g; ++j) {
            b200v_set_device(j);
            b200v_event_destroy(e0[j]);
            b200v_event_destroy(e1[j]);
            b200v_stream_destroy(strm[j]);
        }
    }
    const char* const pair_name[kMethods] = {"memcpyPeer", "peer-load", "peer-store"};
    const char* const ring_name[kMethods] = {"all memcpyPeer", "all peer-load", "all peer-store"};
    Stats ps[kMethods], rs[kMethods];
    for (int m = 0; m < kMethods; ++m) {
        ps[m] = stats_of(pair_gbs[m]);
        rs[m] = stats_of(ring_gbs[m]);
    }
    auto summary = [&](const char* name, const Stats& s) {
        if (!s.any) return;
        printf("[p2p] %-15s min %.1f  median %.1f  max %.1f GB/s per direction (%.1f%% of the %.0f GB/s NVLink 5 figure at the median)\n",
               name, s.min, s.median, s.max, 100.0 * s.median / kNvlinkGBs, kNvlinkGBs);
    };
    for (int m = 0; m < kMethods; ++m) summary(pair_name[m], ps[m]);
    for (int m = 0; m < kMethods; ++m) summary(ring_name[m], rs[m]);
    if (!a.json_path.empty()) {
        if (FILE* js = fopen((a.json_path + (a.mode == "full" ? ".p2p" : "")).c_str(), "w")) {
            fprintf(js, "{\"mode\": \"p2p\", \"gpus\": %d, \"bytes\": %zu, \"ok\": %s, \"nvlink_gbs_per_direction\": %.0f, \"pairs\": [", g,
                    bytes, ok ? "true" : "false", kNvlinkGBs);
            size_t k = 0;
            for (int i = 0; i < g; ++i)
                for (int j = 0; j < g; ++j)
                    if (i != j) {
                        fprintf(js, "%s{\"src\": %d, \"dst\": %d, \"memcpy_peer_gbs\": %.1f, \"peer_load_gbs\": %.1f, \"peer_store_gbs\": %.1f}",
                                k ? ", " : "", i, j, pair_gbs[COPY][k], pair_gbs[PULL][k], pair_gbs[PUSH][k]);
                        ++k;
                    }
            fprintf(js, "]");
            // concurrent rounds, one figure per shift: GB/s into every GPU at once
            const char* const ring_key[kMethods] = {"concurrent_shift_push_gbs_per_gpu", "concurrent_shift_gbs_per_gpu",
                                                    "concurrent_shift_push_kernel_gbs_per_gpu"};
            for (int m = 0; m < kMethods; ++m) {
                fprintf(js, ", \"%s\": [", ring_key[m]);
                for (size_t q = 0; q < ring_gbs[m].size(); ++q) fprintf(js, "%s%.1f", q ? ", " : "", ring_gbs[m][q]);
                fprintf(js, "]");
            }
            fprintf(js, ", \"summary\": {");
            for (int m = 0; m < kMethods; ++m)
                fprintf(js, "%s\"%s\": {\"min\": %.1f, \"median\": %.1f, \"max\": %.1f, \"frac_of_nvlink_at_median\": %.4f}", m ? ", " : "",
                        kMethodKey[m], ps[m].min, ps[m].median, ps[m].max, ps[m].median / kNvlinkGBs);
            for (int m = 0; m < kMethods; ++m)
                fprintf(js, ", \"concurrent_%s\": {\"min\": %.1f, \"median\": %.1f, \"max\": %.1f, \"frac_of_nvlink_at_median\": %.4f}",
                        kMethodKey[m], rs[m].min, rs[m].median, rs[m].max, rs[m].median / kNvlinkGBs);
            fprintf(js, "}}\n");
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

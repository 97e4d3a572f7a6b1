// mode_node.cpp — `gpu-validate --mode node|burnin`: one process per GPU (BASELINE.json configs[3] and
// the burn-in half of configs[4]).  The parent forks and execs one child per GPU before it touches CUDA,
// releases them together once all are initialised, and reduces max(ms) / AND(verified) over pipes.
#include "launcher_common.h"

namespace gpuvalidate {

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
    long long ecc_corrected_new = -1, ecc_uncorrected_new = -1;   // ECC errors that appeared during the run
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
    const bool whole_hbm = burnin && a.hbm_fraction > 0;
    if (whole_hbm) {
        // memory-coverage burn-in: the three arrays take the requested share of this GPU's free HBM
        // (tens of billions of elements: every index computation in the kernels runs past 2^32)
        b200v_device_info fi;
        memset(&fi, 0, sizeof fi);
        fi.struct_size = sizeof fi;
        CTRY(b200v_device_info_get(a.device, &fi));
        count = (uint64_t)((double)fi.hbm_free_bytes * a.hbm_fraction / 3.0 / (double)esize(dtype));
        count &= ~(((uint64_t)1 << 20) - 1);
        if (count == 0) return child_fail(a, B200V_ERR_NOMEM, "no free HBM to cover");
        begin = (uint64_t)a.rank * count;   // every rank gets its own stretch of the generator's index space
    }
    // (a bf16 shard may begin on an odd global index: the buffer starts at its own element 0 and the
    // checksum's word indices are shard-local, so every element of [0, total) is covered exactly once)
    const size_t bytes = (size_t)count * esize(dtype);
    DevArrays v;
    CTRY(alloc_arrays(v, bytes));
    CTRY(b200v_fill(dtype, v.x, begin, count, a.seed, 0, B200V_GEN_UNIT, nullptr));
    CTRY(b200v_fill(dtype, v.y, begin, count, a.seed, 1, B200V_GEN_UNIT, nullptr));
    const bool two = op == B200V_OP_ADD || op == B200V_OP_TRIAD;
    for (int i = 0; i < a.warmup; ++i) CTRY(b200v_stream_op(op, dtype, v.d, v.x, two ? v.y : nullptr, kScalars[0], count, nullptr));
    // the warm-up leaves the right answer in the output: scrub it, so the checksum taken after the timed
    // launches proves that THEY wrote the result
    CTRY(b200v_memset(v.d, 0x5A, bytes, nullptr));
    CTRY(b200v_device_sync());

    // expected checksums from the host loop (word indices are shard-local)
    // whole-HBM runs keep to one scalar (the host's expected checksum over ~10^10 elements is the slow part)
    // and scrub the output between passes instead, so a stale result still cannot pass
    const int nscal = burnin && !whole_hbm ? 4 : 1;
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
    long long ecc_new[2] = {-1, -1};
    if (!burnin) {
        b200v_timing t;
        memset(&t, 0, sizeof t);
        t.struct_size = sizeof t;
        CTRY(b200v_time_op(op, dtype, v.d, v.x, two ? v.y : nullptr, kScalars[0], count, 0, a.iters, nullptr, &t));
        best_ms = t.best_ms;
        mean_ms = t.mean_ms;
        passes = 2L * a.iters;
        if (a.inject_fault && a.rank == a.world - 1 && bytes >= 4) {
            // flip one bit of one output word on the device: the checksum comparison itself must catch it
            const size_t at = (bytes / 4 / 2) * 4;
            uint32_t w = 0;
            CTRY(b200v_memcpy_d2h(&w, (const char*)v.d + at, 4, nullptr));
            CTRY(b200v_device_sync());
            w ^= 1u;
            CTRY(b200v_memcpy_h2d((char*)v.d + at, &w, 4, nullptr));
            CTRY(b200v_device_sync());
            child_say(a, "NOTE %d --inject-fault: flipped the low bit of output word %zu on the device", a.rank, at / 4);
        }
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
        nvmlprobe::Health h_before;
        const bool have_health = have_nvml && nvmlprobe::health(dinfo.pci_bus_id, &h_before);
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
            const int k = (int)(passes % nscal);
            if (whole_hbm) CTRY(b200v_memset(v.d, 0x5A, bytes, nullptr));
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
        nvmlprobe::Health h_after;
        if (have_health && nvmlprobe::health(dinfo.pci_bus_id, &h_after) && h_before.ecc_uncorrected >= 0 &&
            h_after.ecc_uncorrected >= 0) {
            ecc_new[0] = h_after.ecc_corrected - h_before.ecc_corrected;
            ecc_new[1] = h_after.ecc_uncorrected - h_before.ecc_uncorrected;
            if (ecc_new[1] > 0) ok = 0;   // memory that produced an uncorrectable error during the burn-in has failed it
        }
        b200v_event_destroy(e0);
        b200v_event_destroy(e1);
    }
    if (burnin && a.inject_fault && a.rank == a.world - 1) ok = 0;
    v.release();
    if (nv_max_temp >= 0)
        child_say(a, "NVML %d %d %d %.1f %llu %lld %lld", a.rank, nv_min_mhz, nv_max_temp, nv_max_power, nv_reasons, ecc_new[0], ecc_new[1]);
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

// Pod deletion sends SIGTERM to PID 1: pass it on to the shard children so no GPU process outlives the pod.
volatile sig_atomic_t g_child_pids[64];
volatile sig_atomic_t g_child_count = 0;

void forward_signal(int sig)
{
    for (int i = 0; i < g_child_count; ++i)
        if (g_child_pids[i] > 0) kill((pid_t)g_child_pids[i], SIGTERM);
    signal(sig, SIG_DFL);
    raise(sig);
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
    g_child_count = 0;
    signal(SIGTERM, forward_signal);
    signal(SIGINT, forward_signal);
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
            snprintf(b[12], 64, "%.4f", a.hbm_fraction);
            setenv("OMP_NUM_THREADS", b[11], 1);
            std::vector<const char*> av = {self, "--mode", child_mode, "--rank", b[0], "--world", b[1], "--device", b[2],
                                           "--go-fd", b[3], "--out-fd", b[4], "--log2n", b[5], "--n", b[6], "--iters", b[7],
                                           "--warmup", b[8], "--seconds", b[9], "--seed", b[10], "--hbm-fraction", b[12], "--dtype", a.dtype.c_str()};
            if (!a.op.empty()) { av.push_back("--op"); av.push_back(a.op.c_str()); }
            if (a.inject_fault) av.push_back("--inject-fault");
            av.push_back(nullptr);
            execv(self, const_cast<char* const*>(av.data()));
            _exit(127);
        }
        close(go[0]);
        close(out[1]);
        if (r < 64) {
            g_child_pids[r] = (sig_atomic_t)pid;
            g_child_count = r + 1;
        }
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
                    sscanf(line.c_str(), "NVML %d %d %d %lf %llu %lld %lld", &rk, &q.sm_min_mhz, &q.temp_max_c, &q.power_max_w, &q.reasons,
                           &q.ecc_corrected_new, &q.ecc_uncorrected_new);
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
    g_child_count = 0;
    signal(SIGTERM, SIG_DFL);
    signal(SIGINT, SIG_DFL);
    return EX_PASS;
}

int mode_node(const Args& a, bool burnin, bool print_verdict)
{
    const int visible = probe_device_count();
    printf("[gpu-validate] %s: one process per GPU, %s (seed %u)\n", burnin ? "burn-in" : "sharded run",
           burnin ? "every pass verified" : "device-timed, max over ranks", a.seed);
    if (visible <= 0) return fail(EX_NO_DEVICE, "no CUDA device is visible to this container");
    const int world = a.gpus > 0 ? a.gpus : visible;
    if (world > visible) return fail(EX_NO_DEVICE, "%d GPUs requested, %d visible", world, visible);
    const int dtype = dtype_id(a.dtype);
    const int op = a.op.empty() ? B200V_OP_TRIAD : op_id(a.op);
    Args b = a;
    if (b.log2n < 0 && !b.n_set) b.log2n = burnin ? 30 : 32;   // configs[3]: 2^32 fp32 elements in total
    const uint64_t total = b.log2n >= 0 ? (1ull << b.log2n) : b.n;
    if (burnin && a.hbm_fraction > 0)
        printf("[node] %s %s over %.0f%% of each GPU's free HBM, %d GPU%s (%d visible)\n", op_name(op), a.dtype.c_str(),
               100.0 * a.hbm_fraction, world, world == 1 ? "" : "s", visible);
    else
        printf("[node] %s %s, %" PRIu64 " elements in total over %d GPU%s (%d visible)\n", op_name(op), a.dtype.c_str(), total,
               world, world == 1 ? "" : "s", visible);
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
            if (s.ecc_uncorrected_new >= 0)
                printf("          ECC errors during the run: %lld corrected, %lld uncorrected%s\n", s.ecc_corrected_new, s.ecc_uncorrected_new,
                       s.ecc_uncorrected_new > 0 ? "  <-- FAILED" : "");
        }
        if (!s.text.empty()) printf("%s", s.text.c_str());
        max_best = std::max(max_best, s.best_ms);
        max_mean = std::max(max_mean, s.mean_ms);
        elems += s.count;
    }
    // The node figure is bytes / (max over ranks of the MEAN launch time): each rank times all its launches
    // back to back under one event pair, the ranks start together, so the slowest rank's mean is the node's
    // time per pass ("device-timed as the max over ranks", BASELINE.json).  The max over ranks of each
    // rank's fastest single launch is printed second, as the STREAM-convention figure.
    double node_gbs = 0, node_best_gbs = 0;
    if (max_mean > 0 && max_best > 0) {
        const double bytes = (double)b200v_op_bytes(op, dtype, (size_t)elems);
        node_gbs = bytes / max_mean * 1e-6;
        node_best_gbs = bytes / max_best * 1e-6;
        printf("[node] %d GPU%s: %.1f GB/s (bytes / max over ranks of mean ms = %.4f); best single launches: %.1f GB/s (max best ms = %.4f)\n",
               world, world == 1 ? "" : "s", node_gbs, max_mean, node_best_gbs, max_best);
        printf("[node] per GPU %.1f GB/s = %.1f%% of the %.0f GB/s nominal HBM peak\n", node_gbs / world,
               100.0 * node_gbs / world / kNominalGBs, kNominalGBs);
    }
    if (!a.json_path.empty()) {
        if (FILE* js = fopen(a.json_path.c_str(), "w")) {
            fprintf(js, "{\"mode\": \"%s\", \"op\": \"%s\", \"dtype\": \"%s\", \"gpus\": %d, \"elements\": %" PRIu64
                        ", \"max_best_ms\": %.6f, \"max_mean_ms\": %.6f, \"node_gbs\": %.1f, \"node_gbs_from\": \"max over ranks of mean ms\""
                        ", \"node_best_gbs\": %.1f, \"frac_of_nominal_per_gpu\": %.4f, \"ok\": %s, \"ranks\": [",
                    burnin ? "burnin" : "node", op_name(op), a.dtype.c_str(), world, elems, max_best, max_mean, node_gbs, node_best_gbs,
                    node_gbs / world / kNominalGBs, ok ? "true" : "false");
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

}  // namespace gpuvalidate

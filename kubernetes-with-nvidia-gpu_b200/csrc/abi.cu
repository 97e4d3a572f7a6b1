// abi.cu — host side of libb200v: status plumbing, device discovery, memory, timing, the
// host-buffer vectorAdd pipeline, shard arithmetic and the peer-to-peer helpers.
// Contract and reference citations: include/b200v.h.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

#include "internal.h"

namespace b200v {

static thread_local char g_err[512] = "";

int set_error(int status, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return status;
}

int cuda_status(cudaError_t e, const char* what, const char* file, int line)
{
    int status = B200V_ERR_CUDA;
    switch (e) {
    case cudaErrorNoDevice:
    case cudaErrorInsufficientDriver:
    case cudaErrorInitializationError:
    case cudaErrorSystemDriverMismatch:
    case cudaErrorCompatNotSupportedOnDevice:
    case cudaErrorNotSupported:
    case cudaErrorInvalidDevice:
        status = B200V_ERR_NO_DEVICE;
        break;
    case cudaErrorMemoryAllocation:
        status = B200V_ERR_NOMEM;
        break;
    case cudaErrorNoKernelImageForDevice:
    case cudaErrorInvalidDeviceFunction:
    case cudaErrorUnsupportedPtxVersion:
        status = B200V_ERR_ARCH;
        break;
    case cudaErrorPeerAccessUnsupported:
        status = B200V_ERR_PEER;
        break;
    default:
        break;
    }
    const char* base = strrchr(file, '/');
    snprintf(g_err, sizeof g_err, "%s: %s (%s) at %s:%d", what, cudaGetErrorString(e), cudaGetErrorName(e),
             base ? base + 1 : file, line);
    // leave the runtime's sticky "last error" clean for the next call
    if (e != cudaErrorNoDevice && e != cudaErrorInsufficientDriver) (void)cudaGetLastError();
    return status;
}

}  // namespace b200v

using namespace b200v;

extern "C" {

int b200v_abi_version(void) { return B200V_ABI_VERSION; }

const char* b200v_last_error(void) { return g_err; }

const char* b200v_status_name(int status)
{
    switch (status) {
    case B200V_OK: return "B200V_OK";
    case B200V_ERR_INVALID: return "B200V_ERR_INVALID";
    case B200V_ERR_NO_DEVICE: return "B200V_ERR_NO_DEVICE";
    case B200V_ERR_CUDA: return "B200V_ERR_CUDA";
    case B200V_ERR_ARCH: return "B200V_ERR_ARCH";
    case B200V_ERR_ALIGN: return "B200V_ERR_ALIGN";
    case B200V_ERR_PEER: return "B200V_ERR_PEER";
    case B200V_ERR_MISMATCH: return "B200V_ERR_MISMATCH";
    case B200V_ERR_NOMEM: return "B200V_ERR_NOMEM";
    default: return "B200V_ERR_UNKNOWN";
    }
}

// ------------------------------------------------------------------------------------------
// discovery

int b200v_device_count(int* count)
{
    B200V_REQUIRE(count != nullptr, B200V_ERR_INVALID, "NULL count");
    *count = 0;
    int n = 0;
    B200V_CUDA(cudaGetDeviceCount(&n));
    if (n <= 0) return set_error(B200V_ERR_NO_DEVICE, "no CUDA device is visible to this process");
    *count = n;
    return B200V_OK;
}

int b200v_device_info_get(int device, b200v_device_info* info)
{
    B200V_REQUIRE(info != nullptr, B200V_ERR_INVALID, "NULL info");
    B200V_REQUIRE(info->struct_size >= sizeof(b200v_device_info), B200V_ERR_INVALID,
                  "struct_size %u is smaller than this library's b200v_device_info (%zu)", info->struct_size,
                  sizeof(b200v_device_info));
    int n = 0;
    if (int rc = b200v_device_count(&n)) return rc;
    B200V_REQUIRE(device >= 0 && device < n, B200V_ERR_INVALID, "device %d not in [0, %d)", device, n);
    cudaDeviceProp p;
    B200V_CUDA(cudaGetDeviceProperties(&p, device));
    const uint32_t keep = info->struct_size;
    memset(info, 0, sizeof *info);
    info->struct_size = keep;
    info->device = device;
    snprintf(info->name, sizeof info->name, "%.127s", p.name);
    B200V_CUDA(cudaDeviceGetPCIBusId(info->pci_bus_id, sizeof info->pci_bus_id, device));
    const unsigned char* u = reinterpret_cast<const unsigned char*>(p.uuid.bytes);
    snprintf(info->uuid, sizeof info->uuid,
             "GPU-%02x%02x%02x%02x-%02x%02x-%02x%02x-%02x%02x-%02x%02x%02x%02x%02x%02x", u[0], u[1], u[2], u[3], u[4],
             u[5], u[6], u[7], u[8], u[9], u[10], u[11], u[12], u[13], u[14], u[15]);
    info->cc_major = p.major;
    info->cc_minor = p.minor;
    info->sm_count = p.multiProcessorCount;
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrClockRate, device) == cudaSuccess) info->clock_khz = v;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMemoryClockRate, device) == cudaSuccess) info->mem_clock_khz = v;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrGlobalMemoryBusWidth, device) == cudaSuccess) info->mem_bus_bits = v;
    (void)cudaGetLastError();
    info->hbm_bytes = p.totalGlobalMem;
    info->l2_bytes = (uint64_t)p.l2CacheSize;
    info->ecc_enabled = p.ECCEnabled;
    B200V_CUDA(cudaDriverGetVersion(&info->driver_version));
    B200V_CUDA(cudaRuntimeGetVersion(&info->runtime_version));
    // Free memory needs a context on the device.  Only the current device is asked, so that listing
    // the GPUs of an 8-GPU node does not create eight contexts (the default pod run must stay well
    // inside the 15 s of reference README.md:326); hbm_free_bytes stays 0 = "not queried" elsewhere.
    int cur = -1;
    B200V_CUDA(cudaGetDevice(&cur));
    size_t free_b = 0, total_b = 0;
    if (cur == device) B200V_CUDA(cudaMemGetInfo(&free_b, &total_b));
    info->hbm_free_bytes = free_b;
    return B200V_OK;
}

int b200v_set_device(int device)
{
    B200V_CUDA(cudaSetDevice(device));
    return B200V_OK;
}

int b200v_get_device(int* device)
{
    B200V_REQUIRE(device != nullptr, B200V_ERR_INVALID, "NULL device");
    B200V_CUDA(cudaGetDevice(device));
    return B200V_OK;
}

int b200v_peer_matrix(int ndev, int* can_access)
{
    B200V_REQUIRE(can_access != nullptr && ndev > 0, B200V_ERR_INVALID, "bad peer matrix arguments");
    int n = 0;
    if (int rc = b200v_device_count(&n)) return rc;
    B200V_REQUIRE(ndev <= n, B200V_ERR_INVALID, "asked for %d devices, %d visible", ndev, n);
    for (int i = 0; i < ndev; ++i)
        for (int j = 0; j < ndev; ++j) {
            int ok = 1;
            if (i != j) B200V_CUDA(cudaDeviceCanAccessPeer(&ok, i, j));
            can_access[i * ndev + j] = ok;
        }
    return B200V_OK;
}

int b200v_require_sm100(int device)
{
    int major = 0, minor = 0;
    B200V_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    B200V_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
    if (major != 10)
        return set_error(B200V_ERR_ARCH, "device %d is sm_%d%d; this image carries sm_100a kernels (B200) only", device,
                         major, minor);
    return B200V_OK;
}

// ------------------------------------------------------------------------------------------
// memory, streams

int b200v_malloc(void** dptr, size_t bytes)
{
    B200V_REQUIRE(dptr != nullptr, B200V_ERR_INVALID, "NULL out-pointer");
    *dptr = nullptr;
    if (bytes == 0) return B200V_OK;
    B200V_CUDA(cudaMalloc(dptr, bytes));
    return B200V_OK;
}

int b200v_free(void* dptr)
{
    if (dptr) B200V_CUDA(cudaFree(dptr));
    return B200V_OK;
}

int b200v_host_alloc(void** hptr, size_t bytes)
{
    B200V_REQUIRE(hptr != nullptr, B200V_ERR_INVALID, "NULL out-pointer");
    *hptr = nullptr;
    if (bytes == 0) return B200V_OK;
    B200V_CUDA(cudaHostAlloc(hptr, bytes, cudaHostAllocPortable));
    return B200V_OK;
}

int b200v_host_free(void* hptr)
{
    if (hptr) B200V_CUDA(cudaFreeHost(hptr));
    return B200V_OK;
}

int b200v_host_register(void* hptr, size_t bytes)
{
    B200V_REQUIRE(hptr != nullptr && bytes > 0, B200V_ERR_INVALID, "NULL or empty host range");
    B200V_CUDA(cudaHostRegister(hptr, bytes, cudaHostRegisterPortable));
    return B200V_OK;
}

int b200v_host_unregister(void* hptr)
{
    if (hptr) B200V_CUDA(cudaHostUnregister(hptr));
    return B200V_OK;
}

int b200v_memcpy_h2d(void* dptr, const void* hptr, size_t bytes, void* stream)
{
    if (bytes == 0) return B200V_OK;
    B200V_REQUIRE(dptr && hptr, B200V_ERR_INVALID, "NULL pointer");
    B200V_CUDA(cudaMemcpyAsync(dptr, hptr, bytes, cudaMemcpyHostToDevice, as_stream(stream)));
    return B200V_OK;
}

int b200v_memcpy_d2h(void* hptr, const void* dptr, size_t bytes, void* stream)
{
    if (bytes == 0) return B200V_OK;
    B200V_REQUIRE(dptr && hptr, B200V_ERR_INVALID, "NULL pointer");
    B200V_CUDA(cudaMemcpyAsync(hptr, dptr, bytes, cudaMemcpyDeviceToHost, as_stream(stream)));
    return B200V_OK;
}

int b200v_memcpy_d2d(void* dst, const void* src, size_t bytes, void* stream)
{
    if (bytes == 0) return B200V_OK;
    B200V_REQUIRE(dst && src, B200V_ERR_INVALID, "NULL pointer");
    B200V_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, as_stream(stream)));
    return B200V_OK;
}

int b200v_memset(void* dptr, int byte, size_t bytes, void* stream)
{
    if (bytes == 0) return B200V_OK;
    B200V_REQUIRE(dptr != nullptr, B200V_ERR_INVALID, "NULL pointer");
    B200V_CUDA(cudaMemsetAsync(dptr, byte, bytes, as_stream(stream)));
    return B200V_OK;
}

int b200v_stream_create(void** stream)
{
    B200V_REQUIRE(stream != nullptr, B200V_ERR_INVALID, "NULL out-pointer");
    cudaStream_t s = nullptr;
    B200V_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    *stream = s;
    return B200V_OK;
}

int b200v_stream_destroy(void* stream)
{
    if (stream) B200V_CUDA(cudaStreamDestroy(as_stream(stream)));
    return B200V_OK;
}

int b200v_stream_sync(void* stream)
{
    B200V_CUDA(cudaStreamSynchronize(as_stream(stream)));
    return B200V_OK;
}

int b200v_device_sync(void)
{
    B200V_CUDA(cudaDeviceSynchronize());
    return B200V_OK;
}

// ------------------------------------------------------------------------------------------
// timing

int b200v_event_create(void** event)
{
    B200V_REQUIRE(event != nullptr, B200V_ERR_INVALID, "NULL out-pointer");
    cudaEvent_t e = nullptr;
    B200V_CUDA(cudaEventCreate(&e));
    *event = e;
    return B200V_OK;
}

int b200v_event_destroy(void* event)
{
    if (event) B200V_CUDA(cudaEventDestroy(reinterpret_cast<cudaEvent_t>(event)));
    return B200V_OK;
}

int b200v_event_record(void* event, void* stream)
{
    B200V_REQUIRE(event != nullptr, B200V_ERR_INVALID, "NULL event");
    B200V_CUDA(cudaEventRecord(reinterpret_cast<cudaEvent_t>(event), as_stream(stream)));
    return B200V_OK;
}

int b200v_event_sync(void* event)
{
    B200V_REQUIRE(event != nullptr, B200V_ERR_INVALID, "NULL event");
    B200V_CUDA(cudaEventSynchronize(reinterpret_cast<cudaEvent_t>(event)));
    return B200V_OK;
}

int b200v_event_elapsed_ms(void* start, void* stop, float* ms)
{
    B200V_REQUIRE(start && stop && ms, B200V_ERR_INVALID, "NULL argument");
    B200V_CUDA(cudaEventElapsedTime(ms, reinterpret_cast<cudaEvent_t>(start), reinterpret_cast<cudaEvent_t>(stop)));
    return B200V_OK;
}

int b200v_time_op(int op, int dtype, void* dst, const void* x, const void* y, float s, size_t n, int warmup, int iters,
                  void* stream, b200v_timing* out)
{
    B200V_REQUIRE(out != nullptr && out->struct_size >= sizeof(b200v_timing), B200V_ERR_INVALID,
                  "NULL or undersized b200v_timing");
    B200V_REQUIRE(iters > 0 && iters <= 100000 && warmup >= 0, B200V_ERR_INVALID, "bad warmup/iters");
    cudaStream_t st = as_stream(stream);
    for (int i = 0; i < warmup; ++i)
        if (int rc = b200v_stream_op(op, dtype, dst, x, y, s, n, stream)) return rc;

    std::vector<cudaEvent_t> ev(2 * (size_t)iters + 2, nullptr);
    int rc = B200V_OK;
    cudaError_t ce = cudaSuccess;
    for (auto& e : ev)
        if ((ce = cudaEventCreate(&e)) != cudaSuccess) break;
    std::vector<float> ms((size_t)iters, 0.0f);
    float total = 0.0f;
    if (ce == cudaSuccess) {
        for (int i = 0; i < iters && rc == B200V_OK && ce == cudaSuccess; ++i) {
            ce = cudaEventRecord(ev[2 * i], st);
            rc = b200v_stream_op(op, dtype, dst, x, y, s, n, stream);
            if (ce == cudaSuccess) ce = cudaEventRecord(ev[2 * i + 1], st);
        }
        if (rc == B200V_OK && ce == cudaSuccess) ce = cudaEventRecord(ev[2 * iters], st);
        for (int i = 0; i < iters && rc == B200V_OK && ce == cudaSuccess; ++i)
            rc = b200v_stream_op(op, dtype, dst, x, y, s, n, stream);
        if (rc == B200V_OK && ce == cudaSuccess) ce = cudaEventRecord(ev[2 * iters + 1], st);
        if (rc == B200V_OK && ce == cudaSuccess) ce = cudaStreamSynchronize(st);
        for (int i = 0; i < iters && rc == B200V_OK && ce == cudaSuccess; ++i)
            ce = cudaEventElapsedTime(&ms[i], ev[2 * i], ev[2 * i + 1]);
        if (rc == B200V_OK && ce == cudaSuccess) ce = cudaEventElapsedTime(&total, ev[2 * iters], ev[2 * iters + 1]);
    }
    for (auto e : ev)
        if (e) cudaEventDestroy(e);
    if (rc != B200V_OK) return rc;
    if (ce != cudaSuccess) return cuda_status(ce, "b200v_time_op", __FILE__, __LINE__);
    std::sort(ms.begin(), ms.end());
    out->iters = iters;
    out->best_ms = ms.front();
    out->median_ms = ms[ms.size() / 2];
    out->mean_ms = total / (float)iters;
    out->bytes_per_launch = b200v_op_bytes(op, dtype, n);
    return B200V_OK;
}

// ------------------------------------------------------------------------------------------
// sharding

int b200v_shard_range(uint64_t n, int rank, int world, uint64_t* begin, uint64_t* count)
{
    B200V_REQUIRE(begin && count, B200V_ERR_INVALID, "NULL out-pointer");
    B200V_REQUIRE(world > 0 && rank >= 0 && rank < world, B200V_ERR_INVALID, "rank %d not in [0, %d)", rank, world);
    const uint64_t q = n / (uint64_t)world, r = n % (uint64_t)world;
    const uint64_t k = (uint64_t)rank;
    *begin = k * q + (k < r ? k : r);
    *count = q + (k < r ? 1 : 0);
    return B200V_OK;
}

// ------------------------------------------------------------------------------------------
// vectorAdd with host buffers: chunked H2D -> kernel -> D2H pipeline.  Each slot owns a stream
// and three device chunks; consecutive chunks go to different slots, so chunk k+1's H2D runs
// on the copy engine while chunk k computes and chunk k-1 drains on the other engine.

namespace {

constexpr int kMaxSlots = 8;
// defaults from the round-1 sweep (profiles/r01_e2e_sweep.json): see b200v_vector_add_host_config
constexpr int kDefaultSlots = 3;
constexpr size_t kDefaultChunkElems = (size_t)8 << 20;  // 8 Mi fp32 = 32 MiB per array per slot

// One pipeline per device ordinal; its shape (want_*) is set per device too, so a process driving
// several GPUs can shape each one's pipeline on its own.
struct HostPipe {
    bool ready = false;
    int want_slots = kDefaultSlots;
    size_t want_chunk_elems = kDefaultChunkElems;
    int slots = 0;
    size_t chunk_elems = 0;
    cudaStream_t stream[kMaxSlots] = {};
    float* a[kMaxSlots] = {};
    float* b[kMaxSlots] = {};
    float* c[kMaxSlots] = {};
};
HostPipe g_pipe[64];

int pipe_release(HostPipe& p)
{
    for (int i = 0; i < kMaxSlots; ++i) {
        if (p.a[i]) cudaFree(p.a[i]);
        if (p.b[i]) cudaFree(p.b[i]);
        if (p.c[i]) cudaFree(p.c[i]);
        if (p.stream[i]) cudaStreamDestroy(p.stream[i]);
        p.a[i] = p.b[i] = p.c[i] = nullptr;
        p.stream[i] = nullptr;
    }
    p.ready = false;
    (void)cudaGetLastError();
    return B200V_OK;
}

int pipe_init(HostPipe& p)
{
    if (p.ready && p.slots == p.want_slots && p.chunk_elems == p.want_chunk_elems) return B200V_OK;
    if (p.ready) pipe_release(p);
    p.slots = p.want_slots;
    p.chunk_elems = p.want_chunk_elems;
    for (int i = 0; i < p.slots; ++i) {
        cudaError_t e = cudaStreamCreateWithFlags(&p.stream[i], cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaMalloc((void**)&p.a[i], p.chunk_elems * sizeof(float));
        if (e == cudaSuccess) e = cudaMalloc((void**)&p.b[i], p.chunk_elems * sizeof(float));
        if (e == cudaSuccess) e = cudaMalloc((void**)&p.c[i], p.chunk_elems * sizeof(float));
        if (e != cudaSuccess) {
            pipe_release(p);
            return cuda_status(e, "staging buffers for b200v_vector_add_host", __FILE__, __LINE__);
        }
    }
    p.ready = true;
    return B200V_OK;
}

int current_pipe(HostPipe** out)
{
    int dev = 0;
    B200V_CUDA(cudaGetDevice(&dev));
    B200V_REQUIRE(dev >= 0 && dev < 64, B200V_ERR_INVALID, "device index %d out of range", dev);
    *out = &g_pipe[dev];
    return B200V_OK;
}

// one chunk: H2D of both inputs, the add kernel, D2H of the result, all on the slot's stream
int enqueue_chunk(HostPipe& p, int slot, const float* hA, const float* hB, float* hC, size_t m)
{
    cudaStream_t st = p.stream[slot];
    // stream order makes the slot's previous D2H finish before its buffers are overwritten
    B200V_CUDA(cudaMemcpyAsync(p.a[slot], hA, m * sizeof(float), cudaMemcpyHostToDevice, st));
    B200V_CUDA(cudaMemcpyAsync(p.b[slot], hB, m * sizeof(float), cudaMemcpyHostToDevice, st));
    if (int rc = b200v_add_f32(p.c[slot], p.a[slot], p.b[slot], m, st)) return rc;
    B200V_CUDA(cudaMemcpyAsync(hC, p.c[slot], m * sizeof(float), cudaMemcpyDeviceToHost, st));
    return B200V_OK;
}

}  // namespace

int b200v_vector_add_host_config(size_t chunk_elems, int slots)
{
    B200V_REQUIRE(slots >= 1 && slots <= kMaxSlots, B200V_ERR_INVALID, "slots %d not in [1, %d]", slots, kMaxSlots);
    B200V_REQUIRE(chunk_elems >= 1024 && chunk_elems <= ((size_t)1 << 30), B200V_ERR_INVALID,
                  "chunk_elems %zu not in [1024, 2^30]", chunk_elems);
    HostPipe* p = nullptr;
    if (int rc = current_pipe(&p)) return rc;
    p->want_slots = slots;
    p->want_chunk_elems = chunk_elems;
    return B200V_OK;
}

int b200v_vector_add_host(const float* hA, const float* hB, float* hC, size_t n)
{
    B200V_REQUIRE(hA && hB && hC, B200V_ERR_INVALID, "NULL host array");
    if (n == 0) return B200V_OK;
    HostPipe* pp = nullptr;
    if (int rc = current_pipe(&pp)) return rc;
    HostPipe& p = *pp;
    if (int rc = pipe_init(p)) return rc;

    // Full-size chunks through the body; the last two chunks' worth is cut into halves of what remains, down to
    // 1/32 of a chunk.  The inbound copies run back to back whatever the chunking, so the call ends one last
    // kernel + one last outbound copy after them: the smaller the last chunk, the shorter that unoverlapped tail
    // (32 MiB at PCIe speed is 0.6 ms of a 42 ms step; measured gain ~0.5 %, profiles/r02_e2e_pipeline_notes.md).
    const size_t floor_elems = std::max<size_t>(p.chunk_elems / 32, 1024);
    int rc = B200V_OK;
    size_t k = 0;
    for (size_t off = 0; off < n && rc == B200V_OK; ++k) {
        const size_t rem = n - off;
        size_t m = std::min(p.chunk_elems, rem);
        if (rem < 2 * p.chunk_elems && rem > floor_elems) m = std::min(m, std::max(floor_elems, (rem / 2 + 7) & ~(size_t)7));
        rc = enqueue_chunk(p, (int)(k % p.slots), hA + off, hB + off, hC + off, m);
        off += m;
    }
    // The call is synchronous on every path: whatever was queued — also before a failure — has finished
    // with the caller's host buffers by the time it returns, so they may be freed or reused at once.
    if (rc != B200V_OK) {
        char first_error[sizeof g_err];
        snprintf(first_error, sizeof first_error, "%s", g_err);
        for (int i = 0; i < p.slots; ++i) (void)cudaStreamSynchronize(p.stream[i]);   // drain; the first error is the one reported
        (void)cudaGetLastError();
        snprintf(g_err, sizeof g_err, "%s", first_error);
        return rc;
    }
    for (int i = 0; i < p.slots; ++i) {
        const cudaError_t e = cudaStreamSynchronize(p.stream[i]);
        if (e != cudaSuccess && rc == B200V_OK) rc = cuda_status(e, "cudaStreamSynchronize (host pipeline)", __FILE__, __LINE__);
    }
    return rc;
}

int b200v_vector_add_host_release(void)
{
    int dev = 0;
    B200V_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return B200V_OK;
    return pipe_release(g_pipe[dev]);
}

// ------------------------------------------------------------------------------------------
// peer to peer

int b200v_peer_enable(int device, int peer)
{
    if (device == peer) return B200V_OK;
    int ok = 0;
    B200V_CUDA(cudaDeviceCanAccessPeer(&ok, device, peer));
    if (!ok) return set_error(B200V_ERR_PEER, "device %d cannot map device %d's memory", device, peer);
    int prev = 0;
    B200V_CUDA(cudaGetDevice(&prev));
    B200V_CUDA(cudaSetDevice(device));
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
        (void)cudaGetLastError();
        e = cudaSuccess;
    }
    (void)cudaSetDevice(prev);
    if (e != cudaSuccess) return cuda_status(e, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
    return B200V_OK;
}

int b200v_peer_copy(void* dst, int dst_device, const void* src, int src_device, size_t bytes, void* stream)
{
    if (bytes == 0) return B200V_OK;
    B200V_REQUIRE(dst && src, B200V_ERR_INVALID, "NULL pointer");
    B200V_CUDA(cudaMemcpyPeerAsync(dst, dst_device, src, src_device, bytes, as_stream(stream)));
    return B200V_OK;
}

}  // extern "C"

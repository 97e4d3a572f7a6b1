/*
 * b200v.h — thin C ABI of the B200 GPU validation workload (libb200v.so).
 *
 * What this replaces.  The reference (mysticrenji/kubernetes-with-nvidia-gpu) has no
 * plugin / operator / FFI for its GPU path: the whole path is the payload of the Step 9
 * pod `cuda-vector-add`, `command: ["/bin/sh","-c"] args: ["nvidia-smi"]`
 * (/root/reference/README.md:302-319), whose stdout is read with `kubectl logs`
 * (README:332) and whose exit code drives `restartPolicy: OnFailure` (README:309).
 * BASELINE.json asks for that payload to become a real cuda-samples-style vectorAdd plus
 * STREAM copy/scale/add/triad, "host code ... calling CUDA through a thin C ABI".  This
 * header is that ABI; each group names the reference behaviour it stands in for.
 *
 * Conventions.
 *  - C linkage, plain pointers and sizes, no C++ types, nothing thrown across the boundary.
 *  - Every function returns an int status: B200V_OK (0) or a stable B200V_ERR_* code; the
 *    library never aborts or exits — the launcher decides the process exit code
 *    (README:309: exit code is the machine-readable verdict).
 *  - b200v_last_error() returns a thread-local description of the last failure.
 *  - Device pointers are ordinary CUDA device addresses (cudaMalloc, torch, cuMemAlloc ...);
 *    `stream` is a cudaStream_t passed as void* (NULL = the default stream).  Kernel entry
 *    points only enqueue work; they do not synchronise.
 *  - "Current device" is the only hidden state.  One host thread per device; the shard
 *    launcher uses processes, not threads.
 *  - There is no CPU fallback anywhere: without a usable GPU every compute entry point
 *    returns B200V_ERR_NO_DEVICE / B200V_ERR_CUDA.
 */
#ifndef B200V_H
#define B200V_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200V_ABI_VERSION 1

#if defined(__GNUC__)
#define B200V_API __attribute__((visibility("default")))
#else
#define B200V_API
#endif

/* ---- status ---------------------------------------------------------------------------- */
enum {
    B200V_OK = 0,
    B200V_ERR_INVALID = 1,     /* bad argument (NULL, overlap, unknown op/dtype/mode, size)  */
    B200V_ERR_NO_DEVICE = 2,   /* no CUDA device visible / driver missing or too old         */
    B200V_ERR_CUDA = 3,        /* any other CUDA runtime error (text in b200v_last_error)    */
    B200V_ERR_ARCH = 4,        /* device is not sm_100 (the kernels are built for sm_100a)   */
    B200V_ERR_ALIGN = 5,       /* pointer not aligned to its element type                    */
    B200V_ERR_PEER = 6,        /* peer access between the two devices is not available       */
    B200V_ERR_MISMATCH = 7,    /* a verify step found differing data                         */
    B200V_ERR_NOMEM = 8        /* device or pinned-host allocation failed                    */
};

B200V_API int b200v_abi_version(void);
B200V_API const char *b200v_last_error(void);
B200V_API const char *b200v_status_name(int status);

/* ---- discovery -------------------------------------------------------------------------
 * Stands in for the `nvidia-smi` "is the GPU there" gates (README:76-84 on the host,
 * README:313-314 / README:335 inside the pod): enumerate what the container was granted
 * by `resources.limits.nvidia.com/gpu` (README:315-317), never assume a count.            */
typedef struct b200v_device_info {
    uint32_t struct_size;          /* caller sets to sizeof(b200v_device_info)               */
    int32_t  device;
    char     name[128];
    char     pci_bus_id[32];
    char     uuid[48];
    int32_t  cc_major, cc_minor;
    int32_t  sm_count;
    int32_t  clock_khz;            /* SM clock as the runtime reports it                     */
    int32_t  mem_clock_khz;
    int32_t  mem_bus_bits;
    uint64_t hbm_bytes;
    uint64_t hbm_free_bytes;       /* filled for the CURRENT device only (needs a context); else 0 */
    uint64_t l2_bytes;
    int32_t  driver_version;       /* cudaDriverGetVersion, e.g. 13000                       */
    int32_t  runtime_version;      /* cudaRuntimeGetVersion, e.g. 12090                      */
    int32_t  ecc_enabled;
    int32_t  reserved;
} b200v_device_info;

B200V_API int b200v_device_count(int *count);
B200V_API int b200v_device_info_get(int device, b200v_device_info *info);
B200V_API int b200v_set_device(int device);
B200V_API int b200v_get_device(int *device);
/* can_access[i*ndev + j] = 1 if device i can map device j's memory (diagonal = 1) */
B200V_API int b200v_peer_matrix(int ndev, int *can_access);
/* B200V_OK only on compute capability 10.x (the kernels are sm_100a SASS, no PTX JIT path) */
B200V_API int b200v_require_sm100(int device);

/* ---- memory, streams --------------------------------------------------------------------
 * Whoever allocates frees.  cuda-samples vectorAdd's cudaMalloc / cudaMemcpy steps.        */
B200V_API int b200v_malloc(void **dptr, size_t bytes);
B200V_API int b200v_free(void *dptr);
B200V_API int b200v_host_alloc(void **hptr, size_t bytes);      /* pinned */
B200V_API int b200v_host_free(void *hptr);
/* pin host memory the caller already owns (malloc, mmap, a numpy array), in place, so that copies
 * from / to it run at PCIe speed without a staging pass; undo with b200v_host_unregister before freeing it */
B200V_API int b200v_host_register(void *hptr, size_t bytes);
B200V_API int b200v_host_unregister(void *hptr);
B200V_API int b200v_memcpy_h2d(void *dptr, const void *hptr, size_t bytes, void *stream);
B200V_API int b200v_memcpy_d2h(void *hptr, const void *dptr, size_t bytes, void *stream);
/* the driver's own device-to-device copy: the same-box comparator the sweep prints beside the kernels */
B200V_API int b200v_memcpy_d2d(void *dst, const void *src, size_t bytes, void *stream);
B200V_API int b200v_memset(void *dptr, int byte, size_t bytes, void *stream);
B200V_API int b200v_stream_create(void **stream);
B200V_API int b200v_stream_destroy(void *stream);
B200V_API int b200v_stream_sync(void *stream);
B200V_API int b200v_device_sync(void);

/* ---- deterministic inputs ----------------------------------------------------------------
 * dst[j] = value of global index start+j of (seed, stream_id); same integer hash on host
 * and device so any shard of a 2^32-element vector can be synthesised where it is needed.  */
enum { B200V_GEN_UNIT = 0, B200V_GEN_BITS = 1, B200V_GEN_DENORM = 2 };
enum { B200V_F32 = 0, B200V_BF16 = 1 };

B200V_API int b200v_fill(int dtype, void *d_dst, uint64_t start, size_t n, uint32_t seed, uint32_t stream_id,
               int mode, void *stream);

/* ---- the workload kernels -----------------------------------------------------------------
 * The payload that takes the place of `args: ["nvidia-smi"]` (README:313-314).
 * n counts elements.  Arrays must not overlap.  bf16 arrays are uint16_t bit patterns.
 *   add   : c[i] = a[i] + b[i]          (cuda-samples vectorAdd / STREAM add)
 *   copy  : c[i] = a[i]
 *   scale : b[i] = s * c[i]
 *   triad : a[i] = fmaf(s, c[i], b[i])  (one rounding)
 * Launches carry cudaLaunchAttributeProgrammaticStreamSerialization and the kernels wait
 * (griddepcontrol.wait) before touching memory: consecutive launches on one stream overlap
 * their drain/placement, the order of the data on the stream is the usual one.              */
B200V_API int b200v_add_f32(float *c, const float *a, const float *b, size_t n, void *stream);
B200V_API int b200v_copy_f32(float *c, const float *a, size_t n, void *stream);
B200V_API int b200v_scale_f32(float *b, float s, const float *c, size_t n, void *stream);
B200V_API int b200v_triad_f32(float *a, float s, const float *c, const float *b, size_t n, void *stream);
B200V_API int b200v_add_bf16(uint16_t *c, const uint16_t *a, const uint16_t *b, size_t n, void *stream);
B200V_API int b200v_copy_bf16(uint16_t *c, const uint16_t *a, size_t n, void *stream);
B200V_API int b200v_scale_bf16(uint16_t *b, float s, const uint16_t *c, size_t n, void *stream);
B200V_API int b200v_triad_bf16(uint16_t *a, float s, const uint16_t *c, const uint16_t *b, size_t n, void *stream);

/* Same kernels by number: dst = op(x, y, s).  add: x+y; copy: x; scale: s*x; triad: fma(s,x,y). */
enum { B200V_OP_ADD = 0, B200V_OP_COPY = 1, B200V_OP_SCALE = 2, B200V_OP_TRIAD = 3 };
B200V_API int b200v_stream_op(int op, int dtype, void *dst, const void *x, const void *y, float s, size_t n,
                    void *stream);
/* algorithmic bytes one launch moves (STREAM convention: 12/8/8/12 B per fp32 element) */
B200V_API uint64_t b200v_op_bytes(int op, int dtype, size_t n);

/* ---- vectorAdd with HOST buffers, the cuda-samples call shape ------------------------------
 * hC = hA + hB for n fp32 elements held in host memory: H2D, kernel, D2H, all inside the
 * call, chunked and pipelined over copy engines.  Pinned buffers (b200v_host_alloc) get
 * PCIe speed; pageable ones work but are staged by the driver.  Synchronous on every path: when
 * the call returns — with B200V_OK or with an error — no copy from hA/hB or into hC is still in
 * flight, so the caller may free or reuse the host buffers at once.                          */
B200V_API int b200v_vector_add_host(const float *hA, const float *hB, float *hC, size_t n);
B200V_API int b200v_vector_add_host_release(void);   /* frees the per-device staging buffers */
/* pipeline shape of the CURRENT device: elements per staged chunk and number of chunk slots
 * (streams) in flight; takes effect at that device's next b200v_vector_add_host call.
 * Defaults: 8 Mi elements, 3 slots.                                                          */
B200V_API int b200v_vector_add_host_config(size_t chunk_elems, int slots);

/* ---- checksum (verifies outputs too large to bring back) -----------------------------------
 * Order-independent, position-sensitive checksum of a device buffer read as little-endian
 * 32-bit words w_j, j counted from word_offset (1..3 trailing bytes are zero-extended):
 *   out[0] = sum w_j ;  out[1] = sum w_j * ((uint32_t)j | 1)      (both mod 2^64)
 * Shards' checksums add up to the whole vector's.  d_workspace is scratch the caller owns
 * (b200v_checksum_workspace_bytes() bytes, 16-byte aligned); d_out is 2 x uint64 on device. */
B200V_API size_t b200v_checksum_workspace_bytes(void);
B200V_API int b200v_checksum(const void *d_buf, size_t nbytes, uint64_t word_offset, void *d_workspace,
                   size_t workspace_bytes, uint64_t *d_out, void *stream);

/* b200v_checksum_static has no counterpart in the reference (nor among the entry points above
 * that mirror its pod): it is b200v_checksum for callers that cannot supply a workspace because
 * they cannot allocate device memory when they need the checksum — a burn-in whose arrays already
 * occupy all free HBM, a probe inside a process whose allocator has reserved the whole GPU, a
 * hook that must avoid cudaMalloc/cudaFree because they synchronise the device.  The scratch is a
 * fixed array inside the library image (about 16 KiB on every GPU that loads it); the call
 * allocates nothing.  Results are identical to b200v_checksum.
 *  - at most B200V_CHECKSUM_STATIC_MAX_BYTES per call, else B200V_ERR_INVALID (split the buffer:
 *    pieces add up; or use b200v_checksum);
 *  - the first call on each device also looks up the array's device address: make it at start-up,
 *    with that device current, before any stream capture begins, and from one host thread (the
 *    look-up is not locked; it falls under this header's one-host-thread-per-device rule);
 *  - the address is remembered for the life of the process: do not call cudaDeviceReset (or tear
 *    down the primary context) between calls — the remembered address would point into a dead
 *    context and the next call would fail or write through it;
 *  - all calls on a device share the array: calls that overlap on the GPU race, so keep them on
 *    one stream or order them with events.                                                      */
#define B200V_CHECKSUM_STATIC_MAX_BYTES (1ull << 32)
B200V_API int b200v_checksum_static(const void *d_buf, size_t nbytes, uint64_t word_offset, uint64_t *d_out,
                                    void *stream);

/* ---- timing (device-timed only; no host clocks in reported GPU numbers) -------------------- */
B200V_API int b200v_event_create(void **event);
B200V_API int b200v_event_destroy(void *event);
B200V_API int b200v_event_record(void *event, void *stream);
B200V_API int b200v_event_sync(void *event);
B200V_API int b200v_event_elapsed_ms(void *start, void *stop, float *ms);

typedef struct b200v_timing {
    uint32_t struct_size;      /* caller sets to sizeof(b200v_timing) */
    int32_t  iters;
    float    best_ms;          /* fastest single launch                                   */
    float    median_ms;        /* median single launch                                    */
    float    mean_ms;          /* all `iters` launches back to back, one event pair / iters */
    uint64_t bytes_per_launch; /* b200v_op_bytes                                           */
} b200v_timing;
/* warm-up launches, then `iters` launches each bracketed by its own event pair, then the same
 * `iters` launches back to back under one event pair. */
B200V_API int b200v_time_op(int op, int dtype, void *dst, const void *x, const void *y, float s, size_t n,
                  int warmup, int iters, void *stream, b200v_timing *out);

/* ---- sharding (host arithmetic only; no GPU needed) ------------------------------------------
 * Contiguous equal split: rank r of world G owns [begin, begin+count), remainders to the
 * lowest ranks.  The vectors shard with no communication, so this is the whole "plan".      */
B200V_API int b200v_shard_range(uint64_t n, int rank, int world, uint64_t *begin, uint64_t *count);

/* ---- NVLink peer-to-peer check (no counterpart in the reference, which requests one GPU:
 * README:317; BASELINE.json configs[4]) -------------------------------------------------------- */
B200V_API int b200v_peer_enable(int device, int peer);      /* idempotent */
/* cudaMemcpyPeerAsync on `stream` of the current device */
B200V_API int b200v_peer_copy(void *dst, int dst_device, const void *src, int src_device, size_t bytes, void *stream);
/* kernel on the CURRENT device (the destination) pulling `bytes` from a peer's buffer with direct
 * 256-bit loads: read requests go out, data comes back */
B200V_API int b200v_peer_pull(void *d_local_dst, const void *d_peer_src, size_t bytes, void *stream);
/* kernel on the CURRENT device (the source) pushing `bytes` of its own buffer into a peer's with direct
 * 256-bit stores: posted writes, nothing but acknowledgements on the return path, so it keeps its
 * rate when every GPU sends and receives at once.  Peer access current -> peer must be enabled.   */
B200V_API int b200v_peer_push(void *d_peer_dst, const void *d_local_src, size_t bytes, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200V_H */

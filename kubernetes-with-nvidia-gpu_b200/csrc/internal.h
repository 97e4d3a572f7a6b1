// internal.h — error plumbing shared by the translation units of libb200v (not part of the ABI).
#pragma once

#include <cstdarg>
#include <cstdio>
#include <cuda_runtime.h>

#include "b200v.h"

namespace b200v {

int set_error(int status, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
int cuda_status(cudaError_t e, const char* what, const char* file, int line);

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

}  // namespace b200v

#define B200V_CUDA(call)                                                                  \
    do {                                                                                  \
        cudaError_t e__ = (call);                                                         \
        if (e__ != cudaSuccess) return ::b200v::cuda_status(e__, #call, __FILE__, __LINE__); \
    } while (0)

#define B200V_REQUIRE(cond, status, ...)                                \
    do {                                                                \
        if (!(cond)) return ::b200v::set_error((status), __VA_ARGS__);  \
    } while (0)

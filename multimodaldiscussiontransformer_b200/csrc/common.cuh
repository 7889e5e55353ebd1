// Shared helpers for libmdt_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/mdt_b200.h"

namespace mdt {

// Thread-local text of the last failure; returned by mdt_last_error_string().
char *last_error_buffer();
int fail(int code, const char *fmt, ...);

#define MDT_CHECK_CUDA(expr)                                                                     \
    do {                                                                                         \
        cudaError_t _e = (expr);                                                                 \
        if (_e != cudaSuccess)                                                                   \
            return ::mdt::fail(MDT_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                               __FILE__, __LINE__);                                              \
    } while (0)

#define MDT_REQUIRE(cond, ...)                                  \
    do {                                                        \
        if (!(cond)) return ::mdt::fail(MDT_EINVAL, __VA_ARGS__); \
    } while (0)

// Refuses to run on anything but sm_100: there is no other code path in this library.
int require_sm100();

inline cudaStream_t as_stream(mdt_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

inline int num_sms() {
    static int cached[64] = {0};  // per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (cached[dev] == 0) {
        int n = 0;
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        cached[dev] = n > 0 ? n : 148;
    }
    return cached[dev];
}

__device__ __forceinline__ float neg_inf() { return __int_as_float(0xff800000); }

template <typename T> __device__ __forceinline__ float to_float(T v);
template <> __device__ __forceinline__ float to_float<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_float<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_float(float v);
template <> __device__ __forceinline__ float from_float<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

}  // namespace mdt

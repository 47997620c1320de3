// orbx_common.cuh -- shared helpers for liborbx.so (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include "../../include/orbx.h"

namespace orbx {

// thread-local last-error text behind orbx_last_error()
std::string& last_error();
void set_error(const char* fmt, ...);

#define ORBX_CUDA_TRY(call)                                                                          \
    do {                                                                                             \
        cudaError_t _e = (call);                                                                     \
        if (_e != cudaSuccess) {                                                                     \
            orbx::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e));   \
            return ORBX_ERR_CUDA;                                                                    \
        }                                                                                            \
    } while (0)

// Device plane layout: interior pixel (0,0) of a level sits at buffer row PAD_TOP, buffer column PAD_LEFT.
// PAD_LEFT = 32 keeps every interior row start 32-byte (sector) aligned because pitches are multiples of 128;
// the 19-pixel REFLECT_101 frame occupies buffer columns [13, 32) and [32+w, 32+w+19).
constexpr int EDGE = ORBX_EDGE;
constexpr int PAD_LEFT = 32;
constexpr int PAD_TOP = EDGE;

__host__ __device__ inline int reflect101(int p, int len)
{
    // valid for -len < p < 2*len-1 (callers clamp to the 19-px frame first); len >= 2
    if (p < 0) p = -p;
    if (p >= len) p = 2 * (len - 1) - p;
    return p;
}

__host__ __device__ inline int align_up(int v, int a) { return (v + a - 1) / a * a; }

}  // namespace orbx

// Host-side helpers shared by the C-ABI entry points: error capture, TMA descriptor encoding.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

namespace swh {

// Last error text, readable through sw_last_error().
char* err_buf();
int fail(int code, const char* fmt, ...);

#define SW_CUDA_OK(expr)                                                              \
    do {                                                                              \
        cudaError_t _e = (expr);                                                      \
        if (_e != cudaSuccess)                                                        \
            return swh::fail(int(_e), "%s failed: %s", #expr, cudaGetErrorString(_e)); \
    } while (0)

// Encode a tiled fp16/bf16 tensor map with SWIZZLE_128B. dims/box are innermost-first;
// strides_bytes[i] is the byte stride of dim i+1. Returns 0 on success.
int encode_tmap_f16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box);

int sm_count();

}  // namespace swh

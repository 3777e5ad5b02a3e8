#include "host_util.h"

#include <stdarg.h>
#include <stdlib.h>

#include <mutex>

namespace swh {

static thread_local char g_err[512];

char* err_buf() { return g_err; }

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code ? code : -1;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        // The runtime hands out the driver symbol, so the library never links libcuda itself.
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
                cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

int encode_tmap_f16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box) {
    return encode_tmap_f16_ex(out, base, rank, dims, strides_bytes, box, true);
}

static int encode_impl(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                       const uint64_t* strides_bytes, const uint32_t* box, bool swizzle128, const uint32_t* elem_strides);

int encode_tmap_f16_ex(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                       const uint64_t* strides_bytes, const uint32_t* box, bool swizzle128) {
    return encode_impl(out, base, rank, dims, strides_bytes, box, swizzle128, nullptr);
}

int encode_tmap_f16_strided(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                            const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* elem_strides) {
    return encode_impl(out, base, rank, dims, strides_bytes, box, true, elem_strides);
}

static int encode_impl(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                       const uint64_t* strides_bytes, const uint32_t* box, bool swizzle128, const uint32_t* elem_strides) {
    EncodeTiledFn enc = get_encode();
    if (!enc) return fail(-2, "cuTensorMapEncodeTiled entry point unavailable");
    cuuint64_t gd[5];
    cuuint64_t gs[4];
    cuuint32_t bx[5];
    cuuint32_t es[5];
    for (int i = 0; i < rank; ++i) {
        gd[i] = dims[i];
        bx[i] = box[i];
        es[i] = elem_strides ? elem_strides[i] : 1;
        if (i + 1 < rank) gs[i] = strides_bytes[i];
    }
    if ((reinterpret_cast<uintptr_t>(base) & 15) != 0)
        return fail(-3, "tensor map base %p is not 16-byte aligned", base);
    for (int i = 0; i + 1 < rank; ++i)
        if (gs[i] % 16 != 0)
            return fail(-3, "tensor map stride[%d]=%llu is not a multiple of 16 bytes", i,
                        (unsigned long long)gs[i]);
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, (cuuint32_t)rank,
                     const_cast<void*>(base), gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        return fail(-4,
                    "cuTensorMapEncodeTiled failed (%d): rank=%d dims=[%llu,%llu,%llu,%llu] "
                    "box=[%u,%u,%u,%u]",
                    int(r), rank, (unsigned long long)gd[0], (unsigned long long)(rank > 1 ? gd[1] : 0),
                    (unsigned long long)(rank > 2 ? gd[2] : 0),
                    (unsigned long long)(rank > 3 ? gd[3] : 0), bx[0], rank > 1 ? bx[1] : 0,
                    rank > 2 ? bx[2] : 0, rank > 3 ? bx[3] : 0);
    }
    return 0;
}

int sm_count() {
    static int n[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (n[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0)
            v = 148;
        n[dev] = v;
    }
    return n[dev];
}

bool pdl_enabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("SW_PDL");
        v = (e && e[0] == '0') ? 0 : 1;
    }
    return v == 1;
}

}  // namespace swh

extern "C" const char* sw_last_error() { return swh::err_buf(); }

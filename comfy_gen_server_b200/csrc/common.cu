// Host-side helpers shared by every launcher in the C-ABI library: last-error string, the
// cuTensorMapEncodeTiled entry point (fetched through the CUDA runtime, so the library does not
// link libcuda), and per-device SM count.
#include "common.cuh"

#include <stdarg.h>
#include <string.h>

namespace cgs {

static thread_local char g_last_error[512] = "";

void set_last_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
    va_end(ap);
}

EncodeTiledFn get_encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    if (fn) return fn;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
        set_last_error("cuTensorMapEncodeTiled entry point unavailable (%s)",
                       cudaGetErrorString(e));
        return nullptr;
    }
    fn = reinterpret_cast<EncodeTiledFn>(p);
    return fn;
}

int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                   const uint64_t* strides_bytes, const uint32_t* box) {
    EncodeTiledFn enc = get_encode_tiled();
    if (!enc) return kDriverEntryMissing;
    cuuint64_t gdim[5];
    cuuint64_t gstr[5];
    cuuint32_t bx[5];
    cuuint32_t estr[5];
    for (int i = 0; i < rank; ++i) {
        gdim[i] = dims[i];
        bx[i] = box[i];
        estr[i] = 1;
    }
    for (int i = 0; i + 1 < rank; ++i) {
        gstr[i] = strides_bytes[i];
        if (gstr[i] % 16 != 0) {
            set_last_error("tensor-map stride %d = %llu bytes is not a multiple of 16", i,
                           (unsigned long long)gstr[i]);
            return kBadArgument;
        }
    }
    if ((reinterpret_cast<uintptr_t>(base) & 0xF) != 0) {
        set_last_error("tensor-map base pointer %p is not 16-byte aligned", base);
        return kBadArgument;
    }
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank,
                     const_cast<void*>(base), gdim, gstr, bx, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_last_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu %llu)",
                       (int)r, rank, (unsigned long long)dims[0], (unsigned long long)dims[1]);
        return kCudaError;
    }
    return kOk;
}

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

}  // namespace cgs

extern "C" const char* cgs_last_error(void) { return cgs::g_last_error; }

extern "C" int cgs_abi_version(void) { return 1; }

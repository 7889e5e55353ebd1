// Host-side TMA tensor-map encoder (cuTensorMapEncodeTiled fetched through the runtime: no libcuda link dependency).
#include "umma.cuh"

namespace mdt {

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// The encoder is a DRIVER entry point: it needs a current context on the calling thread.  The runtime binds the primary
// context lazily, and a thread that has made no runtime call yet (autograd's backward worker calling straight into a
// C-ABI entry whose first action is to build a map) has none: CUDA_ERROR_INVALID_CONTEXT.  cudaFree(0) binds it.
static void ensure_context() {
    static thread_local bool ready = false;
    if (!ready) {
        cudaFree(nullptr);
        ready = true;
    }
}

static EncodeTiledFn encode_fn() {
    ensure_context();
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

int make_tensor_map_bf16(CUtensorMap *map, const void *base, int rank, const uint64_t *dims, const uint64_t *strides_bytes,
                         const uint32_t *box) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(MDT_ECUDA, "cuTensorMapEncodeTiled is not available from the driver");
    MDT_REQUIRE(((uintptr_t)base & 15) == 0, "TMA base pointer must be 16-byte aligned");
    cuuint64_t gdim[5], gstr[5];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) {
        gdim[i] = dims[i];
        bx[i] = box[i];
        es[i] = 1;
        if (i > 0) {
            MDT_REQUIRE(strides_bytes[i] % 16 == 0, "TMA stride %d (%llu bytes) must be a multiple of 16", i,
                        (unsigned long long)strides_bytes[i]);
            gstr[i - 1] = strides_bytes[i];
        }
    }
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void *>(base), gdim, gstr, bx, es,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(MDT_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return MDT_OK;
}

// uint8 2-D tensor (rows x pitch bytes), 128-byte swizzle, box {128 bytes, box_rows}
int make_tensor_map_u8_2d(CUtensorMap *map, const void *base, uint64_t cols, uint64_t rows, uint64_t pitch_bytes,
                          uint32_t box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(MDT_ECUDA, "cuTensorMapEncodeTiled is not available from the driver");
    MDT_REQUIRE(((uintptr_t)base & 15) == 0 && pitch_bytes % 16 == 0, "TMA needs a 16-byte aligned base and pitch");
    const cuuint64_t gdim[2] = {cols, rows}, gstr[1] = {pitch_bytes};
    const cuuint32_t bx[2] = {128, box_rows}, es[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), gdim, gstr, bx, es,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(MDT_ECUDA, "cuTensorMapEncodeTiled(u8) failed with CUresult %d", (int)r);
    return MDT_OK;
}

}  // namespace mdt

// Self-test of the tcgen05 operand layouts used by the
// attention kernels: K-major / MN-major shared-memory descriptors, TMA 128-byte swizzle,
// thread-written swizzled tiles, TMEM read-back.  Test infrastructure for tests/ only.
#include "umma.cuh"

namespace mdt {

using namespace sm100;

// smem map (bytes, 1024-aligned): A 0..16K | B 16K..32K | P 32K..64K | V 64K..80K | At 80K..112K | Bk 112K..128K
__global__ void __launch_bounds__(128, 1)
    umma_selftest_kernel(const __grid_constant__ CUtensorMap ma, const __grid_constant__ CUtensorMap mb,
                         const __grid_constant__ CUtensorMap mv, const __grid_constant__ CUtensorMap mat,
                         const __grid_constant__ CUtensorMap mbk, const __nv_bfloat16 *__restrict__ p_g, float *d1,
                         float *d2, float *d3, float *d4) {
    extern __shared__ __align__(1024) uint8_t smem_dyn[];
    uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t bar_tma, bar_mma;
    __shared__ uint32_t tmem_base_s;
    uint8_t *sA = smem, *sB = smem + 16384, *sP = smem + 32768, *sV = smem + 65536, *sAt = smem + 81920,
            *sBk = smem + 114688;
    const int tid = threadIdx.x, warp = tid >> 5;

    if (tid == 0) {
        mbar_init(&bar_tma, 1);
        mbar_init(&bar_mma, 1);
        fence_barrier_init();
    }
    if (warp == 0) {
        tmem_alloc(&tmem_base_s, 512);
        tmem_relinquish();
    }
    // P [128 x 128] row-major in global -> two K-major SW128 sub-tiles written by threads
    for (int idx = tid; idx < 128 * 16; idx += 128) {
        const int r = idx >> 4, ch = idx & 15;  // ch: 16-byte chunk (8 bf16) along the row
        const uint4 v = *reinterpret_cast<const uint4 *>(p_g + r * 128 + ch * 8);
        *reinterpret_cast<uint4 *>(sP + (ch >> 3) * 16384 + sw128_offset(r, ch & 7)) = v;
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    {   // P row `tid` -> TMEM columns 256..319 of lane `tid`: 32-bit column j holds elements 2j, 2j+1 (K order)
        uint32_t w[64];
        const uint32_t *src = reinterpret_cast<const uint32_t *>(p_g + tid * 128);
        for (int j = 0; j < 64; ++j) w[j] = src[j];
        const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
        tmem_st_32x32(lane_addr + 256, w);
        tmem_st_32x32(lane_addr + 288, w + 32);
        tmem_wait_st();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    if (tid == 0) {
        mbar_expect_tx(&bar_tma, 16384 * 3 + 32768 + 16384);
        tma_load_2d(sA, &ma, &bar_tma, 0, 0);
        tma_load_2d(sB, &mb, &bar_tma, 0, 0);
        tma_load_2d(sV, &mv, &bar_tma, 0, 0);
        tma_load_2d(sAt, &mat, &bar_tma, 0, 0);           // columns 0..63 of At   (M chunk 0)
        tma_load_2d(sAt + 16384, &mat, &bar_tma, 64, 0);  // columns 64..127       (M chunk 1)
        tma_load_2d(sBk, &mbk, &bar_tma, 0, 0);           // Bk[:, 0..63]  -> K chunk 0 (64 rows x 128 B)
        tma_load_2d(sBk + 8192, &mbk, &bar_tma, 64, 0);   // Bk[:, 64..127] -> K chunk 1
        mbar_wait(&bar_tma, 0);
        tc_fence_after();
        // test 1: D1[128,128] = A[128,64] * B[128,64]^T      (both K-major)     -> TMEM cols 0..127
        {
            const uint32_t id = idesc_bf16(128, 128, 0, 0);
            for (int k = 0; k < 4; ++k)
                umma_bf16(tmem, smem_desc_sw128(smem_u32(sA) + k * 32, 16, 1024), smem_desc_sw128(smem_u32(sB) + k * 32, 16, 1024),
                          id, k > 0);
        }
        // test 4: D4[128,64] = P[128,128] * V[128,64] with A = P read from TENSOR MEMORY (bf16 pairs packed into 32-bit
        // columns 256..319 of lane = row by the threads, tcgen05.st) -> cols 320..383
        {
            const uint32_t id = idesc_bf16(128, 64, 0, 1);
            for (int k = 0; k < 8; ++k)
                umma_bf16_ts(tmem + 320, tmem + 256 + k * 8, smem_desc_sw128(smem_u32(sV) + k * 2048, 8192, 1024), id, k > 0);
        }
        // test 2: D2[128,64] = P[128,128] * V[128,64]         (A K-major, B MN-major) -> cols 128..191
        {
            const uint32_t id = idesc_bf16(128, 64, 0, 1);
            for (int k = 0; k < 8; ++k)
                umma_bf16(tmem + 128, smem_desc_sw128(smem_u32(sP) + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                          smem_desc_sw128(smem_u32(sV) + k * 2048, 8192, 1024), id, k > 0);
        }
        // test 3: D3[128,64] = At[128(K),128(M)]^T * Bk[64(N),128(K)]^T   (A MN-major, B K-major) -> cols 192..255
        {
            const uint32_t id = idesc_bf16(128, 64, 1, 0);
            for (int k = 0; k < 8; ++k)
                umma_bf16(tmem + 192, smem_desc_sw128(smem_u32(sAt) + k * 2048, 16384, 1024),
                          smem_desc_sw128(smem_u32(sBk) + (k >> 2) * 8192 + (k & 3) * 32, 16, 1024), id, k > 0);
        }
        umma_commit(&bar_mma);
    }
    __syncwarp();
    mbar_wait(&bar_mma, 0);
    tc_fence_after();
    const int row = tid;  // TMEM lane == accumulator row
    const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
    uint32_t r[32];
    for (int c = 0; c < 4; ++c) {
        tmem_ld_32x32(lane_base + c * 32, r);
        tmem_wait_ld();
        for (int j = 0; j < 32; ++j) d1[row * 128 + c * 32 + j] = __uint_as_float(r[j]);
    }
    for (int c = 0; c < 2; ++c) {
        tmem_ld_32x32(lane_base + 128 + c * 32, r);
        tmem_wait_ld();
        for (int j = 0; j < 32; ++j) d2[row * 64 + c * 32 + j] = __uint_as_float(r[j]);
    }
    for (int c = 0; c < 2; ++c) {
        tmem_ld_32x32(lane_base + 192 + c * 32, r);
        tmem_wait_ld();
        for (int j = 0; j < 32; ++j) d3[row * 64 + c * 32 + j] = __uint_as_float(r[j]);
    }
    for (int c = 0; c < 2; ++c) {
        tmem_ld_32x32(lane_base + 320 + c * 32, r);
        tmem_wait_ld();
        for (int j = 0; j < 32; ++j) d4[row * 64 + c * 32 + j] = __uint_as_float(r[j]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace mdt

using namespace mdt;

extern "C" int mdt_umma_selftest(const void *a, const void *b, const void *p, const void *v, const void *at, const void *bk,
                                 float *d1, float *d2, float *d3, float *d4, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(a && b && p && v && at && bk && d1 && d2 && d3 && d4, "null pointer");
    CUtensorMap ma, mb, mv, mat, mbk;
    const uint32_t box64x128[2] = {64, 128}, box64x64[2] = {64, 64};
    {
        const uint64_t dims[2] = {64, 128}, str[2] = {2, 128};
        if (int rc = make_tensor_map_bf16(&ma, a, 2, dims, str, box64x128)) return rc;
        if (int rc = make_tensor_map_bf16(&mb, b, 2, dims, str, box64x128)) return rc;
        if (int rc = make_tensor_map_bf16(&mv, v, 2, dims, str, box64x128)) return rc;
    }
    {
        const uint64_t dims[2] = {128, 128}, str[2] = {2, 256};
        if (int rc = make_tensor_map_bf16(&mat, at, 2, dims, str, box64x128)) return rc;
    }
    {
        const uint64_t dims[2] = {128, 64}, str[2] = {2, 256};
        if (int rc = make_tensor_map_bf16(&mbk, bk, 2, dims, str, box64x64)) return rc;
    }
    const int smem = 131072 + 1024;
    MDT_CHECK_CUDA(cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    umma_selftest_kernel<<<1, 128, smem, as_stream(stream)>>>(ma, mb, mv, mat, mbk, (const __nv_bfloat16 *)p, d1, d2, d3, d4);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

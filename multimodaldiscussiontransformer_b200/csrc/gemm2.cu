// N2: the GEMM of gemm.cu on CTA PAIRS (tcgen05 cta_group::2):  C[M,N] = A[M,K] * W[N,K]^T (+ bias) (-> exact GELU)
//   mDT/src/modules/graphormer_graph_encoder_layer.py:132-136 (fc1 -> gelu), Hugging Face BertIntermediate / ViTIntermediate
//   called from mDT/src/modules/multi_graphormer_fusion_layer.py:94-98,138-148.
//
// Why pairs: a 128 x 256 tile per CTA (gemm.cu) re-reads its operands from L2 at 85 FLOP per byte and stops at 63 % of the tensor
// peak; a 256 x 256 tile per CTA leaves no second TMEM accumulator, so its epilogue cannot hide under the next tile.  A pair of
// CTAs (one cluster of 2, the two SMs of a TPC) computes a 256 x 256 tile with ONE tcgen05.mma.cta_group::2 stream: each CTA
// stages its own 128 rows of A and its own 128 rows of W (32 KB per 64-wide k-block instead of 48 KB), holds the 128 x 256
// half of the accumulator in its own tensor memory -- twice, so the epilogue of tile t runs under the MMAs of tile t + 1 --
// and writes its own 128 output rows.
//   warp 0      TMA producer (both CTAs): A rows / W rows of this CTA into a 6-stage ring; the bytes of BOTH CTAs complete on
//               the leader's `full` barrier (cp.async.bulk.tensor ... cta_group::2 with the leader's barrier address)
//   warp 1      MMA issuer (leader CTA only): 4 x tcgen05.mma.cta_group::2 256x256x16 per k-block; tcgen05.commit multicasts
//               the `empty` / `acc_full` arrivals to both CTAs
//   warps 2-9   epilogue (both CTAs): tensor memory -> registers -> (+ bias, GELU) -> a 128-byte-swizzled [32 rows x 64 columns]
//               staging tile per warp -> ONE TMA store per tile (row-per-thread 16-byte global stores hit L2 as quarter-filled
//               sectors: the single-CTA kernel and the first version of this one stopped at 1000 TFLOP/s on them);
//               `acc_empty` arrivals of both CTAs land on the leader's barrier
#include "umma.cuh"

namespace mdt {
namespace gemm2 {
using namespace sm100;

constexpr int kBM = 128, kBN = 256, kBK = 64, kStages = 6;   // per CTA: 128 rows of A, 128 of the tile's 256 W rows
constexpr int kStageBytes = kBM * kBK * 2 + (kBN / 2) * kBK * 2;  // 32 KB
constexpr int kOutTile = 32 * 128;                                // per epilogue warp: 32 rows x 64 bf16 columns
constexpr int kThreads = 320;
enum Epi : int { kEpiBias = 0, kEpiBiasGelu = 1, kEpiGeluBwd = 2 };
// kEpiGeluBwd: the input-gradient GEMM of the layer behind a GELU, dU = (dY W) * gelu'(U):  A = dY [M, K], W = [K, N] ROW-major
// (the forward weight of that layer as it is stored; consumed MN-major, no transpose), aux = U [M, N] (read), out = dU.

struct Ctrl {
    uint64_t full[kStages], empty[kStages], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};
constexpr int kSmem = kStages * kStageBytes + 8 * kOutTile + (int)sizeof(Ctrl);

struct Args {
    int M, N, K;
    const __nv_bfloat16 *bias;
    __nv_bfloat16 *out, *aux;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local_smem_addr` in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t mapa(uint32_t local_smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t *dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish2() { asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// TMA tile into THIS CTA's shared memory; the bytes complete on `bar_cluster_addr` (the leader's barrier)
__device__ __forceinline__ void tma_load_2d_pair(void *dst, const CUtensorMap *m, uint32_t bar_cluster_addr, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrives on the barrier at this shared-memory offset in BOTH CTAs once the MMAs issued so far have completed
__device__ __forceinline__ void umma_commit_pair(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
                 "h"((uint16_t)3)
                 : "memory");
}

// [32 rows x 64 columns] bf16 box from shared memory to (column c0, row r0) of the tensor behind `m`; rows / columns past the
// tensor's extent are clipped by the hardware
__device__ __forceinline__ void tma_store_2d(const CUtensorMap *m, const void *src, int c0, int r0) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(reinterpret_cast<uint64_t>(m)),
                 "r"(smem_u32(src)), "r"(c0), "r"(r0)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

__device__ __forceinline__ float gelu_exact(float x) {  // x * Phi(x), Abramowitz-Stegun 7.1.26 (see gelu.cu)
    const float z = fabsf(x) * 0.70710678118654752f;
    float t, e;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.f)));
    float poly = fmaf(1.061405429f, t, -1.453152027f);
    poly = fmaf(poly, t, 1.421413741f);
    poly = fmaf(poly, t, -0.284496736f);
    poly = fmaf(poly, t, 0.254829592f);
    poly *= t;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-0.72134752044448170f * x * x));
    const float half_erfc = 0.5f * poly * e;
    return x * (x >= 0.f ? 1.f - half_erfc : half_erfc);
}

// d/dx [x Phi(x)] = Phi(x) + x phi(x)
__device__ __forceinline__ float gelu_grad(float x) {
    const float z = fabsf(x) * 0.70710678118654752f;
    float t, e;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.f)));
    float poly = fmaf(1.061405429f, t, -1.453152027f);
    poly = fmaf(poly, t, 1.421413741f);
    poly = fmaf(poly, t, -0.284496736f);
    poly = fmaf(poly, t, 0.254829592f);
    poly *= t;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-0.72134752044448170f * x * x));
    const float half_erfc = 0.5f * poly * e;
    const float cdf = x >= 0.f ? 1.f - half_erfc : half_erfc;
    return fmaf(x * 0.39894228040143268f, e, cdf);
}

template <int EPI>
__global__ void __launch_bounds__(kThreads, 1)
    gemm_pair_kernel(const __grid_constant__ CUtensorMap ma, const __grid_constant__ CUtensorMap mw, const __grid_constant__ CUtensorMap mo,
                     const __grid_constant__ CUtensorMap mx, const Args a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *out_tiles = smem + kStages * kStageBytes;
    Ctrl &ctl = *reinterpret_cast<Ctrl *>(out_tiles + 8 * kOutTile);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
    const int n_tiles_n = a.N / kBN, n_tiles_m = (a.M + 2 * kBM - 1) / (2 * kBM);
    const int n_tiles = n_tiles_m * n_tiles_n, n_kb = a.K / kBK;

    if (tid == 0) {
        if (smem_u32(smem) & 1023u) __trap();
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&ctl.full[s], 1);    // used in the leader only: its own arrive.expect_tx; the bytes come from both CTAs
            mbar_init(&ctl.empty[s], 1);   // the multicast commit
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&ctl.acc_full[s], 1);     // the multicast commit
            mbar_init(&ctl.acc_empty[s], 512);  // used in the leader only: the epilogue threads of both CTAs
        }
        fence_barrier_init();
        tma_prefetch_desc(&ma);
        tma_prefetch_desc(&mw);
        tma_prefetch_desc(&mo);
        if (EPI == kEpiBiasGelu) tma_prefetch_desc(&mx);
    }
    if (warp == 1) {
        tmem_alloc2(&ctl.tmem_base, 512);
        tmem_relinquish2();
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();   // the peer's barriers are initialised and its tensor memory is allocated
    tc_fence_after();
    const uint32_t tmem = ctl.tmem_base;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t full_leader[kStages];
#pragma unroll
            for (int s = 0; s < kStages; ++s) full_leader[s] = mapa(smem_u32(&ctl.full[s]), 0);
            uint32_t g = 0;
            for (int t = pair; t < n_tiles; t += n_pairs) {
                const int m0 = (t / n_tiles_n) * 2 * kBM + (int)rank * kBM;
                const int n0 = (t % n_tiles_n) * kBN + (int)rank * (kBN / 2);
                for (int kb = 0; kb < n_kb; ++kb, ++g) {
                    const int s = g % kStages;
                    if (g >= kStages) mbar_wait(&ctl.empty[s], ((g / kStages) - 1) & 1);
                    uint8_t *st = smem + s * kStageBytes;
                    if (rank == 0) mbar_expect_tx(&ctl.full[s], 2 * kStageBytes);
                    tma_load_2d_pair(st, &ma, full_leader[s], kb * kBK, m0);
                    if (EPI == kEpiGeluBwd) {   // W [K, N] row-major: two boxes of [64 k-rows x 64 n-columns], consumed MN-major
                        tma_load_2d_pair(st + kBM * kBK * 2, &mw, full_leader[s], n0, kb * kBK);
                        tma_load_2d_pair(st + kBM * kBK * 2 + 8192, &mw, full_leader[s], n0 + 64, kb * kBK);
                    } else {
                        tma_load_2d_pair(st + kBM * kBK * 2, &mw, full_leader[s], kb * kBK, n0);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            const uint32_t idesc = idesc_bf16(2 * kBM, kBN, 0, EPI == kEpiGeluBwd ? 1 : 0);
            uint32_t g = 0;
            int it = 0;
            for (int t = pair; t < n_tiles; t += n_pairs, ++it) {
                const int acc = it & 1;
                if (it >= 2) mbar_wait(&ctl.acc_empty[acc], ((it >> 1) - 1) & 1);  // both CTAs have drained this accumulator
                tc_fence_after();
                for (int kb = 0; kb < n_kb; ++kb, ++g) {
                    const int s = g % kStages;
                    mbar_wait(&ctl.full[s], (g / kStages) & 1);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + s * kStageBytes), sb = sa + kBM * kBK * 2;
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16_pair(tmem + acc * kBN, smem_desc_sw128(sa + k * 32, 16, 1024),
                                       EPI == kEpiGeluBwd ? smem_desc_sw128(sb + k * 2048, 8192, 1024) : smem_desc_sw128(sb + k * 32, 16, 1024),
                                       idesc, (kb > 0 || k > 0) ? 1u : 0u);
                    umma_commit_pair(&ctl.empty[s]);
                }
                umma_commit_pair(&ctl.acc_full[acc]);
            }
        }
    } else {
        // ------------------------------------------------------------------ epilogue warps (both CTAs)
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        const int row = q * 32 + lane;
        const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
        const uint32_t acc_empty_leader[2] = {mapa(smem_u32(&ctl.acc_empty[0]), 0), mapa(smem_u32(&ctl.acc_empty[1]), 0)};
        uint8_t *tile = out_tiles + (warp - 2) * kOutTile;      // this warp's staging tile: row = lane, 8 chunks of 16 bytes
        int it = 0;
        for (int t = pair; t < n_tiles; t += n_pairs, ++it) {
            const int acc = it & 1;
            const int m0 = (t / n_tiles_n) * 2 * kBM + (int)rank * kBM, n0 = (t % n_tiles_n) * kBN + half * 128;
            mbar_wait(&ctl.acc_full[acc], (it >> 1) & 1);
            tc_fence_after();
            __syncwarp();
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {          // 64 output columns per pass
                float v[64];
                tmem_ld_32x32(lane_base + acc * kBN + half * 128 + c * 64, reinterpret_cast<uint32_t *>(v));
                tmem_ld_32x32(lane_base + acc * kBN + half * 128 + c * 64 + 32, reinterpret_cast<uint32_t *>(v) + 32);
                tmem_wait_ld();
                const int n = n0 + c * 64;
                if (EPI == kEpiGeluBwd) {   // times gelu'(u): this thread's row of U, 128 contiguous bytes (L1 keeps the line)
                    const int m = m0 + row;
                    if (m < a.M) {
                        const uint4 *up = reinterpret_cast<const uint4 *>(a.aux + (size_t)m * a.N + n);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const uint4 uw = __ldg(up + j);
                            const uint32_t uu[4] = {uw.x, uw.y, uw.z, uw.w};
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                v[j * 8 + 2 * e] *= gelu_grad(__uint_as_float(uu[e] << 16));
                                v[j * 8 + 2 * e + 1] *= gelu_grad(__uint_as_float(uu[e] & 0xffff0000u));
                            }
                        }
                    }
                } else if (a.bias != nullptr) {
                    const uint4 *bp = reinterpret_cast<const uint4 *>(a.bias + n);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const uint4 bw = __ldg(bp + j);
                        const __nv_bfloat162 *b2 = reinterpret_cast<const __nv_bfloat162 *>(&bw);
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float2 f = __bfloat1622float2(b2[e]);
                            v[j * 8 + 2 * e] += f.x;
                            v[j * 8 + 2 * e + 1] += f.y;
                        }
                    }
                }
                uint32_t w[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) w[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
                tma_store_wait_read();             // the previous store has finished reading the staging tile
                __syncwarp();
                if (EPI == kEpiBiasGelu) {
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        *reinterpret_cast<uint4 *>(tile + sw128_offset(lane, j)) = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
                    fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
                        tma_store_2d(&mx, tile, n, m0 + q * 32);
                        tma_store_commit();
                    }
#pragma unroll
                    for (int j = 0; j < 32; ++j) {   // gelu of the ROUNDED pre-activation: what the unfused path computes
                        const float x0 = __uint_as_float(w[j] << 16), x1 = __uint_as_float(w[j] & 0xffff0000u);
                        w[j] = pack_bf16x2(gelu_exact(x0), gelu_exact(x1));
                    }
                    tma_store_wait_read();
                    __syncwarp();
                }
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<uint4 *>(tile + sw128_offset(lane, j)) = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
                    tma_store_2d(&mo, tile, n, m0 + q * 32);
                    tma_store_commit();
                }
            }
            tc_fence_before();
            mbar_arrive_cluster(acc_empty_leader[acc]);
        }
        tma_store_wait_read();
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();   // the leader's MMAs read the peer's shared memory / write its tensor memory: leave together
    if (warp == 1) tmem_dealloc2(tmem, 512);
}

}  // namespace gemm2
}  // namespace mdt

using namespace mdt;

// Same contract as mdt_gemm_bias_act (gemm.cu); called from there when the shape qualifies and MDT_GEMM_PAIR is not 0.
int mdt_gemm_pair_launch(const void *a, int64_t lda, const void *w, int64_t ldw, const void *bias, void *out, void *aux, int M, int N,
                         int K, int act, cudaStream_t stream) {
    CUtensorMap ma, mw;
    {
        const uint64_t dims[2] = {(uint64_t)K, (uint64_t)M}, str[2] = {2, (uint64_t)lda * 2};
        const uint32_t box[2] = {(uint32_t)gemm2::kBK, (uint32_t)gemm2::kBM};
        if (int rc = make_tensor_map_bf16(&ma, a, 2, dims, str, box)) return rc;
    }
    if (act == 2) {   // W [K, N] row-major
        const uint64_t dims[2] = {(uint64_t)N, (uint64_t)K}, str[2] = {2, (uint64_t)ldw * 2};
        const uint32_t box[2] = {64, 64};
        if (int rc = make_tensor_map_bf16(&mw, w, 2, dims, str, box)) return rc;
    } else {
        const uint64_t dims[2] = {(uint64_t)K, (uint64_t)N}, str[2] = {2, (uint64_t)ldw * 2};
        const uint32_t box[2] = {(uint32_t)gemm2::kBK, (uint32_t)(gemm2::kBN / 2)};
        if (int rc = make_tensor_map_bf16(&mw, w, 2, dims, str, box)) return rc;
    }
    CUtensorMap mo, mx;
    {
        const uint64_t dims[2] = {(uint64_t)N, (uint64_t)M}, str[2] = {2, (uint64_t)N * 2};
        const uint32_t box[2] = {64, 32};
        if (int rc = make_tensor_map_bf16(&mo, out, 2, dims, str, box)) return rc;
        if (int rc = make_tensor_map_bf16(&mx, (aux && act == 1) ? aux : out, 2, dims, str, box)) return rc;
    }
    gemm2::Args g{};
    g.M = M, g.N = N, g.K = K;
    g.bias = reinterpret_cast<const __nv_bfloat16 *>(bias);
    g.out = reinterpret_cast<__nv_bfloat16 *>(out), g.aux = reinterpret_cast<__nv_bfloat16 *>(aux);
    const int tiles = ((M + 2 * gemm2::kBM - 1) / (2 * gemm2::kBM)) * (N / gemm2::kBN);
    void (*kern)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap, const gemm2::Args) =
        act == 2 ? gemm2::gemm_pair_kernel<gemm2::kEpiGeluBwd>
                 : (act == 1 ? gemm2::gemm_pair_kernel<gemm2::kEpiBiasGelu> : gemm2::gemm_pair_kernel<gemm2::kEpiBias>);
    MDT_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm2::kSmem));
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2, attr[0].val.clusterDim.y = 1, attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr, cfg.numAttrs = 1;
    cfg.blockDim = dim3(gemm2::kThreads);
    cfg.dynamicSmemBytes = gemm2::kSmem;
    cfg.stream = stream;
    // pairs that can be resident at once (some GPCs hold an odd number of SMs)
    static int max_pairs[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (max_pairs[dev] == 0) {
        cfg.gridDim = dim3(2 * (num_sms() / 2));
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n <= 0) n = num_sms() / 2 - 4;
        max_pairs[dev] = n;
    }
    const int pairs = tiles < max_pairs[dev] ? tiles : max_pairs[dev];
    cfg.gridDim = dim3(2 * pairs);
    MDT_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mw, mo, mx, g));
    return MDT_OK;
}

// dU[M, N] = (dY[M, K] * W[K, N]) * gelu'(U[M, N]): the input-gradient GEMM of the dense layer behind an exact GELU with the
// activation's derivative in its epilogue (no separate pass over [M, N]).  bf16; W row-major [K, N] (the layer's weight as
// stored: out_features x in_features of the layer AFTER the activation); N % 256 == 0, K % 64 == 0.
extern "C" int mdt_gemm_dgrad_gelu(const void *dy, int64_t ldy, const void *w, int64_t ldw, const void *u, void *du, int M, int N, int K,
                                   mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(dy && w && u && du && M >= 1, "mdt_gemm_dgrad_gelu: null pointer or M < 1");
    MDT_REQUIRE(N % gemm2::kBN == 0 && K % gemm2::kBK == 0 && K >= gemm2::kBK, "mdt_gemm_dgrad_gelu: N must be a multiple of 256 and K of 64 (got N=%d K=%d)", N, K);
    MDT_REQUIRE(ldy % 8 == 0 && ldw % 8 == 0 && ldy >= K && ldw >= N, "mdt_gemm_dgrad_gelu: leading dimensions must be multiples of 8, ldy >= K, ldw >= N");
    MDT_REQUIRE((((uintptr_t)dy | (uintptr_t)w | (uintptr_t)u | (uintptr_t)du) & 15) == 0, "mdt_gemm_dgrad_gelu: buffers must be 16-byte aligned");
    return mdt_gemm_pair_launch(dy, ldy, w, ldw, nullptr, du, const_cast<void *>(u), M, N, K, 2, as_stream(stream));
}

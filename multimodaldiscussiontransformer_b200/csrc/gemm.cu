// N2: bf16 GEMM with fused epilogues for the feed-forward blocks  C[M,N] = A[M,K] * W[N,K]^T (+ bias) (-> exact GELU)
//   mDT/src/modules/graphormer_graph_encoder_layer.py:132-136 (fc1 -> gelu), Hugging Face BertIntermediate / ViTIntermediate
//   called from mDT/src/modules/multi_graphormer_fusion_layer.py:94-98,138-148.
//
// Persistent, warp-specialised tcgen05 kernel, one CTA per SM, 128 x 256 output tiles (N fastest, so the CTAs of a wave share
// an A row band and W stays in L2):
//   warp 0      TMA producer: per 64-wide k-block the A tile (128 x 64, 16 KB) and the W tile (256 x 64, 32 KB) into a
//               4-stage ring (128-byte swizzle, both operands K-major)
//   warp 1      MMA issuer: 4 x tcgen05.mma 128x256x16 per k-block into one of TWO 256-column TMEM accumulators; the tensor
//               pipe starts tile t+1 while the epilogue warps drain tile t
//   warps 2-9   epilogue: thread = (row of the tile = TMEM lane, 128-column half): tcgen05.ld 32 columns at a time,
//               + bias, [round to bf16 = the pre-activation `aux`, exact GELU in fp32], bf16 stores of 64 contiguous bytes
// Epilogue modes: kEpiBias (out = acc + bias) and kEpiBiasGelu (aux = acc + bias, out = gelu(aux): the activation costs one
// extra write of `out` instead of a separate read + write pass over [M, N]).
// Shared memory 4 x 48 KB; TMEM 512 columns; rows past M are zero-filled by TMA on the way in and masked on the way out.
#include <cstdlib>

#include "umma.cuh"

namespace mdt {
namespace gemm {
using namespace sm100;

constexpr int kBM = 128, kBN = 256, kBK = 64, kStages = 4;
constexpr int kStageBytes = kBM * kBK * 2 + kBN * kBK * 2;  // 48 KB
constexpr int kThreads = 320;                               // TMA, MMA, 8 epilogue warps
enum Epi : int { kEpiBias = 0, kEpiBiasGelu = 1 };

struct Ctrl {
    uint64_t full[kStages], empty[kStages], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};
constexpr int kSmem = kStages * kStageBytes + (int)sizeof(Ctrl);

struct Args {
    int M, N, K;
    const __nv_bfloat16 *bias;  // [N] or null
    __nv_bfloat16 *out, *aux;   // [M, N] row-major (ld = N); aux only for kEpiBiasGelu
};

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

template <int EPI>
__global__ void __launch_bounds__(kThreads, 1)
    gemm_bf16_kernel(const __grid_constant__ CUtensorMap ma, const __grid_constant__ CUtensorMap mw, const Args a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    Ctrl &ctl = *reinterpret_cast<Ctrl *>(smem + kStages * kStageBytes);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_tiles_n = a.N / kBN, n_tiles_m = (a.M + kBM - 1) / kBM;
    const int n_tiles = n_tiles_m * n_tiles_n, n_kb = a.K / kBK;

    if (tid == 0) {
        if (smem_u32(smem) & 1023u) __trap();
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&ctl.full[s], 1);
            mbar_init(&ctl.empty[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&ctl.acc_full[s], 1);
            mbar_init(&ctl.acc_empty[s], 256);
        }
        fence_barrier_init();
        tma_prefetch_desc(&ma);
        tma_prefetch_desc(&mw);
    }
    if (warp == 1) {
        tmem_alloc(&ctl.tmem_base, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctl.tmem_base;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t g = 0;  // k-block sequence number across tiles
            for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                const int m0 = (t / n_tiles_n) * kBM, n0 = (t % n_tiles_n) * kBN;
                for (int kb = 0; kb < n_kb; ++kb, ++g) {
                    const int s = g % kStages;
                    if (g >= kStages) mbar_wait(&ctl.empty[s], ((g / kStages) - 1) & 1);
                    uint8_t *st = smem + s * kStageBytes;
                    mbar_expect_tx(&ctl.full[s], kStageBytes);
                    tma_load_2d(st, &ma, &ctl.full[s], kb * kBK, m0);
                    tma_load_2d(st + kBM * kBK * 2, &mw, &ctl.full[s], kb * kBK, n0);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = idesc_bf16(kBM, kBN, 0, 0);
            uint32_t g = 0;
            int it = 0;
            for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
                const int acc = it & 1;
                if (it >= 2) mbar_wait(&ctl.acc_empty[acc], ((it >> 1) - 1) & 1);  // the epilogue has drained this accumulator
                tc_fence_after();
                for (int kb = 0; kb < n_kb; ++kb, ++g) {
                    const int s = g % kStages;
                    mbar_wait(&ctl.full[s], (g / kStages) & 1);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + s * kStageBytes), sb = sa + kBM * kBK * 2;
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_bf16(tmem + acc * kBN, smem_desc_sw128(sa + k * 32, 16, 1024), smem_desc_sw128(sb + k * 32, 16, 1024), idesc,
                                  (kb > 0 || k > 0) ? 1u : 0u);
                    umma_commit(&ctl.empty[s]);  // the stage may be refilled once these MMAs have read it
                }
                umma_commit(&ctl.acc_full[acc]);
            }
        }
    } else {
        // ------------------------------------------------------------------ epilogue warps
        const int q = warp & 3;            // TMEM lane quarter this warp may read
        const int half = (warp - 2) >> 2;  // columns [128 half, 128 half + 128) of the tile
        const int row = q * 32 + lane;
        const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
        int it = 0;
        for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
            const int acc = it & 1;
            const int m0 = (t / n_tiles_n) * kBM, n0 = (t % n_tiles_n) * kBN + half * 128;
            const int m = m0 + row;
            mbar_wait(&ctl.acc_full[acc], (it >> 1) & 1);
            tc_fence_after();
            __syncwarp();
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                float v[32];
                tmem_ld_32x32(lane_base + acc * kBN + half * 128 + c * 32, reinterpret_cast<uint32_t *>(v));
                tmem_wait_ld();
                const int n = n0 + c * 32;
                if (a.bias != nullptr) {
                    const uint4 *bp = reinterpret_cast<const uint4 *>(a.bias + n);
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
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
                uint32_t w[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) w[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
                if (EPI == kEpiBiasGelu) {
                    if (m < a.M) {
                        uint4 *pa = reinterpret_cast<uint4 *>(a.aux + (size_t)m * a.N + n);
#pragma unroll
                        for (int j = 0; j < 4; ++j) pa[j] = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
                    }
#pragma unroll
                    for (int j = 0; j < 16; ++j) {   // gelu of the ROUNDED pre-activation: what the unfused path computes
                        const float x0 = __uint_as_float(w[j] << 16), x1 = __uint_as_float(w[j] & 0xffff0000u);
                        w[j] = pack_bf16x2(gelu_exact(x0), gelu_exact(x1));
                    }
                }
                if (m < a.M) {
                    uint4 *po = reinterpret_cast<uint4 *>(a.out + (size_t)m * a.N + n);
#pragma unroll
                    for (int j = 0; j < 4; ++j) po[j] = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
                }
            }
            tc_fence_before();
            mbar_arrive(&ctl.acc_empty[acc]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, 512);
}

}  // namespace gemm
}  // namespace mdt

using namespace mdt;

int mdt_gemm_pair_launch(const void *a, int64_t lda, const void *w, int64_t ldw, const void *bias, void *out, void *aux, int M, int N,
                         int K, int act, cudaStream_t stream);   // gemm2.cu

// out[M,N] = a[M,K] * w[N,K]^T + bias, optionally followed by the exact GELU (then aux receives the pre-activation).
// bf16 row-major tensors with the given leading dimensions of a and w (elements); out / aux are dense [M,N].
extern "C" int mdt_gemm_bias_act(const void *a, int64_t lda, const void *w, int64_t ldw, const void *bias, void *out, void *aux,
                                 int M, int N, int K, int act, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(a && w && out && M >= 1, "mdt_gemm_bias_act: a, w, out must be non-null and M >= 1");
    MDT_REQUIRE(N % gemm::kBN == 0 && K % gemm::kBK == 0 && K >= gemm::kBK, "mdt_gemm_bias_act: N must be a multiple of 256 and K of 64 (got N=%d K=%d)", N, K);
    MDT_REQUIRE(act == 0 || (act == 1 && aux != nullptr), "mdt_gemm_bias_act: act is 0 (bias) or 1 (bias + GELU, needs aux)");
    MDT_REQUIRE(lda % 8 == 0 && ldw % 8 == 0 && lda >= K && ldw >= K, "mdt_gemm_bias_act: leading dimensions must be multiples of 8 and >= K");
    MDT_REQUIRE((((uintptr_t)out | (uintptr_t)aux | (uintptr_t)bias) & 15) == 0, "mdt_gemm_bias_act: out, aux, bias must be 16-byte aligned");
    // CTA-pair kernel (gemm2.cu): MDT_GEMM_PAIR=0 keeps the single-CTA kernel below (A/B runs)
    static const bool use_pair = []() { const char *e = getenv("MDT_GEMM_PAIR"); return !(e && e[0] == '0'); }();
    if (use_pair) return mdt_gemm_pair_launch(a, lda, w, ldw, bias, out, aux, M, N, K, act, as_stream(stream));
    CUtensorMap ma, mw;
    {
        const uint64_t dims[2] = {(uint64_t)K, (uint64_t)M}, str[2] = {2, (uint64_t)lda * 2};
        const uint32_t box[2] = {(uint32_t)gemm::kBK, (uint32_t)gemm::kBM};
        if (int rc = make_tensor_map_bf16(&ma, a, 2, dims, str, box)) return rc;
    }
    {
        const uint64_t dims[2] = {(uint64_t)K, (uint64_t)N}, str[2] = {2, (uint64_t)ldw * 2};
        const uint32_t box[2] = {(uint32_t)gemm::kBK, (uint32_t)gemm::kBN};
        if (int rc = make_tensor_map_bf16(&mw, w, 2, dims, str, box)) return rc;
    }
    gemm::Args g{};
    g.M = M, g.N = N, g.K = K;
    g.bias = reinterpret_cast<const __nv_bfloat16 *>(bias);
    g.out = reinterpret_cast<__nv_bfloat16 *>(out), g.aux = reinterpret_cast<__nv_bfloat16 *>(aux);
    const int tiles = ((M + gemm::kBM - 1) / gemm::kBM) * (N / gemm::kBN);
    const int grid = tiles < num_sms() ? tiles : num_sms();
    if (act == 1) {
        MDT_CHECK_CUDA(cudaFuncSetAttribute(gemm::gemm_bf16_kernel<gemm::kEpiBiasGelu>, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm::kSmem));
        gemm::gemm_bf16_kernel<gemm::kEpiBiasGelu><<<grid, gemm::kThreads, gemm::kSmem, as_stream(stream)>>>(ma, mw, g);
    } else {
        MDT_CHECK_CUDA(cudaFuncSetAttribute(gemm::gemm_bf16_kernel<gemm::kEpiBias>, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm::kSmem));
        gemm::gemm_bf16_kernel<gemm::kEpiBias><<<grid, gemm::kThreads, gemm::kSmem, as_stream(stream)>>>(ma, mw, g);
    }
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

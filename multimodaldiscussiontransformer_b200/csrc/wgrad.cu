// N2: weight-gradient GEMM of the dense layers,  dW[n][c] (+)= sum_m dY[m][n] * X[m][c]   (torch: dY.t() @ X)
//   Hugging Face BertSelfOutput / BertOutput / BertSelfAttention dense layers called from
//   mDT/src/modules/multi_graphormer_fusion_layer.py:94-98,138-148; graphormer_graph_encoder_layer.py:124-141 (backward of them).
//
// The reduction runs over the ROWS of both operands (m = 238 806 packed tokens in a cfg3 step) while the output is tiny
// (768 x 768 .. 2304 x 768).  cuBLAS runs these as 128 x 128 two-CTA tiles at 620 TFLOP/s (39 % of the bf16 peak; 285 TFLOP/s on
// the ViT's 82 410 rows): every operand byte is re-read from L2 by 6 tiles, 85 FLOP per L2 byte against the ~200 the tensor pipe
// needs.  Here a CTA owns a 256 x 256 output tile (128 FLOP per L2 byte) and a slice of the rows (split-K: tiles x splits ~ one
// CTA per SM), and both operands are consumed MN-MAJOR straight from their row-major tensors -- no transposes:
//   warp 0     TMA producer: per 64-row block 4 + 4 boxes of [64 rows x 64 columns] (128-byte swizzle), 64 KB per stage, 3 stages
//   warp 1     MMA issuer: per block 4 k-steps x 2 tcgen05.mma 128 x 256 x 16 (A = dY columns as M, B = X columns as N, both
//              MN-major) into two 256-column TMEM accumulators (all 512 columns)
//   warps 2-9  epilogue, once per CTA: TMEM -> registers -> red.global.add.v4.f32 into the fp32 output (the caller zeroes it;
//              the splits of a tile meet there).  While the main loop runs, two of them (in the CTAs of the first column tile)
//              optionally add up the dY columns from the staged tiles: colsum[n] += sum_m dY[m][n], the bias gradient of the
//              same dense layer, without another pass over dY
// Rows past M are zero-filled by TMA and add nothing.
#include "umma.cuh"

namespace mdt {
namespace wgrad {
using namespace sm100;

constexpr int kT = 256;                       // output tile edge
constexpr int kBK = 64, kStages = 3;
constexpr int kAtom = 64 * 128;               // one [64 rows x 64 columns] box: 8 KB
constexpr int kOperand = 4 * kAtom;           // 256 columns of one operand: 32 KB
constexpr int kStageBytes = 2 * kOperand;     // 64 KB
constexpr int kThreads = 320;

struct Ctrl {
    uint64_t full[kStages], empty[kStages], acc_full;
    uint32_t tmem_base;
};
constexpr int kSmem = kStages * kStageBytes + (int)sizeof(Ctrl);

struct Args {
    int M, N, C;        // rows of both operands, columns of dY (= rows of dW), columns of X (= columns of dW)
    int splits, kb_per_split;
    float *out;         // [N, ldo] fp32, accumulated into
    int64_t ldo;
    float *colsum;      // [N] fp32 or null: += column sums of dy
};

__global__ void __launch_bounds__(kThreads, 1)
    wgrad_kernel(const __grid_constant__ CUtensorMap my, const __grid_constant__ CUtensorMap mx, const Args a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    Ctrl &ctl = *reinterpret_cast<Ctrl *>(smem + kStages * kStageBytes);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int tiles_c = a.C / kT;
    const int tile = blockIdx.x / a.splits, split = blockIdx.x % a.splits;
    const int n0 = (tile / tiles_c) * kT, c0 = (tile % tiles_c) * kT;
    const int n_kb_all = (a.M + kBK - 1) / kBK;
    const int kb0 = split * a.kb_per_split;
    const int kb1 = min(n_kb_all, kb0 + a.kb_per_split);
    const int n_kb = max(kb1 - kb0, 0);

    const bool do_colsum = a.colsum != nullptr && c0 == 0;   // CTA-uniform
    if (tid == 0) {
        if (smem_u32(smem) & 1023u) __trap();
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&ctl.full[s], 1);
            mbar_init(&ctl.empty[s], do_colsum ? 3 : 1);   // the MMA commit (+ the two column-sum warps)
        }
        mbar_init(&ctl.acc_full, 1);
        fence_barrier_init();
        tma_prefetch_desc(&my);
        tma_prefetch_desc(&mx);
    }
    if (warp == 1) {
        tmem_alloc(&ctl.tmem_base, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctl.tmem_base;

    if (n_kb > 0) {   // uniform: a split past the last row block has nothing to add
        if (warp == 0) {
            if (lane == 0) {
                for (int g = 0; g < n_kb; ++g) {
                    const int s = g % kStages;
                    if (g >= kStages) mbar_wait(&ctl.empty[s], ((g / kStages) - 1) & 1);
                    uint8_t *st = smem + s * kStageBytes;
                    const int m0 = (kb0 + g) * kBK;
                    mbar_expect_tx(&ctl.full[s], kStageBytes);
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        tma_load_2d(st + j * kAtom, &my, &ctl.full[s], n0 + 64 * j, m0);
                        tma_load_2d(st + kOperand + j * kAtom, &mx, &ctl.full[s], c0 + 64 * j, m0);
                    }
                }
            }
        } else if (warp == 1) {
            if (lane == 0) {
                const uint32_t idesc = idesc_bf16(128, kT, 1, 1);
                for (int g = 0; g < n_kb; ++g) {
                    const int s = g % kStages;
                    mbar_wait(&ctl.full[s], (g / kStages) & 1);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + s * kStageBytes), sb = sa + kOperand;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {       // 16 rows (K) per step: 16 x 128 B inside every atom
                        const uint64_t db = smem_desc_sw128(sb + k * 2048, kAtom, 1024);
#pragma unroll
                        for (int h = 0; h < 2; ++h)     // dY columns [128 h, 128 h + 128) of the tile = atoms 2h, 2h + 1
                            umma_bf16(tmem + h * kT, smem_desc_sw128(sa + h * 2 * kAtom + k * 2048, kAtom, 1024), db, idesc,
                                      (g > 0 || k > 0) ? 1u : 0u);
                    }
                    umma_commit(&ctl.empty[s]);
                }
                umma_commit(&ctl.acc_full);
            }
        } else {
            if (do_colsum && warp < 4) {
                // warps 2, 3: thread t owns dY columns [4 t, 4 t + 4) of the tile; 64 rows per stage as 8-byte reads
                const int t = tid - 64;
                const uint32_t atom = (uint32_t)(t >> 4) * kAtom, c16 = (uint32_t)(t & 15) >> 1, sub = (uint32_t)(t & 1) * 8;
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
                for (int g = 0; g < n_kb; ++g) {
                    const int s = g % kStages;
                    mbar_wait(&ctl.full[s], (g / kStages) & 1);
                    const uint32_t base = smem_u32(smem + s * kStageBytes) + atom + sub;
#pragma unroll 8
                    for (int r = 0; r < kBK; ++r) {
                        uint32_t w0, w1;
                        asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(w0), "=r"(w1) : "r"(base + r * 128 + ((c16 ^ (r & 7)) << 4)));
                        acc[0] += __uint_as_float(w0 << 16), acc[1] += __uint_as_float(w0 & 0xffff0000u);
                        acc[2] += __uint_as_float(w1 << 16), acc[3] += __uint_as_float(w1 & 0xffff0000u);
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&ctl.empty[s]);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) atomicAdd(a.colsum + n0 + 4 * t + i, acc[i]);
            }
            const int q = warp & 3;              // TMEM lane quarter this warp may read
            const int h = (warp - 2) >> 2;       // accumulator (128-row half of the tile)
            const int n = n0 + h * 128 + q * 32 + lane;
            const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16) + h * kT;
            float *orow = a.out + (size_t)n * a.ldo + c0;
            mbar_wait(&ctl.acc_full, 0);
            tc_fence_after();
            __syncwarp();
#pragma unroll 1
            for (int c = 0; c < kT / 32; ++c) {
                float v[32];
                tmem_ld_32x32(lane_base + c * 32, reinterpret_cast<uint32_t *>(v));
                tmem_wait_ld();
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    atomicAdd(reinterpret_cast<float4 *>(orow + c * 32) + j, make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]));
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, 512);
}

}  // namespace wgrad
}  // namespace mdt

using namespace mdt;

// out[N, C] (fp32, leading dimension ldo, ACCUMULATED into: the caller zeroes it) += dy[M, N]^T * x[M, C]; dy, x bf16 row-major
// with leading dimensions ldy, ldx (elements, multiples of 8).  N and C multiples of 256.  colsum (fp32 [N], may be null) +=
// the column sums of dy.
extern "C" int mdt_wgrad_gemm(const void *dy, int64_t ldy, const void *x, int64_t ldx, float *out, int64_t ldo, float *colsum,
                              int64_t M, int N, int C, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(dy && x && out && M >= 1 && M < (int64_t)1 << 31, "mdt_wgrad_gemm: dy, x, out must be non-null and 1 <= M < 2^31");
    MDT_REQUIRE(N >= wgrad::kT && C >= wgrad::kT && N % wgrad::kT == 0 && C % wgrad::kT == 0,
                "mdt_wgrad_gemm: N and C must be multiples of 256 (got N=%d C=%d)", N, C);
    MDT_REQUIRE(ldy % 8 == 0 && ldx % 8 == 0 && ldy >= N && ldx >= C && ldo % 4 == 0 && ldo >= C,
                "mdt_wgrad_gemm: leading dimensions: ldy, ldx multiples of 8, ldo of 4, each >= its row length");
    MDT_REQUIRE((((uintptr_t)dy | (uintptr_t)x | (uintptr_t)out) & 15) == 0, "mdt_wgrad_gemm: buffers must be 16-byte aligned");
    CUtensorMap my, mx;
    const uint32_t box[2] = {64, 64};
    {
        const uint64_t dims[2] = {(uint64_t)N, (uint64_t)M}, str[2] = {2, (uint64_t)ldy * 2};
        if (int rc = make_tensor_map_bf16(&my, dy, 2, dims, str, box)) return rc;
    }
    {
        const uint64_t dims[2] = {(uint64_t)C, (uint64_t)M}, str[2] = {2, (uint64_t)ldx * 2};
        if (int rc = make_tensor_map_bf16(&mx, x, 2, dims, str, box)) return rc;
    }
    wgrad::Args g{};
    g.M = (int)M, g.N = N, g.C = C, g.out = out, g.ldo = ldo, g.colsum = colsum;
    const int tiles = (N / wgrad::kT) * (C / wgrad::kT);
    const int n_kb = (int)((M + wgrad::kBK - 1) / wgrad::kBK);
    int splits = num_sms() / tiles;
    if (splits < 1) splits = 1;
    if (splits > n_kb) splits = n_kb;
    g.kb_per_split = (n_kb + splits - 1) / splits;
    g.splits = (n_kb + g.kb_per_split - 1) / g.kb_per_split;   // no empty splits
    MDT_CHECK_CUDA(cudaFuncSetAttribute(wgrad::wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, wgrad::kSmem));
    wgrad::wgrad_kernel<<<tiles * g.splits, wgrad::kThreads, wgrad::kSmem, as_stream(stream)>>>(my, mx, g);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

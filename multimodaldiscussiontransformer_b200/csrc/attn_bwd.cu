// A6 backward of O = softmax(Q K^T + bias) V                       multihead_attention.py:168-199
//
//   delta[q] = sum_c dO[q,c] O[q,c]                                  (attn_delta_kernel)
//   one CTA per (batch b, head h, 128-key tile); 256 threads: threads r and r+128 own key row r (= TMEM
//   lane r) and split the query columns of every tile between them; loop over
//   128-query tiles i (Q_i, dO_i double-buffered by TMA):
//     S^T  = K Q_i^T          tcgen05.mma 128x128x64                       -> TMEM cols [0,128)
//     dP^T = V dO_i^T         tcgen05.mma 128x128x64                       -> TMEM cols [128,256)
//     P^T  = exp(S^T + bias^T - lse[q]);  dS^T = P^T * (dP^T - delta[q])    (registers, fp32)
//     P^T, dS^T -> bf16 swizzled tiles [key][query]
//     dV += P^T dO_i          tcgen05.mma 128x64x128 (B = dO_i, MN-major)  -> cols [256,320)
//     dK += dS^T Q_i          tcgen05.mma 128x64x128 (B = Q_i,  MN-major)  -> cols [320,384)
//     dQ_i = dS K             tcgen05.mma 128x64x128 (A = dS^T tile read MN-major, B = K MN-major)
//                                                                          -> cols [384,448)
//     dQ_i is reduced across key tiles with fp32 vector atomics into dq_accum (attn_dq_convert_kernel
//     turns it into bf16).  d-bias is either stored ([B,H,T,T], coalesced along keys) or binned by
//     bias code into per-thread shared-memory bins (conflict free, no atomics) -> grad_lut[h][code].
#include <stdlib.h>

#include "attn_common.cuh"

namespace mdt {
namespace attn {
using namespace sm100;

constexpr int kBins = 24;       // bias codes 0..22 have their own bin (the collate kernel emits <= 21), MDT_CODE_VIRTUAL uses bin 23
constexpr int kBwdThreads = 256;  // two warpgroups: each owns 64 of the 128 query columns of a tile

struct BwdArgs {
    int B, H, T;
    int Tmain;  // keys [0, Tmain) are this kernel's (all T queries); keys [Tmain, T) belong to attn_bwd_tail_k_kernel (attn_tail.cu)
    const void *bias;
    const uint8_t *code;
    int code_pitch;
    const float *lut;
    const uint8_t *key_pad;
    const float *lse;
    const float *delta;
    float *dq_accum;
    __nv_bfloat16 *dk, *dv;
    int64_t dk_sb, dk_st, dk_sh, dv_sb, dv_st, dv_sh;
    void *grad_bias;
    float *grad_lut;
    DropArgs drop;
};

// ---------------------------------------------------------------- delta = rowsum(dO * O)
__global__ void attn_delta_kernel(const __nv_bfloat16 *__restrict__ o, const __nv_bfloat16 *__restrict__ go, int B, int H, int T,
                                  int64_t o_sb, int64_t o_st, int64_t o_sh, int64_t g_sb, int64_t g_st, int64_t g_sh,
                                  float *__restrict__ delta) {
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 3) + (threadIdx.x >> 3);  // (b, h, t) flattened
    const int part = threadIdx.x & 7;
    const int64_t total = (int64_t)B * H * T;
    float acc = 0.f;
    if (row < total) {
        const int t = (int)(row % T);
        const int h = (int)((row / T) % H);
        const int b = (int)(row / ((int64_t)T * H));
        const uint4 a = *reinterpret_cast<const uint4 *>(o + b * o_sb + t * o_st + h * o_sh + part * 8);
        const uint4 g = *reinterpret_cast<const uint4 *>(go + b * g_sb + t * g_st + h * g_sh + part * 8);
        const __nv_bfloat162 *pa = reinterpret_cast<const __nv_bfloat162 *>(&a);
        const __nv_bfloat162 *pg = reinterpret_cast<const __nv_bfloat162 *>(&g);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float2 x = __bfloat1622float2(pa[i]), y = __bfloat1622float2(pg[i]);
            acc = fmaf(x.x, y.x, acc);
            acc = fmaf(x.y, y.y, acc);
        }
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    acc += __shfl_xor_sync(0xffffffffu, acc, 4);
    if (row < total && part == 0) delta[row] = acc;
}

// ---------------------------------------------------------------- dq_accum (f32 [B,H,T,64]) -> dq (bf16, strided)
__global__ void attn_dq_convert_kernel(const float *__restrict__ acc, int B, int H, int T, __nv_bfloat16 *__restrict__ dq,
                                       int64_t sb, int64_t st, int64_t sh) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // one thread per 8 elements
    const int64_t total = (int64_t)B * H * T * 8;
    if (idx >= total) return;
    const int part = (int)(idx & 7);
    const int64_t row = idx >> 3;
    const int t = (int)(row % T);
    const int h = (int)((row / T) % H);
    const int b = (int)(row / ((int64_t)T * H));
    const float4 x = *reinterpret_cast<const float4 *>(acc + row * 64 + part * 8);
    const float4 y = *reinterpret_cast<const float4 *>(acc + row * 64 + part * 8 + 4);
    uint4 v;
    v.x = pack_bf16x2(x.x, x.y);
    v.y = pack_bf16x2(x.z, x.w);
    v.z = pack_bf16x2(y.x, y.y);
    v.w = pack_bf16x2(y.z, y.w);
    *reinterpret_cast<uint4 *>(dq + b * sb + t * st + h * sh + part * 8) = v;
}

// ---------------------------------------------------------------- main backward kernel
constexpr int kBwdSmemBase = 16384 * 2 /*K,V*/ + 2 * 16384 /*Q x2*/ + 2 * 16384 /*dO x2*/ + 32768 /*P^T*/ + 32768 /*dS^T*/;
constexpr int kBwdCodeBytes = 2 * 16384;  // byte-code tiles [query][key] of two query tiles (TMA, 128-byte swizzle)
constexpr int kBwdBinsBytes = kBins * kBwdThreads * 4;

template <int MODE>
__global__ void __launch_bounds__(kBwdThreads, 1)
    attn_bwd_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                    const __grid_constant__ CUtensorMap mv, const __grid_constant__ CUtensorMap mdo,
                    const __grid_constant__ CUtensorMap mc, const BwdArgs a) {
    extern __shared__ uint8_t smem_dyn[];
    uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    uint8_t *sK = smem, *sV = smem + 16384, *sQ = smem + 32768, *sdO = smem + 65536, *sPt = smem + 98304,
            *sdSt = smem + 131072;
    uint8_t *sCode = smem + kBwdSmemBase;                                           // 2 x 16 KB, only with codes
    float *bins = reinterpret_cast<float *>(smem + kBwdSmemBase + kBwdCodeBytes);  // [kBins][256], only with grad_lut
    __shared__ __align__(8) uint64_t bar_kv, bar_qdo[2], bar_sp, bar_dq, bar_fin;
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(16) float s_nlse[2][128];   // -lse * log2(e); -inf for queries >= T
    __shared__ __align__(16) float s_delta[2][128];
    __shared__ float s_lut[256];        // materialised-free modes: lut * log2(e), -inf for MDT_CODE_MASKED
    __shared__ uint8_t s_bin[256];      // bias code -> gradient bin

    const int tid = threadIdx.x, warp = tid >> 5;
    const int r = tid & 127, half = tid >> 7;  // key row inside the tile, which half of the query columns
    const int h = blockIdx.y, b = blockIdx.z;
    const int T = a.T;
    const int k0 = blockIdx.x * kTile;
    const int key = k0 + r;
    const int Tm = a.Tmain;
    const bool key_ok = key < Tm && !(a.key_pad != nullptr && a.key_pad[(size_t)b * T + key] != 0);
    const int nq = (T + kTile - 1) / kTile;
    const bool want_gbias = (MODE == kBiasF32 || MODE == kBiasBF16) && a.grad_bias != nullptr;
    const bool want_glut = MODE == kBiasCode && a.grad_lut != nullptr;
    const uint32_t salt = drop_salt(a.drop, (uint32_t)(b * a.H + h));
    const bool warp_keys = __any_sync(0xffffffffu, key_ok);  // a warp whose 32 keys are all masked only writes zeros

    // A key tile without a single live key has dK = dV = 0 and adds nothing to dQ or the LUT gradient.
    // (The materialised-bias modes still have to write their d-bias zeros, so they take the long way.)
    if ((MODE == kBiasCode || MODE == kBiasNone) && !__syncthreads_or(key_ok ? 1 : 0)) {
        if (key < Tm) {
            uint4 *pk = reinterpret_cast<uint4 *>(a.dk + (size_t)b * a.dk_sb + (size_t)key * a.dk_st + (size_t)h * a.dk_sh + half * 32);
            uint4 *pv = reinterpret_cast<uint4 *>(a.dv + (size_t)b * a.dv_sb + (size_t)key * a.dv_st + (size_t)h * a.dv_sh + half * 32);
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) pk[ch] = pv[ch] = make_uint4(0, 0, 0, 0);
        }
        return;
    }

    auto stage_stats = [&](int i) {  // lse / delta of query tile i -> shared memory
        if (half != 0) return;
        const int qq = i * kTile + r;
        float nl = neg_inf(), dl = 0.f;
        if (qq < T) {
            const size_t row = ((size_t)b * a.H + h) * T + qq;
            nl = -a.lse[row] * kLog2e;
            dl = a.delta[row];
        }
        s_nlse[i & 1][r] = nl;
        s_delta[i & 1][r] = dl;
    };

    if (tid == 0) {
        mbar_init(&bar_kv, 1);
        mbar_init(&bar_qdo[0], 1);
        mbar_init(&bar_qdo[1], 1);
        mbar_init(&bar_sp, 1);
        mbar_init(&bar_dq, 1);
        mbar_init(&bar_fin, 1);
        fence_barrier_init();
        tma_prefetch_desc(&mq);
        tma_prefetch_desc(&mk);
        tma_prefetch_desc(&mv);
        tma_prefetch_desc(&mdo);
        if (MODE == kBiasCode) tma_prefetch_desc(&mc);
    }
    if (warp == 0) {
        tmem_alloc(&tmem_base_s, 512);
        tmem_relinquish();
    }
    if (MODE == kBiasCode) {
        s_lut[tid] = (tid == MDT_CODE_MASKED) ? neg_inf() : a.lut[(size_t)h * 256 + tid] * kLog2e;
        s_bin[tid] = tid == MDT_CODE_VIRTUAL ? (uint8_t)(kBins - 1) : (uint8_t)min(tid, kBins - 2);
        if (want_glut)
            for (int c = 0; c < kBins; ++c) bins[c * kBwdThreads + tid] = 0.f;
    }
    stage_stats(0);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    constexpr uint32_t cS = 0, cDP = 128, cDV = 256, cDK = 320, cDQ = 384;
    const uint32_t idesc_s = idesc_bf16(128, 128, 0, 0);
    const uint32_t idesc_kv = idesc_bf16(128, 64, 0, 1);
    const uint32_t idesc_dq = idesc_bf16(128, 64, 1, 1);

    auto issue_s_dp = [&](int i) {  // tid 0: S^T and dP^T of query tile i (operands already landed)
        const uint32_t ak = smem_u32(sK), av = smem_u32(sV);
        const uint32_t bq = smem_u32(sQ + (i & 1) * 16384), bo = smem_u32(sdO + (i & 1) * 16384);
#pragma unroll
        for (int k = 0; k < 4; ++k)
            umma_bf16(tmem + cS, smem_desc_sw128(ak + k * 32, 16, 1024), smem_desc_sw128(bq + k * 32, 16, 1024), idesc_s, k > 0);
#pragma unroll
        for (int k = 0; k < 4; ++k)
            umma_bf16(tmem + cDP, smem_desc_sw128(av + k * 32, 16, 1024), smem_desc_sw128(bo + k * 32, 16, 1024), idesc_s, k > 0);
        umma_commit(&bar_sp);
    };
    auto load_qdo = [&](int i) {  // tid 0
        mbar_expect_tx(&bar_qdo[i & 1], MODE == kBiasCode ? 49152 : 32768);
        tma_load_4d(sQ + (i & 1) * 16384, &mq, &bar_qdo[i & 1], 0, i * kTile, h, b);
        tma_load_4d(sdO + (i & 1) * 16384, &mdo, &bar_qdo[i & 1], 0, i * kTile, h, b);
        if (MODE == kBiasCode) tma_load_2d_raw(sCode + (i & 1) * 16384, &mc, &bar_qdo[i & 1], k0, b * T + i * kTile);
    };

    if (tid == 0) {
        mbar_expect_tx(&bar_kv, 32768);
        tma_load_4d(sK, &mk, &bar_kv, 0, k0, h, b);
        tma_load_4d(sV, &mv, &bar_kv, 0, k0, h, b);
        load_qdo(0);
        mbar_wait(&bar_kv, 0);
        mbar_wait(&bar_qdo[0], 0);
        tc_fence_after();
        issue_s_dp(0);
    }
    __syncwarp();

    for (int i = 0; i < nq; ++i) {
        const int q0 = i * kTile;
        const int qvalid = min(kTile, T - q0);
        mbar_wait(&bar_sp, i & 1);
        tc_fence_after();
        // buffer of tile i-1: bar_sp(i) was committed after every MMA that read it (dV, dK, S^T, dP^T of tile i-1)
        if (tid == 0 && i + 1 < nq) load_qdo(i + 1);
        __syncwarp();
        if (MODE == kBiasCode) mbar_wait(&bar_qdo[i & 1], (i >> 1) & 1);  // the code tile was written by TMA

        const float *nl = s_nlse[i & 1];
        const float *dl = s_delta[i & 1];
        float nl_next = neg_inf(), dl_next = 0.f;  // lse / delta of the next query tile: the global loads fly under this tile's work
        if (i + 1 < nq && half == 0 && (i + 1) * kTile + r < T) {
            const size_t row = ((size_t)b * a.H + h) * T + (i + 1) * kTile + r;
            nl_next = -a.lse[row] * kLog2e;
            dl_next = a.delta[row];
        }
#pragma unroll 1
        for (int cc = half * 2; cc < half * 2 + 2; ++cc) {
            if ((!warp_keys || cc * 32 >= qvalid) && !want_gbias) {  // 32 masked keys, or 32 query columns past T: P^T = dS^T = 0
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const uint32_t off = (cc >> 1) * 16384 + sw128_offset(r, (cc & 1) * 4 + u);
                    *reinterpret_cast<uint4 *>(sPt + off) = make_uint4(0, 0, 0, 0);
                    *reinterpret_cast<uint4 *>(sdSt + off) = make_uint4(0, 0, 0, 0);
                }
                continue;
            }
            if (MODE == kBiasCode || MODE == kBiasNone) {
                // ---- lean path.  No per-element validity tests: queries past T carry -lse = -inf (p = 0), keys past
                // T or padded zero the whole thread, rows of the next batch inside the tile are killed by the same -inf.
                float st[32], dp[32];
                tmem_ld_32x32(lane_base + cS + cc * 32, reinterpret_cast<uint32_t *>(st));   // .sync.aligned: whole warp
                tmem_ld_32x32(lane_base + cDP + cc * 32, reinterpret_cast<uint32_t *>(dp));
                tmem_wait_ld();
                if (key_ok) {
                    const uint32_t cbase = smem_u32(sCode + (i & 1) * 16384) + cc * 4096 + (r & 15);
                    const uint32_t rchunk = (uint32_t)(r >> 4) << 4;
                    const uint32_t lut_base = smem_u32(s_lut), bin_base = smem_u32(s_bin);
                    float *my_bins = bins + tid;
                    uint32_t bin_of[32];
#pragma unroll
                    for (int i2 = 0; i2 < 32; i2 += 4) {
                        const float4 n4 = *reinterpret_cast<const float4 *>(nl + cc * 32 + i2);
                        const float4 d4 = *reinterpret_cast<const float4 *>(dl + cc * 32 + i2);
                        const float nn[4] = {n4.x, n4.y, n4.z, n4.w}, dd[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int e = i2 + u;
                            float x;
                            uint32_t c = 0;
                            if (MODE == kBiasCode) {
                                // byte (query cc*32+e, key r) of the [query][key] tile: 128-byte rows, 16-byte chunks XOR-swizzled by row & 7
                                asm("ld.shared.u8 %0, [%1];" : "=r"(c) : "r"(cbase + e * 128 + (rchunk ^ ((e & 7) << 4))));
                                x = fmaf(st[e], kLog2e, lds_f32(lut_base + c * 4));
                            } else {
                                x = st[e] * kLog2e;
                            }
                            const float p = ex2(x + nn[u]);
                            float keepf = 1.f;  // dropout on P: dV takes keep * P / (1 - p_drop), dS = P * (keep * dP / (1 - p_drop) - delta)
                            if (a.drop.thr16) keepf = drop_keep_one(a.drop, salt, (uint32_t)(q0 + cc * 32 + e), (uint32_t)key) ? a.drop.scale : 0.f;
                            const float ds = p * (dp[e] * keepf - dd[u]);
                            st[e] = p * keepf;
                            dp[e] = ds;
                            if (MODE == kBiasCode && want_glut) asm("ld.shared.u8 %0, [%1];" : "=r"(bin_of[e]) : "r"(bin_base + c));
                        }
                    }
                    // ---- LUT gradient: per-thread bins [bin][tid] (conflict free, no atomics; masked codes carry dS = 0).
                    // The read-modify-writes of one thread form a dependent chain, so they go in pairs: both loads
                    // first, and when the two elements share a bin the second store carries both addends.
                    if (MODE == kBiasCode && want_glut) {
#pragma unroll
                        for (int e = 0; e < 32; e += 2) {
                            float *p0 = my_bins + bin_of[e] * kBwdThreads, *p1 = my_bins + bin_of[e + 1] * kBwdThreads;
                            const float a0 = *p0 + dp[e];
                            const float a1 = (bin_of[e] == bin_of[e + 1] ? a0 : *p1) + dp[e + 1];
                            *p0 = a0;
                            *p1 = a1;
                        }
                    }
                } else {
#pragma unroll
                    for (int e = 0; e < 32; ++e) st[e] = dp[e] = 0.f;
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    uint4 vp, vd;
                    vp.x = pack_bf16x2(st[u * 8 + 0], st[u * 8 + 1]);
                    vp.y = pack_bf16x2(st[u * 8 + 2], st[u * 8 + 3]);
                    vp.z = pack_bf16x2(st[u * 8 + 4], st[u * 8 + 5]);
                    vp.w = pack_bf16x2(st[u * 8 + 6], st[u * 8 + 7]);
                    vd.x = pack_bf16x2(dp[u * 8 + 0], dp[u * 8 + 1]);
                    vd.y = pack_bf16x2(dp[u * 8 + 2], dp[u * 8 + 3]);
                    vd.z = pack_bf16x2(dp[u * 8 + 4], dp[u * 8 + 5]);
                    vd.w = pack_bf16x2(dp[u * 8 + 6], dp[u * 8 + 7]);
                    const uint32_t off = (cc >> 1) * 16384 + sw128_offset(r, (cc & 1) * 4 + u);
                    *reinterpret_cast<uint4 *>(sPt + off) = vp;
                    *reinterpret_cast<uint4 *>(sdSt + off) = vd;
                }
                continue;
            }
            float st[32], dp[32];
            tmem_ld_32x32(lane_base + cS + cc * 32, reinterpret_cast<uint32_t *>(st));
            tmem_ld_32x32(lane_base + cDP + cc * 32, reinterpret_cast<uint32_t *>(dp));
            tmem_wait_ld();
            uint32_t code_w[32];
            // ---- bias^T: for a fixed query column the 128 threads read consecutive keys
            if (MODE == kBiasF32 || MODE == kBiasBF16) {
                const size_t base = (((size_t)b * a.H + h) * T + (q0 + cc * 32)) * (size_t)T + key;
#pragma unroll
                for (int i2 = 0; i2 < 32; ++i2) {
                    if (key < T && cc * 32 + i2 < qvalid) {
                        const size_t off = base + (size_t)i2 * T;
                        st[i2] += (MODE == kBiasF32) ? __ldg(reinterpret_cast<const float *>(a.bias) + off)
                                                     : __bfloat162float(reinterpret_cast<const __nv_bfloat16 *>(a.bias)[off]);
                    }
                }
            } else if (MODE == kBiasCode) {
                // byte (query cc*32+i2, key r) of the [query][key] tile: 128-byte rows, 16-byte chunks XOR-swizzled by row & 7
                const uint32_t cbase = smem_u32(sCode + (i & 1) * 16384) + cc * 4096 + (r & 15);
                const uint32_t rchunk = (uint32_t)(r >> 4) << 4;
#pragma unroll
                for (int i2 = 0; i2 < 32; ++i2) {
                    uint32_t c;
                    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(c) : "r"(cbase + i2 * 128 + (rchunk ^ ((i2 & 7) << 4))) : "memory");
                    if (!(key < T && cc * 32 + i2 < qvalid)) c = MDT_CODE_MASKED;
                    code_w[i2] = c;
                    st[i2] += s_lut[c];
                }
            }
            // ---- P^T and dS^T
#pragma unroll
            for (int i2 = 0; i2 < 32; i2 += 4) {
                const float4 n4 = *reinterpret_cast<const float4 *>(nl + cc * 32 + i2);
                const float4 d4 = *reinterpret_cast<const float4 *>(dl + cc * 32 + i2);
                const float nn[4] = {n4.x, n4.y, n4.z, n4.w}, dd[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float p = key_ok ? ex2(fmaf(st[i2 + u], kLog2e, nn[u])) : 0.f;
                    float keepf = 1.f;
                    if (a.drop.thr16) keepf = drop_keep_one(a.drop, salt, (uint32_t)(q0 + cc * 32 + i2 + u), (uint32_t)key) ? a.drop.scale : 0.f;
                    st[i2 + u] = p * keepf;
                    dp[i2 + u] = p * (dp[i2 + u] * keepf - dd[u]);
                }
            }
            if (want_gbias) {
                const size_t base = (((size_t)b * a.H + h) * T + (q0 + cc * 32)) * (size_t)T + key;
#pragma unroll
                for (int i2 = 0; i2 < 32; ++i2) {
                    if (key < T && cc * 32 + i2 < qvalid) {
                        const size_t off = base + (size_t)i2 * T;
                        if (MODE == kBiasF32) reinterpret_cast<float *>(a.grad_bias)[off] = dp[i2];
                        else reinterpret_cast<__nv_bfloat16 *>(a.grad_bias)[off] = __float2bfloat16_rn(dp[i2]);
                    }
                }
            }
            if (MODE == kBiasCode && want_glut) {
#pragma unroll
                for (int i2 = 0; i2 < 32; ++i2) {
                    const uint32_t c = code_w[i2];
                    const uint32_t bin = c == MDT_CODE_VIRTUAL ? (uint32_t)(kBins - 1) : min(c, (uint32_t)(kBins - 2));
                    bins[bin * kBwdThreads + tid] += dp[i2];  // masked codes carry dS = 0
                }
            }
            // ---- bf16 tiles [key][query], 128-byte swizzle, two 64-query halves
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                uint4 vp, vd;
                vp.x = pack_bf16x2(st[u * 8 + 0], st[u * 8 + 1]);
                vp.y = pack_bf16x2(st[u * 8 + 2], st[u * 8 + 3]);
                vp.z = pack_bf16x2(st[u * 8 + 4], st[u * 8 + 5]);
                vp.w = pack_bf16x2(st[u * 8 + 6], st[u * 8 + 7]);
                vd.x = pack_bf16x2(dp[u * 8 + 0], dp[u * 8 + 1]);
                vd.y = pack_bf16x2(dp[u * 8 + 2], dp[u * 8 + 3]);
                vd.z = pack_bf16x2(dp[u * 8 + 4], dp[u * 8 + 5]);
                vd.w = pack_bf16x2(dp[u * 8 + 6], dp[u * 8 + 7]);
                const uint32_t off = (cc >> 1) * 16384 + sw128_offset(r, (cc & 1) * 4 + u);
                *reinterpret_cast<uint4 *>(sPt + off) = vp;
                *reinterpret_cast<uint4 *>(sdSt + off) = vd;
            }
        }
        if (i + 1 < nq && half == 0) {  // loaded before the softmax work of this tile, stored after it
            s_nlse[(i + 1) & 1][r] = nl_next;
            s_delta[(i + 1) & 1][r] = dl_next;
        }
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            const uint32_t ap = smem_u32(sPt), ad = smem_u32(sdSt);
            const uint32_t bo = smem_u32(sdO + (i & 1) * 16384), bq = smem_u32(sQ + (i & 1) * 16384), bk = smem_u32(sK);
            const uint32_t acc = i > 0 ? 1u : 0u;
            // dQ first: it is the only product the threads wait for inside the loop (bar_dq); dV / dK accumulate under
            // the dQ read-out and its atomics, and are covered by the next bar_sp (or bar_fin after the last tile)
#pragma unroll
            for (int k = 0; k < 8; ++k)  // dQ_i = dS K   (A: the dS^T tile read MN-major, k-step = 16 keys)
                umma_bf16(tmem + cDQ, smem_desc_sw128(ad + k * 2048, 16384, 1024), smem_desc_sw128(bk + k * 2048, 8192, 1024),
                          idesc_dq, k > 0);
            umma_commit(&bar_dq);
#pragma unroll
            for (int k = 0; k < 8; ++k)  // dV += P^T dO_i
                umma_bf16(tmem + cDV, smem_desc_sw128(ap + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                          smem_desc_sw128(bo + k * 2048, 8192, 1024), idesc_kv, (k > 0) ? 1u : acc);
#pragma unroll
            for (int k = 0; k < 8; ++k)  // dK += dS^T Q_i
                umma_bf16(tmem + cDK, smem_desc_sw128(ad + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                          smem_desc_sw128(bq + k * 2048, 8192, 1024), idesc_kv, (k > 0) ? 1u : acc);
            if (i + 1 < nq) {  // S^T / dP^T of the next tile
                mbar_wait(&bar_qdo[(i + 1) & 1], ((i + 1) >> 1) & 1);
                tc_fence_after();
                issue_s_dp(i + 1);
            } else {
                umma_commit(&bar_fin);
            }
        }
        __syncwarp();
        mbar_wait(&bar_dq, i & 1);
        tc_fence_after();
        __syncwarp();
        {
            float dq[32];  // this half's 32 of the 64 dQ columns of query row r
            tmem_ld_32x32(lane_base + cDQ + half * 32, reinterpret_cast<uint32_t *>(dq));
            tmem_wait_ld();
            if (r < qvalid) {
                float4 *dst = reinterpret_cast<float4 *>(a.dq_accum + (((size_t)b * a.H + h) * T + (q0 + r)) * 64 + half * 32);
#pragma unroll
                for (int u = 0; u < 8; ++u) atomicAdd(dst + u, make_float4(dq[u * 4], dq[u * 4 + 1], dq[u * 4 + 2], dq[u * 4 + 3]));
            }
        }
        tc_fence_before();  // the dQ columns are rewritten by the next tile's MMAs after the next __syncthreads
    }

    // ---- epilogue: dK, dV rows of this key tile; each half stores 32 columns
    mbar_wait(&bar_fin, 0);
    tc_fence_after();
    __syncwarp();
    {
        float acc[32];
#pragma unroll
        for (int which = 0; which < 2; ++which) {
            tmem_ld_32x32(lane_base + (which == 0 ? cDV : cDK) + half * 32, reinterpret_cast<uint32_t *>(acc));
            tmem_wait_ld();
            if (key < Tm) {
                __nv_bfloat16 *p = which == 0 ? a.dv + (size_t)b * a.dv_sb + (size_t)key * a.dv_st + (size_t)h * a.dv_sh
                                              : a.dk + (size_t)b * a.dk_sb + (size_t)key * a.dk_st + (size_t)h * a.dk_sh;
                p += half * 32;
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) {
                    uint4 v;
                    v.x = pack_bf16x2(acc[ch * 8 + 0], acc[ch * 8 + 1]);
                    v.y = pack_bf16x2(acc[ch * 8 + 2], acc[ch * 8 + 3]);
                    v.z = pack_bf16x2(acc[ch * 8 + 4], acc[ch * 8 + 5]);
                    v.w = pack_bf16x2(acc[ch * 8 + 6], acc[ch * 8 + 7]);
                    reinterpret_cast<uint4 *>(p)[ch] = v;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
    if (MODE == kBiasCode && want_glut && tid < kBins) {
        // bin `tid`: sum over the 256 per-thread copies (rotated start keeps the banks distinct)
        float sum = 0.f;
        for (int t = 0; t < kBwdThreads; ++t) sum += bins[tid * kBwdThreads + ((t + tid) & (kBwdThreads - 1))];
        const int code = tid == kBins - 1 ? MDT_CODE_VIRTUAL : tid;
        if (sum != 0.f) atomicAdd(a.grad_lut + (size_t)h * 256 + code, sum);
    }
}

template <int MODE>
static int launch_bwd(const CUtensorMap &mq, const CUtensorMap &mk, const CUtensorMap &mv, const CUtensorMap &mdo,
                      const CUtensorMap &mc, const BwdArgs &a, cudaStream_t st) {
    const int smem = kBwdSmemBase + 1024 + (MODE == kBiasCode ? kBwdCodeBytes : 0) + ((MODE == kBiasCode && a.grad_lut) ? kBwdBinsBytes : 0);
    // per launch: the attribute is per device and the call is cheap (no cache keyed on the wrong device)
    MDT_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    dim3 grid((a.Tmain + kTile - 1) / kTile, a.H, a.B);
    attn_bwd_kernel<MODE><<<grid, kBwdThreads, smem, st>>>(mq, mk, mv, mdo, mc, a);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

}  // namespace attn
}  // namespace mdt

using namespace mdt;
using namespace mdt::attn;

extern "C" int mdt_biased_attn_bwd(const void *q, const void *k, const void *v, const void *out, const void *grad_out,
                                   const float *lse, const mdt_attn_dims *d, const void *bias, int bias_dtype,
                                   const uint8_t *code, const uint8_t *code_t, int code_pitch, const float *lut,
                                   const uint8_t *key_pad, void *dq, void *dk, void *dv, const mdt_attn_dims *g, float *dq_accum, float *delta,
                                   void *grad_bias, float *grad_lut, float dropout_p, uint64_t dropout_seed, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    if (int rc = check_dims(d, "mdt_biased_attn_bwd")) return rc;
    if (int rc = check_dims(g, "mdt_biased_attn_bwd(grad_dims)")) return rc;
    MDT_REQUIRE(g->B == d->B && g->H == d->H && g->T == d->T, "mdt_biased_attn_bwd: grad_dims must repeat B, H, T");
    MDT_REQUIRE(q && k && v && out && grad_out && lse && dq && dk && dv && dq_accum && delta,
                "mdt_biased_attn_bwd: q, k, v, out, grad_out, lse, dq, dk, dv, dq_accum, delta must be non-null");
    MDT_REQUIRE(!(bias && code), "mdt_biased_attn_bwd: pass a materialised bias or a code/lut pair, not both");
    MDT_REQUIRE(bias == nullptr || bias_dtype == MDT_F32 || bias_dtype == MDT_BF16, "mdt_biased_attn_bwd: bad bias dtype");
    MDT_REQUIRE(grad_bias == nullptr || bias != nullptr, "mdt_biased_attn_bwd: grad_bias needs bias");
    MDT_REQUIRE(grad_lut == nullptr || code != nullptr, "mdt_biased_attn_bwd: grad_lut needs code");
    MDT_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "mdt_biased_attn_bwd: dropout_p must be in [0, 1)");
    if (code) {
        MDT_REQUIRE(lut != nullptr, "mdt_biased_attn_bwd: code needs lut");
        MDT_REQUIRE(code_pitch >= d->T, "mdt_biased_attn_bwd: code_pitch must be >= T");
        MDT_REQUIRE(code_t == nullptr || (code_pitch % 16 == 0 && ((uintptr_t)code_t & 15) == 0),
                    "mdt_biased_attn_bwd: code_t needs a 16-byte aligned base and a pitch that is a multiple of 16");
    }
    const int64_t strides[] = {d->o_stride_b, d->o_stride_t, d->o_stride_h, g->o_stride_b, g->o_stride_t, g->o_stride_h,
                               g->q_stride_b, g->q_stride_t, g->q_stride_h, g->k_stride_b, g->k_stride_t, g->k_stride_h,
                               g->v_stride_b, g->v_stride_t, g->v_stride_h};
    for (int64_t s : strides) MDT_REQUIRE(s % 8 == 0, "mdt_biased_attn_bwd: strides must be multiples of 8 elements");
    MDT_REQUIRE((((uintptr_t)out | (uintptr_t)grad_out | (uintptr_t)dq | (uintptr_t)dk | (uintptr_t)dv | (uintptr_t)dq_accum) & 15) == 0,
                "mdt_biased_attn_bwd: buffers must be 16-byte aligned");
    cudaStream_t st = as_stream(stream);
    const int B = d->B, H = d->H, T = d->T;

    const int mode = bias_mode(bias, bias_dtype, code);
    // code / no bias: warp-specialised kernel (attn_bwd2.cu); it computes delta itself.  MDT_ATTN_BWD=1 selects the older
    // one-role kernel below (with its tail-key helper for T = 128 k + r) for A/B measurements.
    static const bool use_bwd2 = []() { const char *e = getenv("MDT_ATTN_BWD"); return !(e && e[0] == '1'); }();
    const bool bwd2 = use_bwd2 && (mode == kBiasNone || (mode == kBiasCode && code_t != nullptr));
    // The tail-key split pays off for the one-role kernel only: with bwd2 the almost empty key tile of T = 128 k + r is a CTA
    // whose dead warps skip the arithmetic (T = 129: 62 us without the split, 71 us with it and its two helper launches).
    // ... up to ~6 key tiles.  From T = 768 on the split wins (T = 1025: 884 -> 826 us; T = 513: 285 -> 281; T = 257: 138 -> 142):
    // the one-key tile costs a ninth of the main kernel there.  MDT_ATTN_TAIL_FROM overrides the threshold (A/B runs).
    static const int tail_from = []() { const char *e = getenv("MDT_ATTN_TAIL_FROM"); return e ? atoi(e) : 768; }();
    const int Tmain = tail_split(T, (!bwd2 || T >= tail_from) && (mode == kBiasCode || mode == kBiasNone));
    if (!bwd2 || Tmain < T) {  // delta = rowsum(dO * O)
        const int64_t rows = (int64_t)B * H * T;
        const int threads = 256, rows_per_block = threads / 8;
        attn_delta_kernel<<<(unsigned)((rows + rows_per_block - 1) / rows_per_block), threads, 0, st>>>(
            reinterpret_cast<const __nv_bfloat16 *>(out), reinterpret_cast<const __nv_bfloat16 *>(grad_out), B, H, T,
            d->o_stride_b, d->o_stride_t, d->o_stride_h, g->o_stride_b, g->o_stride_t, g->o_stride_h, delta);
        MDT_CHECK_CUDA(cudaGetLastError());
    }
    CUtensorMap mq, mk, mv, mdo;
    if (int rc = make_bthd_map(&mq, q, B, H, T, d->q_stride_b, d->q_stride_t, d->q_stride_h, kTile)) return rc;
    if (int rc = make_bthd_map(&mk, k, B, H, T, d->k_stride_b, d->k_stride_t, d->k_stride_h, kTile)) return rc;
    if (int rc = make_bthd_map(&mv, v, B, H, T, d->v_stride_b, d->v_stride_t, d->v_stride_h, kTile)) return rc;
    if (int rc = make_bthd_map(&mdo, grad_out, B, H, T, g->o_stride_b, g->o_stride_t, g->o_stride_h, kTile)) return rc;
    BwdArgs a;
    a.B = B, a.H = H, a.T = T;
    a.bias = bias, a.code = code, a.code_pitch = code_pitch, a.lut = lut, a.key_pad = key_pad;
    a.lse = lse, a.delta = delta, a.dq_accum = dq_accum;
    a.dk = reinterpret_cast<__nv_bfloat16 *>(dk), a.dv = reinterpret_cast<__nv_bfloat16 *>(dv);
    a.dk_sb = g->k_stride_b, a.dk_st = g->k_stride_t, a.dk_sh = g->k_stride_h;
    a.dv_sb = g->v_stride_b, a.dv_st = g->v_stride_t, a.dv_sh = g->v_stride_h;
    a.grad_bias = grad_bias, a.grad_lut = grad_lut;
    a.drop = make_drop(dropout_p, dropout_seed);
    a.Tmain = Tmain;
    CUtensorMap mc = mq;  // only read by the code mode: [query][key] tiles of code, or [key][query] tiles of code_t (bwd2)
    if (code)
        if (int rc2 = make_tensor_map_u8_2d(&mc, bwd2 ? code_t : code, (uint64_t)code_pitch, (uint64_t)B * T, (uint64_t)code_pitch, kTile)) return rc2;
    int rc;
    if (bwd2) {
        Bwd2Args a2{};
        a2.B = B, a2.H = H, a2.T = T, a2.Tmain = Tmain;
        a2.key_pad = key_pad, a2.lut = lut, a2.lse = lse;
        a2.delta = Tmain < T ? delta : nullptr;
        a2.out = reinterpret_cast<const __nv_bfloat16 *>(out), a2.dout = reinterpret_cast<const __nv_bfloat16 *>(grad_out);
        a2.o_sb = d->o_stride_b, a2.o_st = d->o_stride_t, a2.o_sh = d->o_stride_h;
        a2.g_sb = g->o_stride_b, a2.g_st = g->o_stride_t, a2.g_sh = g->o_stride_h;
        a2.dq_accum = dq_accum, a2.dk = a.dk, a2.dv = a.dv;
        a2.dk_sb = a.dk_sb, a2.dk_st = a.dk_st, a2.dk_sh = a.dk_sh, a2.dv_sb = a.dv_sb, a2.dv_st = a.dv_st, a2.dv_sh = a.dv_sh;
        a2.grad_lut = grad_lut, a2.drop = a.drop;
        rc = launch_bwd2(mq, mk, mv, mdo, mc, a2, code != nullptr, st);
    } else
    switch (mode) {
        case kBiasNone: rc = launch_bwd<kBiasNone>(mq, mk, mv, mdo, mc, a, st); break;
        case kBiasF32: rc = launch_bwd<kBiasF32>(mq, mk, mv, mdo, mc, a, st); break;
        case kBiasBF16: rc = launch_bwd<kBiasBF16>(mq, mk, mv, mdo, mc, a, st); break;
        default: rc = launch_bwd<kBiasCode>(mq, mk, mv, mdo, mc, a, st); break;
    }
    if (rc) return rc;
    if (a.Tmain < T) {  // keys [Tmain, T) on CUDA cores (attn_tail.cu)
        TailArgs t{};
        t.B = B, t.H = H, t.T = T, t.Tmain = a.Tmain;
        t.q = reinterpret_cast<const __nv_bfloat16 *>(q), t.k = reinterpret_cast<const __nv_bfloat16 *>(k);
        t.v = reinterpret_cast<const __nv_bfloat16 *>(v);
        t.q_sb = d->q_stride_b, t.q_st = d->q_stride_t, t.q_sh = d->q_stride_h;
        t.k_sb = d->k_stride_b, t.k_st = d->k_stride_t, t.k_sh = d->k_stride_h;
        t.v_sb = d->v_stride_b, t.v_st = d->v_stride_t, t.v_sh = d->v_stride_h;
        t.code = code, t.code_pitch = code_pitch, t.lut = lut, t.key_pad = key_pad;
        t.lse = const_cast<float *>(lse);
        t.dout = reinterpret_cast<const __nv_bfloat16 *>(grad_out);
        t.g_sb = g->o_stride_b, t.g_st = g->o_stride_t, t.g_sh = g->o_stride_h;
        t.delta = delta, t.dq_accum = dq_accum, t.dk = a.dk, t.dv = a.dv;
        t.dk_sb = a.dk_sb, t.dk_st = a.dk_st, t.dk_sh = a.dk_sh, t.dv_sb = a.dv_sb, t.dv_st = a.dv_st, t.dv_sh = a.dv_sh;
        t.grad_lut = grad_lut;
        t.drop = a.drop;
        if (int rc3 = launch_bwd_tails(t, st)) return rc3;
    }
    {
        const int64_t n = (int64_t)B * H * T * 8;
        attn_dq_convert_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(dq_accum, B, H, T, reinterpret_cast<__nv_bfloat16 *>(dq),
                                                                            g->q_stride_b, g->q_stride_t, g->q_stride_h);
        MDT_CHECK_CUDA(cudaGetLastError());
    }
    return MDT_OK;
}

// A6 forward, warp-specialised (bias synthesised from byte codes, or no bias)   multihead_attention.py:168-199
//
// One CTA per (batch b, head h, 256 queries) = two 128-query tiles that share every K/V tile.
//   warps 0-3  softmax warpgroup of query tile 0   (thread r <-> query row r <-> TMEM lane r)
//   warps 4-7  softmax warpgroup of query tile 1
//   warp  8    TMA producer: Q0/Q1 once, then per key tile K, V and the two 128x128 byte-code tiles
//   warp  9    MMA issuer:   S_t = Q_t K^T (128x128x64) and O_t += P_t V (128x64x128), tcgen05 -> TMEM
//   warps 10-11 idle (setmaxnreg works per warpgroup)
// Issue order of the tensor pipe:  S0(0) S1(0) | PV0(0) S0(1) PV1(0) S1(1) | PV0(1) S0(2) ...
// so the MMAs of one query tile run under the softmax of the other (ping-pong).  All hand-offs are
// mbarriers; nothing uses __syncthreads after the prologue.
// TMEM columns: S0 [0,128) S1 [128,256) O0 [256,320) O1 [320,384).
// Shared memory: Q0 Q1 | 2 stages x (K, V, code0, code1) | P0 P1  = 224 KB.
#include "attn_common.cuh"

namespace mdt {
namespace attn {
using namespace sm100;


constexpr int kF2Threads = 384;
constexpr int kF2Stage = 4 * 16384;                                 // K, V, code0, code1
constexpr int kF2Smem = 2 * 16384 + 2 * kF2Stage + 2 * 32768 + 1024; // + alignment slack
constexpr float kF2Rescale = 8.0f;   // log2 units: P may reach 2^8 before O is rescaled
constexpr int kF2MaxT = 8192;         // key-mask bitmap of the whole row lives in shared memory

struct F2Bars {
    uint64_t q_full, k_full[2], v_full[2], c_full[2], kv_empty[2], c_empty[2];
    uint64_t s_full[2], s_free[2], p_full[2], o_done[2];
};

template <bool HAS_CODE>
__global__ void __launch_bounds__(kF2Threads, 1)
    attn_fwd2_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                     const __grid_constant__ CUtensorMap mv, const __grid_constant__ CUtensorMap mc, const Fwd2Args a) {
    extern __shared__ uint8_t smem_dyn[];
    uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    uint8_t *sQ = smem;                       // 2 x 16 KB
    uint8_t *sStage = smem + 2 * 16384;       // 2 x 64 KB
    uint8_t *sP = sStage + 2 * kF2Stage;      // 2 x 32 KB
    __shared__ __align__(8) F2Bars bars;
    __shared__ uint32_t tmem_base_s;
    // 64-entry table in log2 units, indexed by (code & 63): spatial codes 0..61, 62 = MDT_CODE_VIRTUAL & 63,
    // 63 = MDT_CODE_MASKED & 63 (-inf).  Codes 62..253 are reserved (the collate kernel emits <= 21).
    __shared__ float s_lut[64];
    __shared__ uint32_t s_kmask[kF2MaxT / 32];  // bit set = key masked (padding or >= T)
    __shared__ int s_last_word;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int h = blockIdx.y, b = blockIdx.z;
    const int T = a.T;
    const int q_base = blockIdx.x * 2 * kTile;
    const int nkv_all = (T + kTile - 1) / kTile;
    const bool tile1_live = q_base + kTile < T;  // a CTA at the tail may own a single query tile

    if (tid == 0) {
        mbar_init(&bars.q_full, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&bars.k_full[s], 1);
            mbar_init(&bars.v_full[s], 1);
            mbar_init(&bars.c_full[s], 1);
            mbar_init(&bars.kv_empty[s], 1);
            mbar_init(&bars.c_empty[s], 256);
            mbar_init(&bars.s_full[s], 1);
            mbar_init(&bars.s_free[s], 128);
            mbar_init(&bars.p_full[s], 128);
            mbar_init(&bars.o_done[s], 1);
        }
        fence_barrier_init();
        tma_prefetch_desc(&mq);
        tma_prefetch_desc(&mk);
        tma_prefetch_desc(&mv);
        if (HAS_CODE) tma_prefetch_desc(&mc);
    }
    if (warp == 8) {
        tmem_alloc(&tmem_base_s, 512);
        tmem_relinquish();
    }
    if (HAS_CODE && tid < 64) {
        const int src = tid == 62 ? MDT_CODE_VIRTUAL : tid;
        s_lut[tid] = tid == 63 ? neg_inf() : a.lut[(size_t)h * 256 + src] * kLog2e;
    }
    if (tid == 0) s_last_word = -1;
    __syncthreads();
    for (int k0 = warp * 32; k0 < nkv_all * kTile; k0 += kF2Threads) {
        const int kk = k0 + lane;
        const bool masked = kk >= T || (a.key_pad != nullptr && a.key_pad[(size_t)b * T + kk] != 0);
        const uint32_t bal = __ballot_sync(0xffffffffu, masked);
        if (lane == 0) {
            s_kmask[k0 >> 5] = bal;
            if (bal != 0xffffffffu) atomicMax(&s_last_word, k0 >> 5);
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    // key tiles after the last unmasked key contribute exactly 0 to every row: never loaded or multiplied
    const int nkv = s_last_word < 0 ? 1 : (s_last_word >> 2) + 1;

    if (warp >= 8) {
        // ------------------------------------------------------------------ producer / MMA warpgroup
        reg_dealloc<40>();
        if (warp == 8 && lane == 0) {
            mbar_expect_tx(&bars.q_full, 32768);
            tma_load_4d(sQ, &mq, &bars.q_full, 0, q_base, h, b);
            tma_load_4d(sQ + 16384, &mq, &bars.q_full, 0, q_base + kTile, h, b);
            for (int j = 0; j < nkv; ++j) {
                const int s = j & 1, k0 = j * kTile;
                uint8_t *st = sStage + s * kF2Stage;
                if (j >= 2) {
                    const uint32_t par = ((j >> 1) - 1) & 1;
                    mbar_wait(&bars.kv_empty[s], par);
                    if (HAS_CODE) mbar_wait(&bars.c_empty[s], par);
                }
                mbar_expect_tx(&bars.k_full[s], 16384);
                tma_load_4d(st, &mk, &bars.k_full[s], 0, k0, h, b);
                if (HAS_CODE) {
                    mbar_expect_tx(&bars.c_full[s], 32768);
                    tma_load_2d_raw(st + 32768, &mc, &bars.c_full[s], k0, b * T + q_base);
                    tma_load_2d_raw(st + 49152, &mc, &bars.c_full[s], k0, b * T + q_base + kTile);
                }
                mbar_expect_tx(&bars.v_full[s], 16384);
                tma_load_4d(st + 16384, &mv, &bars.v_full[s], 0, k0, h, b);
            }
        } else if (warp == 9 && lane == 0) {
            const uint32_t idesc_s = idesc_bf16(128, 128, 0, 0);
            const uint32_t idesc_o = idesc_bf16(128, 64, 0, 1);
            auto issue_s = [&](int t, int stage) {
                const uint32_t aq = smem_u32(sQ + t * 16384), bk = smem_u32(sStage + stage * kF2Stage);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16(tmem + t * 128, smem_desc_sw128(aq + k * 32, 16, 1024), smem_desc_sw128(bk + k * 32, 16, 1024), idesc_s,
                              k > 0);
                umma_commit(&bars.s_full[t]);
            };
            mbar_wait(&bars.q_full, 0);
            mbar_wait(&bars.k_full[0], 0);
            tc_fence_after();
            issue_s(0, 0);
            issue_s(1, 0);
            for (int j = 0; j < nkv; ++j) {
                const int s = j & 1;
                for (int t = 0; t < 2; ++t) {
                    mbar_wait(&bars.p_full[t], j & 1);
                    if (t == 0) mbar_wait(&bars.v_full[s], (j >> 1) & 1);
                    tc_fence_after();
                    const uint32_t ap = smem_u32(sP + t * 32768), bv = smem_u32(sStage + s * kF2Stage + 16384);
#pragma unroll
                    for (int k = 0; k < 8; ++k)
                        umma_bf16(tmem + 256 + t * 64, smem_desc_sw128(ap + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                                  smem_desc_sw128(bv + k * 2048, 8192, 1024), idesc_o, (j > 0 || k > 0) ? 1u : 0u);
                    umma_commit(&bars.o_done[t]);
                    if (t == 1) umma_commit(&bars.kv_empty[s]);
                    if (j + 1 < nkv) {
                        if (t == 0) mbar_wait(&bars.k_full[(j + 1) & 1], ((j + 1) >> 1) & 1);
                        mbar_wait(&bars.s_free[t], j & 1);
                        tc_fence_after();
                        issue_s(t, (j + 1) & 1);
                    }
                }
            }
        }
    } else {
        // ------------------------------------------------------------------ softmax warpgroups
        reg_alloc<232>();
        const int t = warp >> 2;                       // query tile of this warpgroup
        const int r = tid & 127;                       // row inside the tile == TMEM lane
        const int q = q_base + t * kTile + r;
        const bool row_ok = q < T;
        const bool warp_live = __any_sync(0xffffffffu, row_ok);  // warps past the end only keep the barriers moving
        const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        const uint32_t tS = lane_base + t * 128, tO = lane_base + 256 + t * 64;
        const uint32_t myP = smem_u32(sP + t * 32768);
        float m_used = neg_inf(), l = 0.f;   // m_used in log2 units (scores are scaled by log2 e on the fly)
        const uint32_t lut_base = smem_u32(s_lut);

        for (int j = 0; j < nkv; ++j) {
            const int s = j & 1;
            uint32_t kmask[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) kmask[c] = s_kmask[j * 4 + c];
            float sc[128];
            mbar_wait(&bars.s_full[t], j & 1);
            tc_fence_after();
            __syncwarp();
#pragma unroll
            for (int c = 0; c < 4; ++c) tmem_ld_32x32(tS + c * 32, reinterpret_cast<uint32_t *>(sc) + c * 32);
            tmem_wait_ld();
            tc_fence_before();
            mbar_arrive(&bars.s_free[t]);              // S_t may be overwritten by the next Q_t K^T
            if (HAS_CODE) mbar_wait(&bars.c_full[s], (j >> 1) & 1);
            // ---- pass 1: x = s * log2(e) + lut[code] (log2 units), key mask, row maximum.
            // Chunks of 32 keys that are masked for every row are skipped in both passes.
            float mx0 = neg_inf(), mx1 = neg_inf();
            if (warp_live) {
                const uint32_t crow = smem_u32(sStage + s * kF2Stage + 32768 + t * 16384);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    if (kmask[c] != 0xffffffffu) {
#pragma unroll
                        for (int u = 2 * c; u < 2 * c + 2; ++u) {
                            if (HAS_CODE) {
                                const uint4 w = lds_u4(crow + sw128_offset(r, u));
                                const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                                for (int g = 0; g < 4; ++g) {
                                    const uint32_t wm = (ww[g] << 2) & 0xfcfcfcfcu;   // byte -> 4 * (code & 63)
                                    const float l0 = lds_f32(lut_base + __byte_perm(wm, 0, 0x4440));
                                    const float l1 = lds_f32(lut_base + __byte_perm(wm, 0, 0x4441));
                                    const float l2 = lds_f32(lut_base + __byte_perm(wm, 0, 0x4442));
                                    const float l3 = lds_f32(lut_base + __byte_perm(wm, 0, 0x4443));
                                    float *x = sc + u * 16 + g * 4;
                                    ffma2(x[0], x[1], x[0], x[1], kLog2e, kLog2e, l0, l1);
                                    ffma2(x[2], x[3], x[2], x[3], kLog2e, kLog2e, l2, l3);
                                }
                            } else {
#pragma unroll
                                for (int g = 0; g < 8; ++g) {
                                    float *x = sc + u * 16 + g * 2;
                                    ffma2(x[0], x[1], x[0], x[1], kLog2e, kLog2e, 0.f, 0.f);
                                }
                            }
                        }
                        if (kmask[c] != 0) {
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                                if (kmask[c] & (1u << i)) sc[c * 32 + i] = neg_inf();
                        }
#pragma unroll
                        for (int i = 0; i < 32; i += 4) {
                            mx0 = fmax3(mx0, sc[c * 32 + i], sc[c * 32 + i + 1]);
                            mx1 = fmax3(mx1, sc[c * 32 + i + 2], sc[c * 32 + i + 3]);
                        }
                    }
                }
            }
            if (HAS_CODE) mbar_arrive(&bars.c_empty[s]);
            float alpha = 1.f;
            bool grow = false;
            if (warp_live) {
                const float mx = fmaxf(mx0, mx1);
                grow = mx > m_used + kF2Rescale;       // also true for the first finite maximum
                if (grow) {
                    alpha = ex2(m_used - mx);
                    m_used = mx;
                    l *= alpha;
                }
            }
            // P_t (shared) and O_t (TMEM) are free once P V of the previous key tile has completed
            if (j > 0) {
                mbar_wait(&bars.o_done[t], (j - 1) & 1);
                tc_fence_after();
                __syncwarp();
                if (__any_sync(0xffffffffu, grow)) {
                    uint32_t o[32];
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        tmem_ld_32x32(tO + c * 32, o);
                        tmem_wait_ld();
#pragma unroll
                        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                        tmem_st_32x32(tO + c * 32, o);
                    }
                    tmem_wait_st();
                }
            }
            // ---- pass 2: p = 2^(x - m), row sum, bf16 P tile (128-byte swizzle, K-major A operand of P V)
            if (warp_live) {
                const float nm = (m_used == neg_inf()) ? 0.f : -m_used;
                float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    if (kmask[c] != 0xffffffffu) {
#pragma unroll
                        for (int ch = 4 * c; ch < 4 * c + 4; ++ch) {
                            float *x = sc + ch * 8;
#pragma unroll
                            for (int i = 0; i < 8; i += 2) {
                                fadd2(x[i], x[i + 1], x[i], x[i + 1], nm, nm);
                                x[i] = ex2(x[i]);
                                x[i + 1] = ex2(x[i + 1]);
                                fadd2(sum0, sum1, sum0, sum1, x[i], x[i + 1]);
                            }
                            uint4 v;
                            v.x = pack_bf16x2(x[0], x[1]);
                            v.y = pack_bf16x2(x[2], x[3]);
                            v.z = pack_bf16x2(x[4], x[5]);
                            v.w = pack_bf16x2(x[6], x[7]);
                            sts_u4(myP + (ch >> 3) * 16384 + sw128_offset(r, ch & 7), v);
                        }
                    } else {
#pragma unroll
                        for (int ch = 4 * c; ch < 4 * c + 4; ++ch)
                            sts_u4(myP + (ch >> 3) * 16384 + sw128_offset(r, ch & 7), make_uint4(0, 0, 0, 0));
                    }
                }
                l += sum0 + sum1;
            }
            fence_proxy_async();
            tc_fence_before();
            mbar_arrive(&bars.p_full[t]);
        }
        // ---- epilogue: O / l -> bf16, log-sum-exp
        mbar_wait(&bars.o_done[t], (nkv - 1) & 1);
        tc_fence_after();
        __syncwarp();
        if (warp_live) {
            float o[64];
            tmem_ld_32x32(tO, reinterpret_cast<uint32_t *>(o));
            tmem_ld_32x32(tO + 32, reinterpret_cast<uint32_t *>(o) + 32);
            tmem_wait_ld();
            if (row_ok) {
                const float inv = l > 0.f ? 1.f / l : 0.f;  // a fully masked row yields zeros (the reference yields NaN)
                __nv_bfloat16 *op = a.out + (size_t)b * a.o_sb + (size_t)q * a.o_st + (size_t)h * a.o_sh;
#pragma unroll
                for (int ch = 0; ch < 8; ++ch) {
                    uint4 v;
                    v.x = pack_bf16x2(o[ch * 8 + 0] * inv, o[ch * 8 + 1] * inv);
                    v.y = pack_bf16x2(o[ch * 8 + 2] * inv, o[ch * 8 + 3] * inv);
                    v.z = pack_bf16x2(o[ch * 8 + 4] * inv, o[ch * 8 + 5] * inv);
                    v.w = pack_bf16x2(o[ch * 8 + 6] * inv, o[ch * 8 + 7] * inv);
                    reinterpret_cast<uint4 *>(op)[ch] = v;
                }
                if (a.lse) a.lse[((size_t)b * a.H + h) * T + q] = l > 0.f ? m_used * kLn2 + logf(l) : __int_as_float(0x7f800000);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) tmem_dealloc(tmem, 512);
}

int launch_fwd2(const CUtensorMap &mq, const CUtensorMap &mk, const CUtensorMap &mv, const CUtensorMap &mc, const Fwd2Args &a,
                cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        MDT_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kF2Smem));
        MDT_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kF2Smem));
        configured = true;
    }
    dim3 grid((a.T + 2 * kTile - 1) / (2 * kTile), a.H, a.B);
    if (a.has_code)
        attn_fwd2_kernel<true><<<grid, kF2Threads, kF2Smem, st>>>(mq, mk, mv, mc, a);
    else
        attn_fwd2_kernel<false><<<grid, kF2Threads, kF2Smem, st>>>(mq, mk, mv, mc, a);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

}  // namespace attn
}  // namespace mdt

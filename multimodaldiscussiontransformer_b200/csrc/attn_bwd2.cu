// A6 backward, warp-specialised (bias from byte codes, or none)            multihead_attention.py:168-199
//
// One CTA per (batch b, head h, 128-key tile), looping over 128-query tiles i; 20 warps with separate jobs, so that the
// tensor pipe, the softmax arithmetic, the LUT-gradient binning and the dQ reduction of neighbouring query tiles overlap
// instead of taking turns:
//   warps 0-7   softmax: thread (r, half) owns key row r (= TMEM lane r) and query columns [64 half, 64 half + 64) of
//               every tile.  It pulls its S^T / dP^T values into registers in one go and hands the TMEM columns back at
//               once (S^T / dP^T of tile i+1 are then computed under the arithmetic of tile i), forms
//               P^T = 2^(S^T + bias^T - lse), dS^T = P^T (keep dP^T / (1-p) - delta) and writes both bf16 tiles
//               [key][query] (128-byte swizzle) to shared memory.  Nothing but loads until the final tile stores: the
//               compiler is free to batch the table look-ups and exponentials of many elements.
//   warps 8-11  dQ: thread q owns query row q of the tile.  One tile ahead it stages -lse log2(e) and
//               delta = rowsum(dO * O) (computed here from the O / dO rows unless the caller supplies it) in shared
//               memory; then it drains dQ_i = dS K from TMEM into the fp32 accumulator with vector reductions
//               (red.global.add.v4.f32) -- off the softmax warps' critical path
//   warps 12-15 LUT gradient (only with grad_lut): thread r re-reads row r of the finished bf16 dS^T tile and of the
//               transposed code tile (16-byte loads) and adds each element into its private shared-memory bin of that
//               code: [2 streams][24 bins][128 threads] floats, two independent streams so that four read-modify-writes
//               are in flight; one reduction per CTA at the end
//   warp  16    MMA issuer (one lane): S^T = K Q_i^T, dP^T = V dO_i^T (128x128x64), dQ_i = dS K, dV += P^T dO_i,
//               dK += dS^T Q_i (128x64x128)
//   warp  17    TMA producer (one lane): K, V once; Q_i, dO_i and the TRANSPOSED byte-code tile [key][query] double-buffered
//   warps 18-19 idle (register donors: setmaxnreg works per warpgroup)
// TMEM columns: S^T [0,128) dP^T [128,256) dV [256,320) dK [320,384) dQ [384,448) P^T (bf16 pairs) [448,512).
// P^T never touches shared memory: it is the A operand of dV += P^T dO_i read from tensor memory (lane = key row,
// 32-bit column j = queries 2j, 2j+1), written there by the softmax threads with tcgen05.st.  dS^T goes to shared memory
// (dQ_i = dS K reads it MN-major, dK += dS^T Q_i K-major, the LUT-gradient warps re-read it).
// The code tile is read from code_t[b][k][q] = code[b][q][k] (one transpose per batch, shared by every layer and head):
// a thread needs the codes of ONE key against 64 queries, which is a row of the transposed tile (4 x 16-byte loads)
// instead of 64 byte loads strided by the row pitch.
#include "attn_common.cuh"

namespace mdt {
namespace attn {
using namespace sm100;

constexpr int kB2Threads = 640;
constexpr int kB2Bins = 24;  // bias codes 0..22 have their own bin, MDT_CODE_VIRTUAL uses bin 23 (include/mdt_b200.h)
constexpr int kB2Tiles = 16384 * 2 /*K,V*/ + 2 * 16384 /*Q x2*/ + 2 * 16384 /*dO x2*/ + 32768 /*dS^T*/;
constexpr int kB2CodeBytes = 2 * 16384;
constexpr int kB2BinsBytes = 2 * kB2Bins * 128 * 4;

struct B2Ctrl {
    alignas(256) float lut[64];       // lut * log2(e) indexed by (code & 63): 62 = MDT_CODE_VIRTUAL & 63, 63 = MDT_CODE_MASKED & 63 (-inf)
    alignas(16) float nlse[2][128];   // -lse * log2(e) of the tile's queries; -inf for queries >= T
    alignas(16) float delta[2][128];
    alignas(8) uint64_t kv_full, qdo_full[2], qdo_empty[2], sp_full, sp_free, pds_full, pds_free, dq_full, dq_free,
        stats_full[2], stats_free[2], bin_free[2], fin;
    uint32_t tmem_base;
};

__device__ __forceinline__ uint32_t byte_of(uint32_t w, int k) { return (w >> (8 * k)) & 0xffu; }

// HAS_DROP is a template parameter, not a run-time test: a (uniform) branch per element inside the unrolled softmax loop
// cuts it into 64 basic blocks, and the look-ups / exponentials of different elements can then no longer be interleaved.
template <bool HAS_CODE, bool HAS_DROP>
__global__ void __launch_bounds__(kB2Threads, 1)
    attn_bwd2_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                     const __grid_constant__ CUtensorMap mv, const __grid_constant__ CUtensorMap mdo,
                     const __grid_constant__ CUtensorMap mc, const Bwd2Args a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *sK = smem, *sV = smem + 16384, *sQ = smem + 32768, *sdO = smem + 65536, *sdSt = smem + 98304;
    uint8_t *sCode = smem + kB2Tiles;  // 2 x [128 keys][128 queries] bytes, 128-byte swizzle
    const bool want_glut = HAS_CODE && a.grad_lut != nullptr;
    float *bins = reinterpret_cast<float *>(smem + kB2Tiles + (HAS_CODE ? kB2CodeBytes : 0));  // [2][kB2Bins][128]
    B2Ctrl &ctl = *reinterpret_cast<B2Ctrl *>(smem + kB2Tiles + (HAS_CODE ? kB2CodeBytes : 0) + (want_glut ? kB2BinsBytes : 0));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int r = tid & 127;
    const int h = blockIdx.y, b = blockIdx.z;
    const int T = a.T, Tm = a.Tmain;
    const int k0 = blockIdx.x * kTile;
    const int key = k0 + r;
    const bool key_ok = key < Tm && !(a.key_pad != nullptr && a.key_pad[(size_t)b * T + key] != 0);
    const int nq = (T + kTile - 1) / kTile;

    // A key tile without a single live key has dK = dV = 0 and adds nothing to dQ or the LUT gradient.
    if (!__syncthreads_or(key_ok ? 1 : 0)) {
        if (tid < 256 && key < Tm) {
            const int half = tid >> 7;
            uint4 *pk = reinterpret_cast<uint4 *>(a.dk + (size_t)b * a.dk_sb + (size_t)key * a.dk_st + (size_t)h * a.dk_sh + half * 32);
            uint4 *pv = reinterpret_cast<uint4 *>(a.dv + (size_t)b * a.dv_sb + (size_t)key * a.dv_st + (size_t)h * a.dv_sh + half * 32);
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) pk[ch] = pv[ch] = make_uint4(0, 0, 0, 0);
        }
        return;
    }

    if (tid == 0) {
        if (smem_u32(smem) & 1023u) __trap();
        mbar_init(&ctl.kv_full, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&ctl.qdo_full[s], 1);
            mbar_init(&ctl.qdo_empty[s], 1);
            mbar_init(&ctl.stats_full[s], 128);
            mbar_init(&ctl.stats_free[s], 256);
            mbar_init(&ctl.bin_free[s], 128);
        }
        mbar_init(&ctl.sp_full, 1);
        mbar_init(&ctl.sp_free, 256);
        mbar_init(&ctl.pds_full, 256);
        mbar_init(&ctl.pds_free, 1);
        mbar_init(&ctl.dq_full, 1);
        mbar_init(&ctl.dq_free, 128);
        mbar_init(&ctl.fin, 1);
        fence_barrier_init();
        tma_prefetch_desc(&mq);
        tma_prefetch_desc(&mk);
        tma_prefetch_desc(&mv);
        tma_prefetch_desc(&mdo);
        if (HAS_CODE) tma_prefetch_desc(&mc);
    }
    if (warp == 16) {
        tmem_alloc(&ctl.tmem_base, 512);
        tmem_relinquish();
    }
    if (HAS_CODE && tid < 256) {
        if (tid < 64) {
            const int src = tid == 62 ? MDT_CODE_VIRTUAL : tid;
            ctl.lut[tid] = tid == 63 ? neg_inf() : a.lut[(size_t)h * 256 + src] * kLog2e;
        }
        if (want_glut)
            for (int c = 0; c < kB2Bins; ++c) bins[c * 256 + tid] = 0.f;  // 2 x 24 x 128 words in all
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctl.tmem_base;
    constexpr uint32_t cS = 0, cDP = 128, cDV = 256, cDK = 320, cDQ = 384, cPT = 448;

    if (warp >= 16) {
        // ------------------------------------------------------------------ MMA issuer / TMA producer / idle
        reg_dealloc<56>();  // 640 threads x 96 registers at launch = 256 x 152 + 128 x (56 + 64 + 56)
        if (warp == 17 && lane == 0) {
            mbar_expect_tx(&ctl.kv_full, 32768);
            tma_load_4d(sK, &mk, &ctl.kv_full, 0, k0, h, b);
            tma_load_4d(sV, &mv, &ctl.kv_full, 0, k0, h, b);
            for (int i = 0; i < nq; ++i) {
                const int s = i & 1;
                if (i >= 2) {
                    mbar_wait(&ctl.qdo_empty[s], ((i >> 1) - 1) & 1);          // the MMAs of tile i-2 have read Q / dO
                    if (want_glut) mbar_wait(&ctl.bin_free[s], ((i >> 1) - 1) & 1);  // the bin warps have read its code tile
                }
                mbar_expect_tx(&ctl.qdo_full[s], HAS_CODE ? 49152 : 32768);
                tma_load_4d(sQ + s * 16384, &mq, &ctl.qdo_full[s], 0, i * kTile, h, b);
                tma_load_4d(sdO + s * 16384, &mdo, &ctl.qdo_full[s], 0, i * kTile, h, b);
                if (HAS_CODE) tma_load_2d_raw(sCode + s * 16384, &mc, &ctl.qdo_full[s], i * kTile, b * T + k0);  // [key][query]
            }
        } else if (warp == 16 && lane == 0) {
            const uint32_t idesc_s = idesc_bf16(128, 128, 0, 0);
            const uint32_t idesc_kv = idesc_bf16(128, 64, 0, 1);
            const uint32_t idesc_dq = idesc_bf16(128, 64, 1, 1);
            const uint32_t ak = smem_u32(sK), av = smem_u32(sV), ad = smem_u32(sdSt);
            auto issue_s_dp = [&](int i) {
                const uint32_t bq = smem_u32(sQ + (i & 1) * 16384), bo = smem_u32(sdO + (i & 1) * 16384);
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16(tmem + cS, smem_desc_sw128(ak + k * 32, 16, 1024), smem_desc_sw128(bq + k * 32, 16, 1024), idesc_s, k > 0);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16(tmem + cDP, smem_desc_sw128(av + k * 32, 16, 1024), smem_desc_sw128(bo + k * 32, 16, 1024), idesc_s, k > 0);
                umma_commit(&ctl.sp_full);
            };
            mbar_wait(&ctl.kv_full, 0);
            mbar_wait(&ctl.qdo_full[0], 0);
            issue_s_dp(0);
            for (int i = 0; i < nq; ++i) {
                if (i + 1 < nq) {  // S^T / dP^T of the next tile: computed under the softmax arithmetic of this one
                    mbar_wait(&ctl.sp_free, i & 1);
                    mbar_wait(&ctl.qdo_full[(i + 1) & 1], ((i + 1) >> 1) & 1);
                    issue_s_dp(i + 1);
                }
                mbar_wait(&ctl.pds_full, i & 1);
                if (i > 0) mbar_wait(&ctl.dq_free, (i - 1) & 1);
                tc_fence_after();
                const uint32_t bo = smem_u32(sdO + (i & 1) * 16384), bq = smem_u32(sQ + (i & 1) * 16384);
                const uint32_t acc = i > 0 ? 1u : 0u;
#pragma unroll
                for (int k = 0; k < 8; ++k)  // dQ_i = dS K   (A: the dS^T tile read MN-major, k-step = 16 keys)
                    umma_bf16(tmem + cDQ, smem_desc_sw128(ad + k * 2048, 16384, 1024), smem_desc_sw128(ak + k * 2048, 8192, 1024),
                              idesc_dq, k > 0);
                umma_commit(&ctl.dq_full);
#pragma unroll
                for (int k = 0; k < 8; ++k)  // dV += P^T dO_i   (A = P^T from tensor memory: 8 columns per 16 queries)
                    umma_bf16_ts(tmem + cDV, tmem + cPT + k * 8, smem_desc_sw128(bo + k * 2048, 8192, 1024), idesc_kv, (k > 0) ? 1u : acc);
#pragma unroll
                for (int k = 0; k < 8; ++k)  // dK += dS^T Q_i
                    umma_bf16(tmem + cDK, smem_desc_sw128(ad + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                              smem_desc_sw128(bq + k * 2048, 8192, 1024), idesc_kv, (k > 0) ? 1u : acc);
                umma_commit(&ctl.pds_free);            // P^T / dS^T may be rewritten
                umma_commit(&ctl.qdo_empty[i & 1]);    // Q_i / dO_i may be reloaded
                if (i + 1 == nq) umma_commit(&ctl.fin);
            }
        }
    } else if (warp >= 12) {
        // ------------------------------------------------------------------ LUT-gradient warpgroup
        reg_dealloc<64>();
        if (want_glut) {
            float *b0 = bins + r, *b1 = bins + kB2Bins * 128 + r;   // two independent streams
            for (int i = 0; i < nq; ++i) {
                mbar_wait(&ctl.pds_full, i & 1);  // dS^T of tile i is in shared memory
                mbar_wait(&ctl.qdo_full[i & 1], (i >> 1) & 1);  // (its code tile has long landed: acquire)
                const uint32_t crow = smem_u32(sCode + (i & 1) * 16384);
                const uint32_t drow = smem_u32(sdSt);
                if (key_ok) {
#pragma unroll 2
                    for (int g = 0; g < 8; ++g) {  // 16 queries per step: one 16-byte code chunk, two 16-byte dS chunks
                        const uint4 cw = lds_u4(crow + sw128_offset(r, g));
                        const uint32_t cws[4] = {cw.x, cw.y, cw.z, cw.w};
#pragma unroll
                        for (int hh = 0; hh < 2; ++hh) {
                            const int ch = 2 * g + hh;  // 8-query chunk of the row: half ch >> 3, chunk ch & 7
                            const uint4 dw = lds_u4(drow + (ch >> 3) * 16384 + sw128_offset(r, ch & 7));
                            const uint32_t dws[4] = {dw.x, dw.y, dw.z, dw.w};
#pragma unroll
                            for (int pr = 0; pr < 4; pr += 2) {  // 4 elements: two per stream, loads first
                                const uint32_t w0 = dws[pr], w1 = dws[pr + 1], cc = cws[hh * 2 + (pr >> 1)];
                                const float d0 = __uint_as_float(w0 << 16), d1 = __uint_as_float(w0 & 0xffff0000u);
                                const float d2 = __uint_as_float(w1 << 16), d3 = __uint_as_float(w1 & 0xffff0000u);
                                // spatial codes 0..22 keep their bin, wider ones fold into bin 22 (include/mdt_b200.h),
                                // MDT_CODE_VIRTUAL (and the masked code, whose dS is exactly 0) go to bin 23 -- four bytes at once
                                const uint32_t nb = __vminu4(cc, 0x16161616u) + (__vcmpgeu4(cc, 0xfefefefeu) & 0x01010101u);
                                const uint32_t n0 = nb & 0xffu, n1 = (nb >> 8) & 0xffu, n2 = (nb >> 16) & 0xffu, n3 = nb >> 24;
                                float *p0 = b0 + n0 * 128, *p1 = b0 + n1 * 128, *p2 = b1 + n2 * 128, *p3 = b1 + n3 * 128;
                                const float v0 = *p0, v1 = *p1, v2 = *p2, v3 = *p3;
                                const float a0 = v0 + d0, a2 = v2 + d2;
                                const float a1 = (n0 == n1 ? a0 : v1) + d1, a3 = (n2 == n3 ? a2 : v3) + d3;
                                *p0 = a0;
                                *p1 = a1;
                                *p2 = a2;
                                *p3 = a3;
                            }
                        }
                    }
                }
                mbar_arrive(&ctl.bin_free[i & 1]);
            }
        }
    } else if (warp >= 8) {
        // ------------------------------------------------------------------ dQ warpgroup
        reg_dealloc<56>();
        const int ql = tid - 256;  // query row inside the tile == TMEM lane
        const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        auto stage_stats = [&](int i) {  // -lse log2(e) and delta of query tile i -> shared memory
            const int qq = i * kTile + ql;
            float nl = neg_inf(), dl = 0.f;
            if (qq < T) {
                const size_t row = ((size_t)b * a.H + h) * T + qq;
                nl = -a.lse[row] * kLog2e;
                if (a.delta != nullptr) {
                    dl = a.delta[row];
                } else {
                    const uint4 *po = reinterpret_cast<const uint4 *>(a.out + (size_t)b * a.o_sb + (size_t)qq * a.o_st + (size_t)h * a.o_sh);
                    const uint4 *pg = reinterpret_cast<const uint4 *>(a.dout + (size_t)b * a.g_sb + (size_t)qq * a.g_st + (size_t)h * a.g_sh);
#pragma unroll
                    for (int ch = 0; ch < 8; ++ch) {
                        const uint4 x = __ldg(po + ch), y = __ldg(pg + ch);
                        const __nv_bfloat162 *px = reinterpret_cast<const __nv_bfloat162 *>(&x);
                        const __nv_bfloat162 *py = reinterpret_cast<const __nv_bfloat162 *>(&y);
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float2 fx = __bfloat1622float2(px[e]), fy = __bfloat1622float2(py[e]);
                            dl = fmaf(fx.x, fy.x, dl);
                            dl = fmaf(fx.y, fy.y, dl);
                        }
                    }
                }
            }
            ctl.nlse[i & 1][ql] = nl;
            ctl.delta[i & 1][ql] = dl;
            mbar_arrive(&ctl.stats_full[i & 1]);
        };
        stage_stats(0);
        for (int i = 0; i < nq; ++i) {
            if (i + 1 < nq) {
                if (i + 1 >= 2) mbar_wait(&ctl.stats_free[(i + 1) & 1], (((i + 1) >> 1) - 1) & 1);
                stage_stats(i + 1);
            }
            mbar_wait(&ctl.dq_full, i & 1);
            tc_fence_after();
            __syncwarp();
            const int qq = i * kTile + ql;
            float4 *dst = reinterpret_cast<float4 *>(a.dq_accum + (((size_t)b * a.H + h) * T + (qq < T ? qq : 0)) * 64);
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {  // two 32-column halves: 32 live registers at a time
                float dq[32];
                tmem_ld_32x32(lane_base + cDQ + hh * 32, reinterpret_cast<uint32_t *>(dq));
                tmem_wait_ld();
                if (hh == 1) {
                    tc_fence_before();
                    mbar_arrive(&ctl.dq_free);  // the dQ columns may be overwritten by the next tile's MMAs
                }
                if (qq < T) {
#pragma unroll
                    for (int u = 0; u < 8; ++u) atomicAdd(dst + hh * 8 + u, make_float4(dq[u * 4], dq[u * 4 + 1], dq[u * 4 + 2], dq[u * 4 + 3]));
                }
            }
        }
    } else {
        // ------------------------------------------------------------------ softmax warpgroups
        reg_alloc<152>();
        const int half = tid >> 7;  // which 64 query columns of every tile
        const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        const uint32_t salt = drop_salt(a.drop, (uint32_t)(b * a.H + h)), thr32 = a.drop.thr16 << 16;
        const uint32_t drop_k = drop_tile(salt, (uint32_t)key >> 6, 1u) ^ drop_lane(64u + ((uint32_t)key & 63u));
        const bool warp_keys = __any_sync(0xffffffffu, key_ok);  // a warp whose 32 keys are all masked only writes zeros
        const uint32_t lut_base = smem_u32(ctl.lut);

        for (int i = 0; i < nq; ++i) {
            const int q0 = i * kTile;
            const int qvalid = min(kTile, T - q0);
            const bool live = warp_keys && half * 64 < qvalid;   // warp-uniform
            const uint32_t drop_w = HAS_DROP ? drop_k ^ drop_tile(salt, (uint32_t)(i * 2 + half), 0u) : 0u;  // 64 queries of this half
            float st[64], dp[64];
            mbar_wait(&ctl.sp_full, i & 1);
            tc_fence_after();
            __syncwarp();
            tmem_ld_32x32(lane_base + cS + half * 64, reinterpret_cast<uint32_t *>(st));
            tmem_ld_32x32(lane_base + cS + half * 64 + 32, reinterpret_cast<uint32_t *>(st) + 32);
            tmem_ld_32x32(lane_base + cDP + half * 64, reinterpret_cast<uint32_t *>(dp));
            tmem_ld_32x32(lane_base + cDP + half * 64 + 32, reinterpret_cast<uint32_t *>(dp) + 32);
            tmem_wait_ld();
            tc_fence_before();
            mbar_arrive(&ctl.sp_free);  // S^T / dP^T of the next tile may overwrite the columns
            mbar_wait(&ctl.stats_full[i & 1], (i >> 1) & 1);
            if (HAS_CODE) mbar_wait(&ctl.qdo_full[i & 1], (i >> 1) & 1);  // the code tile was written by TMA
            if (live && key_ok) {
                const uint32_t nl = smem_u32(ctl.nlse[i & 1] + half * 64), dl = smem_u32(ctl.delta[i & 1] + half * 64);
                const uint32_t crow = smem_u32(sCode + (i & 1) * 16384);
#pragma unroll
                for (int g = 0; g < 4; ++g) {  // 16 query columns per step: one 16-byte chunk of this key's code row
                    uint4 cw = make_uint4(0, 0, 0, 0);
                    if (HAS_CODE) cw = lds_u4(crow + sw128_offset(r, half * 4 + g));
                    const uint32_t cws[4] = {cw.x, cw.y, cw.z, cw.w};
#pragma unroll
                    for (int v4 = 0; v4 < 4; ++v4) {
                        const uint4 n4 = lds_u4(nl + (g * 16 + v4 * 4) * 4), d4 = lds_u4(dl + (g * 16 + v4 * 4) * 4);
                        float *x = st + g * 16 + v4 * 4, *y = dp + g * 16 + v4 * 4;  // 4 query columns
                        if (HAS_CODE) {
                            // byte -> 4 * (code & 63), merged into the 256-byte aligned table address by PRMT
                            const uint32_t wm = (cws[v4] << 2) & 0xfcfcfcfcu;
                            const float l0 = lds_f32(__byte_perm(wm, lut_base, 0x7650)), l1 = lds_f32(__byte_perm(wm, lut_base, 0x7651));
                            const float l2 = lds_f32(__byte_perm(wm, lut_base, 0x7652)), l3 = lds_f32(__byte_perm(wm, lut_base, 0x7653));
                            ffma2(x[0], x[1], x[0], x[1], kLog2e, kLog2e, l0, l1);
                            ffma2(x[2], x[3], x[2], x[3], kLog2e, kLog2e, l2, l3);
                            fadd2(x[0], x[1], x[0], x[1], __uint_as_float(n4.x), __uint_as_float(n4.y));
                            fadd2(x[2], x[3], x[2], x[3], __uint_as_float(n4.z), __uint_as_float(n4.w));
                        } else {
                            ffma2(x[0], x[1], x[0], x[1], kLog2e, kLog2e, __uint_as_float(n4.x), __uint_as_float(n4.y));
                            ffma2(x[2], x[3], x[2], x[3], kLog2e, kLog2e, __uint_as_float(n4.z), __uint_as_float(n4.w));
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) x[u] = ex2(x[u]);   // P^T
                        if (HAS_DROP) {  // dropout on P: dV takes keep * P / (1 - p_drop), dS = P * (keep * dP / (1 - p_drop) - delta)
                            float kf[4];
#pragma unroll
                            for (int u = 0; u < 4; ++u)
                                kf[u] = drop_keep_w(drop_w, drop_lane(g * 16 + v4 * 4 + u), thr32) ? a.drop.scale : 0.f;
                            fmul2(y[0], y[1], y[0], y[1], kf[0], kf[1]);
                            fmul2(y[2], y[3], y[2], y[3], kf[2], kf[3]);
                            fadd2(y[0], y[1], y[0], y[1], -__uint_as_float(d4.x), -__uint_as_float(d4.y));
                            fadd2(y[2], y[3], y[2], y[3], -__uint_as_float(d4.z), -__uint_as_float(d4.w));
                            fmul2(y[0], y[1], y[0], y[1], x[0], x[1]);   // dS^T = P (keep dP - delta)
                            fmul2(y[2], y[3], y[2], y[3], x[2], x[3]);
                            fmul2(x[0], x[1], x[0], x[1], kf[0], kf[1]);  // what dV sees
                            fmul2(x[2], x[3], x[2], x[3], kf[2], kf[3]);
                        } else {
                            fadd2(y[0], y[1], y[0], y[1], -__uint_as_float(d4.x), -__uint_as_float(d4.y));
                            fadd2(y[2], y[3], y[2], y[3], -__uint_as_float(d4.z), -__uint_as_float(d4.w));
                            fmul2(y[0], y[1], y[0], y[1], x[0], x[1]);   // dS^T = P (dP - delta)
                            fmul2(y[2], y[3], y[2], y[3], x[2], x[3]);
                        }
                    }
                }
            }
            mbar_arrive(&ctl.stats_free[i & 1]);
            if (i > 0) {
                mbar_wait(&ctl.pds_free, (i - 1) & 1);                      // the MMAs of tile i-1 have read the P^T / dS^T tiles
                if (want_glut) mbar_wait(&ctl.bin_free[(i - 1) & 1], ((i - 1) >> 1) & 1);  // ... and so have the bin warps
            }
            {
                // masked key, or query columns past T: P^T = dS^T = 0.  The TMEM stores are warp-collective, so every lane
                // executes them and the per-lane case is a select, not a branch.
                const bool on = live && key_ok;
                const uint32_t tP = lane_base + cPT + half * 32;
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    uint32_t pw[16];
#pragma unroll
                    for (int u = 0; u < 16; ++u) pw[u] = on ? pack_bf16x2(st[hh * 32 + 2 * u], st[hh * 32 + 2 * u + 1]) : 0u;
                    tmem_st_32x16(tP + hh * 16, pw);
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    uint4 vd = make_uint4(0, 0, 0, 0);
                    if (on) {
                        vd.x = pack_bf16x2(dp[u * 8 + 0], dp[u * 8 + 1]);
                        vd.y = pack_bf16x2(dp[u * 8 + 2], dp[u * 8 + 3]);
                        vd.z = pack_bf16x2(dp[u * 8 + 4], dp[u * 8 + 5]);
                        vd.w = pack_bf16x2(dp[u * 8 + 6], dp[u * 8 + 7]);
                    }
                    *reinterpret_cast<uint4 *>(sdSt + half * 16384 + sw128_offset(r, u)) = vd;
                }
                tmem_wait_st();
                tc_fence_before();
            }
            fence_proxy_async();
            mbar_arrive(&ctl.pds_full);
        }

        // ---- epilogue: dK, dV rows of this key tile; each half stores 32 columns
        mbar_wait(&ctl.fin, 0);
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
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 16) tmem_dealloc(tmem, 512);
    if (want_glut && tid < kB2Bins) {
        // bin `tid`: sum over the 2 x 128 per-thread copies (rotated start keeps the banks distinct)
        float sum = 0.f;
        for (int t = 0; t < 128; ++t) {
            const int tt = (t + tid) & 127;
            sum += bins[tid * 128 + tt] + bins[(kB2Bins + tid) * 128 + tt];
        }
        const int code = tid == kB2Bins - 1 ? MDT_CODE_VIRTUAL : tid;
        if (sum != 0.f) atomicAdd(a.grad_lut + (size_t)h * 256 + code, sum);
    }
}

int launch_bwd2(const CUtensorMap &mq, const CUtensorMap &mk, const CUtensorMap &mv, const CUtensorMap &mdo, const CUtensorMap &mc,
                const Bwd2Args &a, bool has_code, cudaStream_t st) {
    const int smem = kB2Tiles + (has_code ? kB2CodeBytes : 0) + ((has_code && a.grad_lut) ? kB2BinsBytes : 0) + (int)sizeof(B2Ctrl);
    dim3 grid((a.Tmain + kTile - 1) / kTile, a.H, a.B);
    const bool drop = a.drop.thr16 != 0;
#define MDT_LAUNCH_BWD2(C, D)                                                                                                  \
    do {                                                                                                                       \
        MDT_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd2_kernel<C, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));      \
        attn_bwd2_kernel<C, D><<<grid, kB2Threads, smem, st>>>(mq, mk, mv, mdo, mc, a);                                        \
    } while (0)
    if (has_code && drop) MDT_LAUNCH_BWD2(true, true);
    else if (has_code) MDT_LAUNCH_BWD2(true, false);
    else if (drop) MDT_LAUNCH_BWD2(false, true);
    else MDT_LAUNCH_BWD2(false, false);
#undef MDT_LAUNCH_BWD2
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

}  // namespace attn
}  // namespace mdt

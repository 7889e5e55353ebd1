// A6 forward, persistent + warp-specialised (bias synthesised from byte codes, or no bias)   multihead_attention.py:168-199
//
// One CTA per SM loops over work items (batch b, head h, 256 queries = two 128-query tiles sharing every K/V tile).
//   warps 0-7   softmax of query tile 0: warpgroup 0 owns key columns [0,64) of every S tile, warpgroup 1 columns
//               [64,128); thread r of either <-> query row r <-> TMEM lane r; the two halves of a row exchange
//               their maximum through shared memory (named barrier per tile)
//   warps 8-15  softmax of query tile 1, same split -- 16 softmax warps = 4 per scheduler hide the LDS / MUFU /
//               tcgen05.ld latencies that 8 warps could not (issue utilisation was 30 %)
//   warp  16    TMA producer: per item Q0/Q1, per key tile K, V and the two 128x128 byte-code tiles
//   warp  17    MMA issuer:   S_t = Q_t K^T (128x128x64) and O_t += P_t V (128x64x128), tcgen05 -> TMEM
//   warp  18    aux: one item ahead, builds the key-mask bitmap of batch b, the number of live key tiles and the
//               64-entry log2-domain LUT of head h in shared memory
//   warp  19    idle (setmaxnreg works per warpgroup)
// The k-steps of all items form ONE sequence: the K/V/code ring, the S/P/O hand-offs and their mbarrier parities
// run straight across item boundaries, so the loads, the first S MMA and the prologue of item n+1 overlap the
// softmax and the epilogue of item n.  S_t(g+1) is issued as soon as the softmax warpgroup has pulled S_t(g)
// into registers, so it executes under softmax(g).
// TMEM columns: S0 [0,128) S1 [128,256) O0 [256,320) O1 [320,384) P0 [384,448) P1 [448,512).
// P never touches shared memory: the softmax threads pack their probabilities to bf16 pairs and store them into tensor
// memory (tcgen05.st; lane = query row, 32-bit column j = keys 2j, 2j+1), and O_t += P_t V reads its A operand from there
// (tcgen05.mma with A in TMEM).  That removes 32 KB of shared-memory stores and 32 KB of operand fetches per 128x128 tile
// -- a third of the kernel's shared-memory traffic, which (not the tensor pipe) bounds it at head dim 64.
// Shared memory (all dynamic, 1024-byte aligned): Q0 Q1 | 2 stages x (K, V, code0, code1) | control.
#include "attn_common.cuh"

namespace mdt {
namespace attn {
using namespace sm100;

constexpr int kF3Threads = 640;
constexpr int kF3SoftmaxWarps = 16;
constexpr int kF3Stage = 4 * 16384;  // K, V, code0, code1
constexpr int kF3Tiles = 2 * 16384 + 2 * kF3Stage;
constexpr float kF3Rescale = 8.0f;   // log2 units: P may reach 2^8 before O is rescaled
constexpr int kF3MaskWords = kFwd3MaxT / 32;

struct F3Ctrl {
    // 64-entry table in log2 units, indexed by (code & 63): spatial codes 0..61, 62 = MDT_CODE_VIRTUAL & 63,
    // 63 = MDT_CODE_MASKED & 63 (-inf).  Codes 62..253 are reserved (the collate kernel emits <= 21).
    // Each table is 256-byte aligned so that one PRMT merges a byte offset into its address.
    float lut[2][64];
    uint64_t q_full, q_empty, k_full[2], v_full[2], c_full[2], k_empty[2], v_empty[2], c_empty[2];
    uint64_t s_full[2], s_free[2], p_full[2], o_done[2], aux_full[2], aux_empty[2];
    uint32_t tmem_base;
    int nkv[2];
    uint32_t kmask[2][kF3MaskWords];  // bit set = key masked (padding or >= T)
    float xmax[2][2][2][128];         // [k-step parity][query tile][column half][row]: row maximum exchange between the halves
    float xsum[2][2][2][128];         // [item parity][query tile][column half][row]: the row sums, once per item
};
constexpr int kF3Smem = kF3Tiles + (int)sizeof(F3Ctrl);

// diagnostics: 64 stamps per CTA, slot 2 + 12 * item + k for the first four items of the CTA
__device__ __forceinline__ void f3_trace(const Fwd2Args &a, int slot) {
    if (a.trace != nullptr && slot < 64) {
        unsigned long long tm;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tm));
        a.trace[(size_t)blockIdx.x * 64 + slot] = tm;
    }
}

struct F3Item {
    int b, h, q_base;
};
__device__ __forceinline__ F3Item f3_item(int item, int nqp, int H) {
    F3Item it;
    const int bh = item / nqp;
    it.q_base = (item - bh * nqp) * 2 * kTile;
    it.b = bh / H;
    it.h = bh - it.b * H;
    return it;
}

// HAS_DROP is a template parameter: a run-time (uniform) test per element pair would cut the unrolled softmax loop into
// basic blocks and keep the exponentials of different elements from being interleaved.
template <bool HAS_CODE, bool HAS_DROP>
__global__ void __launch_bounds__(kF3Threads, 1)
    attn_fwd3_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                     const __grid_constant__ CUtensorMap mv, const __grid_constant__ CUtensorMap mc, const Fwd2Args a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *sQ = smem;                       // 2 x 16 KB
    uint8_t *sStage = smem + 2 * 16384;       // 2 x 64 KB
    F3Ctrl &ctl = *reinterpret_cast<F3Ctrl *>(smem + kF3Tiles);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int T = a.T, H = a.H, Tm = a.Tmain;  // T: row pitch of lse / code / key_pad; Tm: queries and keys of this kernel
    const int nqp = (Tm + 2 * kTile - 1) / (2 * kTile);
    const int n_items = a.B * H * nqp;
    const int nkv_all = (Tm + kTile - 1) / kTile;

    if (tid == 0) {
        f3_trace(a, 0);
        if (smem_u32(smem) & 1023u) __trap();  // the 128-byte swizzle atoms need 1024-byte aligned tiles
        mbar_init(&ctl.q_full, 1);
        mbar_init(&ctl.q_empty, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&ctl.k_full[s], 1);
            mbar_init(&ctl.v_full[s], 1);
            mbar_init(&ctl.c_full[s], 1);
            mbar_init(&ctl.k_empty[s], 1);
            mbar_init(&ctl.v_empty[s], 1);
            mbar_init(&ctl.c_empty[s], 512);
            mbar_init(&ctl.s_full[s], 1);
            mbar_init(&ctl.s_free[s], 256);
            mbar_init(&ctl.p_full[s], 256);
            mbar_init(&ctl.o_done[s], 1);
            mbar_init(&ctl.aux_full[s], 32);
            mbar_init(&ctl.aux_empty[s], 514);  // 512 softmax threads + producer + MMA issuer
        }
        fence_barrier_init();
        tma_prefetch_desc(&mq);
        tma_prefetch_desc(&mk);
        tma_prefetch_desc(&mv);
        if (HAS_CODE) tma_prefetch_desc(&mc);
    }
    if (warp == kF3SoftmaxWarps) {
        tmem_alloc(&ctl.tmem_base, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctl.tmem_base;
    if (tid == 0) f3_trace(a, 1);

    if (warp >= kF3SoftmaxWarps) {
        // ------------------------------------------------------------------ producer / MMA / aux warpgroup
        reg_dealloc<64>();  // 640 x 96 registers at launch = 512 x 104 + 128 x 64
        if (warp == kF3SoftmaxWarps + 2) {
            // ---- aux: key-mask bitmap, live key tiles and LUT of item n into buffer n & 1, one item ahead
            int n = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++n) {
                const int buf = n & 1;
                const F3Item it = f3_item(item, nqp, H);
                if (n >= 2) mbar_wait(&ctl.aux_empty[buf], ((n >> 1) - 1) & 1);
                // lane l builds words l, l + 32, ...: 32 independent byte loads per word, no cross-lane traffic
                int last = -1;
                const uint8_t *kp = a.key_pad ? a.key_pad + (size_t)it.b * T : nullptr;
                for (int w = lane; w < nkv_all * 4; w += 32) {
                    uint32_t bits = 0;
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const int kk = w * 32 + i;
                        const bool masked = kk >= Tm || (kp != nullptr && kp[kk] != 0);
                        bits |= (masked ? 1u : 0u) << i;
                    }
                    ctl.kmask[buf][w] = bits;
                    if (bits != 0xffffffffu) last = w;
                }
                last = __reduce_max_sync(0xffffffffu, last);
                if (lane == 0) ctl.nkv[buf] = last < 0 ? 1 : (last >> 2) + 1;  // later key tiles contribute exactly 0
                if (HAS_CODE) {
                    for (int e = lane; e < 64; e += 32) {
                        const int src = e == 62 ? MDT_CODE_VIRTUAL : e;
                        ctl.lut[buf][e] = e == 63 ? neg_inf() : a.lut[(size_t)it.h * 256 + src] * kLog2e;
                    }
                }
                mbar_arrive(&ctl.aux_full[buf]);
            }
        } else if (warp == kF3SoftmaxWarps && lane == 0) {
            // ---- TMA producer
            int n = 0;
            uint32_t g = 0;  // k-step sequence number across items
            for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++n) {
                const int buf = n & 1;
                const F3Item it = f3_item(item, nqp, H);
                mbar_wait(&ctl.aux_full[buf], (n >> 1) & 1);
                const int nkv = ctl.nkv[buf];
                mbar_arrive(&ctl.aux_empty[buf]);
                if (n >= 1) mbar_wait(&ctl.q_empty, (n - 1) & 1);
                mbar_expect_tx(&ctl.q_full, 32768);
                tma_load_4d(sQ, &mq, &ctl.q_full, 0, it.q_base, it.h, it.b);
                tma_load_4d(sQ + 16384, &mq, &ctl.q_full, 0, it.q_base + kTile, it.h, it.b);
                f3_trace(a, 2 + 12 * n + 6);
                for (int j = 0; j < nkv; ++j, ++g) {
                    const int s = g & 1, k0 = j * kTile;
                    uint8_t *st = sStage + s * kF3Stage;
                    const bool reuse = g >= 2;
                    const uint32_t par = ((g >> 1) - 1) & 1;
                    // K, code and V slots are released separately: K right after both S MMAs of the step, the code
                    // tiles after pass 1 of both softmax warpgroups, V only after both P V MMAs.
                    if (reuse) mbar_wait(&ctl.k_empty[s], par);
                    mbar_expect_tx(&ctl.k_full[s], 16384);
                    tma_load_4d(st, &mk, &ctl.k_full[s], 0, k0, it.h, it.b);
                    if (HAS_CODE) {
                        if (reuse) mbar_wait(&ctl.c_empty[s], par);
                        mbar_expect_tx(&ctl.c_full[s], 32768);
                        tma_load_2d_raw(st + 32768, &mc, &ctl.c_full[s], k0, it.b * T + it.q_base);
                        tma_load_2d_raw(st + 49152, &mc, &ctl.c_full[s], k0, it.b * T + it.q_base + kTile);
                    }
                    if (reuse) mbar_wait(&ctl.v_empty[s], par);
                    mbar_expect_tx(&ctl.v_full[s], 16384);
                    tma_load_4d(st + 16384, &mv, &ctl.v_full[s], 0, k0, it.h, it.b);
                    if (j == 0) f3_trace(a, 2 + 12 * n + 7);
                }
            }
        } else if (warp == kF3SoftmaxWarps + 1 && lane == 0) {
            // ---- MMA issuer.  Per k-step g three jobs are pending: S_0/S_1 of step g+1 (need S_t(g) pulled into
            // registers, the K tile and -- across an item boundary -- the next Q) and P_t V of step g for both tiles
            // (need P_t from the softmax).  They are issued in whatever order their inputs arrive (non-blocking
            // mbarrier tests), so a late load of the next item never holds back P V of the current one.
            const uint32_t idesc_s = idesc_bf16(128, 128, 0, 0);
            const uint32_t idesc_o = idesc_bf16(128, 64, 0, 1);
            auto mma_s = [&](int t, int stage) {
                tc_fence_after();
                const uint32_t aq = smem_u32(sQ + t * 16384), bk = smem_u32(sStage + stage * kF3Stage);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_bf16(tmem + t * 128, smem_desc_sw128(aq + k * 32, 16, 1024), smem_desc_sw128(bk + k * 32, 16, 1024),
                              idesc_s, k > 0);
                umma_commit(&ctl.s_full[t]);
            };
            int n = 0;
            uint32_t g = 0;
            int nkv = 0;
            if (blockIdx.x < n_items) {  // first S of the first item
                mbar_wait(&ctl.aux_full[0], 0);
                nkv = ctl.nkv[0];
                mbar_arrive(&ctl.aux_empty[0]);
                mbar_wait(&ctl.q_full, 0);
                mbar_wait(&ctl.k_full[0], 0);
                mma_s(0, 0);
                mma_s(1, 0);
                umma_commit(&ctl.k_empty[0]);
                if (nkv == 1) umma_commit(&ctl.q_empty);
            }
            for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++n) {
                const bool has_next = item + (int)gridDim.x < n_items;
                int nkv_next = 0;
                for (int j = 0; j < nkv; ++j, ++g) {
                    const int s = g & 1, s1 = (g + 1) & 1;
                    const bool same_item = j + 1 < nkv;
                    const bool want_s = same_item || has_next;
                    // readiness of S(g+1): 0 aux of the next item, 1 its Q, 2 the K tile, 3 ready
                    int s_stage = want_s ? (same_item ? 2 : 0) : 4;
                    bool s_done[2] = {!want_s, !want_s}, pv_done[2] = {false, false}, v_ready = false;
                    long long t0 = 0;
                    while (!(s_done[0] && s_done[1] && pv_done[0] && pv_done[1])) {
                        if (s_stage == 0 && mbar_try_wait(&ctl.aux_full[(n + 1) & 1], ((n + 1) >> 1) & 1)) {
                            nkv_next = ctl.nkv[(n + 1) & 1];
                            mbar_arrive(&ctl.aux_empty[(n + 1) & 1]);
                            s_stage = 1;
                        }
                        if (s_stage == 1 && mbar_try_wait(&ctl.q_full, (n + 1) & 1)) {
                            f3_trace(a, 2 + 12 * (n + 1) + 8);
                            s_stage = 2;
                        }
                        if (s_stage == 2 && mbar_try_wait(&ctl.k_full[s1], ((g + 1) >> 1) & 1)) s_stage = 3;
                        if (s_stage == 3) {
#pragma unroll
                            for (int t = 0; t < 2; ++t) {
                                if (!s_done[t] && mbar_try_wait(&ctl.s_free[t], g & 1)) {
                                    mma_s(t, s1);
                                    s_done[t] = true;
                                    if (s_done[t ^ 1]) {
                                        umma_commit(&ctl.k_empty[s1]);
                                        // Q0/Q1 are free after the last S of their item
                                        if (same_item ? (j + 2 == nkv) : (nkv_next == 1)) umma_commit(&ctl.q_empty);
                                    }
                                }
                            }
                        }
                        if (!v_ready) v_ready = mbar_try_wait(&ctl.v_full[s], (g >> 1) & 1);
                        if (v_ready) {
#pragma unroll
                            for (int t = 0; t < 2; ++t) {
                                if (!pv_done[t] && mbar_try_wait(&ctl.p_full[t], g & 1)) {
                                    tc_fence_after();
                                    const uint32_t bv = smem_u32(sStage + s * kF3Stage + 16384);
#pragma unroll
                                    for (int k = 0; k < 8; ++k)  // A = P_t from tensor memory: 8 columns per 16 keys
                                        umma_bf16_ts(tmem + 256 + t * 64, tmem + 384 + t * 64 + k * 8,
                                                     smem_desc_sw128(bv + k * 2048, 8192, 1024), idesc_o, (j > 0 || k > 0) ? 1u : 0u);
                                    umma_commit(&ctl.o_done[t]);
                                    if (j == 0 && t == 0) f3_trace(a, 2 + 12 * n + 9);
                                    pv_done[t] = true;
                                    if (pv_done[t ^ 1]) umma_commit(&ctl.v_empty[s]);
                                }
                            }
                        }
                        // bounded: a pipeline bug traps instead of hanging the GPU
                        if (t0 == 0) t0 = clock64();
                        else if (clock64() - t0 > 4000000000LL) __trap();
                    }
                }
                nkv = nkv_next;
            }
        }
    } else {
        // ------------------------------------------------------------------ softmax warpgroups
        reg_alloc<104>();  // funded by the producer warpgroup: 4 x 128 x 8 <= 128 x (96 - 40)
        const int wg = warp >> 2;
        const int t = wg >> 1;                         // query tile of this warpgroup pair
        const int hf = wg & 1;                         // key-column half [64 hf, 64 hf + 64) of every S tile
        const int r = tid & 127;                       // row inside the tile == TMEM lane
        const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        const uint32_t tS = lane_base + t * 128 + hf * 64, tO = lane_base + 256 + t * 64 + hf * 32;
        const uint32_t tP = lane_base + 384 + t * 64 + hf * 32;       // this half = keys [64 hf, 64 hf + 64) of P_t, bf16 pairs
        const uint32_t bar_id = 1 + t;                 // named barrier of the two warpgroups sharing tile t
        int n = 0;
        uint32_t g = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++n) {
            const int buf = n & 1;
            const F3Item it = f3_item(item, nqp, H);
            const int q = it.q_base + t * kTile + r;
            const bool row_ok = q < Tm;
            const bool warp_live = __any_sync(0xffffffffu, row_ok);  // warps past the end only keep the barriers moving
            mbar_wait(&ctl.aux_full[buf], (n >> 1) & 1);
            const int nkv = ctl.nkv[buf];
            if (tid == 0) f3_trace(a, 2 + 12 * n + 0);
            const uint32_t lut_base = smem_u32(ctl.lut[buf]);
            float m_used = neg_inf(), l = 0.f;   // m_used in log2 units (scores are scaled by log2 e on the fly)
            const uint32_t salt = drop_salt(a.drop, (uint32_t)(it.b * H + it.h)), thr32 = a.drop.thr16 << 16;
            const uint32_t drop_q = HAS_DROP ? drop_tile(salt, (uint32_t)q >> 6, 0u) ^ drop_lane((uint32_t)q & 63u) : 0u;

            for (int j = 0; j < nkv; ++j, ++g) {
                const int s = g & 1;
                uint32_t kmask[2];
#pragma unroll
                for (int c = 0; c < 2; ++c) kmask[c] = ctl.kmask[buf][j * 4 + hf * 2 + c];
                const uint32_t drop_w = HAS_DROP ? drop_q ^ drop_tile(salt, (uint32_t)(j * 2 + hf), 1u) : 0u;  // 64 keys of this half
                float sc[64];
                mbar_wait(&ctl.s_full[t], g & 1);
                if (tid == 0 && j == 0) f3_trace(a, 2 + 12 * n + 1);
                tc_fence_after();
                __syncwarp();
#pragma unroll
                for (int c = 0; c < 2; ++c) tmem_ld_32x32(tS + c * 32, reinterpret_cast<uint32_t *>(sc) + c * 32);
                tmem_wait_ld();
                tc_fence_before();
                mbar_arrive(&ctl.s_free[t]);              // S_t may be overwritten by the next Q_t K^T
                if (HAS_CODE) mbar_wait(&ctl.c_full[s], (g >> 1) & 1);
                // ---- pass 1: x = s * log2(e) + lut[code] (log2 units), key mask, row maximum.
                // Chunks of 32 keys that are masked for every row are skipped in both passes.
                const uint32_t crow = smem_u32(sStage + s * kF3Stage + 32768 + t * 16384);
                float mx0 = neg_inf(), mx1 = neg_inf();
                if (warp_live) {
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        if (kmask[c] != 0xffffffffu) {
#pragma unroll
                            for (int u = 2 * c; u < 2 * c + 2; ++u) {
                                if (HAS_CODE) {
                                    const uint4 w = lds_u4(crow + sw128_offset(r, hf * 4 + u));
                                    const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                                    for (int e = 0; e < 4; ++e) {
                                        // byte -> 4 * (code & 63), merged into the 256-byte aligned table address by PRMT
                                        const uint32_t wm = (ww[e] << 2) & 0xfcfcfcfcu;
                                        const float l0 = lds_f32(__byte_perm(wm, lut_base, 0x7650));
                                        const float l1 = lds_f32(__byte_perm(wm, lut_base, 0x7651));
                                        const float l2 = lds_f32(__byte_perm(wm, lut_base, 0x7652));
                                        const float l3 = lds_f32(__byte_perm(wm, lut_base, 0x7653));
                                        float *x = sc + u * 16 + e * 4;
                                        ffma2(x[0], x[1], x[0], x[1], kLog2e, kLog2e, l0, l1);
                                        ffma2(x[2], x[3], x[2], x[3], kLog2e, kLog2e, l2, l3);
                                    }
                                } else {
#pragma unroll
                                    for (int e = 0; e < 8; ++e) {
                                        float *x = sc + u * 16 + e * 2;
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
                // ---- the two halves of a row agree on the maximum: each publishes its own in a code chunk of the
                // row that only it reads (dead after pass 1), then reads the other's
                float mx = fmaxf(mx0, mx1);
                if (tid == 0 && j == 0) f3_trace(a, 2 + 12 * n + 2);
                {
                    const uint32_t mine = smem_u32(&ctl.xmax[g & 1][t][hf][r]), other = smem_u32(&ctl.xmax[g & 1][t][hf ^ 1][r]);
                    asm volatile("st.shared.f32 [%0], %1;" ::"r"(mine), "f"(mx) : "memory");
                    asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory");
                    float omx;
                    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(omx) : "r"(other) : "memory");
                    mx = fmaxf(mx, omx);
                }
                if (HAS_CODE) mbar_arrive(&ctl.c_empty[s]);  // the code tile has been read (pass 1)
                float alpha = 1.f;
                bool grow = false;
                if (warp_live) {
                    grow = mx > m_used + kF3Rescale;       // also true for the first finite maximum
                    if (grow) {
                        alpha = ex2(m_used - mx);
                        m_used = mx;
                        l *= alpha;
                    }
                }
                // P_t (shared) and O_t (TMEM) are free once P V of the previous k-step has completed (across an
                // item boundary the epilogue below has already waited for it)
                if (j > 0) {
                    mbar_wait(&ctl.o_done[t], (g - 1) & 1);
                    tc_fence_after();
                    __syncwarp();
                    if (__any_sync(0xffffffffu, grow)) {   // this half rescales its 32 columns of O_t
                        uint32_t o[32];
                        tmem_ld_32x32(tO, o);
                        tmem_wait_ld();
#pragma unroll
                        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                        tmem_st_32x32(tO, o);
                        tmem_wait_st();
                    }
                }
                // ---- pass 2: p = 2^(x - m), row sum, bf16 pairs -> tensor memory (A operand of P V: lane = row, column = key pair)
                if (warp_live) {
                    const float nm = (m_used == neg_inf()) ? 0.f : -m_used;
                    float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        uint32_t pw[16];  // 32 keys of this row as bf16 pairs
                        if (kmask[c] != 0xffffffffu) {
#pragma unroll
                            for (int ch = 4 * c; ch < 4 * c + 4; ++ch) {
                                float *x = sc + ch * 8;
#pragma unroll
                                for (int i = 0; i < 8; i += 2) {
                                    fadd2(x[i], x[i + 1], x[i], x[i + 1], nm, nm);
                                    x[i] = ex2(x[i]);
                                    x[i + 1] = ex2(x[i + 1]);
                                    fadd2(sum0, sum1, sum0, sum1, x[i], x[i + 1]);   // the row sum is of the undropped p
                                    if (HAS_DROP) {  // dropout on P (multihead_attention.py:196)
                                        x[i] = drop_keep_w(drop_w, drop_lane(64u + ch * 8 + i), thr32) ? x[i] * a.drop.scale : 0.f;
                                        x[i + 1] = drop_keep_w(drop_w, drop_lane(65u + ch * 8 + i), thr32) ? x[i + 1] * a.drop.scale : 0.f;
                                    }
                                    pw[(ch - 4 * c) * 4 + (i >> 1)] = pack_bf16x2(x[i], x[i + 1]);
                                }
                            }
                        } else {
#pragma unroll
                            for (int w = 0; w < 16; ++w) pw[w] = 0u;
                        }
                        tmem_st_32x16(tP + c * 16, pw);
                    }
                    l += sum0 + sum1;
                    tmem_wait_st();
                }
                tc_fence_before();
                mbar_arrive(&ctl.p_full[t]);
                if (tid == 0 && j == 0) f3_trace(a, 2 + 12 * n + 3);
            }
            mbar_arrive(&ctl.aux_empty[buf]);          // key mask and LUT of this item are no longer read
            // ---- epilogue: O / l -> bf16, log-sum-exp; the halves add their row sums
            mbar_wait(&ctl.o_done[t], (g - 1) & 1);
            if (tid == 0) f3_trace(a, 2 + 12 * n + 4);
            tc_fence_after();
            __syncwarp();
            {
                const uint32_t mine = smem_u32(&ctl.xsum[n & 1][t][hf][r]), other = smem_u32(&ctl.xsum[n & 1][t][hf ^ 1][r]);
                asm volatile("st.shared.f32 [%0], %1;" ::"r"(mine), "f"(l) : "memory");
                asm volatile("bar.sync %0, 256;" ::"r"(bar_id) : "memory");
                float ol;
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(ol) : "r"(other) : "memory");
                l += ol;
            }
            if (warp_live) {
                float o[32];
                tmem_ld_32x32(tO, reinterpret_cast<uint32_t *>(o));
                tmem_wait_ld();
                if (row_ok) {
                    const float inv = l > 0.f ? 1.f / l : 0.f;  // a fully masked row yields zeros (the reference yields NaN)
                    __nv_bfloat16 *op = a.out + (size_t)it.b * a.o_sb + (size_t)q * a.o_st + (size_t)it.h * a.o_sh + hf * 32;
#pragma unroll
                    for (int ch = 0; ch < 4; ++ch) {
                        uint4 v;
                        v.x = pack_bf16x2(o[ch * 8 + 0] * inv, o[ch * 8 + 1] * inv);
                        v.y = pack_bf16x2(o[ch * 8 + 2] * inv, o[ch * 8 + 3] * inv);
                        v.z = pack_bf16x2(o[ch * 8 + 4] * inv, o[ch * 8 + 5] * inv);
                        v.w = pack_bf16x2(o[ch * 8 + 6] * inv, o[ch * 8 + 7] * inv);
                        reinterpret_cast<uint4 *>(op)[ch] = v;
                    }
                    if (a.lse && hf == 0)
                        a.lse[((size_t)it.b * H + it.h) * T + q] = l > 0.f ? m_used * kLn2 + logf(l) : __int_as_float(0x7f800000);
                }
            }
            if (tid == 0) f3_trace(a, 2 + 12 * n + 5);
            tc_fence_before();  // the O_t reads above are ordered before the p_full arrive that lets P V overwrite O_t
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kF3SoftmaxWarps) tmem_dealloc(tmem, 512);
}

int launch_fwd3(const CUtensorMap &mq, const CUtensorMap &mk, const CUtensorMap &mv, const CUtensorMap &mc, const Fwd2Args &a,
                cudaStream_t st) {
    const bool drop = a.drop.thr16 != 0;
#define MDT_LAUNCH_F3(C, D)                                                                                               \
    do {                                                                                                                  \
        MDT_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd3_kernel<C, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, kF3Smem)); \
        attn_fwd3_kernel<C, D><<<grid, kF3Threads, kF3Smem, st>>>(mq, mk, mv, mc, a);                                      \
    } while (0)
    const int nqp = (a.Tmain + 2 * kTile - 1) / (2 * kTile);
    const long long n_items = (long long)a.B * a.H * nqp;
    const int grid = (int)(n_items < (long long)num_sms() ? n_items : (long long)num_sms());
    if (a.has_code && drop) MDT_LAUNCH_F3(true, true);
    else if (a.has_code) MDT_LAUNCH_F3(true, false);
    else if (drop) MDT_LAUNCH_F3(false, true);
    else MDT_LAUNCH_F3(false, false);
#undef MDT_LAUNCH_F3
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

}  // namespace attn
}  // namespace mdt

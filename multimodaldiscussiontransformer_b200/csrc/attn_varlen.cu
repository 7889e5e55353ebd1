// N1: self-attention of the fusion layers' BERT blocks over PACKED (unpadded) sequences
//     mDT/src/modules/multi_graphormer_fusion_layer.py:105-149 (HF BertSelfAttention: softmax(q k^T / sqrt(d) + key mask) v,
//     dropout on the probabilities), driven from multigraphormer_graph_encoder.py:321-362,413-439.
//
// Layout: tokens of all comments laid end to end, q / k / v rows of `row_stride` elements (slices of one [total, 3, H, 64]
// projection), sequence s = rows [cu[s], cu[s+1]), every sequence <= 128 rows (4 bottleneck slots + <= 100 text tokens).
// A work unit is (item, head).  An ITEM is a run of consecutive sequences that fits one 128-row tile (built once per batch
// by mdt_varlen_attn_plan and shared by all layers, forward and backward): the tile is loaded with ONE TMA box per operand
// (2-D map over [total, row_stride], 128-byte swizzle), S = Q K^T is one 128x128x64 tcgen05.mma, and attention stays
// inside each sequence through a per-row column window [lo, hi) -- the block-diagonal mask.  Short comments (mean 62 rows)
// would otherwise leave half of every tile, and half of every TMEM lane quarter's threads, idle.
//
// forward   S = Q K^T -> TMEM; thread r = query row r (TMEM lane r): two passes over the 32-column chunks that hold live
//           keys (row maximum; p = 2^(x - m), row sum, dropout, bf16 P tile in shared memory); O = P V (k-steps only over
//           the live key chunks); O / l -> out, lse.  2 CTAs / SM; the next unit's Q, K (after S) and V (after P V) are
//           prefetched under the current unit's softmax / epilogue.
// backward  S = Q K^T, dP = dO V^T -> TMEM; delta = rowsum(dO * O) from global under the TMA latency; P = 2^(S - lse),
//           dS = P * (keep * dP / (1 - p) - delta) -> bf16 tiles [query][key]; dV = P^T dO, dK = dS^T Q (A read MN-major),
//           dQ = dS K.  No atomics, no dq scratch: a tile holds whole sequences, so each gradient row is written once.
#include <initializer_list>

#include "attn_common.cuh"

namespace mdt {
namespace attn {
using namespace sm100;

constexpr int kVlThreads = 128;

struct VarlenArgs {
    int total, H;
    const int32_t *cu;    // [n_seq + 1]
    const int32_t *plan;  // [0] = n_items, [1] = error flag, [2 + i] = first sequence of item i, [2 + n_items] = n_seq
    __nv_bfloat16 *out;   // forward: O rows; backward: unused
    int64_t o_stride;
    float *lse;           // [H, total]
    float scale, scale_log2;  // softmax scale (> 0) and scale * log2(e): scores are scaled inside the exponent
    DropArgs drop;
    // backward only
    const __nv_bfloat16 *o_in, *dout;
    int64_t g_stride;
    __nv_bfloat16 *dq, *dk, *dv;
    int64_t dq_stride, dk_stride, dv_stride;
};

// bit i set <=> column 32 c + i lies inside [lo, hi)
__device__ __forceinline__ uint32_t window_bits(int lo, int hi, int c) {
    const int a = lo - 32 * c, b = hi - 32 * c;
    if (b <= 0 || a >= 32) return 0u;
    const uint32_t below_hi = b >= 32 ? 0xffffffffu : ((1u << b) - 1u);
    const uint32_t below_lo = a <= 0 ? 0u : ((1u << a) - 1u);
    return below_hi & ~below_lo;
}

struct VlUnit {
    int row0, nrows, lo, hi, h;
    bool valid;
};
// item / head of work unit w and this thread's row window
__device__ __forceinline__ VlUnit vl_unit(const VarlenArgs &a, int w, int r) {
    VlUnit u;
    const int item = w / a.H;
    u.h = w - item * a.H;
    const int s0 = a.plan[2 + item], s1 = a.plan[3 + item];
    u.row0 = __ldg(a.cu + s0);
    u.nrows = min(__ldg(a.cu + s1) - u.row0, 128);
    u.valid = r < u.nrows;
    u.lo = u.hi = 0;
    if (u.valid) {
        const int g = u.row0 + r;
        int start = u.row0;
        for (int s = s0; s < s1; ++s) {
            const int end = __ldg(a.cu + s + 1);
            if (g < end) {
                u.lo = start - u.row0;
                u.hi = end - u.row0;
                break;
            }
            start = end;
        }
    }
    return u;
}

// ------------------------------------------------------------------------------------------------ plan
// Greedy grouping of consecutive sequences into <= 128-row items.  One thread: the loop-carried state is a register, the
// cu loads do not depend on it, and the plan is built once per batch (shared by every layer, forward and backward).
__global__ void varlen_plan_kernel(const int32_t *__restrict__ cu, int n_seq, int group, int32_t *__restrict__ plan) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int n_items = 0, err = 0;
    int start_row = cu[0], s_first = 0;
    for (int s = 0; s < n_seq; ++s) {
        const int end = cu[s + 1];
        if (end - cu[s] > 128) err = 1;  // a sequence longer than one tile: not this kernel's case
        if (s > s_first && (!group || end - start_row > 128)) {  // sequence s does not fit any more: close the item before it
            plan[2 + n_items++] = s_first;
            s_first = s;
            start_row = cu[s];
        }
    }
    if (n_seq > 0) plan[2 + n_items++] = s_first;
    plan[2 + n_items] = n_seq;
    plan[0] = n_items;
    plan[1] = err;
}

// ------------------------------------------------------------------------------------------------ forward
struct VlFwdCtrl {
    uint64_t bar_qk, bar_v, bar_s, bar_o;
    uint32_t tmem_base;
};
constexpr int kVlFwdTiles = 3 * 16384 + 32768;  // Q K V | P
constexpr int kVlFwdSmem = kVlFwdTiles + 64;

__global__ void __launch_bounds__(kVlThreads, 2)
    attn_varlen_fwd_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                           const __grid_constant__ CUtensorMap mv, const VarlenArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *sQ = smem, *sK = smem + 16384, *sV = smem + 32768, *sP = smem + 49152;
    VlFwdCtrl &ctl = *reinterpret_cast<VlFwdCtrl *>(smem + kVlFwdTiles);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int n_items = a.plan[0];
    const int n_units = a.plan[1] ? 0 : n_items * a.H;  // a flagged plan (sequence > 128 rows) runs nothing

    if (tid == 0) {
        if (smem_u32(smem) & 1023u) __trap();
        mbar_init(&ctl.bar_qk, 1);
        mbar_init(&ctl.bar_v, 1);
        mbar_init(&ctl.bar_s, 1);
        mbar_init(&ctl.bar_o, 1);
        fence_barrier_init();
        tma_prefetch_desc(&mq);
        tma_prefetch_desc(&mk);
        tma_prefetch_desc(&mv);
    }
    if (warp == 0) {
        tmem_alloc(&ctl.tmem_base, 256);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctl.tmem_base;
    const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
    const uint32_t idesc_s = idesc_bf16(128, 128, 0, 0);
    const uint32_t idesc_o = idesc_bf16(128, 64, 0, 1);

    auto load_qk = [&](int w) {  // tid 0
        const int item = w / a.H, h = w - item * a.H;
        const int row0 = __ldg(a.cu + a.plan[2 + item]);
        mbar_expect_tx(&ctl.bar_qk, 32768);
        tma_load_2d(sQ, &mq, &ctl.bar_qk, h * kD, row0);
        tma_load_2d(sK, &mk, &ctl.bar_qk, h * kD, row0);
    };
    auto load_v = [&](int w) {
        const int item = w / a.H, h = w - item * a.H;
        const int row0 = __ldg(a.cu + a.plan[2 + item]);
        mbar_expect_tx(&ctl.bar_v, 16384);
        tma_load_2d(sV, &mv, &ctl.bar_v, h * kD, row0);
    };

    if (tid == 0 && (int)blockIdx.x < n_units) {
        load_qk(blockIdx.x);
        load_v(blockIdx.x);
    }

    int it = 0;
    for (int w = blockIdx.x; w < n_units; w += gridDim.x, ++it) {
        const uint32_t par = it & 1;
        const VlUnit u = vl_unit(a, w, tid);
        const int nch = (u.nrows + 31) >> 5;
        const bool warp_live = __any_sync(0xffffffffu, u.valid);
        const int w_next = w + gridDim.x;

        if (tid == 0) {
            mbar_wait(&ctl.bar_qk, par);
            tc_fence_after();
            const uint32_t aq = smem_u32(sQ), bk = smem_u32(sK);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                umma_bf16(tmem, smem_desc_sw128(aq + k * 32, 16, 1024), smem_desc_sw128(bk + k * 32, 16, 1024), idesc_s, k > 0);
            umma_commit(&ctl.bar_s);
        }
        __syncwarp();
        mbar_wait(&ctl.bar_s, par);
        tc_fence_after();
        if (tid == 0 && w_next < n_units) load_qk(w_next);  // Q, K are free: the next unit's tiles fly under this softmax
        __syncwarp();

        float m = neg_inf(), l = 0.f;
        if (warp_live) {
            // ---- pass 1: row maximum over the row's window (log2 units)
            for (int c = 0; c < nch; ++c) {
                float x[32];
                tmem_ld_32x32(lane_base + c * 32, reinterpret_cast<uint32_t *>(x));
                tmem_wait_ld();
                const uint32_t bits = window_bits(u.lo, u.hi, c);
                if (bits == 0u) continue;
                float mx0 = neg_inf(), mx1 = neg_inf();
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    mx0 = fmaxf(mx0, (bits >> i) & 1u ? x[i] : neg_inf());
                    mx1 = fmaxf(mx1, (bits >> (i + 1)) & 1u ? x[i + 1] : neg_inf());
                }
                m = fmaxf(m, fmaxf(mx0, mx1));
            }
            m *= a.scale_log2;
            const float nm = (m == neg_inf()) ? 0.f : -m;
            const float sl2 = a.scale_log2;
            const uint32_t salt = drop_salt(a.drop, (uint32_t)w);
            // ---- pass 2: p = 2^(x log2e - m), row sum, dropout, bf16 P tile (K-major A operand of P V)
            for (int c = 0; c < nch; ++c) {
                float x[32];
                tmem_ld_32x32(lane_base + c * 32, reinterpret_cast<uint32_t *>(x));
                tmem_wait_ld();
                const uint32_t bits = window_bits(u.lo, u.hi, c);
                float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    float p0 = (bits >> i) & 1u ? ex2(fmaf(x[i], sl2, nm)) : 0.f;
                    float p1 = (bits >> (i + 1)) & 1u ? ex2(fmaf(x[i + 1], sl2, nm)) : 0.f;
                    sum0 += p0;  // the row sum is of the undropped p
                    sum1 += p1;
                    if (a.drop.thr16) {
                        const uint32_t hp = drop_hash(salt, (uint32_t)(tid * 128 + c * 32 + i));
                        p0 = (hp >> 16) >= a.drop.thr16 ? p0 * a.drop.scale : 0.f;
                        p1 = (hp & 0xffffu) >= a.drop.thr16 ? p1 * a.drop.scale : 0.f;
                    }
                    x[i] = p0;
                    x[i + 1] = p1;
                }
                l += sum0 + sum1;
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) {
                    uint4 v;
                    v.x = pack_bf16x2(x[ch * 8 + 0], x[ch * 8 + 1]);
                    v.y = pack_bf16x2(x[ch * 8 + 2], x[ch * 8 + 3]);
                    v.z = pack_bf16x2(x[ch * 8 + 4], x[ch * 8 + 5]);
                    v.w = pack_bf16x2(x[ch * 8 + 6], x[ch * 8 + 7]);
                    *reinterpret_cast<uint4 *>(sP + (c >> 1) * 16384 + sw128_offset(tid, (c & 1) * 4 + ch)) = v;
                }
            }
        }
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            mbar_wait(&ctl.bar_v, par);
            tc_fence_after();
            const uint32_t ap = smem_u32(sP), bv = smem_u32(sV);
            for (int k = 0; k < 2 * nch; ++k)  // 16 keys per k-step, live key chunks only
                umma_bf16(tmem + 128, smem_desc_sw128(ap + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                          smem_desc_sw128(bv + k * 2048, 8192, 1024), idesc_o, k > 0 ? 1u : 0u);
            umma_commit(&ctl.bar_o);
        }
        __syncwarp();
        mbar_wait(&ctl.bar_o, par);
        tc_fence_after();
        if (tid == 0 && w_next < n_units) load_v(w_next);  // V and P are free
        __syncwarp();
        if (warp_live) {
            float o[64];
            tmem_ld_32x32(lane_base + 128, reinterpret_cast<uint32_t *>(o));
            tmem_ld_32x32(lane_base + 160, reinterpret_cast<uint32_t *>(o) + 32);
            tmem_wait_ld();
            if (u.valid) {
                const float inv = l > 0.f ? 1.f / l : 0.f;
                const size_t g = (size_t)(u.row0 + tid);
                __nv_bfloat16 *op = a.out + g * a.o_stride + (size_t)u.h * kD;
#pragma unroll
                for (int ch = 0; ch < 8; ++ch) {
                    uint4 v;
                    v.x = pack_bf16x2(o[ch * 8 + 0] * inv, o[ch * 8 + 1] * inv);
                    v.y = pack_bf16x2(o[ch * 8 + 2] * inv, o[ch * 8 + 3] * inv);
                    v.z = pack_bf16x2(o[ch * 8 + 4] * inv, o[ch * 8 + 5] * inv);
                    v.w = pack_bf16x2(o[ch * 8 + 6] * inv, o[ch * 8 + 7] * inv);
                    reinterpret_cast<uint4 *>(op)[ch] = v;
                }
                if (a.lse) a.lse[(size_t)u.h * a.total + g] = l > 0.f ? m * kLn2 + logf(l) : __int_as_float(0x7f800000);
            }
        }
        tc_fence_before();  // S / O columns are rewritten by the next unit's MMAs (issued after the next __syncthreads / by tid 0 in order)
        __syncthreads();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}

// ------------------------------------------------------------------------------------------------ backward
struct VlBwdCtrl {
    uint64_t bar_ld, bar_s, bar_o;
    uint32_t tmem_base;
};
// Q | K | dO | P (32 KB: its second half is the V tile, dead once dP = dO V^T has completed) | dS
constexpr int kVlBwdTiles = 3 * 16384 + 32768 + 32768;
constexpr int kVlBwdSmem = kVlBwdTiles + 64;

__global__ void __launch_bounds__(kVlThreads, 2)
    attn_varlen_bwd_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                           const __grid_constant__ CUtensorMap mv, const __grid_constant__ CUtensorMap mdo, const VarlenArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *sQ = smem, *sK = smem + 16384, *sdO = smem + 32768, *sP = smem + 49152, *sV = smem + 65536, *sdS = smem + 81920;
    VlBwdCtrl &ctl = *reinterpret_cast<VlBwdCtrl *>(smem + kVlBwdTiles);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int n_items = a.plan[0];
    const int n_units = a.plan[1] ? 0 : n_items * a.H;

    if (tid == 0) {
        if (smem_u32(smem) & 1023u) __trap();
        mbar_init(&ctl.bar_ld, 1);
        mbar_init(&ctl.bar_s, 1);
        mbar_init(&ctl.bar_o, 1);
        fence_barrier_init();
        tma_prefetch_desc(&mq);
        tma_prefetch_desc(&mk);
        tma_prefetch_desc(&mv);
        tma_prefetch_desc(&mdo);
    }
    if (warp == 0) {
        tmem_alloc(&ctl.tmem_base, 256);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = ctl.tmem_base;
    const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
    constexpr uint32_t cS = 0, cDP = 128, cDV = 0, cDK = 64, cDQ = 128;  // dV / dK / dQ reuse the S / dP columns
    const uint32_t idesc_s = idesc_bf16(128, 128, 0, 0);
    const uint32_t idesc_kv = idesc_bf16(128, 64, 1, 1);  // A = P / dS read MN-major (M = key), B = dO / Q MN-major
    const uint32_t idesc_dq = idesc_bf16(128, 64, 0, 1);  // A = dS K-major (M = query), B = K MN-major

    int it = 0;
    for (int w = blockIdx.x; w < n_units; w += gridDim.x, ++it) {
        const uint32_t par = it & 1;
        const VlUnit u = vl_unit(a, w, tid);
        const int nch = (u.nrows + 31) >> 5;
        const bool row_used = tid < 32 * nch;  // rows the dV / dK k-steps read: live or zero
        const bool warp_used = (warp * 32) < 32 * nch;
        if (tid == 0) {
            mbar_expect_tx(&ctl.bar_ld, 65536);
            tma_load_2d(sQ, &mq, &ctl.bar_ld, u.h * kD, u.row0);
            tma_load_2d(sK, &mk, &ctl.bar_ld, u.h * kD, u.row0);
            tma_load_2d(sV, &mv, &ctl.bar_ld, u.h * kD, u.row0);
            tma_load_2d(sdO, &mdo, &ctl.bar_ld, u.h * kD, u.row0);
        }
        // ---- delta = rowsum(dO * O), -lse in log2 units: global loads under the TMA latency
        float delta = 0.f, nlse = neg_inf();
        if (u.valid) {
            const size_t g = (size_t)(u.row0 + tid);
            const uint4 *po = reinterpret_cast<const uint4 *>(a.o_in + g * a.o_stride + (size_t)u.h * kD);
            const uint4 *pg = reinterpret_cast<const uint4 *>(a.dout + g * a.g_stride + (size_t)u.h * kD);
#pragma unroll
            for (int ch = 0; ch < 8; ++ch) {
                const uint4 x = __ldg(po + ch), y = __ldg(pg + ch);
                const __nv_bfloat162 *px = reinterpret_cast<const __nv_bfloat162 *>(&x);
                const __nv_bfloat162 *py = reinterpret_cast<const __nv_bfloat162 *>(&y);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float2 fx = __bfloat1622float2(px[i]), fy = __bfloat1622float2(py[i]);
                    delta = fmaf(fx.x, fy.x, delta);
                    delta = fmaf(fx.y, fy.y, delta);
                }
            }
            const float lse = a.lse[(size_t)u.h * a.total + g];
            nlse = (lse == __int_as_float(0x7f800000)) ? neg_inf() : -lse * kLog2e;  // +inf marks a row without keys
        }
        if (tid == 0) {
            mbar_wait(&ctl.bar_ld, par);
            tc_fence_after();
            const uint32_t aq = smem_u32(sQ), bk = smem_u32(sK), ao = smem_u32(sdO), bv = smem_u32(sV);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                umma_bf16(tmem + cS, smem_desc_sw128(aq + k * 32, 16, 1024), smem_desc_sw128(bk + k * 32, 16, 1024), idesc_s, k > 0);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                umma_bf16(tmem + cDP, smem_desc_sw128(ao + k * 32, 16, 1024), smem_desc_sw128(bv + k * 32, 16, 1024), idesc_s, k > 0);
            umma_commit(&ctl.bar_s);
        }
        __syncwarp();
        mbar_wait(&ctl.bar_s, par);
        tc_fence_after();
        __syncwarp();
        if (warp_used) {
            const uint32_t salt = drop_salt(a.drop, (uint32_t)w);
            const float sl2 = a.scale_log2, sc = a.scale;
            for (int c = 0; c < nch; ++c) {
                float st[32], dp[32];
                tmem_ld_32x32(lane_base + cS + c * 32, reinterpret_cast<uint32_t *>(st));
                tmem_ld_32x32(lane_base + cDP + c * 32, reinterpret_cast<uint32_t *>(dp));
                tmem_wait_ld();
                const uint32_t bits = window_bits(u.lo, u.hi, c);  // 0 for rows past the item: P = dS = 0
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    float p0 = (bits >> i) & 1u ? ex2(fmaf(st[i], sl2, nlse)) : 0.f;
                    float p1 = (bits >> (i + 1)) & 1u ? ex2(fmaf(st[i + 1], sl2, nlse)) : 0.f;
                    float k0 = 1.f, k1 = 1.f;
                    if (a.drop.thr16) {
                        const uint32_t hp = drop_hash(salt, (uint32_t)(tid * 128 + c * 32 + i));
                        k0 = (hp >> 16) >= a.drop.thr16 ? a.drop.scale : 0.f;
                        k1 = (hp & 0xffffu) >= a.drop.thr16 ? a.drop.scale : 0.f;
                    }
                    dp[i] = p0 * (dp[i] * k0 - delta) * sc;       // d(scaled score) * scale = d(q k^T)
                    dp[i + 1] = p1 * (dp[i + 1] * k1 - delta) * sc;
                    st[i] = p0 * k0;
                    st[i + 1] = p1 * k1;
                }
                if (row_used) {
#pragma unroll
                    for (int ch = 0; ch < 4; ++ch) {
                        uint4 vp, vd;
                        vp.x = pack_bf16x2(st[ch * 8 + 0], st[ch * 8 + 1]);
                        vp.y = pack_bf16x2(st[ch * 8 + 2], st[ch * 8 + 3]);
                        vp.z = pack_bf16x2(st[ch * 8 + 4], st[ch * 8 + 5]);
                        vp.w = pack_bf16x2(st[ch * 8 + 6], st[ch * 8 + 7]);
                        vd.x = pack_bf16x2(dp[ch * 8 + 0], dp[ch * 8 + 1]);
                        vd.y = pack_bf16x2(dp[ch * 8 + 2], dp[ch * 8 + 3]);
                        vd.z = pack_bf16x2(dp[ch * 8 + 4], dp[ch * 8 + 5]);
                        vd.w = pack_bf16x2(dp[ch * 8 + 6], dp[ch * 8 + 7]);
                        const uint32_t off = (c >> 1) * 16384 + sw128_offset(tid, (c & 1) * 4 + ch);
                        *reinterpret_cast<uint4 *>(sP + off) = vp;   // half 1 of P overwrites the V tile (dP has completed)
                        *reinterpret_cast<uint4 *>(sdS + off) = vd;
                    }
                }
            }
        }
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            const uint32_t ap = smem_u32(sP), ad = smem_u32(sdS);
            const uint32_t bo = smem_u32(sdO), bq = smem_u32(sQ), bk = smem_u32(sK);
            const int nks = 2 * nch;  // 16 rows (queries for dV / dK, keys for dQ) per k-step: live chunks only
            for (int k = 0; k < nks; ++k)  // dV = P^T dO
                umma_bf16(tmem + cDV, smem_desc_sw128(ap + k * 2048, 16384, 1024), smem_desc_sw128(bo + k * 2048, 8192, 1024),
                          idesc_kv, k > 0 ? 1u : 0u);
            for (int k = 0; k < nks; ++k)  // dK = dS^T Q
                umma_bf16(tmem + cDK, smem_desc_sw128(ad + k * 2048, 16384, 1024), smem_desc_sw128(bq + k * 2048, 8192, 1024),
                          idesc_kv, k > 0 ? 1u : 0u);
            for (int k = 0; k < nks; ++k)  // dQ = dS K
                umma_bf16(tmem + cDQ, smem_desc_sw128(ad + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                          smem_desc_sw128(bk + k * 2048, 8192, 1024), idesc_dq, k > 0 ? 1u : 0u);
            umma_commit(&ctl.bar_o);
        }
        __syncwarp();
        mbar_wait(&ctl.bar_o, par);
        tc_fence_after();
        __syncwarp();
        if (__any_sync(0xffffffffu, u.valid)) {
            const size_t g = (size_t)(u.row0 + tid);
#pragma unroll
            for (int which = 0; which < 3; ++which) {
                float acc[64];
                const uint32_t col = which == 0 ? cDV : (which == 1 ? cDK : cDQ);
                tmem_ld_32x32(lane_base + col, reinterpret_cast<uint32_t *>(acc));
                tmem_ld_32x32(lane_base + col + 32, reinterpret_cast<uint32_t *>(acc) + 32);
                tmem_wait_ld();
                if (u.valid) {
                    __nv_bfloat16 *p = which == 0 ? a.dv + g * a.dv_stride : (which == 1 ? a.dk + g * a.dk_stride : a.dq + g * a.dq_stride);
                    p += (size_t)u.h * kD;
#pragma unroll
                    for (int ch = 0; ch < 8; ++ch) {
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
        __syncthreads();  // every operand tile and TMEM column of this unit is free for the next one
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}

// TMA view of packed rows: dims {H * 64, total}, box {64, 128}; rows past `total` read as zeros
static int make_rows_map(CUtensorMap *map, const void *base, int total, int H, int64_t row_stride) {
    MDT_REQUIRE(row_stride % 8 == 0 && row_stride >= (int64_t)H * kD, "row stride must be a multiple of 8 elements and >= H * 64");
    const uint64_t dims[2] = {(uint64_t)H * kD, (uint64_t)total};
    const uint64_t strides[2] = {2, (uint64_t)row_stride * 2};
    const uint32_t box[2] = {(uint32_t)kD, (uint32_t)kTile};
    return make_tensor_map_bf16(map, base, 2, dims, strides, box);
}

static int units_grid(int n_seq, int H) {
    const long long cap = 2LL * num_sms(), want = (long long)n_seq * H;
    return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace attn
}  // namespace mdt

using namespace mdt;
using namespace mdt::attn;

extern "C" int mdt_varlen_attn_plan(const int32_t *cu_seqlens, int n_seq, int group, int32_t *plan, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(cu_seqlens && plan && n_seq >= 1, "mdt_varlen_attn_plan: cu_seqlens, plan must be non-null and n_seq >= 1");
    varlen_plan_kernel<<<1, 32, 0, as_stream(stream)>>>(cu_seqlens, n_seq, group, plan);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

extern "C" int mdt_varlen_attn_fwd(const void *q, const void *k, const void *v, int64_t q_stride, int64_t k_stride,
                                   int64_t v_stride, int total, int H, const int32_t *cu_seqlens, int n_seq,
                                   const int32_t *plan, void *out, int64_t o_stride, float *lse, float softmax_scale,
                                   float dropout_p, uint64_t dropout_seed, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(q && k && v && out && lse && cu_seqlens && plan, "mdt_varlen_attn_fwd: null pointer");
    MDT_REQUIRE(total >= 1 && H >= 1 && n_seq >= 1, "mdt_varlen_attn_fwd: total, H, n_seq must be >= 1");
    MDT_REQUIRE(((uintptr_t)out & 15) == 0 && o_stride % 8 == 0, "mdt_varlen_attn_fwd: out must be 16-byte aligned, stride a multiple of 8");
    MDT_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "mdt_varlen_attn_fwd: dropout_p must be in [0, 1)");
    MDT_REQUIRE(softmax_scale > 0.f, "mdt_varlen_attn_fwd: softmax_scale must be positive");
    CUtensorMap mq, mk, mv;
    if (int rc = make_rows_map(&mq, q, total, H, q_stride)) return rc;
    if (int rc = make_rows_map(&mk, k, total, H, k_stride)) return rc;
    if (int rc = make_rows_map(&mv, v, total, H, v_stride)) return rc;
    VarlenArgs a{};
    a.total = total, a.H = H, a.cu = cu_seqlens, a.plan = plan;
    a.out = reinterpret_cast<__nv_bfloat16 *>(out), a.o_stride = o_stride, a.lse = lse;
    a.scale = softmax_scale, a.scale_log2 = softmax_scale * kLog2e;
    a.drop = make_drop(dropout_p, dropout_seed);
    MDT_CHECK_CUDA(cudaFuncSetAttribute(attn_varlen_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kVlFwdSmem));
    attn_varlen_fwd_kernel<<<units_grid(n_seq, H), kVlThreads, kVlFwdSmem, as_stream(stream)>>>(mq, mk, mv, a);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

extern "C" int mdt_varlen_attn_bwd(const void *q, const void *k, const void *v, int64_t q_stride, int64_t k_stride,
                                   int64_t v_stride, const void *out, int64_t o_stride, const void *grad_out,
                                   int64_t g_stride, const float *lse, int total, int H, const int32_t *cu_seqlens,
                                   int n_seq, const int32_t *plan, void *dq, void *dk, void *dv, int64_t dq_stride,
                                   int64_t dk_stride, int64_t dv_stride, float softmax_scale, float dropout_p,
                                   uint64_t dropout_seed, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(q && k && v && out && grad_out && lse && cu_seqlens && plan && dq && dk && dv, "mdt_varlen_attn_bwd: null pointer");
    MDT_REQUIRE(total >= 1 && H >= 1 && n_seq >= 1, "mdt_varlen_attn_bwd: total, H, n_seq must be >= 1");
    for (int64_t s : {o_stride, g_stride, dq_stride, dk_stride, dv_stride})
        MDT_REQUIRE(s % 8 == 0, "mdt_varlen_attn_bwd: strides must be multiples of 8 elements");
    MDT_REQUIRE((((uintptr_t)out | (uintptr_t)grad_out | (uintptr_t)dq | (uintptr_t)dk | (uintptr_t)dv) & 15) == 0,
                "mdt_varlen_attn_bwd: buffers must be 16-byte aligned");
    MDT_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "mdt_varlen_attn_bwd: dropout_p must be in [0, 1)");
    MDT_REQUIRE(softmax_scale > 0.f, "mdt_varlen_attn_bwd: softmax_scale must be positive");
    CUtensorMap mq, mk, mv, mdo;
    if (int rc = make_rows_map(&mq, q, total, H, q_stride)) return rc;
    if (int rc = make_rows_map(&mk, k, total, H, k_stride)) return rc;
    if (int rc = make_rows_map(&mv, v, total, H, v_stride)) return rc;
    if (int rc = make_rows_map(&mdo, grad_out, total, H, g_stride)) return rc;
    VarlenArgs a{};
    a.total = total, a.H = H, a.cu = cu_seqlens, a.plan = plan;
    a.lse = const_cast<float *>(lse);
    a.scale = softmax_scale, a.scale_log2 = softmax_scale * kLog2e;
    a.drop = make_drop(dropout_p, dropout_seed);
    a.o_in = reinterpret_cast<const __nv_bfloat16 *>(out), a.o_stride = o_stride;
    a.dout = reinterpret_cast<const __nv_bfloat16 *>(grad_out), a.g_stride = g_stride;
    a.dq = reinterpret_cast<__nv_bfloat16 *>(dq), a.dk = reinterpret_cast<__nv_bfloat16 *>(dk), a.dv = reinterpret_cast<__nv_bfloat16 *>(dv);
    a.dq_stride = dq_stride, a.dk_stride = dk_stride, a.dv_stride = dv_stride;
    MDT_CHECK_CUDA(cudaFuncSetAttribute(attn_varlen_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kVlBwdSmem));
    attn_varlen_bwd_kernel<<<units_grid(n_seq, H), kVlThreads, kVlBwdSmem, as_stream(stream)>>>(mq, mk, mv, mdo, a);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

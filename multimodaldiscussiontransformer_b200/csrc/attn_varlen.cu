// N1: self-attention of the fusion layers' BERT blocks over PACKED (unpadded) sequences
//     mDT/src/modules/multi_graphormer_fusion_layer.py:105-149 (HF BertSelfAttention: softmax(q k^T / sqrt(d) + key mask) v,
//     dropout on the probabilities), driven from multigraphormer_graph_encoder.py:321-362,413-439.
//
// Layout: tokens of all comments laid end to end, q / k / v rows of `row_stride` elements (slices of one [total, 3, H, 64]
// projection), sequence s = rows [cu[s], cu[s+1]), every sequence <= 128 rows (4 bottleneck slots + <= 100 text tokens).
// A work unit is (item, head).  An ITEM is a run of consecutive sequences that fits one 128-row tile (built once per batch
// by mdt_varlen_attn_plan and shared by all layers, forward and backward): the tile is loaded with ONE TMA box per operand
// (2-D map over [total, row_stride], 128-byte swizzle), S = Q K^T is one 128x128x64 tcgen05.mma, and attention stays
// inside each sequence through a per-row column window [lo, hi) -- the block-diagonal mask, precomputed per packed row by
// the plan kernel.  Short comments (mean 62 rows) would otherwise leave half of every tile, and half of every TMEM lane
// quarter's threads, idle.
//
// 256 threads per CTA, 2 CTAs per SM: threads r and r + 128 share query row r (= TMEM lane r) and split its key columns
// ([0,64) / [64,128)) and its 64 output columns, so a unit's softmax / epilogue latency is half of the one-thread-per-row
// form and 16 warps per SM hide the TMEM / shared-memory / MUFU latencies.
//
// forward   S = Q K^T -> TMEM; each thread holds its <= 64 scores in registers: masking only in the 32-column chunks the
//           row's window cuts through, zero stores in chunks of other sequences; the halves exchange row maximum and row
//           sum through shared memory; bf16 P tile -> O = P V (k-steps over live key chunks only); O / l -> out, lse.
//           The next unit's Q, K (after S) and V (after P V) are prefetched under the current unit's softmax / epilogue.
// backward  S = Q K^T, dP = dO V^T -> TMEM; delta = rowsum(dO * O) = rowsum(keep P dP / (1 - p)) from the two accumulators
//           (no O / dO rows from global memory); P = 2^(S - lse),
//           dS = P * (keep * dP / (1 - p) - delta) -> bf16 tiles [query][key]; dV = P^T dO, dK = dS^T Q (A read MN-major),
//           dQ = dS K.  No atomics, no dq scratch: a tile holds whole sequences, so each gradient row is written once.
//           The next unit's four tiles are requested as soon as the last MMAs have completed, under the epilogue.
#include <initializer_list>

#include "attn_common.cuh"

namespace mdt {
namespace attn {
using namespace sm100;

constexpr int kVlThreads = 256;

// plan layout (int32): [0] n_items  [1] error flag  [2] n_seq  [3] total
//   [4 + 2 i] first row of item i, [5 + 2 i] its first sequence, i = 0 .. n_items (entry n_items = {total, n_seq})
//   then n_seq words of scratch, then `total` row windows: lo | hi << 16, relative to the item's first row
__host__ __device__ constexpr int plan_items_off() { return 4; }
__host__ __device__ inline int plan_scratch_off(int n_seq) { return 4 + 2 * (n_seq + 1); }
__host__ __device__ inline int plan_rows_off(int n_seq) { return plan_scratch_off(n_seq) + n_seq; }

struct VarlenArgs {
    int total, H, n_seq;
    const int32_t *plan;
    __nv_bfloat16 *out;   // forward: O rows
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
    int row0, nrows, h;
};
__device__ __forceinline__ VlUnit vl_unit(const VarlenArgs &a, int w) {
    VlUnit u;
    const int item = w / a.H;
    u.h = w - item * a.H;
    const int row0 = __ldg(a.plan + plan_items_off() + 2 * item);
    const int next_row0 = __ldg(a.plan + plan_items_off() + 2 * (item + 1));
    u.row0 = row0;
    u.nrows = min(next_row0 - row0, 128);
    return u;
}

// ------------------------------------------------------------------------------------------------ plan
// Greedy grouping of consecutive sequences into <= 128-row items by thread 0 (the loop-carried state is a register and the
// cu loads do not depend on it), then all threads fill the per-row windows.  Built once per batch.
constexpr int kPlanSmemSeqs = 11 * 1024;  // cu_seqlens of up to 11263 sequences staged in shared memory (44 KB)

__global__ void __launch_bounds__(1024) varlen_plan_kernel(const int32_t *__restrict__ cu, int n_seq, int group, int32_t *__restrict__ plan) {
    __shared__ int32_t s_cu[kPlanSmemSeqs];
    int32_t *items = plan + plan_items_off();
    int32_t *seq_row0 = plan + plan_scratch_off(n_seq);
    int32_t *rowwin = plan + plan_rows_off(n_seq);
    const bool staged = n_seq + 1 <= kPlanSmemSeqs;   // the serial pass then reads shared memory (29 cycles) instead of L2
    if (staged)
        for (int i = threadIdx.x; i <= n_seq; i += blockDim.x) s_cu[i] = cu[i];
    __syncthreads();
    const int32_t *c = staged ? s_cu : cu;
    if (threadIdx.x == 0) {
        int n_items = 0, err = 0;
        int prev = c[0], start_row = prev, s_first = 0;
        for (int s = 0; s < n_seq; ++s) {
            const int end = c[s + 1];
            if (end - prev > 128) err = 1;  // a sequence longer than one tile: not this kernel's case
            if (s > s_first && (!group || end - start_row > 128)) {  // sequence s does not fit any more: close the item before it
                items[2 * n_items] = start_row;
                items[2 * n_items + 1] = s_first;
                ++n_items;
                s_first = s;
                start_row = prev;
            }
            seq_row0[s] = start_row;
            prev = end;
        }
        items[2 * n_items] = start_row;
        items[2 * n_items + 1] = s_first;
        ++n_items;
        items[2 * n_items] = prev;
        items[2 * n_items + 1] = n_seq;
        plan[0] = n_items;
        plan[1] = err;
        plan[2] = n_seq;
        plan[3] = prev;
    }
    __syncthreads();
    for (int s = threadIdx.x; s < n_seq; s += blockDim.x) {
        const int a = c[s], b = c[s + 1], r0 = seq_row0[s];
        const int32_t win = (a - r0) | ((b - r0) << 16);
        for (int g = a; g < b; ++g) rowwin[g] = win;
    }
}

// The same plan without the serial pass (n_seq <= kPlanParSeqs; the kernel above stays for longer batches): the greedy grouping
// is a chain s -> nxt[s] (first sequence that no longer fits into the item that starts at s: a binary search per sequence, all
// in parallel); the chain from sequence 0 is marked by pointer doubling (log2 n rounds), item numbers are a block scan of the
// marks, and a second, multi-CTA launch fills the per-row windows, one warp per sequence.  350 us -> ~15 us per plan.
constexpr int kPlanParSeqs = 11 * 1024;

__global__ void __launch_bounds__(1024) varlen_plan_items_kernel(const int32_t *__restrict__ cu, int n_seq, int group, int32_t *__restrict__ plan) {
    __shared__ uint16_t s_jump[kPlanParSeqs + 1];   // pointer-doubling table; entry n_seq is the end of the chain
    __shared__ uint8_t s_nxt[kPlanParSeqs + 1];     // nxt[s] - s (<= 129)
    __shared__ uint8_t s_mark[kPlanParSeqs + 1];    // 1 = an item starts at this sequence
    __shared__ int s_warp[32], s_err, s_carry;
    int32_t *items = plan + plan_items_off();
    int32_t *seq_row0 = plan + plan_scratch_off(n_seq);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_err = 0, s_carry = 0;
    __syncthreads();
    for (int s = tid; s <= n_seq; s += 1024) {
        int t = n_seq;
        if (s < n_seq) {
            const int base = cu[s];
            if (cu[s + 1] - base > 128) s_err = 1;   // a sequence longer than one tile: not this kernel's case
            if (!group) {
                t = s + 1;
            } else {
                // largest t in (s, min(n_seq, s + 129)] with cu[t] - cu[s] <= 128 (cu is non-decreasing); at least s + 1
                int lo = s + 1, hi = min(n_seq, s + 129);
                while (lo < hi) {
                    const int mid = (lo + hi + 1) >> 1;
                    if (cu[mid] - base <= 128) lo = mid; else hi = mid - 1;
                }
                t = lo;
            }
        }
        s_jump[s] = (uint16_t)t;
        s_nxt[s] = (uint8_t)(t - s);
        s_mark[s] = s == 0;
    }
    __syncthreads();
    for (int span = 1; span < n_seq; span <<= 1) {   // after round r the marks cover the first 2^r chain nodes
        for (int s = tid; s < n_seq; s += 1024)
            if (s_mark[s]) s_mark[s_jump[s]] = 1;     // racing writers all write 1 (entry n_seq absorbs the chain's end)
        __syncthreads();
        uint16_t j2[(kPlanParSeqs + 1023) / 1024];
        int q = 0;
        for (int s = tid; s < n_seq; s += 1024, ++q) j2[q] = s_jump[s_jump[s]];
        __syncthreads();
        q = 0;
        for (int s = tid; s < n_seq; s += 1024, ++q) s_jump[s] = j2[q];
        __syncthreads();
    }
    // item number of every chain node = exclusive prefix sum of the marks (1024 sequences per sweep)
    for (int base = 0; base < n_seq; base += 1024) {
        const int s = base + tid;
        const int m = (s < n_seq) ? s_mark[s] : 0;
        int incl = m;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            int w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += v;
            }
            s_warp[lane] = w;   // inclusive over warps
        }
        __syncthreads();
        const int carry = s_carry;
        const int idx = carry + (warp ? s_warp[warp - 1] : 0) + incl - m;
        if (m) {
            const int row0 = cu[s];
            items[2 * idx] = row0;
            items[2 * idx + 1] = s;
            const int e = s + s_nxt[s];
            for (int u = s; u < e; ++u) seq_row0[u] = row0;
        }
        __syncthreads();
        if (tid == 0) s_carry = carry + s_warp[31];
        __syncthreads();
    }
    if (tid == 0) {
        const int n_items = s_carry, total = cu[n_seq];
        items[2 * n_items] = total;
        items[2 * n_items + 1] = n_seq;
        plan[0] = n_items;
        plan[1] = s_err;
        plan[2] = n_seq;
        plan[3] = total;
    }
}

__global__ void __launch_bounds__(256) varlen_plan_rows_kernel(const int32_t *__restrict__ cu, int n_seq, int32_t *__restrict__ plan) {
    const int32_t *seq_row0 = plan + plan_scratch_off(n_seq);
    int32_t *rowwin = plan + plan_rows_off(n_seq);
    const int s = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (s >= n_seq) return;
    const int a = cu[s], b = cu[s + 1], r0 = seq_row0[s];
    const int32_t win = (a - r0) | ((b - r0) << 16);
    for (int g = a + lane; g < b; g += 32) rowwin[g] = win;
}

// ------------------------------------------------------------------------------------------------ forward
struct VlFwdCtrl {
    uint64_t bar_qk, bar_v, bar_s, bar_o;
    uint32_t tmem_base;
    float xmax[2][128], xsum[2][128];
};
constexpr int kVlFwdTiles = 3 * 16384 + 32768;  // Q K V | P
constexpr int kVlFwdSmem = kVlFwdTiles + (int)sizeof(VlFwdCtrl);

__global__ void __launch_bounds__(kVlThreads, 2)
    attn_varlen_fwd_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                           const __grid_constant__ CUtensorMap mv, const VarlenArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *sQ = smem, *sK = smem + 16384, *sV = smem + 32768, *sP = smem + 49152;
    VlFwdCtrl &ctl = *reinterpret_cast<VlFwdCtrl *>(smem + kVlFwdTiles);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int r = tid & 127, hf = tid >> 7;  // query row, which half of its key / output columns
    const int n_items = a.plan[0];
    const int n_units = a.plan[1] ? 0 : n_items * a.H;  // a flagged plan (sequence > 128 rows) runs nothing
    const int32_t *rowwin = a.plan + plan_rows_off(a.n_seq);

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
    const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const uint32_t idesc_s = idesc_bf16(128, 128, 0, 0);
    const uint32_t idesc_o = idesc_bf16(128, 64, 0, 1);

    VlUnit u{0, 0, 0}, un{0, 0, 0};
    if ((int)blockIdx.x < n_units) u = vl_unit(a, blockIdx.x);
    if (tid == 0 && (int)blockIdx.x < n_units) {
        mbar_expect_tx(&ctl.bar_qk, 32768);
        tma_load_2d(sQ, &mq, &ctl.bar_qk, u.h * kD, u.row0);
        tma_load_2d(sK, &mk, &ctl.bar_qk, u.h * kD, u.row0);
        mbar_expect_tx(&ctl.bar_v, 16384);
        tma_load_2d(sV, &mv, &ctl.bar_v, u.h * kD, u.row0);
    }

    int it = 0;
    for (int w = blockIdx.x; w < n_units; w += gridDim.x, ++it, u = un) {
        const uint32_t par = it & 1;
        const int w_next = w + gridDim.x;
        const bool has_next = w_next < n_units;
        if (has_next) un = vl_unit(a, w_next);  // the next unit's header: its loads complete under this unit's work
        const bool valid = r < u.nrows;
        int lo = 0, hi = 0;
        if (valid) {
            const int32_t win = __ldg(rowwin + u.row0 + r);
            lo = win & 0xffff, hi = win >> 16;
        }
        const int nch = (u.nrows + 31) >> 5;
        const bool warp_live = (warp & 3) * 32 < u.nrows;

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
        if (tid == 0 && has_next) {  // Q, K are free: the next unit's tiles fly under this softmax
            mbar_expect_tx(&ctl.bar_qk, 32768);
            tma_load_2d(sQ, &mq, &ctl.bar_qk, un.h * kD, un.row0);
            tma_load_2d(sK, &mk, &ctl.bar_qk, un.h * kD, un.row0);
        }
        __syncwarp();

        // ---- this thread's <= 64 scores (chunks 2 hf, 2 hf + 1): masking only in chunks the window [lo, hi) cuts through
        float x[64];
        uint32_t bits[2] = {0u, 0u};
        float mloc = neg_inf();
        if (warp_live) {
#pragma unroll
            for (int j = 0; j < 2; ++j)
                if (2 * hf + j < nch) tmem_ld_32x32(lane_base + (2 * hf + j) * 32, reinterpret_cast<uint32_t *>(x) + j * 32);
            tmem_wait_ld();
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int c = 2 * hf + j;
                bits[j] = c < nch ? window_bits(lo, hi, c) : 0u;
                if (bits[j] == 0u) continue;
                if (bits[j] != 0xffffffffu) {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (!((bits[j] >> i) & 1u)) x[j * 32 + i] = neg_inf();
                }
                float mx0 = neg_inf(), mx1 = neg_inf();
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    mx0 = fmax3(mx0, x[j * 32 + i], x[j * 32 + i + 1]);
                    mx1 = fmax3(mx1, x[j * 32 + i + 2], x[j * 32 + i + 3]);
                }
                mloc = fmax3(mloc, mx0, mx1);
            }
        }
        ctl.xmax[hf][r] = mloc;
        __syncthreads();
        const float m = fmaxf(mloc, ctl.xmax[hf ^ 1][r]) * a.scale_log2;
        float lpart = 0.f;
        if (warp_live) {
            const float nm = (m == neg_inf()) ? 0.f : -m;
            const float sl2 = a.scale_log2;
            const uint32_t salt = drop_salt(a.drop, (uint32_t)w), thr32 = a.drop.thr16 << 16;
            float sum0 = 0.f, sum1 = 0.f;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int c = 2 * hf + j;
                if (c >= nch) continue;
                uint8_t *prow = sP + hf * 16384;
                if (bits[j] == 0u) {
#pragma unroll
                    for (int ch = 0; ch < 4; ++ch)
                        *reinterpret_cast<uint4 *>(prow + sw128_offset(r, j * 4 + ch)) = make_uint4(0, 0, 0, 0);
                    continue;
                }
                float *xc = x + j * 32;
                const uint32_t dw = drop_chunk(salt, (uint32_t)r, (uint32_t)c);
#pragma unroll
                for (int i = 0; i < 32; i += 2) {   // masked entries are -inf: 2^(-inf) = 0, no select
                    ffma2(xc[i], xc[i + 1], xc[i], xc[i + 1], sl2, sl2, nm, nm);
                    xc[i] = ex2(xc[i]);
                    xc[i + 1] = ex2(xc[i + 1]);
                    fadd2(sum0, sum1, sum0, sum1, xc[i], xc[i + 1]);   // the row sum is of the undropped p
                    if (a.drop.thr16) {
                        xc[i] = drop_keep_w(dw, drop_lane(i), thr32) ? xc[i] * a.drop.scale : 0.f;
                        xc[i + 1] = drop_keep_w(dw, drop_lane(i + 1), thr32) ? xc[i + 1] * a.drop.scale : 0.f;
                    }
                }
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) {
                    uint4 v;
                    v.x = pack_bf16x2(xc[ch * 8 + 0], xc[ch * 8 + 1]);
                    v.y = pack_bf16x2(xc[ch * 8 + 2], xc[ch * 8 + 3]);
                    v.z = pack_bf16x2(xc[ch * 8 + 4], xc[ch * 8 + 5]);
                    v.w = pack_bf16x2(xc[ch * 8 + 6], xc[ch * 8 + 7]);
                    *reinterpret_cast<uint4 *>(prow + sw128_offset(r, j * 4 + ch)) = v;
                }
            }
            lpart = sum0 + sum1;
        }
        ctl.xsum[hf][r] = lpart;
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
        const float l = lpart + ctl.xsum[hf ^ 1][r];
        mbar_wait(&ctl.bar_o, par);
        tc_fence_after();
        if (tid == 0 && has_next) {  // V and P are free
            mbar_expect_tx(&ctl.bar_v, 16384);
            tma_load_2d(sV, &mv, &ctl.bar_v, un.h * kD, un.row0);
        }
        __syncwarp();
        if (warp_live) {
            float o[32];
            tmem_ld_32x32(lane_base + 128 + hf * 32, reinterpret_cast<uint32_t *>(o));
            tmem_wait_ld();
            if (valid) {
                const float inv = l > 0.f ? 1.f / l : 0.f;
                const size_t g = (size_t)(u.row0 + r);
                __nv_bfloat16 *op = a.out + g * a.o_stride + (size_t)u.h * kD + hf * 32;
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) {
                    uint4 v;
                    v.x = pack_bf16x2(o[ch * 8 + 0] * inv, o[ch * 8 + 1] * inv);
                    v.y = pack_bf16x2(o[ch * 8 + 2] * inv, o[ch * 8 + 3] * inv);
                    v.z = pack_bf16x2(o[ch * 8 + 4] * inv, o[ch * 8 + 5] * inv);
                    v.w = pack_bf16x2(o[ch * 8 + 6] * inv, o[ch * 8 + 7] * inv);
                    reinterpret_cast<uint4 *>(op)[ch] = v;
                }
                if (a.lse && hf == 0) a.lse[(size_t)u.h * a.total + g] = l > 0.f ? m * kLn2 + logf(l) : __int_as_float(0x7f800000);
            }
        }
        tc_fence_before();  // the S / O columns and the exchange words are rewritten by the next unit
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
constexpr int kVlBwdSmem = kVlBwdTiles + 64;  // 2 CTAs / SM: 2 x (112 KB + 64 B + 1 KB reserved) <= 228 KB

__global__ void __launch_bounds__(kVlThreads, 2)
    attn_varlen_bwd_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                           const __grid_constant__ CUtensorMap mv, const __grid_constant__ CUtensorMap mdo, const VarlenArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *sQ = smem, *sK = smem + 16384, *sdO = smem + 32768, *sP = smem + 49152, *sV = smem + 65536, *sdS = smem + 81920;
    VlBwdCtrl &ctl = *reinterpret_cast<VlBwdCtrl *>(smem + kVlBwdTiles);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int r = tid & 127, hf = tid >> 7;
    const int n_items = a.plan[0];
    const int n_units = a.plan[1] ? 0 : n_items * a.H;
    const int32_t *rowwin = a.plan + plan_rows_off(a.n_seq);

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
    const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    constexpr uint32_t cS = 0, cDP = 128, cDV = 0, cDK = 64, cDQ = 128;  // dV / dK / dQ reuse the S / dP columns
    const uint32_t idesc_s = idesc_bf16(128, 128, 0, 0);
    const uint32_t idesc_kv = idesc_bf16(128, 64, 1, 1);  // A = P / dS read MN-major (M = key), B = dO / Q MN-major
    const uint32_t idesc_dq = idesc_bf16(128, 64, 0, 1);  // A = dS K-major (M = query), B = K MN-major

    auto load_unit = [&](const VlUnit &x) {  // tid 0: the four operand tiles of a unit
        mbar_expect_tx(&ctl.bar_ld, 65536);
        tma_load_2d(sQ, &mq, &ctl.bar_ld, x.h * kD, x.row0);
        tma_load_2d(sK, &mk, &ctl.bar_ld, x.h * kD, x.row0);
        tma_load_2d(sV, &mv, &ctl.bar_ld, x.h * kD, x.row0);
        tma_load_2d(sdO, &mdo, &ctl.bar_ld, x.h * kD, x.row0);
    };
    VlUnit u{0, 0, 0}, un{0, 0, 0};
    if ((int)blockIdx.x < n_units) u = vl_unit(a, blockIdx.x);
    if (tid == 0 && (int)blockIdx.x < n_units) load_unit(u);
    int32_t win_c = 0;
    float lse_c = 0.f;
    if ((int)blockIdx.x < n_units && r < u.nrows) {
        win_c = __ldg(rowwin + u.row0 + r);
        lse_c = __ldg(a.lse + (size_t)u.h * a.total + u.row0 + r);
    }

    int it = 0;
    for (int w = blockIdx.x; w < n_units; w += gridDim.x, ++it, u = un) {
        const uint32_t par = it & 1;
        const int w_next = w + gridDim.x;
        const bool has_next = w_next < n_units;
        if (has_next) un = vl_unit(a, w_next);
        const bool valid = r < u.nrows;
        const int nch = (u.nrows + 31) >> 5;
        const bool warp_used = (warp & 3) < nch;  // rows the dV / dK k-steps read (live, or zero-filled): rows [0, 32 nch)
        // ---- row window and -lse (log2 units) of this row: requested one unit ahead (win_n / lse_n below)
        const int lo = valid ? (win_c & 0xffff) : 0, hi = valid ? (win_c >> 16) : 0;
        const float nlse = (!valid || lse_c == __int_as_float(0x7f800000)) ? neg_inf() : -lse_c * kLog2e;  // +inf: a row without keys
        int32_t win_n = 0;
        float lse_n = 0.f;
        if (has_next && r < un.nrows) {
            const size_t gn = (size_t)(un.row0 + r);
            win_n = __ldg(rowwin + gn);
            lse_n = __ldg(a.lse + (size_t)un.h * a.total + gn);
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
        // ---- pass A: delta = rowsum(dO * O) = sum_k keep P dP / (1 - p_drop) over the row's window -- from S and dP alone,
        // no O / dO rows from global memory (their DRAM latency sat on every unit's critical path, and 25 % of the traffic)
        const uint32_t salt = drop_salt(a.drop, (uint32_t)w), thr32 = a.drop.thr16 << 16;
        const float sl2 = a.scale_log2, sc = a.scale;
        float dpart = 0.f;
        if (warp_used) {
#pragma unroll 1
            for (int j = 0; j < 2; ++j) {
                const int c = 2 * hf + j;
                if (c >= nch) break;   // uniform across the warp
                const uint32_t bits = window_bits(lo, hi, c);
                if (!__any_sync(0xffffffffu, bits != 0u)) continue;   // warp-uniform
                float st[32], dp[32];
                tmem_ld_32x32(lane_base + cS + c * 32, reinterpret_cast<uint32_t *>(st));
                tmem_ld_32x32(lane_base + cDP + c * 32, reinterpret_cast<uint32_t *>(dp));
                tmem_wait_ld();
                if (bits != 0u) {
                    float a0 = 0.f, a1 = 0.f;
                    const uint32_t dw = drop_chunk(salt, (uint32_t)r, (uint32_t)c);
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        float p0 = (bits >> i) & 1u ? ex2(fmaf(st[i], sl2, nlse)) : 0.f;
                        float p1 = (bits >> (i + 1)) & 1u ? ex2(fmaf(st[i + 1], sl2, nlse)) : 0.f;
                        if (a.drop.thr16) {
                            p0 = drop_keep_w(dw, drop_lane(i), thr32) ? p0 * a.drop.scale : 0.f;
                            p1 = drop_keep_w(dw, drop_lane(i + 1), thr32) ? p1 * a.drop.scale : 0.f;
                        }
                        a0 = fmaf(p0, dp[i], a0);
                        a1 = fmaf(p1, dp[i + 1], a1);
                    }
                    dpart += a0 + a1;
                }
                __syncwarp();
            }
        }
        // the halves of a row exchange their partial sums through the dS tile (free until pass B): each thread's word lies
        // inside the bytes only its READER overwrites later (row r of the reader's half), so one barrier is enough
        *reinterpret_cast<float *>(sdS + (hf ^ 1) * 16384 + sw128_offset(r, 0)) = dpart;
        __syncthreads();
        const float delta = dpart + *reinterpret_cast<const float *>(sdS + hf * 16384 + sw128_offset(r, 0));
        if (warp_used) {
            const float nds = -delta * sc;   // dS * scale = p * (keep * dP * scale - delta * scale): d(q k^T), not d(scaled score)
#pragma unroll 1
            for (int j = 0; j < 2; ++j) {
                const int c = 2 * hf + j;
                if (c >= nch) break;   // uniform across the warp
                const uint32_t bits = window_bits(lo, hi, c);  // 0 for rows past the item and for the other sequences' columns
                const uint32_t off0 = hf * 16384;
                float st[32], dp[32];   // warp-collective TMEM reads: every lane, whatever its window
                tmem_ld_32x32(lane_base + cS + c * 32, reinterpret_cast<uint32_t *>(st));
                tmem_ld_32x32(lane_base + cDP + c * 32, reinterpret_cast<uint32_t *>(dp));
                tmem_wait_ld();
                if (bits == 0u) {   // per-thread branch: no warp-collective operation until both sides have rejoined
#pragma unroll
                    for (int ch = 0; ch < 4; ++ch) {
                        const uint32_t off = off0 + sw128_offset(r, j * 4 + ch);
                        *reinterpret_cast<uint4 *>(sP + off) = make_uint4(0, 0, 0, 0);
                        *reinterpret_cast<uint4 *>(sdS + off) = make_uint4(0, 0, 0, 0);
                    }
                } else {
                    const uint32_t dw = drop_chunk(salt, (uint32_t)r, (uint32_t)c);
                    if (bits != 0xffffffffu) {
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (!((bits >> i) & 1u)) st[i] = neg_inf();   // p = 2^(-inf) = 0 and dS = 0 * finite = 0
                    }
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        ffma2(st[i], st[i + 1], st[i], st[i + 1], sl2, sl2, nlse, nlse);
                        st[i] = ex2(st[i]);
                        st[i + 1] = ex2(st[i + 1]);
                        if (a.drop.thr16) {
                            const float k0 = drop_keep_w(dw, drop_lane(i), thr32) ? a.drop.scale : 0.f;
                            const float k1 = drop_keep_w(dw, drop_lane(i + 1), thr32) ? a.drop.scale : 0.f;
                            ffma2(dp[i], dp[i + 1], dp[i], dp[i + 1], k0 * sc, k1 * sc, nds, nds);
                            fmul2(dp[i], dp[i + 1], dp[i], dp[i + 1], st[i], st[i + 1]);
                            fmul2(st[i], st[i + 1], st[i], st[i + 1], k0, k1);
                        } else {
                            ffma2(dp[i], dp[i + 1], dp[i], dp[i + 1], sc, sc, nds, nds);
                            fmul2(dp[i], dp[i + 1], dp[i], dp[i + 1], st[i], st[i + 1]);
                        }
                    }
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
                        const uint32_t off = off0 + sw128_offset(r, j * 4 + ch);
                        *reinterpret_cast<uint4 *>(sP + off) = vp;   // half 1 of P overwrites the V tile (dP has completed)
                        *reinterpret_cast<uint4 *>(sdS + off) = vd;
                    }
                }
                __syncwarp();
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
        // every shared-memory tile has been consumed by the MMAs: the next unit's loads fly under this unit's epilogue
        if (tid == 0 && has_next) load_unit(un);
        __syncwarp();
        if ((warp & 3) * 32 < u.nrows) {
            const size_t g = (size_t)(u.row0 + r);
#pragma unroll
            for (int which = 0; which < 3; ++which) {
                float acc[32];
                tmem_ld_32x32(lane_base + (which == 0 ? cDV : (which == 1 ? cDK : cDQ)) + hf * 32, reinterpret_cast<uint32_t *>(acc));
                tmem_wait_ld();
                if (valid) {
                    __nv_bfloat16 *p = which == 0 ? a.dv + g * a.dv_stride : (which == 1 ? a.dk + g * a.dk_stride : a.dq + g * a.dq_stride);
                    uint4 *dst = reinterpret_cast<uint4 *>(p + (size_t)u.h * kD + hf * 32);
#pragma unroll
                    for (int ch = 0; ch < 4; ++ch)
                        dst[ch] = make_uint4(pack_bf16x2(acc[ch * 8 + 0], acc[ch * 8 + 1]), pack_bf16x2(acc[ch * 8 + 2], acc[ch * 8 + 3]),
                                             pack_bf16x2(acc[ch * 8 + 4], acc[ch * 8 + 5]), pack_bf16x2(acc[ch * 8 + 6], acc[ch * 8 + 7]));
                }
            }
        }
        win_c = win_n, lse_c = lse_n;
        tc_fence_before();
        __syncthreads();  // every TMEM column and the exchange words of this unit are free for the next one
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

extern "C" int64_t mdt_varlen_attn_plan_words(int n_seq, int total) { return (int64_t)plan_rows_off(n_seq) + total; }

extern "C" int mdt_varlen_attn_plan(const int32_t *cu_seqlens, int n_seq, int total, int group, int32_t *plan, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(cu_seqlens && plan && n_seq >= 1 && total >= 1, "mdt_varlen_attn_plan: cu_seqlens, plan must be non-null, n_seq and total >= 1");
    if (n_seq <= kPlanParSeqs) {
        varlen_plan_items_kernel<<<1, 1024, 0, as_stream(stream)>>>(cu_seqlens, n_seq, group, plan);
        MDT_CHECK_CUDA(cudaGetLastError());
        varlen_plan_rows_kernel<<<(n_seq + 7) / 8, 256, 0, as_stream(stream)>>>(cu_seqlens, n_seq, plan);
    } else {
        varlen_plan_kernel<<<1, 1024, 0, as_stream(stream)>>>(cu_seqlens, n_seq, group, plan);
    }
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

extern "C" int mdt_varlen_attn_fwd(const void *q, const void *k, const void *v, int64_t q_stride, int64_t k_stride,
                                   int64_t v_stride, int total, int H, int n_seq, const int32_t *plan, void *out,
                                   int64_t o_stride, float *lse, float softmax_scale, float dropout_p,
                                   uint64_t dropout_seed, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(q && k && v && out && lse && plan, "mdt_varlen_attn_fwd: null pointer");
    MDT_REQUIRE(total >= 1 && H >= 1 && n_seq >= 1, "mdt_varlen_attn_fwd: total, H, n_seq must be >= 1");
    MDT_REQUIRE(((uintptr_t)out & 15) == 0 && o_stride % 8 == 0, "mdt_varlen_attn_fwd: out must be 16-byte aligned, stride a multiple of 8");
    MDT_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "mdt_varlen_attn_fwd: dropout_p must be in [0, 1)");
    MDT_REQUIRE(softmax_scale > 0.f, "mdt_varlen_attn_fwd: softmax_scale must be positive");
    CUtensorMap mq, mk, mv;
    if (int rc = make_rows_map(&mq, q, total, H, q_stride)) return rc;
    if (int rc = make_rows_map(&mk, k, total, H, k_stride)) return rc;
    if (int rc = make_rows_map(&mv, v, total, H, v_stride)) return rc;
    VarlenArgs a{};
    a.total = total, a.H = H, a.n_seq = n_seq, a.plan = plan;
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
                                   int64_t g_stride, const float *lse, int total, int H, int n_seq, const int32_t *plan,
                                   void *dq, void *dk, void *dv, int64_t dq_stride, int64_t dk_stride, int64_t dv_stride,
                                   float softmax_scale, float dropout_p, uint64_t dropout_seed, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(q && k && v && out && grad_out && lse && plan && dq && dk && dv, "mdt_varlen_attn_bwd: null pointer");
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
    a.total = total, a.H = H, a.n_seq = n_seq, a.plan = plan;
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

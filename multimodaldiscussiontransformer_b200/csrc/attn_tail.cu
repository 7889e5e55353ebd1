// Tail keys of the biased-attention backward (A6) for T = 128 k + r with a small r -- in this model T = comments + 1 graph
// token, so a batch of 128-comment graphs has T = 129 and one key sticks out of the 128-wide tcgen05 tiles.  Through the
// tile kernel those r keys cost a whole CTA per (b, h) that runs every query tile for an almost empty key tile (half of
// all CTAs at T = 129).  Here the tile kernel covers keys [0, Tmain), Tmain = T - r, and one small CUDA-core kernel does
// keys [Tmain, T) for all T queries: dK, dV of those keys, their dQ contributions (fp32 atomics into dq_accum) and their
// share of the LUT gradient.  Each element is a 64-long dot product and there are only r * T of them per (b, h):
// a warp per query, two head columns per lane.
// (The same split for tail *queries* and for the forward was measured and dropped: the extra launches and their strided
// row traffic cost more than the almost-empty tiles they remove -- DESIGN.md section 4.)
#include "attn_common.cuh"

namespace mdt {
namespace attn {

constexpr int kTailThreads = 1024;  // 32 warps: each (query, key) element is a chain of dependent global accesses, so width hides it
constexpr int kTailBins = 24;  // codes 0..22, bin 23 = MDT_CODE_VIRTUAL (same binning as attn_bwd.cu)

__device__ __forceinline__ float tail_warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int tail_bin(uint32_t c) { return c == MDT_CODE_VIRTUAL ? kTailBins - 1 : (int)min(c, (uint32_t)(kTailBins - 2)); }
__device__ __forceinline__ float2 load_pair(const __nv_bfloat16 *row, int lane) {
    return __bfloat1622float2(reinterpret_cast<const __nv_bfloat162 *>(row)[lane]);
}

// ------------------------------------------------------------------------------------------------ backward, tail keys
// grid (H, B): all T queries against keys [Tmain, T).  A warp takes one query at a time and lane l owns head columns 2l, 2l+1:
// every row access (q, dO, the dQ atomics) is one coalesced 128 / 256-byte transaction, the two dot products are two warp
// sums, and p, dS come out uniform across the warp so the LUT-gradient bin of this (query, key) is a single predicated add.
template <bool HAS_CODE>
__global__ void __launch_bounds__(kTailThreads) attn_bwd_tail_k_kernel(const TailArgs a) {
    constexpr int kWarps = kTailThreads / 32;
    __shared__ float s_lut[256], s_red[kWarps][128], s_bins[kWarps][kTailBins];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int h = blockIdx.x, b = blockIdx.y, T = a.T;
    const bool want_glut = HAS_CODE && a.grad_lut != nullptr;
    if (HAS_CODE && tid < 256) s_lut[tid] = tid == MDT_CODE_MASKED ? neg_inf() : a.lut[(size_t)h * 256 + tid] * kLog2e;  // log2 domain
    __syncthreads();
    const uint32_t salt = drop_salt(a.drop, (uint32_t)(b * a.H + h));
    const __nv_bfloat16 *qb = a.q + (size_t)b * a.q_sb + (size_t)h * a.q_sh;
    const __nv_bfloat16 *gb = a.dout + (size_t)b * a.g_sb + (size_t)h * a.g_sh;
    const size_t row0 = ((size_t)b * a.H + h) * T;
    float mybin = 0.f;  // lane l keeps bin l of this warp
    for (int k = a.Tmain; k < T; ++k) {
        __nv_bfloat16 *dkp = a.dk + (size_t)b * a.dk_sb + (size_t)k * a.dk_st + (size_t)h * a.dk_sh;
        __nv_bfloat16 *dvp = a.dv + (size_t)b * a.dv_sb + (size_t)k * a.dv_st + (size_t)h * a.dv_sh;
        if (a.key_pad != nullptr && a.key_pad[(size_t)b * T + k] != 0) {  // uniform: a padded key has dK = dV = 0
            if (tid < 64) dkp[tid] = dvp[tid] = __float2bfloat16_rn(0.f);
            continue;
        }
        const float2 kk = load_pair(a.k + (size_t)b * a.k_sb + (size_t)k * a.k_st + (size_t)h * a.k_sh, lane);
        const float2 vv = load_pair(a.v + (size_t)b * a.v_sb + (size_t)k * a.v_st + (size_t)h * a.v_sh, lane);
        float2 accv = make_float2(0.f, 0.f), acck = make_float2(0.f, 0.f);
#pragma unroll 4
        for (int q = warp; q < T; q += kWarps) {
            const float2 qv = load_pair(qb + (size_t)q * a.q_st, lane);
            const float2 gv = load_pair(gb + (size_t)q * a.g_st, lane);
            const float lse = a.lse[row0 + q], delta = a.delta[row0 + q];
            uint32_t code = 0;
            if (HAS_CODE) code = a.code[((size_t)b * T + q) * a.code_pitch + k];
            const float s = tail_warp_sum(fmaf(qv.x, kk.x, qv.y * kk.y));
            const float dp = tail_warp_sum(fmaf(gv.x, vv.x, gv.y * vv.y));
            float x = (s - lse) * kLog2e;
            if (HAS_CODE) x += s_lut[code];
            const float p0 = sm100::ex2(x);
            float keepf = 1.f;
            if (a.drop.thr16) keepf = drop_keep_one(a.drop, salt, (uint32_t)q, (uint32_t)k) ? a.drop.scale : 0.f;
            const float ds = p0 * (dp * keepf - delta);
            const float p = p0 * keepf;  // what dV sees
            if (p0 != 0.f) {
                accv.x = fmaf(p, gv.x, accv.x);
                accv.y = fmaf(p, gv.y, accv.y);
                acck.x = fmaf(ds, qv.x, acck.x);
                acck.y = fmaf(ds, qv.y, acck.y);
                atomicAdd(reinterpret_cast<float2 *>(a.dq_accum + (row0 + q) * 64) + lane, make_float2(ds * kk.x, ds * kk.y));
                if (want_glut && lane == tail_bin(code)) mybin += ds;
            }
        }
        __syncthreads();  // s_red of the previous key has been read
        s_red[warp][2 * lane] = accv.x;
        s_red[warp][2 * lane + 1] = accv.y;
        s_red[warp][64 + 2 * lane] = acck.x;
        s_red[warp][64 + 2 * lane + 1] = acck.y;
        __syncthreads();
        if (tid < 128) {
            float sum = 0.f;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) sum += s_red[w][tid];
            (tid < 64 ? dvp : dkp)[tid & 63] = __float2bfloat16_rn(sum);
        }
    }
    if (want_glut) {
        if (lane < kTailBins) s_bins[warp][lane] = mybin;
        __syncthreads();
        if (tid < kTailBins) {
            float sum = 0.f;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) sum += s_bins[w][tid];
            const int code = tid == kTailBins - 1 ? MDT_CODE_VIRTUAL : tid;
            if (sum != 0.f) atomicAdd(a.grad_lut + (size_t)h * 256 + code, sum);
        }
    }
}

// ------------------------------------------------------------------------------------------------ launchers
int launch_bwd_tails(const TailArgs &a, cudaStream_t st) {
    const int r = a.T - a.Tmain;
    if (r <= 0) return MDT_OK;
    const dim3 gk(a.H, a.B);
    if (a.code)
        attn_bwd_tail_k_kernel<true><<<gk, kTailThreads, 0, st>>>(a);
    else
        attn_bwd_tail_k_kernel<false><<<gk, kTailThreads, 0, st>>>(a);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

}  // namespace attn
}  // namespace mdt

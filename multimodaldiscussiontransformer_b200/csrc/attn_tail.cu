// Tail keys of the biased-attention backward (A6) for T = 128 k + r with a small r -- in this model T = comments + 1 graph
// token, so a batch of 128-comment graphs has T = 129 and one key sticks out of the 128-wide tcgen05 tiles.  Through the
// tile kernel those r keys cost a whole CTA per (b, h) that runs every query tile for an almost empty key tile (half of
// all CTAs at T = 129).  Here the tile kernel covers keys [0, Tmain), Tmain = T - r, and one small CUDA-core kernel does
// keys [Tmain, T) for all T queries: dK, dV of those keys, their dQ contributions (fp32 atomics into dq_accum) and their
// share of the LUT gradient.  Each element is a 64-long dot product and there are only r * T of them per (b, h).
// (The same split for tail *queries* and for the forward was measured and dropped: the extra launches and their strided
// row traffic cost more than the almost-empty tiles they remove -- DESIGN.md section 4.)
#include "attn_common.cuh"

namespace mdt {
namespace attn {

constexpr int kTailThreads = 256;
constexpr int kTailBins = 24;  // codes 0..22, bin 23 = MDT_CODE_VIRTUAL (same binning as attn_bwd.cu)

__device__ __forceinline__ void load_row64(const __nv_bfloat16 *p, float *f) {
    const uint4 *v = reinterpret_cast<const uint4 *>(p);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const uint4 w = v[i];
        const __nv_bfloat162 *h = reinterpret_cast<const __nv_bfloat162 *>(&w);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 t = __bfloat1622float2(h[j]);
            f[i * 8 + 2 * j] = t.x;
            f[i * 8 + 2 * j + 1] = t.y;
        }
    }
}
__device__ __forceinline__ float tail_warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int tail_bin(uint32_t c) { return c == MDT_CODE_VIRTUAL ? kTailBins - 1 : (int)min(c, (uint32_t)(kTailBins - 2)); }

// log2-domain LUT of head h in shared memory (256 entries, -inf for MDT_CODE_MASKED)
template <bool HAS_CODE> __device__ __forceinline__ void tail_stage_lut(float *s_lut, const TailArgs &a, int h) {
    if (HAS_CODE) s_lut[threadIdx.x] = threadIdx.x == MDT_CODE_MASKED ? neg_inf() : a.lut[(size_t)h * 256 + threadIdx.x] * kLog2e;
}
// every warp adds `ds` of its lanes into per-code sums: lane b of the warp keeps bin b
__device__ __forceinline__ void tail_bin_add(float &mybin, uint32_t code, float ds, int lane) {
    const int bin = tail_bin(code);
#pragma unroll
    for (int b = 0; b < kTailBins; ++b) {
        const float s = tail_warp_sum(bin == b ? ds : 0.f);
        if (lane == b) mybin += s;
    }
}
__device__ __forceinline__ void tail_bin_flush(float mybin, float (*s_bins)[kTailBins], float *grad_lut, int h) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane < kTailBins) s_bins[warp][lane] = mybin;
    __syncthreads();
    if (threadIdx.x < kTailBins) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < kTailThreads / 32; ++w) s += s_bins[w][threadIdx.x];
        const int code = threadIdx.x == kTailBins - 1 ? MDT_CODE_VIRTUAL : (int)threadIdx.x;
        if (s != 0.f) atomicAdd(grad_lut + (size_t)h * 256 + code, s);
    }
}

// ------------------------------------------------------------------------------------------------ backward, tail keys
// grid (H, B): all T queries against keys [Tmain, T)
template <bool HAS_CODE>
__global__ void __launch_bounds__(kTailThreads, 1) attn_bwd_tail_k_kernel(const TailArgs a) {
    __shared__ float s_k[64], s_v[64], s_lut[256], s_red[kTailThreads / 32][128], s_bins[kTailThreads / 32][kTailBins];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int h = blockIdx.x, b = blockIdx.y, T = a.T;
    const bool want_glut = HAS_CODE && a.grad_lut != nullptr;
    tail_stage_lut<HAS_CODE>(s_lut, a, h);
    const uint32_t salt = drop_salt(a.drop, (uint32_t)(b * a.H + h));
    float mybin = 0.f;
    for (int k = a.Tmain; k < T; ++k) {
        __nv_bfloat16 *dkp = a.dk + (size_t)b * a.dk_sb + (size_t)k * a.dk_st + (size_t)h * a.dk_sh;
        __nv_bfloat16 *dvp = a.dv + (size_t)b * a.dv_sb + (size_t)k * a.dv_st + (size_t)h * a.dv_sh;
        if (a.key_pad != nullptr && a.key_pad[(size_t)b * T + k] != 0) {  // uniform: a padded key has dK = dV = 0
            if (tid < 64) dkp[tid] = dvp[tid] = __float2bfloat16_rn(0.f);
            continue;
        }
        __syncthreads();
        if (tid < 64) {
            s_k[tid] = __bfloat162float(a.k[(size_t)b * a.k_sb + (size_t)k * a.k_st + (size_t)h * a.k_sh + tid]);
            s_v[tid] = __bfloat162float(a.v[(size_t)b * a.v_sb + (size_t)k * a.v_st + (size_t)h * a.v_sh + tid]);
        }
        __syncthreads();
        float accv[64], acck[64];
#pragma unroll
        for (int c = 0; c < 64; ++c) accv[c] = acck[c] = 0.f;
        for (int q0 = 0; q0 < T; q0 += kTailThreads) {
            const int q = q0 + tid;
            float ds = 0.f;
            uint32_t code = MDT_CODE_MASKED;
            if (q < T) {
                float qr[64], gr[64];
                load_row64(a.q + (size_t)b * a.q_sb + (size_t)q * a.q_st + (size_t)h * a.q_sh, qr);
                load_row64(a.dout + (size_t)b * a.g_sb + (size_t)q * a.g_st + (size_t)h * a.g_sh, gr);
                float s = 0.f, dp = 0.f;
#pragma unroll
                for (int c = 0; c < 64; ++c) {
                    s = fmaf(qr[c], s_k[c], s);
                    dp = fmaf(gr[c], s_v[c], dp);
                }
                const size_t row = ((size_t)b * a.H + h) * T + q;
                float x = s * kLog2e - a.lse[row] * kLog2e;
                if (HAS_CODE) {
                    code = a.code[((size_t)b * T + q) * a.code_pitch + k];
                    x += s_lut[code];
                }
                const float p0 = sm100::ex2(x);
                float keepf = 1.f;
                if (a.drop.thr16) keepf = drop_keep_one(a.drop, salt, (uint32_t)q, (uint32_t)T, (uint32_t)k) ? a.drop.scale : 0.f;
                ds = p0 * (dp * keepf - a.delta[row]);
                const float p = p0 * keepf;  // what dV sees
                if (p0 != 0.f) {
                    float4 *dq = reinterpret_cast<float4 *>(a.dq_accum + row * 64);
#pragma unroll
                    for (int c = 0; c < 64; c += 4) {
                        accv[c] = fmaf(p, gr[c], accv[c]);
                        accv[c + 1] = fmaf(p, gr[c + 1], accv[c + 1]);
                        accv[c + 2] = fmaf(p, gr[c + 2], accv[c + 2]);
                        accv[c + 3] = fmaf(p, gr[c + 3], accv[c + 3]);
                        acck[c] = fmaf(ds, qr[c], acck[c]);
                        acck[c + 1] = fmaf(ds, qr[c + 1], acck[c + 1]);
                        acck[c + 2] = fmaf(ds, qr[c + 2], acck[c + 2]);
                        acck[c + 3] = fmaf(ds, qr[c + 3], acck[c + 3]);
                        atomicAdd(dq + (c >> 2), make_float4(ds * s_k[c], ds * s_k[c + 1], ds * s_k[c + 2], ds * s_k[c + 3]));
                    }
                }
            }
            if (want_glut) tail_bin_add(mybin, code, ds, lane);
        }
#pragma unroll
        for (int c = 0; c < 64; ++c) {
            const float sv = tail_warp_sum(accv[c]), sk = tail_warp_sum(acck[c]);
            if (lane == 0) {
                s_red[warp][c] = sv;
                s_red[warp][64 + c] = sk;
            }
        }
        __syncthreads();
        if (tid < 128) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < kTailThreads / 32; ++w) s += s_red[w][tid];
            (tid < 64 ? dvp : dkp)[tid & 63] = __float2bfloat16_rn(s);
        }
    }
    if (want_glut) tail_bin_flush(mybin, s_bins, a.grad_lut, h);
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

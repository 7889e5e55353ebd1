// A6 forward: O = softmax_fp32(Q K^T + bias [+ key padding]) V      multihead_attention.py:168-199
//
// One CTA per (batch b, head h, 128-query tile); 4 warps, thread r owns query row r (= TMEM
// lane r).  Per 128-key tile:
//   S  = Q K^T      tcgen05.mma 128x128x64  (Q, K tiles: TMA, 128-byte swizzle, K-major)  -> TMEM cols [0,128)
//   softmax         tcgen05.ld -> registers, + bias, masks, online max / sum (fp32), lazy rescale of O
//   P  -> bf16      written by the threads into a swizzled K-major tile (A operand)
//   O += P V        tcgen05.mma 128x64x128  (V tile: TMA, MN-major B operand)           -> TMEM cols [128,192)
// K/V tiles stream through a 3-slot ring (K_j, V_j, K_j+1 ...); two CTAs share an SM so that one
// CTA's MMAs overlap the other's softmax.  The bias is either a materialised [B,H,T,T] tensor or
// one byte per (b,q,k) expanded through lut[h][256] held in shared memory.
#include "attn_common.cuh"

namespace mdt {
namespace attn {
using namespace sm100;

struct FwdArgs {
    int B, H, T;
    const void *bias;
    const uint8_t *code;
    int code_pitch;
    const float *lut;
    const uint8_t *key_pad;
    __nv_bfloat16 *out;
    int64_t o_sb, o_st, o_sh;
    float *lse;
    DropArgs drop;
};

constexpr int kFwdSmem = 16384 /*Q*/ + 3 * 16384 /*K/V ring*/ + 32768 /*P*/ + 1024 /*alignment slack*/;
constexpr float kRescaleThreshold = 8.0f * kLn2;  // P may reach 2^8 before O is rescaled

template <int MODE>
__global__ void __launch_bounds__(128, 2)
    attn_fwd_kernel(const __grid_constant__ CUtensorMap mq, const __grid_constant__ CUtensorMap mk,
                    const __grid_constant__ CUtensorMap mv, const FwdArgs a) {
    extern __shared__ uint8_t smem_dyn[];
    uint8_t *smem = reinterpret_cast<uint8_t *>(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    uint8_t *sQ = smem, *sRing = smem + 16384, *sP = smem + 4 * 16384;
    __shared__ __align__(8) uint64_t bar_q, bar_full[3], bar_s, bar_pv;
    __shared__ uint32_t tmem_base_s;
    __shared__ uint32_t s_kmask[2][4];  // per key tile: bit set = key masked (padding or >= T)
    __shared__ float s_lut[256];

    const int tid = threadIdx.x, warp = tid >> 5;
    const int h = blockIdx.y, b = blockIdx.z;
    const int T = a.T;
    const int q0 = blockIdx.x * kTile;
    const int q = q0 + tid;
    const bool row_ok = q < T;
    const int nkv = (T + kTile - 1) / kTile;

    auto key_masked = [&](int kk) -> bool { return kk >= T || (a.key_pad != nullptr && a.key_pad[(size_t)b * T + kk] != 0); };

    if (tid == 0) {
        mbar_init(&bar_q, 1);
        mbar_init(&bar_full[0], 1);
        mbar_init(&bar_full[1], 1);
        mbar_init(&bar_full[2], 1);
        mbar_init(&bar_s, 1);
        mbar_init(&bar_pv, 1);
        fence_barrier_init();
        tma_prefetch_desc(&mq);
        tma_prefetch_desc(&mk);
        tma_prefetch_desc(&mv);
    }
    if (warp == 0) {
        tmem_alloc(&tmem_base_s, 256);
        tmem_relinquish();
    }
    if (MODE == kBiasCode) {
        s_lut[tid] = a.lut[(size_t)h * 256 + tid];
        s_lut[tid + 128] = (tid + 128 == MDT_CODE_MASKED) ? neg_inf() : a.lut[(size_t)h * 256 + tid + 128];
    }
    {
        const uint32_t bal = __ballot_sync(0xffffffffu, key_masked(tid));
        if ((tid & 31) == 0) s_kmask[0][warp] = bal;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);

    if (tid == 0) {
        mbar_expect_tx(&bar_q, 16384);
        tma_load_4d(sQ, &mq, &bar_q, 0, q0, h, b);
        mbar_expect_tx(&bar_full[0], 16384);
        tma_load_4d(sRing, &mk, &bar_full[0], 0, 0, h, b);
        mbar_expect_tx(&bar_full[1], 16384);
        tma_load_4d(sRing + 16384, &mv, &bar_full[1], 0, 0, h, b);
    }

    float m_used = neg_inf();  // the max the stored P / O are relative to (natural-log units)
    float l = 0.f;
    const uint32_t idesc_s = idesc_bf16(128, 128, 0, 0);
    const uint32_t idesc_o = idesc_bf16(128, 64, 0, 1);

    for (int j = 0; j < nkv; ++j) {
        const int k0 = j * kTile;
        const int nk = 2 * j, nv = 2 * j + 1;  // ring sequence numbers of K_j and V_j
        const int kslot = nk % 3, vslot = nv % 3;
        if (tid == 0) {
            if (j == 0) mbar_wait(&bar_q, 0);
            mbar_wait(&bar_full[kslot], (nk / 3) & 1);
            tc_fence_after();
            const uint32_t aq = smem_u32(sQ), bk = smem_u32(sRing + kslot * 16384);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                umma_bf16(tmem, smem_desc_sw128(aq + k * 32, 16, 1024), smem_desc_sw128(bk + k * 32, 16, 1024), idesc_s, k > 0);
            umma_commit(&bar_s);  // also orders after P V of tile j-1 (the tensor pipe runs in issue order)
        }
        __syncwarp();
        mbar_wait(&bar_s, j & 1);
        tc_fence_after();
        if (tid == 0 && j + 1 < nkv) {  // slots of V_{j-1} and K_j are free now
            const int s1 = (nk + 2) % 3, s2 = (nk + 3) % 3;
            mbar_expect_tx(&bar_full[s1], 16384);
            tma_load_4d(sRing + s1 * 16384, &mk, &bar_full[s1], 0, k0 + kTile, h, b);
            mbar_expect_tx(&bar_full[s2], 16384);
            tma_load_4d(sRing + s2 * 16384, &mv, &bar_full[s2], 0, k0 + kTile, h, b);
        }
        __syncwarp();

        float s[128];
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld_32x32(lane_base + c * 32, reinterpret_cast<uint32_t *>(s) + c * 32);
        tmem_wait_ld();

        // ---- bias
        if (MODE == kBiasF32 || MODE == kBiasBF16) {
            if (row_ok) {
                const size_t row = (((size_t)b * a.H + h) * T + q) * (size_t)T + k0;
                const int valid = min(kTile, T - k0);
                if (MODE == kBiasF32) {
                    const float *bp = reinterpret_cast<const float *>(a.bias) + row;
#pragma unroll
                    for (int i = 0; i < 128; ++i)
                        if (i < valid) s[i] += __ldg(bp + i);
                } else {
                    const __nv_bfloat16 *bp = reinterpret_cast<const __nv_bfloat16 *>(a.bias) + row;
#pragma unroll
                    for (int i = 0; i < 128; ++i)
                        if (i < valid) s[i] += __bfloat162float(bp[i]);
                }
            }
        } else if (MODE == kBiasCode) {
            if (row_ok) {
                const uint8_t *cp = a.code + ((size_t)b * T + q) * (size_t)a.code_pitch + k0;
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    if (k0 + u * 16 < a.code_pitch) {
                        const uint4 w = __ldg(reinterpret_cast<const uint4 *>(cp) + u);
                        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                        for (int i = 0; i < 16; ++i) s[u * 16 + i] += s_lut[(ww[i >> 2] >> ((i & 3) * 8)) & 0xffu];
                    }
                }
            }
        }
        // ---- key masks (padding, tail past T)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const uint32_t w = s_kmask[j & 1][c];
            if (w != 0) {
#pragma unroll
                for (int i = 0; i < 32; ++i)
                    if (w & (1u << i)) s[c * 32 + i] = neg_inf();
            }
        }
        // ---- online softmax with lazy rescale
        float mx = neg_inf();
#pragma unroll
        for (int i = 0; i < 128; ++i) mx = fmaxf(mx, s[i]);
        float alpha = 1.f;
        const bool grow = mx > m_used + kRescaleThreshold;  // also true for the first finite maximum
        if (grow) {
            alpha = ex2((m_used - mx) * kLog2e);
            m_used = mx;
            l *= alpha;
        }
        if (j > 0 && __any_sync(0xffffffffu, grow)) {  // O (TMEM) is complete up to tile j-1: see bar_s above
            uint32_t o[32];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                tmem_ld_32x32(lane_base + 128 + c * 32, o);
                tmem_wait_ld();
#pragma unroll
                for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                tmem_st_32x32(lane_base + 128 + c * 32, o);
            }
            tmem_wait_st();
        }
        const float msc = (m_used == neg_inf()) ? 0.f : m_used * kLog2e;
        float sum = 0.f;
#pragma unroll
        for (int i = 0; i < 128; ++i) {
            s[i] = ex2(fmaf(s[i], kLog2e, -msc));
            sum += s[i];
        }
        l += sum;
        if (a.drop.thr16) {  // dropout on P (multihead_attention.py:196); the row sum above is of the undropped p
            const uint32_t salt = drop_salt(a.drop, (uint32_t)(b * a.H + h)), thr32 = a.drop.thr16 << 16;
            const uint32_t wq = drop_tile(salt, (uint32_t)q >> 6, 0u) ^ drop_lane((uint32_t)q & 63u);
#pragma unroll
            for (int hk = 0; hk < 2; ++hk) {   // k0 is a multiple of 128: two hash tiles of 64 keys
                const uint32_t w = wq ^ drop_tile(salt, (uint32_t)(k0 >> 6) + hk, 1u);
#pragma unroll
                for (int i = 0; i < 64; ++i)
                    s[hk * 64 + i] = drop_keep_w(w, drop_lane(64u + i), thr32) ? s[hk * 64 + i] * a.drop.scale : 0.f;
            }
        }
        // ---- P (bf16) -> swizzled K-major A tile: two 64-key halves of 16 KB
#pragma unroll
        for (int ch = 0; ch < 16; ++ch) {
            uint4 v;
            v.x = pack_bf16x2(s[ch * 8 + 0], s[ch * 8 + 1]);
            v.y = pack_bf16x2(s[ch * 8 + 2], s[ch * 8 + 3]);
            v.z = pack_bf16x2(s[ch * 8 + 4], s[ch * 8 + 5]);
            v.w = pack_bf16x2(s[ch * 8 + 6], s[ch * 8 + 7]);
            *reinterpret_cast<uint4 *>(sP + (ch >> 3) * 16384 + sw128_offset(tid, ch & 7)) = v;
        }
        if (j + 1 < nkv) {  // key mask of the next tile
            const uint32_t bal = __ballot_sync(0xffffffffu, key_masked(k0 + kTile + tid));
            if ((tid & 31) == 0) s_kmask[(j + 1) & 1][warp] = bal;
        }
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            mbar_wait(&bar_full[vslot], (nv / 3) & 1);
            tc_fence_after();
            const uint32_t ap = smem_u32(sP), bv = smem_u32(sRing + vslot * 16384);
#pragma unroll
            for (int k = 0; k < 8; ++k)
                umma_bf16(tmem + 128, smem_desc_sw128(ap + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                          smem_desc_sw128(bv + k * 2048, 8192, 1024), idesc_o, (j > 0 || k > 0) ? 1u : 0u);
            if (j + 1 == nkv) umma_commit(&bar_pv);
        }
        __syncwarp();
    }

    // ---- epilogue: O / l -> bf16, log-sum-exp
    mbar_wait(&bar_pv, 0);
    tc_fence_after();
    __syncwarp();
    {
        float o[64];
        tmem_ld_32x32(lane_base + 128, reinterpret_cast<uint32_t *>(o));
        tmem_ld_32x32(lane_base + 160, reinterpret_cast<uint32_t *>(o) + 32);
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
            if (a.lse) a.lse[((size_t)b * a.H + h) * T + q] = l > 0.f ? m_used + logf(l) : __int_as_float(0x7f800000);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}

template <int MODE>
static int launch_fwd(const CUtensorMap &mq, const CUtensorMap &mk, const CUtensorMap &mv, const FwdArgs &a, cudaStream_t st) {
    MDT_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem));  // per device
    dim3 grid((a.T + kTile - 1) / kTile, a.H, a.B);
    attn_fwd_kernel<MODE><<<grid, 128, kFwdSmem, st>>>(mq, mk, mv, a);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

}  // namespace attn
}  // namespace mdt

using namespace mdt;
using namespace mdt::attn;

// diagnostics only: per-CTA globaltimer stamps of the persistent forward (64 x u64 per CTA, device memory), or null
static unsigned long long *g_attn_trace = nullptr;
extern "C" int mdt_debug_attn_trace(unsigned long long *dev_buffer) {
    g_attn_trace = dev_buffer;
    return MDT_OK;
}

extern "C" int mdt_biased_attn_fwd(const void *q, const void *k, const void *v, const mdt_attn_dims *d, const void *bias,
                                   int bias_dtype, const uint8_t *code, int code_pitch, const float *lut,
                                   const uint8_t *key_pad, void *out, float *lse, float dropout_p, uint64_t dropout_seed,
                                   mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    if (int rc = check_dims(d, "mdt_biased_attn_fwd")) return rc;
    MDT_REQUIRE(q && k && v && out, "mdt_biased_attn_fwd: q, k, v, out must be non-null");
    MDT_REQUIRE(((uintptr_t)out & 15) == 0 && d->o_stride_b % 8 == 0 && d->o_stride_t % 8 == 0 && d->o_stride_h % 8 == 0,
                "mdt_biased_attn_fwd: out must be 16-byte aligned with strides that are multiples of 8");
    MDT_REQUIRE(!(bias && code), "mdt_biased_attn_fwd: pass a materialised bias or a code/lut pair, not both");
    MDT_REQUIRE(bias == nullptr || bias_dtype == MDT_F32 || bias_dtype == MDT_BF16, "mdt_biased_attn_fwd: bad bias dtype");
    MDT_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "mdt_biased_attn_fwd: dropout_p must be in [0, 1)");
    if (code) {
        MDT_REQUIRE(lut != nullptr, "mdt_biased_attn_fwd: code needs lut");
        MDT_REQUIRE(code_pitch >= d->T && code_pitch % 16 == 0 && ((uintptr_t)code & 15) == 0,
                    "mdt_biased_attn_fwd: code_pitch must be >= T and a multiple of 16, code 16-byte aligned");
    }
    CUtensorMap mq, mk, mv;
    if (int rc = make_bthd_map(&mq, q, d->B, d->H, d->T, d->q_stride_b, d->q_stride_t, d->q_stride_h, kTile)) return rc;
    if (int rc = make_bthd_map(&mk, k, d->B, d->H, d->T, d->k_stride_b, d->k_stride_t, d->k_stride_h, kTile)) return rc;
    if (int rc = make_bthd_map(&mv, v, d->B, d->H, d->T, d->v_stride_b, d->v_stride_t, d->v_stride_h, kTile)) return rc;
    FwdArgs a;
    a.B = d->B, a.H = d->H, a.T = d->T;
    a.bias = bias, a.code = code, a.code_pitch = code_pitch, a.lut = lut, a.key_pad = key_pad;
    a.out = reinterpret_cast<__nv_bfloat16 *>(out);
    a.o_sb = d->o_stride_b, a.o_st = d->o_stride_t, a.o_sh = d->o_stride_h;
    a.lse = lse;
    a.drop = make_drop(dropout_p, dropout_seed);
    cudaStream_t st = as_stream(stream);
    const int mode = bias_mode(bias, bias_dtype, code);
    if ((mode == kBiasNone || mode == kBiasCode) && d->T > kTile && d->T <= kFwd3MaxT) {  // persistent warp-specialised kernel (attn_fwd3.cu)
        CUtensorMap mc = mq;
        if (code)
            if (int rc = make_tensor_map_u8_2d(&mc, code, (uint64_t)code_pitch, (uint64_t)d->B * d->T, (uint64_t)code_pitch, kTile))
                return rc;
        Fwd2Args a2;
        a2.B = d->B, a2.H = d->H, a2.T = d->T, a2.has_code = code != nullptr, a2.lut = lut, a2.key_pad = key_pad;
        a2.out = a.out, a2.o_sb = a.o_sb, a2.o_st = a.o_st, a2.o_sh = a.o_sh, a2.lse = lse;
        a2.trace = g_attn_trace;
        a2.drop = a.drop;
        a2.Tmain = d->T;
        return launch_fwd3(mq, mk, mv, mc, a2, st);
    }
    switch (mode) {
        case kBiasNone: return launch_fwd<kBiasNone>(mq, mk, mv, a, st);
        case kBiasF32: return launch_fwd<kBiasF32>(mq, mk, mv, a, st);
        case kBiasBF16: return launch_fwd<kBiasBF16>(mq, mk, mv, a, st);
        default: return launch_fwd<kBiasCode>(mq, mk, mv, a, st);
    }
}

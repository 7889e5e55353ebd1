// Shared pieces of the biased-attention kernels (A6, multihead_attention.py:168-199):
// argument validation, the 4-D TMA maps over strided [b, t, h, 64] bf16 tensors and the
// bias-source modes.
#pragma once
#include "umma.cuh"

namespace mdt {
namespace attn {

constexpr int kD = 64;       // head dim (embed 768 / 12 heads)
constexpr int kTile = 128;   // queries per CTA tile == keys per tile == UMMA M
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// where the additive bias of S comes from
enum BiasMode : int {
    kBiasNone = 0,
    kBiasF32 = 1,   // materialised [B,H,T,T] float32
    kBiasBF16 = 2,  // materialised [B,H,T,T] bfloat16
    kBiasCode = 3,  // one byte per (b,q,k) + lut[H,256]: the [B,H,T,T] tensor never exists
};

inline int bias_mode(const void *bias, int bias_dtype, const uint8_t *code) {
    if (code) return kBiasCode;
    if (!bias) return kBiasNone;
    return bias_dtype == MDT_F32 ? kBiasF32 : kBiasBF16;
}

// ---- attention dropout (multihead_attention.py:196): the keep mask is a counter hash, never stored.
//   salt(b,h)  = fmix32(seed_lo ^ (b*H+h) * 0x9e3779b1) + seed_hi
//   tile(t, w) = fmix32(salt ^ (2 t + w) * 0x9e3779b1)            one strong hash per 64 queries (w = 0) / 64 keys (w = 1)
//   lane(c)    = fmix32((c + 1) * 0x9e3779b1)                      a constant per position inside the 64 (c + 64 for keys)
//   x(q, k)    = (tile(q >> 6, 0) ^ lane(q & 63) ^ tile(k >> 6, 1) ^ lane(64 + (k & 63))) * 0x9e3779b1
//   keep <=> x >= thr16 << 16                                      (drop probability thr16 / 65536)
// The strong hashes cost ~10 integer operations per 64 elements; inside an unrolled loop lane() folds to an immediate, so
// an element costs xor + multiply + compare + select whether a thread walks along the keys of one query (forward) or
// along the queries of one key (backward): both regenerate the same mask from (seed, b, h, q, k).  thr16 == 0 switches
// dropout off.  The packed-sequence kernels (attn_varlen.cu) use drop_chunk below: there both passes walk along keys.
struct DropArgs {
    uint32_t thr16;   // round(p * 65536)
    float scale;      // 1 / (1 - p)
    uint32_t seed_lo, seed_hi;
};
inline DropArgs make_drop(float p, uint64_t seed) {
    DropArgs d;
    const float t = p * 65536.f + 0.5f;
    d.thr16 = p <= 0.f ? 0u : (t >= 65535.f ? 65535u : (uint32_t)t);
    d.scale = 1.f / (1.f - p);
    d.seed_lo = (uint32_t)seed, d.seed_hi = (uint32_t)(seed >> 32);
    return d;
}
__host__ __device__ constexpr uint32_t drop_fmix(uint32_t h) {
    h ^= h >> 16;
    h *= 0x85ebca6bu;
    h ^= h >> 13;
    h *= 0xc2b2ae35u;
    h ^= h >> 16;
    return h;
}
__device__ __forceinline__ uint32_t drop_salt(const DropArgs &d, uint32_t bh) {
    return drop_fmix(d.seed_lo ^ (bh * 0x9e3779b1u)) + d.seed_hi;
}
constexpr uint32_t kDropMul = 0x9e3779b1u;
// which = 0: a tile of 64 queries, 1: a tile of 64 keys
__device__ __forceinline__ uint32_t drop_tile(uint32_t salt, uint32_t tile, uint32_t which) {
    return drop_fmix(salt ^ ((2u * tile + which) * kDropMul));
}
// c = q & 63 for a query, 64 + (k & 63) for a key; constant-folds when c is a compile-time value
__host__ __device__ constexpr uint32_t drop_lane(uint32_t c) { return drop_fmix((c + 1u) * kDropMul); }
// w = the xor of the two tile hashes and of the lane constant of the coordinate the thread owns
__device__ __forceinline__ bool drop_keep_w(uint32_t w, uint32_t lane_const, uint32_t thr32) {
    return (w ^ lane_const) * kDropMul >= thr32;
}
__device__ __forceinline__ bool drop_keep_one(const DropArgs &d, uint32_t salt, uint32_t q, uint32_t k) {
    const uint32_t w = drop_tile(salt, q >> 6, 0u) ^ drop_lane(q & 63u) ^ drop_tile(salt, k >> 6, 1u);
    return drop_keep_w(w, drop_lane(64u + (k & 63u)), d.thr16 << 16);
}
// packed-sequence kernels (attn_varlen.cu; salt per (head, unit of 128 packed rows), both passes walk along the keys of
// a row): one strong hash per (row r, 32-key chunk c), then  x = (chunk ^ lane(i)) * 0x9e3779b1  for key 32 c + i
__device__ __forceinline__ uint32_t drop_chunk(uint32_t salt, uint32_t r, uint32_t c) {
    return drop_fmix(salt ^ ((r * 4u + c) * kDropMul));
}

// TMA view of a strided bf16 tensor addressed (b, t, h, c): dims innermost-first {64, T, H, B},
// box {64, rows, 1, 1}, 128-byte swizzle.  Rows past T read as zeros.
inline int make_bthd_map(CUtensorMap *map, const void *base, int B, int H, int T, int64_t sb, int64_t st, int64_t sh,
                         int box_rows) {
    MDT_REQUIRE(sb % 8 == 0 && st % 8 == 0 && sh % 8 == 0, "q/k/v/o strides must be multiples of 8 elements");
    const uint64_t dims[4] = {(uint64_t)kD, (uint64_t)T, (uint64_t)H, (uint64_t)B};
    // a dimension of extent 1 still needs a legal (multiple of 16, non-zero) stride
    const uint64_t strides[4] = {2, (uint64_t)(st > 0 ? st : kD) * 2, (uint64_t)(sh > 0 ? sh : kD) * 2,
                                 (uint64_t)(sb > 0 ? sb : kD) * 2};
    const uint32_t box[4] = {(uint32_t)kD, (uint32_t)box_rows, 1, 1};
    return make_tensor_map_bf16(map, base, 4, dims, strides, box);
}

inline int check_dims(const mdt_attn_dims *d, const char *who) {
    MDT_REQUIRE(d != nullptr, "%s: dims is null", who);
    MDT_REQUIRE(d->B >= 1 && d->H >= 1 && d->T >= 1, "%s: B, H, T must be >= 1 (got %d, %d, %d)", who, d->B, d->H, d->T);
    MDT_REQUIRE(d->T <= 65535 && d->H <= 65535 && d->B <= 65535, "%s: B, H, T must fit a CUDA grid dimension", who);
    return MDT_OK;
}

// persistent warp-specialised forward (attn_fwd3.cu): bias from byte codes or none
struct Fwd2Args {
    int B, H, T;
    int has_code;
    const float *lut;
    const uint8_t *key_pad;
    __nv_bfloat16 *out;
    int64_t o_sb, o_st, o_sh;
    float *lse;
    DropArgs drop;
    int Tmain;                  // queries / keys processed: [0, Tmain) (= T; kept separate from the row pitch T)
    unsigned long long *trace;  // diagnostics: per-CTA globaltimer stamps (mdt_debug_attn_trace), else null
};
constexpr int kFwd3MaxT = 4096;  // the key-mask bitmap of a whole row lives in shared memory (attn_fwd3.cu)
int launch_fwd3(const CUtensorMap &mq, const CUtensorMap &mk, const CUtensorMap &mv, const CUtensorMap &mc, const Fwd2Args &a,
                cudaStream_t st);


// warp-specialised backward for bias from codes / no bias (attn_bwd2.cu)
struct Bwd2Args {
    int B, H, T, Tmain;
    const uint8_t *key_pad;
    const float *lut;
    const float *lse;
    const float *delta;  // precomputed rowsum(dO * O), or null: computed here from out / dout
    const __nv_bfloat16 *out, *dout;
    int64_t o_sb, o_st, o_sh, g_sb, g_st, g_sh;
    float *dq_accum;
    __nv_bfloat16 *dk, *dv;
    int64_t dk_sb, dk_st, dk_sh, dv_sb, dv_st, dv_sh;
    float *grad_lut;
    DropArgs drop;
};

int launch_bwd2(const CUtensorMap &mq, const CUtensorMap &mk, const CUtensorMap &mv, const CUtensorMap &mdo, const CUtensorMap &mc,
                const Bwd2Args &a, bool has_code, cudaStream_t st);

// tail keys of T = 128 k + r in the backward (attn_tail.cu): keys [Tmain, T) on CUDA cores
constexpr int kTailMax = 8;
inline int tail_split(int T, bool eligible) {  // -> Tmain
    const int r = T % kTile;
    return (eligible && T > kTile && r >= 1 && r <= kTailMax) ? T - r : T;
}
struct TailArgs {
    int B, H, T, Tmain;
    const __nv_bfloat16 *q, *k, *v;
    int64_t q_sb, q_st, q_sh, k_sb, k_st, k_sh, v_sb, v_st, v_sh;
    const uint8_t *code;
    int code_pitch;
    const float *lut;
    const uint8_t *key_pad;
    const float *lse;
    const __nv_bfloat16 *dout;
    int64_t g_sb, g_st, g_sh;
    const float *delta;
    float *dq_accum;
    __nv_bfloat16 *dk, *dv;
    int64_t dk_sb, dk_st, dk_sh, dv_sb, dv_st, dv_sh;
    float *grad_lut;
    DropArgs drop;
};
int launch_bwd_tails(const TailArgs &a, cudaStream_t st);

}  // namespace attn
}  // namespace mdt

// Fused  y = LayerNorm(dropout(x) + residual)  forward / backward over token rows (bf16 in, fp32 math).
//
// The post-LN blocks of the path (Hugging Face BertLayer inside the fusion layers, multi_graphormer_fusion_layer.py:105-149;
// GraphormerGraphEncoderLayer, graphormer_graph_encoder_layer.py:115-141) end in dense -> dropout -> + residual -> LayerNorm.
// As separate ATen kernels that is 5 passes forward and 6 backward over [tokens, C], and the gamma/beta gradient
// (a column reduction over ~2.4e5 rows) alone took 15 % of a training step.  Here: one pass each way.
//   forward : z = keep(x) / (1 - p) + residual;  y = (z - mean) * rstd * gamma + beta;   saves z (bf16), mean, rstd
//   backward: dz = rstd * (g - mean(g) - xhat * mean(g * xhat)),  g = dy * gamma;  d_residual = dz;
//             d_x = keep * dz / (1 - p);  d_gamma += dy * xhat,  d_beta += dy  (per-lane fp32 partials over a
//             grid-stride row loop, one shared-memory reduction per CTA, then fp32 atomics: 2 * C per CTA)
// HBM-bound: one warp per row, 16-byte vector accesses, lanes interleaved so every access is 512 contiguous bytes.
// The dropout mask is never stored: it is a counter-based hash of (seed, element index), regenerated in the backward.
#include "common.cuh"

namespace mdt {
namespace ln {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

__device__ __forceinline__ uint32_t fmix32(uint32_t h) {
    h ^= h >> 16;
    h *= 0x85ebca6bu;
    h ^= h >> 13;
    h *= 0xc2b2ae35u;
    h ^= h >> 16;
    return h;
}
// 8 keep flags (bit i = element i of the 8-element vector `vec`) for drop probability thr16 / 65536
__device__ __forceinline__ uint32_t keep8(uint64_t seed, uint64_t vec, uint32_t thr16) {
    const uint32_t lo = (uint32_t)seed, hi = (uint32_t)(seed >> 32);
    const uint32_t base = (uint32_t)vec * 4u, up = (uint32_t)(vec >> 30) * 0x9e3779b1u + hi;
    uint32_t bits = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t r = fmix32(fmix32((base + k) ^ lo) + up);
        bits |= ((r & 0xffffu) >= thr16 ? 1u : 0u) << (2 * k);
        bits |= ((r >> 16) >= thr16 ? 1u : 0u) << (2 * k + 1);
    }
    return bits;
}

__device__ __forceinline__ void unpack8(const uint4 &v, float *f) {
    const __nv_bfloat162 *p = reinterpret_cast<const __nv_bfloat162 *>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 t = __bfloat1622float2(p[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}
__device__ __forceinline__ uint4 pack8(const float *f) {
    uint4 v;
    __nv_bfloat162 *p = reinterpret_cast<__nv_bfloat162 *>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) p[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return v;
}
// streaming 16-byte load (read once: keep it out of L1)
__device__ __forceinline__ uint4 ldg_stream(const uint4 *p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

struct FwdArgs {
    const __nv_bfloat16 *x, *res, *gamma, *beta;
    __nv_bfloat16 *y, *z;
    float *mean, *rstd;
    int64_t rows;
    float eps, scale;
    uint32_t thr16;
    uint64_t seed;
};

// VEC = C / 256: 16-byte vectors per lane
template <int VEC>
__global__ void __launch_bounds__(kThreads) ln_fwd_kernel(const FwdArgs a) {
    constexpr int C = VEC * 256;
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (int64_t)blockIdx.x * kWarps + (threadIdx.x >> 5), nwarps = (int64_t)gridDim.x * kWarps;
    float g[VEC][8], bta[VEC][8];
#pragma unroll
    for (int v = 0; v < VEC; ++v) {
        unpack8(reinterpret_cast<const uint4 *>(a.gamma)[v * 32 + lane], g[v]);
        unpack8(reinterpret_cast<const uint4 *>(a.beta)[v * 32 + lane], bta[v]);
    }
    // the raw vectors of the next row are requested before this row is reduced: two rows of loads in flight per warp
    uint4 nx[VEC], nr[VEC];
    auto fetch = [&](int64_t row) {
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
            nx[v] = ldg_stream(reinterpret_cast<const uint4 *>(a.x + row * C) + v * 32 + lane);
            if (a.res) nr[v] = ldg_stream(reinterpret_cast<const uint4 *>(a.res + row * C) + v * 32 + lane);
        }
    };
    if (warp0 < a.rows) fetch(warp0);
    for (int64_t row = warp0; row < a.rows; row += nwarps) {
        uint4 cx[VEC], cr[VEC];
#pragma unroll
        for (int v = 0; v < VEC; ++v) cx[v] = nx[v], cr[v] = nr[v];
        if (row + nwarps < a.rows) fetch(row + nwarps);
        float z[VEC][8];
        float s = 0.f;
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
            unpack8(cx[v], z[v]);
            if (a.thr16) {
                const uint32_t keep = keep8(a.seed, (uint64_t)row * (C / 8) + v * 32 + lane, a.thr16);
#pragma unroll
                for (int i = 0; i < 8; ++i) z[v][i] = (keep >> i) & 1u ? z[v][i] * a.scale : 0.f;
            }
            if (a.res) {
                float r[8];
                unpack8(cr[v], r);
#pragma unroll
                for (int i = 0; i < 8; ++i) z[v][i] += r[i];
            }
            if (a.z) {  // the saved pre-norm sum is what the backward normalises again: round it first
                const uint4 zv = pack8(z[v]);
                reinterpret_cast<uint4 *>(a.z + row * C)[v * 32 + lane] = zv;
                unpack8(zv, z[v]);
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) s += z[v][i];
        }
        const float mean = warp_sum(s) * (1.f / C);
        float q = 0.f;
#pragma unroll
        for (int v = 0; v < VEC; ++v)
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float d = z[v][i] - mean;
                q = fmaf(d, d, q);
            }
        const float rstd = rsqrtf(warp_sum(q) * (1.f / C) + a.eps);
        if (lane == 0) {
            a.mean[row] = mean;
            a.rstd[row] = rstd;
        }
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
            float o[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) o[i] = fmaf((z[v][i] - mean) * rstd, g[v][i], bta[v][i]);
            reinterpret_cast<uint4 *>(a.y + row * C)[v * 32 + lane] = pack8(o);
        }
    }
}

struct BwdArgs {
    const __nv_bfloat16 *dy, *z, *gamma;
    const float *mean, *rstd;
    __nv_bfloat16 *dx, *dres;
    float *dgamma, *dbeta;
    int64_t rows;
    float scale;
    uint32_t thr16;
    uint64_t seed;
};

template <int VEC>
__global__ void __launch_bounds__(kThreads, VEC <= 3 ? 2 : 1) ln_bwd_kernel(const BwdArgs a) {
    constexpr int C = VEC * 256;
    __shared__ float red[kWarps][C];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t warp0 = (int64_t)blockIdx.x * kWarps + warp, nwarps = (int64_t)gridDim.x * kWarps;
    // Register budget (two CTAs per SM => 128 registers): the column partials dg / db stay in fp32 registers, the row itself is
    // kept as the packed bf16 vectors it arrived in and unpacked twice (before and after the row reduction), gamma is re-read
    // from L1 per row, and the next row's vectors are requested before this row is reduced.
    float dg[VEC][8], db[VEC][8];
#pragma unroll
    for (int v = 0; v < VEC; ++v)
#pragma unroll
        for (int i = 0; i < 8; ++i) dg[v][i] = db[v][i] = 0.f;
    uint4 ny[VEC], nz[VEC];
    float nmean = 0.f, nrstd = 0.f;
    auto fetch = [&](int64_t row) {
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
            ny[v] = ldg_stream(reinterpret_cast<const uint4 *>(a.dy + row * C) + v * 32 + lane);
            nz[v] = ldg_stream(reinterpret_cast<const uint4 *>(a.z + row * C) + v * 32 + lane);
        }
        nmean = a.mean[row], nrstd = a.rstd[row];
    };
    if (warp0 < a.rows) fetch(warp0);
    for (int64_t row = warp0; row < a.rows; row += nwarps) {
        uint4 cy[VEC], cz[VEC];
#pragma unroll
        for (int v = 0; v < VEC; ++v) cy[v] = ny[v], cz[v] = nz[v];
        const float mean = nmean, rstd = nrstd;
        if (row + nwarps < a.rows) fetch(row + nwarps);
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
            float dy[8], xh[8], g[8];
            unpack8(cy[v], dy);
            unpack8(cz[v], xh);
            unpack8(__ldg(reinterpret_cast<const uint4 *>(a.gamma) + v * 32 + lane), g);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                xh[i] = (xh[i] - mean) * rstd;
                dg[v][i] = fmaf(dy[i], xh[i], dg[v][i]);
                db[v][i] += dy[i];
                const float gy = dy[i] * g[i];
                s1 += gy;
                s2 = fmaf(gy, xh[i], s2);
            }
        }
        const float c1 = warp_sum(s1) * (1.f / C), c2 = warp_sum(s2) * (1.f / C);
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
            float dy[8], xh[8], g[8], dz[8];
            unpack8(cy[v], dy);
            unpack8(cz[v], xh);
            unpack8(__ldg(reinterpret_cast<const uint4 *>(a.gamma) + v * 32 + lane), g);
#pragma unroll
            for (int i = 0; i < 8; ++i) dz[i] = (dy[i] * g[i] - c1 - (xh[i] - mean) * rstd * c2) * rstd;
            if (a.dres) reinterpret_cast<uint4 *>(a.dres + row * C)[v * 32 + lane] = pack8(dz);
            if (a.dx) {
                if (a.thr16) {
                    const uint32_t keep = keep8(a.seed, (uint64_t)row * (C / 8) + v * 32 + lane, a.thr16);
#pragma unroll
                    for (int i = 0; i < 8; ++i) dz[i] = (keep >> i) & 1u ? dz[i] * a.scale : 0.f;
                }
                reinterpret_cast<uint4 *>(a.dx + row * C)[v * 32 + lane] = pack8(dz);
            }
        }
    }
    // ---- per-CTA column sums of the lane partials, then one fp32 atomic per column and CTA
#pragma unroll
    for (int which = 0; which < 2; ++which) {
        __syncthreads();
#pragma unroll
        for (int v = 0; v < VEC; ++v)
#pragma unroll
            for (int i = 0; i < 8; ++i) red[warp][(v * 32 + lane) * 8 + i] = which == 0 ? dg[v][i] : db[v][i];
        __syncthreads();
        float *dst = which == 0 ? a.dgamma : a.dbeta;
        for (int c = threadIdx.x; c < C; c += kThreads) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) s += red[w][c];
            atomicAdd(dst + c, s);
        }
    }
}

inline int grid_for(int64_t rows) {
    const int64_t want = (rows + kWarps - 1) / kWarps;
    const int64_t cap = (int64_t)num_sms() * 4;
    return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace ln
}  // namespace mdt

using namespace mdt;

static int check_c(int C, const char *who) {
    MDT_REQUIRE(C == 256 || C == 512 || C == 768 || C == 1024, "%s: C must be 256, 512, 768 or 1024 (got %d)", who, C);
    return MDT_OK;
}
static uint32_t thr16_of(float p) {
    if (p <= 0.f) return 0;
    const float t = p * 65536.f + 0.5f;
    return t >= 65535.f ? 65535u : (uint32_t)t;
}

extern "C" int mdt_dropout_add_ln_fwd(const void *x, const void *residual, const void *gamma, const void *beta, void *y, void *z,
                                      float *mean, float *rstd, int64_t rows, int C, float p, uint64_t seed, float eps,
                                      mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    if (int rc = check_c(C, "mdt_dropout_add_ln_fwd")) return rc;
    MDT_REQUIRE(x && gamma && beta && y && mean && rstd, "mdt_dropout_add_ln_fwd: x, gamma, beta, y, mean, rstd must be non-null");
    MDT_REQUIRE(p >= 0.f && p < 1.f, "mdt_dropout_add_ln_fwd: p must be in [0, 1)");
    MDT_REQUIRE(z != nullptr || (residual == nullptr && p == 0.f), "mdt_dropout_add_ln_fwd: z is required with a residual or dropout");
    if (rows <= 0) return MDT_OK;
    ln::FwdArgs a;
    a.x = (const __nv_bfloat16 *)x, a.res = (const __nv_bfloat16 *)residual, a.gamma = (const __nv_bfloat16 *)gamma;
    a.beta = (const __nv_bfloat16 *)beta, a.y = (__nv_bfloat16 *)y, a.z = (__nv_bfloat16 *)z, a.mean = mean, a.rstd = rstd;
    a.rows = rows, a.eps = eps, a.thr16 = thr16_of(p), a.scale = 1.f / (1.f - p), a.seed = seed;
    cudaStream_t st = as_stream(stream);
    const int grid = ln::grid_for(rows);
    switch (C / 256) {
        case 1: ln::ln_fwd_kernel<1><<<grid, ln::kThreads, 0, st>>>(a); break;
        case 2: ln::ln_fwd_kernel<2><<<grid, ln::kThreads, 0, st>>>(a); break;
        case 3: ln::ln_fwd_kernel<3><<<grid, ln::kThreads, 0, st>>>(a); break;
        default: ln::ln_fwd_kernel<4><<<grid, ln::kThreads, 0, st>>>(a); break;
    }
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

extern "C" int mdt_dropout_add_ln_bwd(const void *dy, const void *z, const float *mean, const float *rstd, const void *gamma,
                                      void *dx, void *dres, float *dgamma, float *dbeta, int64_t rows, int C, float p,
                                      uint64_t seed, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    if (int rc = check_c(C, "mdt_dropout_add_ln_bwd")) return rc;
    MDT_REQUIRE(dy && z && mean && rstd && gamma && dgamma && dbeta, "mdt_dropout_add_ln_bwd: null argument");
    MDT_REQUIRE(dx || dres, "mdt_dropout_add_ln_bwd: dx or dres must be given");
    MDT_REQUIRE(p >= 0.f && p < 1.f, "mdt_dropout_add_ln_bwd: p must be in [0, 1)");
    if (rows <= 0) return MDT_OK;
    ln::BwdArgs a;
    a.dy = (const __nv_bfloat16 *)dy, a.z = (const __nv_bfloat16 *)z, a.gamma = (const __nv_bfloat16 *)gamma, a.mean = mean, a.rstd = rstd;
    a.dx = (__nv_bfloat16 *)dx, a.dres = (__nv_bfloat16 *)dres, a.dgamma = dgamma, a.dbeta = dbeta;
    a.rows = rows, a.thr16 = thr16_of(p), a.scale = 1.f / (1.f - p), a.seed = seed;
    cudaStream_t st = as_stream(stream);
    const int grid = ln::grid_for(rows);
    switch (C / 256) {
        case 1: ln::ln_bwd_kernel<1><<<grid, ln::kThreads, 0, st>>>(a); break;
        case 2: ln::ln_bwd_kernel<2><<<grid, ln::kThreads, 0, st>>>(a); break;
        case 3: ln::ln_bwd_kernel<3><<<grid, ln::kThreads, 0, st>>>(a); break;
        default: ln::ln_bwd_kernel<4><<<grid, ln::kThreads, 0, st>>>(a); break;
    }
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

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
// HBM-bound, 16-byte vector accesses, a warp touches 512 contiguous bytes.  Forward: one warp per row, the next row's vectors
// requested before this row is reduced (measured 5.9 TB/s = 90 % of the copy peak at 240 k x 768).  Backward: C / 256 warps per
// row with cp.async staging (see ln_bwd_kernel; 5.0 TB/s = 77 %).
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
// Dropout decisions of the 8-element vector `vec`: one fmix32 of (seed, vec) per vector, then one multiply-xorshift round per
// element pair; element 2k keeps if the low half of r[k] >= thr16, element 2k+1 if the high half does (drop probability
// thr16 / 65536).  Forward and backward call the same function, nothing is stored.
struct Rand4 {
    uint32_t r[4];
};
__device__ __forceinline__ Rand4 rand4(uint64_t seed, uint64_t vec) {
    const uint32_t lo = (uint32_t)seed, hi = (uint32_t)(seed >> 32);
    const uint32_t h = fmix32(((uint32_t)vec ^ lo) * 0x9e3779b1u + (uint32_t)(vec >> 32) * 0x85ebca77u + hi);
    Rand4 o;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        uint32_t r = (h ^ (0x632be5abu * (uint32_t)(k + 1))) * 0x2c1b3c6du;
        o.r[k] = r ^ (r >> 15);
    }
    return o;
}
// z[i] <- keep ? z[i] * scale : 0
__device__ __forceinline__ void apply_keep(float *z, const Rand4 &rn, uint32_t thr16, float scale) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        z[2 * k] = (rn.r[k] & 0xffffu) >= thr16 ? z[2 * k] * scale : 0.f;
        z[2 * k + 1] = (rn.r[k] >> 16) >= thr16 ? z[2 * k + 1] * scale : 0.f;
    }
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
                apply_keep(z[v], rand4(a.seed, (uint64_t)row * (C / 8) + v * 32 + lane), a.thr16, a.scale);
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
    const __nv_bfloat16 *dadd;   // optional [rows, C]: added to dz before it is written / masked (a residual gradient that joins here)
    const float *mean, *rstd;
    __nv_bfloat16 *dx, *dres;
    float *dgamma, *dbeta;
    float *dxsum;   // column sums of the (rounded) tensor written to dx, or to dres without dx: the bias gradient of the dense layer in front
    int64_t rows;
    float scale;
    uint32_t thr16;
    uint64_t seed;
};

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Backward.  The row (dy and z) is needed twice -- once for the row sums, once for dz -- and the column partials dg / db pin 16
// fp32 registers per 16-byte vector.  With a warp per row (3 vectors per lane at C = 768) the partials, the row and a
// prefetched row did not fit 128 registers: the spilt version ran at 61 % of the HBM peak against 89 % for the forward, its
// stalls on the local-memory reloads.  So here VEC warps share a row, one vector per lane: 16 partials per lane, no spills.
//   * a lane stages its own dy / z vectors in shared memory with 16-byte cp.async copies, kBwdStages rows deep; it later reads
//     only the slots it copied itself, so the staging needs no barrier
//   * the row sums (sum gy, sum gy*xhat) are warp sums + one named barrier over the VEC warps of the row group
//     (partials double-buffered in shared memory by row parity)
constexpr int kBwdThreads = 384;  // 12 warps: 12 / VEC row groups
constexpr int kBwdStages = 4;
constexpr int kBwdSmem = kBwdStages * 2 * kBwdThreads * 16;  // 48 KB of dynamic shared memory (6 stages measured no faster)

template <int VEC, bool XSUM>
__global__ void __launch_bounds__(kBwdThreads, 2) ln_bwd_kernel(const BwdArgs a) {
    constexpr int C = VEC * 256;
    constexpr int kW = kBwdThreads / 32, kGroups = kW / VEC;
    extern __shared__ uint4 ln_smem[];
    uint4(*stage_mem)[2][kBwdThreads] = reinterpret_cast<uint4(*)[2][kBwdThreads]>(ln_smem);  // [stage][dy | z][thread]
    __shared__ float2 s_part[2][kGroups][VEC];
    static_assert(kBwdSmem >= sizeof(float) * kGroups * C, "the final reduction reuses the staging memory");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int group = warp / VEC, sub = warp % VEC;  // this lane owns columns [sub * 256 + lane * 8, + 8)
    const int64_t row0 = (int64_t)blockIdx.x * kGroups + group, nrows = (int64_t)gridDim.x * kGroups;
    const int col = sub * 32 + lane;                  // vector index inside the row
    float g[8], dg[8], db[8], xs[XSUM ? 8 : 1];
    unpack8(__ldg(reinterpret_cast<const uint4 *>(a.gamma) + col), g);
#pragma unroll
    for (int i = 0; i < 8; ++i) dg[i] = db[i] = 0.f;
    if (XSUM) {
#pragma unroll
        for (int i = 0; i < 8; ++i) xs[i] = 0.f;
    }
    auto issue = [&](int64_t row, int st) {
        if (row < a.rows) {
            cp_async16(&stage_mem[st][0][threadIdx.x], reinterpret_cast<const uint4 *>(a.dy + row * C) + col);
            cp_async16(&stage_mem[st][1][threadIdx.x], reinterpret_cast<const uint4 *>(a.z + row * C) + col);
        }
        cp_async_commit();  // always: the group count per iteration stays uniform
    };
#pragma unroll
    for (int s = 0; s < kBwdStages - 1; ++s) issue(row0 + s * nrows, s);
    int st = 0, par = 0;
    for (int64_t row = row0; row < a.rows; row += nrows, st = st + 1 == kBwdStages ? 0 : st + 1, par ^= 1) {
        const float mean = a.mean[row], rstd = a.rstd[row];
        uint4 addv = make_uint4(0, 0, 0, 0);
        if (a.dadd) addv = ldg_stream(reinterpret_cast<const uint4 *>(a.dadd + row * C) + col);   // in flight under the row sums
        issue(row + (kBwdStages - 1) * nrows, st == 0 ? kBwdStages - 1 : st - 1);
        cp_async_wait<kBwdStages - 1>();  // this row's copies have landed
        float dy[8], z[8], xh[8];
        unpack8(stage_mem[st][0][threadIdx.x], dy);
        unpack8(stage_mem[st][1][threadIdx.x], z);
        float s1 = 0.f, s2 = 0.f;
        const float nmr = -mean * rstd;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            xh[i] = fmaf(z[i], rstd, nmr);
            dg[i] = fmaf(dy[i], xh[i], dg[i]);
            db[i] += dy[i];
            dy[i] *= g[i];  // gy
            s1 += dy[i];
            s2 = fmaf(dy[i], xh[i], s2);
        }
        s1 = warp_sum(s1), s2 = warp_sum(s2);
        if (VEC > 1) {
            if (lane == 0) s_part[par][group][sub] = make_float2(s1, s2);
            asm volatile("bar.sync %0, %1;" ::"r"(1 + group), "n"(VEC * 32) : "memory");
            s1 = s2 = 0.f;
#pragma unroll
            for (int w = 0; w < VEC; ++w) {
                const float2 t = s_part[par][group][w];
                s1 += t.x, s2 += t.y;
            }
        }
        const float c1 = s1 * (1.f / C), c2 = s2 * (1.f / C);
        float dz[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) dz[i] = (dy[i] - c1 - xh[i] * c2) * rstd;
        if (a.dadd) {
            float ad[8];
            unpack8(addv, ad);
#pragma unroll
            for (int i = 0; i < 8; ++i) dz[i] += ad[i];
        }
        uint4 packed = pack8(dz);
        if (a.dres) reinterpret_cast<uint4 *>(a.dres + row * C)[col] = packed;
        if (a.dx) {
            if (a.thr16) {
                apply_keep(dz, rand4(a.seed, (uint64_t)row * (C / 8) + col), a.thr16, a.scale);
                packed = pack8(dz);
            }
            reinterpret_cast<uint4 *>(a.dx + row * C)[col] = packed;
        }
        if (XSUM) {   // sums of the rounded values: the tensor the weight-gradient GEMM reads
            float q[8];
            unpack8(packed, q);
#pragma unroll
            for (int i = 0; i < 8; ++i) xs[i] += q[i];
        }
    }
    cp_async_wait<0>();
    // ---- per-CTA column sums over the row groups (the staging buffers are dead now), then one fp32 atomic per column and CTA
    float(*red)[C] = reinterpret_cast<float(*)[C]>(ln_smem);
#pragma unroll
    for (int which = 0; which < (XSUM ? 3 : 2); ++which) {
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 8; ++i) red[group][col * 8 + i] = which == 0 ? dg[i] : which == 1 ? db[i] : xs[XSUM ? i : 0];
        __syncthreads();
        float *dst = which == 0 ? a.dgamma : which == 1 ? a.dbeta : a.dxsum;
        for (int c = threadIdx.x; c < C; c += kBwdThreads) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < kGroups; ++w) s += red[w][c];
            atomicAdd(dst + c, s);
        }
    }
}

template <int VEC>
static int launch_bwd(const BwdArgs &a, cudaStream_t st) {
    constexpr int kGroups = kBwdThreads / 32 / VEC;
    const int64_t want = (a.rows + kGroups - 1) / kGroups, cap = (int64_t)num_sms() * 2;
    const int grid = (int)(want < cap ? want : cap);
    if (a.dxsum) {
        MDT_CHECK_CUDA(cudaFuncSetAttribute(ln_bwd_kernel<VEC, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem));
        ln_bwd_kernel<VEC, true><<<grid, kBwdThreads, kBwdSmem, st>>>(a);
    } else {
        MDT_CHECK_CUDA(cudaFuncSetAttribute(ln_bwd_kernel<VEC, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem));  // per device
        ln_bwd_kernel<VEC, false><<<grid, kBwdThreads, kBwdSmem, st>>>(a);
    }
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

// ---- y = dropout_p(x) + residual without a LayerNorm behind it (the second residual of a pre-LN block: Hugging Face ViTOutput),
// and its backward dx = keep * dy / (1 - p) with the column sums of the rounded dx (= the bias gradient of the dense layer
// that produced x).  Same keep mask as the LayerNorm kernels: rand4(seed, row * C / 8 + vector).  Threads own one 16-byte
// vector column and walk down the rows, two rows in flight.
constexpr int kDaTx = 32, kDaTy = 8;

__global__ void __launch_bounds__(kDaTx *kDaTy)
    dropout_add_fwd_kernel(const uint4 *__restrict__ x, const uint4 *__restrict__ res, uint4 *__restrict__ y, int64_t rows, int vec_per_row,
                           uint32_t thr16, float scale, uint64_t seed) {
    const int cx = blockIdx.x * kDaTx + threadIdx.x;
    if (cx >= vec_per_row) return;
    const int64_t step = (int64_t)gridDim.y * kDaTy;
    for (int64_t r = (int64_t)blockIdx.y * kDaTy + threadIdx.y; r < rows; r += 2 * step) {
        const bool two = r + step < rows;
        const int64_t r1 = two ? r + step : r;
        const uint4 x0 = ldg_stream(x + r * vec_per_row + cx), x1 = ldg_stream(x + r1 * vec_per_row + cx);
        uint4 q0 = make_uint4(0, 0, 0, 0), q1 = q0;
        if (res) q0 = ldg_stream(res + r * vec_per_row + cx), q1 = ldg_stream(res + r1 * vec_per_row + cx);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 1 && !two) break;
            const int64_t rr = h == 0 ? r : r1;
            float v[8], w[8];
            unpack8(h == 0 ? x0 : x1, v);
            unpack8(h == 0 ? q0 : q1, w);
            if (thr16) apply_keep(v, rand4(seed, (uint64_t)rr * vec_per_row + cx), thr16, scale);
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] += w[i];
            y[rr * vec_per_row + cx] = pack8(v);
        }
    }
}

__global__ void __launch_bounds__(kDaTx *kDaTy)
    dropout_bwd_kernel(const uint4 *__restrict__ dy, uint4 *__restrict__ dx, float *__restrict__ dxsum, int64_t rows, int vec_per_row,
                       uint32_t thr16, float scale, uint64_t seed) {
    __shared__ float red[kDaTy][kDaTx * 8 + 1];
    const int cx = blockIdx.x * kDaTx + threadIdx.x;
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    if (cx < vec_per_row) {
        const int64_t step = (int64_t)gridDim.y * kDaTy;
        for (int64_t r = (int64_t)blockIdx.y * kDaTy + threadIdx.y; r < rows; r += 2 * step) {
            const bool two = r + step < rows;
            const int64_t r1 = two ? r + step : r;
            const uint4 g0 = ldg_stream(dy + r * vec_per_row + cx), g1 = ldg_stream(dy + r1 * vec_per_row + cx);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (h == 1 && !two) break;
                const int64_t rr = h == 0 ? r : r1;
                float v[8];
                unpack8(h == 0 ? g0 : g1, v);
                if (thr16) apply_keep(v, rand4(seed, (uint64_t)rr * vec_per_row + cx), thr16, scale);
                const uint4 out = pack8(v);
                if (dx) dx[rr * vec_per_row + cx] = out;
                if (dxsum) {
                    float q[8];
                    unpack8(out, q);
#pragma unroll
                    for (int i = 0; i < 8; ++i) acc[i] += q[i];
                }
            }
        }
    }
    if (dxsum == nullptr) return;
#pragma unroll
    for (int i = 0; i < 8; ++i) red[threadIdx.y][threadIdx.x * 8 + i] = acc[i];
    __syncthreads();
    const int t = threadIdx.y * kDaTx + threadIdx.x;
    const int colo = blockIdx.x * kDaTx * 8 + t;
    if (colo < vec_per_row * 8) {
        float sum = 0.f;
#pragma unroll
        for (int yy = 0; yy < kDaTy; ++yy) sum += red[yy][t];
        atomicAdd(dxsum + colo, sum);
    }
}

static dim3 da_grid(int64_t rows, int vec_per_row) {
    const int gx = (vec_per_row + kDaTx - 1) / kDaTx;
    int64_t gy = ((int64_t)num_sms() * 8 + gx - 1) / gx;
    const int64_t need = (rows + kDaTy - 1) / kDaTy;
    if (gy > need) gy = need;
    if (gy < 1) gy = 1;
    return dim3((unsigned)gx, (unsigned)gy);
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
    return mdt_dropout_add_ln_bwd_colsum(dy, z, mean, rstd, gamma, nullptr, dx, dres, dgamma, dbeta, nullptr, rows, C, p, seed, stream);
}

extern "C" int mdt_dropout_add_ln_bwd_colsum(const void *dy, const void *z, const float *mean, const float *rstd, const void *gamma,
                                             const void *dadd, void *dx, void *dres, float *dgamma, float *dbeta, float *dxsum,
                                             int64_t rows, int C, float p, uint64_t seed, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    if (int rc = check_c(C, "mdt_dropout_add_ln_bwd")) return rc;
    MDT_REQUIRE(dy && z && mean && rstd && gamma && dgamma && dbeta, "mdt_dropout_add_ln_bwd: null argument");
    MDT_REQUIRE(dx || dres, "mdt_dropout_add_ln_bwd: dx or dres must be given");
    MDT_REQUIRE(p >= 0.f && p < 1.f, "mdt_dropout_add_ln_bwd: p must be in [0, 1)");
    if (rows <= 0) return MDT_OK;
    ln::BwdArgs a;
    a.dy = (const __nv_bfloat16 *)dy, a.z = (const __nv_bfloat16 *)z, a.gamma = (const __nv_bfloat16 *)gamma, a.mean = mean, a.rstd = rstd;
    a.dx = (__nv_bfloat16 *)dx, a.dres = (__nv_bfloat16 *)dres, a.dgamma = dgamma, a.dbeta = dbeta, a.dxsum = dxsum;
    a.dadd = (const __nv_bfloat16 *)dadd;
    a.rows = rows, a.thr16 = thr16_of(p), a.scale = 1.f / (1.f - p), a.seed = seed;
    cudaStream_t st = as_stream(stream);
    switch (C / 256) {
        case 1: return ln::launch_bwd<1>(a, st);
        case 2: return ln::launch_bwd<2>(a, st);
        case 3: return ln::launch_bwd<3>(a, st);
        default: return ln::launch_bwd<4>(a, st);
    }
}

extern "C" int mdt_dropout_add_fwd(const void *x, const void *residual, void *y, int64_t rows, int C, float p, uint64_t seed,
                                   mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(x && y && rows >= 0 && C >= 8 && C % 8 == 0, "mdt_dropout_add_fwd: x, y non-null, C a multiple of 8");
    MDT_REQUIRE(p >= 0.f && p < 1.f, "mdt_dropout_add_fwd: p must be in [0, 1)");
    MDT_REQUIRE((((uintptr_t)x | (uintptr_t)residual | (uintptr_t)y) & 15) == 0, "mdt_dropout_add_fwd: buffers must be 16-byte aligned");
    if (rows == 0) return MDT_OK;
    ln::dropout_add_fwd_kernel<<<ln::da_grid(rows, C / 8), dim3(ln::kDaTx, ln::kDaTy), 0, as_stream(stream)>>>(
        (const uint4 *)x, (const uint4 *)residual, (uint4 *)y, rows, C / 8, thr16_of(p), 1.f / (1.f - p), seed);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

extern "C" int mdt_dropout_bwd(const void *dy, void *dx, float *dxsum, int64_t rows, int C, float p, uint64_t seed,
                               mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(dy && (dx || dxsum) && rows >= 0 && C >= 8 && C % 8 == 0, "mdt_dropout_bwd: dy and one of dx / dxsum, C a multiple of 8");
    MDT_REQUIRE(p >= 0.f && p < 1.f, "mdt_dropout_bwd: p must be in [0, 1)");
    MDT_REQUIRE((((uintptr_t)dy | (uintptr_t)dx) & 15) == 0, "mdt_dropout_bwd: buffers must be 16-byte aligned");
    if (rows == 0) return MDT_OK;
    ln::dropout_bwd_kernel<<<ln::da_grid(rows, C / 8), dim3(ln::kDaTx, ln::kDaTy), 0, as_stream(stream)>>>(
        (const uint4 *)dy, (uint4 *)dx, dxsum, rows, C / 8, thr16_of(p), 1.f / (1.f - p), seed);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

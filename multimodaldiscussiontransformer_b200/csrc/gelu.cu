// N2: the GELU of the feed-forward blocks as two HBM-bound passes of this library, the backward one fused with the
// bias-gradient column sums of the dense layer in front of it.
//   mDT/src/modules/graphormer_graph_encoder_layer.py:132-136 (fc1 -> gelu in fp32 -> dropout), and the Hugging Face
//   BertIntermediate / ViTIntermediate blocks called from mDT/src/modules/multi_graphormer_fusion_layer.py:94-98,138-148.
//
//   forward : g = u * Phi(u)                      u, g bf16 [rows, C]; arithmetic in fp32 (fairseq: gelu(x.float()).type_as(x))
//   backward: du = dg * (Phi(u) + u * phi(u))      and  dbias[c] += sum_rows du[r, c]   (f32, accumulated; may be null)
// Phi is the exact (erf) Gaussian CDF, evaluated as 1 - 0.5 * poly(t) * exp(-u^2 / 2) with t = 1 / (1 + p |u| / sqrt 2)
// (Abramowitz-Stegun 7.1.26, |error| < 1.5e-7: far below a bf16 ulp); the same exponential serves phi in the backward, so
// an element costs one MUFU.EX2 and one MUFU.RCP.  ATen runs the two GELU kernels at ~55 % of the HBM peak (erff + expf per
// element) and needs a third full pass over du for the bias gradient.
// Threads own 8 consecutive columns (16-byte accesses, a warp covers 512 contiguous bytes) and walk down the rows, so the
// column sums stay in registers until one shared-memory reduction and 8 atomics per thread column at the end.
#include "common.cuh"

namespace mdt {
namespace gelu {

constexpr int kTx = 32, kTy = 8;   // 32 vector columns (256 elements) x 8 row lanes per CTA

__device__ __forceinline__ void unpack8(const uint4 &v, float *f) {
    const __nv_bfloat162 *p = reinterpret_cast<const __nv_bfloat162 *>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 t = __bfloat1622float2(p[i]);
        f[2 * i] = t.x, f[2 * i + 1] = t.y;
    }
}
__device__ __forceinline__ uint4 pack8(const float *f) {
    uint4 v;
    __nv_bfloat162 *p = reinterpret_cast<__nv_bfloat162 *>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) p[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return v;
}

// Phi(x) and exp(-x^2 / 2)
__device__ __forceinline__ void phi_cdf(float x, float &cdf, float &e) {
    const float ax = fabsf(x);
    const float z = ax * 0.70710678118654752f;
    float t;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.f)));   // one MUFU.RCP (relative error 2^-23)
    float poly = fmaf(1.061405429f, t, -1.453152027f);
    poly = fmaf(poly, t, 1.421413741f);
    poly = fmaf(poly, t, -0.284496736f);
    poly = fmaf(poly, t, 0.254829592f);
    poly *= t;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-0.72134752044448170f * x * x));   // exp(-x^2 / 2): one MUFU.EX2
    const float half_erfc = 0.5f * poly * e;          // 0.5 * erfc(|x| / sqrt 2)
    cdf = x >= 0.f ? 1.f - half_erfc : half_erfc;
}

__global__ void __launch_bounds__(kTx *kTy) gelu_fwd_kernel(const uint4 *__restrict__ u, uint4 *__restrict__ g, int64_t rows, int vec_per_row) {
    const int cx = blockIdx.x * kTx + threadIdx.x;
    if (cx >= vec_per_row) return;
    const int64_t step = (int64_t)gridDim.y * kTy;
    for (int64_t r = (int64_t)blockIdx.y * kTy + threadIdx.y; r < rows; r += 2 * step) {   // two rows in flight per thread
        const bool two = r + step < rows;
        const uint4 v0 = __ldg(u + r * vec_per_row + cx);
        const uint4 v1 = two ? __ldg(u + (r + step) * vec_per_row + cx) : v0;
        float x[8], y[8];
        unpack8(v0, x);
        unpack8(v1, y);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float cdf, e;
            phi_cdf(x[i], cdf, e);
            x[i] *= cdf;
            phi_cdf(y[i], cdf, e);
            y[i] *= cdf;
        }
        g[r * vec_per_row + cx] = pack8(x);
        if (two) g[(r + step) * vec_per_row + cx] = pack8(y);
    }
}

__global__ void __launch_bounds__(kTx *kTy) gelu_bwd_kernel(const uint4 *dg, const uint4 *__restrict__ u, uint4 *du,
                                                             float *__restrict__ dbias, int64_t rows, int vec_per_row) {
    __shared__ float red[kTy][kTx * 8 + 1];
    const int cx = blockIdx.x * kTx + threadIdx.x;
    const bool col_ok = cx < vec_per_row;
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    if (col_ok) {
        const int64_t step = (int64_t)gridDim.y * kTy;
        // two rows in flight per thread; all four loads are issued before the first store (du may be dg: in-place)
        for (int64_t r = (int64_t)blockIdx.y * kTy + threadIdx.y; r < rows; r += 2 * step) {
            const bool two = r + step < rows;
            const int64_t r1 = two ? r + step : r;
            const uint4 u0 = __ldg(u + r * vec_per_row + cx), u1 = __ldg(u + r1 * vec_per_row + cx);
            const uint4 g0 = dg[r * vec_per_row + cx], g1 = dg[r1 * vec_per_row + cx];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (h == 1 && !two) break;
                float x[8], d[8];
                unpack8(h == 0 ? u0 : u1, x);
                unpack8(h == 0 ? g0 : g1, d);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    float cdf, e;
                    phi_cdf(x[i], cdf, e);
                    d[i] *= fmaf(x[i] * 0.39894228040143268f, e, cdf);   // Phi(x) + x * phi(x)
                }
                const uint4 out = pack8(d);
                du[(h == 0 ? r : r1) * vec_per_row + cx] = out;
                if (dbias != nullptr) {   // the sums are of the ROUNDED gradient, the tensor the weight-gradient GEMM reads
                    float q[8];
                    unpack8(out, q);
#pragma unroll
                    for (int i = 0; i < 8; ++i) acc[i] += q[i];
                }
            }
        }
    }
    if (dbias == nullptr) return;
#pragma unroll
    for (int i = 0; i < 8; ++i) red[threadIdx.y][threadIdx.x * 8 + i] = acc[i];
    __syncthreads();
    const int t = threadIdx.y * kTx + threadIdx.x;   // 256 threads: one column each
    const int col = blockIdx.x * kTx * 8 + t;
    if (col < vec_per_row * 8) {
        float s = 0.f;
#pragma unroll
        for (int y = 0; y < kTy; ++y) s += red[y][t];
        atomicAdd(dbias + col, s);
    }
}

static dim3 grid_for(int64_t rows, int vec_per_row) {
    const int gx = (vec_per_row + kTx - 1) / kTx;
    int64_t gy = ((int64_t)num_sms() * 8 + gx - 1) / gx;            // ~8 CTAs of 256 threads per SM in all
    const int64_t need = (rows + kTy - 1) / kTy;
    if (gy > need) gy = need;
    if (gy < 1) gy = 1;
    return dim3((unsigned)gx, (unsigned)gy);
}

}  // namespace gelu
}  // namespace mdt

using namespace mdt;

extern "C" int mdt_gelu_fwd(const void *u, void *g, int64_t rows, int C, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(u && g && rows >= 0 && C >= 8 && C % 8 == 0, "mdt_gelu_fwd: u, g non-null, C a multiple of 8");
    MDT_REQUIRE((((uintptr_t)u | (uintptr_t)g) & 15) == 0, "mdt_gelu_fwd: buffers must be 16-byte aligned");
    if (rows == 0) return MDT_OK;
    gelu::gelu_fwd_kernel<<<gelu::grid_for(rows, C / 8), dim3(gelu::kTx, gelu::kTy), 0, as_stream(stream)>>>(
        reinterpret_cast<const uint4 *>(u), reinterpret_cast<uint4 *>(g), rows, C / 8);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

extern "C" int mdt_gelu_bwd(const void *dg, const void *u, void *du, float *dbias, int64_t rows, int C, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(dg && u && du && rows >= 0 && C >= 8 && C % 8 == 0, "mdt_gelu_bwd: dg, u, du non-null, C a multiple of 8");
    MDT_REQUIRE((((uintptr_t)u | (uintptr_t)dg | (uintptr_t)du) & 15) == 0, "mdt_gelu_bwd: buffers must be 16-byte aligned");
    if (rows == 0) return MDT_OK;
    gelu::gelu_bwd_kernel<<<gelu::grid_for(rows, C / 8), dim3(gelu::kTx, gelu::kTy), 0, as_stream(stream)>>>(
        reinterpret_cast<const uint4 *>(dg), reinterpret_cast<const uint4 *>(u), reinterpret_cast<uint4 *>(du), dbias, rows, C / 8);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

// Row-wise HBM-bound kernels: GraphNodeFeature and the node-major <-> padded row moves (sm_100a).
//
// Replaces   mDT/src/modules/graphormer_layers.py:39-50                       GraphNodeFeature.forward (A4)
//            mDT/src/modules/multigraphormer_graph_encoder.py:363-371,425,435  bool-mask gather/scatter (K12)
//
// One warp per row, 16-byte accesses, grid sized from the row count; no host sync anywhere.
#include "common.cuh"

namespace mdt {

template <typename T> struct RowVec;
template <> struct RowVec<float> {
    static constexpr int n = 4;
    using type = float4;
};
template <> struct RowVec<__nv_bfloat16> {
    static constexpr int n = 8;
    using type = uint4;
};

template <typename T> __device__ __forceinline__ void unpack(const typename RowVec<T>::type &v, float *f);
template <> __device__ __forceinline__ void unpack<float>(const float4 &v, float *f) {
    f[0] = v.x, f[1] = v.y, f[2] = v.z, f[3] = v.w;
}
template <> __device__ __forceinline__ void unpack<__nv_bfloat16>(const uint4 &v, float *f) {
    const __nv_bfloat162 *p = reinterpret_cast<const __nv_bfloat162 *>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float2 t = __bfloat1622float2(p[i]);
        f[2 * i] = t.x, f[2 * i + 1] = t.y;
    }
}
template <typename T> __device__ __forceinline__ typename RowVec<T>::type pack(const float *f);
template <> __device__ __forceinline__ float4 pack<float>(const float *f) { return make_float4(f[0], f[1], f[2], f[3]); }
template <> __device__ __forceinline__ uint4 pack<__nv_bfloat16>(const float *f) {
    uint4 v;
    __nv_bfloat162 *p = reinterpret_cast<__nv_bfloat162 *>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) p[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return v;
}

// out[b,0,:] = token ; out[b,1+i,:] = x[b,i,:] + Ein[in[b,i]] + Eout[out[b,i]]   (sum order as the reference: (x+in)+out)
template <typename T>
__global__ void node_feature_fwd_kernel(const T *__restrict__ x, const int64_t *__restrict__ in_deg,
                                        const int64_t *__restrict__ out_deg, const T *__restrict__ w_in,
                                        const T *__restrict__ w_out, const T *__restrict__ token, int B, int N, int C,
                                        int num_in, int num_out, T *__restrict__ out) {
    using V = typename RowVec<T>::type;
    constexpr int VN = RowVec<T>::n;
    const int warps_per_block = blockDim.x >> 5;
    const int lane = threadIdx.x & 31;
    const long rows = (long)B * (N + 1);
    for (long row = (long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < rows; row += (long)gridDim.x * warps_per_block) {
        const int b = (int)(row / (N + 1)), t = (int)(row - (long)b * (N + 1));
        V *o = reinterpret_cast<V *>(out + row * C);
        if (t == 0) {
            const V *tk = reinterpret_cast<const V *>(token);
            for (int c = lane; c < C / VN; c += 32) o[c] = tk[c];
            continue;
        }
        const long node = (long)b * N + (t - 1);
        long di = in_deg[node], dq_ = out_deg[node];
        di = di < 0 ? 0 : (di >= num_in ? num_in - 1 : di);
        dq_ = dq_ < 0 ? 0 : (dq_ >= num_out ? num_out - 1 : dq_);
        const V *xv = reinterpret_cast<const V *>(x + node * C);
        const V *a = reinterpret_cast<const V *>(w_in + di * C);
        const V *c2 = reinterpret_cast<const V *>(w_out + dq_ * C);
        for (int c = lane; c < C / VN; c += 32) {
            float fx[VN], fa[VN], fb[VN];
            unpack<T>(xv[c], fx);
            unpack<T>(a[c], fa);
            unpack<T>(c2[c], fb);
#pragma unroll
            for (int q = 0; q < VN; ++q) fx[q] = to_float(from_float<T>(to_float(from_float<T>(fx[q] + fa[q])) + fb[q]));
            o[c] = pack<T>(fx);
        }
    }
}

// grad_x[b,i,:] = go[b,1+i,:]; grad tables via f32 atomics (degree values are few, rows many:
// the reduction is done per warp-row, contention is spread over C columns); grad_token = sum_b go[b,0,:]
template <typename T>
__global__ void node_feature_bwd_kernel(const T *__restrict__ go, const int64_t *__restrict__ in_deg,
                                        const int64_t *__restrict__ out_deg, int B, int N, int C, int num_in, int num_out,
                                        T *__restrict__ gx, float *__restrict__ g_in, float *__restrict__ g_out,
                                        float *__restrict__ g_tok) {
    using V = typename RowVec<T>::type;
    constexpr int VN = RowVec<T>::n;
    const int warps_per_block = blockDim.x >> 5;
    const int lane = threadIdx.x & 31;
    const long rows = (long)B * (N + 1);
    for (long row = (long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < rows; row += (long)gridDim.x * warps_per_block) {
        const int b = (int)(row / (N + 1)), t = (int)(row - (long)b * (N + 1));
        const V *g = reinterpret_cast<const V *>(go + row * C);
        if (t == 0) {
            if (g_tok)
                for (int c = lane; c < C / VN; c += 32) {
                    float f[VN];
                    unpack<T>(g[c], f);
#pragma unroll
                    for (int q = 0; q < VN; ++q) atomicAdd(&g_tok[c * VN + q], f[q]);
                }
            continue;
        }
        const long node = (long)b * N + (t - 1);
        const long di = in_deg[node], dq_ = out_deg[node];
        V *gxv = gx ? reinterpret_cast<V *>(gx + node * C) : nullptr;
        for (int c = lane; c < C / VN; c += 32) {
            const V v = g[c];
            if (gxv) gxv[c] = v;
            float f[VN];
            unpack<T>(v, f);
            if (g_in && di > 0 && di < num_in)  // padding_idx = 0 gets no gradient (graphormer_layers.py:30-35)
#pragma unroll
                for (int q = 0; q < VN; ++q) atomicAdd(&g_in[di * C + c * VN + q], f[q]);
            if (g_out && dq_ > 0 && dq_ < num_out)
#pragma unroll
                for (int q = 0; q < VN; ++q) atomicAdd(&g_out[dq_ * C + c * VN + q], f[q]);
        }
    }
}

template <bool SCATTER>
__global__ void move_rows_kernel(const uint4 *__restrict__ src, long src_pitch16, const int32_t *__restrict__ index,
                                 int n_rows, int row16, uint4 *__restrict__ dst, long dst_pitch16) {
    const int warps_per_block = blockDim.x >> 5;
    const int lane = threadIdx.x & 31;
    for (long r = (long)blockIdx.x * warps_per_block + (threadIdx.x >> 5); r < n_rows; r += (long)gridDim.x * warps_per_block) {
        const long ix = index[r];
        const uint4 *s = src + (SCATTER ? r : ix) * src_pitch16;
        uint4 *d = dst + (SCATTER ? ix : r) * dst_pitch16;
        for (int c = lane; c < row16; c += 32) d[c] = s[c];
    }
}

static inline int row_grid(long rows, int warps_per_block) {
    long g = (rows + warps_per_block - 1) / warps_per_block;
    const long cap = (long)num_sms() * 16;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace mdt

using namespace mdt;

extern "C" int mdt_node_feature_fwd(const void *x, const int64_t *in_degree, const int64_t *out_degree, const void *w_in,
                                    const void *w_out, const void *graph_token, int B, int N, int C, int num_in,
                                    int num_out, int dtype, void *out, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(x && in_degree && out_degree && w_in && w_out && graph_token && out, "null pointer");
    MDT_REQUIRE(dtype == MDT_F32 || dtype == MDT_BF16, "unknown dtype %d", dtype);
    const int vn = dtype == MDT_F32 ? 4 : 8;
    MDT_REQUIRE(B >= 1 && N >= 1 && C >= vn && C % vn == 0, "bad shape B=%d N=%d C=%d", B, N, C);
    const int grid = row_grid((long)B * (N + 1), 8);
    if (dtype == MDT_F32)
        node_feature_fwd_kernel<float><<<grid, 256, 0, as_stream(stream)>>>((const float *)x, in_degree, out_degree, (const float *)w_in,
                                                                             (const float *)w_out, (const float *)graph_token, B, N,
                                                                             C, num_in, num_out, (float *)out);
    else
        node_feature_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, as_stream(stream)>>>(
            (const __nv_bfloat16 *)x, in_degree, out_degree, (const __nv_bfloat16 *)w_in, (const __nv_bfloat16 *)w_out,
            (const __nv_bfloat16 *)graph_token, B, N, C, num_in, num_out, (__nv_bfloat16 *)out);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

extern "C" int mdt_node_feature_bwd(const void *grad_out, const int64_t *in_degree, const int64_t *out_degree, int B, int N,
                                    int C, int num_in, int num_out, int dtype, void *grad_x, float *grad_w_in,
                                    float *grad_w_out, float *grad_token, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(grad_out && in_degree && out_degree, "null pointer");
    MDT_REQUIRE(dtype == MDT_F32 || dtype == MDT_BF16, "unknown dtype %d", dtype);
    const int vn = dtype == MDT_F32 ? 4 : 8;
    MDT_REQUIRE(B >= 1 && N >= 1 && C >= vn && C % vn == 0, "bad shape B=%d N=%d C=%d", B, N, C);
    const int grid = row_grid((long)B * (N + 1), 8);
    if (dtype == MDT_F32)
        node_feature_bwd_kernel<float><<<grid, 256, 0, as_stream(stream)>>>((const float *)grad_out, in_degree, out_degree, B, N, C,
                                                                             num_in, num_out, (float *)grad_x, grad_w_in, grad_w_out,
                                                                             grad_token);
    else
        node_feature_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, as_stream(stream)>>>(
            (const __nv_bfloat16 *)grad_out, in_degree, out_degree, B, N, C, num_in, num_out, (__nv_bfloat16 *)grad_x, grad_w_in,
            grad_w_out, grad_token);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

static int move_rows(bool scatter, const void *src, int64_t src_pitch, const int32_t *index, int n_rows, int row_elems,
                     int dtype, void *dst, int64_t dst_pitch, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(src && index && dst, "null pointer");
    MDT_REQUIRE(dtype == MDT_F32 || dtype == MDT_BF16, "unknown dtype %d", dtype);
    const int eb = dtype == MDT_F32 ? 4 : 2;
    MDT_REQUIRE((row_elems * eb) % 16 == 0 && (src_pitch * eb) % 16 == 0 && (dst_pitch * eb) % 16 == 0,
                "rows and pitches must be multiples of 16 bytes");
    MDT_REQUIRE((((uintptr_t)src) & 15) == 0 && (((uintptr_t)dst) & 15) == 0, "pointers must be 16-byte aligned");
    if (n_rows <= 0) return MDT_OK;
    const int grid = row_grid(n_rows, 8);
    if (scatter)
        move_rows_kernel<true><<<grid, 256, 0, as_stream(stream)>>>((const uint4 *)src, src_pitch * eb / 16, index, n_rows,
                                                                     row_elems * eb / 16, (uint4 *)dst, dst_pitch * eb / 16);
    else
        move_rows_kernel<false><<<grid, 256, 0, as_stream(stream)>>>((const uint4 *)src, src_pitch * eb / 16, index, n_rows,
                                                                      row_elems * eb / 16, (uint4 *)dst, dst_pitch * eb / 16);
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

extern "C" int mdt_gather_rows(const void *src, int64_t src_pitch, const int32_t *index, int n_rows, int row_elems,
                               int dtype, void *dst, int64_t dst_pitch, mdt_stream_t stream) {
    return move_rows(false, src, src_pitch, index, n_rows, row_elems, dtype, dst, dst_pitch, stream);
}
extern "C" int mdt_scatter_rows(const void *src, int64_t src_pitch, const int32_t *index, int n_rows, int row_elems,
                                int dtype, void *dst, int64_t dst_pitch, mdt_stream_t stream) {
    return move_rows(true, src, src_pitch, index, n_rows, row_elems, dtype, dst, dst_pitch, stream);
}

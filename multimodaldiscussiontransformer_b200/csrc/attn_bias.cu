// GraphAttnBias forward / backward, materialising variant (sm_100a).
//
// Replaces   mDT/src/modules/graphormer_layers.py:86-110   GraphAttnBias.forward   (A5)
//
//   out[b,h,i,j] = 2*attn_bias[b,i,j] + (i>=1 && j>=1 ? table[spatial_pos[b,i-1,j-1]][h] : virt[h])
//
// HBM-bound: B*T*T*(e+4) bytes in, B*H*T*T*e bytes out.  One CTA owns a chunk of one
// graph's [T*T] plane: it stages one BYTE per position in shared memory (spatial index,
// 254 = graph-token row/column, 255 = -inf) plus a [H][256] value table, then streams the
// H output planes with 16-byte stores.  T = N+1 is odd, so plane starts are only
// element-aligned: vectors are aligned on the global address and the byte codes are read
// with a funnel shift.  Inputs whose attn_bias holds anything but 0 / -inf (or indices
// above 253) take the element-wise path of the same kernel.
#include "common.cuh"

namespace mdt {

constexpr int kChunk = 8192;
constexpr int kBiasThreads = 256;

template <typename T> struct Vec16;
template <> struct Vec16<float> { static constexpr int n = 4; };
template <> struct Vec16<__nv_bfloat16> { static constexpr int n = 8; };

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&v);
}

template <typename T>
__global__ void __launch_bounds__(kBiasThreads)
    attn_bias_fwd_kernel(const T *__restrict__ ab, const int32_t *__restrict__ sp, const float *__restrict__ table,
                         const float *__restrict__ virt, int H, int N, int num_spatial, T *__restrict__ out) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint8_t *code = smem;                                          // [kChunk + 16]
    float *lut = reinterpret_cast<float *>(smem + kChunk + 16);    // [H][256]
    const int T_ = N + 1;
    const size_t plane = (size_t)T_ * T_;
    const int b = blockIdx.y;
    const size_t r0 = (size_t)blockIdx.x * kChunk;
    const int len = (int)min((size_t)kChunk, plane - r0);
    const T *abp = ab + (size_t)b * plane + r0;
    const int32_t *spp = sp + (size_t)b * N * N;

    int ok = 1;
    for (int e = threadIdx.x; e < len; e += kBiasThreads) {
        const size_t r = r0 + e;
        const int i = (int)(r / T_), j = (int)(r - (size_t)i * T_);
        const float a = to_float(abp[e]);
        int c = MDT_CODE_VIRTUAL;
        if (i >= 1 && j >= 1) {
            c = spp[(size_t)(i - 1) * N + (j - 1)];
            ok &= (c >= 0 && c < 254);
        }
        if (a == neg_inf())
            c = MDT_CODE_MASKED;
        else
            ok &= (a == 0.f);
        code[e] = (uint8_t)c;
    }
    for (int t = threadIdx.x; t < H * 256; t += kBiasThreads) {
        const int h = t >> 8, c = t & 255;
        float v = 0.f;
        if (c == MDT_CODE_MASKED)
            v = neg_inf();
        else if (c == MDT_CODE_VIRTUAL)
            v = virt[h];
        else if (c < num_spatial)
            v = to_float(from_float<T>(table[(size_t)c * H + h]));
        lut[t] = v;
    }
    const int all_ok = __syncthreads_and(ok);

    constexpr int V = Vec16<T>::n;
    if (all_ok) {
        for (int h = 0; h < H; ++h) {
            const float *l = lut + h * 256;
            const size_t e0 = ((size_t)b * H + h) * plane + r0;
            T *o = out + e0;
            int head = (int)((V - (e0 % V)) % V);
            if (head > len) head = len;
            if ((int)threadIdx.x < head) o[threadIdx.x] = from_float<T>(l[code[threadIdx.x]]);
            const int nvec = (len - head) / V;
            for (int v = threadIdx.x; v < nvec; v += kBiasThreads) {
                const int e = head + v * V;
                // V consecutive code bytes starting at byte offset e (unaligned)
                const uint32_t *w = reinterpret_cast<const uint32_t *>(code + (e & ~3));
                const int sh = (e & 3) * 8;
                const uint32_t w0 = w[0], w1 = w[1];
                const uint32_t c0 = __funnelshift_r(w0, w1, sh);
                if constexpr (V == 4) {
                    float4 r;
                    r.x = l[c0 & 255];
                    r.y = l[(c0 >> 8) & 255];
                    r.z = l[(c0 >> 16) & 255];
                    r.w = l[c0 >> 24];
                    __stcs(reinterpret_cast<float4 *>(o + e), r);
                } else {
                    const uint32_t w2 = w[2];
                    const uint32_t c1 = __funnelshift_r(w1, w2, sh);
                    uint4 r;
                    r.x = pack_bf16(l[c0 & 255], l[(c0 >> 8) & 255]);
                    r.y = pack_bf16(l[(c0 >> 16) & 255], l[c0 >> 24]);
                    r.z = pack_bf16(l[c1 & 255], l[(c1 >> 8) & 255]);
                    r.w = pack_bf16(l[(c1 >> 16) & 255], l[c1 >> 24]);
                    __stcs(reinterpret_cast<uint4 *>(o + e), r);
                }
            }
            const int tail0 = head + nvec * V;
            if ((int)threadIdx.x < len - tail0) o[tail0 + threadIdx.x] = from_float<T>(l[code[tail0 + threadIdx.x]]);
        }
    } else {
        // element-wise path: arbitrary attn_bias values / indices; same rounding steps as
        // the reference (add in the tensor dtype, then add attn_bias again, :93-108)
        for (int h = 0; h < H; ++h) {
            T *o = out + ((size_t)b * H + h) * plane + r0;
            for (int e = threadIdx.x; e < len; e += kBiasThreads) {
                const size_t r = r0 + e;
                const int i = (int)(r / T_), j = (int)(r - (size_t)i * T_);
                const float a = to_float(abp[e]);
                float add;
                if (i >= 1 && j >= 1) {
                    const int c = spp[(size_t)(i - 1) * N + (j - 1)];
                    add = (c >= 0 && c < num_spatial) ? to_float(from_float<T>(table[(size_t)c * H + h])) : 0.f;
                } else {
                    add = to_float(from_float<T>(virt[h]));
                }
                const float s1 = to_float(from_float<T>(a + add));
                o[e] = from_float<T>(s1 + a);
            }
        }
    }
}

// backward: grad_table[c][h] += sum of grad_out over positions with spatial index c (c != 0),
// grad_virt[h] += sum over the graph-token row and column.  Reads grad_out once.  A thread owns 32 positions of the
// chunk (stride kBiasThreads, so a warp load covers whole sectors); their bin codes are the same for every head, so
// they are packed into 8 registers once.  Per head the 32 loads are issued back to back, and the next head's loads
// are in flight while this head's values go into per-thread private shared-memory bins (indices 0..31; mDT uses
// 1..21) - the first version issued one dependent load per bin update and ran at the latency of 32 x H loads per CTA.
constexpr int kPerThread = kChunk / kBiasThreads;  // 32
constexpr int kBins = 33;                          // spatial indices 0..31 + the graph-token (virtual distance) bin
constexpr uint32_t kCodeVirt = 32, kCodeFar = 254;

template <typename T>
__device__ __forceinline__ void bias_bwd_load(T (&v)[kPerThread], const T *__restrict__ g, int len, bool full) {
    if (full) {
#pragma unroll
        for (int k = 0; k < kPerThread; ++k) v[k] = __ldcs(g + k * kBiasThreads + threadIdx.x);
    } else {
#pragma unroll
        for (int k = 0; k < kPerThread; ++k) {
            const int e = k * kBiasThreads + threadIdx.x;
            v[k] = e < len ? __ldcs(g + e) : from_float<T>(0.f);
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(kBiasThreads)
    attn_bias_bwd_kernel(const T *__restrict__ go, const int32_t *__restrict__ sp, int H, int N, int num_spatial,
                         float *__restrict__ grad_table, float *__restrict__ grad_virt) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint16_t *code = reinterpret_cast<uint16_t *>(smem);                              // [kChunk]
    float *bins = reinterpret_cast<float *>(smem + kChunk * sizeof(uint16_t));        // [kBins][kBiasThreads]
    const int T_ = N + 1;
    const size_t plane = (size_t)T_ * T_;
    const int b = blockIdx.y;
    const size_t r0 = (size_t)blockIdx.x * kChunk;
    const int len = (int)min((size_t)kChunk, plane - r0);
    const bool full = len == kChunk;
    const int32_t *spp = sp + (size_t)b * N * N;
    const T *g0 = go + (size_t)b * H * plane + r0;
    T cur[kPerThread];
    bias_bwd_load(cur, g0, len, full);  // head 0 travels while the codes are built
    uint32_t cw[kPerThread / 4];
#pragma unroll
    for (int k = 0; k < kPerThread; ++k) {
        const int e = k * kBiasThreads + threadIdx.x;
        uint32_t byte = 0;  // past the chunk: value 0 into bin 0 (discarded)
        if (e < len) {
            const size_t r = r0 + e;
            const int i = (int)(r / T_), j = (int)(r - (size_t)i * T_);
            int c = 0xffff;  // virtual
            if (i >= 1 && j >= 1) {
                c = spp[(size_t)(i - 1) * N + (j - 1)];
                if (c < 0 || c >= num_spatial || c >= 0xfffe) c = 0;  // treated like padding: no gradient
            }
            code[e] = (uint16_t)c;
            byte = c < 32 ? (uint32_t)c : (c == 0xffff ? kCodeVirt : kCodeFar);
        }
        if ((k & 3) == 0) cw[k >> 2] = 0;
        cw[k >> 2] |= byte << (8 * (k & 3));
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int h = 0; h < H; ++h) {
        T nxt[kPerThread];
        if (h + 1 < H) bias_bwd_load(nxt, g0 + (size_t)(h + 1) * plane, len, full);
#pragma unroll
        for (int c = 0; c < kBins; ++c) bins[c * kBiasThreads + threadIdx.x] = 0.f;
#pragma unroll
        for (int k = 0; k < kPerThread; ++k) {
            const uint32_t c = (cw[k >> 2] >> (8 * (k & 3))) & 255u;
            const float v = to_float(cur[k]);
            if (c != kCodeFar)  // the only branch: no mDT batch takes the far side (indices >= 32)
                bins[c * kBiasThreads + threadIdx.x] += v;
            else
                atomicAdd(&grad_table[(size_t)code[k * kBiasThreads + threadIdx.x] * H + h], v);
        }
        __syncthreads();
        for (int c = warp; c < kBins; c += kBiasThreads / 32) {
            float s = 0.f;
#pragma unroll
            for (int q = 0; q < kBiasThreads / 32; ++q) s += bins[c * kBiasThreads + q * 32 + lane];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0 && s != 0.f) {
                if (c == (int)kCodeVirt)
                    atomicAdd(&grad_virt[h], s);
                else if (c != 0 && c < num_spatial)
                    atomicAdd(&grad_table[(size_t)c * H + h], s);
            }
        }
        __syncthreads();
        if (h + 1 < H) {
#pragma unroll
            for (int k = 0; k < kPerThread; ++k) cur[k] = nxt[k];
        }
    }
}


// ---- backward, bf16, H <= 16: the histogram as a matrix product on the tensor cores.
//   grad[c][h] = sum_p onehot[c][p] * go[h][p]        c = spatial index 0..31 (row 0 = the graph-token bin: index 0 is padding
//                                                     and gets no gradient), p = position of the [T*T] plane, h = head
// mma.sync.m16n8k16 (bf16 x bf16 -> f32): A = one-hot rows built in registers from the position's code byte (one SIMD byte compare
// + two byte permutes per pair of rows: 0 / 1.0 are exact in bf16, the products are exact, the sums are f32), B = 16 positions x 8
// heads read from a shared-memory copy of the head planes.  Nothing depends on the previous element (the private-bin kernel above
// runs at the latency of its read-modify-write chains: 15 % of the HBM peak), the global loads are 16-byte vectors.
// T * T is odd, so the H planes of a chunk start at different 2-byte alignments: rows are copied as ALIGNED vectors (position p of
// head h sits at element p + mis[h] of its row) and a B register is two 32-bit reads joined by a funnel shift.
// Indices >= 32 (no mDT batch has them) are added with global atomics after the product.
constexpr int kMmaChunk = 1024;                       // positions per step of a CTA
constexpr int kMmaRowBytes = (kMmaChunk + 8) * 2 + 16;  // + one vector of misalignment, + 16: rows 8 banks apart
constexpr int kMmaMaxH = 16;

__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void ab_cp_async16(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}

// Stage layout (two stages: the next chunk's planes and codes travel while this chunk is multiplied):
//   [H][kMmaRowBytes] head rows | code8 [kMmaChunk] one-hot row, 0xff = none | code16 [kMmaChunk] raw index (far path)
__global__ void __launch_bounds__(256)
    attn_bias_bwd_mma_kernel(const __nv_bfloat16 *__restrict__ go, const int32_t *__restrict__ sp, int H, int N, int num_spatial,
                             int chunks_per_cta, float *__restrict__ grad_table, float *__restrict__ grad_virt) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int stage_bytes = H * kMmaRowBytes + kMmaChunk * 3;
    const int T_ = N + 1;
    const size_t plane = (size_t)T_ * T_;
    const int b = blockIdx.y;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    const int32_t *spp = sp + (size_t)b * N * N;
    const __nv_bfloat16 *gb = go + (size_t)b * H * plane;
    float acc[2][2][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) (&acc[0][0][0])[i] = 0.f;
    const int h0 = min(g, H - 1), h1 = min(g + 8, H - 1);   // heads of this lane's B columns (columns >= H: duplicates, never stored)
    // elements past a chunk's end carry one-hot weight 0, but 0 * NaN = NaN: nothing unwritten may ever be multiplied
    for (int i = tid; i < 2 * stage_bytes / 16; i += 256) reinterpret_cast<uint4 *>(smem)[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();

    const size_t chunk0 = (size_t)blockIdx.x * chunks_per_cta;
    // the H plane slices of a chunk as ALIGNED 16-byte vectors: warp w copies heads w and w + 8
    auto issue_rows = [&](size_t r0, int len, uint8_t *stage) {
        for (int h = warp; h < H; h += 8) {
            const __nv_bfloat16 *src = gb + (size_t)h * plane + r0;
            const int mis = (int)((reinterpret_cast<uintptr_t>(src) >> 1) & 7);
            const int nvec = (mis + len + 7) >> 3;
            const uint4 *vsrc = reinterpret_cast<const uint4 *>(src - mis);
            uint4 *dst = reinterpret_cast<uint4 *>(stage + (size_t)h * kMmaRowBytes);
            for (int v = lane; v < nvec; v += 32) ab_cp_async16(dst + v, vsrc + v);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    // the last vector of a row may reach past the chunk (the very last one past the tensor): zero what is not ours
    auto fix_tail = [&](size_t r0, int len, uint8_t *stage) {
        for (int h = warp; h < H; h += 8) {
            const int mis = (int)((reinterpret_cast<uintptr_t>(gb + (size_t)h * plane + r0) >> 1) & 7);
            const int end = mis + len;                                 // first element of the row that is not ours
            uint16_t *row = reinterpret_cast<uint16_t *>(stage + (size_t)h * kMmaRowBytes);
            const int stop = min((end + 7) & ~7, kMmaChunk + 8);
            if (end + lane < stop) row[end + lane] = 0;                // <= 7 elements
        }
    };
    // spatial index of 4 positions per thread (head independent); -1 = the graph-token row / column, -2 = outside the plane
    auto gather_codes = [&](size_t r0, int len, int (&c)[4]) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int e = tid + 256 * k;
            c[k] = -2;
            if (e < len) {
                const uint32_t r = (uint32_t)r0 + e;                   // plane < 2^32
                const uint32_t i = r / (uint32_t)T_, j = r - i * (uint32_t)T_;
                c[k] = (i >= 1 && j >= 1) ? __ldg(spp + (size_t)(i - 1) * N + (j - 1)) : -1;
            }
        }
    };

    int codes[4];
    {
        const size_t r0 = chunk0 * kMmaChunk;
        if (r0 < plane) {
            const int len = (int)min((size_t)kMmaChunk, plane - r0);
            issue_rows(r0, len, smem);
            gather_codes(r0, len, codes);
        }
    }
    for (int cc = 0; cc < chunks_per_cta; ++cc) {
        const size_t r0 = (chunk0 + cc) * kMmaChunk;
        if (r0 >= plane) break;                                        // uniform
        const int len = (int)min((size_t)kMmaChunk, plane - r0);
        uint8_t *stage = smem + (size_t)(cc & 1) * stage_bytes;
        uint8_t *rows = stage;
        uint8_t *code8 = stage + (size_t)H * kMmaRowBytes;
        uint16_t *code16 = reinterpret_cast<uint16_t *>(code8 + kMmaChunk);
        int far = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int c = codes[k];
            int row = 0xff, raw = 0;
            if (c == -1) row = 0;
            else if (c >= 1 && c < num_spatial && c < 0xfffe) {
                if (c < 32) row = c;
                else raw = c, far = 1;
            }
            code8[tid + 256 * k] = (uint8_t)row;
            code16[tid + 256 * k] = (uint16_t)raw;
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
        fix_tail(r0, len, stage);
        const int any_far = __syncthreads_or(far);    // this stage is complete; the other one is no longer being read
        {
            const size_t rn = r0 + kMmaChunk;
            if (cc + 1 < chunks_per_cta && rn < plane) {
                const int ln = (int)min((size_t)kMmaChunk, plane - rn);
                issue_rows(rn, ln, smem + (size_t)((cc + 1) & 1) * stage_bytes);
                gather_codes(rn, ln, codes);
            }
        }
        // ---- 8 steps of 16 positions per warp
        const int mis0 = (int)((reinterpret_cast<uintptr_t>(gb + (size_t)h0 * plane + r0) >> 1) & 7);
        const int mis1 = (int)((reinterpret_cast<uintptr_t>(gb + (size_t)h1 * plane + r0) >> 1) & 7);
        const uint32_t *row0 = reinterpret_cast<const uint32_t *>(rows + (size_t)h0 * kMmaRowBytes);
        const uint32_t *row1 = reinterpret_cast<const uint32_t *>(rows + (size_t)h1 * kMmaRowBytes);
#pragma unroll 4
        for (int s = 0; s < kMmaChunk / 16 / 8; ++s) {
            const int p = (warp * (kMmaChunk / 16 / 8) + s) * 16 + 2 * t;   // this lane's positions: p, p+1, p+8, p+9
            const uint32_t cb = (uint32_t) * reinterpret_cast<const uint16_t *>(code8 + p) |
                                ((uint32_t) * reinterpret_cast<const uint16_t *>(code8 + p + 8) << 16);
            uint32_t bq[2][2];
            {
                const int e = p + mis0;
                const uint32_t w0 = row0[e >> 1], w1 = row0[(e >> 1) + 1], w2 = row0[(e >> 1) + 4], w3 = row0[(e >> 1) + 5];
                bq[0][0] = __funnelshift_r(w0, w1, (e & 1) * 16);
                bq[0][1] = __funnelshift_r(w2, w3, (e & 1) * 16);
            }
            {
                const int e = p + mis1;
                const uint32_t w0 = row1[e >> 1], w1 = row1[(e >> 1) + 1], w2 = row1[(e >> 1) + 4], w3 = row1[(e >> 1) + 5];
                bq[1][0] = __funnelshift_r(w0, w1, (e & 1) * 16);
                bq[1][1] = __funnelshift_r(w2, w3, (e & 1) * 16);
            }
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                uint32_t a[4];
#pragma unroll
                for (int hi = 0; hi < 2; ++hi) {
                    const uint32_t r = (uint32_t)(mt * 16 + hi * 8 + g);
                    const uint32_t m = __vcmpeq4(cb, r * 0x01010101u);          // 0xff where the position's row is r
                    a[hi] = __byte_perm(m, 0, 0x1100) & 0x3f803f80u;            // positions p, p+1   -> bf16 1.0 / 0
                    a[hi + 2] = __byte_perm(m, 0, 0x3322) & 0x3f803f80u;        // positions p+8, p+9
                }
                mma_bf16_16816(acc[mt][0], a, bq[0][0], bq[0][1]);
                mma_bf16_16816(acc[mt][1], a, bq[1][0], bq[1][1]);
            }
        }
        if (any_far) {                                                // uniform; never taken by an mDT batch
            for (int e = tid; e < len; e += 256) {
                const int c = code16[e];
                if (c == 0) continue;
                for (int h = 0; h < H; ++h) {
                    const int mis = (int)((reinterpret_cast<uintptr_t>(gb + (size_t)h * plane + r0) >> 1) & 7);
                    const __nv_bfloat16 v = reinterpret_cast<const __nv_bfloat16 *>(rows + (size_t)h * kMmaRowBytes)[e + mis];
                    atomicAdd(&grad_table[(size_t)c * H + h], __bfloat162float(v));
                }
            }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    // ---- the 8 warps' partial [32 rows][16 heads] tiles -> one atomic per (row, head) and CTA
    __syncthreads();
    float *red = reinterpret_cast<float *>(smem);                     // [8][32][16]
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int row = mt * 16 + g + (i >> 1) * 8, col = nt * 8 + 2 * t + (i & 1);
                red[(warp * 32 + row) * 16 + col] = acc[mt][nt][i];
            }
    __syncthreads();
    for (int o = tid; o < 32 * 16; o += 256) {
        const int row = o >> 4, col = o & 15;
        float sum = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) sum += red[(w * 32 + row) * 16 + col];
        if (col < H && sum != 0.f) {
            if (row == 0)
                atomicAdd(&grad_virt[col], sum);
            else if (row < num_spatial)
                atomicAdd(&grad_table[(size_t)row * H + col], sum);
        }
    }
}

}  // namespace mdt

using namespace mdt;

extern "C" int mdt_attn_bias_fwd(const void *attn_bias, const int32_t *spatial_pos, const float *table, const float *virt,
                                 int B, int H, int N, int num_spatial, int dtype, void *out, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(attn_bias && spatial_pos && table && virt && out, "null pointer");
    MDT_REQUIRE(B >= 1 && H >= 1 && N >= 1 && num_spatial >= 1, "bad shape");
    MDT_REQUIRE(H <= 32, "H=%d > 32 heads not supported by the bias kernel", H);
    MDT_REQUIRE(dtype == MDT_F32 || dtype == MDT_BF16, "unknown dtype %d", dtype);
    const size_t plane = (size_t)(N + 1) * (N + 1);
    dim3 grid((unsigned)((plane + kChunk - 1) / kChunk), B);
    const size_t smem = kChunk + 16 + (size_t)H * 256 * sizeof(float);
    if (dtype == MDT_F32) {
        MDT_CHECK_CUDA(cudaFuncSetAttribute(attn_bias_fwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attn_bias_fwd_kernel<float><<<grid, kBiasThreads, smem, as_stream(stream)>>>(
            (const float *)attn_bias, spatial_pos, table, virt, H, N, num_spatial, (float *)out);
    } else {
        MDT_CHECK_CUDA(cudaFuncSetAttribute(attn_bias_fwd_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attn_bias_fwd_kernel<__nv_bfloat16><<<grid, kBiasThreads, smem, as_stream(stream)>>>(
            (const __nv_bfloat16 *)attn_bias, spatial_pos, table, virt, H, N, num_spatial, (__nv_bfloat16 *)out);
    }
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

extern "C" int mdt_attn_bias_bwd(const void *grad_out, const int32_t *spatial_pos, int B, int H, int N, int num_spatial,
                                 int dtype, float *grad_table, float *grad_virt, mdt_stream_t stream) {
    if (int rc = require_sm100()) return rc;
    MDT_REQUIRE(grad_out && spatial_pos && grad_table && grad_virt, "null pointer");
    MDT_REQUIRE(B >= 1 && H >= 1 && N >= 1 && num_spatial >= 1, "bad shape");
    MDT_REQUIRE(dtype == MDT_F32 || dtype == MDT_BF16, "unknown dtype %d", dtype);
    const size_t plane = (size_t)(N + 1) * (N + 1);
    if (dtype == MDT_BF16 && H <= kMmaMaxH && (reinterpret_cast<uintptr_t>(grad_out) & 1) == 0) {
        const size_t chunks = (plane + kMmaChunk - 1) / kMmaChunk;
        // enough chunks per CTA to bound the atomics, few enough to fill the SMs several times over
        int per_cta = (int)((chunks * (size_t)B) / ((size_t)num_sms() * 8));
        per_cta = per_cta < 1 ? 1 : (per_cta > 8 ? 8 : per_cta);
        size_t smem = 2 * ((size_t)H * kMmaRowBytes + kMmaChunk * 3);   // two stages
        if (smem < 8 * 32 * 16 * sizeof(float)) smem = 8 * 32 * 16 * sizeof(float);
        MDT_CHECK_CUDA(cudaFuncSetAttribute(attn_bias_bwd_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        dim3 grid((unsigned)((chunks + per_cta - 1) / per_cta), B);
        attn_bias_bwd_mma_kernel<<<grid, 256, smem, as_stream(stream)>>>((const __nv_bfloat16 *)grad_out, spatial_pos, H, N,
                                                                          num_spatial, per_cta, grad_table, grad_virt);
        MDT_CHECK_CUDA(cudaGetLastError());
        return MDT_OK;
    }
    dim3 grid((unsigned)((plane + kChunk - 1) / kChunk), B);
    const size_t smem = kChunk * sizeof(uint16_t) + (size_t)kBins * kBiasThreads * sizeof(float);
    if (dtype == MDT_F32) {
        MDT_CHECK_CUDA(cudaFuncSetAttribute(attn_bias_bwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attn_bias_bwd_kernel<float><<<grid, kBiasThreads, smem, as_stream(stream)>>>((const float *)grad_out, spatial_pos, H, N,
                                                                                    num_spatial, grad_table, grad_virt);
    } else {
        MDT_CHECK_CUDA(cudaFuncSetAttribute(attn_bias_bwd_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attn_bias_bwd_kernel<__nv_bfloat16><<<grid, kBiasThreads, smem, as_stream(stream)>>>(
            (const __nv_bfloat16 *)grad_out, spatial_pos, H, N, num_spatial, grad_table, grad_virt);
    }
    MDT_CHECK_CUDA(cudaGetLastError());
    return MDT_OK;
}

// Memory-bound hot ops of the fused GPT-2 engine and of the security layer:
//   K10 LayerNorm fwd / bwd (bwd fuses the residual-gradient add and accumulates dgamma/dbeta)
//   K11 bias gradient (column sums; GELU itself lives in the GEMM epilogue)
//   K14 fused AdamW (fp32 master + moments, bf16 working copy, optional device-side grad scale)
//   K15 cross-entropy fwd+bwd in place over the padded vocabulary
//   K1  bucket pack (fp32 main grad -> bf16 bucket) + local sum of squares
//   K16 fault-injection (scale / sign flip / Gaussian noise, counter-based RNG)
//   K9  building blocks: 64-bit shard checksum, drift norm with snapshot refresh
//   embedding fwd / bwd, multi-tensor squared norms
// All are single-pass, 16-byte vectorised, fp32 math, bf16 storage.  None of them exist in the
// reference (torch eager ops + per-tensor .item() syncs).
#include "common.cuh"

namespace {

__device__ __forceinline__ void bf16x8_to_f32(const uint4& u, float (&f)[8]) {
    float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 f32_to_bf16x8(const float (&f)[8]) {
    uint4 u;
    u.x = pack_bf16x2(f[0], f[1]); u.y = pack_bf16x2(f[2], f[3]); u.z = pack_bf16x2(f[4], f[5]); u.w = pack_bf16x2(f[6], f[7]);
    return u;
}

// ------------------------------------------------------------------------------------ LayerNorm
constexpr int LN_MAX_ITERS = 8;            // C <= 2048, C % 8 == 0; one warp per row

// Templated on the number of 256-column slices per row so the per-lane register arrays are exactly as
// large as the row needs (C = 1024 -> 4 slices); the generic 8-slice version spilled.
template <int ITERS>
__global__ void __launch_bounds__(128) layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x,
        const float* __restrict__ w, const float* __restrict__ b, __nv_bfloat16* __restrict__ y,
        float* __restrict__ mean_out, float* __restrict__ rstd_out, int M, int C, float eps) {
    pdl_enter();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = blockIdx.x * 4 + warp;
    if (row >= M) return;
    const __nv_bfloat16* xr = x + (long long)row * C;
    float v[ITERS][8];
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < ITERS; ++i) {
        const int col = i * 256 + lane * 8;
        if (col < C) {
            bf16x8_to_f32(*reinterpret_cast<const uint4*>(xr + col), v[i]);
#pragma unroll
            for (int j = 0; j < 8; ++j) sum += v[i][j];
        }
    }
    const float mean = warp_sum(sum) / C;
    float var = 0.f;
#pragma unroll
    for (int i = 0; i < ITERS; ++i) {
        if (i * 256 + lane * 8 < C) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { const float d = v[i][j] - mean; var += d * d; }
        }
    }
    const float rstd = rsqrtf(warp_sum(var) / C + eps);
    __nv_bfloat16* yr = y + (long long)row * C;
#pragma unroll
    for (int i = 0; i < ITERS; ++i) {
        const int col = i * 256 + lane * 8;
        if (col < C) {
            float o[8];
            const float4 w0 = *reinterpret_cast<const float4*>(w + col), w1 = *reinterpret_cast<const float4*>(w + col + 4);
            const float4 b0 = *reinterpret_cast<const float4*>(b + col), b1 = *reinterpret_cast<const float4*>(b + col + 4);
            const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
            const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * ww[j] + bb[j];
            *reinterpret_cast<uint4*>(yr + col) = f32_to_bf16x8(o);
        }
    }
    if (lane == 0) { mean_out[row] = mean; rstd_out[row] = rstd; }
}

// dx = dres + LN'(dy) ; dgamma/dbeta accumulated (fp32 atomics) into dw/db.
// Register budget: the row stays in registers as raw bf16 (dy, x) and is unpacked twice; only the
// per-lane dgamma/dbeta partials are fp32 arrays -> 2 blocks of 256 threads per SM.
template <int ITERS>
__global__ void __launch_bounds__(256, 2) layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy,
        const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ mean,
        const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx,
        float* __restrict__ dw, float* __restrict__ db, int M, int C) {
    extern __shared__ float sh[];             // [2][C] block accumulators
    float* sh_dw = sh;
    float* sh_db = sh + C;
    for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) sh[i] = 0.f;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    float aw[ITERS][8], ab[ITERS][8];
#pragma unroll
    for (int i = 0; i < ITERS; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) { aw[i][j] = 0.f; ab[i][j] = 0.f; }
    for (int row = blockIdx.x * nwarps + warp; row < M; row += gridDim.x * nwarps) {
        const long long off = (long long)row * C;
        const float mu = mean[row], rs = rstd[row];
        uint4 rdy[ITERS], rx[ITERS];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < ITERS; ++i) {
            const int col = i * 256 + lane * 8;
            if (col < C) {
                rdy[i] = ld_nc_v4(dy + off + col);
                rx[i] = ld_nc_v4(x + off + col);
            }
        }
#pragma unroll
        for (int i = 0; i < ITERS; ++i) {
            const int col = i * 256 + lane * 8;
            if (col < C) {
                float dyv[8], xv[8];
                bf16x8_to_f32(rdy[i], dyv);
                bf16x8_to_f32(rx[i], xv);
                const float4 w0 = __ldg(reinterpret_cast<const float4*>(w + col)), w1 = __ldg(reinterpret_cast<const float4*>(w + col + 4));
                const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float xh = (xv[j] - mu) * rs, g = dyv[j] * ww[j];
                    s1 += g;
                    s2 += g * xh;
                    aw[i][j] += dyv[j] * xh;
                    ab[i][j] += dyv[j];
                }
            }
        }
        s1 = warp_sum(s1) / C;
        s2 = warp_sum(s2) / C;
#pragma unroll
        for (int i = 0; i < ITERS; ++i) {
            const int col = i * 256 + lane * 8;
            if (col < C) {
                float dyv[8], xv[8], o[8];
                bf16x8_to_f32(rdy[i], dyv);
                bf16x8_to_f32(rx[i], xv);
                const float4 w0 = __ldg(reinterpret_cast<const float4*>(w + col)), w1 = __ldg(reinterpret_cast<const float4*>(w + col + 4));
                const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
                if (dres != nullptr) bf16x8_to_f32(ld_nc_v4(dres + off + col), o);
                else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) o[j] = 0.f;
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float xh = (xv[j] - mu) * rs;
                    o[j] += (dyv[j] * ww[j] - s1 - xh * s2) * rs;
                }
                *reinterpret_cast<uint4*>(dx + off + col) = f32_to_bf16x8(o);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < ITERS; ++i) {
        const int col = i * 256 + lane * 8;
        if (col < C) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { atomicAdd(&sh_dw[col + j], aw[i][j]); atomicAdd(&sh_db[col + j], ab[i][j]); }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < C; i += blockDim.x) { atomicAdd(&dw[i], sh_dw[i]); atomicAdd(&db[i], sh_db[i]); }
}

// Four warps per row (8 or 16 columns per lane): the dgamma / dbeta / column-sum partials shrink to 24 registers per
// thread, which leaves room to have the three input rows (dy, x, dres) in flight together and to keep 4 blocks (32
// warps) per SM.  Row sums cross the four warps through shared memory and one named barrier per row.
// dxsum (optional, fp32 [C]) += column sums of the bf16 dx that is written: the bias gradient of the layer below,
// which used to cost a separate pass over dx.
template <int IT>
__global__ void __launch_bounds__(256, IT == 1 ? 3 : 2) layernorm_bwd4_kernel(const __nv_bfloat16* __restrict__ dy,
        const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ mean,
        const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx,
        float* __restrict__ dw, float* __restrict__ db, float* __restrict__ dxsum, int M, int C) {
    extern __shared__ float sh[];             // [3][C] block accumulators (dw, db, dxsum)
    __shared__ float part[2][2][4][2];        // [row group][parity][warp][s1, s2]
    for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) sh[i] = 0.f;
    __syncthreads();
    pdl_enter();
    const int grp = threadIdx.x >> 7, t = threadIdx.x & 127, wq = t >> 5, lane = threadIdx.x & 31;
    float aw[IT][8], ab[IT][8], ao[IT][8], ww[IT][8];
#pragma unroll
    for (int i = 0; i < IT; ++i) {
        const int col = i * 1024 + t * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) { aw[i][j] = 0.f; ab[i][j] = 0.f; ao[i][j] = 0.f; ww[i][j] = 0.f; }
        if (col < C) {
            const float4 w0 = __ldg(reinterpret_cast<const float4*>(w + col)), w1 = __ldg(reinterpret_cast<const float4*>(w + col + 4));
            ww[i][0] = w0.x; ww[i][1] = w0.y; ww[i][2] = w0.z; ww[i][3] = w0.w;
            ww[i][4] = w1.x; ww[i][5] = w1.y; ww[i][6] = w1.z; ww[i][7] = w1.w;
        }
    }
    const float inv_c = 1.0f / C;
    int par = 0;
    for (int row = blockIdx.x * 2 + grp; row < M; row += gridDim.x * 2) {
        const long long off = (long long)row * C;
        uint4 rdy[IT], rx[IT], rr[IT];
#pragma unroll
        for (int i = 0; i < IT; ++i) {
            const int col = i * 1024 + t * 8;
            if (col < C) {
                rdy[i] = ld_nc_v4(dy + off + col);
                rx[i] = ld_nc_v4(x + off + col);
                if (dres != nullptr) rr[i] = ld_nc_v4(dres + off + col);
            }
        }
        const float mu = mean[row], rs = rstd[row];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < IT; ++i) {
            const int col = i * 1024 + t * 8;
            if (col < C) {
                float dyv[8], xv[8];
                bf16x8_to_f32(rdy[i], dyv);
                bf16x8_to_f32(rx[i], xv);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float xh = (xv[j] - mu) * rs, g = dyv[j] * ww[i][j];
                    s1 += g;
                    s2 += g * xh;
                    aw[i][j] += dyv[j] * xh;
                    ab[i][j] += dyv[j];
                }
            }
        }
        s1 = warp_sum(s1);
        s2 = warp_sum(s2);
        if (lane == 0) { part[grp][par][wq][0] = s1; part[grp][par][wq][1] = s2; }
        asm volatile("bar.sync %0, 128;" :: "r"(1 + grp) : "memory");
        s1 = (part[grp][par][0][0] + part[grp][par][1][0] + part[grp][par][2][0] + part[grp][par][3][0]) * inv_c;
        s2 = (part[grp][par][0][1] + part[grp][par][1][1] + part[grp][par][2][1] + part[grp][par][3][1]) * inv_c;
        par ^= 1;                                  // the other slot is free: everyone passed this row's barrier
#pragma unroll
        for (int i = 0; i < IT; ++i) {
            const int col = i * 1024 + t * 8;
            if (col < C) {
                float dyv[8], xv[8], o[8];
                bf16x8_to_f32(rdy[i], dyv);
                bf16x8_to_f32(rx[i], xv);
                if (dres != nullptr) bf16x8_to_f32(rr[i], o);
                else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) o[j] = 0.f;
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float xh = (xv[j] - mu) * rs;
                    o[j] += (dyv[j] * ww[i][j] - s1 - xh * s2) * rs;
                }
                const uint4 packed = f32_to_bf16x8(o);
                *reinterpret_cast<uint4*>(dx + off + col) = packed;
                if (dxsum != nullptr) {
                    float ob[8];
                    bf16x8_to_f32(packed, ob);
#pragma unroll
                    for (int j = 0; j < 8; ++j) ao[i][j] += ob[j];
                }
            }
        }
    }
    // block combine (two row groups) in shared memory, then one vector reduction per 4 columns
#pragma unroll
    for (int i = 0; i < IT; ++i) {
        const int col = i * 1024 + t * 8;
        if (col < C) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                atomicAdd(&sh[col + j], aw[i][j]);
                atomicAdd(&sh[C + col + j], ab[i][j]);
                if (dxsum != nullptr) atomicAdd(&sh[2 * C + col + j], ao[i][j]);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x * 4; i < C; i += blockDim.x * 4) {
        asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(dw + i), "f"(sh[i]), "f"(sh[i + 1]), "f"(sh[i + 2]), "f"(sh[i + 3]) : "memory");
        asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(db + i), "f"(sh[C + i]), "f"(sh[C + i + 1]), "f"(sh[C + i + 2]), "f"(sh[C + i + 3]) : "memory");
        if (dxsum != nullptr)
            asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(dxsum + i), "f"(sh[2 * C + i]), "f"(sh[2 * C + i + 1]), "f"(sh[2 * C + i + 2]), "f"(sh[2 * C + i + 3]) : "memory");
    }
}

// ------------------------------------------------------------------------------------ bias grad
// db[n] += sum_m dy[m, n];  block = 32 column-groups x 8 row lanes, 256 columns per block
__global__ void __launch_bounds__(256) bias_grad_kernel(const __nv_bfloat16* __restrict__ dy, float* __restrict__ db,
                                                        int M, int N, long long ld, int rows_per_block) {
    __shared__ float sh[8][256 + 8];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int col = blockIdx.x * 256 + tx * 8;
    const int r0 = blockIdx.y * rows_per_block, r1 = min(M, r0 + rows_per_block);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (col < N) {
        for (int r = r0 + ty; r < r1; r += 8) {
            float f[8];
            bf16x8_to_f32(ld_nc_v4(dy + (long long)r * ld + col), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += f[j];
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) sh[ty][tx * 8 + j] = acc[j];
    __syncthreads();
    const int c = threadIdx.x;
    if (blockIdx.x * 256 + c < N) {
        float s = 0.f;
#pragma unroll
        for (int y = 0; y < 8; ++y) s += sh[y][c];
        atomicAdd(&db[blockIdx.x * 256 + c], s);
    }
}

// ------------------------------------------------------------------------------------ embedding
__global__ void __launch_bounds__(128) embedding_fwd_kernel(const long long* __restrict__ idx,
        const __nv_bfloat16* __restrict__ wte, const __nv_bfloat16* __restrict__ wpe, __nv_bfloat16* __restrict__ out,
        int BT, int T, int C) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = blockIdx.x * 4 + warp;
    if (row >= BT) return;
    const long long tok = idx[row];
    const int pos = row % T;
    for (int col = lane * 8; col < C; col += 256) {
        float a[8], b[8];
        bf16x8_to_f32(*reinterpret_cast<const uint4*>(wte + tok * C + col), a);
        bf16x8_to_f32(*reinterpret_cast<const uint4*>(wpe + (long long)pos * C + col), b);
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] += b[j];
        *reinterpret_cast<uint4*>(out + (long long)row * C + col) = f32_to_bf16x8(a);
    }
}

__global__ void __launch_bounds__(128) embedding_bwd_kernel(const long long* __restrict__ idx,
        const __nv_bfloat16* __restrict__ dx, float* __restrict__ dwte, float* __restrict__ dwpe, int BT, int T, int C) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = blockIdx.x * 4 + warp;
    if (row >= BT) return;
    const long long tok = idx[row];
    const int pos = row % T;
    for (int col = lane * 8; col < C; col += 256) {
        float g[8];
        bf16x8_to_f32(*reinterpret_cast<const uint4*>(dx + (long long)row * C + col), g);
        float* a = dwte + tok * C + col;
        float* b = dwpe + (long long)pos * C + col;
        asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(a), "f"(g[0]), "f"(g[1]), "f"(g[2]), "f"(g[3]) : "memory");
        asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(a + 4), "f"(g[4]), "f"(g[5]), "f"(g[6]), "f"(g[7]) : "memory");
        asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(b), "f"(g[0]), "f"(g[1]), "f"(g[2]), "f"(g[3]) : "memory");
        asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(b + 4), "f"(g[4]), "f"(g[5]), "f"(g[6]), "f"(g[7]) : "memory");
    }
}

// ------------------------------------------------------------------------------------ AdamW
template <typename G>
__global__ void __launch_bounds__(256) adamw_kernel(float* __restrict__ p32, __nv_bfloat16* __restrict__ p16,
        const G* __restrict__ g, float* __restrict__ m, float* __restrict__ v, long long n, long long n_decay,
        float lr, float b1, float b2, float eps, float wd, float bc1, float bc2, const float* __restrict__ gscale,
        const float* __restrict__ hyper) {
    const float gs = gscale != nullptr ? *gscale : 1.0f;
    if (hyper != nullptr) { lr = hyper[0]; bc1 = hyper[1]; bc2 = hyper[2]; }
    const long long stride = (long long)gridDim.x * blockDim.x * 4;
    for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
        float gg[4];
        if constexpr (sizeof(G) == 4) {
            const float4 t = *reinterpret_cast<const float4*>(g + i);
            gg[0] = t.x; gg[1] = t.y; gg[2] = t.z; gg[3] = t.w;
        } else {
            const uint2 t = *reinterpret_cast<const uint2*>(g + i);
            const float2 a = unpack_bf16x2(t.x), b = unpack_bf16x2(t.y);
            gg[0] = a.x; gg[1] = a.y; gg[2] = b.x; gg[3] = b.y;
        }
        float4 pp = *reinterpret_cast<float4*>(p32 + i), mm = *reinterpret_cast<float4*>(m + i), vv = *reinterpret_cast<float4*>(v + i);
        float* P = reinterpret_cast<float*>(&pp); float* Mv = reinterpret_cast<float*>(&mm); float* V = reinterpret_cast<float*>(&vv);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float gj = gg[j] * gs;
            Mv[j] = b1 * Mv[j] + (1.f - b1) * gj;
            V[j] = b2 * V[j] + (1.f - b2) * gj * gj;
            const float mh = Mv[j] / bc1, vh = V[j] / bc2;
            const float decay = (i + j) < n_decay ? wd : 0.f;
            P[j] -= lr * (mh / (sqrtf(vh) + eps) + decay * P[j]);
        }
        *reinterpret_cast<float4*>(p32 + i) = pp;
        *reinterpret_cast<float4*>(m + i) = mm;
        *reinterpret_cast<float4*>(v + i) = vv;
        if (p16 != nullptr) {
            uint2 o; o.x = pack_bf16x2(P[0], P[1]); o.y = pack_bf16x2(P[2], P[3]);
            *reinterpret_cast<uint2*>(p16 + i) = o;
        }
    }
}

// ------------------------------------------------------------------------------------ cross entropy
// one CTA per row; logits bf16 [M, Vp] overwritten with dlogits; only columns < V are real.
// Pass 1: online (max, sum); pass 2 re-reads the row (mostly from L2) and writes the gradient.
// Both passes move XENT_U vectors per thread per trip so that enough loads are in flight (one 16-byte load per
// thread per trip left the kernel latency-bound at ~0.6 of the HBM rate).
constexpr int XENT_U = 4;
__device__ __forceinline__ uint4 xent_mask_tail(uint4 raw, int nv) {        // columns >= V of a vector become -inf
    uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        if (2 * q >= nv) w[q] = 0xff80ff80u;
        else if (2 * q + 1 >= nv) w[q] = (w[q] & 0xffffu) | 0xff800000u;
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}
__global__ void __launch_bounds__(1024) xent_kernel(__nv_bfloat16* __restrict__ logits, const long long* __restrict__ targets,
        float* __restrict__ row_loss, float* __restrict__ loss_sum, int V, int Vp, float grad_scale) {
    __shared__ float shm[32], shs[32];
    const int row = blockIdx.x;
    __nv_bfloat16* lr = logits + (long long)row * Vp;
    const int nvec = Vp / 8;
    const uint4 ninf = make_uint4(0xff80ff80u, 0xff80ff80u, 0xff80ff80u, 0xff80ff80u);
    float mx = -INFINITY, sum = 0.f;
    for (int base = threadIdx.x; base < nvec; base += blockDim.x * XENT_U) {
        uint4 raw[XENT_U];
#pragma unroll
        for (int u = 0; u < XENT_U; ++u) {
            const int i = base + u * blockDim.x;
            raw[u] = i < nvec ? ld_nc_v4(lr + i * 8) : ninf;
        }
#pragma unroll
        for (int u = 0; u < XENT_U; ++u) {
            const int i = base + u * blockDim.x;
            if (i < nvec && i * 8 + 8 > V) raw[u] = xent_mask_tail(raw[u], V - i * 8);
            float f[8];
            bf16x8_to_f32(raw[u], f);
            float lm = fmaxf(fmaxf(fmaxf(f[0], f[1]), fmaxf(f[2], f[3])), fmaxf(fmaxf(f[4], f[5]), fmaxf(f[6], f[7])));
            if (lm > mx) { sum *= __expf(mx - lm); mx = lm; }
            if (mx != -INFINITY) {
#pragma unroll
                for (int j = 0; j < 8; ++j) sum += __expf(f[j] - mx);      // exp(-inf) = 0 for masked columns
            }
        }
    }
    // block combine of (max, sum)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float wm = warp_max(mx);
    sum = warp_sum(mx == -INFINITY ? 0.f : sum * __expf(mx - wm));      // idle lanes (Vp/8 < threads) hold -inf
    if (lane == 0) { shm[warp] = wm; shs[warp] = sum; }
    __syncthreads();
    float bm = -INFINITY, bs = 0.f;
    for (int w2 = 0; w2 < (int)(blockDim.x >> 5); ++w2) bm = fmaxf(bm, shm[w2]);
    for (int w2 = 0; w2 < (int)(blockDim.x >> 5); ++w2) bs += shm[w2] == -INFINITY ? 0.f : shs[w2] * __expf(shm[w2] - bm);
    const long long tgt = targets[row];
    const float inv = grad_scale / bs;
    if (threadIdx.x == 0) {
        const float loss = logf(bs) + bm - __bfloat162float(lr[tgt]);
        if (row_loss != nullptr) row_loss[row] = loss;
        if (loss_sum != nullptr) atomicAdd(loss_sum, loss);
    }
    __syncthreads();       // the target logit is read before anyone overwrites it
    const int tvec = (int)(tgt >> 3), tj = (int)(tgt & 7);
    for (int base = threadIdx.x; base < nvec; base += blockDim.x * XENT_U) {
        uint4 raw[XENT_U];
#pragma unroll
        for (int u = 0; u < XENT_U; ++u) {
            const int i = base + u * blockDim.x;
            raw[u] = i < nvec ? *reinterpret_cast<const uint4*>(lr + i * 8) : ninf;
        }
#pragma unroll
        for (int u = 0; u < XENT_U; ++u) {
            const int i = base + u * blockDim.x;
            if (i < nvec) {
                if (i * 8 + 8 > V) raw[u] = xent_mask_tail(raw[u], V - i * 8);
                float f[8];
                bf16x8_to_f32(raw[u], f);
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = __expf(f[j] - bm) * inv;     // padding: exp(-inf) = 0
                if (i == tvec) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) if (j == tj) f[j] -= grad_scale;
                }
                *reinterpret_cast<uint4*>(lr + i * 8) = f32_to_bf16x8(f);
            }
        }
    }
}

// Cluster variant (default when the row fits): one row per cluster of 4 CTAs.  Each CTA keeps its quarter of the bf16 row
// in shared memory (25 KB for the GPT-2 vocabulary, so 6-8 CTAs stay resident per SM), the four partial (max, sum) pairs
// are exchanged through distributed shared memory, and the gradient is produced from the cached quarter: the logits are
// read from HBM once and written once.  (The streaming kernel's second pass was measured to miss L2: with enough rows in
// flight to hide latency the working set is ~100 MB.)
constexpr int XENT_CL = 4;
__device__ __forceinline__ uint32_t xent_cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void xent_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float xent_ld_peer(const float* local_smem, uint32_t peer) {   // the same variable in CTA `peer` of the cluster
    uint32_t a;
    float v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(a) : "r"(smem_u32(local_smem)), "r"(peer));
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(a) : "memory");
    return v;
}
__global__ void __cluster_dims__(XENT_CL, 1, 1) __launch_bounds__(256)
xent_cluster_kernel(__nv_bfloat16* __restrict__ logits, const long long* __restrict__ targets, float* __restrict__ row_loss,
                    float* __restrict__ loss_sum, int V, int Vp, int per, float grad_scale) {
    extern __shared__ uint4 srow[];            // this CTA's `per` vectors of the row
    __shared__ float shm[8], shs[8];
    __shared__ float stat[2];                  // this CTA's (max, sum), read by the peers
    const uint32_t rank = xent_cluster_rank();
    const int row = blockIdx.x / XENT_CL;
    __nv_bfloat16* lr = logits + (long long)row * Vp;
    const int nvec = Vp / 8;
    const int v0 = (int)rank * per, v1 = min(nvec, v0 + per);
    float mx = -INFINITY, sum = 0.f;
    for (int base = v0 + threadIdx.x; base < v1; base += 256 * XENT_U) {
        uint4 raw[XENT_U];
#pragma unroll
        for (int u = 0; u < XENT_U; ++u) {
            const int i = base + u * 256;
            if (i < v1) raw[u] = ld_nc_v4(lr + i * 8);
        }
#pragma unroll
        for (int u = 0; u < XENT_U; ++u) {
            const int i = base + u * 256;
            if (i < v1) {
                if (i * 8 + 8 > V) raw[u] = xent_mask_tail(raw[u], V - i * 8);
                srow[i - v0] = raw[u];
                float f[8];
                bf16x8_to_f32(raw[u], f);
                const float lm = fmaxf(fmaxf(fmaxf(f[0], f[1]), fmaxf(f[2], f[3])), fmaxf(fmaxf(f[4], f[5]), fmaxf(f[6], f[7])));
                if (lm > mx) { sum *= __expf(mx - lm); mx = lm; }
                if (mx != -INFINITY) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) sum += __expf(f[j] - mx);
                }
            }
        }
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float wm = warp_max(mx);
    sum = warp_sum(mx == -INFINITY ? 0.f : sum * __expf(mx - wm));
    if (lane == 0) { shm[warp] = wm; shs[warp] = sum; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float m = -INFINITY, s2 = 0.f;
        for (int w2 = 0; w2 < 8; ++w2) m = fmaxf(m, shm[w2]);
        for (int w2 = 0; w2 < 8; ++w2) s2 += shm[w2] == -INFINITY ? 0.f : shs[w2] * __expf(shm[w2] - m);
        stat[0] = m; stat[1] = s2;
    }
    xent_cluster_sync();                       // all four (max, sum) pairs are published (also orders this CTA's srow writes)
    float pm[XENT_CL], ps[XENT_CL];
#pragma unroll
    for (int q = 0; q < XENT_CL; ++q) { pm[q] = xent_ld_peer(&stat[0], q); ps[q] = xent_ld_peer(&stat[1], q); }
    float bm = -INFINITY, bs = 0.f;
#pragma unroll
    for (int q = 0; q < XENT_CL; ++q) bm = fmaxf(bm, pm[q]);
#pragma unroll
    for (int q = 0; q < XENT_CL; ++q) bs += pm[q] == -INFINITY ? 0.f : ps[q] * __expf(pm[q] - bm);
    const long long tgt = targets[row];
    const int tvec = (int)(tgt >> 3), tj = (int)(tgt & 7);
    const float inv = grad_scale / bs;
    if (threadIdx.x == 0 && tvec >= v0 && tvec < v1) {         // the CTA that holds the target logit reports the loss
        const float lt = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(srow)[tgt - (long long)v0 * 8]);
        const float loss = logf(bs) + bm - lt;
        if (row_loss != nullptr) row_loss[row] = loss;
        if (loss_sum != nullptr) atomicAdd(loss_sum, loss);
    }
    for (int i = v0 + threadIdx.x; i < v1; i += 256) {
        float f[8];
        bf16x8_to_f32(srow[i - v0], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = __expf(f[j] - bm) * inv;          // padding: exp(-inf) = 0
        if (i == tvec) {
#pragma unroll
            for (int j = 0; j < 8; ++j) if (j == tj) f[j] -= grad_scale;
        }
        *reinterpret_cast<uint4*>(lr + i * 8) = f32_to_bf16x8(f);
    }
    xent_cluster_sync();                       // nobody exits while a peer may still read its stat[]
}

// ------------------------------------------------------------------------------------ K1 pack + norm
__global__ void __launch_bounds__(256) pack_norm_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                                        long long n, double* __restrict__ sumsq, const float* __restrict__ gscale) {
    __shared__ double sh[8];
    const float gs = gscale != nullptr ? *gscale : 1.0f;      // fault-injection hook: a poisoned rank scales / flips its bucket
    float acc = 0.f;
    double dacc = 0.0;
    int run = 0;
    const long long stride = (long long)gridDim.x * blockDim.x * 8;
    for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8; i < n; i += stride) {
        const float4 a = *reinterpret_cast<const float4*>(src + i), b = *reinterpret_cast<const float4*>(src + i + 4);
        const float f[8] = {a.x * gs, a.y * gs, a.z * gs, a.w * gs, b.x * gs, b.y * gs, b.z * gs, b.w * gs};
#pragma unroll
        for (int j = 0; j < 8; ++j) acc += f[j] * f[j];
        *reinterpret_cast<uint4*>(dst + i) = f32_to_bf16x8(f);
        if (++run == 16) { dacc += acc; acc = 0.f; run = 0; }
    }
    dacc += acc;
    for (int o = 16; o > 0; o >>= 1) dacc += __shfl_xor_sync(0xffffffffu, dacc, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = dacc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += sh[w];
        atomicAdd(sumsq, t);
    }
}

// sum of squares of an already packed bucket (the eager data-parallel path writes gradients straight into the bucket)
template <typename T>
__global__ void __launch_bounds__(256) sumsq_kernel(const T* __restrict__ x, long long n, double* __restrict__ sumsq) {
    __shared__ double sh[8];
    float acc = 0.f;
    double dacc = 0.0;
    int run = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float v = to_f32<T>(x[i]);
        acc = fmaf(v, v, acc);
        if (++run == 64) { dacc += acc; acc = 0.f; run = 0; }
    }
    dacc += acc;
    for (int o = 16; o > 0; o >>= 1) dacc += __shfl_xor_sync(0xffffffffu, dacc, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = dacc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += sh[w];
        atomicAdd(sumsq, t);
    }
}

// ------------------------------------------------------------------------------------ multi-tensor sq norm
template <typename T>
__global__ void __launch_bounds__(256) multi_sqnorm_kernel(const long long* __restrict__ ptrs, const long long* __restrict__ numels,
                                                           float* __restrict__ out) {
    __shared__ float sh[8];
    const T* x = reinterpret_cast<const T*>(ptrs[blockIdx.y]);
    const long long n = numels[blockIdx.y];
    float acc = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float v = to_f32<T>(x[i]);
        acc += v * v;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0;
        for (int w = 0; w < 8; ++w) t += sh[w];
        if (t != 0.f) atomicAdd(&out[blockIdx.y], t);
    }
}

// ------------------------------------------------------------------------------------ integrity
__global__ void __launch_bounds__(256) checksum_kernel(const uint32_t* __restrict__ w, long long n, unsigned long long* __restrict__ out) {
    __shared__ uint32_t s1s[8], s2s[8];
    uint32_t s1 = 0, s2 = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const uint32_t v = w[i];
        s1 += v;
        s2 += static_cast<uint32_t>((n - i) & 0xFFFFFFFFll) * v;       // all arithmetic mod 2^32
    }
    for (int o = 16; o > 0; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
    if ((threadIdx.x & 31) == 0) { s1s[threadIdx.x >> 5] = s1; s2s[threadIdx.x >> 5] = s2; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t a = 0, b = 0;
        for (int k = 0; k < 8; ++k) { a += s1s[k]; b += s2s[k]; }
        atomicAdd(&out[0], (unsigned long long)a);
        atomicAdd(&out[1], (unsigned long long)b);
    }
}

template <typename T>
__global__ void __launch_bounds__(256) drift_kernel(const T* __restrict__ cur, T* __restrict__ prev, long long n, float* __restrict__ out) {
    __shared__ float sh[8];
    float acc = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float c = to_f32<T>(cur[i]), d = c - to_f32<T>(prev[i]);
        acc += d * d;
        prev[i] = cur[i];
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0;
        for (int w = 0; w < 8; ++w) t += sh[w];
        atomicAdd(out, t);
    }
}

// ------------------------------------------------------------------------------------ fault injection
__device__ __forceinline__ uint32_t mix32(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return static_cast<uint32_t>((z ^ (z >> 31)) >> 16);
}
template <typename T>
__global__ void __launch_bounds__(256) poison_kernel(T* __restrict__ x, long long n, float a, float b, unsigned long long seed) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float v = a * to_f32<T>(x[i]);
        if (b != 0.f) {
            const float u1 = (mix32(seed * 0x100000001B3ull + 2 * i) + 1.0f) * (1.0f / 4294967808.0f);
            const float u2 = mix32(seed * 0x100000001B3ull + 2 * i + 1) * (1.0f / 4294967296.0f);
            v += b * sqrtf(-2.0f * __logf(u1)) * __cosf(6.283185307f * u2);
        }
        x[i] = from_f32<T>(v);
    }
}

// parameter-poisoning fault injection that can sit inside a captured step: does nothing while *level == 0,
// otherwise adds level * N(0,1) to every element (a compromised owner corrupting the shard it publishes)
__global__ void __launch_bounds__(256) poison_if_kernel(__nv_bfloat16* __restrict__ x, long long n, const float* __restrict__ level,
                                                        unsigned long long seed) {
    const float b = *level;
    if (b == 0.f) return;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float u1 = (mix32(seed * 0x100000001B3ull + 2 * i) + 1.0f) * (1.0f / 4294967808.0f);
        const float u2 = mix32(seed * 0x100000001B3ull + 2 * i + 1) * (1.0f / 4294967296.0f);
        x[i] = __float2bfloat16(__bfloat162float(x[i]) + b * sqrtf(-2.0f * __logf(u1)) * __cosf(6.283185307f * u2));
    }
}

__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
    for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n; i += (long long)gridDim.x * blockDim.x * 4) {
        const float4 a = *reinterpret_cast<const float4*>(src + i);
        uint2 o; o.x = pack_bf16x2(a.x, a.y); o.y = pack_bf16x2(a.z, a.w);
        *reinterpret_cast<uint2*>(dst + i) = o;
    }
}

__global__ void __launch_bounds__(256) cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        dst[i] = __bfloat162float(src[i]);
}

// advances the optimizer step on the device and refreshes (lr, 1-b1^t, 1-b2^t): graph-replay safe
__global__ void adam_tick_kernel(float* hyper, float lr, float b1, float b2) {
    if (threadIdx.x == 0) {
        const float step = hyper[3] + 1.0f;
        hyper[3] = step;
        hyper[0] = lr;
        hyper[1] = 1.0f - powf(b1, step);
        hyper[2] = 1.0f - powf(b2, step);
    }
}

// K7 second half: out[r, c] = sum_s staging[s][r][c] (+ residual[r, c])   (bf16 in/out, fp32 sum)
__global__ void __launch_bounds__(256) reduce_slots_kernel(const __nv_bfloat16* __restrict__ staging, int W, long long slot_elems,
        const __nv_bfloat16* __restrict__ residual, __nv_bfloat16* __restrict__ out) {
    for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8; i < slot_elems; i += (long long)gridDim.x * blockDim.x * 8) {
        float acc[8];
        if (residual != nullptr) bf16x8_to_f32(*reinterpret_cast<const uint4*>(residual + i), acc);
        else {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        }
        for (int s2 = 0; s2 < W; ++s2) {
            float f[8];
            bf16x8_to_f32(ld_nc_v4(staging + s2 * slot_elems + i), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += f[j];
        }
        *reinterpret_cast<uint4*>(out + i) = f32_to_bf16x8(acc);
    }
}

// reduce-scatter epilogue of a row-parallel linear: out[r, c] = sum_slots staging[s][r][c] + bias[c] + residual[r, c]
__global__ void __launch_bounds__(256) reduce_slots_bias_kernel(const __nv_bfloat16* __restrict__ staging, int W, long long slot_elems,
        int cols, const float* __restrict__ bias, const __nv_bfloat16* __restrict__ residual, __nv_bfloat16* __restrict__ out) {
    for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8; i < slot_elems; i += (long long)gridDim.x * blockDim.x * 8) {
        float acc[8];
        if (residual != nullptr) bf16x8_to_f32(*reinterpret_cast<const uint4*>(residual + i), acc);
        else {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        }
        if (bias != nullptr) {
            const int c = (int)(i % cols);
            const float4 b0 = *reinterpret_cast<const float4*>(bias + c), b1 = *reinterpret_cast<const float4*>(bias + c + 4);
            acc[0] += b0.x; acc[1] += b0.y; acc[2] += b0.z; acc[3] += b0.w; acc[4] += b1.x; acc[5] += b1.y; acc[6] += b1.z; acc[7] += b1.w;
        }
        for (int s2 = 0; s2 < W; ++s2) {
            float f[8];
            bf16x8_to_f32(ld_nc_v4(staging + s2 * slot_elems + i), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += f[j];
        }
        *reinterpret_cast<uint4*>(out + i) = f32_to_bf16x8(acc);
    }
}

// block size / grid of the memory-bound helper kernels (pack + norm, sum of squares, AdamW): 256 threads, <= 8 blocks per SM
inline int helper_threads() { return 256; }
inline int helper_grid(long long n, int per_thread) {
    long long b = (n + 256ll * per_thread - 1) / (256ll * per_thread);
    const long long cap = (long long)tdl_num_sms() * 8;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

inline int grid_for(long long n, int per_block) {
    long long b = (n + per_block - 1) / per_block;
    const long long cap = (long long)tdl_num_sms() * 8;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace

#define STREAM(s) reinterpret_cast<cudaStream_t>(s)

extern "C" int tdl_layernorm_fwd(const void* x, const float* w, const float* b, void* y, float* mean, float* rstd,
                                 int64_t M, int64_t C, float eps, void* stream) {
    TDL_REQUIRE(C % 8 == 0 && C <= 256 * LN_MAX_ITERS, "LayerNorm width %lld unsupported (need %%8==0, <= %d)", (long long)C, 256 * LN_MAX_ITERS);
#define LN_FWD(I) TDL_CHECK(tdl_launch_pdl(layernorm_fwd_kernel<I>, dim3((unsigned)((M + 3) / 4)), dim3(128), 0, STREAM(stream), \
        (const __nv_bfloat16*)x, w, b, (__nv_bfloat16*)y, mean, rstd, (int)M, (int)C, eps))
    switch ((C + 255) / 256) {
        case 1: LN_FWD(1); break; case 2: LN_FWD(2); break; case 3: LN_FWD(3); break; case 4: LN_FWD(4); break;
        case 5: LN_FWD(5); break; case 6: LN_FWD(6); break; case 7: LN_FWD(7); break; default: LN_FWD(8); break;
    }
#undef LN_FWD
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_bias_grad(const void* dy, float* db, int64_t M, int64_t N, int64_t ld, void* stream);

// dxsum (optional, fp32 [C], 16-byte aligned like dw / db): += column sums of the bf16 dx (bias gradient of the layer below)
extern "C" int tdl_layernorm_bwd_colsum(const void* dy, const void* x, const float* w, const float* mean, const float* rstd,
                                        const void* dres, void* dx, float* dw, float* db, float* dxsum, int64_t M, int64_t C,
                                        void* stream) {
    TDL_REQUIRE(C % 8 == 0 && C <= 256 * LN_MAX_ITERS, "LayerNorm width %lld unsupported", (long long)C);
    static int variant = -1;                   // TDL_LN_BWD=1: one warp per row (previous kernel)
    if (variant < 0) { const char* e = getenv("TDL_LN_BWD"); variant = e ? atoi(e) : 4; }
    const bool aligned = ((reinterpret_cast<uintptr_t>(dw) | reinterpret_cast<uintptr_t>(db) | reinterpret_cast<uintptr_t>(dxsum)) & 15) == 0;
    if (variant == 4 && C % 4 == 0 && aligned) {
        const int grid = (int)std::min<long long>((M + 1) / 2, (long long)tdl_num_sms() * (C <= 1024 ? 3 : 2));   // one resident wave
        const size_t smem = 3 * C * sizeof(float);
#define LN_BWD4(I) TDL_CHECK(tdl_launch_pdl(layernorm_bwd4_kernel<I>, dim3(grid), dim3(256), smem, STREAM(stream), \
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, w, mean, rstd, (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, \
        dw, db, dxsum, (int)M, (int)C))
        if (C <= 1024) LN_BWD4(1); else LN_BWD4(2);
#undef LN_BWD4
        TDL_LAUNCH_CHECK();
        return 0;
    }
    const int grid = (int)std::min<long long>((M + 7) / 8, (long long)tdl_num_sms() * 2);   // one resident wave: fewer dgamma/dbeta atomics
#define LN_BWD(I) layernorm_bwd_kernel<I><<<grid, 256, 2 * C * sizeof(float), STREAM(stream)>>>( \
        (const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, w, mean, rstd, (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, \
        dw, db, (int)M, (int)C)
    switch ((C + 255) / 256) {
        case 1: LN_BWD(1); break; case 2: LN_BWD(2); break; case 3: LN_BWD(3); break; case 4: LN_BWD(4); break;
        case 5: LN_BWD(5); break; case 6: LN_BWD(6); break; case 7: LN_BWD(7); break; default: LN_BWD(8); break;
    }
#undef LN_BWD
    TDL_LAUNCH_CHECK();
    if (dxsum != nullptr) return tdl_bias_grad(dx, dxsum, M, C, C, stream);
    return 0;
}

extern "C" int tdl_layernorm_bwd(const void* dy, const void* x, const float* w, const float* mean, const float* rstd,
                                 const void* dres, void* dx, float* dw, float* db, int64_t M, int64_t C, void* stream) {
    return tdl_layernorm_bwd_colsum(dy, x, w, mean, rstd, dres, dx, dw, db, nullptr, M, C, stream);
}

extern "C" int tdl_bias_grad(const void* dy, float* db, int64_t M, int64_t N, int64_t ld, void* stream) {
    TDL_REQUIRE(N % 8 == 0 && ld % 8 == 0, "bias_grad needs N, ld multiples of 8");
    const int gx = (int)((N + 255) / 256);
    int gy = std::max(1, (tdl_num_sms() * 4) / gx);
    int rows = (int)((M + gy - 1) / gy);
    rows = (rows + 7) / 8 * 8;
    gy = (int)((M + rows - 1) / rows);
    bias_grad_kernel<<<dim3(gx, gy), 256, 0, STREAM(stream)>>>((const __nv_bfloat16*)dy, db, (int)M, (int)N, ld, rows);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_embedding_fwd(const int64_t* idx, const void* wte, const void* wpe, void* out, int64_t BT, int64_t T,
                                 int64_t C, void* stream) {
    TDL_REQUIRE(C % 8 == 0, "embedding width must be a multiple of 8");
    embedding_fwd_kernel<<<(unsigned)((BT + 3) / 4), 128, 0, STREAM(stream)>>>(
        (const long long*)idx, (const __nv_bfloat16*)wte, (const __nv_bfloat16*)wpe, (__nv_bfloat16*)out, (int)BT, (int)T, (int)C);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_embedding_bwd(const int64_t* idx, const void* dx, float* dwte, float* dwpe, int64_t BT, int64_t T,
                                 int64_t C, void* stream) {
    TDL_REQUIRE(C % 8 == 0, "embedding width must be a multiple of 8");
    embedding_bwd_kernel<<<(unsigned)((BT + 3) / 4), 128, 0, STREAM(stream)>>>(
        (const long long*)idx, (const __nv_bfloat16*)dx, dwte, dwpe, (int)BT, (int)T, (int)C);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_adamw(float* p32, void* p16, const void* grad, int grad_dtype, float* m, float* v, int64_t n,
                         int64_t n_decay, float lr, float beta1, float beta2, float eps, float wd, int64_t step,
                         const float* grad_scale, const float* hyper_dev, void* stream) {
    TDL_REQUIRE(n % 4 == 0, "AdamW length must be a multiple of 4 (pad the flat buffer)");
    const float bc1 = 1.0f - powf(beta1, (float)step), bc2 = 1.0f - powf(beta2, (float)step);
    const int grid = helper_grid(n, 4 * 4), th = helper_threads();
    if (grad_dtype == TDL_F32)
        adamw_kernel<float><<<grid, th, 0, STREAM(stream)>>>(p32, (__nv_bfloat16*)p16, (const float*)grad, m, v, n, n_decay,
                                                              lr, beta1, beta2, eps, wd, bc1, bc2, grad_scale, hyper_dev);
    else if (grad_dtype == TDL_BF16)
        adamw_kernel<__nv_bfloat16><<<grid, th, 0, STREAM(stream)>>>(p32, (__nv_bfloat16*)p16, (const __nv_bfloat16*)grad, m, v, n,
                                                                      n_decay, lr, beta1, beta2, eps, wd, bc1, bc2, grad_scale, hyper_dev);
    else return tdl_set_error("AdamW grads must be fp32 or bf16");
    TDL_LAUNCH_CHECK();
    return 0;
}

static int xent_threads() {          // TDL_XENT_THREADS=1024: fewer rows in flight (more of the re-read hits L2)
    static int v = -1;
    if (v < 0) { const char* e = getenv("TDL_XENT_THREADS"); v = e ? atoi(e) : 256; if (v != 512 && v != 1024) v = 256; }
    return v;
}

extern "C" int tdl_xent_fwd_bwd(void* logits, const int64_t* targets, float* row_loss, float* loss_sum, int64_t M,
                                int64_t V, int64_t Vp, float grad_scale, void* stream) {
    TDL_REQUIRE(Vp % 8 == 0 && V <= Vp, "padded vocab must be a multiple of 8 and >= vocab");
    static int variant = -1;                   // TDL_XENT=0: streaming kernel (two reads of the logits)
    if (variant < 0) { const char* e = getenv("TDL_XENT"); variant = e ? atoi(e) : 1; }
    const int nvec = (int)(Vp / 8);
    const int per = (nvec + XENT_CL - 1) / XENT_CL;
    const size_t smem = (size_t)per * 16;
    if (variant == 1 && smem <= 64 * 1024 && M * XENT_CL < (1ll << 31)) {
        static bool configured = false;
        if (!configured) {
            TDL_CHECK(cudaFuncSetAttribute(xent_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
            configured = true;
        }
        xent_cluster_kernel<<<(unsigned)(M * XENT_CL), 256, smem, STREAM(stream)>>>((__nv_bfloat16*)logits, (const long long*)targets,
                                                                                  row_loss, loss_sum, (int)V, (int)Vp, per, grad_scale);
    } else {
        xent_kernel<<<(unsigned)M, xent_threads(), 0, STREAM(stream)>>>((__nv_bfloat16*)logits, (const long long*)targets, row_loss,
                                                                        loss_sum, (int)V, (int)Vp, grad_scale);
    }
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_pack_norm(const float* src, void* dst_bf16, int64_t n, double* sumsq, const float* scale_dev, void* stream) {
    TDL_REQUIRE(n % 8 == 0, "bucket length must be a multiple of 8");
    pack_norm_kernel<<<helper_grid(n, 8 * 4), helper_threads(), 0, STREAM(stream)>>>(src, (__nv_bfloat16*)dst_bf16, n, sumsq, scale_dev);
    TDL_LAUNCH_CHECK();
    return 0;
}

// sumsq (device double) += sum x^2 ; the cross-rank exchange re-zeroes the accumulator after publishing it
extern "C" int tdl_sumsq(const void* x, int dtype, int64_t n, double* sumsq, void* stream) {
    const int grid = helper_grid(n, 16), th = helper_threads();
    if (dtype == TDL_F32) sumsq_kernel<float><<<grid, th, 0, STREAM(stream)>>>((const float*)x, n, sumsq);
    else if (dtype == TDL_BF16) sumsq_kernel<__nv_bfloat16><<<grid, th, 0, STREAM(stream)>>>((const __nv_bfloat16*)x, n, sumsq);
    else if (dtype == TDL_F16) sumsq_kernel<__half><<<grid, th, 0, STREAM(stream)>>>((const __half*)x, n, sumsq);
    else return tdl_set_error("unsupported dtype %d", dtype);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_multi_sqnorm(const int64_t* ptrs_dev, const int64_t* numels_dev, int num, int dtype, float* out,
                                void* stream) {
    if (num <= 0) return 0;
    dim3 grid(32, num);
    if (dtype == TDL_F32) multi_sqnorm_kernel<float><<<grid, 256, 0, STREAM(stream)>>>((const long long*)ptrs_dev, (const long long*)numels_dev, out);
    else if (dtype == TDL_BF16) multi_sqnorm_kernel<__nv_bfloat16><<<grid, 256, 0, STREAM(stream)>>>((const long long*)ptrs_dev, (const long long*)numels_dev, out);
    else if (dtype == TDL_F16) multi_sqnorm_kernel<__half><<<grid, 256, 0, STREAM(stream)>>>((const long long*)ptrs_dev, (const long long*)numels_dev, out);
    else return tdl_set_error("unsupported dtype %d", dtype);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_checksum64(const void* words, int64_t nwords, void* out2_u64, void* stream) {
    checksum_kernel<<<grid_for(nwords, 256 * 16), 256, 0, STREAM(stream)>>>((const uint32_t*)words, nwords, (unsigned long long*)out2_u64);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_drift_update(const void* cur, void* prev, int64_t n, int dtype, float* out, void* stream) {
    const int grid = grid_for(n, 256 * 8);
    if (dtype == TDL_F32) drift_kernel<float><<<grid, 256, 0, STREAM(stream)>>>((const float*)cur, (float*)prev, n, out);
    else if (dtype == TDL_BF16) drift_kernel<__nv_bfloat16><<<grid, 256, 0, STREAM(stream)>>>((const __nv_bfloat16*)cur, (__nv_bfloat16*)prev, n, out);
    else return tdl_set_error("unsupported dtype %d", dtype);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_poison_inplace(void* x, int64_t n, int dtype, float a, float b, int64_t seed, void* stream) {
    const int grid = grid_for(n, 256 * 8);
    if (dtype == TDL_F32) poison_kernel<float><<<grid, 256, 0, STREAM(stream)>>>((float*)x, n, a, b, (unsigned long long)seed);
    else if (dtype == TDL_BF16) poison_kernel<__nv_bfloat16><<<grid, 256, 0, STREAM(stream)>>>((__nv_bfloat16*)x, n, a, b, (unsigned long long)seed);
    else if (dtype == TDL_F16) poison_kernel<__half><<<grid, 256, 0, STREAM(stream)>>>((__half*)x, n, a, b, (unsigned long long)seed);
    else return tdl_set_error("unsupported dtype %d", dtype);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_poison_if(void* x_bf16, int64_t n, const float* level_dev, int64_t seed, void* stream) {
    poison_if_kernel<<<grid_for(n, 256 * 8), 256, 0, STREAM(stream)>>>((__nv_bfloat16*)x_bf16, n, level_dev, (unsigned long long)seed);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream) {
    TDL_REQUIRE(n % 4 == 0, "length must be a multiple of 4");
    cast_f32_bf16_kernel<<<grid_for(n, 256 * 4 * 4), 256, 0, STREAM(stream)>>>(src, (__nv_bfloat16*)dst, n);
    TDL_LAUNCH_CHECK();
    return 0;
}

// hyper (device float[4]): lr, 1-beta1^t, 1-beta2^t, t.  One tick per optimizer step.
extern "C" int tdl_adam_tick(float* hyper_dev, float lr, float beta1, float beta2, void* stream) {
    adam_tick_kernel<<<1, 32, 0, STREAM(stream)>>>(hyper_dev, lr, beta1, beta2);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_reduce_slots(const void* staging, int W, int64_t slot_elems, const void* residual, void* out, void* stream) {
    TDL_REQUIRE(slot_elems % 8 == 0, "slot size must be a multiple of 8 elements");
    reduce_slots_kernel<<<grid_for(slot_elems, 256 * 8 * 2), 256, 0, STREAM(stream)>>>((const __nv_bfloat16*)staging, W, slot_elems,
                                                                                    (const __nv_bfloat16*)residual, (__nv_bfloat16*)out);
    TDL_LAUNCH_CHECK();
    return 0;
}

// as tdl_reduce_slots for a [rows, cols] slot, plus the (replicated, fp32) bias of the row-parallel layer
extern "C" int tdl_reduce_slots_bias(const void* staging, int W, int64_t rows, int64_t cols, const float* bias, const void* residual,
                                     void* out, void* stream) {
    TDL_REQUIRE(cols % 8 == 0, "columns must be a multiple of 8");
    const long long n = rows * cols;
    reduce_slots_bias_kernel<<<grid_for(n, 256 * 8 * 2), 256, 0, STREAM(stream)>>>((const __nv_bfloat16*)staging, W, n, (int)cols, bias,
                                                                                (const __nv_bfloat16*)residual, (__nv_bfloat16*)out);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_cast_bf16_f32(const void* src, float* dst, int64_t n, void* stream) {
    if (n <= 0) return 0;
    cast_bf16_f32_kernel<<<grid_for(n, 256 * 4), 256, 0, STREAM(stream)>>>((const __nv_bfloat16*)src, dst, n);
    TDL_LAUNCH_CHECK();
    return 0;
}

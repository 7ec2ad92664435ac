// K6: persistent, warp-specialised bf16 GEMM on 5th-gen tensor cores (tcgen05.mma, accumulators in
// TMEM, operands staged by TMA with 128B swizzle) with fused epilogues.
//
//   C[M,N] (+)= op(A)[M,K] * op(B)[N,K]^T      fp32 accumulate
//
// Operand layouts (so forward, dgrad and wgrad all run without materialising a transpose):
//   a_mn_major = 0 : A stored [M, K] row-major (K contiguous)       "K-major"
//   a_mn_major = 1 : A stored [K, M] row-major (M contiguous)       "MN-major"  (wgrad: dY^T)
//   b_mn_major = 0 : B stored [N, K] row-major (nn.Linear weight)   "K-major"
//   b_mn_major = 1 : B stored [K, N] row-major                      "MN-major"  (dgrad: W, wgrad: X)
//
// Epilogues (runtime flags): +bias[n], GELU(tanh) with the pre-activation saved to aux_out,
// +residual[m,n], *GELU'(aux_in[m,n]) (dgrad through the MLP activation), bf16 store, or
// split-K fp32 accumulation with red.global.add (wgrad into the fp32 main-grad buffer).
//
// CTA = 320 threads: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane) + TMEM owner,
// warps 2..9 = epilogue (each owns one 32-lane TMEM quarter x one column half).  Two accumulator stages in TMEM
// (2 x BLOCK_N columns) let the epilogue of tile i overlap the MMAs of tile i+1.
// The reference has no GEMM of its own: it would call cuBLAS through torch.
#include "common.cuh"
#include <cstdlib>

extern "C" int tdl_bias_grad(const void* dy, float* db, int64_t M, int64_t N, int64_t ld, void* stream);   // elementwise.cu

__device__ __forceinline__ void red_add_v4_f32(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;                 // 64 bf16 = 128 B = one swizzle row
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 320;            // TMA warp + MMA warp + 8 epilogue warps

enum Epilogue { EPI_NONE = 0, EPI_GELU = 1, EPI_RESIDUAL = 2, EPI_DGELU = 3,
                EPI_GELU_DG = 4,     // out = gelu(pre), aux_out = gelu'(pre)   (forward of a layer whose backward uses EPI_MUL)
                EPI_MUL = 5 };       // out = acc * aux_in

struct GemmParams {
    int M, N, K;
    int m_blocks, n_blocks, k_blocks_per_split, split_k;
    int a_mn_major, b_mn_major;
    int epilogue, out_f32_atomic, bias_f32;
    void* C; long long ldc;
    const void* bias;
    const void* aux_in; void* aux_out; long long ld_aux;
    int raster_group;              // 2-CTA kernel: m-blocks per rasterisation group (<= 1: row-major tile order)
    float* colsum;                 // optional [N] fp32: += column sums of the bf16 output (staged 2-CTA epilogue only)
    // K8 (all-gather -> GEMM): A is row-sharded over `a_shards` peer buffers of `a_shard_rows` rows each, every
    // shard has its own tensor map; m-blocks are visited starting at `m_block_offset` (own shard first).
    int a_shards, a_shard_rows, m_block_offset;
    // K7 (GEMM -> reduce-scatter): output row r belongs to rank r / c_shard_rows; the partial tile is stored
    // into that rank's staging buffer c_ptrs[owner] at slot `c_src_slot` (P2P stores from the epilogue).
    int c_scatter, c_shard_rows, c_src_slot;
    void* c_ptrs[8];
};

struct TmapArray { CUtensorMap m[8]; };

template <int BLOCK_N, int STAGES_> struct Cfg {
    static constexpr int STAGES = STAGES_;
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
    static constexpr int B_BYTES = BLOCK_N * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int TMEM_COLS = 2 * BLOCK_N;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

// one 32-column chunk of one accumulator row: bias / activation / residual / store.
// `aux` holds the 32 bf16 of aux_in for this chunk (prefetched one chunk ahead by the caller).
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, int row, int col0, const uint32_t (&r)[32],
                                               const uint4 (&aux)[4]) {
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
    if (p.bias != nullptr) {
        if (p.bias_f32) {
            const float4* b = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.bias) + col0);
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const float4 t = __ldg(b + q);
                v[q * 4 + 0] += t.x; v[q * 4 + 1] += t.y; v[q * 4 + 2] += t.z; v[q * 4 + 3] += t.w;
            }
        } else {
            const uint4* b = reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p.bias) + col0);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                uint4 u = __ldg(b + q);
                float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y), f2 = unpack_bf16x2(u.z), f3 = unpack_bf16x2(u.w);
                v[q * 8 + 0] += f0.x; v[q * 8 + 1] += f0.y; v[q * 8 + 2] += f1.x; v[q * 8 + 3] += f1.y;
                v[q * 8 + 4] += f2.x; v[q * 8 + 5] += f2.y; v[q * 8 + 6] += f3.x; v[q * 8 + 7] += f3.y;
            }
        }
    }
    if (p.out_f32_atomic) {
        float* out = reinterpret_cast<float*>(p.C) + static_cast<long long>(row) * p.ldc + col0;
#pragma unroll
        for (int q = 0; q < 8; ++q)
            asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};"
                         :: "l"(out + q * 4), "f"(v[q * 4]), "f"(v[q * 4 + 1]), "f"(v[q * 4 + 2]), "f"(v[q * 4 + 3]) : "memory");
        return;
    }
    if (p.epilogue == EPI_GELU || p.epilogue == EPI_GELU_DG) {
        __nv_bfloat16* pre = reinterpret_cast<__nv_bfloat16*>(p.aux_out) + static_cast<long long>(row) * p.ld_aux + col0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {            // aux_out gets the pre-activation, or gelu' of it; v becomes gelu(pre)
            float sv[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (p.epilogue == EPI_GELU_DG) { float g2; gelu_tanh_both_f(v[q * 8 + j], g2, sv[j]); v[q * 8 + j] = g2; }
                else { sv[j] = v[q * 8 + j]; v[q * 8 + j] = gelu_tanh_f(v[q * 8 + j]); }
            }
            uint4 u;
            u.x = pack_bf16x2(sv[0], sv[1]); u.y = pack_bf16x2(sv[2], sv[3]);
            u.z = pack_bf16x2(sv[4], sv[5]); u.w = pack_bf16x2(sv[6], sv[7]);
            *reinterpret_cast<uint4*>(pre + q * 8) = u;
        }
    } else if (p.epilogue == EPI_RESIDUAL || p.epilogue == EPI_DGELU || p.epilogue == EPI_MUL) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint4 u = aux[q];
            float2 f[4] = {unpack_bf16x2(u.x), unpack_bf16x2(u.y), unpack_bf16x2(u.z), unpack_bf16x2(u.w)};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (p.epilogue == EPI_RESIDUAL) {
                    v[q * 8 + 2 * e] += f[e].x; v[q * 8 + 2 * e + 1] += f[e].y;
                } else if (p.epilogue == EPI_MUL) {
                    v[q * 8 + 2 * e] *= f[e].x; v[q * 8 + 2 * e + 1] *= f[e].y;
                } else {
                    v[q * 8 + 2 * e] *= gelu_tanh_grad_f(f[e].x); v[q * 8 + 2 * e + 1] *= gelu_tanh_grad_f(f[e].y);
                }
            }
        }
    }
    __nv_bfloat16* out;
    if (p.c_scatter) {
        const int owner = row / p.c_shard_rows;
        const long long lrow = static_cast<long long>(p.c_src_slot) * p.c_shard_rows + (row - owner * p.c_shard_rows);
        out = reinterpret_cast<__nv_bfloat16*>(p.c_ptrs[owner]) + lrow * p.ldc + col0;
    } else {
        out = reinterpret_cast<__nv_bfloat16*>(p.C) + static_cast<long long>(row) * p.ldc + col0;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]); u.y = pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]);
        u.z = pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]); u.w = pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]);
        *reinterpret_cast<uint4*>(out + q * 8) = u;
    }
}

__device__ __forceinline__ void load_aux(const GemmParams& p, int row, int col0, bool ok, uint4 (&aux)[4]) {
    if (ok && (p.epilogue == EPI_RESIDUAL || p.epilogue == EPI_DGELU || p.epilogue == EPI_MUL)) {
        const uint4* a = reinterpret_cast<const uint4*>(
            reinterpret_cast<const __nv_bfloat16*>(p.aux_in) + static_cast<long long>(row) * p.ld_aux + col0);
#pragma unroll
        for (int q = 0; q < 4; ++q) aux[q] = __ldg(a + q);
    }
}

template <int BLOCK_N, int STAGES_>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ TmapArray tmaps_a, const __grid_constant__ CUtensorMap tmap_b,
                         const __grid_constant__ GemmParams p) {
    const CUtensorMap& tmap_a = tmaps_a.m[0];
    using C = Cfg<BLOCK_N, STAGES_>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + C::STAGES * C::A_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);
    uint64_t* full = bars;
    uint64_t* empty = bars + C::STAGES;
    uint64_t* tmem_full = bars + 2 * C::STAGES;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        for (int i = 1; i < p.a_shards; ++i) tma_prefetch_desc(&tmaps_a.m[i]);
        for (int s = 0; s < C::STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&tmem_full[s], 1); mbar_init(&tmem_empty[s], 8); }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, C::TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int tiles_mn = p.m_blocks * p.n_blocks;
    const int total_work = tiles_mn * p.split_k;
    const int total_k_blocks = (p.K + BLOCK_K - 1) / BLOCK_K;

    if (warp == 0) {
        // ================================ TMA producer ================================
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
                const int split = w / tiles_mn, t = w % tiles_mn;
                int m_blk = t / p.n_blocks + p.m_block_offset;
                if (m_blk >= p.m_blocks) m_blk -= p.m_blocks;
                const int m_idx = m_blk * BLOCK_M, n_idx = (t % p.n_blocks) * BLOCK_N;
                const CUtensorMap* ta = &tmap_a;
                int m_local = m_idx;
                if (p.a_shards > 1) {
                    const int sh = m_idx / p.a_shard_rows;
                    ta = &tmaps_a.m[sh];
                    m_local = m_idx - sh * p.a_shard_rows;
                }
                const int kb0 = split * p.k_blocks_per_split;
                const int kb1 = min(kb0 + p.k_blocks_per_split, total_k_blocks);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    mbar_arrive_expect_tx(&full[stage], C::STAGE_BYTES);
                    uint8_t* sa = smem_a + stage * C::A_BYTES;
                    uint8_t* sb = smem_b + stage * C::B_BYTES;
                    if (!p.a_mn_major) {
                        tma_load_2d(sa, ta, &full[stage], kb * BLOCK_K, m_local);
                    } else {
#pragma unroll
                        for (int j = 0; j < BLOCK_M / 64; ++j)
                            tma_load_2d(sa + j * (BLOCK_K * 128), &tmap_a, &full[stage], m_idx + j * 64, kb * BLOCK_K);
                    }
                    if (!p.b_mn_major) {
                        tma_load_2d(sb, &tmap_b, &full[stage], kb * BLOCK_K, n_idx);
                    } else {
#pragma unroll
                        for (int j = 0; j < BLOCK_N / 64; ++j)
                            tma_load_2d(sb + j * (BLOCK_K * 128), &tmap_b, &full[stage], n_idx + j * 64, kb * BLOCK_K);
                    }
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==================================
        const uint32_t idesc = umma_idesc_bf16(BLOCK_M, BLOCK_N, p.a_mn_major, p.b_mn_major);
        // per-UMMA_K advance of the descriptor start address (bytes)
        const uint32_t a_kstep = p.a_mn_major ? UMMA_K * 128 : UMMA_K * 2;
        const uint32_t b_kstep = p.b_mn_major ? UMMA_K * 128 : UMMA_K * 2;
        const uint32_t a_lbo = p.a_mn_major ? BLOCK_K * 128 : 16;
        const uint32_t b_lbo = p.b_mn_major ? BLOCK_K * 128 : 16;
        int stage = 0; uint32_t phase = 0;
        int acc = 0; uint32_t acc_phase = 0;
        for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
            const int split = w / tiles_mn;
            const int kb0 = split * p.k_blocks_per_split;
            const int kb1 = min(kb0 + p.k_blocks_per_split, total_k_blocks);
            mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(&full[stage], phase);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t sa = smem_u32(smem_a + stage * C::A_BYTES);
                    const uint32_t sb = smem_u32(smem_b + stage * C::B_BYTES);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                        const uint64_t da = umma_smem_desc(sa + k * a_kstep, a_lbo, 1024);
                        const uint64_t db = umma_smem_desc(sb + k * b_kstep, b_lbo, 1024);
                        umma_bf16_ss(tmem_d, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                    }
                    umma_commit(&empty[stage]);                     // smem slot free once these MMAs retire
                    if (kb == kb1 - 1) umma_commit(&tmem_full[acc]); // accumulator complete
                }
                __syncwarp();
                if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    } else {
        // ================================ epilogue ====================================
        // 8 warps: warp (2+e) reads TMEM lane quarter (warp & 3) and column half (e >> 2).  TMEM loads are
        // double-buffered in registers (chunk c+1 is in flight while chunk c is processed), aux_in is
        // prefetched one chunk ahead, and the accumulator stage is released as soon as it is in registers.
        const int e = warp - 2;
        const int quarter = warp & 3;
        const int half = e >> 2;
        constexpr int NC = BLOCK_N / 64;                              // 32-column chunks per warp
        int acc = 0; uint32_t acc_phase = 0;
        for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
            const int t = w % tiles_mn;
            int m_blk = t / p.n_blocks + p.m_block_offset;
            if (m_blk >= p.m_blocks) m_blk -= p.m_blocks;
            const int m_idx = m_blk * BLOCK_M, n_idx = (t % p.n_blocks) * BLOCK_N + half * (BLOCK_N / 2);
            const int row = m_idx + quarter * 32 + lane;
            const bool row_ok = row < p.M;
            uint4 aux[2][4];
            load_aux(p, row, n_idx, row_ok && n_idx < p.N, aux[0]);
            mbar_wait(&tmem_full[acc], acc_phase);
            __syncwarp();
            tc_fence_after();
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BLOCK_N + half * (BLOCK_N / 2);
            uint32_t r[2][32];
            tmem_ld_32x32(taddr, r[0]);
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                tmem_ld_wait();
                if (c + 1 < NC) {
                    tmem_ld_32x32(taddr + (c + 1) * 32, r[(c + 1) & 1]);
                    load_aux(p, row, n_idx + (c + 1) * 32, row_ok && n_idx + (c + 1) * 32 < p.N, aux[(c + 1) & 1]);
                } else {
                    // everything this warp needs from the accumulator stage is in registers: hand it back
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tmem_empty[acc]);
                }
                const int col0 = n_idx + c * 32;
                if (row_ok && col0 < p.N) epilogue_chunk(p, row, col0, r[c & 1], aux[c & 1]);
                __syncwarp();
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, C::TMEM_COLS);
    }
}


// =================================================================================================
// 2-CTA variant: a cluster of two CTAs (one TPC) computes a 256 x BLOCK_N tile with cta_group::2 MMAs.
// Each CTA stages its own 128 rows of A and only HALF of the B tile (the tensor core reads the other half
// from the peer's shared memory), so L2->SM traffic per FLOP drops by a third and a stage shrinks from
// 48 KB to 32 KB (6 stages instead of 4 in flight).  Only the leader CTA (cluster rank 0) issues MMAs; both
// CTAs run TMA producers (their transaction bytes land on the leader's `full` barrier) and epilogues (each
// drains its own 128 TMEM lanes).  tcgen05.commit multicasts the `empty` / `tmem_full` arrivals to both CTAs.
// =================================================================================================
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same smem offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t cta) {
    asm volatile(
        "{\n\t.reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}"
        :: "r"(smem_u32(bar)), "r"(cta) : "memory");
}
// TMA load whose completion bytes are credited to the LEADER CTA's barrier (peer bit cleared)
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    const uint32_t mbar = smem_u32(bar) & 0xFEFFFFFFu;
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        :: "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(mbar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void umma_bf16_ss_2cta(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        :: "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit_2cta(uint64_t* bar) {      // arrives on `bar` in both CTAs of the pair
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 :: "r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3)) : "memory");
}

constexpr int PANEL_BYTES = 128 * 64 * 2;        // one [128 rows x 64 cols] bf16 staging panel (128B-swizzled)

template <int BLOCK_N, int STAGES_, bool AUX> struct Cfg2 {
    static constexpr int STAGES = STAGES_;
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;                 // this CTA's 128 rows
    static constexpr int B_BYTES = (BLOCK_N / 2) * BLOCK_K * 2;           // this CTA's half of the B tile
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int TMEM_COLS = 2 * BLOCK_N;
    // epilogue staging: one output panel per column half (+ one aux panel per half when the epilogue reads or
    // writes a second tensor); the output leaves through TMA stores, aux_in arrives through TMA loads
    static constexpr int STAGING_BYTES = (AUX ? 4 : 2) * PANEL_BYTES;
    static constexpr int COLSUM_BYTES = 8 * 16 * 4;                       // per epilogue warp: 16 column sums
    static constexpr int BIAS_BYTES = 2 * (BLOCK_N / 2) * 4;              // per column half: this tile's bias as fp32
    // No alignment slack: the kernel has no static shared memory, so the dynamic window starts 1024-aligned (checked at
    // kernel start).  The KB saved is what lets one small block of another stream (the 64-thread trust-reduce, the
    // 128-thread pack / AdamW helpers: <= 256 B of shared memory, < 12 K registers) become resident beside a GEMM CTA:
    // 231,168 B here + 1 KB system reservation per block leaves 1,280 B of the SM's 233,472 B.
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + STAGING_BYTES + COLSUM_BYTES + BIAS_BYTES + 256;
};

struct EpiMaps { CUtensorMap c, aux; };

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* smem_src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 :: "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(nthreads) : "memory");
}
// byte offset of 16-byte piece q8 (0..7) of row r in a 128B-swizzled [rows x 64] bf16 panel
__device__ __forceinline__ uint32_t panel_off(uint32_t r, uint32_t q8) { return r * 128u + ((q8 ^ (r & 7u)) << 4); }


// Tile order of the persistent 2-CTA kernel.  Row-major (n fastest) streams the whole B operand once per m-block, which is
// fine while B fits in L2 (layer weights) but not for the LM head (N = 50304: B = 103 MB against 3.3 GB of output
// passing through L2).  Grouped order walks `group` m-blocks x all n-blocks with m fastest: the A slab of the group
// stays in L2 and B is streamed once per group instead of once per m-block.
__device__ __forceinline__ void tile_mn(int t, int n_blocks, int m_blocks, int group, int& m, int& n) {
    if (group <= 1) { m = t / n_blocks; n = t - m * n_blocks; return; }
    const int per = group * n_blocks;
    const int g = t / per, r = t - g * per;
    const int gm = min(group, m_blocks - g * group);
    n = r / gm;
    m = g * group + (r - n * gm);
}

template <int BLOCK_N, int STAGES_, bool AUX, bool COLSUM = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
gemm_bf16_tcgen05_2cta_kernel(const __grid_constant__ TmapArray tmaps_a, const __grid_constant__ CUtensorMap tmap_b,
                              const __grid_constant__ EpiMaps emaps, const __grid_constant__ TmapArray cmaps,
                              const __grid_constant__ GemmParams p) {
    using C = Cfg2<BLOCK_N, STAGES_, AUX>;
    const CUtensorMap& tmap_a = tmaps_a.m[0];
    extern __shared__ __align__(1024) uint8_t smem[];
    if (threadIdx.x == 0 && (smem_u32(smem) & 1023u)) __trap();          // swizzled TMA tiles need the 1024-byte alignment
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + C::STAGES * C::A_BYTES;
    uint8_t* staging = smem + C::STAGES * C::STAGE_BYTES;                 // [half][out | aux] panels
    float* colsum_smem = reinterpret_cast<float*>(staging + C::STAGING_BYTES);
    float* bias_smem = reinterpret_cast<float*>(staging + C::STAGING_BYTES + C::COLSUM_BYTES);     // [2][BLOCK_N / 2]
    uint64_t* bars = reinterpret_cast<uint64_t*>(staging + C::STAGING_BYTES + C::COLSUM_BYTES + C::BIAS_BYTES);
    uint64_t* full = bars;                       // used in the leader only
    uint64_t* empty = bars + C::STAGES;          // one per CTA
    uint64_t* tmem_full = bars + 2 * C::STAGES;  // one per CTA
    uint64_t* tmem_empty = tmem_full + 2;        // used in the leader only
    uint64_t* aux_bar = tmem_empty + 2;          // [2]: aux_in panel landed (one per column half)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(aux_bar + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t cta = cluster_ctarank();
    const bool leader = cta == 0;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        for (int i = 1; i < p.a_shards; ++i) tma_prefetch_desc(&tmaps_a.m[i]);
        for (int s = 0; s < C::STAGES; ++s) { mbar_init(&full[s], 2); mbar_init(&empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&tmem_full[s], 1); mbar_init(&tmem_empty[s], 16); mbar_init(&aux_bar[s], 1); }
        fence_barrier_init();
    }
    cluster_sync_all();                          // peers' barriers exist before anyone arrives remotely
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32(tmem_slot)), "r"(C::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int m_blocks2 = (p.M + 2 * BLOCK_M - 1) / (2 * BLOCK_M);
    const int tiles_mn = m_blocks2 * p.n_blocks;
    const int total_work = tiles_mn * p.split_k;
    const int total_k_blocks = (p.K + BLOCK_K - 1) / BLOCK_K;
    const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
    const int m2_offset = p.m_block_offset >> 1;

    // Programmatic dependent launch: everything above (barrier init, TMEM allocation, cluster handshake) may run while
    // the previous kernel in the stream is still draining on other SMs; global memory is touched only below this point.
    pdl_enter();

    if (warp == 0) {
        // ================================ TMA producer (both CTAs) ================================
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int w = cluster_id; w < total_work; w += n_clusters) {
                const int split = w / tiles_mn, t = w % tiles_mn;
                int m2, nb;
                tile_mn(t, p.n_blocks, m_blocks2, p.raster_group, m2, nb);
                m2 += m2_offset;
                if (m2 >= m_blocks2) m2 -= m_blocks2;
                const int m_idx = m2 * 2 * BLOCK_M + (int)cta * BLOCK_M;
                const int n_idx = nb * BLOCK_N + (int)cta * (BLOCK_N / 2);
                const CUtensorMap* ta = &tmap_a;
                int m_local = m_idx;
                if (p.a_shards > 1) {
                    const int sh = m_idx / p.a_shard_rows;
                    ta = &tmaps_a.m[sh < p.a_shards ? sh : p.a_shards - 1];
                    m_local = m_idx - sh * p.a_shard_rows;
                }
                const int kb0 = split * p.k_blocks_per_split;
                const int kb1 = min(kb0 + p.k_blocks_per_split, total_k_blocks);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    if (leader) mbar_arrive_expect_tx(&full[stage], 2 * C::STAGE_BYTES);
                    else mbar_arrive_remote(&full[stage], 0);
                    uint8_t* sa = smem_a + stage * C::A_BYTES;
                    uint8_t* sb = smem_b + stage * C::B_BYTES;
                    if (!p.a_mn_major) {
                        tma_load_2d_2sm(sa, ta, &full[stage], kb * BLOCK_K, m_local);
                    } else {
#pragma unroll
                        for (int j = 0; j < BLOCK_M / 64; ++j)
                            tma_load_2d_2sm(sa + j * (BLOCK_K * 128), &tmap_a, &full[stage], m_idx + j * 64, kb * BLOCK_K);
                    }
                    if (!p.b_mn_major) {
                        tma_load_2d_2sm(sb, &tmap_b, &full[stage], kb * BLOCK_K, n_idx);
                    } else {
#pragma unroll
                        for (int j = 0; j < BLOCK_N / 2 / 64; ++j)
                            tma_load_2d_2sm(sb + j * (BLOCK_K * 128), &tmap_b, &full[stage], n_idx + j * 64, kb * BLOCK_K);
                    }
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer (leader CTA only) ============================
        if (leader) {
            const uint32_t idesc = umma_idesc_bf16(2 * BLOCK_M, BLOCK_N, p.a_mn_major, p.b_mn_major);
            const uint32_t a_kstep = p.a_mn_major ? UMMA_K * 128 : UMMA_K * 2;
            const uint32_t b_kstep = p.b_mn_major ? UMMA_K * 128 : UMMA_K * 2;
            const uint32_t a_lbo = p.a_mn_major ? BLOCK_K * 128 : 16;
            const uint32_t b_lbo = p.b_mn_major ? BLOCK_K * 128 : 16;
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            for (int w = cluster_id; w < total_work; w += n_clusters) {
                const int split = w / tiles_mn;
                const int kb0 = split * p.k_blocks_per_split;
                const int kb1 = min(kb0 + p.k_blocks_per_split, total_k_blocks);
                mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    if (lane == 0) {
                        const uint32_t sa = smem_u32(smem_a + stage * C::A_BYTES);
                        const uint32_t sb = smem_u32(smem_b + stage * C::B_BYTES);
#pragma unroll
                        for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                            const uint64_t da = umma_smem_desc(sa + k * a_kstep, a_lbo, 1024);
                            const uint64_t db = umma_smem_desc(sb + k * b_kstep, b_lbo, 1024);
                            umma_bf16_ss_2cta(tmem_d, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                        }
                        umma_commit_2cta(&empty[stage]);
                        if (kb == kb1 - 1) umma_commit_2cta(&tmem_full[acc]);
                    }
                    __syncwarp();
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else {
        // ================================ epilogue (both CTAs) ====================================
        const int e = warp - 2;
        const int quarter = warp & 3;
        const int half = e >> 2;
        constexpr int NC = BLOCK_N / 64;
        // coalesced path: smem panels + TMA stores.  c_scatter == 2 (K7): the finished panel of a 128-row tile is stored by
        // TMA straight into the OWNER rank's staging slot over NVLink (one tensor map per owner in `cmaps`)
        const bool staged = !p.out_f32_atomic && p.c_scatter != 1;
        const bool aux_in = AUX && (p.epilogue == EPI_RESIDUAL || p.epilogue == EPI_DGELU || p.epilogue == EPI_MUL);
        const bool aux_out = AUX && (p.epilogue == EPI_GELU || p.epilogue == EPI_GELU_DG);
        uint8_t* out_panel = staging + half * (AUX ? 2 : 1) * PANEL_BYTES;
        uint8_t* aux_panel = out_panel + PANEL_BYTES;                   // only dereferenced when AUX
        const bool group_leader = (e & 3) == 0 && lane == 0;            // one thread per column half issues TMA
        const int bar_id = 1 + half;
        const uint32_t lrow = quarter * 32 + lane;                      // row inside the 128-row tile
        uint32_t aux_phase = 0;
        int acc = 0; uint32_t acc_phase = 0;
        for (int w = cluster_id; w < total_work; w += n_clusters) {
            const int t = w % tiles_mn;
            int m2, nb;
            tile_mn(t, p.n_blocks, m_blocks2, p.raster_group, m2, nb);
            m2 += m2_offset;
            if (m2 >= m_blocks2) m2 -= m_blocks2;
            const int m_idx = m2 * 2 * BLOCK_M + (int)cta * BLOCK_M;
            const int n_idx = nb * BLOCK_N + half * (BLOCK_N / 2);
            const int row = m_idx + quarter * 32 + lane;
            const bool row_ok = row < p.M;
            if (!staged) {
                // ---- direct path (split-K atomics, scatter to peers): per-thread global accesses -----------
                uint4 aux[2][4];
                load_aux(p, row, n_idx, row_ok && n_idx < p.N, aux[0]);
                mbar_wait(&tmem_full[acc], acc_phase);
                __syncwarp();
                tc_fence_after();
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BLOCK_N + half * (BLOCK_N / 2);
                uint32_t r[2][32];
                tmem_ld_32x32(taddr, r[0]);
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    tmem_ld_wait();
                    if (c + 1 < NC) {
                        tmem_ld_32x32(taddr + (c + 1) * 32, r[(c + 1) & 1]);
                        load_aux(p, row, n_idx + (c + 1) * 32, row_ok && n_idx + (c + 1) * 32 < p.N, aux[(c + 1) & 1]);
                    } else {
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) {
                            if (leader) mbar_arrive(&tmem_empty[acc]);
                            else mbar_arrive_remote(&tmem_empty[acc], 0);
                        }
                    }
                    const int col0 = n_idx + c * 32;
                    if (row_ok && col0 < p.N) epilogue_chunk(p, row, col0, r[c & 1], aux[c & 1]);
                    __syncwarp();
                }
            } else {
                // ---- staged path: 2 panels of 64 columns per half; outputs leave by TMA store, aux_in arrives by
                //      TMA load into the aux panel (issued before the accumulator is even ready) -------------
                if (aux_in && group_leader) {
                    mbar_arrive_expect_tx(&aux_bar[half], PANEL_BYTES);
                    tma_load_2d(aux_panel, &emaps.aux, &aux_bar[half], n_idx, m_idx);
                }
                // this half's 128 bias values -> shared memory (one per thread), read back as broadcasts after the first
                // panel barrier: a global load right before each chunk's bias add stalled the two epilogue warps of a
                // scheduler for the L2 latency.  Safe to overwrite: every warp passed the previous tile's last barrier.
                float* bias_s = bias_smem + half * (BLOCK_N / 2);
                if (p.bias != nullptr) {
                    const int bc = n_idx + (int)lrow;
                    float bvl = 0.f;
                    if (bc < p.N) bvl = p.bias_f32 ? __ldg(reinterpret_cast<const float*>(p.bias) + bc)
                                                   : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.bias)[bc]);
                    bias_s[lrow] = bvl;
                }
                mbar_wait(&tmem_full[acc], acc_phase);
                __syncwarp();
                tc_fence_after();
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * BLOCK_N + half * (BLOCK_N / 2);
                uint32_t r[2][32];
                tmem_ld_32x32(taddr, r[0]);
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    const int pi = c >> 1, cc = c & 1;                  // panel index, 32-col chunk inside the panel
                    tmem_ld_wait();
                    if (c + 1 < NC) {
                        tmem_ld_32x32(taddr + (c + 1) * 32, r[(c + 1) & 1]);
                    } else {
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) {
                            if (leader) mbar_arrive(&tmem_empty[acc]);
                            else mbar_arrive_remote(&tmem_empty[acc], 0);
                        }
                    }
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[c & 1][j]);
                    if (cc == 0) {
                        // the panel buffers are about to be rewritten: the previous TMA stores must have read them
                        if (group_leader) tma_store_wait_read();
                        if (aux_in) mbar_wait(&aux_bar[half], aux_phase);
                        named_bar_sync(bar_id, 128);
                    }
                    uint4 ua[4];
                    if (aux_in) {                                       // issued ahead of the bias adds that hide their latency
#pragma unroll
                        for (int q = 0; q < 4; ++q) ua[q] = *reinterpret_cast<const uint4*>(aux_panel + panel_off(lrow, cc * 4 + q));
                    }
                    if (p.bias != nullptr) {
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const float4 tb4 = *reinterpret_cast<const float4*>(bias_s + c * 32 + q * 4);
                            v[q * 4 + 0] += tb4.x; v[q * 4 + 1] += tb4.y; v[q * 4 + 2] += tb4.z; v[q * 4 + 3] += tb4.w;
                        }
                    }
                    if (aux_in) {
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const uint4 u = ua[q];
                            const float2 f[4] = {unpack_bf16x2(u.x), unpack_bf16x2(u.y), unpack_bf16x2(u.z), unpack_bf16x2(u.w)};
#pragma unroll
                            for (int e2 = 0; e2 < 4; ++e2) {
                                if (p.epilogue == EPI_RESIDUAL) {
                                    v[q * 8 + 2 * e2] += f[e2].x; v[q * 8 + 2 * e2 + 1] += f[e2].y;
                                } else if (p.epilogue == EPI_MUL) {
                                    v[q * 8 + 2 * e2] *= f[e2].x; v[q * 8 + 2 * e2 + 1] *= f[e2].y;
                                } else {
                                    v[q * 8 + 2 * e2] *= gelu_tanh_grad_f(f[e2].x); v[q * 8 + 2 * e2 + 1] *= gelu_tanh_grad_f(f[e2].y);
                                }
                            }
                        }
                    }
                    if (aux_out) {
                        if (p.epilogue == EPI_GELU_DG) {
                            // gelu and gelu' share the tanh; gelu' goes to the aux panel 8 values at a time
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                float dg[8];
#pragma unroll
                                for (int j = 0; j < 8; ++j) { float g2; gelu_tanh_both_f(v[q * 8 + j], g2, dg[j]); v[q * 8 + j] = g2; }
                                uint4 u;
                                u.x = pack_bf16x2(dg[0], dg[1]); u.y = pack_bf16x2(dg[2], dg[3]);
                                u.z = pack_bf16x2(dg[4], dg[5]); u.w = pack_bf16x2(dg[6], dg[7]);
                                *reinterpret_cast<uint4*>(aux_panel + panel_off(lrow, cc * 4 + q)) = u;
                            }
                        } else {
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                uint4 u;
                                u.x = pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]); u.y = pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]);
                                u.z = pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]); u.w = pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]);
                                *reinterpret_cast<uint4*>(aux_panel + panel_off(lrow, cc * 4 + q)) = u;
                            }
#pragma unroll
                            for (int j = 0; j < 32; ++j) v[j] = gelu_tanh_f(v[j]);  // from the fp32 pre-activation (its bf16 copy went to aux_out)
                        }
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        uint4 u;
                        u.x = pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]); u.y = pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]);
                        u.z = pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]); u.w = pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]);
                        *reinterpret_cast<uint4*>(out_panel + panel_off(lrow, cc * 4 + q)) = u;
                    }
                    if (cc == 1) {
                        fence_proxy_async_smem();                       // generic-proxy writes -> visible to TMA
                        named_bar_sync(bar_id, 128);
                        if (group_leader) {
                            const int pc = n_idx + pi * 64;
                            if (pc < p.N) {
                                if (p.c_scatter) {
                                    const int owner = m_idx / p.c_shard_rows;
                                    tma_store_2d(&cmaps.m[owner], out_panel, pc, p.c_src_slot * p.c_shard_rows + (m_idx - owner * p.c_shard_rows));
                                } else {
                                    tma_store_2d(&emaps.c, out_panel, pc, m_idx);
                                }
                                if (aux_out) tma_store_2d(&emaps.aux, aux_panel, pc, m_idx);
                            }
                            tma_store_commit();
                            if (aux_in && pi + 1 < NC / 2) {            // next panel's aux tile (aux panel is free: all 4 warps passed the barrier)
                                mbar_arrive_expect_tx(&aux_bar[half], PANEL_BYTES);
                                tma_load_2d(aux_panel, &emaps.aux, &aux_bar[half], pc + 64, m_idx);
                            }
                        }
                        if (aux_in) aux_phase ^= 1;
                        if (COLSUM && p.colsum != nullptr) {
                            // bias gradient of the consumer of this output: column sums of the finished bf16 panel
                            // (intact until the next panel's first barrier).  Warp wq owns 16 columns: lane = piece
                            // 2wq + (lane & 1) of rows (lane >> 1) + 16 i.  The 16 sums leave through an asynchronous
                            // bulk reduction (smem -> global add): plain red.global here would make the following
                            // bar.sync wait for the L2 round trip.
                            const uint32_t wq = e & 3;
                            const uint32_t q8 = 2u * wq + (lane & 1u);
                            float cs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const uint32_t r2 = (lane >> 1) + 16u * i;
                                if (m_idx + (int)r2 < p.M) {
                                    const uint4 u = *reinterpret_cast<const uint4*>(out_panel + panel_off(r2, q8));
                                    const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y), f2 = unpack_bf16x2(u.z), f3 = unpack_bf16x2(u.w);
                                    cs[0] += f0.x; cs[1] += f0.y; cs[2] += f1.x; cs[3] += f1.y;
                                    cs[4] += f2.x; cs[5] += f2.y; cs[6] += f3.x; cs[7] += f3.y;
                                }
                            }
#pragma unroll
                            for (int off = 2; off <= 16; off <<= 1) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], off);
                            }
                            float* slot = colsum_smem + e * 16;
                            if (lane == 1) tma_store_wait_read();       // this warp's previous reduction has read the slot
                            __syncwarp();
                            if (lane < 2) {
                                *reinterpret_cast<float4*>(slot + lane * 8) = make_float4(cs[0], cs[1], cs[2], cs[3]);
                                *reinterpret_cast<float4*>(slot + lane * 8 + 4) = make_float4(cs[4], cs[5], cs[6], cs[7]);
                                fence_proxy_async_smem();
                            }
                            __syncwarp();
                            const int gc = n_idx + pi * 64 + (int)wq * 16;
                            if (lane == 1 && gc < p.N) {
                                asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], 64;"
                                             :: "l"(p.colsum + gc), "r"(smem_u32(slot)) : "memory");
                                tma_store_commit();
                            }
                        }
                    }
                }
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (staged && (group_leader || (COLSUM && p.colsum != nullptr && lane == 1))) tma_store_wait_all();
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"(C::TMEM_COLS) : "memory");
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn == nullptr) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

}  // namespace

// 2-D bf16 tensor map, 128B swizzle, zero fill out of bounds. dims/box are (inner, outer).
int tdl_make_tmap_2d_bf16(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer, uint64_t ld_elems,
                          uint32_t box_inner, uint32_t box_outer) {
    EncodeTiledFn fn = get_encode_fn();
    TDL_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled entry point not available");
    TDL_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (ld_elems * 2) % 16 == 0,
                "TMA needs 16-byte aligned base and row pitch (base=%p ld=%llu)", base, (unsigned long long)ld_elems);
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {ld_elems * 2};
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    TDL_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (inner=%llu outer=%llu ld=%llu box=%ux%u)",
                (int)r, (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)ld_elems, box_inner, box_outer);
    return 0;
}

template <int BLOCK_N, int STAGES_>
static int launch_gemm(const TmapArray& ta, const CUtensorMap& tb, GemmParams& p, cudaStream_t stream, int sm_reserve = 0) {
    using C = Cfg<BLOCK_N, STAGES_>;
    static bool configured = false;
    if (!configured) {
        TDL_CHECK(cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BLOCK_N, STAGES_>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       C::SMEM_BYTES));
        configured = true;
    }
    const int work = p.m_blocks * p.n_blocks * p.split_k;
    const int sms = tdl_num_sms() - sm_reserve;
    const int grid = work < sms ? work : sms;
    gemm_bf16_tcgen05_kernel<BLOCK_N, STAGES_><<<grid, NUM_THREADS, C::SMEM_BYTES, stream>>>(ta, tb, p);
    TDL_LAUNCH_CHECK();
    return 0;
}

template <int BLOCK_N, int STAGES_, bool AUX, bool COLSUM = false>
static int launch_gemm_2cta(const TmapArray& ta, const CUtensorMap& tb, const EpiMaps& em, const TmapArray& cm, GemmParams& p,
                            cudaStream_t stream, int sm_reserve = 0) {
    using C = Cfg2<BLOCK_N, STAGES_, AUX>;
    static bool configured = false;
    if (!configured) {
        TDL_CHECK(cudaFuncSetAttribute(gemm_bf16_tcgen05_2cta_kernel<BLOCK_N, STAGES_, AUX, COLSUM>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
        configured = true;
    }
    const int m_blocks2 = (p.M + 2 * BLOCK_M - 1) / (2 * BLOCK_M);
    const int work = m_blocks2 * p.n_blocks * p.split_k;
    const int max_clusters = (tdl_num_sms() - sm_reserve) / 2;
    const int clusters = work < max_clusters ? work : max_clusters;
    TDL_CHECK(tdl_launch_pdl(gemm_bf16_tcgen05_2cta_kernel<BLOCK_N, STAGES_, AUX, COLSUM>, dim3(2 * clusters), dim3(NUM_THREADS),
                             C::SMEM_BYTES, stream, ta, tb, em, cm, p));
    TDL_LAUNCH_CHECK();
    return 0;
}

static int gemm_common(const void* const* a_shard_ptrs, int a_shards, int64_t a_shard_rows, int m_block_offset,
                       const void* B, void* Cout, void* const* c_ptrs, int c_scatter, int64_t c_shard_rows, int c_src_slot,
                       int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc, int a_mn_major, int b_mn_major,
                       const void* bias, int bias_f32, int epilogue, const void* aux_in, void* aux_out, int64_t ld_aux,
                       int out_f32_atomic, int split_k, int block_n, float* colsum, void* stream) {
    TDL_REQUIRE(M > 0 && N > 0 && K > 0, "empty GEMM");
    TDL_REQUIRE(colsum == nullptr || (!out_f32_atomic && !c_scatter), "column sums need the bf16 output");
    TDL_REQUIRE(N % 32 == 0, "N (%lld) must be a multiple of 32", (long long)N);
    TDL_REQUIRE(ldc % 8 == 0 && (ld_aux % 8 == 0), "ldc/ld_aux must be multiples of 8 elements");
    if (block_n != 128 && block_n != 256) block_n = (N % 256 == 0 || N > 2048) ? 256 : 128;
    GemmParams p{};
    p.M = (int)M; p.N = (int)N; p.K = (int)K;
    p.m_blocks = (int)((M + BLOCK_M - 1) / BLOCK_M);
    p.n_blocks = (int)((N + block_n - 1) / block_n);
    const int kblocks = (int)((K + BLOCK_K - 1) / BLOCK_K);
    if (split_k <= 0 && out_f32_atomic) {
        // auto split-K: pick the split whose work-unit count fills whole waves of the persistent grid
        // (74 CTA pairs for 256-row tiles), keeping at least 16 k-blocks per unit; ties go to the smaller split
        const bool pair = block_n == 256 && M >= 512;
        const int units = pair ? tdl_num_sms() / 2 : tdl_num_sms();
        const int tiles = (int)((M + (pair ? 255 : 127)) / (pair ? 256 : 128)) * p.n_blocks;
        double best = -1.0;
        for (int sk = 1; sk <= 16 && kblocks / sk >= 16; ++sk) {
            const int work = tiles * sk;
            const double eff = (double)work / (double)(((work + units - 1) / units) * units);
            if (eff > best + 0.02) { best = eff; split_k = sk; }
        }
    }
    if (split_k < 1) split_k = 1;
    if (split_k > kblocks) split_k = kblocks;
    TDL_REQUIRE(split_k == 1 || out_f32_atomic, "split-K needs the fp32 atomic epilogue");
    TDL_REQUIRE(!(out_f32_atomic && bias != nullptr), "the fp32 atomic epilogue takes no bias");
    p.k_blocks_per_split = (kblocks + split_k - 1) / split_k;
    p.split_k = (kblocks + p.k_blocks_per_split - 1) / p.k_blocks_per_split;
    p.a_mn_major = a_mn_major; p.b_mn_major = b_mn_major;
    p.epilogue = epilogue; p.out_f32_atomic = out_f32_atomic; p.bias_f32 = bias_f32;
    p.C = Cout; p.ldc = ldc; p.bias = bias; p.aux_in = aux_in; p.aux_out = aux_out; p.ld_aux = ld_aux;
    p.colsum = nullptr;
    p.raster_group = 1;
    TDL_REQUIRE(epilogue >= EPI_NONE && epilogue <= EPI_MUL, "unknown epilogue %d", epilogue);
    TDL_REQUIRE(!(epilogue == EPI_GELU || epilogue == EPI_GELU_DG) || aux_out != nullptr, "GELU epilogue needs aux_out");
    TDL_REQUIRE(!(epilogue == EPI_RESIDUAL || epilogue == EPI_DGELU || epilogue == EPI_MUL) || aux_in != nullptr, "epilogue needs aux_in");
    p.a_shards = a_shards; p.a_shard_rows = (int)a_shard_rows; p.m_block_offset = m_block_offset % p.m_blocks;
    p.c_scatter = c_scatter; p.c_shard_rows = (int)c_shard_rows; p.c_src_slot = c_src_slot;
    if (a_shards > 1) {
        TDL_REQUIRE(!a_mn_major && a_shards <= 8 && a_shard_rows % BLOCK_M == 0 && a_shard_rows * a_shards == M,
                    "sharded A needs K-major layout, <= 8 shards and shard rows that are a multiple of %d", BLOCK_M);
    }
    if (c_scatter) {
        TDL_REQUIRE(!out_f32_atomic && c_shard_rows > 0 && (M + c_shard_rows - 1) / c_shard_rows <= 8, "bad scatter arguments");
        for (int i = 0; i < (int)((M + c_shard_rows - 1) / c_shard_rows); ++i) p.c_ptrs[i] = c_ptrs[i];
    }
    TmapArray ta;
    CUtensorMap tb;
    int rc = 0;
    if (a_shards > 1) {
        for (int i = 0; i < a_shards && !rc; ++i)
            rc = tdl_make_tmap_2d_bf16(&ta.m[i], a_shard_ptrs[i], (uint64_t)K, (uint64_t)a_shard_rows, (uint64_t)lda, BLOCK_K, BLOCK_M);
    } else if (!a_mn_major) {
        rc = tdl_make_tmap_2d_bf16(&ta.m[0], a_shard_ptrs[0], (uint64_t)K, (uint64_t)M, (uint64_t)lda, BLOCK_K, BLOCK_M);
    } else {
        rc = tdl_make_tmap_2d_bf16(&ta.m[0], a_shard_ptrs[0], (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, BLOCK_K);
    }
    if (rc) return rc;
    // 2-CTA (cta_group::2) path: 256-row tiles; chosen when M is large enough to keep all SM pairs busy
    static int use_2cta = -1;
    if (use_2cta < 0) { const char* e = getenv("TDL_GEMM_2CTA"); use_2cta = e ? atoi(e) : 1; }
    const bool two_cta = use_2cta && block_n == 256 && M >= 512 && (a_shards <= 1 || a_shard_rows % 256 == 0) &&
                         (p.m_block_offset % 2 == 0);
    const uint32_t b_box_rows = two_cta ? (uint32_t)block_n / 2 : (uint32_t)block_n;
    if (!b_mn_major) rc = tdl_make_tmap_2d_bf16(&tb, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BLOCK_K, b_box_rows);
    else             rc = tdl_make_tmap_2d_bf16(&tb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, BLOCK_K);
    if (rc) return rc;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (two_cta) {
        static int raster = -1;                 // TDL_GEMM_RASTER=<m-blocks per group> (0/1 = row-major everywhere)
        if (raster < 0) { const char* e = getenv("TDL_GEMM_RASTER"); raster = e ? atoi(e) : 16; }
        if (p.n_blocks >= 32 && p.split_k == 1 && p.m_block_offset == 0) p.raster_group = raster;
        EpiMaps em;
        TmapArray cm = ta;                               // only read when c_scatter == 2
        // K7 through TMA stores needs every 128-row tile to belong to one owner; otherwise per-thread peer stores
        static int scatter_tma = -1;
        if (scatter_tma < 0) { const char* e = getenv("TDL_SCATTER_TMA"); scatter_tma = e ? atoi(e) : 1; }
        if (c_scatter && scatter_tma && c_shard_rows % BLOCK_M == 0 && M % BLOCK_M == 0) {
            p.c_scatter = 2;
            const int owners = (int)(M / c_shard_rows);
            for (int i = 0; i < owners && !rc; ++i)
                rc = tdl_make_tmap_2d_bf16(&cm.m[i], c_ptrs[i], (uint64_t)N, (uint64_t)(owners * c_shard_rows), (uint64_t)ldc, 64, 128);
            if (rc) return rc;
        }
        const bool staged = !out_f32_atomic && p.c_scatter != 1;
        const bool needs_aux = epilogue != EPI_NONE;
        if (staged && p.c_scatter == 2) {
            em.c = tb; em.aux = tb;
        } else if (staged) {
            rc = tdl_make_tmap_2d_bf16(&em.c, Cout, (uint64_t)N, (uint64_t)M, (uint64_t)ldc, 64, 128);
            if (rc) return rc;
            const void* auxp = (epilogue == EPI_GELU || epilogue == EPI_GELU_DG) ? aux_out : aux_in;
            rc = tdl_make_tmap_2d_bf16(&em.aux, needs_aux ? auxp : Cout, (uint64_t)N, (uint64_t)M, (uint64_t)(needs_aux ? ld_aux : ldc), 64, 128);
            if (rc) return rc;
        } else {
            em.c = tb; em.aux = tb;                     // unused on the direct path
        }
        const bool fuse_colsum = colsum != nullptr && staged && (reinterpret_cast<uintptr_t>(colsum) & 15) == 0;
        if (fuse_colsum) p.colsum = colsum;             // column sums leave from the staged epilogue panels
        if (fuse_colsum)
            rc = needs_aux ? launch_gemm_2cta<256, 5, true, true>(ta, tb, em, cm, p, s)
                           : launch_gemm_2cta<256, 6, false, true>(ta, tb, em, cm, p, s);
        else
            rc = (staged && needs_aux) ? launch_gemm_2cta<256, 5, true>(ta, tb, em, cm, p, s)
                                       : launch_gemm_2cta<256, 6, false>(ta, tb, em, cm, p, s);
        if (!rc && colsum != nullptr && !fuse_colsum) rc = tdl_bias_grad(Cout, colsum, M, N, ldc, stream);
        return rc;
    }
    static int exp_stages = -1;                  // TDL_GEMM_STAGES=3 selects the shallow pipeline (experiments only)
    if (exp_stages < 0) { const char* e = getenv("TDL_GEMM_STAGES"); exp_stages = e ? atoi(e) : 0; }
    if (block_n == 256) rc = exp_stages == 3 ? launch_gemm<256, 3>(ta, tb, p, s) : launch_gemm<256, 4>(ta, tb, p, s);
    else                rc = exp_stages == 3 ? launch_gemm<128, 3>(ta, tb, p, s) : launch_gemm<128, 6>(ta, tb, p, s);
    // the 1-CTA kernel has no fused column sums: follow with the stand-alone reduction over the output
    if (!rc && colsum != nullptr) rc = tdl_bias_grad(Cout, colsum, M, N, ldc, stream);
    return rc;
}

extern "C" int tdl_gemm_bf16(const void* A, const void* B, void* Cout, int64_t M, int64_t N, int64_t K,
                             int64_t lda, int64_t ldb, int64_t ldc, int a_mn_major, int b_mn_major,
                             const void* bias, int bias_f32, int epilogue, const void* aux_in, void* aux_out,
                             int64_t ld_aux, int out_f32_atomic, int split_k, int block_n, void* stream) {
    const void* a1[1] = {A};
    return gemm_common(a1, 1, M, 0, B, Cout, nullptr, 0, 0, 0, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, bias, bias_f32,
                       epilogue, aux_in, aux_out, ld_aux, out_f32_atomic, split_k, block_n, nullptr, stream);
}

// Same GEMM, and colsum[n] += sum_m bf16(C[m, n]) accumulated from the epilogue's staged output panels -- the
// bias gradient of the layer that produced this GEMM's output gradient, without re-reading C from HBM.
extern "C" int tdl_gemm_bf16_colsum(const void* A, const void* B, void* Cout, int64_t M, int64_t N, int64_t K,
                                    int64_t lda, int64_t ldb, int64_t ldc, int a_mn_major, int b_mn_major,
                                    const void* bias, int bias_f32, int epilogue, const void* aux_in, void* aux_out,
                                    int64_t ld_aux, int block_n, float* colsum, void* stream) {
    const void* a1[1] = {A};
    return gemm_common(a1, 1, M, 0, B, Cout, nullptr, 0, 0, 0, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, bias, bias_f32,
                       epilogue, aux_in, aux_out, ld_aux, 0, 1, block_n, colsum, stream);
}

// K8: all-gather -> GEMM.  A = concat of `W` row shards [shard_rows, K] that live in peer-mapped buffers;
// the TMA producer reads each tile straight from the owning peer while the tensor core works on earlier
// tiles.  `my_rank` rotates the tile order so every rank starts on its own shard.
extern "C" int tdl_allgather_gemm_bf16(const void* const* a_shard_ptrs, int W, int my_rank, int64_t shard_rows, const void* B,
                                       void* Cout, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc,
                                       const void* bias, int bias_f32, int epilogue, const void* aux_in, void* aux_out,
                                       int64_t ld_aux, int block_n, void* stream) {
    const int64_t M = shard_rows * W;
    const int offset = (int)((shard_rows / BLOCK_M) * my_rank);
    return gemm_common(a_shard_ptrs, W, shard_rows, offset, B, Cout, nullptr, 0, 0, 0, M, N, K, lda, ldb, ldc, 0, 0, bias,
                       bias_f32, epilogue, aux_in, aux_out, ld_aux, 0, 1, block_n, nullptr, stream);
}

// K8 with the operand layouts of the backward pass: B may be stored [K, N] (dgrad of a row-parallel layer reads the
// weight MN-major) and the column sums of the bf16 output (a bias gradient) can leave from the epilogue panels.
extern "C" int tdl_allgather_gemm_bf16_ex(const void* const* a_shard_ptrs, int W, int my_rank, int64_t shard_rows, const void* B,
                                          void* Cout, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc, int b_mn_major,
                                          const void* bias, int bias_f32, int epilogue, const void* aux_in, void* aux_out,
                                          int64_t ld_aux, int block_n, float* colsum, void* stream) {
    const int64_t M = shard_rows * W;
    const int offset = (int)((shard_rows / BLOCK_M) * my_rank);
    return gemm_common(a_shard_ptrs, W, shard_rows, offset, B, Cout, nullptr, 0, 0, 0, M, N, K, lda, ldb, ldc, 0, b_mn_major, bias,
                       bias_f32, epilogue, aux_in, aux_out, ld_aux, 0, 1, block_n, colsum, stream);
}

// K7: GEMM -> reduce-scatter (first half).  The partial product tile for rows owned by rank o is stored from
// the epilogue directly into o's staging buffer c_ptrs[o] ([W slots][shard_rows][ldc] bf16) at slot `my_rank`;
// after the cross-rank barrier the owner sums its W slots (tdl_reduce_slots, usually fused with residual+LN).
extern "C" int tdl_gemm_scatter_bf16(const void* A, const void* B, void* const* c_ptrs, int W, int my_rank,
                                     int64_t shard_rows, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc,
                                     int b_mn_major, int block_n, void* stream) {
    const int64_t M = shard_rows * W;
    const void* a1[1] = {A};
    const int offset = (int)(((shard_rows + BLOCK_M - 1) / BLOCK_M) * ((my_rank + 1) % W));   // start on a peer: spread the traffic
    return gemm_common(a1, 1, M, offset, B, nullptr, c_ptrs, 1, shard_rows, my_rank, M, N, K, lda, ldb, ldc, 0, b_mn_major,
                       nullptr, 0, 0, nullptr, nullptr, ldc, 0, 1, block_n, nullptr, stream);
}

// K12: causal softmax-attention forward / backward for head_dim 64 on tcgen05 tensor cores.
//
// Layout: qkv is the fused projection output [B*T, 3C] (bf16), head h uses columns
// [h*64, h*64+64) of each third; the output / its gradient are [B*T, C].  TMA pulls 128x64 tiles
// straight out of those buffers (no head split / transpose copies).
//
// Forward (one CTA = one 128-query block of one head, 2 CTAs per SM):
//   S = Q K^T      tcgen05.mma, K-major operands, fp32 S in TMEM (128 lanes x 128 columns)
//   softmax        each of the 128 threads owns one query row (tcgen05.ld 32x32b), online max/sum in
//                  registers, exp2 with the scale folded in, P written back to tensor memory as bf16 pairs
//                  (tcgen05.st) -- no shared-memory round trip
//   O += P V       tcgen05.mma with A = P from TMEM and V consumed MN-major exactly as it sits in memory; partial
//                  products land in TMEM and are folded into the register accumulator with the
//                  running rescale
// Backward = two persistent tensor-core kernels (one CTA per SM walking a longest-first tile list):
//   dQ kernel      per query block, loops over key blocks; dQ += dS K; also computes delta = rowsum(dO*O) from the tiles it
//                  already holds and writes it out for the second kernel
//   dK/dV kernel   per key block, loops over query blocks on the TRANSPOSED tiles S^T = K Q^T, dP^T = V dO^T, so that
//                  dV += P^T dO and dK += dS^T Q take P^T / dS^T as A operands straight from tensor memory
//   Roles inside a CTA: one TMA warp, one MMA-issue warp (elect.sync, precomputed descriptor words), 8 warps that only
//   turn S into P (exp2), 8 warps that turn (P, dP) into dS and write finished tiles out through TMA stores (and add
//   the qkv bias gradient).  S is double-buffered in TMEM; P / dS are written back into the S buffer they came from.
// The reference reaches attention only through torch/HF eager ops.
#include "common.cuh"
#include <cstdlib>

#ifdef TDL_ATTN_TRACE
// poor man's tracer (bench/attn_trace.cu, bench/attn_bwd_trace.cu): one thread of one CTA stamps clock64() at phase
// boundaries.  The stamp index lives in a register (TRACE_DECL) so a stamp costs a clock read and two stores.
__device__ long long g_trace[8192];
__device__ int g_trace_n;
#define TRACE_DECL int _trace_i = 0
#define TRACE_AT(tag)                                                                                   \
    do {                                                                                                \
        if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == TDL_ATTN_TRACE_TID) { \
            if (_trace_i < 4096) { g_trace[2 * _trace_i] = (tag); g_trace[2 * _trace_i + 1] = clock64(); g_trace_n = ++_trace_i; } \
        }                                                                                               \
    } while (0)
#else
#define TRACE_DECL do {} while (0)
#define TRACE_AT(tag) do {} while (0)
#endif

namespace {

constexpr int D = 64;            // head dim
constexpr int BQ = 128;          // query rows per tile
constexpr int BK = 128;          // key rows per tile
constexpr int TILE_BYTES = 128 * 64 * 2;   // 16 KB: one [128 x 64] bf16 tile

struct AttnParams {
    int B, T, H, C;              // C = H * 64
    float scale_log2;            // softmax scale * log2(e)
    float scale;                 // softmax scale
    __nv_bfloat16* out;          // fwd: O [B*T, C]
    float* lse;                  // [B, H, T]  natural-log logsumexp of the scaled scores
    const float* delta;          // bwd: [B, H, T]
    __nv_bfloat16* dqkv;         // bwd: [B*T, 3C]
    float* dbias;                // bwd, optional: [3C] += column sums of dqkv
    int bwd_chunk;               // bwd (persistent): (batch, head) pairs per scheduling chunk (working set <= L2)
    unsigned long long mg_tile, mg_chunk, mg_head;   // ceil(2^40 / d) for d = nblk * chunk, chunk, H (exact for n, d < 2^20)
};

// byte offset of the 16-byte piece `q8` (0..7) of row r inside a [rows x 64] bf16 tile with 128B swizzle
__device__ __forceinline__ uint32_t swz(uint32_t r, uint32_t q8) { return r * 128u + ((q8 ^ (r & 7u)) << 4); }

__device__ __forceinline__ void tma_store_2d_tile(const CUtensorMap* map, const void* smem_src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 :: "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
}

// =================================================================================================
// forward
// =================================================================================================
__global__ void __launch_bounds__(128, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const AttnParams p) {
    TRACE_DECL;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sQ = smem;
    uint8_t* sK = smem + TILE_BYTES;              // 2 buffers
    uint8_t* sV = smem + 3 * TILE_BYTES;          // 2 buffers
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 5 * TILE_BYTES);
    uint64_t* bar_q = bars;
    uint64_t* bar_kv = bars + 1;                  // [2]
    uint64_t* bar_s = bars + 3;
    uint64_t* bar_o = bars + 4;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5);

    const int tid = threadIdx.x, warp = tid >> 5;
    // heaviest (longest causal row) tiles first
    const int qb = (gridDim.x - 1) - blockIdx.x;
    const int h = blockIdx.y, b = blockIdx.z;
    const int row0 = b * p.T;
    const int n_kv = qb + 1;

    if (tid == 0) {
        if (smem_u32(smem) & 1023u) __trap();
        tma_prefetch_desc(&tmap_qkv);
        mbar_init(bar_q, 1); mbar_init(&bar_kv[0], 1); mbar_init(&bar_kv[1], 1);
        mbar_init(bar_s, 1); mbar_init(bar_o, 1);
        fence_barrier_init();
    }
    if (warp == 0) { tmem_alloc(tmem_slot, 256); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    // S [0,128) fp32 | O block product [128,192) | P [192,256) as bf16 pairs: the A operand of P V straight from TMEM
    const uint32_t tmem_s = tmem, tmem_o = tmem + 128, tmem_p = tmem + 192;
    const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;
    pdl_enter();                                     // global memory is touched only below

    if (tid == 0) {
        mbar_arrive_expect_tx(bar_q, TILE_BYTES);
        tma_load_2d(sQ, &tmap_qkv, bar_q, h * D, row0 + qb * BQ);
        mbar_arrive_expect_tx(&bar_kv[0], 2 * TILE_BYTES);
        tma_load_2d(sK, &tmap_qkv, &bar_kv[0], p.C + h * D, row0);
        tma_load_2d(sV, &tmap_qkv, &bar_kv[0], 2 * p.C + h * D, row0);
    }

    const uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);
    const uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);
    float o_acc[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) o_acc[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;
    const int r = tid;                               // query row inside the tile

    // S(j+1) = Q K_{j+1}^T is queued on the tensor pipe right behind P_j V_j (the S accumulator is already in
    // registers by then), so the next block's scores are ready when the softmax warps come back for them.
    // MMA issue: warp 0 waits as a whole, one elected lane issues with precomputed descriptor low words (a lone thread
    // that rebuilds descriptors under `if (tid == 0)` spends ~70 clk per tcgen05.mma, which sat on the critical path).
    const uint32_t q_lo = umma_desc_lo(smem_u32(sQ), 16);
    auto issue_s = [&](int jj) {                     // warp 0, converged
        const int b2 = jj & 1;
        mbar_wait(&bar_kv[b2], (jj >> 1) & 1);
        tc_fence_after();
        const uint32_t k_lo = umma_desc_lo(smem_u32(sK + b2 * TILE_BYTES), 16);
        if (elect_one()) {
#pragma unroll
            for (int k = 0; k < D / 16; ++k) umma_bf16_ss_lo(tmem_s, q_lo + 2 * k, k_lo + 2 * k, idesc_s, k > 0);
            umma_commit(bar_s);
        }
        __syncwarp();
    };
    if (warp == 0) { mbar_wait(bar_q, 0); issue_s(0); }
    for (int j = 0; j < n_kv; ++j) {
        const int buf = j & 1;
        if (tid == 0 && j + 1 < n_kv) {              // prefetch next K/V into the other buffer
            mbar_arrive_expect_tx(&bar_kv[buf ^ 1], 2 * TILE_BYTES);
            tma_load_2d(sK + (buf ^ 1) * TILE_BYTES, &tmap_qkv, &bar_kv[buf ^ 1], p.C + h * D, row0 + (j + 1) * BK);
            tma_load_2d(sV + (buf ^ 1) * TILE_BYTES, &tmap_qkv, &bar_kv[buf ^ 1], 2 * p.C + h * D, row0 + (j + 1) * BK);
        }
        TRACE_AT(1);
        mbar_wait(bar_s, j & 1);
        __syncwarp();
        tc_fence_after();
        TRACE_AT(2);

        // ---- online softmax over this thread's row --------------------------------------------
        // The whole S row (128 fp32) is pulled into registers with one round of TMEM loads; max and
        // exp are then pure register passes (1 FMNMX, then FFMA + EX2 + FADD per element).
        const bool diag = (j == qb);
        uint32_t sraw[128];
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld_32x32(tmem_s + lane_base + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&sraw[c * 32]));
        tmem_ld_wait();
        TRACE_AT(3);
        // causal mask only exists in the diagonal block: fold it into the scores there (one FSEL per element,
        // warp-uniform branch) so the common path carries no predicate arithmetic at all
        if (diag) {
#pragma unroll
            for (int i = 0; i < 128; ++i) if (i > r) sraw[i] = 0xff800000u;           // -inf
        }
        // four independent chains: a single running max/sum would serialise 128 dependent ops
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int i = 0; i < 128; ++i) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(sraw[i]));
        const float mraw = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
        const float mx = fmaxf(m_run, mraw * p.scale_log2);
        const float alpha = exp2f(m_run - mx);
        TRACE_AT(4);
        float rs4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float pv[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const float e = exp2f(fmaf(__uint_as_float(sraw[c * 32 + i]), p.scale_log2, -mx));   // exp2(-inf) = 0 under the mask
                pv[i] = e;
                rs4[i & 3] += e;
            }
            uint32_t pk[16];                         // this row's 32 probabilities as 16 bf16 pairs -> TMEM
#pragma unroll
            for (int i = 0; i < 16; ++i) pk[i] = pack_bf16x2(pv[2 * i], pv[2 * i + 1]);
            tmem_st_32x16(tmem_p + lane_base + c * 16, pk);
        }
        const float rs = (rs4[0] + rs4[1]) + (rs4[2] + rs4[3]);
        TRACE_AT(5);
        l_run = l_run * alpha + rs;
        m_run = mx;
#pragma unroll
        for (int i = 0; i < 64; ++i) o_acc[i] *= alpha;

        TRACE_AT(6);
        tmem_st_wait();                              // P is in tensor memory
        tc_fence_before();
        __syncthreads();
        TRACE_AT(7);
        if (warp == 0) {
            tc_fence_after();
            const uint32_t v_lo = umma_desc_lo(smem_u32(sV + buf * TILE_BYTES), TILE_BYTES);
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < BK / 16; ++k)    // A: P from TMEM, 8 packed columns per K slice; B: V MN-major, 2048 B per K slice
                    umma_bf16_ts_lo(tmem_o, tmem_p + 8 * k, v_lo + 128 * k, idesc_o, k > 0);
                umma_commit(bar_o);
            }
            __syncwarp();
            if (j + 1 < n_kv) issue_s(j + 1);
        }
        TRACE_AT(8);
        mbar_wait(bar_o, j & 1);
        __syncwarp();
        tc_fence_after();
        TRACE_AT(9);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            uint32_t raw[32];
            tmem_ld_32x32(tmem_o + lane_base + c * 32, raw);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o_acc[c * 32 + i] += __uint_as_float(raw[i]);
        }
        tc_fence_before();
        TRACE_AT(10);
    }

    // ---- epilogue -----------------------------------------------------------------------------
    const float inv = 1.0f / l_run;
    const long long grow = (long long)(row0 + qb * BQ + r);
    __nv_bfloat16* orow = p.out + grow * p.C + h * D;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        uint4 u;
        u.x = pack_bf16x2(o_acc[q * 8 + 0] * inv, o_acc[q * 8 + 1] * inv); u.y = pack_bf16x2(o_acc[q * 8 + 2] * inv, o_acc[q * 8 + 3] * inv);
        u.z = pack_bf16x2(o_acc[q * 8 + 4] * inv, o_acc[q * 8 + 5] * inv); u.w = pack_bf16x2(o_acc[q * 8 + 6] * inv, o_acc[q * 8 + 7] * inv);
        *reinterpret_cast<uint4*>(orow + q * 8) = u;
    }
    p.lse[((long long)b * p.H + h) * p.T + qb * BQ + r] = (m_run + log2f(l_run)) * 0.6931471805599453f;

    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 256); }
}

// =================================================================================================
// backward: persistent, software-pipelined, P / dS kept in tensor memory (default).
// Same math as attn_bwd_kernel, restructured around what bounded it there:
//   * the tensor core and the CUDA cores took turns            -> warp 8 (one elected lane) issues every tcgen05.mma,
//     warp 9 every TMA load, the 8 math warps only talk to them through mbarriers, so
//         tensor pipe: acc(j) | dP(j+1) | S(j+2)     runs under     math: exp(j+1), dS(j+1);
//   * shared-memory bandwidth went into writing P / dS and reading them back as A operands
//                                                              -> S is double-buffered in TMEM and the bf16 P / dS tiles are
//     written back INTO the S buffer they came from (tcgen05.st) and consumed from there (A-from-TMEM tcgen05.mma).
//     MODE 0 computes the transposed tiles S^T = K Q^T, dP^T = V dO^T so that P^T / dS^T have the key index on the
//     TMEM lanes, as dV = P^T dO and dK = dS^T Q need; the per-query statistics then run along the columns and are
//     staged in shared memory once per query block;
//   * at T = 1024 a (block, head) tile is only 4.5 iterations long and 4.6 us of launch / prologue / drain per CTA was
//     43 % of the kernel                                       -> one CTA per SM walks a static, longest-first list of
//     tiles; barriers, TMEM and the load / MMA pipelines run on across tile boundaries (global iteration counter j), so
//     the next tile's loads and first MMAs are in flight while the math warps write the previous tile's accumulators.
// smem (14 tiles = 224 KB): 2 fixed pairs (ring) | 4 stream stages (MODE 0: {Q,dO}; MODE 1: {K,V}) | 2 output staging
//       tiles (leave by TMA store) | column statistics
// TMEM: S0 [0,128)  S1 [128,256)  dP [256,384)  acc0 [384,448)  acc1 [448,512)
//       inside S[b], the thread that owns columns [64h, 64h+64) stores P at [64h, 64h+32) and dS at [64h+32, 64h+64)
// =================================================================================================
constexpr int BWD_WS_NCG = 4;                       // threads per tile row (column groups): 4 -> 16 math warps
constexpr int BWD_WS_MATH = BWD_WS_NCG * 128;        // math threads
constexpr int BWD_WS_THREADS = BWD_WS_MATH + 64;     // + MMA-issue warp + TMA warp
constexpr int BWD_WS_NSTR = 4;       // a stage is refilled when acc(j) retires and needed again by S(j+4): > 1 block of slack
constexpr int BWD_WS_NFIX = 2;
constexpr int BWD_WS_SMEM = (2 * BWD_WS_NFIX + 2 * BWD_WS_NSTR + 2) * TILE_BYTES + 2048 + 256;

struct BwdCursor {
    int t;        // global tile id; >= total when exhausted
    int k;        // ordinal of the tile inside this CTA
    int it, n;    // iteration inside the tile, iterations of the tile
    int j;        // iteration counter across tiles (drives every ring buffer and barrier parity)
    int ob, h, b; // decoded tile: outer block, head, batch (integer divisions only when the tile changes)
};


// =================================================================================================
// backward: persistent, warp-specialised (TMA warp, MMA-issue warp, 16 math warps), the math split by PHASE rather than
// by column: warps 0-7 only turn S into P (the MUFU-bound stage), warps 8-15 turn (P, dP) into dS and own
// the tile epilogue and the delta computation.  The exp of block j+1 therefore runs under the dS arithmetic, the dK/dQ
// MMAs and the S MMA of block j instead of taking turns with them (the single dP buffer no longer serialises anything:
// only the dS group touches dP).  dS is formed from the bf16 P that the tensor core consumes.
// =================================================================================================
template <int MODE>
__global__ void __launch_bounds__(BWD_WS_THREADS, 1)
attn_bwd_ps_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_do,
                   const __grid_constant__ CUtensorMap tmap_dqkv, const __grid_constant__ CUtensorMap tmap_o, const AttnParams p) {
    constexpr int NSTR = BWD_WS_NSTR, NFIX = BWD_WS_NFIX;
    TRACE_DECL;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sFix = smem;                                   // [NFIX][2] : MODE0: K,V   MODE1: Q,dO
    uint8_t* sStr = smem + 2 * NFIX * TILE_BYTES;           // [NSTR][2] : MODE0: Q,dO  MODE1: K,V
    uint8_t* sOut = smem + (2 * NFIX + 2 * NSTR) * TILE_BYTES;                                  // [2] finished dK,dV / dQ tiles
    float* sStat = reinterpret_cast<float*>(smem + (2 * NFIX + 2 * NSTR + 2) * TILE_BYTES);     // [2][2][128]: lse*log2e, delta (MODE 0)
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (2 * NFIX + 2 * NSTR + 2) * TILE_BYTES + 2048);
    uint64_t* fix_full = bars;                     // [<= 3]
    uint64_t* fix_empty = bars + 3;                // [<= 3]
    uint64_t* str_full = bars + 6;                 // [4]
    uint64_t* str_empty = bars + 10;               // [4]
    uint64_t* s_full = bars + 14;                  // [2]
    uint64_t* dp_full = bars + 16;
    uint64_t* dp_free = bars + 17;
    uint64_t* p_full = bars + 18;                  // [2]
    uint64_t* acc_done = bars + 20;                // [2]
    uint64_t* acc_free = bars + 22;
    uint64_t* o_full = bars + 23;                  // MODE 1: the tile's O rows landed (for delta = rowsum(dO * O))
    uint64_t* o_free = bars + 24;
    uint64_t* ds_full = bars + 25;                 // [2]: dS(j) is in S[j & 1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 27);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nblk = p.T / 128;
    const int BH = p.B * p.H;
    const int total = nblk * BH;
    // Tile order: (batch, head) pairs in chunks of p.bwd_chunk whose Q/K/V/dO fit in L2; inside a chunk longest tiles
    // first.  CTA c takes tiles c, c + gridDim.x, ...: the chunk size is not a multiple of the grid, so which CTA gets
    // the long tiles rotates from chunk to chunk.
    const int CH = p.bwd_chunk;
    auto fdiv = [](int n, unsigned long long magic) { return (int)(((unsigned long long)(unsigned)n * magic) >> 40); };
    auto cur_decode = [&](BwdCursor& c) {                     // runs when a cursor enters a new tile; no hardware division
        if (c.t >= total) { c.n = 0; return; }
        const int chunk = fdiv(c.t, p.mg_tile);               // t / (nblk * CH); CH divides B*H
        const int tl = c.t - chunk * nblk * CH;
        const int rank = fdiv(tl, p.mg_chunk);                // tl / CH
        const int bh = chunk * CH + tl - rank * CH;
        c.ob = MODE == 0 ? rank : nblk - 1 - rank;
        c.n = MODE == 0 ? nblk - c.ob : c.ob + 1;
        c.b = fdiv(bh, p.mg_head);                            // bh / H
        c.h = bh - c.b * p.H;
    };
    auto cur_init = [&]() {
        BwdCursor c; c.t = blockIdx.x; c.k = 0; c.it = 0; c.j = 0; c.ob = c.h = c.b = 0;
        cur_decode(c);
        return c;
    };
    auto cur_next = [&](BwdCursor& c) {
        ++c.j;
        if (++c.it == c.n) { c.it = 0; c.t += gridDim.x; ++c.k; cur_decode(c); }
    };

    if (tid == 0) {
        if (smem_u32(smem) & 1023u) __trap();
        tma_prefetch_desc(&tmap_qkv); tma_prefetch_desc(&tmap_do); tma_prefetch_desc(&tmap_dqkv);
        for (int s = 0; s < NFIX; ++s) { mbar_init(&fix_full[s], 1); mbar_init(&fix_empty[s], 1); }
        for (int s = 0; s < NSTR; ++s) { mbar_init(&str_full[s], 1); mbar_init(&str_empty[s], 1); }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&s_full[s], 1); mbar_init(&p_full[s], 8); mbar_init(&ds_full[s], 8); mbar_init(&acc_done[s], 1);
        }
        mbar_init(dp_full, 1); mbar_init(dp_free, 8); mbar_init(acc_free, 8);
        mbar_init(o_full, 1); mbar_init(o_free, 8);
        fence_barrier_init();
    }
    if (warp == 0) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t tmem_dp = tmem + 256, tmem_a0 = tmem + 384, tmem_a1 = tmem + 448;
    pdl_enter();                                     // global memory is touched only below

    if (warp == BWD_WS_MATH / 32 + 1) {
        // ======================================= TMA producer ====================================
        if (lane == 0) {
            for (BwdCursor c = cur_init(); c.t < total; cur_next(c)) {
                const int ob = c.ob, h = c.h;
                const int row0 = c.b * p.T;
                if (c.it == 0) {
                    const int fb = c.k % NFIX;
                    if (c.k >= NFIX) mbar_wait(&fix_empty[fb], ((c.k / NFIX) - 1) & 1);
                    uint8_t* f0 = sFix + fb * 2 * TILE_BYTES;
                    mbar_arrive_expect_tx(&fix_full[fb], 2 * TILE_BYTES);
                    if (MODE == 0) {
                        tma_load_2d(f0, &tmap_qkv, &fix_full[fb], p.C + h * D, row0 + ob * 128);                  // K
                        tma_load_2d(f0 + TILE_BYTES, &tmap_qkv, &fix_full[fb], 2 * p.C + h * D, row0 + ob * 128); // V
                    } else {
                        tma_load_2d(f0, &tmap_qkv, &fix_full[fb], h * D, row0 + ob * 128);                        // Q
                        tma_load_2d(f0 + TILE_BYTES, &tmap_do, &fix_full[fb], h * D, row0 + ob * 128);            // dO
                        // O rows of the tile -> the staging tile MODE 1 does not use for output (single buffer)
                        if (c.k >= 1) mbar_wait(o_free, (c.k - 1) & 1);
                        mbar_arrive_expect_tx(o_full, TILE_BYTES);
                        tma_load_2d(sOut + TILE_BYTES, &tmap_o, o_full, h * D, row0 + ob * 128);
                    }
                }
                const int st = c.j % NSTR;
                const int i = (MODE == 0 ? ob : 0) + c.it;
                if (c.j >= NSTR) mbar_wait(&str_empty[st], ((c.j / NSTR) - 1) & 1);
                uint8_t* s0 = sStr + st * 2 * TILE_BYTES;
                mbar_arrive_expect_tx(&str_full[st], 2 * TILE_BYTES);
                if (MODE == 0) {
                    tma_load_2d(s0, &tmap_qkv, &str_full[st], h * D, row0 + i * 128);                  // Q_i
                    tma_load_2d(s0 + TILE_BYTES, &tmap_do, &str_full[st], h * D, row0 + i * 128);      // dO_i
                } else {
                    tma_load_2d(s0, &tmap_qkv, &str_full[st], p.C + h * D, row0 + i * 128);            // K_i
                    tma_load_2d(s0 + TILE_BYTES, &tmap_qkv, &str_full[st], 2 * p.C + h * D, row0 + i * 128);   // V_i
                }
            }
        }
        __syncwarp();
    } else if (warp == BWD_WS_MATH / 32) {
        // ======================================= MMA issue =======================================
        // the whole warp walks the schedule and waits; one elected lane issues
        {
            const uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);       // S, dP : K-major x K-major
            const uint32_t idesc_ts = umma_idesc_bf16(128, 64, 0, 1);       // acc   : A from TMEM (K along columns) x MN-major B
            // K-slice k (16 values = 8 packed columns) of P lives in the S columns of the thread group that owns them
            auto a_col = [](int k) { return (uint32_t)((k * 16 / 64) * 64 + ((k * 16) % 64) / 2); };   // two threads per row: 64 columns each
            // MODE 0: S^T = K Q_i^T, dP^T = V dO_i^T (keys on the lanes); MODE 1: S = Q K_i^T, dP = dO V_i^T
            auto issue_s = [&](const BwdCursor& c) {
                if (c.it == 0) mbar_wait(&fix_full[c.k % NFIX], (c.k / NFIX) & 1);
                mbar_wait(&str_full[c.j % NSTR], (c.j / NSTR) & 1);
                tc_fence_after();
                const uint32_t a_lo = umma_desc_lo(smem_u32(sFix + (c.k % NFIX) * 2 * TILE_BYTES), 16);
                const uint32_t b_lo = umma_desc_lo(smem_u32(sStr + (c.j % NSTR) * 2 * TILE_BYTES), 16);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < D / 16; ++k)      // K-major operands: 32 bytes (2 descriptor units) per K slice
                        umma_bf16_ss_lo(tmem + (c.j & 1) * 128, a_lo + 2 * k, b_lo + 2 * k, idesc_s, k > 0);
                    umma_commit(&s_full[c.j & 1]);
                }
                __syncwarp();
            };
            auto issue_dp = [&](const BwdCursor& c) {       // its fixed pair and stream stage were waited for by issue_s
                const uint32_t a_lo = umma_desc_lo(smem_u32(sFix + (c.k % NFIX) * 2 * TILE_BYTES + TILE_BYTES), 16);
                const uint32_t b_lo = umma_desc_lo(smem_u32(sStr + (c.j % NSTR) * 2 * TILE_BYTES + TILE_BYTES), 16);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < D / 16; ++k) umma_bf16_ss_lo(tmem_dp, a_lo + 2 * k, b_lo + 2 * k, idesc_s, k > 0);
                    umma_commit(dp_full);
                }
                __syncwarp();
            };
            BwdCursor cs = cur_init(), cd = cur_init(), ca = cur_init();
            if (cs.t < total) { issue_s(cs); cur_next(cs); }
            if (cd.t < total) { issue_dp(cd); cur_next(cd); }
            if (cs.t < total) { issue_s(cs); cur_next(cs); }
            for (; ca.t < total; cur_next(ca)) {
                const int bb = ca.j & 1;
                const uint8_t* q0 = sStr + (ca.j % NSTR) * 2 * TILE_BYTES;
                const uint32_t ts = tmem + bb * 128;
                const uint32_t accum = ca.it > 0;
                const bool first = ca.it == 0 && ca.k > 0;
                // MN-major B tiles: 2048 bytes (128 descriptor units) per K slice
                const uint32_t bq_lo = umma_desc_lo(smem_u32(q0), TILE_BYTES);                  // MODE 0: Q_i    MODE 1: K_i
                const uint32_t bo_lo = umma_desc_lo(smem_u32(q0 + TILE_BYTES), TILE_BYTES);     // MODE 0: dO_i
                TRACE_AT(30);
                if (MODE == 0) {
                    mbar_wait(&p_full[bb], (ca.j >> 1) & 1);            // P(j) sits in S[bb]
                    if (first) mbar_wait(acc_free, (ca.k - 1) & 1);     // previous tile's accumulators were read out
                    tc_fence_after();
                    if (elect_one()) {                                  // dV[k, d] += sum_q P^T[k, q] dO[q, d]
#pragma unroll
                        for (int k = 0; k < 8; ++k)
                            umma_bf16_ts_lo(tmem_a0, ts + a_col(k), bo_lo + 128 * k, idesc_ts, (accum || k > 0));
                    }
                    __syncwarp();
                }
                TRACE_AT(31);
                if (cd.t < total) {
                    mbar_wait(dp_free, ca.j & 1);                       // the dS group has pulled dP(j) out of TMEM
                    tc_fence_after();
                    issue_dp(cd); cur_next(cd);
                }
                TRACE_AT(32);
                mbar_wait(&ds_full[bb], (ca.j >> 1) & 1);               // dS(j) sits in S[bb]
                if (MODE == 1 && first) mbar_wait(acc_free, (ca.k - 1) & 1);
                tc_fence_after();
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        // MODE 0: dK[k, d] += sum_q dS^T[k, q] Q[q, d];  MODE 1: dQ[q, d] += sum_k dS[q, k] K[k, d]
                        umma_bf16_ts_lo(MODE == 0 ? tmem_a1 : tmem_a0, ts + a_col(k) + 32, bq_lo + 128 * k, idesc_ts, (accum || k > 0));
                    }
                    umma_commit(&acc_done[bb]);
                    umma_commit(&str_empty[ca.j % NSTR]);               // the stream stage can be refilled
                    if (ca.it == ca.n - 1) umma_commit(&fix_empty[ca.k % NFIX]);
                }
                __syncwarp();
                TRACE_AT(33);
                if (cs.t < total) {
                    mbar_wait(&acc_done[bb], (ca.j >> 1) & 1);          // the MMAs of block j are done with S[bb]
                    TRACE_AT(34);
                    issue_s(cs); cur_next(cs);
                }
                TRACE_AT(35);
            }
        }
        __syncwarp();
    } else if (warp < 8) {
        // ============================ exp group (warps 0-7): S -> P =================================
        // Two threads per tile row: warp w reads TMEM lane quarter (w & 3) and owns column half (w >> 2).  This group
        // only turns scores into probabilities (FFMA + MUFU.EX2 + pack) and writes them back into the S buffer; it is
        // the MUFU-bound stage and never waits for dP or for the accumulating MMAs.
        const int quarter = warp & 3, half = warp >> 2;
        const uint32_t lane_base = static_cast<uint32_t>(quarter * 32) << 16;
        const int r = quarter * 32 + lane;
        auto fetch_lse = [&](const BwdCursor& c) {           // MODE 0: statistics run along the columns (queries)
            return p.lse[(long long)(c.b * p.H + c.h) * p.T + (long long)(c.ob + c.it) * 128 + (tid & 127)];
        };
        BwdCursor c = cur_init(), cn = cur_init();
        if (cn.t < total) cur_next(cn);
        if (MODE == 0 && c.t < total) {
            if (tid < 128) sStat[tid] = fetch_lse(c) * 1.4426950408889634f;
            asm volatile("bar.sync 1, 256;" ::: "memory");
        }
        float lse2 = 0.f;
        for (; c.t < total; cur_next(c)) {
            if (MODE == 1 && c.it == 0)                        // this thread's query row is fixed for the whole tile
                lse2 = p.lse[(long long)(c.b * p.H + c.h) * p.T + c.ob * 128 + r] * 1.4426950408889634f;
            const int bb = c.j & 1;
            const bool diag = MODE == 0 ? (c.it == 0) : (c.it == c.n - 1);
            const uint32_t ts = tmem + bb * 128 + lane_base + half * 64;
            const float4* st_l = reinterpret_cast<const float4*>(sStat + bb * 128 + half * 64);
            float lse_next = 0.f;
            if (MODE == 0 && tid < 128 && cn.t < total) lse_next = fetch_lse(cn);
            TRACE_AT(20);
            mbar_wait(&s_full[bb], (c.j >> 1) & 1);
            tc_fence_after();
            TRACE_AT(21);
            uint32_t rs_[64];
            tmem_ld_32x32(ts, *reinterpret_cast<uint32_t(*)[32]>(&rs_[0]));
            tmem_ld_32x32(ts + 32, *reinterpret_cast<uint32_t(*)[32]>(&rs_[32]));
            tmem_ld_wait();
            if (diag) {                                        // causal mask only in the diagonal block (warp-uniform branch)
#pragma unroll
                for (int k = 0; k < 64; ++k) {
                    const int cc = half * 64 + k;              // MODE 0: row = key, column = query; MODE 1: the transpose
                    if (MODE == 0 ? (cc < r) : (cc > r)) rs_[k] = 0xff800000u;   // -inf -> P = 0
                }
            }
            uint32_t pk[32];
#pragma unroll
            for (int k4 = 0; k4 < 16; ++k4) {
                float4 l4 = make_float4(lse2, lse2, lse2, lse2);
                if (MODE == 0) l4 = st_l[k4];
                const float e0 = exp2f(fmaf(__uint_as_float(rs_[k4 * 4 + 0]), p.scale_log2, -l4.x));
                const float e1 = exp2f(fmaf(__uint_as_float(rs_[k4 * 4 + 1]), p.scale_log2, -l4.y));
                const float e2 = exp2f(fmaf(__uint_as_float(rs_[k4 * 4 + 2]), p.scale_log2, -l4.z));
                const float e3 = exp2f(fmaf(__uint_as_float(rs_[k4 * 4 + 3]), p.scale_log2, -l4.w));
                pk[k4 * 2 + 0] = pack_bf16x2(e0, e1);
                pk[k4 * 2 + 1] = pack_bf16x2(e2, e3);
            }
            tmem_st_32x32(ts, pk);                             // P (bf16 pairs along the K index) -> S[bb] columns [64h, 64h+32)
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[bb]);
            TRACE_AT(22);
            if (MODE == 0) {                                   // park the next block's lse; everyone is done with this block's
                if (tid < 128) sStat[(bb ^ 1) * 128 + tid] = lse_next * 1.4426950408889634f;
                asm volatile("bar.sync 1, 256;" ::: "memory");
            }
            if (cn.t < total) cur_next(cn);
        }
    } else if (warp < 16) {
        // ============================ dS group (warps 8-15): (P, dP) -> dS, and the tile epilogue ====
        // Same row / column ownership as the exp group.  P comes back from tensor memory as the bf16 the MMA consumes,
        // dS = P (dP - delta) goes next to it.  The group has slack (no MUFU work), so it also computes delta for the
        // dQ kernel and writes finished accumulator tiles out.
        const int gw = warp - 8, lt = tid - 256;
        const int quarter = gw & 3, half = gw >> 2;
        const uint32_t lane_base = static_cast<uint32_t>(quarter * 32) << 16;
        const int r = quarter * 32 + lane;
        float* sDel = sStat + 256;                            // MODE 0: delta [2][128]; MODE 1: partial sums [2][128]
        auto fetch_delta = [&](const BwdCursor& c) {
            return p.delta[(long long)(c.b * p.H + c.h) * p.T + (long long)(c.ob + c.it) * 128 + (lt & 127)];
        };
        // ---- finished tile: accumulators -> bf16 -> swizzled staging tiles -> TMA store; the accumulators are released
        //      (acc_free) as soon as they are in registers
        auto write_tile = [&](const BwdCursor& c) {
            const int bb = c.j & 1;
            mbar_wait(&acc_done[bb], (c.j >> 1) & 1);
            __syncwarp();
            tc_fence_after();
            TRACE_AT(27);
            if (lt == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // previous tile's stores have read the staging
            asm volatile("bar.sync 2, 256;" ::: "memory");
            auto stage_acc = [&](uint32_t taddr, uint8_t* tile, float mul) {
                uint32_t raw[32];                               // each of the two row threads handles 32 of the 64 columns
                tmem_ld_32x32(taddr + lane_base + half * 32, raw);
                tmem_ld_wait();
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    uint4 u;
                    u.x = pack_bf16x2(__uint_as_float(raw[q * 8 + 0]) * mul, __uint_as_float(raw[q * 8 + 1]) * mul);
                    u.y = pack_bf16x2(__uint_as_float(raw[q * 8 + 2]) * mul, __uint_as_float(raw[q * 8 + 3]) * mul);
                    u.z = pack_bf16x2(__uint_as_float(raw[q * 8 + 4]) * mul, __uint_as_float(raw[q * 8 + 5]) * mul);
                    u.w = pack_bf16x2(__uint_as_float(raw[q * 8 + 6]) * mul, __uint_as_float(raw[q * 8 + 7]) * mul);
                    *reinterpret_cast<uint4*>(tile + swz(r, half * 4 + q)) = u;
                }
            };
            if (MODE == 0) {
                stage_acc(tmem_a1, sOut, p.scale);                    // dK
                stage_acc(tmem_a0, sOut + TILE_BYTES, 1.0f);          // dV
            } else {
                stage_acc(tmem_a0, sOut, p.scale);                    // dQ
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_free);              // the accumulators are in registers / shared memory
            fence_proxy_async_smem();
            asm volatile("bar.sync 2, 256;" ::: "memory");
            if (lt == 0) {
                const int grow = c.b * p.T + c.ob * 128;
                if (MODE == 0) {
                    tma_store_2d_tile(&tmap_dqkv, sOut, p.C + c.h * D, grow);
                    tma_store_2d_tile(&tmap_dqkv, sOut + TILE_BYTES, 2 * p.C + c.h * D, grow);
                } else {
                    tma_store_2d_tile(&tmap_dqkv, sOut, c.h * D, grow);
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            if (p.dbias != nullptr && (MODE == 0 || lt < 128)) {
                // qkv bias gradient = column sums of the staged bf16 tiles (they stay intact until the next tile's first
                // barrier).  128 threads per tile: thread = 16-byte piece (t & 7) of rows (t >> 3) + 16 i; the four row
                // groups of a warp combine by shuffle, the four warps of a tile by red.global.
                const int t = lt & 127, which = lt >> 7;              // MODE 0: tile 0 = dK, tile 1 = dV
                const uint8_t* tile = sOut + which * TILE_BYTES;
                const uint32_t q8 = t & 7;
                float cs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const uint4 u = *reinterpret_cast<const uint4*>(tile + swz((t >> 3) + 16 * i, q8));
                    const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y), f2 = unpack_bf16x2(u.z), f3 = unpack_bf16x2(u.w);
                    cs[0] += f0.x; cs[1] += f0.y; cs[2] += f1.x; cs[3] += f1.y;
                    cs[4] += f2.x; cs[5] += f2.y; cs[6] += f3.x; cs[7] += f3.y;
                }
#pragma unroll
                for (int j2 = 0; j2 < 8; ++j2) {
                    cs[j2] += __shfl_xor_sync(0xffffffffu, cs[j2], 8);
                    cs[j2] += __shfl_xor_sync(0xffffffffu, cs[j2], 16);
                }
                if (lane < 8) {
                    float* dst = p.dbias + c.h * D + (MODE == 0 ? (1 + which) * p.C : 0) + q8 * 8;
                    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(dst), "f"(cs[0]), "f"(cs[1]), "f"(cs[2]), "f"(cs[3]) : "memory");
                    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(dst + 4), "f"(cs[4]), "f"(cs[5]), "f"(cs[6]), "f"(cs[7]) : "memory");
                }
            }
            TRACE_AT(28);
        };
        BwdCursor c = cur_init(), cn = cur_init();
        if (cn.t < total) cur_next(cn);
        if (MODE == 0 && c.t < total) {
            if (lt < 128) sDel[lt] = fetch_delta(c);
            asm volatile("bar.sync 3, 256;" ::: "memory");
        }
        float dlt = 0.f;
        for (; c.t < total; cur_next(c)) {
            if (MODE == 1 && c.it == 0) {
                // delta = rowsum(dO * O) of this thread's query row, from the two tiles in shared memory (it used to be a
                // separate kernel and a separate pass over O and dO); written out for the dK/dV kernel that runs next
                const int fb = c.k % NFIX;
                mbar_wait(&fix_full[fb], (c.k / NFIX) & 1);
                mbar_wait(o_full, c.k & 1);
                const uint8_t* tdo = sFix + fb * 2 * TILE_BYTES + TILE_BYTES;
                const uint8_t* to = sOut + TILE_BYTES;
                float part = 0.f;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const uint32_t piece = half * 4 + q;
                    const uint4 ua = *reinterpret_cast<const uint4*>(to + swz(r, piece));
                    const uint4 ug = *reinterpret_cast<const uint4*>(tdo + swz(r, piece));
                    const uint32_t wa[4] = {ua.x, ua.y, ua.z, ua.w}, wg[4] = {ug.x, ug.y, ug.z, ug.w};
#pragma unroll
                    for (int j2 = 0; j2 < 4; ++j2) {
                        const float2 fa = unpack_bf16x2(wa[j2]), fg = unpack_bf16x2(wg[j2]);
                        part = fmaf(fa.x, fg.x, part);
                        part = fmaf(fa.y, fg.y, part);
                    }
                }
                sDel[half * 128 + r] = part;
                __syncwarp();
                if (lane == 0) mbar_arrive(o_free);
                asm volatile("bar.sync 3, 256;" ::: "memory");
                dlt = sDel[r] + sDel[128 + r];
                if (half == 0) const_cast<float*>(p.delta)[(long long)(c.b * p.H + c.h) * p.T + c.ob * 128 + r] = dlt;
            }
            const int bb = c.j & 1;
            const uint32_t ts = tmem + bb * 128 + lane_base + half * 64;
            const float4* st_d = reinterpret_cast<const float4*>(sDel + bb * 128 + half * 64);
            float del_next = 0.f;
            if (MODE == 0 && lt < 128 && cn.t < total) del_next = fetch_delta(cn);
            TRACE_AT(23);
            mbar_wait(&p_full[bb], (c.j >> 1) & 1);            // P(j) is in S[bb] (written by the exp group)
            mbar_wait(dp_full, c.j & 1);
            tc_fence_after();
            TRACE_AT(24);
#pragma unroll
            for (int ch = 0; ch < 2; ++ch) {
                uint32_t pp[16], rd_[32];
                tmem_ld_32x16(ts + ch * 16, pp);               // 32 probabilities of this chunk, as bf16 pairs
                tmem_ld_32x32(tmem_dp + lane_base + half * 64 + ch * 32, rd_);
                tmem_ld_wait();
                if (ch == 1) {                                 // dP(j) is in registers: the tensor core may overwrite it
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(dp_free);
                }
                uint32_t pk[16];
#pragma unroll
                for (int k4 = 0; k4 < 8; ++k4) {
                    float4 d4 = make_float4(dlt, dlt, dlt, dlt);
                    if (MODE == 0) d4 = st_d[ch * 8 + k4];
                    const float2 p01 = unpack_bf16x2(pp[k4 * 2 + 0]), p23 = unpack_bf16x2(pp[k4 * 2 + 1]);
                    const float s0 = p01.x * (__uint_as_float(rd_[k4 * 4 + 0]) - d4.x);
                    const float s1 = p01.y * (__uint_as_float(rd_[k4 * 4 + 1]) - d4.y);
                    const float s2 = p23.x * (__uint_as_float(rd_[k4 * 4 + 2]) - d4.z);
                    const float s3 = p23.y * (__uint_as_float(rd_[k4 * 4 + 3]) - d4.w);
                    pk[k4 * 2 + 0] = pack_bf16x2(s0, s1);
                    pk[k4 * 2 + 1] = pack_bf16x2(s2, s3);
                }
                tmem_st_32x16(ts + 32 + ch * 16, pk);          // dS -> S[bb] columns [64h+32, 64h+64)
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ds_full[bb]);
            TRACE_AT(25);
            if (MODE == 0) {                                   // park the next block's delta; everyone is done with this block's
                if (lt < 128) sDel[(bb ^ 1) * 128 + lt] = del_next;
                asm volatile("bar.sync 3, 256;" ::: "memory");
            }
            if (c.it == c.n - 1) write_tile(c);
            if (cn.t < total) cur_next(cn);
        }
        if (lt == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");      // staging tiles are read before the CTA exits
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}


}  // namespace

int tdl_make_tmap_2d_bf16(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer, uint64_t ld_elems,
                          uint32_t box_inner, uint32_t box_outer);

static int check_attn_shapes(int64_t B, int64_t T, int64_t H, int64_t Dh) {
    TDL_REQUIRE(Dh == 64, "attention kernels are specialised for head_dim 64 (got %lld)", (long long)Dh);
    TDL_REQUIRE(T % 128 == 0 && T > 0, "sequence length must be a multiple of 128 (got %lld)", (long long)T);
    TDL_REQUIRE(B > 0 && H > 0 && B <= 65535 && H <= 65535, "bad batch/head count");
    return 0;
}

extern "C" int tdl_attention_fwd(const void* qkv, void* out, float* lse, int64_t B, int64_t T, int64_t H, int64_t Dh,
                                 void* stream) {
    if (int rc = check_attn_shapes(B, T, H, Dh)) return rc;
    const int C = (int)(H * 64);
    CUtensorMap tm;
    if (int rc = tdl_make_tmap_2d_bf16(&tm, qkv, 3ull * C, (uint64_t)(B * T), 3ull * C, 64, 128)) return rc;
    AttnParams p{};
    p.B = (int)B; p.T = (int)T; p.H = (int)H; p.C = C;
    p.scale = 1.0f / sqrtf(64.0f);
    p.scale_log2 = p.scale * 1.4426950408889634f;
    p.out = (__nv_bfloat16*)out; p.lse = lse;
    const int smem = 5 * TILE_BYTES + 64;                 // Q, 2 x K, 2 x V (P lives in tensor memory)
    static bool configured = false;
    if (!configured) {
        TDL_CHECK(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured = true;
    }
    TDL_CHECK(tdl_launch_pdl(attn_fwd_kernel, dim3((unsigned)(T / 128), (unsigned)H, (unsigned)B), dim3(128), smem,
                             reinterpret_cast<cudaStream_t>(stream), tm, p));
    TDL_LAUNCH_CHECK();
    return 0;
}

// dbias (optional, fp32 [3C]): += column sums of dqkv, i.e. the qkv projection's bias gradient, taken from the
// accumulator tiles as they are written instead of re-reading dqkv.
extern "C" int tdl_attention_bwd_dbias(const void* qkv, const void* out, const void* d_out, const float* lse, float* delta,
                                       void* dqkv, float* dbias, int64_t B, int64_t T, int64_t H, int64_t Dh, void* stream) {
    if (int rc = check_attn_shapes(B, T, H, Dh)) return rc;
    const int C = (int)(H * 64);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    CUtensorMap tq, tdo;
    if (int rc = tdl_make_tmap_2d_bf16(&tq, qkv, 3ull * C, (uint64_t)(B * T), 3ull * C, 64, 128)) return rc;
    if (int rc = tdl_make_tmap_2d_bf16(&tdo, d_out, (uint64_t)C, (uint64_t)(B * T), (uint64_t)C, 64, 128)) return rc;
    AttnParams p{};
    p.B = (int)B; p.T = (int)T; p.H = (int)H; p.C = C;
    p.scale = 1.0f / sqrtf(64.0f);
    p.scale_log2 = p.scale * 1.4426950408889634f;
    p.lse = const_cast<float*>(lse); p.delta = delta; p.dqkv = (__nv_bfloat16*)dqkv; p.dbias = dbias;
    TDL_REQUIRE((reinterpret_cast<uintptr_t>(dbias) & 15) == 0, "attention backward: dbias must be 16-byte aligned");
    const int smem_ws = BWD_WS_SMEM;
    static bool configured = false;
    if (!configured) {
        TDL_CHECK(cudaFuncSetAttribute(attn_bwd_ps_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_ws));
        TDL_CHECK(cudaFuncSetAttribute(attn_bwd_ps_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_ws));
        configured = true;
    }
    const long long tiles = (T / 128) * B * H;
    long long ch = (96 * 1024) / T;                      // 4 tensors x T x 64 x 2 B per pair: keep a chunk under ~48 MB
    ch = ch < 1 ? 1 : (ch > B * H ? B * H : ch);
    while ((B * H) % ch) --ch;                           // chunks are all the same size
    p.bwd_chunk = (int)ch;
    TDL_REQUIRE(tiles < (1 << 20) && H < (1 << 20), "attention backward: too many tiles for the persistent schedule");
    auto magic = [](long long d) { return (unsigned long long)(((1ull << 40) + d - 1) / d); };
    p.mg_tile = magic((T / 128) * ch); p.mg_chunk = magic(ch); p.mg_head = magic(H);
    const unsigned ctas = (unsigned)(tiles < tdl_num_sms() ? tiles : tdl_num_sms());      // persistent: one CTA per SM
    CUtensorMap tdq;
    if (int rc = tdl_make_tmap_2d_bf16(&tdq, dqkv, 3ull * C, (uint64_t)(B * T), 3ull * C, 64, 128)) return rc;
    CUtensorMap to;
    if (int rc = tdl_make_tmap_2d_bf16(&to, out, (uint64_t)C, (uint64_t)(B * T), (uint64_t)C, 64, 128)) return rc;
    // dQ first: it also produces delta = rowsum(dO * O), which the dK/dV kernel reads
    TDL_CHECK(tdl_launch_pdl(attn_bwd_ps_kernel<1>, dim3(ctas), dim3(BWD_WS_THREADS), smem_ws, s, tq, tdo, tdq, to, p));
    TDL_CHECK(tdl_launch_pdl(attn_bwd_ps_kernel<0>, dim3(ctas), dim3(BWD_WS_THREADS), smem_ws, s, tq, tdo, tdq, to, p));
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_attention_bwd(const void* qkv, const void* out, const void* d_out, const float* lse, float* delta,
                                 void* dqkv, int64_t B, int64_t T, int64_t H, int64_t Dh, void* stream) {
    return tdl_attention_bwd_dbias(qkv, out, d_out, lse, delta, dqkv, nullptr, B, T, H, Dh, stream);
}

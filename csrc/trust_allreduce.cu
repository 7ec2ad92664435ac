// K2 / K3 / K4: trust-weighted gradient all-reduce with the anomaly statistics fused into the
// reduction, over peer-mapped (symmetric) memory on NVLink 5 / NVSwitch.
//
// "Gradient aggregation" is only named by the reference (README.md:31, distributed_trainer.py:442)
// and its gradient checks are a post-hoc host pass (attack_detector.py:109-141, 225-239).  Here:
//
//   two-shot reduction: rank r owns slice r of the bucket.  It P2P-loads that slice from every peer
//   (16-byte vectors; individual values are needed, so the in-switch NVLS sum cannot be used),
//   and per element runs a W-input sorting network -> coordinate-wise median, accumulates
//   dot(g_i, median) per peer and |median|^2, forms sum_i c_i g_i with the per-rank coefficients
//   c_i (trust weight x norm validity x clip), and stores the reduced vector into every peer's
//   bucket.  Wire cost equals a plain two-shot all-reduce: (W-1)/W x bytes each way.
//
//   cross-rank synchronisation is kept out of the math kernel: a tiny "exchange" kernel all-gathers
//   a few doubles per rank through the peers' pads and doubles as the barrier (st.release.sys flag
//   per peer, ld.acquire.sys spin, monotonically increasing epochs kept on the device so the whole
//   sequence can be replayed from a CUDA graph).  With ranks emulated on one GPU the math kernel is
//   launched once per emulated rank and the exchange kernel is not used (B200_PROFILING.md: never
//   run kernels that wait on each other on one device).
//
//   coefficients and the trust state machine run on the device (K3): the trust vector never leaves
//   HBM on the step path; the host mirrors it lazily for logging.
#include "common.cuh"

namespace {

constexpr int MAX_W = 8;
constexpr int EXCH_MAX = 16;                 // doubles per rank per exchange

struct PeerPtrs { void* p[MAX_W]; };

template <typename T> struct Vec;
template <> struct Vec<__nv_bfloat16> { static constexpr int N = 8; };
template <> struct Vec<float> { static constexpr int N = 4; };

template <typename T> __device__ __forceinline__ void unpack_vec(const uint4& u, float* f);
template <> __device__ __forceinline__ void unpack_vec<__nv_bfloat16>(const uint4& u, float* f) {
    float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
template <> __device__ __forceinline__ void unpack_vec<float>(const uint4& u, float* f) {
    f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y); f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
}
template <typename T> __device__ __forceinline__ uint4 pack_vec(const float* f);
template <> __device__ __forceinline__ uint4 pack_vec<__nv_bfloat16>(const float* f) {
    uint4 u; u.x = pack_bf16x2(f[0], f[1]); u.y = pack_bf16x2(f[2], f[3]); u.z = pack_bf16x2(f[4], f[5]); u.w = pack_bf16x2(f[6], f[7]);
    return u;
}
template <> __device__ __forceinline__ uint4 pack_vec<float>(const float* f) {
    uint4 u; u.x = __float_as_uint(f[0]); u.y = __float_as_uint(f[1]); u.z = __float_as_uint(f[2]); u.w = __float_as_uint(f[3]);
    return u;
}

__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {       // plain (coherent) 16-byte load
    uint4 r;
    asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void st_peer_v4(void* p, const uint4& v) {
    asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <int W> __device__ __forceinline__ float median_w(float (&a)[W]) {
    // odd-even transposition network: W rounds sort W values held in registers
#pragma unroll
    for (int round = 0; round < W; ++round) {
#pragma unroll
        for (int i = round & 1; i + 1 < W; i += 2) {
            const float lo = fminf(a[i], a[i + 1]), hi = fmaxf(a[i], a[i + 1]);
            a[i] = lo; a[i + 1] = hi;
        }
    }
    return (W & 1) ? a[W / 2] : 0.5f * (a[W / 2 - 1] + a[W / 2]);
}

// NVLS store: one 16-byte store to the multicast address lands in every GPU bound to the multicast object
template <typename T> __device__ __forceinline__ void st_multimem_v4(void* p, const uint4& v);
template <> __device__ __forceinline__ void st_multimem_v4<__nv_bfloat16>(void* p, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
template <> __device__ __forceinline__ void st_multimem_v4<float>(void* p, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)),
                 "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w)) : "memory");
}

// stats_partial: [W] dot(g_i, med) then [1] |med|^2   (doubles, accumulated with atomics)
// store_rank < 0: the reduced vector goes to every peer (W stores, or ONE multimem.st when `mc_out` is the bucket's
// multicast address); store_rank >= 0: only to out.p[store_rank] (ZeRO-1: the owner alone consumes its slice).
template <typename T, int W, int TPB = 256>
__global__ void __launch_bounds__(TPB) trust_reduce_kernel(PeerPtrs in, PeerPtrs out, long long vec_begin, long long vec_end,
                                                           const float* __restrict__ coeff, double* __restrict__ stats_partial,
                                                           int store_rank, void* mc_out) {
    constexpr int VN = Vec<T>::N;
    __shared__ double sh[TPB / 32][W + 1];
    float c[W];
#pragma unroll
    for (int i = 0; i < W; ++i) c[i] = coeff[i];
    float dot[W];
#pragma unroll
    for (int i = 0; i < W; ++i) dot[i] = 0.f;
    float med2 = 0.f;
    double ddot[W], dmed2 = 0.0;
#pragma unroll
    for (int i = 0; i < W; ++i) ddot[i] = 0.0;
    int run = 0;
    // U vectors per thread per trip: W * U peer loads are in flight before anything is consumed.  Loaded NVLink latency is
    // tens of microseconds (a plain peer copy needs ~100 KB in flight per SM to reach 730 GB/s), so one vector per trip
    // leaves the links mostly idle.
    constexpr int U = W <= 2 ? 8 : (W <= 4 ? 4 : 2);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v0 = vec_begin + (long long)blockIdx.x * blockDim.x + threadIdx.x; v0 < vec_end; v0 += stride * U) {
        uint4 rawu[U][W];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long v = v0 + u * stride;
#pragma unroll
            for (int i = 0; i < W; ++i)
                rawu[u][i] = v < vec_end ? ld_peer_v4(reinterpret_cast<const uint4*>(in.p[i]) + v) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
        const long long v = v0 + u * stride;
        if (v < vec_end) {
        uint4 (&raw)[W] = rawu[u];
        float g[W][VN];
#pragma unroll
        for (int i = 0; i < W; ++i) {
            unpack_vec<T>(raw[i], g[i]);
            // a rank that ships NaN/Inf gets coefficient 0 (non-finite norm), but 0 * NaN is still NaN: non-finite
            // elements are read as 0 so they can reach neither the aggregate nor the median / dot statistics
#pragma unroll
            for (int e = 0; e < VN; ++e) g[i][e] = fabsf(g[i][e]) <= 3.3895314e38f ? g[i][e] : 0.f;
        }
        float o[VN];
#pragma unroll
        for (int e = 0; e < VN; ++e) {
            float col[W];
            float acc = 0.f;
#pragma unroll
            for (int i = 0; i < W; ++i) { col[i] = g[i][e]; acc = fmaf(c[i], g[i][e], acc); }
            const float m = median_w<W>(col);
            med2 = fmaf(m, m, med2);
#pragma unroll
            for (int i = 0; i < W; ++i) dot[i] = fmaf(g[i][e], m, dot[i]);
            o[e] = acc;
        }
        const uint4 packed = pack_vec<T>(o);
        if (store_rank >= 0) {
            st_peer_v4(reinterpret_cast<uint4*>(out.p[store_rank]) + v, packed);
        } else if (mc_out != nullptr) {
            st_multimem_v4<T>(reinterpret_cast<uint4*>(mc_out) + v, packed);
        } else {
#pragma unroll
            for (int i = 0; i < W; ++i) st_peer_v4(reinterpret_cast<uint4*>(out.p[i]) + v, packed);
        }
        if (++run == 32) {
#pragma unroll
            for (int i = 0; i < W; ++i) { ddot[i] += dot[i]; dot[i] = 0.f; }
            dmed2 += med2; med2 = 0.f; run = 0;
        }
        }
        }
    }
#pragma unroll
    for (int i = 0; i < W; ++i) ddot[i] += dot[i];
    dmed2 += med2;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < W; ++i) {
        double t = ddot[i];
        for (int o2 = 16; o2 > 0; o2 >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o2);
        if (lane == 0) sh[warp][i] = t;
    }
    {
        double t = dmed2;
        for (int o2 = 16; o2 > 0; o2 >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o2);
        if (lane == 0) sh[warp][W] = t;
    }
    __syncthreads();
    if (threadIdx.x <= W) {
        double t = 0;
        for (int w2 = 0; w2 < TPB / 32; ++w2) t += sh[w2][threadIdx.x];
        atomicAdd(&stats_partial[threadIdx.x], t);
    }
}

// ------------------------------------------------------------------ K4: plain sum all-reduce (no-security arm)
// Two-shot over the same symmetric bucket.  NVLS form: multimem.ld_reduce pulls the slice already summed inside the
// switch (fp32 accumulation), the scaled result goes back with one multimem.st.  P2P form (no multicast object): W
// peer loads + W peer stores, i.e. K2 without the median network -- the arm the detection overhead is measured against.
template <typename T> __device__ __forceinline__ uint4 ld_reduce_multimem_v4(const void* p);
template <> __device__ __forceinline__ uint4 ld_reduce_multimem_v4<__nv_bfloat16>(const void* p) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
template <> __device__ __forceinline__ uint4 ld_reduce_multimem_v4<float>(const void* p) {
    float a, b, c, d;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(a), "=f"(b), "=f"(c), "=f"(d) : "l"(p) : "memory");
    return make_uint4(__float_as_uint(a), __float_as_uint(b), __float_as_uint(c), __float_as_uint(d));
}

template <typename T, bool NVLS>
__global__ void __launch_bounds__(256) sum_allreduce_kernel(PeerPtrs bufs, void* mc, int W, long long vec_begin, long long vec_end,
                                                            float scale) {
    constexpr int VN = Vec<T>::N;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long v = vec_begin + (long long)blockIdx.x * blockDim.x + threadIdx.x; v < vec_end; v += stride) {
        float acc[VN];
        if (NVLS) {
            const uint4 u = ld_reduce_multimem_v4<T>(reinterpret_cast<const uint4*>(mc) + v);
            unpack_vec<T>(u, acc);
        } else {
#pragma unroll
            for (int e = 0; e < VN; ++e) acc[e] = 0.f;
            for (int i = 0; i < W; ++i) {
                float g[VN];
                unpack_vec<T>(ld_peer_v4(reinterpret_cast<const uint4*>(bufs.p[i]) + v), g);
#pragma unroll
                for (int e = 0; e < VN; ++e) acc[e] += g[e];
            }
        }
#pragma unroll
        for (int e = 0; e < VN; ++e) acc[e] *= scale;
        const uint4 packed = pack_vec<T>(acc);
        if (NVLS) st_multimem_v4<T>(reinterpret_cast<uint4*>(mc) + v, packed);
        else for (int i = 0; i < W; ++i) st_peer_v4(reinterpret_cast<uint4*>(bufs.p[i]) + v, packed);
    }
}

// ------------------------------------------------------------------ exchange (= tiny all-gather + barrier)
// pads[i]  : peer i's pad: uint32 flags[MAX_W] at byte 0, doubles gather[MAX_W][EXCH_MAX] at byte 64+slot*...
// my rank writes its `count` doubles into row `rank` of every peer's gather area `which`, then raises
// flag[rank] on every peer to the new epoch and waits until its own flags[*] reached that epoch.
struct ExchPad {
    unsigned int flags[2][MAX_W];                 // one flag row per gather area
    double gather[2][MAX_W][EXCH_MAX];
};

__global__ void exchange_kernel(PeerPtrs pads, int rank, int W, int which, const double* __restrict__ vals, int count,
                                unsigned int* __restrict__ epoch_counter, int zero_vals) {
    __shared__ unsigned int s_epoch;
    if (threadIdx.x == 0) s_epoch = ++epoch_counter[which];
    __syncthreads();
    const unsigned int epoch = s_epoch;
    const int t = threadIdx.x;
    if (t < W) {
        ExchPad* peer = reinterpret_cast<ExchPad*>(pads.p[t]);
        for (int k = 0; k < count; ++k) peer->gather[which][rank][k] = vals[k];
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(&peer->flags[which][rank]), "r"(epoch) : "memory");
        const ExchPad* mine = reinterpret_cast<const ExchPad*>(pads.p[rank]);
        unsigned int seen;
        // A peer that never raises its flag (crashed or wedged process) must not hang this GPU for good: after ~2 minutes of
        // cycles the kernel traps with the missing rank named, which the host sees as a launch failure (liveness signal).
        const long long t0 = clock64();
        unsigned int spins = 0;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(&mine->flags[which][t]) : "memory");
            if ((int)(seen - epoch) < 0 && (++spins & 0xFFFu) == 0 && clock64() - t0 > 200000000000ll) {
                printf("tdl: exchange %d: rank %d never reached epoch %u (seen %u) -- peer lost?\n", which, t, epoch, seen);
                __trap();
            }
        } while ((int)(seen - epoch) < 0);
    }
    __syncthreads();
    if (zero_vals && t < count) const_cast<double*>(vals)[t] = 0.0;      // accumulators start at zero next time
}

// ------------------------------------------------------------------ K3: coefficients + trust state machine
struct TrustParams {
    float alpha, threshold, compromised_below, recovering_above, retrust_above, penalty;
    float clip_ratio, norm_ratio_max, cosine_min;
    float w_dev, w_gc, base_score;          // base_score: contribution of the four metrics held at their defaults
};

// norms_sq[W] (from the exchange pad), trust[W], status[W] -> coeff[W]
// acc_sq[W] / acc_bad[W] (optional): a step that reduces several buckets accumulates the squared norms and the
// "rejected in some bucket" marks here; finalize then judges the step as a whole.
__global__ void coefficients_kernel(const double* __restrict__ gathered /*[W][EXCH_MAX], col 0 = sumsq*/, int W,
                                    const float* __restrict__ trust, const int* __restrict__ status,
                                    TrustParams tp, float* __restrict__ coeff, float* __restrict__ norms_out,
                                    double* __restrict__ acc_sq, int* __restrict__ acc_bad) {
    if (threadIdx.x != 0) return;
    if (acc_sq != nullptr) for (int i = 0; i < W; ++i) acc_sq[i] += gathered[i * EXCH_MAX];
    float n[MAX_W], w[MAX_W];
    bool live[MAX_W];
    float sorted[MAX_W];
    int k = 0;
    for (int i = 0; i < W; ++i) {
        n[i] = sqrtf((float)gathered[i * EXCH_MAX]);
        norms_out[i] = n[i];
        w[i] = (status[i] == 2 || status[i] == 4) ? 0.f : trust[i];       // COMPROMISED / OFFLINE contribute nothing
        live[i] = w[i] > 0.f && isfinite(n[i]);
        if (live[i]) sorted[k++] = n[i];
    }
    if (k == 0) { for (int i = 0; i < W; ++i) { coeff[i] = 0.f; if (acc_bad != nullptr && w[i] > 0.f) acc_bad[i] = 1; } return; }
    for (int a = 1; a < k; ++a) { float v = sorted[a]; int b = a - 1; while (b >= 0 && sorted[b] > v) { sorted[b + 1] = sorted[b]; --b; } sorted[b + 1] = v; }
    const float med = sorted[(k - 1) / 2];       // lower median: robust against one exploding norm even for k = 2
    float tot = 0.f;
    bool valid[MAX_W];
    for (int i = 0; i < W; ++i) { valid[i] = live[i] && (n[i] <= tp.norm_ratio_max * med || med <= 0.f); if (valid[i]) tot += w[i]; }
    for (int i = 0; i < W; ++i) {
        float c = 0.f;
        if (valid[i] && tot > 0.f) {
            float clip = 1.f;
            if (tp.clip_ratio > 0.f && med > 0.f) clip = fminf(1.f, tp.clip_ratio * med / fmaxf(n[i], 1e-30f));
            c = w[i] * clip / tot;
        }
        coeff[i] = c;
        if (acc_bad != nullptr && !valid[i] && w[i] > 0.f) acc_bad[i] = 1;       // live rank rejected in this bucket
    }
}

// gathered partials [W ranks][W dots + med2] -> stats[W][4] = (norm, dot, |med|^2, cos); then the trust EMA + FSM.
// flags_out[i]: bit0 = failed verification this step (norm ratio / cosine / non-finite).
// acc_sq / acc_bad (optional, see coefficients_kernel): whole-step norms come from the accumulated squared bucket
// norms, a rank rejected in any bucket fails verification; both are cleared for the next step.
__global__ void finalize_trust_kernel(const double* __restrict__ gathered, int W, const float* __restrict__ norms_in,
                                      const float* __restrict__ out_dev, TrustParams tp, float* __restrict__ stats,
                                      float* __restrict__ trust, int* __restrict__ status, int* __restrict__ flags_out,
                                      int update_trust, double* __restrict__ acc_sq, int* __restrict__ acc_bad) {
    if (threadIdx.x != 0) return;
    float norms[MAX_W];
    bool bad[MAX_W];
    for (int i = 0; i < W; ++i) {
        norms[i] = acc_sq != nullptr ? sqrtf((float)acc_sq[i]) : norms_in[i];
        bad[i] = acc_bad != nullptr && acc_bad[i] != 0;
        if (acc_sq != nullptr) acc_sq[i] = 0.0;
        if (acc_bad != nullptr) acc_bad[i] = 0;
    }
    double med2 = 0.0, dot[MAX_W];
    for (int i = 0; i < W; ++i) dot[i] = 0.0;
    for (int r = 0; r < W; ++r) {
        for (int i = 0; i < W; ++i) dot[i] += gathered[r * EXCH_MAX + i];
        med2 += gathered[r * EXCH_MAX + W];
    }
    const float mn = sqrtf((float)med2);
    float srt[MAX_W]; int k = 0;
    for (int i = 0; i < W; ++i) if (isfinite(norms[i])) srt[k++] = norms[i];
    for (int a = 1; a < k; ++a) { float v = srt[a]; int b = a - 1; while (b >= 0 && srt[b] > v) { srt[b + 1] = srt[b]; --b; } srt[b + 1] = v; }
    const float ref = k ? srt[(k - 1) / 2] : 0.f;
    // gradient-consistency input of the trust formula is the cosine *relative to the typical rank*: with noisy
    // mini-batch gradients honest ranks all sit at a modest absolute cosine to the median, and must not lose
    // trust for that; a rank far below the typical value does.
    float cs[MAX_W]; int kc = 0;
    for (int i = 0; i < W; ++i) {
        const float n = norms[i];
        const float c0 = (n > 0.f && mn > 0.f) ? (float)(dot[i] / ((double)n * mn)) : 1.f;
        if (isfinite(c0)) cs[kc++] = c0;
    }
    for (int a = 1; a < kc; ++a) { float v = cs[a]; int b = a - 1; while (b >= 0 && cs[b] > v) { cs[b + 1] = cs[b]; --b; } cs[b + 1] = v; }
    const float med_cos = kc ? cs[(kc - 1) / 2] : 1.f;
    for (int i = 0; i < W; ++i) {
        const float n = norms[i];
        const float cosv = (n > 0.f && mn > 0.f) ? (float)(dot[i] / ((double)n * mn)) : 1.f;
        stats[i * 4 + 0] = n; stats[i * 4 + 1] = (float)dot[i]; stats[i * 4 + 2] = (float)med2; stats[i * 4 + 3] = cosv;
        bool ok = isfinite(n) && isfinite(cosv) && !bad[i];
        if (W >= 3) ok = ok && cosv >= tp.cosine_min;
        if (W >= 2 && ref > 0.f) ok = ok && n <= tp.norm_ratio_max * ref;
        int fl = ok ? 0 : 1;
        if (update_trust) {
            if (!ok && status[i] != 2) { trust[i] = tp.penalty; status[i] = 2; fl |= 2; }      // mark_compromised
            const float dev = out_dev != nullptr ? out_dev[i] : 0.f;
            const float gc = !ok ? 0.f : (W >= 3 && med_cos > 0.f ? fminf(fmaxf(cosv / med_cos, 0.f), 1.f) : 1.f);
            const float inst = fminf(fmaxf(tp.w_dev * (1.f - fminf(1.f, dev)) + tp.w_gc * gc + tp.base_score, 0.f), 1.f);
            const float t = fminf(fmaxf((1.f - tp.alpha) * trust[i] + tp.alpha * inst, 0.f), 1.f);
            trust[i] = t;
            const int cur = status[i];
            int nxt = cur;
            if (t < tp.compromised_below) nxt = 2;
            else if (t < tp.threshold) nxt = 1;
            else if (cur == 2 && t > tp.recovering_above) nxt = 3;
            else if (cur == 3 && t > tp.retrust_above) nxt = 0;
            else if (t >= tp.threshold) nxt = 0;
            status[i] = nxt;
        }
        flags_out[i] = fl;
    }
}

}  // namespace

#define STREAM(s) reinterpret_cast<cudaStream_t>(s)

extern "C" size_t tdl_exchange_pad_bytes(void) { return sizeof(ExchPad); }

// One rank's share of the fused reduction.  in_ptrs/out_ptrs: W device pointers (peer-mapped or local).
// The slice [vec_begin, vec_end) is in 16-byte vectors.  coeff: device [W].  stats_partial: device doubles
// [W+1], accumulated (caller zeroes; the exchange kernel re-zeroes it after publishing).
// store_rank: -1 = every peer, r = only out_ptrs[r].  mc_out: multicast (NVLS) address of the bucket or NULL.
// threads: 256; max_blocks <= 0: 4 x #SMs.  (A 64-thread variant meant to sit beside resident tensor-core CTAs was built
// and dropped: the block scheduler never placed it there)
extern "C" int tdl_trust_reduce_ex(void* const* in_ptrs, void* const* out_ptrs, int W, int dtype, int64_t vec_begin,
                                   int64_t vec_end, const float* coeff, double* stats_partial, int store_rank, void* mc_out,
                                   int threads, int max_blocks, void* stream) {
    TDL_REQUIRE(W >= 1 && W <= MAX_W, "world size %d unsupported (1..%d)", W, MAX_W);
    TDL_REQUIRE(store_rank < W, "store_rank %d out of range", store_rank);
    TDL_REQUIRE(threads == 256, "trust_reduce: 256 threads per block");
    PeerPtrs in{}, out{};
    for (int i = 0; i < W; ++i) { in.p[i] = in_ptrs[i]; out.p[i] = out_ptrs[i]; }
    const long long nvec = vec_end - vec_begin;
    if (nvec <= 0) return 0;
    long long blocks = (nvec + threads - 1) / threads;
    static int per_sm = -1;                      // TDL_TR_BLOCKS: resident blocks per SM for the reduction
    if (per_sm < 0) { const char* e = getenv("TDL_TR_BLOCKS"); per_sm = e ? atoi(e) : 4; if (per_sm < 1) per_sm = 1; }
    const long long cap = max_blocks > 0 ? max_blocks : (long long)tdl_num_sms() * per_sm;
    const int grid = (int)(blocks > cap ? cap : blocks);
#define LAUNCH(T, WW) trust_reduce_kernel<T, WW, 256><<<grid, 256, 0, STREAM(stream)>>>(in, out, vec_begin, vec_end, coeff, stats_partial, store_rank, mc_out)
#define BY_W(T) switch (W) { case 1: LAUNCH(T, 1); break; case 2: LAUNCH(T, 2); break; case 3: LAUNCH(T, 3); break; \
    case 4: LAUNCH(T, 4); break; case 5: LAUNCH(T, 5); break; case 6: LAUNCH(T, 6); break; case 7: LAUNCH(T, 7); break; \
    case 8: LAUNCH(T, 8); break; }
    if (dtype == TDL_BF16) { BY_W(__nv_bfloat16) }
    else if (dtype == TDL_F32) { BY_W(float) }
    else return tdl_set_error("trust_reduce: dtype %d unsupported", dtype);
#undef BY_W
#undef LAUNCH
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_trust_reduce(void* const* in_ptrs, void* const* out_ptrs, int W, int dtype, int64_t vec_begin,
                                int64_t vec_end, const float* coeff, double* stats_partial, void* stream) {
    return tdl_trust_reduce_ex(in_ptrs, out_ptrs, W, dtype, vec_begin, vec_end, coeff, stats_partial, -1, nullptr, 256, 0, stream);
}

// K4: plain (sum x scale) all-reduce of slice [vec_begin, vec_end) of a symmetric bucket: through the switch when
// `mc_ptr` (the multicast address) is given, over P2P loads/stores otherwise.  The caller brackets it with tdl_exchange
// barriers exactly like tdl_trust_reduce.
extern "C" int tdl_sum_allreduce(void* const* buf_ptrs, void* mc_ptr, int W, int dtype, int64_t vec_begin, int64_t vec_end,
                                 float scale, int max_blocks, void* stream) {
    TDL_REQUIRE(W >= 1 && W <= MAX_W, "world size %d unsupported (1..%d)", W, MAX_W);
    PeerPtrs b{};
    for (int i = 0; i < W; ++i) b.p[i] = buf_ptrs[i];
    const long long nvec = vec_end - vec_begin;
    if (nvec <= 0) return 0;
    long long blocks = (nvec + 255) / 256;
    const long long cap = max_blocks > 0 ? max_blocks : (long long)tdl_num_sms() * 4;
    const int grid = (int)(blocks > cap ? cap : blocks);
    if (dtype == TDL_BF16) {
        if (mc_ptr) sum_allreduce_kernel<__nv_bfloat16, true><<<grid, 256, 0, STREAM(stream)>>>(b, mc_ptr, W, vec_begin, vec_end, scale);
        else        sum_allreduce_kernel<__nv_bfloat16, false><<<grid, 256, 0, STREAM(stream)>>>(b, mc_ptr, W, vec_begin, vec_end, scale);
    } else if (dtype == TDL_F32) {
        if (mc_ptr) sum_allreduce_kernel<float, true><<<grid, 256, 0, STREAM(stream)>>>(b, mc_ptr, W, vec_begin, vec_end, scale);
        else        sum_allreduce_kernel<float, false><<<grid, 256, 0, STREAM(stream)>>>(b, mc_ptr, W, vec_begin, vec_end, scale);
    } else return tdl_set_error("sum_allreduce: dtype %d unsupported", dtype);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_exchange(void* const* pad_ptrs, int rank, int W, int which, const double* vals, int count,
                            unsigned int* epoch_counter, int zero_vals, void* stream) {
    TDL_REQUIRE(W >= 1 && W <= MAX_W && count <= EXCH_MAX && (which == 0 || which == 1), "bad exchange arguments");
    PeerPtrs pads{};
    for (int i = 0; i < W; ++i) pads.p[i] = pad_ptrs[i];
    static bool configured = false;
    if (!configured) {
        // These one-block kernels are issued on the side stream while persistent tensor-core CTAs hold every SM with the
        // maximum shared-memory split; asking for the same split lets them start without waiting for an SM to go idle
        // and be reconfigured.  (The configuration the 8-GPU numbers in profiles/r2 were measured with.)
        cudaFuncSetAttribute(exchange_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaFuncSetAttribute(coefficients_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        configured = true;
    }
    exchange_kernel<<<1, 32, 0, STREAM(stream)>>>(pads, rank, W, which, vals, count, epoch_counter, zero_vals);
    TDL_LAUNCH_CHECK();
    return 0;
}

// copies what the peers published in gather area `which` (count 8-byte words per rank) into out[W][count]
__global__ void exchange_read_kernel(const ExchPad* pad, int W, int which, int count, unsigned long long* __restrict__ out) {
    const int t = threadIdx.x;
    if (t < W * count) out[t] = reinterpret_cast<const unsigned long long*>(&pad->gather[which][t / count][0])[t % count];
}

// K9 verdict on the device: shard i is intact iff the checksum this rank folded while pulling it equals the one its
// owner published (both mod 2^32 per word).  flags[i] |= 1 on mismatch.
__global__ void integrity_compare_kernel(const unsigned long long* __restrict__ seen, const unsigned long long* __restrict__ pub,
                                         int W, int* __restrict__ flags) {
    const int i = threadIdx.x;
    if (i < W) {
        const bool same = (seen[2 * i] & 0xFFFFFFFFull) == (pub[2 * i] & 0xFFFFFFFFull) &&
                          (seen[2 * i + 1] & 0xFFFFFFFFull) == (pub[2 * i + 1] & 0xFFFFFFFFull);
        if (!same) flags[i] |= 1;
    }
}

extern "C" int tdl_exchange_read(const void* my_pad, int W, int which, int count, void* out_u64, void* stream) {
    TDL_REQUIRE(W >= 1 && W <= MAX_W && count >= 1 && count <= EXCH_MAX && (which == 0 || which == 1), "bad exchange_read arguments");
    exchange_read_kernel<<<1, 128, 0, STREAM(stream)>>>(reinterpret_cast<const ExchPad*>(my_pad), W, which, count,
                                                        reinterpret_cast<unsigned long long*>(out_u64));
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_integrity_compare(const void* seen_u64, const void* published_u64, int W, int* flags, void* stream) {
    TDL_REQUIRE(W >= 1 && W <= MAX_W, "world size %d unsupported", W);
    integrity_compare_kernel<<<1, 32, 0, STREAM(stream)>>>(reinterpret_cast<const unsigned long long*>(seen_u64),
                                                           reinterpret_cast<const unsigned long long*>(published_u64), W, flags);
    TDL_LAUNCH_CHECK();
    return 0;
}

static TrustParams make_tp(const float* f) {
    TrustParams tp;
    tp.alpha = f[0]; tp.threshold = f[1]; tp.compromised_below = f[2]; tp.recovering_above = f[3]; tp.retrust_above = f[4];
    tp.penalty = f[5]; tp.clip_ratio = f[6]; tp.norm_ratio_max = f[7]; tp.cosine_min = f[8]; tp.w_dev = f[9]; tp.w_gc = f[10];
    tp.base_score = f[11];
    return tp;
}

// params12 (host floats): alpha, threshold, compromised_below, recovering_above, retrust_above, penalty,
//                         clip_ratio, norm_ratio_max, cosine_min, w_dev, w_gc, base_score
extern "C" int tdl_trust_coefficients_acc(const void* my_pad, int W, const float* trust, const int* status,
                                          const float* params12, float* coeff, float* norms_out, double* acc_sq,
                                          int* acc_bad, void* stream) {
    const ExchPad* pad = reinterpret_cast<const ExchPad*>(my_pad);
    coefficients_kernel<<<1, 32, 0, STREAM(stream)>>>(&pad->gather[0][0][0], W, trust, status, make_tp(params12), coeff, norms_out,
                                                      acc_sq, acc_bad);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_trust_coefficients(const void* my_pad, int W, const float* trust, const int* status,
                                      const float* params12, float* coeff, float* norms_out, void* stream) {
    return tdl_trust_coefficients_acc(my_pad, W, trust, status, params12, coeff, norms_out, nullptr, nullptr, stream);
}

extern "C" int tdl_trust_finalize_acc(const void* my_pad, int W, const float* norms, const float* out_dev,
                                      const float* params12, float* stats, float* trust, int* status, int* flags_out,
                                      int update_trust, double* acc_sq, int* acc_bad, void* stream) {
    const ExchPad* pad = reinterpret_cast<const ExchPad*>(my_pad);
    finalize_trust_kernel<<<1, 32, 0, STREAM(stream)>>>(&pad->gather[1][0][0], W, norms, out_dev, make_tp(params12), stats, trust,
                                                        status, flags_out, update_trust, acc_sq, acc_bad);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_trust_finalize(const void* my_pad, int W, const float* norms, const float* out_dev,
                                  const float* params12, float* stats, float* trust, int* status, int* flags_out,
                                  int update_trust, void* stream) {
    return tdl_trust_finalize_acc(my_pad, W, norms, out_dev, params12, stats, trust, status, flags_out, update_trust, nullptr,
                                  nullptr, stream);
}

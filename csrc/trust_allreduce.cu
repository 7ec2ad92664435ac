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

// stats_partial: [W] dot(g_i, med) then [1] |med|^2   (doubles, accumulated with atomics)
template <typename T, int W>
__global__ void __launch_bounds__(256) trust_reduce_kernel(PeerPtrs in, PeerPtrs out, long long vec_begin, long long vec_end,
                                                           const float* __restrict__ coeff, double* __restrict__ stats_partial) {
    constexpr int VN = Vec<T>::N;
    __shared__ double sh[8][W + 1];
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
        for (int i = 0; i < W; ++i) unpack_vec<T>(raw[i], g[i]);
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
#pragma unroll
        for (int i = 0; i < W; ++i) st_peer_v4(reinterpret_cast<uint4*>(out.p[i]) + v, packed);
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
        for (int w2 = 0; w2 < 8; ++w2) t += sh[w2][threadIdx.x];
        atomicAdd(&stats_partial[threadIdx.x], t);
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
__global__ void coefficients_kernel(const double* __restrict__ gathered /*[W][EXCH_MAX], col 0 = sumsq*/, int W,
                                    const float* __restrict__ trust, const int* __restrict__ status,
                                    TrustParams tp, float* __restrict__ coeff, float* __restrict__ norms_out) {
    if (threadIdx.x != 0) return;
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
    if (k == 0) { for (int i = 0; i < W; ++i) coeff[i] = 0.f; return; }
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
    }
}

// gathered partials [W ranks][W dots + med2] -> stats[W][4] = (norm, dot, |med|^2, cos); then the trust EMA + FSM.
// flags_out[i]: bit0 = failed verification this step (norm ratio / cosine / non-finite).
__global__ void finalize_trust_kernel(const double* __restrict__ gathered, int W, const float* __restrict__ norms,
                                      const float* __restrict__ out_dev, TrustParams tp, float* __restrict__ stats,
                                      float* __restrict__ trust, int* __restrict__ status, int* __restrict__ flags_out,
                                      int update_trust) {
    if (threadIdx.x != 0) return;
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
        bool ok = isfinite(n) && isfinite(cosv);
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
extern "C" int tdl_trust_reduce(void* const* in_ptrs, void* const* out_ptrs, int W, int dtype, int64_t vec_begin,
                                int64_t vec_end, const float* coeff, double* stats_partial, void* stream) {
    TDL_REQUIRE(W >= 1 && W <= MAX_W, "world size %d unsupported (1..%d)", W, MAX_W);
    PeerPtrs in{}, out{};
    for (int i = 0; i < W; ++i) { in.p[i] = in_ptrs[i]; out.p[i] = out_ptrs[i]; }
    const long long nvec = vec_end - vec_begin;
    if (nvec <= 0) return 0;
    long long blocks = (nvec + 255) / 256;
    static int per_sm = -1;                      // TDL_TR_BLOCKS: resident blocks per SM for the reduction
    if (per_sm < 0) { const char* e = getenv("TDL_TR_BLOCKS"); per_sm = e ? atoi(e) : 4; if (per_sm < 1) per_sm = 1; }
    const long long cap = (long long)tdl_num_sms() * per_sm;
    const int grid = (int)(blocks > cap ? cap : blocks);
#define LAUNCH(T, WW) trust_reduce_kernel<T, WW><<<grid, 256, 0, STREAM(stream)>>>(in, out, vec_begin, vec_end, coeff, stats_partial)
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

extern "C" int tdl_exchange(void* const* pad_ptrs, int rank, int W, int which, const double* vals, int count,
                            unsigned int* epoch_counter, int zero_vals, void* stream) {
    TDL_REQUIRE(W >= 1 && W <= MAX_W && count <= EXCH_MAX && (which == 0 || which == 1), "bad exchange arguments");
    PeerPtrs pads{};
    for (int i = 0; i < W; ++i) pads.p[i] = pad_ptrs[i];
    exchange_kernel<<<1, 32, 0, STREAM(stream)>>>(pads, rank, W, which, vals, count, epoch_counter, zero_vals);
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
extern "C" int tdl_trust_coefficients(const void* my_pad, int W, const float* trust, const int* status,
                                      const float* params12, float* coeff, float* norms_out, void* stream) {
    const ExchPad* pad = reinterpret_cast<const ExchPad*>(my_pad);
    coefficients_kernel<<<1, 32, 0, STREAM(stream)>>>(&pad->gather[0][0][0], W, trust, status, make_tp(params12), coeff, norms_out);
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_trust_finalize(const void* my_pad, int W, const float* norms, const float* out_dev,
                                  const float* params12, float* stats, float* trust, int* status, int* flags_out,
                                  int update_trust, void* stream) {
    const ExchPad* pad = reinterpret_cast<const ExchPad*>(my_pad);
    finalize_trust_kernel<<<1, 32, 0, STREAM(stream)>>>(&pad->gather[1][0][0], W, norms, out_dev, make_tp(params12), stats, trust,
                                                        status, flags_out, update_trust);
    TDL_LAUNCH_CHECK();
    return 0;
}

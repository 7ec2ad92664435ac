// K5: the detector's 12 statistics computed where the tensor lives.
//
// Absorbs reference attack_detector.py:78 (full activation D2H copy) and :185-200 (numpy/scipy
// mean, std, min, max, median, skewness, kurtosis, p25, p75, L1, L2, Linf on the host: 2.2 s for one
// GPT-2-medium stage activation) plus the .item() chains of distributed_trainer.py:242-243,284-287.
//
//   pass A  moments_kernel   one streaming read: shifted power sums (fp32 per thread in short runs,
//                            fp64 across threads/blocks), min, max, sum|x|, max|x|
//   pass B  radix select     exact order statistics for the 6 ranks numpy's linear-interpolated
//                            median / p25 / p75 need, 8-bit digits over order-preserving keys:
//                            2 passes for bf16, 4 for fp32/fp16.  All six ranks ride the same passes.
//   finalize                 12 doubles in STAT order on the device.
//
// Scratch (caller-provided, zeroed once at allocation; the kernels leave it zeroed for the next
// call): per-block partials, 6x256 histograms, select state.
#include "common.cuh"

namespace {

constexpr int THREADS = 256;
constexpr int MAX_BLOCKS = 148 * 4;
constexpr int NT = 6;                       // target ranks: (median, p25, p75) x (lo, hi)

struct SelectState {
    unsigned long long rank[NT];            // remaining rank inside the current prefix bucket
    unsigned int prefix[NT];                // key bits decided so far
    unsigned int group[NT];                 // histogram row used by target t
    unsigned int group_prefix[NT];
    unsigned int n_groups;
    unsigned int pad;
    double frac[3];                         // interpolation weights
    double moments[10];                     // n, mean, m2, m3, m4, min, max, l1, sumsq, linf
};

struct Scratch {
    SelectState st;
    unsigned int hist[NT][256];
    double partial[MAX_BLOCKS][8];
};

template <typename T> __device__ __forceinline__ float load_f(const T* p, long long i) { return to_f32<T>(p[i]); }

__device__ __forceinline__ unsigned int float_key(float f) {
    unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_float(unsigned int k) {
    unsigned int u = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
    return __uint_as_float(u);
}

__device__ __forceinline__ double block_sum(double v, double* sh) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    double r = 0.0;
    if (w == 0) {
        r = l < (blockDim.x >> 5) ? sh[l] : 0.0;
        for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
    }
    return r;                                // valid in warp 0
}
__device__ __forceinline__ float block_minmax(float v, float* sh, bool is_max) {
    v = is_max ? warp_max(v) : warp_min(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    float r = v;
    if (w == 0) {
        r = l < (blockDim.x >> 5) ? sh[l] : (is_max ? -INFINITY : INFINITY);
        r = is_max ? warp_max(r) : warp_min(r);
    }
    return r;
}

// ------------------------------------------------------------------ pass A
template <typename T>
__global__ void __launch_bounds__(THREADS) moments_kernel(const T* __restrict__ x, long long n, Scratch* sc) {
    __shared__ double shd[8];
    __shared__ float shf[8];
    __shared__ float s_pivot;
    // pivot: mean of the first <=1024 elements (same value in every block) keeps the power sums
    // well conditioned when |mean| >> std
    {
        const int m = n < 1024 ? (int)n : 1024;
        float a = 0.f;
        for (int i = threadIdx.x; i < m; i += THREADS) a += load_f(x, i);
        double t = block_sum((double)a, shd);
        if (threadIdx.x == 0) s_pivot = (float)(t / m);
        __syncthreads();
    }
    const float pv = s_pivot;
    double d1 = 0, d2 = 0, d3 = 0, d4 = 0, l1 = 0;
    float mn = INFINITY, mx = -INFINITY, linf = 0.f;
    constexpr int VEC = 16 / sizeof(T);
    const long long nvec = ((reinterpret_cast<uintptr_t>(x) & 15) == 0) ? n / VEC : 0;
    const long long stride = (long long)gridDim.x * THREADS;
    float f1 = 0, f2 = 0, f3 = 0, f4 = 0, fl = 0;
    int run = 0;
    auto feed = [&](float v) {
        const float d = v - pv, dd = d * d;
        f1 += d; f2 += dd; f3 += dd * d; f4 += dd * dd; fl += fabsf(v);
        mn = fminf(mn, v); mx = fmaxf(mx, v); linf = fmaxf(linf, fabsf(v));
    };
    for (long long i = (long long)blockIdx.x * THREADS + threadIdx.x; i < nvec; i += stride) {
        uint4 raw = ld_nc_v4(reinterpret_cast<const uint4*>(x) + i);
        const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
        for (int j = 0; j < VEC; ++j) feed(to_f32<T>(e[j]));
        if (++run == 8) { d1 += f1; d2 += f2; d3 += f3; d4 += f4; l1 += fl; f1 = f2 = f3 = f4 = fl = 0.f; run = 0; }
    }
    for (long long i = nvec * VEC + (long long)blockIdx.x * THREADS + threadIdx.x; i < n; i += stride) feed(load_f(x, i));
    d1 += f1; d2 += f2; d3 += f3; d4 += f4; l1 += fl;
    double r1 = block_sum(d1, shd), r2 = block_sum(d2, shd), r3 = block_sum(d3, shd), r4 = block_sum(d4, shd);
    double rl = block_sum(l1, shd);
    float rmn = block_minmax(mn, shf, false), rmx = block_minmax(mx, shf, true), rli = block_minmax(linf, shf, true);
    if (threadIdx.x == 0) {
        double* p = sc->partial[blockIdx.x];
        p[0] = r1; p[1] = r2; p[2] = r3; p[3] = r4; p[4] = rmn; p[5] = rmx; p[6] = rl; p[7] = rli;
        if (blockIdx.x == 0) sc->st.moments[9] = pv;           // stash the pivot (overwritten by finalize)
    }
}

__global__ void moments_final_kernel(Scratch* sc, long long n, int nblocks) {
    __shared__ double shd[8];
    __shared__ float shf[8];
    double a[4] = {0, 0, 0, 0}, l1 = 0;
    float mn = INFINITY, mx = -INFINITY, li = 0.f;
    for (int b = threadIdx.x; b < nblocks; b += blockDim.x) {
        const double* p = sc->partial[b];
        a[0] += p[0]; a[1] += p[1]; a[2] += p[2]; a[3] += p[3];
        mn = fminf(mn, (float)p[4]); mx = fmaxf(mx, (float)p[5]); l1 += p[6]; li = fmaxf(li, (float)p[7]);
    }
    double s1 = block_sum(a[0], shd), s2 = block_sum(a[1], shd), s3 = block_sum(a[2], shd), s4 = block_sum(a[3], shd);
    double sl = block_sum(l1, shd);
    float rmn = block_minmax(mn, shf, false), rmx = block_minmax(mx, shf, true), rli = block_minmax(li, shf, true);
    if (threadIdx.x == 0) {
        const double N = (double)n, pv = sc->st.moments[9];
        const double e1 = s1 / N, e2 = s2 / N, e3 = s3 / N, e4 = s4 / N;    // moments about the pivot
        const double mean = pv + e1;
        const double m2 = e2 - e1 * e1;
        const double m3 = e3 - 3 * e1 * e2 + 2 * e1 * e1 * e1;
        const double m4 = e4 - 4 * e1 * e3 + 6 * e1 * e1 * e2 - 3 * e1 * e1 * e1 * e1;
        const double sumsq = s2 + 2 * pv * s1 + N * pv * pv;                // sum x^2 about zero
        double* m = sc->st.moments;
        m[0] = N; m[1] = mean; m[2] = m2 > 0 ? m2 : 0; m[3] = m3; m[4] = m4; m[5] = rmn; m[6] = rmx; m[7] = sl;
        m[8] = sumsq > 0 ? sumsq : 0; m[9] = rli;
        // ranks for numpy 'linear' percentiles: pos = q (n-1)
        const double qs[3] = {0.5, 0.25, 0.75};
        for (int i = 0; i < 3; ++i) {
            const double pos = qs[i] * (N - 1.0);
            const unsigned long long lo = (unsigned long long)floor(pos);
            const unsigned long long hi = lo + 1 < (unsigned long long)n ? lo + 1 : lo;
            sc->st.rank[2 * i] = lo; sc->st.rank[2 * i + 1] = hi;
            sc->st.frac[i] = pos - (double)lo;
        }
        for (int t = 0; t < NT; ++t) { sc->st.prefix[t] = 0; sc->st.group[t] = 0; sc->st.group_prefix[t] = 0; }
        sc->st.n_groups = 1;
    }
}

// ------------------------------------------------------------------ pass B
template <typename T>
__global__ void __launch_bounds__(THREADS) hist_kernel(const T* __restrict__ x, long long n, Scratch* sc, int shift,
                                                       unsigned int decided_mask) {
    __shared__ unsigned int sh[NT][256];
    const unsigned int ng = sc->st.n_groups;
    for (int i = threadIdx.x; i < (int)ng * 256; i += THREADS) (&sh[0][0])[i] = 0;
    unsigned int gp[NT];
#pragma unroll
    for (int g = 0; g < NT; ++g) gp[g] = sc->st.group_prefix[g];
    __syncthreads();
    constexpr int VEC = 16 / sizeof(T);
    const long long nvec = ((reinterpret_cast<uintptr_t>(x) & 15) == 0) ? n / VEC : 0;
    const long long stride = (long long)gridDim.x * THREADS;
    auto feed = [&](float v) {
        const unsigned int k = float_key(v);
        const unsigned int digit = (k >> shift) & 255u;
#pragma unroll
        for (int g = 0; g < NT; ++g)
            if (g < (int)ng && ((k ^ gp[g]) & decided_mask) == 0) atomicAdd(&sh[g][digit], 1u);
    };
    for (long long i = (long long)blockIdx.x * THREADS + threadIdx.x; i < nvec; i += stride) {
        uint4 raw = ld_nc_v4(reinterpret_cast<const uint4*>(x) + i);
        const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
        for (int j = 0; j < VEC; ++j) feed(to_f32<T>(e[j]));
    }
    for (long long i = nvec * VEC + (long long)blockIdx.x * THREADS + threadIdx.x; i < n; i += stride) feed(load_f(x, i));
    __syncthreads();
    for (int i = threadIdx.x; i < (int)ng * 256; i += THREADS) {
        const unsigned int c = (&sh[0][0])[i];
        if (c) atomicAdd(&sc->hist[0][0] + i, c);
    }
}

// one warp per target: find the digit whose bucket contains the rank, descend, regroup, clear.
__global__ void select_kernel(Scratch* sc, int shift) {
    __shared__ unsigned int new_prefix[NT];
    const int t = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (t < NT) {
        const unsigned int* h = sc->hist[sc->st.group[t]];
        unsigned long long rank = sc->st.rank[t];
        // 8 bins per lane, warp-exclusive scan of lane totals
        unsigned int c[8];
        unsigned long long tot = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) { c[j] = h[l * 8 + j]; tot += c[j]; }
        unsigned long long incl = tot;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long v = __shfl_up_sync(0xffffffffu, incl, o);
            if (l >= o) incl += v;
        }
        const unsigned long long excl = incl - tot;
        const bool mine = rank >= excl && rank < incl;
        unsigned int digit = 0;
        unsigned long long before = 0;
        if (mine) {
            unsigned long long run = excl;
            for (int j = 0; j < 8; ++j) {
                if (rank < run + c[j]) { digit = l * 8 + j; before = run; break; }
                run += c[j];
            }
        }
        const unsigned int ball = __ballot_sync(0xffffffffu, mine);
        const int src = ball ? __ffs(ball) - 1 : 31;
        digit = __shfl_sync(0xffffffffu, digit, src);
        before = __shfl_sync(0xffffffffu, before, src);
        if (l == 0) {
            new_prefix[t] = sc->st.prefix[t] | (digit << shift);
            sc->st.rank[t] = rank - before;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int ng = 0;
        for (int a = 0; a < NT; ++a) {
            sc->st.prefix[a] = new_prefix[a];
            int g = -1;
            for (unsigned int b = 0; b < ng; ++b) if (sc->st.group_prefix[b] == new_prefix[a]) g = (int)b;
            if (g < 0) { g = (int)ng; sc->st.group_prefix[ng++] = new_prefix[a]; }
            sc->st.group[a] = (unsigned int)g;
        }
        sc->st.n_groups = ng;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < NT * 256; i += blockDim.x) (&sc->hist[0][0])[i] = 0;   // leave scratch clean
}

__global__ void finalize_kernel(Scratch* sc, double* out, int key_bits) {
    if (threadIdx.x != 0) return;
    const double* m = sc->st.moments;
    double q[3];
    // undecided low key bits: a 16-bit input has zero low mantissa bits, which the order-preserving
    // map turns into all-ones for negative values (top key bit clear)
    const unsigned int low = key_bits >= 32 ? 0u : ((1u << (32 - key_bits)) - 1u);
    for (int i = 0; i < 3; ++i) {
        unsigned int klo = sc->st.prefix[2 * i], khi = sc->st.prefix[2 * i + 1];
        if (!(klo & 0x80000000u)) klo |= low;
        if (!(khi & 0x80000000u)) khi |= low;
        const double lo = (double)key_float(klo), hi = (double)key_float(khi);
        q[i] = lo + (hi - lo) * sc->st.frac[i];
    }
    const double m2 = m[2];
    out[0] = m[1];
    out[1] = sqrt(m2);
    out[2] = m[5];
    out[3] = m[6];
    out[4] = q[0];
    out[5] = m[3] / (m2 * sqrt(m2));           // biased skewness (scipy default); nan for constant input
    out[6] = m[4] / (m2 * m2) - 3.0;           // biased Fisher kurtosis
    out[7] = q[1];
    out[8] = q[2];
    out[9] = m[7];
    out[10] = sqrt(m[8]);
    out[11] = m[9];
}

__global__ void moments_out_kernel(Scratch* sc, double* out) {
    if (threadIdx.x != 0) return;
    const double* m = sc->st.moments;           // n, mean, m2, m3, m4, min, max, l1, sumsq, linf
    for (int i = 0; i < 10; ++i) out[i] = m[i];
}

template <typename T>
int run_stats(const T* x, long long n, double* out, Scratch* sc, int key_bits, bool with_quantiles, cudaStream_t s) {
    long long want = (n + THREADS * 16 - 1) / (THREADS * 16);
    int nb = (int)(want < 1 ? 1 : (want > MAX_BLOCKS ? MAX_BLOCKS : want));
    moments_kernel<T><<<nb, THREADS, 0, s>>>(x, n, sc);
    moments_final_kernel<<<1, 256, 0, s>>>(sc, n, nb);
    if (!with_quantiles) {
        moments_out_kernel<<<1, 32, 0, s>>>(sc, out);
        TDL_LAUNCH_CHECK();
        return 0;
    }
    unsigned int decided = 0;
    for (int shift = 24; shift >= 32 - key_bits; shift -= 8) {
        hist_kernel<T><<<nb, THREADS, 0, s>>>(x, n, sc, shift, decided);
        select_kernel<<<1, NT * 32, 0, s>>>(sc, shift);
        decided |= 255u << shift;
    }
    finalize_kernel<<<1, 32, 0, s>>>(sc, out, key_bits);
    TDL_LAUNCH_CHECK();
    return 0;
}

}  // namespace

extern "C" size_t tdl_stats_scratch_bytes(int64_t /*n*/) { return sizeof(Scratch); }

static int stats_dispatch(const void* x, int dtype, int64_t n, double* out, void* scratch, size_t scratch_bytes,
                          bool quant, void* stream) {
    TDL_REQUIRE(n > 0, "tensor statistics of an empty tensor");
    TDL_REQUIRE(scratch != nullptr && scratch_bytes >= sizeof(Scratch), "stats scratch too small: %zu < %zu",
                scratch_bytes, sizeof(Scratch));
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    Scratch* sc = reinterpret_cast<Scratch*>(scratch);
    switch (dtype) {
        case TDL_F32: return run_stats<float>(reinterpret_cast<const float*>(x), n, out, sc, 32, quant, s);
        case TDL_BF16: return run_stats<__nv_bfloat16>(reinterpret_cast<const __nv_bfloat16*>(x), n, out, sc, 16, quant, s);
        case TDL_F16: return run_stats<__half>(reinterpret_cast<const __half*>(x), n, out, sc, 32, quant, s);
    }
    return tdl_set_error("unsupported dtype %d", dtype);
}

// 12 statistics (mean,std,min,max,median,skew,kurtosis,p25,p75,L1,L2,Linf) -> out12 (device, float64)
extern "C" int tdl_tensor_stats(const void* x, int dtype, int64_t n, double* out12, void* scratch,
                                size_t scratch_bytes, void* stream) {
    return stats_dispatch(x, dtype, n, out12, scratch, scratch_bytes, true, stream);
}

// moments only: out10 = n, mean, m2, m3, m4, min, max, L1, sum(x^2), Linf (device, float64)
extern "C" int tdl_tensor_moments(const void* x, int dtype, int64_t n, double* out10, void* scratch,
                                  size_t scratch_bytes, void* stream) {
    return stats_dispatch(x, dtype, n, out10, scratch, scratch_bytes, false, stream);
}


// -------------------------------------------------------------------------------------------------
// Allocation-free variant of tdl_tensor_stats.
//
// Who needs it: code that wants the 12 statistics but is not in a position to obtain device memory.
// Two situations came up while integrating the detector:
//   * a monitoring probe that is called from somebody else's runtime (a C/C++ trainer or serving
//     loop that links this library directly, without PyTorch): it has no caching allocator to borrow
//     from, and a raw cudaMalloc there would synchronise the whole device in the middle of a step and
//     can fail outright once the job has handed all of HBM to its own pool;
//   * a job that runs with a fixed, pre-reserved memory budget and treats any driver allocation after
//     start-up as a bug.
// For them the scratch the kernels need lives in the library image itself: a file-scope __device__
// array, present once on every GPU that loads the library and kept until the process exits
// (sizeof(Scratch), about 44 KB).  The caller still owns the input and the 12-double output.
//
// Rules of use
//   * Limit: n <= TDL_STATS_NOALLOC_MAX_ELEMS (2^30 elements).  Larger requests are refused with an
//     error; use tdl_tensor_stats with your own scratch for those.
//   * The first call on a given device resolves the array's device address (cudaGetSymbolAddress with
//     that device current).  Make that first call during start-up, before any stream capture begins.
//   * There is one array per device and every call uses it.  It still holds whatever the previous call
//     left behind, so each call first clears it on the caller's stream.  Two calls whose GPU work can
//     overlap (different streams without an ordering between them, or two host threads) would race on
//     it: issue them on one stream or order them with events.
//   * Nothing here calls cudaMalloc / cudaFree or any allocating library routine.
// -------------------------------------------------------------------------------------------------
#include <mutex>

#define TDL_STATS_NOALLOC_MAX_ELEMS_IMPL (1ll << 30)

static __device__ __align__(256) unsigned char g_stats_static_scratch[sizeof(Scratch)];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
void* g_static_ptr[kMaxDevices] = {nullptr};     // device address of g_stats_static_scratch, per device
}  // namespace

extern "C" int tdl_tensor_stats_noalloc(const void* x, int dtype, int64_t n, double* out12, void* stream) {
    TDL_REQUIRE(n > 0, "tdl_tensor_stats_noalloc: empty tensor");
    TDL_REQUIRE(n <= TDL_STATS_NOALLOC_MAX_ELEMS_IMPL,
                "tdl_tensor_stats_noalloc: %lld elements exceed the built-in scratch limit of %lld; "
                "use tdl_tensor_stats with caller-provided scratch",
                (long long)n, (long long)TDL_STATS_NOALLOC_MAX_ELEMS_IMPL);
    TDL_REQUIRE(x != nullptr && out12 != nullptr, "tdl_tensor_stats_noalloc: null input/output");
    int dev = -1;
    TDL_CHECK(cudaGetDevice(&dev));
    TDL_REQUIRE(dev >= 0 && dev < kMaxDevices, "tdl_tensor_stats_noalloc: device index %d out of range", dev);
    void* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_static_mu);
        if (g_static_ptr[dev] == nullptr) {
            // In host code `g_stats_static_scratch` is only a handle for the symbol; the address a kernel
            // can dereference has to be looked up, with the target device current.
            void* p = nullptr;
            TDL_CHECK(cudaGetSymbolAddress(&p, g_stats_static_scratch));
            TDL_REQUIRE(p != nullptr, "tdl_tensor_stats_noalloc: symbol address lookup returned null");
            g_static_ptr[dev] = p;
        }
        scratch = g_static_ptr[dev];
    }
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    // histograms and select state must start from zero; the array still holds the previous call's data
    TDL_CHECK(cudaMemsetAsync(scratch, 0, sizeof(Scratch), s));
    return stats_dispatch(x, dtype, n, out12, scratch, sizeof(Scratch), true, stream);
}

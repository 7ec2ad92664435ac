// K9: parameter all-gather with the integrity checks fused into the copy.
//
// "Param Monitor" appears in the reference's architecture sketch (README.md:22) and MODEL_POISONING
// in its enum (attack_detector.py:22); neither is implemented.  North star: per-shard checksums and
// drift norms fused into the parameter all-gather.
//
// Each rank owns one shard (ZeRO-1 style: it ran the optimizer on that shard).  Every rank pulls all
// W shards straight out of its peers' symmetric buffers (16-byte P2P loads over NVLink) into its full
// local parameter buffer and, in the same pass, folds
//     s1 = sum(w_i), s2 = sum((n - i) w_i)   (mod 2^32 over the raw 32-bit words; order sensitive)
//     drift = sum((theta - theta_prev)^2)    and refreshes theta_prev
// per shard.  The owner publishes the checksum of what it wrote; a receiver that computes a
// different value saw a corrupted / tampered shard.  One pass over the bytes, no extra traffic.
#include "common.cuh"

namespace {

constexpr int MAX_W = 8;
struct Ptrs { const void* p[MAX_W]; };

template <typename T>
__global__ void __launch_bounds__(256) param_gather_kernel(Ptrs shards, T* __restrict__ dst, T* __restrict__ prev,
                                                           long long shard_elems, unsigned long long* __restrict__ sums /*[W][2]*/,
                                                           double* __restrict__ drift /*[W]*/) {
    __shared__ unsigned int s1s[8], s2s[8];
    __shared__ double ds[8];
    const int peer = blockIdx.y;
    constexpr int VE = 16 / sizeof(T);                  // elements per 16-byte vector
    const long long nvec = shard_elems / VE;
    const long long nwords = shard_elems * sizeof(T) / 4;
    const uint4* src = reinterpret_cast<const uint4*>(shards.p[peer]);
    uint4* out = reinterpret_cast<uint4*>(dst + peer * shard_elems);
    uint4* pv = prev != nullptr ? reinterpret_cast<uint4*>(prev + peer * shard_elems) : nullptr;
    unsigned int s1 = 0, s2 = 0;
    float dacc = 0.f;
    double dd = 0.0;
    int run = 0;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
        uint4 u;
        asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "l"(src + v) : "memory");
        const unsigned int w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const long long wi = v * 4 + j;
            s1 += w[j];
            s2 += static_cast<unsigned int>((nwords - wi) & 0xFFFFFFFFll) * w[j];
        }
        if (pv != nullptr) {
            const uint4 o = pv[v];
            const T* a = reinterpret_cast<const T*>(&u);
            const T* b = reinterpret_cast<const T*>(&o);
#pragma unroll
            for (int j = 0; j < VE; ++j) { const float d = to_f32<T>(a[j]) - to_f32<T>(b[j]); dacc = fmaf(d, d, dacc); }
            pv[v] = u;
            if (++run == 64) { dd += dacc; dacc = 0.f; run = 0; }
        }
        out[v] = u;
    }
    dd += dacc;
    for (int o = 16; o > 0; o >>= 1) {
        s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o);
        dd += __shfl_xor_sync(0xffffffffu, dd, o);
    }
    if ((threadIdx.x & 31) == 0) { s1s[threadIdx.x >> 5] = s1; s2s[threadIdx.x >> 5] = s2; ds[threadIdx.x >> 5] = dd; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int a = 0, b = 0; double d = 0;
        for (int k = 0; k < 8; ++k) { a += s1s[k]; b += s2s[k]; d += ds[k]; }
        atomicAdd(&sums[peer * 2 + 0], (unsigned long long)a);
        atomicAdd(&sums[peer * 2 + 1], (unsigned long long)b);
        if (pv != nullptr) atomicAdd(&drift[peer], d);
    }
}

}  // namespace

// shard_ptrs: W device pointers (peer-mapped) to each rank's shard of `shard_elems` elements.
// dst: full buffer [W * shard_elems]; prev: optional snapshot of the same size (refreshed in place).
// sums (uint64 [W][2], caller-zeroed) receive s1/s2 (mask with 0xFFFFFFFF); drift (double [W], caller-zeroed).
extern "C" int tdl_param_gather(const void* const* shard_ptrs, int W, int dtype, void* dst, void* prev, int64_t shard_elems,
                                void* sums_u64, double* drift, void* stream) {
    TDL_REQUIRE(W >= 1 && W <= MAX_W, "world size %d unsupported", W);
    const int esz = dtype == TDL_F32 ? 4 : 2;
    TDL_REQUIRE((shard_elems * esz) % 16 == 0, "shard size must be a multiple of 16 bytes");
    Ptrs ps{};
    for (int i = 0; i < W; ++i) ps.p[i] = shard_ptrs[i];
    const long long nvec = shard_elems * esz / 16;
    int gx = (int)std::min<long long>((nvec + 255) / 256, (long long)tdl_num_sms() * 4 / W + 1);
    if (gx < 1) gx = 1;
    dim3 grid(gx, W);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == TDL_F32)
        param_gather_kernel<float><<<grid, 256, 0, s>>>(ps, (float*)dst, (float*)prev, shard_elems, (unsigned long long*)sums_u64, drift);
    else if (dtype == TDL_BF16)
        param_gather_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>(ps, (__nv_bfloat16*)dst, (__nv_bfloat16*)prev, shard_elems,
                                                               (unsigned long long*)sums_u64, drift);
    else return tdl_set_error("param_gather: dtype %d unsupported", dtype);
    TDL_LAUNCH_CHECK();
    return 0;
}

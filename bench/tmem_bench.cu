// Microbenchmark: tcgen05.ld throughput (TMEM -> registers) and MUFU.EX2 throughput per SM on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I csrc bench/tmem_bench.cu -o bench/bin/tmem_bench
#include "common.cuh"
#include <cstdio>

int tdl_set_error(const char*, ...) { return 1; }

template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1) tmem_ld_kernel(int iters, long long* cycles, unsigned* sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t base = slot + (static_cast<uint32_t>((warp & 3) * 32) << 16) + (warp >> 2) * 128;
    unsigned acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        uint32_t r[32];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            tmem_ld_32x32(base + c * 32, r);
            tmem_ld_wait();
            acc += r[0] ^ r[31];
        }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    __syncthreads();
    if (warp == 0) tmem_dealloc(slot, 512);
}

template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1) ex2_kernel(int iters, long long* cycles, float* sink) {
    float x[8];
    for (int j = 0; j < 8; ++j) x[j] = -0.001f * (threadIdx.x + j);
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[j]));
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
    float s = 0; for (int j = 0; j < 8; ++j) s += x[j];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
    long long* cyc; unsigned* sink; float* fs;
    cudaMalloc(&cyc, 148 * 8); cudaMalloc(&sink, 148 * 1024 * 4); cudaMalloc(&fs, 148 * 1024 * 4);
    long long h[148];
    const int iters = 2000;
#define RUN_LD(W) { tmem_ld_kernel<W><<<148, W * 32>>>(iters, cyc, sink); cudaError_t e = cudaDeviceSynchronize(); \
    cudaMemcpy(h, cyc, 148 * 8, cudaMemcpyDeviceToHost); double bytes = (double)iters * 4 * W * 32 * 32 * 4; \
    printf("tcgen05.ld 32x32b.x32  %2d warps/SM: %8lld cycles  -> %.1f B/clk/SM  (%s)\n", W, h[0], bytes / h[0], cudaGetErrorString(e)); }
    RUN_LD(1) RUN_LD(4) RUN_LD(8)
#define RUN_EX(W) { ex2_kernel<W><<<148, W * 32>>>(iters, cyc, fs); cudaError_t e = cudaDeviceSynchronize(); \
    cudaMemcpy(h, cyc, 148 * 8, cudaMemcpyDeviceToHost); double ops = (double)iters * 8 * W * 32; \
    printf("ex2.approx             %2d warps/SM: %8lld cycles  -> %.1f ops/clk/SM  (%s)\n", W, h[0], ops / h[0], cudaGetErrorString(e)); }
    RUN_EX(4) RUN_EX(8) RUN_EX(16)
    return 0;
}

// Microbenchmark: issue-to-retire rate of tcgen05.mma (cta_group::1, bf16, M = 128) per instruction shape, as the
// attention kernels use it: operands from shared memory (SS) or A from tensor memory (TS), N = 128 or 64, one
// accumulator (dependent chain) or two alternating ones.  Operand contents are irrelevant (uninitialised smem).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I csrc bench/mma_rate_bench.cu -o bench/bin/mma_rate_bench
#include "common.cuh"
#include <cstdio>

int tdl_set_error(const char*, ...) { return 1; }

// mode: 0 = SS N128 one D, 1 = SS N128 two D, 2 = TS N64 one D, 3 = TS N64 two D, 4 = SS N64 (MN-major A,B) two D,
//       5 = SS N128 one D while 8 other warps hammer tcgen05.ld on the same TMEM, 6 = TS N64 two D under the same load
template <int MODE>
__global__ void __launch_bounds__(320, 1) mma_rate_kernel(int n_mma, long long* cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t slot;
    __shared__ uint64_t bar;
    __shared__ volatile int stop;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); stop = 0; }
    if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
    tc_fence_before(); __syncthreads(); tc_fence_after();
    const uint32_t tmem = slot;
    constexpr bool TS = (MODE == 2 || MODE == 3 || MODE == 6);
    constexpr bool TWO = (MODE == 1 || MODE == 3 || MODE == 4 || MODE == 6);
    constexpr int N = (MODE == 0 || MODE == 1 || MODE == 5) ? 128 : 64;
    constexpr bool LOAD = (MODE == 5 || MODE == 6);
    const uint32_t idesc = MODE == 4 ? umma_idesc_bf16(128, 64, 1, 1) : (TS ? umma_idesc_bf16(128, 64, 0, 1) : umma_idesc_bf16(128, N, 0, 0));
    if (warp == 9) {
        long long t0 = 0, t1 = 0;
        if (elect_one()) {
            const uint32_t a_lo = umma_desc_lo(smem_u32(smem), MODE == 4 ? 16384 : 16);
            const uint32_t b_lo = umma_desc_lo(smem_u32(smem + 32768), (TS || MODE == 4) ? 16384 : 16);
            t0 = clock64();
            for (int i = 0; i < n_mma; i += 8) {
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint32_t d = tmem + 256 + ((TWO && (k & 1)) ? 128 : 0);
                    if (TS) umma_bf16_ts_lo(d, tmem + (k & 3) * 8, b_lo + 128 * (k & 7), idesc, 1);
                    else if (MODE == 4) umma_bf16_ss_lo(d, a_lo + 128 * k, b_lo + 128 * k, idesc, 1);
                    else umma_bf16_ss_lo(d, a_lo + 2 * (k & 3), b_lo + 2 * (k & 3), idesc, 1);
                }
            }
            umma_commit(&bar);
            t1 = clock64();
            mbar_wait(&bar, 0);
            const long long t2 = clock64();
            cycles[blockIdx.x * 2] = t1 - t0;          // issue time
            cycles[blockIdx.x * 2 + 1] = t2 - t0;      // until the last MMA retired
            stop = 1;
        }
        __syncwarp();
    } else if (LOAD && warp < 8) {
        const uint32_t base = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16) + (warp >> 2) * 64;
        unsigned acc = 0;
        while (!stop) {
            uint32_t r[32];
            tmem_ld_32x32(base, r);
            tmem_ld_32x32(base + 32, r);
            tmem_ld_wait();
            acc += r[0];
        }
        if (acc == 0x12345678u) cycles[0] = 0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 512); }
    (void)lane;
}

template <int MODE>
static void run(const char* name, int n_mma, double flops_per_mma, long long* cyc) {
    const int smem = 160 * 1024;
    cudaFuncSetAttribute(mma_rate_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    long long h[2];
    for (int rep = 0; rep < 2; ++rep) {
        mma_rate_kernel<MODE><<<148, 320, smem>>>(n_mma, cyc);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    }
    cudaMemcpy(h, cyc, 16, cudaMemcpyDeviceToHost);
    printf("%-46s issue %6.1f clk/MMA   retire %6.1f clk/MMA   (%.0f%% of 8192 flop/clk/SM)\n", name, (double)h[0] / n_mma,
           (double)h[1] / n_mma, 100.0 * flops_per_mma / ((double)h[1] / n_mma) / 8192.0);
}

int main() {
    long long* cyc; cudaMalloc(&cyc, 148 * 16);
    const int n = 4096;
    run<0>("SS  M128 N128 K16, one accumulator", n, 2.0 * 128 * 128 * 16, cyc);
    run<1>("SS  M128 N128 K16, two accumulators", n, 2.0 * 128 * 128 * 16, cyc);
    run<2>("TS  M128 N64  K16, one accumulator", n, 2.0 * 128 * 64 * 16, cyc);
    run<3>("TS  M128 N64  K16, two accumulators", n, 2.0 * 128 * 64 * 16, cyc);
    run<4>("SS  M128 N64  K16 (MN-major A,B), two acc.", n, 2.0 * 128 * 64 * 16, cyc);
    run<5>("SS  M128 N128, one acc., 8 warps in tcgen05.ld", n, 2.0 * 128 * 128 * 16, cyc);
    run<6>("TS  M128 N64, two acc., 8 warps in tcgen05.ld", n, 2.0 * 128 * 64 * 16, cyc);
    return 0;
}

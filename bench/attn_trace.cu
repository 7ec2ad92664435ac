// Phase-level trace of the attention forward kernel: unity build of the library sources with TDL_ATTN_TRACE.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 --use_fast_math -I csrc bench/attn_trace.cu -o bench/bin/attn_trace
#ifndef TDL_ATTN_TRACE_TID
#define TDL_ATTN_TRACE_TID 33
#endif
#define TDL_ATTN_TRACE 1
#include <cstdint>
extern "C" int tdl_bias_grad(const void*, float*, int64_t, int64_t, int64_t, void*) { return 1; }   // not linked here
#include "../csrc/runtime.cu"
#include "../csrc/gemm_tcgen05.cu"
#include "../csrc/attention.cu"
#include <cstdio>
#include <vector>

int main() {
    const int B = 32, T = 1024, H = 16, C = H * 64;
    __nv_bfloat16 *qkv, *out; float* lse;
    cudaMalloc(&qkv, (size_t)B * T * 3 * C * 2); cudaMalloc(&out, (size_t)B * T * C * 2); cudaMalloc(&lse, (size_t)B * H * T * 4);
    std::vector<__nv_bfloat16> h((size_t)B * T * 3 * C);
    for (size_t i = 0; i < h.size(); ++i) h[i] = __float2bfloat16((float)((i * 2654435761u) % 2000) / 1000.f - 1.f);
    cudaMemcpy(qkv, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
    for (int rep = 0; rep < 3; ++rep) {
        int zero = 0;
        cudaMemcpyToSymbol(g_trace_n, &zero, sizeof(int));
        int rc = tdl_attention_fwd(qkv, out, lse, B, T, H, 64, nullptr);
        cudaError_t e = cudaDeviceSynchronize();
        if (rc || e != cudaSuccess) { printf("failed: %d %s %s\n", rc, cudaGetErrorString(e), tdl_last_error()); return 1; }
    }
    int n; cudaMemcpyFromSymbol(&n, g_trace_n, sizeof(int));
    std::vector<long long> tr(2 * n);
    cudaMemcpyFromSymbol(tr.data(), g_trace, sizeof(long long) * 2 * n);
    const char* names[] = {"", "loop top->", "S ready (wait bar_s)", "S in regs (tmem ld)", "max+alpha", "exp+P store", "o_acc scale", "fence+syncthreads", "PV issue (t0) / pass", "PV done (wait bar_o)", "O ld + accumulate"};
    printf("thread %d of CTA (0,0,0): %d stamps (blockIdx.x=0 is the LAST q block: %d key blocks)\n", TDL_ATTN_TRACE_TID, n, T / 128);
    for (int i = 1; i < n; ++i)
        printf("  %-28s %6lld cycles\n", names[tr[2 * i]], tr[2 * i + 1] - tr[2 * i - 1]);
    return 0;
}

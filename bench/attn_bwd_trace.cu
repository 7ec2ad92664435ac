// Phase-level trace of the persistent attention backward kernel (MODE 0 then MODE 1): unity build with TDL_ATTN_TRACE.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 --use_fast_math -I csrc -DTDL_ATTN_TRACE_TID=33 bench/attn_bwd_trace.cu -o bench/bin/attn_bwd_trace -lcuda
// TID < BWD_WS_MATH = a math thread, BWD_WS_MATH (512 with 4 threads per row) = the MMA-issue thread
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
    __nv_bfloat16 *qkv, *out, *dout, *dqkv; float *lse, *delta;
    cudaMalloc(&qkv, (size_t)B * T * 3 * C * 2); cudaMalloc(&dqkv, (size_t)B * T * 3 * C * 2);
    cudaMalloc(&out, (size_t)B * T * C * 2); cudaMalloc(&dout, (size_t)B * T * C * 2);
    cudaMalloc(&lse, (size_t)B * H * T * 4); cudaMalloc(&delta, (size_t)B * H * T * 4);
    std::vector<__nv_bfloat16> h((size_t)B * T * 3 * C);
    for (size_t i = 0; i < h.size(); ++i) h[i] = __float2bfloat16((float)((i * 2654435761u) % 2000) / 1000.f - 1.f);
    cudaMemcpy(qkv, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dout, h.data(), (size_t)B * T * C * 2, cudaMemcpyHostToDevice);
    tdl_attention_fwd(qkv, out, lse, B, T, H, 64, nullptr);
    for (int rep = 0; rep < 3; ++rep) {
        int zero = 0;
        cudaMemcpyToSymbol(g_trace_n, &zero, sizeof(int));
        int rc = tdl_attention_bwd_dbias(qkv, out, dout, lse, delta, dqkv, nullptr, B, T, H, 64, nullptr);
        cudaError_t e = cudaDeviceSynchronize();
        if (rc || e != cudaSuccess) { printf("failed: %d %s %s\n", rc, cudaGetErrorString(e), tdl_last_error()); return 1; }
    }
    int n; cudaMemcpyFromSymbol(&n, g_trace_n, sizeof(int));
    std::vector<long long> tr(2 * n);
    cudaMemcpyFromSymbol(tr.data(), g_trace, sizeof(long long) * 2 * n);
    printf("thread %d of CTA 0: %d stamps (both kernels back to back; tag: cycles since previous stamp)\n", TDL_ATTN_TRACE_TID, n);
    printf("math tags: 20 loop top | 21 S ready | 22 exp done | 23 dP ready | 24 dS computed+stored | 25 p_full arrived | 26 stats barrier | 27 acc ready | 28 acc written\n");
    printf("mma tags : 30 loop top | 32 p_full seen | 33 acc(j)+dP(j+1) issued | 34 acc(j) done | 35 S(j+2) issued\n");
    for (int i = 1; i < n && i < 400; ++i) {
        printf(" %2lld:%5lld", tr[2 * i], tr[2 * i + 1] - tr[2 * i - 1]);
        if (tr[2 * i] == 26 || tr[2 * i] == 28 || tr[2 * i] == 35) printf("\n");
    }
    printf("\n");
    return 0;
}

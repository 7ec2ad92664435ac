// K13: convolution epilogue for the ResNet / VGG stages: per-channel affine (conv bias or folded
// BatchNorm scale/shift) + optional residual + ReLU in one pass, and its backward (dReLU, per-channel
// d_scale / d_shift reductions).  NCHW and channels-last layouts, bf16 / fp16 / fp32.
// The reference leaves ResNet/VGG stages unimplemented (distributed_trainer.py:137-144); eager PyTorch
// would run conv -> BN -> add -> ReLU as four kernels with three extra round trips through HBM.
#include "common.cuh"

namespace {

template <typename T>
__global__ void __launch_bounds__(256) affine_act_fwd_kernel(const T* __restrict__ x, const float* __restrict__ scale,
        const float* __restrict__ shift, const T* __restrict__ res, T* __restrict__ y, long long total, int C, int HW,
        int channels_last, int relu) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = channels_last ? (int)(i % C) : (int)((i / HW) % C);
        float v = to_f32<T>(x[i]) * (scale != nullptr ? scale[c] : 1.f) + (shift != nullptr ? shift[c] : 0.f);
        if (res != nullptr) v += to_f32<T>(res[i]);
        if (relu) v = fmaxf(v, 0.f);
        y[i] = from_f32<T>(v);
    }
}

// dx = dy * [y > 0] * scale ; d_res = dy * [y > 0] ; d_shift[c] += sum dy*[y>0] ; d_scale[c] += sum dy*[y>0]*x
template <typename T>
__global__ void __launch_bounds__(256) affine_act_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ y,
        const T* __restrict__ x, const float* __restrict__ scale, T* __restrict__ dx, T* __restrict__ dres,
        float* __restrict__ dscale, float* __restrict__ dshift, int N, int C, int HW, int channels_last, int relu) {
    // one block per (channel, slab of the batch) keeps the channel reduction local
    const int c = blockIdx.x;
    const float sc = scale != nullptr ? scale[c] : 1.f;
    float a_shift = 0.f, a_scale = 0.f;
    const long long per_c = (long long)N * HW;
    for (long long j = (long long)blockIdx.y * blockDim.x + threadIdx.x; j < per_c; j += (long long)gridDim.y * blockDim.x) {
        const long long n = j / HW, s = j % HW;
        const long long i = channels_last ? (n * HW + s) * C + c : (n * C + c) * (long long)HW + s;
        float g = to_f32<T>(dy[i]);
        if (relu && !(to_f32<T>(y[i]) > 0.f)) g = 0.f;
        a_shift += g;
        if (x != nullptr) a_scale += g * to_f32<T>(x[i]);
        if (dres != nullptr) dres[i] = from_f32<T>(g);
        dx[i] = from_f32<T>(g * sc);
    }
    __shared__ float s0[8], s1[8];
    a_shift = warp_sum(a_shift); a_scale = warp_sum(a_scale);
    if ((threadIdx.x & 31) == 0) { s0[threadIdx.x >> 5] = a_shift; s1[threadIdx.x >> 5] = a_scale; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float t0 = 0, t1 = 0;
        for (int k = 0; k < 8; ++k) { t0 += s0[k]; t1 += s1[k]; }
        if (dshift != nullptr) atomicAdd(&dshift[c], t0);
        if (dscale != nullptr) atomicAdd(&dscale[c], t1);
    }
}

}  // namespace

#define DISPATCH(dtype, CALL)                                                          \
    switch (dtype) {                                                                   \
        case TDL_F32: { using T = float; CALL; break; }                                \
        case TDL_BF16: { using T = __nv_bfloat16; CALL; break; }                       \
        case TDL_F16: { using T = __half; CALL; break; }                               \
        default: return tdl_set_error("unsupported dtype %d", dtype);                  \
    }

extern "C" int tdl_conv_epilogue_fwd(const void* x, const float* scale, const float* shift, const void* residual, void* y,
                                     int64_t N, int64_t C, int64_t HW, int dtype, int channels_last, int relu, void* stream) {
    const long long total = N * C * HW;
    const int grid = (int)std::min<long long>((total + 255) / 256, (long long)tdl_num_sms() * 8);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    DISPATCH(dtype, (affine_act_fwd_kernel<T><<<grid, 256, 0, s>>>((const T*)x, scale, shift, (const T*)residual, (T*)y, total,
                                                                  (int)C, (int)HW, channels_last, relu)));
    TDL_LAUNCH_CHECK();
    return 0;
}

extern "C" int tdl_conv_epilogue_bwd(const void* dy, const void* y, const void* x, const float* scale, void* dx, void* dres,
                                     float* dscale, float* dshift, int64_t N, int64_t C, int64_t HW, int dtype,
                                     int channels_last, int relu, void* stream) {
    const long long per_c = N * HW;
    int gy = (int)std::min<long long>((per_c + 255) / 256, std::max<long long>(1, (long long)tdl_num_sms() * 8 / C));
    if (gy < 1) gy = 1;
    dim3 grid((unsigned)C, (unsigned)gy);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    DISPATCH(dtype, (affine_act_bwd_kernel<T><<<grid, 256, 0, s>>>((const T*)dy, (const T*)y, (const T*)x, scale, (T*)dx, (T*)dres,
                                                                  dscale, dshift, (int)N, (int)C, (int)HW, channels_last, relu)));
    TDL_LAUNCH_CHECK();
    return 0;
}

// reduce.cu -- small helpers around the CTC loss: deterministic sum of nll[B], gradient scaling.
#include "common.cuh"
#include "ctc_params.cuh"

namespace mdd {
namespace {

// fixed-order reduction: thread i sums elements i, i+1024, ... in order, then a fixed tree -> bitwise reproducible
__global__ void __launch_bounds__(1024) sum_nll_kernel(const float* __restrict__ nll, int B, float* __restrict__ out) {
    __shared__ float part[1024];
    const int tid = threadIdx.x;
    float s = 0.f;
    for (int i = tid; i < B; i += 1024) s += nll[i];
    part[tid] = s;
    __syncthreads();
    for (int d = 512; d > 0; d >>= 1) {
        if (tid < d) part[tid] += part[tid + d];
        __syncthreads();
    }
    if (tid == 0) *out = part[0];
}

__global__ void scale_grad_kernel(float* __restrict__ grad, long long n, int B, int C, const float* __restrict__ scale,
                                  float scalar) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int b = (int)((i / C) % B);
        grad[i] *= (scale ? scale[b] : 1.f) * scalar;
    }
}

}  // namespace

int launch_sum_nll(const float* nll, int B, float* out, cudaStream_t st) {
    sum_nll_kernel<<<1, 1024, 0, st>>>(nll, B, out);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

int launch_scale_grad(float* grad, int T, int B, int C, const float* scale, float scalar, cudaStream_t st) {
    const long long n = (long long)T * B * C;
    if (n == 0) return MDD_OK;
    const int grid = (int)min((n + 255) / 256, (long long)kNumSM * 16);
    scale_grad_kernel<<<grid, 256, 0, st>>>(grad, n, B, C, scale, scalar);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

}  // namespace mdd

// micro-benchmark: latency / issue rate of legacy mma.sync shapes on sm_100a, one warp per SM and 8 warps per SM
#include <cstdio>
#include <cuda_runtime.h>
template <int KIND, int CH>
__global__ void k(float* out, long long* cyc, int iters) {
    float c[CH][4];
    for (int i = 0; i < CH; ++i) for (int j = 0; j < 4; ++j) c[i][j] = 0.f;
    unsigned a0 = threadIdx.x, a1 = threadIdx.x * 3, a2 = 5, a3 = 7, b0 = 0x3f800000u, b1 = 0;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (KIND == 0)
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                    : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else if (KIND == 1)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                    : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else if (KIND == 2)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                    : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else
                asm volatile("mma.sync.aligned.m16n8k4.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                    : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(b0));
        }
    }
    long long t1 = clock64();
    float s = 0; for (int i = 0; i < CH; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int KIND, int CH> void run(const char* name, int warps) {
    float* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    const int iters = 1000;
    k<KIND, CH><<<148, 32 * warps>>>(out, cyc, iters); cudaDeviceSynchronize();
    k<KIND, CH><<<148, 32 * warps>>>(out, cyc, iters); cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-22s chains=%d warps/SM=%d: %.1f cycles per mma (per warp)  err=%s\n", name, CH, warps, (double)h / iters / CH, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}
int main() {
    run<0, 1>("m16n8k8 tf32", 1); run<0, 4>("m16n8k8 tf32", 1); run<0, 4>("m16n8k8 tf32", 8);
    run<3, 1>("m16n8k4 tf32", 1); run<3, 4>("m16n8k4 tf32", 1);
    run<1, 1>("m16n8k16 bf16", 1); run<1, 4>("m16n8k16 bf16", 1); run<1, 4>("m16n8k16 bf16", 8);
    run<2, 1>("m16n8k16 f16", 1); run<2, 4>("m16n8k16 f16", 1);
    return 0;
}

// Throughput probe (tools only): legacy mma.sync.m16n8k8 TF32 on this GPU, next to plain FFMA, to decide whether an error-compensated
// 3xTF32 version of k_ppo_grad (three MMAs per product, fp32-level accuracy) could beat the FFMA kernel.  No memory traffic: every warp
// keeps NACC independent accumulator fragments and issues MMAs on register operands.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/mma_tf32_probe tools/mma_tf32_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int NACC>
__global__ void __launch_bounds__(256) k_mma(float* out, int iters) {
    float c[NACC][4];
    unsigned a[4], b[2];
    for (int k = 0; k < 4; ++k) a[k] = __float_as_uint(1.0f + 0.001f * (threadIdx.x + k));
    for (int k = 0; k < 2; ++k) b[k] = __float_as_uint(0.5f + 0.001f * (threadIdx.x + k));
#pragma unroll
    for (int j = 0; j < NACC; ++j) for (int k = 0; k < 4; ++k) c[j][k] = 0.f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < NACC; ++j)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                         : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3])
                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NACC; ++j) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) k_ffma(float* out, int iters) {
    float c[16];
    const float a = 1.0f + 1e-6f * threadIdx.x, b = 0.5f;
#pragma unroll
    for (int j = 0; j < 16; ++j) c[j] = (float)j;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) c[j] = fmaf(c[j], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) s += c[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount, ctas = sms * 8, iters = 20000;
    float* out; cudaMalloc(&out, sizeof(float) * ctas * 256);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto time = [&](auto launch) { launch(); cudaDeviceSynchronize(); cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
                                   float ms; cudaEventElapsedTime(&ms, e0, e1); return ms; };
    const double warps = (double)ctas * 8;
    float ms = time([&] { k_mma<8><<<ctas, 256>>>(out, iters); });
    printf("%s, %d SMs\n", p.name, sms);
    printf("mma.sync m16n8k8 tf32, 8 accumulators/warp: %.1f TFLOP/s (%.3f ms)\n", warps * iters * 8 * 2.0 * 16 * 8 * 8 / (ms * 1e-3) / 1e12, ms);
    ms = time([&] { k_mma<4><<<ctas, 256>>>(out, iters); });
    printf("mma.sync m16n8k8 tf32, 4 accumulators/warp: %.1f TFLOP/s (%.3f ms)\n", warps * iters * 4 * 2.0 * 16 * 8 * 8 / (ms * 1e-3) / 1e12, ms);
    ms = time([&] { k_ffma<<<ctas, 256>>>(out, iters); });
    printf("ffma, 16 independent chains/thread: %.1f TFLOP/s (%.3f ms)\n", (double)ctas * 256 * iters * 16 * 2.0 / (ms * 1e-3) / 1e12, ms);
    printf("last error: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}

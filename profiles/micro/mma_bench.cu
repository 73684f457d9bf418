// Issue rate of the warp-level mma.sync forms on this GPU (the tensor-core path a warp-per-frame
// kernel can use without TMEM): TF32 m16n8k8, BF16 m16n8k16.  Prints MMA per clock per SM and TFLOP/s.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_bench mma_bench.cu && ./mma_bench
#include <cstdio>
#include <cuda_runtime.h>

template <int KIND, int ILP>
__global__ void k(float *out, int iters) {
    float c[ILP][4];
    unsigned a[4] = {threadIdx.x, threadIdx.x * 3u, 7u, 9u}, b[2] = {threadIdx.x + 1u, 5u};
#pragma unroll
    for (int i = 0; i < ILP; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.0f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            if (KIND == 0)
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                             : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
            else
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                             : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int KIND, int ILP>
void run(const char *name, double flop_per_mma, int warps) {
    int dev = 0, sms = 0, khz = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    float *out;
    cudaMalloc(&out, sizeof(float) * sms * warps * 32);
    const int iters = 20000;
    k<KIND, ILP><<<sms, warps * 32>>>(out, 100);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<KIND, ILP><<<sms, warps * 32>>>(out, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double mmas = (double)sms * warps * iters * ILP;
    printf("{\"op\": \"%s\", \"warps_per_sm\": %d, \"ilp\": %d, \"mma_per_s\": %.4g, \"mma_per_clk_per_sm\": %.3f, \"tflops\": %.1f}\n",
           name, warps, ILP, mmas / (ms * 1e-3), mmas / (ms * 1e-3) / sms / (khz * 1e3), mmas * flop_per_mma / (ms * 1e-3) / 1e12);
    cudaFree(out);
}

int main() {
    run<0, 4>("mma.sync m16n8k8 tf32", 2.0 * 16 * 8 * 8, 4);
    run<0, 4>("mma.sync m16n8k8 tf32", 2.0 * 16 * 8 * 8, 8);
    run<0, 8>("mma.sync m16n8k8 tf32", 2.0 * 16 * 8 * 8, 16);
    run<1, 4>("mma.sync m16n8k16 bf16", 2.0 * 16 * 8 * 16, 4);
    run<1, 8>("mma.sync m16n8k16 bf16", 2.0 * 16 * 8 * 16, 16);
    return 0;
}

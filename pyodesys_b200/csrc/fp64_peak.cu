// FP64 roofline denominator: sustained DFMA throughput of the whole chip, measured live.
// MEASURED_PEAKS.json (driver-written) has HBM and bf16 figures only; the batched ODE kernels are
// bound by the FP64 pipe, so bench.py measures this peak on the same box right before reporting
// roofline.frac.  16 independent fma.rn.f64 chains per thread, 8 warps per block, every SM filled.
#include <cuda_runtime.h>
#include <stdint.h>

__global__ void __launch_bounds__(256) fp64_fma_kernel(double *out, int iters, double a, double b) {
    double x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = 1.0 + 1e-3 * (threadIdx.x + i);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) x[i] = fma(x[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += x[i];
    if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;  // never true; keeps the chains alive
}

extern "C" int pyodesys_cuda_fp64_peak(int device, double seconds, double *tflops_best, double *tflops_sustained) {
    if (cudaSetDevice(device) != cudaSuccess) return -1;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return -1;
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 14;
    double *out = nullptr;
    if (cudaMalloc(&out, (size_t)blocks * threads * 8) != cudaSuccess) return -1;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const double flops = 2.0 * 16.0 * iters * (double)blocks * threads;
    double best = 0, total_flops = 0, total_s = 0;
    for (int w = 0; w < 3; ++w) fp64_fma_kernel<<<blocks, threads>>>(out, iters, 0.999999, 1e-6);
    cudaDeviceSynchronize();
    while (total_s < seconds) {
        cudaEventRecord(e0);
        fp64_fma_kernel<<<blocks, threads>>>(out, iters, 0.999999, 1e-6);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) return -1;
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double s = ms * 1e-3;
        if (flops / s > best) best = flops / s;
        total_flops += flops;
        total_s += s;
    }
    *tflops_best = best * 1e-12;
    *tflops_sustained = total_flops / total_s * 1e-12;
    cudaFree(out);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return 0;
}

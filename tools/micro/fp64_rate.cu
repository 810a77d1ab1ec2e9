// Throughput of non-fused FP64 (DADD / DMUL) and of FP32 on one GPU: N independent chains per thread.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 --fmad=false -o tools/micro/fp64_rate tools/micro/fp64_rate.cu
#include <cstdio>
#include <cuda_runtime.h>

template <class T, int CH>
__global__ void chain(T *out, T a, T b, int iters)
{
    T v[CH];
    for (int c = 0; c < CH; ++c) v[c] = a + (T)(threadIdx.x + c);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int c = 0; c < CH; ++c) v[c] = v[c] * b + a;      // --fmad=false: one multiply + one add
    }
    T s = 0;
    for (int c = 0; c < CH; ++c) s += v[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class T>
static void run(const char *name, int sms)
{
    T *out;
    cudaMalloc(&out, sizeof(T) * sms * 8 * 256);
    const int iters = 20000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    chain<T, 8><<<sms * 8, 256>>>(out, (T)1.0000001, (T)0.9999999, 100);
    cudaEventRecord(e0);
    chain<T, 8><<<sms * 8, 256>>>(out, (T)1.0000001, (T)0.9999999, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double ops = 2.0 * 8 * iters * (double)sms * 8 * 256;
    printf("%s: %.3f ms, %.2f Tops/s (mul and add counted separately), %.1f ops/clk/SM at 1.9 GHz\n", name, ms, ops / ms * 1e-9, ops / (ms * 1e-3) / sms / 1.9e9);
    cudaFree(out);
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs\n", p.name, p.multiProcessorCount);
    run<double>("fp64", p.multiProcessorCount);
    run<float>("fp32", p.multiProcessorCount);
    return 0;
}

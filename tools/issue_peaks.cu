// issue_peaks.cu — measures the FP64 and INT32 issue-rate peaks of the device (MEASURED_PEAKS.json has no
// FP64/INT entry, and the AD path's relevant roofline is the issue rate).  Dependency-free chains of DFMA / DADD /
// DSETP+select (min) / IMAD, 8 independent accumulators per thread, enough resident warps to hide the pipe latency.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/issue_peaks tools/issue_peaks.cu && tools/issue_peaks
// Prints one JSON object; bench.py reads profiles/issue_peaks.json (a committed copy of that output).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

constexpr int kIters = 4096;
constexpr int kAcc = 8;

template <int OP>
__global__ void __launch_bounds__(256) k_issue(double* out, double a, double b, int iters)
{
    double v[kAcc];
    int w[kAcc];
#pragma unroll
    for (int i = 0; i < kAcc; ++i) { v[i] = (double) (threadIdx.x + i) * 1e-3; w[i] = threadIdx.x + i; }
    const int ia = (int) a + 3, ib = (int) b + 7;
    for (int it = 0; it < iters; ++it)
    {
#pragma unroll
        for (int i = 0; i < kAcc; ++i)
        {
            if (OP == 0) v[i] = __fma_rn(v[i], a, b);
            else if (OP == 1) v[i] = __dadd_rn(v[i], b);
            else if (OP == 2) v[i] = fmin(v[i], __longlong_as_double(__double_as_longlong(a) + it));   // DSETP + 2 SEL, plus one IADD
            else if (OP == 3) w[i] = w[i] * ia + ib;
            else if (OP == 4) v[i] = __dmul_rn(v[i], a);
        }
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < kAcc; ++i) s += v[i] + (double) w[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
static double run(int sms, double* d_out, int ctas_per_sm)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = sms * ctas_per_sm;
    k_issue<OP><<<grid, 256>>>(d_out, 1.0000001, 1e-9, kIters);       // warm-up
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep)
    {
        cudaEventRecord(e0);
        k_issue<OP><<<grid, 256>>>(d_out, 1.0000001, 1e-9, kIters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0.f; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double ops = (double) grid * 256 * kIters * kAcc;
    return ops / (best * 1e-3);          // thread-level instructions per second
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    double* d_out = nullptr;
    cudaMalloc(&d_out, sizeof(double) * p.multiProcessorCount * 8 * 256);
    const int sms = p.multiProcessorCount;
    const double dfma = run<0>(sms, d_out, 8), dadd = run<1>(sms, d_out, 8), dmin = run<2>(sms, d_out, 8), imad = run<3>(sms, d_out, 8),
                 dmul = run<4>(sms, d_out, 8);
    if (cudaDeviceSynchronize() != cudaSuccess) { fprintf(stderr, "CUDA error\n"); return 1; }
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_khz_max\": %d, \"dfma_per_s\": %.4e, \"dadd_per_s\": %.4e, \"dmul_per_s\": %.4e, "
           "\"dmin_per_s\": %.4e, \"imad_per_s\": %.4e, \"fp64_tflops_fma\": %.3f, \"dfma_per_sm_per_clk_at_max\": %.2f, "
           "\"how\": \"tools/issue_peaks.cu: 8 independent accumulators per thread, 8 CTAs x 256 threads per SM, %d iterations, best of 5, CUDA events\"}\n",
           p.name, sms, clk, dfma, dadd, dmul, dmin, imad, 2.0 * dfma / 1e12, dfma / sms / (clk * 1e3), kIters);
    return 0;
}

// fp32_pipes.cu — issue-rate microbenchmark behind DESIGN.md §6 (development helper, not part of the library):
// how many warp instructions per cycle and scheduler the FP32 pipes of an sm_100a SM accept for scalar FFMA, packed FFMA2
// (two FP32 lanes per thread) and mixes of them with each other and with ALU-pipe work (FSETP/LOP3), measured with
// clock64() on resident CTAs.  Answers: does a packed instruction occupy the FMA pipe for one issue slot or for two, and
// can scalar FFMA (or compares) issue in the shadow of packed ones.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp32_pipes tools/fp32_pipes.cu && ./fp32_pipes
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kIters = 2048;

// P packed chains, S scalar chains, C compare+select chains per thread, each advanced once per inner step.
template <int P, int S, int C>
__global__ void __launch_bounds__(256) k_mix(float* out, float a, float b, unsigned long long* cyc) {
    float2 p[P > 0 ? P : 1];
    float s[S > 0 ? S : 1];
    float c[C > 0 ? C : 1];
    for (int i = 0; i < P; ++i) p[i] = make_float2(threadIdx.x + i, threadIdx.x - i);
    for (int i = 0; i < S; ++i) s[i] = threadIdx.x + 0.5f * i;
    for (int i = 0; i < C; ++i) c[i] = threadIdx.x * 0.25f + i;
    const float2 a2 = make_float2(a, a), b2 = make_float2(b, b);
    __syncthreads();
    unsigned long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < kIters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < P; ++i) p[i] = __ffma2_rn(p[i], a2, b2);
#pragma unroll
            for (int i = 0; i < S; ++i) s[i] = __fmaf_rn(s[i], a, b);
#pragma unroll
            for (int i = 0; i < C; ++i) c[i] = fmaxf(c[i] > a ? b : c[i], s[0] * 0.0f + b);   // FSETP + FSEL + FMNMX (ALU pipe)
        }
    }
    unsigned long long t1 = clock64();
    __syncthreads();
    float acc = 0.f;
    for (int i = 0; i < P; ++i) acc += p[i].x + p[i].y;
    for (int i = 0; i < S; ++i) acc += s[i];
    for (int i = 0; i < C; ++i) acc += c[i];
    if (acc == 12345.678f) out[0] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int P, int S, int C>
void run(const char* name, int sms) {
    float* d; unsigned long long* cyc;
    const int grid = sms * 4;   // 4 CTAs of 256 threads per SM = 8 warps per scheduler
    cudaMalloc(&d, 4); cudaMalloc(&cyc, 8 * grid);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_mix<P, S, C><<<grid, 256>>>(d, 0.999f, 0.001f, cyc);
    cudaEventRecord(e0);
    k_mix<P, S, C><<<grid, 256>>>(d, 0.999f, 0.001f, cyc);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    unsigned long long* h = new unsigned long long[grid];
    cudaMemcpy(h, cyc, 8 * grid, cudaMemcpyDeviceToHost);
    double mean = 0; for (int i = 0; i < grid; ++i) mean += (double)h[i]; mean /= grid;
    // per scheduler: 8 warps, each issuing kIters * 4 * (P + S [+ ~3C]) instructions
    const double wi_p = 8.0 * kIters * 4 * P, wi_s = 8.0 * kIters * 4 * S;
    const double flops = (double)grid * 256 * kIters * 4 * (4.0 * P + 2.0 * S);
    printf("%-28s P=%d S=%d C=%d  cycles %.0f  packed/clk/sched %.3f  scalar/clk/sched %.3f  flop/clk/SM %.1f  %.2f TFLOP/s (%.3f ms)\n", name, P, S, C,
           mean, wi_p / mean, wi_s / mean, (4.0 * P + 2.0 * S) * 32 * 8 * 4 * kIters * 4 / mean, flops / (ms * 1e-3) / 1e12, ms);
    delete[] h; cudaFree(d); cudaFree(cyc);
}

int main() {
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    printf("%s, %d SMs\n", pr.name, pr.multiProcessorCount);
    const int sms = pr.multiProcessorCount;
    run<0, 8, 0>("scalar FFMA", sms);
    run<8, 0, 0>("packed FFMA2", sms);
    run<4, 4, 0>("FFMA2 : FFMA = 1 : 1", sms);
    run<4, 8, 0>("FFMA2 : FFMA = 1 : 2", sms);
    run<6, 2, 0>("FFMA2 : FFMA = 3 : 1", sms);
    run<0, 8, 2>("FFMA + compares", sms);
    run<8, 0, 2>("FFMA2 + compares", sms);
    run<8, 0, 4>("FFMA2 + more compares", sms);
    run<0, 0, 8>("compares only", sms);
    return 0;
}

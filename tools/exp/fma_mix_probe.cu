// How do the packed (FFMA2) and scalar (FFMA) fp32 multiply-adds share the FMA pipes of sm_100?  ncu shows FFMA2 on the
// "fmaheavy" pipe only.  Per variant: independent register chains, 8 warps per scheduler, FMAs per clock and SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fma_mix_probe fma_mix_probe.cu && ./fma_mix_probe
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    float2 d;
    asm volatile("{.reg .b64 ra, rb, rc, rd; mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %5}; mov.b64 rc, {%6, %7};\n\t"
                 "fma.rn.f32x2 rd, ra, rb, rc; mov.b64 {%0, %1}, rd;}"
                 : "=f"(d.x), "=f"(d.y)
                 : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return d;
}

__device__ __forceinline__ float ffma1(float a, float b, float c) {  // asm: ptxas would pair scalar chains into FFMA2
    float d;
    asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// P packed chains and S scalar chains per thread, interleaved
template <int P, int S>
__global__ void __launch_bounds__(1024) probe(float* out, int iters, float w) {
    float2 p[P > 0 ? P : 1];
    float s[S > 0 ? S : 1];
    for (int i = 0; i < P; ++i) p[i] = make_float2(threadIdx.x * 1e-3f + i, i * 0.5f);
    for (int i = 0; i < S; ++i) s[i] = threadIdx.x * 1e-3f - i;
    const float2 w2 = make_float2(w, w);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
#pragma unroll
            for (int i = 0; i < (P > S ? P : S); ++i) {
                if (i < P) p[i] = ffma2(p[i], w2, w2);
                if (i < S) s[i] = ffma1(s[i], w, w);
            }
        }
    }
    float acc = 0.f;
    for (int i = 0; i < P; ++i) acc += p[i].x + p[i].y;
    for (int i = 0; i < S; ++i) acc += s[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int P, int S>
static void run(const char* name, int sms, double ghz) {
    float* out;
    cudaMalloc(&out, (size_t)sms * 1024 * sizeof(float));
    const int iters = 4000;
    probe<P, S><<<sms, 1024>>>(out, 10, 0.999f);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    cudaEventRecord(e0);
    probe<P, S><<<sms, 1024>>>(out, iters, 0.999f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fmas = (double)sms * 1024 * iters * 8 * (2.0 * P + S);
    printf("%-28s %8.3f ms  %7.1f FMA / clk / SM at %.3f GHz  (%5.1f TFLOP/s)\n", name, ms, fmas / (ms * 1e-3) / sms / (ghz * 1e9),
           ghz, 2 * fmas / (ms * 1e-3) / 1e12);
    cudaFree(out);
}

int main() {
    cudaDeviceProp pr;
    cudaGetDeviceProperties(&pr, 0);
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double ghz = khz / 1e6;
    const int sms = pr.multiProcessorCount;
    run<0, 8>("scalar FFMA only", sms, ghz);
    run<8, 0>("packed FFMA2 only", sms, ghz);
    run<4, 4>("FFMA2 : FFMA = 1 : 1", sms, ghz);
    run<6, 3>("FFMA2 : FFMA = 2 : 1", sms, ghz);
    run<4, 8>("FFMA2 : FFMA = 1 : 2", sms, ghz);
    run<8, 2>("FFMA2 : FFMA = 4 : 1", sms, ghz);
    return 0;
}

// HBM ceiling for the read:write mix of the Zero-DCE++ full-resolution pass (1 stream in, 2 streams out) next to a
// plain copy (1 in, 1 out), same access pattern as dcepp_apply_vec4_kernel: persistent grid, 128-bit streaming
// accesses.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/exp/rw_mix_probe tools/exp/rw_mix_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ float4 ld4(const float* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void st4(float* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

template <int NOUT, int UNROLL>
__global__ void __launch_bounds__(256) mix(const float* __restrict__ in, float* __restrict__ o1, float* __restrict__ o2, size_t n4) {
    const size_t stride = (size_t)gridDim.x * 256 * UNROLL;
    for (size_t i = (size_t)blockIdx.x * 256 * UNROLL + threadIdx.x; i < n4; i += stride) {
        float4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
            if (i + u * 256 < n4) v[u] = ld4(in + 4 * (i + u * 256));
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
            if (i + u * 256 < n4) {
                float4 a = v[u];
                st4(o1 + 4 * (i + u * 256), a);
                if (NOUT == 2) {
                    a.x += 1.f, a.y += 1.f, a.z += 1.f, a.w += 1.f;
                    st4(o2 + 4 * (i + u * 256), a);
                }
            }
    }
}

template <int NOUT, int UNROLL>
static void run(const char* name, const float* in, float* o1, float* o2, size_t n, int ctas_per_sm) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = 148 * ctas_per_sm;
    for (int i = 0; i < 3; ++i) mix<NOUT, UNROLL><<<grid, 256>>>(in, o1, o2, n / 4);
    cudaEventRecord(e0);
    const int reps = 20;
    for (int i = 0; i < reps; ++i) mix<NOUT, UNROLL><<<grid, 256>>>(in, o1, o2, n / 4);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    ms /= reps;
    const double bytes = (double)n * 4 * (1 + NOUT);
    printf("%-28s grid=%4d  %.4f ms  %.1f GB/s  (%s)\n", name, grid, ms, bytes / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    const size_t n = (size_t)16 * 3 * 1080 * 1920;  // floats per stream
    float *in, *o1, *o2;
    cudaMalloc(&in, n * 4);
    cudaMalloc(&o1, n * 4);
    cudaMalloc(&o2, n * 4);
    cudaMemset(in, 0, n * 4);
    for (int rep = 0; rep < 2; ++rep)
        for (int c : {2, 4, 8}) {
            run<2, 2>("mix 1:2 unroll 2", in, o1, o2, n, c);
            run<2, 4>("mix 1:2 unroll 4", in, o1, o2, n, c);
            run<2, 8>("mix 1:2 unroll 8", in, o1, o2, n, c);
            run<2, 16>("mix 1:2 unroll 16", in, o1, o2, n, c);
        }
    float ms;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    for (int i = 0; i < 20; ++i) cudaMemcpyAsync(o1, in, n * 4, cudaMemcpyDeviceToDevice);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    printf("cudaMemcpy D2D               %.4f ms  %.1f GB/s\n", ms / 20, (double)n * 8 / (ms / 20) / 1e6);
    return 0;
}

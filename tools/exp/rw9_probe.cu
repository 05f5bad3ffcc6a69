// Compute-free model of dcepp_apply_vec4_kernel's traffic: x [B][3][H][W] in, enh and adj [B][3][H][W] out, a CTA owns
// strips of R rows.  Which strip order / plane order does HBM like?
//   V0: as the kernel (strided strips; per thread 3 loads then 6 stores)      V1: plane after plane inside a strip
//   V2: V0 with a contiguous range of strips per CTA                           V3: V0 with 2 groups per thread in flight
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/exp/rw9_probe tools/exp/rw9_probe.cu
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

template <int V>
__global__ void __launch_bounds__(256) k(const float* __restrict__ x, float* __restrict__ enh, float* __restrict__ adj,
                                         int B, int H, int W, int R) {
    const int spf = H / R, total = B * spf, W4 = W / 4, t = threadIdx.x;
    const size_t plane = (size_t)H * W;
    int s0 = blockIdx.x, s1 = total, ds = gridDim.x;
    if (V == 2) {
        const int per = (total + gridDim.x - 1) / gridDim.x;
        s0 = blockIdx.x * per;
        s1 = min(total, s0 + per);
        ds = 1;
    }
    for (int strip = s0; strip < s1; strip += ds) {
        const int b = strip / spf, Y0 = (strip - b * spf) * R;
        const size_t fb = (size_t)b * 3 * plane + (size_t)Y0 * W;
        const int n = R * W4;
        if (V == 1) {
            for (int c = 0; c < 3; ++c)
                for (int it = t; it < n; it += 256) {
                    const size_t o = fb + c * plane + (size_t)it * 4;
                    float4 v = ld4(x + o);
                    st4(enh + o, v);
                    v.x += 1.f;
                    st4(adj + o, v);
                }
        } else if (V == 3) {
            for (int it = t; it < n; it += 512) {
                float4 v[2][3];
                const bool two = it + 256 < n;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    v[0][c] = ld4(x + fb + c * plane + (size_t)it * 4);
                    if (two) v[1][c] = ld4(x + fb + c * plane + (size_t)(it + 256) * 4);
                }
#pragma unroll
                for (int u = 0; u < 2; ++u)
                    if (u == 0 || two)
#pragma unroll
                        for (int c = 0; c < 3; ++c) {
                            const size_t o = fb + c * plane + (size_t)(it + u * 256) * 4;
                            float4 a = v[u][c];
                            st4(enh + o, a);
                            a.x += 1.f;
                            st4(adj + o, a);
                        }
            }
        } else {
            for (int it = t; it < n; it += 256) {
                float4 v[3];
#pragma unroll
                for (int c = 0; c < 3; ++c) v[c] = ld4(x + fb + c * plane + (size_t)it * 4);
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const size_t o = fb + c * plane + (size_t)it * 4;
                    float4 a = v[c];
                    st4(enh + o, a);
                    a.x += 1.f;
                    st4(adj + o, a);
                }
            }
        }
    }
}

template <int V>
static void run(const float* x, float* e, float* a, int B, int H, int W, int R, int cps) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = 148 * cps;
    for (int i = 0; i < 3; ++i) k<V><<<grid, 256>>>(x, e, a, B, H, W, R);
    cudaEventRecord(e0);
    for (int i = 0; i < 20; ++i) k<V><<<grid, 256>>>(x, e, a, B, H, W, R);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    ms /= 20;
    printf("V%d R=%d ctas/SM=%d  %.4f ms  %.1f GB/s (%s)\n", V, R, cps, ms, (double)B * 3 * H * W * 12 / ms / 1e6,
           cudaGetErrorString(cudaGetLastError()));
}

int main() {
    const int B = 16, H = 1080, W = 1920;
    const size_t n = (size_t)B * 3 * H * W;
    float *x, *e, *a;
    cudaMalloc(&x, n * 4);
    cudaMalloc(&e, n * 4);
    cudaMalloc(&a, n * 4);
    cudaMemset(x, 0, n * 4);
    for (int rep = 0; rep < 2; ++rep)
        for (int R : {1, 2, 4, 8})
            for (int cps : {2, 4, 8}) {
                run<0>(x, e, a, B, H, W, R, cps);
                run<1>(x, e, a, B, H, W, R, cps);
                run<2>(x, e, a, B, H, W, R, cps);
                run<3>(x, e, a, B, H, W, R, cps);
            }
    return 0;
}

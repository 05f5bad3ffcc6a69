// Probe 2: SS-mode tcgen05.mma with the A operand read straight from a TMA-SWIZZLE_64B pixel-major tile at a start
// address shifted by dx pixels; correctness (D = A * I) and pacing vs N.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_bf16.h>
#include "../../etss_07_imageenhance_b200/csrc/tcgen05.cuh"
using namespace b200::tc;

__device__ __forceinline__ uint64_t desc_sw64(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(512 >> 4) << 32) | (1ull << 46) | (4ull << 61);
}
__device__ __forceinline__ void mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc), "r"(0u) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr) : "memory");
}

// which: 0 -> A value = pixel index, 1 -> A value = channel index
__global__ void __launch_bounds__(128) probe(float* out, int which) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t sbase = (smem_u32(smem) + 1023u) & ~1023u;
    uint8_t* tile = smem + (sbase - smem_u32(smem));
    for (int i = tid; i < 130 * 32; i += 128) {
        const int p = i >> 5, c = i & 31;
        const int off = p * 64 + (((c >> 3) ^ ((p >> 1) & 3)) << 4) + (c & 7) * 2;
        *reinterpret_cast<__nv_bfloat16*>(tile + off) = __float2bfloat16_rn((float)(which ? c : p));
    }
    // B = [N=16][K=16] identity, K-major no-swizzle core matrices (LBO 128, SBO 256) at tile + 12288
    __nv_bfloat16* Bm = reinterpret_cast<__nv_bfloat16*>(tile + 12288);
    for (int i = tid; i < 256; i += 128) {
        const int n = i >> 4, kk = i & 15;
        Bm[(n >> 3) * 128 + (kk >> 3) * 64 + (n & 7) * 8 + (kk & 7)] = __float2bfloat16_rn(n == kk ? 1.f : 0.f);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 0) {
        if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
        __syncwarp();
        tmem_alloc(&s_tmem, 128);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = s_tmem;
    if (tid == 0) {
        const uint32_t idesc = idesc_bf16_f32(128, 16);
        const uint64_t bdesc = smem_desc_kmajor(sbase + 12288, 128, 256);
        for (int dx = 0; dx < 3; ++dx)
            for (int kc = 0; kc < 2; ++kc)
                mma_ss(tb + (dx * 2 + kc) * 16, desc_sw64(sbase + dx * 64 + kc * 32), bdesc, idesc, 0u);
        tc_commit(&bar);
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    const uint32_t tl = tb + ((uint32_t)(warp * 32) << 16);
    for (int j = 0; j < 6; ++j) {
        uint32_t v[16];
        tmem_ld16(tl + j * 16, v);
        tc_wait_ld();
        for (int i = 0; i < 16; ++i) out[(j * 128 + tid) * 16 + i] = __uint_as_float(v[i]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tb, 128);
}


template <int MODE, int N, int NMMA>
__global__ void __launch_bounds__(128) pace(long long* out, int reps) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t sbase = (smem_u32(smem) + 1023u) & ~1023u;
    for (int i = tid; i < 150000 / 4; i += 128) reinterpret_cast<uint32_t*>(smem + (sbase - smem_u32(smem)))[i] = 0;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 0) {
        if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
        __syncwarp();
        tmem_alloc(&s_tmem, 512);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = s_tmem;
    if (tid == 0) {
        constexpr uint32_t idesc = idesc_bf16_f32(128, N);
        const uint64_t bdesc = smem_desc_kmajor(sbase + 32768, 128, 256);
        const uint64_t adesc = desc_sw64(sbase);
        long long t0 = clock64();
#pragma unroll 1
        for (int r = 0; r < reps; ++r) {
            const uint32_t dacc = tb + (r & 1) * 96;
#pragma unroll
            for (int j = 0; j < NMMA; ++j) {
                if (MODE == 0) mma_ts(dacc, tb + 256 + j * 8, bdesc + (uint64_t)((j * 3072) >> 4), idesc, 1u);
                else mma_ss(dacc, adesc + (uint64_t)(((j % 3) * 64 + ((j / 3) & 1) * 32 + (j / 6) * 8704) >> 4),
                            bdesc + (uint64_t)((j * 3072) >> 4), idesc, 1u);
            }
        }
        tc_commit(&bar);
        mbar_wait(&bar, 0);
        out[0] = clock64() - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc(tb, 512); }
}
template <int MODE, int N, int NMMA>
void run_pace(long long* dt) {
    cudaFuncSetAttribute(pace<MODE, N, NMMA>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160000);
    long long t = 0;
    for (int it = 0; it < 2; ++it) {
        pace<MODE, N, NMMA><<<1, 128, 160000>>>(dt, 2000);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("pace: CUDA error %s\n", cudaGetErrorString(e)); exit(1); }
        cudaMemcpy(&t, dt, 8, cudaMemcpyDeviceToHost);
    }
    printf("pace %s N=%d nmma=%d: %.1f clk/row = %.1f clk/mma\n", MODE ? "SS" : "TS", N, NMMA, t / 2000.0, t / 2000.0 / NMMA);
}
int main() {
    float* d; cudaMalloc(&d, 6 * 128 * 16 * 4);
    std::vector<float> h(6 * 128 * 16);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 20480);
    for (int which = 0; which < 2; ++which) {
        probe<<<1, 128, 20480>>>(d, which);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("probe: CUDA error %s\n", cudaGetErrorString(e)); return 1; }
        cudaMemcpy(h.data(), d, h.size() * 4, cudaMemcpyDeviceToHost);
        for (int dx = 0; dx < 3; ++dx) for (int kc = 0; kc < 2; ++kc) {
            int bad = 0, first = -1;
            for (int m = 0; m < 128; ++m) for (int i = 0; i < 16; ++i) {
                const float exp = which ? (float)(kc * 16 + i) : (float)(m + dx);
                if (h[((dx * 2 + kc) * 128 + m) * 16 + i] != exp) { if (first < 0) first = m * 16 + i; ++bad; }
            }
            printf("SS which %d dx %d kc %d: %s (%d bad", which, dx, kc, bad ? "MISMATCH" : "ok", bad);
            if (bad) printf(", first lane %d col %d got %g", first / 16, first % 16, h[((dx * 2 + kc) * 128 + first / 16) * 16 + first % 16]);
            printf(")\n");
        }
    }
    long long* dt; cudaMalloc(&dt, 8);
    run_pace<0, 96, 6>(dt); run_pace<0, 96, 12>(dt); run_pace<1, 96, 6>(dt); run_pace<1, 96, 12>(dt);
    run_pace<0, 32, 12>(dt); run_pace<1, 32, 12>(dt); run_pace<0, 64, 12>(dt); run_pace<1, 64, 12>(dt);
    run_pace<0, 128, 12>(dt); run_pace<1, 128, 12>(dt); run_pace<0, 192, 12>(dt); run_pace<1, 192, 12>(dt);
    return 0;
}

// Probe: does tcgen05.cp.128x256b from a TMA-SWIZZLE_64B pixel-major tile honour a start address shifted by 1 or 2
// pixels (64 / 128 B)?  And how do cp + TS-mode MMA pipeline?  Build: nvcc -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_bf16.h>
#include "../../etss_07_imageenhance_b200/csrc/tcgen05.cuh"
using namespace b200::tc;

__device__ __forceinline__ uint64_t desc_sw64(uint32_t saddr, uint32_t base_off) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(512 >> 4) << 32) | (1ull << 46) |
           ((uint64_t)(base_off & 7) << 49) | (4ull << 61);
}
__device__ __forceinline__ void tc_cp_128x256b(uint32_t taddr, uint64_t desc) {
    asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(taddr), "l"(desc) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr) : "memory");
}

// mode 0: sw64 pixel-major, base_offset 0; mode 1: base_offset=(addr>>7)&7; mode 2: no-swizzle chunk-planar
__global__ void __launch_bounds__(128) probe(uint32_t* out, int mode) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t sbase = (smem_u32(smem) + 1023u) & ~1023u;
    uint8_t* tile = smem + (sbase - smem_u32(smem));
    // value of (pixel p, channel c) = p*32 + c as bf16 bits (distinct 16-bit ids: p < 130 -> < 4160)
    for (int i = tid; i < 130 * 32; i += 128) {
        const int p = i >> 5, c = i & 31;
        const uint16_t v = (uint16_t)(p * 32 + c);
        int off;
        if (mode < 2) off = p * 64 + (((c >> 3) ^ ((p >> 1) & 3)) << 4) + (c & 7) * 2;
        else off = (c >> 3) * 2080 + p * 16 + (c & 7) * 2;
        *reinterpret_cast<uint16_t*>(tile + off) = v;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 0) {
        if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
        __syncwarp();
        tmem_alloc(&s_tmem, 64);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = s_tmem;
    if (tid == 0) {
        for (int dx = 0; dx < 3; ++dx)
            for (int kc = 0; kc < 2; ++kc) {
                uint64_t d;
                if (mode < 2) {
                    const uint32_t a = sbase + dx * 64 + kc * 32;
                    d = desc_sw64(a, mode == 1 ? (a >> 7) & 7 : 0);
                } else {
                    d = smem_desc_kmajor(sbase + kc * 2 * 2080 + dx * 16, 2080, 128);
                }
                tc_cp_128x256b(tb + (dx * 2 + kc) * 8, d);
            }
        tc_commit(&bar);
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    const uint32_t tl = tb + ((uint32_t)(warp * 32) << 16);
    for (int j = 0; j < 6; ++j) {
        uint32_t v[8];
        tmem_ld8(tl + j * 8, v);
        tc_wait_ld();
        for (int i = 0; i < 8; ++i) out[(j * 128 + tid) * 8 + i] = v[i];
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tb, 64);
}

// timing: per "row": NCP cp's (128x256b) then NMMA TS MMAs (M128 N96 K16) ; reps rows; clock64 around
__global__ void __launch_bounds__(128) pace(long long* out, int ncp, int nmma, int reps) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t s_tmem;
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t sbase = (smem_u32(smem) + 1023u) & ~1023u;
    for (int i = tid; i < 40000 / 4; i += 128) reinterpret_cast<uint32_t*>(smem + (sbase - smem_u32(smem)))[i] = 0;
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
        const uint32_t idesc = idesc_bf16_f32(128, 96);
        const uint64_t bdesc = smem_desc_kmajor(sbase + 16384, 128, 256);
        long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            const uint32_t slot = tb + 192 + (r % 3) * 96;
            for (int j = 0; j < ncp; ++j) tc_cp_128x256b(slot + (j % 12) * 8, desc_sw64(sbase + (j % 3) * 64 + ((j / 3) & 1) * 32, 0));
            for (int j = 0; j < nmma; ++j) mma_ts(tb + (r % 2) * 96, slot + (j % 12) * 8, bdesc + (uint64_t)((j * 3072) >> 4), idesc, 1u);
        }
        tc_commit(&bar);
        mbar_wait(&bar, 0);
        long long t1 = clock64();
        out[0] = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc(tb, 512); }
}

int main() {
    uint32_t* d; cudaMalloc(&d, 6 * 128 * 8 * 4);
    std::vector<uint32_t> h(6 * 128 * 8);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
    for (int mode = 0; mode < 3; ++mode) {
        cudaMemset(d, 0xff, h.size() * 4);
        probe<<<1, 128, 16384>>>(d, mode);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("mode %d: CUDA error %s\n", mode, cudaGetErrorString(e)); return 1; }
        cudaMemcpy(h.data(), d, h.size() * 4, cudaMemcpyDeviceToHost);
        for (int dx = 0; dx < 3; ++dx) for (int kc = 0; kc < 2; ++kc) {
            int bad = 0, first = -1;
            for (int m = 0; m < 128; ++m) for (int i = 0; i < 8; ++i) {
                const int p = m + dx, c = kc * 16 + 2 * i;
                const uint32_t exp = (uint32_t)(p * 32 + c) | ((uint32_t)(p * 32 + c + 1) << 16);
                const uint32_t got = h[((dx * 2 + kc) * 128 + m) * 8 + i];
                if (got != exp) { if (first < 0) first = m * 8 + i; ++bad; }
            }
            printf("mode %d dx %d kc %d: %s (%d bad", mode, dx, kc, bad ? "MISMATCH" : "ok", bad);
            if (bad) { const int m = first / 8, i = first % 8; const uint32_t g = h[((dx * 2 + kc) * 128 + m) * 8 + i];
                printf(", first at lane %d col %d: got px %d ch %d / px %d ch %d", m, i, (g & 0xffff) >> 5, g & 31, (g >> 16) >> 5, (g >> 16) & 31); }
            printf(")\n");
        }
    }
    long long* dt; cudaMalloc(&dt, 8);
    cudaFuncSetAttribute(pace, cudaFuncAttributeMaxDynamicSharedMemorySize, 80000);
    const int cfg[][2] = {{0, 6}, {6, 0}, {6, 6}, {0, 12}, {12, 0}, {12, 12}, {0, 3}, {3, 3}};
    for (auto& c : cfg) {
        long long t = 0;
        for (int it = 0; it < 2; ++it) {
            pace<<<1, 128, 80000>>>(dt, c[0], c[1], 2000);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("pace: CUDA error %s\n", cudaGetErrorString(e)); return 1; }
            cudaMemcpy(&t, dt, 8, cudaMemcpyDeviceToHost);
        }
        printf("pace ncp=%d nmma=%d: %.1f clk/row\n", c[0], c[1], t / 2000.0);
    }
    return 0;
}

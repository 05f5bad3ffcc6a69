// Generic TMA tile-load probe: ./tma_cfg_probe type rank bw bh bp promo swz cx cy cz [w h planes]
//   type 0 = UINT64, 1 = FLOAT32 (8-byte pixels seen as 2 floats), 2 = UINT32 ... ; the tensor is [planes][h][w] pixels of 8 bytes
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include "../../etss_07_imageenhance_b200/csrc/tcgen05.cuh"
#include "../../etss_07_imageenhance_b200/csrc/tma_host.cuh"
using namespace b200;

__global__ void probe(const __grid_constant__ CUtensorMap tm, uint32_t* out, int rank, int c0, int c1, int c2, int bytes) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 65536);
    if (threadIdx.x == 0) {
        tc::mbar_init(bar, 1);
        tc::mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        tc::mbar_arrive_expect_tx(bar, bytes);
        if (rank == 4) tc::tma_load_4d(tc::smem_u32(smem), &tm, c0, c1, c2, 0, bar);
        else tc::tma_load_3d(tc::smem_u32(smem), &tm, c0, c1, c2, bar);
    }
    tc::mbar_wait(bar, 0);
    for (int i = threadIdx.x; i < bytes / 4; i += blockDim.x) out[i] = reinterpret_cast<uint32_t*>(smem)[i];
}

int main(int argc, char** argv) {
    if (argc < 11) return printf("usage\n"), 2;
    const int type = atoi(argv[1]), rank = atoi(argv[2]), bw = atoi(argv[3]), bh = atoi(argv[4]), bp = atoi(argv[5]);
    const int promo = atoi(argv[6]), swz = atoi(argv[7]), cx = atoi(argv[8]), cy = atoi(argv[9]), cz = atoi(argv[10]);
    const int w = argc > 11 ? atoi(argv[11]) : 44, h = argc > 12 ? atoi(argv[12]) : 36, planes = argc > 13 ? atoi(argv[13]) : 32;
    const int es = type == 0 ? 8 : 4, per = 8 / es;  // elements per 8-byte pixel
    std::vector<uint32_t> host((size_t)planes * h * w * 2);
    for (size_t i = 0; i < host.size(); ++i) host[i] = (uint32_t)i + 1;
    uint32_t *d, *o;
    cudaMalloc(&d, host.size() * 4);
    cudaMalloc(&o, 65536);
    cudaMemcpy(d, host.data(), host.size() * 4, cudaMemcpyHostToDevice);
    CUtensorMap m;
    tma_encode_fn enc = tma_encoder();
    const cuuint64_t dims[4] = {(cuuint64_t)w * per, (cuuint64_t)h, (cuuint64_t)planes, 1};
    const cuuint64_t strides[3] = {(cuuint64_t)w * 8, (cuuint64_t)h * w * 8, (cuuint64_t)planes * h * w * 8};
    const cuuint32_t box[4] = {(cuuint32_t)(bw * per), (cuuint32_t)bh, (cuuint32_t)bp, 1}, est[4] = {1, 1, 1, 1};
    const CUtensorMapDataType ty = type == 0 ? CU_TENSOR_MAP_DATA_TYPE_UINT64 : (type == 1 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_UINT32);
    CUresult r = enc(&m, ty, rank, d, dims, strides, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)swz,
                     (CUtensorMapL2promotion)promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("cfg type=%d rank=%d box=%dx%dx%d promo=%d swz=%d at (%d,%d,%d) map %dx%dx%d: encode %d", type, rank, bw, bh, bp, promo, swz, cx, cy, cz, w, h, planes, (int)r);
    if (r != CUDA_SUCCESS) return printf("\n"), 1;
    const int bytes = bw * bh * bp * 8;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536 + 64);
    probe<<<1, 128, 65536 + 64>>>(m, o, rank, cx * per, cy, cz, bytes);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return printf(" -> kernel: %s\n", cudaGetErrorString(e)), 1;
    std::vector<uint32_t> got(bytes / 4);
    cudaMemcpy(got.data(), o, bytes, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int p = 0; p < bp; ++p)
        for (int rr = 0; rr < bh; ++rr)
            for (int cc = 0; cc < bw; ++cc)
                for (int e2 = 0; e2 < 2; ++e2) {
                    const int x = cx + cc, y = cy + rr, pl = cz + p;
                    uint32_t want = 0;
                    if (x >= 0 && x < w && y >= 0 && y < h && pl >= 0 && pl < planes) want = host[(((size_t)pl * h + y) * w + x) * 2 + e2];
                    if (got[((p * bh + rr) * bw + cc) * 2 + e2] != want) ++bad;
                }
    printf(" -> %s (%d mismatches)\n", bad ? "FAIL" : "ok", bad);
    return 0;
}

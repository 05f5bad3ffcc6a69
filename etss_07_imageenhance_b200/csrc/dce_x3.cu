// Zero-DCE conv stack (ZeroDCE_RE.forward, zero_dce.py:173-191) at fp32-class accuracy on the tensor cores:
// precision B200_PREC_BF16X3.  Every product of the 3x3 convolutions is evaluated as the 3-term bf16 split
//     a.w ~= a_hi.w_hi + a_lo.w_hi + a_hi.w_lo        (a = a_hi + a_lo, w = w_hi + w_lo, |err| ~ 2^-16 per product)
// with fp32 accumulation (mma.sync m16n8k16), which meets the path's fp32-mode contract (max-abs <= 1e-3 against the
// fp32 reference) at ~8x the throughput of the CUDA-core fp32 kernels in dce.cu.
//
// Layout.  Feature maps are private to the path, so they are stored ALREADY SPLIT and channel-pair interleaved:
// [B][16 pairs][H][Wp] of uint2 {hi, lo}, each word a bf16x2 of channels (2p, 2p+1) of one pixel (Wp = W rounded up
// to even; 128 B per pixel per map, the same bytes as fp32).  A word is exactly one MMA A-fragment register, so the
// implicit-GEMM operand needs no conversion, no im2col and no register shuffles: for tap (dy, dx) and a 16-channel
// block, the four A registers (hi and lo at once) are four 64-bit shared loads at a shifted address of the halo tile.
// The split is done once per OUTPUT element in the epilogue (64-bit store per pixel and pair).
//
// Schedule.  One launch per layer, one persistent CTA per SM: a producer warp streams 16-channel halo stages
// (8 pairs x 18 rows x 38 columns, one cp.async.bulk.tensor.3d each, zero fill outside the image = the conv's zero
// padding) through a 3-slot mbarrier ring, 16 consumer warps own a 16 x 32 output tile (two m16 tiles of 2 rows x 8
// columns each, all 32 output channels).  The layer's B fragments (weights, hi / lo) stay in shared memory for the
// whole launch.  Layer 1 (3 input channels) runs on the CUDA cores; the curve is `b200_curve_per_iter`'s kernel.
#include <cuda.h>
#include <cuda_bf16.h>
#include <string.h>

#include "common.cuh"
#include "dce_x3.cuh"
#include "tcgen05.cuh"
#include "tma_host.cuh"

namespace b200 {

constexpr int X3_TH = 16, X3_TW = 32;
constexpr int X3_HR = X3_TH + 2, X3_HC = 38;  // halo rows; box columns (x0 - 2 .. x0 + 35: TMA wants a 16-byte aligned start)
constexpr int X3_XOFF = 2;
constexpr int X3_PS = X3_HR * X3_HC;          // uint2 per pair plane of a stage
constexpr int X3_STAGE_BYTES = 8 * X3_PS * 8;
constexpr int X3_NS = 3;
constexpr int X3_CW = 16;                     // consumer warps
constexpr int X3_THREADS = (X3_CW + 1) * 32;
constexpr int X3_FRAMES_PER_PASS = 8;         // workspace = 5 maps x 8 frames (5.3 GB at 1080p)
static_assert(X3_STAGE_BYTES % 128 == 0, "TMA destinations stay 128-byte aligned");
static_assert(X3_PS % 16 == 4 || X3_PS % 16 == 12, "pair-plane stride keeps the (g, c) 64-bit loads conflict free");

// packed block (32-bit words).  layer 0: w[27][32] fp32 (k = ci*9 + tap) | bias[32].
// layers 1..6: B fragments [9 taps][KT][NT][32 lanes][4] = {hi b0, hi b1, lo b0, lo b1} | bias[32]
__host__ __device__ constexpr int x3_kt(int l) { return l <= 3 ? 2 : 4; }
__host__ __device__ constexpr int x3_nt(int l) { return l == 6 ? 3 : 4; }
__host__ __device__ constexpr int x3_words(int l) { return l == 0 ? 27 * 32 + 32 : 9 * x3_kt(l) * x3_nt(l) * 128 + 32; }
__host__ __device__ constexpr int x3_offset(int l) {
    int o = 0;
    for (int i = 0; i < l; ++i) o += x3_words(i);
    return o;
}
constexpr int X3_TOTAL_WORDS = x3_offset(7);

__device__ __forceinline__ uint32_t x3_pack_bf16x2(float lo_elem, float hi_elem) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo_elem, hi_elem);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint2 x3_split2(float v0, float v1) {  // {hi, lo} words of a channel pair
    const uint32_t hi = x3_pack_bf16x2(v0, v1);
    const float h0 = __uint_as_float(hi << 16), h1 = __uint_as_float(hi & 0xffff0000u);
    return make_uint2(hi, x3_pack_bf16x2(v0 - h0, v1 - h1));
}
__device__ __forceinline__ void x3_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

struct X3PackArgs {
    const float* w[7];
    const float* b[7];
};

__global__ void dce_x3_pack_kernel(X3PackArgs a, uint32_t* __restrict__ packed, int cout_last) {
    const int l = blockIdx.x;
    uint32_t* out = packed + x3_offset(l);
    float* f = reinterpret_cast<float*>(out);
    if (l == 0) {
        for (int i = threadIdx.x; i < 27 * 32; i += blockDim.x) {
            const int k = i >> 5, co = i & 31;  // k = ci*9 + tap; e_conv1.weight is [32][3][3][3]
            f[i] = a.w[0][co * 27 + k];
        }
        for (int i = threadIdx.x; i < 32; i += blockDim.x) f[27 * 32 + i] = a.b[0][i];
        return;
    }
    const int kt_n = x3_kt(l), nt_n = x3_nt(l), cin = kt_n * 16, cout = l == 6 ? cout_last : 32;
    const int n_fr = 9 * kt_n * nt_n * 128;
    for (int i = threadIdx.x; i < n_fr; i += blockDim.x) {
        const int j = i & 3, lane = (i >> 2) & 31, tile = i >> 7;
        const int nt = tile % nt_n, kt = (tile / nt_n) % kt_n, tap = tile / (nt_n * kt_n);
        const int g = lane >> 2, c = lane & 3;
        const int n = nt * 8 + g;
        const int k0 = kt * 16 + 2 * c + ((j & 1) ? 8 : 0);
        float w0 = 0.f, w1 = 0.f;
        if (n < cout) {  // OIHW: [cout][cin][3][3]
            w0 = a.w[l][((size_t)n * cin + k0) * 9 + tap];
            w1 = a.w[l][((size_t)n * cin + k0 + 1) * 9 + tap];
        }
        const uint2 s = x3_split2(w0, w1);
        out[i] = j < 2 ? s.x : s.y;
    }
    for (int i = threadIdx.x; i < 32; i += blockDim.x) f[n_fr + i] = i < cout ? a.b[l][i] : 0.f;
}

// ------------------------------------------------------------------------------------------------
// layer 1: e_conv1 (3 -> 32) + ReLU on the CUDA cores, one pixel per thread, output already split
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
dce_x3_first_kernel(const float* __restrict__ x, uint2* __restrict__ dst, const float* __restrict__ wl, int B, int H,
                    int W, int Wp) {
    __shared__ __align__(16) float s_w[27 * 32 + 32];
    for (int i = threadIdx.x; i < 27 * 32 + 32; i += 256) s_w[i] = wl[i];
    __syncthreads();
    const size_t plane = (size_t)H * W;
    const long long total = (long long)B * H * W;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const int xx = (int)(i % W);
        const long long r1 = i / W;
        const int yy = (int)(r1 % H), b = (int)(r1 / H);
        float acc[32];
#pragma unroll
        for (int co = 0; co < 32; ++co) acc[co] = s_w[27 * 32 + co];
        const float* xb = x + (size_t)b * 3 * plane;
#pragma unroll
        for (int ci = 0; ci < 3; ++ci) {
#pragma unroll
            for (int dy = 0; dy < 3; ++dy) {
                const int y2 = yy + dy - 1;
#pragma unroll
                for (int dx = 0; dx < 3; ++dx) {
                    const int x2 = xx + dx - 1;
                    float v = 0.f;
                    if (y2 >= 0 && y2 < H && x2 >= 0 && x2 < W) v = __ldg(xb + ci * plane + (size_t)y2 * W + x2);
                    const float4* wr = reinterpret_cast<const float4*>(s_w + (ci * 9 + dy * 3 + dx) * 32);
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const float4 w4 = wr[q];
                        acc[4 * q] = fmaf(w4.x, v, acc[4 * q]);
                        acc[4 * q + 1] = fmaf(w4.y, v, acc[4 * q + 1]);
                        acc[4 * q + 2] = fmaf(w4.z, v, acc[4 * q + 2]);
                        acc[4 * q + 3] = fmaf(w4.w, v, acc[4 * q + 3]);
                    }
                }
            }
        }
        uint2* dp = dst + ((size_t)b * 16 * H + yy) * Wp + xx;
#pragma unroll
        for (int p = 0; p < 16; ++p)
            dp[(size_t)p * H * Wp] = x3_split2(fmaxf(acc[2 * p], 0.f), fmaxf(acc[2 * p + 1], 0.f));
    }
}

// ------------------------------------------------------------------------------------------------
// layers 2..7
// ------------------------------------------------------------------------------------------------
struct X3Args {
    void* dst;           // !LAST: uint2 [B][16][H][Wp];  LAST: float adjust [B][3*iters][H][W]
    const uint32_t* wl;  // this layer's packed block
    int B, H, W, cout;
};

template <int NSRC, int NT>
constexpr int x3_smem_bytes() {
    return X3_NS * X3_STAGE_BYTES + (9 * 2 * NSRC * NT * 128 + 32) * 4 + 2 * X3_NS * 8;
}

template <int NSRC, int NT, bool LAST>
__global__ void __launch_bounds__(X3_THREADS, 1)
dce_x3_layer_kernel(const X3Args a, const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB) {
    constexpr int KT = 2 * NSRC;
    constexpr int WWORDS = 9 * KT * NT * 128 + 32;
    extern __shared__ __align__(128) uint8_t x3_smem[];
    uint8_t* ring = x3_smem;
    uint32_t* s_w = reinterpret_cast<uint32_t*>(x3_smem + X3_NS * X3_STAGE_BYTES);
    const uint4* s_fr = reinterpret_cast<const uint4*>(s_w);
    const float* s_b = reinterpret_cast<const float*>(s_w + 9 * KT * NT * 128);
    uint64_t* full = reinterpret_cast<uint64_t*>(s_w + WWORDS);
    uint64_t* empty = full + X3_NS;

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int H = a.H, W = a.W;
    const int ntx = (W + X3_TW - 1) / X3_TW, nty = (H + X3_TH - 1) / X3_TH;
    const int tpf = ntx * nty, total = a.B * tpf;
    const int Wp = (W + 1) & ~1;

    if (t == 0) {
        for (int i = 0; i < X3_NS; ++i) {
            tc::mbar_init(&full[i], 1);
            tc::mbar_init(&empty[i], X3_CW);
        }
        tc::mbar_fence_init();
    }
    for (int i = t; i < WWORDS / 4; i += X3_THREADS)
        reinterpret_cast<uint4*>(s_w)[i] = __ldg(reinterpret_cast<const uint4*>(a.wl) + i);
    __syncthreads();

    if (warp == X3_CW) {
        // ================= producer warp =================
        if (lane == 0) {
            tc::tma_prefetch_desc(&tmA);
            if (NSRC == 2) tc::tma_prefetch_desc(&tmB);
        }
        int pk = 0;
        for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
            const int b = tile / tpf, r2 = tile - b * tpf;
            const int tyi = r2 / ntx, txi = r2 - tyi * ntx;
            for (int i = 0; i < KT; ++i, ++pk) {
                const int slot = pk % X3_NS;
                if (pk >= X3_NS) tc::mbar_wait(&empty[slot], ((pk / X3_NS) - 1) & 1);
                if (lane == 0) {
                    tc::mbar_arrive_expect_tx(&full[slot], X3_STAGE_BYTES);
                    tc::tma_load_3d(tc::smem_u32(ring + slot * X3_STAGE_BYTES), i < 2 ? &tmA : &tmB,
                                    txi * X3_TW - X3_XOFF, tyi * X3_TH - 1, b * 16 + (i & 1) * 8, &full[slot]);
                }
                __syncwarp();
            }
        }
        return;
    }

    // ================= consumer warps =================
    const int g = lane >> 2, c = lane & 3;
    const int wy = (warp >> 2) * 4, wx = (warp & 3) * 8;
    int ck = 0;
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int b = tile / tpf, r2 = tile - b * tpf;
        const int tyi = r2 / ntx, txi = r2 - tyi * ntx;
        const int y0 = tyi * X3_TH, x0 = txi * X3_TW;

        // accumulators start at the bias: acc[mt][nt] = {(row g, n 2c), (g, 2c+1), (g+8, 2c), (g+8, 2c+1)}
        float acc[2][NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const float2 pb = *reinterpret_cast<const float2*>(s_b + nt * 8 + 2 * c);
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
                acc[mt][nt][0] = acc[mt][nt][2] = pb.x, acc[mt][nt][1] = acc[mt][nt][3] = pb.y;
        }
#pragma unroll 1
        for (int kt = 0; kt < KT; ++kt) {
            const int slot = ck % X3_NS;
            tc::mbar_wait(&full[slot], (ck / X3_NS) & 1);
            // halo row 0 = image row y0 - 1, halo column 0 = image column x0 - X3_XOFF
            const uint2* base = reinterpret_cast<const uint2*>(ring + slot * X3_STAGE_BYTES) + c * X3_PS + wy * X3_HC +
                                wx + g + (X3_XOFF - 1);
#pragma unroll
            for (int dy = 0; dy < 3; ++dy) {
#pragma unroll
                for (int dx = 0; dx < 3; ++dx) {
                    uint4 bf[NT];  // {hi b0, hi b1, lo b0, lo b1}
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) bf[nt] = s_fr[(((dy * 3 + dx) * KT + kt) * NT + nt) * 32 + lane];
#pragma unroll
                    for (int mt = 0; mt < 2; ++mt) {
                        const uint2* p = base + (2 * mt + dy) * X3_HC + dx;
                        const uint2 u00 = p[0], u10 = p[X3_HC];                          // rows g / g+8, pairs c
                        const uint2 u01 = p[4 * X3_PS], u11 = p[4 * X3_PS + X3_HC];      // pairs c + 4
                        const uint32_t ah[4] = {u00.x, u10.x, u01.x, u11.x};
                        const uint32_t al[4] = {u00.y, u10.y, u01.y, u11.y};
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt) {
                            x3_mma(acc[mt][nt], al, bf[nt].x, bf[nt].y);
                            x3_mma(acc[mt][nt], ah, bf[nt].z, bf[nt].w);
                            x3_mma(acc[mt][nt], ah, bf[nt].x, bf[nt].y);
                        }
                    }
                }
            }
            __syncwarp();  // every lane's shared loads of this stage have returned
            if (lane == 0) tc::mbar_arrive(&empty[slot]);
            ++ck;
        }

        // ---- epilogue: image row y0 + wy + r = m-tile r / 2, half r % 2; column x0 + wx + g ----
        const int xx = x0 + wx + g;
        const int rows = H - (y0 + wy);
        if (xx < W) {
            if (!LAST) {  // ReLU, split, one 64-bit store per (pixel, channel pair nt*4 + c)
                const size_t plane = (size_t)H * Wp;
                uint2* dp = reinterpret_cast<uint2*>(a.dst) + ((size_t)b * 16 + c) * plane + (size_t)(y0 + wy) * Wp + xx;
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    uint2* q = dp + (size_t)(nt * 4) * plane;
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const uint2 u = x3_split2(fmaxf(acc[r >> 1][nt][(r & 1) * 2], 0.f),
                                                  fmaxf(acc[r >> 1][nt][(r & 1) * 2 + 1], 0.f));
                        if (r < rows) q[(size_t)r * Wp] = u;
                    }
                }
            } else {  // x_r = tanh(.): planar [B][cout][H][W], channels nt*8 + 2c, + 1 (zero_dce.py:182)
                const size_t plane = (size_t)H * W;
                float* dp = reinterpret_cast<float*>(a.dst) + (size_t)b * a.cout * plane + (size_t)(y0 + wy) * W + xx;
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int ch = nt * 8 + 2 * c + e;
                        if (ch >= a.cout) continue;
#pragma unroll
                        for (int r = 0; r < 4; ++r)
                            if (r < rows) dp[(size_t)ch * plane + (size_t)r * W] = tanhf(acc[r >> 1][nt][(r & 1) * 2 + e]);
                    }
                }
            }
        }
    }
}

static size_t x3_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
static size_t x3_map_bytes(int frames, int H, int W) {
    return x3_align_up((size_t)frames * 16 * H * ((W + 1) & ~1) * sizeof(uint2), 256);
}

// [frames][16 pairs][H][Wp] uint2 as a 3-D tensor of 8-byte elements (x, y, pair + 16*b); box = one stage
static int x3_make_map(CUtensorMap* m, const void* base, int frames, int H, int W) {
    tma_encode_fn enc = tma_encoder();
    if (!enc) return B200_ERR_LAUNCH;
    const int Wp = (W + 1) & ~1;
    const cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)16 * frames};
    const cuuint64_t strides[2] = {(cuuint64_t)Wp * 8, (cuuint64_t)H * Wp * 8};
    const cuuint32_t box[3] = {X3_HC, X3_HR, 8};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<void*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? B200_OK : B200_ERR_LAUNCH;
}

template <int NSRC, int NT, bool LAST>
static int x3_launch(const X3Args& a, const CUtensorMap& mA, const CUtensorMap& mB, cudaStream_t s) {
    auto kern = dce_x3_layer_kernel<NSRC, NT, LAST>;
    constexpr int smem = x3_smem_bytes<NSRC, NT>();
    static_assert(smem <= 227 * 1024, "shared memory budget");
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;
    }
    int dev = 0, sms = B200_NUM_SMS;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long long total = (long long)a.B * ((a.H + X3_TH - 1) / X3_TH) * ((a.W + X3_TW - 1) / X3_TW);
    const int grid = (int)(total < sms ? total : sms);
    kern<<<grid, X3_THREADS, smem, s>>>(a, mA, mB);
    return launch_status();
}

size_t dce_x3_packed_bytes(int iters) { return iters < 1 || iters > 8 ? 0 : (size_t)X3_TOTAL_WORDS * 4; }

int dce_x3_pack_weights(const float* const* w, const float* const* b, void* packed, int iters, cudaStream_t s) {
    X3PackArgs a;
    for (int l = 0; l < 7; ++l) {
        a.w[l] = w[l];
        a.b[l] = b[l];
    }
    dce_x3_pack_kernel<<<7, 256, 0, s>>>(a, (uint32_t*)packed, 3 * iters);
    return launch_status();
}

// 5 split feature maps for one pass of <= 8 frames + that pass's curve maps when the caller does not want them
size_t dce_x3_workspace_bytes(int B, int H, int W) {
    const int fb = B < X3_FRAMES_PER_PASS ? B : X3_FRAMES_PER_PASS;
    return 5 * x3_map_bytes(fb, H, W) + x3_align_up((size_t)fb * 24 * H * W * sizeof(float), 256);
}

int dce_x3_forward(const float* x, const void* packed, float* adjust, float* enhanced, int B, int H, int W, int iters,
                   void* workspace, size_t workspace_bytes, cudaStream_t s) {
    if (workspace_bytes < dce_x3_workspace_bytes(B, H, W)) return B200_ERR_WORKSPACE;
    const int fb = B < X3_FRAMES_PER_PASS ? B : X3_FRAMES_PER_PASS;
    const size_t map = x3_map_bytes(fb, H, W);
    uint8_t* ws = (uint8_t*)workspace;
    void* xm[5];
    for (int i = 0; i < 5; ++i) xm[i] = ws + i * map;  // x1 x2 x3 x4 x5 (x6 reuses x4's buffer)
    float* a_ws = (float*)(ws + 5 * map);
    const uint32_t* pk = (const uint32_t*)packed;
    const size_t plane = (size_t)H * W;
    const int cout7 = 3 * iters, Wp = (W + 1) & ~1;
    for (int b0 = 0; b0 < B; b0 += fb) {
        const int nb = B - b0 < fb ? B - b0 : fb;
        const float* xb = x + (size_t)b0 * 3 * plane;
        float* ab = adjust ? adjust + (size_t)b0 * cout7 * plane : a_ws;
        CUtensorMap m[5];
        memset(m, 0, sizeof(m));
        int rc;
        for (int i = 0; i < 5; ++i)
            if ((rc = x3_make_map(&m[i], xm[i], nb, H, W))) return rc;
        {
            long long blocks = ((long long)nb * H * W + 255) / 256;
            if (blocks > B200_NUM_SMS * 16) blocks = B200_NUM_SMS * 16;
            dce_x3_first_kernel<<<(unsigned)blocks, 256, 0, s>>>(xb, (uint2*)xm[0], (const float*)(pk + x3_offset(0)),
                                                                 nb, H, W, Wp);
            if ((rc = launch_status())) return rc;
        }
        X3Args a{};
        a.B = nb;
        a.H = H;
        a.W = W;
        a.cout = 32;
        auto layer = [&](int l, void* dst) {
            a.dst = dst;
            a.wl = pk + x3_offset(l);
        };
        // zero_dce.py:176-182: x2 x3 x4 | cat(x3,x4) -> x5 | cat(x2,x5) -> x6 | cat(x1,x6) -> x_r
        layer(1, xm[1]);
        if ((rc = x3_launch<1, 4, false>(a, m[0], m[0], s))) return rc;
        layer(2, xm[2]);
        if ((rc = x3_launch<1, 4, false>(a, m[1], m[1], s))) return rc;
        layer(3, xm[3]);
        if ((rc = x3_launch<1, 4, false>(a, m[2], m[2], s))) return rc;
        layer(4, xm[4]);
        if ((rc = x3_launch<2, 4, false>(a, m[2], m[3], s))) return rc;
        layer(5, xm[3]);  // x6 -> x4's buffer (x4 is dead once layer 5 has run)
        if ((rc = x3_launch<2, 4, false>(a, m[1], m[4], s))) return rc;
        layer(6, ab);
        a.cout = cout7;
        if ((rc = x3_launch<2, 3, true>(a, m[0], m[3], s))) return rc;
        if ((rc = curve_per_iter_unaligned_ok(xb, ab, enhanced + (size_t)b0 * 3 * plane, nb, 3, H, W, iters, s)))
            return rc;
    }
    return B200_OK;
}

}  // namespace b200

// Zero-DCE conv stack (ZeroDCE_RE.forward, zero_dce.py:173-191) on the sm_100a tensor cores.
//
// One launch per layer.  Per layer the 3x3 convolution is an implicit GEMM
//     D[M = 128 pixels of one image row, N = 32] += A_tap[M, K = Cin] * W_tap[K, N]      (9 taps)
// issued as tcgen05.mma.cta_group::1.kind::f16 (bf16 x bf16 -> fp32 in TMEM), M=128 N=32 K=16, in **TS mode**:
// the activation operand A lives in TMEM (lane = pixel, one 32-bit column = two bf16 channels), only the 1 KB
// weight tile B comes from shared memory.  With N = 32 an SS-mode MMA would need 256 B/clk of shared-memory reads
// for A — twice what the SM has — so the A rows are staged once per input row instead:
//     global NHWC bf16 --cp.async--> smem (80 B pixel pitch, bank-conflict-free) --LDS.128--> registers
//       --tcgen05.st--> three lane-shifted copies (dx = -1, 0, +1) in a 4-slot TMEM ring of input rows.
// dy shifts are ring slots, so rows stream top to bottom with no recompute inside a strip.
//
// Warp roles (288 threads, 1 persistent CTA per SM, all 512 TMEM columns):
//   warps 0-3  loader: stage + fill TMEM ring slot, arrive on full[slot]
//   warp  4    lane 0 issues the 9 * Cin/16 MMAs per output row, tcgen05.commit -> acc_full[buf], empty[slot]
//   warps 5-8  epilogue: tcgen05.ld accumulator row -> bias + ReLU -> bf16 NHWC store; layer 7: tanh -> `adjust`
//              (NCHW fp32) and the 8-iteration curve on x -> `enhanced`   (zero_dce.py:182-188)
// Layer 1 (Cin = 3) uses the same pipeline with an im2col loader: K = 27 (pad 32), one "tap", 2 MMAs per row.
#include <cuda_bf16.h>

#include "common.cuh"
#include "dce_tc.cuh"
#include "tcgen05.cuh"

namespace b200 {
using namespace tc;

constexpr int TC_M = 128;        // pixels per strip = MMA M = TMEM lanes
constexpr int TC_N = 32;         // output channels (24 padded to 32 in layer 7)
constexpr int TC_NSLOT = 4;      // TMEM ring slots (input rows)
constexpr int TC_NSTAGE = 8;     // smem staging buffers (input rows)
constexpr int TC_PF = 6;         // cp.async prefetch distance (rows); NSTAGE >= PF + 2
constexpr int TC_NSTAGE1 = 12;   // layer 1: staged fp32 input rows, each feeds 3 output rows; NSTAGE1 >= PF1 + 4
constexpr int TC_PF1 = 8;
constexpr int TC_ROW1 = 3 * 132; // layer 1: floats per staged row ([3 channels][132])
constexpr int TC_NACC = 4;       // accumulator buffers
constexpr int TC_PXP = 80;       // staging pixel pitch in bytes (64 B of channels + 16 B pad)
constexpr int TC_SPX = TC_M + 2; // staged pixels per row (halo of 1 on each side)
constexpr int TC_TILE = 1024;    // bytes of one B tile (N=32 x K=16 bf16)
constexpr int TC_THREADS = 288;
constexpr int TC_ACC_COLS = 128; // 4 accumulator buffers x 32 columns

__host__ __device__ constexpr int tc_ntap(int l) { return l == 0 ? 1 : 9; }
__host__ __device__ constexpr int tc_kc(int l) { return l == 0 ? 2 : (l <= 3 ? 2 : 4); }
__host__ __device__ constexpr int tc_wbytes(int l) { return tc_ntap(l) * tc_kc(l) * TC_TILE; }
__host__ __device__ constexpr int tc_layer_bytes(int l) { return tc_wbytes(l) + 128; }  // + 32 fp32 biases
__host__ __device__ constexpr int tc_layer_offset(int l) {
    int o = 0;
    for (int i = 0; i < l; ++i) o += tc_layer_bytes(i);
    return o;
}
constexpr int TC_PACKED_BYTES = tc_layer_offset(7);

// ------------------------------------------------------------------------------------------------
// weight packing: OIHW fp32 -> per (tap, k-chunk) UMMA K-major core-matrix tiles, bf16
//   element (n, kk) of a tile at byte (n/8)*256 + (kk/8)*128 + (n%8)*16 + (kk%8)*2      (LBO = 128, SBO = 256)
// ------------------------------------------------------------------------------------------------
struct TcPackArgs {
    const float* w[7];
    const float* b[7];
};

__global__ void dce_tc_pack_kernel(TcPackArgs a, uint8_t* __restrict__ packed, int cout_last) {
    const int l = blockIdx.x;
    const int ntap = tc_ntap(l), kc_n = tc_kc(l);
    const int cin = l == 0 ? 3 : (l <= 3 ? 32 : 64);
    const int cout = l == 6 ? cout_last : 32;
    uint8_t* base = packed + tc_layer_offset(l);
    __nv_bfloat16* tiles = reinterpret_cast<__nv_bfloat16*>(base);
    const int n_el = ntap * kc_n * 512;
    for (int i = threadIdx.x; i < n_el; i += blockDim.x) {
        const int tile = i >> 9, e = i & 511;
        const int tap = tile / kc_n, kc = tile - tap * kc_n;
        // e enumerates (n, kk) row-major: n = e / 16, kk = e % 16
        const int n = e >> 4, kk = e & 15;
        float v = 0.f;
        if (n < cout) {
            if (l == 0) {
                const int k = kc * 16 + kk;  // k = (ky*3+kx)*3 + c
                if (k < 27) {
                    const int tp = k / 3, c = k - tp * 3;
                    v = a.w[0][(n * 3 + c) * 9 + tp];
                }
            } else {
                const int ci = kc * 16 + kk;
                v = a.w[l][(n * cin + ci) * 9 + tap];
            }
        }
        const int off = (n >> 3) * 128 + (kk >> 3) * 64 + (n & 7) * 8 + (kk & 7);  // in bf16 elements
        tiles[tile * 512 + off] = __float2bfloat16_rn(v);
    }
    float* bias = reinterpret_cast<float*>(base + tc_wbytes(l));
    for (int i = threadIdx.x; i < 32; i += blockDim.x) bias[i] = i < cout ? a.b[l][i] : 0.f;
}

// ------------------------------------------------------------------------------------------------
struct TcArgs {
    const void* src0;       // NHWC bf16 [B,H,W,32]   (FIRST: x, NCHW fp32)
    const void* src1;       // second concat source or nullptr
    __nv_bfloat16* dst;     // NHWC bf16 [B,H,W,32]   (!LAST)
    const float* x;         // LAST: network input (NCHW fp32) for the curve
    float* adjust;          // LAST: [B,3*iters,H,W] or nullptr
    float* enhanced;        // LAST: [B,3,H,W]
    const uint8_t* wpack;   // this layer's packed block
    int B, H, W, R, nseg, nstrip, iters;
};

template <int NSRC, bool FIRST>
constexpr int tc_smem_bytes() {
    return (FIRST ? 2 : 9 * 2 * NSRC) * TC_TILE + 128 + 128 + 16 +
           (FIRST ? TC_NSTAGE1 * TC_ROW1 * 4 : TC_NSTAGE * NSRC * TC_SPX * TC_PXP);
}

template <int NSRC, bool FIRST, bool LAST>
__global__ void __launch_bounds__(TC_THREADS, 1) dce_tc_layer_kernel(const TcArgs args) {
    constexpr int KC = FIRST ? 2 : 2 * NSRC;        // K chunks of 16 per tap
    constexpr int NTAP = FIRST ? 1 : 9;
    constexpr int WBYTES = NTAP * KC * TC_TILE;
    constexpr int COPY_COLS = FIRST ? 16 : 16 * NSRC;  // TMEM columns of one dx copy
    constexpr int SLOT_COLS = FIRST ? 16 : 3 * COPY_COLS;
    constexpr int ROWS = FIRST ? 1 : 3;             // input rows per output row
    constexpr int SRC_STAGE = TC_SPX * TC_PXP;      // bytes per source per staged row
    static_assert(TC_ACC_COLS + TC_NSLOT * SLOT_COLS <= 512, "TMEM budget");

    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* s_w = smem;
    float* s_bias = reinterpret_cast<float*>(smem + WBYTES);
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(smem + WBYTES + 128);
    uint64_t* full = s_bar;                   // [NSLOT]  loader -> MMA      (128 arrivals)
    uint64_t* empty = s_bar + TC_NSLOT;       // [NSLOT]  MMA commit -> loader (1 arrival)
    uint64_t* acc_full = s_bar + 2 * TC_NSLOT;              // [NACC] MMA commit -> epilogue (1)
    uint64_t* acc_empty = s_bar + 2 * TC_NSLOT + TC_NACC;   // [NACC] epilogue -> MMA (128)
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + WBYTES + 256);
    uint8_t* s_stage = smem + WBYTES + 256 + 16;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int B = args.B, H = args.H, W = args.W, R = args.R;
    const int n_items = B * args.nseg * args.nstrip;

    // ---- one-time setup ----
    {
        const uint4* g = reinterpret_cast<const uint4*>(args.wpack);
        uint4* s = reinterpret_cast<uint4*>(smem);
        for (int i = tid; i < (WBYTES + 128) / 16; i += TC_THREADS) s[i] = __ldg(g + i);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // weights are read by the tensor core (async proxy)
    }
    if (warp == 4) {
        if (lane == 0) {
            for (int i = 0; i < TC_NSLOT; ++i) {
                mbar_init(&full[i], 128);
                mbar_init(&empty[i], 1);
            }
            for (int i = 0; i < TC_NACC; ++i) {
                mbar_init(&acc_full[i], 1);
                mbar_init(&acc_empty[i], 128);
            }
            mbar_fence_init();
        }
        __syncwarp();
        tmem_alloc(s_tmem, 512);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = *s_tmem;

    if (warp < 4) {
        // =========================== LOADER ===========================
        const int t = tid;  // 0..127 = pixel in strip = TMEM lane
        const uint32_t tlane = tbase + ((uint32_t)(warp * 32) << 16);
        uint32_t fill = 0, g = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int strip = item % args.nstrip;
            const int seg = (item / args.nstrip) % args.nseg;
            const int b = item / (args.nstrip * args.nseg);
            const int y0 = seg * R, x0 = strip * TC_M;
            const int nout = min(R, H - y0);
            const int nrows = nout + ROWS - 1;
            if constexpr (FIRST) {
                // layer 1: fp32 NCHW input rows staged once ([3][132] floats per row), each feeds 3 output rows;
                // the im2col row (27 taps x channels, padded to K = 32) is gathered from shared memory.
                const float* xin = reinterpret_cast<const float*>(args.src0) + (size_t)b * 3 * H * W;
                float* stg = reinterpret_cast<float*>(s_stage);
                const int nin = nout + 2;
                auto issue_row1 = [&](int y, uint32_t stage) {
                    const bool yok = y >= 0 && y < H;
                    const uint32_t sdst = smem_u32(stg) + stage * (TC_ROW1 * 4);
                    for (int i = t; i < 3 * TC_SPX; i += 128) {
                        const int c = i / TC_SPX, px = i - c * TC_SPX;
                        const int xs = x0 - 1 + px;
                        const bool ok = yok && xs >= 0 && xs < W;
                        const float* gp = xin + (ok ? ((size_t)c * H + y) * W + xs : (size_t)0);
                        cp_async4(sdst + (c * 132 + px) * 4, gp, ok);
                    }
                };
#pragma unroll
                for (int k = 0; k < TC_PF1; ++k) {
                    if (k < nin) issue_row1(y0 - 1 + k, (g + k) % TC_NSTAGE1);
                    cp_async_commit();
                }
                for (int ri = 0; ri < nin; ++ri) {
                    if (ri + TC_PF1 < nin) issue_row1(y0 - 1 + ri + TC_PF1, (g + ri + TC_PF1) % TC_NSTAGE1);
                    cp_async_commit();
                    cp_async_wait<TC_PF1>();
                    named_bar_sync(1, 128);
                    if (ri < 2) continue;
                    float v[28];
#pragma unroll
                    for (int ky = 0; ky < 3; ++ky) {
                        const float* sp = stg + ((g + ri - 2 + ky) % TC_NSTAGE1) * TC_ROW1 + t;
#pragma unroll
                        for (int kx = 0; kx < 3; ++kx)
#pragma unroll
                            for (int c = 0; c < 3; ++c) v[(ky * 3 + kx) * 3 + c] = sp[c * 132 + kx];
                    }
                    v[27] = 0.f;
                    uint32_t pk[16];
#pragma unroll
                    for (int j = 0; j < 14; ++j) {
                        __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
                        pk[j] = *reinterpret_cast<uint32_t*>(&h);
                    }
                    pk[14] = pk[15] = 0u;
                    const uint32_t slot = fill % TC_NSLOT, use = fill / TC_NSLOT;
                    if (use > 0) mbar_wait(&empty[slot], (use - 1) & 1);
                    tc_fence_after();
                    tmem_st16(tlane + TC_ACC_COLS + slot * SLOT_COLS, pk);
                    tc_wait_st();
                    tc_fence_before();
                    mbar_arrive(&full[slot]);
                    ++fill;
                }
                g += nin;
            } else {
                const uint8_t* srcs[2] = {reinterpret_cast<const uint8_t*>(args.src0),
                                          reinterpret_cast<const uint8_t*>(args.src1)};
                auto issue_row = [&](int y, uint32_t stage) {
                    const bool yok = y >= 0 && y < H;
                    const uint32_t sdst = smem_u32(s_stage) + stage * (NSRC * SRC_STAGE);
                    for (int i = t; i < NSRC * TC_SPX * 4; i += 128) {
                        const int s = i / (TC_SPX * 4);
                        const int rem = i - s * (TC_SPX * 4);
                        const int px = rem >> 2, ch = rem & 3;
                        const int xs = x0 - 1 + px;
                        const bool ok = yok && xs >= 0 && xs < W;
                        const uint8_t* gp =
                            srcs[s] + (ok ? (((size_t)b * H + y) * W + xs) * 64 + ch * 16 : (size_t)0);
                        cp_async16(sdst + s * SRC_STAGE + px * TC_PXP + ch * 16, gp, ok);
                    }
                };
#pragma unroll
                for (int k = 0; k < TC_PF; ++k) {
                    if (k < nrows) issue_row(y0 - 1 + k, (g + k) % TC_NSTAGE);
                    cp_async_commit();
                }
                for (int r = 0; r < nrows; ++r) {
                    if (r + TC_PF < nrows) issue_row(y0 - 1 + r + TC_PF, (g + r + TC_PF) % TC_NSTAGE);
                    cp_async_commit();
                    cp_async_wait<TC_PF>();
                    named_bar_sync(1, 128);
                    const uint32_t slot = fill % TC_NSLOT, use = fill / TC_NSLOT;
                    if (use > 0) mbar_wait(&empty[slot], (use - 1) & 1);
                    tc_fence_after();
                    const uint8_t* st = s_stage + ((g + r) % TC_NSTAGE) * (NSRC * SRC_STAGE);
#pragma unroll
                    for (int s = 0; s < NSRC; ++s) {
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx) {
                            const uint4* p = reinterpret_cast<const uint4*>(st + s * SRC_STAGE + (t + dx) * TC_PXP);
                            const uint4 q0 = p[0], q1 = p[1], q2 = p[2], q3 = p[3];
                            const uint32_t pk[16] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w,
                                                     q2.x, q2.y, q2.z, q2.w, q3.x, q3.y, q3.z, q3.w};
                            tmem_st16(tlane + TC_ACC_COLS + slot * SLOT_COLS + dx * COPY_COLS + s * 16, pk);
                        }
                    }
                    tc_wait_st();
                    tc_fence_before();
                    mbar_arrive(&full[slot]);
                    ++fill;
                }
                g += nrows;
            }
        }
    } else if (warp == 4) {
        // =========================== MMA ISSUER (one thread) ===========================
        if (lane == 0) {
            constexpr uint32_t IDESC = idesc_bf16_f32(TC_M, TC_N);
            const uint64_t bdesc0 = smem_desc_kmajor(smem_u32(s_w), 128, 256);
            uint32_t f0 = 0, o = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                const int seg = (item / args.nstrip) % args.nseg;
                const int nout = min(R, H - seg * R);
                for (int j = 0; j < nout; ++j) {
                    if (j == 0) {
                        for (int dy = 0; dy < ROWS - 1; ++dy) {
                            const uint32_t f = f0 + dy;
                            mbar_wait(&full[f % TC_NSLOT], (f / TC_NSLOT) & 1);
                        }
                    }
                    {
                        const uint32_t f = f0 + j + ROWS - 1;
                        mbar_wait(&full[f % TC_NSLOT], (f / TC_NSLOT) & 1);
                    }
                    const uint32_t buf = o % TC_NACC;
                    if (o >= TC_NACC) mbar_wait(&acc_empty[buf], ((o / TC_NACC) - 1) & 1);
                    tc_fence_after();
                    const uint32_t d = tbase + buf * TC_N;
#pragma unroll
                    for (int dy = 0; dy < ROWS; ++dy) {
                        const uint32_t slot = (f0 + j + dy) % TC_NSLOT;
                        const uint32_t abase = tbase + TC_ACC_COLS + slot * SLOT_COLS;
#pragma unroll
                        for (int dx = 0; dx < (FIRST ? 1 : 3); ++dx) {
#pragma unroll
                            for (int kc = 0; kc < KC; ++kc) {
                                const int tile = (dy * 3 + dx) * KC + kc;
                                mma_ts(d, abase + dx * COPY_COLS + kc * 8, bdesc0 + (uint64_t)(tile * (TC_TILE >> 4)),
                                       IDESC, (dy | dx | kc) != 0);
                            }
                        }
                    }
                    tc_commit(&acc_full[buf]);
                    tc_commit(&empty[(f0 + j) % TC_NSLOT]);
                    ++o;
                }
                for (int dy = 1; dy < ROWS; ++dy) tc_commit(&empty[(f0 + nout - 1 + dy) % TC_NSLOT]);
                f0 += nout + ROWS - 1;
            }
        }
    } else {
        // =========================== EPILOGUE ===========================
        const int q = warp & 3;
        const int m = q * 32 + lane;
        const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
        float bias[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) bias[i] = s_bias[i];
        uint32_t o = 0;
        const size_t plane = (size_t)H * W;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int strip = item % args.nstrip;
            const int seg = (item / args.nstrip) % args.nseg;
            const int b = item / (args.nstrip * args.nseg);
            const int y0 = seg * R, xx = strip * TC_M + m;
            const int nout = min(R, H - y0);
            for (int j = 0; j < nout; ++j) {
                const uint32_t buf = o % TC_NACC;
                mbar_wait(&acc_full[buf], (o / TC_NACC) & 1);
                tc_fence_after();
                uint32_t v[32];
                tmem_ld32(tlane + buf * TC_N, v);
                tc_wait_ld();
                tc_fence_before();
                mbar_arrive(&acc_empty[buf]);
                ++o;
                const int y = y0 + j;
                if (xx >= W) continue;
                if constexpr (!LAST) {
                    uint32_t pk[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float a0 = fmaxf(__uint_as_float(v[2 * i]) + bias[2 * i], 0.f);
                        const float a1 = fmaxf(__uint_as_float(v[2 * i + 1]) + bias[2 * i + 1], 0.f);
                        __nv_bfloat162 h = __floats2bfloat162_rn(a0, a1);
                        pk[i] = *reinterpret_cast<uint32_t*>(&h);
                    }
                    uint4* dp = reinterpret_cast<uint4*>(args.dst + (((size_t)b * H + y) * W + xx) * 32);
                    dp[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    dp[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                    dp[2] = make_uint4(pk[8], pk[9], pk[10], pk[11]);
                    dp[3] = make_uint4(pk[12], pk[13], pk[14], pk[15]);
                } else {
                    const int iters = args.iters;
                    const size_t pix = (size_t)y * W + xx;
                    float a[24];
#pragma unroll
                    for (int c = 0; c < 24; ++c) a[c] = tanh_approx(__uint_as_float(v[c]) + bias[c]);
                    if (args.adjust != nullptr) {
                        float* ap = args.adjust + (size_t)b * 3 * iters * plane + pix;
#pragma unroll
                        for (int c = 0; c < 24; ++c)
                            if (c < 3 * iters) ap[(size_t)c * plane] = a[c];
                    }
                    const float* xp = args.x + (size_t)b * 3 * plane + pix;
                    float* ep = args.enhanced + (size_t)b * 3 * plane + pix;
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        float yv = __ldg(xp + (size_t)c * plane);
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            if (i < iters) yv = curve_step(yv, a[i * 3 + c]);
                        ep[(size_t)c * plane] = yv;
                    }
                }
            }
        }
    }

    // ---- teardown ----
    tc_fence_before();
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        tmem_dealloc(tbase, 512);
    }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static size_t tc_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

template <int NSRC, bool FIRST, bool LAST>
static int tc_launch(const TcArgs& a, cudaStream_t s) {
    auto kern = dce_tc_layer_kernel<NSRC, FIRST, LAST>;
    constexpr int smem = tc_smem_bytes<NSRC, FIRST>();
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;
    }
    const long long items = (long long)a.B * a.nseg * a.nstrip;
    const int grid = (int)(items < B200_NUM_SMS ? items : B200_NUM_SMS);
    kern<<<grid, TC_THREADS, smem, s>>>(a);
    return launch_status();
}

size_t dce_tc_packed_bytes(int precision, int iters) {
    if (precision != B200_PREC_BF16 || iters < 1 || iters > 8) return 0;
    return (size_t)TC_PACKED_BYTES;
}

int dce_tc_pack_weights(const float* const* w, const float* const* b, void* packed, int precision, int iters,
                        cudaStream_t s) {
    if (precision != B200_PREC_BF16) return B200_ERR_UNSUPPORTED;
    TcPackArgs a;
    for (int l = 0; l < 7; ++l) {
        a.w[l] = w[l];
        a.b[l] = b[l];
    }
    dce_tc_pack_kernel<<<7, 256, 0, s>>>(a, (uint8_t*)packed, 3 * iters);
    return launch_status();
}

// 5 NHWC bf16 feature maps [B,H,W,32] (x4's buffer is reused for x6)
size_t dce_tc_workspace_bytes(int B, int H, int W, int precision) {
    if (precision != B200_PREC_BF16) return 0;
    return 5 * tc_align_up((size_t)B * H * W * 64, 256);
}

int dce_tc_forward(const float* x, const void* packed, float* adjust, float* enhanced, int B, int H, int W, int iters,
                   int precision, void* workspace, size_t workspace_bytes, cudaStream_t s) {
    if (precision != B200_PREC_BF16) return B200_ERR_UNSUPPORTED;
    const size_t fmap = tc_align_up((size_t)B * H * W * 64, 256);
    if (workspace_bytes < 5 * fmap) return B200_ERR_WORKSPACE;
    uint8_t* ws = (uint8_t*)workspace;
    __nv_bfloat16* x1 = (__nv_bfloat16*)(ws + 0 * fmap);
    __nv_bfloat16* x2 = (__nv_bfloat16*)(ws + 1 * fmap);
    __nv_bfloat16* x3 = (__nv_bfloat16*)(ws + 2 * fmap);
    __nv_bfloat16* x4 = (__nv_bfloat16*)(ws + 3 * fmap);
    __nv_bfloat16* x5 = (__nv_bfloat16*)(ws + 4 * fmap);
    __nv_bfloat16* x6 = x4;
    const uint8_t* pk = (const uint8_t*)packed;

    TcArgs a{};
    a.B = B;
    a.H = H;
    a.W = W;
    a.iters = iters;
    a.R = H < 40 ? H : 40;
    a.nseg = (H + a.R - 1) / a.R;
    a.nstrip = (W + TC_M - 1) / TC_M;
    int rc;
    auto layer = [&](int l, const void* s0, const void* s1, __nv_bfloat16* dst) { 
        a.src0 = s0;
        a.src1 = s1;
        a.dst = dst;
        a.wpack = pk + tc_layer_offset(l);
    };
    layer(0, x, nullptr, x1);
    if ((rc = tc_launch<1, true, false>(a, s))) return rc;
    layer(1, x1, nullptr, x2);
    if ((rc = tc_launch<1, false, false>(a, s))) return rc;
    layer(2, x2, nullptr, x3);
    if ((rc = tc_launch<1, false, false>(a, s))) return rc;
    layer(3, x3, nullptr, x4);
    if ((rc = tc_launch<1, false, false>(a, s))) return rc;
    layer(4, x3, x4, x5);
    if ((rc = tc_launch<2, false, false>(a, s))) return rc;
    layer(5, x2, x5, x6);
    if ((rc = tc_launch<2, false, false>(a, s))) return rc;
    layer(6, x1, x6, nullptr);
    a.x = x;
    a.adjust = adjust;
    a.enhanced = enhanced;
    if ((rc = tc_launch<2, false, true>(a, s))) return rc;
    return B200_OK;
}

}  // namespace b200

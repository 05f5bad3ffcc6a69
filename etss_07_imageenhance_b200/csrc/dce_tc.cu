// Zero-DCE conv stack (ZeroDCE_RE.forward, zero_dce.py:173-191) on the sm_100a tensor cores.
//
// One launch per layer.  Per layer the 3x3 convolution is an implicit GEMM in "input-row-stationary" form:
// an input row r (128 pixels of one strip, all Cin channels) contributes to the three output rows r+1, r, r-1
// through the filter rows dy = 0, 1, 2, so for every (dx, 16-channel chunk) ONE
//     tcgen05.mma.cta_group::1.kind::f16   M = 128 pixels, N = 96 = 3 output rows x 32 couts, K = 16
// accumulates into three ADJACENT 32-column fp32 accumulator blocks of a TMEM ring (bf16 x bf16 -> fp32).
// The MMA runs in **TS mode**: the activation operand A lives in TMEM (lane = pixel, one 32-bit column = two
// bf16 channels); only the 3 KB weight tile B is read from shared memory.  With N this small an SS-mode MMA would
// need more shared-memory bandwidth for A than an SM has.  A 3x3 tap's dx shift cannot be a TMEM lane shift, so the
// loader warps write three lane-shifted copies (dx = -1, 0, +1) of each input row:
//     global NHWC bf16 --TMA (4-D tile, 64B swizzle, zero fill outside the image)--> smem ring
//       --LDS.128 (conflict-free through the swizzle)--> registers --tcgen05.st--> TMEM A slot (3 copies).
// Rows stream top to bottom inside a strip: no halo recompute in y, one input row lives in TMEM for one MMA batch.
// An accumulator block is complete after the input row below it has been consumed; the epilogue drains it,
// re-zeroes it (every MMA accumulates) and hands it back.
//
// Warp roles (576 threads, 1 persistent CTA per SM, 512 TMEM columns):
//   warps 0-7    loaders, two per TMEM lane quarter, alternating input rows
//   warp  8      MMA issuer: warp-uniform control flow, one elected lane issues tcgen05.mma / tcgen05.commit
//                (-> empty[slot], acc_full[block])
//   warp  9      TMA producer: one elected lane, expect_tx + cp.async.bulk.tensor per input row and source
//   warps 10-17  epilogue, two per lane quarter, alternating output rows: tcgen05.ld -> bias + ReLU -> bf16 NHWC
//                store; layer 7: tanh -> `adjust` (NCHW fp32) + the 8-iteration curve on x -> `enhanced`
// Layer 1 (Cin = 3) runs the same pipeline with K padded to 16 (one chunk): its loaders read the fp32 NCHW input
// directly (coalesced along x, one row ahead) and rewrite only the two non-zero TMEM columns of each dx copy.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "dce_tc.cuh"
#include "tcgen05.cuh"

namespace b200 {
using namespace tc;

constexpr int TC_M = 128;            // pixels per strip = MMA M = TMEM lanes
constexpr int TC_NSLOT = 3;          // TMEM A slots (input rows)
constexpr int TC_NB = 6;             // accumulator blocks (32 columns each) in the TMEM ring; even
constexpr int TC_ACC_COLS = TC_NB * 32;
constexpr int TC_NST = 8;            // TMA smem stages (input rows); even
constexpr int TC_SPX = TC_M + 2;     // staged pixels per row (halo of 1 on each side)
constexpr int TC_SRC_BYTES = TC_SPX * 64;          // TMA box bytes per source per row (8320)
constexpr int TC_SRC_STRIDE = 8704;                // smem bytes per source per stage (17 x 512: keeps the swizzle phase)
constexpr int TC_THREADS = 576;
static_assert((2 * TC_NST + 2 * TC_NSLOT + 2 * TC_NB) * 8 <= 384, "mbarrier area");

// packed weights per layer: B tiles then 32 fp32 biases.
//   per (dx, kc) one tile [N=96 = dy*32 + cout][K=16] (3 KB); layer 0 has one K chunk (3 channels, zero padded)
// element (n, kk) of a tile at byte (n/8)*256 + (kk/8)*128 + (n%8)*16 + (kk%8)*2   (K-major core matrices,
// LBO = 128, SBO = 256)
__host__ __device__ constexpr int tc_kc(int l) { return l == 0 ? 1 : (l <= 3 ? 2 : 4); }
__host__ __device__ constexpr int tc_wbytes(int l) { return 3 * tc_kc(l) * 3072; }
__host__ __device__ constexpr int tc_layer_bytes(int l) { return tc_wbytes(l) + 128; }
__host__ __device__ constexpr int tc_layer_offset(int l) {
    int o = 0;
    for (int i = 0; i < l; ++i) o += tc_layer_bytes(i);
    return o;
}
constexpr int TC_PACKED_BYTES = tc_layer_offset(7);

struct TcPackArgs {
    const float* w[7];
    const float* b[7];
};

__global__ void dce_tc_pack_kernel(TcPackArgs a, uint8_t* __restrict__ packed, int cout_last) {
    const int l = blockIdx.x;
    const int kc_n = tc_kc(l);
    const int cin = l == 0 ? 3 : (l <= 3 ? 32 : 64);
    const int cout = l == 6 ? cout_last : 32;
    uint8_t* base = packed + tc_layer_offset(l);
    __nv_bfloat16* tiles = reinterpret_cast<__nv_bfloat16*>(base);
    const int n_el = 3 * kc_n * 1536;
    for (int i = threadIdx.x; i < n_el; i += blockDim.x) {
        const int tile = i / 1536, e = i - tile * 1536;
        const int dx = tile / kc_n, kc = tile - dx * kc_n;
        const int n = e >> 4, kk = e & 15;  // n = dy*32 + cout
        const int dy = n >> 5, oc = n & 31;
        const int ci = kc * 16 + kk;
        const float v = (oc < cout && ci < cin) ? a.w[l][(oc * cin + ci) * 9 + dy * 3 + dx] : 0.f;
        tiles[tile * 1536 + (n >> 3) * 128 + (kk >> 3) * 64 + (n & 7) * 8 + (kk & 7)] = __float2bfloat16_rn(v);
    }
    float* bias = reinterpret_cast<float*>(base + tc_wbytes(l));
    for (int i = threadIdx.x; i < 32; i += blockDim.x) bias[i] = i < cout ? a.b[l][i] : 0.f;
}

// ------------------------------------------------------------------------------------------------
struct TcArgs {
    const float* x;        // network input, NCHW fp32 (layer 1 source; layer 7 curve input)
    __nv_bfloat16* dst;    // NHWC bf16 [B,H,W,32]   (!LAST)
    float* adjust;         // LAST: [B,3*iters,H,W] or nullptr
    float* enhanced;       // LAST: [B,3,H,W]
    const uint8_t* wpack;  // this layer's packed block
    int B, H, W, R, nseg, nstrip, iters;
};

template <int NSRC, bool FIRST>
__host__ __device__ constexpr int tc_wbytes_t() {
    return FIRST ? 3 * 3072 : 3 * 2 * NSRC * 3072;
}
template <int NSRC, bool FIRST>
constexpr int tc_smem_bytes() {
    // weights | bias 128 | barriers 384 | tmem addr 16 | pad to 1024 | stages
    return tc_wbytes_t<NSRC, FIRST>() + 128 + 384 + 16 + 1024 +
           (FIRST ? 0 : TC_NST * NSRC * TC_SRC_STRIDE);
}

__device__ __forceinline__ int tc_pos(uint32_t o) { return TC_NB - 1 - (int)(o % TC_NB); }

template <int NSRC, bool FIRST, bool LAST>
__global__ void __launch_bounds__(TC_THREADS, 1)
dce_tc_layer_kernel(const TcArgs args, const __grid_constant__ CUtensorMap tm0, const __grid_constant__ CUtensorMap tm1) {
    constexpr int KC = FIRST ? 1 : 2 * NSRC;            // K chunks of 16 per (dy, dx)
    constexpr int NDY = 3;                              // output rows one input row feeds
    constexpr int NDX = 3;
    constexpr int WBYTES = tc_wbytes_t<NSRC, FIRST>();
    constexpr int TILE = 3072;
    constexpr int COPY_COLS = FIRST ? 8 : 16 * NSRC;    // TMEM columns of one dx copy
    constexpr int SLOT_COLS = NDX * COPY_COLS;
    static_assert(TC_ACC_COLS + TC_NSLOT * SLOT_COLS <= 512, "TMEM budget");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* s_w = smem_raw;
    float* s_bias = reinterpret_cast<float*>(smem_raw + WBYTES);
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(smem_raw + WBYTES + 128);
    uint64_t* stage_full = s_bar;                                   // [NST]   TMA complete_tx (1 arrival + bytes)
    uint64_t* stage_empty = s_bar + TC_NST;                         // [NST]   loaders of one parity (128)
    uint64_t* full = s_bar + 2 * TC_NST;                            // [NSLOT] loaders of one parity (128)
    uint64_t* empty = s_bar + 2 * TC_NST + TC_NSLOT;                // [NSLOT] MMA commit (1)
    uint64_t* acc_full = s_bar + 2 * TC_NST + 2 * TC_NSLOT;         // [NB]    MMA commit (1)
    uint64_t* acc_empty = s_bar + 2 * TC_NST + 2 * TC_NSLOT + TC_NB;  // [NB]  epilogue warps of one parity (128)
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem_raw + WBYTES + 128 + 384);
    // stage area: 1024-byte aligned in the shared window (swizzle phase = address bits 7-8)
    const uint32_t stage_u32 = (smem_u32(smem_raw + WBYTES + 128 + 384 + 16) + 1023u) & ~1023u;
    uint8_t* s_stage = smem_raw + (stage_u32 - smem_u32(smem_raw));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int H = args.H, W = args.W, R = args.R;
    const int n_items = args.B * args.nseg * args.nstrip;

    // ---- one-time setup ----
    {
        const uint4* g = reinterpret_cast<const uint4*>(args.wpack);
        uint4* s = reinterpret_cast<uint4*>(smem_raw);
        for (int i = tid; i < (WBYTES + 128) / 16; i += TC_THREADS) s[i] = __ldg(g + i);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // B tiles are read through the async proxy
    }
    if (warp == 8) {
        if (lane == 0) {
            for (int i = 0; i < TC_NST; ++i) {
                mbar_init(&stage_full[i], 1);
                mbar_init(&stage_empty[i], 128);
            }
            for (int i = 0; i < TC_NSLOT; ++i) {
                mbar_init(&full[i], 128);
                mbar_init(&empty[i], 1);
            }
            for (int i = 0; i < TC_NB; ++i) {
                mbar_init(&acc_full[i], 1);
                mbar_init(&acc_empty[i], 128);
            }
            mbar_fence_init();
        }
        __syncwarp();
        tmem_alloc(s_tmem, 512);
    }
    if (!FIRST && warp == 9 && lane == 0) {
        tma_prefetch_desc(&tm0);
        if (NSRC == 2) tma_prefetch_desc(&tm1);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = *s_tmem;
    if (warp >= 10 && warp < 14) {  // every MMA accumulates: start from zeroed accumulator blocks
        const uint32_t tl = tbase + ((uint32_t)((warp & 3) * 32) << 16);
        for (int i = 0; i < TC_NB; ++i) tmem_zero32(tl + i * 32);
        tc_wait_st();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    if (warp < 8) {
        // =========================== LOADERS ===========================
        const int q = warp & 3, par = warp >> 2;
        const int t = q * 32 + lane;  // pixel in strip = TMEM lane
        const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16) + TC_ACC_COLS;
        uint32_t g_base = 0;
        if constexpr (FIRST) {
            // Layer 1: Cin = 3, K padded to 16 (8 TMEM columns per dx copy, only the first two are non-zero and are
            // the only ones rewritten).  Each lane reads its 3 x 3 (dx, channel) fp32 values straight from the NCHW
            // input (coalesced along x), one row ahead of the one it is converting.
            for (int sl = par; sl < TC_NSLOT; sl += 2) {  // zero the padding columns of every slot once
                const uint32_t z[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
                tmem_st16(tlane + sl * SLOT_COLS, z);
                tmem_st16(tlane + sl * SLOT_COLS + 8, z);
            }
            tc_wait_st();
            named_bar_sync(1, 256);
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                const int strip = item % args.nstrip;
                const int seg = (item / args.nstrip) % args.nseg;
                const int b = item / (args.nstrip * args.nseg);
                const int y0 = seg * R, xx = strip * TC_M + t;
                const int nin = min(R, H - y0) + 2;
                const float* xin = args.x + (size_t)b * 3 * H * W;
                auto load_row = [&](int y, float (&v)[9]) {
                    const bool yok = y >= 0 && y < H;
#pragma unroll
                    for (int dx = 0; dx < 3; ++dx) {
                        const int xs = xx + dx - 1;
                        const bool ok = yok && xs >= 0 && xs < W;
#pragma unroll
                        for (int c = 0; c < 3; ++c) v[dx * 3 + c] = ok ? __ldg(xin + ((size_t)c * H + y) * W + xs) : 0.f;
                    }
                };
                int ri = (int)((g_base ^ (uint32_t)par) & 1);  // first input row of this item with (g & 1) == par
                float nxt[9];
                if (ri < nin) load_row(y0 - 1 + ri, nxt);
                for (; ri < nin; ri += 2) {
                    float cur[9];
#pragma unroll
                    for (int i = 0; i < 9; ++i) cur[i] = nxt[i];
                    if (ri + 2 < nin) load_row(y0 - 1 + ri + 2, nxt);
                    const uint32_t g = g_base + ri;
                    const uint32_t slot = g % TC_NSLOT, use = g / TC_NSLOT;
                    if (use > 0) mbar_wait(&empty[slot], (use - 1) & 1);
                    tc_fence_after();
#pragma unroll
                    for (int dx = 0; dx < 3; ++dx) {
                        __nv_bfloat162 h0 = __floats2bfloat162_rn(cur[dx * 3], cur[dx * 3 + 1]);
                        __nv_bfloat162 h1 = __floats2bfloat162_rn(cur[dx * 3 + 2], 0.f);
                        tmem_st2(tlane + slot * SLOT_COLS + dx * COPY_COLS, *reinterpret_cast<uint32_t*>(&h0),
                                 *reinterpret_cast<uint32_t*>(&h1));
                    }
                    tc_wait_st();
                    tc_fence_before();
                    mbar_arrive(&full[slot]);
                }
                g_base += nin;
            }
        } else {
            // per-thread swizzled smem offsets of the three pixels (dx copies) this lane feeds:
            // TMA SWIZZLE_64B stores 16-byte chunk c of pixel p at  p*64 + ((c ^ ((p >> 1) & 3)) << 4)
            uint32_t poff[3], psw[3];
#pragma unroll
            for (int dx = 0; dx < 3; ++dx) {
                const int p = t + dx;
                poff[dx] = p * 64;
                psw[dx] = (p >> 1) & 3;
            }
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                const int seg = (item / args.nstrip) % args.nseg;
                const int nin = min(R, H - seg * R) + 2;
                for (int ri = 0; ri < nin; ++ri) {
                    const uint32_t g = g_base + ri;
                    if ((int)(g & 1) != par) continue;
                    const uint32_t s = g % TC_NST, us = g / TC_NST;
                    mbar_wait(&stage_full[s], us & 1);
                    const uint32_t slot = g % TC_NSLOT, use = g / TC_NSLOT;
                    if (use > 0) mbar_wait(&empty[slot], (use - 1) & 1);
                    tc_fence_after();
                    const uint8_t* st = s_stage + s * (NSRC * TC_SRC_STRIDE);
#pragma unroll
                    for (int src = 0; src < NSRC; ++src) {
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx) {
                            const uint8_t* pp = st + src * TC_SRC_STRIDE + poff[dx];
                            const uint4 q0 = *reinterpret_cast<const uint4*>(pp + ((0 ^ psw[dx]) << 4));
                            const uint4 q1 = *reinterpret_cast<const uint4*>(pp + ((1 ^ psw[dx]) << 4));
                            const uint4 q2 = *reinterpret_cast<const uint4*>(pp + ((2 ^ psw[dx]) << 4));
                            const uint4 q3 = *reinterpret_cast<const uint4*>(pp + ((3 ^ psw[dx]) << 4));
                            const uint32_t pk[16] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w,
                                                     q2.x, q2.y, q2.z, q2.w, q3.x, q3.y, q3.z, q3.w};
                            tmem_st16(tlane + slot * SLOT_COLS + dx * COPY_COLS + src * 16, pk);
                        }
                    }
                    tc_wait_st();
                    tc_fence_before();
                    mbar_arrive(&full[slot]);
                    mbar_arrive(&stage_empty[s]);
                }
                g_base += nin;
            }
        }
    } else if (warp == 9) {
        // =========================== TMA PRODUCER ===========================
        // warp-uniform control flow, one elected lane issues expect_tx + the bulk tensor copies
        if constexpr (!FIRST) {
            const bool leader = elect_one();
            uint32_t s = 0, spar = 0;  // stage ring position / phase parity of the NEXT fill
            bool first_round = true;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                const int strip = item % args.nstrip;
                const int seg = (item / args.nstrip) % args.nseg;
                const int b = item / (args.nstrip * args.nseg);
                const int y0 = seg * R, x0 = strip * TC_M;
                const int nin = min(R, H - y0) + 2;
                for (int ri = 0; ri < nin; ++ri) {
                    if (!first_round) mbar_wait(&stage_empty[s], spar ^ 1);
                    if (leader) {
                        mbar_arrive_expect_tx(&stage_full[s], NSRC * TC_SRC_BYTES);
                        const uint32_t dst = stage_u32 + s * (NSRC * TC_SRC_STRIDE);
                        tma_load_4d(dst, &tm0, 0, x0 - 1, y0 - 1 + ri, b, &stage_full[s]);
                        if (NSRC == 2)
                            tma_load_4d(dst + TC_SRC_STRIDE, &tm1, 0, x0 - 1, y0 - 1 + ri, b, &stage_full[s]);
                    }
                    __syncwarp();
                    if (++s == TC_NST) {
                        s = 0;
                        spar ^= 1;
                        first_round = false;
                    }
                }
            }
        }
    } else if (warp == 8) {
        // =========================== MMA ISSUER ===========================
        // The whole warp runs the (warp-uniform) control flow so that addresses and descriptors stay in uniform
        // registers; one elected lane issues every tcgen05.mma / tcgen05.commit.
        constexpr uint32_t IDESC0 = idesc_bf16_f32(TC_M, 0);
        constexpr uint32_t IDESC_FULL = IDESC0 | ((uint32_t)(NDY * 32 >> 3) << 17);
        const uint64_t bdesc0 = smem_desc_kmajor(smem_u32(s_w), 128, 256);
        const bool leader = elect_one();
        uint32_t slot = 0, spar = 0;  // A-slot ring position and phase parity
        uint32_t o_base = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int seg = (item / args.nstrip) % args.nseg;
            const int nout = min(R, H - seg * R);
            const int nin = nout + NDY - 1;
            uint32_t on = o_base % TC_NB, upar = (o_base / TC_NB) & 1;  // block ring position / use parity of row o_base+ri
            for (int ri = 0; ri < nin; ++ri) {
                mbar_wait(&full[slot], spar);
                const int pn = TC_NB - 1 - (int)on;  // block opened by this input row (output row ri, dy = 0)
                if (ri < nout && o_base + ri >= TC_NB) mbar_wait(&acc_empty[pn], upar ^ 1);
                tc_fence_after();
                const uint32_t abase = tbase + TC_ACC_COLS + slot * SLOT_COLS;
                if (ri >= NDY - 1 && ri < nout && pn + NDY <= TC_NB) {
                    // interior row, no ring wrap: one N = 32*NDY MMA per (dx, kc)
                    if (leader) {
                        const uint32_t d = tbase + pn * 32;
#pragma unroll
                        for (int dx = 0; dx < NDX; ++dx)
#pragma unroll
                            for (int kc = 0; kc < KC; ++kc)
                                mma_ts(d, abase + dx * COPY_COLS + kc * 8,
                                       bdesc0 + (uint64_t)(((dx * KC + kc) * TILE) >> 4), IDESC_FULL, 1u);
                    }
                } else if (leader) {
                    // segment borders (output row j = ri - dy must lie in [0, nout)) and ring wrap: split into groups
                    int dy = max(0, ri - (nout - 1));
                    const int dy_hi = min(NDY - 1, ri);
                    while (dy <= dy_hi) {
                        const int p0 = (pn + dy) % TC_NB;
                        const int cnt = min(dy_hi - dy + 1, TC_NB - p0);
                        const uint32_t idesc = IDESC0 | ((uint32_t)(cnt * 32 >> 3) << 17);
                        const uint32_t d = tbase + p0 * 32;
                        const uint64_t bd = bdesc0 + (uint64_t)((dy * 1024) >> 4);
#pragma unroll
                        for (int dx = 0; dx < NDX; ++dx)
#pragma unroll
                            for (int kc = 0; kc < KC; ++kc)
                                mma_ts(d, abase + dx * COPY_COLS + kc * 8,
                                       bd + (uint64_t)(((dx * KC + kc) * TILE) >> 4), idesc, 1u);
                        dy += cnt;
                    }
                }
                if (leader) {
                    tc_commit(&empty[slot]);
                    if (ri >= NDY - 1) tc_commit(&acc_full[(pn + NDY - 1) % TC_NB]);
                }
                __syncwarp();
                if (++slot == TC_NSLOT) {
                    slot = 0;
                    spar ^= 1;
                }
                if (++on == TC_NB) {
                    on = 0;
                    upar ^= 1;
                }
            }
            o_base += nout;
        }
    } else {
        // =========================== EPILOGUE ===========================
        const int q = warp & 3, par = (warp - 10) >> 2;
        const int m = q * 32 + lane;
        const uint32_t tlane = tbase + ((uint32_t)(q * 32) << 16);
        float bias[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) bias[i] = s_bias[i];
        uint32_t o_base = 0;
        const size_t plane = (size_t)H * W;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int strip = item % args.nstrip;
            const int seg = (item / args.nstrip) % args.nseg;
            const int b = item / (args.nstrip * args.nseg);
            const int y0 = seg * R, xx = strip * TC_M + m;
            const int nout = min(R, H - y0);
            for (int j = 0; j < nout; ++j) {
                const uint32_t o = o_base + j;
                if ((int)(o & 1) != par) continue;
                const int p = tc_pos(o);
                mbar_wait(&acc_full[p], (o / TC_NB) & 1);
                tc_fence_after();
                uint32_t v[32];
                tmem_ld32(tlane + p * 32, v);
                tc_wait_ld();
                tmem_zero32(tlane + p * 32);
                tc_wait_st();
                tc_fence_before();
                mbar_arrive(&acc_empty[p]);
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
                    float* ap = args.adjust ? args.adjust + (size_t)b * 3 * iters * plane + pix : nullptr;
                    const float* xp = args.x + (size_t)b * 3 * plane + pix;
                    float* ep = args.enhanced + (size_t)b * 3 * plane + pix;
                    float yv[3];
#pragma unroll
                    for (int c = 0; c < 3; ++c) yv[c] = __ldg(xp + (size_t)c * plane);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        if (i < iters) {
#pragma unroll
                            for (int c = 0; c < 3; ++c) {
                                const float a = tanh_approx(__uint_as_float(v[i * 3 + c]) + bias[i * 3 + c]);
                                if (ap) ap[(size_t)(i * 3 + c) * plane] = a;
                                yv[c] = curve_step(yv[c], a);
                            }
                        }
                    }
#pragma unroll
                    for (int c = 0; c < 3; ++c) ep[(size_t)c * plane] = yv[c];
                }
            }
            o_base += nout;
        }
    }

    // ---- teardown ----
    tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc_fence_after();
        tmem_dealloc(tbase, 512);
    }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static size_t tc_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

typedef CUresult (*tc_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda dependency)
static tc_encode_fn tc_encoder() {
    static tc_encode_fn fn = []() -> tc_encode_fn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            (void)cudaGetLastError();
            return nullptr;
        }
        return reinterpret_cast<tc_encode_fn>(p);
    }();
    return fn;
}

// NHWC bf16 [B,H,W,32] as a 4-D tensor (c, x, y, b); box = 32 channels x 130 pixels of one row; 64-byte swizzle;
// coordinates outside the image (x = -1, x >= W, y = -1, y >= H) are zero-filled = the conv's zero padding.
static int tc_make_map(CUtensorMap* m, const void* base, int B, int H, int W) {
    tc_encode_fn enc = tc_encoder();
    if (!enc) return B200_ERR_LAUNCH;
    const cuuint64_t dims[4] = {32, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    const cuuint64_t strides[3] = {64, (cuuint64_t)W * 64, (cuuint64_t)H * W * 64};
    const cuuint32_t box[4] = {32, (cuuint32_t)TC_SPX, 1, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? B200_OK : B200_ERR_LAUNCH;
}

template <int NSRC, bool FIRST, bool LAST>
static int tc_launch(const TcArgs& a, const void* s0, const void* s1, cudaStream_t s) {
    auto kern = dce_tc_layer_kernel<NSRC, FIRST, LAST>;
    constexpr int smem = tc_smem_bytes<NSRC, FIRST>();
    static_assert(smem <= 227 * 1024, "shared memory budget");
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;
    }
    CUtensorMap m0, m1;
    memset(&m0, 0, sizeof(m0));
    memset(&m1, 0, sizeof(m1));
    if (!FIRST) {
        int rc;
        if ((rc = tc_make_map(&m0, s0, a.B, a.H, a.W))) return rc;
        if (NSRC == 2 && (rc = tc_make_map(&m1, s1, a.B, a.H, a.W))) return rc;
    }
    const long long items = (long long)a.B * a.nseg * a.nstrip;
    const int grid = (int)(items < B200_NUM_SMS ? items : B200_NUM_SMS);
    kern<<<grid, TC_THREADS, smem, s>>>(a, m0, m1);
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
    a.x = x;
    a.B = B;
    a.H = H;
    a.W = W;
    a.iters = iters;
    a.R = H < 40 ? H : 40;
    a.nseg = (H + a.R - 1) / a.R;
    a.nstrip = (W + TC_M - 1) / TC_M;
    int rc;
    auto layer = [&](int l, __nv_bfloat16* dst) {
        a.dst = dst;
        a.wpack = pk + tc_layer_offset(l);
    };
    layer(0, x1);
    if ((rc = tc_launch<1, true, false>(a, nullptr, nullptr, s))) return rc;
    layer(1, x2);
    if ((rc = tc_launch<1, false, false>(a, x1, nullptr, s))) return rc;
    layer(2, x3);
    if ((rc = tc_launch<1, false, false>(a, x2, nullptr, s))) return rc;
    layer(3, x4);
    if ((rc = tc_launch<1, false, false>(a, x3, nullptr, s))) return rc;
    layer(4, x5);
    if ((rc = tc_launch<2, false, false>(a, x3, x4, s))) return rc;
    layer(5, x6);
    if ((rc = tc_launch<2, false, false>(a, x2, x5, s))) return rc;
    layer(6, nullptr);
    a.adjust = adjust;
    a.enhanced = enhanced;
    if ((rc = tc_launch<2, false, true>(a, x1, x6, s))) return rc;
    return B200_OK;
}

}  // namespace b200

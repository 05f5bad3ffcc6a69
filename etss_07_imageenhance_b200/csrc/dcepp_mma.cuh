// Zero-DCE++ layers 2..7 (depthwise 3x3 -> pointwise 1x1, zero_dcepp.py:191-196 / dsconv.py:145-149) for sm_100a,
// as ONE persistent, TMA-fed launch.
//
// Arithmetic.  The depthwise stage stays on the fp32 CUDA cores, two channels per instruction (fma.rn.f32x2); the
// pointwise stage is a [pixels x Cin] . [Cin x Cout] GEMM on warp-level tensor-core MMAs (mma.sync m16n8k16, bf16
// operands, fp32 accumulate) with a 3-term split so the result keeps fp32-class accuracy (the path's contract is
// max-abs <= 1e-3 against the fp32 reference):
//     a = a_hi + a_lo,  w = w_hi + w_lo  (bf16 each)      a.w ~= a_hi.w_hi + a_lo.w_hi + a_hi.w_lo      (|err| ~ 2^-16)
// There is no operand exchange between the stages: every lane computes exactly the depthwise outputs that ARE its
// A fragment -- for an m16 tile laid out as 2 image rows x 8 columns, lane (g = lane/4, c = lane%4) owns pixels
// (row 0/1, column g) x channel pairs {c, c+4} of a 16-channel block; two stacked m-tiles give it 4 vertically
// adjacent pixels, i.e. a 6-row sliding window per channel pair.
//
// Layout.  The feature maps x1..x6 are private to the path (caller's workspace), so they are stored the way both
// stages want them: CHANNEL-PAIR INTERLEAVED, [B][16 pairs][h][wp] float2 (wp = w rounded up to even, so rows are
// 16-byte multiples).  A float2 is (channel 2p, channel 2p+1) of one pixel = one FFMA2 operand = one A-fragment
// register pair before the split = one C-fragment register pair after the MMA (stored with one 64-bit store).
//
// Data movement.  A 3-D tensor map per feature map (8-byte elements: x, y, pair + 16*b) lets one elected lane fetch a
// stage = 8 pairs x 10 rows x 38 columns (a tile's halo for one k16 block, 24.3 KB) with a single
// cp.async.bulk.tensor; coordinates outside the map (x < 0, x >= w, ...) are zero-filled = the conv's zero padding,
// so there is no per-thread halo addressing at all.  With 380 float2 per pair plane the (g, c) access pattern of the
// 64-bit shared loads is bank-conflict free.
//
// Schedule.  One CTA per SM holds two independent HALVES (a producer warp + 8 consumer warps each, own 4-slot ring and
// barriers; 18 warps are allocated as 20, i.e. 96 registers per thread, like two 9-warp CTAs).  The 6 x (B x tiles)
// output tiles form one work list ordered in frame groups (see below), layer-major inside a group.  A half's producer
// draws tile numbers from a global counter, waits until the (up to) nine layer l-1 tiles under the tile's halo are
// published (acquire), and streams the tile's 2 or 4 stages through the ring (full / empty mbarriers) -- running up
// to a whole tile ahead of its consumers, so tile-fetch, dependency and load latencies are off the critical path.
// Every stage carries, behind its feature-map box, the 2.8 KB weight blob of its k16 block (a bulk copy into the same
// slot on the same barrier): a half holds no resident weights and changes layers from one tile to the next for
// free.  A tile is published by the LAST consumer warp to finish it (shared-memory arrival counter, then a release
// at GPU scope); there is no CTA barrier anywhere in the loop.  No launch boundary or wave tail between layers.
// Deadlock-free: a producer only waits on tiles with smaller numbers, drawn earlier by halves that are resident
// (grid <= one resident wave) and whose consumers never wait on anything but their own producer.
//
// Whole forward in one launch (b200_dcepp_forward; W % 4 == 0 and at most 64 low-res columns under a block).  The
// full-resolution pass (bilinear upsample, align_corners=True, + the 8-iteration curve; zero_dcepp.py:198-203) is cut
// into blocks of `br` rows x `bw` pixels x 3 planes of x that fit one ring slot together with the tile of the
// low-resolution map under them, and joins the same queue: a block of frame f becomes runnable when all layer-7
// tiles of f have published (per-frame counter, release / acquire); the producer fetches x AND the low-res tile with
// two bulk tensor copies (up to four blocks = 96 KB in flight per half, so the HBM read stream never depends on
// registers or occupancy), the consumer warps work from shared memory only, run the curve and store `enhanced` /
// `adjust` with 128-bit streaming stores.  Blocks are claimed four at a time with one atomicAdd whether or not their
// frame is ready; a claimed block whose frame is not ready stays pending while the producer goes on with tiles, so
// nobody ever waits for a block and the deadlock argument above is unchanged.  One half per SM prefers blocks (falls
// back to tiles while none is ready), the other prefers tiles (falls back to blocks when the tiles are exhausted).
// Frame groups: tiles are ordered group by group so that the blocks of early groups can run while later groups are
// still in the net; measured (profiles/r02_summary.md), what pays is a group of ~3.25 tiles per half and layer --
// smaller groups start the blocks earlier but stall producers on predecessors still in flight, and the sum of
// half-time (tiles + blocks) is conserved: both paths are latency-bound at 18 warps per SM.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "tcgen05.cuh"
#include "tma_host.cuh"

namespace b200 {

constexpr int MM_TH = 8, MM_TW = 32;                 // output tile
// halo rows / box columns.  TMA wants the box's first byte 16-byte aligned in the innermost dimension (measured:
// tools/exp/tma_cfg_probe.cu -- an odd x start with 8-byte elements raises an illegal-instruction fault), so the box
// starts at x0 - 2 instead of x0 - 1; 38 columns (2 + 32 + 1, +3 unused) keep the pair-plane stride = 4 mod 16.
constexpr int MM_HR = MM_TH + 2, MM_HC = 38;
constexpr int MM_XOFF = 2;                           // image column of box column 0 = x0 - MM_XOFF
constexpr int MM_PS = MM_HR * MM_HC;                 // float2 per pair plane of a stage
constexpr int MM_STAGE_PAIRS = 8;                    // one k16 block
constexpr int MM_STAGE_BYTES = MM_STAGE_PAIRS * MM_PS * 8;
constexpr int MM_NS = 4;          // ring slots
constexpr int MM_HALF = 288;      // one half of a CTA: 8 consumer warps + the producer warp, own ring and barriers
constexpr int MM_THREADS = 2 * MM_HALF;  // two independent halves per CTA, one CTA per SM (18 warps are allocated as
                                         // 20: 96 registers per thread, as with two 9-warp CTAs)
static_assert(MM_STAGE_BYTES % 128 == 0, "TMA destinations stay 128-byte aligned");
static_assert(MM_PS % 16 == 4 || MM_PS % 16 == 12, "pair-plane stride keeps the (g, c) 64-bit loads conflict free");

// per-layer packed block (32-bit words): one blob per k16 block (= per stage), shipped into the ring slot behind the
// stage's feature-map box by a bulk copy, so a CTA holds no resident weights and changes layers for free:
//   blob kt: dw [8 pairs][20] | frags [NT][32 lanes][4] | pw bias [32]
//   dw row of pair p: 9 x (tap of channel 2p, tap of channel 2p+1), then (bias 2p, bias 2p+1)
__host__ __device__ constexpr int mm_cin(int l) { return l <= 3 ? 32 : 64; }  // l = 1..6 (e_conv2..7)
__host__ __device__ constexpr int mm_nt(int l) { return l == 6 ? 1 : 4; }
__host__ __device__ constexpr int mm_blob_words(int l) { return 160 + mm_nt(l) * 128 + 32; }
__host__ __device__ constexpr int mm_words(int l) { return (mm_cin(l) / 16) * mm_blob_words(l); }
__host__ __device__ constexpr int mm_offset(int l) {
    int o = 0;
    for (int i = 1; i < l; ++i) o += mm_words(i);
    return o;
}
constexpr int MM_TOTAL_WORDS = mm_offset(7);
constexpr int MM_BLOB_BYTES_MAX = mm_blob_words(1) * 4;
constexpr int MM_SLOT_BYTES = MM_STAGE_BYTES + MM_BLOB_BYTES_MAX;  // feature-map box | weight blob
static_assert(MM_SLOT_BYTES % 128 == 0 && (mm_blob_words(1) * 4) % 16 == 0 && (mm_blob_words(6) * 4) % 16 == 0,
              "bulk-copy sizes / TMA destinations");
constexpr int MM_VCOLS = 64;                      // low-res columns under one block of the full-resolution pass
constexpr int MM_VROW_FLOATS = 8 * 3 * MM_VCOLS;  // per CTA: [consumer warp][channel][column]
// ring | full / empty barriers | 4 item descriptors | 4 tile arrival counters (+ pad) | per-warp strips of the full-res pass
constexpr int MM_HALF_SMEM = MM_NS * MM_SLOT_BYTES + 2 * MM_NS * 8 + 4 * 16 + 128 + MM_VROW_FLOATS * 4;
constexpr int MM_NET_SMEM = 2 * MM_HALF_SMEM;
static_assert(MM_HALF_SMEM % 128 == 0, "the second half's ring stays 128-byte aligned");
static_assert(MM_NET_SMEM <= 227 * 1024, "one CTA per SM");

__device__ __forceinline__ uint32_t mm_pack_bf16x2(float lo_elem, float hi_elem) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo_elem, hi_elem);  // .x (low 16 bits) = lo_elem
    return *reinterpret_cast<uint32_t*>(&h);
}
// split two fp32 values into packed bf16 hi and lo parts
__device__ __forceinline__ void mm_split2(float v0, float v1, uint32_t& hi, uint32_t& lo) {
    hi = mm_pack_bf16x2(v0, v1);
    const float h0 = __uint_as_float(hi << 16), h1 = __uint_as_float(hi & 0xffff0000u);
    lo = mm_pack_bf16x2(v0 - h0, v1 - h1);
}

// one layer's block; called by dcepp_pack_kernel (block l = 1..6)
__device__ void mm_pack_layer(int l, const float* dw_w, const float* dw_b, const float* pw_w, const float* pw_b,
                              uint32_t* out) {
    const int cin = mm_cin(l), nt_n = mm_nt(l), cout = l == 6 ? 3 : 32;
    const int bw = mm_blob_words(l), kt_n = cin / 16;
    for (int i = threadIdx.x; i < kt_n * bw; i += blockDim.x) {
        const int kt = i / bw, e = i - kt * bw;
        float* f = reinterpret_cast<float*>(out);
        if (e < 160) {  // depthwise taps + bias of pairs kt*8 .. kt*8+7
            const int p = kt * 8 + e / 20, j = e % 20;
            const int tap = j >> 1, ch = 2 * p + (j & 1);
            f[i] = tap < 9 ? dw_w[ch * 9 + tap] : dw_b[ch];
        } else if (e < 160 + nt_n * 128) {  // pointwise B fragments of this k16 block, split {hi, lo}
            const int fi = e - 160;
            const int j = fi & 3, lane = (fi >> 2) & 31, nt = fi >> 7;
            const int g = lane >> 2, c = lane & 3;
            const int n = nt * 8 + g;
            const int k0 = kt * 16 + 2 * c + ((j & 1) ? 8 : 0);
            float w0 = 0.f, w1 = 0.f;
            if (n < cout) {
                w0 = pw_w[n * cin + k0];
                w1 = pw_w[n * cin + k0 + 1];
            }
            uint32_t hi, lo;
            mm_split2(w0, w1, hi, lo);
            out[i] = j < 2 ? hi : lo;
        } else {  // pointwise bias (every blob carries it; the tile's first stage is the one that is read)
            const int o = e - 160 - nt_n * 128;
            f[i] = o < cout ? pw_b[o] : 0.f;
        }
    }
}

__device__ __forceinline__ void mm_mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ int mm_ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st_release(int* p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void mm_red_release_add(int* p, int v) {
    asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// sync words (caller's workspace, zeroed by layer 1's launch):
//   [0] tile counter | [1] block counter (full-resolution pass) | [2] abort flag (a spin guard tripped) |
//   [MM_FDONE0 + b] layer-7 tiles of frame b published | [mm_flag0(B) + q] tile q of layers 2..6 published
constexpr int MM_FDONE0 = 32;
__host__ __device__ inline int mm_flag0(int B) { return (MM_FDONE0 + B + 31) & ~31; }
__host__ __device__ inline size_t mm_sync_words(int B, int h, int w) {
    const size_t tiles = (size_t)B * ((h + MM_TH - 1) / MM_TH) * ((w + MM_TW - 1) / MM_TW);
    return mm_flag0(B) + 5 * tiles;
}
__host__ __device__ inline int mm_pitch(int w) { return (w + 1) & ~1; }  // float2 per feature-map row

struct MmNetArgs {
    float* dst[6];          // x2 x3 x4 x5 x6 (pair interleaved) | a_lr (planar [B,3,h,w])
    const uint32_t* wl[6];  // packed blocks of e_conv2..7
};

// the full-resolution pass as queue items (total == 0: this launch runs the net only)
struct MmApplyArgs {
    float* enh;     // [B,3,H,W]
    float* adj;     // [B,3,H,W] or nullptr
    int H, W, iters;
    float sh, sw;   // ATen area_pixel_compute_scale(align_corners=True): (h-1)/(H-1), (w-1)/(W-1)
    int bw, br;     // block: bw pixels (multiple of 4) x br rows x 3 planes of x ...
    int ac, ar;     // ... + the ac columns x ar rows x 3 channels of the low-res map under it <= one ring stage
    int aoff;       // float offset of the low-res tile inside the stage
    int wq;         // row pitch of the low-res map (floats, multiple of 4: a tensor map reads it)
    int nbx, nby;   // blocks per row band, bands per frame
    int total;      // B * nbx * nby
    int prefer;     // CTAs per SM that prefer blocks over tiles (0..2)
    int group;      // frames per tile-ordering group
    uint32_t* trace;  // debug: per-item timeline records {cta << 16 | smid << 4 | kind, t_wait, t_start, t_end} (ns), or nullptr
    int trace_cap;    // records that fit (record 0 holds the running count)
};
__device__ __forceinline__ uint32_t mm_now_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return (uint32_t)t;
}

// Spin guard: every wait of the launch gives up after ~2 s and raises the abort word; once it is up, every wait of
// every CTA returns at once (checked every 256 spins of a slow path only) and the control flow runs to its normal end
// on garbage -- a scheduling bug must show up as wrong numbers in a test, never as a hung device.
constexpr long long MM_SPIN_CLOCKS = 4000000000LL;
__device__ __forceinline__ bool mm_spin_expired(long long t0, int& n, int* sync) {
    if ((++n & 255) != 0) return false;
    if (*reinterpret_cast<volatile int*>(sync + 2)) return true;
    if (clock64() - t0 < MM_SPIN_CLOCKS) return false;
    atomicExch(sync + 2, 1);
    return true;
}
__device__ __forceinline__ void mm_mbar_wait(uint64_t* bar, uint32_t parity, int* sync) {
    if (tc::mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    int n = 0;
    while (!tc::mbar_try_wait(bar, parity))
        if (mm_spin_expired(t0, n, sync)) return;
}
__device__ __forceinline__ float mm_ld_cg(const float* p) {  // L2 (coherent): data written earlier in this launch
    float v;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
// w0*v0 + w1*v1 with the contraction spelled out (the unfused full-resolution kernel uses the same helper, so both
// paths round identically)
__device__ __forceinline__ float mm_lerp(float w0, float v0, float w1, float v1) {
    return __fmaf_rn(w0, v0, __fmul_rn(w1, v1));
}

// One 8 x 32 output tile, executed by the 8 consumer warps.  `ck` = this CTA's running stage counter.
template <int NSRC, int NT, bool TANH>
__device__ __forceinline__ void mm_tile(int t, const uint8_t* ring, uint64_t* full, uint64_t* empty, int& ck,
                                        float* __restrict__ dst, int h, int w, int wp, int b, int y0, int x0,
                                        int wq = 0) {
    constexpr int CIN = 32 * NSRC, KT = CIN / 16;

    const int lane = t & 31, warp = t >> 5;
    const int g = lane >> 2, c = lane & 3;
    const int wy = (warp >> 2) * 4, wx = (warp & 3) * 8;

    // acc[mt][nt] = {(row g, n 2c), (g, 2c+1), (g+8, 2c), (g+8, 2c+1)}; starts at the pointwise bias, read from the
    // blob of the tile's first stage (which the caller has already waited for)
    float acc[2][NT][4];
    {
        const float* s_pb = reinterpret_cast<const float*>(ring + (ck % MM_NS) * MM_SLOT_BYTES + MM_STAGE_BYTES) + 160 +
                            NT * 128;  // [32]
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const float2 pb = *reinterpret_cast<const float2*>(s_pb + nt * 8 + 2 * c);
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) acc[mt][nt][0] = acc[mt][nt][2] = pb.x, acc[mt][nt][1] = acc[mt][nt][3] = pb.y;
        }
    }

#pragma unroll 1
    for (int kt = 0; kt < KT; ++kt) {
        const int slot = ck % MM_NS;
        // not guarded: once a tile's first stage has landed, the producer issues the others without waiting for
        // anything but ring slots, i.e. for these consumers
        tc::mbar_wait(&full[slot], (ck / MM_NS) & 1);
        const float2* st = reinterpret_cast<const float2*>(ring + slot * MM_SLOT_BYTES);
        const float* s_dw = reinterpret_cast<const float*>(ring + slot * MM_SLOT_BYTES + MM_STAGE_BYTES);  // [8][20]
        const uint4* s_fr = reinterpret_cast<const uint4*>(s_dw + 160);                                    // [NT][32]

        // ---- depthwise 3x3 for this lane's fragment entries: 2 channel pairs x 4 vertically adjacent pixels ----
        float2 val[2][4];  // [pair c / c+4][output row]
#pragma unroll
        for (int j2 = 0; j2 < 2; ++j2) {
            const int p = j2 * 4 + c;
            const float4* wq = reinterpret_cast<const float4*>(s_dw + p * 20);
            const float4 q0 = wq[0], q1 = wq[1], q2 = wq[2], q3 = wq[3], q4 = wq[4];
            const float2 wk[9] = {make_float2(q0.x, q0.y), make_float2(q0.z, q0.w), make_float2(q1.x, q1.y),
                                  make_float2(q1.z, q1.w), make_float2(q2.x, q2.y), make_float2(q2.z, q2.w),
                                  make_float2(q3.x, q3.y), make_float2(q3.z, q3.w), make_float2(q4.x, q4.y)};
            // three independent partial sums per output (one per filter column): 12 short dependency chains instead
            // of 4 long ones keep the packed-FMA pipe fed at 4-5 warps per scheduler
            float2 o[4][3];
            const float2* hp = st + p * MM_PS + wy * MM_HC + wx + g + (MM_XOFF - 1);
#pragma unroll
            for (int r = 0; r < 6; ++r) {
                const float2 v[3] = {hp[r * MM_HC], hp[r * MM_HC + 1], hp[r * MM_HC + 2]};
#pragma unroll
                for (int dy = 0; dy < 3; ++dy) {
                    const int oo = r - dy;
                    if (oo >= 0 && oo < 4) {
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx) {
                            if (dy == 0)
                                o[oo][dx] = dx == 0 ? __ffma2_rn(wk[dx], v[dx], make_float2(q4.z, q4.w))  // + bias
                                                    : __fmul2_rn(wk[dx], v[dx]);
                            else
                                o[oo][dx] = __ffma2_rn(wk[dy * 3 + dx], v[dx], o[oo][dx]);
                        }
                    }
                }
            }
#pragma unroll
            for (int oo = 0; oo < 4; ++oo) val[j2][oo] = __fadd2_rn(__fadd2_rn(o[oo][0], o[oo][1]), o[oo][2]);
        }
        // ---- A fragments (hi / lo) of the two m-tiles, then 3 MMAs per n-tile (B fragments from the stage's blob) ----
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            uint32_t ah[4], al[4];
            mm_split2(val[0][2 * mt].x, val[0][2 * mt].y, ah[0], al[0]);          // row g    , k = 2c, 2c+1
            mm_split2(val[0][2 * mt + 1].x, val[0][2 * mt + 1].y, ah[1], al[1]);  // row g + 8, k = 2c, 2c+1
            mm_split2(val[1][2 * mt].x, val[1][2 * mt].y, ah[2], al[2]);          // row g    , k = 2c+8, 2c+9
            mm_split2(val[1][2 * mt + 1].x, val[1][2 * mt + 1].y, ah[3], al[3]);  // row g + 8, k = 2c+8, 2c+9
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const uint4 bf = s_fr[nt * 32 + lane];  // {hi b0, hi b1, lo b0, lo b1}
                mm_mma_bf16(acc[mt][nt], al, bf.x, bf.y);
                mm_mma_bf16(acc[mt][nt], ah, bf.z, bf.w);
                mm_mma_bf16(acc[mt][nt], ah, bf.x, bf.y);
            }
        }
        __syncwarp();  // every lane's shared loads of this stage (halo box and blob) have returned
        if (lane == 0) tc::mbar_arrive(&empty[slot]);
        ++ck;
    }

    // ---- epilogue: activation + stores; image row y0 + wy + r = m-tile r / 2, half r % 2 ----
    const int xx = x0 + wx + g;
    const int rows = h - (y0 + wy);  // valid rows of this warp's 4
    if (xx < w) {
        if (!TANH) {  // pair plane nt*4 + c, one 64-bit store per (pixel, pair)
            const size_t plane = (size_t)h * wp;
            float2* dp = reinterpret_cast<float2*>(dst) + ((size_t)b * 16 + c) * plane + (size_t)(y0 + wy) * wp + xx;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                float2* q = dp + (size_t)(nt * 4) * plane;
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const float2 u = make_float2(fmaxf(acc[r >> 1][nt][(r & 1) * 2], 0.f),
                                                 fmaxf(acc[r >> 1][nt][(r & 1) * 2 + 1], 0.f));
                    if (r < rows) q[(size_t)r * wp] = u;
                }
            }
        } else {  // x_r: planar [B,3,h,wq] (row pitch wq = w for the caller's dense map), channels 2c and 2c+1
            const size_t plane = (size_t)h * wq;
            float* dp = dst + ((size_t)b * 3 + 2 * c) * plane + (size_t)(y0 + wy) * wq + xx;
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                if (2 * c + e >= 3) continue;
#pragma unroll
                for (int r = 0; r < 4; ++r)
                    if (r < rows) dp[(size_t)e * plane + (size_t)r * wq] = tanhf(acc[r >> 1][0][(r & 1) * 2 + e]);
            }
        }
    }
}

// One block of the full-resolution pass, executed by the 8 consumer warps.  The ring slot ck % MM_NS holds, both put
// there by the producer's bulk tensor copies, br rows x bw pixels x 3 planes of x and the ar x ac x 3 tile of the
// low-resolution map under them -- the consumers touch shared memory only (no global load, so no L2 round trip on a
// warp's critical path; with 16 consumer warps per SM there is nothing to hide one behind).  Warp i owns rows i,
// i + 8, ... of the block, a lane the 4-pixel groups lane, lane + 32 of a row.
// Upsample = ATen upsample_bilinear2d(align_corners=True), vertical lerp of the two low-res rows first, horizontal
// second (the order of the unfused kernel): per row, the lanes lerp the <= MM_VCOLS low-res columns under the block
// vertically into the warp's strip of shared memory (column index clamped at w - 1, so the "+1" neighbour of the last
// column is itself, as in ATen); a group then needs 2 shared loads per pixel and channel.  The slot goes back to the
// producer when the warp has read its last x values.
template <bool ITERS8>
__device__ __forceinline__ void mm_curve4(float4& xv, const float4& av, int iters) {
    // y <- y + a * (y*y - y) (zero_dcepp.py:203), ATen's rounding sequence, two pixels per instruction:
    // rn(p - y) with p = rn(y*y) is the FFMA2 by -1 (one rounding), then FMUL2, FADD2
    float2 y0 = make_float2(xv.x, xv.y), y1 = make_float2(xv.z, xv.w);
    const float2 a0 = make_float2(av.x, av.y), a1 = make_float2(av.z, av.w);
    const float2 m1 = make_float2(-1.f, -1.f);
    if (ITERS8) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            y0 = __fadd2_rn(y0, __fmul2_rn(a0, __ffma2_rn(y0, m1, __fmul2_rn(y0, y0))));
            y1 = __fadd2_rn(y1, __fmul2_rn(a1, __ffma2_rn(y1, m1, __fmul2_rn(y1, y1))));
        }
    } else {
        for (int i = 0; i < iters; ++i) {
            y0 = __fadd2_rn(y0, __fmul2_rn(a0, __ffma2_rn(y0, m1, __fmul2_rn(y0, y0))));
            y1 = __fadd2_rn(y1, __fmul2_rn(a1, __ffma2_rn(y1, m1, __fmul2_rn(y1, y1))));
        }
    }
    xv = make_float4(y0.x, y0.y, y1.x, y1.y);
}

template <bool ITERS8>
__device__ __forceinline__ void mm_apply_block(int t, const uint8_t* ring, uint64_t* empty, int& ck,
                                               const MmApplyArgs& ap, float* s_vrow, int h, int w, int b, int Y0, int X0) {
    const int lane = t & 31, warp = t >> 5;
    const int slot = ck % MM_NS;
    const float* st = reinterpret_cast<const float*>(ring + slot * MM_SLOT_BYTES);
    const float* sa = st + ap.aoff;  // [3][ar][ac]
    const int gw = ap.bw >> 2, ps = ap.br * ap.bw, pa = ap.ar * ap.ac;
    const size_t plane = (size_t)ap.H * ap.W;
    float* vr = s_vrow + warp * (3 * MM_VCOLS);
    // low-res columns under this block: cb .. cb + ncol - 1 (the last one is the "+1" neighbour of the last pixel);
    // the tile in shared memory starts at column cb & ~3 and row yb
    const int xlast = min(X0 + ap.bw, ap.W) - 1;
    const int cb = tap_align_corners(X0, ap.sw, w).i0;
    const int ncol = tap_align_corners(xlast, ap.sw, w).i0 - cb + 2;
    const int xb = cb & ~3;
    const int yb = tap_align_corners(Y0, ap.sh, h).i0;
    const int nrows = min(ap.br, ap.H - Y0);
    const int npass = (gw + 31) >> 5;                    // 1 or 2
    for (int r = warp; r < nrows; r += 8) {
        const int Y = Y0 + r;
        // ---- vertical lerp of the low-res rows under image row Y into the warp's strip ----
        const Tap ty = tap_align_corners(Y, ap.sh, h);
        const float* a0 = sa + (ty.i0 - yb) * ap.ac;
        const float* a1 = sa + (ty.i1 - yb) * ap.ac;
        __syncwarp();  // the previous row's strip has been read by every lane
        for (int col = lane; col < ncol; col += 32) {
            const int cc = min(cb + col, w - 1) - xb;
#pragma unroll
            for (int c = 0; c < 3; ++c) vr[c * MM_VCOLS + col] = mm_lerp(ty.w0, a0[c * pa + cc], ty.w1, a1[c * pa + cc]);
        }
        __syncwarp();
        float* erow = ap.enh + (size_t)b * 3 * plane + (size_t)Y * ap.W + X0;
        float* arow = ap.adj ? ap.adj + (size_t)b * 3 * plane + (size_t)Y * ap.W + X0 : nullptr;
#pragma unroll 1
        for (int pss = 0; pss < npass; ++pss) {
            // ---- x values -> registers; the warp's last pass over the stage releases the slot ----
            const int c4 = lane + 32 * pss;
            const int X = X0 + 4 * c4;
            const bool ok = c4 < gw && X < ap.W;
            if (!ok) continue;
            // ---- horizontal lerp + curve + stores ----
            // tap_align_corners without its clamps: src = rn(sw * X) <= (w - 1)(1 + 2^-23)^2 < w, so the index never
            // exceeds w - 1 and the weight stays in [0, 1) -- the same bits with 6 instead of 12 instructions a pixel
            Tap t0, t1, t2, t3;
            auto tap = [&](Tap& tp, int xo) {
                const float src = __fmul_rn(ap.sw, (float)(X + xo));
                tp.i0 = (int)src;
                tp.w1 = __fsub_rn(src, (float)tp.i0);
                tp.w0 = __fsub_rn(1.f, tp.w1);
            };
            tap(t0, 0), tap(t1, 1), tap(t2, 2), tap(t3, 3);
            const int d0 = t0.i0 - cb, d1 = t1.i0 - cb, d2 = t2.i0 - cb, d3 = t3.i0 - cb;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float* v = vr + c * MM_VCOLS;
                float4 xvc = *reinterpret_cast<const float4*>(st + c * ps + r * ap.bw + 4 * c4);
                float4 av;
                av.x = mm_lerp(t0.w0, v[d0], t0.w1, v[d0 + 1]);
                av.y = mm_lerp(t1.w0, v[d1], t1.w1, v[d1 + 1]);
                av.z = mm_lerp(t2.w0, v[d2], t2.w1, v[d2 + 1]);
                av.w = mm_lerp(t3.w0, v[d3], t3.w1, v[d3 + 1]);
                mm_curve4<ITERS8>(xvc, av, ap.iters);
                st_stream4(erow + c * plane + 4 * c4, xvc);
                if (arow) st_stream4(arow + c * plane + 4 * c4, av);
            }
        }
    }
    __syncwarp();  // every lane's shared loads of this stage have returned
    if (lane == 0) tc::mbar_arrive(&empty[slot]);
    ++ck;
}

constexpr int MM_CLAIM = 4;  // blocks of the full-resolution pass claimed per atomic (= ring depth)

// FUSED = false: the net only (b200_dcepp_lowres); without the block path in the consumer loop the tile code keeps
// all its values in the 96 registers.
template <bool FUSED, bool TRACE>
__global__ void __launch_bounds__(MM_THREADS, 1)
dcepp_net_kernel(const MmNetArgs a, const MmApplyArgs ap, const __grid_constant__ CUtensorMap tm1,
                 const __grid_constant__ CUtensorMap tm2, const __grid_constant__ CUtensorMap tm3,
                 const __grid_constant__ CUtensorMap tm4, const __grid_constant__ CUtensorMap tm5,
                 const __grid_constant__ CUtensorMap tmx, const __grid_constant__ CUtensorMap tma, int B, int h, int w,
                 int* __restrict__ sync) {
    extern __shared__ __align__(128) uint8_t mm_smem_all[];
    const int half = threadIdx.x >= MM_HALF ? 1 : 0;  // two independent halves: own ring, barriers, producer
    uint8_t* mm_smem = mm_smem_all + half * MM_HALF_SMEM;
    uint8_t* ring = mm_smem;
    uint64_t* full = reinterpret_cast<uint64_t*>(mm_smem + MM_NS * MM_SLOT_BYTES);
    uint64_t* empty = full + MM_NS;
    // item descriptors, producer -> consumers: {item number q (-1 = end), kind + 8 * frame, y0, x0};
    // kind 0..5 = tile of e_conv2..7, 6 = block of the full-resolution pass
    volatile int4* s_tq = reinterpret_cast<volatile int4*>(empty + MM_NS);
    int* s_cnt = reinterpret_cast<int*>(mm_smem + MM_NS * MM_SLOT_BYTES + 2 * MM_NS * 8 + 64);  // [4] warps done with a tile
    float* s_vrow = reinterpret_cast<float*>(mm_smem + MM_NS * MM_SLOT_BYTES + 2 * MM_NS * 8 + 64 + 128);

    const int t = threadIdx.x - half * MM_HALF, lane = t & 31, warp = t >> 5;
    const int ntx = (w + MM_TW - 1) / MM_TW, nty = (h + MM_TH - 1) / MM_TH;
    const int tpf = ntx * nty, total = 6 * B * tpf;
    const int wp = mm_pitch(w);
    int* flags = sync + mm_flag0(B);
    const int tpl = B * tpf;  // flags of one layer

    if (t == 0) {
        for (int i = 0; i < MM_NS; ++i) {
            tc::mbar_init(&full[i], 1);   // the producer's expect_tx arrive (+ the stage's bytes)
            tc::mbar_init(&empty[i], 8);  // one arrive per consumer warp
            s_cnt[i] = 0;
        }
        tc::mbar_fence_init();
    }
    __syncthreads();

    if (warp == 8) {
        // ================= producer warp =================
        if (lane == 0) {
            tc::tma_prefetch_desc(&tm1);
            tc::tma_prefetch_desc(&tm2);
            tc::tma_prefetch_desc(&tm3);
            tc::tma_prefetch_desc(&tm4);
            tc::tma_prefetch_desc(&tm5);
            if (ap.total > 0) {
                tc::tma_prefetch_desc(&tmx);
                tc::tma_prefetch_desc(&tma);
            }
        }
        // which of this SM's two halves prefer blocks of the full-resolution pass
        const bool prefer_blocks = ap.total > 0 && (1 - half) < ap.prefer;
        const int bpf = ap.nbx * ap.nby;        // blocks per frame
        const int G = ap.total > 0 ? ap.group : B;  // frames per tile-ordering group (the last group may be short)
        // Blocks are claimed MM_CLAIM at a time with one atomicAdd, whether or not their frame is complete yet; a claimed
        // block whose frame is not ready stays pending while this producer goes on with tiles, so nobody ever waits for
        // a block while tiles are left (and once they are all claimed, their claimers finish them without help).
        bool tiles_left = true, blocks_left = ap.total > 0;
        int pend_i = 0, pend_n = 0, ready_frame = -1;
        int pk = 0, tk = 0;
        for (;;) {
            int q = -1, kind = -1;
            if ((blocks_left || pend_n > 0) && (prefer_blocks || !tiles_left)) {
                if (pend_n == 0) {
                    if (lane == 0) q = atomicAdd(sync + 1, MM_CLAIM);
                    q = __shfl_sync(0xffffffffu, q, 0);
                    if (q >= ap.total)
                        blocks_left = false;
                    else
                        pend_i = q, pend_n = min(MM_CLAIM, ap.total - q);
                }
                if (pend_n > 0) {
                    const int fb = pend_i / bpf;
                    if (fb != ready_frame) {
                        int done = 0;
                        if (lane == 0) done = mm_ld_acquire(sync + MM_FDONE0 + fb);
                        done = __shfl_sync(0xffffffffu, done, 0);
                        if (done >= tpf) ready_frame = fb;
                    }
                    if (fb == ready_frame) {
                        q = pend_i++;
                        --pend_n;
                        kind = 6;
                    }
                }
            }
            if (kind < 0 && tiles_left) {
                if (lane == 0) q = atomicAdd(sync, 1);
                q = __shfl_sync(0xffffffffu, q, 0);
                if (q < total)
                    kind = 0;
                else
                    tiles_left = false;
            }
            if (kind < 0) {
                if (!tiles_left && !blocks_left && pend_n == 0) break;
                if (!tiles_left) {  // the pending blocks belong to a frame whose last tiles are still in flight elsewhere
                    __nanosleep(256);
                    if (*reinterpret_cast<volatile int*>(sync + 2)) break;
                }
                continue;
            }
            if (kind == 6) {
                // ---- one block of the full-resolution pass: a single stage ----
                const int b = q / bpf, r = q - b * bpf;
                const int by = r / ap.nbx, bx = r - by * ap.nbx;
                const int slot = pk % MM_NS;
                if (pk >= MM_NS) mm_mbar_wait(&empty[slot], ((pk / MM_NS) - 1) & 1, sync);
                if (lane == 0) {
                    volatile int4* d = &s_tq[tk & 3];
                    d->x = q, d->y = 6 + 8 * b, d->z = by * ap.br, d->w = bx * ap.bw;
                    // the low-res map of frame b was written through the generic proxy (acquired above)
                    tc::fence_proxy_async_all();
                    tc::mbar_arrive_expect_tx(&full[slot], (uint32_t)(12 * (ap.br * ap.bw + ap.ar * ap.ac)));
                    const uint32_t dst = tc::smem_u32(ring + slot * MM_SLOT_BYTES);
                    tc::tma_load_3d(dst, &tmx, bx * ap.bw, by * ap.br, 3 * b, &full[slot]);
                    tc::tma_load_3d(dst + 4 * ap.aoff, &tma, tap_align_corners(bx * ap.bw, ap.sw, w).i0 & ~3,
                                    tap_align_corners(by * ap.br, ap.sh, h).i0, 3 * b, &full[slot]);
                }
                __syncwarp();
                ++pk;
                ++tk;
                continue;
            }
            // ---- one tile of the net.  Order: frame group (G frames), layer, frame, tile row, tile column ----
            const int gsz = 6 * G * tpf;              // tiles of a full group
            const int grp = q / gsz, rg = q - grp * gsz;
            const int gfr = min(G, B - grp * G);      // frames of this group
            const int l = rg / (gfr * tpf), r = rg - l * (gfr * tpf);
            const int b = grp * G + r / tpf, r2 = r % tpf;
            const int tyi = r2 / ntx, txi = r2 - tyi * ntx;
            if (l > 0) {
                if (lane < 9) {  // the nine layer l-1 tiles under this tile's halo
                    const int ny = tyi + lane / 3 - 1, nx = txi + lane % 3 - 1;
                    if (ny >= 0 && ny < nty && nx >= 0 && nx < ntx) {
                        const int* f = flags + (size_t)(l - 1) * tpl + b * tpf + ny * ntx + nx;
                        if (mm_ld_acquire(f) == 0) {
                            const long long t0 = clock64();
                            int n = 0;
                            while (mm_ld_acquire(f) == 0) {
                                __nanosleep(32);
                                if (mm_spin_expired(t0, n, sync)) break;
                            }
                        }
                    }
                }
                __syncwarp();
            }
            // sources (zero_dcepp.py:191-196): x1 | x2 | x3 | cat(x3,x4) | cat(x2,x5) | cat(x1,x6 = x4's buffer)
            const CUtensorMap* mA = l == 0 || l == 5 ? &tm1 : (l == 1 || l == 4 ? &tm2 : &tm3);
            const CUtensorMap* mB = l == 4 ? &tm5 : &tm4;
            const int nst = l < 3 ? 2 : 4;
            const uint32_t* wl = nullptr;  // selected with constant indices: the kernel parameters stay in the constant
#pragma unroll                             // bank (a dynamic index would copy them to the stack)
            for (int i = 0; i < 6; ++i)
                if (i == l) wl = a.wl[i];
            const uint32_t blob = (l == 5 ? mm_blob_words(6) : mm_blob_words(1)) * 4;
            for (int i = 0; i < nst; ++i, ++pk) {
                const int slot = pk % MM_NS;
                if (pk >= MM_NS) mm_mbar_wait(&empty[slot], ((pk / MM_NS) - 1) & 1, sync);
                if (lane == 0) {
                    if (i == 0) {
                        volatile int4* d = &s_tq[tk & 3];
                        d->x = l * tpl + b * tpf + r2, d->y = l + 8 * b, d->z = tyi * MM_TH, d->w = txi * MM_TW;
                        tc::fence_proxy_async_all();  // the acquired tiles were written through the generic proxy
                    }
                    tc::mbar_arrive_expect_tx(&full[slot], MM_STAGE_BYTES + blob);
                    const uint32_t dst = tc::smem_u32(ring + slot * MM_SLOT_BYTES);
                    tc::tma_load_3d(dst, i < 2 ? mA : mB, txi * MM_TW - MM_XOFF, tyi * MM_TH - 1, b * 16 + (i & 1) * 8,
                                    &full[slot]);
                    tc::bulk_load(dst + MM_STAGE_BYTES, reinterpret_cast<const uint8_t*>(wl) + i * blob, blob, &full[slot]);
                }
                __syncwarp();
            }
            ++tk;
        }
        // end marker: an empty stage whose item number is -1
        {
            const int slot = pk % MM_NS;
            if (pk >= MM_NS) mm_mbar_wait(&empty[slot], ((pk / MM_NS) - 1) & 1, sync);
            if (lane == 0) {
                s_tq[tk & 3].x = -1;
                tc::mbar_arrive(&full[slot]);
            }
        }
        return;
    }

    // ================= consumer warps =================
    int ck = 0, tk = 0, nk = 0;  // stages, items, net tiles consumed so far
    uint32_t tr_wait = 0, tr_start = 0, tr_id = 0;
    const bool tracing = TRACE && t == 0;
    if (tracing) {
        uint32_t smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        tr_id = ((2 * blockIdx.x + half) << 16) | ((smid & 0xfff) << 4);
    }
    auto trace_item = [&](int kind) {
        const uint32_t i = atomicAdd(ap.trace, 1u) + 1;
        if ((int)i < ap.trace_cap)
            reinterpret_cast<uint4*>(ap.trace)[i] = make_uint4(tr_id | (uint32_t)kind, tr_wait, tr_start, mm_now_ns());
    };
    for (;;) {
        if (tracing) tr_wait = mm_now_ns();
        mm_mbar_wait(&full[ck % MM_NS], (ck / MM_NS) & 1, sync);  // the item's first stage (or the end marker)
        if (tracing) tr_start = mm_now_ns();
        const int q = s_tq[tk & 3].x;
        if (q < 0) break;
        const int lb = s_tq[tk & 3].y, y0 = s_tq[tk & 3].z, x0 = s_tq[tk & 3].w;
        const int l = lb & 7, b = lb >> 3;
        if (FUSED && l == 6) {
            if (ap.iters == 8)
                mm_apply_block<true>(t, ring, empty, ck, ap, s_vrow, h, w, b, y0, x0);
            else
                mm_apply_block<false>(t, ring, empty, ck, ap, s_vrow, h, w, b, y0, x0);
            ++tk;
            if (tracing) trace_item(6);
            continue;
        }
        float* ds = nullptr;  // selected with constant indices: the kernel parameters stay in the constant bank
#pragma unroll
        for (int i = 0; i < 6; ++i)
            if (i == l) ds = a.dst[i];
        if (l < 3)
            mm_tile<1, 4, false>(t, ring, full, empty, ck, ds, h, w, wp, b, y0, x0);
        else if (l < 5)
            mm_tile<2, 4, false>(t, ring, full, empty, ck, ds, h, w, wp, b, y0, x0);
        else
            mm_tile<2, 1, true>(t, ring, full, empty, ck, ds, h, w, wp, b, y0, x0, ap.total > 0 ? ap.wq : w);
        // Publication without a CTA barrier: every warp counts itself done (acq_rel at CTA scope behind its own stores),
        // the LAST one resets the counter and publishes the tile (release at GPU scope, cumulative over what it acquired);
        // the other seven go straight on to the next item.  At most two tiles are in flight per half (4 ring slots, >= 2
        // stages per tile), so four counters cannot be lapped.
        __syncwarp();
        if (lane == 0) {
            int old;
            asm volatile("atom.acq_rel.cta.shared.add.s32 %0, [%1], 1;"
                         : "=r"(old)
                         : "r"(tc::smem_u32(s_cnt + (nk & 3)))
                         : "memory");
            if (old == 7) {
                s_cnt[nk & 3] = 0;
                if (l < 5)
                    mm_st_release(flags + q, 1);
                else if (ap.total > 0)
                    mm_red_release_add(sync + MM_FDONE0 + b, 1);  // one more layer-7 tile of frame b
            }
        }
        ++nk;
        ++tk;
        if (tracing) trace_item(l);
    }
}

// [B][16 pairs][h][wp] float2 as a 3-D tensor of 8-byte elements (x, y, pair + 16*b); box = one stage
static int mm_make_map(CUtensorMap* m, const void* base, int B, int h, int w) {
    tma_encode_fn enc = tma_encoder();
    if (!enc) return B200_ERR_LAUNCH;
    const int wp = mm_pitch(w);
    const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)16 * B};
    const cuuint64_t strides[2] = {(cuuint64_t)wp * 8, (cuuint64_t)h * wp * 8};
    const cuuint32_t box[3] = {MM_HC, MM_HR, MM_STAGE_PAIRS};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<void*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? B200_OK : B200_ERR_LAUNCH;
}

// planar fp32 [planes][rows][pitch] as a 3-D tensor (x, y, plane); `cols` <= pitch are the valid columns
static int mm_make_plane_map(CUtensorMap* m, const void* base, int planes, int rows, int cols, int pitch, int bx,
                             int by) {
    tma_encode_fn enc = tma_encoder();
    if (!enc) return B200_ERR_LAUNCH;
    const cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)planes};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch * 4, (cuuint64_t)rows * pitch * 4};
    const cuuint32_t box[3] = {(cuuint32_t)bx, (cuuint32_t)by, 3};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? B200_OK : B200_ERR_LAUNCH;
}

// Can the full-resolution pass ride in the net launch?  Fills the block geometry of `ap` (H, W, sh, sw set by the
// caller; w = width of the low-res map).  Needs 128-bit rows (W % 4 == 0), at most MM_VCOLS low-res columns under a
// block, and a stage that holds the block of x plus the low-res tile under it.
static bool mm_plan_blocks(MmApplyArgs& ap, int B, int w) {
    const int H = ap.H, W = ap.W;
    if ((W & 3) != 0 || W < 4) return false;
    // widest block: 64 groups of 4 pixels (two passes of a warp) with <= MM_VCOLS low-res columns below
    int bw_max = 252;
    if (ap.sw > 0.f) {
        const long long lim = (long long)((MM_VCOLS - 4) / ap.sw);
        if (lim < bw_max) bw_max = (int)lim & ~3;
    }
    if (bw_max < 4) return false;
    ap.nbx = (W + bw_max - 1) / bw_max;
    ap.bw = (((W + ap.nbx - 1) / ap.nbx) + 3) & ~3;
    if (ap.bw > bw_max) ap.bw = bw_max;
    // low-res tile: columns cb & ~3 .. i0(last pixel) + 1, rows i0(first row) .. i0(last row) + 1 (safe upper bounds)
    ap.ac = ((int)(ap.sw * (float)(ap.bw - 1)) + 3 + 3 + 3) & ~3;
    if (ap.ac > 256) return false;
    ap.wq = (w + 3) & ~3;
    auto tile = [&](int br) {
        ap.br = br;
        ap.ar = (int)(ap.sh * (float)(br - 1)) + 3;
        ap.aoff = (br * ap.bw * 3 + 31) & ~31;  // the tile starts 128-byte aligned
        return ap.ar <= 256 && (ap.aoff + 3 * ap.ar * ap.ac) * 4 <= MM_STAGE_BYTES;
    };
    int br = MM_STAGE_BYTES / (12 * ap.bw);
    if (br > 256) br = 256;
    if (br > H) br = H;
    while (br >= 1 && !tile(br)) --br;
    if (br < 1) return false;
    if (br > 8 && (br & 7) && tile(br & ~7)) br &= ~7;  // whole rounds of the 8 consumer warps
    tile(br);
    ap.nbx = (W + ap.bw - 1) / ap.bw;
    ap.nby = (H + ap.br - 1) / ap.br;
    const long long total = (long long)B * ap.nbx * ap.nby;
    if (total > 0x3fffffffLL) return false;
    ap.total = (int)total;
    return true;
}

// x[0..4] = x1 x2 x3 x4 x5 (x6 reuses x4's buffer); `ap.total` > 0 adds the full-resolution pass over `xin`
static int launch_mm_net(const MmNetArgs& a, MmApplyArgs ap, const float* xin, float* const* x, int B, int h, int w,
                         int* sync, cudaStream_t s) {
    auto kern = ap.total > 0 ? (ap.trace ? dcepp_net_kernel<true, true> : dcepp_net_kernel<true, false>)
                             : dcepp_net_kernel<false, false>;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, MM_NET_SMEM) != cudaSuccess) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;
    }
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, MM_THREADS, MM_NET_SMEM) != cudaSuccess || occ < 1) {
        (void)cudaGetLastError();
        return B200_ERR_LAUNCH;  // the wait-on-earlier-tiles scheme needs every CTA resident
    }
    int dev = 0, sms = B200_NUM_SMS;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    CUtensorMap m[7];
    memset(m, 0, sizeof(m));
    for (int i = 0; i < 5; ++i) {
        const int rc = mm_make_map(&m[i], x[i], B, h, w);
        if (rc) return rc;
    }
    const long long tpf = (long long)((h + MM_TH - 1) / MM_TH) * ((w + MM_TW - 1) / MM_TW);
    const long long total = 6LL * B * tpf + ap.total;
    long long grid = (long long)sms * occ;  // x 2 halves = workers
    if (2 * grid > total) grid = (total + 1) / 2;
    if (ap.total > 0) {
        int rc = mm_make_plane_map(&m[5], xin, 3 * B, ap.H, ap.W, ap.W, ap.bw, ap.br);
        if (rc) return rc;
        if ((rc = mm_make_plane_map(&m[6], a.dst[5], 3 * B, h, w, ap.wq, ap.ac, ap.ar))) return rc;
        // frames per tile-ordering group: enough tiles per layer that a tile's nine predecessors (claimed one layer =
        // group * tpf claims earlier) are finished rather than in flight when it is claimed
        // (measured, 16 x 1080p / 8 x 2160p: ~3.25 tiles per worker and layer is the knee -- smaller groups start the
        // full-resolution pass earlier but stall producers on predecessors and starve the tile phase of parallelism)
        long long g = (13 * 2 * grid + 4 * tpf - 1) / (4 * tpf);
        if (const char* e = getenv("B200DCE_DCEPP_GROUP"))  // tuning aid; an empty / non-positive value means "unset"
            if (atoi(e) > 0) g = atoi(e);
        g = g < 1 ? 1 : (g > B ? B : g);
        const long long ngroups = (B + g - 1) / g;
        ap.group = (int)((B + ngroups - 1) / ngroups);  // equal groups: 8 frames with g = 5 run as 4 + 4, not 5 + 3
    }
    kern<<<(unsigned)grid, MM_THREADS, MM_NET_SMEM, s>>>(a, ap, m[0], m[1], m[2], m[3], m[4], m[5], m[6], B, h, w, sync);
    return launch_status();
}

}  // namespace b200

// GuidedFilter(radius, eps).forward(image = x, guide = y) at full resolution, as GCENet applies it to the result of its
// curve loop (src/mon/vision/filtering/guided_filter.py:99-146, src/mon/vision/enhance/llie/gcenet/gcenet.py:391,485):
//   N = box(1); mean_x = box(x)/N; mean_y = box(y)/N; cov = box(x*y)/N - mean_x*mean_y; var = box(x*x)/N - mean_x^2
//   A = cov / (var + eps); b = mean_y - A*mean_x; out = (box(A)/N) * x + box(b)/N
// with box(t) = diff_y(diff_x(t.cumsum(2), r).cumsum(3), r) (box_filter.py:26-117): window sums as DIFFERENCES OF
// ROUNDED RUNNING SUMS.  ATen's CPU cumsum accumulates fp32 input in double and rounds every prefix to fp32, and the
// kernels below do the same (sums of <= 2^12 fp32 values >= 2^-18 are exact in double in any order), because A is a
// quotient of two such differences and amplifies their rounding (see dcev_tail.cu / DESIGN.md).
//
// Passes over planes of B*C*H*W fp32 (default): chunk totals of the columns | column pass: D1 = diff_x(cumsum(2)) from two
// running prefixes, no prefix plane stored | row pass: cumsum(3) of a row in shared memory, diff_y, and the stage's
// arithmetic (A, b / the output) | the same three for A, b.  ~80 B per plane pixel instead of the ~250 B of the first
// version (B200DCE_GF_FUSED=0: column scan, row-scan totals, row scan, A / b, ... as separate HBM round trips; kept
// for A/B, bit-identical).  16 x 1080p: 1.9 ms (was 7.2), one frame 0.17 ms (was 0.56).
#include <math.h>
#include <stdlib.h>

#include "common.cuh"
#include "guided_filter.cuh"

namespace b200 {

__host__ __device__ inline size_t gfl_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// diff_x / diff_y of box_filter.py:26-66 on a prefix array `p` with element stride `st`
__device__ __forceinline__ float gfl_diff(const float* __restrict__ p, int i, int n, int r, size_t st) {
    if (i <= r) return p[(size_t)(i + r) * st];
    const float lo = p[(size_t)(i - r - 1) * st];
    const float hi = i < n - r ? p[(size_t)(i + r) * st] : p[(size_t)(n - 1) * st];
    return __fsub_rn(hi, lo);
}

// Column scan in chunks of GFL_CH rows (a frame's columns alone are too few threads to hide the latency of 1080
// dependent steps): pass 1 = every chunk's total, pass 2 = the chunk's running sums on top of the totals above it.
// Both accumulate in double: the partial sums are exact (see the header), so the chunked order gives the prefixes of
// ATen's sequential loop.  MODE 0: sources x (plane index px) and y -> maps x, y, x*y, x*x; MODE 1: sources A, b -> A, b.
// grid (ceil(W / 256), planes, chunks); out planes [map][plane][H][W]; tot [map][plane][chunk][W] doubles
constexpr int GFL_CH = 72;

template <int MODE, bool TOTALS>
__global__ void __launch_bounds__(256) gfl_colscan_kernel(const float* __restrict__ s0, const float* __restrict__ s1,
                                                          float* __restrict__ out, double* __restrict__ tot, int C, int Cx,
                                                          int H, int W, size_t map_stride) {
    constexpr int NM = MODE == 0 ? 4 : 2;
    const int w = blockIdx.x * 256 + threadIdx.x;
    if (w >= W) return;
    const int pl = blockIdx.y, ch = blockIdx.z, planes = gridDim.y, nch = gridDim.z;
    const size_t hw = (size_t)H * W;
    const int px = MODE == 0 ? (pl / C) * Cx + (Cx == 1 ? 0 : pl % C) : pl;
    const float* a = s0 + (size_t)px * hw + w;
    const float* b = s1 + (size_t)pl * hw + w;
    float* o = out + (size_t)pl * hw + w;
    double acc[NM];
#pragma unroll
    for (int m = 0; m < NM; ++m) acc[m] = 0.0;
    if (!TOTALS) {
        for (int c = 0; c < ch; ++c)
#pragma unroll
            for (int m = 0; m < NM; ++m) acc[m] += tot[(((size_t)m * planes + pl) * nch + c) * W + w];
    }
    auto step = [&](float av, float bv, size_t off) {
        float v[NM];
        v[0] = av, v[1] = bv;
        if (MODE == 0) v[2 % NM] = __fmul_rn(av, bv), v[3 % NM] = __fmul_rn(av, av);
#pragma unroll
        for (int m = 0; m < NM; ++m) {
            acc[m] += (double)v[m];
            if (!TOTALS) o[m * map_stride + off] = (float)acc[m];
        }
    };
    const int chrows = gridDim.z == 1 ? H : GFL_CH;  // one chunk = the whole column
    const int h1 = min(H, (ch + 1) * chrows);
    int h = ch * chrows;
    for (; h + 8 <= h1; h += 8) {
        float av[8], bv[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) av[k] = a[(size_t)(h + k) * W], bv[k] = b[(size_t)(h + k) * W];
#pragma unroll
        for (int k = 0; k < 8; ++k) step(av[k], bv[k], (size_t)(h + k) * W);
    }
    for (; h < h1; ++h) step(a[(size_t)h * W], b[(size_t)h * W], (size_t)h * W);
    if (TOTALS) {
#pragma unroll
        for (int m = 0; m < NM; ++m) tot[(((size_t)m * planes + pl) * nch + ch) * W + w] = acc[m];
    }
}

// Row scan of diff_x(P1) in chunks of GFL_CW columns, two passes like the column scan (chunk totals, then the running
// sums on top of the totals to the left).  grid (ceil(chunks / 8), maps * planes, ceil(H / 32)): the 8 warps of a CTA
// take 8 adjacent column chunks of the same 32 rows (4 KB contiguous per row and CTA); a warp moves 32 x 32 blocks
// through shared memory: coalesced 128-byte rows in and out, conflict-free scan along the rows.  Measured 1.5-2.2 TB/s
// per pass at 16 x 1080p (~29 instructions per 32 elements: address arithmetic, fp32 <-> fp64 conversions, the two
// shared-memory transposes): the slowest passes of the filter, not tuned further.
// tot [map * planes][H][chunks] doubles
constexpr int GFL_CW = 128;

template <bool TOTALS>
__global__ void __launch_bounds__(256) gfl_rowscan_kernel(const float* __restrict__ p1, float* __restrict__ p2,
                                                          double* __restrict__ tot, int H, int W, int r, int planes,
                                                          size_t map_stride, int nch) {
    __shared__ float tile[8][32][33];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t hw = (size_t)H * W;
    const size_t base = (size_t)(blockIdx.y / planes) * map_stride + (size_t)(blockIdx.y % planes) * hw;
    const float* src = p1 + base;
    float* dst = p2 + base;
    const int r0 = blockIdx.z * 32;
    const int ch = blockIdx.x * 8 + warp;
    if (ch >= nch) return;
    const int row = r0 + lane;  // the row this lane scans
    double* trow = tot + ((size_t)blockIdx.y * H + min(row, H - 1)) * nch;
    double acc = 0.0;
    if (!TOTALS && row < H)
        for (int c = 0; c < ch; ++c) acc += trow[c];
    const int c1 = min(W, (ch + 1) * GFL_CW);
    for (int c0 = ch * GFL_CW; c0 < c1; c0 += 32) {
        const int w = c0 + lane;
#pragma unroll 16
        for (int k = 0; k < 32; ++k)
            tile[warp][k][lane] = (r0 + k < H && w < c1) ? gfl_diff(src + w, r0 + k, H, r, (size_t)W) : 0.f;
        __syncwarp();
        float o[32];
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            acc += (double)tile[warp][lane][k];
            o[k] = (float)acc;
        }
        __syncwarp();
        if (!TOTALS) {
#pragma unroll
            for (int k = 0; k < 32; ++k) tile[warp][lane][k] = o[k];
            __syncwarp();
#pragma unroll 16
            for (int k = 0; k < 32; ++k)
                if (r0 + k < H && w < c1) dst[(size_t)(r0 + k) * W + w] = tile[warp][k][lane];
            __syncwarp();
        }
    }
    if (TOTALS && row < H) trow[ch] = acc;
}

__device__ __forceinline__ float gfl_count(int h, int w, int H, int W, int r) {  // box(1): exact
    const int nh = min(h + r, H - 1) - max(h - r, 0) + 1, nw = min(w + r, W - 1) - max(w - r, 0) + 1;
    return (float)(nh * nw);
}

// window sums = diff_y(P2) -> A, b.  p2: [4][planes][H][W]; A, b: [planes][H][W]
__global__ void __launch_bounds__(256) gfl_ab_kernel(const float* __restrict__ p2, float* __restrict__ A,
                                                     float* __restrict__ Bc, int H, int W, int r, float eps,
                                                     size_t map_stride, long long total) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / W;
        const int w = (int)(i - row * W), h = (int)(row % H);
        const float n = gfl_count(h, w, H, W, r);
        float b[4];
#pragma unroll
        for (int m = 0; m < 4; ++m) b[m] = gfl_diff(p2 + m * map_stride + (size_t)row * W, w, W, r, 1);
        const float mean_x = __fdiv_rn(b[0], n), mean_y = __fdiv_rn(b[1], n);
        const float cov = __fsub_rn(__fdiv_rn(b[2], n), __fmul_rn(mean_x, mean_y));
        const float var = __fsub_rn(__fdiv_rn(b[3], n), __fmul_rn(mean_x, mean_x));
        const float a = __fdiv_rn(cov, __fadd_rn(var, eps));
        A[i] = a;
        Bc[i] = __fsub_rn(mean_y, __fmul_rn(a, mean_x));
    }
}

// out = (box(A)/N) * x + box(b)/N.  p2: [2][planes][H][W]
__global__ void __launch_bounds__(256) gfl_out_kernel(const float* __restrict__ p2, const float* __restrict__ x,
                                                      float* __restrict__ out, int C, int Cx, int H, int W, int r,
                                                      size_t map_stride, long long total) {
    const long long hw = (long long)H * W;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / W;
        const int w = (int)(i - row * W), h = (int)(row % H);
        const long long pl = row / H;
        const float n = gfl_count(h, w, H, W, r);
        const float ma = __fdiv_rn(gfl_diff(p2 + (size_t)row * W, w, W, r, 1), n);
        const float mb = __fdiv_rn(gfl_diff(p2 + map_stride + (size_t)row * W, w, W, r, 1), n);
        const long long px = (pl / C) * Cx + (Cx == 1 ? 0 : pl % C);
        out[i] = __fadd_rn(__fmul_rn(ma, x[px * hw + (long long)h * W + w]), mb);
    }
}


// ---- the same arithmetic in four passes instead of six (default; B200DCE_GF_FUSED=0 keeps the six above for A/B) ----
// Column pass: D1 = diff_x(t.cumsum(2)) without storing the prefixes.  A thread walks its column chunk with TWO exact
// double running sums, the prefix through row h + r and the prefix through row h - r - 1 (both start from the chunk
// totals of gfl_colscan_kernel<MODE, true>; the trailing rows are re-read from L1 / L2), rounds each to fp32 as the
// cumsum does and subtracts.  grid (ceil(W / 256), planes, chunks); out planes [map][plane][H][W]
template <int MODE>
__global__ void __launch_bounds__(256) gfl_coldiff_kernel(const float* __restrict__ s0, const float* __restrict__ s1,
                                                          float* __restrict__ out, const double* __restrict__ tot, int C,
                                                          int Cx, int H, int W, int r, size_t map_stride) {
    constexpr int NM = MODE == 0 ? 4 : 2;
    const int w = blockIdx.x * 256 + threadIdx.x;
    if (w >= W) return;
    const int pl = blockIdx.y, ch = blockIdx.z, planes = gridDim.y, nch = gridDim.z;
    const size_t hw = (size_t)H * W;
    const int px = MODE == 0 ? (pl / C) * Cx + (Cx == 1 ? 0 : pl % C) : pl;
    const float* a = s0 + (size_t)px * hw + w;
    const float* b = s1 + (size_t)pl * hw + w;
    float* o = out + (size_t)pl * hw + w;
    double lead[NM], trail[NM];
#pragma unroll
    for (int m = 0; m < NM; ++m) lead[m] = 0.0;
    for (int c = 0; c < ch; ++c)
#pragma unroll
        for (int m = 0; m < NM; ++m) lead[m] += tot[(((size_t)m * planes + pl) * nch + c) * W + w];
#pragma unroll
    for (int m = 0; m < NM; ++m) trail[m] = lead[m];
    auto maps = [&](float av, float bv, float (&v)[NM]) {
        v[0] = av, v[1] = bv;
        if (MODE == 0) v[2 % NM] = __fmul_rn(av, bv), v[3 % NM] = __fmul_rn(av, av);
    };
    const int chrows = gridDim.z == 1 ? H : GFL_CH;  // one chunk = the whole column
    const int h0 = ch * chrows, h1 = min(H, h0 + chrows);
    for (int k = h0; k < min(h0 + r, H); ++k) {  // lead = prefix through row h0 + r - 1
        float v[NM];
        maps(a[(size_t)k * W], b[(size_t)k * W], v);
#pragma unroll
        for (int m = 0; m < NM; ++m) lead[m] += (double)v[m];
    }
    for (int k = max(h0 - r - 1, 0); k < h0; ++k) {  // trail = prefix through row h0 - r - 2 (exact: see the header)
        float v[NM];
        maps(a[(size_t)k * W], b[(size_t)k * W], v);
#pragma unroll
        for (int m = 0; m < NM; ++m) trail[m] -= (double)v[m];
    }
    for (int h = h0; h < h1; h += 4) {
        float la[4], lb[4], ta[4], tb[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {  // rows past the ends are read clamped and not added
            const size_t lo = (size_t)min(h + k + r, H - 1) * W, to = (size_t)max(h + k - r - 1, 0) * W;
            la[k] = a[lo], lb[k] = b[lo], ta[k] = a[to], tb[k] = b[to];
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = h + k;
            if (i >= h1) break;
            float v[NM], u[NM];
            maps(la[k], lb[k], v);
            maps(ta[k], tb[k], u);
            const bool add_lead = i + r < H, has_lo = i - r - 1 >= 0;
#pragma unroll
            for (int m = 0; m < NM; ++m) {
                if (add_lead) lead[m] += (double)v[m];
                if (has_lo) trail[m] += (double)u[m];
                // box_filter.py:26-45: rows <= r take the prefix itself (hi - 0 is the same value)
                o[m * map_stride + (size_t)i * W] = __fsub_rn((float)lead[m], has_lo ? (float)trail[m] : 0.f);
            }
        }
    }
}

// The same pass for the window radii that occur (GCENet: radius 3 -> r = 1), with ONE running prefix: the rounded prefix
// of row h + r is the one row h + 2r + 1 subtracts, so the last 2r + 1 rounded prefixes stay in a register ring (static
// indices: the row loop is unrolled over whole turns of the ring) -- half the loads, additions and fp32 <-> fp64
// conversions (16 lanes per SM) of the two-prefix form above; the same values.
template <int MODE, int R>
__global__ void __launch_bounds__(256) gfl_coldiff_ring_kernel(const float* __restrict__ s0, const float* __restrict__ s1,
                                                               float* __restrict__ out, const double* __restrict__ tot,
                                                               int C, int Cx, int H, int W, size_t map_stride) {
    constexpr int NM = MODE == 0 ? 4 : 2;
    constexpr int N = 2 * R + 1;                 // ring entries
    constexpr int U = N >= 7 ? 1 : (N == 5 ? 2 : 3);  // ring turns per loop trip (7 .. 10 rows of loads in flight)
    const int w = blockIdx.x * 256 + threadIdx.x;
    if (w >= W) return;
    const int pl = blockIdx.y, ch = blockIdx.z, planes = gridDim.y, nch = gridDim.z;
    const size_t hw = (size_t)H * W;
    const int px = MODE == 0 ? (pl / C) * Cx + (Cx == 1 ? 0 : pl % C) : pl;
    const float* a = s0 + (size_t)px * hw + w;
    const float* b = s1 + (size_t)pl * hw + w;
    float* o = out + (size_t)pl * hw + w;
    double base[NM], d[NM];
#pragma unroll
    for (int m = 0; m < NM; ++m) base[m] = 0.0;
    for (int c = 0; c < ch; ++c)
#pragma unroll
        for (int m = 0; m < NM; ++m) base[m] += tot[(((size_t)m * planes + pl) * nch + c) * W + w];
    auto maps = [&](float av, float bv, float (&v)[NM]) {
        v[0] = av, v[1] = bv;
        if (MODE == 0) v[2 % NM] = __fmul_rn(av, bv), v[3 % NM] = __fmul_rn(av, av);
    };
    const int chrows = gridDim.z == 1 ? H : GFL_CH;  // one chunk = the whole column
    const int h0 = ch * chrows, h1 = min(H, h0 + chrows);
    // ring[m][s] = rounded prefix through row h0 - R - 1 + s (slots of rows < 0 are never read)
    float ring[NM][N];
#pragma unroll
    for (int m = 0; m < NM; ++m) d[m] = base[m];
#pragma unroll
    for (int s2 = R; s2 >= 0; --s2) {  // downwards from the chunk base: P[j - 1] = P[j] - t[j], exact
        const int j = h0 - R - 1 + s2;
        float v[NM];
        maps(a[(size_t)max(j, 0) * W], b[(size_t)max(j, 0) * W], v);
#pragma unroll
        for (int m = 0; m < NM; ++m) {
            ring[m][s2] = (float)d[m];
            if (j >= 0) d[m] -= (double)v[m];
        }
    }
#pragma unroll
    for (int m = 0; m < NM; ++m) d[m] = base[m];
#pragma unroll
    for (int s2 = R + 1; s2 < N; ++s2) {  // upwards: rows h0 .. h0 + R - 1 (the prefix stops growing at row H - 1)
        const int j = h0 - R - 1 + s2;
        float v[NM];
        maps(a[(size_t)min(j, H - 1) * W], b[(size_t)min(j, H - 1) * W], v);
#pragma unroll
        for (int m = 0; m < NM; ++m) {
            if (j < H) d[m] += (double)v[m];
            ring[m][s2] = (float)d[m];
        }
    }
    for (int h = h0; h < h1; h += N * U) {
        float la[N * U], lb[N * U];
#pragma unroll
        for (int k = 0; k < N * U; ++k) {  // rows past the end are read clamped and not added
            const size_t lo = (size_t)min(h + k + R, H - 1) * W;
            la[k] = a[lo], lb[k] = b[lo];
        }
#pragma unroll
        for (int k = 0; k < N * U; ++k) {
            const int i = h + k;
            if (i >= h1) break;
            float v[NM];
            maps(la[k], lb[k], v);
            const bool add_lead = i + R < H, has_lo = i - R - 1 >= 0;
#pragma unroll
            for (int m = 0; m < NM; ++m) {
                if (add_lead) d[m] += (double)v[m];
                const float hi = (float)d[m], lo = ring[m][k % N];
                ring[m][k % N] = hi;
                // box_filter.py:26-45: rows <= r take the prefix itself
                o[m * map_stride + (size_t)i * W] = has_lo ? __fsub_rn(hi, lo) : hi;
            }
        }
    }
}
template <int MODE>
static void gfl_coldiff(const float* s0, const float* s1, float* out, const double* tot, int C, int Cx, int H, int W,
                        int r, size_t map_stride, dim3 grid, cudaStream_t s) {
    switch (r) {
        case 1: gfl_coldiff_ring_kernel<MODE, 1><<<grid, 256, 0, s>>>(s0, s1, out, tot, C, Cx, H, W, map_stride); break;
        case 2: gfl_coldiff_ring_kernel<MODE, 2><<<grid, 256, 0, s>>>(s0, s1, out, tot, C, Cx, H, W, map_stride); break;
        case 3: gfl_coldiff_ring_kernel<MODE, 3><<<grid, 256, 0, s>>>(s0, s1, out, tot, C, Cx, H, W, map_stride); break;
        case 4: gfl_coldiff_ring_kernel<MODE, 4><<<grid, 256, 0, s>>>(s0, s1, out, tot, C, Cx, H, W, map_stride); break;
        default: gfl_coldiff_kernel<MODE><<<grid, 256, 0, s>>>(s0, s1, out, tot, C, Cx, H, W, r, map_stride);
    }
}

// Row pass: one CTA per (image row, plane) runs cumsum(3) of the row's D1 values -- a lane sums its run of E contiguous
// elements in double, a shuffle scan and eight warp totals give every run its exact base, the rounded prefixes go to
// shared memory -- then takes diff_y from there and finishes the stage: MODE 0: the four window sums -> A, b;
// MODE 1: box(A), box(b), x -> out.  No totals pass, no prefix plane in HBM.  grid (H, planes); 8 * 32 * E >= W
// (tried: 512 threads with runs of 4 and every element converted to double once -- 0.90 instead of 0.71 ms)
// RN(x / n) for a window count n (an integer <= 4096), rn = RN(1 / n): q = RN(x rn), the exact remainder x - n q by
// FMA, one correction.  q + rem rn differs from x / n by < 2^-47 relative while x / n stays >= 2^-25 / n away from every
// rounding boundary (x has 24 bits; n times a 25-bit midpoint never has), so the result is the IEEE quotient --
// 3 instructions instead of the ~11 of the general division (reciprocal, range check, slow-path branch).  Tiny
// dividends (the remainder would be subnormal) take the general path.
__device__ __forceinline__ float gfl_div_count(float x, float n, float rn) {
    if (fabsf(x) < 1e-25f) return __fdiv_rn(x, n);
    const float q = __fmul_rn(x, rn);
    return __fmaf_rn(__fmaf_rn(-q, n, x), rn, q);
}
__device__ __forceinline__ double gfl_shfl_up(double v, int o) {
    return __hiloint2double(__shfl_up_sync(0xffffffffu, __double2hiint(v), o),
                            __shfl_up_sync(0xffffffffu, __double2loint(v), o));
}
template <int MODE, int E>
__global__ void __launch_bounds__(256) gfl_rowbox_kernel(const float* __restrict__ d1, float* __restrict__ o0,
                                                         float* __restrict__ o1, const float* __restrict__ x, int C,
                                                         int Cx, int H, int W, int r, float eps, size_t map_stride) {
    constexpr int NM = MODE == 0 ? 4 : 2;
    extern __shared__ __align__(16) unsigned char gfl_smem[];
    double* wtot = reinterpret_cast<double*>(gfl_smem);   // [NM][8]
    float* srow = reinterpret_cast<float*>(wtot + NM * 8);  // [NM][Wp]
    const int Wp = (W + 3) & ~3;
    const int h = blockIdx.x, pl = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t row = ((size_t)pl * H + h) * W;
    const int w0 = threadIdx.x * E;
    const bool vec = (W & 3) == 0;
    float v[NM][E];
    double excl[NM];
#pragma unroll
    for (int m = 0; m < NM; ++m) {
        const float* src = d1 + m * map_stride + row;
        double acc = 0.0;
#pragma unroll
        for (int j = 0; j < E; j += 4) {
            float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
            if (vec) {
                if (w0 + j < W) q = *reinterpret_cast<const float4*>(src + w0 + j);
            } else {
                if (w0 + j < W) q.x = src[w0 + j];
                if (w0 + j + 1 < W) q.y = src[w0 + j + 1];
                if (w0 + j + 2 < W) q.z = src[w0 + j + 2];
                if (w0 + j + 3 < W) q.w = src[w0 + j + 3];
            }
            v[m][j] = q.x, v[m][j + 1] = q.y, v[m][j + 2] = q.z, v[m][j + 3] = q.w;
            acc += (double)q.x, acc += (double)q.y, acc += (double)q.z, acc += (double)q.w;
        }
        double s = acc;  // inclusive scan over the lanes (exact sums: any order gives the sequential loop's prefixes)
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double t = gfl_shfl_up(s, o);
            if (lane >= o) s += t;
        }
        excl[m] = s - acc;
        if (lane == 31) wtot[m * 8 + warp] = s;
    }
    __syncthreads();
#pragma unroll
    for (int m = 0; m < NM; ++m) {
        double acc = excl[m];
        for (int k = 0; k < warp; ++k) acc += wtot[m * 8 + k];
        float* dst = srow + m * Wp;
#pragma unroll
        for (int j = 0; j < E; j += 4) {
            float4 q;
            acc += (double)v[m][j], q.x = (float)acc;
            acc += (double)v[m][j + 1], q.y = (float)acc;
            acc += (double)v[m][j + 2], q.z = (float)acc;
            acc += (double)v[m][j + 3], q.w = (float)acc;
            if (w0 + j < Wp) *reinterpret_cast<float4*>(dst + w0 + j) = q;
        }
    }
    __syncthreads();
    const long long hw = (long long)H * W;
    const float nh = (float)(min(h + r, H - 1) - max(h - r, 0) + 1), n_in = nh * (float)(2 * r + 1);  // exact small integers
    const float rn_in = __frcp_rn(n_in);
    for (int w = threadIdx.x; w < W; w += 256) {
        const bool inner = w > r && w < W - r;  // the window is whole and both prefixes exist
        float n = n_in, rn = rn_in;
        float bx[NM];
        if (inner) {
#pragma unroll
            for (int m = 0; m < NM; ++m) bx[m] = __fsub_rn(srow[m * Wp + w + r], srow[m * Wp + w - r - 1]);
        } else {
            n = gfl_count(h, w, H, W, r), rn = __frcp_rn(n);
#pragma unroll
            for (int m = 0; m < NM; ++m) bx[m] = gfl_diff(srow + m * Wp, w, W, r, 1);
        }
        if (MODE == 0) {
            const float mean_x = gfl_div_count(bx[0], n, rn), mean_y = gfl_div_count(bx[1], n, rn);
            const float cov = __fsub_rn(gfl_div_count(bx[2 % NM], n, rn), __fmul_rn(mean_x, mean_y));
            const float var = __fsub_rn(gfl_div_count(bx[3 % NM], n, rn), __fmul_rn(mean_x, mean_x));
            const float a = __fdiv_rn(cov, __fadd_rn(var, eps));
            o0[row + w] = a;
            o1[row + w] = __fsub_rn(mean_y, __fmul_rn(a, mean_x));
        } else {
            const float ma = gfl_div_count(bx[0], n, rn), mb = gfl_div_count(bx[1], n, rn);
            const long long px = (long long)(pl / C) * Cx + (Cx == 1 ? 0 : pl % C);
            o0[row + w] = __fadd_rn(__fmul_rn(ma, x[px * hw + (long long)h * W + w]), mb);
        }
    }
}

template <int MODE, int E>
static void gfl_launch_rowbox(const float* d1, float* o0, float* o1, const float* x, int planes, int C, int Cx, int H,
                              int W, int r, float eps, size_t map_stride, cudaStream_t s) {
    constexpr int NM = MODE == 0 ? 4 : 2;
    const size_t smem = NM * 8 * sizeof(double) + (size_t)NM * ((W + 3) & ~3) * sizeof(float);
    if (smem > 48 * 1024)  // only rows beyond ~3000 pixels; per device, so not cached
        cudaFuncSetAttribute(gfl_rowbox_kernel<MODE, E>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    gfl_rowbox_kernel<MODE, E><<<dim3(H, planes), 256, smem, s>>>(d1, o0, o1, x, C, Cx, H, W, r, eps, map_stride);
}
template <int MODE>
static void gfl_rowbox(const float* d1, float* o0, float* o1, const float* x, int planes, int C, int Cx, int H, int W,
                       int r, float eps, size_t map_stride, cudaStream_t s) {
    if (W <= 1024)
        gfl_launch_rowbox<MODE, 4>(d1, o0, o1, x, planes, C, Cx, H, W, r, eps, map_stride, s);
    else if (W <= 2048)
        gfl_launch_rowbox<MODE, 8>(d1, o0, o1, x, planes, C, Cx, H, W, r, eps, map_stride, s);
    else
        gfl_launch_rowbox<MODE, 16>(d1, o0, o1, x, planes, C, Cx, H, W, r, eps, map_stride, s);
}

// The filter's first stage alone -- A, b of every window of (x, y), planes of H x W -- for FastGuidedFilter's
// low-resolution stage (dcev_tail.cu): d1 = 4 maps of `map_stride` floats, tot = gfl_tot_doubles() doubles.
size_t gfl_tot_doubles(int planes, int H, int W) { return 4 * (size_t)planes * ((H + GFL_CH - 1) / GFL_CH) * W; }
int gfl_coefficients(const float* x, const float* y, float* A, float* Bc, int planes, int H, int W, int r, float eps,
                     float* d1, size_t map_stride, double* tot, cudaStream_t s) {
    const dim3 cgrid((W + 255) / 256, planes, (H + GFL_CH - 1) / GFL_CH);
    gfl_colscan_kernel<0, true><<<cgrid, 256, 0, s>>>(x, y, d1, tot, 1, 1, H, W, map_stride);
    gfl_coldiff<0>(x, y, d1, tot, 1, 1, H, W, r, map_stride, cgrid, s);
    gfl_rowbox<0>(d1, A, Bc, nullptr, planes, 1, 1, H, W, r, eps, map_stride, s);
    return launch_status();
}

}  // namespace b200

// scratch: 4 + 4 prefix planes and A, b, each planes * H * W fp32, and the chunk totals of the column scans
extern "C" size_t b200_guided_filter_workspace_bytes(int B, int C, int H, int W) {
    if (B <= 0 || C <= 0 || H <= 0 || W <= 0) return 0;
    const size_t nch = (size_t)(H + b200::GFL_CH - 1) / b200::GFL_CH, ncw = (size_t)(W + b200::GFL_CW - 1) / b200::GFL_CW;
    const size_t per_plane = nch * W > ncw * H ? nch * W : ncw * H;  // column-scan totals, later row-scan totals
    return 10 * b200::gfl_align_up((size_t)B * C * H * W * sizeof(float), 256) +
           b200::gfl_align_up(4 * (size_t)B * C * per_plane * sizeof(double), 256);
}

extern "C" int b200_guided_filter(const float* x, const float* y, float* out, int B, int Cx, int C, int H, int W,
                                  int box_radius, float eps, void* workspace, size_t workspace_bytes, void* stream) {
    using namespace b200;
    if (!x || !y || !out || !workspace) return B200_ERR_NULL;
    if (B <= 0 || C <= 0 || H <= 0 || W <= 0) return B200_ERR_BAD_SHAPE;
    if (Cx != 1 && Cx != C) return B200_ERR_BAD_SHAPE;  // guided_filter.py:124
    const int r = box_radius;
    // box_filter's slicing needs 2r + 1 <= size; r = 0 fails in the reference (empty slice), the double-accumulated
    // prefixes are exact for <= 4096 addends
    if (r < 1 || H < 2 * r + 1 || W < 2 * r + 1 || H > 4096 || W > 4096 || (long long)B * C * 4 > 65535) return B200_ERR_UNSUPPORTED;  // grid.y / grid.z limits
    if (workspace_bytes < b200_guided_filter_workspace_bytes(B, C, H, W)) return B200_ERR_WORKSPACE;
    if ((reinterpret_cast<uintptr_t>(workspace) & 15u) != 0) return B200_ERR_MISALIGNED;
    cudaStream_t s = (cudaStream_t)stream;
    const int planes = B * C;
    const size_t plane_set = gfl_align_up((size_t)planes * H * W * sizeof(float), 256) / sizeof(float);
    float* p1 = (float*)workspace;        // 4 maps
    float* p2 = p1 + 4 * plane_set;       // 4 maps
    float* A = p2 + 4 * plane_set;
    float* Bc = A + plane_set;
    double* tot = (double*)(Bc + plane_set);
    const int nch = (H + GFL_CH - 1) / GFL_CH;
    const long long total = (long long)planes * H * W;
    long long eb = (total + 255) / 256;
    if (eb > B200_NUM_SMS * 32) eb = B200_NUM_SMS * 32;
    const dim3 cgrid((W + 255) / 256, planes, nch);
    const char* fused = getenv("B200DCE_GF_FUSED");
    if (!fused || fused[0] != '0') {  // four passes (+ two totals passes when the columns are chunked)
        // enough columns in flight (>= ~400 threads per SM, 18 loads each) to stream HBM without chunking: one chunk per
        // column, no totals pass (B200DCE_GF_FUSED=2 / 3 force the chunks / one chunk: A/B aid)
        const bool whole = fused ? fused[0] == '3' : (long long)planes * W >= 60000;  // 3 / 2: force one chunk / chunks
        const dim3 fgrid = whole ? dim3(cgrid.x, cgrid.y, 1) : cgrid;
        if (fgrid.z > 1) gfl_colscan_kernel<0, true><<<fgrid, 256, 0, s>>>(x, y, p1, tot, C, Cx, H, W, plane_set);
        gfl_coldiff<0>(x, y, p1, tot, C, Cx, H, W, r, plane_set, fgrid, s);
        gfl_rowbox<0>(p1, A, Bc, nullptr, planes, C, Cx, H, W, r, eps, plane_set, s);
        if (fgrid.z > 1) gfl_colscan_kernel<1, true><<<fgrid, 256, 0, s>>>(A, Bc, p1, tot, C, C, H, W, plane_set);
        gfl_coldiff<1>(A, Bc, p1, tot, C, C, H, W, r, plane_set, fgrid, s);
        gfl_rowbox<1>(p1, out, nullptr, x, planes, C, Cx, H, W, r, eps, plane_set, s);
        return launch_status();
    }
    gfl_colscan_kernel<0, true><<<cgrid, 256, 0, s>>>(x, y, p1, tot, C, Cx, H, W, plane_set);
    gfl_colscan_kernel<0, false><<<cgrid, 256, 0, s>>>(x, y, p1, tot, C, Cx, H, W, plane_set);
    const int ncw = (W + GFL_CW - 1) / GFL_CW;
    gfl_rowscan_kernel<true><<<dim3((ncw + 7) / 8, planes * 4, (H + 31) / 32), 256, 0, s>>>(p1, p2, tot, H, W, r, planes, plane_set, ncw);
    gfl_rowscan_kernel<false><<<dim3((ncw + 7) / 8, planes * 4, (H + 31) / 32), 256, 0, s>>>(p1, p2, tot, H, W, r, planes, plane_set, ncw);
    gfl_ab_kernel<<<(unsigned)eb, 256, 0, s>>>(p2, A, Bc, H, W, r, eps, plane_set, total);
    gfl_colscan_kernel<1, true><<<cgrid, 256, 0, s>>>(A, Bc, p1, tot, C, C, H, W, plane_set);
    gfl_colscan_kernel<1, false><<<cgrid, 256, 0, s>>>(A, Bc, p1, tot, C, C, H, W, plane_set);
    gfl_rowscan_kernel<true><<<dim3((ncw + 7) / 8, planes * 2, (H + 31) / 32), 256, 0, s>>>(p1, p2, tot, H, W, r, planes, plane_set, ncw);
    gfl_rowscan_kernel<false><<<dim3((ncw + 7) / 8, planes * 2, (H + 31) / 32), 256, 0, s>>>(p1, p2, tot, H, W, r, planes, plane_set, ncw);
    gfl_out_kernel<<<(unsigned)eb, 256, 0, s>>>(p2, x, out, C, Cx, H, W, r, plane_set, total);
    return launch_status();
}

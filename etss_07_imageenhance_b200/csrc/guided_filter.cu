// GuidedFilter(radius, eps).forward(image = x, guide = y) at full resolution, as GCENet applies it to the result of its
// curve loop (src/mon/vision/filtering/guided_filter.py:99-146, src/mon/vision/enhance/llie/gcenet/gcenet.py:391,485):
//   N = box(1); mean_x = box(x)/N; mean_y = box(y)/N; cov = box(x*y)/N - mean_x*mean_y; var = box(x*x)/N - mean_x^2
//   A = cov / (var + eps); b = mean_y - A*mean_x; out = (box(A)/N) * x + box(b)/N
// with box(t) = diff_y(diff_x(t.cumsum(2), r).cumsum(3), r) (box_filter.py:26-117): window sums as DIFFERENCES OF
// ROUNDED RUNNING SUMS.  ATen's CPU cumsum accumulates fp32 input in double and rounds every prefix to fp32, and the
// kernels below do the same (sums of <= 2^12 fp32 values >= 2^-18 are exact in double in any order), because A is a
// quotient of two such differences and amplifies their rounding (see dcev_tail.cu / DESIGN.md).
//
// Six HBM-bound passes over planes of B*C*H*W fp32:
//   column scan (x, y -> 4 prefix planes) | row scan of diff_x | A, b | column scan (A, b) | row scan | output
// Column scans: one thread per column and 72-row chunk (coalesced across the warp), 8 rows of loads in flight.  Row scans: a warp owns
// 32 rows and moves 32 x 32 blocks through shared memory (coalesced 128-byte rows in and out, conflict-free scan).
#include <math.h>

#include "common.cuh"

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
    const int h1 = min(H, (ch + 1) * GFL_CH);
    int h = ch * GFL_CH;
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

// =====================================================================================
// cvx_kernels.cu -- hand-written sm_100a kernels of the 3D Curvix edge-sketch hot path.
//
// Compile with: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -lineinfo
// (-fmad=false keeps every a*b+c un-fused so accept/reject decisions match an un-fused CPU build).
//
// Pipeline of one hypothesis view pair (Run_3D_Edge_Sketch, EdgeSketch_Core.cpp:168-573):
//   stage_views   per validation view: two epipolar angles per edge (a2); two counting-sorted angular indexes
//   h2_angles     epipolar angle of every view-2 edge (a2)          h1_wedges: wedge of every view-1 edge (a3)
//   form_count / form_fill   one warp per view-1 edge: wedge scan of view 2 (a5), epipolar correction (a6),
//                 greedy clustering in shared memory (a7) -> hypothesis table + per-edge cluster ranks
//   uniq_fill     one entry per cluster (members of a cluster carry the same averaged edge) + hoisted reprojection terms
//   wedge31       wedge of the view-1 edge in each validation view (a3, row 0 only)
//   validate      per (unique hypothesis, view): reprojected tangent (a8), wedge of the view-2 edge (a3), parallel-wedge
//                 test (a10), wedge31 n wedge32 enumeration from the narrower of the two angular indexes (a5, a11),
//                 orientation arg-max (a12) -> support index
//   count_u / select / select_fill   vote + ratio test (a13);  finalize: correction + 2-view DLT + tangent (a15)
// No tensor cores: there is no dense contraction anywhere on this path.
// =====================================================================================
#include "cvx_kernels.cuh"

#include <float.h>
#include <stdlib.h>
#include <stdio.h>

#include <algorithm>

namespace cvx {

#define FULL 0xffffffffu

__device__ __forceinline__ unsigned lanemask_lt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}
__device__ __forceinline__ int grid_bin(double a, double lo, double sc, int G) {   // monotone in a: a <= b => bin(a) <= bin(b)
    const int b = int((a - lo) * sc);
    return b < 0 ? 0 : (b >= G ? G - 1 : b);
}
__device__ __forceinline__ int grid_bins(double a1, double a2, const ViewGrid& g) {   // (b1 << 16) | b2, -1 for NaN
    if (a1 != a1 || a2 != a2) return -1;
    return (grid_bin(a1, g.lo1, g.s1, g.G1) << 16) | grid_bin(a2, g.lo2, g.s2, g.G2);
}
__device__ __forceinline__ int cell_A(int packed, const ViewGrid& g) { return packed < 0 ? g.G1 * g.G2 : (packed >> 16) * g.G2 + (packed & 0xffff); }
__device__ __forceinline__ int cell_B(int packed, const ViewGrid& g) { return packed < 0 ? g.G1 * g.G2 : (packed & 0xffff) * g.G1 + (packed >> 16); }

// ----------------------------------------------------------------------------- upload-time
__global__ void k_edge_trig(const double* __restrict__ th, double* __restrict__ cs, double* __restrict__ sn, int64_t n) {
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < n) { cs[i] = cos(th[i]); sn[i] = sin(th[i]); }   // EdgeSketch_Core.cpp:321-322, hoisted: depends on the view only
}
void launch_edge_trig(const double* th, double* cs, double* sn, int64_t n, cudaStream_t st) {
    if (n <= 0) return;
    k_edge_trig<<<unsigned((n + 255) / 256), 256, 0, st>>>(th, cs, sn, n);
}

// All views in one launch: packed host-order columns (already on the device in a staging buffer) -> per-view arena slots, with
// cos/sin of the orientation.  tab = [src_off | dst_off | n] (3 x n_views int64).
__global__ void __launch_bounds__(256) k_ingest_views(const double* __restrict__ sx, const double* __restrict__ sy, const double* __restrict__ st,
                                                      const int64_t* __restrict__ tab, int n_views, EdgeArena ea) {
    const int v = blockIdx.y;
    const int64_t n = tab[2 * n_views + v];
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    const int64_t s = tab[v] + i, d = tab[n_views + v] + i;
    const double th = st[s];
    ea.x[d] = sx[s]; ea.y[d] = sy[s]; ea.th[d] = th;
    ea.cs[d] = cos(th); ea.sn[d] = sin(th);                       // EdgeSketch_Core.cpp:321-322, hoisted
}
void launch_ingest_views(const double* sx, const double* sy, const double* st_, const int64_t* tab, int n_views, int64_t max_n, EdgeArena ea,
                         cudaStream_t st) {
    if (max_n <= 0 || n_views <= 0) return;
    dim3 grid(unsigned((max_n + 255) / 256), n_views);
    k_ingest_views<<<grid, 256, 0, st>>>(sx, sy, st_, tab, n_views, ea);
}

// ----------------------------------------------------------------------------- per-pass view staging
#define GRID_MAX_BINS 16384
#define GRID_CELLS_PER_WEDGE 2.0   // a wedge spans 3 (at most 4) rows of cells: what a lane keeps in registers (VAL_LIGHT_ROWS)
int pick_cell_stride(int64_t max_edges) {
    int64_t per_edge = 16;   // cell budget per edge of the largest view (tuning knob for experiments: CVX_GRID_CELLS_PER_EDGE)
    if (const char* e = getenv("CVX_GRID_CELLS_PER_EDGE")) per_edge = std::max(1, atoi(e));
    int64_t budget = std::max<int64_t>(per_edge * max_edges, 4096);
    budget = std::min<int64_t>(budget, int64_t(1) << 26);
    return int((budget + 1 + 3) & ~int64_t(3));
}
// order-preserving map double -> u64 (so that atomicMin / atomicMax on the keys order like the doubles)
__device__ __forceinline__ unsigned long long ord_key(double d) {
    const unsigned long long u = (unsigned long long)__double_as_longlong(d);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ULL);
}
__device__ __forceinline__ double ord_val(unsigned long long k) {
    return __longlong_as_double((long long)((k >> 63) ? (k & 0x7fffffffffffffffULL) : ~k));
}
// warp-reduced min / max of the two angles into minmax[v][4]; NaNs are skipped by the callers
__device__ __forceinline__ void minmax_commit(double mn1, double mx1, double mn2, double mx2, unsigned long long* mm) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mn1 = fmin(mn1, __shfl_xor_sync(FULL, mn1, o)); mx1 = fmax(mx1, __shfl_xor_sync(FULL, mx1, o));
        mn2 = fmin(mn2, __shfl_xor_sync(FULL, mn2, o)); mx2 = fmax(mx2, __shfl_xor_sync(FULL, mx2, o));
    }
    if ((threadIdx.x & 31) == 0 && mn1 <= mx1) {
        atomicMin(&mm[0], ord_key(mn1)); atomicMax(&mm[1], ord_key(mx1));
        atomicMin(&mm[2], ord_key(mn2)); atomicMax(&mm[3], ord_key(mx2));
    }
}
__global__ void k_minmax_init(unsigned long long* mm, int n_vali) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n_vali) { mm[4 * v] = mm[4 * v + 2] = 0xfff0000000000000ULL; mm[4 * v + 1] = mm[4 * v + 3] = 0x000fffffffffffffULL; }   // +inf / -inf
}
// getOreListVali for both hypothesis views (EdgeSketch_Core.cpp:337,342), hoisted out of the H1 loop, + per-view angle ranges
__global__ void __launch_bounds__(256) k_stage_angles(EdgeArena ea, const int64_t* __restrict__ edge_off, const int32_t* __restrict__ edge_n,
                                                      const int32_t* __restrict__ vali_view, const ViewConst* __restrict__ vc) {
    const int v = blockIdx.y;
    const int view = vali_view[v];
    const int n = edge_n[view];
    const int64_t off = edge_off[view];
    const double e31x = vc[v].e2_31[0], e31y = vc[v].e2_31[1], e32x = vc[v].e2_32[0], e32y = vc[v].e2_32[1];
    double mn1 = DBL_MAX, mx1 = -DBL_MAX, mn2 = DBL_MAX, mx2 = -DBL_MAX;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double x = ea.x[off + i], y = ea.y[off + i];
        const double a31 = ore_angle(x, y, e31x, e31y);
        const double a32 = ore_angle(x, y, e32x, e32y);
        ea.ang[off + i] = make_double2(a31, a32);
        if (a31 == a31 && a32 == a32) { mn1 = fmin(mn1, a31); mx1 = fmax(mx1, a31); mn2 = fmin(mn2, a32); mx2 = fmax(mx2, a32); }
    }
    minmax_commit(mn1, mx1, mn2, mx2, ea.gi.minmax + 4 * v);
}
// Grid sizing, one warp per validation view.  The lanes sample wedges (64 of each kind): their typical (geometric mean)
// width sets the cell size in that angle -- half a wedge -- within the per-view cell budget.  A wrong estimate only
// costs time: membership is always decided by the exact tests on the stored angles.
template <bool GIVEN>
__global__ void __launch_bounds__(128) k_grid_params(GridIndex gi, int n_vali, const __grid_constant__ GridSample gs, const ViewConst* __restrict__ vc) {
    const int v = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (v >= n_vali) return;
    double sl1 = 0, sl2 = 0;
    int n1 = 0, n2 = 0;
    for (int k = lane; k < 64; k += 32) {
        double wd1 = nan(""), wd2 = nan("");
        if (GIVEN) {
            if (gs.n31 > 0) { const double2 w = gs.w31[int64_t(v) * gs.n31 + (gs.n31 * k) / 64]; wd1 = w.y - w.x; }
            if (gs.n32 > 0) { const double2 w = gs.w32[int64_t(v) * gs.n32 + (gs.n32 * k) / 64]; wd2 = w.y - w.x; }
        } else {
            const ViewConst& c = vc[v];
            double lo, hi;
            if (gs.E1 > 0) {
                const int i = int((int64_t(gs.E1) * k) / 64);
                wedge_row(gs.x1[i], gs.y1[i], c.F31, c.e1_31[0], c.e1_31[1], c.e2_31[0], c.e2_31[1], gs.delta, gs.rows, gs.cols, &lo, &hi);
                wd1 = hi - lo;
            }
            if (gs.E2 > 0) {
                const int i = int((int64_t(gs.E2) * k) / 64);
                wedge_row(gs.x2[i], gs.y2[i], c.F32, c.e1_32[0], c.e1_32[1], c.e2_32[0], c.e2_32[1], gs.delta, gs.rows, gs.cols, &lo, &hi);
                wd2 = hi - lo;
            }
        }
        if (wd1 == wd1) { sl1 += log(fmax(wd1, 1e-6)); n1++; }
        if (wd2 == wd2) { sl2 += log(fmax(wd2, 1e-6)); n2++; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sl1 += __shfl_xor_sync(FULL, sl1, o); sl2 += __shfl_xor_sync(FULL, sl2, o);
        n1 += __shfl_xor_sync(FULL, n1, o); n2 += __shfl_xor_sync(FULL, n2, o);
    }
    if (lane != 0) return;
    const unsigned long long* mm = gi.minmax + 4 * v;
    ViewGrid g;
    g.lo1 = ord_val(mm[0]); g.hi1 = ord_val(mm[1]); g.lo2 = ord_val(mm[2]); g.hi2 = ord_val(mm[3]);
    if (!(g.lo1 <= g.hi1)) { g.lo1 = g.hi1 = g.lo2 = g.hi2 = 0.0; }   // view without (finite) edges
    const double range1 = g.hi1 - g.lo1, range2 = g.hi2 - g.lo2;
    const double cw1 = n1 ? exp(sl1 / n1) / GRID_CELLS_PER_WEDGE : range1 / 64, cw2 = n2 ? exp(sl2 / n2) / GRID_CELLS_PER_WEDGE : range2 / 64;
    double G1 = range1 > 0 ? fmin(ceil(range1 / cw1), double(GRID_MAX_BINS)) : 1.0;
    double G2 = range2 > 0 ? fmin(ceil(range2 / cw2), double(GRID_MAX_BINS)) : 1.0;
    const double budget = double(gi.CS - 4);
    if (G1 * G2 > budget) {
        const double f = sqrt(budget / (G1 * G2));
        G1 = fmax(1.0, floor(G1 * f)); G2 = fmax(1.0, floor(G2 * f));
        if (G1 * G2 > budget) G2 = fmax(1.0, floor(budget / G1));
        if (G1 * G2 > budget) G1 = fmax(1.0, floor(budget / G2));
    }
    g.G1 = int(G1); g.G2 = int(G2);
    g.s1 = range1 > 0 ? G1 / range1 : 0.0;
    g.s2 = range2 > 0 ? G2 / range2 : 0.0;
    gi.vg[v] = g;
}
__global__ void __launch_bounds__(256) k_stage_hist(EdgeArena ea, const int64_t* __restrict__ edge_off, const int32_t* __restrict__ edge_n,
                                                    const int32_t* __restrict__ vali_view) {
    const int v = blockIdx.y;
    const int view = vali_view[v];
    const int n = edge_n[view];
    const int64_t off = edge_off[view];
    const ViewGrid g = ea.gi.vg[v];
    int32_t* ceA = ea.gi.A.cell_end + int64_t(v) * ea.gi.CS;
    int32_t* ceB = ea.gi.B.cell_end + int64_t(v) * ea.gi.CS;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double2 a = ea.ang[off + i];
        const int c = grid_bins(a.x, a.y, g);
        ea.cell[off + i] = c;
        atomicAdd(&ceA[cell_A(c, g)], 1);
        atomicAdd(&ceB[cell_B(c, g)], 1);
    }
}
// Segmented exclusive scan of the cell histograms (one segment per view and layout, G1 * G2 + 1 used cells each), in place:
// tile sums -> scan of the tile sums of a segment -> tile-local scan + tile offset.  Coalesced int4 traffic throughout.
#define SCAN_TILE 4096   // 1024 threads x int4
int scan_tiles(int CS) { return (CS + SCAN_TILE - 1) / SCAN_TILE; }
__device__ __forceinline__ int32_t* scan_segment(int32_t* cellsA, int32_t* cellsB, int n_vali, int CS, const ViewGrid* vg, int* n_used) {
    const int seg = blockIdx.y;
    const int v = seg < n_vali ? seg : seg - n_vali;
    *n_used = vg[v].G1 * vg[v].G2 + 1;
    return (seg < n_vali ? cellsA : cellsB) + int64_t(v) * CS;
}
__device__ __forceinline__ int block_sum_1024(int x, int* sh) {   // returns the block total; sh[32]
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(FULL, x, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = x;
    __syncthreads();
    int t = (threadIdx.x < 32) ? sh[threadIdx.x] : 0;
    if (threadIdx.x < 32) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(FULL, t, o);
        if (threadIdx.x == 0) sh[0] = t;
    }
    __syncthreads();
    return sh[0];
}
__global__ void __launch_bounds__(1024) k_cell_tile_sums(int32_t* cellsA, int32_t* cellsB, int n_vali, int CS, const ViewGrid* __restrict__ vg,
                                                         int32_t* __restrict__ tile_sums, int tiles) {
    __shared__ int sh[32];
    int n;
    const int32_t* b = scan_segment(cellsA, cellsB, n_vali, CS, vg, &n);
    const int base = blockIdx.x * SCAN_TILE;
    if (base >= n) return;
    const int i = base + threadIdx.x * 4;
    int s = 0;
    if (i + 3 < n) { const int4 q = *reinterpret_cast<const int4*>(b + i); s = q.x + q.y + q.z + q.w; }
    else for (int k = i; k < n; k++) s += b[k];
    const int tot = block_sum_1024(s, sh);
    if (threadIdx.x == 0) tile_sums[blockIdx.y * tiles + blockIdx.x] = tot;
}
__global__ void __launch_bounds__(1024) k_cell_tile_scan(int32_t* tile_sums, int tiles, int n_vali, int CS, const ViewGrid* __restrict__ vg) {
    // one block per segment; tiles <= 16384 (CS <= 2^26): each thread owns a contiguous run
    __shared__ int part[1024];
    const int seg = blockIdx.x;
    const int v = seg < n_vali ? seg : seg - n_vali;
    const int used = (vg[v].G1 * vg[v].G2 + 1 + SCAN_TILE - 1) / SCAN_TILE;
    int32_t* t = tile_sums + seg * tiles;
    const int per = (used + 1023) / 1024;
    const int lo = min(int(threadIdx.x) * per, used), hi = min(lo + per, used);
    int s = 0;
    for (int i = lo; i < hi; i++) s += t[i];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        const int x = (threadIdx.x >= d) ? part[threadIdx.x - d] : 0;
        __syncthreads();
        part[threadIdx.x] += x;
        __syncthreads();
    }
    int run = part[threadIdx.x] - s;
    for (int i = lo; i < hi; i++) { const int c = t[i]; t[i] = run; run += c; }
}
__global__ void __launch_bounds__(1024) k_cell_scan_apply(int32_t* cellsA, int32_t* cellsB, int n_vali, int CS, const ViewGrid* __restrict__ vg,
                                                          const int32_t* __restrict__ tile_sums, int tiles) {
    __shared__ int wsum[32];
    int n;
    int32_t* b = scan_segment(cellsA, cellsB, n_vali, CS, vg, &n);
    const int base = blockIdx.x * SCAN_TILE;
    if (base >= n) return;
    const int i = base + threadIdx.x * 4;
    int q[4] = {0, 0, 0, 0};
    if (i + 3 < n) { const int4 t = *reinterpret_cast<const int4*>(b + i); q[0] = t.x; q[1] = t.y; q[2] = t.z; q[3] = t.w; }
    else for (int k = 0; k < 4; k++) if (i + k < n) q[k] = b[i + k];
    const int mine = q[0] + q[1] + q[2] + q[3];
    int inc = mine;                                     // inclusive scan across the warp
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int w = wsum[lane], winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULL, winc, o); if (lane >= o) winc += t; }
        wsum[lane] = winc - w;                          // exclusive offset of each warp
    }
    __syncthreads();
    int run = tile_sums[blockIdx.y * tiles + blockIdx.x] + wsum[wid] + inc - mine;
    int o4[4];
#pragma unroll
    for (int k = 0; k < 4; k++) { o4[k] = run; run += q[k]; }
    if (i + 3 < n) *reinterpret_cast<int4*>(b + i) = make_int4(o4[0], o4[1], o4[2], o4[3]);
    else for (int k = 0; k < 4; k++) if (i + k < n) b[i + k] = o4[k];
}
static void launch_cell_scan(const GridIndex& gi, int n_vali, int32_t* scan_tmp, cudaStream_t st) {
    const int tiles = scan_tiles(gi.CS);
    dim3 grid(tiles, 2 * n_vali);
    k_cell_tile_sums<<<grid, 1024, 0, st>>>(gi.A.cell_end, gi.B.cell_end, n_vali, gi.CS, gi.vg, scan_tmp, tiles);
    k_cell_tile_scan<<<2 * n_vali, 1024, 0, st>>>(scan_tmp, tiles, n_vali, gi.CS, gi.vg);
    k_cell_scan_apply<<<grid, 1024, 0, st>>>(gi.A.cell_end, gi.B.cell_end, n_vali, gi.CS, gi.vg, scan_tmp, tiles);
}
// counting-sort scatter; afterwards cell_end[v][c] holds the END of cell c (start of cell c is cell_end[v][c-1])
__global__ void __launch_bounds__(256) k_stage_scatter(EdgeArena ea, const int64_t* __restrict__ edge_off, const int32_t* __restrict__ edge_n,
                                                       const int32_t* __restrict__ vali_view) {
    const int v = blockIdx.y;
    const int view = vali_view[v];
    const int n = edge_n[view];
    const int64_t off = edge_off[view];
    const ViewGrid g = ea.gi.vg[v];
    int32_t* ceA = ea.gi.A.cell_end + int64_t(v) * ea.gi.CS;
    int32_t* ceB = ea.gi.B.cell_end + int64_t(v) * ea.gi.CS;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int c = ea.cell[off + i];
        const double2 a = ea.ang[off + i];
        const double4 rec = make_double4(a.x, a.y, ea.cs[off + i], ea.sn[off + i]);
        const int pa = atomicAdd(&ceA[cell_A(c, g)], 1);
        const int pb = atomicAdd(&ceB[cell_B(c, g)], 1);
        ea.gi.A.rec[off + pa] = rec; ea.gi.A.idx[off + pa] = i;
        ea.gi.B.rec[off + pb] = rec; ea.gi.B.idx[off + pb] = i;
    }
}
void launch_stage_views(const EdgeArena& ea, const int64_t* edge_off, const int32_t* edge_n, const int32_t* vali_view, int n_vali,
                        int max_edges, const ViewConst* vc, const GridSample& gs, int32_t* scan_tmp, cudaStream_t st, int* launches) {
    if (n_vali <= 0) return;
    cudaMemsetAsync(ea.gi.A.cell_end, 0, size_t(n_vali) * ea.gi.CS * sizeof(int32_t), st);
    cudaMemsetAsync(ea.gi.B.cell_end, 0, size_t(n_vali) * ea.gi.CS * sizeof(int32_t), st);
    int bx = (max_edges + 255) / 256;
    if (bx < 1) bx = 1;
    if (bx > 1024) bx = 1024;
    dim3 grid(bx, n_vali);
    k_minmax_init<<<(n_vali + 127) / 128, 128, 0, st>>>(ea.gi.minmax, n_vali);
    k_stage_angles<<<grid, 256, 0, st>>>(ea, edge_off, edge_n, vali_view, vc);
    k_grid_params<false><<<(n_vali + 3) / 4, 128, 0, st>>>(ea.gi, n_vali, gs, vc);
    k_stage_hist<<<grid, 256, 0, st>>>(ea, edge_off, edge_n, vali_view);
    launch_cell_scan(ea.gi, n_vali, scan_tmp, st);
    k_stage_scatter<<<grid, 256, 0, st>>>(ea, edge_off, edge_n, vali_view);
    if (launches) *launches += 8;
}

// stand-alone support test: the per-view angle / tangent arrays are given; only the grid is built here
__global__ void __launch_bounds__(256) k_given_minmax(const double* __restrict__ a31, const double* __restrict__ a32,
                                                      const int64_t* __restrict__ view_off, unsigned long long* minmax) {
    const int v = blockIdx.y;
    const int64_t b0 = view_off[v], n = view_off[v + 1] - b0;
    double mn1 = DBL_MAX, mx1 = -DBL_MAX, mn2 = DBL_MAX, mx2 = -DBL_MAX;
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        const double x = a31[b0 + i], y = a32[b0 + i];
        if (x == x && y == y) { mn1 = fmin(mn1, x); mx1 = fmax(mx1, x); mn2 = fmin(mn2, y); mx2 = fmax(mx2, y); }
    }
    minmax_commit(mn1, mx1, mn2, mx2, minmax + 4 * v);
}
__global__ void __launch_bounds__(256) k_given_hist(const double* __restrict__ a31, const double* __restrict__ a32,
                                                    const int64_t* __restrict__ view_off, int32_t* __restrict__ cell, GridIndex gi) {
    const int v = blockIdx.y;
    const int64_t b0 = view_off[v], n = view_off[v + 1] - b0;
    const ViewGrid g = gi.vg[v];
    int32_t* ceA = gi.A.cell_end + int64_t(v) * gi.CS;
    int32_t* ceB = gi.B.cell_end + int64_t(v) * gi.CS;
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        const int c = grid_bins(a31[b0 + i], a32[b0 + i], g);
        cell[b0 + i] = c;
        atomicAdd(&ceA[cell_A(c, g)], 1);
        atomicAdd(&ceB[cell_B(c, g)], 1);
    }
}
__global__ void __launch_bounds__(256) k_given_scatter(const double* __restrict__ a31, const double* __restrict__ a32, const double* __restrict__ cs,
                                                       const double* __restrict__ sn, const int64_t* __restrict__ view_off,
                                                       const int32_t* __restrict__ cell, GridIndex gi) {
    const int v = blockIdx.y;
    const int64_t b0 = view_off[v], n = view_off[v + 1] - b0;
    const ViewGrid g = gi.vg[v];
    int32_t* ceA = gi.A.cell_end + int64_t(v) * gi.CS;
    int32_t* ceB = gi.B.cell_end + int64_t(v) * gi.CS;
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        const int c = cell[b0 + i];
        const double4 rec = make_double4(a31[b0 + i], a32[b0 + i], cs[b0 + i], sn[b0 + i]);
        const int pa = atomicAdd(&ceA[cell_A(c, g)], 1);
        const int pb = atomicAdd(&ceB[cell_B(c, g)], 1);
        gi.A.rec[b0 + pa] = rec; gi.A.idx[b0 + pa] = int32_t(i);
        gi.B.rec[b0 + pb] = rec; gi.B.idx[b0 + pb] = int32_t(i);
    }
}
void launch_stage_given(const double* a31, const double* a32, const double* cs, const double* sn, const int64_t* view_off, int n_vali,
                        int64_t max_edges, int32_t* cell, GridIndex gi, const GridSample& gs, int32_t* scan_tmp, cudaStream_t st, int* launches) {
    cudaMemsetAsync(gi.A.cell_end, 0, size_t(n_vali) * gi.CS * sizeof(int32_t), st);
    cudaMemsetAsync(gi.B.cell_end, 0, size_t(n_vali) * gi.CS * sizeof(int32_t), st);
    int64_t bx = (max_edges + 255) / 256;
    if (bx < 1) bx = 1;
    if (bx > 2048) bx = 2048;
    dim3 grid((unsigned)bx, n_vali);
    k_minmax_init<<<(n_vali + 127) / 128, 128, 0, st>>>(gi.minmax, n_vali);
    k_given_minmax<<<grid, 256, 0, st>>>(a31, a32, view_off, gi.minmax);
    k_grid_params<true><<<(n_vali + 3) / 4, 128, 0, st>>>(gi, n_vali, gs, nullptr);
    k_given_hist<<<grid, 256, 0, st>>>(a31, a32, view_off, cell, gi);
    launch_cell_scan(gi, n_vali, scan_tmp, st);
    k_given_scatter<<<grid, 256, 0, st>>>(a31, a32, cs, sn, view_off, cell, gi);
    if (launches) *launches += 8;
}

// ----------------------------------------------------------------------------- pass-level tables
__global__ void k_h2_angles(const double* __restrict__ x2, const double* __restrict__ y2, int n, double ex, double ey, double* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = ore_angle(x2[i], y2[i], ex, ey);   // getOreList, EdgeSketch_Core.cpp:161
}
void launch_h2_angles(const double* x2, const double* y2, int n, double ex, double ey, double* out, cudaStream_t st) {
    if (n > 0) k_h2_angles<<<(n + 255) / 256, 256, 0, st>>>(x2, y2, n, ex, ey, out);
}
int pick_angle_bins(int n) {
    int B = 1 << 12;
    while (B < 4 * n && B < (1 << 20)) B <<= 1;
    return B;
}
__device__ __forceinline__ int angle_bin(double a, double scale, int B) {   // monotone in a
    if (a != a) return B;
    const int b = int(a * scale);
    return b < 0 ? 0 : (b >= B ? B - 1 : b);
}
__global__ void k_h2_hist(const double* __restrict__ ang, int n, AngleIndex ix, int32_t* __restrict__ bin_of) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int b = angle_bin(ang[i], ix.scale, ix.B);
    bin_of[i] = b;
    atomicAdd(&ix.bin_end[b], 1);
}
// exclusive scan of the 1024 per-thread partial sums of a block: two levels of warp shuffles, two barriers
template <typename T>
__device__ __forceinline__ T block_excl_scan_1024(T mine, T* sh /* 32 */) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    T inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const T t = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) sh[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        const T w = sh[lane];
        T winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const T t = __shfl_up_sync(FULL, winc, o); if (lane >= o) winc += t; }
        sh[lane] = winc - w;
    }
    __syncthreads();
    return sh[wid] + inc - mine;
}
__global__ void __launch_bounds__(1024) k_scan_bins1(int32_t* b, int n) {   // in-place exclusive scan, one block, int4 traffic
    __shared__ int sh[32];
    const int per = (((n + 1023) / 1024) + 3) & ~3;                          // multiple of 4: every thread's run is 16-byte aligned
    const int lo = min(int(threadIdx.x) * per, n), hi = min(lo + per, n);
    int s = 0;
    int i = lo;
    for (; i + 3 < hi; i += 4) { const int4 q = *reinterpret_cast<const int4*>(b + i); s += (q.x + q.y) + (q.z + q.w); }
    for (; i < hi; i++) s += b[i];
    int run = block_excl_scan_1024<int>(s, sh);
    i = lo;
    for (; i + 3 < hi; i += 4) {
        const int4 q = *reinterpret_cast<const int4*>(b + i);
        int4 o;
        o.x = run; o.y = o.x + q.x; o.z = o.y + q.y; o.w = o.z + q.z; run = o.w + q.w;
        *reinterpret_cast<int4*>(b + i) = o;
    }
    for (; i < hi; i++) { const int c = b[i]; b[i] = run; run += c; }
}
__global__ void k_h2_scatter(const double* __restrict__ ang, int n, AngleIndex ix, const int32_t* __restrict__ bin_of) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int p = atomicAdd(&ix.bin_end[bin_of[i]], 1);   // afterwards bin_end[b] = END of bin b
    ix.ang[p] = ang[i]; ix.idx[p] = i;
}
void launch_h2_index(const double* ang2, int n, int32_t* bin_scratch, AngleIndex ix, cudaStream_t st, int* launches) {
    cudaMemsetAsync(ix.bin_end, 0, size_t(ix.B + 1) * sizeof(int32_t), st);
    if (n <= 0) return;
    k_h2_hist<<<(n + 255) / 256, 256, 0, st>>>(ang2, n, ix, bin_scratch);
    k_scan_bins1<<<1, 1024, 0, st>>>(ix.bin_end, ix.B + 1);
    k_h2_scatter<<<(n + 255) / 256, 256, 0, st>>>(ang2, n, ix, bin_scratch);
    if (launches) *launches += 3;
}
__global__ void k_h1_wedges(const __grid_constant__ PassConst pc, const double* __restrict__ x1, const double* __restrict__ y1, int n,
                            double2* __restrict__ wedge, int32_t* __restrict__ flag) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double lo, hi;
    int f = wedge_row(x1[i], y1[i], pc.F21, pc.e1_21[0], pc.e1_21[1], pc.e2_21[0], pc.e2_21[1], pc.delta, pc.rows, pc.cols, &lo, &hi);
    wedge[i] = make_double2(lo, hi);   // getOreListBar, EdgeSketch_Core.cpp:163
    flag[i] = f;
}
void launch_h1_wedges(const PassConst& pc, const double* x1, const double* y1, int n, double2* wedge, int32_t* flag, cudaStream_t st) {
    if (n > 0) k_h1_wedges<<<(n + 127) / 128, 128, 0, st>>>(pc, x1, y1, n, wedge, flag);
}

// ----------------------------------------------------------------------------- scan
__global__ void __launch_bounds__(1024) k_scan_i32_i64(const int32_t* __restrict__ in, int64_t* __restrict__ out, int n) {
    __shared__ long long sh[32];
    const int per = (n + 1023) / 1024;
    const int lo = min(int(threadIdx.x) * per, n), hi = min(lo + per, n);
    long long s = 0;
    for (int i = lo; i < hi; i++) s += in[i];
    long long run = block_excl_scan_1024<long long>(s, sh);
    for (int i = lo; i < hi; i++) { out[i] = run; run += in[i]; }
    if (threadIdx.x == 1023) out[n] = run;
}
void launch_scan_i32_i64(const int32_t* in, int64_t* out, int n, cudaStream_t st) { k_scan_i32_i64<<<1, 1024, 0, st>>>(in, out, n); }

// ----------------------------------------------------------------------------- clustering (one warp per view-1 edge)
// EdgeClusterer.cpp:127-293.  The reference is a sequential greedy merge with restart: scan i ascending, take the nearest
// j of another cluster (dist < 1 px, cluster orientations within 20 deg, first minimum), merge if the sizes allow, restart.
// Distances never change, so the warp keeps an adjacency bit matrix (dist < threshold) in shared memory and evaluates
// "would edge i merge?" for 32 edges at a time; the first such i (ballot + ffs) is the reference's next merge.
template <int N>
struct ClusterSm {
    double x[N], y[N], thn[N], orient[N];   // corrected location, normalised orientation, per-label average orientation
    uint32_t adj[N * (N / 32)];
    uint32_t dirty[N / 32], canw[N / 32];   // rows whose merge decision must be re-evaluated / rows that can merge (bit per row)
    int16_t label[N], size[N], rank[N], mem[32];   // rank doubles as the cached merge partner during the merge phase
};
#define CL_N 128           // size classes: <= CL_N rows (a warp per edge), <= CL_MID_N (4 warps per edge), <= CL_BIG_N (8 warps per edge) --
#define CL_MID_N 256       // these three share one launch, k_cluster_classes; <= CL_HUGE_N (8 warps per edge, 171 KB of shared memory: one
#define CL_BIG_N 512       // block per SM, k_cluster_coop); beyond that, or MAX_CLUSTER_SIZE > 32: global-memory fallback (k_cluster_huge)
#define CL_HUGE_N 1024
#define FORM_WARPS 4
// The <= 128 / 256 / 512-row classes share one launch (k_cluster_classes): every block has the same shared-memory size and holds
// one 512-row edge (8 warps), two 256-row edges (4 warps each) or eight <= 128-row edges (a warp each).
#define CL_SMEM_QUANTUM (56 * 1024)   // 4 blocks per SM
#define CL_SMALL_WARPS 8             // warps per block; a <= 128-row edge per warp
#define CL_MID_WARPS 4               // warps that share one <= 256-row edge
#define CL_MID_EDGES (CL_SMALL_WARPS / CL_MID_WARPS)

// Gaussian-weighted average over the members s.mem[0..cnt) (ascending index): EdgeClusterer.cpp:60-125.  Lane k holds member k;
// every sum runs in member order on every lane through shuffles (bit-identical to the sequential code: the products are formed
// per member first, the additions are the reference's left-to-right chain); sqrt / exp of the members are lane-parallel.
__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(FULL, v, src); }
template <class SM, bool ORIENT_ONLY>   // ORIENT_ONLY: the merge step needs the averaged orientation alone (EdgeClusterer.cpp:168-176)
__device__ __noinline__ void gauss_avg_list_t(SM& s, const int16_t* mem, int cnt, double sigma, double sigma_inv, int lane, double* gx,
                                              double* gy, double* gth) {
    double xi = 0, yi = 0, ti = 0;
    if (lane < cnt) { const int i = mem[lane]; xi = s.x[i]; yi = s.y[i]; ti = s.thn[i]; }
    double sx = 0, sy = 0;
    for (int k = 0; k < cnt; k++) { sx += shfl_d(xi, k); sy += shfl_d(yi, k); }
    // x / cnt: a power-of-two count (2 members is the most frequent merge) divides exactly by a multiplication (warp-uniform branch)
    const bool p2 = (cnt & (cnt - 1)) == 0;
    const double ic = __hiloint2double((1023 - (31 - __clz(cnt))) << 20, 0);   // 2^-floor(log2 cnt)
    const double cx = p2 ? sx * ic : sx / cnt, cy = p2 ? sy * ic : sy / cnt;
    const double dx = xi - cx, dy = yi - cy;
    const double di = sqrt(dx * dx + dy * dy);
    double tot = 0.0;
    for (int k = 0; k < cnt; k++) tot += shfl_d(di, k);
    const double mean_shift = p2 ? tot * ic : tot / cnt;
    const double q = sigma_inv != 0.0 ? (di - mean_shift) * sigma_inv : (di - mean_shift) / sigma;   // exact when sigma is a power of two
    const double w = exp(-0.5 * (q * q));
    const double wo = w * ti;
    double swo = 0, tw = 0;
    if (ORIENT_ONLY) {
        for (int k = 0; k < cnt; k++) { swo += shfl_d(wo, k); tw += shfl_d(w, k); }
    } else {
        const double wx = w * xi, wy = w * yi;
        double swx = 0, swy = 0;
        for (int k = 0; k < cnt; k++) { swx += shfl_d(wx, k); swy += shfl_d(wy, k); swo += shfl_d(wo, k); tw += shfl_d(w, k); }
        *gx = swx / tw; *gy = swy / tw;
    }
    *gth = swo / tw;
}
template <int N, bool ORIENT_ONLY>
__device__ __forceinline__ void gauss_avg_list(ClusterSm<N>& s, const int16_t* mem, int cnt, double sigma, double sigma_inv, int lane, double* gx,
                                               double* gy, double* gth) {
    gauss_avg_list_t<ClusterSm<N>, ORIENT_ONLY>(s, mem, cnt, sigma, sigma_inv, lane, gx, gy, gth);
}
// members with label la or lb, ascending, into s.mem; returns the count (<= 32 by construction)
template <int N>
__device__ __forceinline__ int member_list(ClusterSm<N>& s, int16_t* mem, int n, int la, int lb, int lane, unsigned lt) {
    int cnt = 0;
    for (int base = 0; base < n; base += 32) {
        const int k = base + lane;
        const bool in = k < n && (s.label[k] == la || s.label[k] == lb);
        const unsigned m = __ballot_sync(FULL, in);
        if (in) { const int p = cnt + __popc(m & lt); if (p < 32) mem[p] = int16_t(k); }
        cnt += __popc(m);
    }
    __syncwarp();
    return cnt < 32 ? cnt : 32;
}

// Would row i merge now?  Nearest row of another cluster within the distance threshold whose cluster orientation is within the
// orientation threshold (first minimum, j ascending), and the two clusters together within the size limit
// (EdgeClusterer.cpp:143-166).  The partner is cached in s.rank[i].
template <int N>
__device__ __forceinline__ bool merge_candidate(ClusterSm<N>& s, int i, int nw, const PassConst& pc) {
    constexpr int NW = N / 32;
    const int li = s.label[i];
    const double oi = s.orient[li], xi = s.x[i], yi = s.y[i];
    double bd = DBL_MAX, bs = DBL_MAX;                     // best distance and its square
    int bj = -1;
    for (int ww = 0; ww < nw; ww++) {
        uint32_t bits = s.adj[i * NW + ww];
        while (bits) {
            const int b = __ffs(bits) - 1;
            bits &= bits - 1;
            const int j = ww * 32 + b;
            const int lj = s.label[j];
            if (lj != li && fabs(oi - s.orient[lj]) < pc.cl_orient_rad) {          // :147-154
                const double ddx = xi - s.x[j], ddy = yi - s.y[j];
                const double sq = ddx * ddx + ddy * ddy;
                if (sq < bs) {                                                      // sqrt is monotone: sq >= bs cannot give d < bd
                    const double d = sqrt(sq);
                    if (d < bd) { bd = d; bs = sq; bj = j; }                        // first minimum (j ascending)
                }
            }
        }
    }
    const bool found = bj >= 0 && int(s.size[s.label[bj]]) + int(s.size[li]) <= pc.max_cluster;   // :166
    s.rank[i] = int16_t(found ? bj : -1);
    return found;
}

template <int N>
__device__ void cluster_warp_sm(int n, const PassConst& pc, ClusterSm<N>& s, const double* __restrict__ xc, const double* __restrict__ yc,
                                const double* __restrict__ th, const int32_t* __restrict__ h2orig, double* xbar, double* ybar, double* thbar,
                                int32_t* label_out, int32_t* h2rep, int32_t* urank, int32_t* n_clusters, int lane) {
    constexpr int NW = N / 32;
    const unsigned lt = lanemask_lt();
    const int nw = (n + 31) >> 5;
    for (int k = lane; k < n; k += 32) {
        const double t = normalize_orientation(th[k], pc.cl_wrap_deg);   // :131-135
        s.x[k] = xc[k]; s.y[k] = yc[k]; s.thn[k] = t; s.orient[k] = t; s.label[k] = int16_t(k); s.size[k] = 1;
    }
    __syncwarp();
    for (int i = lane; i < n; i += 32) {                                  // adjacency: dist(i, j) < CLUSTER_DIST_THRESH
        const double xi = s.x[i], yi = s.y[i];
        for (int w = 0; w < nw; w++) {
            uint32_t bits = 0;
            const int jend = min(32, n - w * 32);
            for (int b = 0; b < jend; b++) {
                const int j = w * 32 + b;
                const double ddx = xi - s.x[j], ddy = yi - s.y[j];
                if (ddx * ddx + ddy * ddy < pc.cl_dist_sq && j != i) bits |= (1u << b);   // == sqrt(.) < cl_dist, see cl_dist_sq
            }
            s.adj[i * NW + w] = bits;
        }
    }
    __syncwarp();
    // :139-189.  The reference restarts its scan after every merge and takes the first row that can merge.  A row's decision
    // depends only on its own cluster and on the clusters of its neighbours (adjacency), so after a merge only the members of
    // the merged cluster and the rows adjacent to them are re-evaluated; every other row keeps its cached decision and partner.
    // The first set bit of the "can merge" mask is exactly the row the restarted scan would reach first.
    for (int w = lane; w < nw; w += 32) { const int rem = n - w * 32; s.dirty[w] = rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u); s.canw[w] = 0u; }
    __syncwarp();
    while (true) {
        for (int w = 0; w < nw; w++) {
            const uint32_t dw = s.dirty[w];
            if (!dw) continue;
            const int i = w * 32 + lane;
            const bool found = ((dw >> lane) & 1u) && merge_candidate<N>(s, i, nw, pc);
            const unsigned m = __ballot_sync(FULL, found);
            if (lane == 0) { s.canw[w] = (s.canw[w] & ~dw) | m; s.dirty[w] = 0u; }
        }
        __syncwarp();
        int mi = -1;
        for (int w = 0; w < nw; w++) { const uint32_t cw = s.canw[w]; if (cw) { mi = w * 32 + __ffs(cw) - 1; break; } }
        if (mi < 0) break;
        const int mj = s.rank[mi];
        const int old_l = s.label[mj], new_l = s.label[mi];
        const int cnt = member_list(s, s.mem, n, old_l, new_l, lane, lt);
        double gx, gy, gth;
        gauss_avg_list<N, true>(s, s.mem, cnt, pc.cl_sigma, pc.cl_sigma_inv, lane, &gx, &gy, &gth);       // :168
        if (lane == 0) { s.orient[new_l] = normalize_orientation(gth, pc.cl_wrap_deg); s.size[new_l] = int16_t(cnt); }   // :175-176
        for (int k = lane; k < n; k += 32) if (s.label[k] == old_l) s.label[k] = int16_t(new_l);   // :179-183
        for (int w = lane; w < nw; w += 32) {                            // members + everything adjacent to a member
            uint32_t d = 0u;
            for (int k = 0; k < cnt; k++) { const int mrow = s.mem[k]; d |= s.adj[mrow * NW + w]; if ((mrow >> 5) == w) d |= 1u << (mrow & 31); }
            s.dirty[w] = d;
        }
        __syncwarp();
    }
    // cluster ranks (clusters in ascending label order = the reference's std::map order); a label in use is the index
    // of one of its members, so the cluster roots are the rows with label[k] == k
    int nroots = 0;
    for (int base = 0; base < n; base += 32) {
        const int k = base + lane;
        const bool root = k < n && s.label[k] == k;
        const unsigned m = __ballot_sync(FULL, root);
        if (root) s.rank[k] = int16_t(nroots + __popc(m & lt));
        nroots += __popc(m);
    }
    __syncwarp();
    // :250-290 per cluster: Gaussian-weighted average written to every member, representative = member closest to it.
    // A one-member cluster averages to its member exactly (weight exp(-0) = 1, x / 1 = x): written directly, lane-parallel.
    for (int k = lane; k < n; k += 32)
        if (s.label[k] == k && s.size[k] == 1) { xbar[k] = s.x[k]; ybar[k] = s.y[k]; thbar[k] = s.thn[k]; h2rep[k] = h2orig[k]; }
    for (int base = 0; base < n; base += 32) {
        const int kk = base + lane;
        unsigned roots = __ballot_sync(FULL, kk < n && s.label[kk] == kk && s.size[kk] > 1);
        while (roots) {
            const int i = base + __ffs(roots) - 1;
            roots &= roots - 1;
            const int cnt = member_list(s, s.mem, n, i, i, lane, lt);
            double gx, gy, gth;
            gauss_avg_list<N, false>(s, s.mem, cnt, pc.cl_sigma, pc.cl_sigma_inv, lane, &gx, &gy, &gth);
            double d = DBL_MAX;
            int pos = lane;
            if (lane < cnt) {
                const int j = s.mem[lane];
                const double ddx = s.x[j] - gx, ddy = s.y[j] - gy;
                d = sqrt(ddx * ddx + ddy * ddy);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {                            // first minimum in member order (:270)
                const double od = __shfl_xor_sync(FULL, d, o);
                const int op = __shfl_xor_sync(FULL, pos, o);
                if (od < d || (od == d && op < pos)) { d = od; pos = op; }
            }
            const int rep = h2orig[s.mem[pos]];                           // corrected(closest, 8), :282
            if (lane < cnt) {
                const int j = s.mem[lane];
                xbar[j] = gx; ybar[j] = gy; thbar[j] = gth; h2rep[j] = rep;
            }
            __syncwarp();
        }
    }
    for (int k = lane; k < n; k += 32) { const int l = s.label[k]; label_out[k] = l; urank[k] = s.rank[l]; }
    if (lane == 0) *n_clusters = nroots;
}

// ---- the same clustering by W warps per view-1 edge (the > 128-row classes).  The merge sequence is inherently serial, but every
// merge is made of data-parallel pieces: re-evaluating the dirty rows (words of 32 rows strided over the warps), collecting the
// member mask, relabelling, and the dirty-set update; the Gaussian average stays on one warp and runs beside the relabel (other
// warps) and the dirty-set update (warp 1).  Three named barriers per merge.  Results are identical to cluster_warp_sm: the same
// decisions in the same order, the same arithmetic per member in the same order.
template <int N>
struct ClusterSmB {
    ClusterSm<N> s;
    uint32_t wmask[N / 32];       // member mask of the merge in flight / root mask in the final phase
    int16_t memw[8][32];          // per-warp member lists of the final per-cluster averages
};
static_assert(CL_SMALL_WARPS * sizeof(ClusterSm<CL_N>) <= CL_SMEM_QUANTUM && CL_MID_EDGES * sizeof(ClusterSmB<CL_MID_N>) <= CL_SMEM_QUANTUM &&
              sizeof(ClusterSmB<CL_BIG_N>) <= CL_SMEM_QUANTUM, "clustering classes must fit the common shared-memory size");
__device__ __forceinline__ void bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }
template <int N, int W>
__device__ void cluster_block_sm(int n, const PassConst& pc, ClusterSmB<N>& sb, const double* __restrict__ xc, const double* __restrict__ yc,
                                 const double* __restrict__ th, const int32_t* __restrict__ h2orig, double* xbar, double* ybar, double* thbar,
                                 int32_t* label_out, int32_t* h2rep, int32_t* urank, int32_t* n_clusters, int gw, int lane, int bar_id) {
    static_assert(W >= 4 && W <= 8 && N / 32 <= 32, "cluster_block_sm: 4..8 warps, at most 32 row words");
    constexpr int NW = N / 32, T = W * 32;
    ClusterSm<N>& s = sb.s;
    const unsigned lt = lanemask_lt();
    const int t = gw * 32 + lane;
    const int nw = (n + 31) >> 5;
    for (int k = t; k < n; k += T) {
        const double o = normalize_orientation(th[k], pc.cl_wrap_deg);   // :131-135
        s.x[k] = xc[k]; s.y[k] = yc[k]; s.thn[k] = o; s.orient[k] = o; s.label[k] = int16_t(k); s.size[k] = 1;
    }
    if (t < nw) { const int rem = n - t * 32; s.dirty[t] = rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u); s.canw[t] = 0u; }
    bar_sync(bar_id, T);
    for (int i = t; i < n; i += T) {                                      // adjacency: dist(i, j) < CLUSTER_DIST_THRESH
        const double xi = s.x[i], yi = s.y[i];
        for (int w = 0; w < nw; w++) {
            uint32_t bits = 0;
            const int jend = min(32, n - w * 32);
            for (int b = 0; b < jend; b++) {
                const int j = w * 32 + b;
                const double ddx = xi - s.x[j], ddy = yi - s.y[j];
                if (ddx * ddx + ddy * ddy < pc.cl_dist_sq && j != i) bits |= (1u << b);
            }
            s.adj[i * NW + w] = bits;
        }
    }
    bar_sync(bar_id, T);
    while (true) {
        for (int w = gw; w < nw; w += W) {                                // dirty rows, a word of 32 rows per warp step
            const uint32_t dw = s.dirty[w];
            if (!dw) continue;
            const int i = w * 32 + lane;
            const bool found = ((dw >> lane) & 1u) && merge_candidate<N>(s, i, nw, pc);
            const unsigned m = __ballot_sync(FULL, found);
            if (lane == 0) { s.canw[w] = (s.canw[w] & ~dw) | m; s.dirty[w] = 0u; }
        }
        bar_sync(bar_id, T);
        // first row that can merge (every warp finds the same one)
        const uint32_t cw = lane < nw ? s.canw[lane] : 0u;
        const unsigned nz = __ballot_sync(FULL, cw != 0u);
        if (!nz) break;
        const int fw = __ffs(nz) - 1;
        const int mi = fw * 32 + __ffs(__shfl_sync(FULL, cw, fw)) - 1;
        const int mj = s.rank[mi];
        const int old_l = s.label[mj], new_l = s.label[mi];
        for (int w = gw; w < nw; w += W) {                                // members of the two clusters, a mask word per warp step
            const int k = w * 32 + lane;
            const int l = k < n ? int(s.label[k]) : -1;
            const unsigned m = __ballot_sync(FULL, l == old_l || l == new_l);
            if (lane == 0) sb.wmask[w] = m;
        }
        bar_sync(bar_id, T);
        if (gw == 0) {                                                    // ascending member list, Gaussian average (:168-176)
            const uint32_t wm = lane < nw ? sb.wmask[lane] : 0u;
            int incl = __popc(wm);
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += v; }
            int pos = incl - __popc(wm);
            uint32_t bits = wm;
            while (bits) {
                const int b = __ffs(bits) - 1;
                bits &= bits - 1;
                if (pos < 32) s.mem[pos] = int16_t(lane * 32 + b);
                pos++;
            }
            const int cnt = min(32, __shfl_sync(FULL, incl, 31));
            __syncwarp();
            double gx, gy, gth;
            gauss_avg_list<N, true>(s, s.mem, cnt, pc.cl_sigma, pc.cl_sigma_inv, lane, &gx, &gy, &gth);
            if (lane == 0) { s.orient[new_l] = normalize_orientation(gth, pc.cl_wrap_deg); s.size[new_l] = int16_t(cnt); }
        } else if (gw == 1) {                                             // dirty set: the members and every row adjacent to one
            uint32_t d = lane < nw ? sb.wmask[lane] : 0u;
            for (int ww = 0; ww < nw; ww++) {
                uint32_t bits = sb.wmask[ww];
                while (bits) {
                    const int mrow = ww * 32 + __ffs(bits) - 1;
                    bits &= bits - 1;
                    if (lane < nw) d |= s.adj[mrow * NW + lane];
                }
            }
            if (lane < nw) s.dirty[lane] = d;
        } else {                                                          // relabel (:179-183)
            for (int w = gw - 2; w < nw; w += W - 2)
                if ((sb.wmask[w] >> lane) & 1u) s.label[w * 32 + lane] = int16_t(new_l);
        }
        bar_sync(bar_id, T);
    }
    // ranks of the cluster roots (label[k] == k) in ascending label order
    for (int w = gw; w < nw; w += W) {
        const int k = w * 32 + lane;
        const unsigned m = __ballot_sync(FULL, k < n && s.label[k] == k);
        if (lane == 0) sb.wmask[w] = m;
    }
    bar_sync(bar_id, T);
    const uint32_t rm = lane < nw ? sb.wmask[lane] : 0u;
    int rincl = __popc(rm);
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(FULL, rincl, o); if (lane >= o) rincl += v; }
    const int nroots = __shfl_sync(FULL, rincl, 31);
    const int rexcl = rincl - __popc(rm);
    for (int w = gw; w < nw; w += W) {
        const int base = __shfl_sync(FULL, rexcl, w);
        const uint32_t m = __shfl_sync(FULL, rm, w);
        if ((m >> lane) & 1u) s.rank[w * 32 + lane] = int16_t(base + __popc(m & lt));
    }
    // :250-290 per cluster (see cluster_warp_sm): one-member clusters directly, the others a warp each
    for (int k = t; k < n; k += T)
        if (s.label[k] == k && s.size[k] == 1) { xbar[k] = s.x[k]; ybar[k] = s.y[k]; thbar[k] = s.thn[k]; h2rep[k] = h2orig[k]; }
    int16_t* mem = sb.memw[gw];
    for (int w = gw; w < nw; w += W) {
        const int kk = w * 32 + lane;
        unsigned roots = __ballot_sync(FULL, kk < n && s.label[kk] == kk && s.size[kk] > 1);
        while (roots) {
            const int i = w * 32 + __ffs(roots) - 1;
            roots &= roots - 1;
            const int cnt = member_list(s, mem, n, i, i, lane, lt);
            double gx, gy, gth;
            gauss_avg_list<N, false>(s, mem, cnt, pc.cl_sigma, pc.cl_sigma_inv, lane, &gx, &gy, &gth);
            double d = DBL_MAX;
            int pos = lane;
            if (lane < cnt) {
                const int j = mem[lane];
                const double ddx = s.x[j] - gx, ddy = s.y[j] - gy;
                d = sqrt(ddx * ddx + ddy * ddy);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {                            // first minimum in member order (:270)
                const double od = __shfl_xor_sync(FULL, d, o);
                const int op = __shfl_xor_sync(FULL, pos, o);
                if (od < d || (od == d && op < pos)) { d = od; pos = op; }
            }
            const int rep = h2orig[mem[pos]];                             // corrected(closest, 8), :282
            if (lane < cnt) {
                const int j = mem[lane];
                xbar[j] = gx; ybar[j] = gy; thbar[j] = gth; h2rep[j] = rep;
            }
            __syncwarp();
        }
    }
    bar_sync(bar_id, T);                                                  // ranks complete
    for (int k = t; k < n; k += T) { const int l = s.label[k]; label_out[k] = l; urank[k] = s.rank[l]; }
    if (t == 0) *n_clusters = nroots;
}

// ---- global-memory fallback for groups larger than CL_N rows (same semantics, the plain sequential scan order)
__device__ void gauss_avg_warp(int n, const double* xc, const double* yc, const double* thn, const int32_t* label, int l1, int l2,
                               double sigma, double* gx, double* gy, double* gth) {
    double sx = 0, sy = 0;
    int cnt = 0;
    for (int i = 0; i < n; i++) {
        const int l = label[i];
        if (l == l1 || l == l2) { sx += xc[i]; sy += yc[i]; cnt++; }
    }
    const double cx = sx / cnt, cy = sy / cnt;
    double tot = 0.0;
    for (int i = 0; i < n; i++) {
        const int l = label[i];
        if (l == l1 || l == l2) { const double dx = xc[i] - cx, dy = yc[i] - cy; tot += sqrt(dx * dx + dy * dy); }
    }
    const double mean_shift = tot / cnt;
    double swx = 0, swy = 0, swo = 0, tw = 0;
    for (int i = 0; i < n; i++) {
        const int l = label[i];
        if (l == l1 || l == l2) {
            const double dx = xc[i] - cx, dy = yc[i] - cy;
            const double dist = sqrt(dx * dx + dy * dy);
            const double q = (dist - mean_shift) / sigma;
            const double w = exp(-0.5 * (q * q));
            swx += w * xc[i]; swy += w * yc[i]; swo += w * thn[i]; tw += w;
        }
    }
    *gx = swx / tw; *gy = swy / tw; *gth = swo / tw;
}
__device__ void cluster_warp_global(int n, const PassConst& pc, const double* xc, const double* yc, const double* th, double* thn, double* orient,
                                    int32_t* label, const int32_t* h2orig, double* xbar, double* ybar, double* thbar, int32_t* h2rep,
                                    int32_t* urank, int32_t* n_clusters, int lane) {
    const unsigned lt = lanemask_lt();
    for (int k = lane; k < n; k += 32) {
        const double t = normalize_orientation(th[k], pc.cl_wrap_deg);
        thn[k] = t; orient[k] = t; label[k] = k;
    }
    __syncwarp();
    bool merged = true;
    while (merged) {
        merged = false;
        for (int i = 0; i < n; i++) {
            const int li = label[i];
            const double oi = orient[li], xi = xc[i], yi = yc[i];
            double bd = DBL_MAX;
            int bj = -1;
            for (int j = lane; j < n; j += 32) {
                const int lj = label[j];
                if (lj != li) {
                    const double ddx = xi - xc[j], ddy = yi - yc[j];
                    const double dist = sqrt(ddx * ddx + ddy * ddy);
                    if (dist < bd && dist < pc.cl_dist && fabs(oi - orient[lj]) < pc.cl_orient_rad) { bd = dist; bj = j; }
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double od = __shfl_xor_sync(FULL, bd, o);
                const int oj = __shfl_xor_sync(FULL, bj, o);
                if (oj >= 0 && (bj < 0 || od < bd || (od == bd && oj < bj))) { bd = od; bj = oj; }
            }
            if (bj >= 0) {
                const int old_l = label[bj], new_l = li;
                int so = 0, sn = 0;
                for (int j = lane; j < n; j += 32) { const int l = label[j]; so += (l == old_l); sn += (l == new_l); }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) { so += __shfl_xor_sync(FULL, so, o); sn += __shfl_xor_sync(FULL, sn, o); }
                if (so + sn <= pc.max_cluster) {
                    double gx, gy, gth;
                    gauss_avg_warp(n, xc, yc, thn, label, old_l, new_l, pc.cl_sigma, &gx, &gy, &gth);
                    __syncwarp();
                    if (lane == 0) orient[new_l] = normalize_orientation(gth, pc.cl_wrap_deg);
                    for (int j = lane; j < n; j += 32) if (label[j] == old_l) label[j] = new_l;
                    __syncwarp();
                    merged = true;
                    break;
                }
            }
        }
    }
    int nroots = 0;
    for (int base = 0; base < n; base += 32) {
        const int k = base + lane;
        const bool root = k < n && label[k] == k;
        const unsigned m = __ballot_sync(FULL, root);
        if (root) urank[k] = nroots + __popc(m & lt);
        nroots += __popc(m);
    }
    __syncwarp();
    for (int i = 0; i < n; i++) {
        if (label[i] != i) continue;
        double gx, gy, gth;
        gauss_avg_warp(n, xc, yc, thn, label, i, i, pc.cl_sigma, &gx, &gy, &gth);
        double md = DBL_MAX;
        int closest = -1;
        for (int j = 0; j < n; j++) {                    // a cluster keeps the label of the row that absorbed it: members may precede the root
            if (label[j] == i) {
                const double ddx = xc[j] - gx, ddy = yc[j] - gy;
                const double dist = sqrt(ddx * ddx + ddy * ddy);
                if (dist < md) { md = dist; closest = j; }
            }
        }
        const int rep = h2orig[closest];
        for (int j = lane; j < n; j += 32)
            if (label[j] == i) { xbar[j] = gx; ybar[j] = gy; thbar[j] = gth; h2rep[j] = rep; }
    }
    __syncwarp();
    for (int k = lane; k < n; k += 32) { const int l = label[k]; if (l != k) urank[k] = urank[l]; }
    if (lane == 0) *n_clusters = nroots;
}

// ----------------------------------------------------------------------------- formation
// One warp per view-1 edge: Skip_this_Edge (hpp:92-110); getHYPO2_idx_Ore (PairEdgeHypo.cpp:88-102) as a bin-range read of the
// angle-sorted view-2 index with the reference's exact strict test; edgelsHYPO2_epipolar_correction (PairEdgeHypo.cpp:225-293);
// then clustering.  Clustering is order dependent, so the fill pass first restores the reference's candidate order
// (ascending view-2 index) by rank-sorting the gathered candidates in shared memory; edges with more candidates than the
// shared-memory list holds fall back to the ordered linear scan of the whole angle list.
#define FORM_CAND_CAP 512

// processes 32 candidates (lane: view-2 edge i, valid when in) of an ordered candidate stream; returns through mbase / nbase
template <bool FILL>
__device__ __forceinline__ void form_chunk(bool in, int i, const double* coef, const PassConst& pc, const double* __restrict__ x2,
                                           const double* __restrict__ y2, const double* __restrict__ th2, int h1, int n_total, int64_t row0,
                                           RowTable& rt, unsigned lt, int& mbase, int& nbase) {
    const unsigned mm = __ballot_sync(FULL, in);
    if (mm == 0) return;
    bool ok = false;
    double cx = 0, cy = 0, t2 = 0;
    if (in) {
        t2 = th2[i];
        ok = correct_one(coef, x2[i], y2[i], t2, pc.loc_pert, pc.tang_thresh, pc.orient_pert, &cx, &cy);
    }
    const unsigned mo = __ballot_sync(FULL, ok);
    if (FILL && in) {
        const int k = mbase + __popc(mm & lt);                           // pre-correction position (col 9)
        if (k < n_total) rt.rawat[row0 + k] = i;                         // HYPO2_idx_raw(local row): EdgeClusterer.cpp:221 quirk
        if (ok) {
            const int64_t r = row0 + nbase + __popc(mo & lt);
            rt.h1[r] = h1; rt.h2orig[r] = i; rt.local[r] = k;
            rt.xc[r] = cx; rt.yc[r] = cy; rt.th[r] = t2;                 // theta written back is the ORIGINAL one (:288)
        }
    }
    mbase += __popc(mm); nbase += __popc(mo);
}

template <bool FILL>
__global__ void __launch_bounds__(FORM_WARPS * 32) k_form(const __grid_constant__ PassConst pc, const double* __restrict__ x1,
                                                          const double* __restrict__ y1, int h1_begin, int h1_end,
                                                          const double2* __restrict__ wedge_h1, const double* __restrict__ x2,
                                                          const double* __restrict__ y2, const double* __restrict__ th2,
                                                          const double* __restrict__ ang2, int E2, const AngleIndex ix, int32_t* m_out,
                                                          int32_t* n_out, const int64_t* __restrict__ row_off, RowTable rt, int32_t* n_clusters) {
    extern __shared__ __align__(16) unsigned char form_smem[];
    const int lane = threadIdx.x & 31;
    const int h1 = h1_begin + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (h1 >= h1_end) return;
    const double x = x1[h1], y = y1[h1];
    const bool skip = (x < 10 || x > double(pc.cols - 10) || y < 10 || y > double(pc.rows - 10));   // hpp:94
    const double2 w = wedge_h1[h1];
    int n_total = 0, m_total = 0;
    int64_t row0 = 0;
    if (FILL) {
        n_total = n_out[h1];
        if (n_total == 0) { if (lane == 0) n_clusters[h1] = 0; return; }
        m_total = m_out[h1];
        row0 = row_off[h1];
    } else if (skip || w.x != w.x) {
        if (lane == 0) { m_out[h1] = 0; n_out[h1] = 0; }
        return;
    }
    double coef[3];
    epip_line_coeffs(pc.F21, x, y, coef);
    const unsigned lt = lanemask_lt();
    int mbase = 0, nbase = 0;
    // bin range of the wedge in the angle-sorted index (a superset; membership is the exact strict test below)
    const int b0 = angle_bin(w.x, ix.scale, ix.B), b1 = angle_bin(w.y, ix.scale, ix.B);
    const int beg = b0 ? ix.bin_end[b0 - 1] : 0, end = ix.bin_end[b1];
    if (!FILL) {
        // counting needs no order
        for (int base = beg; base < end; base += 32) {
            const int j = base + lane;
            bool in = false;
            int i = 0;
            if (j < end) { const double a = ix.ang[j]; in = (a > w.x && a < w.y); i = ix.idx[j]; }   // PairEdgeHypo.cpp:94 (strict both sides)
            form_chunk<false>(in, i, coef, pc, x2, y2, th2, h1, 0, 0, rt, lt, mbase, nbase);
        }
        if (lane == 0) { m_out[h1] = mbase; n_out[h1] = nbase; }
        return;
    }
    if (m_total <= FORM_CAND_CAP) {
        // gather (unordered), rank-sort by view-2 index, then the reference's ordered pass over the candidates
        int32_t* cand = reinterpret_cast<int32_t*>(form_smem + size_t(threadIdx.x >> 5) * (2 * FORM_CAND_CAP * sizeof(int32_t)));
        int32_t* sorted = cand + FORM_CAND_CAP;
        int m = 0;
        for (int base = beg; base < end; base += 32) {
            const int j = base + lane;
            bool in = false;
            if (j < end) { const double a = ix.ang[j]; in = (a > w.x && a < w.y); }
            const unsigned mm = __ballot_sync(FULL, in);
            if (in) cand[m + __popc(mm & lt)] = ix.idx[j];
            m += __popc(mm);
        }
        __syncwarp();
        for (int c = lane; c < m; c += 32) {
            const int mine = cand[c];
            int rank = 0;
            for (int k = 0; k < m; k++) rank += (cand[k] < mine);
            sorted[rank] = mine;
        }
        __syncwarp();
        for (int base = 0; base < m; base += 32) {
            const int c = base + lane;
            const bool in = c < m;
            form_chunk<true>(in, in ? sorted[c] : 0, coef, pc, x2, y2, th2, h1, n_total, row0, rt, lt, mbase, nbase);
        }
    } else {
        for (int base = 0; base < E2; base += 32) {
            const int i = base + lane;
            bool in = false;
            if (i < E2) { const double a = ang2[i]; in = (a > w.x && a < w.y); }
            form_chunk<true>(in, i, coef, pc, x2, y2, th2, h1, n_total, row0, rt, lt, mbase, nbase);
        }
    }
}
// ---- clustering as its own launches (the formation kernel stays small enough for the instruction cache, and the clustering
// kernels see the view-1 edges in DESCENDING order of their hypothesis count, so the long sequential merges start first).
// k_order_by_n: one block; order[] = edges with 1 <= n <= CL_N sorted by n descending, then those with CL_N < n <= CL_BIG_N, then
// then CL_MID_N < n <= CL_BIG_N, then CL_BIG_N < n <= CL_HUGE_N, then the rest; cnts[0..4] = the five counts.
#define ORDER_BUCKETS (CL_N + 4)
__global__ void __launch_bounds__(1024) k_order_by_n(const int32_t* __restrict__ n_of, int h1_begin, int h1_end, const int32_t* __restrict__ end_dev,
                                                     int max_cluster, int small_max, int32_t* __restrict__ order, int32_t* __restrict__ cnts) {
    __shared__ int hist[ORDER_BUCKETS];
    if (end_dev) h1_end = min(h1_end, *end_dev);          // a count that only the device knows (virtual edges of the run-based clustering)
    for (int k = threadIdx.x; k < ORDER_BUCKETS; k += blockDim.x) hist[k] = 0;
    __syncthreads();
    const bool sm_ok = max_cluster <= 32;          // the shared-memory algorithm keeps member lists of at most 32
    auto key = [&](int n) { return (!sm_ok || n > CL_HUGE_N) ? CL_N + 3 : n > CL_BIG_N ? CL_N + 2 : n > CL_MID_N ? CL_N + 1 : n > small_max ? CL_N : CL_N - n; };
    for (int i = h1_begin + threadIdx.x; i < h1_end; i += blockDim.x) { const int n = n_of[i]; if (n > 0) atomicAdd(&hist[key(n)], 1); }
    __syncthreads();
    if (threadIdx.x == 0) {
        int run = 0, ns = 0;
        for (int k = 0; k < ORDER_BUCKETS; k++) { const int c = hist[k]; hist[k] = run; run += c; if (k == CL_N - 1) ns = run; }
        cnts[0] = ns; cnts[1] = hist[CL_N + 1] - ns; cnts[2] = hist[CL_N + 2] - hist[CL_N + 1];
        cnts[3] = hist[CL_N + 3] - hist[CL_N + 2]; cnts[4] = run - hist[CL_N + 3];
        cnts[5] = run;                             // view-1 edges with at least one hypothesis
    }
    __syncthreads();
    for (int i = h1_begin + threadIdx.x; i < h1_end; i += blockDim.x) { const int n = n_of[i]; if (n > 0) order[atomicAdd(&hist[key(n)], 1)] = i; }
}
// ---- clustering by RUNS (the > CL_N-row classes).  Every corrected candidate of a view-1 edge lies on that edge's epipolar line
// (snapped to it or moved along its tangent onto it, PairEdgeHypo.cpp:243-282), so "dist < CLUSTER_DIST_THRESH" links only
// candidates that are neighbours along the line: sorted by their position t along the line, the candidates fall into runs separated by
// gaps wider than the threshold, and no merge of EdgeClusterer.cpp:138-189 ever crosses a gap.  A row's merge decision reads only
// its own cluster and the clusters of the rows within the distance threshold, so the reference's restarted scan, restricted to a run, is
// the reference's scan of that run alone: the runs are clustered independently, each in the reference's order (~48 runs of ~5 rows
// per edge on cfg4, against one sequential chain of ~250 rows) -- thousands of independent chains in flight instead of one per
// block.  Same decisions in the same order and the same per-member arithmetic as cluster_warp_sm.
//   k_runs_prepare  block per edge: position along the line, bitonic sort, run boundaries, rows of a run back into ascending row
//                   order.  Single-row runs are finished here; runs of 2..32 rows go to a work list; longer runs are copied, as
//                   "virtual edges", into a scratch hypothesis table
//   k_runs_cluster  warp per short run, the run's state in 5 lanes-worth of registers and 1.3 KB of shared memory
//   (long runs)     the per-edge kernels below (k_order_by_n, k_cluster_classes, k_cluster_coop) on the scratch table, then
//   k_runs_scatter  results of the long runs back to their rows
//   k_runs_finish   warp per edge: cluster ranks in ascending label order, cluster count
// An edge whose rows are not collinear to 1e-6 px or not finite is left to the per-edge kernels (n_clusters = -1).
#define RUN_NMAX CL_HUGE_N     // rows of an edge
#define RUN_SHORT 32           // rows of a run handled by one warp in registers
#define RUN_WARPS 8
#define RUN_START_FLAG (1 << 30)   // run_ord entry: first row of a run inside its pack
struct RunPrepSm {
    double key[RUN_NMAX];                  // position along the epipolar line (sort key); reused as scratch afterwards
    int16_t ord[RUN_NMAX];                 // rows sorted by key; then, run by run, by ascending row index
    int16_t start[RUN_NMAX + 1];           // first sorted position of every run
    int16_t pbase[RUN_NMAX];               // short runs: first position in the edge's compacted list of pack rows
    int2 pack[RUN_NMAX / 2];               // packs of the edge: (first compacted position, rows)
    int wcnt[RUN_WARPS];
    int bad, n_long, n_short, rows_long, base_long, base_short, base_rows;
    double lineA, lineB, lineC, lineInv;
};
__global__ void __launch_bounds__(RUN_WARPS * 32) k_runs_prepare(const __grid_constant__ PassConst pc, const int32_t* __restrict__ order,
                                                                  const int32_t* __restrict__ cnts, const double* __restrict__ x1,
                                                                  const double* __restrict__ y1, const int32_t* __restrict__ n_of,
                                                                  const int64_t* __restrict__ row_off, RowTable rt, int32_t* n_clusters,
                                                                  const __grid_constant__ RunScratch rs) {
    __shared__ RunPrepSm s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt = lanemask_lt();
    const int base = cnts[0], total = cnts[1] + cnts[2] + cnts[3];        // the CL_N < n <= CL_HUGE_N classes
    int32_t* run_ord = rt.urank;                                          // scratch until k_runs_finish writes the ranks
    for (int e = blockIdx.x; e < total; e += gridDim.x) {
        const int h1 = order[base + total - 1 - e];                       // longest edges first
        const int n = n_of[h1];
        const int64_t row0 = row_off[h1];
        __syncthreads();                                                  // the previous edge is done with the shared memory
        if (tid == 0) {
            double coef[3];
            mat_vec(pc.F21, x1[h1], y1[h1], 1.0, coef);                   // the epipolar line of the view-1 edge: A x + B y + C = 0
            s.lineA = coef[0]; s.lineB = coef[1]; s.lineC = coef[2]; s.lineInv = 1.0 / sqrt(coef[0] * coef[0] + coef[1] * coef[1]);
            s.bad = 0; s.n_long = 0; s.n_short = 0; s.rows_long = 0;
        }
        __syncthreads();
        int P = 32;
        while (P < n) P <<= 1;
        {
            const double A = s.lineA, B = s.lineB, C = s.lineC, inv = s.lineInv;
            int bad = 0;
            for (int k = tid; k < P; k += blockDim.x) {
                if (k < n) {
                    const double xk = rt.xc[row0 + k], yk = rt.yc[row0 + k];
                    const double t = (B * xk - A * yk) * inv, res = fabs((A * xk + B * yk) + C) * inv;
                    s.key[k] = t;
                    if (!(res <= 1e-6) || !(fabs(t) < 1e300)) bad = 1;
                } else {
                    s.key[k] = DBL_MAX;
                }
                s.ord[k] = int16_t(k);
            }
            if (bad) s.bad = 1;
        }
        __syncthreads();
        if (s.bad || pc.max_cluster > 32) { if (tid == 0) n_clusters[h1] = -1; continue; }
        for (int kk = 2; kk <= P; kk <<= 1) {                             // bitonic sort of (key, row) over P slots
            for (int j = kk >> 1; j > 0; j >>= 1) {
                for (int i = tid; i < P; i += blockDim.x) {
                    const int l = i ^ j;
                    if (l > i) {
                        const double a = s.key[i], b = s.key[l];
                        const bool up = (i & kk) == 0;
                        if ((a > b) == up) { s.key[i] = b; s.key[l] = a; const int16_t t = s.ord[i]; s.ord[i] = s.ord[l]; s.ord[l] = t; }
                    }
                }
                __syncthreads();
            }
        }
        // run starts: a gap wider than the distance threshold (plus the collinearity tolerance) cannot be bridged by any pair
        int run_total = 0;
        for (int b0 = 0; b0 < P; b0 += blockDim.x) {
            const int k = b0 + tid;
            const bool st = k < n && (k == 0 || s.key[k] - s.key[k - 1] > pc.cl_dist + 4e-6);
            const unsigned m = __ballot_sync(FULL, st);
            if (lane == 0) s.wcnt[warp] = __popc(m);
            __syncthreads();
            int off = run_total, tot = 0;
            for (int w = 0; w < RUN_WARPS; w++) { const int cw = s.wcnt[w]; if (w < warp) off += cw; tot += cw; }
            if (st) s.start[off + __popc(m & lt)] = int16_t(k);
            run_total += tot;
            __syncthreads();
        }
        if (tid == 0) s.start[run_total] = int16_t(n);
        __syncthreads();
        {
            int nl = 0, rl = 0;
            for (int r = tid; r < run_total; r += blockDim.x) {
                const int c = s.start[r + 1] - s.start[r];
                if (c > RUN_SHORT) { nl++; rl += c; }
            }
            if (nl) { atomicAdd(&s.n_long, nl); atomicAdd(&s.rows_long, rl); }
        }
        if (tid == 0) {
            // packs: consecutive short runs (2 .. RUN_SHORT rows; single rows and long runs are skipped) while they fit one warp
            int pos = 0, p0 = 0, prow = 0, np = 0;
            for (int r = 0; r < run_total; r++) {
                const int c = s.start[r + 1] - s.start[r];
                if (c < 2 || c > RUN_SHORT) continue;
                if (prow + c > RUN_SHORT) { s.pack[np++] = make_int2(p0, prow); p0 = pos; prow = 0; }
                s.pbase[r] = int16_t(pos);
                pos += c; prow += c;
            }
            if (prow) s.pack[np++] = make_int2(p0, prow);
            s.n_short = np;
        }
        __syncthreads();
        if (tid == 0) {
            s.base_long = s.n_long ? atomicAdd(&rs.ctr[0], s.n_long) : 0;
            s.base_short = s.n_short ? atomicAdd(&rs.ctr[1], s.n_short) : 0;
            s.base_rows = s.rows_long ? atomicAdd(&rs.ctr[2], s.rows_long) : 0;
            if (s.base_long + s.n_long > rs.cap_long || int64_t(s.base_rows) + s.rows_long > rs.cap_rows) s.bad = 1;   // cannot happen with the sizes of cvx_api.cu
            s.n_long = 0; s.rows_long = 0;
        }
        int16_t* tmp = reinterpret_cast<int16_t*>(s.key);                // the keys are dead from here on
        for (int k = tid; k < n; k += blockDim.x) tmp[k] = s.ord[k];
        __syncthreads();
        if (s.bad) { if (tid == 0) n_clusters[h1] = -1; continue; }
        for (int r = warp; r < run_total; r += RUN_WARPS) {               // rows of a run into ascending row order; work-list entries
            const int s0 = s.start[r], c = s.start[r + 1] - s0;
            if (c == 1) {
                if (lane == 0) {
                    const int i = tmp[s0];
                    rt.label[row0 + i] = i;
                    rt.xbar[row0 + i] = rt.xc[row0 + i]; rt.ybar[row0 + i] = rt.yc[row0 + i];
                    rt.thbar[row0 + i] = normalize_orientation(rt.th[row0 + i], pc.cl_wrap_deg);
                    rt.h2rep[row0 + i] = rt.h2orig[row0 + i];
                }
                continue;
            }
            int vid = 0, voff = 0;
            if (c > RUN_SHORT) {                                          // a virtual edge of the scratch table
                if (lane == 0) { vid = s.base_long + atomicAdd(&s.n_long, 1); voff = s.base_rows + atomicAdd(&s.rows_long, c); }
                vid = __shfl_sync(FULL, vid, 0); voff = __shfl_sync(FULL, voff, 0);
            }
            for (int k = lane; k < c; k += 32) {
                const int mine = tmp[s0 + k];
                int rk = 0;
                for (int j = 0; j < c; j++) rk += (tmp[s0 + j] < mine);
                if (c > RUN_SHORT) {
                    const int64_t q = int64_t(voff) + rk;
                    rs.vrt.xc[q] = rt.xc[row0 + mine]; rs.vrt.yc[q] = rt.yc[row0 + mine]; rs.vrt.th[q] = rt.th[row0 + mine];
                    rs.vrt.h2orig[q] = rt.h2orig[row0 + mine]; rs.v_oi[q] = mine;
                } else {
                    run_ord[row0 + s.pbase[r] + rk] = mine | (rk == 0 ? RUN_START_FLAG : 0);
                }
            }
            if (lane == 0 && c > RUN_SHORT) { rs.desc_long[vid] = make_int4(h1, voff, c, 0); rs.v_n[vid] = c; rs.v_off[vid] = voff; }
        }
        for (int k = tid; k < s.n_short; k += blockDim.x) rs.desc_short[s.base_short + k] = make_int4(h1, s.pack[k].x, s.pack[k].y, 0);
        if (tid == 0) n_clusters[h1] = 0;                                 // marks "clustered by runs" for k_runs_finish
    }
}
// A PACK of short runs of one view-1 edge: up to 32 rows, lane k = row k of the pack, every run a contiguous SEGMENT of lanes in ascending
// row order.  Runs are independent (no pair of rows of different runs is within the distance threshold), so every segment performs its own
// next merge of the reference's scan in the same step of the loop below, and all clusters of the pack are averaged at once at the end:
// the step count is the largest merge count of a run instead of the sum over the runs.  Sums over the members of a cluster run in ascending
// member order on every lane of that cluster (the reference's left-to-right chain), members read through shuffles from a per-lane source.
struct RunWarpSm {
    double x[RUN_SHORT], y[RUN_SHORT], orient[RUN_SHORT];   // row location; per-label averaged orientation (label = lane of the cluster's first row)
    int16_t size[RUN_SHORT];
};
// sum of v over the lanes of mask m in ascending lane order, on every lane for its own mask; steps = warp-wide largest member count
__device__ __forceinline__ double seg_sum(double v, unsigned m, int steps, int lane) {
    double acc = 0.0;
    for (int it = 0; it < steps; it++) {
        const bool has = m != 0u;
        const int src = has ? __ffs(m) - 1 : lane;
        m &= m - 1u;
        const double t = shfl_d(v, src);
        if (has) acc += t;
    }
    return acc;
}
// EdgeClusterer.cpp:60-125 for the cluster whose member lanes are mm (cnt = popc(mm)), evaluated by every lane for its own cluster
template <bool ORIENT_ONLY>
__device__ __forceinline__ void gauss_avg_seg(unsigned mm, int cnt, int steps, double xi, double yi, double ti, double sigma, double sigma_inv, int lane,
                                              double* gx, double* gy, double* gth) {
    const int c1 = max(cnt, 1);
    const double sx = seg_sum(xi, mm, steps, lane), sy = seg_sum(yi, mm, steps, lane);
    const bool p2 = (c1 & (c1 - 1)) == 0;                                       // see gauss_avg_list_t
    const double ic = __hiloint2double((1023 - (31 - __clz(c1))) << 20, 0);
    const double cx = p2 ? sx * ic : sx / c1, cy = p2 ? sy * ic : sy / c1;
    const double dx = xi - cx, dy = yi - cy;
    const double di = sqrt(dx * dx + dy * dy);
    const double tot = seg_sum(di, mm, steps, lane);
    const double mean_shift = p2 ? tot * ic : tot / c1;
    const double q = sigma_inv != 0.0 ? (di - mean_shift) * sigma_inv : (di - mean_shift) / sigma;
    const double w = exp(-0.5 * (q * q));
    const double wo = w * ti;
    double swo = 0.0, tw = 0.0, swx = 0.0, swy = 0.0;
    const double wx = w * xi, wy = w * yi;
    unsigned m = mm;
    for (int it = 0; it < steps; it++) {
        const bool has = m != 0u;
        const int src = has ? __ffs(m) - 1 : lane;
        m &= m - 1u;
        const double a = shfl_d(wo, src), b = shfl_d(w, src);
        if (has) { swo += a; tw += b; }
        if (!ORIENT_ONLY) {
            const double c = shfl_d(wx, src), d = shfl_d(wy, src);
            if (has) { swx += c; swy += d; }
        }
    }
    if (!ORIENT_ONLY) { *gx = swx / tw; *gy = swy / tw; }
    *gth = swo / tw;
}
__device__ void cluster_pack_warp(RunWarpSm& s, int nrows, bool run_start, int oi, double xi, double yi, double thn, const PassConst& pc,
                                  const int32_t* __restrict__ h2orig, double* xbar, double* ybar, double* thbar, int32_t* h2rep, int32_t* label_out,
                                  int lane) {
    const bool me = lane < nrows;
    // my run: lanes [lo, hi)
    const unsigned starts = __ballot_sync(FULL, me && run_start) | 1u;
    const unsigned le = lanemask_lt() | (1u << lane);
    const int lo = 31 - __clz(starts & le);
    const unsigned above = starts & ~le;
    const int hi = above ? __ffs(above) - 1 : nrows;
    const unsigned segmask = me ? (((hi >= 32) ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u)) : 0u;
    uint32_t adj = 0u;
    for (int j = 0; j < nrows; j++) {
        const double ddx = xi - s.x[j], ddy = yi - s.y[j];
        if (ddx * ddx + ddy * ddy < pc.cl_dist_sq && j != lane) adj |= 1u << j;   // == sqrt(.) < cl_dist, see cl_dist_sq
    }
    adj &= segmask;
    int label = lane;
    const int maxdeg = __reduce_max_sync(FULL, __popc(adj));
    while (true) {
        // would my row merge now?  (EdgeClusterer.cpp:143-166; nearest row of another cluster, first minimum in ascending row order)
        int partner = -1;
        {
            const double o_i = s.orient[label];
            double bd = DBL_MAX, bs = DBL_MAX;
            uint32_t bits = adj;
            for (int it = 0; it < maxdeg; it++) {                         // every lane walks its own neighbours, in step (a neighbour's label
                const bool has = bits != 0u;                              // is read from its lane: the shuffle needs all lanes)
                const int j = has ? __ffs(bits) - 1 : lane;
                bits &= bits - 1u;
                const int lj = __shfl_sync(FULL, label, j);
                if (has && lj != label && fabs(o_i - s.orient[lj]) < pc.cl_orient_rad) {   // :147-154
                    const double ddx = xi - s.x[j], ddy = yi - s.y[j];
                    const double sq = ddx * ddx + ddy * ddy;
                    if (sq < bs) {
                        const double d = sqrt(sq);
                        if (d < bd) { bd = d; bs = sq; partner = j; }
                    }
                }
            }
        }
        const int lpartner = __shfl_sync(FULL, label, partner >= 0 ? partner : lane);
        if (partner >= 0 && int(s.size[lpartner]) + int(s.size[label]) > pc.max_cluster) partner = -1;   // :166
        const unsigned can = __ballot_sync(FULL, partner >= 0);
        if (!can) break;
        // the next merge of MY run: its first row that can merge (the row the reference's restarted scan reaches first)
        const unsigned cseg = can & segmask;
        const bool act = cseg != 0u;
        const int mi = act ? __ffs(cseg) - 1 : lane;
        const int pj = __shfl_sync(FULL, partner, mi);
        const int mj = act ? pj : lane;
        const int new_l = __shfl_sync(FULL, label, mi), old_l = __shfl_sync(FULL, label, mj);
        const bool in = act && (label == old_l || label == new_l);
        const unsigned mm = __ballot_sync(FULL, in) & segmask;           // members of the merged cluster of my run (0: my run rests)
        const int cnt = __popc(mm);
        const int steps = __reduce_max_sync(FULL, cnt);
        double gx, gy, gth;
        gauss_avg_seg<true>(mm, cnt, steps, xi, yi, thn, pc.cl_sigma, pc.cl_sigma_inv, lane, &gx, &gy, &gth);   // :168
        __syncwarp();
        if (act && lane == new_l) { s.orient[new_l] = normalize_orientation(gth, pc.cl_wrap_deg); s.size[new_l] = int16_t(cnt); }   // :175-176
        if (act && label == old_l) label = new_l;                         // :179-183
        __syncwarp();
    }
    // :250-290 every cluster at once: Gaussian-weighted average written to every member, representative = member closest to it
    const unsigned cm = __match_any_sync(FULL, me ? label : 32 + lane);
    const int cnt = me ? __popc(cm) : 0;
    const int steps = __reduce_max_sync(FULL, cnt);
    const int label_row = __shfl_sync(FULL, oi, label);                   // row (within the edge) of the label
    double gx = xi, gy = yi, gth = thn;
    int rep_oi = oi;
    if (steps > 1) {
        double ax, ay, ath;
        gauss_avg_seg<false>(cm, cnt, steps, xi, yi, thn, pc.cl_sigma, pc.cl_sigma_inv, lane, &ax, &ay, &ath);
        const double ddx = xi - ax, ddy = yi - ay;
        const double d = sqrt(ddx * ddx + ddy * ddy);
        double best = DBL_MAX;
        int rep = lane;
        unsigned m = cm;
        for (int it = 0; it < steps; it++) {                              // first minimum in member order (:270)
            const bool has = m != 0u;
            const int src = has ? __ffs(m) - 1 : lane;
            m &= m - 1u;
            const double dv = shfl_d(d, src);
            if (has && dv < best) { best = dv; rep = src; }
        }
        const int r_oi = __shfl_sync(FULL, oi, rep);
        if (cnt > 1) { gx = ax; gy = ay; gth = ath; rep_oi = r_oi; }
    }
    if (me) {
        label_out[oi] = label_row;
        xbar[oi] = gx; ybar[oi] = gy; thbar[oi] = gth;
        h2rep[oi] = h2orig[rep_oi];                                       // corrected(closest, 8), :282
    }
}
__global__ void __launch_bounds__(RUN_WARPS * 32, 4) k_runs_cluster(const __grid_constant__ PassConst pc, const int64_t* __restrict__ row_off,
                                                                     RowTable rt, const __grid_constant__ RunScratch rs) {
    __shared__ RunWarpSm sm[RUN_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    RunWarpSm& s = sm[warp];
    const int n_packs = rs.ctr[1];
    const int32_t* run_ord = rt.urank;
    const int stride = gridDim.x * RUN_WARPS;
    for (int r = blockIdx.x * RUN_WARPS + warp; r < n_packs; r += stride) {
        const int4 d = rs.desc_short[r];                                  // (view-1 edge, first compacted position, rows)
        const int64_t row0 = row_off[d.x];
        const int c = d.z;
        __syncwarp();
        int oi = 0;
        bool start = false;
        double xi = 0.0, yi = 0.0, t = 0.0;
        if (lane < c) {
            const int e = run_ord[row0 + d.y + lane];
            oi = e & (RUN_START_FLAG - 1); start = (e & RUN_START_FLAG) != 0;
            t = normalize_orientation(rt.th[row0 + oi], pc.cl_wrap_deg);   // :131-135
            xi = rt.xc[row0 + oi]; yi = rt.yc[row0 + oi];
        }
        s.x[lane] = xi; s.y[lane] = yi; s.orient[lane] = t; s.size[lane] = lane < c ? 1 : 0;
        __syncwarp();
        cluster_pack_warp(s, c, start, oi, xi, yi, t, pc, rt.h2orig + row0, rt.xbar + row0, rt.ybar + row0, rt.thbar + row0, rt.h2rep + row0,
                          rt.label + row0, lane);
    }
}
// results of the long runs (virtual edges of the scratch table) back to their rows: warp per run
__global__ void __launch_bounds__(RUN_WARPS * 32) k_runs_scatter(const int64_t* __restrict__ row_off, RowTable rt, const __grid_constant__ RunScratch rs) {
    const int lane = threadIdx.x & 31;
    const int n_long = rs.ctr[0];
    const int stride = gridDim.x * RUN_WARPS;
    for (int r = blockIdx.x * RUN_WARPS + (threadIdx.x >> 5); r < n_long; r += stride) {
        const int4 d = rs.desc_long[r];
        const int64_t row0 = row_off[d.x], q0 = d.y;
        for (int k = lane; k < d.z; k += 32) {
            const int64_t q = q0 + k;
            const int64_t i = row0 + rs.v_oi[q];
            rt.xbar[i] = rs.vrt.xbar[q]; rt.ybar[i] = rs.vrt.ybar[q]; rt.thbar[i] = rs.vrt.thbar[q]; rt.h2rep[i] = rs.vrt.h2rep[q];
            rt.label[i] = rs.v_oi[q0 + rs.vrt.label[q]];
        }
    }
}
// cluster ranks in ascending label order (the reference's std::map order) and the cluster count of every edge clustered by runs
__global__ void __launch_bounds__(RUN_WARPS * 32) k_runs_finish(const int32_t* __restrict__ order, const int32_t* __restrict__ cnts,
                                                                 const int32_t* __restrict__ n_of, const int64_t* __restrict__ row_off, RowTable rt,
                                                                 int32_t* n_clusters) {
    const int lane = threadIdx.x & 31;
    const unsigned lt = lanemask_lt();
    const int base = cnts[0], total = cnts[1] + cnts[2] + cnts[3];
    const int e = blockIdx.x * RUN_WARPS + (threadIdx.x >> 5);
    if (e >= total) return;
    const int h1 = order[base + e];
    if (n_clusters[h1] != 0) return;                                      // left to the per-edge kernels
    const int n = n_of[h1];
    const int64_t row0 = row_off[h1];
    int nroots = 0;
    for (int b0 = 0; b0 < n; b0 += 32) {
        const int k = b0 + lane;
        const bool root = k < n && rt.label[row0 + k] == k;
        const unsigned m = __ballot_sync(FULL, root);
        if (root) rt.urank[row0 + k] = nroots + __popc(m & lt);
        nroots += __popc(m);
    }
    __syncwarp();
    for (int k = lane; k < n; k += 32) { const int l = rt.label[row0 + k]; if (l != k) rt.urank[row0 + k] = rt.urank[row0 + l]; }
    if (lane == 0) n_clusters[h1] = nroots;
}
// The <= 128 / 256 / 512-row classes in ONE launch.  A block is a "unit" of one shared-memory size (CL_SMEM_QUANTUM): one 512-row edge
// (all 8 warps), two 256-row edges (4 warps each) or eight <= 128-row edges (a warp each); units are numbered longest class first, so the hardware
// hands out the long sequential chains before the many short ones.  (As separate launches on side streams the long classes
// starved: the block scheduler finishes handing out the blocks of the grid that arrived first -- the small class, a few
// microseconds ahead -- before it starts the next grid, so the longest chains began 0.3 - 0.9 ms late on cfg2.)
template <int N>
__device__ __forceinline__ void cluster_edge(int h1, const PassConst& pc, ClusterSm<N>& s, const int32_t* __restrict__ n_of,
                                             const int64_t* __restrict__ row_off, const RowTable& rt, int32_t* n_clusters, int lane) {
    const int64_t row0 = row_off[h1];
    cluster_warp_sm<N>(n_of[h1], pc, s, rt.xc + row0, rt.yc + row0, rt.th + row0, rt.h2orig + row0, rt.xbar + row0, rt.ybar + row0,
                       rt.thbar + row0, rt.label + row0, rt.h2rep + row0, rt.urank + row0, n_clusters + h1, lane);
}
template <int N, int W>
__device__ __forceinline__ void cluster_edge_block(int h1, const PassConst& pc, ClusterSmB<N>& sb, const int32_t* __restrict__ n_of,
                                                   const int64_t* __restrict__ row_off, const RowTable& rt, int32_t* n_clusters, int gw, int lane,
                                                   int bar_id) {
    const int64_t row0 = row_off[h1];
    cluster_block_sm<N, W>(n_of[h1], pc, sb, rt.xc + row0, rt.yc + row0, rt.th + row0, rt.h2orig + row0, rt.xbar + row0, rt.ybar + row0,
                           rt.thbar + row0, rt.label + row0, rt.h2rep + row0, rt.urank + row0, n_clusters + h1, gw, lane, bar_id);
}
__global__ void __launch_bounds__(CL_SMALL_WARPS * 32, 4) k_cluster_classes(const __grid_constant__ PassConst pc, const int32_t* __restrict__ order,
                                                                         const int32_t* __restrict__ cnts, const int32_t* __restrict__ n_of,
                                                                         const int64_t* __restrict__ row_off, RowTable rt, int32_t* n_clusters, int mode) {
    // mode 3: every class; 1: the <= CL_N-row class only; 2: only the > CL_N-row edges that k_cluster_runs left behind (n_clusters == -1)
    extern __shared__ __align__(16) unsigned char cl_smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_small = cnts[0], n_mid = cnts[1], n_big = cnts[2];
    const int u_big = (mode & 2) ? n_big : 0, u_mid = (mode & 2) ? (n_mid + CL_MID_EDGES - 1) / CL_MID_EDGES : 0;
    const int u_small = (mode & 1) ? (n_small + CL_SMALL_WARPS - 1) / CL_SMALL_WARPS : 0;
    const int units = u_big + u_mid + u_small;
    const bool left_only = mode == 2;
    for (int unit = blockIdx.x; unit < units; unit += gridDim.x) {
        if (unit < u_big) {                                               // one edge, the whole block
            const int h1 = order[n_small + n_mid + unit];
            if (!left_only || n_clusters[h1] == -1)
                cluster_edge_block<CL_BIG_N, CL_SMALL_WARPS>(h1, pc, *reinterpret_cast<ClusterSmB<CL_BIG_N>*>(cl_smem), n_of, row_off, rt, n_clusters, warp, lane, 1);
        } else if (unit < u_big + u_mid) {                                // CL_MID_EDGES edges, CL_MID_WARPS warps each
            const int g = warp / CL_MID_WARPS;
            const int e = (unit - u_big) * CL_MID_EDGES + g;
            if (e < n_mid && (!left_only || n_clusters[order[n_small + e]] == -1))
                cluster_edge_block<CL_MID_N, CL_MID_WARPS>(order[n_small + e], pc, reinterpret_cast<ClusterSmB<CL_MID_N>*>(cl_smem)[g], n_of, row_off, rt,
                                                           n_clusters, warp - g * CL_MID_WARPS, lane, 1 + g);
        } else {                                                          // one <= 128-row edge per warp
            const int e = (unit - u_big - u_mid) * CL_SMALL_WARPS + warp;
            if (e < n_small) cluster_edge<CL_N>(order[e], pc, reinterpret_cast<ClusterSm<CL_N>*>(cl_smem)[warp], n_of, row_off, rt, n_clusters, lane);
        }
        if (unit + int(gridDim.x) < units) __syncthreads();   // the next unit may lay the shared memory out differently
    }
}
// the 513..1024-row class: one edge per block of W warps, one block per SM (171 KB of shared memory)
template <int N, int W>
__global__ void __launch_bounds__(W * 32) k_cluster_coop(const __grid_constant__ PassConst pc, const int32_t* __restrict__ order,
                                                         const int32_t* __restrict__ cnts, int which, const int32_t* __restrict__ n_of,
                                                         const int64_t* __restrict__ row_off, RowTable rt, int32_t* n_clusters, int left_only) {
    extern __shared__ __align__(16) unsigned char cl_smem[];
    int base = 0;
    for (int k = 0; k < which; k++) base += cnts[k];
    const int cnt = cnts[which];
    for (int b = blockIdx.x; b < cnt; b += gridDim.x) {
        const int h1 = order[base + b];
        if (left_only && n_clusters[h1] != -1) continue;                  // already clustered by k_cluster_runs
        cluster_edge_block<N, W>(h1, pc, *reinterpret_cast<ClusterSmB<N>*>(cl_smem), n_of, row_off, rt, n_clusters, threadIdx.x >> 5,
                                 threadIdx.x & 31, 1);
        __syncthreads();
    }
}
__global__ void __launch_bounds__(32) k_cluster_huge(const __grid_constant__ PassConst pc, const int32_t* __restrict__ order,
                                                     const int32_t* __restrict__ cnts, const int32_t* __restrict__ n_of,
                                                     const int64_t* __restrict__ row_off, RowTable rt, int32_t* n_clusters) {
    const int base = cnts[0] + cnts[1] + cnts[2] + cnts[3], cnt = cnts[4];
    for (int b = blockIdx.x; b < cnt; b += gridDim.x) {
        const int h1 = order[base + b];
        const int64_t row0 = row_off[h1];
        cluster_warp_global(n_of[h1], pc, rt.xc + row0, rt.yc + row0, rt.th + row0, rt.thn + row0, rt.orient + row0, rt.label + row0,
                            rt.h2orig + row0, rt.xbar + row0, rt.ybar + row0, rt.thbar + row0, rt.h2rep + row0, rt.urank + row0,
                            n_clusters + h1, threadIdx.x);
        __syncwarp();
    }
}
void launch_form_count(const PassConst& pc, const double* x1, const double* y1, int h1_begin, int h1_end, const double2* wedge_h1,
                       const double* x2, const double* y2, const double* th2, const double* ang2, int E2, AngleIndex ix, int32_t* m_out,
                       int32_t* n_out, cudaStream_t st) {
    const int n = h1_end - h1_begin;
    if (n <= 0) return;
    RowTable rt = {};
    k_form<false><<<(n + FORM_WARPS - 1) / FORM_WARPS, FORM_WARPS * 32, 0, st>>>(pc, x1, y1, h1_begin, h1_end, wedge_h1, x2, y2, th2, ang2, E2, ix,
                                                                                m_out, n_out, nullptr, rt, nullptr);
}
void launch_form_fill(const PassConst& pc, const double* x1, const double* y1, int h1_begin, int h1_end, const double2* wedge_h1,
                      const double* x2, const double* y2, const double* th2, const double* ang2, int E2, AngleIndex ix, const int32_t* m_of,
                      const int32_t* n_of, const int64_t* row_off, RowTable rt, int32_t* n_clusters, int32_t* big_list, int32_t* big_count,
                      const RunScratch* rs, cudaStream_t st, cudaStream_t side, cudaEvent_t ev_fork, cudaEvent_t ev_join) {
    const int n = h1_end - h1_begin;
    if (n <= 0) return;
    const size_t smem = FORM_WARPS * 2 * FORM_CAND_CAP * sizeof(int32_t);
    k_form<true><<<(n + FORM_WARPS - 1) / FORM_WARPS, FORM_WARPS * 32, smem, st>>>(pc, x1, y1, h1_begin, h1_end, wedge_h1, x2, y2, th2, ang2, E2, ix,
                                                                                  const_cast<int32_t*>(m_of), const_cast<int32_t*>(n_of), row_off, rt,
                                                                                  n_clusters);
    // clustering: order by hypothesis count, then the shared-memory classes (one launch) beside the rare larger groups
    int32_t* order = big_list;
    int32_t* cnts = big_count;
    static const int small_max = getenv("CVX_CL_SMALL_MAX") ? std::min(CL_N, std::max(1, atoi(getenv("CVX_CL_SMALL_MAX")))) : CL_N;
    k_order_by_n<<<1, 1024, 0, st>>>(n_of, h1_begin, h1_end, nullptr, pc.max_cluster, small_max, order, cnts);
    const size_t sm_q = CL_SMEM_QUANTUM, sm_huge = sizeof(ClusterSmB<CL_HUGE_N>);
    cudaFuncSetAttribute(k_cluster_classes, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sm_q));   // per device, cheap
    cudaFuncSetAttribute(k_cluster_coop<CL_HUGE_N, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sm_huge));
    // > CL_N-row edges: clustering by runs along the epipolar line (k_cluster_runs), then the per-edge kernels for whatever it left
    // behind; all of it on a side stream beside the <= CL_N-row class.  CVX_CL_RUNS=0 restores the per-edge kernels for every class.
    static const bool runs_on = !(getenv("CVX_CL_RUNS") && atoi(getenv("CVX_CL_RUNS")) == 0);
    const bool use_runs = runs_on && rs != nullptr;
    cudaStream_t s_huge = side ? side : st;
    static const bool dbg = getenv("CVX_FORM_DEBUG") != nullptr;
    cudaEvent_t de[4];
    if (dbg) for (auto& e : de) cudaEventCreate(&e);
    if (side) { cudaEventRecord(ev_fork, st); cudaStreamWaitEvent(side, ev_fork, 0); }
    if (dbg) cudaEventRecord(de[0], s_huge);
    const int grid = std::min(n, (n + CL_SMALL_WARPS - 1) / CL_SMALL_WARPS + 148 * 4);
    if (use_runs) {
        cudaMemsetAsync(rs->ctr, 0, 4 * sizeof(int32_t), s_huge);
        k_runs_prepare<<<std::min(n, 148 * 8), RUN_WARPS * 32, 0, s_huge>>>(pc, order, cnts, x1, y1, n_of, row_off, rt, n_clusters, *rs);
        k_runs_cluster<<<148 * 8, RUN_WARPS * 32, 0, s_huge>>>(pc, row_off, rt, *rs);
        // the long runs as virtual edges of the scratch table, through the per-edge kernels (count known on the device only)
        cudaMemsetAsync(rs->v_order + rs->cap_long, 0, 8 * sizeof(int32_t), s_huge);
        k_order_by_n<<<1, 1024, 0, s_huge>>>(rs->v_n, 0, rs->cap_long, rs->ctr, pc.max_cluster, small_max, rs->v_order, rs->v_order + rs->cap_long);
        k_cluster_classes<<<148 * 4, CL_SMALL_WARPS * 32, sm_q, s_huge>>>(pc, rs->v_order, rs->v_order + rs->cap_long, rs->v_n, rs->v_off, rs->vrt, rs->v_ncl, 3);
        k_cluster_coop<CL_HUGE_N, 8><<<148, 256, sm_huge, s_huge>>>(pc, rs->v_order, rs->v_order + rs->cap_long, 3, rs->v_n, rs->v_off, rs->vrt, rs->v_ncl, 0);
        k_runs_scatter<<<148 * 4, RUN_WARPS * 32, 0, s_huge>>>(row_off, rt, *rs);
        k_runs_finish<<<(n + RUN_WARPS - 1) / RUN_WARPS, RUN_WARPS * 32, 0, s_huge>>>(order, cnts, n_of, row_off, rt, n_clusters);
        k_cluster_classes<<<std::min(n, 148 * 4), CL_SMALL_WARPS * 32, sm_q, s_huge>>>(pc, order, cnts, n_of, row_off, rt, n_clusters, 2);
    }
    k_cluster_coop<CL_HUGE_N, 8><<<std::min(n, 148), 256, sm_huge, s_huge>>>(pc, order, cnts, 3, n_of, row_off, rt, n_clusters, use_runs ? 1 : 0);
    k_cluster_huge<<<std::min(n, 148 * 4), 32, 0, s_huge>>>(pc, order, cnts, n_of, row_off, rt, n_clusters);
    if (dbg) { cudaEventRecord(de[1], s_huge); cudaEventRecord(de[2], st); }
    if (side) cudaEventRecord(ev_join, side);
    // one block per unit in the common case (units ~ n / 8 plus the few long edges); beyond that the blocks stride
    k_cluster_classes<<<grid, CL_SMALL_WARPS * 32, sm_q, st>>>(pc, order, cnts, n_of, row_off, rt, n_clusters, use_runs ? 1 : 3);
    if (dbg) {
        cudaEventRecord(de[3], st);
        cudaDeviceSynchronize();
        float t[4];
        for (int i = 0; i < 4; i++) cudaEventElapsedTime(&t[i], de[2], de[i]);
        fprintf(stderr, "[form] relative to the k_cluster_classes launch: > %d-row classes %.3f..%.3f, <= %d-row class 0..%.3f ms\n", CL_N, t[0], t[1], CL_N, t[3]);
        for (auto& e : de) cudaEventDestroy(e);
    }
    if (side) cudaStreamWaitEvent(st, ev_join, 0);
}

// ----------------------------------------------------------------------------- unique-hypothesis table
// Every member row of a cluster was overwritten with the same (xbar, ybar, thbar) (EdgeClusterer.cpp:273-276), so
// getGamma3LocationAndTangent / getOreListBar / getSupportIdx give identical results for all of them: they are evaluated
// once per cluster ("unique hypothesis") and the member rows read the result through row_u.
__global__ void __launch_bounds__(256) k_uniq_fill(const __grid_constant__ PassConst pc, const double* __restrict__ x1, const double* __restrict__ y1,
                                                   const double* __restrict__ th1, RowTable rt, int64_t H, const int64_t* __restrict__ row_off,
                                                   const int64_t* __restrict__ u_off, UniqTable ut) {
    const int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (r >= H) return;
    const int h1 = rt.h1[r];
    const int k = int(r - row_off[h1]);
    const int u = int(u_off[h1]) + rt.urank[r];
    rt.row_u[r] = u;
    atomicAdd(&ut.mult[u], 1);
    if (rt.label[r] != k) return;
    const double xb = rt.xbar[r], yb = rt.ybar[r];
    RowGeom g;
    row_geometry(x1[h1], y1[h1], th1[h1], xb, yb, rt.thbar[r], pc, &g);   // view-independent part of getGamma3LocationAndTangent
    ut.h1[u] = h1; ut.row[u] = int32_t(r);
    ut.xy[u] = make_double2(xb, yb);
    ut.geom[u] = make_double4(g.rho1, g.Tv1[0], g.Tv1[1], g.Tv1[2]);
}
void launch_uniq_fill(const PassConst& pc, const double* x1, const double* y1, const double* th1, RowTable rt, int64_t H,
                      const int64_t* row_off, const int64_t* u_off, UniqTable ut, cudaStream_t st) {
    if (H <= 0) return;
    k_uniq_fill<<<unsigned((H + 255) / 256), 256, 0, st>>>(pc, x1, y1, th1, rt, H, row_off, u_off, ut);
}

// ----------------------------------------------------------------------------- validation
// wedge of the view-1 edge in every validation view (getOreListBar row 0, EdgeSketch_Core.cpp:334,345)
__global__ void __launch_bounds__(128) k_wedge31(const __grid_constant__ PassConst pc, const ViewConst* __restrict__ vc, const double* __restrict__ x1,
                                                 const double* __restrict__ y1, const int32_t* __restrict__ n_of, int h1_begin, int h1_end, int E1,
                                                 double2* __restrict__ w31, DevCounters* ctr) {
    const int h1 = h1_begin + blockIdx.x * blockDim.x + threadIdx.x;
    const int v = blockIdx.y;
    if (h1 >= h1_end || n_of[h1] == 0) return;
    const ViewConst& c = vc[v];
    double lo, hi;
    const int f = wedge_row(x1[h1], y1[h1], c.F31, c.e1_31[0], c.e1_31[1], c.e2_31[0], c.e2_31[1], pc.delta, pc.rows, pc.cols, &lo, &hi);
    w31[int64_t(v) * E1 + h1] = make_double2(lo, hi);
    if (f) atomicAdd(&ctr->n_degenerate[(blockIdx.x + blockIdx.y) & (CVX_CTR_SLOTS - 1)], 1ULL);
}
void launch_wedge31(const PassConst& pc, const ViewConst* vc, const double* x1, const double* y1, const int32_t* n_of, int h1_begin, int h1_end,
                    int E1, double2* w31, DevCounters* ctr, cudaStream_t st) {
    const int n = h1_end - h1_begin;
    if (n <= 0 || pc.n_vali <= 0) return;
    dim3 grid((n + 127) / 128, pc.n_vali);
    k_wedge31<<<grid, 128, 0, st>>>(pc, vc, x1, y1, n_of, h1_begin, h1_end, E1, w31, ctr);
}

// One thread per (unique hypothesis, validation view):
//   getGamma3LocationAndTangent (EdgeSketch_Core.cpp:330, view-dependent part), getOreListBar of the clustered view-2 edge
//   (:340), is_Epipolar_Wedges_in_Parallel (hpp:112-128), getHYPO2_idx_Ore x2 + set_intersection (:348-374) as a query of the
//   view's (a31, a32) bin grid -- the cells overlapped by wedge31 x wedge32 -- with the exact strict tests of
//   PairEdgeHypo.cpp:94 on both angles, and getSupportIdx (getSupportedEdgels.cpp:32-47: d >= thr && d > prev in ascending
//   index order == maximum, ties to the lowest index).  Threads of a warp are consecutive hypotheses of (mostly) one view-1
//   edge and one view: they share wedge31 and hit neighbouring cells.  Queries that overlap many cells or long cells (wedges
//   that wrap around 0/180 degrees, epipole inside the image: SURVEY.md a3) are walked by the whole warp instead of one lane.
// Launch shape per variant (profiles/r01l_validate_variants.txt): the fused kernel (FP64 wedge + walk, issue bound) likes 256-thread blocks
// and three independent record loads per walk step; the stand-alone support test (walk only, L1-lookup bound) 128-thread blocks and two.
// 64 registers either way: occupancy beats a deeper batch (profiles/r01f_variants.txt).
#ifndef VAL_THREADS
#define VAL_THREADS 256
#endif
#ifndef VAL_THREADS_GIVEN
#define VAL_THREADS_GIVEN 128
#endif
#ifndef VAL_BATCH
#define VAL_BATCH 3           // entries a lane loads per step of its walk (independent loads in flight)
#endif
#ifndef VAL_BATCH_GIVEN
#define VAL_BATCH_GIVEN 2
#endif
#ifndef VAL_REGS_THREADS
#define VAL_REGS_THREADS 1024  // resident threads per SM at 64 registers
#endif
#define VAL_LIGHT_ROWS 4       // row ranges a lane keeps in registers
#ifndef VAL_COOP_MIN
#define VAL_COOP_MIN 32        // a query becomes a whole-warp job when its work exceeds this and VAL_COOP_RATIO x the warp's mean
#endif
#ifndef VAL_COOP_RATIO
#define VAL_COOP_RATIO 2
#endif
struct SupportAcc {
    double best;   // prev_prod, getSupportedEdgels.cpp:33
    int bj;        // position (in the walked layout) of the best entry so far, -1 = none
    unsigned n_in, n_vis;
};
__device__ __forceinline__ double4 ld_rec(const double4* p) {   // one 256-bit load (LDG.E.ENL2.256) instead of two 128-bit ones:
    double4 r;                                                  // half the L1 tag lookups of the walk, which is L1-lookup bound
    asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
    return r;
}
// One candidate entry.  The edge index is only needed to break exact ties (equal |t . t^|: the reference keeps the first in
// ascending index order), so it is not loaded on the common path: the walk carries the POSITION of the best entry.
__device__ __forceinline__ void support_visit(SupportAcc& s, const double4 e, int j, const double2 w1, const double2 w2, const double2 tg,
                                              double thr, const int32_t* __restrict__ idx_) {
    if (e.x > w1.x && e.x < w1.y && e.y > w2.x && e.y < w2.y) {          // PairEdgeHypo.cpp:94, both wedges
        s.n_in++;
        const double d = fabs(tg.x * e.z + tg.y * e.w);                   // getSupportedEdgels.cpp:38
        if (d >= thr) {
            if (d > s.best) { s.best = d; s.bj = j; }
            else if (d == s.best && s.bj >= 0 && idx_[j] < idx_[s.bj]) s.bj = j;
        }
    }
}
// GIVEN: wedge32 / tangent are inputs (stand-alone support test on pre-formed hypotheses).
template <bool GIVEN, bool STRICT>
__global__ void __launch_bounds__(GIVEN ? VAL_THREADS_GIVEN : VAL_THREADS, VAL_REGS_THREADS / (GIVEN ? VAL_THREADS_GIVEN : VAL_THREADS))
    k_validate(const __grid_constant__ PassConst pc, const __grid_constant__ ValidateArgs a) {
    constexpr int THREADS = GIVEN ? VAL_THREADS_GIVEN : VAL_THREADS, BATCH = GIVEN ? VAL_BATCH_GIVEN : VAL_BATCH;
    const int64_t u = blockIdx.x * int64_t(THREADS) + threadIdx.x;
    const int v = blockIdx.y;
    const int lane = threadIdx.x & 31;
    unsigned long long c_sup = 0, c_par = 0, c_inl = 0, c_deg = 0, c_vis = 0, c_hvy = 0, mult = 0;
    double2 w1 = make_double2(0, 0), w2 = make_double2(0, 0), tg = make_double2(0, 0);
    bool active = false;
    if (u < a.U) {
        const int h1 = a.ut.h1[u];
        mult = a.ut.mult ? (unsigned long long)a.ut.mult[u] : 1ULL;
        w1 = a.w31[int64_t(v) * a.w31_stride + h1];
        if (GIVEN) {
            w2 = a.w32_in[int64_t(v) * a.U + u];
            tg = a.tgt_in[int64_t(v) * a.U + u];
        } else {
            const ViewConst& c = a.vc[v];
            const double2 xy = a.ut.xy[u];
            const double4 ge = a.ut.geom[u];
            RowGeom g;
            mat_vec(pc.K1inv, a.x1[h1], a.y1[h1], 1.0, g.G1);
            g.rho1 = ge.x; g.Tv1[0] = ge.y; g.Tv1[1] = ge.z; g.Tv1[2] = ge.w;
            reproject_tangent(g, c.R31, c.T31, &tg.x, &tg.y);
            if (STRICT) wedge_row(xy.x, xy.y, c.F32, c.e1_32[0], c.e1_32[1], c.e2_32[0], c.e2_32[1], a.delta, a.rows, a.cols, &w2.x, &w2.y);
            else wedge_row_fast(xy.x, xy.y, c.F32, c.e1_32[0], c.e1_32[1], c.e2_32[0], c.e2_32[1], a.delta, a.rows, a.cols, &w2.x, &w2.y);
            if (a.w32_out) { a.w32_out[int64_t(v) * a.U + u] = w2; a.tgt_out[int64_t(v) * a.U + u] = tg; }
        }
        const bool w1_bad = (w1.x != w1.x), w2_bad = (w2.x != w2.x);
        if (w2_bad) c_deg = mult;
        if (!w1_bad && !w2_bad) {
            const double d0 = fabs(w1.x - w2.x), d1 = fabs(w1.x - w2.y), d2 = fabs(w1.y - w2.x), d3 = fabs(w1.y - w2.y);
            const double dmax = fmax(fmax(d0, d1), fmax(d2, d3));
            if (dmax <= a.parallel_deg) c_par = mult;                       // hpp:120 (inclusive)
            else active = true;
        }
    }
    // ---- wedge31 n wedge32 through the view's bin grid
    const ViewGrid vg = a.gi.vg[v];
    const int64_t off = a.edge_off[a.vali_view ? a.vali_view[v] : v];
    const int64_t ce_off = int64_t(v) * a.gi.CS;
    SupportAcc acc = {0.0, -1, 0u, 0u};
    int r0 = 0, r1 = -1, c0 = 0, c1 = 0, work = 0, G = 1;
    int beg[VAL_LIGHT_ROWS], end[VAL_LIGHT_ROWS];
    bool useB = false;
    // every edge angle of the view lies in [lo, hi]: a wedge outside that range has no member (strict tests)
    if (active && w1.y > vg.lo1 && w1.x < vg.hi1 && w2.y > vg.lo2 && w2.x < vg.hi2) {
        const int p0 = grid_bin(w1.x, vg.lo1, vg.s1, vg.G1), p1 = grid_bin(w1.y, vg.lo1, vg.s1, vg.G1);
        const int q0 = grid_bin(w2.x, vg.lo2, vg.s2, vg.G2), q1 = grid_bin(w2.y, vg.lo2, vg.s2, vg.G2);
        useB = (p1 - p0) > (q1 - q0);                   // rows along the angle in which the rectangle is narrower
        r0 = useB ? q0 : p0; r1 = useB ? q1 : p1; c0 = useB ? p0 : q0; c1 = useB ? p1 : q1;
        G = useB ? vg.G1 : vg.G2;                       // row stride of the chosen layout
        const int32_t* __restrict__ ce = (useB ? a.gi.B.cell_end : a.gi.A.cell_end) + ce_off;
        if (r1 - r0 < VAL_LIGHT_ROWS) {
#pragma unroll
            for (int r = 0; r < VAL_LIGHT_ROWS; r++) {
                beg[r] = end[r] = 0;
                if (r0 + r <= r1) {
                    const int cc = (r0 + r) * G + c0;
                    beg[r] = cc ? ce[cc - 1] : 0;
                    end[r] = ce[(r0 + r) * G + c1];
                    work += end[r] - beg[r];
                }
            }
        } else {
            work = 8 * (r1 - r0 + 1);                   // not looked up yet: an estimate is enough for the split below
        }
    } else {
        active = false;
    }
    // a query much heavier than its warp's mean is walked by the whole warp; the rest by their own lane
    const int wsum = __reduce_add_sync(FULL, work), wcnt = __popc(__ballot_sync(FULL, active));   // REDUX: one instruction each
    const bool heavy = active && work > VAL_COOP_MIN && work * max(wcnt, 1) > VAL_COOP_RATIO * wsum;
    if (active && !heavy) {
        const GridLayout& L = useB ? a.gi.B : a.gi.A;
        const double4* __restrict__ rec = L.rec + off;
        const int32_t* __restrict__ idx_ = L.idx + off;
        if (r1 - r0 < VAL_LIGHT_ROWS) {
            // the (at most VAL_LIGHT_ROWS) ranges are walked as one flattened list, BATCH entries per step, so that several
            // independent 32-byte loads are in flight per lane instead of one (the walk is latency bound)
            const int l0 = end[0] - beg[0], l1 = l0 + (end[1] - beg[1]), l2 = l1 + (end[2] - beg[2]);
            const int g0 = beg[1] - end[0], g1 = beg[2] - end[1], g2 = beg[3] - end[2];   // jumps between consecutive ranges
            int k = 0;
            for (; k + BATCH <= work; k += BATCH) {     // full steps: no clamping, no tail predicate
                int j[BATCH];
                double4 e[BATCH];
#pragma unroll
                for (int t = 0; t < BATCH; t++) {
                    const int kk = k + t;
                    j[t] = beg[0] + kk + (kk >= l0 ? g0 : 0) + (kk >= l1 ? g1 : 0) + (kk >= l2 ? g2 : 0);
                }
#pragma unroll
                for (int t = 0; t < BATCH; t++) e[t] = ld_rec(rec + j[t]);
#pragma unroll
                for (int t = 0; t < BATCH; t++) support_visit(acc, e[t], j[t], w1, w2, tg, a.cos_thresh, idx_);
            }
            if (k < work) {                             // the last, partial step (indices clamped to the last entry)
                int j[BATCH - 1];
                double4 e[BATCH - 1];
#pragma unroll
                for (int t = 0; t < BATCH - 1; t++) {
                    const int kk = min(k + t, work - 1);
                    j[t] = beg[0] + kk + (kk >= l0 ? g0 : 0) + (kk >= l1 ? g1 : 0) + (kk >= l2 ? g2 : 0);
                }
#pragma unroll
                for (int t = 0; t < BATCH - 1; t++) e[t] = ld_rec(rec + j[t]);
#pragma unroll
                for (int t = 0; t < BATCH - 1; t++)
                    if (t == 0 || k + t < work) support_visit(acc, e[t], j[t], w1, w2, tg, a.cos_thresh, idx_);
            }
            acc.n_vis = unsigned(work);                 // entries visited (statistics)
        } else {
            const int32_t* __restrict__ ce = L.cell_end + ce_off;
            for (int b = r0; b <= r1; b++) {
                const int cc = b * G + c0;
                const int jb = cc ? ce[cc - 1] : 0, je = ce[b * G + c1];
                for (int j = jb; j < je; j++) support_visit(acc, ld_rec(rec + j), j, w1, w2, tg, a.cos_thresh, idx_);
                acc.n_vis += unsigned(je - jb);
            }
        }
    }
    // ---- warp-cooperative walk of the heavy queries, one at a time
    int bidx = -2;
    unsigned hm = __ballot_sync(FULL, heavy);
    while (hm) {
        const int src = __ffs(hm) - 1;
        hm &= hm - 1;
        const double2 sw1 = make_double2(__shfl_sync(FULL, w1.x, src), __shfl_sync(FULL, w1.y, src));
        const double2 sw2 = make_double2(__shfl_sync(FULL, w2.x, src), __shfl_sync(FULL, w2.y, src));
        const double2 stg = make_double2(__shfl_sync(FULL, tg.x, src), __shfl_sync(FULL, tg.y, src));
        const int sr0 = __shfl_sync(FULL, r0, src), sr1 = __shfl_sync(FULL, r1, src);
        const int sc0 = __shfl_sync(FULL, c0, src), sc1 = __shfl_sync(FULL, c1, src);
        const bool sB = __shfl_sync(FULL, int(useB), src) != 0;
        const int sG = __shfl_sync(FULL, G, src);
        const GridLayout& L = sB ? a.gi.B : a.gi.A;
        const double4* __restrict__ rec = L.rec + off;
        const int32_t* __restrict__ idx_ = L.idx + off;
        const int32_t* __restrict__ ce = L.cell_end + ce_off;
        SupportAcc ca = {0.0, -1, 0u, 0u};
        if (sr1 - sr0 + 1 >= 8) {                       // many rows: one lane per row
            for (int b = sr0 + lane; b <= sr1; b += 32) {
                const int cc = b * sG + sc0;
                const int jb = cc ? ce[cc - 1] : 0, je = ce[b * sG + sc1];
                for (int j = jb; j < je; j++) support_visit(ca, ld_rec(rec + j), j, sw1, sw2, stg, a.cos_thresh, idx_);
                ca.n_vis += unsigned(je - jb);
            }
        } else {                                        // few long rows: the lanes stride over the entries
            for (int b = sr0; b <= sr1; b++) {
                const int cc = b * sG + sc0;
                const int jb = cc ? ce[cc - 1] : 0, je = ce[b * sG + sc1];
                for (int j = jb + lane; j < je; j += 32) support_visit(ca, ld_rec(rec + j), j, sw1, sw2, stg, a.cos_thresh, idx_);
                if (je - jb > lane) ca.n_vis += unsigned((je - jb - lane + 31) >> 5);
            }
        }
        int ci = ca.bj >= 0 ? idx_[ca.bj] : -2;         // edge index of the lane's best
        {                                               // maximum, ties to the lowest edge index: |t . t^| > 0 orders as its bit
            const unsigned long long kb = ci != -2 ? (unsigned long long)__double_as_longlong(ca.best) : 0ULL;   // pattern (REDUX x 3)
            const unsigned hi = unsigned(kb >> 32), lo = unsigned(kb);
            const unsigned mh = __reduce_max_sync(FULL, hi);
            const unsigned ml = __reduce_max_sync(FULL, hi == mh ? lo : 0u);
            const unsigned mi_ = __reduce_min_sync(FULL, (ci != -2 && hi == mh && lo == ml) ? unsigned(ci) : 0xffffffffu);
            ci = mi_ == 0xffffffffu ? -2 : int(mi_);
        }
        ca.n_in = __reduce_add_sync(FULL, ca.n_in); ca.n_vis = __reduce_add_sync(FULL, ca.n_vis);
        if (lane == src) { acc = ca; bidx = ci; c_hvy = 1; }
    }
    if (active && !heavy && acc.bj >= 0) bidx = ((useB ? a.gi.B.idx : a.gi.A.idx) + off)[acc.bj];
    if (u < a.U) {
        c_inl = mult * acc.n_in; c_vis = acc.n_vis;
        if (bidx != -2) c_sup = mult;
        a.support[int64_t(v) * a.sup_stride + u] = bidx;
    }
    if (a.ctr == nullptr) return;
    // warp-aggregated statistics -> one counter slot per warp (no block barrier: warps retire independently).  Per-warp sums of
    // everything but the weighted inlier count fit 32 bits (at most 32 lanes x edges of a view / rows of a group): one REDUX each.
    c_sup = __reduce_add_sync(FULL, unsigned(c_sup)); c_par = __reduce_add_sync(FULL, unsigned(c_par));
    c_deg = __reduce_add_sync(FULL, unsigned(c_deg)); c_vis = __reduce_add_sync(FULL, unsigned(c_vis));
    c_hvy = __reduce_add_sync(FULL, unsigned(c_hvy));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c_inl += __shfl_xor_sync(FULL, c_inl, o);
    if (lane == 0) {
        const int slot = int((blockIdx.x * (THREADS / 32) + (threadIdx.x >> 5) + blockIdx.y) & (CVX_CTR_SLOTS - 1));
        if (c_sup) atomicAdd(&a.ctr->n_supported[slot], c_sup);
        if (c_par) atomicAdd(&a.ctr->n_parallel[slot], c_par);
        if (c_inl) atomicAdd(&a.ctr->n_inliers[slot], c_inl);
        if (c_deg) atomicAdd(&a.ctr->n_degenerate[slot], c_deg);
        if (c_vis) atomicAdd(&a.ctr->n_visited[slot], c_vis);
        if (c_hvy) atomicAdd(&a.ctr->n_heavy[slot], c_hvy);
    }
}
void launch_validate(const PassConst& pc, const ValidateArgs& a, bool given, cudaStream_t st) {
    if (a.U <= 0 || a.n_vali <= 0) return;
    const int threads = given ? VAL_THREADS_GIVEN : VAL_THREADS;
    dim3 grid(unsigned((a.U + threads - 1) / threads), a.n_vali);
    if (given) k_validate<true, false><<<grid, threads, 0, st>>>(pc, a);
    else if (a.strict) k_validate<false, true><<<grid, threads, 0, st>>>(pc, a);
    else k_validate<false, false><<<grid, threads, 0, st>>>(pc, a);
}

// ----------------------------------------------------------------------------- vote + ratio test
// EdgeSketch_Core.cpp:414-427: number of validation views that support a hypothesis (per unique hypothesis).
__global__ void __launch_bounds__(256) k_count_u(const int32_t* __restrict__ support, int64_t sup_stride, int64_t U, int n_vali,
                                                 int32_t* __restrict__ count_u) {
    const int64_t u = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (u >= U) return;
    int c = 0;
    for (int v = 0; v < n_vali; v++) c += (support[int64_t(v) * sup_stride + u] != -2);
    count_u[u] = c;
}
void launch_count_u(const int32_t* support, int64_t sup_stride, int64_t U, int n_vali, int32_t* count_u, cudaStream_t st) {
    if (U > 0) k_count_u<<<unsigned((U + 255) / 256), 256, 0, st>>>(support, sup_stride, U, n_vali, count_u);
}
// EdgeSketch_Core.cpp:430-508.  One warp per view-1 edge.  flags bit0 = valid (count >= N), bit1 = passes count > ratio * max.
__global__ void __launch_bounds__(128) k_select(const __grid_constant__ PassConst pc, const int64_t* __restrict__ row_off, int h1_begin, int h1_end,
                                                const int32_t* __restrict__ count_u, RowTable rt, int32_t* n_valid, int32_t* n_sel,
                                                int32_t* max_sel) {
    const int lane = threadIdx.x & 31;
    const int h1 = h1_begin + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (h1 >= h1_end) return;
    const int64_t r0 = row_off[h1], r1 = row_off[h1 + 1];
    int maxc = 0, nv = 0;
    for (int64_t r = r0 + lane; r < r1; r += 32) {
        const int c = count_u[rt.row_u[r]];
        rt.count[r] = c;
        if (c > 0 && c >= pc.max_support) { nv++; maxc = max(maxc, c); }   // :432
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { maxc = max(maxc, __shfl_xor_sync(FULL, maxc, o)); nv += __shfl_xor_sync(FULL, nv, o); }
    const double thr = maxc * pc.lowes_ratio;   // :498
    int ns = 0;
    for (int64_t r = r0 + lane; r < r1; r += 32) {
        const int c = rt.count[r];
        int f = 0;
        if (c > 0 && c >= pc.max_support) { f = 1; if (double(c) > thr) { f |= 2; ns++; } }   // :501
        rt.flags[r] = f;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ns += __shfl_xor_sync(FULL, ns, o);
    if (lane == 0) {
        n_valid[h1] = nv; n_sel[h1] = ns;
        // the reference keeps at most n_views rows per view-1 edge (paired_edge capacity, EdgeSketch_Core.cpp:513-526): the host
        // applies that cut only when some edge exceeds it
        if (ns > *max_sel) atomicMax(max_sel, ns);
    }
}
void launch_select(const PassConst& pc, const int64_t* row_off, int h1_begin, int h1_end, const int32_t* count_u, RowTable rt,
                   int32_t* n_valid, int32_t* n_sel, int32_t* max_sel, cudaStream_t st) {
    const int n = h1_end - h1_begin;
    if (n <= 0) return;
    k_select<<<(n + 3) / 4, 128, 0, st>>>(pc, row_off, h1_begin, h1_end, count_u, rt, n_valid, n_sel, max_sel);
}
__global__ void __launch_bounds__(128) k_select_fill(const int64_t* __restrict__ row_off, int h1_begin, int h1_end, RowTable rt,
                                                     const int64_t* __restrict__ valid_off, const int64_t* __restrict__ pair_off, int32_t* v_h1,
                                                     int32_t* v_row, int32_t* v_count, int32_t* v_sel, int32_t* p_h1, int32_t* p_h2rep,
                                                     int32_t* p_row, int32_t* p_count) {
    const int lane = threadIdx.x & 31;
    const int h1 = h1_begin + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (h1 >= h1_end) return;
    const int64_t r0 = row_off[h1], r1 = row_off[h1 + 1];
    int64_t vo = valid_off[h1], po = pair_off[h1];
    if (valid_off[h1 + 1] == vo) return;
    const unsigned lt = lanemask_lt();
    for (int64_t rc = r0; rc < r1; rc += 32) {
        const int64_t r = rc + lane;
        const int f = (r < r1) ? rt.flags[r] : 0;
        const unsigned mv = __ballot_sync(FULL, f & 1), ms = __ballot_sync(FULL, f & 2);
        if (f & 1) {
            const int64_t q = vo + __popc(mv & lt);
            v_h1[q] = h1; v_row[q] = int32_t(r); v_count[q] = rt.count[r]; v_sel[q] = (f >> 1) & 1;
        }
        if (f & 2) {
            const int64_t q = po + __popc(ms & lt);
            p_h1[q] = h1; p_h2rep[q] = rt.h2rep[r]; p_row[q] = int32_t(r); p_count[q] = rt.count[r];
        }
        vo += __popc(mv); po += __popc(ms);
    }
}
void launch_select_fill(const int64_t* row_off, int h1_begin, int h1_end, RowTable rt, const int64_t* valid_off, const int64_t* pair_off,
                        int32_t* v_h1, int32_t* v_row, int32_t* v_count, int32_t* v_sel, int32_t* p_h1, int32_t* p_h2rep, int32_t* p_row,
                        int32_t* p_count, cudaStream_t st) {
    const int n = h1_end - h1_begin;
    if (n <= 0) return;
    k_select_fill<<<(n + 3) / 4, 128, 0, st>>>(row_off, h1_begin, h1_end, rt, valid_off, pair_off, v_h1, v_row, v_count, v_sel, p_h1, p_h2rep, p_row,
                                              p_count);
}
__global__ void k_gather_supports(const int32_t* __restrict__ p_row, const int32_t* __restrict__ row_u, int64_t P, int n_vali,
                                  const int32_t* __restrict__ support, int64_t sup_stride, int32_t* __restrict__ out) {
    const int64_t t = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (t >= P * n_vali) return;
    const int64_t p = t / n_vali;
    const int v = int(t % n_vali);
    out[t] = support[int64_t(v) * sup_stride + row_u[p_row[p]]];
}
void launch_gather_supports(const int32_t* p_row, const int32_t* row_u, int64_t P, int n_vali, const int32_t* support, int64_t sup_stride,
                            int32_t* out, cudaStream_t st) {
    const int64_t n = P * n_vali;
    if (n > 0) k_gather_supports<<<unsigned((n + 255) / 256), 256, 0, st>>>(p_row, row_u, P, n_vali, support, sup_stride, out);
}
// out[r][v] = support[v][row_u[r]] through a 32x32 shared-memory tile (reads coalesced over rows, writes over views)
__global__ void k_expand_support(const int32_t* __restrict__ support, int64_t sup_stride, const int32_t* __restrict__ row_u, int64_t H, int n_vali,
                                 int32_t* __restrict__ out) {
    __shared__ int32_t tile[32][33];
    const int64_t r0 = blockIdx.x * 32LL;
    const int v0 = blockIdx.y * 32;
    const int64_t rr = r0 + threadIdx.x;
    const int64_t u = rr < H ? row_u[rr] : 0;
    for (int k = threadIdx.y; k < 32; k += blockDim.y) {
        const int v = v0 + k;
        tile[k][threadIdx.x] = (v < n_vali && rr < H) ? support[int64_t(v) * sup_stride + u] : 0;
    }
    __syncthreads();
    for (int k = threadIdx.y; k < 32; k += blockDim.y) {
        const int64_t r = r0 + k;
        const int v = v0 + threadIdx.x;
        if (r < H && v < n_vali) out[r * n_vali + v] = tile[threadIdx.x][k];
    }
}
void launch_expand_support(const int32_t* support, int64_t sup_stride, const int32_t* row_u, int64_t H, int n_vali, int32_t* out,
                           cudaStream_t st) {
    if (H <= 0 || n_vali <= 0) return;
    dim3 grid(unsigned((H + 31) / 32), (n_vali + 31) / 32), block(32, 8);
    k_expand_support<<<grid, block, 0, st>>>(support, sup_stride, row_u, H, n_vali, out);
}
// ----------------------------------------------------------------------------- finalize
// Finalize_Edge_Pairs_and_Reconstruct_3D_Edges geometry (EdgeSketch_Core.cpp:625-676) + Stack_3D_Edges transform (:742-748)
__global__ void __launch_bounds__(128) k_finalize(const __grid_constant__ PassConst pc, const double* __restrict__ x1, const double* __restrict__ y1,
                                                  const double* __restrict__ th1, const double* __restrict__ x2, const double* __restrict__ y2,
                                                  const double* __restrict__ th2, const int32_t* __restrict__ p_h1, const int32_t* __restrict__ p_h2rep,
                                                  int64_t P, double* gamma_cam, double* gamma_world, double* tangent_world, double* h2c,
                                                  int32_t* nan_flag) {
    const int64_t p = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (p >= P) return;
    const int a = p_h1[p], b = p_h2rep[p];
    const double xa = x1[a], ya = y1[a], ta = th1[a], xb = x2[b], yb = y2[b], tb = th2[b];
    double coef[3];
    epip_line_coeffs(pc.F21, xa, ya, coef);
    double cx = nan(""), cy = nan("");
    const bool ok = correct_one(coef, xb, yb, tb, pc.loc_pert, pc.tang_thresh, pc.orient_pert, &cx, &cy);   // :636
    double G[3] = {0, 0, 0};
    int nf = 0;
    if (ok) triangulate2(xa, ya, cx, cy, pc.R21, pc.T21, pc.K1inv, G);   // :656 (K_HYPO1 for both points)
    else nf = 2;
    if (G[0] != G[0] || G[1] != G[1] || G[2] != G[2]) { G[0] = G[1] = G[2] = 0; nf |= 1; }   // :658-663
    double W[3], tw[3];
    to_world(G, pc.R1, pc.T1, W);                                         // :666, :745
    tangent3d_world(xa, ya, ta, xb, yb, tb, pc, tw);                      // :675 (uncorrected original rows)
    for (int k = 0; k < 3; k++) { gamma_cam[p * 3 + k] = G[k]; gamma_world[p * 3 + k] = W[k]; tangent_world[p * 3 + k] = tw[k]; }
    h2c[p * 2] = cx; h2c[p * 2 + 1] = cy;
    nan_flag[p] = nf;
}
void launch_finalize(const PassConst& pc, const double* x1, const double* y1, const double* th1, const double* x2, const double* y2,
                     const double* th2, const int32_t* p_h1, const int32_t* p_h2rep, int64_t P, double* gamma_cam, double* gamma_world,
                     double* tangent_world, double* h2c, int32_t* nan_flag, cudaStream_t st) {
    if (P <= 0) return;
    k_finalize<<<unsigned((P + 127) / 128), 128, 0, st>>>(pc, x1, y1, th1, x2, y2, th2, p_h1, p_h2rep, P, gamma_cam, gamma_world, tangent_world,
                                                         h2c, nan_flag);
}

// ----------------------------------------------------------------------------- claimed edges (next-view selection input)
// The reference walks edge_3D_to_supporting_edges and inserts (image, edge) of every supporting record into one std::set per
// view (EdgeSketch_Core.cpp:785-797).  The records of a pass are, per accepted pair that reached the insertions
// (EdgeSketch_Core.cpp:682-726): the view-1 edge, the view-2 representative, the members of its cluster as looked up through
// hypo2_clusters[(h1, h2rep)] (key = HYPO2_idx_raw(local row), EdgeClusterer.cpp:216-236; default {h2rep}), and the
// supporting edge of every validation view.  One warp per accepted pair sets the corresponding bits.
__device__ __forceinline__ void claim_bit(uint32_t* bits, int64_t off, int idx) {
    const int64_t e = off + idx;
    atomicOr(&bits[e >> 5], 1u << (e & 31));
}
__global__ void __launch_bounds__(128) k_claim_pairs(const __grid_constant__ ClaimArgs a) {
    const int lane = threadIdx.x & 31;
    const int64_t p = blockIdx.x * int64_t(blockDim.x >> 5) + (threadIdx.x >> 5);
    if (p >= a.P) return;
    if (a.keep && !a.keep[p]) return;                     // beyond the n_views rows paired_edge holds per view-1 edge (:513-526)
    if (a.nan_flag[p] != 0) return;                       // `continue` before the insertions (:639-642, :658-663)
    const int h1 = a.p_h1[p], h2 = a.p_h2rep[p];
    const int64_t off1 = a.edge_off[a.view1], off2 = a.edge_off[a.view2];
    if (lane == 0) { claim_bit(a.bits, off1, h1); claim_bit(a.bits, off2, h2); }
    for (int v = lane; v < a.n_vali; v += 32) {           // :694-724
        const int s = a.p_sup[p * a.n_vali + v];
        if (s != -2) claim_bit(a.bits, a.edge_off[a.vali_view[v]], s);
    }
    // cluster members: find the row whose raw_at_local equals h2rep (the map key), then every row with its label
    const int64_t r0 = a.row_off[h1], r1 = a.row_off[h1 + 1];
    int key_label = -1;
    for (int64_t rc = r0; rc < r1; rc += 32) {
        const int64_t r = rc + lane;
        const bool hit = r < r1 && a.rawat[r] == h2;
        const unsigned m = __ballot_sync(FULL, hit);
        if (m) { key_label = __shfl_sync(FULL, hit ? a.label[r] : 0, __ffs(m) - 1); break; }
    }
    if (key_label < 0) return;                            // not a key: get_edges_in_same_cluster returns {h2rep} (hpp:193-202)
    for (int64_t r = r0 + lane; r < r1; r += 32)
        if (a.label[r] == key_label) { const int mi = a.rawat[r]; if (mi >= 0 && mi < a.E2) claim_bit(a.bits, off2, mi); }   // :686
}
void launch_claim_pairs(const ClaimArgs& a, cudaStream_t st) {
    if (a.P > 0) k_claim_pairs<<<unsigned((a.P + 3) / 4), 128, 0, st>>>(a);
}
__global__ void __launch_bounds__(256) k_claim_count(const uint32_t* __restrict__ bits, const int64_t* __restrict__ edge_off,
                                                     const int32_t* __restrict__ edge_n, unsigned long long* out) {
    const int v = blockIdx.y;
    const int64_t w0 = edge_off[v] >> 5;
    const int n = edge_n[v], nw = (n + 31) >> 5;
    int c = 0;
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < nw; w += gridDim.x * blockDim.x) {
        uint32_t b = bits[w0 + w];
        if (w == nw - 1 && (n & 31)) b &= (1u << (n & 31)) - 1u;
        c += __popc(b);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(FULL, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&out[v], (unsigned long long)c);
}
void launch_claim_count(const uint32_t* bits, const int64_t* edge_off, const int32_t* edge_n, int n_views, unsigned long long* out, cudaStream_t st) {
    cudaMemsetAsync(out, 0, size_t(n_views) * 8, st);
    dim3 grid(8, n_views);
    k_claim_count<<<grid, 256, 0, st>>>(bits, edge_off, edge_n, out);
}
__global__ void k_claim_expand(const uint32_t* __restrict__ bits, int64_t off, int n, uint8_t* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const int64_t e = off + i; out[i] = (bits[e >> 5] >> (e & 31)) & 1u; }
}
void launch_claim_expand(const uint32_t* bits, int64_t off, int n, uint8_t* out, cudaStream_t st) {
    if (n > 0) k_claim_expand<<<(n + 255) / 256, 256, 0, st>>>(bits, off, n, out);
}

}  // namespace cvx

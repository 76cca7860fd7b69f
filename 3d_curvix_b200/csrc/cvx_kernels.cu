// =====================================================================================
// cvx_kernels.cu -- hand-written sm_100a kernels of the 3D Curvix edge-sketch hot path.
//
// Compile with: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -lineinfo
// (-fmad=false keeps every a*b+c un-fused so accept/reject decisions match an un-fused CPU build).
//
// Pipeline of one hypothesis view pair (Run_3D_Edge_Sketch, EdgeSketch_Core.cpp:168-573):
//   stage_views   per validation view: two epipolar angles per edge (a2), binned by the first one
//   h2_angles     epipolar angle of every view-2 edge (a2)          h1_wedges: wedge of every view-1 edge (a3)
//   form_count / form_fill   one warp per view-1 edge: wedge scan of view 2 (a5), epipolar correction (a6),
//                 greedy clustering (a7) -> hypothesis table
//   wedge31       wedge of the view-1 edge in each validation view (a3, row 0 only)
//   reproject     per hypothesis x view: reprojected tangent (a8) + wedge of the view-2 edge (a3)
//   support       one warp per (view-1 edge, view): walks the bins covered by wedge31, intersects with
//                 wedge32, orientation arg-max (a10, a11, a12)  <- the HBM-bound support test
//   select        vote + ratio test (a13), finalize: correction + 2-view DLT + tangent (a15)
// No tensor cores: there is no dense contraction anywhere on this path.
// =====================================================================================
#include "cvx_kernels.cuh"

#include <float.h>

namespace cvx {

#define FULL 0xffffffffu

__device__ __forceinline__ unsigned lanemask_lt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}
__device__ __forceinline__ int bin_of(double a, double scale, int B) {
    if (a != a) return B;   // NaN bin, never queried
    int b = int(a * scale);
    return b < 0 ? 0 : (b >= B ? B - 1 : b);
}
__device__ __forceinline__ double2 ldg2(const double2* p) { return __ldg(p); }

// ----------------------------------------------------------------------------- upload-time
__global__ void k_edge_trig(const double* __restrict__ th, double* __restrict__ cs, double* __restrict__ sn, int64_t n) {
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < n) { cs[i] = cos(th[i]); sn[i] = sin(th[i]); }   // EdgeSketch_Core.cpp:321-322, hoisted: depends on the view only
}
void launch_edge_trig(const double* th, double* cs, double* sn, int64_t n, cudaStream_t st) {
    if (n <= 0) return;
    k_edge_trig<<<unsigned((n + 255) / 256), 256, 0, st>>>(th, cs, sn, n);
}

// ----------------------------------------------------------------------------- per-pass view staging
// getOreListVali for both hypothesis views (EdgeSketch_Core.cpp:337,342), hoisted out of the H1 loop.
__global__ void k_stage_angles(EdgeArena ea, const int64_t* __restrict__ edge_off, const int32_t* __restrict__ edge_n,
                               const int32_t* __restrict__ vali_view, const ViewConst* __restrict__ vc, BinGrid bg) {
    const int v = blockIdx.y;
    const int view = vali_view[v];
    const int n = edge_n[view];
    const int64_t off = edge_off[view];
    const double e31x = vc[v].e2_31[0], e31y = vc[v].e2_31[1], e32x = vc[v].e2_32[0], e32y = vc[v].e2_32[1];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double x = ea.x[off + i], y = ea.y[off + i];
        const double a31 = ore_angle(x, y, e31x, e31y);
        const double a32 = ore_angle(x, y, e32x, e32y);
        const int b = bin_of(a31, bg.scale, bg.B);
        ea.a31[off + i] = a31; ea.a32[off + i] = a32; ea.bin[off + i] = b;
        atomicAdd(&bg.bins[int64_t(v) * bg.BS + b], 1);
    }
}
// in-place exclusive scan of one view's bin histogram (one block per view)
__global__ void __launch_bounds__(1024) k_scan_bins(BinGrid bg) {
    __shared__ int part[1024];
    int32_t* b = bg.bins + int64_t(blockIdx.x) * bg.BS;
    const int per = (bg.BS + 1023) / 1024;
    const int lo = threadIdx.x * per, hi = min(lo + per, bg.BS);
    int s = 0;
    for (int i = lo; i < hi; i++) s += b[i];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        int t = (threadIdx.x >= d) ? part[threadIdx.x - d] : 0;
        __syncthreads();
        part[threadIdx.x] += t;
        __syncthreads();
    }
    int run = part[threadIdx.x] - s;
    for (int i = lo; i < hi; i++) { int c = b[i]; b[i] = run; run += c; }
}
// counting-sort scatter; afterwards bins[v][b] holds the END of bin b (start of bin b is bins[v][b-1])
__global__ void k_stage_scatter(EdgeArena ea, const int64_t* __restrict__ edge_off, const int32_t* __restrict__ edge_n,
                                const int32_t* __restrict__ vali_view, BinGrid bg) {
    const int v = blockIdx.y;
    const int view = vali_view[v];
    const int n = edge_n[view];
    const int64_t off = edge_off[view];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int b = ea.bin[off + i];
        const int pos = atomicAdd(&bg.bins[int64_t(v) * bg.BS + b], 1);
        const int64_t d = off + pos;
        ea.s_a31[d] = ea.a31[off + i]; ea.s_a32[d] = ea.a32[off + i];
        ea.s_cs[d] = ea.cs[off + i]; ea.s_sn[d] = ea.sn[off + i]; ea.s_idx[d] = i;
    }
}
void launch_stage_views(const EdgeArena& ea, const int64_t* edge_off, const int32_t* edge_n, const int32_t* vali_view, int n_vali,
                        int max_edges, const ViewConst* vc, const BinGrid& bg, cudaStream_t st, int* launches) {
    if (n_vali <= 0) return;
    cudaMemsetAsync(bg.bins, 0, size_t(n_vali) * bg.BS * sizeof(int32_t), st);
    int bx = (max_edges + 255) / 256;
    if (bx < 1) bx = 1;
    if (bx > 1024) bx = 1024;
    dim3 grid(bx, n_vali);
    k_stage_angles<<<grid, 256, 0, st>>>(ea, edge_off, edge_n, vali_view, vc, bg);
    k_scan_bins<<<n_vali, 1024, 0, st>>>(bg);
    k_stage_scatter<<<grid, 256, 0, st>>>(ea, edge_off, edge_n, vali_view, bg);
    if (launches) *launches += 3;
}

// stand-alone support test: the per-view angle / tangent arrays are given; only the bin grid is built here
__global__ void k_given_hist(const double* __restrict__ a31, const int64_t* __restrict__ view_off, int32_t* __restrict__ bin, BinGrid bg) {
    const int v = blockIdx.y;
    const int64_t b0 = view_off[v], n = view_off[v + 1] - b0;
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        const int b = bin_of(a31[b0 + i], bg.scale, bg.B);
        bin[b0 + i] = b;
        atomicAdd(&bg.bins[int64_t(v) * bg.BS + b], 1);
    }
}
__global__ void k_given_scatter(const double* __restrict__ a31, const double* __restrict__ a32, const double* __restrict__ cs,
                                const double* __restrict__ sn, const int64_t* __restrict__ view_off, const int32_t* __restrict__ bin, BinGrid bg,
                                double* s_a31, double* s_a32, double* s_cs, double* s_sn, int32_t* s_idx) {
    const int v = blockIdx.y;
    const int64_t b0 = view_off[v], n = view_off[v + 1] - b0;
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        const int pos = atomicAdd(&bg.bins[int64_t(v) * bg.BS + bin[b0 + i]], 1);
        const int64_t d = b0 + pos;
        s_a31[d] = a31[b0 + i]; s_a32[d] = a32[b0 + i]; s_cs[d] = cs[b0 + i]; s_sn[d] = sn[b0 + i]; s_idx[d] = int32_t(i);
    }
}
__global__ void k_iota(int32_t* p, int n) { int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) p[i] = i; }
void launch_stage_given(const double* a31, const double* a32, const double* cs, const double* sn, const int64_t* view_off, int n_vali,
                        int64_t max_edges, int32_t* bin_tmp, const BinGrid& bg, double* s_a31, double* s_a32, double* s_cs, double* s_sn,
                        int32_t* s_idx, int32_t* vali_iota, cudaStream_t st) {
    k_iota<<<(n_vali + 255) / 256, 256, 0, st>>>(vali_iota, n_vali);
    int64_t bx = (max_edges + 255) / 256;
    if (bx < 1) bx = 1;
    if (bx > 2048) bx = 2048;
    dim3 grid((unsigned)bx, n_vali);
    k_given_hist<<<grid, 256, 0, st>>>(a31, view_off, bin_tmp, bg);
    k_scan_bins<<<n_vali, 1024, 0, st>>>(bg);
    k_given_scatter<<<grid, 256, 0, st>>>(a31, a32, cs, sn, view_off, bin_tmp, bg, s_a31, s_a32, s_cs, s_sn, s_idx);
}

// ----------------------------------------------------------------------------- pass-level tables
__global__ void k_h2_angles(const double* __restrict__ x2, const double* __restrict__ y2, int n, double ex, double ey, double* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = ore_angle(x2[i], y2[i], ex, ey);   // getOreList, EdgeSketch_Core.cpp:161
}
void launch_h2_angles(const double* x2, const double* y2, int n, double ex, double ey, double* out, cudaStream_t st) {
    if (n > 0) k_h2_angles<<<(n + 255) / 256, 256, 0, st>>>(x2, y2, n, ex, ey, out);
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
    __shared__ long long part[1024];
    const int per = (n + 1023) / 1024;
    const int lo = min(threadIdx.x * per, n), hi = min(lo + per, n);
    long long s = 0;
    for (int i = lo; i < hi; i++) s += in[i];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        long long t = (threadIdx.x >= d) ? part[threadIdx.x - d] : 0;
        __syncthreads();
        part[threadIdx.x] += t;
        __syncthreads();
    }
    long long run = part[threadIdx.x] - s;
    for (int i = lo; i < hi; i++) { out[i] = run; run += in[i]; }
    if (threadIdx.x == 1023) out[n] = part[1023];
}
void launch_scan_i32_i64(const int32_t* in, int64_t* out, int n, cudaStream_t st) { k_scan_i32_i64<<<1, 1024, 0, st>>>(in, out, n); }

// ----------------------------------------------------------------------------- clustering (one warp)
// EdgeClusterer.cpp:60-125; every lane runs the same sequential sums (ascending member index, as the reference).
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

// EdgeClusterer.cpp:127-293 for the n rows of one view-1 edge.
__device__ void cluster_warp(int n, const PassConst& pc, const double* xc, const double* yc, const double* th, double* thn, double* orient,
                             int32_t* label, const int32_t* h2orig, double* xbar, double* ybar, double* thbar, int32_t* h2rep, int lane) {
    for (int k = lane; k < n; k += 32) {
        const double t = normalize_orientation(th[k], pc.cl_wrap_deg);   // :131-135
        thn[k] = t; orient[k] = t; label[k] = k;
    }
    __syncwarp();
    bool merged = true;
    while (merged) {                                                       // :139-189
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
                    if (dist < bd && dist < pc.cl_dist && fabs(oi - orient[lj]) < pc.cl_orient_rad) { bd = dist; bj = j; }   // :154
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {   // arg-min, ties -> lowest j (the sequential scan keeps the first minimum)
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
                if (so + sn <= pc.max_cluster) {                           // :166
                    double gx, gy, gth;
                    gauss_avg_warp(n, xc, yc, thn, label, old_l, new_l, pc.cl_sigma, &gx, &gy, &gth);
                    __syncwarp();
                    if (lane == 0) orient[new_l] = normalize_orientation(gth, pc.cl_wrap_deg);   // :175-176
                    for (int j = lane; j < n; j += 32) if (label[j] == old_l) label[j] = new_l;   // :179-183
                    __syncwarp();
                    merged = true;
                    break;
                }
            }
        }
    }
    // :250-290 per cluster: Gaussian-weighted average written to every member, representative = member closest to it
    for (int i = 0; i < n; i++) {
        const int li = label[i];
        bool first = true;
        for (int j = lane; j < i; j += 32) if (label[j] == li) first = false;
        if (!__all_sync(FULL, first)) continue;
        double gx, gy, gth;
        gauss_avg_warp(n, xc, yc, thn, label, li, li, pc.cl_sigma, &gx, &gy, &gth);
        double md = DBL_MAX;
        int closest = -1;
        for (int j = i; j < n; j++) {
            if (label[j] == li) {
                const double ddx = xc[j] - gx, ddy = yc[j] - gy;
                const double dist = sqrt(ddx * ddx + ddy * ddy);
                if (dist < md) { md = dist; closest = j; }                 // :270
            }
        }
        const int rep = h2orig[closest];                                   // corrected(closest, 8), :282
        for (int j = i + lane; j < n; j += 32)
            if (label[j] == li) { xbar[j] = gx; ybar[j] = gy; thbar[j] = gth; h2rep[j] = rep; }
    }
}

// ----------------------------------------------------------------------------- formation
// One warp per view-1 edge: Skip_this_Edge (hpp:92-110), getHYPO2_idx_Ore (PairEdgeHypo.cpp:88-102) as a
// ballot-compacted scan over the view-2 angle list (keeps the reference's ascending-index order for free),
// edgelsHYPO2_epipolar_correction (PairEdgeHypo.cpp:225-293), then clustering.
template <bool FILL>
__global__ void __launch_bounds__(128) k_form(const __grid_constant__ PassConst pc, const double* __restrict__ x1, const double* __restrict__ y1,
                                              int h1_begin, int h1_end, const double2* __restrict__ wedge_h1, const double* __restrict__ x2,
                                              const double* __restrict__ y2, const double* __restrict__ th2, const double* __restrict__ ang2,
                                              int E2, int32_t* m_out, int32_t* n_out, const int64_t* __restrict__ row_off, RowTable rt) {
    const int lane = threadIdx.x & 31;
    const int h1 = h1_begin + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (h1 >= h1_end) return;
    const double x = x1[h1], y = y1[h1];
    const bool skip = (x < 10 || x > double(pc.cols - 10) || y < 10 || y > double(pc.rows - 10));   // hpp:94
    const double2 w = wedge_h1[h1];
    int n_total = 0;
    int64_t row0 = 0;
    if (FILL) {
        n_total = n_out[h1];
        if (n_total == 0) return;
        row0 = row_off[h1];
    } else if (skip || w.x != w.x) {
        if (lane == 0) { m_out[h1] = 0; n_out[h1] = 0; }
        return;
    }
    double coef[3];
    epip_line_coeffs(pc.F21, x, y, coef);
    const unsigned lt = lanemask_lt();
    int mbase = 0, nbase = 0;
    for (int base = 0; base < E2; base += 32) {
        const int i = base + lane;
        bool in = false, ok = false;
        double cx = 0, cy = 0, t2 = 0;
        if (i < E2) {
            const double a = ang2[i];
            in = (a > w.x && a < w.y);                                       // PairEdgeHypo.cpp:94 (strict both sides)
        }
        if (in) {
            t2 = th2[i];
            ok = correct_one(coef, x2[i], y2[i], t2, pc.loc_pert, pc.tang_thresh, pc.orient_pert, &cx, &cy);
        }
        const unsigned mm = __ballot_sync(FULL, in), mo = __ballot_sync(FULL, ok);
        if (FILL) {
            if (in) {
                const int k = mbase + __popc(mm & lt);                       // pre-correction position (col 9)
                if (k < n_total) rt.rawat[row0 + k] = i;                     // HYPO2_idx_raw(local row): EdgeClusterer.cpp:221 quirk
                if (ok) {
                    const int64_t r = row0 + nbase + __popc(mo & lt);
                    rt.h1[r] = h1; rt.h2orig[r] = i; rt.local[r] = k;
                    rt.xc[r] = cx; rt.yc[r] = cy; rt.th[r] = t2;             // theta written back is the ORIGINAL one (:288)
                }
            }
        }
        mbase += __popc(mm); nbase += __popc(mo);
    }
    if (!FILL) {
        if (lane == 0) { m_out[h1] = mbase; n_out[h1] = nbase; }
        return;
    }
    __syncwarp();
    cluster_warp(n_total, pc, rt.xc + row0, rt.yc + row0, rt.th + row0, rt.thn + row0, rt.orient + row0, rt.label + row0, rt.h2orig + row0,
                 rt.xbar + row0, rt.ybar + row0, rt.thbar + row0, rt.h2rep + row0, lane);
}
void launch_form_count(const PassConst& pc, const double* x1, const double* y1, int h1_begin, int h1_end, const double2* wedge_h1,
                       const double* x2, const double* y2, const double* th2, const double* ang2, int E2, int32_t* m_out, int32_t* n_out,
                       cudaStream_t st) {
    const int n = h1_end - h1_begin;
    if (n <= 0) return;
    RowTable rt = {};
    k_form<false><<<(n + 3) / 4, 128, 0, st>>>(pc, x1, y1, h1_begin, h1_end, wedge_h1, x2, y2, th2, ang2, E2, m_out, n_out, nullptr, rt);
}
void launch_form_fill(const PassConst& pc, const double* x1, const double* y1, int h1_begin, int h1_end, const double2* wedge_h1,
                      const double* x2, const double* y2, const double* th2, const double* ang2, int E2, const int32_t* n_of,
                      const int64_t* row_off, RowTable rt, cudaStream_t st) {
    const int n = h1_end - h1_begin;
    if (n <= 0) return;
    k_form<true><<<(n + 3) / 4, 128, 0, st>>>(pc, x1, y1, h1_begin, h1_end, wedge_h1, x2, y2, th2, ang2, E2, nullptr, const_cast<int32_t*>(n_of),
                                             row_off, rt);
}

// ----------------------------------------------------------------------------- validation: geometry
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
    if (f) atomicAdd(&ctr->n_degenerate, 1ULL);
}
void launch_wedge31(const PassConst& pc, const ViewConst* vc, const double* x1, const double* y1, const int32_t* n_of, int h1_begin, int h1_end,
                    int E1, double2* w31, DevCounters* ctr, cudaStream_t st) {
    const int n = h1_end - h1_begin;
    if (n <= 0 || pc.n_vali <= 0) return;
    dim3 grid((n + 127) / 128, pc.n_vali);
    k_wedge31<<<grid, 128, 0, st>>>(pc, vc, x1, y1, n_of, h1_begin, h1_end, E1, w31, ctr);
}

// per hypothesis row x validation view: reprojected tangent (getGamma3LocationAndTangent, EdgeSketch_Core.cpp:330)
// and wedge of the clustered view-2 edge (getOreListBar, :340).  FP64-transcendental bound (7 atan, asin, 6 sin/cos, ~24 div).
__global__ void __launch_bounds__(128) k_reproject(const __grid_constant__ PassConst pc, const ViewConst* __restrict__ vc, const double* __restrict__ x1,
                                                   const double* __restrict__ y1, const double* __restrict__ th1, RowTable rt, int64_t row_begin,
                                                   int64_t row_end, double2* __restrict__ w32, double2* __restrict__ tgt, DevCounters* ctr) {
    const int64_t r = row_begin + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    const int v = blockIdx.y;
    if (r >= row_end) return;
    const int h1 = rt.h1[r];
    const double xb = rt.xbar[r], yb = rt.ybar[r], tb = rt.thbar[r];
    RowGeom g;
    row_geometry(x1[h1], y1[h1], th1[h1], xb, yb, tb, pc, &g);
    const ViewConst& c = vc[v];
    double t3x, t3y;
    reproject_tangent(g, c.R31, c.T31, &t3x, &t3y);
    double lo, hi;
    const int f = wedge_row(xb, yb, c.F32, c.e1_32[0], c.e1_32[1], c.e2_32[0], c.e2_32[1], pc.delta, pc.rows, pc.cols, &lo, &hi);
    const int64_t o = int64_t(v) * (row_end - row_begin) + (r - row_begin);
    w32[o] = make_double2(lo, hi);
    tgt[o] = make_double2(t3x, t3y);
    if (f) atomicAdd(&ctr->n_degenerate, 1ULL);
}
void launch_reproject(const PassConst& pc, const ViewConst* vc, const double* x1, const double* y1, const double* th1, RowTable rt, int64_t row_begin,
                      int64_t row_end, double2* w32, double2* tgt, DevCounters* ctr, cudaStream_t st) {
    const int64_t n = row_end - row_begin;
    if (n <= 0 || pc.n_vali <= 0) return;
    dim3 grid(unsigned((n + 127) / 128), pc.n_vali);
    k_reproject<<<grid, 128, 0, st>>>(pc, vc, x1, y1, th1, rt, row_begin, row_end, w32, tgt, ctr);
}

// ----------------------------------------------------------------------------- validation: the support test
// One warp per (group = view-1 edge, validation view).  The members of wedge31 are enumerated from the view's
// angular bin grid (bins covered by [lo31, hi31], exact test on the staged angle), compacted into shared memory,
// and every lane owns one hypothesis row: wedge32 membership (PairEdgeHypo.cpp:94), parallel-wedge test
// (EdgeSketch_Core.hpp:112-128), orientation arg-max with the reference's first-maximum rule
// (getSupportedEdgels.cpp:32-47: d >= thr && d > prev, ascending index -> ties go to the lowest index).
#define SUP_WARPS 4
#define SUP_TILE 64
__global__ void __launch_bounds__(SUP_WARPS * 32) k_support(const __grid_constant__ SupportArgs a) {
    __shared__ double sm_a32[SUP_WARPS][SUP_TILE];
    __shared__ double sm_cs[SUP_WARPS][SUP_TILE];
    __shared__ double sm_sn[SUP_WARPS][SUP_TILE];
    __shared__ int32_t sm_idx[SUP_WARPS][SUP_TILE];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int g = a.g_begin + blockIdx.x * SUP_WARPS + wid;
    const int v = blockIdx.y;
    if (g >= a.g_end) return;
    const int64_t r_beg = a.group_off[g], r_end = a.group_off[g + 1];
    if (r_end <= r_beg) return;
    const double2 w1 = a.w31[int64_t(v) * a.w31_stride + g];
    const bool w1_bad = (w1.x != w1.x);
    int beg = 0, end = 0;
    int64_t off = 0;
    if (!w1_bad) {
        const int view = a.vali_view[v];
        off = a.edge_off[view];
        const int32_t* bins = a.bg.bins + int64_t(v) * a.bg.BS;
        const int b0 = bin_of(w1.x, a.bg.scale, a.bg.B), b1 = bin_of(w1.y, a.bg.scale, a.bg.B);
        beg = (b0 == 0) ? 0 : bins[b0 - 1];
        end = bins[b1];
    }
    const unsigned lt = lanemask_lt();
    unsigned long long n_sup = 0, n_par = 0, n_deg = 0, n_inl = 0;
    for (int64_t rc = r_beg; rc < r_end; rc += 32) {
        const int64_t r = rc + lane;
        const bool valid = r < r_end;
        double2 w2 = make_double2(0, 0), tg = make_double2(0, 0);
        bool active = false;
        if (valid) {
            const int64_t o = int64_t(v) * a.chunk_rows + (r - a.row_base);
            w2 = ldg2(a.w32 + o);
            tg = ldg2(a.tgt + o);
            const bool w2_bad = (w2.x != w2.x);
            if (w2_bad) n_deg++;
            if (!w1_bad && !w2_bad) {
                const double d0 = fabs(w1.x - w2.x), d1 = fabs(w1.x - w2.y), d2 = fabs(w1.y - w2.x), d3 = fabs(w1.y - w2.y);
                const double dmax = fmax(fmax(d0, d1), fmax(d2, d3));
                if (dmax <= a.parallel_deg) n_par++;                        // hpp:120 (inclusive)
                else active = true;
            }
        }
        double best = 0.0;   // prev_prod, getSupportedEdgels.cpp:33
        int bidx = -2;
        if (__any_sync(FULL, active)) {
            for (int tb = beg; tb < end; tb += SUP_TILE) {
                int cnt = 0;
#pragma unroll
                for (int s = 0; s < SUP_TILE / 32; s++) {
                    const int j = tb + s * 32 + lane;
                    bool in = false;
                    if (j < end) { const double a31 = a.s_a31[off + j]; in = (a31 > w1.x && a31 < w1.y); }
                    const unsigned m = __ballot_sync(FULL, in);
                    if (in) {
                        const int p = cnt + __popc(m & lt);
                        sm_a32[wid][p] = a.s_a32[off + j]; sm_cs[wid][p] = a.s_cs[off + j]; sm_sn[wid][p] = a.s_sn[off + j];
                        sm_idx[wid][p] = a.s_idx[off + j];
                    }
                    cnt += __popc(m);
                }
                __syncwarp();
                if (active) {
                    for (int k = 0; k < cnt; k++) {
                        const double a32 = sm_a32[wid][k];
                        if (a32 > w2.x && a32 < w2.y) {
                            n_inl++;
                            const double d = fabs(tg.x * sm_cs[wid][k] + tg.y * sm_sn[wid][k]);
                            const int idx = sm_idx[wid][k];
                            if (d >= a.cos_thresh && (d > best || (d == best && bidx != -2 && idx < bidx))) { best = d; bidx = idx; }
                        }
                    }
                }
                __syncwarp();
            }
        }
        if (valid) {
            a.support[int64_t(v) * a.sup_stride + r] = bidx;
            if (bidx != -2) n_sup++;
        }
    }
    // warp-aggregated statistics
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        n_sup += __shfl_xor_sync(FULL, n_sup, o); n_par += __shfl_xor_sync(FULL, n_par, o); n_deg += __shfl_xor_sync(FULL, n_deg, o);
        n_inl += __shfl_xor_sync(FULL, n_inl, o);
    }
    if (lane == 0 && a.ctr) {
        if (n_sup) atomicAdd(&a.ctr->n_supported, n_sup);
        if (n_par) atomicAdd(&a.ctr->n_parallel, n_par);
        if (n_inl) atomicAdd(&a.ctr->n_inliers, n_inl);
        (void)n_deg;
    }
}
void launch_support(const SupportArgs& a, cudaStream_t st) {
    const int n = a.g_end - a.g_begin;
    if (n <= 0 || a.n_vali <= 0) return;
    dim3 grid((n + SUP_WARPS - 1) / SUP_WARPS, a.n_vali);
    k_support<<<grid, SUP_WARPS * 32, 0, st>>>(a);
}

// ----------------------------------------------------------------------------- vote + ratio test
// EdgeSketch_Core.cpp:414-508.  One warp per view-1 edge.  flags bit0 = valid (count >= N), bit1 = passes count > ratio * max.
__global__ void __launch_bounds__(128) k_select(const __grid_constant__ PassConst pc, const int64_t* __restrict__ row_off, int h1_begin, int h1_end,
                                                const int32_t* __restrict__ support, int64_t sup_stride, RowTable rt, int32_t* n_valid, int32_t* n_sel) {
    const int lane = threadIdx.x & 31;
    const int h1 = h1_begin + blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (h1 >= h1_end) return;
    const int64_t r0 = row_off[h1], r1 = row_off[h1 + 1];
    int maxc = 0, nv = 0;
    for (int64_t r = r0 + lane; r < r1; r += 32) {
        int c = 0;
        for (int v = 0; v < pc.n_vali; v++) c += (support[int64_t(v) * sup_stride + r] != -2);
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
    if (lane == 0) { n_valid[h1] = nv; n_sel[h1] = ns; }
}
void launch_select(const PassConst& pc, const int64_t* row_off, int h1_begin, int h1_end, const int32_t* support, int64_t sup_stride, RowTable rt,
                   int32_t* n_valid, int32_t* n_sel, cudaStream_t st) {
    const int n = h1_end - h1_begin;
    if (n <= 0) return;
    k_select<<<(n + 3) / 4, 128, 0, st>>>(pc, row_off, h1_begin, h1_end, support, sup_stride, rt, n_valid, n_sel);
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
__global__ void k_gather_supports(const int32_t* __restrict__ p_row, int64_t P, int n_vali, const int32_t* __restrict__ support, int64_t sup_stride,
                                  int32_t* __restrict__ out) {
    const int64_t t = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (t >= P * n_vali) return;
    const int64_t p = t / n_vali;
    const int v = int(t % n_vali);
    out[t] = support[int64_t(v) * sup_stride + p_row[p]];
}
void launch_gather_supports(const int32_t* p_row, int64_t P, int n_vali, const int32_t* support, int64_t sup_stride, int32_t* out, cudaStream_t st) {
    const int64_t n = P * n_vali;
    if (n > 0) k_gather_supports<<<unsigned((n + 255) / 256), 256, 0, st>>>(p_row, P, n_vali, support, sup_stride, out);
}
// [V][stride] -> [H][V] (the reference's supported_indices layout) through a 32x32 shared-memory tile
__global__ void k_transpose_support(const int32_t* __restrict__ support, int64_t sup_stride, int64_t H, int n_vali, int32_t* __restrict__ out) {
    __shared__ int32_t tile[32][33];
    const int64_t r0 = blockIdx.x * 32LL;
    const int v0 = blockIdx.y * 32;
    for (int k = threadIdx.y; k < 32; k += blockDim.y) {
        const int v = v0 + k;
        const int64_t r = r0 + threadIdx.x;
        tile[k][threadIdx.x] = (v < n_vali && r < H) ? support[int64_t(v) * sup_stride + r] : 0;
    }
    __syncthreads();
    for (int k = threadIdx.y; k < 32; k += blockDim.y) {
        const int64_t r = r0 + k;
        const int v = v0 + threadIdx.x;
        if (r < H && v < n_vali) out[r * n_vali + v] = tile[threadIdx.x][k];
    }
}
void launch_transpose_support(const int32_t* support, int64_t sup_stride, int64_t H, int n_vali, int32_t* out, cudaStream_t st) {
    if (H <= 0 || n_vali <= 0) return;
    dim3 grid(unsigned((H + 31) / 32), (n_vali + 31) / 32), block(32, 8);
    k_transpose_support<<<grid, block, 0, st>>>(support, sup_stride, H, n_vali, out);
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

}  // namespace cvx

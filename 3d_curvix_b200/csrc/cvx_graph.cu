// =====================================================================================
// cvx_graph.cu -- the data-parallel front half of the reference's EdgeMapping::consolidate_3D_edges, behind the
// C-ABI (include/curvix_b200.h, "3-D edge graph"):
//
//   build3DEdgeWeightedGraph                       edge_mapping.cpp:107-192   cvx_graph_build_links   k_pair_count / k_pair_emit + radix sort + unique
//   pruneEdgeGraph_by_3DProximityAndOrientation    edge_mapping.cpp:229-335   cvx_graph_prune         k_link_metrics + reductions + k_link_keep
//   pruneEdgeGraphbyProjections                    edge_mapping.cpp:337-405   cvx_graph_prune         k_link_project
//   buildEdgeNodeGraph (neighbour lists)           edge_mapping.cpp:411-458   cvx_graph_prune         k_directed + radix sort -> CSR
//   align3DEdgesUsingEdgeNodes                     edge_mapping.cpp:460-620   cvx_graph_align         k_align_step x NUM_OF_ITERATIONS
//
// The reference keys everything on the exact doubles of a 3-D point in unordered_maps; here a 3-D edge is a NODE INDEX (the caller
// numbers the distinct 3-D points), a link is a pair of node indices, and the hash maps become a sort + unique of 64-bit keys.  The
// reference's results depend on the iteration order of its hash maps only through floating-point summation order (link statistics,
// per-node neighbour sums) and through the order-dependent host stages that follow (connectivity graph, curve tracing: host code of
// the drop-in, 3d_curvix_b200/host/edge_mapping.cpp); here every order is the ascending index order.
// Arithmetic is FP64 in the reference's operation order (-fmad=false); only libm calls differ (CUDA libdevice vs glibc, <= 2 ulp).
// =====================================================================================
#include <cub/cub.cuh>

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "cvx_graph_math.cuh"
#include "cvx_internal.h"

namespace {

using namespace cvxg;

template <typename T>
struct GBuf {   // grow-only device buffer
    T* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        const size_t want = n + n / 8 + 64;
        const cudaError_t e = cudaMalloc(&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    ~GBuf() { if (p) cudaFree(p); }
};

}  // namespace

struct cvx_graph {
    cvx_ctx* ctx = nullptr;
    int device = 0;
    cudaStream_t st = nullptr;
    int64_t n = 0;                                   // nodes
    GBuf<double> loc, tan0, ori, loc2, ori2;         // n x 3 each: location, tangent of the node's first support, orientation (+ ping-pong)
    GBuf<unsigned long long> keys, keys2;            // links as (min node << 32 | max node); sorted, unique
    GBuf<unsigned long long> pos, pos2;              // position of a candidate pair in the curvelet-order stream (first occurrence after the unique)
    std::vector<unsigned long long> h_pos;           // first occurrence of every stage-0 link
    std::vector<unsigned char> h_keep3d, h_keepproj; // per stage-0 link: survives the 3-D pruning / the projection test
    GBuf<unsigned char> tmp;                         // cub scratch
    GBuf<double> met;                                // 3 x L link metrics: normal distance, tangential distance, angle
    GBuf<double> dev2;                               // L scratch for the squared deviations
    GBuf<unsigned char> flag;                        // L keep flags
    GBuf<long long> off;                             // scans
    GBuf<int32_t> adj;                               // neighbour lists (CSR over nodes), ascending neighbour index
    GBuf<long long> adj_off;                         // n + 1
    GBuf<double> red;                                // reduction outputs
    GBuf<long long> cnt;                             // device counters
    int64_t L0 = 0, L1 = 0, L2 = 0;                  // links: built / after the 3-D proximity + orientation pruning / after the projection pruning
    std::vector<unsigned long long> h_links[3];      // host copies per stage (fetched on demand)
    double stats[6] = {0, 0, 0, 0, 0, 0};            // mu_t, sigma_t, mu_n, sigma_n, mu_theta, sigma_theta
    bool have_adj = false;
};

namespace {

#define GFAIL(code, msg) cvx_internal_fail(g ? g->ctx : nullptr, (code), (msg))
#define GCU(call)                                                                                           \
    do {                                                                                                    \
        cudaError_t _e = (call);                                                                            \
        if (_e != cudaSuccess) return GFAIL(CVX_ERR_CUDA, (std::string(#call) + ": " + cudaGetErrorString(_e)).c_str()); \
    } while (0)

// ---- links from curvelets (edge_mapping.cpp:122-187).  A curvelet names two 2-D edges of one image; the records of both (the 3-D edges they
// support, in the caller's record order) are concatenated and every pair (i < j) of records links its two 3-D edges unless the two points share
// a coordinate (:182).  One warp per curvelet.
__device__ __forceinline__ int rec_count(const long long* __restrict__ key_off, int k) { return k < 0 ? 0 : int(key_off[k + 1] - key_off[k]); }
__global__ void __launch_bounds__(256) k_pair_count(int64_t n_cv, const int32_t* __restrict__ cv_self, const int32_t* __restrict__ cv_nb,
                                                    const long long* __restrict__ key_off, long long* __restrict__ out) {
    const int64_t c = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (c >= n_cv) return;
    const long long r = rec_count(key_off, cv_self[c]) + rec_count(key_off, cv_nb[c]);
    out[c] = r >= 2 ? r * (r - 1) / 2 : 0;
}
__global__ void __launch_bounds__(256) k_pair_emit(int64_t n_cv, const int32_t* __restrict__ cv_self, const int32_t* __restrict__ cv_nb,
                                                   const long long* __restrict__ key_off, const int32_t* __restrict__ key_nodes,
                                                   const double* __restrict__ loc, const long long* __restrict__ pair_off,
                                                   unsigned long long* __restrict__ keys) {
    const int64_t c = blockIdx.x * int64_t(blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (c >= n_cv) return;
    const int ks = cv_self[c], kn = cv_nb[c];
    const int rs = rec_count(key_off, ks), rn = rec_count(key_off, kn);
    const int r = rs + rn;
    if (r < 2) return;
    const long long bs = ks < 0 ? 0 : key_off[ks], bn = kn < 0 ? 0 : key_off[kn];
    unsigned long long* out = keys + pair_off[c];
    for (int i = 0; i < r - 1; i++) {
        const int a = i < rs ? key_nodes[bs + i] : key_nodes[bn + (i - rs)];
        const D3 pa = ld3(loc, a);
        const long long row = (long long)i * r - (long long)i * (i + 1) / 2;     // pairs of the rows before i
        for (int j = i + 1 + lane; j < r; j += 32) {
            const int b = j < rs ? key_nodes[bs + j] : key_nodes[bn + (j - rs)];
            const D3 pb = ld3(loc, b);
            unsigned long long k = ~0ULL;                                        // dropped: sorts to the end
            if (pa.x != pb.x && pa.y != pb.y && pa.z != pb.z)                    // :182
                k = ((unsigned long long)(unsigned)min(a, b) << 32) | (unsigned)max(a, b);
            out[row + (j - i - 1)] = k;
        }
    }
}
// the reference orders a pair by coordinates (:174-178): first = the lexicographically smaller point
__device__ __forceinline__ void link_ends(unsigned long long k, const double* __restrict__ loc, int* first, int* second) {
    const int a = int(k >> 32), b = int(k & 0xffffffffu);
    const D3 pa = ld3(loc, a), pb = ld3(loc, b);
    const bool swap = (pa.x > pb.x) || (pa.x == pb.x && pa.y > pb.y) || (pa.x == pb.x && pa.y == pb.y && pa.z > pb.z);
    *first = swap ? b : a; *second = swap ? a : b;
}
// edge_mapping.cpp:247-277
__global__ void __launch_bounds__(256) k_link_metrics(int64_t L, const unsigned long long* __restrict__ keys, const double* __restrict__ loc,
                                                      const double* __restrict__ tan0, double* __restrict__ met) {
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= L) return;
    int a, b;
    link_ends(keys[i], loc, &a, &b);
    const D3 p1 = ld3(loc, a), p2 = ld3(loc, b);
    const D3 t1 = normalized3(ld3(tan0, a)), t2 = normalized3(ld3(tan0, b));
    double d = fabs(dot3(t1, t2));
    d = fmin(1.0, fmax(-1.0, d));
    met[2 * L + i] = acos(d) * 180 / kPI;
    met[i] = norm3(closest_vector_to_line(p2, t2, p1));       // normal distance
    met[L + i] = norm3(sub3(p1, p2));                         // "tangential" distance (:272: the distance of the two points)
}
__global__ void __launch_bounds__(256) k_sq_dev(int64_t L, const double* __restrict__ v, const double* __restrict__ sum, double* __restrict__ out) {
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= L) return;
    const double mean = *sum / double(L);
    out[i] = (v[i] - mean) * (v[i] - mean);
}
// edge_mapping.cpp:297-323 (thresholds on the host from the reductions) and :362-395 (projection into every view)
struct PruneArgs {
    double thr_n, thr_t, thr_a, proj_thresh;
    int n_views;
    const double *R, *T, *K;      // n_views x 9, n_views x 3, 9 (row-major)
};
__global__ void __launch_bounds__(256) k_link_keep(int64_t L, const double* __restrict__ met, PruneArgs pa, unsigned char* __restrict__ flag) {
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= L) return;
    flag[i] = (met[i] < pa.thr_n && met[L + i] < pa.thr_t && met[2 * L + i] < pa.thr_a) ? 1 : 0;
}
__global__ void __launch_bounds__(256) k_link_project(int64_t L, const unsigned long long* __restrict__ keys, const double* __restrict__ loc,
                                                      PruneArgs pa, unsigned char* __restrict__ flag) {
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= L) return;
    int a, b;
    link_ends(keys[i], loc, &a, &b);
    const D3 X1 = ld3(loc, a), X2 = ld3(loc, b);
    M33 K;
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) K.m[r][c] = pa.K[3 * r + c];
    bool prune = false;
    for (int v = 0; v < pa.n_views && !prune; v++) {
        M33 R;
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) R.m[r][c] = pa.R[9 * v + 3 * r + c];
        const D3 T = ld3(pa.T, v);
        D3 q1 = mat_vec3(K, add3(mat_vec3(R, X1), T)), q2 = mat_vec3(K, add3(mat_vec3(R, X2), T));
        q1.x /= q1.z; q1.y /= q1.z; q1.z = 1.0;                 // util.cpp:169-174
        q2.x /= q2.z; q2.y /= q2.z; q2.z = 1.0;
        prune = norm3(sub3(q1, q2)) > pa.proj_thresh;           // :381
    }
    flag[i] = prune ? 0 : 1;
}
__global__ void __launch_bounds__(256) k_directed(int64_t L, const unsigned long long* __restrict__ keys, unsigned long long* __restrict__ out) {
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= L) return;
    const unsigned long long k = keys[i];
    out[2 * i] = k;
    out[2 * i + 1] = (k << 32) | (k >> 32);
}
__global__ void __launch_bounds__(256) k_adj_fill(int64_t M, const unsigned long long* __restrict__ dk, int32_t* __restrict__ adj,
                                                  long long* __restrict__ adj_off, int64_t n) {
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i > M) return;
    // adj_off[s] = first position whose source is >= s: written at every boundary of the sorted list
    const long long src = i < M ? (long long)(dk[i] >> 32) : n;
    const long long prev = i > 0 ? (long long)(dk[i - 1] >> 32) : -1;
    for (long long s = prev + 1; s <= src; s++) adj_off[s] = i;
    if (i < M) adj[i] = int32_t(dk[i] & 0xffffffffu);
}
// edge_mapping.cpp:488-585, one synchronous (Jacobi) iteration: one thread per node, neighbours in list order
__global__ void __launch_bounds__(128) k_align_step(int64_t n, const long long* __restrict__ adj_off, const int32_t* __restrict__ adj,
                                                    const double* __restrict__ loc, const double* __restrict__ ori, double* __restrict__ loc_out,
                                                    double* __restrict__ ori_out, double step_force, double step_torque) {
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    const D3 B = ld3(loc, i), tB = ld3(ori, i);
    const long long e0 = adj_off[i], e1 = adj_off[i + 1];
    const int deg = int(e1 - e0);
    if (deg == 0) {
        loc_out[3 * i] = B.x; loc_out[3 * i + 1] = B.y; loc_out[3 * i + 2] = B.z;
        ori_out[3 * i] = tB.x; ori_out[3 * i + 1] = tB.y; ori_out[3 * i + 2] = tB.z;
        return;
    }
    // mean / standard deviation of the distances to the neighbours (:506-520; std::accumulate order)
    double s = 0.0;
    for (long long e = e0; e < e1; e++) s += norm3(sub3(ld3(loc, adj[e]), B));
    const double mu = s / double(deg);
    double sq = 0.0;
    for (long long e = e0; e < e1; e++) { const double d = norm3(sub3(ld3(loc, adj[e]), B)); sq = sq + (d - mu) * (d - mu); }
    const double sigma = fmax(sqrt(sq / double(deg)), 0.001);
    D3 force{0.0, 0.0, 0.0}, euler{0.0, 0.0, 0.0};
    double tw = 0.0;
    for (long long e = e0; e < e1; e++) {
        const int j = adj[e];
        const D3 p = ld3(loc, j), tn = ld3(ori, j);
        const D3 tang = closest_vector_to_line(p, tn, B);                        // :541
        const D3 ea = shortest_align_euler_deg(tB, tn);                          // :544
        const double d = norm3(sub3(p, B));
        const double q = (d - mu) / sigma;
        const double w = exp(-0.5 * (q * q));                                    // :548
        force = add3(force, mul3(w, tang));
        euler = add3(euler, mul3(w, ea));
        tw += w;
    }
    if (tw > 0) {                                                                // :560-565: divided by the neighbour count, not by the weight
        force = D3{force.x / double(deg), force.y / double(deg), force.z / double(deg)};
        euler = D3{euler.x / double(deg), euler.y / double(deg), euler.z / double(deg)};
    }
    const D3 nl = add3(B, mul3(step_force, force));                              // :568
    euler = D3{euler.x * step_torque, euler.y * step_torque, euler.z * step_torque};
    euler = D3{euler.x * kMPI / 180.0, euler.y * kMPI / 180.0, euler.z * kMPI / 180.0};
    const M33 R = euler_to_rotation(euler.x, euler.y, euler.z);
    const D3 no = mat_vec3(R, tB);                                               // :575
    loc_out[3 * i] = nl.x; loc_out[3 * i + 1] = nl.y; loc_out[3 * i + 2] = nl.z;
    ori_out[3 * i] = no.x; ori_out[3 * i + 1] = no.y; ori_out[3 * i + 2] = no.z;
}

__global__ void __launch_bounds__(256) k_iota(int64_t N, unsigned long long* __restrict__ v) {
    const int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (i < N) v[i] = (unsigned long long)i;
}
// keys (the candidate pairs in curvelet order) -> sorted distinct keys in g->keys with the stream position of their FIRST occurrence in g->pos
// (the radix sort is stable, UniqueByKey keeps the head of every run); dropped pairs (~0) are removed
int sort_unique(cvx_graph* g, int64_t N, int64_t* n_out) {
    size_t t1 = 0, t2 = 0;
    GCU(g->keys2.reserve(size_t(N))); GCU(g->pos.reserve(size_t(N))); GCU(g->pos2.reserve(size_t(N)));
    GCU(g->cnt.reserve(4));
    k_iota<<<unsigned((N + 255) / 256), 256, 0, g->st>>>(N, g->pos.p);
    cub::DeviceRadixSort::SortPairs(nullptr, t1, g->keys.p, g->keys2.p, g->pos.p, g->pos2.p, N, 0, 64, g->st);
    cub::DeviceSelect::UniqueByKey(nullptr, t2, g->keys2.p, g->pos2.p, g->keys.p, g->pos.p, g->cnt.p, N, g->st);
    GCU(g->tmp.reserve(std::max(t1, t2)));
    t1 = t2 = g->tmp.cap;
    GCU(cub::DeviceRadixSort::SortPairs(g->tmp.p, t1, g->keys.p, g->keys2.p, g->pos.p, g->pos2.p, N, 0, 64, g->st));
    GCU(cub::DeviceSelect::UniqueByKey(g->tmp.p, t2, g->keys2.p, g->pos2.p, g->keys.p, g->pos.p, g->cnt.p, N, g->st));
    long long u = 0;
    unsigned long long last = 0;
    GCU(cudaMemcpyAsync(&u, g->cnt.p, 8, cudaMemcpyDeviceToHost, g->st));
    GCU(cudaStreamSynchronize(g->st));
    if (u > 0) {
        GCU(cudaMemcpyAsync(&last, g->keys.p + (u - 1), 8, cudaMemcpyDeviceToHost, g->st));
        GCU(cudaStreamSynchronize(g->st));
        if (last == ~0ULL) u--;
    }
    g->h_pos.resize(size_t(u));
    if (u > 0) {
        GCU(cudaMemcpyAsync(g->h_pos.data(), g->pos.p, size_t(u) * 8, cudaMemcpyDeviceToHost, g->st));
        GCU(cudaStreamSynchronize(g->st));
    }
    *n_out = u;
    return CVX_OK;
}
int fetch_links(cvx_graph* g, int stage, int64_t L) {
    g->h_links[stage].resize(size_t(L));
    if (L > 0) {
        GCU(cudaMemcpyAsync(g->h_links[stage].data(), g->keys.p, size_t(L) * 8, cudaMemcpyDeviceToHost, g->st));
        GCU(cudaStreamSynchronize(g->st));
    }
    return CVX_OK;
}
// mean and population standard deviation of v[0..L) as the reference's mean_std lambda (edge_mapping.cpp:279-285)
int mean_std(cvx_graph* g, const double* v, int64_t L, double* mu, double* sigma) {
    *mu = *sigma = 0.0;
    if (L <= 0) return CVX_OK;
    size_t t = 0;
    GCU(g->red.reserve(2));
    GCU(g->dev2.reserve(size_t(L)));
    cub::DeviceReduce::Sum(nullptr, t, v, g->red.p, L, g->st);
    GCU(g->tmp.reserve(t));
    t = g->tmp.cap;
    GCU(cub::DeviceReduce::Sum(g->tmp.p, t, v, g->red.p, L, g->st));
    k_sq_dev<<<unsigned((L + 255) / 256), 256, 0, g->st>>>(L, v, g->red.p, g->dev2.p);
    t = g->tmp.cap;
    GCU(cub::DeviceReduce::Sum(g->tmp.p, t, g->dev2.p, g->red.p + 1, L, g->st));
    double h[2];
    GCU(cudaMemcpyAsync(h, g->red.p, 16, cudaMemcpyDeviceToHost, g->st));
    GCU(cudaStreamSynchronize(g->st));
    *mu = h[0] / double(L);
    *sigma = std::sqrt(h[1] / double(L));
    cvx_internal_count_launch(g->ctx, 3);
    return CVX_OK;
}

}  // namespace

extern "C" {

int cvx_graph_create(cvx_ctx* ctx, int64_t n_nodes, const double* location, const double* first_tangent, cvx_graph** out) {
    cvx_graph* g = nullptr;
    if (!ctx || !out || n_nodes < 0 || (n_nodes > 0 && (!location || !first_tangent))) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_graph_create: bad arguments");
    if (n_nodes > 0x7fffffffLL) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_graph_create: more than 2^31 3-D edges");
    g = new cvx_graph();
    g->ctx = ctx; g->n = n_nodes;
    if (cvx_internal_device(ctx, &g->device, &g->st) != CVX_OK) { delete g; return CVX_ERR_INVALID; }
    cudaSetDevice(g->device);
    const size_t m = size_t(std::max<int64_t>(n_nodes, 1)) * 3;
    cudaError_t e = g->loc.reserve(m);
    if (e == cudaSuccess) e = g->tan0.reserve(m);
    if (e == cudaSuccess && n_nodes) e = cudaMemcpyAsync(g->loc.p, location, size_t(n_nodes) * 24, cudaMemcpyHostToDevice, g->st);
    if (e == cudaSuccess && n_nodes) e = cudaMemcpyAsync(g->tan0.p, first_tangent, size_t(n_nodes) * 24, cudaMemcpyHostToDevice, g->st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(g->st);
    if (e != cudaSuccess) { const int rc = GFAIL(CVX_ERR_CUDA, (std::string("cvx_graph_create: ") + cudaGetErrorString(e)).c_str()); delete g; return rc; }
    *out = g;
    return CVX_OK;
}

int cvx_graph_destroy(cvx_graph* g) {
    if (!g) return CVX_OK;
    cudaSetDevice(g->device);
    delete g;
    return CVX_OK;
}

int cvx_graph_build_links(cvx_graph* g, int64_t n_keys, const int64_t* key_off, const int32_t* key_nodes, int64_t n_curvelets,
                          const int32_t* cv_self_key, const int32_t* cv_neighbor_key, int64_t* n_links) {
    if (!g || n_keys < 0 || n_curvelets < 0 || !key_off || (n_curvelets > 0 && (!cv_self_key || !cv_neighbor_key)))
        return GFAIL(CVX_ERR_INVALID, "cvx_graph_build_links: bad arguments");
    GCU(cudaSetDevice(g->device));
    const int64_t n_rec = key_off[n_keys];
    for (int64_t c = 0; c < n_curvelets; c++)
        if (cv_self_key[c] >= n_keys || cv_neighbor_key[c] >= n_keys) return GFAIL(CVX_ERR_INVALID, "cvx_graph_build_links: curvelet names a 2-D edge beyond n_keys");
    GBuf<long long> d_key_off, d_pairs;
    GBuf<int32_t> d_nodes, d_self, d_nb;
    GCU(d_key_off.reserve(size_t(n_keys) + 1)); GCU(d_nodes.reserve(size_t(std::max<int64_t>(n_rec, 1))));
    GCU(d_self.reserve(size_t(std::max<int64_t>(n_curvelets, 1)))); GCU(d_nb.reserve(size_t(std::max<int64_t>(n_curvelets, 1))));
    GCU(d_pairs.reserve(size_t(n_curvelets) + 1));
    GCU(cudaMemcpyAsync(d_key_off.p, key_off, size_t(n_keys + 1) * 8, cudaMemcpyHostToDevice, g->st));
    if (n_rec) GCU(cudaMemcpyAsync(d_nodes.p, key_nodes, size_t(n_rec) * 4, cudaMemcpyHostToDevice, g->st));
    g->L0 = g->L1 = g->L2 = 0; g->have_adj = false;
    if (n_curvelets == 0 || g->n == 0) { if (n_links) *n_links = 0; return fetch_links(g, 0, 0); }
    GCU(cudaMemcpyAsync(d_self.p, cv_self_key, size_t(n_curvelets) * 4, cudaMemcpyHostToDevice, g->st));
    GCU(cudaMemcpyAsync(d_nb.p, cv_neighbor_key, size_t(n_curvelets) * 4, cudaMemcpyHostToDevice, g->st));
    k_pair_count<<<unsigned((n_curvelets + 255) / 256), 256, 0, g->st>>>(n_curvelets, d_self.p, d_nb.p, d_key_off.p, d_pairs.p);
    {   // exclusive scan of the pair counts (n_curvelets + 1 entries: the last one is the total)
        size_t t = 0;
        GCU(cudaMemsetAsync(d_pairs.p + n_curvelets, 0, 8, g->st));
        cub::DeviceScan::ExclusiveSum(nullptr, t, d_pairs.p, d_pairs.p, n_curvelets + 1, g->st);
        GCU(g->tmp.reserve(t));
        t = g->tmp.cap;
        GCU(cub::DeviceScan::ExclusiveSum(g->tmp.p, t, d_pairs.p, d_pairs.p, n_curvelets + 1, g->st));
    }
    long long total = 0;
    GCU(cudaMemcpyAsync(&total, d_pairs.p + n_curvelets, 8, cudaMemcpyDeviceToHost, g->st));
    GCU(cudaStreamSynchronize(g->st));
    if (total > (1LL << 32)) return GFAIL(CVX_ERR_INVALID, "cvx_graph_build_links: more than 2^32 candidate pairs in one call");
    int64_t L = 0;
    if (total > 0) {
        GCU(g->keys.reserve(size_t(total)));
        k_pair_emit<<<unsigned((n_curvelets + 7) / 8), 256, 0, g->st>>>(n_curvelets, d_self.p, d_nb.p, d_key_off.p, d_nodes.p, g->loc.p, d_pairs.p, g->keys.p);
        const int rc = sort_unique(g, total, &L);
        if (rc != CVX_OK) return rc;
    }
    GCU(cudaGetLastError());
    cvx_internal_count_launch(g->ctx, 5);
    g->L0 = L;
    if (n_links) *n_links = L;
    return fetch_links(g, 0, L);
}

int cvx_graph_prune(cvx_graph* g, double lambda1, double lambda2, double lambda3, double proj_thresh, int n_views, const double* R,
                    const double* T, const double* K, double* stats6, int64_t* n_after_3d, int64_t* n_after_projection) {
    if (!g || n_views < 0 || (n_views > 0 && (!R || !T || !K))) return GFAIL(CVX_ERR_INVALID, "cvx_graph_prune: bad arguments");
    GCU(cudaSetDevice(g->device));
    const int64_t L = g->L0;
    int64_t L1 = 0, L2 = 0;
    PruneArgs pa;
    std::memset(&pa, 0, sizeof(pa));
    GBuf<double> d_R, d_T, d_K;
    GBuf<unsigned char> f2;
    g->h_keep3d.assign(size_t(L), 0); g->h_keepproj.assign(size_t(L), 1);
    if (L > 0) {
        // the sorted distinct links of stage 0 live in h_links[0]; g->keys may have been compacted by an earlier call
        GCU(g->keys.reserve(size_t(L)));
        GCU(cudaMemcpyAsync(g->keys.p, g->h_links[0].data(), size_t(L) * 8, cudaMemcpyHostToDevice, g->st));
        GCU(g->met.reserve(size_t(L) * 3)); GCU(g->flag.reserve(size_t(L))); GCU(f2.reserve(size_t(L)));
        k_link_metrics<<<unsigned((L + 255) / 256), 256, 0, g->st>>>(L, g->keys.p, g->loc.p, g->tan0.p, g->met.p);
        int rc = mean_std(g, g->met.p + L, L, &g->stats[0], &g->stats[1]);              // tangential
        if (rc == CVX_OK) rc = mean_std(g, g->met.p, L, &g->stats[2], &g->stats[3]);     // normal
        if (rc == CVX_OK) rc = mean_std(g, g->met.p + 2 * L, L, &g->stats[4], &g->stats[5]);   // angle
        if (rc != CVX_OK) return rc;
        pa.thr_n = g->stats[2] + g->stats[3] * lambda1; pa.thr_t = g->stats[0] + g->stats[1] * lambda3; pa.thr_a = g->stats[4] + g->stats[5] * lambda2;   // :319-320
        k_link_keep<<<unsigned((L + 255) / 256), 256, 0, g->st>>>(L, g->met.p, pa, g->flag.p);
        GCU(cudaMemcpyAsync(g->h_keep3d.data(), g->flag.p, size_t(L), cudaMemcpyDeviceToHost, g->st));
        if (n_views > 0) {   // the projection test of every stage-0 link (the reference runs it on the survivors; the result per link is the same)
            GCU(d_R.reserve(size_t(n_views) * 9)); GCU(d_T.reserve(size_t(n_views) * 3)); GCU(d_K.reserve(9));
            GCU(cudaMemcpyAsync(d_R.p, R, size_t(n_views) * 72, cudaMemcpyHostToDevice, g->st));
            GCU(cudaMemcpyAsync(d_T.p, T, size_t(n_views) * 24, cudaMemcpyHostToDevice, g->st));
            GCU(cudaMemcpyAsync(d_K.p, K, 72, cudaMemcpyHostToDevice, g->st));
            pa.proj_thresh = proj_thresh; pa.n_views = n_views; pa.R = d_R.p; pa.T = d_T.p; pa.K = d_K.p;
            k_link_project<<<unsigned((L + 255) / 256), 256, 0, g->st>>>(L, g->keys.p, g->loc.p, pa, f2.p);
            GCU(cudaMemcpyAsync(g->h_keepproj.data(), f2.p, size_t(L), cudaMemcpyDeviceToHost, g->st));
        }
        GCU(cudaStreamSynchronize(g->st));
        g->h_links[1].clear(); g->h_links[2].clear();
        for (int64_t i = 0; i < L; i++) {
            if (!g->h_keep3d[size_t(i)]) continue;
            g->h_links[1].push_back(g->h_links[0][size_t(i)]);
            if (g->h_keepproj[size_t(i)]) g->h_links[2].push_back(g->h_links[0][size_t(i)]);
        }
        L1 = int64_t(g->h_links[1].size()); L2 = int64_t(g->h_links[2].size());
    } else {
        g->h_links[1].clear(); g->h_links[2].clear();
    }
    g->L1 = L1; g->L2 = L2;
    // default neighbour lists of the surviving links (edge_mapping.cpp:440-455): ascending neighbour index per node
    GCU(g->adj_off.reserve(size_t(g->n) + 2)); GCU(g->adj.reserve(size_t(std::max<int64_t>(2 * L2, 1))));
    if (L2 > 0) {
        size_t t = 0;
        GCU(g->keys.reserve(size_t(L2))); GCU(g->keys2.reserve(size_t(2 * L2)));
        GBuf<unsigned long long> dk;
        GCU(dk.reserve(size_t(2 * L2)));
        GCU(cudaMemcpyAsync(g->keys.p, g->h_links[2].data(), size_t(L2) * 8, cudaMemcpyHostToDevice, g->st));
        k_directed<<<unsigned((L2 + 255) / 256), 256, 0, g->st>>>(L2, g->keys.p, g->keys2.p);
        cub::DeviceRadixSort::SortKeys(nullptr, t, g->keys2.p, dk.p, 2 * L2, 0, 64, g->st);
        GCU(g->tmp.reserve(t));
        t = g->tmp.cap;
        GCU(cub::DeviceRadixSort::SortKeys(g->tmp.p, t, g->keys2.p, dk.p, 2 * L2, 0, 64, g->st));
        k_adj_fill<<<unsigned((2 * L2 + 1 + 255) / 256), 256, 0, g->st>>>(2 * L2, dk.p, g->adj.p, g->adj_off.p, g->n);
        GCU(cudaStreamSynchronize(g->st));
    } else {
        GCU(cudaMemsetAsync(g->adj_off.p, 0, size_t(g->n + 1) * 8, g->st));
    }
    GCU(cudaStreamSynchronize(g->st));
    GCU(cudaGetLastError());
    cvx_internal_count_launch(g->ctx, 6);
    g->have_adj = true;
    if (stats6) std::memcpy(stats6, g->stats, sizeof(g->stats));
    if (n_after_3d) *n_after_3d = L1;
    if (n_after_projection) *n_after_projection = L2;
    return CVX_OK;
}

int cvx_graph_get_links(cvx_graph* g, int stage, int32_t* node_a, int32_t* node_b, int64_t* n) {
    if (!g || stage < 0 || stage > 2) return GFAIL(CVX_ERR_INVALID, "cvx_graph_get_links: bad arguments");
    const std::vector<unsigned long long>& h = g->h_links[stage];
    if (n) *n = int64_t(h.size());
    for (size_t i = 0; i < h.size(); i++) {
        if (node_a) node_a[i] = int32_t(h[i] >> 32);
        if (node_b) node_b[i] = int32_t(h[i] & 0xffffffffu);
    }
    return CVX_OK;
}

int cvx_graph_get_link_details(cvx_graph* g, int64_t* first_occurrence, uint8_t* keep_3d, uint8_t* keep_projection) {
    if (!g) return CVX_ERR_INVALID;
    const size_t L = g->h_links[0].size();
    if (first_occurrence) for (size_t i = 0; i < L && i < g->h_pos.size(); i++) first_occurrence[i] = int64_t(g->h_pos[i]);
    if (keep_3d && g->h_keep3d.size() == L) std::memcpy(keep_3d, g->h_keep3d.data(), L);
    if (keep_projection && g->h_keepproj.size() == L) std::memcpy(keep_projection, g->h_keepproj.data(), L);
    return CVX_OK;
}

int cvx_graph_set_neighbors(cvx_graph* g, const int64_t* offsets, const int32_t* neighbors) {
    if (!g || !offsets) return GFAIL(CVX_ERR_INVALID, "cvx_graph_set_neighbors: bad arguments");
    GCU(cudaSetDevice(g->device));
    const int64_t m = offsets[g->n];
    if (m < 0 || (m > 0 && !neighbors)) return GFAIL(CVX_ERR_INVALID, "cvx_graph_set_neighbors: bad offsets");
    for (int64_t i = 0; i < m; i++) if (neighbors[i] < 0 || neighbors[i] >= g->n) return GFAIL(CVX_ERR_INVALID, "cvx_graph_set_neighbors: neighbour out of range");
    GCU(g->adj_off.reserve(size_t(g->n) + 2)); GCU(g->adj.reserve(size_t(std::max<int64_t>(m, 1))));
    GCU(cudaMemcpyAsync(g->adj_off.p, offsets, size_t(g->n + 1) * 8, cudaMemcpyHostToDevice, g->st));
    if (m) GCU(cudaMemcpyAsync(g->adj.p, neighbors, size_t(m) * 4, cudaMemcpyHostToDevice, g->st));
    GCU(cudaStreamSynchronize(g->st));
    g->have_adj = true;
    return CVX_OK;
}

int cvx_graph_get_neighbors(cvx_graph* g, int64_t* offsets, int32_t* neighbors) {
    if (!g || !g->have_adj) return GFAIL(CVX_ERR_INVALID, "cvx_graph_get_neighbors: cvx_graph_prune has not run");
    GCU(cudaSetDevice(g->device));
    long long m = 0;
    GCU(cudaMemcpyAsync(&m, g->adj_off.p + g->n, 8, cudaMemcpyDeviceToHost, g->st));
    GCU(cudaStreamSynchronize(g->st));
    if (offsets) GCU(cudaMemcpyAsync(offsets, g->adj_off.p, size_t(g->n + 1) * 8, cudaMemcpyDeviceToHost, g->st));
    if (neighbors && m > 0) GCU(cudaMemcpyAsync(neighbors, g->adj.p, size_t(m) * 4, cudaMemcpyDeviceToHost, g->st));
    GCU(cudaStreamSynchronize(g->st));
    return CVX_OK;
}

int cvx_graph_align(cvx_graph* g, const double* orientation, int iterations, double step_force, double step_torque, double* location_out,
                    double* orientation_out, float* kernel_ms) {
    if (!g || !g->have_adj || iterations < 0 || (g->n > 0 && !orientation)) return GFAIL(CVX_ERR_INVALID, "cvx_graph_align: bad arguments or cvx_graph_prune has not run");
    GCU(cudaSetDevice(g->device));
    const int64_t n = g->n;
    if (n == 0) { if (kernel_ms) *kernel_ms = 0.f; return CVX_OK; }
    const size_t m = size_t(n) * 3;
    GCU(g->ori.reserve(m)); GCU(g->loc2.reserve(m)); GCU(g->ori2.reserve(m));
    GBuf<double> cur;   // the node locations stay untouched in g->loc (the link pruning may be repeated); iterate on a copy
    GCU(cur.reserve(m));
    GCU(cudaMemcpyAsync(cur.p, g->loc.p, m * 8, cudaMemcpyDeviceToDevice, g->st));
    GCU(cudaMemcpyAsync(g->ori.p, orientation, m * 8, cudaMemcpyHostToDevice, g->st));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, g->st);
    double *la = cur.p, *oa = g->ori.p, *lb = g->loc2.p, *ob = g->ori2.p;
    for (int it = 0; it < iterations; it++) {
        k_align_step<<<unsigned((n + 127) / 128), 128, 0, g->st>>>(n, g->adj_off.p, g->adj.p, la, oa, lb, ob, step_force, step_torque);
        std::swap(la, lb); std::swap(oa, ob);
    }
    cudaEventRecord(e1, g->st);
    if (location_out) GCU(cudaMemcpyAsync(location_out, la, m * 8, cudaMemcpyDeviceToHost, g->st));
    if (orientation_out) GCU(cudaMemcpyAsync(orientation_out, oa, m * 8, cudaMemcpyDeviceToHost, g->st));
    GCU(cudaStreamSynchronize(g->st));
    GCU(cudaGetLastError());
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (kernel_ms) *kernel_ms = ms;
    cvx_internal_count_launch(g->ctx, iterations);
    return CVX_OK;
}

}  // extern "C"

// =====================================================================================
// cvx_graph_align_no_alloc -- a variant of cvx_graph_align for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
// loops, harnesses that forbid it).  No counterpart in the reference.  The alignment iterates
// between two buffers; here one of them is the caller's output and the other a file-scope device array of a fixed maximum size, which
// every GPU that loads the library holds until exit (CVX_ALIGN_NO_ALLOC_MAX_NODES x 6 doubles = 3 MiB).  The entry point calls no
// cudaMalloc / cudaFree / allocating library routine, creates no stream or event and does not synchronise: it only enqueues kernels (and, for
// iterations == 0, two device-to-device copies) on the caller's stream.  Every call on a device shares the array: calls that overlap on the
// GPU (different streams, no ordering between them) race.  Nothing in it has to start at zero -- every iteration writes the row of every
// node before the next one reads it -- so the contents left by an earlier call are never read.
// =====================================================================================
#define CVX_ALIGN_NO_ALLOC_MAX_NODES_ 65536
static __device__ __align__(256) double g_align_scratch[size_t(CVX_ALIGN_NO_ALLOC_MAX_NODES_) * 6];   // locations [0, 3 max), orientations [3 max, 6 max)

namespace {
constexpr int kMaxDevices = 64;
double* g_align_scratch_addr[kMaxDevices];   // device address of g_align_scratch on every device, resolved once per device
bool g_align_scratch_ready[kMaxDevices];
std::mutex g_align_scratch_mu;

int no_alloc_fail(char* errbuf, int errbuf_len, int code, const std::string& msg) {
    if (errbuf && errbuf_len > 0) { std::snprintf(errbuf, size_t(errbuf_len), "%s", msg.c_str()); }
    return code;
}
}  // namespace

extern "C" int cvx_graph_align_no_alloc(int device, void* stream, int64_t n_nodes, const int64_t* d_offsets, const int32_t* d_neighbors,
                                        const double* d_location, const double* d_orientation, int iterations, double step_force,
                                        double step_torque, double* d_location_out, double* d_orientation_out, char* errbuf, int errbuf_len) {
    if (n_nodes < 0 || iterations < 0 || device < 0 || device >= kMaxDevices)
        return no_alloc_fail(errbuf, errbuf_len, CVX_ERR_INVALID, "cvx_graph_align_no_alloc: bad arguments");
    if (n_nodes > CVX_ALIGN_NO_ALLOC_MAX_NODES_)
        return no_alloc_fail(errbuf, errbuf_len, CVX_ERR_INVALID,
                             "cvx_graph_align_no_alloc: " + std::to_string(n_nodes) + " nodes exceed the fixed scratch of " +
                                 std::to_string(CVX_ALIGN_NO_ALLOC_MAX_NODES_) + " nodes (use cvx_graph_align, which allocates)");
    if (n_nodes == 0) return CVX_OK;
    if (!d_offsets || !d_neighbors || !d_location || !d_orientation || !d_location_out || !d_orientation_out)
        return no_alloc_fail(errbuf, errbuf_len, CVX_ERR_INVALID, "cvx_graph_align_no_alloc: null device pointer");
    if (d_location_out == d_location || d_orientation_out == d_orientation)
        return no_alloc_fail(errbuf, errbuf_len, CVX_ERR_INVALID, "cvx_graph_align_no_alloc: outputs must not alias the inputs");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return no_alloc_fail(errbuf, errbuf_len, CVX_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    double* scratch = nullptr;
    {   // set-up, once per device, with that device current: in host code the array's name is a stand-in, not a device address
        std::lock_guard<std::mutex> lock(g_align_scratch_mu);
        if (!g_align_scratch_ready[device]) {
            double* p = nullptr;
            e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_align_scratch);
            if (e != cudaSuccess || !p)
                return no_alloc_fail(errbuf, errbuf_len, CVX_ERR_CUDA, std::string("cudaGetSymbolAddress(g_align_scratch): ") + cudaGetErrorString(e));
            g_align_scratch_addr[device] = p;
            g_align_scratch_ready[device] = true;
        }
        scratch = g_align_scratch_addr[device];
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t bytes = size_t(n_nodes) * 24;
    if (iterations == 0) {
        e = cudaMemcpyAsync(d_location_out, d_location, bytes, cudaMemcpyDeviceToDevice, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(d_orientation_out, d_orientation, bytes, cudaMemcpyDeviceToDevice, st);
        if (e != cudaSuccess) return no_alloc_fail(errbuf, errbuf_len, CVX_ERR_CUDA, std::string("cudaMemcpyAsync: ") + cudaGetErrorString(e));
        return CVX_OK;
    }
    double* s_loc = scratch;
    double* s_ori = scratch + size_t(CVX_ALIGN_NO_ALLOC_MAX_NODES_) * 3;
    // ping-pong between the caller's outputs and the scratch so that the LAST iteration writes the outputs
    const double *la = d_location, *oa = d_orientation;
    for (int it = 0; it < iterations; it++) {
        const bool to_out = ((iterations - 1 - it) & 1) == 0;
        double* lb = to_out ? d_location_out : s_loc;
        double* ob = to_out ? d_orientation_out : s_ori;
        k_align_step<<<unsigned((n_nodes + 127) / 128), 128, 0, st>>>(n_nodes, reinterpret_cast<const long long*>(d_offsets), d_neighbors, la, oa, lb, ob,
                                                                     step_force, step_torque);
        la = lb; oa = ob;
    }
    e = cudaGetLastError();
    if (e != cudaSuccess) return no_alloc_fail(errbuf, errbuf_len, CVX_ERR_CUDA, std::string("k_align_step: ") + cudaGetErrorString(e));
    return CVX_OK;
}

/* =====================================================================================
 * curvix_b200.h -- C-ABI of the B200-native 3D Curvix edge-sketch hot path.
 *
 * The reference (C-H-Chien/3D_Curvix) has no FFI/plugin layer: its boundary is the C++ class
 * EdgeSketch_Core as driven by cmd/main.cpp:114-170.  This header is what a re-implementation of that
 * class binds instead of the in-process Eigen/OpenMP code; each entry point cites the reference
 * interface it replaces (paths relative to the reference root).  Plain pointers and sizes only; the
 * caller owns every host buffer, the context owns all device memory and its stream; variable-length
 * outputs use a size-then-fill pattern (cvx_get_counts first).  Every function returns 0 on success
 * or a negative cvx_status; the message is available through cvx_last_error().  Nothing here ever
 * calls exit() or throws, and there is no CPU fallback: without a CUDA device cvx_create fails.
 *
 * One context per GPU, used from one host thread at a time.  Multi-GPU = one context per device, each
 * given a disjoint [h1_begin, h1_end) range (or a disjoint set of hypothesis view pairs).
 * ===================================================================================== */
#ifndef CURVIX_B200_H
#define CURVIX_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum cvx_status {
    CVX_OK = 0,
    CVX_ERR_INVALID = -1,    /* bad argument / call order */
    CVX_ERR_CUDA = -2,       /* CUDA runtime error (message has the cudaError string) */
    CVX_ERR_NO_DEVICE = -3,  /* no CUDA device: there is deliberately no CPU path */
    CVX_ERR_NOMEM = -4
} cvx_status;

/* Run-time YAML knobs (EdgeSketch_Core.cpp:37-59) and compile-time macros (Edge_Reconst/definitions.h:14-26,54)
 * of the reference, passed as one POD so the thresholds travel with the call. */
typedef struct cvx_config {
    double delta;                      /* YAML delta: edge location perturbation      EdgeSketch_Core.cpp:40 */
    double delta_theta_deg;            /* YAML delta_theta                            EdgeSketch_Core.cpp:41 */
    double parallel_angle_deg;         /* YAML Parallel_Epipolar_Line_Angle           EdgeSketch_Core.cpp:44 */
    double lowes_ratio;                /* LOWES_RATIO_THRESHOLD (0.6), must be in [0,1) definitions.h:14 (>= 1 would need the
                                        * reference's unstable-sort fallback, EdgeSketch_Core.cpp:505-507: cvx_create rejects it) */
    double epip_tangency_displ_thresh; /* EPIP_TANGENCY_DISPL_THRESH (3)              definitions.h:20 */
    double location_perturbation;      /* LOCATION_PERTURBATION (0.4)                 definitions.h:21 */
    double orient_perturbation;        /* ORIENT_PERTURBATION (0.174533)              definitions.h:22 */
    double cluster_dist_thresh;        /* CLUSTER_DIST_THRESH (1)                     definitions.h:23 */
    double cluster_orient_thresh_deg;  /* CLUSTER_ORIENT_THRESH (20.0)                definitions.h:24 */
    double cluster_gauss_sigma;        /* CLUSTER_ORIENT_GAUSS_SIGMA (2.0)            definitions.h:26 */
    double pi_truncated;               /* PI (3.1415926, truncated on purpose)        definitions.h:54 */
    int32_t n_views;                   /* YAML Total_Num_Of_Images */
    int32_t img_rows;                  /* YAML Image_Rows */
    int32_t img_cols;                  /* YAML Image_Cols */
    int32_t max_support_views;         /* YAML Max_Num_Of_Support_Views (must be >= 1) */
    int32_t max_cluster_size;          /* MAX_CLUSTER_SIZE (10)                       definitions.h:25 */
    int32_t strict_libm_order;   /* 0 (default): the validation kernel evaluates the view-2 wedge with the libm round trips
                                  * removed (wedge_row_fast, bounds equal to a few ulp); 1: the reference's libm call order */
} cvx_config;

/* Sizes of the variable-length outputs of the last cvx_run_edge_sketch + run statistics. */
typedef struct cvx_counts {
    int64_t n_h1_total;          /* edges in hypothesis view 1 */
    int64_t n_h1_processed;      /* H1 edges that reached validation (n >= 1) */
    int64_t n_candidates;        /* sum of m: raw epipolar-wedge candidates        PairEdgeHypo.cpp:88-102 */
    int64_t n_hypotheses;        /* sum of n: rows surviving epipolar correction   PairEdgeHypo.cpp:225-293 */
    int64_t n_validations;       /* sum of n * V: the metric's unit                EdgeSketch_Core.cpp:355 */
    int64_t n_valid_rows;        /* rows with >= max_support_views supports        EdgeSketch_Core.cpp:430-435 */
    int64_t n_pairs;             /* accepted rows after the ratio test             EdgeSketch_Core.cpp:510-526 */
    int64_t n_degenerate_wedges; /* wedges with != 2 border crossings or NaN (reference reads stale memory there) */
    int64_t n_supported;         /* validations that returned a support index != -2 */
    int64_t n_parallel;          /* validations short-circuited by is_Epipolar_Wedges_in_Parallel */
    int64_t n_inliers;           /* sum over validations of |wedge31 ∩ wedge32| (edges that reached getSupportIdx) */
    int64_t n_unique_hypotheses; /* clusters: rows of one cluster carry the same averaged edge and are validated once */
    int64_t n_entries_visited;   /* bin-grid entries tested by the validation kernel (per unique hypothesis x view, unweighted) */
    int64_t n_heavy_queries;     /* grid queries walked by a whole warp (wide wedges / long cells) */
} cvx_counts;

typedef struct cvx_ctx cvx_ctx;

/* Fills *c with the values of configs/3D_Curvix_ABC_NEF.yaml:4-27 and definitions.h. */
void cvx_default_config(cvx_config* c);

/* replaces EdgeSketch_Core::EdgeSketch_Core(YAML::Node)             EdgeSketch_Core.cpp:32-93 (device side) */
int cvx_create(const cvx_config* cfg, int device, cvx_ctx** out);
int cvx_destroy(cvx_ctx* ctx);
const char* cvx_last_error(const cvx_ctx* ctx);

/* replaces Read_Camera_Data                                         EdgeSketch_Core.cpp:95-113
 * R: n x 9 row-major world->camera rotations; T: n x 3; K: 9 doubles (k_per_view = 0) or n x 9 (k_per_view = 1) */
int cvx_set_cameras(cvx_ctx* ctx, int n, const double* R, const double* T, const double* K, int k_per_view);

/* replaces Read_Edgels_Data / file_reader::read_All_Edgels           EdgeSketch_Core.hpp:87-90
 * Host SoA columns of Edges/Edge_<view>_t<thr>.txt; copied to the device once and kept resident across passes.
 * strength may be NULL (it is carried through, never computed on). */
int cvx_set_view_edges(cvx_ctx* ctx, int view, int n, const double* x, const double* y, const double* theta,
                       const double* strength);

/* Same, all views in one call (what Read_Edgels_Data does: it loads every view, EdgeSketch_Core.hpp:87-90).
 * x, y, theta are the per-view columns concatenated in view order (view v starts at n[0] + ... + n[v-1]); n_views must equal
 * cfg.n_views.  Three host->device copies and one kernel instead of 3 copies + 1 kernel per view; pinned host memory streams at
 * link speed. */
int cvx_set_all_view_edges(cvx_ctx* ctx, int n_views, const int32_t* n, const double* x, const double* y, const double* theta);

/* replaces Set_Hypothesis_Views_Camera + Set_Hypothesis_Views_Edgels EdgeSketch_Core.cpp:115-166
 * Computes the per-pass constants (R21,T21,F21, per-view F/epipoles), the epipolar angles of the view-2 edges
 * (getOreList), the wedges of the view-1 edges (getOreListBar) and stages every validation view
 * (two epipolar angles, cos/sin of the orientation) into its angular bin grid. */
int cvx_set_hypothesis_pair(cvx_ctx* ctx, int h1, int h2);

/* replaces Run_3D_Edge_Sketch for view-1 edges [h1_begin, h1_end)    EdgeSketch_Core.cpp:168-573
 * (h1_end < 0 means "to the last edge").  Asynchronous on the context stream until a getter is called. */
int cvx_run_edge_sketch(cvx_ctx* ctx, int h1_begin, int h1_end);

int cvx_get_counts(cvx_ctx* ctx, cvx_counts* out);

/* Pass-level arrays, for parity checks: OreListdegree (E2) and OreListBardegree (E1 x 2, NaN when flagged). */
int cvx_get_h2_angles(cvx_ctx* ctx, double* out);
int cvx_get_h1_wedges(cvx_ctx* ctx, double* out, int32_t* flag);

/* Per view-1 edge (E1 entries): m, n and the first row of the edge in the hypothesis table (-1 when n == 0). */
int cvx_get_h1_table(cvx_ctx* ctx, int32_t* m, int32_t* n, int64_t* row_off);

/* The hypothesis table (n_hypotheses rows, grouped by view-1 edge, ascending view-2 index inside a group):
 * h2_orig = corrected(:,8), local_idx = corrected(:,9), raw_at_local = HYPO2_idx_raw(local row) (the reference's
 * cluster-map key quirk, EdgeClusterer.cpp:216-221), (xc,yc) epipolar-corrected location, (xbar,ybar,thbar)
 * cluster average written back to every member, label = cluster label, h2_rep = HYPO2_idx (EdgeClusterer.cpp:282),
 * count = number of supporting validation views.  Any pointer may be NULL.   replaces hypo2_clusters (hpp:47-48) */
int cvx_get_rows(cvx_ctx* ctx, int32_t* h1, int32_t* h2_orig, int32_t* local_idx, int32_t* raw_at_local, double* xc,
                 double* yc, double* theta_orig, double* xbar, double* ybar, double* thbar, int32_t* label,
                 int32_t* h2_rep, int32_t* count);

/* replaces all_supported_indices (EdgeSketch_Core.hpp:160): H x V int32 row-major, -2 = no support */
int cvx_get_support_matrix(cvx_ctx* ctx, int32_t* out);
/* Intermediates of the validation stage (debug/parity): wedge31 (E1 x V x 2), wedge32 and tgt (H x V x 2).  wedge32 / tgt are
 * only materialised when cvx_set_keep_intermediates(ctx, 1) was called before the run (the production path keeps them in
 * registers). */
int cvx_set_keep_intermediates(cvx_ctx* ctx, int on);
int cvx_get_intermediates(cvx_ctx* ctx, double* wedge31, double* wedge32, double* tgt);

/* Rows with count >= max_support_views, ascending (h1,row): the input of the reference's std::sort + ratio test
 * (EdgeSketch_Core.cpp:480-508); selected = passes count > lowes_ratio * max. */
int cvx_get_valid_rows(cvx_ctx* ctx, int32_t* h1, int32_t* row, int32_t* count, int32_t* selected);

/* replaces paired_edge / paired_edge_final (EdgeSketch_Core.cpp:510-526,582-601): accepted rows in ascending
 * (h1, row) order; supports is P x V.  The reference's within-edge order (std::sort by count) and its per-edge
 * capacity of n_views rows are applied by the host shim from cvx_get_valid_rows. */
int cvx_get_pairs(cvx_ctx* ctx, int32_t* h1, int32_t* h2rep, int32_t* row, int32_t* count, int32_t* supports);

/* The reference's host-side ordering of the accepted rows of each view-1 edge (EdgeSketch_Core.cpp:480-526): std::sort of
 * the valid rows by support count (descending; literally std::sort, so libstdc++'s unstable order above 16 elements is
 * reproduced), ratio-test filter, capacity of n_views rows per edge.  Writes indices into the arrays of cvx_get_pairs /
 * cvx_finalize in paired_edge order; idx must hold n_pairs entries; *n_out <= n_pairs (smaller only if the capacity cut). */
int cvx_get_reference_order(cvx_ctx* ctx, int64_t* idx, int64_t* n_out);

/* replaces Finalize_Edge_Pairs_and_Reconstruct_3D_Edges (geometry) + the transform of Stack_3D_Edges
 * EdgeSketch_Core.cpp:625-676,742-748.  P x 3 each, h2_corrected_xy P x 2, nan_flag P (bit0 NaN point -> zero row,
 * bit1 representative dropped by the correction).  Any pointer may be NULL. */
int cvx_finalize(cvx_ctx* ctx, double* gamma_cam, double* gamma_world, double* tangent_world,
                 double* h2_corrected_xy, int32_t* nan_flag);

/* Device-resident results of the last run/finalize for a device-to-device gather (NCCL): pointers stay valid
 * until the next cvx_run_edge_sketch on this context. */
typedef struct cvx_device_results {
    int64_t n_pairs; int32_t n_vali_views; int32_t pad;
    const int32_t* pair_h1; const int32_t* pair_h2rep; const int32_t* pair_count; const int32_t* pair_supports;
    const double* gamma_cam; const double* gamma_world; const double* tangent_world;
} cvx_device_results;
int cvx_get_device_results(cvx_ctx* ctx, cvx_device_results* out);

/* Balanced split of the view-1 edges of the current pair into `parts` contiguous ranges of about equal HYPOTHESIS-ROW count
 * (the candidate count pass of Run_3D_Edge_Sketch, EdgeSketch_Core.cpp:223-247, run over all of view 1; an equal-edge split is
 * unbalanced when the edges come in scan order).  bounds has parts + 1 entries: range k = [bounds[k], bounds[k+1]).  This is the
 * only split a data-dependent pass schedule (configs 1) can use across GPUs: every rank runs its range of the same pair. */
int cvx_split_h1_by_rows(cvx_ctx* ctx, int parts, int32_t* bounds);

/* ---- multi-GPU: the path's only exchange step, the gather of a unit's results to rank 0 ----------------------------
 * One context per GPU (one process each, or one thread each in a single process).  The consumer is the rank-0 loop that feeds
 * EdgeMapping::add3DToSupportingEdgesMapping (EdgeSketch_Core.cpp:682-726) and the output writers (:735-774).
 * A "unit" is one cvx_run_edge_sketch (+ finalize) call: a whole hypothesis view pair, or a view-1 range of one.  Gathers are
 * collective ROUNDS: in every round each rank contributes one unit (or nothing); counts travel with one ncclAllGather of a
 * 64-byte record per rank, payloads with grouped ncclSend / ncclRecv of the exact byte counts to rank 0 on a side stream, so the
 * transfer of round k overlaps the compute of round k + 1.  NCCL is loaded at run time (dlopen of libnccl.so.2); contexts that
 * never call these functions do not need it.
 *
 * cvx_comm_unique_id fills 128 bytes (ncclUniqueId) on one rank; the caller broadcasts them by any means and every rank calls
 * cvx_comm_init.  cvx_comm_attach adopts a communicator the caller created (ncclComm_t as void*, e.g. from ncclCommInitAll). */
int cvx_comm_unique_id(void* id128);
int cvx_comm_init(cvx_ctx* ctx, int world, int rank, const void* id128);
int cvx_comm_attach(cvx_ctx* ctx, void* nccl_comm, int world, int rank);
int cvx_comm_destroy(cvx_ctx* ctx);
/* Starts a new sequence of rounds (rank 0 rewinds its receive arena). */
int cvx_gather_reset(cvx_ctx* ctx);
/* Contributes the results of the last run of this context (empty != 0: this rank has no unit in this round) as round
 * number = calls since cvx_gather_reset.  unit_id is carried to rank 0 untouched.  Returns after enqueueing; the context may be
 * used for the next unit immediately (its results were copied to a staging buffer on the context stream). */
int cvx_gather_post(cvx_ctx* ctx, int64_t unit_id, int empty);
/* Completes every posted round.  With join_stream != 0 the context stream is made to wait for the transfers (so an event
 * recorded on it afterwards times them); with host_sync != 0 the call also blocks until they are done. */
int cvx_gather_wait(cvx_ctx* ctx, int join_stream, int host_sync);
/* Rank 0, after cvx_gather_wait: what `rank` contributed in `round`.  Device pointers into the receive arena, valid until the next
 * cvx_gather_reset.  meta: {n_pairs, view1, view2, h1_begin, h1_end, unit_id, n_vali_views, n_supports}.
 * The supporting edges of the accepted rows (columns 2.. of the reference's paired_edge, EdgeSketch_Core.cpp:510-526) travel as a CSR
 * list: row r owns entries [support_off[r], support_off[r+1]), each entry two int32 {validation-view slot, edge index in that view},
 * ascending slot; a slot without an entry is the reference's -2.  (A row is supported by ~10 % of the views: a quarter of the bytes.) */
typedef struct cvx_gathered_unit {
    int64_t meta[8];
    const int32_t* pair_h1; const int32_t* pair_h2rep; const int32_t* pair_count; const int32_t* nan_flag;
    const int64_t* support_off; const int32_t* support_entries;
    const double* gamma_world; const double* tangent_world; const double* gamma_cam;
    const double* h2_corrected_xy;     /* P x 2: the epipolar-corrected location of the view-2 representative (EdgeSketch_Core.cpp:636,651) */
} cvx_gathered_unit;
int cvx_gather_get(cvx_ctx* ctx, int round, int rank, cvx_gathered_unit* out);
/* Host copies of a gathered unit (any pointer may be NULL; support_off n_pairs + 1 entries, support_entries 2 * n_supports); synchronises. */
int cvx_gather_fetch(cvx_ctx* ctx, int round, int rank, int32_t* h1, int32_t* h2rep, int32_t* count, int32_t* nan_flag, int64_t* support_off,
                     int32_t* support_entries, double* gamma_world, double* tangent_world, double* gamma_cam, double* h2_corrected_xy);
/* Same copies, enqueued on the gather's side stream behind the transfers of `round` and not waited for: the round must have been
 * handed to NCCL (any later cvx_gather_post or cvx_gather_wait did that); pinned destinations then fill while the context computes
 * its next unit.  cvx_gather_wait(ctx, x, 1) completes them. */
int cvx_gather_fetch_async(cvx_ctx* ctx, int round, int rank, int32_t* h1, int32_t* h2rep, int32_t* count, int32_t* nan_flag, int64_t* support_off,
                           int32_t* support_entries, double* gamma_world, double* tangent_world, double* gamma_cam, double* h2_corrected_xy);
/* Bitwise OR of the claimed-edge bitmaps of all ranks into every rank (ncclAllReduce is not defined for OR: the words are gathered
 * and OR-ed), so that cvx_claims_get_counts / next-view selection see the claims of the whole pass when a pair was sharded by
 * view-1 range.  Collective. */
int cvx_claims_allreduce(cvx_ctx* ctx);
/* bytes moved and milliseconds the comm stream was busy since cvx_gather_reset (rank-local; ms needs the rounds complete) */
int cvx_gather_stats(cvx_ctx* ctx, int64_t* bytes_sent, int64_t* bytes_received, float* comm_ms);

/* ---- claimed edges: replaces the bookkeeping half of Calculate_Edge_Support_Ratios_And_Select_Next_Views
 * (EdgeSketch_Core.cpp:776-813).  The reference re-reads all edge files and walks EdgeMapping::edge_3D_to_supporting_edges
 * into one std::set<int> per view; here the context keeps one bit per resident edge.  cvx_claims_accumulate marks, for every
 * accepted pair of the last run that reaches the EdgeMapping insertions (EdgeSketch_Core.cpp:682-726): the view-1 edge, the
 * view-2 representative, its cluster members (hypo2_clusters lookup, hpp:193-202) and the supporting edge of every
 * validation view.  Only the rows the reference's paired_edge holds per view-1 edge (at most n_views, in its std::sort order,
 * EdgeSketch_Core.cpp:513-526) claim.  Bits persist across passes until cvx_claims_reset or a new cvx_set_view_edges of that view. */
int cvx_claims_reset(cvx_ctx* ctx);
int cvx_claims_accumulate(cvx_ctx* ctx);
/* claimed[v] = number of claimed edges of view v, total[v] = edges of view v (n_views entries each; either may be NULL) */
int cvx_claims_get_counts(cvx_ctx* ctx, int64_t* claimed, int64_t* total);
/* out[i] = 1 when edge i of `view` is claimed (edge count of the view entries) */
int cvx_claims_get_mask(cvx_ctx* ctx, int view, uint8_t* out);

/* Timing of the last phases on the context stream (CUDA events, milliseconds). */
typedef struct cvx_timings {
    float set_pair_ms; float run_ms; float finalize_ms;
    float form_ms;      /* candidate search + epipolar correction + clustering */
    float prep_ms;      /* unique-hypothesis table + wedge of the view-1 edges in every validation view */
    float validate_ms;  /* the validation kernel: reprojection + wedge + parallel test + support test */
    float select_ms;    /* vote + ratio test + compaction */
    int32_t kernel_launches; int32_t pad;
} cvx_timings;
int cvx_get_timings(cvx_ctx* ctx, cvx_timings* out);
int cvx_synchronize(cvx_ctx* ctx);
/* The cudaStream_t every call of this context is ordered on (lets the caller record its own CUDA events / order its own
 * copies against the context's work).  No reference counterpart: the reference is single-process host code. */
int cvx_get_stream(cvx_ctx* ctx, void** stream);

/* ---- stand-alone support test on pre-formed hypotheses (BASELINE.json config 5) -------------------------
 * replaces the inner body of LOOP 2/3: getHYPO2_idx_Ore x2 + set_intersection + getSupportIdx
 * (EdgeSketch_Core.cpp:348-384, getSupportedEdgels.cpp:32-47).  Hypotheses are grouped (one group = one
 * view-1 edge, sharing wedge31).  Per-view edge arrays are concatenated with offsets view_off[V+1].
 * wedge31: V x G x 2, wedge32 and tgt: V x H x 2 (view-major), support out: V x H. */
typedef struct cvx_support_problem cvx_support_problem;
int cvx_support_problem_create(cvx_ctx* ctx, int n_groups, const int64_t* group_off, int n_vali, const int64_t* view_off,
                               const double* ang31, const double* ang32, const double* cos_t, const double* sin_t,
                               const double* wedge31, const double* wedge32, const double* tgt, double cos_thresh,
                               double parallel_deg, cvx_support_problem** out);
int cvx_support_problem_run(cvx_support_problem* p, float* kernel_ms);
/* Counted re-run (untimed): wedge-intersection members (the n of SURVEY.md 8d "52 + 16 n" bytes per validation), grid entries
 * visited, supported and parallel-wedge validations, summed over the whole problem.  Any pointer may be NULL. */
int cvx_support_problem_stats(cvx_support_problem* p, int64_t* n_inliers, int64_t* n_visited, int64_t* n_supported,
                              int64_t* n_parallel);
int cvx_support_problem_result(cvx_support_problem* p, int32_t* support);
int cvx_support_problem_destroy(cvx_support_problem* p);

/* ---- 3-D edge graph: the data-parallel front half of EdgeMapping::consolidate_3D_edges -------------------------------
 * The reference keys its maps on the exact doubles of a 3-D point; across this ABI a 3-D edge is a NODE INDEX in [0, n_nodes) chosen
 * by the caller (3d_curvix_b200/host/edge_mapping.cpp numbers the distinct 3-D points in first-insertion order), a 2-D edge is a KEY
 * INDEX in [0, n_keys) (view offset + edge index), and a link is a pair of node indices.  Host stages that stay on the host (connectivity
 * graph, curve tracing, merging, writers: edge_mapping.cpp:639-1164) consume cvx_graph_get_neighbors + cvx_graph_align.
 *
 * cvx_graph_create        nodes: location (world) and tangents_3D_world of the FIRST support of every 3-D edge (what the link statistics use,
 *                         edge_mapping.cpp:259-260), n_nodes x 3 doubles each (host pointers; copied)
 * cvx_graph_build_links   build3DEdgeWeightedGraph (edge_mapping.cpp:107-192).  key_off / key_nodes: for every 2-D edge the 3-D edges it supports
 *                         (CSR, the records of map_Uncorrected2DEdge_To_SupportingData :36-62, duplicates kept); curvelet c names the 2-D edges
 *                         cv_self_key[c], cv_neighbor_key[c] (-1: none).  Every pair of records of a curvelet links its two 3-D edges unless the two
 *                         points share a coordinate (:182).  Result: the distinct links (the reference's weights are never read).
 * cvx_graph_prune         pruneEdgeGraph_by_3DProximityAndOrientation (:229-335: normal distance, point distance and tangent angle of a link below
 *                         mean + lambda x standard deviation over all links) then pruneEdgeGraphbyProjections (:337-405: the two points project within
 *                         proj_thresh pixels of each other in EVERY view; R n_views x 9 row-major, T n_views x 3, K 9), then the neighbour lists of
 *                         buildEdgeNodeGraph (:440-455).  stats6 = {mu_t, sigma_t, mu_n, sigma_n, mu_theta, sigma_theta} as the reference prints them.
 * cvx_graph_get_links     links of stage 0 (built), 1 (after the 3-D pruning), 2 (after the projection pruning); node_a < node_b, ascending
 * cvx_graph_get_neighbors CSR neighbour lists of stage 2 (offsets n_nodes + 1, neighbours 2 x links), ascending neighbour index
 * cvx_graph_align         align3DEdgesUsingEdgeNodes (:460-620): `iterations` synchronous updates of every node's location (Gaussian-weighted
 *                         pull onto its neighbours' tangent lines) and orientation (weighted Euler-angle alignment), from the node locations and
 *                         `orientation` (n_nodes x 3: normalised sum of the supports' tangents, :426-433); one kernel launch per iteration */
typedef struct cvx_graph cvx_graph;
int cvx_graph_create(cvx_ctx* ctx, int64_t n_nodes, const double* location, const double* first_tangent, cvx_graph** out);
int cvx_graph_build_links(cvx_graph* g, int64_t n_keys, const int64_t* key_off, const int32_t* key_nodes, int64_t n_curvelets,
                          const int32_t* cv_self_key, const int32_t* cv_neighbor_key, int64_t* n_links);
int cvx_graph_prune(cvx_graph* g, double lambda1, double lambda2, double lambda3, double proj_thresh, int n_views, const double* R,
                    const double* T, const double* K, double* stats6, int64_t* n_after_3d, int64_t* n_after_projection);
int cvx_graph_get_links(cvx_graph* g, int stage, int32_t* node_a, int32_t* node_b, int64_t* n);
/* per stage-0 link (order of cvx_graph_get_links(0)): position of its first occurrence in the curvelet-order stream of candidate pairs (the
 * insertion order of the reference's graph map, edge_mapping.cpp:183) and whether it survives the 3-D pruning / the projection test */
int cvx_graph_get_link_details(cvx_graph* g, int64_t* first_occurrence, uint8_t* keep_3d, uint8_t* keep_projection);
int cvx_graph_get_neighbors(cvx_graph* g, int64_t* offsets, int32_t* neighbors);
/* replaces the neighbour lists (a caller that wants the reference's neighbour ORDER, which follows its hash-map iteration, builds them itself) */
int cvx_graph_set_neighbors(cvx_graph* g, const int64_t* offsets, const int32_t* neighbors);
int cvx_graph_align(cvx_graph* g, const double* orientation, int iterations, double step_force, double step_torque, double* location_out,
                    double* orientation_out, float* kernel_ms);
int cvx_graph_destroy(cvx_graph* g);

/* cvx_graph_align_no_alloc -- NO COUNTERPART IN THE REFERENCE; added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same iterations as cvx_graph_align
 * (align3DEdgesUsingEdgeNodes, one k_align_step launch each), but every pointer is a DEVICE pointer owned by the caller and the one
 * intermediate buffer the iteration needs lives in a file-scope device array of fixed size inside the library:
 *   limit   n_nodes <= CVX_ALIGN_NO_ALLOC_MAX_NODES (65 536; 3 MiB per GPU that loads the library, held until exit); larger sizes are rejected
 *           with CVX_ERR_INVALID and a message in errbuf -- use cvx_graph_align there
 *   calls   no cudaMalloc / cudaFree / allocating library routine, no stream or event creation, no synchronisation: kernels (and two
 *           device-to-device copies when iterations == 0) are enqueued on `stream` (a cudaStream_t, NULL = the default stream) of `device`
 *   sharing every call on a device uses the same array: calls that may overlap on the GPU (different streams without an ordering) race.  The
 *           array keeps the previous call's contents; nothing in it has to start at zero (each iteration writes every node's row before the
 *           next reads it), so stale contents are never read
 *   inputs  d_offsets n_nodes + 1 (int64), d_neighbors offsets[n_nodes] (int32) as cvx_graph_get_neighbors / cvx_graph_set_neighbors;
 *           d_location, d_orientation n_nodes x 3 doubles; outputs d_location_out, d_orientation_out n_nodes x 3 doubles, not aliasing the inputs
 *   errbuf  optional, receives the error text (there is no context to hold it) */
#define CVX_ALIGN_NO_ALLOC_MAX_NODES 65536
int cvx_graph_align_no_alloc(int device, void* stream, int64_t n_nodes, const int64_t* d_offsets, const int32_t* d_neighbors,
                             const double* d_location, const double* d_orientation, int iterations, double step_force, double step_torque,
                             double* d_location_out, double* d_orientation_out, char* errbuf, int errbuf_len);

#ifdef __cplusplus
}
#endif
#endif /* CURVIX_B200_H */

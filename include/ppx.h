/*
 * ppx.h -- C ABI of libppx.so: the B200 (sm_100a) implementation of PathPlanning's
 * collision-checked expansion path.
 *
 * This is the drop-in boundary.  Every entry point replaces one in-process C++ virtual call of
 * the reference (cited as path:line under the reference tree) and is what the reference-side
 * host classes (pathplanning_b200/host/, same names as the reference's) bind to.  See
 * INTEGRATION.md for the binding a reference maintainer would add.
 *
 * Conventions
 *   - plain C, no exceptions; every function returns 0 on success and a negative ppx_status on
 *     error; ppx_last_error() gives the text.  There is NO CPU fallback: without a CUDA device
 *     ppx_ctx_create fails with PPX_ERR_CUDA.
 *   - one opaque context per GPU; a context is not thread safe; all work of a context is issued
 *     on one CUDA stream (its own, or the caller's via ppx_ctx_set_stream).
 *   - functions without suffix take HOST buffers and include the host<->device copies and a
 *     stream synchronisation; functions ending in _dev take DEVICE pointers, enqueue on the
 *     context stream and return without synchronising.
 *   - poses are 3 consecutive doubles (x, y, theta) and are used verbatim (no implicit wrap of
 *     theta) unless stated.
 *   - enums have the reference's numeric values: Steer {Left=0, Straight=1, Right=2},
 *     Direction {Forward=0, Backward=1, NoMotion=2}  (paths/path.h:9-19).
 */
#ifndef PPX_H
#define PPX_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PPX_VERSION 100

typedef enum ppx_status {
	PPX_OK = 0,
	PPX_ERR_INVALID = -1, /* bad argument */
	PPX_ERR_CUDA = -2, /* CUDA runtime / driver error, or no device */
	PPX_ERR_STATE = -3, /* call order (e.g. no map uploaded) */
	PPX_ERR_CAPACITY = -4, /* a device-side work buffer overflowed */
	PPX_ERR_NCCL = -5
} ppx_status;

typedef struct ppx_ctx ppx_ctx;

/* ------------------------------------------------------------------------------------------ */
/* Context                                                                                     */
int ppx_version(void);
int ppx_ctx_create(int device, ppx_ctx** out);
int ppx_ctx_destroy(ppx_ctx* ctx);
/* Use the caller's cudaStream_t (NULL = back to the context's own stream). */
int ppx_ctx_set_stream(ppx_ctx* ctx, void* cuda_stream);
int ppx_ctx_synchronize(ppx_ctx* ctx);
const char* ppx_last_error(const ppx_ctx* ctx);
/* Number of kernels this context launched since creation (bench.py's gpu_launches). */
int64_t ppx_ctx_launch_count(const ppx_ctx* ctx);
/* Raw device memory helpers so that non-CUDA hosts (ctypes, cgo, JNI) can use the _dev entry
 * points without another CUDA binding. */
int ppx_dev_alloc(ppx_ctx* ctx, uint64_t bytes, void** dptr);
int ppx_dev_free(ppx_ctx* ctx, void* dptr);
int ppx_memcpy_h2d(ppx_ctx* ctx, void* dst, const void* src, uint64_t bytes);
int ppx_memcpy_d2h(ppx_ctx* ctx, void* dst, const void* src, uint64_t bytes);
/* Page-locked host buffers: the host-buffer entry points accept any host pointer, but copies from
 * pinned memory run at full PCIe rate. */
int ppx_host_alloc(ppx_ctx* ctx, uint64_t bytes, void** hptr);
int ppx_host_free(ppx_ctx* ctx, void* hptr);

/* ------------------------------------------------------------------------------------------ */
/* World model (read-only inputs of every kernel)                                              */
/*
 * Mirrors what StateValidatorOccupancyMap reads:
 *   OccupancyMap frames  state_validator/occupancy_map.h:101-117,175-193, occupancy_map.cpp:6-14
 *   StateSpaceSE2 bounds state_space/state_space_se2.cpp:15-25
 *   Grid<T> row-major    utils/grid.h:87   (row <-> x, col <-> y)
 */
typedef struct ppx_map_desc {
	int32_t rows, cols;
	float resolution; /* OccupancyMap::resolution */
	float _pad;
	double world_origin[2]; /* OccupancyMap::m_worldGridOrigin */
	double local_origin[2]; /* OccupancyMap::m_localOrigin (world -> local translation) */
	double lb[3], ub[3]; /* StateSpaceSE2::bounds (x, y, theta) */
} ppx_map_desc;

/*
 * Upload (or refresh after OccupancyMap::Update / GVD::Update) the grids.  Any pointer may be
 * NULL to keep the previous device copy.
 *   sqdist    GVD::ObstacleDistanceMap::m_distance, squared cell distance, int32 (gvd.h:38-39)
 *   occupancy OccupancyMap::m_occupancyMatrix, >=0 means occupied (occupancy_map.h:129-130,
 *             obstacle_list_occupancy_map.cpp:63-69)
 *   pathcost  GVD::PathCostMap, Voronoi potential, float (gvd.h:102-111,146)
 */
int ppx_map_upload(ppx_ctx* ctx, const ppx_map_desc* desc, const int32_t* sqdist, const int32_t* occupancy,
	const float* pathcost);

/* StateValidatorOccupancyMap public fields (state_validator_occupancy_map.h:27-28). */
typedef struct ppx_validator_params {
	float min_safe_radius;
	float min_path_interpolation_distance;
} ppx_validator_params;

/* ------------------------------------------------------------------------------------------ */
/* Edge descriptors                                                                            */
/*
 * A pre-solved Reeds-Shepp edge = PathReedsShepp{init, PathSegment, minTurningRadius}
 * (paths/path_reeds_shepp.h:8-37, geometry/reeds_shepp.h:51-104): 72 bytes.
 *   code: 4 bits per motion i at bits [4i, 4i+4): steer in bits 0-1, direction in bits 2-3.
 *   len:  normalised motion lengths (unit turning radius); direction==NoMotion or len==+inf
 *         terminates the segment (reeds_shepp.cpp:305-308).
 */
typedef struct ppx_rs_edge {
	double start[3];
	double len[5];
	uint64_t code;
} ppx_rs_edge;

#define PPX_RS_CODE(steer, dir, i) ((uint64_t)(((steer) & 3) | (((dir) & 3) << 2)) << (4 * (i)))

/* A constant-steer primitive = PathConstantSteer (paths/path_constant_steer.cpp:5-25) over
 * KinematicBicycleModel::ConstantSteer (models/kinematic_bicycle_model.cpp:5-32).  The three
 * model constants are evaluated on the host by the KinematicBicycleModel mirror so that they
 * are the same bits the reference computes:
 *   beta = atan(rearToCenter*tan(steering)/wheelbase), cos_beta = cos(beta),
 *   dtheta_ddist = cos_beta*tan(steering)/wheelbase. */
typedef struct ppx_csteer_edge {
	double start[3];
	double length; /* Path::m_length */
	double beta, cos_beta, dtheta_ddist;
	int32_t direction; /* Direction */
	int32_t _pad;
} ppx_csteer_edge;

/* A straight SE(2) segment = PathSE2 (paths/path_se2.cpp:5-22). */
typedef struct ppx_line_edge {
	double from[3];
	double to[3];
} ppx_line_edge;

/* ------------------------------------------------------------------------------------------ */
/* A7: StateValidatorOccupancyMap::IsStateValid / IsPathValid, batched                         */
/*     state_validator/state_validator_occupancy_map.cpp:15-26, 28-71                          */
/*
 * valid[i]   the bool IsPathValid returns
 * last[i]    the float written through `last` (ratio of the last valid sample; 1.0f if valid)
 * samples[i] optional (may be NULL): number of IsStateValid evaluations the adaptive stepping
 *            made -- the S of the 80+4S algorithmic bytes per edge.
 */
int ppx_is_state_valid(ppx_ctx* ctx, int64_t n, const double* poses, const ppx_validator_params* vp, uint8_t* valid);

int ppx_check_edges_rs(ppx_ctx* ctx, int64_t n, const ppx_rs_edge* edges, double min_turning_radius,
	const ppx_validator_params* vp, uint8_t* valid, float* last, uint32_t* samples);
int ppx_check_edges_rs_dev(ppx_ctx* ctx, int64_t n, const ppx_rs_edge* d_edges, double min_turning_radius,
	const ppx_validator_params* vp, uint8_t* d_valid, float* d_last, uint32_t* d_samples);

int ppx_check_edges_csteer(ppx_ctx* ctx, int64_t n, const ppx_csteer_edge* edges, const ppx_validator_params* vp,
	uint8_t* valid, float* last, double* final_pose /* 3n, after Truncate(last) when invalid; may be NULL */);
int ppx_check_edges_line(ppx_ctx* ctx, int64_t n, const ppx_line_edge* edges, const ppx_validator_params* vp,
	uint8_t* valid, float* last);

/* ------------------------------------------------------------------------------------------ */
/* A3: ReedsShepp::Solver::GetOptimalPath, batched  (geometry/reeds_shepp.cpp:654-683)         */
typedef struct ppx_rs_cost_params {
	double min_turning_radius;
	float reverse_cost_multiplier;
	float forward_cost_multiplier;
	float direction_switching_cost;
	int32_t shortest; /* 0: GetOptimalPath (float cost, :654-683); 1: GetShortestPath (double length, :608-652) */
} ppx_rs_cost_params;

/*
 * starts/goals: 3n doubles each, used verbatim (the Pose2d constructor's theta wrap,
 * geometry/2dplane.h:19-22, is the caller's job: Hybrid A* passes un-wrapped headings, trap H11).
 * Outputs (any may be NULL): edges[i] = solved segment with start (ready for
 * ppx_check_edges_rs), word[i] = PathWords value (-1 NoPath), length[i] = GetLength(r),
 * cost[i] = PathSegment::ComputeCost (float).
 */
int ppx_rs_solve(ppx_ctx* ctx, int64_t n, const double* starts, const double* goals, const ppx_rs_cost_params* cp,
	ppx_rs_edge* edges, int32_t* word, double* length, float* cost);
int ppx_rs_solve_dev(ppx_ctx* ctx, int64_t n, const double* d_starts, const double* d_goals,
	const ppx_rs_cost_params* cp, ppx_rs_edge* d_edges, int32_t* d_word, double* d_length, float* d_cost);

/* Fused PathConnectionReedsShepp::Connect + IsPathValid (paths/path_reeds_shepp.cpp:174-179 then
 * state_validator_occupancy_map.cpp:28-71): the "solve+check" edge of BASELINE config C3 and of
 * GetReedsSheppChild (algo/hybrid_a_star.cpp:149-160). */
int ppx_connect_check_rs(ppx_ctx* ctx, int64_t n, const double* starts, const double* goals,
	const ppx_rs_cost_params* cp, const ppx_validator_params* vp, uint8_t* valid, float* last, int32_t* word,
	double* length, float* cost);
int ppx_connect_check_rs_dev(ppx_ctx* ctx, int64_t n, const double* d_starts, const double* d_goals,
	const ppx_rs_cost_params* cp, const ppx_validator_params* vp, uint8_t* d_valid, float* d_last, int32_t* d_word,
	double* d_length, float* d_cost);

/* Path<Pose2d>::Interpolate for RS edges (paths/path_reeds_shepp.cpp:12-47): ratios holds
 * n*nratios values (nratios per edge, non-decreasing within an edge is NOT required), poses out =
 * 3*n*nratios doubles. */
int ppx_rs_interpolate(ppx_ctx* ctx, int64_t n, const ppx_rs_edge* edges, double min_turning_radius, int32_t nratios,
	const double* ratios, double* poses);

/* ------------------------------------------------------------------------------------------ */
/* A8: ObstaclesHeuristic::Update, batched over goals  (algo/heuristics.cpp:106-153)           */
/*
 * For each goal pose (3 doubles, only x,y are used) runs the holonomic-with-obstacles wavefront
 * over the uploaded occupancy grid with the reference's exact semantics (first reach wins, pops
 * in ascending float cost with LIFO ties, utils/frontier.h:39-48,83-91).
 *   cost      n * rows * cols floats: ObstaclesHeuristic::m_cost (+inf where never reached)
 *   explored  optional n * rows * cols bytes: ObstaclesHeuristic::m_explored
 * A goal outside the map leaves its grid all +inf (heuristics.cpp:115-117).
 */
int ppx_wavefront(ppx_ctx* ctx, int64_t n, const double* goals, float* cost, uint8_t* explored);
int ppx_wavefront_dev(ppx_ctx* ctx, int64_t n, const double* d_goals, float* d_cost, uint8_t* d_explored);

#ifdef __cplusplus
}
#endif
#endif /* PPX_H */

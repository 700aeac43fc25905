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
/* Options: "rs_filter" (default 1; 0 = the Reeds-Shepp edge checker evaluates every sample with the
 * exact FP64 restatement, for A/B tests of the FP32 cell filter).    "rs_overlap" (default 0; 1 = consecutive ppx_check_edges_rs_dev calls on this context may overlap: a call's kernel
 * starts filling the SMs while the last long sample chains of the previous call drain (programmatic dependent
 * launch; up to three calls are resident on the GPU at once).  The caller guarantees that the output arrays of a call do
 * not alias those of the previous THREE calls and that its inputs are not produced by them; any other work submitted to
 * the stream still waits for every earlier call to complete.)
 * "rs_engine" (default 1: k_check_rs2, knot records + bulk-copy staging; 0: the round-1 lane machine k_check_rs; 2: as 1, and
 * ppx_rs_solve uses the quad-cooperative solver) and "search_engine" (default 1: one CTA per Hybrid A* query, master + worker
 * warps; 0: one warp per query) are A/B switches: the results are identical (tests/test_edges_gpu.py::test_rs_engines_identical,
 * tests/test_hybrid_gpu.py::test_search_engines_identical). */
int ppx_ctx_set_option(ppx_ctx* ctx, const char* name, int64_t value);
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
 * Upload (or refresh after OccupancyMap::Update / GVD::Update) the grids.  sqdist / occupancy may be
 * NULL to keep the previous device copy.  pathcost == NULL keeps the previous potential only when
 * neither of the other grids changes in this call; with a new sqdist or occupancy the potential is
 * reset to zero (it belongs to the map it was computed for).
 *   sqdist    GVD::ObstacleDistanceMap::m_distance, squared cell distance, int32 (gvd.h:38-39)
 *   occupancy OccupancyMap::m_occupancyMatrix, >=0 means occupied (occupancy_map.h:129-130,
 *             obstacle_list_occupancy_map.cpp:63-69)
 *   pathcost  GVD::PathCostMap, Voronoi potential, float (gvd.h:102-111,146)
 */
int ppx_map_upload(ppx_ctx* ctx, const ppx_map_desc* desc, const int32_t* sqdist, const int32_t* occupancy,
	const float* pathcost);

/* N1: cold build of GVD::ObstacleDistanceMap (state_validator/gvd.cpp:30-89,105-118) on the device:
 * uploads the occupancy grid (>= 0 occupied) and computes the squared cell distance to the nearest
 * occupied cell (exact Euclidean distance transform, INT_MAX on a map without obstacles), then
 * proceeds like ppx_map_upload.  sqdist_out (rows*cols, may be NULL) receives the distance grid. */
int ppx_map_build(ppx_ctx* ctx, const ppx_map_desc* desc, const int32_t* occupancy, const float* pathcost,
	int32_t* sqdist_out);

/* N1, all of it: the GVD of the occupancy grid on the device -- GVD::ObstacleDistanceMap (m_distance, m_obstacle;
 * state_validator/gvd.cpp:30-89), the Voronoi-edge set CheckVoro leaves behind (:105-131), GVD::VoronoiDistanceMap
 * (m_distance, m_edge; :200-264) and GVD::PathCostMap (:266-283, alpha 20, dMax 30 as gvd.h:180) -- then proceeds like
 * ppx_map_upload with the grids it built (the context's sqdist, occupancy and pathcost).  Incremental
 * SetObstacle / UnsetObstacle (:74-89) are a rebuild here: ppx_map_rasterize the change, call this again (3 ms on
 * 2000 x 2000).
 *   flags  PPX_GVD_VORONOI: Voronoi edges as the reference marks them when the GVD (and with it the Voronoi map) exists
 *          while the obstacle brushfire runs -- `GVD gvd(map); gvd.Update();` before or instead of `map.Update()`
 *          (interfaces/python/scripts/example.py:47-48).  Without the flag: the Voronoi map the reference has when
 *          OccupancyMap::Update() ran BEFORE the GVD was created (HybridAStar::Initialize after map->Update()): no edge
 *          cell, Voronoi distance INT_MAX everywhere, and a potential that is a function of the obstacle distance only.
 * Outputs (host, each may be NULL): sqdist / voro_sqdist rows*cols int32, pathcost rows*cols float,
 * nearest_obstacle / nearest_edge rows*cols*2 int32 (row, col) pairs, (-1, -1) where there is none.
 * Parity: the reference computes the two distance maps by a priority-queue brushfire through
 * 8-neighbours, whose result depends on how std::priority_queue orders equal keys; it equals the exact transform used
 * here on axis-aligned maps (C2: every grid identical) and differs in a few cells per thousand on arbitrarily rotated
 * obstacles, where the brushfire itself is not exact.  The Voronoi rule is order-free: applied to the reference's own
 * final grids it reproduces the reference's edge set cell for cell.  ppx_map_upload with reference-built grids stays
 * the strict-parity path. */
#define PPX_GVD_VORONOI 1
int ppx_gvd_build(ppx_ctx* ctx, const ppx_map_desc* desc, const int32_t* occupancy, int32_t flags, int32_t* sqdist_out,
	int32_t* voro_sqdist_out, float* pathcost_out, int32_t* nearest_obstacle_out, int32_t* nearest_edge_out);

/*
 * N4: outline rasterisation of polygon obstacles into the occupancy grid -- Shape::RasterizeLine
 * (state_validator/obstacle.cpp:7-61) over PolygonShape::GetGridCellsPosition (:77-85) with the occupancy write
 * of ObstacleListOccupancyMap::AddObstacle (obstacle_list_occupancy_map.cpp:29-45), for n_polygons obstacles
 * added in order.  vertices: world coordinates (x, y) of every outline vertex as
 * Obstacle::GetBoundaryWorldPosition returns them (the rotation by the pose stays on the host: its libm decides the
 * cell of a vertex on a cell border); offsets[p] .. offsets[p+1] are polygon p's vertices.  occupancy (host, rows *
 * cols, in/out): cells of polygon p receive id first_id + p, a later polygon overwrites an earlier one like the
 * reference's sequential adds; first_id must exceed every id already in the grid.  first_id = -1 removes instead:
 * the outline cells of the given polygons are set free whoever owns them, as ObstacleListOccupancyMap::RemoveObstacle
 * does (obstacle_list_occupancy_map.cpp:47-61).  Follow with ppx_map_build.
 */
int ppx_map_rasterize(ppx_ctx* ctx, const ppx_map_desc* desc, int32_t n_polygons, const int32_t* offsets,
	const double* vertices, int32_t first_id, int32_t* occupancy);

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

/* Rectangular car footprint (BASELINE config 3 "rectangular car footprint").  NOT a reference behaviour: the reference
 * checks one disc of radius minSafeRadius at the pose (two `// TODO add offset`, state_validator_occupancy_map.cpp:23,60).
 * Defined here: the length x width rectangle, its centre rear_to_center ahead of the pose along the heading, is covered by
 * n_discs discs of one radius rho = sqrt((length / n_discs)^2 + width^2) / 2 centred on the slices' centres; a pose is
 * valid iff every disc centre passes IsStateValid with minSafeRadius = rho.  Along a path: the reference's sphere tracing
 * (state_validator_occupancy_map.cpp:42-66) with step = max(min(slack, border) / gamma, min_step), slack = the smallest
 * (clearance - rho) over the discs, border = the smallest distance of a disc centre to the bounds,
 * gamma = sqrt(1 + (o_max / min_turning_radius)^2), o_max the largest disc offset (a disc centre travels up to gamma times
 * as far as the pose).  Its checker is a restatement over the reference's own Interpolate and distance map
 * (oracle/ref_harness.cpp, kind "port", parity unpinned by definition); results are reported apart from the disc check. */
typedef struct ppx_footprint_params {
	double length, width; /* e.g. 4.5 x 2.0 m */
	double rear_to_center; /* offset of the rectangle's centre from the pose along the heading (m) */
	int32_t n_discs; /* 1..16 */
	float min_step; /* minPathInterpolationDistance */
} ppx_footprint_params;

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

/* Diagnostic twin of ppx_check_edges_rs: same results, and for every sample the FP32 cell filter
 * certified it also evaluates the exact FP64 position.  counters[0..2] = certified samples, samples
 * sent to the exact path, certified samples whose exact cell differs (must be 0);
 * *max_dev_cells = max |filtered - exact| position over certified samples, in cells. */
int ppx_check_edges_rs_audit(ppx_ctx* ctx, int64_t n, const ppx_rs_edge* edges, double min_turning_radius,
	const ppx_validator_params* vp, uint8_t* valid, float* last, uint32_t* samples, uint64_t* counters,
	float* max_dev_cells);

/* The rectangular-footprint variant of ppx_is_state_valid / ppx_check_edges_rs[_dev] (see ppx_footprint_params). */
int ppx_is_state_valid_rect(ppx_ctx* ctx, int64_t n, const double* poses, const ppx_footprint_params* fp, uint8_t* valid);
int ppx_check_edges_rs_rect(ppx_ctx* ctx, int64_t n, const ppx_rs_edge* edges, double min_turning_radius,
	const ppx_footprint_params* fp, uint8_t* valid, float* last, uint32_t* samples);
int ppx_check_edges_rs_rect_dev(ppx_ctx* ctx, int64_t n, const ppx_rs_edge* d_edges, double min_turning_radius,
	const ppx_footprint_params* fp, uint8_t* d_valid, float* d_last, uint32_t* d_samples);
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
/*
 * ppx_wavefront_dev with NO allocation at call time.  It has no counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same kernel, same results bit for bit; the kernel's scratch (per goal in flight: a tag word per
 * cell + the four-slot bucket ring) lives in a file-scope `static __device__` array of PPX_WAVEFRONT_STATIC_BYTES, which
 * every GPU that loads the library holds until the process exits.  The call performs no cudaMalloc / cudaFree and calls no
 * allocating library routine; it only enqueues on the context's stream (two 8-byte memsets and the kernel).
 *   limit     the scratch of at least one goal must fit: 4 bytes per cell + 48 bytes x (the power of two >= 8 (rows + cols)).
 *             Any map of up to PPX_WAVEFRONT_STATIC_MAX_CELLS cells with rows + cols <= 65536 does (C3's 2000 x 2000
 *             included); a larger map may not, and is then rejected with PPX_ERR_CAPACITY and a message that states
 *             the bytes needed.  n is not limited: the goals in flight at once are as many as fit (C2, 400 x 400: 64),
 *             the rest queue behind them, so big batches are slower than through ppx_wavefront_dev.
 *   d_status  optional caller-owned device word: 0 ok, 1 a frontier bucket overflowed its ring slot, 2 internal
 *             (what ppx_wavefront turns into PPX_ERR_CAPACITY); NULL keeps it in the static array.
 *   set-up    the first call on a device resolves the array's device address (cudaGetSymbolAddress, once per device, may
 *             load the module): make one call outside any graph capture.
 *   sharing   every call on a device shares the one array, which keeps the previous call's contents (the call clears what
 *             must start at zero).  Calls that overlap on the GPU -- two contexts or two streams on one device -- race:
 *             serialise them.  Calls on one context's stream are ordered and safe.
 */
#define PPX_WAVEFRONT_STATIC_BYTES (64 << 20)
#define PPX_WAVEFRONT_STATIC_MAX_CELLS (4 << 20)
int ppx_wavefront_dev_static(ppx_ctx* ctx, int64_t n, const double* d_goals, float* d_cost, uint8_t* d_explored, int32_t* d_status);

/* ------------------------------------------------------------------------------------------ */
/* A1-A6, A9: Hybrid A* successor generation, heuristics and the batched on-device graph search  */
/*
 * HybridAStar::SearchParameters (algo/hybrid_a_star.h:29-50).  The 8-argument constructor of the
 * reference cannot set wheelbase (it stays 2.6); the field is here so the host mirror passes
 * whatever its SearchParameters holds.
 */
typedef struct ppx_hybrid_params {
	double wheelbase;
	double min_turning_radius;
	double direction_switching_cost;
	double reverse_cost_multiplier;
	double forward_cost_multiplier;
	double voronoi_cost_multiplier;
	uint32_t num_generated_motion;
	uint32_t _pad;
	double spatial_resolution;
	double angular_resolution;
} ppx_hybrid_params;

typedef struct ppx_hybrid ppx_hybrid;

/*
 * HybridAStar::Initialize (algo/hybrid_a_star.cpp:206-235) for the map currently uploaded to ctx:
 * derives the steering set (hybrid_a_star.cpp:13-29), builds the NonHolonomicHeuristic table on the
 * device (algo/heuristics.cpp:36-76, one Reeds-Shepp solve per entry) and keeps the validator
 * parameters.  The planner reads the context's map at search time, so ppx_map_upload may refresh the
 * grids between searches (same size and bounds).
 */
int ppx_hybrid_create(ppx_ctx* ctx, const ppx_hybrid_params* params, const ppx_validator_params* vp, ppx_hybrid** out);
int ppx_hybrid_destroy(ppx_hybrid* h);
/* NonHolonomicHeuristic: dims = {numSpatialX, numSpatialY, numAngular}, offsets = {offsetX, offsetY},
 * values (may be NULL) = m_values[i][j][k] row-major. */
int ppx_hybrid_table(ppx_hybrid* h, int32_t* dims, double* offsets, double* values);
/*
 * AStarCombinedHeuristic / NonHolonomicHeuristic / ObstaclesHeuristic ::GetHeuristicValue
 * (algo/a_star.h:102-109, algo/heuristics.cpp:78-95,155-165) for n poses and one goal; runs the
 * wavefront of that goal first (ObstaclesHeuristic::Update).  Any output may be NULL.
 * Reads of the table with a negative heading index (out of bounds in the reference)
 * are defined as index + numAngular.
 */
int ppx_hybrid_heuristic(ppx_hybrid* h, const double* goal, int64_t n, const double* poses, double* combined,
	double* nonholo, double* obstacles);

#define PPX_MAX_CHILDREN 19
typedef struct ppx_hybrid_child {
	double pose[3]; /* State::continuousPose */
	double cost; /* transition cost */
	double heuristic; /* combined heuristic of the child */
	double length; /* action path length (after truncation) */
	int32_t key[3]; /* State::discretePose (x, y, theta bin) */
	int32_t direction; /* State::direction */
	int32_t kind; /* 0 constant steer, 1 Reeds-Shepp */
	float last; /* lastValidRatio of the primitive (1 when fully valid) */
} ppx_hybrid_child;

/*
 * StatePropagator::GetNeighborStates (algo/hybrid_a_star.cpp:59-91) for n arbitrary parents and one
 * goal: children[i*PPX_MAX_CHILDREN + c], c < n_children[i], in the reference's order.  seeds[i] seeds
 * the std::mt19937_64 that decides the Reeds-Shepp shot of parent i (hybrid_a_star.cpp:82, trap H4).
 */
int ppx_hybrid_expand(ppx_hybrid* h, const double* goal, int64_t n, const double* parents, const uint64_t* seeds,
	int32_t* n_children, ppx_hybrid_child* children, double* parent_heuristic);

#define PPX_ERR_LIMIT (-6) /* search stopped at ppx_hybrid_limits.max_expansions */

typedef struct ppx_hybrid_limits {
	int32_t max_nodes; /* node pool per query; 0 = 2 x number of discrete states inside the bounds */
	int32_t max_expansions; /* 0 = unlimited (the reference has no limit) */
	int32_t max_path; /* capacity of the per-query path output (states) */
	int32_t max_log; /* capacity of the per-query expansion log (entries); 0 = no log */
} ppx_hybrid_limits;

typedef struct ppx_hybrid_result {
	int32_t status; /* 0 Success, -1 Failure (frontier exhausted) like Planner::Status; PPX_ERR_CAPACITY; PPX_ERR_LIMIT */
	int32_t n_expanded; /* nodes popped and expanded */
	int32_t n_nodes; /* nodes created (pushed to the frontier) */
	int32_t n_children; /* children generated */
	int32_t n_rs_attempts; /* Reeds-Shepp shots attempted */
	int32_t path_len; /* states on the solution chain root..goal (0 without solution) */
	double cost; /* GraphSearch::GetOptimalCost (+inf without solution) */
	uint64_t expanded_hash; /* order-sensitive FNV hash of the expanded discrete states */
} ppx_hybrid_result;

/* One state of the solution chain (AStar::GetPath / GetActions): the action that reaches it is
 * reconstructed by the host mirror as PathConstantSteer(parent pose, delta[slot/2], dir) truncated to
 * `last`, or as the Reeds-Shepp connection parent -> goal. */
typedef struct ppx_hybrid_path_state {
	double pose[3];
	double length;
	float last;
	int8_t kind; /* -1 root, 0 constant steer, 1 Reeds-Shepp */
	int8_t direction;
	int8_t slot; /* child slot in GetNeighborStates order */
	int8_t _pad;
} ppx_hybrid_path_state;

typedef struct ppx_hybrid_log_entry {
	int32_t key[3];
	int32_t direction;
	double pose[3];
} ppx_hybrid_log_entry;

/*
 * The graph-search part of HybridAStar::SearchPath (algo/hybrid_a_star.cpp:248-257) for nq
 * independent queries: ObstaclesHeuristic::Update(goal) (one wavefront per distinct goal cell), then
 * AStar::SearchPath (algo/a_star.h:326-346) with the Frontier's tie order, entirely on the device,
 * one warp per query.  seeds[q] seeds query q's private std::mt19937_64 (trap H4).
 *   results  nq records
 *   paths    optional, nq * limits->max_path states
 *   log      optional, nq * limits->max_log entries: the expanded nodes in expansion order
 */
int ppx_hybrid_search(ppx_hybrid* h, int64_t nq, const double* starts, const double* goals, const uint64_t* seeds,
	const ppx_hybrid_limits* limits, ppx_hybrid_result* results, ppx_hybrid_path_state* paths, ppx_hybrid_log_entry* log);
/* Device-resident variant: no host copies, no synchronisation, one wavefront per query. */
int ppx_hybrid_search_dev(ppx_hybrid* h, int64_t nq, const double* d_starts, const double* d_goals,
	const uint64_t* d_seeds, const ppx_hybrid_limits* limits, ppx_hybrid_result* d_results,
	ppx_hybrid_path_state* d_paths, ppx_hybrid_log_entry* d_log);

/* Device time of the two phases of the last ppx_hybrid_search[_dev] call on this handle -- the wavefronts
 * (ObstaclesHeuristic::Update per distinct goal cell) and the graph searches -- in milliseconds, from CUDA events
 * on the context's stream.  Waits for that call to finish.  PPX_ERR_STATE when the call was split into chunks. */
int ppx_hybrid_timing(ppx_hybrid* h, double* wavefront_ms, double* search_ms);

/*
 * N3: the "Process graph search path" step of HybridAStar::SearchPath (algo/hybrid_a_star.cpp:260-291) for
 * the solution chains a search returned: the composite path (paths/path_composite.h:43-100) is resampled
 * every `path_interpolation` (HybridAStar::pathInterpolation, 0.1f) with its cusp points snapped in
 * (path_composite.cpp:4-27) -- the pose array the smoother takes.  goals/results/paths are the arrays of
 * ppx_hybrid_search (paths with capacity max_path per query).  Outputs per query: poses[q*max_points*3 ..],
 * cusp[q*max_points ..] (1 = cuspIndices member), count[q] = number of samples (0 without a solution, -1 when
 * max_path cut the chain or max_points does not suffice), lengths[q] (may be NULL) = composite path length.
 */
int ppx_hybrid_resample(ppx_hybrid* h, int64_t nq, const double* goals, const ppx_hybrid_result* results,
	const ppx_hybrid_path_state* paths, int32_t max_path, float path_interpolation, int32_t max_points, double* poses,
	uint8_t* cusp, int32_t* count, double* lengths);

/* ------------------------------------------------------------------------------------------ */
/* A10: Tree::GetNearestNode / GetNearestNodes  (utils/tree.h:73-116) as exact brute-force k-NN  */
/*
 * Replaces the flann::Index<L2_Simple<double>> kd-tree (utils/tree.h:44,174): squared Euclidean
 * distance over the raw state coordinates (dim = 2 for Point2d, 3 for Pose2d: x, y, theta), exact,
 * neighbours by increasing distance, equal distances by increasing point index.
 *   indices  nqueries * k point indices (-1 where fewer than k points exist)
 *   dist2    optional nqueries * k squared distances
 */
int ppx_knn(ppx_ctx* ctx, int32_t dim, int64_t npoints, const double* points, int64_t nqueries, const double* queries,
	int32_t k, int32_t* indices, double* dist2);
int ppx_knn_dev(ppx_ctx* ctx, int32_t dim, int64_t npoints, const double* d_points, int64_t nqueries,
	const double* d_queries, int32_t k, int32_t* d_indices, double* d_dist2);

/* ------------------------------------------------------------------------------------------ */
/* A11: RRT / RRT* (algo/rrt.h:55-95,125-140, algo/rrt_star.h:53-112,136-160), batched over trees  */
/*
 * One tree per CTA, the reference's loop run verbatim: goal-biased sample from the tree's private
 * std::mt19937_64 (state_space_se2.cpp:27-38, utils/random.h, trap H4) -> nearest node (exact
 * brute-force scan, ties to the lower index) -> SteerTowards (Connect + Truncate to
 * max_connection_distance) -> IsPathValid -> [RRT*: k = max(1, (unsigned)log(size)) nearest nodes,
 * SteerExactly + cost + IsPathValid per candidate, best parent; NO rewiring, trap H12] -> Tree::Extend
 * with its duplicate-state rule (utils/tree.h:124-146) -> goal test.
 */
typedef struct ppx_rrt_params {
	int32_t dim; /* 2: Point2d states (RRTR2 / RRTStarR2); 3: Pose2d states */
	int32_t connection; /* 0 PathConnectionR2 / PathConnectionSE2 (straight), 1 PathConnectionReedsShepp (dim 3) */
	int32_t star; /* 0 RRT (goal test |state - goal| < 1, dim 2 only), 1 RRT* (exact-equality goal test) */
	int32_t validator; /* 0 StateValidatorFree (every path valid), 1 StateValidatorOccupancyMap on the uploaded map (dim 3) */
	uint32_t max_iteration; /* RRTParameters::maxIteration */
	uint32_t max_nodes; /* RRTParameters::maxNumberTreeNode */
	double max_connection_distance;
	double goal_bias;
	double min_turning_radius; /* PathConnectionReedsShepp(minTurningRadius), unit cost multipliers */
	double lb[3], ub[3]; /* StateSpace bounds used by SampleUniform */
} ppx_rrt_params;

typedef struct ppx_rrt_result {
	int32_t status; /* 0 Success, -1 Failure (iteration or node limit), PPX_ERR_CAPACITY */
	int32_t iterations; /* loop turns taken (the reference's count) */
	int32_t n_nodes; /* Tree::Size() */
	int32_t solution; /* node index of the solution, -1 */
	int32_t path_len; /* nodes on the chain root..solution */
	int32_t n_edges_checked; /* IsPathValid calls */
	double solution_cost; /* RRT*: costFromGoal of the solution node */
} ppx_rrt_result;

/*
 * starts / goals: ntrees * 3 doubles (x, y, theta; theta ignored for dim 2); seeds: ntrees.
 * Optional outputs, each with node_capacity entries per tree (node_capacity >= max_nodes + 2 to
 * receive whole trees): nodes (3 doubles per node, insertion order = flann point order),
 * parents (int32, -1 for the root), costs (RRT*: NodeInfo::costFromGoal).
 */
int ppx_rrt_run(ppx_ctx* ctx, const ppx_rrt_params* params, const ppx_validator_params* vp, int64_t ntrees,
	const double* starts, const double* goals, const uint64_t* seeds, ppx_rrt_result* results, int64_t node_capacity,
	double* nodes, int32_t* parents, double* costs);

/* ---------------------------------------------------------------------------------------------
 * N2: Smoother (algo/smoother.h:17-92, algo/smoother.cpp:33-224) for a batch of pose arrays.
 *
 * Replaces Smoother::Initialize + Smoother::Smooth(path, cuspIndices) + GetPath(), the step that follows the
 * graph search in HybridAStar::SearchPath (hybrid_a_star.cpp:293-303).  The gradient descent runs on the
 * device, one CTA per path; the two GVD lookups it makes (GVD::GetNearestObstacleCell /
 * GetNearestVoronoiEdgeCell, gvd.h:36,76 -- smoother.cpp:110,118) read cell grids the caller uploads with
 * ppx_smoother_set_gvd after GVD::Update(), or takes from the context's device-built GVD (ppx_gvd_build) with
 * ppx_smoother_use_gvd.
 */
typedef struct ppx_smoother ppx_smoother;

typedef struct ppx_smoother_params { /* Smoother::Parameters (smoother.h:29-62), same defaults on the reference side */
	float step_tolerance; /* 1e-3 */
	int32_t max_iterations; /* 2000 */
	float learning_rate; /* 0.01 */
	float path_weight; /* 0 */
	float smooth_weight; /* 0.4 */
	float voronoi_weight; /* 0.02 */
	float collision_weight; /* 0.2 */
	float curvature_weight; /* 0.4 */
	float collision_ratio; /* 0.2 */
	float max_curvature; /* Initialize(.., maxCurvature): 1 / minTurningRadius (hybrid_a_star.cpp:218) */
	float alpha, d_max; /* GVD::alpha = 20, GVD::dMax = 30 (gvd.h:181) */
	float min_safe_radius; /* StateValidatorOccupancyMap::minSafeRadius */
} ppx_smoother_params;

int ppx_smoother_create(ppx_ctx* ctx, const ppx_smoother_params* params, ppx_smoother** out);
int ppx_smoother_destroy(ppx_smoother* s);
/* nearest_obstacle / nearest_edge: rows * cols (row, col) int32 pairs, row-major like the map grids */
int ppx_smoother_set_gvd(ppx_smoother* s, const int32_t* nearest_obstacle, const int32_t* nearest_edge);
/* the same two grids from the GVD ppx_gvd_build left on the device (PPX_ERR_STATE without one) */
int ppx_smoother_use_gvd(ppx_smoother* s);
/*
 * poses / cusp / count: the arrays ppx_hybrid_resample returns (max_points entries per path; count[q] <= 0
 * = no path).  out_poses[q] receives Smoother::GetPath() (x, y moved; theta unchanged, smoother.cpp:152),
 * status[q] Smoother::Status: 0 MaxIteration, 1 StepTolerance, 2 PathSize, -1 Failure (unsafe result or no
 * path; the caller then keeps the unsmoothed path, hybrid_a_star.cpp:295-298).  iterations (may be NULL):
 * gradient steps taken.
 */
int ppx_smoother_smooth(ppx_smoother* s, int64_t nq, const double* poses, const uint8_t* cusp, const int32_t* count,
	int32_t max_points, double* out_poses, int32_t* status, int32_t* iterations);

/* ---------------------------------------------------------------------------------------------
 * N4 / BASELINE config C1: grid A* -- AStarN2 (algo/a_star_n2.h:29-34 over algo/a_star.h:326-434) with
 * AStarStatePropagatorFcnN2::GetNeighborStates (algo/a_star_n2.cpp:12-28), for a batch of (start, goal) cells on
 * the context's occupancy grid.  The reference's step-cost and heuristic callbacks (std::function) are fixed to
 * the Euclidean cell distance (interfaces/python/scripts/example_a_star_grid.py:46-52).
 */
typedef struct ppx_astar_result {
	int32_t status; /* 0 Success, -1 Failure (frontier exhausted, or an end point outside the grid), PPX_ERR_CAPACITY */
	int32_t n_explored; /* GetExploredStates().size() */
	int32_t n_nodes; /* search-tree nodes created */
	int32_t path_len; /* cells on GetPath(), start and goal included */
	double cost; /* GetOptimalCost() */
} ppx_astar_result;

/* starts / goals: nq (row, col) pairs; paths (may be NULL with max_path 0): max_path (row, col) pairs per query,
 * -1 past path_len; max_nodes 0 = 4 * cells. */
int ppx_astar_n2(ppx_ctx* ctx, int64_t nq, const int32_t* starts, const int32_t* goals, int32_t max_nodes, int32_t max_path,
	ppx_astar_result* results, int32_t* paths);

/* The same with a choice of functors, and BidirectionalAStarN2 (algo/bidirectional_a_star.h:41-204, a_star_n2.h:36-42).
 * The reference's callbacks are std::function objects (Python lambdas through pyplanning.cpp:140-154); a callback cannot
 * cross a C ABI into a kernel, so the device offers the table of functors those callbacks are in practice:
 *   cost_kind       0 Euclidean cell distance, 1 Manhattan, 2 octile (max + (sqrt 2 - 1) min), 3 Euclidean x cell_cost[neighbour]
 *   heuristic_kind  0 Euclidean, 1 Manhattan, 2 octile to the search's goal; 3 zero
 * bidirectional != 0: forward and reverse searches advance in turn (pop + Expand + FindIntersection each) until the two
 * frontier tops prove the best meeting point optimal; the forward search's heuristic aims at the goal, the reverse one at
 * the start.  (The reference's AverageHeuristic pair never sets the goal of the heuristics it averages -- built from one
 * heuristic object, as example_a_star_grid.py:97 does, it is the constant zero: that is heuristic_kind 3.)
 * Results: n_explored / n_nodes are the sums over both searches, cost = GetOptimalCost(), the path is GetPath(): the forward
 * path followed by the reverse search's path reversed, the meeting cell twice (bidirectional_a_star.h:66-72). */
typedef struct ppx_astar_params {
	int32_t cost_kind, heuristic_kind;
	const float* cell_cost; /* host, rows * cols, cost_kind 3 only */
	int32_t bidirectional;
	int32_t _pad;
} ppx_astar_params;
int ppx_astar_n2_ex(ppx_ctx* ctx, int64_t nq, const int32_t* starts, const int32_t* goals, const ppx_astar_params* params,
	int32_t max_nodes, int32_t max_path, ppx_astar_result* results, int32_t* paths);

/* ---------------------------------------------------------------------------------------------
 * Multi-GPU.  The path shards by independent units -- edge batches, start/goal queries -- with the world model
 * replicated; one context per GPU, one process (or thread) per context.  The reference has no counterpart (it is
 * single-threaded, no communication backend); these are the two collectives a sharded caller needs, over NCCL / NVLink:
 *   ppx_comm_broadcast_map     the grids of the root's context (what ppx_map_upload / ppx_gvd_build installed: description,
 *                              sqdist, occupancy, pathcost) into every rank's context, once per map update;
 *   ppx_comm_gather_results    fixed-size records of every rank (e.g. ppx_hybrid_result of its shard), uneven shards
 *                              allowed, concatenated in rank order on every rank;
 *   ppx_comm_gather_paths      variable-length pose arrays (ppx_hybrid_resample's count / poses) by counts -> offsets:
 *                              every rank receives offsets[total_queries + 1] and the packed poses of all queries, rank
 *                              order, query order within a rank.
 * All device pointers; collectives are enqueued on the context's stream, the calls return after the host-side
 * bookkeeping they need (counts) -- ppx_ctx_synchronize before reading the outputs.  NCCL is loaded at run time
 * (libnccl.so.2); PPX_ERR_STATE when it is not there.
 * Bootstrap: rank 0 calls ppx_comm_unique_id and hands the 128 bytes to the other ranks by any means (MPI, a file,
 * torch.distributed's store); every rank then calls ppx_comm_init. */
#define PPX_COMM_ID_BYTES 128
int ppx_comm_unique_id(void* id_out /* PPX_COMM_ID_BYTES */);
int ppx_comm_init(ppx_ctx* ctx, const void* id, int32_t rank, int32_t world);
int ppx_comm_destroy(ppx_ctx* ctx);
int ppx_comm_broadcast_map(ppx_ctx* ctx, int32_t root);
/* d_all: room for `capacity` records; counts (host, world entries, may be NULL): records per rank */
int ppx_comm_gather_results(ppx_ctx* ctx, const void* d_local, int64_t n_local, int32_t record_bytes, void* d_all, int64_t capacity,
	int64_t* counts);
/* d_count[n_local], d_poses[n_local][max_points][3] as ppx_hybrid_resample lays them out (on the device);
 * query_counts (host, world entries, may be NULL); d_offsets_all: total_queries + 1 int64; d_poses_all: total poses x 3 */
int ppx_comm_gather_paths(ppx_ctx* ctx, int64_t n_local, const int32_t* d_count, const double* d_poses, int32_t max_points,
	int64_t* query_counts, int64_t* d_offsets_all, int64_t offsets_capacity, double* d_poses_all, int64_t poses_capacity,
	int64_t* total_poses);

#ifdef __cplusplus
}
#endif
#endif /* PPX_H */

/* hg_b200.h — C-ABI of libhg_b200.so: HectorGrapher's 3D scan-matching hot path on B200.
 *
 * Plain C: opaque handles, plain pointers and sizes, int status codes. No Eigen,
 * Ceres, protobuf or torch types cross this boundary. Every entry point names
 * the reference interface it stands in for (paths relative to
 * /root/reference/cartographer). Host-side C++ adapters that present the
 * reference's own class surfaces on top of these calls are in
 * hectorgrapher_b200/csrc/host/ and INTEGRATION.md shows the reference-side glue.
 *
 * Conventions
 *  - status: 0 = HG_OK; anything else is an error, hg_last_error() gives the text
 *    (thread-local). The reference aborts on these conditions (CHECK / LOG(FATAL)).
 *  - poses are double t[3], q[4] with q = (w, x, y, z)
 *    (mapping/internal/optimization/ceres_pose.h:49-50).
 *  - point clouds are float xyz with a stride in floats: 3 = sensor::PointCloud,
 *    4 = sensor::TimedPointCloud (sensor/rangefinder_point.h:31-40).
 *  - voxel lists are the content of the grid wire formats
 *    (mapping/proto/3d/hybrid_grid.proto:19-29, hybrid_grid_tsdf.proto:19-32):
 *    int32 xyz triples + raw uint16 values exactly as HybridGrid stores them.
 *  - all handles may be used from several host threads for evaluate / match
 *    calls (each call takes its own stream and scratch); create / destroy are
 *    not to race with use of the same handle.
 *  - there is no CPU fallback: without a CUDA device every compute entry point
 *    fails with HG_ERR_CUDA.
 */
#ifndef HG_B200_H_
#define HG_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HG_OK 0
#define HG_ERR_INVALID 1
#define HG_ERR_CUDA 2
#define HG_ERR_LIMIT 3

typedef struct hg_grid hg_grid;   /* HybridGrid or HybridGridTSDF as a device brick table */
typedef struct hg_cost hg_cost;   /* one scan-to-grid residual block (a ceres::CostFunction) */
typedef struct hg_batch hg_batch; /* many residual blocks evaluated in one launch */
typedef struct hg_stack hg_stack; /* PrecomputationGridStack3D */
typedef struct hg_fcsm hg_fcsm;   /* FastCorrelativeScanMatcher3D */

/* Text of the last error on this thread ("" if none). */
const char* hg_last_error(void);
/* Library version string. */
const char* hg_version(void);
/* Number of kernels this library has launched in this process (all threads). */
uint64_t hg_kernel_launch_count(void);
/* Selects the CUDA device used by subsequent calls on this thread (cudaSetDevice). */
int hg_set_device(int device);

/* ---------------------------------------------------------------- grids -- */

/* mapping/3d/hybrid_grid.h:46-140 (HybridGrid). 'grid_size' is the reference
 * grid's current grid_size() (64 << bits, hybrid_grid_base.h:260); it only
 * feeds MatchFullSubmap's window (fast_correlative_scan_matcher_3d.cc:186-188).
 * Pass 0 to have it derived the way DynamicGrid grows (smallest 64<<bits, bits>=1,
 * containing every voxel). Values may carry the update marker (bit 15). */
int hg_grid_upload_prob(float resolution, int grid_size, int64_t num_voxels, const int32_t* xyz,
                        const uint16_t* values, hg_grid** out);

/* mapping/3d/hybrid_grid_tsdf.h:57-150 (HybridGridTSDF). max_tsd =
 * relative_truncation_distance * resolution (hybrid_grid_tsdf.h:65-67). */
int hg_grid_upload_tsdf(float resolution, float relative_truncation_distance, float max_weight,
                        int grid_size, int64_t num_voxels, const int32_t* xyz,
                        const uint16_t* tsd_values, const uint16_t* weight_values, hg_grid** out);

void hg_grid_destroy(hg_grid* grid);
int hg_grid_size(const hg_grid* grid);
int64_t hg_grid_num_bricks(const hg_grid* grid);
/* Bytes of the grid's packed 2x2x2 stencil copy that is valid right now, 0 if there is none. The
 * copy (32 bytes per voxel of every brick; no reference counterpart: a read-side layout of the
 * same values HybridGridTSDF::GetTSD / HybridGrid::GetProbability return, hybrid_grid_tsdf.h:105-113)
 * is built when a cost function over >= 32768 points is created on the grid (hg_batch_create,
 * hg_cost_create), marked stale when the grid is written and rebuilt by that cost function's next
 * evaluation. */
int64_t hg_grid_stencil_bytes(const hg_grid* grid);

/* HybridGrid::GetProbability / HybridGridTSDF::GetTSD / GetWeight at integer
 * cells (hybrid_grid.h:133-135, hybrid_grid_tsdf.h:105-113), evaluated on the
 * device. what: 0 probability, 1 tsd, 2 weight. Host pointers. */
int hg_grid_lookup(const hg_grid* grid, int what, int64_t n, const int32_t* xyz, float* out);

/* InterpolatedTSDF::GetWeight (interpolated_tsdf.h:129-156) at real-valued
 * points (world frame of the grid, doubles), on the device: the trilinear blend
 * of the eight corner weights around the lower voxel. The multi-resolution
 * variant (interpolated_multi_resolution_tsdf.h:139-171) returns from the first
 * iteration of its loop, i.e. it is this call on the pyramid's finest grid.
 * TSDF grids only. Host pointers. */
int hg_grid_interpolated_weight(const hg_grid* grid, int64_t n, const double* points_xyz, double* out);

/* HybridGridBase::GetCellIndex (hybrid_grid_base.h:428-433) on the device. */
int hg_get_cell_index(float resolution, int64_t n, const float* points_xyz, int32_t* out_xyz);

/* The grid's non-default voxels, i.e. what HybridGrid::ToProto / HybridGridTSDF::ToProto
 * iterate (hybrid_grid.h:88-99, hybrid_grid_tsdf.h:119-134; order: brick by brick, z-major
 * inside a brick — the reference's iterator order is a property of its tree and is not
 * reproduced). values_b is the TSDF weight array (ignored for probability grids). Host
 * pointers; *num_voxels receives the count and nothing is written when it exceeds
 * max_voxels (call with max_voxels = 0 to size the arrays). */
int hg_grid_export(const hg_grid* grid, int64_t max_voxels, int64_t* num_voxels, int32_t* xyz,
                   uint16_t* values_a, uint16_t* values_b);

/* ------------------------------------------- grid wire format (N3) -- */

/* Scalar fields of proto::HybridGrid / proto::HybridGridTSDF
 * (mapping/proto/3d/hybrid_grid.proto:19-29, hybrid_grid_tsdf.proto:19-32). */
typedef struct hg_grid_proto_info {
  float resolution;
  float relative_truncation_distance; /* TSDF only (field 8, as stored — see hg_grid_to_proto) */
  float max_weight;                   /* TSDF only (field 9) */
  int32_t is_tsdf;
  int64_t num_voxels;
} hg_grid_proto_info;

/* Decodes a serialized proto::HybridGrid (is_tsdf = 0) or proto::HybridGridTSDF (is_tsdf = 1)
 * message on the host: packed or unpacked repeated fields, zig-zag indices, unknown fields
 * skipped; the index and value arrays must have equal lengths like the CHECKs of the
 * constructors (hybrid_grid.h:63-66, hybrid_grid_tsdf.h:73-77). Fills *info; writes the
 * voxels only if capacity >= info->num_voxels (arrays may be NULL when capacity is 0).
 * values_b receives the weights of a TSDF grid. No device work. */
int hg_grid_proto_decode(const void* bytes, int64_t size, int is_tsdf, hg_grid_proto_info* info,
                         int64_t capacity, int32_t* xyz, uint16_t* values_a, uint16_t* values_b);

/* HybridGrid(const proto::HybridGrid&) / HybridGridTSDF(const proto::HybridGridTSDF&)
 * (hybrid_grid.h:61-73, hybrid_grid_tsdf.h:69-83): decode + upload. Values go through the
 * constructors' SetProbability(ValueToProbability(v)) / SetCell(ValueToTSD(t), ValueToWeight(w))
 * round trip, i.e. the update marker of a stored TSD value is dropped and re-added and value
 * 0 becomes 1. */
int hg_grid_from_proto(const void* bytes, int64_t size, int is_tsdf, hg_grid** out);

/* HybridGrid::ToProto / HybridGridTSDF::ToProto (hybrid_grid.h:88-99, hybrid_grid_tsdf.h:119-134)
 * serialized to wire format (packed repeated fields, the encoding protobuf emits for proto3).
 * Like the reference, the TSDF message's relative_truncation_distance field receives
 * ValueConverter().getMaxTSD() — the truncation distance in metres, not the relative one
 * (hybrid_grid_tsdf.h:131) — so a grid that is written and read back truncates at
 * max_tsd * resolution. *size receives the message size; nothing is written if it exceeds
 * capacity. Voxel order: see hg_grid_export. */
int hg_grid_to_proto(const hg_grid* grid, void* buffer, int64_t capacity, int64_t* size);

/* A .pbstream file (io/proto_stream.cc:27-106: magic, then size-prefixed gzip chunks holding
 * a SerializationHeader and SerializedData messages, mapping/proto/serialization.proto:72-88),
 * read for its 3D submaps. */
typedef struct hg_pbstream hg_pbstream;

typedef struct hg_submap_info {
  int32_t trajectory_id;   /* SubmapId (pose_graph.proto:22-25) */
  int32_t submap_index;
  double local_pose[7];    /* Submap3D.local_pose, t3 q4 (w, x, y, z) */
  int32_t num_range_data;
  int32_t finished;
  int32_t high_is_tsdf;    /* which member of the oneofs is set; -1 = none */
  int32_t low_is_tsdf;
  int32_t histogram_size;  /* rotational_scan_matcher_histogram */
  int32_t reserved;
} hg_submap_info;

int hg_pbstream_open(const char* path, hg_pbstream** out);
/* Advances to the next SerializedData that holds a Submap3D (everything else is skipped).
 * *has_submap = 0 at the end of the stream. */
int hg_pbstream_next_submap(hg_pbstream* stream, int* has_submap, hg_submap_info* info);
/* Uploads the current submap's high- (which = 0) or low-resolution (which = 1) grid. */
int hg_pbstream_submap_grid(hg_pbstream* stream, int which, hg_grid** out);
/* Copies the current submap's rotational scan matcher histogram (capacity in floats). */
int hg_pbstream_submap_histogram(hg_pbstream* stream, int capacity, float* histogram);
void hg_pbstream_close(hg_pbstream* stream);

/* ------------------------------------------ range-data insertion (N1) -- */

/* proto::ProbabilityGridRangeDataInserterOptions3D
 * (mapping/proto/3d/probability_grid_range_data_inserter_options_3d.proto). */
typedef struct hg_occupancy_inserter_options {
  double hit_probability;
  double miss_probability;
  int32_t num_free_space_voxels;
  int32_t reserved;
} hg_occupancy_inserter_options;

/* OccupancyGridRangeDataInserter3D::Insert (mapping/3d/range_data_inserter_3d.cc:27-101)
 * on a device-resident probability grid: every return's cell gets the hit odds, the last
 * num_free_space_voxels cells of every ray the miss odds, each cell at most once per call
 * with hits taking precedence (the update-marker rule of HybridGrid::ApplyLookupTable,
 * hybrid_grid.h:119-130, including FinishUpdate). The grid grows as DynamicGrid does
 * (hybrid_grid_base.h:385-403). 'returns' are host points, 'stride_floats' floats apart.
 * Like the reference's Insert on a submap's grid, an insertion must not run concurrently
 * with any other use of the same grid. Cost functions created before an insertion read the
 * grid's new content afterwards; a precomputation stack / matcher is a snapshot of the grid
 * at hg_stack_build / hg_fcsm_create time, as in the reference (built per finished submap). */
int hg_grid_insert_occupancy(hg_grid* grid, const hg_occupancy_inserter_options* options,
                             const float origin[3], const float* returns, int stride_floats,
                             int64_t num_returns);

/* proto::TSDFRangeDataInserterOptions3D, the fields read by the branch without normal
 * projection (mapping/proto/3d/tsdf_range_data_inserter_options_3d.proto). */
typedef struct hg_tsdf_inserter_options {
  double relative_truncation_distance;
  double maximum_weight;
  double weight_function_epsilon;
  double weight_function_sigma;
  double min_range;
  double max_range;
  double insertion_ratio;
  int32_t num_free_space_voxels;
  int32_t reserved;
} hg_tsdf_inserter_options;

/* TSDFRangeDataInserter3D::Insert with project_sdf_distance_to_scan_normal = false
 * (mapping/3d/tsdf_range_data_inserter_3d.cc:700-722): insertion-ratio counter, NaN and
 * range gates, then InsertHit's default ray cast (:294-343) and UpdateCell (:725-737) for
 * every return. A cell's updates are applied in the reference's order (return by return,
 * sample by sample), re-quantised to uint16 after each one like HybridGridTSDF::SetCell, so
 * the stored voxels are the reference's bit for bit. num_updates (may be NULL) receives
 * the number of (return, sample) cell updates applied. */
int hg_grid_insert_tsdf(hg_grid* grid, const hg_tsdf_inserter_options* options,
                        const float origin[3], const float* returns, int stride_floats,
                        int64_t num_returns, int64_t* num_updates);

/* ------------------------------------------------------- cost functions -- */

/* The eight functors of mapping/internal/3d/scan_matching/, numbered in the
 * order SURVEY.md lists them. */
#define HG_COST_OCCUPIED_SPACE 0                                   /* occupied_space_cost_function_3d.h:36-80 */
#define HG_COST_INTERPOLATED_OCCUPIED_SPACE 1                      /* interpolated_occupied_space_cost_function_3d.h:37-92 */
#define HG_COST_TSDF_SPACE 2                                       /* tsdf_space_cost_function_3d.h:37-79 */
#define HG_COST_INTERPOLATED_TSDF_SPACE 3                          /* interpolated_tsdf_space_cost_function_3d.h:38-96 */
#define HG_COST_INTERPOLATED_TSDF_PER_POINT_SPACE 4                /* interpolated_tsdf_per_point_space_cost_function_3d.h:38-74 */
#define HG_COST_MULTI_RESOLUTION_TSDF_SPACE 5                      /* multi_resolution_tsdf_space_cost_function_3d.h:37-82 */
#define HG_COST_INTERPOLATED_MULTI_RESOLUTION_TSDF_SPACE 6         /* interpolated_multi_resolution_tsdf_space_cost_function_3d.h:38-98 */
#define HG_COST_INTERPOLATED_MULTI_RESOLUTION_TSDF_PER_POINT_SPACE 7 /* ..._per_point_space_cost_function_3d.h:38-74 */

#define HG_MAX_PYRAMID_LEVELS 4

/* XxxCostFunction3D::CreateAutoDiffCostFunction(scaling_factor, point_cloud,
 * grid | grid pyramid [, interpolation_ratio]). The cloud is copied to the
 * device. 'grids': one HybridGrid (kinds 0-1), one HybridGridTSDF (2-4) or an
 * ascending-resolution TSDF pyramid (5-7; the handles are copied, unlike
 * interpolated_multi_resolution_tsdf.h:218 which keeps a reference). One-pose
 * kinds have parameter blocks {3,4}; interpolated kinds {3,4,3,4}. */
int hg_cost_create(int kind, double scaling_factor, const float* points, int stride_floats,
                   int num_points, const hg_grid* const* grids, int num_grids,
                   double interpolation_ratio, hg_cost** out);
void hg_cost_destroy(hg_cost* cost);
int hg_cost_num_residuals(const hg_cost* cost);
int hg_cost_num_parameter_blocks(const hg_cost* cost);

/* ceres::CostFunction::Evaluate(parameters, residuals, jacobians) (Ceres 1.13
 * cost_function.h; instantiations at tsdf_space_cost_function_3d.h:40-44 and
 * interpolated_tsdf_space_cost_function_3d.h:45-51). All pointers are HOST
 * memory. 'jacobians' may be NULL, and any jacobians[b] may be NULL (block held
 * constant). jacobians[b] is row-major num_residuals x block_size. */
int hg_cost_evaluate(const hg_cost* cost, const double* const* parameters, double* residuals,
                     double** jacobians);

/* Same evaluation with DEVICE output pointers (any may be NULL); parameters
 * stay host pointers. Runs on 'cuda_stream' (a cudaStream_t, NULL = default
 * stream) and does not synchronise. */
int hg_cost_evaluate_device(const hg_cost* cost, const double* const* parameters,
                            double* d_residuals, double* const* d_jacobians, void* cuda_stream);

/* Gauss-Newton terms of one block without materialising the Jacobian:
 * JtJ (n x n row-major, n = 7 or 14 ambient parameters in block order),
 * Jtr (n) and cost = 1/2 sum r^2. Host outputs. */
int hg_cost_normal_equations(const hg_cost* cost, const double* const* parameters, double* JtJ,
                             double* Jtr, double* cost_out);

/* No counterpart in the reference. Same result as hg_cost_normal_equations for one
 * residual block, but the call allocates nothing: no cudaMalloc / cudaFree, no
 * allocating library routine. It exists for callers that must not use the CUDA
 * allocator once they are running — a real-time scan-matching loop (allocation
 * synchronises the device and has no time bound) or a process whose device memory is
 * already owned by its own arena. Its scratch memory is a fixed array that is part of
 * the library image on every GPU (984 KB per GPU, held until exit), which bounds the
 * block size: at most 1,048,576 points, larger blocks return HG_ERR_LIMIT. The cost
 * function's device must be the current device. The first call on a device looks up
 * the array's address (make that call before starting any stream capture). All calls
 * share the one array per GPU and take a process-wide lock; the function returns
 * after the device has finished. Two calls that overlapped on the GPU (which the lock
 * prevents inside one process) would race on the array. The call does NOT take the
 * cost function's own lock: do not run it concurrently with hg_batch_update_points,
 * hg_cost_evaluate or hg_cost_normal_equations on the SAME cost function from another
 * thread (points queued by an earlier hg_batch_update_points are waited for on the
 * stream). Host outputs, any may be NULL. */
int hg_cost_normal_equations_static(const hg_cost* cost, const double* const* parameters,
                                    double* JtJ, double* Jtr, double* cost_out);

/* ------------------------------------------------- batched cost blocks -- */

/* Many residual blocks of one kind over the same grids, evaluated by one
 * launch: the shape OptimizingLocalTrajectoryBuilder builds per optimisation
 * (mapping/internal/3d/optimizing_local_trajectory_builder.cc:323-511: one
 * block per scan of the window, each on one or two of the window's control
 * poses). Block b uses points[cloud_offsets[b] .. cloud_offsets[b+1]). */
int hg_batch_create(int kind, int num_blocks, const float* points, int stride_floats,
                    const int64_t* cloud_offsets /* num_blocks + 1 */,
                    const double* scaling_factors /* num_blocks */,
                    const double* interpolation_ratios /* num_blocks, may be NULL for one-pose */,
                    const hg_grid* const* grids, int num_grids, hg_batch** out);
void hg_batch_destroy(hg_batch* batch);
int64_t hg_batch_num_residuals(const hg_batch* batch);
/* Replaces the point clouds (same offsets) from HOST memory: the next window of scans.
 * The copy is queued chunk by chunk on a stream of the batch and the call returns; the next
 * evaluation of the batch starts on a chunk of blocks as soon as that chunk's points have
 * arrived (with host outputs it also sends each chunk's results back while later chunks are
 * still being copied in and computed). Pinned 'points' must therefore stay unchanged until the
 * next evaluation with host outputs (or hg_batch_normal_equations) has returned, or the device
 * has been synchronised; pageable memory is staged by the CUDA runtime before the call returns.
 * Waits for evaluations with device outputs that are still queued on the old points. */
int hg_batch_update_points(hg_batch* batch, const float* points, int stride_floats);

/* Evaluates every block. 'poses' = num_poses x 7 doubles (t3, q4) on the host;
 * block b uses pose_index0[b] (and pose_index1[b] for interpolated kinds).
 * Outputs are HOST pointers when device_outputs == 0, DEVICE pointers otherwise;
 * any may be NULL. Layout: residuals[N]; jt0[N x 3], jq0[N x 4], jt1[N x 3],
 * jq1[N x 4] row-major, so the slice of block b is exactly the jacobians[k]
 * array Ceres would pass for it. With host outputs the call returns when the outputs
 * are in place. With device outputs it returns once the work is queued on 'cuda_stream'
 * (the batch's own stream when NULL): several such calls may be in flight, each with its
 * own poses — the block descriptors go through a ring of staging slots. */
int hg_batch_evaluate(const hg_batch* batch, const double* poses, int num_poses,
                      const int32_t* pose_index0, const int32_t* pose_index1, int device_outputs,
                      double* residuals, double* jt0, double* jq0, double* jt1, double* jq1,
                      void* cuda_stream);

/* Per-block Gauss-Newton terms for the whole batch: JtJ[num_blocks x n x n],
 * Jtr[num_blocks x n], cost[num_blocks]; n = 7 (one-pose) or 14. Host outputs. */
int hg_batch_normal_equations(const hg_batch* batch, const double* poses, int num_poses,
                              const int32_t* pose_index0, const int32_t* pose_index1, double* JtJ,
                              double* Jtr, double* cost_out);

/* Diagnostic for multi-resolution TSDF batches: how many points each pyramid level answers
 * (InterpolatedMultiResolutionTSDF::GetTSD takes the first level whose eight corners are all
 * observed, interpolated_multi_resolution_tsdf.h:99-106). counts[l] for level l of the batch's
 * grid list, counts[num_grids] for points no level answers (they read the finest level's
 * minimum TSD, :136). counts has num_grids + 1 entries. */
int hg_batch_level_usage(const hg_batch* batch, const double* poses, int num_poses,
                         const int32_t* pose_index0, const int32_t* pose_index1, int64_t* counts);

/* CeresScanMatcher3D::Match (ceres_scan_matcher_3d.cc:72-95, problem set-up
 * :120-172) for num_clouds (cloud, grid) pairs: occupied-space / TSDF blocks
 * scaled by occupied_space_weights[i] / sqrt(N_i), plus the translation and
 * rotation delta priors (translation_delta_cost_functor_3d.h:31-45,
 * rotation_delta_cost_functor_3d.h:35-54). Solved with the dense LM of
 * include/hg_dense_lm.h over the GPU cost blocks. pose arrays are t3 q4.
 * summary_out (may be NULL): initial cost, final cost, iterations, successful
 * steps, cost evaluations, jacobian evaluations. */
int hg_scan_match(int num_clouds, const float* const* clouds, const int32_t* cloud_sizes,
                  const hg_grid* const* grids, const double* occupied_space_weights,
                  double translation_weight, double rotation_weight, int only_optimize_yaw,
                  int max_num_iterations, int use_nonmonotonic_steps,
                  const double* target_translation, const double* initial_pose, double* pose_out,
                  double* summary_out);

/* CeresScanMatcher3D::Evaluate (ceres_scan_matcher_3d.cc:97-118): the problem of hg_scan_match
 * evaluated once at 'initial_pose'. cost = 1/2 |r|^2; residuals (may be NULL): sum of the cloud
 * sizes + 6 doubles — every cloud's residuals in order, then the translation and the rotation
 * delta; gradient (may be NULL): 6 doubles (4 with only_optimize_yaw), the gradient of the cost in
 * the tangent space of (translation, rotation) — what the reference's call receives, because it
 * passes its 'jacobians' vector as ceres::Problem::Evaluate's gradient argument. */
int hg_scan_match_evaluate(int num_clouds, const float* const* clouds, const int32_t* cloud_sizes,
                           const hg_grid* const* grids, const double* occupied_space_weights,
                           double translation_weight, double rotation_weight, int only_optimize_yaw,
                           const double* target_translation, const double* initial_pose, double* cost,
                           double* residuals, double* gradient);

/* ------------------------------------------------ precomputation grids -- */

/* PrecomputationGridStack3D(grid, options)
 * (fast_correlative_scan_matcher_3d.cc:60-94; ConvertToPrecomputationGrid and
 * PrecomputeGrid, precomputation_grid_3d.cc:49-97). */
int hg_stack_build(const hg_grid* grid, int branch_and_bound_depth, int full_resolution_depth,
                   hg_stack** out);
void hg_stack_destroy(hg_stack* stack);
int hg_stack_depth(const hg_stack* stack);
/* PrecomputationGrid3D::value at integer cells of level 'depth' (host pointers). */
int hg_stack_lookup(const hg_stack* stack, int depth, int64_t n, const int32_t* xyz, uint8_t* out);
/* Number of non-zero voxels of a level, and their export (for parity checks). */
int64_t hg_stack_num_nonzero(const hg_stack* stack, int depth);
int hg_stack_export(const hg_stack* stack, int depth, int64_t capacity, int32_t* xyz,
                    uint8_t* values, int64_t* num_written);

/* ------------------------------------------ fast correlative scan matcher -- */

/* proto::FastCorrelativeScanMatcherOptions3D
 * (fast_correlative_scan_matcher_3d.cc:37-58, configuration_files/pose_graph.lua:40-49). */
typedef struct hg_fcsm_options {
  int32_t branch_and_bound_depth;
  int32_t full_resolution_depth;
  int32_t use_rotational_scan_matcher;
  double min_rotational_score;
  double min_low_resolution_score;
  double linear_xy_search_window;
  double linear_z_search_window;
  double angular_search_window;
  /* 1 = Eigen's SSE2 float summation orders (x86-64 reference build), 0 = Eigen's
   * generic scalar code. See DESIGN.md "float flavors". */
  int32_t float_flavor;
  /* 0 (default): search only the candidates the reference's loop can reach and settle
   * cross-subtree ties by a probe — same selected candidate, far less host work.
   * 1: sort and visit every candidate exactly in the reference's order (the
   * num_candidates_visited statistic then equals the reference's own count). */
  int32_t strict_visit_order;
} hg_fcsm_options;

/* TrajectoryNode::Data fields the matcher reads (mapping/trajectory_node.h:45-62). */
typedef struct hg_node_data {
  const float* high_resolution_point_cloud; /* xyz, stride 3 */
  int32_t num_high_resolution_points;
  const float* low_resolution_point_cloud;
  int32_t num_low_resolution_points;
  const float* rotational_scan_matcher_histogram;
  int32_t histogram_size;
  double gravity_alignment[4]; /* (w, x, y, z) */
} hg_node_data;

/* FastCorrelativeScanMatcher3D::Result (fast_correlative_scan_matcher_3d.h:68-73)
 * plus search statistics. found == 0 corresponds to a null result. */
typedef struct hg_match_result {
  int32_t found;
  float score;
  double pose_translation[3];
  double pose_rotation[4]; /* (w, x, y, z) */
  float rotational_score;
  float low_resolution_score;
  int32_t best_scan_index;
  int32_t best_offset[3];
  int32_t num_discrete_scans;
  int64_t num_candidates_scored;     /* candidates the GPU scored (a superset of the DFS's) */
  int64_t num_candidates_visited;    /* candidates the reference's DFS order consumed */
  int32_t num_score_launches;
  int64_t num_lowest_resolution_candidates; /* size of the lowest-resolution lattice (one launch) */
  float lowest_resolution_kernel_ms;        /* its device time, CUDA events on the launching stream */
  /* What a reduce over shards needs to order equal leaf scores the way the reference's search
   * visits them (fast_correlative_scan_matcher_3d.cc:384, 437-468: subtrees in descending order of
   * their lowest-resolution score): score of the winner's lowest-resolution ancestor and that
   * ancestor's position in the unsharded generation order (:328-361). */
  float best_ancestor_score;
  int64_t best_ancestor_index;
  /* Score of the best valid leaf the greedy dive found (0 = none): a lower bound on the answer
   * that shards may exchange before the frontier expansion. */
  float dive_bound;
  /* 1 when the search was one shard of a sweep and several of its leaves tied for the best score
   * below DIFFERENT lowest-resolution candidates of EQUAL score: their order in the reference is
   * that of std::sort over the whole (unsharded) list; the earlier generation index was taken. */
  int32_t order_unresolved;
} hg_match_result;

/* FastCorrelativeScanMatcher3D(grid, low_resolution_grid, histogram, options)
 * (fast_correlative_scan_matcher_3d.cc:129-153). The stack is built here. */
int hg_fcsm_create(const hg_grid* grid, const hg_grid* low_resolution_grid,
                   const float* rotational_scan_matcher_histogram, int histogram_size,
                   const hg_fcsm_options* options, hg_fcsm** out);
void hg_fcsm_destroy(hg_fcsm* matcher);
const hg_stack* hg_fcsm_stack(const hg_fcsm* matcher);

/* FastCorrelativeScanMatcher3D::Match (fast_correlative_scan_matcher_3d.cc:157-174).
 * Poses are t3 q4 doubles. */
int hg_fcsm_match(const hg_fcsm* matcher, const double* global_node_pose,
                  const double* global_submap_pose, const hg_node_data* data, float min_score,
                  hg_match_result* result);

/* FastCorrelativeScanMatcher3D::MatchFullSubmap (:176-200). Rotations are (w,x,y,z).
 * 'shard_index' / 'num_shards' restrict the search to the discrete scans
 * s with s % num_shards == shard_index (multi-GPU sharding by yaw; 0 / 1 = all).
 * best_scan_index is always the index in the unsharded scan list. */
int hg_fcsm_match_full_submap(const hg_fcsm* matcher, const double* global_node_rotation,
                              const double* global_submap_rotation, const hg_node_data* data,
                              float min_score, int shard_index, int num_shards,
                              hg_match_result* result);

/* FastCorrelativeScanMatcher3D::Match for many nodes against this matcher's submap in one
 * device-resident search — what ConstraintBuilder3D's thread pool does one task at a time for the
 * pairs of a submap (constraint_builder_3d.cc:97-106, 213-246). Poses are num_problems x 7
 * doubles (t3 q4); results[p] is what hg_fcsm_match returns for problem p. */
int hg_fcsm_match_batch(const hg_fcsm* matcher, int num_problems, const double* global_node_poses,
                        const double* global_submap_poses, const hg_node_data* data, float min_score,
                        hg_match_result* results);

/* MatchFullSubmap (:176-200) of one shard in two steps, so that the shards of a multi-GPU search
 * can exchange one number in between: 'begin' scores the shard's lowest-resolution candidates and
 * follows the best one of every discrete scan greedily down to a leaf; *dive_bound receives the
 * score of the best such leaf that passes the low-resolution gate (0 = none) — a lower bound on
 * the answer of the WHOLE search. 'finish' takes the largest bound any shard found (or 0) and
 * searches this shard under it; it consumes the handle. Same answer as
 * hg_fcsm_match_full_submap on every shard; a shard whose scans hold nothing that reaches the
 * bound reports found == 0. Both calls must come from the same host thread, with no other search
 * of that thread in between; 'data' must stay valid until 'finish' returns. */
typedef struct hg_fcsm_search hg_fcsm_search;
int hg_fcsm_full_submap_begin(const hg_fcsm* matcher, const double* global_node_rotation,
                              const double* global_submap_rotation, const hg_node_data* data,
                              float min_score, int shard_index, int num_shards,
                              hg_fcsm_search** search, float* dive_bound);
int hg_fcsm_full_submap_finish(hg_fcsm_search* search, float bound, hg_match_result* result);

/* The first half without a host round trip: phase A is only queued — on *stream, a cudaStream_t of
 * the matcher's device that belongs to the calling thread — and *d_bound is the device address of a
 * float that holds this shard's dive bound (0 = none) once the queued work has run. The caller
 * all-reduces (max) that float in place over the shards with its own collective on *stream or on
 * a stream ordered after it (ncclAllReduce(d_bound, d_bound, 1, ncclFloat, ncclMax, comm, stream)),
 * then calls hg_fcsm_full_submap_finish(search, 0, result) from the same thread: phase B reads the
 * exchanged bound from device memory. *d_bound is null when the matcher searches on the host
 * (strict_visit_order): use hg_fcsm_full_submap_begin then. The float lives in the thread's
 * workspace: no other search of this thread may run between the two halves.
 * Same reference step as hg_fcsm_full_submap_begin (fast_correlative_scan_matcher_3d.cc:215-238, 384). */
int hg_fcsm_full_submap_begin_async(const hg_fcsm* matcher, const double* global_node_rotation,
                                    const double* global_submap_rotation, const hg_node_data* data,
                                    float min_score, int shard_index, int num_shards, hg_fcsm_search** search,
                                    float** d_bound, void** stream);
/* Drops a search that will not be finished. */
void hg_fcsm_search_abandon(hg_fcsm_search* search);

/* The one exchange of a sharded search: every shard packs its result into a
 * fixed-size record of HG_BEST_RECORD_DOUBLES doubles, the caller all-gathers the records (for
 * instance ncclAllGather with ncclDouble, or torch.distributed), and every rank selects the same
 * winner. Order: the reference's search keeps the FIRST best leaf it visits and visits the
 * lowest-resolution candidates in descending order of their score
 * (fast_correlative_scan_matcher_3d.cc:384, 437-440, 465-466): higher leaf score first; equal
 * leaf scores go to the leaf whose lowest-resolution ancestor scores higher. Ancestors of equal
 * score are ordered by the reference's std::sort of the whole lowest-resolution list, which only
 * an unsharded search in strict order reproduces: the earlier generation index wins here and
 * *order_unresolved (may be NULL) is set. *winner = index of the winning record, -1 if none. */
#define HG_BEST_RECORD_DOUBLES 20
int hg_pack_best_record(const hg_match_result* result, double* record);
int hg_unpack_best_record(const double* record, hg_match_result* result);
int hg_select_best_record(const double* records, int num_records, int* winner, int* order_unresolved);

/* Stage-level entry points (parity tests and the candidates/s benchmark). */

/* GenerateDiscreteScans (:276-326) for an explicit window. Host outputs, any
 * may be NULL: poses [S x 7 float: t3, q wxyz], rotational scores [S], cells
 * [S x depth x N x 3 int32]. *num_scans receives S; nothing is written if
 * S > max_scans. */
int hg_fcsm_discrete_scans(const hg_fcsm* matcher, int linear_xy_window_size,
                           int linear_z_window_size, double angular_search_window,
                           const double* global_node_pose, const double* global_submap_pose,
                           const hg_node_data* data, int max_scans, int32_t* num_scans,
                           float* poses, float* rotational_scores, int32_t* cells);

/* ScoreCandidates (:363-386) without the sort: integer gather sums and float
 * scores for explicit candidates (scan, ox, oy, oz) at 'depth'. Host pointers.
 * kernel_ms (may be NULL) receives the device time of the scoring kernel(s). */
int hg_fcsm_score_candidates(const hg_fcsm* matcher, int linear_xy_window_size,
                             int linear_z_window_size, double angular_search_window,
                             const double* global_node_pose, const double* global_submap_pose,
                             const hg_node_data* data, int depth, int64_t num_candidates,
                             const int32_t* candidates /* n x 4 */, int32_t* sums, float* scores,
                             float* kernel_ms);

/* CreateLowResolutionMatcher's closure (low_resolution_matcher.cc:24-69) for a
 * batch of float poses [n x 7: t3, q wxyz]; host pointers. */
int hg_low_resolution_scores(const hg_grid* low_resolution_grid, const float* points,
                             int num_points, int64_t num_poses, const float* poses, float* scores);

/* -------------------------------------------------------- voxel filters -- */

/* sensor::VoxelFilter(size).Filter (sensor/internal/voxel_filter.cc:25-73): the first
 * point of every voxel of edge 'size', input order kept. Works for any point type whose
 * first three floats are xyz (stride in floats). Host pointers; writes the indices of
 * the kept points (capacity num_points) and their number. */
int hg_voxel_filter(float size, const float* points, int stride_floats, int64_t num_points,
                    int64_t* num_kept, int32_t* kept_indices);

/* sensor::AdaptiveVoxelFilter(options).Filter (sensor/internal/adaptive_voxel_filter.h:34-111):
 * drops points beyond max_range, then voxel-filters with the largest edge length <= max_length
 * (halving, then bisection to within 10 %) that still leaves min_num_points points. */
int hg_adaptive_voxel_filter(float max_length, int min_num_points, float max_range, const float* points,
                             int stride_floats, int64_t num_points, int64_t* num_kept,
                             int32_t* kept_indices);

/* ------------------------------------------------- constraint batches -- */

/* One (node, submap) pair of ConstraintBuilder3D::ComputeConstraint
 * (mapping/internal/constraints/constraint_builder_3d.cc:192-295). */
typedef struct hg_constraint_job {
  int32_t submap_index;          /* which matcher / grids (index into the arrays below) */
  int32_t match_full_submap;     /* 0 = MaybeAddConstraint path, 1 = MaybeAddGlobalConstraint */
  double global_node_pose[7];    /* t3 q4 */
  double global_submap_pose[7];
  hg_node_data data;
} hg_constraint_job;

typedef struct hg_constraint_options {
  float min_score;                    /* constraint_builder options: min_score / global_localization_min_score */
  float global_localization_min_score;
  double occupied_space_weights[2];   /* ceres_scan_matcher_options_3d: high- and low-resolution cloud */
  double translation_weight;
  double rotation_weight;
  int32_t only_optimize_yaw;
  int32_t max_num_iterations;
  int32_t use_nonmonotonic_steps;
} hg_constraint_options;

typedef struct hg_constraint_result {
  int32_t found;
  float score;
  double pose[7];          /* refined submap-local pose of the node, t3 q4 */
  float rotational_score;
  float low_resolution_score;
  int64_t num_candidates_scored;
} hg_constraint_result;

/* Runs jobs [0, num_jobs): fast correlative match, then the Ceres-style refine
 * on {high-res cloud + grid, low-res cloud + low-res grid}
 * (constraint_builder_3d.cc:210-265). matchers[i], grids[i], low_grids[i]
 * belong to submap i. Only jobs with job_index % num_shards == shard_index are
 * computed (others are left untouched) — the multi-GPU partition. */
int hg_constraints_run_batch(const hg_fcsm* const* matchers, const hg_grid* const* grids,
                             const hg_grid* const* low_resolution_grids, int num_submaps,
                             const hg_constraint_job* jobs, int64_t num_jobs,
                             const hg_constraint_options* options, int shard_index, int num_shards,
                             hg_constraint_result* results);

#ifdef __cplusplus
}
#endif
#endif /* HG_B200_H_ */

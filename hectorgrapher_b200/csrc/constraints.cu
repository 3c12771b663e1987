// ConstraintBuilder3D::ComputeConstraint for a batch of (node, submap) pairs.
//
// Reference behaviour reproduced
// (/root/reference/cartographer/mapping/internal/constraints/constraint_builder_3d.cc):
//   :192-246  Match (local window) or MatchFullSubmap (global), null result => no constraint
//   :257-265  CeresScanMatcher3D::Match on {high-res cloud + grid, low-res cloud + low-res grid},
//             initial pose and target translation = the correlative match's pose
//   :267-272  constraint pose = refined pose
// The thread-pool / task bookkeeping of the reference (:77-190, :297-337) is
// control plane and stays with the caller; this entry point is the per-pair
// compute, partitioned across GPUs by job index (one process per GPU).
#include <chrono>
#include <cstdlib>

#include "hg_internal.cuh"

extern "C" int hg_constraints_run_batch(const hg_fcsm* const* matchers, const hg_grid* const* grids,
                                        const hg_grid* const* low_resolution_grids, int num_submaps,
                                        const hg_constraint_job* jobs, int64_t num_jobs,
                                        const hg_constraint_options* options, int shard_index,
                                        int num_shards, hg_constraint_result* results) {
  HG_REQUIRE(matchers && grids && low_resolution_grids && jobs && options && results && num_jobs >= 0,
             "hg_constraints_run_batch: bad arguments");
  HG_REQUIRE(num_shards >= 1 && shard_index >= 0 && shard_index < num_shards,
             "hg_constraints_run_batch: bad shard");
  for (int64_t j = shard_index; j < num_jobs; j += num_shards) {
    const hg_constraint_job& job = jobs[j];
    HG_REQUIRE(job.submap_index >= 0 && job.submap_index < num_submaps,
               "hg_constraints_run_batch: submap index out of range");
    hg_constraint_result& out = results[j];
    out = hg_constraint_result{};
    out.pose[3] = 1.0;
    hg_match_result match;
    int rc;
    const auto t0 = std::chrono::steady_clock::now();
    if (job.match_full_submap) {
      rc = hg_fcsm_match_full_submap(matchers[job.submap_index], job.global_node_pose + 3,
                                     job.global_submap_pose + 3, &job.data,
                                     options->global_localization_min_score, 0, 1, &match);
    } else {
      rc = hg_fcsm_match(matchers[job.submap_index], job.global_node_pose, job.global_submap_pose,
                         &job.data, options->min_score, &match);
    }
    if (rc != HG_OK) return rc;
    out.num_candidates_scored = match.num_candidates_scored;
    const auto t1 = std::chrono::steady_clock::now();
    if (!match.found) continue;
    out.score = match.score;
    out.rotational_score = match.rotational_score;
    out.low_resolution_score = match.low_resolution_score;
    double initial[7];
    for (int k = 0; k < 3; ++k) initial[k] = match.pose_translation[k];
    for (int k = 0; k < 4; ++k) initial[3 + k] = match.pose_rotation[k];
    const float* clouds[2] = {job.data.high_resolution_point_cloud, job.data.low_resolution_point_cloud};
    const int32_t sizes[2] = {job.data.num_high_resolution_points, job.data.num_low_resolution_points};
    const hg_grid* pair_grids[2] = {grids[job.submap_index], low_resolution_grids[job.submap_index]};
    const int num_clouds = sizes[1] > 0 ? 2 : 1;
    rc = hg_scan_match(num_clouds, clouds, sizes, pair_grids, options->occupied_space_weights,
                       options->translation_weight, options->rotation_weight, options->only_optimize_yaw,
                       options->max_num_iterations, options->use_nonmonotonic_steps, initial, initial,
                       out.pose, nullptr);
    if (rc != HG_OK) return rc;
    out.found = 1;
    if (std::getenv("HG_TRACE") != nullptr) {
      const auto t2 = std::chrono::steady_clock::now();
      std::fprintf(stderr, "[hg trace] constraint job %lld: match %.2f ms, refine %.2f ms\n", static_cast<long long>(j),
                   std::chrono::duration<double, std::milli>(t1 - t0).count(),
                   std::chrono::duration<double, std::milli>(t2 - t1).count());
    }
  }
  return HG_OK;
}

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
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <functional>
#include <mutex>
#include <thread>

#include "hg_internal.cuh"

namespace {

// The reference computes constraints on a thread pool (constraint_builder_3d.cc:97-106,
// one task per pair). One search is a chain of small launches with host decisions in
// between, so a single host thread leaves the GPU idle most of the time; several pairs in
// flight (each search on its own non-blocking stream with its own thread-local scratch,
// see fcsm.cu) fill it. The workers live as long as the process so that their scratch
// (pinned buffers in particular) is allocated once. HG_CONSTRAINT_THREADS sets the number
// of workers (default 3; 1 = compute on the calling thread).
class WorkerPool {
 public:
  static WorkerPool& Get() {
    static WorkerPool* pool = new WorkerPool;  // never destroyed: workers outlive main()
    return *pool;
  }
  int size() const { return static_cast<int>(threads_.size()); }

  // Runs fn(i) for i in [begin, end) with stride 'step' on the workers. Returns the first
  // failure's code and copies its error text to the calling thread.
  int Run(const int64_t begin, const int64_t end, const int64_t step, const int device,
          const std::function<int(int64_t)>& fn) {
    std::unique_lock<std::mutex> batch_lock(batch_mutex_);  // one batch at a time
    {
      std::lock_guard<std::mutex> lock(mutex_);
      fn_ = &fn;
      next_.store(begin);
      end_ = end;
      step_ = step;
      device_ = device;
      rc_ = HG_OK;
      error_.clear();
      pending_ = size();
      ++generation_;
    }
    wake_.notify_all();
    std::unique_lock<std::mutex> lock(mutex_);
    done_.wait(lock, [this] { return pending_ == 0; });
    fn_ = nullptr;
    if (rc_ != HG_OK) hg::SetError(error_);
    return rc_;
  }

 private:
  WorkerPool() {
    // Default: three workers (two when the rank has fewer than three cores). A unit of work is a
    // whole group of pairs — one device-resident search and one refine launch — so a worker's host
    // time is small and a second and third one only have to cover its synchronisation gaps; more
    // concurrent groups make the GPU slower, not faster (measured on C5, 2 000 pairs: 1 worker 61 us
    // per pair, 2-3 workers 51-52, 4: 67, 8: 81, 16: 87).
    int ranks = 1;
    if (const char* env = std::getenv("LOCAL_WORLD_SIZE")) ranks = std::max(1, std::atoi(env));
    int n = static_cast<int>(std::thread::hardware_concurrency()) / ranks;
    n = n < 3 ? 2 : 3;
    if (const char* env = std::getenv("HG_CONSTRAINT_THREADS")) n = std::atoi(env);
    n = n < 1 ? 1 : (n > 64 ? 64 : n);
    for (int i = 0; i < n; ++i) {
      threads_.emplace_back([this] { Loop(); });
      threads_.back().detach();
    }
  }
  void Loop() {
    uint64_t seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> lock(mutex_);
        wake_.wait(lock, [&] { return generation_ != seen; });
        seen = generation_;
      }
      cudaSetDevice(device_);
      for (;;) {
        const int64_t i = next_.fetch_add(step_);
        if (i >= end_) break;
        if (rc_ != HG_OK) continue;  // drain after a failure
        const int rc = (*fn_)(i);
        if (rc != HG_OK) {
          std::lock_guard<std::mutex> lock(mutex_);
          if (rc_ == HG_OK) {
            rc_ = rc;
            error_ = hg_last_error();
          }
        }
      }
      std::lock_guard<std::mutex> lock(mutex_);
      if (--pending_ == 0) done_.notify_all();
    }
  }

  std::vector<std::thread> threads_;
  std::mutex batch_mutex_, mutex_;
  std::condition_variable wake_, done_;
  const std::function<int(int64_t)>* fn_ = nullptr;
  std::atomic<int64_t> next_{0};
  int64_t end_ = 0, step_ = 1;
  int device_ = 0;
  std::atomic<int> rc_{HG_OK};
  std::string error_;
  int pending_ = 0;
  uint64_t generation_ = 0;
};

// One (node, submap) pair: constraint_builder_3d.cc:210-272.
int ComputeConstraint(const hg_fcsm* const* matchers, const hg_grid* const* grids,
                      const hg_grid* const* low_resolution_grids, const hg_constraint_job& job,
                      const hg_constraint_options* options, const int64_t j, hg_constraint_result* result) {
  hg_constraint_result& out = *result;
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
    rc = hg_fcsm_match(matchers[job.submap_index], job.global_node_pose, job.global_submap_pose, &job.data,
                       options->min_score, &match);
  }
  if (rc != HG_OK) return rc;
  out.num_candidates_scored = match.num_candidates_scored;
  const auto t1 = std::chrono::steady_clock::now();
  if (!match.found) return HG_OK;
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
                     options->max_num_iterations, options->use_nonmonotonic_steps, initial, initial, out.pose,
                     nullptr);
  if (rc != HG_OK) return rc;
  out.found = 1;
  static const bool trace = std::getenv("HG_TRACE") != nullptr;
  if (trace) {
    const auto t2 = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[hg trace] constraint job %lld: match %.2f ms, refine %.2f ms\n",
                 static_cast<long long>(j), std::chrono::duration<double, std::milli>(t1 - t0).count(),
                 std::chrono::duration<double, std::milli>(t2 - t1).count());
  }
  return HG_OK;
}

// All the local-window pairs of one submap (a slice of them): ONE device-resident search for the
// whole group (hg_fcsm_match_batch), then ONE launch that refines every match found
// (ScanMatchBatchOnDevice, a CTA per pair) — constraint_builder_3d.cc:210-272 for each pair, without
// a host thread per pair driving its own chain of small launches.
int ComputeConstraintGroup(const hg_fcsm* const* matchers, const hg_grid* const* grids,
                           const hg_grid* const* low_resolution_grids, const hg_constraint_job* jobs,
                           const hg_constraint_options* options, const std::vector<int64_t>& group,
                           hg_constraint_result* results) {
  const int n = static_cast<int>(group.size());
  if (n == 0) return HG_OK;
  const int submap = jobs[group[0]].submap_index;
  std::vector<double> node_poses(7 * static_cast<size_t>(n)), submap_poses(7 * static_cast<size_t>(n));
  std::vector<hg_node_data> data(n);
  std::vector<hg_match_result> matches(n);
  for (int k = 0; k < n; ++k) {
    const hg_constraint_job& job = jobs[group[k]];
    std::copy(job.global_node_pose, job.global_node_pose + 7, &node_poses[7 * k]);
    std::copy(job.global_submap_pose, job.global_submap_pose + 7, &submap_poses[7 * k]);
    data[k] = job.data;
  }
  const auto t0 = std::chrono::steady_clock::now();
  int rc = hg_fcsm_match_batch(matchers[submap], n, node_poses.data(), submap_poses.data(), data.data(),
                               options->min_score, matches.data());
  if (rc != HG_OK) return rc;
  const auto t1 = std::chrono::steady_clock::now();
  std::vector<hg::ScanMatchBatchItem> items;
  std::vector<int> item_of;
  for (int k = 0; k < n; ++k) {
    hg_constraint_result& out = results[group[k]];
    out = hg_constraint_result{};
    out.pose[3] = 1.0;
    out.num_candidates_scored = matches[k].num_candidates_scored;
    if (!matches[k].found) continue;
    out.score = matches[k].score;
    out.rotational_score = matches[k].rotational_score;
    out.low_resolution_score = matches[k].low_resolution_score;
    const hg_constraint_job& job = jobs[group[k]];
    hg::ScanMatchBatchItem it{};
    it.num_clouds = job.data.num_low_resolution_points > 0 ? 2 : 1;
    it.clouds[0] = job.data.high_resolution_point_cloud;
    it.clouds[1] = job.data.low_resolution_point_cloud;
    it.sizes[0] = job.data.num_high_resolution_points;
    it.sizes[1] = job.data.num_low_resolution_points;
    it.grids[0] = grids[submap];
    it.grids[1] = low_resolution_grids[submap];
    for (int c = 0; c < 3; ++c) it.target_translation[c] = it.initial_pose[c] = matches[k].pose_translation[c];
    for (int c = 0; c < 4; ++c) it.initial_pose[3 + c] = matches[k].pose_rotation[c];
    items.push_back(it);
    item_of.push_back(k);
  }
  if (items.empty()) return HG_OK;
  std::vector<double> poses(7 * items.size());
  rc = hg::ScanMatchBatchOnDevice(static_cast<int>(items.size()), items.data(), options->occupied_space_weights,
                                  options->translation_weight, options->rotation_weight, options->only_optimize_yaw,
                                  options->max_num_iterations, options->use_nonmonotonic_steps, poses.data(), nullptr);
  if (rc == HG_ERR_LIMIT) {  // a cloud too large for the one-CTA solve: pair by pair
    for (size_t i = 0; i < items.size(); ++i) {
      const hg::ScanMatchBatchItem& it = items[i];
      rc = hg_scan_match(it.num_clouds, it.clouds, it.sizes, it.grids, options->occupied_space_weights,
                         options->translation_weight, options->rotation_weight, options->only_optimize_yaw,
                         options->max_num_iterations, options->use_nonmonotonic_steps, it.target_translation,
                         it.initial_pose, &poses[7 * i], nullptr);
      if (rc != HG_OK) return rc;
    }
  } else if (rc != HG_OK) {
    return rc;
  }
  for (size_t i = 0; i < items.size(); ++i) {
    hg_constraint_result& out = results[group[item_of[i]]];
    std::copy(&poses[7 * i], &poses[7 * i] + 7, out.pose);
    out.found = 1;
  }
  static const bool trace = std::getenv("HG_TRACE") != nullptr;
  if (trace) {
    const auto t2 = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[hg trace] constraint group of %d pairs (submap %d): match %.2f ms, refine of %zu %.2f ms\n", n,
                 submap, std::chrono::duration<double, std::milli>(t1 - t0).count(), items.size(),
                 std::chrono::duration<double, std::milli>(t2 - t1).count());
  }
  return HG_OK;
}

}  // namespace

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
    HG_REQUIRE(jobs[j].submap_index >= 0 && jobs[j].submap_index < num_submaps,
               "hg_constraints_run_batch: submap index out of range");
  }
  // Local-window pairs are grouped by submap and searched / refined a group at a time; a
  // full-submap pair is a search of its own. HG_CONSTRAINT_GROUP bounds the group size (1 = the
  // pair-at-a-time path of round 1).
  static const int group_limit = [] {
    const char* env = std::getenv("HG_CONSTRAINT_GROUP");
    return env != nullptr ? std::max(1, std::atoi(env)) : 128;
  }();
  std::vector<std::vector<int64_t>> units;
  {
    std::vector<std::vector<int64_t>> open(num_submaps);
    for (int64_t j = shard_index; j < num_jobs; j += num_shards) {
      if (jobs[j].match_full_submap || group_limit == 1) {
        units.push_back({j});
        continue;
      }
      std::vector<int64_t>& g = open[jobs[j].submap_index];
      g.push_back(j);
      if (static_cast<int>(g.size()) >= group_limit) {
        units.push_back(std::move(g));
        g.clear();
      }
    }
    for (std::vector<int64_t>& g : open)
      if (!g.empty()) units.push_back(std::move(g));
  }
  const auto one = [&](const int64_t u) {
    const std::vector<int64_t>& unit = units[u];
    if (unit.size() == 1 && (jobs[unit[0]].match_full_submap || group_limit == 1)) {
      return ComputeConstraint(matchers, grids, low_resolution_grids, jobs[unit[0]], options, unit[0], &results[unit[0]]);
    }
    return ComputeConstraintGroup(matchers, grids, low_resolution_grids, jobs, options, unit, results);
  };
  const int64_t num_units = static_cast<int64_t>(units.size());
  static const bool serial = [] {
    const char* env = std::getenv("HG_CONSTRAINT_THREADS");
    return env != nullptr && std::atoi(env) <= 1;
  }();
  if (serial || num_units <= 1) {
    for (int64_t u = 0; u < num_units; ++u) {
      const int rc = one(u);
      if (rc != HG_OK) return rc;
    }
    return HG_OK;
  }
  int device = 0;
  HG_CUDA(cudaGetDevice(&device));
  // Everything the calling thread has queued (grid uploads, pyramids) must be visible to the workers.
  HG_CUDA(cudaDeviceSynchronize());
  return WorkerPool::Get().Run(0, num_units, 1, device, one);
}

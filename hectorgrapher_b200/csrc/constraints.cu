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
// of pairs in flight (default: host cores per rank, clamped to 4..16; 1 = compute on the
// calling thread).
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
    // Default: the host cores this process can count on (one process per GPU: the cores are
    // shared by LOCAL_WORLD_SIZE ranks), between 4 and 16 — the workers mostly wait on their
    // streams, and 16 in flight is where a B200 stops gaining on the C5 workload.
    int ranks = 1;
    if (const char* env = std::getenv("LOCAL_WORLD_SIZE")) ranks = std::max(1, std::atoi(env));
    int n = static_cast<int>(std::thread::hardware_concurrency()) / ranks;
    n = n < 4 ? 4 : (n > 16 ? 16 : n);
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
  const auto one = [&](const int64_t j) {
    return ComputeConstraint(matchers, grids, low_resolution_grids, jobs[j], options, j, &results[j]);
  };
  const int64_t mine = (num_jobs - shard_index + num_shards - 1) / num_shards;
  static const bool serial = [] {
    const char* env = std::getenv("HG_CONSTRAINT_THREADS");
    return env != nullptr && std::atoi(env) <= 1;
  }();
  if (serial || mine <= 1) {
    for (int64_t j = shard_index; j < num_jobs; j += num_shards) {
      const int rc = one(j);
      if (rc != HG_OK) return rc;
    }
    return HG_OK;
  }
  int device = 0;
  HG_CUDA(cudaGetDevice(&device));
  // Everything the calling thread has queued (grid uploads, pyramids) must be visible to the workers.
  HG_CUDA(cudaDeviceSynchronize());
  return WorkerPool::Get().Run(shard_index, num_jobs, num_shards, device, one);
}

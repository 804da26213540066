#include "LoadBalancer.hpp"

#include <algorithm>
#include <random>
#include <stdexcept>

#include "logging.hpp"

RecordList createWorkListForImage(std::size_t imageWidth, std::size_t imageHeight) {
  RecordList workList;
  workList.reserve(imageWidth * imageHeight);
  for (std::size_t r = 0; r < imageHeight; ++r)
    for (std::size_t c = 0; c < imageWidth; ++c)
      workList.push_back(TraceRecord{(std::uint16_t)c, (std::uint16_t)r, 0.f, 0.f, 0.f, 0, 0});
  return workList;
}

void padWorkList(RecordList& list, std::size_t multiple) {
  if (multiple == 0) return;
  while (list.size() % multiple) list.push_back(TraceRecord{kPadCoord, kPadCoord, 0.f, 0.f, 0.f, 0, 0});
}

std::vector<RecordList> createTracingJobs(std::size_t imageWidth, std::size_t imageHeight, std::size_t numGpus, std::size_t tile,
                                          unsigned seed) {
  if (numGpus < 1) throw std::logic_error("createTracingJobs: need at least one GPU");
  if (tile < 1) throw std::logic_error("createTracingJobs: tile size must be positive");
  const std::size_t tx = (imageWidth + tile - 1) / tile, ty = (imageHeight + tile - 1) / tile;
  std::vector<std::uint32_t> order(tx * ty);
  for (std::size_t i = 0; i < order.size(); ++i) order[i] = (std::uint32_t)i;
  std::mt19937 g(seed);
  std::shuffle(order.begin(), order.end(), g);
  std::vector<RecordList> jobs(numGpus);
  for (std::size_t i = 0; i < order.size(); ++i) {
    RecordList& job = jobs[i % numGpus];
    const std::size_t x0 = (order[i] % tx) * tile, y0 = (order[i] / tx) * tile;
    for (std::size_t y = y0; y < std::min(y0 + tile, imageHeight); ++y)
      for (std::size_t x = x0; x < std::min(x0 + tile, imageWidth); ++x)
        job.push_back(TraceRecord{(std::uint16_t)x, (std::uint16_t)y, 0.f, 0.f, 0.f, 0, 0});
  }
  std::size_t longest = 0;
  for (auto& j : jobs) longest = std::max(longest, j.size());
  longest = std::max<std::size_t>(longest, 1);
  for (auto& j : jobs)
    while (j.size() < longest) j.push_back(TraceRecord{kPadCoord, kPadCoord, 0.f, 0.f, 0.f, 0, 0});
  return jobs;
}

void WorkList::swap() {
  std::swap(activeWork, inactiveWork);
  if (activeWork.empty()) throw std::logic_error("The new active worklist is empty.");
}

LoadBalancer::LoadBalancer(const std::vector<RecordList>& jobs)
    : work(jobs.empty() ? 0 : jobs.size() * jobs.front().size()), numJobs(jobs.size()), itemsPerJob(jobs.empty() ? 0 : jobs.front().size()) {
  std::size_t k = 0;
  for (const auto& j : jobs) {
    if (j.size() != itemsPerJob) throw std::logic_error("LoadBalancer: all jobs must have the same size");
    for (const auto& w : j) work.inactive()[k++] = w;
  }
  work.active() = work.inactive();
}

void LoadBalancer::allocateWorkByPathLength() {
  auto sorted = work.inactive();
  if (sorted.empty()) return;
  std::stable_sort(sorted.begin(), sorted.end(), [](const TraceRecord& a, const TraceRecord& b) { return a.pathLength < b.pathLength; });
  logging::info("Load balancing started ({} work items)", sorted.size());
  logging::info("Path length min/max: {}/{}", sorted.front().pathLength, sorted.back().pathLength);
  std::vector<RecordList> perJob(numJobs);
  for (auto& t : perJob) t.reserve(itemsPerJob + 2);
  // One item from each end of the sorted list per job in turn: the job that takes the longest path of
  // the last step also takes the shortest (the reference reads past the middle when the count is odd;
  // the indices are checked here instead).
  std::size_t lo = 0, hi = sorted.size();
  while (lo < hi) {
    for (auto& t : perJob) {
      if (lo >= hi) break;
      t.push_back(sorted[lo++]);
      if (lo >= hi) break;
      t.push_back(sorted[--hi]);
    }
  }
  logging::info("Load balancing finished");
  // Jobs must keep equal sizes (one device buffer each): hand the overflow of early jobs to later ones.
  RecordList flat;
  flat.reserve(sorted.size());
  for (auto& t : perJob) for (auto& w : t) flat.push_back(w);
  work.inactive() = flat;
}

std::size_t LoadBalancer::clearInactiveAccumulators() {
  auto& list = work.inactive();
  std::size_t sum = 0;
#pragma omp parallel for reduction(+ : sum) schedule(static)
  for (std::size_t i = 0; i < list.size(); ++i) {
    auto& t = list[i];
    sum += t.pathLength;
    t.r = t.g = t.b = 0.f;
    t.pathLength = 0;
    t.sampleCount = 0;
  }
  return sum;
}

void LoadBalancer::clearActiveAccumulators() {
  auto& list = work.active();
#pragma omp parallel for schedule(static)
  for (std::size_t i = 0; i < list.size(); ++i) {
    auto& t = list[i];
    t.r = t.g = t.b = 0.f;
    t.pathLength = 0;
    t.sampleCount = 0;
  }
}

std::uint64_t LoadBalancer::activeOrderDigest() {
  std::uint64_t h = 1469598103934665603ull;
  for (const auto& t : work.active()) {
    h = (h ^ t.u) * 1099511628211ull;
    h = (h ^ t.v) * 1099511628211ull;
  }
  return h;
}

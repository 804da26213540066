// Work lists and their distribution: the reference's LoadBalancer (src/LoadBalancer.hpp:15-56,
// src/LoadBalancer.cpp) re-read for GPUs. The reference deals equal chunks of a shuffled per-pixel
// list to IPU tiles; here the unit dealt to a GPU is a square pixel tile (coherent warps, even mix of
// sky and geometry per GPU), shuffled with the reference's seed 142 (LoadBalancer.cpp:133-135).
#pragma once

#include <cstddef>
#include <cstdint>
#include <vector>

#include "../include/b200pt.h"

using TraceRecord = b200pt_trace_record;      // byte-compatible with src/codelets/TraceRecord.hpp:7-19
using RecordList = std::vector<TraceRecord>;

constexpr std::uint16_t kPadCoord = 65535;    // LoadBalancer.cpp:66-71

/// Row-major list of every pixel (LoadBalancer.cpp:38-52).
RecordList createWorkListForImage(std::size_t imageWidth, std::size_t imageHeight);

/// Pads with null work (u = v = 65535) up to a multiple (LoadBalancer.cpp:62-71).
void padWorkList(RecordList& list, std::size_t multiple);

/// One work list per GPU: tile x tile pixel tiles dealt round-robin after a seeded shuffle; together
/// the lists cover every pixel exactly once. All lists are padded to the same length
/// (createTracingJobs, LoadBalancer.cpp:54-86: "all tiles have same worklist size").
std::vector<RecordList> createTracingJobs(std::size_t imageWidth, std::size_t imageHeight, std::size_t numGpus,
                                          std::size_t tile = 16, unsigned seed = 142u);

/// A double buffered work list (LoadBalancer.hpp:27-42).
struct WorkList {
  explicit WorkList(std::size_t size) : activeWork(size), inactiveWork(size) {}
  void swap();
  RecordList& active() { return activeWork; }
  RecordList& inactive() { return inactiveWork; }

private:
  RecordList activeWork;
  RecordList inactiveWork;
};

struct LoadBalancer {
  /// jobs: one list per GPU, all of the same length; the flattened list is job-major.
  explicit LoadBalancer(const std::vector<RecordList>& jobs);

  WorkList& getWork() { return work; }
  std::size_t jobCount() const { return numJobs; }
  std::size_t jobSize() const { return itemsPerJob; }
  /// The slice of a flattened list that belongs to GPU g.
  TraceRecord* jobSlice(RecordList& list, std::size_t g) { return list.data() + g * itemsPerJob; }

  /// Sort the inactive list by last step's path length and deal one cheap and one expensive item to
  /// every job in turn (LoadBalancer.cpp:141-192). Padding items sort first (path length 0).
  void allocateWorkByPathLength();
  /// Zero the accumulators of the inactive list, return the path-length sum ("rays", :198-214).
  std::size_t clearInactiveAccumulators();
  void clearActiveAccumulators();
  /// FNV-1a over the (u, v) sequence of the active list: which pixel sits in which slot of which job.
  std::uint64_t activeOrderDigest();

private:
  WorkList work;
  std::size_t numJobs, itemsPerJob;
};

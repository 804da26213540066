// The render application: the reference's PathTracerApp (src/PathTracerApp.hpp:32-122,
// src/PathTracerApp.cpp:47-123,484-792) with the Poplar graph/engine replaced by libb200pt contexts,
// one per GPU, driven through the C ABI's five programs. Same command line, same step loop
// (setup / path_trace / read_results with host film processing overlapped on an AsyncTask thread),
// same outputs (tone-mapped PNG + EXR of hdr/step at every save interval), same log lines.
#pragma once

#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "AccumulatedImage.hpp"
#include "InterfaceServer.hpp"
#include "LoadBalancer.hpp"
#include "NifAssets.hpp"
#include "Options.hpp"

struct RenderSummary {
  double elapsedSecs = 0.0;
  double samplesPerSec = 0.0;
  std::size_t steps = 0;
  std::size_t samplesPerPixel = 0;
  std::size_t restarts = 0;            // renders restarted by the remote UI
  std::uint64_t paths = 0, escaped = 0, rays = 0;
  double traceMs = 0.0, nifMs = 0.0, deviceMs = 0.0;
  std::string lastError;
};

/// Adjust samples per pixel to be a multiple of samples per step (PathTracerApp.cpp:19-27).
std::size_t roundSamplesPerPixel(std::size_t samplesPerPixel, std::size_t samplesPerStep);

class PathTracerApp {
public:
  PathTracerApp();
  ~PathTracerApp();
  /// Reads the options and the NIF assets (PathTracerApp.cpp:62-72). Throws std::runtime_error.
  void init(const Options& options);
  /// Creates the GPU contexts, uploads the NIF, renders, saves (execute(), :566-792).
  RenderSummary execute();

private:
  struct Gpu;                         // one context (+ gather staging) per device
  const Options* args = nullptr;
  std::size_t samplesPerPixel = 0;
  std::uint32_t samplesPerStep = 0;
  std::unique_ptr<NifModelData> nif;
  std::vector<std::unique_ptr<Gpu>> gpus;
  struct FilmGather;                  // NCCL communicator group for the per-save-interval film gather
  std::unique_ptr<FilmGather> gather;

  bool loadNifModels(const std::string& assetPath);
  void createContexts(std::uint32_t width, std::uint32_t height, std::size_t itemsPerGpu);
  void uploadNifAll();
  void check(int rc, Gpu& g, const char* what);
  std::vector<int> deviceList() const;
  void initFilmGatherComms();
  void initFilmGather(std::size_t floats);
  void gatherDeviceFilm(AccumulatedImage& film);
  /// Stop / detach / (reload NIF and) restart decision for a changed UI state (PathTracerApp.cpp:531-564).
  InterfaceServer::Status processUserInput(InterfaceServer::State& state);
};

#include "PathTracerApp.hpp"

#include <cuda_runtime.h>
#include <nccl.h>

#include <nvtx3/nvToolsExt.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <sstream>
#include <stdexcept>
#include <thread>

#include "AsyncTask.hpp"
#include "logging.hpp"

namespace {
// NVTX ranges carrying the names of the reference's PVTI tracepoints (PathTracerApp.cpp:313-780), so that an nsys
// timeline of this host reads like a PopVision trace of the reference. No-ops unless a tool is attached.
struct Trace {
  explicit Trace(const char* name) { nvtxRangePushA(name); }
  ~Trace() { nvtxRangePop(); }
};
} // namespace

std::size_t roundSamplesPerPixel(std::size_t samplesPerPixel, std::size_t samplesPerStep) {
  if (samplesPerPixel % samplesPerStep) {
    samplesPerPixel += samplesPerStep - (samplesPerPixel % samplesPerStep);
    logging::info("Rounding SPP to next multiple of {}  (Rounded SPP :=  {})", samplesPerStep, samplesPerPixel);
  }
  return samplesPerPixel;
}

struct PathTracerApp::Gpu {
  int device = 0;
  b200pt_ctx* ctx = nullptr;
  float* staging = nullptr;          // full-frame film copy for the gather (device memory)
  float* reduced = nullptr;          // gather result (root only)
  cudaStream_t stream = nullptr;
  ~Gpu() {
    if (ctx) b200pt_destroy(ctx);
    cudaSetDevice(device);
    if (staging) cudaFree(staging);
    if (reduced) cudaFree(reduced);
    if (stream) cudaStreamDestroy(stream);
  }
};

struct PathTracerApp::FilmGather {
  std::vector<ncclComm_t> comms;
  ~FilmGather() { for (auto c : comms) if (c) ncclCommDestroy(c); }
};

PathTracerApp::PathTracerApp() {}
PathTracerApp::~PathTracerApp() { gather.reset(); gpus.clear(); }

void PathTracerApp::check(int rc, Gpu& g, const char* what) {
  if (rc == B200PT_OK) return;
  std::stringstream ss;
  ss << what << " failed on GPU " << g.device << " (" << rc << "): " << b200pt_last_error(g.ctx);
  throw std::runtime_error(ss.str());
}

void PathTracerApp::init(const Options& options) {
  args = &options;
  samplesPerPixel = options.u32("samples");
  samplesPerStep = options.u32("samples-per-step");
  if (samplesPerStep == 0) throw std::runtime_error("samples-per-step must be positive");
  samplesPerPixel = roundSamplesPerPixel(samplesPerPixel, samplesPerStep);
  for (const char* ignored : {"model", "save-exe", "load-exe", "compile-only", "defer-attach", "codelet-path", "partials-type", "available-memory-proportion"})
    if (options.given(ignored)) logging::warn("Option --{} configures the Poplar graph of the reference and has no effect here.", ignored);
  if (!loadNifModels(options.str("assets"))) throw std::runtime_error("Could not load NIF model.");
}

bool PathTracerApp::loadNifModels(const std::string& assetPath) {
  try {
    nif.reset(new NifModelData(assetPath, args->flag("synthetic-weights")));
  } catch (std::exception& e) {
    logging::error("Could not load NIF model from '{}'. Exception: {}", assetPath, e.what());
    return false;
  }
  return true;
}

std::vector<int> PathTracerApp::deviceList() const {
  const std::size_t numGpus = args->u64("ipus");
  std::vector<int> devices;
  if (args->given("devices")) {
    std::stringstream ss(args->str("devices"));
    std::string tok;
    while (std::getline(ss, tok, ',')) devices.push_back(std::atoi(tok.c_str()));
    if (devices.size() != numGpus) throw std::runtime_error("--devices must list exactly --ipus ordinals");
  } else {
    for (std::size_t g = 0; g < numGpus; ++g) devices.push_back((int)g);
  }
  return devices;
}

void PathTracerApp::createContexts(std::uint32_t width, std::uint32_t height, std::size_t itemsPerGpu) {
  const std::size_t numGpus = args->u64("ipus");
  const std::vector<int> devices = deviceList();
  const std::string aaName = args->str("aa-noise-type");
  int aaType;
  if (aaName == "normal") aaType = B200PT_AA_NORMAL;
  else if (aaName == "uniform") aaType = B200PT_AA_UNIFORM;
  else if (aaName == "truncated-normal") aaType = B200PT_AA_TRUNCATED_NORMAL;
  else throw std::runtime_error("Invalid AA noise type: " + aaName);       // PathTracerApp.cpp:43-45

  // --max-nif-batch-size bounded the serialised NIF batch on the IPU; here it sizes the escaped-ray
  // wavefront: whole samples of every pixel of the GPU per launch (0 = let the library choose).
  std::uint32_t samplesPerLaunch = 0;
  if (args->given("max-nif-batch-size") && args->u64("max-nif-batch-size") > 0)
    samplesPerLaunch = (std::uint32_t)std::max<std::uint64_t>(1, args->u64("max-nif-batch-size") / std::max<std::size_t>(1, itemsPerGpu));

  // One thread per GPU: context creation (CUDA context, kernels, 0.3-1 GB of device buffers) takes ~1 s each and used to
  // run one GPU after the other.
  gpus.resize(numGpus);
  std::vector<std::string> errors(numGpus);
  std::vector<std::thread> workers;
  for (std::size_t g = 0; g < numGpus; ++g) {
    workers.emplace_back([&, g]() {
      Trace t("create_context");
      std::unique_ptr<Gpu> gpu(new Gpu());
      gpu->device = devices[g];
      b200pt_config cfg{};
      cfg.struct_size = sizeof(cfg);
      cfg.device = devices[g];
      cfg.width = width; cfg.height = height;
      cfg.max_path_length = args->u32("max-path-length");
      cfg.roulette_depth = args->u32("roulette-depth");
      cfg.stop_prob = args->f32("stop-prob");
      cfg.refractive_index = args->f32("refractive-index");
      cfg.aa_noise_type = aaType;
      cfg.max_worklist = (std::uint32_t)itemsPerGpu;
      cfg.samples_per_launch = samplesPerLaunch;
      cfg.nif_impl = B200PT_NIF_AUTO;
      std::stringstream ss;
      int rc = b200pt_create(&cfg, &gpu->ctx);
      if (rc != B200PT_OK) {
        ss << "Could not create a render context on GPU " << devices[g] << " (" << rc << "): " << b200pt_last_error(nullptr);
        errors[g] = ss.str();
        return;
      }
      gpus[g] = std::move(gpu);
    });
  }
  for (auto& w : workers) w.join();
  for (auto& e : errors) if (!e.empty()) { gpus.clear(); throw std::runtime_error(e); }
  logging::info("Created {} GPU context(s): {}", gpus.size(), b200pt_version());
}

// init_nif_weights on every GPU (PathTracerApp.cpp:610), one thread per GPU (packing the weight stream and building the
// 3 MB feature table is host work).
void PathTracerApp::uploadNifAll() {
  Trace t("init_nif_weights");
  const b200pt_nif_desc desc = nif->descriptor();
  std::vector<std::string> errors(gpus.size());
  std::vector<std::thread> workers;
  for (std::size_t g = 0; g < gpus.size(); ++g)
    workers.emplace_back([&, g]() {
      const int rc = b200pt_upload_nif(gpus[g]->ctx, &desc);
      if (rc != B200PT_OK) {
        std::stringstream ss;
        ss << "init_nif_weights failed on GPU " << gpus[g]->device << " (" << rc << "): " << b200pt_last_error(gpus[g]->ctx);
        errors[g] = ss.str();
      }
    });
  for (auto& w : workers) w.join();
  for (auto& e : errors) if (!e.empty()) throw std::runtime_error(e);
}

// Communicators and staging buffers of the film gather; created before the render starts so that the first save
// interval does not pay for ncclCommInitAll.
// ncclCommInitAll takes seconds for 8 GPUs; it needs only the device ordinals, so execute() runs it on its own thread
// beside context creation and the NIF upload.
void PathTracerApp::initFilmGatherComms() {
  Trace t("nccl_comm_init");
  std::unique_ptr<FilmGather> fg(new FilmGather());
  std::vector<int> devs = deviceList();
  fg->comms.assign(devs.size(), nullptr);
  const ncclResult_t r = ncclCommInitAll(fg->comms.data(), (int)devs.size(), devs.data());
  if (r != ncclSuccess) throw std::runtime_error(std::string("ncclCommInitAll: ") + ncclGetErrorString(r));
  // The first collective of a communicator sets its channels up (tenths of a second): do it here, on a scratch word per
  // GPU, so that the first save interval of the render does not pay for it.
  std::vector<float*> word(devs.size(), nullptr);
  bool ok = true;
  for (std::size_t i = 0; i < devs.size(); ++i) ok = ok && cudaSetDevice(devs[i]) == cudaSuccess && cudaMalloc(&word[i], 256) == cudaSuccess && cudaMemset(word[i], 0, 256) == cudaSuccess;
  if (ok) {
    ncclGroupStart();
    for (std::size_t i = 0; i < devs.size(); ++i) { cudaSetDevice(devs[i]); ncclReduce(word[i], word[i], 64, ncclFloat, ncclSum, 0, fg->comms[i], nullptr); }
    ncclGroupEnd();
    for (std::size_t i = 0; i < devs.size(); ++i) { cudaSetDevice(devs[i]); cudaStreamSynchronize(nullptr); }
  }
  for (std::size_t i = 0; i < devs.size(); ++i) if (word[i]) { cudaSetDevice(devs[i]); cudaFree(word[i]); }
  gather = std::move(fg);
}

void PathTracerApp::initFilmGather(std::size_t n) {
  auto cuda = [](cudaError_t e, const char* what) { if (e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e)); };
  if (!gather) initFilmGatherComms();
  for (auto& g : gpus) {
    cuda(cudaSetDevice(g->device), "cudaSetDevice");
    cuda(cudaMalloc(&g->staging, n * sizeof(float)), "cudaMalloc");
    cuda(cudaStreamCreateWithFlags(&g->stream, cudaStreamNonBlocking), "cudaStreamCreate");
  }
  cuda(cudaSetDevice(gpus[0]->device), "cudaSetDevice");
  cuda(cudaMalloc(&gpus[0]->reduced, n * sizeof(float)), "cudaMalloc");
}

// One gather of the per-GPU films per save interval. Every GPU holds a full-frame film that is zero
// outside its own pixels, so the gather is an exact sum: ncclReduce to GPU 0 over NVLink, one D->H copy.
void PathTracerApp::gatherDeviceFilm(AccumulatedImage& film) {
  const std::size_t n = film.width() * film.height() * 3;
  if (gpus.size() == 1) {
    check(b200pt_film_read(gpus[0]->ctx, film.hdrData(), 1.f), *gpus[0], "film_read");     // straight into the (page-locked) host film
    return;
  }
  auto cuda = [](cudaError_t e, const char* what) { if (e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e)); };
  auto nccl = [](ncclResult_t r, const char* what) { if (r != ncclSuccess) throw std::runtime_error(std::string(what) + ": " + ncclGetErrorString(r)); };
  if (!gather || !gpus[0]->reduced) initFilmGather(n);
  for (auto& g : gpus) check(b200pt_film_read_device(g->ctx, g->staging, 1.f), *g, "film_read_device");   // synchronises the context
  nccl(ncclGroupStart(), "ncclGroupStart");
  for (std::size_t i = 0; i < gpus.size(); ++i)
    nccl(ncclReduce(gpus[i]->staging, gpus[0]->reduced, n, ncclFloat, ncclSum, 0, gather->comms[i], gpus[i]->stream), "ncclReduce");
  nccl(ncclGroupEnd(), "ncclGroupEnd");
  for (auto& g : gpus) { cuda(cudaSetDevice(g->device), "cudaSetDevice"); cuda(cudaStreamSynchronize(g->stream), "cudaStreamSynchronize"); }
  cuda(cudaSetDevice(gpus[0]->device), "cudaSetDevice");
  cuda(cudaMemcpy(film.hdrData(), gpus[0]->reduced, n * sizeof(float), cudaMemcpyDeviceToHost), "cudaMemcpy");    // page-locked target
}

// The user interaction invalidates all rendering in progress (processUserInput + defunctState,
// PathTracerApp.cpp:510-564). The reference swaps in a fresh host state so that it need not wait for the
// defunct jobs; here a step is a few milliseconds, so the GPUs are simply drained before the reset.
InterfaceServer::Status PathTracerApp::processUserInput(InterfaceServer::State& state) {
  if (state.stop) {
    logging::info("Rendering stopped by remote UI");
    return InterfaceServer::Status::Stop;
  }
  if (state.detach) {
    // If the Remote-UI detaches just continue rendering:
    logging::info("Remote UI disconnected.");
    return InterfaceServer::Status::Disconnected;
  }
  if (!state.newNif.empty()) {
    // Load of a new NIF was requested:
    logging::info("Loading NIF: {}", state.newNif);
    if (loadNifModels(state.newNif)) {
      b200pt_nif_desc desc = nif->descriptor();
      for (auto& g : gpus) {
        check(b200pt_sync(g->ctx), *g, "sync");
        check(b200pt_upload_nif(g->ctx, &desc), *g, "init_nif_weights");
      }
    }
  }
  return InterfaceServer::Status::Restart;
}

RenderSummary PathTracerApp::execute() {
  const auto imageWidth = args->u32("width");
  const auto imageHeight = args->u32("height");
  const auto seed = args->u64("seed");
  const auto antiAliasingScale = args->f32("aa-noise-scale");
  const float fieldOfView = args->f32("fov") * (float)(M_PI / 180.0);
  const auto fileName = args->str("outfile");
  const auto loadBalanceEnabled = args->flag("enable-load-balancing");
  const auto saveInterval = std::max<std::uint32_t>(1, args->u32("save-interval"));
  const bool hostFilm = args->flag("host-film");
  const auto steps = samplesPerPixel / samplesPerStep;
  const float degrees = args->f32("env-map-rotation");
  if (imageWidth == 0 || imageHeight == 0 || imageWidth > 65535 || imageHeight > 65535)
    throw std::runtime_error("Image width and height must be in 1..65535 (TraceRecord coordinates are 16 bit).");
  if (loadBalanceEnabled && !hostFilm) logging::warn("--enable-load-balancing re-deals work items between steps and needs the per-step record stream: enabling --host-film.");
  const bool recordStream = hostFilm || loadBalanceEnabled;

  // ---- device acquisition: what GraphManager::run does before it calls execute() in the reference (attach to the devices,
  // load the executable onto them: ipu_utils.hpp:473-550) - CUDA contexts with their kernels and buffers, one thread per GPU,
  // and the NCCL communicators of the film gather beside them. Like there, it is outside the run-level clock.
  const auto acquireStart = std::chrono::steady_clock::now();
  nvtxRangePushA("device_acquisition");
  std::thread ncclThread;
  std::exception_ptr ncclError;
  if (!recordStream && args->u64("ipus") > 1)
    ncclThread = std::thread([&]() { try { initFilmGatherComms(); } catch (...) { ncclError = std::current_exception(); } });
  struct Joiner { std::thread& t; ~Joiner() { if (t.joinable()) t.join(); } } ncclJoiner{ncclThread};
  // Two host films: the one being encoded by the save task and the one the next gather writes into; page-locked so that
  // the gather's device-to-host copy runs at the full link rate straight into them.
  AccumulatedImage films[2] = {AccumulatedImage(imageWidth, imageHeight), AccumulatedImage(imageWidth, imageHeight)};
  AccumulatedImage& film = films[0];
  std::size_t saveBuffer = 0, skippedSaves = 0;
  const std::size_t jobSize = createTracingJobs(imageWidth, imageHeight, args->u64("ipus")).front().size();
  createContexts(imageWidth, imageHeight, jobSize);
  if (ncclThread.joinable()) ncclThread.join();
  if (ncclError) std::rethrow_exception(ncclError);
  if (!recordStream && gpus.size() > 1) initFilmGather((std::size_t)imageWidth * imageHeight * 3);
  struct HostPin {
    std::vector<float*> p;
    ~HostPin() { for (auto* q : p) cudaHostUnregister(q); }
  } hostPin;
  // This thread has not touched CUDA yet (contexts are created on worker threads): select the first GPU of the device
  // list before registering host memory, or the runtime would bring up a context on GPU 0 even when --devices leaves it out.
  if (const cudaError_t e = cudaSetDevice(gpus[0]->device); e != cudaSuccess) throw std::runtime_error(std::string("cudaSetDevice: ") + cudaGetErrorString(e));
  bool pinFailureLogged = false;
  auto pinFailed = [&](const char* what, cudaError_t e) {
    cudaGetLastError();
    if (!pinFailureLogged) logging::warn("Could not page-lock {} ({}): host transfers take the slower pageable path", what, cudaGetErrorString(e));
    pinFailureLogged = true;
  };
  if (!recordStream) {
    for (auto& f : films) {
      const cudaError_t e = cudaHostRegister(f.hdrData(), f.width() * f.height() * 3 * sizeof(float), cudaHostRegisterPortable);
      if (e == cudaSuccess) hostPin.p.push_back(f.hdrData());
      else pinFailed("the host film", e);
    }
  }
  nvtxRangePop();
  logging::info("Device acquisition took {} seconds", std::chrono::duration<double>(std::chrono::steady_clock::now() - acquireStart).count());

  // ---- the reference's execute() starts its clock here (PathTracerApp.cpp:604-605), before init_nif_weights
  auto startTime = std::chrono::steady_clock::now();
  nvtxRangePushA("initialisation");
  uploadNifAll();
  // Build the tracing jobs (initialiseState, PathTracerApp.cpp:484-497): pixel tiles dealt to GPUs.
  nvtxRangePushA("create_path_tracing_jobs");
  auto jobs = createTracingJobs(imageWidth, imageHeight, args->u64("ipus"));
  nvtxRangePop();
  logging::info("Created worklists for {} GPUs ({} work items each)", jobs.size(), jobs.front().size());
  LoadBalancer work(jobs);
  // Record-stream modes move both work lists over the host link every step: page-lock them (the copies of setup /
  // read_results then run at the full link rate). Declared after `work` so that it unregisters before the lists are freed.
  struct ListPin {
    std::vector<void*> p;
    bool holds(const void* q) const { return std::find(p.begin(), p.end(), q) != p.end(); }
    ~ListPin() { for (auto* q : p) cudaHostUnregister(q); }
  } listPin;
  if (recordStream) {
    for (RecordList* l : {&work.getWork().active(), &work.getWork().inactive()}) {
      const cudaError_t e = cudaHostRegister(l->data(), l->size() * sizeof(TraceRecord), cudaHostRegisterPortable);
      if (e == cudaSuccess) listPin.p.push_back(l->data());
      else pinFailed("the work lists", e);
    }
  }
  // The registration is by address: swapping the two lists or copy-assigning an equal-sized list keeps the addresses, a
  // reallocation would leave registered memory freed and the live list pageable. Checked after every re-deal.
  auto listsStillPinned = [&]() {
    return listPin.p.empty() || (listPin.holds(work.getWork().active().data()) && listPin.holds(work.getWork().inactive().data()));
  };

  // Setup remote user interface (:618-633). Unlike the reference the state a client has not touched
  // yet starts from the command line, with or without a client.
  InterfaceServer::State state;
  state.exposure = args->f32("exposure");
  state.gamma = args->f32("gamma");
  state.fov = fieldOfView;
  state.envRotationDegrees = degrees;
  state.interactiveSamples = std::max<std::uint32_t>(1, args->u32("interactive-samples"));
  std::unique_ptr<InterfaceServer> uiServer;
  if (const int uiPort = (int)args->u32("ui-port")) {
    uiServer.reset(new InterfaceServer(uiPort));
    uiServer->setInitialState(state);
    uiServer->start((int)args->u32("ui-accept-timeout"));
    startTime = std::chrono::steady_clock::now();
  }
  constexpr std::size_t sampleCountReversionStep = 5;

  if (!recordStream)
    for (std::size_t g = 0; g < gpus.size(); ++g) check(b200pt_setup(gpus[g]->ctx, work.jobSlice(work.getWork().active(), g), work.jobSize()), *gpus[g], "setup");
  nvtxRangePop();      // initialisation
  logging::info("Initialisation took {} seconds", std::chrono::duration<double>(std::chrono::steady_clock::now() - startTime).count());
  nvtxRangePushA("rendering");

  AsyncTask hostProcessing;
  RenderSummary summary;
  summary.steps = steps;
  summary.samplesPerPixel = samplesPerPixel;
  logging::info("Render started");
  std::size_t totalRays = 0;

  for (std::size_t step = 1; step <= steps; ++step) {
    auto loopStartTime = std::chrono::steady_clock::now();

    // Do the simple thing and restart the entire render if any state changed (:651-675):
    if (uiServer && !uiServer->connected()) {
      logging::warn("Remote UI connection lost: rendering continues head-less.");
      hostProcessing.waitForCompletion();        // the async task of the last step may still be talking to the server
      uiServer.reset();
    }
    if (uiServer && uiServer->stateChanged()) {
      state = uiServer->consumeState();
      const auto status = processUserInput(state);
      if (status == InterfaceServer::Status::Stop) {
        hostProcessing.waitForCompletion();
        uiServer.reset();
        summary.steps = step - 1;
        break;
      }
      if (status == InterfaceServer::Status::Disconnected) {
        hostProcessing.waitForCompletion();
        uiServer.reset();
      } else if (status == InterfaceServer::Status::Restart) {
        // reset_host_render_state: drain, then clear the film and the work-item accumulators
        hostProcessing.waitForCompletion();
        for (auto& g : gpus) { check(b200pt_sync(g->ctx), *g, "sync"); check(b200pt_film_clear(g->ctx), *g, "film_clear"); }
        film.reset();
        if (recordStream) work.clearActiveAccumulators();
        summary.restarts += 1;
        startTime = loopStartTime;
        step = 1;
        samplesPerStep = state.interactiveSamples;
      }
    } else if (uiServer && step == sampleCountReversionStep && samplesPerStep != args->u32("samples-per-step")) {
      // No UI input for a few steps so revert to performant number of samples:
      samplesPerStep = args->u32("samples-per-step");
      logging::debug("Interaction stopped reverting samples per step to: {}", samplesPerStep);
    }

    // Render settings can only be updated on these steps (:677-686):
    if (step == 1 || (uiServer && step == sampleCountReversionStep)) {
      const float radians = (state.envRotationDegrees / 360.f) * (float)(2.0 * M_PI);
      for (auto& g : gpus) {
        check(b200pt_sync(g->ctx), *g, "sync");
        check(b200pt_set_render_settings(g->ctx, seed, antiAliasingScale, state.fov, radians, samplesPerStep), *g, "init_render_settings");
      }
    }

    // Run path tracing on the GPUs (setup / path_trace / read_results, :690-694). Results go into the
    // active buffer whilst the async host task processes the last result from the inactive buffer.
    nvtxRangePushA("ipu_render");
    if (recordStream && loadBalanceEnabled) logging::debug("Work list order digest at step {}: {}", step, work.activeOrderDigest());
    // One host thread per GPU for the step's programs: enqueueing a path_trace costs a few hundred microseconds of driver
    // calls, and reading a GPU's counters waits for that GPU - done from one thread, GPU 7 would start its step only after
    // the seven enqueues before it and every step would begin with idle devices. A context is still only ever used by one
    // thread at a time (the threads are joined before anything else touches it).
    std::vector<b200pt_counters> ctrs(gpus.size());
    auto renderOn = [&](std::size_t g) {
      Gpu& gpu = *gpus[g];
      if (recordStream) check(b200pt_setup(gpu.ctx, work.jobSlice(work.getWork().active(), g), work.jobSize()), gpu, "setup");
      check(b200pt_path_trace(gpu.ctx), gpu, "path_trace");                      // asynchronous
      if (recordStream) check(b200pt_read_results(gpu.ctx, work.jobSlice(work.getWork().active(), g), work.jobSize(), &ctrs[g]), gpu, "read_results");
      else {
        check(b200pt_film_accumulate(gpu.ctx), gpu, "film_accumulate");
        check(b200pt_get_counters(gpu.ctx, &ctrs[g]), gpu, "sync");
      }
    };
    if (gpus.size() == 1) renderOn(0);
    else {
      std::vector<std::exception_ptr> errors(gpus.size());
      std::vector<std::thread> workers;
      for (std::size_t g = 0; g < gpus.size(); ++g)
        workers.emplace_back([&, g] { try { renderOn(g); } catch (...) { errors[g] = std::current_exception(); } });
      for (auto& w : workers) w.join();
      for (auto& e : errors) if (e) std::rethrow_exception(e);
    }
    std::size_t stepRays = 0;
    for (std::size_t g = 0; g < gpus.size(); ++g) {
      const b200pt_counters& ctr = ctrs[g];
      summary.paths += ctr.paths; summary.escaped += ctr.escaped; summary.rays += ctr.path_segments;
      summary.traceMs += ctr.trace_ms; summary.nifMs += ctr.nif_ms; summary.deviceMs = std::max(summary.deviceMs, 0.0) + ctr.total_ms / gpus.size();
      stepRays += ctr.path_segments;
      logging::debug("GPU {}: path-trace {} ms, NIF {} ms, total {} ms", gpus[g]->device, ctr.trace_ms, ctr.nif_ms, ctr.total_ms);
    }

    nvtxRangePop();      // ipu_render
    // Wait for completion of previous async task before starting the next (:702-706). With the device film
    // the host thread only ever holds a gathered copy, so the wait moves to the next gather.
    if (recordStream) { Trace t("wait_for_host"); hostProcessing.waitForCompletion(); }

    const bool saveNow = (step % saveInterval == 0) || step == steps;
    if (recordStream) {
      // Swap the worklist buffers (PathTracerApp.cpp:707-709): the list the previous async task accumulated, re-dealt
      // and cleared becomes active; the host thread processes the results just read in the now inactive list. Both
      // lists hold every pixel (LoadBalancer's constructor), so the two alternate and a re-deal of one list takes
      // effect at the step after next, exactly as in the reference.
      work.getWork().swap();
      hostProcessing.run([&, step, saveNow, exposure = state.exposure, gamma = state.gamma]() {
        Trace t("async_work");
        { Trace t1("accumulate_framebuffers"); film.accumulate(work.getWork().inactive()); }
        if (loadBalanceEnabled && step > 1) {
          Trace t2("run_load_balancing");
          work.allocateWorkByPathLength();
          if (!listsStillPinned()) throw std::logic_error("load balancing reallocated a page-locked work list");
        }
        { Trace t3("clear_accumulators"); totalRays = work.clearInactiveAccumulators(); }
        if (saveNow) {
          // With a UI server the full uncompressed image is transmitted at the save interval (:748-759);
          // files are written on the final step only.
          if (uiServer) uiServer->startSendingRawImage(film.getHdrImage(), imageWidth, imageHeight, step, step == steps);
          if (!uiServer || step == steps) {
            Trace t4("save_images");
            film.saveImages(fileName, step, exposure, gamma);
            logging::info("Saved images at step {}", step);
          }
        }
      });
    } else {
      totalRays = stepRays;
      if (saveNow) {
        // One gather per save interval; encoding and writing the files overlaps the next steps. A save interval never
        // stalls a step: if the previous encode is still running the interval is skipped (as the reference drops a
        // raw-image request while a transfer is in progress, InterfaceServer.hpp:282-286) - except on the final step.
        const bool last = step == steps;
        if (hostProcessing.isRunning() && !last) {
          skippedSaves += 1;
          logging::debug("Save at step {} skipped: the previous images are still being written", step);
        } else {
          Trace t("save_images");
          hostProcessing.waitForCompletion();
          saveBuffer ^= 1;
          AccumulatedImage& target = films[saveBuffer];
          gatherDeviceFilm(target);
          if (uiServer) uiServer->startSendingRawImage(target.getHdrImage(), imageWidth, imageHeight, step, last);
          if (!uiServer || last)
            hostProcessing.run([&, step, exposure = state.exposure, gamma = state.gamma, filmPtr = &target]() {
              Trace t("async_work");
              filmPtr->saveImages(fileName, step, exposure, gamma);
              logging::info("Saved images at step {}", step);
            });
        }
      }
    }

    auto loopEndTime = std::chrono::steady_clock::now();
    auto secs = std::chrono::duration<double>(loopEndTime - loopStartTime).count();
    const double pixelSamplesPerStep = (double)imageWidth * imageHeight * samplesPerStep;
    logging::info("Completed render step {}/{} in {} seconds (Samples/sec {}) (Rays/sec {})", step, steps, secs, pixelSamplesPerStep / secs,
                  (recordStream ? totalRays : stepRays) / secs);
    if (uiServer) {
      // exposure and gamma are applied by the client to the raw image: pick up the latest for the files
      const auto latest = uiServer->getState();
      state.exposure = latest.exposure; state.gamma = latest.gamma;
      uiServer->sendPreviewImage({}, imageWidth, imageHeight);
      uiServer->updateProgress((int)step, (int)steps);
      uiServer->updateSampleRate((float)(pixelSamplesPerStep / secs), (float)((recordStream ? totalRays : stepRays) / secs));
    }
  }

  hostProcessing.waitForCompletion();
  nvtxRangePop();      // rendering
  if (skippedSaves) logging::info("{} save interval(s) were skipped because the previous images were still being written", skippedSaves);
  if (uiServer) { uiServer->waitForRawImage(); uiServer.reset(); }
  if (recordStream && loadBalanceEnabled) {
    // the inactive list was re-dealt for a step that never ran: nothing to do, the film is complete
  }

  auto endTime = std::chrono::steady_clock::now();
  summary.elapsedSecs = std::chrono::duration<double>(endTime - startTime).count();
  logging::info("Render finished: {} seconds", summary.elapsedSecs);
  const double pixelsPerFrame = (double)imageWidth * imageHeight;
  summary.samplesPerSec = (pixelsPerFrame / summary.elapsedSecs) * (double)samplesPerPixel;
  logging::info("Samples/sec: {}", summary.samplesPerSec);
  logging::info("Samples/sec/GPU: {}", summary.samplesPerSec / gpus.size());
  if (summary.paths) logging::info("Escaped-ray fraction: {} ; device time per step: path-trace {} ms, NIF {} ms", (double)summary.escaped / summary.paths,
                                   summary.traceMs / steps / gpus.size(), summary.nifMs / steps / gpus.size());
  return summary;
}

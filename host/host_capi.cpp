// C entry points over the host components, so that the CPU test-suite (ctypes) can exercise the option
// parser, the work-list code, the image writers and the asset loaders without a GPU. Not part of the
// drop-in boundary (that is include/b200pt.h).
#include <cstring>
#include <algorithm>
#include <cstdio>
#include <exception>
#include <stdexcept>
#include <string>

#include "AccumulatedImage.hpp"
#include "Hdf5File.hpp"
#include "LoadBalancer.hpp"
#include "NifAssets.hpp"
#include "Options.hpp"
#include "PathTracerApp.hpp"
#include "image_io.hpp"

namespace {
thread_local std::string g_error;
template <class F>
int guarded(F&& f) {
  try { f(); return 0; } catch (const std::exception& e) { g_error = e.what(); return -1; }
}
void copyOut(const std::string& s, char* out, size_t cap) {
  if (!out || !cap) return;
  const size_t n = s.size() < cap - 1 ? s.size() : cap - 1;
  std::memcpy(out, s.data(), n);
  out[n] = 0;
}
} // namespace

extern "C" {

const char* b200host_last_error() { return g_error.c_str(); }

int b200host_write_png(const char* path, const uint8_t* bgr8, uint32_t w, uint32_t h) { return guarded([&] { image_io::writePng(path, bgr8, w, h); }); }
int b200host_write_exr(const char* path, const float* bgr, uint32_t w, uint32_t h) { return guarded([&] { image_io::writeExr(path, bgr, w, h); }); }
int b200host_read_png(const char* path, uint8_t* bgr8, size_t cap, uint32_t* w, uint32_t* h) {
  return guarded([&] { std::vector<uint8_t> v; image_io::readPng(path, v, *w, *h); if (v.size() > cap) throw std::runtime_error("buffer too small"); std::memcpy(bgr8, v.data(), v.size()); });
}
int b200host_read_exr(const char* path, float* bgr, size_t capFloats, uint32_t* w, uint32_t* h) {
  return guarded([&] { std::vector<float> v; image_io::readExr(path, v, *w, *h); if (v.size() > capFloats) throw std::runtime_error("buffer too small"); std::memcpy(bgr, v.data(), v.size() * 4); });
}

// AccumulatedImage: accumulate `n` records `steps` times worth (records are one step's result), save.
int b200host_film_accumulate_save(const b200pt_trace_record* rec, size_t n, uint32_t w, uint32_t h, uint32_t step, float exposure, float gamma,
                                  const char* fileName, float* hdrOut, uint8_t* ldrOut) {
  return guarded([&] {
    AccumulatedImage film(w, h);
    RecordList list(rec, rec + n);
    for (uint32_t s = 0; s < step; ++s) film.accumulate(list);
    if (fileName && *fileName) film.saveImages(fileName, step, exposure, gamma);
    if (hdrOut) std::memcpy(hdrOut, film.getHdrImage().data(), film.getHdrImage().size() * 4);
    if (ldrOut) { auto& l = film.updateLdrImage(step, exposure, gamma); std::memcpy(ldrOut, l.data(), l.size()); }
  });
}

// createTracingJobs: returns items per GPU; out (nullable) receives numGpus * itemsPerGpu records.
long b200host_create_jobs(uint32_t w, uint32_t h, uint32_t numGpus, uint32_t tile, b200pt_trace_record* out, size_t cap) {
  long items = -1;
  guarded([&] {
    auto jobs = createTracingJobs(w, h, numGpus, tile);
    items = (long)jobs.front().size();
    if (out) {
      if (cap < jobs.size() * jobs.front().size()) throw std::runtime_error("buffer too small");
      size_t k = 0;
      for (auto& j : jobs) for (auto& r : j) out[k++] = r;
    }
  });
  return items;
}

// LoadBalancer round trip on a flattened list: balance by path length then clear; returns the ray sum.
long b200host_balance(b200pt_trace_record* items, size_t itemsPerJob, size_t numJobs, int balance) {
  long rays = -1;
  guarded([&] {
    std::vector<RecordList> jobs(numJobs);
    for (size_t j = 0; j < numJobs; ++j) jobs[j].assign(items + j * itemsPerJob, items + (j + 1) * itemsPerJob);
    LoadBalancer lb(jobs);
    if (balance) lb.allocateWorkByPathLength();
    std::memcpy(items, lb.getWork().inactive().data(), numJobs * itemsPerJob * sizeof(b200pt_trace_record));
    rays = (long)lb.clearInactiveAccumulators();
  });
  return rays;
}

size_t b200host_round_spp(size_t spp, size_t perStep) { return roundSamplesPerPixel(spp, perStep); }

// Options: parses argv, writes "name=value\n" for every option that has a value.
int b200host_parse_options(int argc, const char* const* argv, char* out, size_t cap) {
  return guarded([&] {
    Options o;
    o.parse(argc, argv);
    std::string s;
    for (auto& spec : o.specs()) {
      if (spec.isFlag) s += spec.name + "=" + (o.flag(spec.name) ? "1" : "0") + "\n";
      else if (o.has(spec.name)) s += spec.name + "=" + o.str(spec.name) + "\n";
    }
    copyOut(s, out, cap);
  });
}

// NIF assets: loads <assets> and reports "layers=..;E=..;max=..;mean=a,b,c;log=..;flops=..;source=..;shapes=in x out r|..".
int b200host_load_assets(const char* assets, int allowSynthetic, char* out, size_t cap, uint16_t* firstKernel, size_t firstKernelCap) {
  return guarded([&] {
    NifModelData m(assets, allowSynthetic != 0);
    std::string s = "layers=" + std::to_string(m.layers.size()) + ";E=" + std::to_string(m.metaData.embeddingDimension) + ";hidden=" +
                    std::to_string(m.metaData.hiddenSize) + ";log=" + std::to_string((int)m.metaData.logToneMap) + ";flops=" + std::to_string(m.flopsPerEval()) +
                    ";source=" + m.source + ";shapes=";
    for (auto& l : m.layers) s += std::to_string(l.in) + "x" + std::to_string(l.out) + (l.relu ? "r" : "n") + (l.bias.empty() ? "" : "b") + (l.halfFloat ? "h" : "f") + "|";
    char buf[160];
    std::snprintf(buf, sizeof(buf), ";max=%.9g;mean=%.9g,%.9g,%.9g;eps=%.9g", m.metaData.max, m.metaData.mean[0], m.metaData.mean[1], m.metaData.mean[2], m.metaData.eps);
    s += buf;
    copyOut(s, out, cap);
    if (firstKernel && !m.layers.empty() && m.layers[0].halfFloat) {
      const size_t n = std::min(firstKernelCap, m.layers[0].kernel.size() / 2);
      std::memcpy(firstKernel, m.layers[0].kernel.data(), n * 2);
    }
  });
}

// HDF5 reader: dataset bytes and string attributes.
int b200host_h5_dataset(const char* file, const char* path, uint8_t* out, size_t cap, size_t* shape, int* rank, size_t* elementSize) {
  return guarded([&] {
    hdf5::File f(file);
    hdf5::Dataset d = f.dataset(path);
    if ((int)d.shape.size() > *rank) throw std::runtime_error("shape buffer too small");
    *rank = (int)d.shape.size();
    for (size_t i = 0; i < d.shape.size(); ++i) shape[i] = d.shape[i];
    *elementSize = d.elementSize;
    if (d.data.size() > cap) throw std::runtime_error("buffer too small");
    std::memcpy(out, d.data.data(), d.data.size());
  });
}
int b200host_h5_attribute(const char* file, const char* object, const char* name, char* out, size_t cap) {
  return guarded([&] { hdf5::File f(file); copyOut(f.stringAttribute(object, name), out, cap); });
}
int b200host_h5_list(const char* file, const char* group, char* out, size_t cap) {
  return guarded([&] { hdf5::File f(file); std::string s; for (auto& n : f.list(group)) s += n + "\n"; copyOut(s, out, cap); });
}

// Remote UI: serves one client on <port> through the same decisions as PathTracerApp::execute, with a
// synthetic "render": every step takes stepMs, the film value of every pixel is (fov, envRotationDegrees,
// interactiveSamples) in B,G,R (so the client sees R = interactiveSamples, G = rotation, B = fov) times the
// step number. Writes "steps=..;restarts=..;stopped=..;detached=..;exposure=..;gamma=..;fov=..;rot=..;nif=..;isamples=..".
int b200host_ui_session(int port, uint32_t w, uint32_t h, uint32_t steps, uint32_t saveInterval, uint32_t stepMs, int acceptTimeoutS, char* out, size_t cap) {
  return guarded([&] {
    std::unique_ptr<InterfaceServer> ui(new InterfaceServer(port));
    InterfaceServer::State state;
    state.fov = 1.f; state.interactiveSamples = 8;
    ui->setInitialState(state);
    ui->start(acceptTimeoutS);
    AccumulatedImage film(w, h);
    std::vector<float> px((size_t)w * h * 3);
    size_t restarts = 0, done = 0;
    bool stopped = false, detached = false;
    std::string lastNif;
    for (uint32_t step = 1; step <= steps; ++step) {
      if (ui && !ui->connected()) ui.reset();
      if (ui && ui->stateChanged()) {
        state = ui->consumeState();
        if (state.stop) { stopped = true; ui.reset(); break; }
        if (state.detach) { detached = true; ui.reset(); }
        else { if (!state.newNif.empty()) lastNif = state.newNif; film.reset(); restarts++; step = 1; }
      }
      std::this_thread::sleep_for(std::chrono::milliseconds(stepMs));
      for (size_t i = 0; i < px.size(); i += 3) { px[i] = state.fov * step; px[i + 1] = state.envRotationDegrees * step; px[i + 2] = (float)state.interactiveSamples * step; }
      film.setHdr(px.data());
      done = step;
      if (ui) {
        const auto latest = ui->getState();
        state.exposure = latest.exposure; state.gamma = latest.gamma;
        if (step % saveInterval == 0 || step == steps) ui->startSendingRawImage(film.getHdrImage(), w, h, step, step == steps);
        ui->updateProgress((int)step, (int)steps);
        ui->updateSampleRate(1000.f * step, 2000.f * step);
      }
    }
    if (ui) { ui->waitForRawImage(); ui.reset(); }
    char buf[512];
    std::snprintf(buf, sizeof(buf), "steps=%zu;restarts=%zu;stopped=%d;detached=%d;exposure=%.9g;gamma=%.9g;fov=%.9g;rot=%.9g;nif=%s;isamples=%u", done, restarts,
                  (int)stopped, (int)detached, state.exposure, state.gamma, state.fov, state.envRotationDegrees, lastNif.c_str(), state.interactiveSamples);
    copyOut(buf, out, cap);
  });
}
} // extern "C"

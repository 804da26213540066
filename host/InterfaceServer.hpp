// Remote user-interface server (src/InterfaceServer.hpp:22-351): the same packet table, payloads and state
// semantics, on host/PacketComms.hpp instead of external/packetcomms + cereal + OpenCV + FFmpeg.
//   client -> server: stop, detach, env_rotation, exposure, gamma, fov (degrees), load_nif, interactive_samples
//   server -> client: progress, sample_rate, hdr_header + hdr_packet (one image row per packet, RGB, throttled 2 ms)
// "render_preview" (FFmpeg FMP4 video of the tone-mapped image, InterfaceServer.hpp:101-111,262-268) is NOT sent:
// there is no libav in this image. The packet id stays in the table so that all other ids keep the reference's values.
#pragma once

#include <atomic>
#include <chrono>
#include <cmath>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "AsyncTask.hpp"
#include "PacketComms.hpp"
#include "logging.hpp"

class InterfaceServer {
public:
  static const std::vector<std::string>& packetTypes() {
    static const std::vector<std::string> t = {"stop", "detach", "progress", "sample_rate", "env_rotation", "exposure", "gamma", "fov",
                                               "load_nif", "render_preview", "hdr_header", "hdr_packet", "interactive_samples"};
    return t;   // order is the wire contract (InterfaceServer.hpp:23-42)
  }

  enum class Status { Stop, Restart, Continue, Disconnected };

  struct State {
    float envRotationDegrees = 0.f;
    float exposure = 0.f;
    float gamma = 2.2f;
    float fov = 90.f;                       // radians once a client has sent it (:147-154), like the reference
    std::uint32_t interactiveSamples = 8;
    std::string newNif;
    bool stop = false;
    bool detach = false;
  };

  explicit InterfaceServer(int portNumber) : port(portNumber) {}
  ~InterfaceServer() {
    try { sendHdrTask.waitForCompletion(); } catch (...) {}
    stop();
  }

  /// Seeds the state a client has not touched yet with the command-line values.
  void setInitialState(const State& s) { std::lock_guard<std::mutex> g(stateLock); state = s; }

  /// Binds, listens and blocks until a client has connected and the Tx/Rx threads run (:88-176, 218-224).
  /// timeoutSeconds <= 0 waits for ever like the reference.
  void start(int timeoutSeconds = 0) {
    listenFd = ::socket(AF_INET, SOCK_STREAM, 0);
    if (listenFd < 0) throw std::runtime_error("User interface server: could not create a socket");
    int one = 1;
    setsockopt(listenFd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in addr{};
    addr.sin_family = AF_INET; addr.sin_addr.s_addr = htonl(INADDR_ANY); addr.sin_port = htons((std::uint16_t)port);
    if (::bind(listenFd, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0 || ::listen(listenFd, 1) != 0) {
      ::close(listenFd); listenFd = -1;
      throw std::runtime_error("User interface server: could not bind/listen on port " + std::to_string(port));
    }
    logging::info("User interface server listening on port {}", port);
    pollfd pfd{listenFd, POLLIN, 0};
    const int r = ::poll(&pfd, 1, timeoutSeconds > 0 ? timeoutSeconds * 1000 : -1);
    if (r <= 0) { ::close(listenFd); listenFd = -1; throw std::runtime_error("User interface server: no client connected"); }
    const int fd = ::accept(listenFd, nullptr, nullptr);
    ::close(listenFd); listenFd = -1;
    if (fd < 0) throw std::runtime_error("User interface server: accept failed");
    logging::info("User interface client connected.");
    link.reset(new packetcomms::Link(fd, packetTypes()));
    using Bytes = std::vector<std::uint8_t>;
    link->subscribe("env_rotation", [this](const Bytes& b) { std::lock_guard<std::mutex> g(stateLock); state.envRotationDegrees = packetcomms::Reader(b).pod<float>(); stateUpdated = true; });
    link->subscribe("detach", [this](const Bytes& b) { std::lock_guard<std::mutex> g(stateLock); state.detach = packetcomms::Reader(b).pod<std::uint8_t>() != 0; stateUpdated = true; });
    link->subscribe("stop", [this](const Bytes& b) { std::lock_guard<std::mutex> g(stateLock); state.stop = packetcomms::Reader(b).pod<std::uint8_t>() != 0; stateUpdated = true; });
    // Tone mapping is not done on the device: exposure and gamma do not restart the render (:131-143).
    link->subscribe("exposure", [this](const Bytes& b) { std::lock_guard<std::mutex> g(stateLock); state.exposure = packetcomms::Reader(b).pod<float>(); });
    link->subscribe("gamma", [this](const Bytes& b) { std::lock_guard<std::mutex> g(stateLock); state.gamma = packetcomms::Reader(b).pod<float>(); });
    link->subscribe("fov", [this](const Bytes& b) { std::lock_guard<std::mutex> g(stateLock); state.fov = packetcomms::Reader(b).pod<float>() * (float)(M_PI / 180.0); stateUpdated = true; });
    link->subscribe("load_nif", [this](const Bytes& b) { std::lock_guard<std::mutex> g(stateLock); state.newNif = packetcomms::Reader(b).str(); stateUpdated = true; });
    link->subscribe("interactive_samples", [this](const Bytes& b) { std::lock_guard<std::mutex> g(stateLock); state.interactiveSamples = packetcomms::Reader(b).pod<std::uint32_t>(); stateUpdated = true; });
    link->start();
    logging::info("User interface server entering Tx/Rx loop.");
  }

  void stop() {
    if (link) { link->stop(); link.reset(); logging::info("User interface server Tx/Rx loop exited."); }
    if (listenFd >= 0) { ::close(listenFd); listenFd = -1; }
  }

  bool connected() const { return link && link->ok(); }

  /// Return a copy of the state and mark it as consumed (:192-197).
  State consumeState() {
    std::lock_guard<std::mutex> g(stateLock);
    State tmp = state;
    stateUpdated = false;
    state.newNif.clear();
    return tmp;
  }
  State getState() { std::lock_guard<std::mutex> g(stateLock); return state; }
  bool stateChanged() const { return stateUpdated; }

  void updateProgress(int step, int totalSteps) {
    if (connected()) link->post("progress", packetcomms::Writer().pod<float>(step / (float)totalSteps).bytes);
  }
  void updateSampleRate(float pathRate, float rayRate) {
    if (connected()) link->post("sample_rate", packetcomms::Writer().pod<float>(pathRate).pod<float>(rayRate).bytes);
  }
  /// The reference encodes the tone-mapped frame into an FMP4 stream; without libav nothing is sent.
  void sendPreviewImage(const std::vector<std::uint8_t>&, std::size_t, std::size_t) {
    if (!warnedNoVideo) { warnedNoVideo = true; logging::warn("Preview video (render_preview) is unavailable in this build; the raw HDR image is sent at every save interval."); }
  }

  /// Sends hdr / step, B,G,R -> R,G,B, one row per packet in the background (:270-331). Returns false if
  /// the previous transfer is still running (the request is dropped, as in the reference).
  /// finalImage waits for a transfer in flight instead, so that the client always ends with the finished render.
  bool startSendingRawImage(const std::vector<float>& bgr, std::size_t width, std::size_t height, std::size_t step, bool finalImage = false) {
    if (!finalImage && sendHdrTask.isRunning()) { logging::debug("Large data transfer still in progress, dropping request"); return false; }
    sendHdrTask.waitForCompletion();
    if (!connected()) { logging::debug("No muxer available: large data transfer aborted."); return false; }
    hdrImage.resize(bgr.size());
    const float scale = 1.f / (float)step;
    for (std::size_t i = 0; i + 2 < bgr.size(); i += 3) { hdrImage[i] = bgr[i + 2] * scale; hdrImage[i + 1] = bgr[i + 1] * scale; hdrImage[i + 2] = bgr[i] * scale; }
    const std::uint32_t chunks = (std::uint32_t)height;
    logging::debug("Initiating large data transfer: {} chunks", chunks);
    link->post("hdr_header", packetcomms::Writer().pod<std::int32_t>((std::int32_t)width).pod<std::int32_t>((std::int32_t)height).pod<std::uint32_t>(chunks).bytes);
    sendHdrTask.run([this, width, chunks]() {
      const auto startTime = std::chrono::steady_clock::now();
      const std::size_t chunkSize = width * 3;
      for (std::uint32_t c = 0; c < chunks && connected(); ++c) {
        link->post("hdr_packet", packetcomms::Writer().pod<std::uint32_t>(c).floats(hdrImage.data() + (std::size_t)c * chunkSize, chunkSize).bytes);
        std::this_thread::sleep_for(std::chrono::milliseconds(2));     // throttle to keep the link interactive
      }
      const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - startTime).count();
      logging::info("{} MiB raw image transmitted in {} seconds", hdrImage.size() * sizeof(float) / (1024.f * 1024.f), secs);
    });
    return true;
  }
  void waitForRawImage() { sendHdrTask.waitForCompletion(); while (connected() && link->queued() > 0) std::this_thread::sleep_for(std::chrono::milliseconds(2)); }

private:
  int port;
  int listenFd = -1;
  std::unique_ptr<packetcomms::Link> link;
  std::atomic<bool> stateUpdated{false};
  State state;
  std::mutex stateLock;
  std::vector<float> hdrImage;
  AsyncTask sendHdrTask;
  bool warnedNoVideo = false;
};

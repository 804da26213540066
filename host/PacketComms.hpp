// Wire protocol of the reference's remote user interface, re-implemented on POSIX sockets:
// external/packetcomms (PacketMuxer.cpp:120-138 framing, IdManager.h:18-31 type ids, ControlMessage.h,
// PacketDemuxer.cpp:105-140,230-249 hello/receive loop) with cereal's portable-less *binary* archive as payload
// encoding (native little-endian scalars; std::string and std::vector prefixed by a 64-bit element count).
//
//   packet  := u32 type (network order) | u32 size (network order) | size bytes
//   type    := 0 invalid, 1 control, 2 + index into the application's packet-name table
//   control := 1 byte: 0 HeartBeat (sent after every idle second), 254 Hello (first packet of a muxer), 255 GoodBye
#pragma once

#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <cerrno>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <deque>
#include <functional>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace packetcomms {

enum : std::uint8_t { HeartBeat = 0, Hello = 254, GoodBye = 255 };
constexpr std::uint32_t kInvalidPacket = 0, kControlPacket = 1;

// cereal::BinaryOutputArchive / BinaryInputArchive for the handful of types the UI exchanges.
struct Writer {
  std::vector<std::uint8_t> bytes;
  template <class T> Writer& pod(const T& v) { const auto* p = reinterpret_cast<const std::uint8_t*>(&v); bytes.insert(bytes.end(), p, p + sizeof(T)); return *this; }
  Writer& str(const std::string& s) { pod<std::uint64_t>(s.size()); bytes.insert(bytes.end(), s.begin(), s.end()); return *this; }
  Writer& floats(const float* p, std::size_t n) { pod<std::uint64_t>(n); const auto* b = reinterpret_cast<const std::uint8_t*>(p); bytes.insert(bytes.end(), b, b + 4 * n); return *this; }
};
struct Reader {
  const std::vector<std::uint8_t>& bytes;
  std::size_t pos = 0;
  explicit Reader(const std::vector<std::uint8_t>& b) : bytes(b) {}
  template <class T> T pod() {
    if (pos + sizeof(T) > bytes.size()) throw std::runtime_error("packet payload too short");
    T v; std::memcpy(&v, bytes.data() + pos, sizeof(T)); pos += sizeof(T); return v;
  }
  std::string str() {
    const std::uint64_t n = pod<std::uint64_t>();
    if (n > bytes.size() - pos) throw std::runtime_error("packet payload too short");
    std::string s(reinterpret_cast<const char*>(bytes.data() + pos), (std::size_t)n); pos += (std::size_t)n; return s;
  }
};

// One accepted TCP connection with a muxer (send thread: Hello first, HeartBeat when idle) and a demuxer
// (receive thread: expects Hello, then dispatches packets to the subscribed callbacks).
class Link {
public:
  using Callback = std::function<void(const std::vector<std::uint8_t>&)>;

  Link(int fd, const std::vector<std::string>& packetNames) : fd_(fd), names_(packetNames) {
    int one = 1;
    setsockopt(fd_, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
  }
  ~Link() { stop(); if (fd_ >= 0) ::close(fd_); }

  std::uint32_t idOf(const std::string& name) const {
    for (std::size_t i = 0; i < names_.size(); ++i) if (names_[i] == name) return (std::uint32_t)i + 2u;
    throw std::out_of_range("unknown packet type '" + name + "'");
  }
  void subscribe(const std::string& name, Callback cb) { callbacks_[idOf(name)] = std::move(cb); }

  /// Starts the two threads (subscribe before).
  void start() {
    error_ = false;
    sendThread_ = std::thread([this] { sendLoop(); });
    recvThread_ = std::thread([this] { receiveLoop(); });
  }
  void stop() {
    { std::lock_guard<std::mutex> g(lock_); error_ = true; }
    ready_.notify_all();
    if (fd_ >= 0) ::shutdown(fd_, SHUT_RDWR);
    if (sendThread_.joinable()) sendThread_.join();
    if (recvThread_.joinable()) recvThread_.join();
  }
  bool ok() const { return !error_; }

  /// PacketMuxer::emplacePacket: queue a packet for the send thread.
  void post(const std::string& name, std::vector<std::uint8_t> payload) { post(idOf(name), std::move(payload)); }
  std::size_t queued() { std::lock_guard<std::mutex> g(lock_); return queue_.size(); }

private:
  int fd_;
  std::vector<std::string> names_;
  std::map<std::uint32_t, Callback> callbacks_;
  std::atomic<bool> error_{false};
  std::mutex lock_;
  std::condition_variable ready_;
  std::deque<std::pair<std::uint32_t, std::vector<std::uint8_t>>> queue_;
  std::thread sendThread_, recvThread_;

  void post(std::uint32_t type, std::vector<std::uint8_t> payload) {
    { std::lock_guard<std::mutex> g(lock_); queue_.emplace_back(type, std::move(payload)); }
    ready_.notify_one();
  }
  bool writeAll(const std::uint8_t* p, std::size_t n) {
    while (n > 0) {
      const ssize_t w = ::send(fd_, p, n, MSG_NOSIGNAL);
      if (w < 0) { if (errno == EINTR) continue; return false; }
      p += w; n -= (std::size_t)w;
    }
    return true;
  }
  bool sendPacket(std::uint32_t type, const std::vector<std::uint8_t>& payload) {
    std::uint32_t hdr[2] = {htonl(type), htonl((std::uint32_t)payload.size())};
    return writeAll(reinterpret_cast<const std::uint8_t*>(hdr), 8) && writeAll(payload.data(), payload.size());
  }
  void sendLoop() {
    if (!sendPacket(kControlPacket, {Hello})) { error_ = true; return; }
    std::unique_lock<std::mutex> g(lock_);
    while (!error_) {
      if (queue_.empty()) {
        if (ready_.wait_for(g, std::chrono::seconds(1)) == std::cv_status::timeout && queue_.empty() && !error_) {
          g.unlock();
          const bool sent = sendPacket(kControlPacket, {HeartBeat});
          g.lock();
          if (!sent) error_ = true;
        }
        continue;
      }
      auto pkt = std::move(queue_.front());
      queue_.pop_front();
      g.unlock();
      const bool sent = sendPacket(pkt.first, pkt.second);
      g.lock();
      if (!sent) error_ = true;
    }
  }
  // returns 1 ok, 0 timeout, -1 error
  int readAll(std::uint8_t* p, std::size_t n, int timeoutMs) {
    while (n > 0) {
      pollfd pfd{fd_, POLLIN, 0};
      const int r = ::poll(&pfd, 1, timeoutMs);
      if (r == 0) return 0;
      if (r < 0) { if (errno == EINTR) continue; return -1; }
      const ssize_t k = ::recv(fd_, p, n, 0);
      if (k == 0) return -1;                      // the other side hung up
      if (k < 0) { if (errno == EINTR || errno == EAGAIN) continue; return -1; }
      p += k; n -= (std::size_t)k;
      timeoutMs = 5000;                           // mid-packet: the rest must follow
    }
    return 1;
  }
  void receiveLoop() {
    bool first = true;
    while (!error_) {
      std::uint32_t hdr[2];
      const int r = readAll(reinterpret_cast<std::uint8_t*>(hdr), 8, first ? 2000 : 1000);
      if (r == 0) continue;
      if (r < 0) { error_ = true; break; }
      const std::uint32_t type = ntohl(hdr[0]), size = ntohl(hdr[1]);
      if (size > (64u << 20)) { error_ = true; break; }
      std::vector<std::uint8_t> payload(size);
      if (size && readAll(payload.data(), size, 5000) != 1) { error_ = true; break; }
      if (first) {
        // Very first packet must be a Hello control packet (PacketDemuxer.cpp:230-249)
        first = false;
        if (type != kControlPacket || payload.empty() || payload[0] != Hello) { error_ = true; break; }
        continue;
      }
      if (type == kControlPacket) {               // heartbeats carry no application data; GoodBye ends the link
        if (!payload.empty() && payload[0] == GoodBye) { error_ = true; break; }
        continue;
      }
      auto it = callbacks_.find(type);
      if (it != callbacks_.end()) {
        try { it->second(payload); } catch (const std::exception&) { /* malformed payload: ignore the packet */ }
      }
    }
    ready_.notify_all();
  }
};

} // namespace packetcomms

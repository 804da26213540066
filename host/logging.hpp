// Levelled logging in the reference's format (src/main.cpp:68-88: spdlog pattern
// "[%H:%M:%S.%f] [%L] [%t] %v", levels trace..off).
#pragma once

#include <chrono>
#include <cstdio>
#include <ctime>
#include <map>
#include <mutex>
#include <sstream>
#include <stdexcept>
#include <string>
#include <thread>

namespace logging {

enum Level { Trace = 0, Debug, Info, Warn, Err, Critical, Off };

inline Level& threshold() { static Level l = Info; return l; }
inline std::mutex& mutex() { static std::mutex m; return m; }

inline void setLevel(const std::string& name) {
  static const std::map<std::string, Level> levels = {{"trace", Trace}, {"debug", Debug}, {"info", Info}, {"warn", Warn},
                                                      {"err", Err},     {"critical", Critical}, {"off", Off}};
  auto it = levels.find(name);
  if (it == levels.end()) throw std::runtime_error("Invalid log-level: '" + name + "'");
  threshold() = it->second;
}

inline void write(Level l, const std::string& msg) {
  if (l < threshold()) return;
  using namespace std::chrono;
  const auto now = system_clock::now();
  const std::time_t t = system_clock::to_time_t(now);
  const long us = (long)(duration_cast<microseconds>(now.time_since_epoch()).count() % 1000000);
  std::tm tm{};
  localtime_r(&t, &tm);
  static const char tags[] = {'T', 'D', 'I', 'W', 'E', 'C'};
  std::ostringstream tid;
  tid << std::this_thread::get_id();
  std::lock_guard<std::mutex> lock(mutex());
  std::fprintf(stderr, "[%02d:%02d:%02d.%06ld] [%c] [%s] %s\n", tm.tm_hour, tm.tm_min, tm.tm_sec, us, tags[l], tid.str().c_str(), msg.c_str());
}

// "{}" placeholders like spdlog/fmt.
inline void formatInto(std::ostringstream& out, const char* fmt) { out << fmt; }
template <class T, class... Rest>
void formatInto(std::ostringstream& out, const char* fmt, const T& v, const Rest&... rest) {
  for (; *fmt; ++fmt) {
    if (fmt[0] == '{' && fmt[1] == '}') { out << v; formatInto(out, fmt + 2, rest...); return; }
    out << *fmt;
  }
}
template <class... Args>
void log(Level l, const char* fmt, const Args&... args) {
  if (l < threshold()) return;
  std::ostringstream out;
  out.precision(9);
  formatInto(out, fmt, args...);
  write(l, out.str());
}
template <class... Args> void trace(const char* f, const Args&... a) { log(Trace, f, a...); }
template <class... Args> void debug(const char* f, const Args&... a) { log(Debug, f, a...); }
template <class... Args> void info(const char* f, const Args&... a) { log(Info, f, a...); }
template <class... Args> void warn(const char* f, const Args&... a) { log(Warn, f, a...); }
template <class... Args> void error(const char* f, const Args&... a) { log(Err, f, a...); }

} // namespace logging

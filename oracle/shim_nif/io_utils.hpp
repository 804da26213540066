// TEST INFRASTRUCTURE ONLY. Stand-in for the reference's src/io_utils.hpp / src/ipu_utils.hpp when its
// neural_networks/IoBuffer.cpp and the host-side twins of NifModel.cpp are compiled unmodified into oracle/_ref:
// those files only need ipu_utils::logger() (spdlog in the reference, absent here).
#pragma once
#include <stdexcept>
#include <string>
namespace ipu_utils {
struct NullLogger {
  template <class... A> void trace(A&&...) {}
  template <class... A> void debug(A&&...) {}
  template <class... A> void info(A&&...) {}
  template <class... A> void warn(A&&...) {}
  template <class... A> void error(A&&...) {}
};
inline NullLogger* logger() { static NullLogger l; return &l; }
}  // namespace ipu_utils

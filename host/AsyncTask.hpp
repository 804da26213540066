// One background job at a time (src/AsyncTask.hpp:13-67): the render loop hands the previous step's
// results to this thread (film accumulation, load balancing, image saving) while the GPUs trace the next.
#pragma once

#include <atomic>
#include <exception>
#include <functional>
#include <memory>
#include <stdexcept>
#include <thread>

#include <nvtx3/nvToolsExt.h>

#include "logging.hpp"

class AsyncTask {
public:
  AsyncTask() : running(false) {}
  ~AsyncTask() { if (job && job->joinable()) job->join(); }

  /// Throws std::logic_error if a task was already in progress.
  void run(std::function<void()>&& f) {
    if (job != nullptr) {
      const char* error = "Attempted to run AsyncTask while a job was in progress.";
      logging::error(error);
      throw std::logic_error(error);
    }
    asyncFunc = std::move(f);
    failure = nullptr;
    nvtxRangePushA("thread_launch");      // the reference's tracepoints (src/AsyncTask.hpp:40,49)
    running = true;                       // set here, not in the thread: isRunning() is true as soon as run() returns
    job.reset(new std::thread([this]() {
      try { asyncFunc(); } catch (...) { failure = std::current_exception(); }
      running = false;
    }));
    nvtxRangePop();
  }

  /// Joins the job; an exception thrown inside it resurfaces here, on the render thread.
  void waitForCompletion() {
    if (job != nullptr) {
      nvtxRangePushA("thread_join");
      job->join();
      nvtxRangePop();
      job.reset();
      if (failure) { auto f = failure; failure = nullptr; std::rethrow_exception(f); }
    }
  }

  bool isRunning() const { return running; }

private:
  std::function<void()> asyncFunc;
  std::unique_ptr<std::thread> job;
  std::atomic<bool> running;
  std::exception_ptr failure;
};

// b200_path_trace: command-line front end, the counterpart of the reference's src/main.cpp:90-102
// (options -> logging -> app.init -> run; any exception is logged and turns into EXIT_FAILURE like
// GraphManager::run, src/ipu_utils.hpp:532-535).
#include <cstdlib>
#include <exception>
#include <string>

#include "Options.hpp"
#include "PathTracerApp.hpp"
#include "logging.hpp"

int main(int argc, char** argv) {
  // Load every kernel when the contexts are created (device acquisition), not inside the first render step.
  setenv("CUDA_MODULE_LOADING", "EAGER", 0);
  try {
    Options opts;
    try {
      opts.parse(argc, argv);
    } catch (const std::runtime_error& e) {
      if (std::string(e.what()) == "Show help") return EXIT_FAILURE;      // the reference's main exits non-zero after --help too
      throw;
    }
    logging::setLevel(opts.str("log-level"));
    PathTracerApp app;
    app.init(opts);
    app.execute();
    return EXIT_SUCCESS;
  } catch (const std::exception& e) {
    logging::error("{}", e.what());
    return EXIT_FAILURE;
  }
}

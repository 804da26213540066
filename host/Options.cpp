#include "Options.hpp"

#include <cstdio>
#include <cstdlib>
#include <sstream>
#include <stdexcept>

Options::Options() {
  auto add = [&](const char* name, char s, bool flag, bool req, const char* def, const char* help) {
    specs_.push_back(OptionSpec{name, s, flag, req, def, help});
  };
  // main.cpp:8-37
  add("help", 0, true, false, "", "Show command help.");
  add("model", 0, true, false, "", "(Poplar IPUModel switch: ignored, there is no CPU path.)");
  add("ipus", 0, false, false, "1", "Number of GPUs to use (the reference's IPU count).");
  add("save-exe", 0, false, false, "", "(Poplar executable caching: ignored.)");
  add("load-exe", 0, false, false, "", "(Poplar executable caching: ignored.)");
  add("compile-only", 0, true, false, "", "(Poplar executable caching: ignored.)");
  add("defer-attach", 0, true, false, "", "(Poplar device attach policy: ignored.)");
  add("log-level", 0, false, false, "info", "Set the log level to one of the following: 'trace', 'debug', 'info', 'warn', 'err', 'critical', 'off'.");
  // PathTracerApp.cpp:794-830
  add("outfile", 'o', false, true, "", "Set output file name.");
  add("save-interval", 0, false, false, "1", "Save images every this many render steps.");
  add("width", 'w', false, false, "256", "Output image width (total pixels).");
  add("height", 'h', false, false, "256", "Output image height (total pixels).");
  add("samples", 's', false, false, "512", "Total samples to take per pixel.");
  add("samples-per-step", 0, false, false, "512", "Samples to take per GPU step.");
  add("interactive-samples", 0, false, false, "8", "Number of samples to take per step during user interaction.");
  add("refractive-index", 'n', false, false, "1.5", "Refractive index.");
  add("roulette-depth", 0, false, false, "3", "Number of bounces before rays are randomly stopped.");
  add("stop-prob", 0, false, false, "0.3", "Probability of a ray being stopped.");
  add("aa-noise-scale", 'a', false, false, "0.3", "Scale of anti-aliasing noise (pixels).");
  add("fov", 0, false, false, "90", "Horizontal field of view (degrees).");
  add("exposure", 0, false, false, "0", "Exposure compensation for tone-mapping.");
  add("gamma", 0, false, false, "2.2", "Gamma correction for tone-mapping.");
  add("env-map-rotation", 0, false, false, "0", "Azimuthal rotation for HDRI environment map (degrees).");
  add("seed", 0, false, false, "1", "Seed for random number generation.");
  add("aa-noise-type", 0, false, false, "normal", "Choose distribution for anti-aliasing noise ['uniform', 'normal', 'truncated-normal'].");
  add("codelet-path", 0, false, false, "./", "(Path to Poplar codelets: ignored.)");
  add("enable-load-balancing", 0, true, false, "", "Run dynamic load balancing algorithm for path tracing.");
  add("max-path-length", 0, false, false, "10", "Maximum number of path segments.");
  add("assets", 0, false, true, "", "Path to the 'assets.extra' directory of the saved keras model.");
  add("partials-type", 0, false, false, "half", "(Poplar matmul partials type: ignored, accumulation is fp32 in tensor memory.)");
  add("available-memory-proportion", 0, false, false, "0.6", "(Poplar matmul memory proportion: ignored.)");
  add("max-nif-batch-size", 0, false, false, "44160",
      "Maximum batch-size for the NIF neural network (sizes the escaped-ray wavefront: samples per launch = "
      "max(1, this / pixels-per-GPU); 0 = automatic).");
  add("ui-port", 0, false, false, "0", "Start a remote user-interface server on the specified port.");
  // additions of this host (no reference counterpart)
  add("host-film", 0, true, false, "",
      "Accumulate the film on the host from the TraceRecord stream of every step, exactly like the reference "
      "(setup / path_trace / read_results per step). Default: device film, one gather per save interval.");
  add("synthetic-weights", 0, true, false, "", "Use seeded synthetic NIF weights in the topology of nif_metadata.txt when the assets hold no weights.");
  add("devices", 0, false, false, "", "Comma separated CUDA device ordinals (default 0..ipus-1).");
  add("ui-accept-timeout", 0, false, false, "0", "Seconds to wait for the remote user interface to connect (0 = wait for ever, like the reference).");
}

const OptionSpec* Options::find(const std::string& name) const {
  for (auto& s : specs_) if (s.name == name) return &s;
  return nullptr;
}
const OptionSpec* Options::findShort(char c) const {
  for (auto& s : specs_) if (s.shortName == c) return &s;
  return nullptr;
}

std::string Options::usage() const {
  std::ostringstream out;
  out << "Options:\n";
  for (auto& s : specs_) {
    std::string left = "  --" + s.name;
    if (s.shortName) left = std::string("  -") + s.shortName + " [ --" + s.name + " ]";
    if (!s.isFlag) left += " arg";
    if (!s.defaultValue.empty()) left += " (=" + s.defaultValue + ")";
    out << left;
    if (left.size() < 42) out << std::string(42 - left.size(), ' '); else out << "\n" << std::string(42, ' ');
    out << s.help << "\n";
  }
  return out.str();
}

void Options::parse(int argc, const char* const* argv) {
  for (int i = 1; i < argc; ++i) {
    const std::string arg = argv[i];
    const OptionSpec* spec = nullptr;
    std::string value;
    bool haveValue = false;
    if (arg.rfind("--", 0) == 0) {
      std::string name = arg.substr(2);
      const size_t eq = name.find('=');
      if (eq != std::string::npos) { value = name.substr(eq + 1); name = name.substr(0, eq); haveValue = true; }
      spec = find(name);
      if (!spec) throw std::runtime_error("unrecognised option '" + arg + "'");
    } else if (arg.size() >= 2 && arg[0] == '-' && !(arg[1] >= '0' && arg[1] <= '9') && arg[1] != '.') {
      spec = findShort(arg[1]);
      if (!spec) throw std::runtime_error("unrecognised option '" + arg + "'");
      if (arg.size() > 2) { value = arg.substr(2); haveValue = true; }      // -w1104
    } else {
      throw std::runtime_error("too many positional options have been specified on the command line: '" + arg + "'");
    }
    if (spec->isFlag) {
      if (haveValue) throw std::runtime_error("option '--" + spec->name + "' does not take any arguments");
      values_[spec->name] = "1";
    } else {
      if (!haveValue) {
        if (i + 1 >= argc) throw std::runtime_error("the required argument for option '--" + spec->name + "' is missing");
        value = argv[++i];
      }
      values_[spec->name] = value;
    }
    given_[spec->name] = true;
  }
  if (flag("help")) {
    std::fputs(usage().c_str(), stdout);
    throw std::runtime_error("Show help");
  }
  for (auto& s : specs_) {
    if (s.required && !given_.count(s.name)) throw std::runtime_error("the option '--" + s.name + "' is required but missing");
    if (!s.isFlag && !values_.count(s.name) && !s.defaultValue.empty()) values_[s.name] = s.defaultValue;
  }
  // main.cpp:58-63
  if (!str("save-exe").empty() && !str("load-exe").empty()) throw std::logic_error("You can not set both save-exe and load-exe.");
}

bool Options::has(const std::string& name) const { return values_.count(name) != 0; }
bool Options::given(const std::string& name) const { return given_.count(name) != 0; }
std::string Options::str(const std::string& name) const {
  auto it = values_.find(name);
  return it == values_.end() ? std::string() : it->second;
}
bool Options::flag(const std::string& name) const { return values_.count(name) != 0 && values_.at(name) == "1" && find(name) && find(name)->isFlag; }

std::uint64_t Options::u64(const std::string& name) const {
  const std::string v = str(name);
  char* end = nullptr;
  if (v.empty() || v[0] == '-') throw std::runtime_error("the argument ('" + v + "') for option '--" + name + "' is invalid");
  const unsigned long long x = std::strtoull(v.c_str(), &end, 10);
  if (!end || *end) throw std::runtime_error("the argument ('" + v + "') for option '--" + name + "' is invalid");
  return x;
}
float Options::f32(const std::string& name) const {
  const std::string v = str(name);
  char* end = nullptr;
  const float x = std::strtof(v.c_str(), &end);
  if (v.empty() || !end || *end) throw std::runtime_error("the argument ('" + v + "') for option '--" + name + "' is invalid");
  return x;
}

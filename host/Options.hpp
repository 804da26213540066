// Command line of the path tracer: the option names, short forms and defaults of the reference
// (src/main.cpp:8-37 "standard" options, src/PathTracerApp.cpp:794-830 tool options), parsed without
// boost::program_options. --ipus is the number of GPUs. Options that only make sense for a Poplar
// graph (--model, --save-exe, --load-exe, --compile-only, --defer-attach, --codelet-path,
// --partials-type, --available-memory-proportion) are accepted and ignored with a warning, so that
// existing command lines keep working.
#pragma once

#include <cstdint>
#include <map>
#include <string>
#include <vector>

struct OptionSpec {
  std::string name;        // long name without the dashes
  char shortName;          // 0 if none
  bool isFlag;             // bool_switch
  bool required;
  std::string defaultValue;
  std::string help;
};

class Options {
public:
  Options();
  /// A parser over another option table (host/trace_main.cpp: the options of external/light/src/trace.cpp).
  explicit Options(std::vector<OptionSpec> specs) : specs_(std::move(specs)) {}
  /// Throws std::runtime_error on unknown options / missing values / missing required options,
  /// and std::runtime_error("Show help") after printing the usage when --help is given (main.cpp:43-46).
  void parse(int argc, const char* const* argv);

  bool has(const std::string& name) const;                 // given on the command line (or defaulted)
  bool given(const std::string& name) const;               // given on the command line
  std::string str(const std::string& name) const;
  std::uint64_t u64(const std::string& name) const;
  std::uint32_t u32(const std::string& name) const { return (std::uint32_t)u64(name); }
  float f32(const std::string& name) const;
  bool flag(const std::string& name) const;

  std::string usage() const;
  const std::vector<OptionSpec>& specs() const { return specs_; }

private:
  const OptionSpec* find(const std::string& name) const;
  const OptionSpec* findShort(char c) const;
  std::vector<OptionSpec> specs_;
  std::map<std::string, std::string> values_;
  std::map<std::string, bool> given_;
};

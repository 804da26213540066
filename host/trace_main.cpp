// b200_trace: the reference's stand-alone renderer of the Cornell-box scene (external/light/src/trace.cpp) with the
// CPU render loop replaced by b200pt_cornell_render. Same options and defaults (trace.cpp:19-35), same outputs:
//   <outfile>                        8-bit image, value = min(radiance, 255) per channel (cvImageFromJobs, trace.cpp:82-98)
//   <outfile>.exr                    radiance (the reference writes tiled FP16; here scan-line FP32)
//   <outfile>.command_line_options   the options in effect (trace.cpp:46-54)
//   <outfile>.hist.txt               (reference: max path length per tile job) - with pixel streams there are no
//                                    jobs; one line with the global maximum
// stdout: "Render time", "Total Rays ... Rays/sec ... Max path length" (trace.cpp:296-310).
// --rng tile reproduces the reference's image for a seed bit for bit (one serial random stream per tile job);
// --rng pixel (default) is the GPU schedule: same estimator, different random numbers.
// The OpenCV window (--no-gui absent) is unavailable on this box: the flag is accepted and ignored.
#include <cfloat>
#include <cmath>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "../include/b200pt.h"
#include "Options.hpp"
#include "image_io.hpp"

int main(int argc, char** argv) {
  try {
    char epsDefault[32], hitDefault[32], aaDefault[32];
    std::snprintf(epsDefault, sizeof(epsDefault), "%.9g", (double)FLT_EPSILON);
    std::snprintf(hitDefault, sizeof(hitDefault), "%.9g", (double)1e-5f);
    std::snprintf(aaDefault, sizeof(aaDefault), "%.9g", (double)(float)(1.0 / 700));
    Options opts(std::vector<OptionSpec>{
        {"help", 0, true, false, "", "Output help and exit."},
        {"outfile", 'o', false, true, "", "Set output file name."},
        {"width", 'w', false, false, "256", "Output image width (total pixels)."},
        {"height", 'h', false, false, "256", "Output image height (total pixels)."},
        {"tile-width", 0, false, false, "16", "Width of tile (pixels)."},
        {"tile-height", 0, false, false, "16", "Height of tile (pixels)."},
        {"samples", 's', false, false, "32", "Samples per pixel."},
        {"refractive-index", 'n', false, false, "1.5", "Refractive index."},
        {"roulette-depth", 0, false, false, "3", "Number of bounces before rays are randomly stopped."},
        {"stop-prob", 0, false, false, "0.2", "Probability of a ray being stopped."},
        {"aa-noise-scale", 'a', false, false, aaDefault, "Scale for pixel space anti-aliasing noise."},
        {"seed", 0, false, false, "1", "Seed for random number generation."},
        {"epsilon", 0, false, false, epsDefault, "Epsilon used in ray tracing, defaults to the machine epsilon."},
        {"intersection-epsilon", 0, false, false, hitDefault, "Tolerance to avoid re-intersections of out-going rays."},
        {"no-gui", 0, true, false, "", "Disable display of render window (there is none here: ignored)."},
        {"rng", 0, false, false, "pixel", "'pixel': one random stream per pixel and sample slice (GPU schedule). 'tile': the reference's one "
                                          "stream per tile job, its image for a seed bit for bit (slow: a job is serial)."},
        {"device", 0, false, false, "0", "CUDA device ordinal."}});
    try {
      opts.parse(argc, argv);
    } catch (const std::runtime_error& e) {
      if (std::string(e.what()) == "Show help") return EXIT_FAILURE;
      throw;
    }
    const std::string fileName = opts.str("outfile");
    const std::string rng = opts.str("rng");
    if (rng != "pixel" && rng != "tile") throw std::runtime_error("--rng must be 'pixel' or 'tile'");
    {
      std::ofstream optsFile(fileName + ".command_line_options");
      for (auto& s : opts.specs())
        if (!s.isFlag && opts.has(s.name)) optsFile << s.name << " " << opts.str(s.name) << "\n";
    }

    b200pt_cornell_config cfg{};
    cfg.struct_size = sizeof(cfg);
    cfg.device = (int)opts.u32("device");
    cfg.width = opts.u32("width"); cfg.height = opts.u32("height");
    cfg.tile_width = opts.u32("tile-width"); cfg.tile_height = opts.u32("tile-height");
    cfg.samples = opts.u32("samples");
    cfg.roulette_depth = opts.u32("roulette-depth");
    cfg.seed = opts.u64("seed");
    cfg.aa_noise_scale = opts.f32("aa-noise-scale");
    cfg.refractive_index = opts.f32("refractive-index");
    cfg.stop_prob = opts.f32("stop-prob");
    cfg.epsilon = opts.f32("epsilon");
    cfg.intersection_epsilon = opts.f32("intersection-epsilon");
    cfg.rng_mode = rng == "tile" ? B200PT_CORNELL_TILE_STREAMS : B200PT_CORNELL_PIXEL_STREAMS;

    const size_t pixels = (size_t)cfg.width * cfg.height;
    std::vector<float> rgb(pixels * 3);
    b200pt_cornell_stats stats{};
    const auto startTime = std::chrono::steady_clock::now();
    if (b200pt_cornell_render(&cfg, rgb.data(), &stats) != B200PT_OK) throw std::runtime_error(b200pt_cornell_last_error());
    const double seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - startTime).count();

    // cvImageFromJobs at the last sample: scale (spp-1)/(spp-1) = 1, cv::Vec3b(float) saturates and rounds to nearest even
    std::vector<float> bgr(pixels * 3);
    std::vector<uint8_t> bgr8(pixels * 3);
    for (size_t i = 0; i < pixels; ++i)
      for (int c = 0; c < 3; ++c) {
        const float v = rgb[i * 3 + (2 - c)];
        bgr[i * 3 + c] = v;
        const float clipped = v < 255.f ? v : 255.f;
        bgr8[i * 3 + c] = (uint8_t)std::lrintf(clipped < 0.f ? 0.f : clipped);
      }
    image_io::writeExr(fileName + ".exr", bgr.data(), cfg.width, cfg.height);
    image_io::writePng(fileName, bgr8.data(), cfg.width, cfg.height);
    { std::ofstream hist(fileName + ".hist.txt"); hist << stats.max_path_length << "\n"; }

    std::cout << "\nRender time: " << seconds << " seconds (device " << stats.device_ms / 1000.0 << ")\n";
    std::cout << "\nTotal Rays: " << stats.total_ray_casts << " Rays/sec: " << stats.total_ray_casts / seconds
              << " Max path length: " << stats.max_path_length << "\n";
    std::cout << "Samples/sec: " << (double)pixels * cfg.samples / seconds << "\n";
    return EXIT_SUCCESS;
  } catch (const std::exception& e) {
    std::cerr << "Error: " << e.what() << "\n";
    return EXIT_FAILURE;
  }
}

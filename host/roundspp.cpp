// roundSamplesPerPixel for the GPU-free test library (PathTracerApp.cpp pulls in CUDA and NCCL).
#include <cstddef>

#include "logging.hpp"

std::size_t roundSamplesPerPixel(std::size_t samplesPerPixel, std::size_t samplesPerStep) {
  if (samplesPerPixel % samplesPerStep) {
    samplesPerPixel += samplesPerStep - (samplesPerPixel % samplesPerStep);
    logging::info("Rounding SPP to next multiple of {}  (Rounded SPP :=  {})", samplesPerStep, samplesPerPixel);
  }
  return samplesPerPixel;
}

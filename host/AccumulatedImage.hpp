// Host film (src/AccumulatedImage.hpp:13-33, src/AccumulatedImage.cpp): an fp32 image in OpenCV's
// channel order B,G,R that sums one estimate per render step and is divided by the step count when it
// is tone-mapped or saved. With the device film of libb200pt the same object only receives the gathered
// sum (setHdr) and does the saving.
#pragma once

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "LoadBalancer.hpp"

struct AccumulatedImage {
  AccumulatedImage(std::size_t w, std::size_t h);

  /// Tone map hdr / step: (x * 2^exposure)^(1/gamma) * 255, saturated (AccumulatedImage.cpp:23-46).
  const std::vector<std::uint8_t>& updateLdrImage(std::size_t step, float exposure, float gamma);

  /// fileName (PNG, tone-mapped) and <fileName without extension>.exr (hdr / step) (:48-56, :11-14).
  void saveImages(const std::string& fileName, std::size_t step, float exposure, float gamma);

  /// hdr(v,u) += (b,g,r) / sampleCount for every record inside the image (:59-74).
  void accumulate(const RecordList& traces);

  void reset();
  void setHdr(const float* bgr);                  // replace the sum (device film gather)
  float* hdrData() { return hdrImage.data(); }    // the gather writes the sum in place (page-locked by the app)
  const std::vector<float>& getHdrImage() const { return hdrImage; }
  std::size_t width() const { return cols; }
  std::size_t height() const { return rows; }

private:
  std::size_t cols, rows;
  std::vector<float> hdrImage;            // rows x cols x 3, B,G,R
  std::vector<std::uint8_t> image;        // tone-mapped, B,G,R
};

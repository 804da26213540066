#include "AccumulatedImage.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>

#include "image_io.hpp"

AccumulatedImage::AccumulatedImage(std::size_t w, std::size_t h) : cols(w), rows(h), hdrImage(w * h * 3, 0.f), image(w * h * 3, 0) {}

namespace {
// cv::Mat::convertTo(CV_8UC3, 255.0): saturate_cast<uchar>(cvRound(x * 255)) ; cvRound rounds half to even.
inline std::uint8_t toByte(float x) {
  const float y = x * 255.f;
  if (!(y > 0.f)) return 0;                // negative and NaN -> 0
  if (y >= 255.f) return 255;
  return (std::uint8_t)std::nearbyintf(y);
}
} // namespace

const std::vector<std::uint8_t>& AccumulatedImage::updateLdrImage(std::size_t step, float exposure, float gamma) {
  const float scale = 1.f / (float)step;
  const float exposureScale = std::pow(2.f, exposure);
  const float invGamma = 1.f / gamma;
  const std::size_t n = hdrImage.size();
#pragma omp parallel for schedule(static)
  for (std::size_t i = 0; i < n; ++i) image[i] = toByte(std::pow(hdrImage[i] * scale * exposureScale, invGamma));
  return image;
}

void AccumulatedImage::saveImages(const std::string& fileName, std::size_t step, float exposure, float gamma) {
  image_io::writePng(fileName, updateLdrImage(step, exposure, gamma).data(), (std::uint32_t)cols, (std::uint32_t)rows);
  // The image accumulated so far is divided by the number of steps so that the integrand is divided
  // by the total sample count:
  std::vector<float> scaled(hdrImage.size());
  const float scale = 1.f / (float)step;
#pragma omp parallel for schedule(static)
  for (std::size_t i = 0; i < scaled.size(); ++i) scaled[i] = hdrImage[i] * scale;
  image_io::writeExr(image_io::exrNameFor(fileName), scaled.data(), (std::uint32_t)cols, (std::uint32_t)rows);
}

void AccumulatedImage::accumulate(const RecordList& traces) {
#pragma omp parallel for schedule(static)
  for (std::size_t i = 0; i < traces.size(); ++i) {
    const auto& t = traces[i];
    const std::size_t c = t.u, r = t.v;
    if (c >= cols || r >= rows) continue;       // worklist padding
    const float scale = 1.f / (float)t.sampleCount;
    float* px = &hdrImage[(r * cols + c) * 3];
    px[0] += t.b * scale; px[1] += t.g * scale; px[2] += t.r * scale;
  }
}

void AccumulatedImage::reset() { std::fill(hdrImage.begin(), hdrImage.end(), 0.f); }
void AccumulatedImage::setHdr(const float* bgr) { std::memcpy(hdrImage.data(), bgr, hdrImage.size() * sizeof(float)); }

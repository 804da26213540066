// Headless image output of the host: 8-bit PNG and 32-bit float OpenEXR, written by hand (zlib only).
// Replaces the two cv::imwrite calls of the reference (src/AccumulatedImage.cpp:11-14,48-56): the
// film is kept in OpenCV's channel order B,G,R like the reference's cv::Mat, and both writers take
// that order and store what OpenCV would store (PNG: R,G,B bytes; EXR: channels named B, G, R).
#pragma once

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace image_io {

// bgr8: height x width x 3 bytes, B,G,R. Writes an 8-bit RGB PNG (filter 0, one zlib stream).
void writePng(const std::string& path, const uint8_t* bgr8, uint32_t width, uint32_t height);

// bgr: height x width x 3 floats, B,G,R. Writes a scan-line OpenEXR 2.0 file with three FLOAT
// channels (B, G, R), ZIP compression in blocks of 16 lines (stored raw where that is not smaller).
void writeExr(const std::string& path, const float* bgr, uint32_t width, uint32_t height);

// Readers for the files above (round-trip tests and the film re-load of a resumed render).
// Both return B,G,R interleaved.
void readPng(const std::string& path, std::vector<uint8_t>& bgr8, uint32_t& width, uint32_t& height);
void readExr(const std::string& path, std::vector<float>& bgr, uint32_t& width, uint32_t& height);

// "<name>.png" -> "<name>.exr" (saveHdrImage, AccumulatedImage.cpp:11-14).
std::string exrNameFor(const std::string& fileName);

} // namespace image_io

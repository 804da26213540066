#include "image_io.hpp"

#include <zlib.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <stdexcept>

namespace image_io {

namespace {

void putBe32(std::vector<uint8_t>& v, uint32_t x) {
  v.push_back((uint8_t)(x >> 24)); v.push_back((uint8_t)(x >> 16)); v.push_back((uint8_t)(x >> 8)); v.push_back((uint8_t)x);
}
uint32_t getBe32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

template <class T>
void putLe(std::vector<uint8_t>& v, T x) {
  uint8_t b[sizeof(T)];
  std::memcpy(b, &x, sizeof(T));       // the host is little-endian (x86-64)
  v.insert(v.end(), b, b + sizeof(T));
}
void putStr(std::vector<uint8_t>& v, const char* s) { while (*s) v.push_back((uint8_t)*s++); v.push_back(0); }

std::vector<uint8_t> deflateAll(const uint8_t* data, size_t n, int level) {
  uLongf cap = compressBound((uLong)n);
  std::vector<uint8_t> out(cap);
  if (compress2(out.data(), &cap, data, (uLong)n, level) != Z_OK) throw std::runtime_error("zlib compress2 failed");
  out.resize(cap);
  return out;
}

// zlib stream of `data`, deflated as independent bands on all cores (the pigz construction): every band is a raw
// deflate stream ended on a byte boundary by Z_SYNC_FLUSH, the last one by Z_FINISH; the Adler-32 of the whole is
// combined from the bands'. Small inputs take the single-stream path.
std::vector<uint8_t> deflateParallel(const uint8_t* data, size_t n, int level, size_t bandBytes) {
  const size_t bands = (n + bandBytes - 1) / bandBytes;
  if (bands < 2) return deflateAll(data, n, level);
  std::vector<std::vector<uint8_t>> parts(bands);
  std::vector<uLong> adlers(bands);
  bool failed = false;
#pragma omp parallel for schedule(dynamic)
  for (long b = 0; b < (long)bands; ++b) {
    const size_t off = (size_t)b * bandBytes, len = std::min(bandBytes, n - off);
    z_stream zs{};
    if (deflateInit2(&zs, level, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) { failed = true; continue; }
    auto& out = parts[(size_t)b];
    out.resize(deflateBound(&zs, (uLong)len) + 16);
    zs.next_in = const_cast<Bytef*>(data + off); zs.avail_in = (uInt)len;
    zs.next_out = out.data(); zs.avail_out = (uInt)out.size();
    const bool last = (size_t)b + 1 == bands;
    const int rc = deflate(&zs, last ? Z_FINISH : Z_SYNC_FLUSH);
    if ((last && rc != Z_STREAM_END) || (!last && (rc != Z_OK || zs.avail_in != 0))) failed = true;
    out.resize(out.size() - zs.avail_out);
    deflateEnd(&zs);
    adlers[(size_t)b] = adler32(adler32(0L, Z_NULL, 0), data + off, (uInt)len);
  }
  if (failed) throw std::runtime_error("zlib deflate failed");
  std::vector<uint8_t> out = {0x78, 0x5e};            // CM 8, 32 K window, FLEVEL 1; (0x785e % 31) == 0
  uLong adler = adler32(0L, Z_NULL, 0);
  for (size_t b = 0; b < bands; ++b) {
    out.insert(out.end(), parts[b].begin(), parts[b].end());
    adler = adler32_combine(adler, adlers[b], (z_off_t)std::min(bandBytes, n - b * bandBytes));
  }
  putBe32(out, (uint32_t)adler);
  return out;
}

std::vector<uint8_t> readFile(const std::string& path) {
  std::ifstream f(path, std::ios::binary);
  if (!f) throw std::runtime_error("Could not open file: '" + path + "'");
  return std::vector<uint8_t>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}

void writeFile(const std::string& path, const std::vector<uint8_t>& bytes) {
  // write to a temporary and rename, so that a viewer polling the file never sees half an image
  const std::string tmp = path + ".tmp";
  {
    std::ofstream f(tmp, std::ios::binary | std::ios::trunc);
    if (!f) throw std::runtime_error("Could not open file for writing: '" + path + "'");
    f.write(reinterpret_cast<const char*>(bytes.data()), (std::streamsize)bytes.size());
    if (!f) throw std::runtime_error("Short write: '" + path + "'");
  }
  if (std::rename(tmp.c_str(), path.c_str()) != 0) throw std::runtime_error("Could not rename '" + tmp + "' to '" + path + "'");
}

void pngChunk(std::vector<uint8_t>& out, const char type[4], const uint8_t* data, size_t n) {
  putBe32(out, (uint32_t)n);
  const size_t start = out.size();
  out.insert(out.end(), type, type + 4);
  if (n) out.insert(out.end(), data, data + n);
  const uint32_t crc = (uint32_t)crc32(0L, out.data() + start, (uInt)(n + 4));
  putBe32(out, crc);
}

const uint8_t kPngSig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};

} // namespace

std::string exrNameFor(const std::string& fileName) {
  return fileName.substr(0, fileName.find_last_of('.')) + ".exr";
}

// ---- PNG ---------------------------------------------------------------------------------------

void writePng(const std::string& path, const uint8_t* bgr8, uint32_t width, uint32_t height) {
  if (!width || !height) throw std::runtime_error("writePng: empty image");
  const size_t stride = (size_t)width * 3 + 1;
  std::vector<uint8_t> raw(stride * height);
#pragma omp parallel for
  for (long y = 0; y < (long)height; ++y) {
    uint8_t* dst = raw.data() + y * stride;
    *dst++ = 0;                                         // filter type 0 (None)
    const uint8_t* src = bgr8 + (size_t)y * width * 3;
    for (uint32_t x = 0; x < width; ++x, src += 3, dst += 3) { dst[0] = src[2]; dst[1] = src[1]; dst[2] = src[0]; }
  }
  std::vector<uint8_t> out(kPngSig, kPngSig + 8);
  std::vector<uint8_t> ihdr;
  putBe32(ihdr, width); putBe32(ihdr, height);
  ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);   // 8-bit, RGB, deflate, adaptive, no interlace
  pngChunk(out, "IHDR", ihdr.data(), ihdr.size());
  const std::vector<uint8_t> z = deflateParallel(raw.data(), raw.size(), 3, stride * std::max<size_t>(1, (512u << 10) / stride));
  pngChunk(out, "IDAT", z.data(), z.size());
  pngChunk(out, "IEND", nullptr, 0);
  writeFile(path, out);
}

void readPng(const std::string& path, std::vector<uint8_t>& bgr8, uint32_t& width, uint32_t& height) {
  const std::vector<uint8_t> f = readFile(path);
  if (f.size() < 8 || std::memcmp(f.data(), kPngSig, 8) != 0) throw std::runtime_error("readPng: not a PNG file: '" + path + "'");
  size_t p = 8;
  std::vector<uint8_t> z;
  width = height = 0;
  while (p + 12 <= f.size()) {
    const uint32_t n = getBe32(&f[p]);
    if (p + 12 + n > f.size()) throw std::runtime_error("readPng: truncated chunk");
    const char* type = reinterpret_cast<const char*>(&f[p + 4]);
    if ((uint32_t)crc32(0L, &f[p + 4], n + 4) != getBe32(&f[p + 8 + n])) throw std::runtime_error("readPng: CRC mismatch");
    if (!std::memcmp(type, "IHDR", 4)) {
      width = getBe32(&f[p + 8]); height = getBe32(&f[p + 12]);
      if (f[p + 16] != 8 || f[p + 17] != 2 || f[p + 20] != 0) throw std::runtime_error("readPng: only 8-bit non-interlaced RGB is supported");
    } else if (!std::memcmp(type, "IDAT", 4)) {
      z.insert(z.end(), &f[p + 8], &f[p + 8] + n);
    } else if (!std::memcmp(type, "IEND", 4)) {
      break;
    }
    p += 12 + n;
  }
  if (!width || !height) throw std::runtime_error("readPng: no IHDR");
  const size_t stride = (size_t)width * 3 + 1;
  std::vector<uint8_t> raw(stride * height);
  uLongf got = (uLongf)raw.size();
  if (uncompress(raw.data(), &got, z.data(), (uLong)z.size()) != Z_OK || got != raw.size()) throw std::runtime_error("readPng: inflate failed");
  bgr8.assign((size_t)width * height * 3, 0);
  std::vector<uint8_t> prev((size_t)width * 3, 0), cur((size_t)width * 3);
  for (uint32_t y = 0; y < height; ++y) {
    const uint8_t* src = raw.data() + y * stride;
    const uint8_t ft = *src++;
    for (size_t i = 0; i < cur.size(); ++i) {
      const int a = i >= 3 ? cur[i - 3] : 0, b = prev[i], c = i >= 3 ? prev[i - 3] : 0;
      int pred = 0;
      switch (ft) {
        case 0: pred = 0; break;
        case 1: pred = a; break;
        case 2: pred = b; break;
        case 3: pred = (a + b) / 2; break;
        case 4: { const int pp = a + b - c, pa = std::abs(pp - a), pb = std::abs(pp - b), pc = std::abs(pp - c);
                  pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); break; }
        default: throw std::runtime_error("readPng: bad filter type");
      }
      cur[i] = (uint8_t)(src[i] + pred);
    }
    uint8_t* dst = bgr8.data() + (size_t)y * width * 3;
    for (uint32_t x = 0; x < width; ++x) { dst[3 * x] = cur[3 * x + 2]; dst[3 * x + 1] = cur[3 * x + 1]; dst[3 * x + 2] = cur[3 * x]; }
    prev.swap(cur);
  }
}

// ---- OpenEXR (scan-line, single part) ---------------------------------------------------------------

namespace {

constexpr uint32_t kExrMagic = 20000630u;
constexpr int kZipLines = 16;

void exrAttr(std::vector<uint8_t>& h, const char* name, const char* type, const std::vector<uint8_t>& value) {
  putStr(h, name); putStr(h, type);
  putLe<int32_t>(h, (int32_t)value.size());
  h.insert(h.end(), value.begin(), value.end());
}

// OpenEXR's ZIP pre-conditioning: split the bytes into even/odd halves, then delta-encode.
std::vector<uint8_t> zipPrecondition(const uint8_t* raw, size_t n) {
  std::vector<uint8_t> t(n);
  uint8_t* t1 = t.data();
  uint8_t* t2 = t.data() + (n + 1) / 2;
  for (size_t i = 0; i < n; ++i) { if (i & 1) *t2++ = raw[i]; else *t1++ = raw[i]; }
  int p = t[0];
  for (size_t i = 1; i < n; ++i) {
    const int d = (int)t[i] - p + (128 + 256);
    p = t[i];
    t[i] = (uint8_t)d;
  }
  return t;
}

void zipUncondition(std::vector<uint8_t>& t, uint8_t* out) {
  const size_t n = t.size();
  for (size_t i = 1; i < n; ++i) t[i] = (uint8_t)((int)t[i - 1] + (int)t[i] - 128);
  const uint8_t* t1 = t.data();
  const uint8_t* t2 = t.data() + (n + 1) / 2;
  for (size_t i = 0; i < n; ++i) out[i] = (i & 1) ? *t2++ : *t1++;
}

} // namespace

void writeExr(const std::string& path, const float* bgr, uint32_t width, uint32_t height) {
  if (!width || !height) throw std::runtime_error("writeExr: empty image");
  std::vector<uint8_t> out;
  putLe<uint32_t>(out, kExrMagic);
  putLe<uint32_t>(out, 2u);                                  // version 2, single-part scan-line, short names
  {
    std::vector<uint8_t> ch;
    for (const char* name : {"B", "G", "R"}) {               // alphabetical, as the format requires
      putStr(ch, name);
      putLe<int32_t>(ch, 2);                                 // FLOAT
      ch.push_back(0); ch.push_back(0); ch.push_back(0); ch.push_back(0);   // pLinear + reserved
      putLe<int32_t>(ch, 1); putLe<int32_t>(ch, 1);          // x/y sampling
    }
    ch.push_back(0);
    exrAttr(out, "channels", "chlist", ch);
  }
  exrAttr(out, "compression", "compression", {3});           // ZIP_COMPRESSION (16 lines per block)
  {
    std::vector<uint8_t> box;
    putLe<int32_t>(box, 0); putLe<int32_t>(box, 0); putLe<int32_t>(box, (int32_t)width - 1); putLe<int32_t>(box, (int32_t)height - 1);
    exrAttr(out, "dataWindow", "box2i", box);
    exrAttr(out, "displayWindow", "box2i", box);
  }
  exrAttr(out, "lineOrder", "lineOrder", {0});               // INCREASING_Y
  { std::vector<uint8_t> v; putLe<float>(v, 1.f); exrAttr(out, "pixelAspectRatio", "float", v); }
  { std::vector<uint8_t> v; putLe<float>(v, 0.f); putLe<float>(v, 0.f); exrAttr(out, "screenWindowCenter", "v2f", v); }
  { std::vector<uint8_t> v; putLe<float>(v, 1.f); exrAttr(out, "screenWindowWidth", "float", v); }
  out.push_back(0);                                          // end of header

  const uint32_t blocks = (height + kZipLines - 1) / kZipLines;
  const size_t tablePos = out.size();
  out.resize(out.size() + (size_t)blocks * 8, 0);

  const size_t lineBytes = (size_t)width * 3 * sizeof(float);
  // blocks are independent: pre-condition and deflate them in parallel, then lay them out in order
  std::vector<std::vector<uint8_t>> packed(blocks);
#pragma omp parallel for schedule(dynamic)
  for (int64_t bi = 0; bi < (int64_t)blocks; ++bi) {
    const uint32_t b = (uint32_t)bi;
    const uint32_t y0 = b * kZipLines, y1 = std::min(height, y0 + kZipLines);
    std::vector<uint8_t> raw(lineBytes * (y1 - y0));
    float* dst = reinterpret_cast<float*>(raw.data());
    for (uint32_t y = y0; y < y1; ++y)
      for (int c = 0; c < 3; ++c) {                           // planar per line: all B, all G, all R
        const float* src = bgr + (size_t)y * width * 3 + c;
        for (uint32_t x = 0; x < width; ++x) *dst++ = src[(size_t)x * 3];
      }
    const std::vector<uint8_t> pre = zipPrecondition(raw.data(), raw.size());
    // fp32 Monte-Carlo noise barely compresses (a 4K render: 0.6 %), and deflating it then costs more than the rest of
    // the save. Probe four 16 KiB pieces spread over the pre-conditioned block at level 1; a block that would shrink by
    // less than 5 % is stored raw, which is the format's own rule for blocks that do not shrink.
    bool worthDeflating = true;
    constexpr size_t kPiece = 16u << 10;
    if (pre.size() >= 16 * kPiece) {
      std::vector<uint8_t> probe(4 * kPiece);
      for (size_t k = 0; k < 4; ++k) std::memcpy(probe.data() + k * kPiece, pre.data() + k * (pre.size() / 4), kPiece);
      worthDeflating = deflateAll(probe.data(), probe.size(), 1).size() < probe.size() - probe.size() / 20;
    }
    if (worthDeflating) {
      std::vector<uint8_t> z = deflateAll(pre.data(), pre.size(), 4);
      packed[b] = z.size() < raw.size() ? std::move(z) : std::move(raw);
    } else {
      packed[b] = std::move(raw);
    }
  }
  for (uint32_t b = 0; b < blocks; ++b) {
    const uint64_t off = out.size();
    std::memcpy(out.data() + tablePos + (size_t)b * 8, &off, 8);
    putLe<int32_t>(out, (int32_t)(b * kZipLines));
    putLe<int32_t>(out, (int32_t)packed[b].size());
    out.insert(out.end(), packed[b].begin(), packed[b].end());
  }
  writeFile(path, out);
}

void readExr(const std::string& path, std::vector<float>& bgr, uint32_t& width, uint32_t& height) {
  const std::vector<uint8_t> f = readFile(path);
  auto need = [&](size_t p, size_t n) { if (p + n > f.size()) throw std::runtime_error("readExr: truncated file: '" + path + "'"); };
  need(0, 8);
  uint32_t magic, version;
  std::memcpy(&magic, &f[0], 4); std::memcpy(&version, &f[4], 4);
  if (magic != kExrMagic) throw std::runtime_error("readExr: not an OpenEXR file: '" + path + "'");
  if ((version & 0xff) != 2 || (version & 0x1a00)) throw std::runtime_error("readExr: only single-part scan-line files are supported");
  size_t p = 8;
  auto cstr = [&]() { std::string s; while (true) { need(p, 1); const char c = (char)f[p++]; if (!c) break; s += c; } return s; };
  int compression = -1, lineOrder = 0;
  int32_t win[4] = {0, 0, -1, -1};
  std::vector<std::pair<std::string, int>> channels;          // name, pixel type
  while (true) {
    const std::string name = cstr();
    if (name.empty()) break;
    const std::string type = cstr();
    need(p, 4);
    int32_t size; std::memcpy(&size, &f[p], 4); p += 4;
    need(p, (size_t)size);
    if (name == "channels") {
      size_t q = p;
      while (f[q]) {
        std::string cn; while (f[q]) cn += (char)f[q++];
        ++q;
        int32_t pt; std::memcpy(&pt, &f[q], 4);
        int32_t xs, ys; std::memcpy(&xs, &f[q + 8], 4); std::memcpy(&ys, &f[q + 12], 4);
        if (xs != 1 || ys != 1) throw std::runtime_error("readExr: sub-sampled channels are not supported");
        channels.emplace_back(cn, pt);
        q += 16;
      }
    } else if (name == "compression") compression = f[p];
    else if (name == "dataWindow") std::memcpy(win, &f[p], 16);
    else if (name == "lineOrder") lineOrder = f[p];
    p += (size_t)size;
  }
  if (compression != 0 && compression != 2 && compression != 3) throw std::runtime_error("readExr: only NONE/ZIPS/ZIP compression is supported");
  if (lineOrder > 1) throw std::runtime_error("readExr: random line order is not supported");
  width = (uint32_t)(win[2] - win[0] + 1); height = (uint32_t)(win[3] - win[1] + 1);
  int slot[3] = {-1, -1, -1};
  size_t pixelBytes = 0;
  std::vector<size_t> chanOff(channels.size());
  for (size_t i = 0; i < channels.size(); ++i) {
    if (channels[i].second != 2 && channels[i].second != 1) throw std::runtime_error("readExr: only HALF/FLOAT channels are supported");
    chanOff[i] = pixelBytes;
    pixelBytes += channels[i].second == 2 ? 4 : 2;
    if (channels[i].first == "B") slot[0] = (int)i;
    if (channels[i].first == "G") slot[1] = (int)i;
    if (channels[i].first == "R") slot[2] = (int)i;
  }
  if (slot[0] < 0 || slot[1] < 0 || slot[2] < 0) throw std::runtime_error("readExr: file has no B, G, R channels");
  const int linesPerBlock = compression == 3 ? kZipLines : 1;
  const uint32_t blocks = (height + linesPerBlock - 1) / linesPerBlock;
  need(p, (size_t)blocks * 8);
  const size_t lineBytes = (size_t)width * pixelBytes;
  bgr.assign((size_t)width * height * 3, 0.f);
  std::vector<uint8_t> raw, tmp;
  for (uint32_t b = 0; b < blocks; ++b) {
    uint64_t off; std::memcpy(&off, &f[p + (size_t)b * 8], 8);
    need((size_t)off, 8);
    int32_t y0, size; std::memcpy(&y0, &f[off], 4); std::memcpy(&size, &f[off + 4], 4);
    need((size_t)off + 8, (size_t)size);
    y0 -= win[1];
    const uint32_t lines = std::min<uint32_t>((uint32_t)linesPerBlock, height - (uint32_t)y0);
    raw.resize(lineBytes * lines);
    if ((size_t)size == raw.size()) {
      std::memcpy(raw.data(), &f[off + 8], raw.size());
    } else {
      tmp.resize(raw.size());
      uLongf got = (uLongf)tmp.size();
      if (uncompress(tmp.data(), &got, &f[off + 8], (uLong)size) != Z_OK || got != tmp.size()) throw std::runtime_error("readExr: inflate failed");
      zipUncondition(tmp, raw.data());
    }
    for (uint32_t l = 0; l < lines; ++l) {
      const uint8_t* line = raw.data() + l * lineBytes;
      for (int c = 0; c < 3; ++c) {
        const size_t ci = (size_t)slot[c];
        // channel planes follow each other inside a line in file (alphabetical) order
        size_t planeOff = 0;
        for (size_t i = 0; i < ci; ++i) planeOff += (size_t)width * (channels[i].second == 2 ? 4 : 2);
        const uint8_t* src = line + planeOff;
        float* dst = bgr.data() + ((size_t)(y0 + (int32_t)l) * width) * 3 + c;
        if (channels[ci].second == 2) {
          for (uint32_t x = 0; x < width; ++x) std::memcpy(dst + (size_t)x * 3, src + (size_t)x * 4, 4);
        } else {
          for (uint32_t x = 0; x < width; ++x) {
            uint16_t h; std::memcpy(&h, src + (size_t)x * 2, 2);
            const uint32_t sign = (uint32_t)(h & 0x8000u) << 16, exp = (h >> 10) & 0x1f, man = h & 0x3ffu;
            uint32_t bits;
            if (exp == 0) {
              if (man == 0) bits = sign;
              else { int e = -1; uint32_t m = man; do { ++e; m <<= 1; } while (!(m & 0x400u)); bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((m & 0x3ffu) << 13); }
            } else if (exp == 31) bits = sign | 0x7f800000u | (man << 13);
            else bits = sign | ((exp + 112u) << 23) | (man << 13);
            std::memcpy(dst + (size_t)x * 3, &bits, 4);
          }
        }
      }
    }
  }
}

} // namespace image_io

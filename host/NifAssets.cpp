#include "NifAssets.hpp"

#include <cmath>
#include <cstring>
#include <fstream>
#include <random>
#include <sstream>
#include <stdexcept>

#include "Hdf5File.hpp"
#include "json.hpp"
#include "logging.hpp"

namespace {

std::string slurp(const std::string& file) {
  std::ifstream f(file, std::ios::binary);
  if (!f) throw std::runtime_error("Could not open file: '" + file + "'");
  std::ostringstream ss;
  ss << f.rdbuf();
  return ss.str();
}

bool exists(const std::string& file) { return std::ifstream(file).good(); }

std::uint16_t floatToHalf(float f) {           // round to nearest even
  std::uint32_t x; std::memcpy(&x, &f, 4);
  const std::uint32_t sign = (x >> 16) & 0x8000u;
  const std::int32_t exp = (std::int32_t)((x >> 23) & 0xff) - 127 + 15;
  std::uint32_t man = x & 0x7fffffu;
  if (((x >> 23) & 0xff) == 0xff) return (std::uint16_t)(sign | 0x7c00u | (man ? 0x200u : 0));
  if (exp >= 31) return (std::uint16_t)(sign | 0x7c00u);
  if (exp <= 0) {
    if (exp < -10) return (std::uint16_t)sign;
    man |= 0x800000u;
    const int shift = 14 - exp;
    std::uint32_t h = man >> shift;
    const std::uint32_t rem = man & ((1u << shift) - 1), halfway = 1u << (shift - 1);
    if (rem > halfway || (rem == halfway && (h & 1))) ++h;
    return (std::uint16_t)(sign | h);
  }
  std::uint32_t h = ((std::uint32_t)exp << 10) | (man >> 13);
  const std::uint32_t rem = man & 0x1fffu;
  if (rem > 0x1000u || (rem == 0x1000u && (h & 1))) ++h;
  return (std::uint16_t)(sign | h);
}

} // namespace

NifMetaData::NifMetaData(const std::string& file) {
  const std::string text = slurp(file);
  json::Value pt;
  try {
    pt = json::parse(text);
  } catch (const std::exception& e) {
    throw std::runtime_error(std::string(e.what()) + " in file: '" + file + "'");
  }
  if (pt.kind != json::Value::Object || pt.members.empty()) throw std::runtime_error("Empty property tree after parsing file: '" + file + "'");
  try {
    embeddingDimension = (std::size_t)pt.at("embedding_dimension").asInt();
    name = pt.at("name").asString();
    for (auto& p : pt.at("original_image_shape").items) imageShape.push_back((int)p.asInt());
    const json::Value& enc = pt.at("encode_params");
    eps = (float)enc.at("eps").asDouble();
    logToneMap = enc.at("log_tone_map").asBool();
    max = (float)enc.at("max").asDouble();
    mean.clear();
    for (auto& p : enc.at("mean").items) mean.push_back((float)p.asDouble());
    if (mean.size() != 3) throw std::runtime_error("encode_params.mean must have 3 entries");
    // If tone-mapping fold the inverse eps into the mean (NifMetaData.cpp:48-53):
    if (logToneMap) { mean[0] -= eps; mean[1] -= eps; mean[2] -= eps; }
    bool next = false;
    for (auto& p : pt.at("train_command").items) {
      if (next) { hiddenSize = (std::size_t)std::atoi(p.asString().c_str()); next = false; }
      if (p.asString() == "--layer-size") next = true;
    }
  } catch (const std::exception& e) {
    std::stringstream ss;
    ss << "Error reading property: " << e.what() << " from file: '" << file << "'";
    throw std::runtime_error(ss.str());
  }
}

NifModelData::NifModelData(const std::string& assetPath, bool allowSynthetic) : metaData(assetPath + "/nif_metadata.txt") {
  logging::info("Loading model metadata from file: '{}'", assetPath + "/nif_metadata.txt");
  logging::debug("Loaded NIF metadata for model name: {}", metaData.name);
  logging::debug("NIF embedding dimension: {}", metaData.embeddingDimension);
  logging::debug("NIF hidden dimension: {}", metaData.hiddenSize);
  const std::string h5File = assetPath + "/converted.hdf5", binFile = assetPath + "/nif.bin";
  if (exists(h5File)) { loadHdf5(h5File); source = h5File; }
  else if (exists(binFile)) { loadNifBin(binFile); source = binFile; }
  else if (allowSynthetic) { makeSynthetic(); source = "synthetic (seed 42)"; }
  else throw std::runtime_error("No weights in '" + assetPath + "': neither converted.hdf5 nor nif.bin (use --synthetic-weights for a seeded stand-in)");
  // shape checks the device upload would also do, with the file named in the message
  std::size_t width = 4 * metaData.embeddingDimension;
  for (auto& l : layers) {
    if (l.in != width && l.in != width + 4 * metaData.embeddingDimension)
      throw std::runtime_error("Layer '" + l.name + "' of " + source + " takes " + std::to_string(l.in) + " inputs, the previous layer gives " + std::to_string(width));
    width = l.out;
  }
  logging::info("NIF weights from {}: {} Dense layers, {} FLOPs per evaluation", source, layers.size(), flopsPerEval());
}

void NifModelData::loadHdf5(const std::string& file) {
  hdf5::File hdf(file);
  logging::info("Reading weights saved from '{}', keras_version {}, backend {}", file, hdf.stringAttribute("/", "keras_version"),
                hdf.stringAttribute("/", "backend"));
  const json::Value cfg = json::parse(hdf.stringAttribute("/", "model_config"));
  // Hdf5Model.cpp:16-58: a Functional model; Dense layers in order, InputLayer/Concatenate ignored.
  if (cfg.at("class_name").asString() != "Functional") throw std::runtime_error("Expdected a Keras 'Functional' Model");
  for (auto& p : cfg.at("config").at("layers").items) {
    const std::string cn = p.at("class_name").asString();
    if (cn == "Dense") {
      const json::Value& c = p.at("config");
      NifLayerData l;
      l.name = c.at("name").asString();
      std::string act = c.at("activation").asString();
      if (act == "linear") act = "none";                        // NifModel.cpp:73-76
      if (act != "relu" && act != "none") throw std::runtime_error("Activation '" + act + "' of layer '" + l.name + "' is not supported.");
      l.relu = act == "relu";
      const bool useBias = c.at("use_bias").asBool();
      const std::string base = "/model_weights/" + l.name + "/" + l.name + "/";
      hdf5::Dataset k = hdf.dataset(base + "kernel:0");
      if (k.shape.size() != 2) throw std::runtime_error("Kernel of layer '" + l.name + "' is not a matrix.");
      if (k.elementSize != 2 && k.elementSize != 4) throw std::runtime_error("Only float32 and float16 weights are supported.");
      l.in = k.shape[0]; l.out = k.shape[1]; l.halfFloat = k.elementSize == 2;
      if (l.out != (std::size_t)c.at("units").asInt()) throw std::runtime_error("Kernel shape of layer '" + l.name + "' disagrees with its config.");
      l.kernel = std::move(k.data);
      if (useBias) {
        hdf5::Dataset b = hdf.dataset(base + "bias:0");
        if (b.elementSize != k.elementSize || b.data.size() != l.out * b.elementSize) throw std::runtime_error("Bias of layer '" + l.name + "' has the wrong size or type.");
        l.bias = std::move(b.data);
      }
      logging::debug("Layer: {} {} (act: {} bias: {})", cn, l.name, act, useBias);
      layers.push_back(std::move(l));
    } else if (cn == "InputLayer" || cn == "Concatenate") {
      logging::warn("Ignoring layer (classname: {})", cn);
    } else {
      throw std::runtime_error("Layer class: '" + cn + "' not supported by Hdf5Model loader.");
    }
  }
  if (layers.empty()) throw std::runtime_error("No Dense layers in '" + file + "'");
}

void NifModelData::loadNifBin(const std::string& file) {
  const std::string d = slurp(file);
  if (d.size() < 36 || d.compare(0, 8, "NIFBIN01") != 0) throw std::runtime_error(file + ": not a nif.bin file");
  std::size_t off = 8;
  auto rd32 = [&]() { if (off + 4 > d.size()) throw std::runtime_error(file + ": truncated"); std::uint32_t v; std::memcpy(&v, d.data() + off, 4); off += 4; return v; };
  const std::uint32_t n = rd32();
  rd32();             // E, max, mean, log_tone_map: the metadata file is authoritative (as in the reference)
  off += 4 * 5;
  for (std::uint32_t i = 0; i < n; ++i) {
    NifLayerData l;
    l.name = "dense_" + std::to_string(i);
    l.in = rd32(); l.out = rd32(); l.relu = rd32() != 0;
    const bool hasBias = rd32() != 0;
    const std::uint32_t dt = rd32();
    if (dt > 1) throw std::runtime_error(file + ": unknown dtype");
    l.halfFloat = dt == 0;
    const std::size_t es = l.halfFloat ? 2 : 4;
    if (off + l.in * l.out * es + (hasBias ? l.out * es : 0) > d.size()) throw std::runtime_error(file + ": truncated");
    l.kernel.assign(d.begin() + off, d.begin() + off + l.in * l.out * es); off += l.in * l.out * es;
    if (hasBias) { l.bias.assign(d.begin() + off, d.begin() + off + l.out * es); off += l.out * es; }
    layers.push_back(std::move(l));
  }
}

// Seeded stand-in in the topology of the shipped model's train_command (--layer-count 6 --layer-size H,
// skip-concat of the features into the 4th hidden layer, 3 outputs): He-uniform kernels, small biases.
void NifModelData::makeSynthetic() {
  const std::size_t nFeat = 4 * metaData.embeddingDimension, H = metaData.hiddenSize ? metaData.hiddenSize : 320;
  std::mt19937 g(42);
  auto uniform = [&](float lim) { return std::uniform_real_distribution<float>(-lim, lim)(g); };
  std::size_t width = nFeat;
  for (int l = 0; l < 7; ++l) {
    NifLayerData d;
    d.name = "dense_" + std::to_string(l);
    const bool last = l == 6;
    d.in = width + (l == 3 ? nFeat : 0); d.out = last ? 3 : H; d.relu = !last; d.halfFloat = true;
    const float lim = (last ? 0.5f : 1.f) * std::sqrt(6.f / (float)d.in);
    d.kernel.resize(d.in * d.out * 2); d.bias.resize(d.out * 2);
    for (std::size_t i = 0; i < d.in * d.out; ++i) { const std::uint16_t h = floatToHalf(uniform(lim)); std::memcpy(&d.kernel[2 * i], &h, 2); }
    for (std::size_t i = 0; i < d.out; ++i) { const std::uint16_t h = floatToHalf(uniform(0.1f)); std::memcpy(&d.bias[2 * i], &h, 2); }
    width = d.out;
    layers.push_back(std::move(d));
  }
}

b200pt_nif_desc NifModelData::descriptor() {
  abiLayers.clear();
  for (auto& l : layers) {
    b200pt_nif_layer a{};
    a.in = (std::int32_t)l.in; a.out = (std::int32_t)l.out; a.relu = l.relu ? 1 : 0; a.dtype = l.halfFloat ? 0 : 1;
    a.kernel = l.kernel.data(); a.bias = l.bias.empty() ? nullptr : l.bias.data();
    abiLayers.push_back(a);
  }
  b200pt_nif_desc d{};
  d.num_layers = (std::int32_t)abiLayers.size();
  d.embedding_dimension = (std::int32_t)metaData.embeddingDimension;
  d.max = metaData.max;
  d.mean[0] = metaData.mean[0]; d.mean[1] = metaData.mean[1]; d.mean[2] = metaData.mean[2];
  d.log_tone_map = metaData.logToneMap ? 1 : 0;
  d.layers = abiLayers.data();
  return d;
}

std::size_t NifModelData::flopsPerEval() const {
  std::size_t f = 0;
  for (auto& l : layers) f += 2 * l.in * l.out + (l.bias.empty() ? 0 : l.out);
  return f;
}

// TEST INFRASTRUCTURE ONLY (oracle/). Built into oracle/_ref/libref_niftwins.so, only where /root/reference exists.
//
// The reference has no CPU inference of the NIF, but it does have HOST-SIDE TWINS of the two ends of it, used by its
// stand-alone image reconstruction: NifModel::generateInputSamples + makeCoefficients + makeGridCoordsUV (the Fourier-feature
// encoding, src/neural_networks/NifModel.cpp:403-434,466-490) and NifModel::decodeSamples (:492-509). Those member functions
// are cut out of NifModel.cpp at build time (oracle/Makefile: sed into oracle/_ref/nif_host_twins.inc - the file itself
// cannot be compiled, it builds Poplar graphs) and compiled UNMODIFIED against a stand-in class that has exactly the
// members they touch; neural_networks/IoBuffer.cpp is compiled unmodified as it lies. This pins the feature order
// [sin u, sin v, cos u, cos v], the coefficients 2^j, the (uv - 1) * 2 mapping, u = row / height as the vertical
// coordinate, and the decode x * max + mean -> exp, against reference code (tests/test_nif_reference_twins.py).
// What it cannot pin is the device graph's binary16 rounding (NifModel.cpp:211-213,314-324).
#include <cmath>
#include <memory>
#include <stdexcept>
#include <string>
#include <tuple>
#include <utility>
#include <vector>

#include <io_utils.hpp>                      // oracle/shim_nif: ipu_utils::logger()
#include <neural_networks/IoBuffer.hpp>      // the reference's, unmodified

// Same field names as src/neural_networks/NifMetaData.hpp (whose .cpp needs boost::property_tree, absent here).
struct NifMetaData {
  std::size_t embeddingDimension = 0;
  std::vector<std::size_t> imageShape;
  float max = 1.f;
  std::vector<float> mean;
  float eps = 0.f;
  bool logToneMap = false;
};

struct NifModel {
  struct Data {
    NifMetaData metaData;
    const NifMetaData& getMetaData() const { return metaData; }
  };
  std::shared_ptr<Data> data;
  std::string name = "oracle";
  bool decodeOnDevice = false;
  std::unique_ptr<IoBuffer> inputBuffer, inputBufferU, inputBufferV, outputBuffer;

  void generateInputSamples();
  bool prepareNextBatch();
  std::vector<float> makeCoefficients();
  std::pair<std::vector<float>, std::vector<float>> makeGridCoordsUV();
  const std::vector<std::vector<float>>& decodeSamples();
};

#include "nif_host_twins.inc"   // reference source, generated into oracle/_ref/ at build time

extern "C" {

// features: [height * width][4 * E] fp32; u, v: the raw grid coordinates (u = row / height, v = column / width)
int refnif_encode_grid(int height, int width, int E, float* features, float* u, float* v) {
  try {
    NifModel m;
    m.data = std::make_shared<NifModel::Data>();
    m.data->metaData.embeddingDimension = (std::size_t)E;
    m.data->metaData.imageShape = {(std::size_t)height, (std::size_t)width, 3};
    const std::size_t n = (std::size_t)height * width;
    m.inputBuffer.reset(new IoBuffer(n, 4 * (std::size_t)E, n));
    m.inputBufferU.reset(new IoBuffer(n, 1, n));
    m.inputBufferV.reset(new IoBuffer(n, 1, n));
    m.generateInputSamples();
    for (std::size_t i = 0; i < n; ++i) {
      for (int k = 0; k < 4 * E; ++k) features[i * 4 * E + k] = m.inputBuffer->data[i][k];
      u[i] = m.inputBufferU->data[i][0];
      v[i] = m.inputBufferV->data[i][0];
    }
    return 0;
  } catch (const std::exception&) { return -1; }
}

// in/out: [n][3] network outputs (B,G,R) -> decoded radiance, by NifModel::decodeSamples
int refnif_decode(int n, const float* in, float maxv, const float* mean3, float eps, int logToneMap, float* out) {
  try {
    NifModel m;
    m.data = std::make_shared<NifModel::Data>();
    m.data->metaData.max = maxv;
    m.data->metaData.mean = {mean3[0], mean3[1], mean3[2]};
    m.data->metaData.eps = eps;
    m.data->metaData.logToneMap = logToneMap != 0;
    m.outputBuffer.reset(new IoBuffer((std::size_t)n, 3, (std::size_t)n));
    for (int i = 0; i < n; ++i) for (int c = 0; c < 3; ++c) m.outputBuffer->data[i][c] = in[3 * i + c];
    const auto& res = m.decodeSamples();
    for (int i = 0; i < n; ++i) for (int c = 0; c < 3; ++c) out[3 * i + c] = res[i][c];
    return 0;
  } catch (const std::exception&) { return -1; }
}

}  // extern "C"

// Host side of the neural image field: what the reference's loader hands to "init_nif_weights".
//   NifMetaData     src/neural_networks/NifMetaData.cpp:11-71 (JSON fields; eps folded into mean :48-53;
//                   hidden size scraped from train_command :56-64)
//   NifModelData    src/neural_networks/NifModel.cpp:40-85 (Dense layers in file order, Keras [in,out]
//                   kernels, "linear" -> none :73-76) fed by src/keras/Hdf5Model.cpp:8-133
// Weights are read from <assets>/converted.hdf5 (Keras HDF5, minimal reader in Hdf5File.hpp) or, if that
// is absent, from <assets>/nif.bin (a flat container written by ipu_path_trace_b200.nif.write_nif_bin).
#pragma once

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "../include/b200pt.h"

struct NifMetaData {
  explicit NifMetaData(const std::string& file);
  NifMetaData() = default;
  std::string name;
  std::size_t embeddingDimension = 0;
  std::size_t hiddenSize = 0;
  std::vector<int> imageShape;
  float eps = 0.f;
  float max = 0.f;
  std::vector<float> mean;        // eps already folded in when logToneMap
  bool logToneMap = false;
};

struct NifLayerData {
  std::string name;
  std::size_t in = 0, out = 0;
  bool relu = false;
  bool halfFloat = true;                 // storage type of kernel/bias
  std::vector<std::uint8_t> kernel;      // [in,out] row-major bytes
  std::vector<std::uint8_t> bias;        // [out] bytes or empty
};

struct NifModelData {
  /// assetPath is the 'assets.extra' directory (PathTracerApp.cpp:106-110). Throws std::runtime_error.
  NifModelData(const std::string& assetPath, bool allowSynthetic);
  NifMetaData metaData;
  std::vector<NifLayerData> layers;
  std::string source;                    // where the weights came from

  /// Descriptor for b200pt_upload_nif; valid while this object is alive.
  b200pt_nif_desc descriptor();
  /// FLOPs of one evaluation (NifModel.cpp:122-134).
  std::size_t flopsPerEval() const;

private:
  std::vector<b200pt_nif_layer> abiLayers;
  void loadNifBin(const std::string& file);
  void loadHdf5(const std::string& file);
  void makeSynthetic();
};

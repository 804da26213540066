// Read-only HDF5 subset, enough for what src/keras/Hdf5Model.cpp:62-133 asks of libhdf5 on a Keras
// "h5" weights file (no HDF5 library exists in this image):
//   * superblock version 0/1 (h5py's default "earliest" format) or 2/3
//   * version 1 object headers (with continuation blocks); version 2 ("OHDR") headers with compact links
//   * old-style groups: symbol-table message -> version 1 B-tree + local heap + SNOD nodes
//   * datasets with contiguous, compact or chunked (version 1 B-tree; deflate and shuffle filters) layout,
//     little-endian IEEE floats of 2, 4 or 8 bytes (any fixed-size type is returned as raw bytes)
//   * string attributes: fixed-length, or variable-length through the global heap
// Anything else (fractal-heap groups, virtual/external storage, other filters) raises std::runtime_error
// naming the feature, never a wrong answer.
#pragma once

#include <cstddef>
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace hdf5 {

struct Dataset {
  std::vector<std::size_t> shape;
  std::size_t elementSize = 0;       // bytes per element
  int typeClass = -1;                // HDF5 datatype class (1 = floating point, 0 = fixed point, 3 = string ...)
  std::vector<std::uint8_t> data;    // row-major, file byte order (little-endian checked for floats)
  std::size_t numElements() const { std::size_t n = 1; for (auto d : shape) n *= d; return n; }
};

class File {
public:
  explicit File(const std::string& path);

  /// String attribute of the object at `objectPath` ("/" = root group). Hdf5Model::readStringAttribute.
  std::string stringAttribute(const std::string& objectPath, const std::string& name) const;
  bool hasAttribute(const std::string& objectPath, const std::string& name) const;
  /// H5File::openDataSet + DataSet::read.
  Dataset dataset(const std::string& path) const;
  /// Names of the links of a group, in B-tree (alphabetical) order.
  std::vector<std::string> list(const std::string& groupPath) const;
  bool exists(const std::string& path) const;

private:
  struct Message { std::uint16_t type; std::uint8_t flags; std::vector<std::uint8_t> body; };
  struct Datatype { int cls = -1; std::size_t size = 0; int vlenKind = -1; bool littleEndian = true; int strPad = 0; };
  struct Dataspace { std::vector<std::size_t> dims; bool scalar = false; bool null = false; };

  std::vector<std::uint8_t> bytes_;
  std::string path_;
  unsigned sizeOfOffsets_ = 8, sizeOfLengths_ = 8;
  std::uint64_t base_ = 0, rootHeader_ = 0;

  [[noreturn]] void fail(const std::string& what) const;
  void need(std::uint64_t off, std::uint64_t n) const;
  std::uint64_t rd(std::uint64_t off, unsigned n) const;
  std::uint64_t rdOffset(std::uint64_t off) const { return rd(off, sizeOfOffsets_); }
  std::uint64_t rdLength(std::uint64_t off) const { return rd(off, sizeOfLengths_); }
  bool undefinedAddress(std::uint64_t a) const;

  std::vector<Message> objectHeader(std::uint64_t addr) const;
  void parseV1Messages(std::uint64_t off, std::uint64_t len, unsigned& remaining, std::vector<Message>& out) const;
  void parseV2Messages(std::uint64_t off, std::uint64_t end, bool creationOrder, std::vector<Message>& out) const;
  std::map<std::string, std::uint64_t> links(std::uint64_t groupHeader, std::vector<std::string>* order) const;
  void walkGroupBtree(std::uint64_t node, std::uint64_t heapData, std::map<std::string, std::uint64_t>& out, std::vector<std::string>* order) const;
  std::uint64_t resolve(const std::string& path) const;
  Datatype parseDatatype(const std::uint8_t* p, std::size_t n) const;
  Dataspace parseDataspace(const std::uint8_t* p, std::size_t n) const;
  std::string globalHeapObject(std::uint64_t collection, std::uint32_t index) const;
  void readChunked(const Message& layout, const std::vector<Message>& msgs, Dataset& ds) const;
  void walkChunkBtree(std::uint64_t node, unsigned rank, const std::vector<std::size_t>& chunkDims, const std::vector<int>& filters, Dataset& ds) const;
};

} // namespace hdf5

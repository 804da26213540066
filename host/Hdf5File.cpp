#include "Hdf5File.hpp"

#include <zlib.h>

#include <algorithm>
#include <cstring>
#include <fstream>
#include <iterator>
#include <stdexcept>

namespace hdf5 {

namespace {
const std::uint8_t kSignature[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
enum : std::uint16_t {
  MSG_NIL = 0x0, MSG_DATASPACE = 0x1, MSG_LINK_INFO = 0x2, MSG_DATATYPE = 0x3, MSG_LINK = 0x6, MSG_LAYOUT = 0x8,
  MSG_FILTERS = 0xB, MSG_ATTRIBUTE = 0xC, MSG_CONTINUATION = 0x10, MSG_SYMBOL_TABLE = 0x11, MSG_ATTRIBUTE_INFO = 0x15
};
inline std::size_t pad8(std::size_t n) { return (n + 7u) & ~(std::size_t)7u; }
} // namespace

void File::fail(const std::string& what) const { throw std::runtime_error("HDF5 '" + path_ + "': " + what); }

void File::need(std::uint64_t off, std::uint64_t n) const {
  if (off > bytes_.size() || n > bytes_.size() - off) fail("address " + std::to_string(off) + "+" + std::to_string(n) + " is outside the file (truncated or corrupt)");
}

std::uint64_t File::rd(std::uint64_t off, unsigned n) const {
  need(off, n);
  std::uint64_t v = 0;
  for (unsigned i = 0; i < n; ++i) v |= (std::uint64_t)bytes_[off + i] << (8 * i);
  return v;
}

bool File::undefinedAddress(std::uint64_t a) const {
  return sizeOfOffsets_ == 8 ? a == ~(std::uint64_t)0 : a == ((std::uint64_t)1 << (8 * sizeOfOffsets_)) - 1;
}

File::File(const std::string& path) : path_(path) {
  std::ifstream f(path, std::ios::binary);
  if (!f) throw std::runtime_error("Could not open HDF5 file: '" + path + "'");
  bytes_.assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
  // The superblock may sit at 0, 512, 1024, ... (user block); Keras files have it at 0.
  std::uint64_t sb = 0;
  bool found = false;
  for (std::uint64_t off = 0; off + 8 <= bytes_.size(); off = off ? off * 2 : 512) {
    if (std::memcmp(&bytes_[off], kSignature, 8) == 0) { sb = off; found = true; break; }
  }
  if (!found) fail("not an HDF5 file (no superblock signature)");
  const unsigned version = (unsigned)rd(sb + 8, 1);
  if (version <= 1) {
    sizeOfOffsets_ = (unsigned)rd(sb + 13, 1);
    sizeOfLengths_ = (unsigned)rd(sb + 14, 1);
    if (sizeOfOffsets_ < 2 || sizeOfOffsets_ > 8 || sizeOfLengths_ < 2 || sizeOfLengths_ > 8) fail("unsupported offset/length size");
    std::uint64_t p = sb + 24 + (version == 1 ? 4 : 0);
    base_ = rdOffset(p); p += sizeOfOffsets_;
    p += sizeOfOffsets_;                     // free-space info
    p += sizeOfOffsets_;                     // end of file
    p += sizeOfOffsets_;                     // driver info
    // root group symbol table entry: link name offset, object header address, ...
    rootHeader_ = rdOffset(p + sizeOfOffsets_);
  } else if (version == 2 || version == 3) {
    sizeOfOffsets_ = (unsigned)rd(sb + 9, 1);
    sizeOfLengths_ = (unsigned)rd(sb + 10, 1);
    std::uint64_t p = sb + 12;
    base_ = rdOffset(p); p += sizeOfOffsets_;
    p += sizeOfOffsets_;                     // superblock extension
    p += sizeOfOffsets_;                     // end of file
    rootHeader_ = rdOffset(p);
  } else {
    fail("unsupported superblock version " + std::to_string(version));
  }
}

// ---- object headers -----------------------------------------------------------------------------------

void File::parseV1Messages(std::uint64_t off, std::uint64_t len, unsigned& remaining, std::vector<Message>& out) const {
  need(off, len);
  const std::uint64_t end = off + len;
  while (remaining > 0 && off + 8 <= end) {
    const std::uint16_t type = (std::uint16_t)rd(off, 2);
    const std::uint16_t size = (std::uint16_t)rd(off + 2, 2);
    const std::uint8_t flags = (std::uint8_t)rd(off + 4, 1);
    off += 8;
    need(off, size);
    --remaining;
    if (type == MSG_CONTINUATION) {
      const std::uint64_t coff = rdOffset(off) + base_, clen = rdLength(off + sizeOfOffsets_);
      parseV1Messages(coff, clen, remaining, out);
    } else if (type != MSG_NIL) {
      Message m{type, flags, std::vector<std::uint8_t>(bytes_.begin() + off, bytes_.begin() + off + size)};
      out.push_back(std::move(m));
    }
    off += size;
  }
}

void File::parseV2Messages(std::uint64_t off, std::uint64_t end, bool creationOrder, std::vector<Message>& out) const {
  while (off + 4 <= end) {
    const std::uint16_t type = (std::uint16_t)rd(off, 1);
    const std::uint16_t size = (std::uint16_t)rd(off + 1, 2);
    const std::uint8_t flags = (std::uint8_t)rd(off + 3, 1);
    off += 4 + (creationOrder ? 2 : 0);
    if (off + size > end) break;                // gap before the checksum
    if (type == MSG_CONTINUATION) {
      const std::uint64_t coff = rdOffset(off) + base_, clen = rdLength(off + sizeOfOffsets_);
      need(coff, clen);
      if (clen < 8 || std::memcmp(&bytes_[coff], "OCHK", 4) != 0) fail("bad object header continuation block");
      parseV2Messages(coff + 4, coff + clen - 4, creationOrder, out);
    } else if (type != MSG_NIL) {
      out.push_back(Message{type, flags, std::vector<std::uint8_t>(bytes_.begin() + off, bytes_.begin() + off + size)});
    }
    off += size;
  }
}

std::vector<File::Message> File::objectHeader(std::uint64_t addr) const {
  addr += base_;
  need(addr, 16);
  std::vector<Message> out;
  if (std::memcmp(&bytes_[addr], "OHDR", 4) == 0) {
    if (rd(addr + 4, 1) != 2) fail("unsupported object header version");
    const unsigned flags = (unsigned)rd(addr + 5, 1);
    std::uint64_t p = addr + 6;
    if (flags & 0x20) p += 16;
    if (flags & 0x10) p += 4;
    const unsigned szBytes = 1u << (flags & 3);
    const std::uint64_t chunk0 = rd(p, szBytes);
    p += szBytes;
    need(p, chunk0);
    parseV2Messages(p, p + chunk0, (flags & 0x04) != 0, out);
    return out;
  }
  if (rd(addr, 1) != 1) fail("unsupported object header version " + std::to_string(rd(addr, 1)));
  unsigned remaining = (unsigned)rd(addr + 2, 2);
  const std::uint64_t size = rd(addr + 8, 4);
  parseV1Messages(addr + 16, size, remaining, out);
  return out;
}

// ---- groups ---------------------------------------------------------------------------------------------

void File::walkGroupBtree(std::uint64_t node, std::uint64_t heapData, std::map<std::string, std::uint64_t>& out,
                          std::vector<std::string>* order) const {
  node += base_;
  need(node, 8);
  if (std::memcmp(&bytes_[node], "TREE", 4) != 0) fail("bad group B-tree node signature");
  if (rd(node + 4, 1) != 0) fail("group B-tree node has the wrong type");
  const unsigned level = (unsigned)rd(node + 5, 1), used = (unsigned)rd(node + 6, 2);
  std::uint64_t p = node + 8 + 2 * sizeOfOffsets_;
  for (unsigned i = 0; i < used; ++i) {
    p += sizeOfLengths_;                               // key i
    const std::uint64_t child = rdOffset(p);
    p += sizeOfOffsets_;
    if (level > 0) { walkGroupBtree(child, heapData, out, order); continue; }
    const std::uint64_t sn = child + base_;
    need(sn, 8);
    if (std::memcmp(&bytes_[sn], "SNOD", 4) != 0) fail("bad symbol table node signature");
    const unsigned n = (unsigned)rd(sn + 6, 2);
    const std::uint64_t entrySize = 2 * sizeOfOffsets_ + 8 + 16;
    for (unsigned k = 0; k < n; ++k) {
      const std::uint64_t e = sn + 8 + k * entrySize;
      const std::uint64_t nameOff = rdOffset(e), header = rdOffset(e + sizeOfOffsets_);
      std::string name;
      for (std::uint64_t q = heapData + nameOff;; ++q) { need(q, 1); if (!bytes_[q]) break; name += (char)bytes_[q]; }
      out[name] = header;
      if (order) order->push_back(name);
    }
  }
}

std::map<std::string, std::uint64_t> File::links(std::uint64_t groupHeader, std::vector<std::string>* order) const {
  std::map<std::string, std::uint64_t> out;
  bool isGroup = false;
  for (const Message& m : objectHeader(groupHeader)) {
    if (m.type == MSG_SYMBOL_TABLE) {
      isGroup = true;
      if (m.body.size() < 2 * sizeOfOffsets_) fail("short symbol table message");
      std::uint64_t btree = 0, heap = 0;
      for (unsigned i = 0; i < sizeOfOffsets_; ++i) { btree |= (std::uint64_t)m.body[i] << (8 * i); heap |= (std::uint64_t)m.body[sizeOfOffsets_ + i] << (8 * i); }
      const std::uint64_t h = heap + base_;
      need(h, 8 + 2 * sizeOfLengths_ + sizeOfOffsets_);
      if (std::memcmp(&bytes_[h], "HEAP", 4) != 0) fail("bad local heap signature");
      const std::uint64_t heapData = rdOffset(h + 8 + 2 * sizeOfLengths_) + base_;
      walkGroupBtree(btree, heapData, out, order);
    } else if (m.type == MSG_LINK) {
      isGroup = true;
      const std::uint8_t* b = m.body.data();
      const std::size_t n = m.body.size();
      if (n < 2 || b[0] != 1) fail("unsupported link message version");
      const unsigned flags = b[1];
      std::size_t p = 2;
      unsigned linkType = 0;
      if (flags & 0x08) linkType = b[p++];
      if (flags & 0x04) p += 8;
      if (flags & 0x10) p += 1;
      const unsigned lenBytes = 1u << (flags & 3);
      std::uint64_t nameLen = 0;
      for (unsigned i = 0; i < lenBytes; ++i) nameLen |= (std::uint64_t)b[p + i] << (8 * i);
      p += lenBytes;
      if (p + nameLen > n) fail("short link message");
      const std::string name(reinterpret_cast<const char*>(b + p), (std::size_t)nameLen);
      p += (std::size_t)nameLen;
      if (linkType != 0) continue;               // soft/external links are not followed
      std::uint64_t addr = 0;
      for (unsigned i = 0; i < sizeOfOffsets_; ++i) addr |= (std::uint64_t)b[p + i] << (8 * i);
      out[name] = addr;
      if (order) order->push_back(name);
    } else if (m.type == MSG_LINK_INFO) {
      isGroup = true;
      const std::uint8_t* b = m.body.data();
      std::size_t p = 2;
      if (b[1] & 1) p += 8;
      std::uint64_t fheap = 0;
      for (unsigned i = 0; i < sizeOfOffsets_; ++i) fheap |= (std::uint64_t)b[p + i] << (8 * i);
      if (!undefinedAddress(fheap)) fail("group with dense link storage (fractal heap) is not supported; re-save with libver='earliest'");
    }
  }
  if (!isGroup) fail("object is not a group");
  return out;
}

std::uint64_t File::resolve(const std::string& path) const {
  std::uint64_t cur = rootHeader_;
  std::size_t p = 0;
  while (p < path.size()) {
    while (p < path.size() && path[p] == '/') ++p;
    if (p >= path.size()) break;
    std::size_t q = path.find('/', p);
    if (q == std::string::npos) q = path.size();
    const std::string part = path.substr(p, q - p);
    const auto l = links(cur, nullptr);
    const auto it = l.find(part);
    if (it == l.end()) fail("object '" + path + "' doesn't exist (no '" + part + "')");
    cur = it->second;
    p = q;
  }
  return cur;
}

bool File::exists(const std::string& path) const {
  try { resolve(path); return true; } catch (const std::runtime_error&) { return false; }
}

std::vector<std::string> File::list(const std::string& groupPath) const {
  std::vector<std::string> order;
  links(resolve(groupPath), &order);
  return order;
}

// ---- datatypes, dataspaces, attributes ------------------------------------------------------------------

File::Datatype File::parseDatatype(const std::uint8_t* p, std::size_t n) const {
  if (n < 8) fail("short datatype message");
  Datatype t;
  t.cls = p[0] & 0x0f;
  t.size = (std::size_t)p[4] | ((std::size_t)p[5] << 8) | ((std::size_t)p[6] << 16) | ((std::size_t)p[7] << 24);
  if (t.cls == 0 || t.cls == 1) t.littleEndian = (p[1] & 1) == 0;
  if (t.cls == 3) t.strPad = p[1] & 0x0f;
  if (t.cls == 9) t.vlenKind = p[1] & 0x0f;          // 0 sequence, 1 string
  return t;
}

File::Dataspace File::parseDataspace(const std::uint8_t* p, std::size_t n) const {
  if (n < 4) fail("short dataspace message");
  Dataspace s;
  const unsigned version = p[0], rank = p[1];
  std::size_t off;
  if (version == 1) off = 8;
  else if (version == 2) { off = 4; if (p[3] == 2) s.null = true; }
  else fail("unsupported dataspace version");
  if (rank == 0) s.scalar = !s.null;
  if (off + (std::size_t)rank * sizeOfLengths_ > n) fail("short dataspace message");
  for (unsigned i = 0; i < rank; ++i) {
    std::uint64_t d = 0;
    for (unsigned k = 0; k < sizeOfLengths_; ++k) d |= (std::uint64_t)p[off + i * sizeOfLengths_ + k] << (8 * k);
    s.dims.push_back((std::size_t)d);
  }
  return s;
}

std::string File::globalHeapObject(std::uint64_t collection, std::uint32_t index) const {
  const std::uint64_t c = collection + base_;
  need(c, 8 + sizeOfLengths_);
  if (std::memcmp(&bytes_[c], "GCOL", 4) != 0) fail("bad global heap collection signature");
  const std::uint64_t size = rdLength(c + 8);
  need(c, size);
  std::uint64_t p = c + 8 + sizeOfLengths_;
  while (p + 8 + sizeOfLengths_ <= c + size) {
    const unsigned idx = (unsigned)rd(p, 2);
    const std::uint64_t osz = rdLength(p + 8);
    if (idx == 0) break;                         // free space
    if (idx == index) {
      need(p + 8 + sizeOfLengths_, osz);
      return std::string(reinterpret_cast<const char*>(&bytes_[p + 8 + sizeOfLengths_]), (std::size_t)osz);
    }
    p += 8 + sizeOfLengths_ + pad8((std::size_t)osz);
  }
  fail("global heap object " + std::to_string(index) + " not found");
}

namespace {
struct AttrView { std::string name; const std::uint8_t* dt; std::size_t dtSize; const std::uint8_t* ds; std::size_t dsSize; const std::uint8_t* data; std::size_t dataSize; };
} // namespace

static bool splitAttribute(const std::vector<std::uint8_t>& body, AttrView& a) {
  if (body.size() < 8) return false;
  const unsigned version = body[0];
  const std::size_t nameSize = body[2] | (body[3] << 8), dtSize = body[4] | (body[5] << 8), dsSize = body[6] | (body[7] << 8);
  std::size_t p = 8;
  if (version == 3) p += 1;
  else if (version != 1 && version != 2) return false;
  const bool padded = version == 1;
  auto adv = [&](std::size_t n) { return padded ? pad8(n) : n; };
  if (p + adv(nameSize) + adv(dtSize) + adv(dsSize) > body.size()) return false;
  a.name.assign(reinterpret_cast<const char*>(&body[p]), nameSize ? nameSize - 1 : 0);
  while (!a.name.empty() && a.name.back() == '\0') a.name.pop_back();
  p += adv(nameSize);
  a.dt = &body[p]; a.dtSize = dtSize; p += adv(dtSize);
  a.ds = &body[p]; a.dsSize = dsSize; p += adv(dsSize);
  a.data = body.data() + p; a.dataSize = body.size() - p;
  return true;
}

bool File::hasAttribute(const std::string& objectPath, const std::string& name) const {
  for (const Message& m : objectHeader(resolve(objectPath))) {
    AttrView a;
    if (m.type == MSG_ATTRIBUTE && splitAttribute(m.body, a) && a.name == name) return true;
  }
  return false;
}

std::string File::stringAttribute(const std::string& objectPath, const std::string& name) const {
  for (const Message& m : objectHeader(resolve(objectPath))) {
    if (m.type == MSG_ATTRIBUTE_INFO) {
      // dense attribute storage: fractal heap address follows flags (+ max creation index)
      const std::uint8_t* b = m.body.data();
      std::size_t p = 2;
      if (b[1] & 1) p += 2;
      std::uint64_t fheap = 0;
      for (unsigned i = 0; i < sizeOfOffsets_ && p + i < m.body.size(); ++i) fheap |= (std::uint64_t)b[p + i] << (8 * i);
      if (!undefinedAddress(fheap)) fail("dense attribute storage (fractal heap) is not supported; re-save with libver='earliest'");
      continue;
    }
    if (m.type != MSG_ATTRIBUTE) continue;
    AttrView a;
    if (!splitAttribute(m.body, a)) fail("unsupported attribute message version");
    if (a.name != name) continue;
    const Datatype t = parseDatatype(a.dt, a.dtSize);
    const Dataspace s = parseDataspace(a.ds, a.dsSize);
    std::size_t count = 1;
    for (auto d : s.dims) count *= d;
    if (count != 1) fail("attribute '" + name + "' is not a scalar string");
    if (t.cls == 3) {
      if (t.size > a.dataSize) fail("short attribute data");
      std::string v(reinterpret_cast<const char*>(a.data), t.size);
      const std::size_t z = v.find('\0');
      if (z != std::string::npos) v.resize(z);
      else if (t.strPad == 2) while (!v.empty() && v.back() == ' ') v.pop_back();
      return v;
    }
    if (t.cls == 9 && t.vlenKind == 1) {
      if (a.dataSize < 4 + sizeOfOffsets_ + 4) fail("short variable-length attribute data");
      std::uint64_t coll = 0;
      for (unsigned i = 0; i < sizeOfOffsets_; ++i) coll |= (std::uint64_t)a.data[4 + i] << (8 * i);
      std::uint32_t idx; std::memcpy(&idx, a.data + 4 + sizeOfOffsets_, 4);
      std::uint32_t len; std::memcpy(&len, a.data, 4);
      std::string v = globalHeapObject(coll, idx);
      if (v.size() > len) v.resize(len);
      return v;
    }
    fail("attribute '" + name + "' is not a string");
  }
  fail("attribute '" + name + "' doesn't exist on '" + objectPath + "'");
}

// ---- datasets ---------------------------------------------------------------------------------------------

void File::walkChunkBtree(std::uint64_t node, unsigned rank, const std::vector<std::size_t>& chunkDims, const std::vector<int>& filters,
                          Dataset& ds) const {
  node += base_;
  need(node, 8);
  if (std::memcmp(&bytes_[node], "TREE", 4) != 0) fail("bad chunk B-tree node signature");
  if (rd(node + 4, 1) != 1) fail("chunk B-tree node has the wrong type");
  const unsigned level = (unsigned)rd(node + 5, 1), used = (unsigned)rd(node + 6, 2);
  const std::uint64_t keySize = 8 + 8ull * (rank + 1);
  std::uint64_t p = node + 8 + 2 * sizeOfOffsets_;
  std::size_t chunkElems = 1;
  for (auto d : chunkDims) chunkElems *= d;
  const std::size_t chunkBytes = chunkElems * ds.elementSize;
  for (unsigned i = 0; i < used; ++i) {
    const std::uint32_t stored = (std::uint32_t)rd(p, 4), mask = (std::uint32_t)rd(p + 4, 4);
    std::vector<std::size_t> origin(rank);
    for (unsigned d = 0; d < rank; ++d) origin[d] = (std::size_t)rd(p + 8 + 8ull * d, 8);
    const std::uint64_t child = rdOffset(p + keySize);
    p += keySize + sizeOfOffsets_;
    if (level > 0) { walkChunkBtree(child, rank, chunkDims, filters, ds); continue; }
    need(child + base_, stored);
    std::vector<std::uint8_t> buf(bytes_.begin() + child + base_, bytes_.begin() + child + base_ + stored);
    // filters are undone in reverse pipeline order; bit k of the mask = filter k was skipped for this chunk
    for (int k = (int)filters.size() - 1; k >= 0; --k) {
      if (mask & (1u << k)) continue;
      if (filters[k] == 1) {                       // deflate
        std::vector<std::uint8_t> outb(chunkBytes);
        uLongf got = (uLongf)outb.size();
        if (uncompress(outb.data(), &got, buf.data(), (uLong)buf.size()) != Z_OK) fail("deflate filter: corrupt chunk");
        outb.resize(got);
        buf.swap(outb);
      } else if (filters[k] == 2) {                // shuffle
        const std::size_t es = ds.elementSize, n = buf.size() / es;
        std::vector<std::uint8_t> outb(buf.size());
        for (std::size_t b = 0; b < es; ++b) for (std::size_t e = 0; e < n; ++e) outb[e * es + b] = buf[b * n + e];
        std::copy(buf.begin() + n * es, buf.end(), outb.begin() + n * es);
        buf.swap(outb);
      } else if (filters[k] == 3) {                // fletcher32: checksum trails the data
        if (buf.size() >= 4) buf.resize(buf.size() - 4);
      } else {
        fail("unsupported filter id " + std::to_string(filters[k]));
      }
    }
    if (buf.size() < chunkBytes) fail("chunk smaller than its declared size");
    // copy the chunk into place, clipping at the dataset edge
    std::vector<std::size_t> idx(rank, 0);
    const std::size_t rowElems = chunkDims[rank - 1];
    std::size_t rows = chunkElems / rowElems;
    for (std::size_t rrow = 0; rrow < rows; ++rrow) {
      std::size_t rem = rrow;
      bool inside = true;
      std::size_t dstIndex = 0;
      for (int d = (int)rank - 2; d >= 0; --d) { idx[d] = rem % chunkDims[d]; rem /= chunkDims[d]; }
      for (unsigned d = 0; d + 1 < rank; ++d) {
        const std::size_t c = origin[d] + idx[d];
        if (c >= ds.shape[d]) { inside = false; break; }
        dstIndex = dstIndex * ds.shape[d] + c;
      }
      if (!inside || origin[rank - 1] >= ds.shape[rank - 1]) continue;
      const std::size_t take = std::min(rowElems, ds.shape[rank - 1] - origin[rank - 1]);
      dstIndex = dstIndex * ds.shape[rank - 1] + origin[rank - 1];
      std::memcpy(&ds.data[dstIndex * ds.elementSize], &buf[rrow * rowElems * ds.elementSize], take * ds.elementSize);
    }
  }
}

void File::readChunked(const Message& layout, const std::vector<Message>& msgs, Dataset& ds) const {
  const std::uint8_t* b = layout.body.data();
  const unsigned dimensionality = b[2];
  if (dimensionality < 2) fail("bad chunked layout");
  const unsigned rank = dimensionality - 1;
  if (rank != ds.shape.size()) fail("chunk rank disagrees with the dataspace");
  std::uint64_t btree = 0;
  for (unsigned i = 0; i < sizeOfOffsets_; ++i) btree |= (std::uint64_t)b[3 + i] << (8 * i);
  std::vector<std::size_t> chunkDims(rank);
  for (unsigned d = 0; d < rank; ++d) { std::uint32_t v; std::memcpy(&v, b + 3 + sizeOfOffsets_ + 4 * d, 4); chunkDims[d] = v; }
  std::vector<int> filters;
  for (const Message& m : msgs) {
    if (m.type != MSG_FILTERS) continue;
    const std::uint8_t* f = m.body.data();
    const unsigned version = f[0], n = f[1];
    std::size_t p = version == 1 ? 8 : 2;
    for (unsigned i = 0; i < n; ++i) {
      const unsigned id = f[p] | (f[p + 1] << 8);
      std::size_t nameLen = 0;
      p += 2;
      if (version == 1 || id >= 256) { nameLen = f[p] | (f[p + 1] << 8); p += 2; }
      p += 2;                                       // flags
      const unsigned ncd = f[p] | (f[p + 1] << 8);
      p += 2;
      p += version == 1 ? pad8(nameLen) : nameLen;
      p += 4 * (std::size_t)ncd;
      if (version == 1 && (ncd & 1)) p += 4;
      filters.push_back((int)id);
    }
  }
  std::fill(ds.data.begin(), ds.data.end(), 0);
  if (!undefinedAddress(btree)) walkChunkBtree(btree, rank, chunkDims, filters, ds);
}

Dataset File::dataset(const std::string& path) const {
  const std::vector<Message> msgs = objectHeader(resolve(path));
  Dataset ds;
  const Message* layout = nullptr;
  bool haveType = false, haveSpace = false;
  for (const Message& m : msgs) {
    if (m.type == MSG_DATATYPE) {
      const Datatype t = parseDatatype(m.body.data(), m.body.size());
      ds.elementSize = t.size; ds.typeClass = t.cls; haveType = true;
      if (t.cls == 1 && !t.littleEndian) fail("big-endian floating point data is not supported");
      if (t.cls == 9) fail("variable-length datasets are not supported");
    } else if (m.type == MSG_DATASPACE) {
      const Dataspace s = parseDataspace(m.body.data(), m.body.size());
      ds.shape = s.dims; haveSpace = true;
    } else if (m.type == MSG_LAYOUT) {
      layout = &m;
    }
  }
  if (!haveType || !haveSpace || !layout) fail("'" + path + "' is not a dataset");
  ds.data.assign(ds.numElements() * ds.elementSize, 0);
  const std::uint8_t* b = layout->body.data();
  const unsigned version = b[0];
  if (version != 3) fail("data layout message version " + std::to_string(version) + " is not supported");
  const unsigned cls = b[1];
  if (cls == 0) {                                   // compact
    const std::size_t size = b[2] | (b[3] << 8);
    if (size < ds.data.size() || 4 + size > layout->body.size()) fail("compact dataset smaller than its dataspace");
    std::memcpy(ds.data.data(), b + 4, ds.data.size());
  } else if (cls == 1) {                            // contiguous
    std::uint64_t addr = 0, size = 0;
    for (unsigned i = 0; i < sizeOfOffsets_; ++i) addr |= (std::uint64_t)b[2 + i] << (8 * i);
    for (unsigned i = 0; i < sizeOfLengths_; ++i) size |= (std::uint64_t)b[2 + sizeOfOffsets_ + i] << (8 * i);
    if (undefinedAddress(addr)) return ds;          // never written: fill value (zeros)
    if (size < ds.data.size()) fail("contiguous dataset smaller than its dataspace");
    need(addr + base_, ds.data.size());
    std::memcpy(ds.data.data(), &bytes_[addr + base_], ds.data.size());
  } else if (cls == 2) {
    readChunked(*layout, msgs, ds);
  } else {
    fail("unsupported data layout class " + std::to_string(cls));
  }
  return ds;
}

} // namespace hdf5

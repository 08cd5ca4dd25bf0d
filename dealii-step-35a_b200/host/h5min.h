// A minimal HDF5 writer / reader for what the reference's h5.h does (h5.h:66-208): contiguous datasets of IEEE
// doubles (1-D vectors, 2-D full matrices) under the root group.  libhdf5 does not exist on the target machines, so the
// file format is written by hand, in the oldest and simplest of its forms (the one H5Fcreate with default properties
// produces): superblock version 0, a root group made of a version-1 B-tree, a local heap and symbol-table nodes,
// version-1 object headers with dataspace, datatype, fill-value and contiguous-layout messages.  Files written here
// open in h5py / h5dump / ParaView; files written by the reference (same layout class) are read back here.
//
// save() follows the reference: an existing file is opened and the dataset added (h5.h:74-93).  The reference never
// checks whether the name exists already (H5Dcreate2 fails silently then and the old data stays); here the new data
// replace the old.  The file is small (a checkpoint: six families of vectors), so adding a dataset rewrites the file.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace h5min {

struct Dataset
{
  std::vector<uint64_t> dims; // rank 1 or 2, row-major
  std::vector<double> data;
};
using File = std::map<std::string, Dataset>; // by name (without the leading '/'); std::map keeps strcmp order

constexpr uint64_t kUndef = ~uint64_t(0);
constexpr int kLeafK = 4, kInternalK = 16; // the library's defaults

namespace detail {

struct Out
{
  std::vector<unsigned char> b;
  void u8(unsigned v) { b.push_back((unsigned char)v); }
  void u16(unsigned v)
  {
    u8(v & 0xff);
    u8((v >> 8) & 0xff);
  }
  void u32(uint32_t v)
  {
    for (int i = 0; i < 4; ++i)
      u8((v >> (8 * i)) & 0xff);
  }
  void u64(uint64_t v)
  {
    for (int i = 0; i < 8; ++i)
      u8((unsigned)((v >> (8 * i)) & 0xff));
  }
  void bytes(const void *p, size_t n) { b.insert(b.end(), (const unsigned char *)p, (const unsigned char *)p + n); }
  void zeros(size_t n) { b.insert(b.end(), n, 0); }
  void pad8()
  {
    while (b.size() % 8)
      u8(0);
  }
  uint64_t at() const { return b.size(); }
  void patch64(size_t where, uint64_t v)
  {
    for (int i = 0; i < 8; ++i)
      b[where + i] = (unsigned char)((v >> (8 * i)) & 0xff);
  }
};

inline uint64_t pad8(uint64_t n) { return (n + 7) & ~uint64_t(7); }

} // namespace detail

// ---------------------------------------------------------------------------------------------------------- writer
// Layout of the file: superblock | root object header | B-tree node | local heap (header + names) | symbol-table
// nodes | per dataset: object header, raw data.
inline std::vector<unsigned char> serialize(const File &file)
{
  using detail::Out;
  using detail::pad8;
  const size_t n_sets = file.size();
  const size_t per_node = 2 * kLeafK;
  const size_t n_nodes = std::max<size_t>(1, (n_sets + per_node - 1) / per_node);
  if (n_nodes > 2 * kInternalK)
    throw std::runtime_error("h5min: more datasets than one B-tree node holds");
  // the names in the heap: offset 0 is the empty string
  std::vector<uint64_t> name_at;
  uint64_t heap_data = 8;
  for (const auto &kv : file)
    {
      name_at.push_back(heap_data);
      heap_data += pad8(kv.first.size() + 1);
    }
  const uint64_t superblock = 96, root_header = 16 + 24, btree = 24 + (2 * kInternalK + 1) * 8 + 2 * kInternalK * 8;
  const uint64_t heap_header = 32, snod = 8 + per_node * 40;
  const uint64_t at_root = superblock, at_btree = at_root + root_header, at_heap = at_btree + btree;
  const uint64_t at_heap_data = at_heap + heap_header, at_snod = at_heap_data + heap_data;
  uint64_t cursor = at_snod + n_nodes * snod;
  std::vector<uint64_t> header_at, data_at;
  for (const auto &kv : file)
    {
      const uint64_t rank = kv.second.dims.size();
      const uint64_t messages = (8 + 8 + 8 * rank) + (8 + 24) + (8 + 8) + (8 + 24);
      header_at.push_back(cursor);
      cursor += 16 + messages;
      data_at.push_back(cursor);
      cursor += pad8(8 * kv.second.data.size());
    }
  const uint64_t eof = cursor;

  Out o;
  // superblock, version 0
  const unsigned char sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
  o.bytes(sig, 8);
  o.u8(0); // superblock version
  o.u8(0); // free-space storage version
  o.u8(0); // root group symbol table entry version
  o.u8(0);
  o.u8(0); // shared header message format version
  o.u8(8); // size of offsets
  o.u8(8); // size of lengths
  o.u8(0);
  o.u16(kLeafK);
  o.u16(kInternalK);
  o.u32(0);      // file consistency flags
  o.u64(0);      // base address
  o.u64(kUndef); // free-space info
  o.u64(eof);    // end of file address
  o.u64(kUndef); // driver information block
  // root group symbol table entry
  o.u64(0);       // link name offset
  o.u64(at_root); // object header
  o.u32(1);       // cache type: group, B-tree and heap addresses in the scratch pad
  o.u32(0);
  o.u64(at_btree);
  o.u64(at_heap);
  // root group object header (version 1): one symbol-table message
  o.u8(1);
  o.u8(0);
  o.u16(1);  // messages
  o.u32(1);  // reference count
  o.u32(24); // header size
  o.u32(0);  // padding to 8
  o.u16(0x0011);
  o.u16(16);
  o.u8(0);
  o.zeros(3);
  o.u64(at_btree);
  o.u64(at_heap);
  // B-tree node of the group: leaf level, one child per symbol-table node
  o.bytes("TREE", 4);
  o.u8(0); // node type: group
  o.u8(0); // level
  o.u16((unsigned)n_nodes);
  o.u64(kUndef);
  o.u64(kUndef);
  {
    size_t written = 0;
    o.u64(0); // key 0: the empty string
    for (size_t k = 0; k < n_nodes; ++k)
      {
        o.u64(at_snod + k * snod);
        const size_t last = std::min(n_sets, (k + 1) * per_node);
        o.u64(n_sets ? name_at[last - 1] : 0); // key k + 1: the largest name in child k
        written += 16;
      }
    o.zeros(btree - 24 - 8 - written);
  }
  // local heap
  o.bytes("HEAP", 4);
  o.u8(0);
  o.zeros(3);
  o.u64(heap_data);
  o.u64(1); // no free block
  o.u64(at_heap_data);
  o.zeros(8);
  for (const auto &kv : file)
    {
      o.bytes(kv.first.data(), kv.first.size());
      o.zeros(pad8(kv.first.size() + 1) - kv.first.size());
    }
  // symbol-table nodes
  {
    size_t i = 0;
    for (size_t k = 0; k < n_nodes; ++k)
      {
        const size_t here = std::min(per_node, n_sets - std::min(n_sets, k * per_node));
        o.bytes("SNOD", 4);
        o.u8(1);
        o.u8(0);
        o.u16((unsigned)here);
        for (size_t e = 0; e < per_node; ++e)
          if (e < here)
            {
              o.u64(name_at[i]);
              o.u64(header_at[i]);
              o.u32(0);
              o.u32(0);
              o.zeros(16);
              ++i;
            }
          else
            o.zeros(40);
      }
  }
  // datasets
  {
    size_t i = 0;
    for (const auto &kv : file)
      {
        const Dataset &d = kv.second;
        const uint64_t rank = d.dims.size();
        uint64_t count = 1;
        for (uint64_t x : d.dims)
          count *= x;
        if (rank < 1 || rank > 2 || count != d.data.size())
          throw std::runtime_error("h5min: dataset " + kv.first + " has inconsistent dimensions");
        const uint64_t messages = (8 + 8 + 8 * rank) + (8 + 24) + (8 + 8) + (8 + 24);
        if (o.at() != header_at[i])
          throw std::logic_error("h5min: layout arithmetic");
        o.u8(1);
        o.u8(0);
        o.u16(4);
        o.u32(1);
        o.u32((uint32_t)messages);
        o.u32(0);
        // dataspace, version 1
        o.u16(0x0001);
        o.u16((unsigned)(8 + 8 * rank));
        o.u8(0);
        o.zeros(3);
        o.u8(1);
        o.u8((unsigned)rank);
        o.u8(0);
        o.zeros(5);
        for (uint64_t x : d.dims)
          o.u64(x);
        // datatype: IEEE double, little endian
        o.u16(0x0003);
        o.u16(24);
        o.u8(1); // constant message
        o.zeros(3);
        o.u8(0x11); // version 1, class 1 (floating point)
        o.u8(0x20); // little endian, mantissa normalisation: implied leading bit
        o.u8(0x3f); // sign bit position
        o.u8(0);
        o.u32(8);   // size in bytes
        o.u16(0);   // bit offset
        o.u16(64);  // precision
        o.u8(52);   // exponent location
        o.u8(11);   // exponent size
        o.u8(0);    // mantissa location
        o.u8(52);   // mantissa size
        o.u32(1023);
        o.zeros(4);
        // fill value, version 2: allocate early, never write a fill value, none defined
        o.u16(0x0005);
        o.u16(8);
        o.u8(1);
        o.zeros(3);
        o.u8(2);
        o.u8(1);
        o.u8(2);
        o.u8(0);
        o.zeros(4);
        // data layout, version 3: contiguous
        o.u16(0x0008);
        o.u16(24);
        o.u8(0);
        o.zeros(3);
        o.u8(3);
        o.u8(1);
        o.u64(data_at[i]);
        o.u64(8 * count);
        o.zeros(6);
        if (o.at() != data_at[i])
          throw std::logic_error("h5min: layout arithmetic");
        o.bytes(d.data.data(), 8 * d.data.size());
        o.pad8();
        ++i;
      }
  }
  if (o.at() != eof)
    throw std::logic_error("h5min: layout arithmetic");
  return o.b;
}

// ---------------------------------------------------------------------------------------------------------- reader
// Reads what the writer above and the reference (libhdf5, default file-creation properties) produce: superblock 0 / 1,
// old-style groups, version-1 object headers (continuation blocks followed), contiguous layout (version 3), IEEE
// little-endian doubles.  Anything else raises std::runtime_error.
namespace detail {

struct In
{
  const std::vector<unsigned char> &b;
  uint64_t base = 0;
  explicit In(const std::vector<unsigned char> &bytes) : b(bytes) {}
  void need(uint64_t at, uint64_t n) const
  {
    if (at + n > b.size() || at + n < at)
      throw std::runtime_error("h5min: truncated or corrupt file");
  }
  uint64_t u(uint64_t at, int n) const
  {
    need(at, n);
    uint64_t v = 0;
    for (int i = n - 1; i >= 0; --i)
      v = (v << 8) | b[at + i];
    return v;
  }
  bool tag(uint64_t at, const char *t) const
  {
    need(at, 4);
    return std::memcmp(&b[at], t, 4) == 0;
  }
};

// the messages of a version-1 object header: (type, offset of the data, size), continuation blocks included
inline void header_messages(const In &in, uint64_t at, std::vector<std::pair<unsigned, std::pair<uint64_t, uint64_t>>> &out)
{
  if (in.u(at, 1) != 1)
    throw std::runtime_error("h5min: only version-1 object headers are supported");
  uint64_t remaining = in.u(at + 2, 2);
  std::vector<std::pair<uint64_t, uint64_t>> blocks{{at + 16, in.u(at + 8, 4)}};
  for (size_t k = 0; k < blocks.size(); ++k)
    {
      uint64_t p = blocks[k].first;
      const uint64_t end = p + blocks[k].second;
      while (p + 8 <= end && remaining > 0)
        {
          const unsigned type = (unsigned)in.u(p, 2);
          const uint64_t size = in.u(p + 2, 2);
          in.need(p + 8, size);
          out.push_back({type, {p + 8, size}});
          if (type == 0x0010) // continuation: offset, length
            blocks.push_back({in.base + in.u(p + 8, 8), in.u(p + 16, 8)});
          p += 8 + size;
          --remaining;
        }
    }
}

inline void read_group_node(const In &in, uint64_t at, uint64_t heap_data, File &file);

inline void read_dataset(const In &in, uint64_t header, const std::string &name, File &file)
{
  std::vector<std::pair<unsigned, std::pair<uint64_t, uint64_t>>> msgs;
  header_messages(in, header, msgs);
  Dataset d;
  uint64_t data_at = kUndef, data_size = 0;
  bool is_double = false;
  for (const auto &m : msgs)
    {
      const uint64_t p = m.second.first;
      if (m.first == 0x0001)
        {
          const unsigned version = (unsigned)in.u(p, 1), rank = (unsigned)in.u(p + 1, 1);
          const uint64_t dims_at = p + (version == 1 ? 8 : 4);
          for (unsigned k = 0; k < rank; ++k)
            d.dims.push_back(in.u(dims_at + 8 * k, 8));
        }
      else if (m.first == 0x0003)
        is_double = (in.u(p, 1) & 0x0f) == 1 && in.u(p + 4, 4) == 8 && (in.u(p + 1, 1) & 1) == 0;
      else if (m.first == 0x0008)
        {
          if (in.u(p, 1) != 3 || in.u(p + 1, 1) != 1)
            throw std::runtime_error("h5min: dataset " + name + " is not stored contiguously (layout version 3)");
          data_at = in.u(p + 2, 8);
          data_size = in.u(p + 10, 8);
        }
    }
  if (!is_double || d.dims.empty())
    return; // not one of ours: groups' other children, other types
  uint64_t count = 1;
  for (uint64_t x : d.dims)
    count *= x;
  if (data_size != 8 * count)
    throw std::runtime_error("h5min: dataset " + name + " has an unexpected storage size");
  d.data.resize(count);
  if (count)
    {
      if (data_at == kUndef)
        throw std::runtime_error("h5min: dataset " + name + " has no storage");
      in.need(in.base + data_at, 8 * count);
      std::memcpy(d.data.data(), &in.b[in.base + data_at], 8 * count);
    }
  file[name] = std::move(d);
}

inline void read_group_node(const In &in, uint64_t at, uint64_t heap_data, File &file)
{
  if (in.tag(at, "TREE"))
    {
      const unsigned entries = (unsigned)in.u(at + 6, 2);
      for (unsigned k = 0; k < entries; ++k)
        read_group_node(in, in.base + in.u(at + 24 + 8 + 16 * k, 8), heap_data, file);
    }
  else if (in.tag(at, "SNOD"))
    {
      const unsigned symbols = (unsigned)in.u(at + 6, 2);
      for (unsigned k = 0; k < symbols; ++k)
        {
          const uint64_t e = at + 8 + 40 * k;
          const uint64_t name_at = heap_data + in.u(e, 8);
          std::string name;
          for (uint64_t q = name_at;; ++q)
            {
              in.need(q, 1);
              if (!in.b[q])
                break;
              name.push_back((char)in.b[q]);
            }
          read_dataset(in, in.base + in.u(e + 8, 8), name, file);
        }
    }
  else
    throw std::runtime_error("h5min: unexpected node in the root group");
}

} // namespace detail

inline File parse(const std::vector<unsigned char> &bytes)
{
  detail::In in(bytes);
  const unsigned char sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
  in.need(0, 8);
  if (std::memcmp(bytes.data(), sig, 8) != 0)
    throw std::runtime_error("h5min: not an HDF5 file");
  const unsigned version = (unsigned)in.u(8, 1);
  if (version > 1)
    throw std::runtime_error("h5min: superblock version " + std::to_string(version) + " is not supported (write the file with the "
                             "library's default, earliest, format)");
  if (in.u(13, 1) != 8 || in.u(14, 1) != 8)
    throw std::runtime_error("h5min: only 8-byte offsets and lengths are supported");
  const uint64_t fixed = version == 0 ? 24 : 28; // version 1 adds the indexed-storage K and two reserved bytes
  in.base = in.u(fixed, 8);
  const uint64_t root_entry = fixed + 32;
  uint64_t btree, heap;
  if (in.u(root_entry + 16, 4) == 1)
    {
      btree = in.u(root_entry + 24, 8);
      heap = in.u(root_entry + 32, 8);
    }
  else
    {
      // not cached in the entry: the root object header's symbol-table message has them
      std::vector<std::pair<unsigned, std::pair<uint64_t, uint64_t>>> msgs;
      detail::header_messages(in, in.base + in.u(root_entry + 8, 8), msgs);
      btree = heap = kUndef;
      for (const auto &m : msgs)
        if (m.first == 0x0011)
          {
            btree = in.u(m.second.first, 8);
            heap = in.u(m.second.first + 8, 8);
          }
      if (btree == kUndef)
        throw std::runtime_error("h5min: the root group is not an old-style group");
    }
  if (!in.tag(in.base + heap, "HEAP"))
    throw std::runtime_error("h5min: local heap not found");
  const uint64_t heap_data = in.base + in.u(in.base + heap + 24, 8);
  File file;
  detail::read_group_node(in, in.base + btree, heap_data, file);
  return file;
}

inline File read_file(const std::string &filename)
{
  std::ifstream in(filename, std::ios::binary);
  if (!in)
    throw std::runtime_error("cannot open file " + filename);
  std::vector<unsigned char> bytes((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
  return parse(bytes);
}

inline void write_file(const std::string &filename, const File &file)
{
  const std::vector<unsigned char> bytes = serialize(file);
  std::ofstream out(filename, std::ios::binary | std::ios::trunc);
  if (!out)
    throw std::runtime_error("cannot open file " + filename);
  out.write((const char *)bytes.data(), (std::streamsize)bytes.size());
  if (!out)
    throw std::runtime_error("cannot write file " + filename);
}

} // namespace h5min

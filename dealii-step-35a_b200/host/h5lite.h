// Stand-in for the reference's h5.h (H5::save / H5::load of 1-D double
// datasets, h5.h:66-124; block vectors as "/a0", "/a1", ..., h5.h:161-183).
// libhdf5 is not available on the target machines, so datasets are kept in a
// minimal container with the same dataset names: a magic string followed by
// records {name length, name, element count, doubles}.  Like the reference,
// save() appends to an existing file.
#pragma once
#include <cstdint>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace H5 {

static const char kMagic[8] = {'S', '3', '5', 'H', '5', 'L', 'T', '1'};

inline void save(const std::string &filename, const std::string &dataset_name, const std::vector<double> &v)
{
  bool fresh = true;
  {
    std::ifstream probe(filename, std::ios::binary);
    char m[8];
    if (probe.read(m, 8) && std::memcmp(m, kMagic, 8) == 0)
      fresh = false;
  }
  std::ofstream out(filename, std::ios::binary | (fresh ? std::ios::trunc : std::ios::app));
  if (!out)
    throw std::runtime_error("cannot open file " + filename);
  if (fresh)
    out.write(kMagic, 8);
  const uint64_t len = dataset_name.size(), n = v.size();
  out.write((const char *)&len, 8);
  out.write(dataset_name.data(), len);
  out.write((const char *)&n, 8);
  out.write((const char *)v.data(), n * sizeof(double));
}

inline void load(const std::string &filename, const std::string &dataset_name, std::vector<double> &v)
{
  std::ifstream in(filename, std::ios::binary);
  char m[8];
  if (!in.read(m, 8) || std::memcmp(m, kMagic, 8) != 0)
    throw std::runtime_error("cannot read checkpoint file " + filename);
  bool found = false;
  for (;;)
    {
      uint64_t len = 0, n = 0;
      if (!in.read((char *)&len, 8))
        break;
      std::string name(len, '\0');
      in.read(&name[0], len);
      in.read((char *)&n, 8);
      if (name == dataset_name)
        {
          // the last record of a name wins (a file may have been appended to)
          v.resize(n);
          in.read((char *)v.data(), n * sizeof(double));
          found = true;
        }
      else
        in.seekg((std::streamoff)(n * sizeof(double)), std::ios::cur);
      if (!in)
        throw std::runtime_error("truncated checkpoint file " + filename);
    }
  if (!found)
    throw std::runtime_error("dataset " + dataset_name + " not found in " + filename);
}

inline void save_block_vector(const std::string &filename, const std::vector<std::vector<double>> &blocks)
{
  for (size_t b = 0; b < blocks.size(); ++b)
    save(filename, "a" + std::to_string(b), blocks[b]);
}

} // namespace H5

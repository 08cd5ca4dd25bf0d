// The reference's h5.h (namespace H5: save / load of one vector as a 1-D double dataset, block vectors as "/a0",
// "/a1", ..., a full matrix as "/a"; h5.h:66-208) over the hand-written HDF5 writer of h5min.h -- real HDF5 files
// with the reference's dataset names, so that checkpoints and snapshots can be exchanged with the reference and
// opened in h5py or ParaView.  std::vector<double> stands for dealii::Vector<double>.
#pragma once
#include "h5min.h"

#include <sys/stat.h>

namespace H5 {

inline std::string plain_name(const std::string &dataset_name)
{
  return !dataset_name.empty() && dataset_name[0] == '/' ? dataset_name.substr(1) : dataset_name;
}

// h5.h:66-93: an existing file is opened and the dataset added, otherwise the file is created
inline void save(const std::string &file_name, const std::string &dataset_name, const std::vector<double> &vector)
{
  h5min::File file;
  struct stat buffer;
  if (stat(file_name.c_str(), &buffer) == 0)
    file = h5min::read_file(file_name);
  h5min::Dataset &d = file[plain_name(dataset_name)];
  d.dims.assign(1, vector.size());
  d.data = vector;
  h5min::write_file(file_name, file);
}

// h5.h:97-121
inline void load(const std::string &file_name, const std::string &dataset_name, std::vector<double> &vector)
{
  const h5min::File file = h5min::read_file(file_name);
  const auto it = file.find(plain_name(dataset_name));
  if (it == file.end())
    throw std::runtime_error("dataset " + dataset_name + " not found in " + file_name);
  vector = it->second.data;
}

// h5.h:161-183: a new file with the blocks as /a0, /a1, ...
inline void save_block_vector(const std::string &file_name, const std::vector<std::vector<double>> &blocks)
{
  h5min::File file;
  for (size_t b = 0; b < blocks.size(); ++b)
    {
      h5min::Dataset &d = file["a" + std::to_string(b)];
      d.dims.assign(1, blocks[b].size());
      d.data = blocks[b];
    }
  h5min::write_file(file_name, file);
}

// h5.h:124-157: one block per object of the file
inline void load_block_vector(const std::string &file_name, std::vector<std::vector<double>> &blocks)
{
  const h5min::File file = h5min::read_file(file_name);
  blocks.assign(file.size(), {});
  for (size_t b = 0; b < blocks.size(); ++b)
    {
      const auto it = file.find("a" + std::to_string(b));
      if (it == file.end())
        throw std::runtime_error("dataset /a" + std::to_string(b) + " not found in " + file_name);
      blocks[b] = it->second.data;
    }
}

// h5.h:187-208: a new file with the m x n matrix (row-major) as /a
inline void save_full_matrix(const std::string &file_name, size_t m, size_t n, const std::vector<double> &row_major)
{
  h5min::File file;
  h5min::Dataset &d = file["a"];
  d.dims = {m, n};
  d.data = row_major;
  h5min::write_file(file_name, file);
}

} // namespace H5

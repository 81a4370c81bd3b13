// engine_util.h -- small helpers shared by the native engines (engine.cu, engine_coll.cu).
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <vector>

#include "check.h"

namespace gmt {

// the engines hop between devices; callers (e.g. a torch process) get their current device back
struct DeviceRestore {
  int dev = -1;
  DeviceRestore()
  {
    if (cudaGetDevice(&dev) != cudaSuccess) dev = -1;
  }
  ~DeviceRestore()
  {
    if (dev >= 0) cudaSetDevice(dev);
  }
};

// bump allocator over one zero-initialised block, 256-byte granules
inline size_t take(size_t& cursor, size_t bytes)
{
  const size_t at = cursor;
  cursor += (bytes + 255) & ~size_t(255);
  return at;
}

// ranks grouped by the device that hosts them, in first-seen order
struct DeviceRanks {
  int device = 0;
  std::vector<int> ranks;
};

inline std::vector<DeviceRanks> group_by_device(const std::vector<int>& device_of_rank)
{
  std::vector<DeviceRanks> groups;
  for (int r = 0; r < (int)device_of_rank.size(); ++r) {
    size_t k = 0;
    while (k < groups.size() && groups[k].device != device_of_rank[r]) ++k;
    if (k == groups.size()) {
      groups.emplace_back();
      groups.back().device = device_of_rank[r];
    }
    groups[k].ranks.push_back(r);
  }
  return groups;
}

}  // namespace gmt

// srt_grid_setup.h -- host loader of the gridded-density file (modelnum 3).  PLACEHOLDER
// until the gridded path lands.
#pragma once
#include "srt_host_setup.h"
namespace srt_host {
struct HostGrid {
  std::vector<double> nodes, xs, ys, zs;
};
inline int load_grid(const char*, SrtModel&, HostGrid&) { return -1; }
}  // namespace srt_host

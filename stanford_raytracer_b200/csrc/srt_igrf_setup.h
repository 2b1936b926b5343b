// srt_igrf_setup.h -- host RECALC_08 (tsyganenko/geopack2008.for:486-1195).  PLACEHOLDER
// until the IGRF path lands.
#pragma once
#include "srt_host_setup.h"
namespace srt_host {
inline void igrf_recalc(int, int, SrtIgrf& ig) { memset(&ig, 0, sizeof ig); }
}  // namespace srt_host

// srt_tricubic.h -- gridded log-density by Lekien-Marsden tricubic interpolation
// (tricubic-for/libtricubic.f95:796-933 called from
// fortran/interp_dens_model_adapter.f95:190-206).  PLACEHOLDER until the gridded
// path lands: returns NaN densities so that a mis-dispatched model cannot pass.
#pragma once
#include "srt_physics.h"

SRT_HD void srt_grid_density(const SrtGrid& g, const double x[3], double Ns[SRT_NSPEC]) {
  (void)g; (void)x;
  for (int i = 0; i < SRT_NSPEC; ++i) Ns[i] = srt_nan();
}

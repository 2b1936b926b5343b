// srt_model.h -- plain-old-data description of a plasma model, as staged by the
// host (srt_host_setup.cpp) and read by the device code (srt_physics.h).
//
// One SrtModel replaces the reference's per-adapter state:
//   * module-global state of ngo_dens_model (fortran/ngo_dens_model.f95:8-24) for
//     the Ngo diffusive-equilibrium density, with every loop-invariant of `dens`
//     (sh(2:4), d2*2) hoisted -- the reference recomputes them on each call
//     because `init` stays -1 (ngo_dens_model.f95:180-186);
//   * interpStateData (fortran/interp_dens_model_adapter.f95:16-39) for the
//     gridded log-density;
//   * the time-only SM<->GSM rotation (xform_double/T4.f95) reduced to cos/sin of
//     the dipole tilt;
//   * Geopack's COMMON /GEOPACK1,2/ after RECALC_08 (tsyganenko/geopack2008.for)
//     for the IGRF field.
#pragma once
#include <stdint.h>

#define SRT_MAX_DUCTS 10
#define SRT_NSPEC 4

enum SrtDensityKind { SRT_DENS_NGO = 0, SRT_DENS_GRID = 1, SRT_DENS_SCAT = 2 };

struct SrtNgo {
  int num;     // number of species incl. electrons (NUM card), 2..4
  int kducts;  // 0: no plasmapause, 1: plasmapause, >=2: + ducts
  double r0;   // 6370. as a default-real literal
  double pi_f; // 3.141592653589793 as a default-real literal (Q14), used by the sinusoidal duct
  double rbase, ane0, alpha0[5];
  double rzero, scbot;
  double lk, expk, ddk, rconsn, scr;
  double sh[5];     // sh(2:4), hoisted
  double d2k_2;     // (ddk*ddk)*2.0, hoisted
  // ducts (k = 2..kducts)
  double l0[SRT_MAX_DUCTS + 1], def[SRT_MAX_DUCTS + 1], dd[SRT_MAX_DUCTS + 1], sidedu[SRT_MAX_DUCTS + 1];
  double rducln[SRT_MAX_DUCTS + 1], rducun[SRT_MAX_DUCTS + 1], rducls[SRT_MAX_DUCTS + 1], rducus[SRT_MAX_DUCTS + 1];
  double hl2n[SRT_MAX_DUCTS + 1], hl2s[SRT_MAX_DUCTS + 1], hu2n[SRT_MAX_DUCTS + 1], hu2s[SRT_MAX_DUCTS + 1];
};

struct SrtGrid {
  int nx, ny, nz, nspec;
  double minx, miny, minz;
  double delx, dely, delz;
  // device (or, in the CPU test build, host) pointers.
  // nodes: [nz][ny][nx][8 arrays][4 species] doubles -- the 32 values one cube corner contributes
  //        (f, dfdx, dfdy, dfdz, d2fdxdy, d2fdxdz, d2fdydz, d3fdxdydz; species fastest) are one
  //        contiguous 256-byte record.
  // coef:  [nz+1][ny+1][nx+1][64 terms][4 species] doubles -- the tricubic coefficients of every cell
  //        of the extended lattice (index = the 1-based lower corner srt_grid_locate returns, 0..n per axis:
  //        edge and outside cells included), a = Amat * (scaled, clamped corner data), precomputed once by srt_coef_kernel in
  //        the reference's accumulation order; term t = 16*i + 4*j + k holds a(i + 4j + 16k), the
  //        order tricubic_eval walks them (libtricubic.f95:677-693).  2 KiB per cell.  NULL: every
  //        query takes the on-the-fly path (CPU test build).
  // xs/ys/zs: node coordinates as the reference regenerates them (interp_dens_model_adapter.f95:90-95)
  const double* nodes;
  const double* coef;
  const double* xs;
  const double* ys;
  const double* zs;
};

// modelnum 4 (csrc/srt_scattered.h): scatteredinterpStateData (fortran/scattered_interp_dens_model_adapter.f95:17-48) with the
// kd-tree replaced by a uniform cell list over the samples' bounding box
struct SrtScat {
  int npts, nspec;
  int order, exact, scaled;  // lsinterp's arguments (driver:709-722: scaled is always 0 there)
  int nmono;                 // J, monomials of degree <= order
  unsigned char mono[64];    // exponents [J][3] in the order of tabular_monomials (lsinterp_mod.f95:73-95)
  double maxnearest, window_scale;
  double radius;             // maxnearest * window_scale, the search radius (adapter :247)
  double local_scale;        // local_window_scale
  double ox, oy, oz;         // corner of cell (0,0,0)
  double inv_cell;           // 1 / cell edge
  int ncx, ncy, ncz;
  // device (or, in the CPU test build, host) pointers
  const double* pts;         // [npts][8]: x, y, z, log-density of species 1..4, distance to the nearest other sample; sorted by cell
  const int* cell_start;     // [ncx*ncy*ncz + 1]: first record of each cell (x fastest)
};

struct SrtIgrf {
  // Schmidt-normalised coefficients and recursion constants after RECALC_08
  // (geopack2008.for:883-1029), fp32 as in the reference
  float g[106], h[106], rec[106];
  // GEO<->GSW rotation (geopack2008.for:1156-1164)
  float a11, a21, a31, a12, a22, a32, a13, a23, a33;
};

// TS05 external field (csrc/srt_ts05.h), use_tsyganenko = 1
struct SrtTs05 {
  int on;            // use_tsyganenko
  float parmod[10];  // Pdyn, Dst, ByIMF, BzIMF, W1..W6 as `real(parmod)` (ngo_dens_model_adapter.f95:184)
  float ps;          // the tilt argument `real(psi)` as the adapter passes it: word 1 of COMMON /GEOPACK1/ = ST0 (Q25)
};

struct SrtModel {
  int density;   // SrtDensityKind
  int use_igrf;  // 0: centred dipole, 1: IGRF
  int nspec;
  double del;    // dF/dx step: 1e-4 (Ngo, driver:251) or 1e-6 (grid, driver:252)
  double qs[SRT_NSPEC], ms[SRT_NSPEC];
  // loop invariants of stix_parameters / dispersion_relation (raytracer.f95:65,92-93)
  double q2[SRT_NSPEC];   // qs**2
  double fs_den;          // minval(ms)*EPS0
  double fs_q2;           // maxval(abs(qs))**2
  double cos_mu, sin_mu;  // cos/sin of the dipole tilt angle mu (T4.f95:13)
  SrtNgo ngo;
  SrtGrid grid;
  SrtScat scat;
  SrtIgrf igrf;
  SrtTs05 ts;
};

// srt_igrf_setup.h -- host side of the IGRF field: what TSY_RECALC -> RECALC_08
// (tsyganenko/geopack0508_adapter.for:21-30, tsyganenko/geopack2008.for:486-1195) and SUN_08
// (:333-380) leave in COMMON /GEOPACK1,2/ for IGRF_GSW_08 and GEOGSW_08: the Schmidt-normalised
// coefficients G, H, the recursion constants REC and the GEO->GSW matrix A11..A33.
//
// The reference re-runs this on EVERY funcPlasmaParams call (ngo_dens_model_adapter.f95:169) although
// it depends on time only; here it runs once per model.  All arithmetic is REAL*4 exactly where the
// Fortran is (implicit typing), DOUBLE PRECISION where SUN_08 declares it; the fp32 elementary
// functions are the host libm's (sinf, cosf, ...), which is what a gfortran build links.
#pragma once
#include <math.h>

#include "srt_host_setup.h"
#include "srt_igrf_tables.h"

namespace srt_host {

// SUN_08: only GST, SRASN, SDEC are used by RECALC_08
inline void sun_08(int iyear, int iday, int ihour, int min, int isec, float& gst, float& slong, float& srasn, float& sdec) {
  const float RAD = 57.295779513f;
  if (iyear < 1901 || iyear > 2099) return;
  double fday = (double)(ihour * 3600 + min * 60 + isec) / 86400.0;
  double dj = (double)(365 * (iyear - 1900) + (iyear - 1901) / 4 + iday) - 0.5 + fday;
  float t = (float)(dj / (double)36525.f);
  float vl = (float)fmod((double)279.696678f + (double)0.9856473354f * dj, 360.0);
  gst = (float)(fmod((double)279.690983f + (double).9856473354f * dj + (double)360.f * fday + (double)180.f, 360.0) / (double)RAD);
  float g = (float)(fmod((double)358.475845f + (double)0.985600267f * dj, 360.0) / (double)RAD);
  slong = (vl + (1.91946f - 0.004789f * t) * sinf(g) + 0.020094f * sinf(2.f * g)) / RAD;
  if (slong > 6.2831853f) slong = slong - 6.2831853f;
  if (slong < 0.f) slong = slong + 6.2831853f;
  float obliq = (23.45229f - 0.0130125f * t) / RAD;
  float sob = sinf(obliq);
  float slp = slong - 9.924E-5f;
  float sind = sob * sinf(slp);
  float cosd = sqrtf(1.f - sind * sind);
  float sc = sind / cosd;
  sdec = atanf(sc);
  srasn = 3.141592654f - atan2f(cosf(obliq) / sob * sc, -cosf(slp) / cosd);
}

// RECALC_08 with VGSE = (-400, 0, 0) (TSY_RECALC).  g/h/rec are 1-based like the Fortran arrays.
inline void recalc_08(int iyear, int iday, int ihour, int min, int isec, SrtIgrf& ig, float* st0_out = nullptr) {
  using namespace srt_igrf_tables;
  memset(&ig, 0, sizeof ig);
  float* G = ig.g;
  float* H = ig.h;
  float* REC = ig.rec;
  int iy = iyear;
  if (iy < 1965) iy = 1965;  // the reference prints a warning (:863-871); the clamp is the behaviour
  if (iy > 2025) iy = 2025;
  for (int n = 1; n <= 14; ++n) {
    int n2 = 2 * n - 1;
    n2 = n2 * (n2 - 2);
    for (int m = 1; m <= n; ++m) {
      int mn = n * (n - 1) / 2 + m;
      REC[mn] = (float)((n - m) * (n + m - 2)) / (float)n2;
    }
  }
  if (iy >= 2020) {
    // extrapolate beyond 2020 with the secular variation (orders 1-8 only), :915-923
    float dt = (float)iy + (float)(iday - 1) / 365.25f - 2020.f;
    for (int n = 1; n <= 105; ++n) {
      G[n] = G20[n - 1];
      H[n] = H20[n - 1];
      if (n > 45) continue;
      G[n] = G[n] + DG20[n - 1] * dt;
      H[n] = H[n] + DH20[n - 1] * dt;
    }
  } else {
    // linear interpolation inside the 5-year bracket [e, e+5), :925-1012
    int b = (iy - 1965) / 5;  // 0: 1965-70 ... 10: 2015-20
    float f2 = ((float)iy + (float)(iday - 1) / 365.25f - (float)(1965 + 5 * b)) / 5.f;
    float f1 = 1.f - f2;
    const float* ga = G_EPOCH[b];
    const float* gb = G_EPOCH[b + 1];
    const float* ha = H_EPOCH[b];
    const float* hb = H_EPOCH[b + 1];
    for (int n = 1; n <= 105; ++n) {
      G[n] = ga[n - 1] * f1 + gb[n - 1] * f2;
      H[n] = ha[n - 1] * f1 + hb[n - 1] * f2;
    }
  }
  // Schmidt normalisation, :1016-1029
  float s = 1.f;
  for (int n = 2; n <= 14; ++n) {
    int mn = n * (n - 1) / 2 + 1;
    s = s * (float)(2 * n - 3) / (float)(n - 1);
    G[mn] = G[mn] * s;
    H[mn] = H[mn] * s;
    float p = s;
    for (int m = 2; m <= n; ++m) {
      float aa = 1.f;
      if (m == 2) aa = 2.f;
      p = p * sqrtf(aa * (float)(n - m + 1) / (float)(n + m - 2));
      int mnn = mn + m - 1;
      G[mnn] = G[mnn] * p;
      H[mnn] = H[mnn] * p;
    }
  }
  // dipole axis, :1031-1046
  float g_10 = -G[2], g_11 = G[3], h_11 = H[3];
  float sq = g_11 * g_11 + h_11 * h_11;
  float sqq = sqrtf(sq);
  float sqr = sqrtf(g_10 * g_10 + sq);
  float sl0 = -h_11 / sqq;
  float cl0 = -g_11 / sqq;
  float st0 = sqq / sqr;
  if (st0_out) *st0_out = st0;  // word 1 of COMMON /GEOPACK1/: what the adapters' `PSI` aliases (Q25)
  float ct0 = g_10 / sqr;
  float stcl = st0 * cl0;
  float stsl = st0 * sl0;
  // Sun direction and the GSW axes in GEI, :1048-1110
  float gst = 0.f, slong = 0.f, srasn = 0.f, sdec = 0.f;
  sun_08(iy, iday, ihour, min, isec, gst, slong, srasn, sdec);
  float s1 = cosf(srasn) * cosf(sdec);
  float s2 = sinf(srasn) * cosf(sdec);
  float s3 = sinf(sdec);
  float dj = (float)(365 * (iy - 1900) + (iy - 1901) / 4 + iday) - 0.5f + (float)(ihour * 3600 + min * 60 + isec) / 86400.f;
  float t = dj / 36525.f;
  float obliq = (23.45229f - 0.0130125f * t) / 57.2957795f;
  float dz1 = 0.f;
  float dz2 = -sinf(obliq);
  float dz3 = cosf(obliq);
  float dy1 = dz2 * s3 - dz3 * s2;
  float dy2 = dz3 * s1 - dz1 * s3;
  float dy3 = dz1 * s2 - dz2 * s1;
  const float vgsex = -400.0f, vgsey = 0.0f, vgsez = 0.0f;
  float v = sqrtf(vgsex * vgsex + vgsey * vgsey + vgsez * vgsez);
  float dx1 = -vgsex / v;
  float dx2 = -vgsey / v;
  float dx3 = -vgsez / v;
  float x1 = dx1 * s1 + dx2 * dy1 + dx3 * dz1;
  float x2 = dx1 * s2 + dx2 * dy2 + dx3 * dz2;
  float x3 = dx1 * s3 + dx2 * dy3 + dx3 * dz3;
  float cgst = cosf(gst);
  float sgst = sinf(gst);
  float dip1 = stcl * cgst - stsl * sgst;
  float dip2 = stcl * sgst + stsl * cgst;
  float dip3 = ct0;
  float y1 = dip2 * x3 - dip3 * x2;
  float y2 = dip3 * x1 - dip1 * x3;
  float y3 = dip1 * x2 - dip2 * x1;
  float y = sqrtf(y1 * y1 + y2 * y2 + y3 * y3);
  y1 = y1 / y;
  y2 = y2 / y;
  y3 = y3 / y;
  float z1 = x2 * y3 - x3 * y2;
  float z2 = x3 * y1 - x1 * y3;
  float z3 = x1 * y2 - x2 * y1;
  // GEO -> GSW, :1156-1164
  ig.a11 = x1 * cgst + x2 * sgst;
  ig.a12 = -x1 * sgst + x2 * cgst;
  ig.a13 = x3;
  ig.a21 = y1 * cgst + y2 * sgst;
  ig.a22 = -y1 * sgst + y2 * cgst;
  ig.a23 = y3;
  ig.a31 = z1 * cgst + z2 * sgst;
  ig.a32 = -z1 * sgst + z2 * cgst;
  ig.a33 = z3;
}

// itime = (yyyyddd, milliseconds of day) -> the five integers the adapters pass to tsy_recalc
// (ngo_dens_model_adapter.f95:149-153)
inline void igrf_recalc(int yearday, int msec, SrtIgrf& ig, float* st0_out = nullptr) {
  int year = yearday / 1000;
  int day = yearday % 1000;
  int hour = msec / (1000 * 60 * 60);
  int min = (msec - hour * (1000 * 60 * 60)) / (1000 * 60);
  int sec = (msec - hour * (1000 * 60 * 60) - min * (1000 * 60)) / 1000;
  recalc_08(year, day, hour, min, sec, ig, st0_out);
}

}  // namespace srt_host

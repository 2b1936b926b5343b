// srt_igrf.h -- IGRF main field in fp32 (IGRF_GSW_08, tsyganenko/geopack2008.for:55-181,
// with GEOGSW_08 :1421-1456 inlined).  Every operation is REAL*4 as in the reference
// (implicit typing), in the reference's order; compile with contraction off.
#pragma once
#include "srt_physics.h"

SRT_HDN void srt_igrf_gsw(const SrtIgrf& ig, float xgsw, float ygsw, float zgsw, float& hxgsw, float& hygsw, float& hzgsw) {
  SRT_COUNT(SRT_T_IGRF);
  // GEOGSW_08, J < 0
  float xgeo = ig.a11 * xgsw + ig.a21 * ygsw + ig.a31 * zgsw;
  float ygeo = ig.a12 * xgsw + ig.a22 * ygsw + ig.a32 * zgsw;
  float zgeo = ig.a13 * xgsw + ig.a23 * ygsw + ig.a33 * zgsw;
  float rho2 = xgeo * xgeo + ygeo * ygeo;
  float r = srt_sqrtf(rho2 + zgeo * zgeo);
  float c = zgeo / r;
  float rho = srt_sqrtf(rho2);
  float s = rho / r;
  float cf, sf;
  const bool polar = s < 1.E-5f;
  if (polar) { cf = 1.f; sf = 0.f; }
  else { cf = xgeo / rho; sf = ygeo / rho; }
  float pp = 1.f / r;
  float p = pp;
  int irp3 = (int)(r + 2.f);
  int nm = 3 + 30 / irp3;
  if (nm > 13) nm = 13;
  const int k = nm + 1;
  float a[15], b[15];
  for (int n = 1; n <= k; ++n) {
    p = p * pp;
    a[n] = p;
    b[n] = p * (float)n;
  }
  p = 1.f;
  float d = 0.f, bbr = 0.f, bbt = 0.f, bbf = 0.f;
  float x = 0.f, y = 1.f;
  for (int m = 1; m <= k; ++m) {
    if (m == 1) { x = 0.f; y = 1.f; }
    else {
      float w = x;
      x = w * cf + y * sf;
      y = y * cf - w * sf;
    }
    float q = p, z = d, bi = 0.f, p2 = 0.f, d2 = 0.f;
    for (int n = m; n <= k; ++n) {
      float an = a[n];
      int mn = n * (n - 1) / 2 + m;
      float e = ig.g[mn], hh = ig.h[mn];
      float w = e * y + hh * x;
      bbr = bbr + b[n] * w * q;
      bbt = bbt - an * w * z;
      if (m != 1) {
        float qq = q;
        if (polar) qq = z;
        bi = bi + an * (e * x - hh * y) * qq;
      }
      float xk = ig.rec[mn];
      float dp = c * z - s * q - xk * d2;
      float pm = c * q - xk * p2;
      d2 = z;
      p2 = q;
      z = dp;
      q = pm;
    }
    d = s * d + c * p;
    p = s * p;
    if (m != 1) {
      bi = bi * (float)(m - 1);
      bbf = bbf + bi;
    }
  }
  float br = bbr, bt = bbt, bf;
  if (!polar) bf = bbf / s;
  else {
    if (c < 0.f) bbf = -bbf;
    bf = bbf;
  }
  float he = br * s + bt * c;
  float hxgeo = he * cf - bf * sf;
  float hygeo = he * sf + bf * cf;
  float hzgeo = br * c - bt * s;
  // GEOGSW_08, J > 0
  hxgsw = ig.a11 * hxgeo + ig.a12 * hygeo + ig.a13 * hzgeo;
  hygsw = ig.a21 * hxgeo + ig.a22 * hygeo + ig.a23 * hzgeo;
  hzgsw = ig.a31 * hxgeo + ig.a32 * hygeo + ig.a33 * hzgeo;
}

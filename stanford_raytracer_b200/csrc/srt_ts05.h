// srt_ts05.h -- the Tsyganenko-Sitnov 2005 external magnetic field (T04_s, tsyganenko/TS05_aka_TS04.for:5-2544), the
// term the adapters add to the base field when use_tsyganenko = 1 (fortran/ngo_dens_model_adapter.f95:183-198,
// interp_dens_model_adapter.f95:249-264).  SURVEY.md section 8 row f4.
//
// One definition for host and device (like srt_math.h): REAL*8 arithmetic inside, REAL*4 at the boundary, operations in
// the reference's order so that results can be compared bit for bit with the reference's own source (translated at
// build time into oracle/_ref, see oracle/f77_to_cpp.py) when both use the elementary functions of srt_math.h.
// Restructured, not transliterated: the unrolled 3x3 / 5x5 harmonic sums are loops over the coefficient tables
// (srt_ts05_tables.h), the four copies of the elliptic-integral vector potential are one function, the two shielding
// sums (BIRK_SHL, RC_SHIELD) are one function, and the Birkeland modes whose amplitude the model never uses (R1 mode 2,
// R2 mode 2: EXTERN adds only BXR11 and BXR21, :318-325) are not evaluated.
// Quirks kept (they change bits): literals without a D exponent are REAL*4 (`COST1=.99994999875`, four coefficients of
// the elliptic-integral polynomials, `DELTADY=4.7`, the exponents `**0.37`), x**3.D0 goes through pow().
#pragma once
#include "srt_math.h"
#include "srt_model.h"
#include "srt_ts05_tables.h"


SRT_HD double ts05_sq(double x) { return x * x; }
SRT_HD double ts05_tan(double x) { double s, c; srt_sincos(x, s, c); return s / c; }

// SHLCAR3X3 (:362-692): shielding field of the dipole, 2 x 9 Cartesian harmonics
SRT_HDN void ts05_shlcar3x3(double x, double y, double z, double ps, double& bx, double& by, double& bz) {
  const double* a = ts05_shlcar3x3_a();
  const double t1 = a[48], t2 = a[49];
  double sps, cps, st1, ct1, st2, ct2;
  srt_sincos(ps, sps, cps);
  const double s2ps = 2.0 * cps;
  srt_sincos(ps * t1, st1, ct1);
  srt_sincos(ps * t2, st2, ct2);
  const double x1 = x * ct1 - z * st1, z1 = x * st1 + z * ct1;
  const double x2 = x * ct2 - z * st2, z2 = x * st2 + z * ct2;
  double gx = 0.0, gy = 0.0, gz = 0.0;
  for (int i = 0; i < 3; ++i) {
    const double p = a[36 + i];
    double syp, cyp;
    srt_sincos(y / p, syp, cyp);
    for (int k = 0; k < 3; ++k) {
      const double r = a[39 + k];
      const double sqpr = srt_sqrt(1.0 / ts05_sq(p) + 1.0 / ts05_sq(r));
      double szr, czr;
      srt_sincos(z1 / r, szr, czr);
      const double expr = srt_exp(sqpr * x1);
      double fx, hy, fz;
      if (k < 2) {
        fx = -sqpr * expr * cyp * szr;
        hy = expr / p * syp * szr;
        fz = -expr * cyp / r * czr;
      } else {
        fx = -expr * cyp * (sqpr * z1 * czr + szr / r * (x1 + 1.0 / sqpr));
        hy = expr / p * syp * (z1 * czr + x1 / r * szr / sqpr);
        fz = -expr * cyp * (czr * (1.0 + x1 / ts05_sq(r) / sqpr) - z1 / r * szr);
      }
      const double hx = fx * ct1 + fz * st1, hz = -fx * st1 + fz * ct1;
      const int n = 3 * i + k;
      const double c = a[2 * n] + a[2 * n + 1] * cps;
      if (n == 0) { gx = c * hx; gy = c * hy; gz = c * hz; }
      else { gx = gx + c * hx; gy = gy + c * hy; gz = gz + c * hz; }
    }
  }
  for (int i = 0; i < 3; ++i) {
    const double q = a[42 + i];
    double syq, cyq;
    srt_sincos(y / q, syq, cyq);
    for (int k = 0; k < 3; ++k) {
      const double s = a[45 + k];
      const double sqqs = srt_sqrt(1.0 / ts05_sq(q) + 1.0 / ts05_sq(s));
      double szs, czs;
      srt_sincos(z2 / s, szs, czs);
      const double exqs = srt_exp(sqqs * x2);
      const double fx = -sqqs * exqs * cyq * czs * sps;
      const double hy = exqs / q * syq * czs * sps;
      const double fz = exqs * cyq / s * szs * sps;
      const double hx = fx * ct2 + fz * st2, hz = -fx * st2 + fz * ct2;
      const int n = 3 * i + k;
      const double c = a[18 + 2 * n] + a[19 + 2 * n] * s2ps;
      gx = gx + c * hx; gy = gy + c * hy; gz = gz + c * hz;
    }
  }
  bx = gx; by = gy; bz = gz;
}

// ------------------------------------------------------------------ tail field
// TAILDISK (:933-1022): the disk-like current sheet, five terms
SRT_HDN void ts05_taildisk(double d0, double deltadx, double deltady, double x, double y, double z, double& bx, double& by, double& bz) {
  const double *f = ts05_taildisk_f(), *b = ts05_taildisk_b(), *c = ts05_taildisk_c();
  const double rho = srt_sqrt(ts05_sq(x) + ts05_sq(y));
  const double drhodx = x / rho, drhody = y / rho;
  const double dex = srt_exp(x / 7.0);
  const double d = d0 + deltady * ts05_sq(y / 20.0) + deltadx * dex;
  const double dddy = deltady * y * 0.005;
  const double dddx = deltadx / 7.0 * dex;
  const double dzeta = srt_sqrt(ts05_sq(z) + ts05_sq(d));
  const double ddzetadx = d * dddx / dzeta, ddzetady = d * dddy / dzeta, ddzetadz = z / dzeta;
  double dbx = 0.0, dby = 0.0, dbz = 0.0;
  for (int i = 0; i < 5; ++i) {
    const double bi = b[i], ci = c[i];
    const double s1 = srt_sqrt(ts05_sq(rho + bi) + ts05_sq(dzeta + ci));
    const double s2 = srt_sqrt(ts05_sq(rho - bi) + ts05_sq(dzeta + ci));
    const double ds1drho = (rho + bi) / s1, ds2drho = (rho - bi) / s2;
    const double ds1ddz = (dzeta + ci) / s1, ds2ddz = (dzeta + ci) / s2;
    const double ds1dx = ds1drho * drhodx + ds1ddz * ddzetadx;
    const double ds1dy = ds1drho * drhody + ds1ddz * ddzetady;
    const double ds1dz = ds1ddz * ddzetadz;
    const double ds2dx = ds2drho * drhodx + ds2ddz * ddzetadx;
    const double ds2dy = ds2drho * drhody + ds2ddz * ddzetady;
    const double ds2dz = ds2ddz * ddzetadz;
    const double s1ts2 = s1 * s2, s1ps2 = s1 + s2;
    const double s1ps2sq = ts05_sq(s1ps2);
    const double fac1 = srt_sqrt(s1ps2sq - ts05_sq(2.0 * bi));
    const double as = fac1 / (s1ts2 * s1ps2sq);
    const double dasds1 = (1.0 / (fac1 * s2) - as / s1ps2 * (s2 * s2 + s1 * (3.0 * s1 + 4.0 * s2))) / (s1 * s1ps2);
    const double dasds2 = (1.0 / (fac1 * s1) - as / s1ps2 * (s1 * s1 + s2 * (3.0 * s2 + 4.0 * s1))) / (s2 * s1ps2);
    const double dasdx = dasds1 * ds1dx + dasds2 * ds2dx;
    const double dasdy = dasds1 * ds1dy + dasds2 * ds2dy;
    const double dasdz = dasds1 * ds1dz + dasds2 * ds2dz;
    dbx = dbx - f[i] * x * dasdz;
    dby = dby - f[i] * y * dasdz;
    dbz = dbz + f[i] * (2.0 * as + x * dasdx + y * dasdy);
  }
  bx = dbx; by = dby; bz = dbz;
}

// SHLCAR5X5 (:1024-1076): shielding field of one tail mode, 25 harmonics with shift-dependent amplitudes
SRT_HDN void ts05_shlcar5x5(const double* a, double x, double y, double z, double dshift, double& hx, double& hy, double& hz) {
  double dhx = 0.0, dhy = 0.0, dhz = 0.0;
  int l = 0;
  for (int i = 0; i < 5; ++i) {
    const double rp = 1.0 / a[50 + i];
    double sypi, cypi;
    srt_sincos(y * rp, sypi, cypi);
    for (int k = 0; k < 5; ++k) {
      const double rr = 1.0 / a[55 + k];
      double szrk, czrk;
      srt_sincos(z * rr, szrk, czrk);
      const double sqpr = srt_sqrt(ts05_sq(rp) + ts05_sq(rr));
      const double epr = srt_exp(x * sqpr);
      const double dbx = -sqpr * epr * cypi * szrk;
      const double dby = rp * epr * sypi * szrk;
      const double dbz = -rr * epr * cypi * czrk;
      const double coef = a[l] + a[l + 1] * dshift;
      l += 2;
      dhx = dhx + coef * dbx;
      dhy = dhy + coef * dby;
      dhz = dhz + coef * dbz;
    }
  }
  hx = dhx; hy = dhy; hz = dhz;
}

struct Ts05Tail {  // COMMON /TAIL/ (:171-175)
  double dxshift1, dxshift2, d, deltady;
};

// UNWARPED (:837-931), both modes (IOPT = 0)
SRT_HD void ts05_unwarped(const Ts05Tail& T, double x, double y, double z, double b1[3], double b2[3]) {
  double f[3], h[3];
  {
    const double alpha = TS05_UNWARPED_ALPHA1, xm = TS05_UNWARPED_XM1;
    const double xsc = (x - TS05_UNWARPED_XSHIFT1 - T.dxshift1) * alpha - xm * (alpha - 1.0);
    ts05_taildisk(T.d * alpha, TS05_UNWARPED_DELTADX1, T.deltady, xsc, y * alpha, z * alpha, f[0], f[1], f[2]);
    ts05_shlcar5x5(ts05_unwarped_a1(), x, y, z, T.dxshift1, h[0], h[1], h[2]);
    for (int i = 0; i < 3; ++i) b1[i] = f[i] + h[i];
  }
  {
    const double alpha = TS05_UNWARPED_ALPHA2, xm = TS05_UNWARPED_XM2;
    const double xsc = (x - TS05_UNWARPED_XSHIFT2 - T.dxshift2) * alpha - xm * (alpha - 1.0);
    ts05_taildisk(T.d * alpha, TS05_UNWARPED_DELTADX2, T.deltady, xsc, y * alpha, z * alpha, f[0], f[1], f[2]);
    ts05_shlcar5x5(ts05_unwarped_a2(), x, y, z, T.dxshift2, h[0], h[1], h[2]);
    for (int i = 0; i < 3; ++i) b2[i] = f[i] + h[i];
  }
}

// WARPED (:764-835): the tail field in the warped (y, z) plane, G = warping parameter (/Gblock/)
SRT_HD void ts05_warped(const Ts05Tail& T, double g, double ps, double x, double y, double z, double b1[3], double b2[3]) {
  const double dgdx = 0.0, xl = 20.0, dxldx = 0.0;
  const double sps = srt_sin(ps);
  const double rho2 = ts05_sq(y) + ts05_sq(z);
  const double rho = srt_sqrt(rho2);
  double phi, cphi, sphi;
  if (y == 0.0 && z == 0.0) { phi = 0.0; cphi = 1.0; sphi = 0.0; }
  else { phi = srt_atan2(z, y); cphi = y / rho; sphi = z / rho; }
  const double xl4 = ts05_sq(xl) * ts05_sq(xl);
  const double rr4l4 = rho / (ts05_sq(rho2) + xl4);
  const double f = phi + g * rho2 * rr4l4 * cphi * sps;
  const double dfdphi = 1.0 - g * rho2 * rr4l4 * sphi * sps;
  const double dfdrho = g * ts05_sq(rr4l4) * (3.0 * xl4 - ts05_sq(rho2)) * cphi * sps;
  const double dfdx = rr4l4 * cphi * sps * (dgdx * rho2 - g * rho * rr4l4 * 4.0 * (xl * xl * xl) * dxldx);
  double sf, cf;
  srt_sincos(f, sf, cf);
  double a1[3], a2[3];
  ts05_unwarped(T, x, rho * cf, rho * sf, a1, a2);
  for (int m = 0; m < 2; ++m) {
    const double* as = m ? a2 : a1;
    double* b = m ? b2 : b1;
    const double brho_as = as[1] * cf + as[2] * sf;
    const double bphi_as = -as[1] * sf + as[2] * cf;
    const double brho_s = brho_as * dfdphi;
    const double bphi_s = bphi_as - rho * (as[0] * dfdx + brho_as * dfdrho);
    b[0] = as[0] * dfdphi;
    b[1] = brho_s * cphi - bphi_s * sphi;
    b[2] = brho_s * sphi + bphi_s * cphi;
  }
}

// DEFORMED (:694-762): tilt-dependent bending of the tail sheet, RH0 = hinging distance (/RH0block/)
SRT_HDN void ts05_deformed(const Ts05Tail& T, double rh0, double g, double ps, double x, double y, double z, double b1[3], double b2[3]) {
  const double rh2 = TS05_DEFORMED_RH2;
  const double sps = srt_sin(ps);
  const double r2 = ts05_sq(x) + ts05_sq(y) + ts05_sq(z);
  const double r = srt_sqrt(r2);
  const double zr = z / r;
  const double rh = rh0 + rh2 * ts05_sq(zr);
  const double drhdr = -zr / r * 2.0 * rh2 * zr;
  const double drhdz = 2.0 * rh2 * zr / r;
  const double rrh = r / rh;
  // IEPS = 3: RRH**IEPS, RRH**(IEPS-1), F**(IEPS+1) are libgcc's integer powers; (..)**(1.D0/IEPS) is pow()
  const double rrh2 = rrh * rrh;
  const double f = 1.0 / srt_pow(1.0 + rrh * rrh2, 1.0 / 3.0);
  const double f2 = f * f;
  const double dfdr = -rrh2 * (f2 * f2) / rh;
  const double dfdrh = -rrh * dfdr;
  const double spsas = sps * f;
  const double cpsas = srt_sqrt(1.0 - ts05_sq(spsas));
  const double xas = x * cpsas - z * spsas;
  const double zas = x * spsas + z * cpsas;
  const double facps = sps / cpsas * (dfdr + dfdrh * drhdr) / r;
  const double psasx = facps * x, psasy = facps * y;
  const double psasz = facps * z + sps / cpsas * dfdrh * drhdz;
  const double dxasdx = cpsas - zas * psasx, dxasdy = -zas * psasy, dxasdz = -spsas - zas * psasz;
  const double dzasdx = spsas + xas * psasx, dzasdy = xas * psasy, dzasdz = cpsas + xas * psasz;
  const double fac1 = dxasdz * dzasdy - dxasdy * dzasdz;
  const double fac2 = dxasdx * dzasdz - dxasdz * dzasdx;
  const double fac3 = dzasdx * dxasdy - dxasdx * dzasdy;
  double a1[3], a2[3];
  ts05_warped(T, g, ps, xas, y, zas, a1, a2);
  for (int m = 0; m < 2; ++m) {
    const double* as = m ? a2 : a1;
    double* b = m ? b2 : b1;
    b[0] = as[0] * dzasdz - as[2] * dxasdz + as[1] * fac1;
    b[1] = as[1] * fac2;
    b[2] = as[2] * dxasdx - as[0] * dzasdx + as[1] * fac3;
  }
}

// ------------------------------------------------------------------ Birkeland (field-aligned) currents
// R_S, THETA_S (:1424-1452): the stretch transformation of the conical harmonics
SRT_HD double ts05_r_s(const double* a, double r, double theta) {
  const double r2 = ts05_sq(r);
  return r + a[1] / r + a[2] * r / srt_sqrt(r2 + ts05_sq(a[10])) + a[3] * r / (r2 + ts05_sq(a[11])) +
         (a[4] + a[5] / r + a[6] * r / srt_sqrt(r2 + ts05_sq(a[12])) + a[7] * r / (r2 + ts05_sq(a[13]))) * srt_cos(theta) +
         (a[8] * r / srt_sqrt(r2 + ts05_sq(a[14])) + a[9] * r / ts05_sq(r2 + ts05_sq(a[15]))) * srt_cos(2.0 * theta);
}
SRT_HD double ts05_theta_s(const double* a, double r, double theta) {
  const double r2 = ts05_sq(r);
  return theta + (a[16] + a[17] / r + a[18] / r2 + a[19] * r / srt_sqrt(r2 + ts05_sq(a[26]))) * srt_sin(theta) +
         (a[20] + a[21] * r / srt_sqrt(r2 + ts05_sq(a[27])) + a[22] * r / (r2 + ts05_sq(a[28]))) * srt_sin(2.0 * theta) +
         (a[23] + a[24] / r + a[25] * r / (r2 + ts05_sq(a[29]))) * srt_sin(3.0 * theta);
}

// FIALCOS (:1454-1530): conical-harmonic field of mode n (1 or 2) at (r, theta, phi); only the n-th harmonic is returned
SRT_HDN void ts05_fialcos(double r, double theta, double phi, double& btheta, double& bphi, int n, double theta0, double dt) {
  double sinte, coste, sinfi, cosfi;
  srt_sincos(theta, sinte, coste);
  srt_sincos(phi, sinfi, cosfi);
  const double ro = r * sinte;
  const double tg = sinte / (1.0 + coste), ctg = sinte / (1.0 - coste);
  const double tetanp = theta0 + dt, tetanm = theta0 - dt;
  double tgp = 0.0, tgm = 0.0, tgm2 = 0.0, tgp2 = 0.0;
  if (!(theta < tetanm)) {
    tgp = ts05_tan(tetanp * 0.5);
    tgm = ts05_tan(tetanm * 0.5);
    tgm2 = tgm * tgm;
    tgp2 = tgp * tgp;
  }
  double cosm1 = 1.0, sinm1 = 0.0, tm = 1.0, tgm2m = 1.0, tgp2m = 1.0, btn = 0.0, bpn = 0.0;
  for (int m = 1; m <= n; ++m) {
    tm = tm * tg;
    const double ccos = cosm1 * cosfi - sinm1 * sinfi;
    const double ssin = sinm1 * cosfi + cosm1 * sinfi;
    cosm1 = ccos;
    sinm1 = ssin;
    const double dm = (double)m;
    double t, dtt;
    if (theta < tetanm) {
      t = tm;
      dtt = 0.5 * dm * tm * (tg + ctg);
    } else if (theta < tetanp) {
      tgm2m = tgm2m * tgm2;
      const double fc = 1.0 / (tgp - tgm);
      const double fc1 = 1.0 / (double)(2 * m + 1);
      const double tgm2m1 = tgm2m * tgm;
      const double tg21 = 1.0 + tg * tg;
      t = fc * (tm * (tgp - tg) + fc1 * (tm * tg - tgm2m1 / tm));
      dtt = 0.5 * dm * fc * tg21 * (tm / tg * (tgp - tg) - fc1 * (tm - tgm2m1 / (tm * tg)));
    } else {
      tgp2m = tgp2m * tgp2;
      tgm2m = tgm2m * tgm2;
      const double fc = 1.0 / (tgp - tgm);
      const double fc1 = 1.0 / (double)(2 * m + 1);
      t = fc * fc1 * (tgp2m * tgp - tgm2m * tgm) / tm;
      dtt = -t * dm * 0.5 * (tg + ctg);
    }
    btn = dm * t * ccos / ro;
    bpn = -dtt * ssin / r;
  }
  btheta = btn * 800.0;
  bphi = bpn * 800.0;
}

// ONE_CONE (:1361-1422): one hemisphere's cone; derivatives of the stretch by central differences (DR = DT = 1e-6)
SRT_HDN void ts05_one_cone(const double* a, int mode, double dtheta, double x, double y, double z, double& bx, double& by, double& bz) {
  const double dr = TS05_ONE_CONE_DR, dt = TS05_ONE_CONE_DT;
  const double theta0 = a[30];
  const double rho2 = ts05_sq(x) + ts05_sq(y);
  const double rho = srt_sqrt(rho2);
  const double r = srt_sqrt(rho2 + ts05_sq(z));
  const double theta = srt_atan2(rho, z);
  const double phi = srt_atan2(y, x);
  const double rs = ts05_r_s(a, r, theta);
  const double thetas = ts05_theta_s(a, r, theta);
  double btast, bfast;
  ts05_fialcos(rs, thetas, phi, btast, bfast, mode, theta0, dtheta);
  const double drsdr = (ts05_r_s(a, r + dr, theta) - ts05_r_s(a, r - dr, theta)) / (2.0 * dr);
  const double drsdt = (ts05_r_s(a, r, theta + dt) - ts05_r_s(a, r, theta - dt)) / (2.0 * dt);
  const double dtsdr = (ts05_theta_s(a, r + dr, theta) - ts05_theta_s(a, r - dr, theta)) / (2.0 * dr);
  const double dtsdt = (ts05_theta_s(a, r, theta + dt) - ts05_theta_s(a, r, theta - dt)) / (2.0 * dt);
  const double stsst = srt_sin(thetas) / srt_sin(theta);
  const double rsr = rs / r;
  const double br = -rsr / r * stsst * btast * drsdt;
  const double btheta = rsr * stsst * btast * drsdr;
  const double bphi = rsr * bfast * (drsdr * dtsdt - drsdt * dtsdr);
  const double s = rho / r, c = z / r, sf = y / rho, cf = x / rho;
  const double be = br * s + btheta * c;
  bx = a[0] * (be * cf - bphi * sf);
  by = a[0] * (be * sf + bphi * cf);
  bz = a[0] * (br * c - btheta * s);
}

// BIRK_1N2 (:1211-1339) with TWOCONES (:1341-1359): region `numb` (1, 2), mode `mode` (1, 2); xkappa = /DPHI_B_RHO0/ XKAPPA
SRT_HDN void ts05_birk_1n2(int numb, int mode, double xkappa, double ps, double x, double y, double z, double b[3]) {
  const double beta = TS05_BIRK_1N2_BETA, rh = TS05_BIRK_1N2_RH, eps = TS05_BIRK_1N2_EPS;
  const double bb = 0.5, rho_0 = 7.0;
  const double dphi = numb == 1 ? 0.055 : 0.030, dtheta = numb == 1 ? 0.06 : 0.09;
  const double xsc = x * xkappa, ysc = y * xkappa, zsc = z * xkappa;
  const double rho = srt_sqrt(ts05_sq(xsc) + ts05_sq(zsc));
  const double rsc = srt_sqrt(ts05_sq(xsc) + ts05_sq(ysc) + ts05_sq(zsc));
  const double rho2 = ts05_sq(rho_0);
  const double phi = (xsc == 0.0 && zsc == 0.0) ? 0.0 : srt_atan2(-zsc, xsc);
  double sphic, cphic;
  srt_sincos(phi, sphic, cphic);
  const double brack = dphi + bb * rho2 / (rho2 + 1.0) * (ts05_sq(rho) - 1.0) / (rho2 + ts05_sq(rho));
  const double r1rh = (rsc - 1.0) / rh;
  const double p_eps = srt_pow(r1rh, eps);                  // R1RH**EPS
  const double p_em1 = srt_pow(r1rh, eps - 1.0);            // R1RH**(EPS-1.D0)
  const double psias = beta * ps / srt_pow(1.0 + p_eps, 1.0 / eps);
  const double phis = phi - brack * sphic - psias;
  const double dphisphi = 1.0 - brack * cphic;
  const double den = rh * rsc * srt_pow(1.0 + p_eps, 1.0 / eps + 1.0);
  const double dphisrho = -2.0 * bb * rho2 * rho / ts05_sq(rho2 + ts05_sq(rho)) * sphic + beta * ps * p_em1 * rho / den;
  const double dphisdy = beta * ps * p_em1 * ysc / den;
  double sphics, cphics;
  srt_sincos(phis, sphics, cphics);
  const double xs = rho * cphics, zs = -rho * sphics;
  const double* a = numb == 1 ? (mode == 1 ? ts05_birk_1n2_a11() : ts05_birk_1n2_a12()) : (mode == 1 ? ts05_birk_1n2_a21() : ts05_birk_1n2_a22());
  double bn[3], bs[3];
  ts05_one_cone(a, mode, dtheta, xs, ysc, zs, bn[0], bn[1], bn[2]);
  ts05_one_cone(a, mode, dtheta, xs, -ysc, -zs, bs[0], bs[1], bs[2]);
  const double bxs = bn[0] - bs[0], byas = bn[1] + bs[1], bzs = bn[2] + bs[2];
  const double brhoas = bxs * cphics - bzs * sphics;
  const double bphias = -bxs * sphics - bzs * cphics;
  const double brho_s = brhoas * dphisphi * xkappa;
  const double bphi_s = (bphias - rho * (byas * dphisdy + brhoas * dphisrho)) * xkappa;
  const double by_s = byas * dphisphi * xkappa;
  b[0] = brho_s * cphic - bphi_s * sphic;
  b[1] = by_s;
  b[2] = -brho_s * sphic - bphi_s * cphic;
}

// BIRK_SHL (:1532-1667) and RC_SHIELD (:2376-2512): the same 2 x 3 x 3 x 2 x 2 harmonic sum over an 86-word table;
// RC_SHIELD multiplies every harmonic by FAC_SC = (X_SC+1)**3, BIRK_SHL does not (fac_sc = 1 changes no bit)
SRT_HDN void ts05_shield86(const double* a, bool rc, double ps, double x_sc, double x, double y, double z, double& bx, double& by, double& bz) {
  const double xs1 = x_sc + 1.0;
  const double fac_sc = rc ? xs1 * xs1 * xs1 : 1.0;
  double sps, cps, st1, ct1, st2, ct2;
  srt_sincos(ps, sps, cps);
  const double s3ps = 2.0 * cps;
  srt_sincos(ps * a[84], st1, ct1);
  srt_sincos(ps * a[85], st2, ct2);
  const double x1 = x * ct1 - z * st1, z1 = x * st1 + z * ct1;
  const double x2 = x * ct2 - z * st2, z2 = x * st2 + z * ct2;
  int l = 0;
  double gx = 0.0, gy = 0.0, gz = 0.0;
  for (int m = 1; m <= 2; ++m) {
    for (int i = 0; i < 3; ++i) {
      const double p = a[72 + i], q = a[78 + i];
      double sy, cy;
      srt_sincos(m == 1 ? y / p : y / q, sy, cy);
      for (int k = 0; k < 3; ++k) {
        const double r = a[75 + k], s = a[81 + k];
        double fx, fy, fz;
        if (m == 1) {
          double szrk, czrk;
          srt_sincos(z1 / r, szrk, czrk);
          const double sqpr = srt_sqrt(1.0 / ts05_sq(p) + 1.0 / ts05_sq(r));
          const double epr = srt_exp(x1 * sqpr);
          fx = -sqpr * epr * cy * szrk * fac_sc;
          fy = epr * sy * szrk / p * fac_sc;
          fz = -epr * cy * czrk / r * fac_sc;
        } else {
          double szsk, czsk;
          srt_sincos(z2 / s, szsk, czsk);
          const double sqqs = srt_sqrt(1.0 / ts05_sq(q) + 1.0 / ts05_sq(s));
          const double eqs = srt_exp(x2 * sqqs);
          fx = -sps * sqqs * eqs * cy * czsk * fac_sc;
          fy = sps / q * eqs * sy * czsk * fac_sc;
          fz = sps / s * eqs * cy * szsk * fac_sc;
        }
        const double tilt = m == 1 ? cps : s3ps;
        for (int n = 1; n <= 2; ++n) {
          for (int nn = 1; nn <= 2; ++nn) {
            double hx = fx, hy = fy, hz = fz;
            if (n == 2) { hx = hx * tilt; hy = hy * tilt; hz = hz * tilt; }
            if (nn == 2) { hx = hx * x_sc; hy = hy * x_sc; hz = hz * x_sc; }
            double hxr, hzr;
            if (m == 1) { hxr = hx * ct1 + hz * st1; hzr = -hx * st1 + hz * ct1; }
            else { hxr = hx * ct2 + hz * st2; hzr = -hx * st2 + hz * ct2; }
            gx = gx + hxr * a[l];
            gy = gy + hy * a[l];
            gz = gz + hzr * a[l];
            ++l;
          }
        }
      }
    }
  }
  bx = gx; by = gy; bz = gz;
}

// ------------------------------------------------------------------ ring current
// the elliptic-integral vector potential of one current loop at (rhos, zs): the block repeated four times in AP and APPRC
// (:1953-1993, :2128-2163).  Three polynomial coefficients are REAL*4 literals in the reference.
SRT_HD double ts05_aphi(double rrc, double dd, double rhos, double zs) {
  const double p = ts05_sq(rrc + rhos) + ts05_sq(zs) + ts05_sq(dd);
  const double xk2 = 4.0 * rrc * rhos / p;
  const double xk = srt_sqrt(xk2);
  const double xkrho12 = xk * srt_sqrt(rhos);
  const double xk2s = 1.0 - xk2;
  const double dl = srt_log(1.0 / xk2s);
  const double elk = 1.38629436112 + xk2s * (0.09666344259 + xk2s * (0.03590092383f + xk2s * (0.03742563713f + xk2s * 0.01451196212f))) +
                     dl * (0.5 + xk2s * (0.12498593597 + xk2s * (0.06880248576 + xk2s * (0.03328355346 + xk2s * 0.00441787012))));
  const double ele = 1.0 + xk2s * (0.44325141463 + xk2s * (0.0626060122 + xk2s * (0.04757383546 + xk2s * 0.01736506451))) +
                     dl * xk2s * (0.2499836831 + xk2s * (0.09200180037 + xk2s * (0.04069697526 + xk2s * 0.00526449639)));
  return ((1.0 - xk2 * 0.5) * elk - ele) / xkrho12;
}

// (alpha_s, gamma_s) -> (rhos, zs): :1938-1951, :2113-2126
SRT_HD void ts05_invert(double alpha_s, double gamma_s, double& rhos, double& zs) {
  const double gammas2 = ts05_sq(gamma_s);
  const double alsqh = ts05_sq(alpha_s) / 2.0;
  const double f = 64.0 / 27.0 * gammas2 + ts05_sq(alsqh);
  const double third = 1.0 / 3.0;
  const double q = srt_pow(srt_sqrt(f) + alsqh, third);
  double c = q - 4.0 * srt_pow(gammas2, third) / (3.0 * q);
  if (c < 0.0) c = 0.0;
  const double g = srt_sqrt(ts05_sq(c) + 4.0 * srt_pow(gammas2, third));
  const double rs = 4.0 / ((srt_sqrt(2.0 * g - c) + srt_sqrt(c)) * (g + c));
  const double costs = gamma_s * ts05_sq(rs);
  const double sints = srt_sqrt(1.0 - ts05_sq(costs));
  rhos = rs * sints;
  zs = rs * costs;
}

SRT_HD double ts05_exp500(double arg) { return arg < -500.0 ? 0.0 : srt_exp(arg); }

// AP (:1891-2006): vector potential of the symmetric ring current
SRT_HDN double ts05_ap(double r, double sint, double cost) {
  bool prox = false;
  double sint1 = sint, cost1 = cost;
  if (sint1 < 1.e-2) { sint1 = 1.e-2; cost1 = .99994999875f; prox = true; }
  const double alpha = ts05_sq(sint1) / r;
  const double gamma = cost1 / ts05_sq(r);
  const double arg1 = -ts05_sq((r - TS05_AP_R1) / TS05_AP_DR1) - ts05_sq(cost1 / TS05_AP_DLA1);
  const double arg2 = -ts05_sq((r - TS05_AP_R2) / TS05_AP_DR2) - ts05_sq(cost1 / TS05_AP_DLA2);
  const double arg3 = -ts05_sq((r - TS05_AP_R3) / TS05_AP_DR3);
  const double dexp1 = ts05_exp500(arg1), dexp2 = ts05_exp500(arg2), dexp3 = ts05_exp500(arg3);
  const double alpha_s = alpha * (1.0 + TS05_AP_P1 * dexp1 + TS05_AP_P2 * dexp2 + TS05_AP_P3 * dexp3);
  double rhos, zs;
  ts05_invert(alpha_s, gamma, rhos, zs);
  const double aphi1 = ts05_aphi(TS05_AP_RRC1, TS05_AP_DD1, rhos, zs);
  const double aphi2 = ts05_aphi(TS05_AP_RRC2, TS05_AP_DD2, rhos, zs);
  double ap = TS05_AP_A1 * aphi1 + TS05_AP_A2 * aphi2;
  if (prox) ap = ap * sint / sint1;
  return ap;
}

// APPRC (:2054-2171): vector potential of the symmetric part of the partial ring current
SRT_HDN double ts05_apprc(double r, double sint, double cost) {
  bool prox = false;
  double sint1 = sint, cost1 = cost;
  if (sint1 < 1.e-2) { sint1 = 1.e-2; cost1 = .99994999875f; prox = true; }
  const double alpha = ts05_sq(sint1) / r;
  const double gamma = cost1 / ts05_sq(r);
  const double arg1 = -ts05_sq(gamma / TS05_APPRC_DG1);
  const double arg2 = -ts05_sq((alpha - TS05_APPRC_ALPHA4) / TS05_APPRC_DAL4) - ts05_sq(gamma / TS05_APPRC_DG4);
  const double dexp1 = ts05_exp500(arg1), dexp2 = ts05_exp500(arg2);
  const double alpha_s =
      alpha * (1.0 + TS05_APPRC_P1 / srt_pow(1.0 + ts05_sq((alpha - TS05_APPRC_ALPHA1) / TS05_APPRC_DAL1), TS05_APPRC_BETA1) * dexp1 +
               TS05_APPRC_P2 * (alpha - TS05_APPRC_ALPHA2) / srt_pow(1.0 + ts05_sq((alpha - TS05_APPRC_ALPHA2) / TS05_APPRC_DAL2), TS05_APPRC_BETA2) /
                   srt_pow(1.0 + ts05_sq(gamma / TS05_APPRC_DG2), TS05_APPRC_BETA3) +
               TS05_APPRC_P3 * ts05_sq(alpha - TS05_APPRC_ALPHA3) / srt_pow(1.0 + ts05_sq((alpha - TS05_APPRC_ALPHA3) / TS05_APPRC_DAL3), TS05_APPRC_BETA4) /
                   srt_pow(1.0 + ts05_sq(gamma / TS05_APPRC_DG3), TS05_APPRC_BETA5));
  const double gamma_s =
      gamma * (1.0 + TS05_APPRC_Q0 + TS05_APPRC_Q1 * (alpha - TS05_APPRC_ALPHA4) * dexp2 +
               TS05_APPRC_Q2 * (alpha - TS05_APPRC_ALPHA5) / srt_pow(1.0 + ts05_sq((alpha - TS05_APPRC_ALPHA5) / TS05_APPRC_DAL5), TS05_APPRC_BETA6) /
                   srt_pow(1.0 + ts05_sq(gamma / TS05_APPRC_DG5), TS05_APPRC_BETA7));
  double rhos, zs;
  ts05_invert(alpha_s, gamma_s, rhos, zs);
  const double aphi1 = ts05_aphi(TS05_APPRC_RRC1, TS05_APPRC_DD1, rhos, zs);
  const double aphi2 = ts05_aphi(TS05_APPRC_RRC2, TS05_APPRC_DD2, rhos, zs);
  double ap = TS05_APPRC_A1 * aphi1 + TS05_APPRC_A2 * aphi2;
  if (prox) ap = ap * sint / sint1;
  return ap;
}

// RC_SYMM (:1846-1889) / PRC_SYMM (:2008-2052): B = curl of (A_phi e_phi), derivatives by central differences (D = 1e-4)
template <int PRC>
SRT_HDN void ts05_symm(double x, double y, double z, double& bx, double& by, double& bz) {
  const double ds = TS05_RC_SYMM_DS, dc = TS05_RC_SYMM_DC, d = TS05_RC_SYMM_D, drd = TS05_RC_SYMM_DRD;
#define TS05_POT(r_, s_, c_) (PRC ? ts05_apprc((r_), (s_), (c_)) : ts05_ap((r_), (s_), (c_)))
  const double rho2 = ts05_sq(x) + ts05_sq(y);
  const double r2 = rho2 + ts05_sq(z);
  const double r = srt_sqrt(r2);
  const double rp = r + d, rm = r - d;
  const double sint = srt_sqrt(rho2) / r, cost = z / r;
  if (sint < ds) {
    const double a = TS05_POT(r, ds, dc) / ds;
    const double dardr = (rp * TS05_POT(rp, ds, dc) - rm * TS05_POT(rm, ds, dc)) * drd;
    const double fxy = z * (2.0 * a - dardr) / (r * r2);
    bx = fxy * x;
    by = fxy * y;
    bz = (2.0 * a * ts05_sq(cost) + dardr * ts05_sq(sint)) / r;
  } else {
    const double theta = srt_atan2(sint, cost);
    double sintp, costp, sintm, costm;
    srt_sincos(theta + d, sintp, costp);
    srt_sincos(theta - d, sintm, costm);
    const double br = (sintp * TS05_POT(r, sintp, costp) - sintm * TS05_POT(r, sintm, costm)) / (r * sint) * drd;
    const double bt = (rm * TS05_POT(rm, sint, cost) - rp * TS05_POT(rp, sint, cost)) / r * drd;
    const double fxy = (br + bt * cost / sint) / r;
    bx = fxy * x;
    by = fxy * y;
    bz = br * cost - bt * sint;
  }
#undef TS05_POT
}

// FFS (:2361-2374)
SRT_HD void ts05_ffs(double a, double a0, double da, double& f, double& fa, double& fs) {
  const double sq1 = srt_sqrt(ts05_sq(a + a0) + ts05_sq(da));
  const double sq2 = srt_sqrt(ts05_sq(a - a0) + ts05_sq(da));
  fa = 2.0 / (sq1 + sq2);
  f = fa * a;
  fs = 0.5 * (sq1 + sq2) / (sq1 * sq2) * (1.0 - f * f);
}
SRT_HD double ts05_p4(double x) { const double x2 = x * x; return x2 * x2; }

// BR_PRC_Q (:2230-2298): radial component of the quadrupole part of the partial ring current
SRT_HDN double ts05_br_prc_q(double r, double sint, double cost) {
#define Q(n) TS05_BR_PRC_Q_##n
  const double sint2 = ts05_sq(sint), cost2 = ts05_sq(cost), sc = sint * cost;
  const double alpha = sint2 / r, gamma = cost / ts05_sq(r);
  double f, fa, fs;
  ts05_ffs(alpha, Q(AL1), Q(DAL1), f, fa, fs);
  const double d1 = sc * srt_pow(f, Q(XK1)) / (srt_pow(r / Q(B1), Q(BE1)) + 1.0);
  const double d2 = d1 * cost2;
  ts05_ffs(alpha, Q(AL2), Q(DAL2), f, fa, fs);
  const double d3 = sc * srt_pow(fs, Q(XK2)) / (srt_pow(r / Q(B2), Q(BE2)) + 1.0);
  const double d4 = d3 * cost2;
  ts05_ffs(alpha, Q(AL3), Q(DAL3), f, fa, fs);
  const double d5 = sc * srt_pow(alpha, Q(XK3)) * srt_pow(fs, Q(XK4)) / (srt_pow(r / Q(B3), Q(BE3)) + 1.0);
  const double d6 = d5 * cost2;
  double arga = ts05_sq((alpha - Q(AL4)) / Q(DAL4)) + 1.0;
  double argg = 1.0 + ts05_sq(gamma / Q(DG1));
  const double d7 = sc / arga / argg, d8 = d7 / arga, d9 = d8 / arga, d10 = d9 / arga;
  arga = ts05_sq((alpha - Q(AL5)) / Q(DAL5)) + 1.0;
  argg = 1.0 + ts05_sq(gamma / Q(DG2));
  const double d11 = sc / arga / argg, d12 = d11 / arga, d13 = d12 / arga, d14 = d13 / arga;
  const double r4 = ts05_p4(r);
  const double d15 = sc / (r4 + ts05_p4(Q(C1)));
  const double d16 = sc / (r4 + ts05_p4(Q(C2))) * cost2;
  const double d17 = sc / (r4 + ts05_p4(Q(C3))) * ts05_sq(cost2);
  ts05_ffs(alpha, Q(AL6), Q(DAL6), f, fa, fs);
  const double d18 = sc * fs / (1.0 + ts05_sq((r - 1.2) / Q(DRM)));
  return Q(A1) * d1 + Q(A2) * d2 + Q(A3) * d3 + Q(A4) * d4 + Q(A5) * d5 + Q(A6) * d6 + Q(A7) * d7 + Q(A8) * d8 + Q(A9) * d9 + Q(A10) * d10 +
         Q(A11) * d11 + Q(A12) * d12 + Q(A13) * d13 + Q(A14) * d14 + Q(A15) * d15 + Q(A16) * d16 + Q(A17) * d17 + Q(A18) * d18;
#undef Q
}

// BT_PRC_Q (:2300-2359): theta component
SRT_HDN double ts05_bt_prc_q(double r, double sint, double cost) {
#define Q(n) TS05_BT_PRC_Q_##n
  const double cost2 = ts05_sq(cost);
  const double alpha = ts05_sq(sint) / r, gamma = cost / ts05_sq(r);
  double f, fa, fs;
  ts05_ffs(alpha, Q(AL1), Q(DAL1), f, fa, fs);
  const double d1 = srt_pow(f, Q(XK1)) / (srt_pow(r / Q(B1), Q(BE1)) + 1.0);
  const double d2 = d1 * cost2;
  ts05_ffs(alpha, Q(AL2), Q(DAL2), f, fa, fs);
  const double d3 = srt_pow(fa, Q(XK2)) / srt_pow(r, Q(BE2));
  const double d4 = d3 * cost2;
  ts05_ffs(alpha, Q(AL3), Q(DAL3), f, fa, fs);
  const double d5 = srt_pow(fs, Q(XK3)) * srt_pow(alpha, Q(XK4)) / (srt_pow(r / Q(B3), Q(BE3)) + 1.0);
  const double d6 = d5 * cost2;
  ts05_ffs(gamma, 0.0, Q(DG1), f, fa, fs);
  const double fcc = 1.0 + ts05_sq((alpha - Q(AL4)) / Q(DAL4));
  const double d7 = 1.0 / fcc * fs, d8 = d7 / fcc, d9 = d8 / fcc, d10 = d9 / fcc;
  const double arg = 1.0 + ts05_sq((alpha - Q(AL5)) / Q(DAL5));
  const double d11 = 1.0 / arg / (1.0 + ts05_sq(gamma / Q(DG2))), d12 = d11 / arg, d13 = d12 / arg, d14 = d13 / arg;
  const double r4 = ts05_p4(r);
  const double d15 = 1.0 / (r4 + ts05_sq(Q(C1)));
  const double d16 = cost2 / (r4 + ts05_sq(Q(C2)));
  const double d17 = ts05_sq(cost2) / (r4 + ts05_sq(Q(C3)));
  return Q(A1) * d1 + Q(A2) * d2 + Q(A3) * d3 + Q(A4) * d4 + Q(A5) * d5 + Q(A6) * d6 + Q(A7) * d7 + Q(A8) * d8 + Q(A9) * d9 + Q(A10) * d10 +
         Q(A11) * d11 + Q(A12) * d12 + Q(A13) * d13 + Q(A14) * d14 + Q(A15) * d15 + Q(A16) * d16 + Q(A17) * d17;
#undef Q
}

// PRC_QUAD (:2173-2228)
SRT_HDN void ts05_prc_quad(double x, double y, double z, double& bx, double& by, double& bz) {
  const double d = TS05_PRC_QUAD_D, dd = TS05_PRC_QUAD_DD, ds = TS05_PRC_QUAD_DS, dc = TS05_PRC_QUAD_DC;
  const double rho2 = ts05_sq(x) + ts05_sq(y);
  const double r = srt_sqrt(rho2 + ts05_sq(z));
  const double rho = srt_sqrt(rho2);
  const double sint = rho / r, cost = z / r;
  const double rp = r + d, rm = r - d;
  if (sint > ds) {
    const double cphi = x / rho, sphi = y / rho;
    const double br = ts05_br_prc_q(r, sint, cost), bt = ts05_bt_prc_q(r, sint, cost);
    const double dbrr = (ts05_br_prc_q(rp, sint, cost) - ts05_br_prc_q(rm, sint, cost)) / dd;
    const double theta = srt_atan2(sint, cost);
    double sintp, costp, sintm, costm;
    srt_sincos(theta + d, sintp, costp);
    srt_sincos(theta - d, sintm, costm);
    const double dbtt = (ts05_bt_prc_q(r, sintp, costp) - ts05_bt_prc_q(r, sintm, costm)) / dd;
    bx = sint * (br + (br + r * dbrr + dbtt) * ts05_sq(sphi)) + cost * bt;
    by = -sint * sphi * cphi * (br + r * dbrr + dbtt);
    bz = (br * cost - bt * sint) * cphi;
  } else {
    const double st = ds;
    const double ct = z < 0.0 ? -dc : dc;
    const double theta = srt_atan2(st, ct);
    double sintp, costp, sintm, costm;
    srt_sincos(theta + d, sintp, costp);
    srt_sincos(theta - d, sintm, costm);
    const double br = ts05_br_prc_q(r, st, ct), bt = ts05_bt_prc_q(r, st, ct);
    const double dbrr = (ts05_br_prc_q(rp, st, ct) - ts05_br_prc_q(rm, st, ct)) / dd;
    const double dbtt = (ts05_bt_prc_q(r, sintp, costp) - ts05_bt_prc_q(r, sintm, costm)) / dd;
    const double fcxy = r * dbrr + dbtt;
    bx = (br * (ts05_sq(x) + 2.0 * ts05_sq(y)) + fcxy * ts05_sq(y)) / ts05_sq(r * st) + bt * cost;
    by = -(br + fcxy) * x * y / ts05_sq(r * st);
    bz = (br * cost / st - bt) * x / r;
  }
}

// SRC_PRC (:1762-1844) + FULL_RC (:1669-1760): shielded symmetric and partial ring currents (IOPR = 0)
SRT_HDN void ts05_full_rc(double sc_sy, double sc_pr, double phi, double ps, double x, double y, double z, double src[3], double prc[3]) {
  double sps, cps;
  srt_sincos(ps, sps, cps);
  const double xt = x * cps - z * sps, zt = z * cps + x * sps;
  const double xts = xt / sc_sy, yts = y / sc_sy, zts = zt / sc_sy;
  const double xta = xt / sc_pr, yta = y / sc_pr, zta = zt / sc_pr;
  double bxs, bys, bzs, bxa_s, bya_s, bza_s, bxa_qr, bya_qr, bza_q;
  ts05_symm<0>(xts, yts, zts, bxs, bys, bzs);
  ts05_symm<1>(xta, yta, zta, bxa_s, bya_s, bza_s);
  double sp, cp;
  srt_sincos(phi, sp, cp);
  const double xr = xta * cp - yta * sp, yr = xta * sp + yta * cp;
  ts05_prc_quad(xr, yr, zta, bxa_qr, bya_qr, bza_q);
  const double bxa_q = bxa_qr * cp + bya_qr * sp, bya_q = -bxa_qr * sp + bya_qr * cp;
  const double bxp = bxa_s + bxa_q, byp = bya_s + bya_q, bzp = bza_s + bza_q;
  const double hs[3] = {bxs * cps + bzs * sps, bys, bzs * cps - bxs * sps};
  const double hp[3] = {bxp * cps + bzp * sps, byp, bzp * cps - bxp * sps};
  double fs[3], fp[3];
  ts05_shield86(ts05_full_rc_c_sy(), true, ps, sc_sy - 1.0, x, y, z, fs[0], fs[1], fs[2]);
  ts05_shield86(ts05_full_rc_c_pr(), true, ps, sc_pr - 1.0, x, y, z, fp[0], fp[1], fp[2]);
  for (int i = 0; i < 3; ++i) { src[i] = hs[i] + fs[i]; prc[i] = hp[i] + fp[i]; }
}

// DIPOLE (:2514-2543)
SRT_HD void ts05_dipole(double ps, double x, double y, double z, double q3[3]) {
  double sps, cps;
  srt_sincos(ps, sps, cps);
  const double p = ts05_sq(x), u = ts05_sq(z), v = 3.0 * z * x, t = ts05_sq(y);
  const double s = srt_sqrt(p + t + u);
  const double s2 = s * s;
  const double q = 30115.0 / (s2 * (s2 * s));  // **5 as gcc's power tree: x^2 * x^3
  q3[0] = q * ((t + u - 2.0 * p) * sps - v * cps);
  q3[1] = -3.0 * y * q * (x * sps + z * cps);
  q3[2] = q * ((p + t - 2.0 * u) * cps - v * sps);
}

// the module fields of EXTERN before their amplitudes, for the tests: total, CF, T1, T2, SRC, PRC, R11, R21
struct Ts05Parts {
  double b[3], cf[3], t1[3], t2[3], src[3], prc[3], r11[3], r21[3];
};

// EXTERN (:118-360) with IOPGEN = IOPT = IOPB = IOPR = 0.  DST is T04_s's DST_AST.
SRT_HDN void ts05_extern(double pdyn, double dst, double byimf, double bzimf, const double w[6], double ps, double x, double y, double z, Ts05Parts& o) {
  const double* a = ts05_t04_s_a();
  const double a0_a = TS05_EXTERN_A0_A, a0_s0 = TS05_EXTERN_A0_S0, a0_x0 = TS05_EXTERN_A0_X0, dsig = TS05_EXTERN_DSIG, rh2 = TS05_EXTERN_RH2;
  const double xappa = srt_pow(pdyn / 2.f, a[22]);
  const double rh0 = 7.5f, g = 35.0f;
  const double xappa3 = xappa * xappa * xappa;
  const double xx = x * xappa, yy = y * xappa, zz = z * xappa;
  const double sps = srt_sin(ps);
  const double x0 = a0_x0 / xappa, am = a0_a / xappa, s0 = a0_s0;
  const double factimf = a[19];
  const double oimfx = 0.0, oimfy = byimf * factimf, oimfz = bzimf * factimf;
  const double r = srt_sqrt(ts05_sq(x) + ts05_sq(y) + ts05_sq(z));
  double xss = x, zss = z, dd;
  do {  // iterative search of the unwarped coordinates
    const double xsold = xss, zsold = zss;
    const double rh = rh0 + rh2 * ts05_sq(zss / r);
    const double rrh = r / rh;
    const double sinpsas = sps / srt_pow(1.0 + rrh * rrh * rrh, 0.33333333);
    const double cospsas = srt_sqrt(1.0 - ts05_sq(sinpsas));
    zss = x * sinpsas + z * cospsas;
    xss = x * cospsas - z * sinpsas;
    dd = srt_fabs(xss - xsold) + srt_fabs(zss - zsold);
  } while (dd > 1.e-6);
  const double rho2 = ts05_sq(y) + ts05_sq(zss);
  const double asq = ts05_sq(am);
  double xmxm = am + xss - x0;
  if (xmxm < 0.f) xmxm = 0.f;
  const double axx0 = ts05_sq(xmxm);
  const double aro = asq + rho2;
  const double sigma = srt_sqrt((aro + axx0 + srt_sqrt(ts05_sq(aro + axx0) - 4.f * asq * axx0)) / (2.f * asq));
  if (sigma < s0 + dsig) {
    double cf[3];
    ts05_shlcar3x3(xx, yy, zz, ps, cf[0], cf[1], cf[2]);
    for (int i = 0; i < 3; ++i) o.cf[i] = cf[i] * xappa3;
    // tail
    double dstt = -20.f;
    if (dst < dstt) dstt = dst;
    double znam = srt_pow(srt_fabs(dstt), (double)0.37f);
    Ts05Tail T;
    T.dxshift1 = a[23] - a[24] / znam;
    T.dxshift2 = a[25] - a[26] / znam;
    T.d = a[35] * srt_exp(-w[0] / a[36]) + a[68];
    T.deltady = 4.7f;
    ts05_deformed(T, rh0, g, ps, xx, yy, zz, o.t1, o.t2);
    // Birkeland
    znam = srt_fabs(dst);
    if (dst >= -20.0) znam = 20.0;
    const double xkappa1 = a[31] * srt_pow(znam / 20.0, a[32]);
    const double xkappa2 = a[33] * srt_pow(znam / 20.0, a[34]);
    double f[3], h[3];
    ts05_birk_1n2(1, 1, xkappa1, ps, xx, yy, zz, f);
    ts05_shield86(ts05_birk_tot_sh11(), false, ps, xkappa1 - 1.1, xx, yy, zz, h[0], h[1], h[2]);
    for (int i = 0; i < 3; ++i) o.r11[i] = f[i] + h[i];
    ts05_birk_1n2(2, 1, xkappa2, ps, xx, yy, zz, f);
    ts05_shield86(ts05_birk_tot_sh21(), false, ps, xkappa2 - 1.0, xx, yy, zz, h[0], h[1], h[2]);
    for (int i = 0; i < 3; ++i) o.r21[i] = f[i] + h[i];
    // ring current
    const double phi = a[37];
    const double sc_sy = a[27] * srt_pow(20.0 / znam, a[28]) * xappa;
    const double sc_as = a[29] * srt_pow(20.0 / znam, a[30]) * xappa;
    ts05_full_rc(sc_sy, sc_as, phi, ps, xx, yy, zz, o.src, o.prc);
    const double himf[3] = {0.0, byimf, bzimf};
    const double dlp1 = srt_pow(pdyn / 2.0, a[20]), dlp2 = srt_pow(pdyn / 2.0, a[21]);
    const double tamp1 = a[1] + a[2] * dlp1 + a[3] * a[38] * w[0] / srt_sqrt(ts05_sq(w[0]) + ts05_sq(a[38])) + a[4] * dst;
    const double tamp2 = a[5] + a[6] * dlp2 + a[7] * a[39] * w[1] / srt_sqrt(ts05_sq(w[1]) + ts05_sq(a[39])) + a[8] * dst;
    const double a_src = a[9] + a[10] * a[40] * w[2] / srt_sqrt(ts05_sq(w[2]) + ts05_sq(a[40])) + a[11] * dst;
    const double a_prc = a[12] + a[13] * a[41] * w[3] / srt_sqrt(ts05_sq(w[3]) + ts05_sq(a[41])) + a[14] * dst;
    const double a_r11 = a[15] + a[16] * a[42] * w[4] / srt_sqrt(ts05_sq(w[4]) + ts05_sq(a[42]));
    const double a_r21 = a[17] + a[18] * a[43] * w[5] / srt_sqrt(ts05_sq(w[5]) + ts05_sq(a[43]));
    double bb[3];
    for (int i = 0; i < 3; ++i)
      bb[i] = a[0] * o.cf[i] + tamp1 * o.t1[i] + tamp2 * o.t2[i] + a_src * o.src[i] + a_prc * o.prc[i] + a_r11 * o.r11[i] + a_r21 * o.r21[i] + a[19] * himf[i];
    if (sigma < s0 - dsig) {
      for (int i = 0; i < 3; ++i) o.b[i] = bb[i];
    } else {  // the boundary layer: interpolation between the inside and the outside field
      const double fint = 0.5f * (1.f - (sigma - s0) / dsig);
      const double fext = 0.5f * (1.f + (sigma - s0) / dsig);
      double q[3];
      ts05_dipole(ps, x, y, z, q);
      const double oimf[3] = {oimfx, oimfy, oimfz};
      for (int i = 0; i < 3; ++i) o.b[i] = (bb[i] + q[i]) * fint + oimf[i] * fext - q[i];
    }
  } else {
    double q[3];
    ts05_dipole(ps, x, y, z, q);
    o.b[0] = oimfx - q[0];
    o.b[1] = oimfy - q[1];
    o.b[2] = oimfz - q[2];
  }
}

// T04_s (:5-116): REAL*4 in and out, REAL*8 inside
SRT_HDN void srt_ts05(const SrtTs05& m, float x, float y, float z, float& bx, float& by, float& bz) {
  const float* parmod = m.parmod;
  const double pdyn = parmod[0];
  const double dst_ast = parmod[1] * 0.8f - 13.f * srt_sqrt(pdyn);
  const double w[6] = {parmod[4], parmod[5], parmod[6], parmod[7], parmod[8], parmod[9]};
  Ts05Parts o;
  for (int i = 0; i < 3; ++i) o.cf[i] = o.t1[i] = o.t2[i] = o.src[i] = o.prc[i] = o.r11[i] = o.r21[i] = 0.0;
  ts05_extern(pdyn, dst_ast, (double)parmod[2], (double)parmod[3], w, (double)m.ps, (double)x, (double)y, (double)z, o);
  bx = (float)o.b[0];
  by = (float)o.b[1];
  bz = (float)o.b[2];
}

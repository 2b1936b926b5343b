// oracle/orc_models.h -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the two
// funcPlasmaParams adapters on the hot path and of what they call:
//   fortran/ngo_dens_model.f95 (readinput, dens), fortran/ngo_dens_model_adapter.f95,
//   fortran/bmodel_dipole.f95, xform_double/{SM_TO_GSM,GSM_TO_SM,T4,Get_q_c,T0,T1,T2,
//   POL_TO_CART,Rotate_*}.f95.
// The gridded adapter and IGRF live in orc_grid.h / orc_igrf.h.  PARITY UNPINNED.
#pragma once
#include "orc_common.h"
#include "orc_raytracer.h"

namespace orc {

// ---------------------------------------------------------------- xform_double
inline void rotate_x_d(double angle, const double a[3], double b[3]) {  // Rotate_x.f95:7-9
  b[0] = a[0];
  b[1] = a[1] * m_cos(angle) + a[2] * m_sin(angle);
  b[2] = a[2] * m_cos(angle) - a[1] * m_sin(angle);
}
inline void rotate_y_d(double angle, const double a[3], double b[3]) {  // Rotate_y.f95:7-9
  b[0] = a[0] * m_cos(angle) + a[2] * m_sin(angle);
  b[1] = a[1];
  b[2] = a[2] * m_cos(angle) - a[0] * m_sin(angle);
}
inline void rotate_z_d(double angle, const double a[3], double b[3]) {  // Rotate_z.f95:7-9
  b[0] = a[0] * m_cos(angle) + a[1] * m_sin(angle);
  b[1] = a[1] * m_cos(angle) - a[0] * m_sin(angle);
  b[2] = a[2];
}
static const double DEGRAD = 3.141592653589793238462643 / 180.0;
inline double t0_d(const int itime[2], double& ut) {  // T0.f95:7-24
  int iyr = itime[0] / 1000;
  int iday = itime[0] - iyr * 1000;
  ut = itime[1] / 3600000.0;
  double fracday = ut / 24.0;
  double rmjd = 45.0 + (double)(float)(iyr - 1859) * 365.0 + ((double)(float)((iyr - 1861) / 4) + 1.0) + (double)(float)(iday) - 1.0 + fracday;
  return (rmjd - 51544.5) / 36525.0;
}
inline void t1_d(const int itime[2], const double a[3], double b[3], int iverse) {  // T1.f95:7-19
  double ut;
  double temp = t0_d(itime, ut);
  double theta = (100.461 + 36000.770 * temp + 15.04107 * ut) * DEGRAD;
  rotate_z_d((double)(float)iverse * theta, a, b);
}
inline void t2_d(const int itime[2], const double a[3], double b[3], int iverse) {  // T2.f95:7-29
  double ut, temp[3];
  double tt0 = t0_d(itime, ut);
  double epsilon = (23.439 - 0.013 * tt0) * DEGRAD;
  double mm = (357.528 + 35999.05 * tt0 + 0.04107 * ut) * DEGRAD;
  double cgamma = 280.46 + 36000.772 * tt0 + 0.04107 * ut;
  double lamdas = (cgamma + (1.915 - 0.0048 * tt0) * m_sin(mm) + 0.02 * m_sin(2.0 * mm)) * DEGRAD;
  if (iverse == 1) {
    rotate_x_d(epsilon, a, temp);
    rotate_z_d(lamdas, temp, b);
  } else {
    rotate_z_d(iverse * lamdas, a, temp);
    rotate_x_d(iverse * epsilon, temp, b);
  }
}
inline void pol_to_cart_d(double lat, double lon, double radial, double out[3]) {  // POL_TO_CART.f95
  double coslat = m_cos(lat);
  out[0] = radial * coslat * m_cos(lon);
  out[1] = radial * coslat * m_sin(lon);
  out[2] = radial * m_sin(lat);
}
inline void get_q_c_d(const int itime[2], double q_c[3]) {  // Get_q_c.f95:4-30
  int iyr = itime[0] / 1000;
  int iday = itime[0] - iyr * 1000;
  double ut = itime[1] / 3600000.0;
  double fracday = ut / 24.0;
  double rmjd = 45.0 + (double)(iyr - 1859) * 365.0 + ((double)((iyr - 1861) / 4) + 1.0) + (double)(iday) - 1.0 + fracday;
  double factor = (rmjd - 46066.0) / 365.25;
  double phi = (78.8 + 4.283e-2 * factor) * DEGRAD;
  double lamda = (289.1 - 1.413e-2 * factor) * DEGRAD;
  double q_g[3], temp[3];
  pol_to_cart_d(phi, lamda, 1.0, q_g);
  t1_d(itime, q_g, temp, -1);
  t2_d(itime, temp, q_c, 1);
}
inline double dipole_tilt_mu(const int itime[2]) {  // T4.f95:13-15
  double q_c[3];
  get_q_c_d(itime, q_c);
  return -m_atan(q_c[0] / m_sqrt(q_c[1] * q_c[1] + q_c[2] * q_c[2]));
}
inline void t4_d(const int itime[2], const double a[3], double b[3], int iverse) {
  double mu = dipole_tilt_mu(itime);
  rotate_y_d(iverse * mu, a, b);
}

// time-only transform, optionally cached (bit-identical: rotate_y_d recomputes
// cos/sin of the same angle on every call)
struct TiltXform {
  int itime[2];
  bool cache = true;
  double mu = 0;
  void init(int yearday, int msec) { itime[0] = yearday; itime[1] = msec; mu = dipole_tilt_mu(itime); }
  void sm_to_gsm(const double a[3], double b[3]) const {
    if (cache) rotate_y_d(-1 * mu, a, b); else t4_d(itime, a, b, -1);
  }
  void gsm_to_sm(const double a[3], double b[3]) const {
    if (cache) rotate_y_d(1 * mu, a, b); else t4_d(itime, a, b, 1);
  }
};

// ---------------------------------------------------------------- dipole
// bmodel_dipole.f95:20-48
inline void bmodel_cartesian(const double x[3], double B[3]) {
  double p[3];
  cartesian_to_spherical(x, p);
  double Bo = .312 / 10000.0;
  double R = p[0] / R_E;
  double Bor3 = Bo * (m_pow(R, -3.0));
  double Brad = -2.0 * Bor3 * m_cos(p[2]);
  double Btheta = -Bor3 * m_sin(p[2]);
  double v[3] = {Brad, 0.0, Btheta};
  spherical_to_cartesian_vec(v, p[1], p[2], B);
}

// field block shared by both adapters (ngo_…adapter.f95:142-202, interp_…:208-267);
// igrf callback is filled in by orc_igrf.h
struct FieldBlock {
  TiltXform tilt;
  int use_igrf = 0;
  void (*igrf_gsm)(void* ctx, float x, float y, float z, float& bx, float& by, float& bz) = nullptr;
  void* igrf_ctx = nullptr;
  void (*recalc_each_call)(void* ctx) = nullptr;  // faithful-cost mode: tsy_recalc on every call
  // use_tsyganenko = 1 (adapter :183-187): T04_s is the reference's own source, translated at build time into
  // oracle/_ref (f77_to_cpp.py); the tests hand its entry point in.  `psi` is what the adapter's `PSI` names: word 1 of
  // COMMON /GEOPACK1/, which in the compiled Geopack-2008 holds ST0 (Q25).
  void (*t04_s)(const float* parmod, float ps, float x, float y, float z, float* b) = nullptr;
  float parmod[10] = {0};
  float psi = 0.0f;
  void eval(const double x[3], double B0[3]) const {
    double x_gsm[3];
    tilt.sm_to_gsm(x, x_gsm);
    if (recalc_each_call) recalc_each_call(igrf_ctx);
    float bx, by, bz;
    if (use_igrf == 1) {
      igrf_gsm(igrf_ctx, (float)(x_gsm[0] / R_E), (float)(x_gsm[1] / R_E), (float)(x_gsm[2] / R_E), bx, by, bz);
    } else {
      double B0tmp[3], B0tmp2[3];
      bmodel_cartesian(x, B0tmp);
      tilt.sm_to_gsm(B0tmp, B0tmp2);
      bx = (float)(1.0e9 * B0tmp2[0]);
      by = (float)(1.0e9 * B0tmp2[1]);
      bz = (float)(1.0e9 * B0tmp2[2]);
    }
    float tx = 0.0f, ty = 0.0f, tz = 0.0f;  // use_tsyganenko = 0
    if (t04_s) {
      float b[3];
      t04_s(parmod, psi, (float)(x_gsm[0] / R_E), (float)(x_gsm[1] / R_E), (float)(x_gsm[2] / R_E), b);
      tx = b[0]; ty = b[1]; tz = b[2];
    }
    double Bt[3];
    Bt[0] = (double)(bx + tx) * 1.0e-9;
    Bt[1] = (double)(by + ty) * 1.0e-9;
    Bt[2] = (double)(bz + tz) * 1.0e-9;
    tilt.gsm_to_sm(Bt, B0);
  }
};

// ---------------------------------------------------------------- list-directed reader
// Enough of Fortran list-directed input for the config cards: items separated by
// blanks / commas / newlines; each READ starts on a new record and continues onto
// following records until its list is satisfied; `d` exponents accepted.
struct ListReader {
  std::vector<std::string> lines;
  size_t cur = 0;
  bool load(const char* path) {
    FILE* f = fopen(path, "r");
    if (!f) return false;
    std::string s;
    int c;
    while ((c = fgetc(f)) != EOF) {
      if (c == '\n') { lines.push_back(s); s.clear(); }
      else if (c != '\r') s.push_back((char)c);
    }
    if (!s.empty()) lines.push_back(s);
    fclose(f);
    return true;
  }
  // read n numeric items; returns false on end of file
  bool read(int n, double* out) {
    int got = 0;
    while (got < n) {
      if (cur >= lines.size()) return false;
      std::string s = lines[cur++];
      size_t i = 0;
      while (i < s.size() && got < n) {
        while (i < s.size() && (s[i] == ' ' || s[i] == '\t' || s[i] == ',')) ++i;
        if (i >= s.size()) break;
        if (s[i] == '/') { while (got < n) { ++got; } break; }
        size_t j = i;
        while (j < s.size() && s[j] != ' ' && s[j] != '\t' && s[j] != ',') ++j;
        std::string tok = s.substr(i, j - i);
        for (auto& ch : tok) if (ch == 'd' || ch == 'D') ch = 'e';
        out[got++] = strtod(tok.c_str(), nullptr);
        i = j;
      }
    }
    return true;
  }
};

// ---------------------------------------------------------------- Ngo density
// module ngo_dens_model: module-global state becomes a struct
struct NgoState {
  double pi, clight, consf0, r0, radgra, grarad, am;
  int intera, numres, nsuppr, num, kskip, mode, kount, kducts, ktape;
  double spelat, refalt, dsrrng, dsrlat, dsdens, egfeq, therm, hm, absb, relb;
  double rbase, ane0, alpha0[5], rzero, scbot, rstop, rdiv, hmin;
  double lk, expk, ddk, rconsn, scr;
  double l0[11], def[11], dd[11], rducln[11], rducun[11], rducls[11], rducus[11], sidedu[11];
  double hl2n[11], hl2s[11], hu2n[11], hu2s[11];
  double latmax, latmin;
  double z[6], latitu, ani[5], alpha[5], l, dlgndr[5], dlgndt[5];
  int init, kinit, lines;
  NgoState() { memset(this, 0, sizeof(*this)); }

  // ngo_dens_model.f95:165-353
  void dens() {
    double exnor[5], qi[5], sh[5];
    double delr = 0.0, arglr = 0.0, delroh = 0.0, frduct = 0.0, delnl = 0.0, anl = 0.0, fac, onedut, onedur;
    double cosz2 = m_cos(z[2]);
    double sinz2 = m_sin(z[2]);
    double cosz22 = cosz2 * cosz2; (void)cosz22;
    double sinz22 = sinz2 * sinz2;
    // init stays -1 forever => recomputed on every call (:180-186)
    double rb7370 = rbase / (double)7370.f;
    sh[2] = (double)1.150600f * therm * rb7370 * rb7370;
    sh[3] = sh[2] / 4.;
    sh[4] = sh[3] / 4.;
    for (int i = 2; i <= 4; ++i) alpha[i] = 0.0;
    double gph = rbase * (1.0 - rbase / z[1]);
    exnor[2] = m_exp(-gph / sh[2]);
    exnor[3] = exnor[2] * exnor[2] * exnor[2] * exnor[2];
    exnor[4] = exnor[3] * exnor[3] * exnor[3] * exnor[3];
    double q = 0.0, sumi = 0.0;
    for (int i = 2; i <= num; ++i) {
      qi[i] = alpha0[i] * exnor[i];
      q = q + qi[i];
      sumi = sumi + qi[i] / sh[i];
    }
    for (int i = 2; i <= num; ++i) alpha[i] = qi[i] / q;
    double anr = m_sqrt(q);
    double arg = (z[1] - rzero) / scbot;
    if (!(arg < 13.0)) arg = 13.0;
    double exarg = m_exp(-arg * arg);
    double anli = 1.0 - exarg;
    double dlnlid = arg * exarg * 2. / (scbot * anli);
    l = z[1] / (r0 * sinz22);
    double dlnldr = 0, dlnldt = 0;
    ani[1] = ane0 * anr;
    ani[1] = ani[1] * anli;
    double cotz2 = 0.0;
    if (kducts != 0) {
      cotz2 = cosz2 / sinz2;
      double deltal = l - lk;
      if (!(deltal < 0.0)) {
        double d2 = ddk * ddk;
        double argl = deltal * deltal / (d2 * 2.0);
        if (!(argl < 80.00)) argl = 80.00;
        double f = m_exp(-argl);
        double trm = m_pow((rconsn / z[1]), expk);
        double argr = (z[1] - rconsn) / scr;
        if (!(argr < 12.50)) argr = 12.5;
        double fr = m_exp(-argr * argr);
        double trmodl = trm + (1. - trm) * fr;
        double dtrmdr = -expk * trm * (1. - fr) / z[1] - (1. - trm) * fr * 2. * argr / scr;
        double anlk = f + trmodl * (1.0 - f);
        double factor = deltal * f * l * (1.0 - trmodl) / (d2 * anlk);
        dlnldr = dlnldr - factor / z[1] + (1. - f) * dtrmdr / anlk;
        dlnldt = dlnldt + 2.0 * factor * cotz2;
        ani[1] = ani[1] * anlk;
      }
      // label 410
      if (kducts != 1) {
        bool skip_to_250 = false;
        if (!(l0[2] > 0.0)) {
          // sinusoidal density perturbation (:243-291)
          kinit = 3;
          deltal = l + l0[2];
          if (!(deltal * sidedu[2] >= 0.0)) deltal = 0;
          double delk = -l0[2] - (lk + ddk) + dd[2] / 2;
          double critl = (lk + ddk) + m_fmod(delk, dd[2]);
          bool to990 = false;
          if (l <= critl) to990 = true;
          if (!to990) {
            double argl = 2.0 * pi * deltal / dd[2];
            delnl = (def[2] / 2.) * (1. + m_cos(argl));
            bool to950 = false, to970 = false, to960 = false;
            if (latitu <= 0 && z[1] <= rducus[2]) to950 = true;
            if (!to950 && latitu >= 0 && z[1] <= rducun[2]) to950 = true;
            if (!to950) {
              if (latitu >= 0) delr = z[1] - rducun[2];
              if (latitu <= 0) delr = z[1] - rducus[2];
              if (latitu <= 0) arglr = delr * delr / hu2s[2];
              if (latitu >= 0) arglr = delr * delr / hu2n[2];
              if (arglr >= 75.0) to990 = true;
              else {
                frduct = m_exp(-arglr);
                if (latitu <= 0) delroh = 2.0 * delr / hu2s[2];
                if (latitu >= 0) delroh = 2.0 * delr / hu2n[2];
                to960 = true;
              }
            } else {
              if (latitu <= 0 && z[1] >= rducls[2]) to970 = true;
              if (!to970 && latitu >= 0 && z[1] >= rducln[2]) to970 = true;
              if (!to970) {
                if (latitu >= 0) delr = z[1] - rducln[2];
                if (latitu <= 0) delr = z[1] - rducls[2];
                if (latitu <= 0) arglr = delr * delr / hl2s[2];
                if (latitu >= 0) arglr = delr * delr / hl2n[2];
                if (arglr >= 75.0) to990 = true;
                else {
                  frduct = m_exp(-arglr);
                  if (latitu >= 0) delroh = 2.0 * delr / hl2n[2];
                  if (latitu <= 0) delroh = 2.0 * delr / hl2s[2];
                  to960 = true;
                }
              }
            }
            if (!to990) {
              if (to960) {
                delnl = delnl * frduct;
                anl = 1.0 + delnl;
                fac = pi * def[2] * l * m_sin(argl) / (dd[2] * anl);
                onedut = fac * 2.0 * cotz2;
                onedur = -fac / z[1] - delnl * delroh / anl;
              } else {  // 970
                anl = 1.0 + delnl;
                fac = pi * def[2] * l * m_sin(argl) / (dd[2] * anl);
                onedut = fac * 2.0 * cotz2;
                onedur = -fac / z[1];
              }
              dlnldr = dlnldr + onedur;  // 980
              dlnldt = dlnldt + onedut;
              ani[1] = ani[1] * anl;
            }
          }
          // 990
          if (kducts == 2) skip_to_250 = true;
        }
        if (!skip_to_250) {
          // 1000: duct(s)
          for (int kduc = kinit; kduc <= kducts; ++kduc) {
            deltal = l - l0[kduc];
            if (!(deltal * sidedu[kduc] >= 0.0)) deltal = 0;
            double d2 = dd[kduc] * dd[kduc];
            double argl = deltal * deltal / (d2 * 2.0);
            if (argl > 80.0) continue;
            delnl = def[kduc] * m_exp(-argl);
            bool to1500 = false, to235 = false;
            if (latitu >= 0 && z[1] <= rducun[kduc]) to1500 = true;
            if (!to1500 && latitu <= 0 && z[1] <= rducus[kduc]) to1500 = true;
            if (!to1500) {
              if (latitu >= 0) delr = z[1] - rducun[kduc];
              if (latitu <= 0) delr = z[1] - rducus[kduc];
              if (latitu >= 0) arglr = delr * delr / hu2n[kduc];
              if (latitu <= 0) arglr = delr * delr / hu2s[kduc];
              if (arglr >= 75.0) continue;
              frduct = m_exp(-arglr);
              if (latitu >= 0) delroh = 2.0 * delr / hu2n[kduc];
              if (latitu <= 0) delroh = 2.0 * delr / hu2s[kduc];
            } else {
              if (latitu >= 0 && z[1] >= rducln[kduc]) to235 = true;
              if (!to235 && latitu <= 0 && z[1] >= rducls[kduc]) to235 = true;
              if (!to235) {
                if (latitu >= 0) delr = z[1] - rducln[kduc];
                if (latitu <= 0) delr = z[1] - rducls[kduc];
                if (latitu >= 0) arglr = delr * delr / hl2n[kduc];
                if (latitu <= 0) arglr = delr * delr / hl2s[kduc];
                if (arglr >= 75.0) continue;
                frduct = m_exp(-arglr);
                if (latitu >= 0) delroh = 2.0 * delr / hl2n[kduc];
                if (latitu <= 0) delroh = 2.0 * delr / hl2s[kduc];
              }
            }
            if (!to235) {  // 1600
              delnl = delnl * frduct;
              anl = 1.0 + delnl;
              fac = delnl * deltal * l / (anl * d2);
              onedut = fac * 2.0 * cotz2;
              onedur = -fac / z[1] - delnl * delroh / anl;
            } else {  // 235
              anl = 1.0 + delnl;
              onedur = -delnl * deltal * l / (anl * d2 * z[1]);
              onedut = -onedur * 2.0 * z[1] * cotz2;
            }
            dlnldr = dlnldr + onedur;  // 1700
            dlnldt = dlnldt + onedut;
            ani[1] = ani[1] * anl;
          }
        }
      }
    }
    // 250
    for (int i = 2; i <= num; ++i) ani[i] = ani[1] * alpha[i];
    double vzs = (rbase / z[1]) * (rbase / z[1]);
    double dlnrdr = -sumi * vzs / (2.0 * q);
    for (int i = 2; i <= num; ++i) {
      dlgndr[i] = -dlnrdr - vzs / sh[i] + dlnldr + dlnlid;
      dlgndt[i] = dlnldt;
    }
    dlgndr[1] = dlnrdr + dlnldr + dlnlid;
    dlgndt[1] = dlnldt;
  }

  // ngo_dens_model.f95:29-160.  Returns 0 on success.
  int readinput(const char* filename) {
    pi = (double)3.141592653589793f;   // Q14: default-real literals
    clight = (double)2.99792501e05f;
    consf0 = (double)80.61638449f;
    r0 = (double)6370.f;
    ListReader rd;
    if (!rd.load(filename)) return -1;
    radgra = 180. / pi;
    grarad = 1. / radgra;
    am = (double)(1.67252e-27f / 9.10955854e-31f);
    double v[16];
    if (!rd.read(4, v)) return -2;
    intera = (int)v[0]; numres = (int)v[1]; nsuppr = (int)v[2]; spelat = v[3];
    kinit = 2;
    init = -1;
    latmax = -100.;
    latmin = 100.;
    for (;;) {  // label 105
      if (!rd.read(2, v)) return 0;  // end=9501
      double distre = v[0];
      latitu = v[1];
      double dummmm = 0.0;
      if (distre <= -1.) dummmm = 99999.0;
      if (latitu <= latmin) latmin = latitu;
      if (latitu >= latmax) latmax = latitu;
      if (dummmm > 99998.0) break;
    }
    if (fabs(latmax - latmin) <= (double)1.e-5f) nsuppr = 0;
    latmax = latmax + 3.;
    latmin = latmin - 3.;
    if (!rd.read(10, v)) return 0;
    num = (int)v[0]; kskip = (int)v[1]; mode = (int)v[2]; kount = (int)v[3]; kducts = (int)v[4]; ktape = (int)v[5];
    refalt = v[6]; dsrrng = v[7]; dsrlat = v[8]; dsdens = v[9];
    if (num == 0) return 0;
    if (!rd.read(5, v)) return 0;
    egfeq = v[0]; therm = v[1]; hm = v[2]; absb = v[3]; relb = v[4];
    if (!rd.read(5, v)) return 0;
    rbase = v[0]; ane0 = v[1]; alpha0[2] = v[2]; alpha0[3] = v[3]; alpha0[4] = v[4];
    if (!rd.read(5, v)) return 0;
    rzero = v[0]; scbot = v[1]; rstop = v[2]; rdiv = v[3]; hmin = v[4];
    if (kducts != 0) {
      if (!rd.read(5, v)) return 0;
      lk = v[0]; expk = v[1]; ddk = v[2]; rconsn = v[3]; scr = v[4];
      if (kducts != 1) {
        if (kducts > 10) return -3;
        for (int k = 2; k <= kducts; ++k) {
          if (!rd.read(12, v)) return 0;
          l0[k] = v[0]; def[k] = v[1]; dd[k] = v[2]; rducln[k] = v[3];
          double hducln = v[4]; rducun[k] = v[5]; double hducun = v[6]; rducls[k] = v[7];
          double hducls = v[8]; rducus[k] = v[9]; double hducus = v[10]; sidedu[k] = v[11];
          hl2n[k] = hducln * hducln;
          hl2s[k] = hducls * hducls;
          hu2n[k] = hducun * hducun;
          hu2s[k] = hducus * hducus;
        }
      }
    }
    if (!rd.read(8, v)) return 0;  // label 500
    double pstalt = v[0], palt1 = v[1], palt2 = v[2], platit = v[3], pstlat = v[4], plat1 = v[5], plat2 = v[6], paltit = v[7];
    // setting ane0 to desired value at dsrrng,dsrlat (:120-123)
    z[2] = (90.00 - dsrlat) * grarad;
    z[1] = dsrrng * r0;
    dens();
    ane0 = ane0 * dsdens / ani[1];
    // density-profile loops (:125-158) only fill plotting arrays; dens() keeps no state
    // besides kinit/alpha/ani/l, which the next adapter call overwrites -- but kinit is
    // sticky, so run them for fidelity.
    if (pstalt != 0.0) {
      z[2] = (90.00 - platit) * grarad;
      z[1] = palt1 * r0;
      double pstep = pstalt * r0, pfinal = palt2 * r0;
      int guard = 0;
      do { dens(); z[1] = z[1] + pstep; } while (z[1] <= pfinal && ++guard < 2000);
    }
    if (pstlat != 0.0) {
      z[1] = paltit + r0;
      z[2] = (90.00 + plat1) * grarad;
      double pstep = pstlat * grarad, pfinal = (90.00 + plat2) * grarad;
      int guard = 0;
      do { dens(); z[2] = z[2] + pstep; } while (z[2] <= pfinal && ++guard < 2000);
    }
    return 0;
  }
};

// fortran/ngo_dens_model_adapter.f95:63-207
struct NgoModel : Model {
  NgoState st;
  FieldBlock field;
  void* field_block_ptr() override { return &field; }
  void plasma(const double x[3], Plasma& o) override {
    ++n_plasma_calls;
    o.nspec = 4;
    double d2r = 2.0 * PI / 360.0;
    double p[3];
    cartesian_to_spherical(x, p);
    double L;
    double sp = m_sin(p[2]);
    if (R_E * (sp * sp) != 0.0) L = p[0] / (R_E * (sp * sp));
    else L = 0.0;
    st.l = L;  // the adapter imports the module's `L` (use ..., only: ... L ...)
    double lam = 90.0 - (p[2] * 360.0 / 2.0 / PI);
    double lamr = d2r * lam;
    double cl = m_cos(lamr);
    double r = st.r0 * L * (cl * cl);
    st.z[1] = r;
    st.z[2] = d2r * (90.0 - lam);
    st.latitu = lam;
    st.dens();
    double ce = st.ani[1], ch = st.ani[2], che = st.ani[3], co = st.ani[4];
    const double sgn[4] = {-1.0, 1.0, 1.0, 1.0};
    for (int i = 0; i < 4; ++i) o.qs[i] = 1.602e-19 * sgn[i];
    o.ms[0] = 9.10938188e-31;
    o.ms[1] = 1.6726e-27;
    o.ms[2] = 4.0 * 1.6726e-27;
    o.ms[3] = 16.0 * 1.6726e-27;
    o.Ns[0] = 1.0e6 * ce; o.Ns[1] = 1.0e6 * ch; o.Ns[2] = 1.0e6 * che; o.Ns[3] = 1.0e6 * co;
    for (int i = 0; i < 4; ++i) o.nus[i] = 0.0;
    field.eval(x, o.B0);
  }
};

}  // namespace orc

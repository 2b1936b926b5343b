// oracle/orc_igrf.h -- TEST INFRASTRUCTURE ONLY (CPU oracle).  PARITY UNPINNED.
//
// Restatement of the Geopack-2008 routines the adapters call when use_igrf = 1:
//   TSY_RECALC  tsyganenko/geopack0508_adapter.for:21-30
//   RECALC_08   tsyganenko/geopack2008.for:486-1195
//   SUN_08      tsyganenko/geopack2008.for:333-380
//   IGRF_GSM    tsyganenko/geopack0508_adapter.for:14-19  -> IGRF_GSW_08 geopack2008.for:55-181
//   GEOGSW_08   tsyganenko/geopack2008.for:1421-1456
// Everything is implicit REAL*4 as in the Fortran (compiled -O3 -fno-automatic, no fast-math);
// COMMON /GEOPACK1,2/ become the Geopack struct.  The coefficient DATA tables are shared with the
// product as data (srt_igrf_tables.h, generated from the DATA statements).
#pragma once
#include "orc_common.h"
#include "orc_models.h"
#include "../stanford_raytracer_b200/csrc/srt_igrf_tables.h"

namespace orc {

struct Geopack {
  // COMMON /GEOPACK1/ (only the members later routines read are kept)
  float ST0, CT0, SL0, CL0, CTCL, STCL, CTSL, STSL, SFI, CFI, SPS, CPS, DS3, CGST, SGST, PSI;
  float A11, A21, A31, A12, A22, A32, A13, A23, A33;
  float E11, E21, E31, E12, E22, E32, E13, E23, E33;
  // COMMON /GEOPACK2/ (index 0 unused)
  float G[106], H[106], REC[106];
  int year, day, hour, min, sec;
};

inline void SUN_08(int IYEAR, int IDAY, int IHOUR, int MIN, int ISEC, float& GST, float& SLONG, float& SRASN, float& SDEC) {
  const float RAD = 57.295779513f;
  if (IYEAR < 1901 || IYEAR > 2099) return;
  double FDAY = (double)(IHOUR * 3600 + MIN * 60 + ISEC) / 86400.0;
  double DJ = (365 * (IYEAR - 1900) + (IYEAR - 1901) / 4 + IDAY) - 0.5 + FDAY;
  float T = (float)(DJ / 36525.f);
  float VL = (float)fmod(279.696678f + 0.9856473354f * DJ, 360.0);
  GST = (float)(fmod(279.690983f + .9856473354f * DJ + 360.f * FDAY + 180.f, 360.0) / RAD);
  float G = (float)(fmod(358.475845f + 0.985600267f * DJ, 360.0) / RAD);
  SLONG = (VL + (1.91946f - 0.004789f * T) * sinf(G) + 0.020094f * sinf(2.f * G)) / RAD;
  if (SLONG > 6.2831853f) SLONG = SLONG - 6.2831853f;
  if (SLONG < 0.f) SLONG = SLONG + 6.2831853f;
  float OBLIQ = (23.45229f - 0.0130125f * T) / RAD;
  float SOB = sinf(OBLIQ);
  float SLP = SLONG - 9.924E-5f;
  float SIND = SOB * sinf(SLP);
  float COSD = sqrtf(1.f - SIND * SIND);
  float SC = SIND / COSD;
  SDEC = atanf(SC);
  SRASN = 3.141592654f - atan2f(cosf(OBLIQ) / SOB * SC, -cosf(SLP) / COSD);
}

inline void RECALC_08(Geopack& c, int IYEAR, int IDAY, int IHOUR, int MIN, int ISEC, float VGSEX, float VGSEY, float VGSEZ) {
  using namespace srt_igrf_tables;
  float* G = c.G;
  float* H = c.H;
  int IY = IYEAR;
  if (IY < 1965) IY = 1965;
  if (IY > 2025) IY = 2025;
  for (int N = 1; N <= 14; ++N) {
    int N2 = 2 * N - 1;
    N2 = N2 * (N2 - 2);
    for (int M = 1; M <= N; ++M) {
      int MN = N * (N - 1) / 2 + M;
      c.REC[MN] = (float)((N - M) * (N + M - 2)) / (float)N2;
    }
  }
  // the GOTO ladder of :890-1012: pick the bracketing epochs
  const float *Ga = nullptr, *Gb = nullptr, *Ha = nullptr, *Hb = nullptr;
  float epoch = 0.f;
  if (IY < 1970) { Ga = G65; Ha = H65; Gb = G70; Hb = H70; epoch = 1965; }
  else if (IY < 1975) { Ga = G70; Ha = H70; Gb = G75; Hb = H75; epoch = 1970; }
  else if (IY < 1980) { Ga = G75; Ha = H75; Gb = G80; Hb = H80; epoch = 1975; }
  else if (IY < 1985) { Ga = G80; Ha = H80; Gb = G85; Hb = H85; epoch = 1980; }
  else if (IY < 1990) { Ga = G85; Ha = H85; Gb = G90; Hb = H90; epoch = 1985; }
  else if (IY < 1995) { Ga = G90; Ha = H90; Gb = G95; Hb = H95; epoch = 1990; }
  else if (IY < 2000) { Ga = G95; Ha = H95; Gb = G00; Hb = H00; epoch = 1995; }
  else if (IY < 2005) { Ga = G00; Ha = H00; Gb = G05; Hb = H05; epoch = 2000; }
  else if (IY < 2010) { Ga = G05; Ha = H05; Gb = G10; Hb = H10; epoch = 2005; }
  else if (IY < 2015) { Ga = G10; Ha = H10; Gb = G15; Hb = H15; epoch = 2010; }
  else if (IY < 2020) { Ga = G15; Ha = H15; Gb = G20; Hb = H20; epoch = 2015; }
  if (Ga) {
    float F2 = ((float)IY + (float)(IDAY - 1) / 365.25f - epoch) / 5.f;
    float F1 = 1.f - F2;
    for (int N = 1; N <= 105; ++N) {
      G[N] = Ga[N - 1] * F1 + Gb[N - 1] * F2;
      H[N] = Ha[N - 1] * F1 + Hb[N - 1] * F2;
    }
  } else {
    float DT = (float)IY + (float)(IDAY - 1) / 365.25f - 2020.f;
    for (int N = 1; N <= 105; ++N) {
      G[N] = G20[N - 1];
      H[N] = H20[N - 1];
      if (N > 45) continue;
      G[N] = G[N] + DG20[N - 1] * DT;
      H[N] = H[N] + DH20[N - 1] * DT;
    }
  }
  float S = 1.f;
  for (int N = 2; N <= 14; ++N) {
    int MN = N * (N - 1) / 2 + 1;
    S = S * (float)(2 * N - 3) / (float)(N - 1);
    G[MN] = G[MN] * S;
    H[MN] = H[MN] * S;
    float P = S;
    for (int M = 2; M <= N; ++M) {
      float AA = 1.f;
      if (M == 2) AA = 2.f;
      P = P * sqrtf(AA * (float)(N - M + 1) / (float)(N + M - 2));
      int MNN = MN + M - 1;
      G[MNN] = G[MNN] * P;
      H[MNN] = H[MNN] * P;
    }
  }
  float G_10 = -G[2], G_11 = G[3], H_11 = H[3];
  float SQ = G_11 * G_11 + H_11 * H_11;
  float SQQ = sqrtf(SQ);
  float SQR = sqrtf(G_10 * G_10 + SQ);
  c.SL0 = -H_11 / SQQ;
  c.CL0 = -G_11 / SQQ;
  c.ST0 = SQQ / SQR;
  c.CT0 = G_10 / SQR;
  c.STCL = c.ST0 * c.CL0;
  c.STSL = c.ST0 * c.SL0;
  c.CTSL = c.CT0 * c.SL0;
  c.CTCL = c.CT0 * c.CL0;
  float GST = 0, SLONG = 0, SRASN = 0, SDEC = 0;
  SUN_08(IY, IDAY, IHOUR, MIN, ISEC, GST, SLONG, SRASN, SDEC);
  float S1 = cosf(SRASN) * cosf(SDEC);
  float S2 = sinf(SRASN) * cosf(SDEC);
  float S3 = sinf(SDEC);
  float DJ = (float)(365 * (IY - 1900) + (IY - 1901) / 4 + IDAY) - 0.5f + (float)(IHOUR * 3600 + MIN * 60 + ISEC) / 86400.f;
  float T = DJ / 36525.f;
  float OBLIQ = (23.45229f - 0.0130125f * T) / 57.2957795f;
  float DZ1 = 0.f, DZ2 = -sinf(OBLIQ), DZ3 = cosf(OBLIQ);
  float DY1 = DZ2 * S3 - DZ3 * S2;
  float DY2 = DZ3 * S1 - DZ1 * S3;
  float DY3 = DZ1 * S2 - DZ2 * S1;
  float V = sqrtf(VGSEX * VGSEX + VGSEY * VGSEY + VGSEZ * VGSEZ);
  float DX1 = -VGSEX / V, DX2 = -VGSEY / V, DX3 = -VGSEZ / V;
  float X1 = DX1 * S1 + DX2 * DY1 + DX3 * DZ1;
  float X2 = DX1 * S2 + DX2 * DY2 + DX3 * DZ2;
  float X3 = DX1 * S3 + DX2 * DY3 + DX3 * DZ3;
  c.CGST = cosf(GST);
  c.SGST = sinf(GST);
  float DIP1 = c.STCL * c.CGST - c.STSL * c.SGST;
  float DIP2 = c.STCL * c.SGST + c.STSL * c.CGST;
  float DIP3 = c.CT0;
  float Y1 = DIP2 * X3 - DIP3 * X2;
  float Y2 = DIP3 * X1 - DIP1 * X3;
  float Y3 = DIP1 * X2 - DIP2 * X1;
  float Y = sqrtf(Y1 * Y1 + Y2 * Y2 + Y3 * Y3);
  Y1 = Y1 / Y; Y2 = Y2 / Y; Y3 = Y3 / Y;
  float Z1 = X2 * Y3 - X3 * Y2;
  float Z2 = X3 * Y1 - X1 * Y3;
  float Z3 = X1 * Y2 - X2 * Y1;
  c.E11 = S1 * X1 + S2 * X2 + S3 * X3;
  c.E12 = S1 * Y1 + S2 * Y2 + S3 * Y3;
  c.E13 = S1 * Z1 + S2 * Z2 + S3 * Z3;
  c.E21 = DY1 * X1 + DY2 * X2 + DY3 * X3;
  c.E22 = DY1 * Y1 + DY2 * Y2 + DY3 * Y3;
  c.E23 = DY1 * Z1 + DY2 * Z2 + DY3 * Z3;
  c.E31 = DZ1 * X1 + DZ2 * X2 + DZ3 * X3;
  c.E32 = DZ1 * Y1 + DZ2 * Y2 + DZ3 * Y3;
  c.E33 = DZ1 * Z1 + DZ2 * Z2 + DZ3 * Z3;
  c.SPS = DIP1 * X1 + DIP2 * X2 + DIP3 * X3;
  c.CPS = sqrtf(1.f - c.SPS * c.SPS);
  c.PSI = asinf(c.SPS);
  c.A11 = X1 * c.CGST + X2 * c.SGST;
  c.A12 = -X1 * c.SGST + X2 * c.CGST;
  c.A13 = X3;
  c.A21 = Y1 * c.CGST + Y2 * c.SGST;
  c.A22 = -Y1 * c.SGST + Y2 * c.CGST;
  c.A23 = Y3;
  c.A31 = Z1 * c.CGST + Z2 * c.SGST;
  c.A32 = -Z1 * c.SGST + Z2 * c.CGST;
  c.A33 = Z3;
}

inline void GEOGSW_08(const Geopack& c, float& XGEO, float& YGEO, float& ZGEO, float& XGSW, float& YGSW, float& ZGSW, int J) {
  if (J > 0) {
    XGSW = c.A11 * XGEO + c.A12 * YGEO + c.A13 * ZGEO;
    YGSW = c.A21 * XGEO + c.A22 * YGEO + c.A23 * ZGEO;
    ZGSW = c.A31 * XGEO + c.A32 * YGEO + c.A33 * ZGEO;
  } else {
    XGEO = c.A11 * XGSW + c.A21 * YGSW + c.A31 * ZGSW;
    YGEO = c.A12 * XGSW + c.A22 * YGSW + c.A32 * ZGSW;
    ZGEO = c.A13 * XGSW + c.A23 * YGSW + c.A33 * ZGSW;
  }
}

inline void IGRF_GSW_08(const Geopack& c, float XGSW, float YGSW, float ZGSW, float& HXGSW, float& HYGSW, float& HZGSW) {
  float A[15], B[15];
  float XGEO, YGEO, ZGEO;
  GEOGSW_08(c, XGEO, YGEO, ZGEO, XGSW, YGSW, ZGSW, -1);
  float RHO2 = XGEO * XGEO + YGEO * YGEO;
  float R = sqrtf(RHO2 + ZGEO * ZGEO);
  float C = ZGEO / R;
  float RHO = sqrtf(RHO2);
  float S = RHO / R;
  float CF, SF;
  if (S < 1.E-5f) { CF = 1.f; SF = 0.f; }
  else { CF = XGEO / RHO; SF = YGEO / RHO; }
  float PP = 1.f / R;
  float P = PP;
  int IRP3 = (int)(R + 2);
  int NM = 3 + 30 / IRP3;
  if (NM > 13) NM = 13;
  int K = NM + 1;
  for (int N = 1; N <= K; ++N) {
    P = P * PP;
    A[N] = P;
    B[N] = P * N;
  }
  P = 1.f;
  float D = 0.f, BBR = 0.f, BBT = 0.f, BBF = 0.f;
  float X = 0.f, Y = 0.f, W, Q, Z, BI, P2, D2;
  int MM = 0;
  for (int M = 1; M <= K; ++M) {
    if (M == 1) {
      X = 0.f;
      Y = 1.f;
    } else {
      MM = M - 1;
      W = X;
      X = W * CF + Y * SF;
      Y = Y * CF - W * SF;
    }
    Q = P;
    Z = D;
    BI = 0.f;
    P2 = 0.f;
    D2 = 0.f;
    for (int N = M; N <= K; ++N) {
      float AN = A[N];
      int MN = N * (N - 1) / 2 + M;
      float E = c.G[MN];
      float HH = c.H[MN];
      W = E * Y + HH * X;
      BBR = BBR + B[N] * W * Q;
      BBT = BBT - AN * W * Z;
      if (M != 1) {
        float QQ = Q;
        if (S < 1.E-5f) QQ = Z;
        BI = BI + AN * (E * X - HH * Y) * QQ;
      }
      float XK = c.REC[MN];
      float DP = C * Z - S * Q - XK * D2;
      float PM = C * Q - XK * P2;
      D2 = Z;
      P2 = Q;
      Z = DP;
      Q = PM;
    }
    D = S * D + C * P;
    P = S * P;
    if (M == 1) continue;
    BI = BI * MM;
    BBF = BBF + BI;
  }
  float BR = BBR, BT = BBT, BF;
  if (S < 1.E-5f) {
    if (C < 0.f) BBF = -BBF;
    BF = BBF;
  } else {
    BF = BBF / S;
  }
  float HE = BR * S + BT * C;
  float HXGEO = HE * CF - BF * SF;
  float HYGEO = HE * SF + BF * CF;
  float HZGEO = BR * C - BT * S;
  GEOGSW_08(c, HXGEO, HYGEO, HZGEO, HXGSW, HYGSW, HZGSW, 1);
}

inline void TSY_RECALC(Geopack& c, int IYEAR, int IDAY, int IHOUR, int MIN, int ISEC) {
  RECALC_08(c, IYEAR, IDAY, IHOUR, MIN, ISEC, -400.0f, 0.0f, 0.0f);
}

// hooks for FieldBlock (orc_models.h): the adapters' date arithmetic, ngo_dens_model_adapter.f95:149-153
inline void orc_igrf_recalc_cb(void* ctx) {
  Geopack* c = (Geopack*)ctx;
  TSY_RECALC(*c, c->year, c->day, c->hour, c->min, c->sec);
}
inline void orc_igrf_gsm_cb(void* ctx, float x, float y, float z, float& bx, float& by, float& bz) {
  IGRF_GSW_08(*(const Geopack*)ctx, x, y, z, bx, by, bz);  // IGRF_GSM
}
inline bool orc_attach_igrf(FieldBlock& f, int yearday, int msec, int faithful_cost) {
  Geopack* c = new Geopack();  // lives as long as the model (test infrastructure: never freed)
  memset(c, 0, sizeof *c);
  c->year = yearday / 1000;
  c->day = yearday % 1000;
  c->hour = msec / (1000 * 60 * 60);
  c->min = (msec - c->hour * (1000 * 60 * 60)) / (1000 * 60);
  c->sec = (msec - c->hour * (1000 * 60 * 60) - c->min * (1000 * 60)) / (1000);
  TSY_RECALC(*c, c->year, c->day, c->hour, c->min, c->sec);
  f.igrf_ctx = c;
  f.psi = c->ST0;  // `PSI` of the adapters = word 1 of /GEOPACK1/ (Q25)
  f.igrf_gsm = orc_igrf_gsm_cb;
  f.recalc_each_call = faithful_cost ? orc_igrf_recalc_cb : nullptr;
  return true;
}

}  // namespace orc

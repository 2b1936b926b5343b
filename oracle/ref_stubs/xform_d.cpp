// oracle/ref_stubs/xform_d.cpp -- TEST INFRASTRUCTURE ONLY.
// C++ stand-ins for the Fortran entry points of libxformd that the reference's damping code links
// (damping/include/coord_transforms.h:31-48; there is no Fortran compiler here): sm_to_mag_d_, geo_to_mag_d_,
// mag_to_geo_d_, cart_to_pol_d_, pol_to_cart_d_, restated statement by statement from
// xform_double/{SM_TO_MAG,GEO_TO_MAG,MAG_TO_GEO,Cart_to_pol,POL_TO_CART,T0,T1,T2,T3,T4,T5,Get_q_c,Rotate_x,Rotate_y,Rotate_z}.f95
// with glibc's libm (what gfortran would call).  `float()` of the small integers involved is exact.
#include <math.h>

namespace {
const double PI_ = 3.141592653589793238462643;
const double DEGRAD = PI_ / 180.0;

void rotate_x(double a, const double x[3], double o[3]) {  // Rotate_x.f95
  o[0] = x[0];
  o[1] = x[1] * cos(a) + x[2] * sin(a);
  o[2] = x[2] * cos(a) - x[1] * sin(a);
}
void rotate_y(double a, const double x[3], double o[3]) {  // Rotate_y.f95
  o[0] = x[0] * cos(a) + x[2] * sin(a);
  o[1] = x[1];
  o[2] = x[2] * cos(a) - x[0] * sin(a);
}
void rotate_z(double a, const double x[3], double o[3]) {  // Rotate_z.f95
  o[0] = x[0] * cos(a) + x[1] * sin(a);
  o[1] = x[1] * cos(a) - x[0] * sin(a);
  o[2] = x[2];
}
double rmjd_of(const int* itime, int& iyr, int& iday, double& ut) {
  iyr = itime[0] / 1000;
  iday = itime[0] - iyr * 1000;
  ut = itime[1] / 3600000.0;
  double fracday = ut / 24.0;
  return 45.0 + (double)(float)(iyr - 1859) * 365.0 + ((double)(float)((iyr - 1861) / 4) + 1.0) + (double)(float)iday - 1.0 + fracday;
}
double t0(const int* itime, int& iyr, int& iday, double& ut) {  // T0.f95
  return (rmjd_of(itime, iyr, iday, ut) - 51544.5) / 36525.0;
}
void t1(const int* itime, const double x[3], double o[3], int iverse) {  // T1.f95
  int iyr, iday;
  double ut;
  double temp = t0(itime, iyr, iday, ut);
  double theta = (100.461 + 36000.770 * temp + 15.04107 * ut) * DEGRAD;
  rotate_z((double)(float)iverse * theta, x, o);
}
void t2(const int* itime, const double x[3], double o[3], int iverse) {  // T2.f95
  int iyr, iday;
  double ut, tmp[3];
  double tt0 = t0(itime, iyr, iday, ut);
  double epsilon = (23.439 - 0.013 * tt0) * DEGRAD;
  double m = (357.528 + 35999.05 * tt0 + 0.04107 * ut) * DEGRAD;
  double cgamma = 280.46 + 36000.772 * tt0 + 0.04107 * ut;
  double lamdas = (cgamma + (1.915 - 0.0048 * tt0) * sin(m) + 0.02 * sin(2.0 * m)) * DEGRAD;
  if (iverse == 1) {
    rotate_x(epsilon, x, tmp);
    rotate_z(lamdas, tmp, o);
  } else {
    rotate_z(iverse * lamdas, x, tmp);
    rotate_x(iverse * epsilon, tmp, o);
  }
}
void pol_to_cart(double lat, double lon, double radial, double o[3]) {  // POL_TO_CART.f95
  double coslat = cos(lat);
  o[0] = radial * coslat * cos(lon);
  o[1] = radial * coslat * sin(lon);
  o[2] = radial * sin(lat);
}
void dipole_axis_angles(const int* itime, double& phi, double& lamda) {  // T5.f95 / Get_q_c.f95
  int iyr, iday;
  double ut;
  double rmjd = rmjd_of(itime, iyr, iday, ut);
  double factor = (rmjd - 46066.0) / 365.25;
  phi = (78.8 + 4.283e-2 * factor) * DEGRAD;
  lamda = (289.1 - 1.413e-2 * factor) * DEGRAD;
}
void get_q_c(const int* itime, double q_c[3]) {  // Get_q_c.f95
  double phi, lamda, q_g[3], tmp[3];
  dipole_axis_angles(itime, phi, lamda);
  pol_to_cart(phi, lamda, 1.0, q_g);
  t1(itime, q_g, tmp, -1);
  t2(itime, tmp, q_c, 1);
}
void t3(const int* itime, const double x[3], double o[3], int iverse) {  // T3.f95
  double q_c[3], psi;
  get_q_c(itime, q_c);
  if (q_c[2] == 0.0) psi = -copysign(PI_ / 2.0, q_c[1]);
  else psi = -atan(q_c[1] / q_c[2]);
  rotate_x(iverse * psi, x, o);
}
void t4(const int* itime, const double x[3], double o[3], int iverse) {  // T4.f95
  double q_c[3];
  get_q_c(itime, q_c);
  double mu = -atan(q_c[0] / sqrt(q_c[1] * q_c[1] + q_c[2] * q_c[2]));
  rotate_y(iverse * mu, x, o);
}
void t5(const int* itime, const double x[3], double o[3], int iverse) {  // T5.f95
  double phi, lamda, tmp[3];
  const double pid2 = PI_ / 2.0;
  dipole_axis_angles(itime, phi, lamda);
  if (iverse == 1) {
    rotate_z(lamda, x, tmp);
    rotate_y(phi - pid2, tmp, o);
  } else {
    rotate_y(iverse * (phi - pid2), x, tmp);
    rotate_z(iverse * lamda, tmp, o);
  }
}
}  // namespace

extern "C" {
void sm_to_mag_d_(int* itime, double* x_in, double* x_out) {  // SM_TO_MAG.f95
  double a[3], b[3], c[3], d[3];
  t4(itime, x_in, a, -1);
  t3(itime, a, b, -1);
  t2(itime, b, c, -1);
  t1(itime, c, d, 1);
  t5(itime, d, x_out, 1);
}
void geo_to_mag_d_(int* itime, double* x_in, double* x_out) { t5(itime, x_in, x_out, 1); }   // GEO_TO_MAG.f95
void mag_to_geo_d_(int* itime, double* x_in, double* x_out) { t5(itime, x_in, x_out, -1); }  // MAG_TO_GEO.f95
void cart_to_pol_d_(double* x_in, double* lat, double* lon, double* radial) {  // Cart_to_pol.f95
  double row = sqrt(x_in[0] * x_in[0] + x_in[1] * x_in[1]);
  *radial = sqrt(row * row + x_in[2] * x_in[2]);
  *lat = atan2(x_in[2], row);
  *lon = atan2(x_in[1], x_in[0]);
  *lon = *lon + (1.0 - copysign(1.0, *lon)) * PI_;
}
void pol_to_cart_d_(double* lat, double* lon, double* radial, double* x_out) { pol_to_cart(*lat, *lon, *radial, x_out); }
}

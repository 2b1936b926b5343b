// TEST INFRASTRUCTURE (oracle/): the I/O statements of the reference's driver, executed by gfortran's OWN run-time library.
//
// The reference is compiled with gfortran; its driver reads the ray file with a list-directed READ
// (fortran/raytracer_driver.f95:1146) and writes output.ray with the edit descriptors i10 and es24.15e3 (:1197-1217).  No
// Fortran compiler exists here, but the run-time library those statements call does: libgfortran.so.5 travels inside the
// image's numpy/scipy wheels.  This shim issues the calls gfortran emits for those statements (st_open / st_read /
// transfer_real / st_write / transfer_*_write / st_close, ABI GFORTRAN_8, parameter blocks as libgfortran/io/io.h lays them
// out), so the product's formatter and ray-file reader are checked against the real digit generation and the real
// list-directed parser.  The library is opened at run time from the path the test hands in; nothing of it is linked or copied.
#include <dlfcn.h>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

// io.h: CHARACTER1(name) = pointer, length;  CHARACTER2(name) = length, pointer
typedef struct { int32_t flags, unit; const char* filename; int32_t line; size_t iomsg_len; char* iomsg; int32_t* iostat; } common_t;
typedef struct {
  common_t common; int64_t rec; int64_t* size; int64_t* iolength; void* internal_unit_desc;
  char* format; size_t format_len; size_t advance_len; char* advance; char* internal_unit; size_t internal_unit_len;
  size_t namelist_name_len; char* namelist_name; int32_t* id; int64_t pos;
  char* asynchronous; size_t asynchronous_len; size_t blank_len; char* blank; char* decimal; size_t decimal_len;
  size_t delim_len; char* delim; char* pad; size_t pad_len; size_t round_len; char* round; char* sign; size_t sign_len;
  char priv[4096];  // the library's private part (the compiler reserves 16 pointers + 32 ints)
} dt_t;
typedef struct {
  common_t common; int32_t recl_in; size_t file_len; char* file; char* status; size_t status_len; size_t access_len; char* access;
  char* form; size_t form_len; size_t blank_len; char* blank; char* position; size_t position_len; size_t action_len; char* action;
  char* delim; size_t delim_len; size_t pad_len; char* pad; char* convert; size_t convert_len; size_t decimal_len; char* decimal;
  char* encoding; size_t encoding_len; size_t round_len; char* round; char* sign; size_t sign_len; size_t asynchronous_len;
  char* asynchronous; int32_t* newunit; int32_t readonly; size_t cc_len; char* cc; char* share; size_t share_len; char spare[256];
} open_t;
typedef struct { common_t common; char* status; size_t status_len; char spare[256]; } close_t;

enum { HAS_IOSTAT = 1 << 5, DT_LIST_FORMAT = 1 << 7, DT_HAS_FORMAT = 1 << 12, DT_HAS_ADVANCE = 1 << 13, DT_HAS_INTERNAL_UNIT = 1 << 14,
       OPEN_HAS_FILE = 1 << 8, OPEN_HAS_STATUS = 1 << 9 };

static void (*st_write)(dt_t*), (*st_write_done)(dt_t*), (*st_read)(dt_t*), (*st_read_done)(dt_t*), (*st_open)(open_t*), (*st_close)(close_t*);
// array descriptors of ABI GFORTRAN_8 (libgfortran.h), for _gfortran_matmul_r8
typedef struct { ptrdiff_t stride, lbound, ubound; } dim_t;
typedef struct { size_t elem_len; int version; signed char rank, type; short attribute; } dtype_t;
typedef struct { double* base; size_t offset; dtype_t dtype; ptrdiff_t span; dim_t dim[2]; } desc_t;
static void (*matmul_r8)(desc_t*, desc_t*, desc_t*, int, int, void*);
static void (*xfer_int)(dt_t*, void*, int);
static void (*xfer_real_w)(dt_t*, void*, int), (*xfer_int_w)(dt_t*, void*, int), (*xfer_real)(dt_t*, void*, int), (*xfer_char_w)(dt_t*, void*, size_t);

int gfio_open_lib(const char* path) {
  void* h = dlopen(path, RTLD_NOW | RTLD_LOCAL);
  if (!h) { fprintf(stderr, "gfio: %s\n", dlerror()); return -1; }
  st_write = dlsym(h, "_gfortran_st_write"); st_write_done = dlsym(h, "_gfortran_st_write_done");
  st_read = dlsym(h, "_gfortran_st_read"); st_read_done = dlsym(h, "_gfortran_st_read_done");
  st_open = dlsym(h, "_gfortran_st_open"); st_close = dlsym(h, "_gfortran_st_close");
  xfer_real_w = dlsym(h, "_gfortran_transfer_real_write"); xfer_int_w = dlsym(h, "_gfortran_transfer_integer_write");
  xfer_real = dlsym(h, "_gfortran_transfer_real"); xfer_char_w = dlsym(h, "_gfortran_transfer_character_write");
  matmul_r8 = dlsym(h, "_gfortran_matmul_r8"); xfer_int = dlsym(h, "_gfortran_transfer_integer");
  return (matmul_r8 && xfer_int && st_write && st_write_done && st_read && st_read_done && st_open && st_close && xfer_real_w && xfer_int_w && xfer_real && xfer_char_w) ? 0 : -2;
}

static void open_unit(int unit, const char* path, const char* status) {
  static open_t o; memset(&o, 0, sizeof o);
  o.common.flags = OPEN_HAS_FILE | OPEN_HAS_STATUS; o.common.unit = unit; o.common.filename = "raytracer_driver.f95"; o.common.line = 1140;
  o.file = (char*)path; o.file_len = strlen(path); o.status = (char*)status; o.status_len = strlen(status);
  st_open(&o);
}
static void close_unit(int unit) {
  static close_t c; memset(&c, 0, sizeof c);
  c.common.unit = unit; c.common.filename = "raytracer_driver.f95"; c.common.line = 1232;
  st_close(&c);
}
static void dt_begin(dt_t* d, int flags, int unit, const char* fmt, int advance_no) {
  memset(d, 0, sizeof *d);
  d->common.flags = flags | (fmt ? DT_HAS_FORMAT : 0) | (advance_no ? DT_HAS_ADVANCE : 0); d->common.unit = unit;
  d->common.filename = "raytracer_driver.f95"; d->common.line = 1198;
  if (fmt) { d->format = (char*)fmt; d->format_len = strlen(fmt); }
  if (advance_no) { d->advance = "no"; d->advance_len = 2; }
}

// write(buf, '(es24.15e3)') x
void gfio_write_es(double x, char* out24) {
  static dt_t d; dt_begin(&d, DT_HAS_INTERNAL_UNIT, -1, "(es24.15e3)", 0);
  d.internal_unit = out24; d.internal_unit_len = 24;
  st_write(&d); xfer_real_w(&d, &x, 8); st_write_done(&d);
}

// the driver's output loop (raytracer_driver.f95:1197-1217) for nrows rows: head[row] = raynum, stopcond; v[row][17] = time, pos,
// vprel, vgrel, n, B0, w; sp[row][4 * nspec] = qs, ms, Ns, nus
void gfio_write_rayfile(const char* path, long nrows, const int* head, const double* v, int nspec, const double* sp) {
  static dt_t d;
  open_unit(61, path, "replace");
  for (long r = 0; r < nrows; ++r) {
    int rn = head[2 * r], sc = head[2 * r + 1];
    dt_begin(&d, 0, 61, "(i10, i10, 17es24.15e3, i10)", 1);
    st_write(&d);
    xfer_int_w(&d, &rn, 4); xfer_int_w(&d, &sc, 4);
    for (int k = 0; k < 17; ++k) { double x = v[17 * r + k]; xfer_real_w(&d, &x, 8); }
    xfer_int_w(&d, &nspec, 4);
    st_write_done(&d);
    for (int k = 0; k < 4 * nspec; ++k) {
      double x = sp[4 * nspec * r + k];
      dt_begin(&d, 0, 61, "(es24.15e3)", 1);
      st_write(&d); xfer_real_w(&d, &x, 8); st_write_done(&d);
    }
    dt_begin(&d, 0, 61, "(a)", 0);
    st_write(&d); xfer_char_w(&d, "", 0); st_write_done(&d);
  }
  close_unit(61);
}

// the driver's input loop (:1140-1150): read(infile, *, iostat=status) pos0, dir0, w until status /= 0.  -> rays read.
// pos0, dir0, w are variables of the main program: they keep their values from one READ to the next (a null value in the
// input leaves an item as it was) and start as zeros.
long gfio_read_rays(const char* path, double* out7, long cap) {
  static dt_t d;
  long n = 0;
  double x[7] = {0, 0, 0, 0, 0, 0, 0};
  open_unit(60, path, "old");
  while (n < cap) {
    int32_t status = 0;
    dt_begin(&d, DT_LIST_FORMAT | HAS_IOSTAT, 60, 0, 0);
    d.common.iostat = &status;
    st_read(&d);
    for (int k = 0; k < 7; ++k) xfer_real(&d, &x[k], 8);
    st_read_done(&d);
    if (status != 0) break;
    memcpy(out7 + 7 * n, x, sizeof x);
    ++n;
  }
  close_unit(60);
  return n;
}

// c = matmul(A, x), A(m,k) in Fortran order, x(k): the library routine gfortran calls for `a = matmul(Amat, x)` of
// tricubic_get_coeff_stacked (libtricubic.f95:719: 64 x 64 is above the inlining limit)
void gfio_matmul_r8(const double* A, const double* x, double* c, int m, int k) {
  desc_t da = {(double*)A, (size_t)(-1 - m), {8, 0, 2, 3, 0}, 8, {{1, 1, m}, {m, 1, k}}};
  desc_t db = {(double*)x, (size_t)-1, {8, 0, 1, 3, 0}, 8, {{1, 1, k}, {0, 0, 0}}};
  desc_t dc = {c, (size_t)-1, {8, 0, 1, 3, 0}, 8, {{1, 1, m}, {0, 0, 0}}};
  matmul_r8(&dc, &da, &db, 0, 0, 0);
}

// the end of dumpmodel.f95 (:1284-1289): write(outfile,'(5i10)') nspec,nx,ny,nz; write(outfile,'(6es24.15e3)') the six limits;
// write(outfile,'(es24.15e3)') f -- the whole array through a one-item format, so one value per record
void gfio_write_dump(const char* path, const int* head4, const double* lims6, const double* f, long n) {
  static dt_t d;
  open_unit(62, path, "replace");
  dt_begin(&d, 0, 62, "(5i10)", 0); st_write(&d);
  for (int k = 0; k < 4; ++k) { int v = head4[k]; xfer_int_w(&d, &v, 4); }
  st_write_done(&d);
  dt_begin(&d, 0, 62, "(6es24.15e3)", 0); st_write(&d);
  for (int k = 0; k < 6; ++k) { double v = lims6[k]; xfer_real_w(&d, &v, 8); }
  st_write_done(&d);
  dt_begin(&d, 0, 62, "(es24.15e3)", 0); st_write(&d);
  for (long i = 0; i < n; ++i) { double v = f[i]; xfer_real_w(&d, &v, 8); }
  st_write_done(&d);
  close_unit(62);
}

// the READ statements of the interp adapter's setup (interp_dens_model_adapter.f95:59-117): header, limits, qs, ms, one READ per
// node for f(1:nspec, j, k, l), and with computederivatives = 1 seven READs of whole arrays.  f and deriv[7] hold
// nspec*nx*ny*nz values each.  -> 0, or the iostat of the READ that failed
int gfio_read_grid(const char* path, int* head5, double* lims6, double* qs, double* ms, double* f, double* deriv, long cap) {
  static dt_t d;
  int32_t status = 0;
  open_unit(63, path, "old");
#define READ_BEGIN dt_begin(&d, DT_LIST_FORMAT | HAS_IOSTAT, 63, 0, 0); d.common.iostat = &status; st_read(&d)
#define READ_END st_read_done(&d); if (status != 0) { close_unit(63); return status; }
  READ_BEGIN; for (int k = 0; k < 5; ++k) xfer_int(&d, &head5[k], 4); READ_END
  READ_BEGIN; for (int k = 0; k < 6; ++k) xfer_real(&d, &lims6[k], 8); READ_END
  const int ns = head5[1];
  const long nodes = (long)head5[2] * head5[3] * head5[4];
  if (ns < 1 || ns > 16 || nodes * ns > cap) { close_unit(63); return -1000; }
  READ_BEGIN; for (int k = 0; k < ns; ++k) xfer_real(&d, &qs[k], 8); READ_END
  READ_BEGIN; for (int k = 0; k < ns; ++k) xfer_real(&d, &ms[k], 8); READ_END
  for (long node = 0; node < nodes; ++node) { READ_BEGIN; for (int k = 0; k < ns; ++k) xfer_real(&d, &f[node * ns + k], 8); READ_END }
  if (head5[0] == 1)
    for (int a = 0; a < 7; ++a) { READ_BEGIN; for (long i = 0; i < nodes * ns; ++i) xfer_real(&d, &deriv[a * nodes * ns + i], 8); READ_END }
  close_unit(63);
  return 0;
}

// a sequence of list-directed READs of REAL(8) items from one file, counts[r] items each, every READ with its own iostat (e.g. the
// cards of readinput, ngo_dens_model.f95:61-115).  out holds the items READ after READ (in/out: a null value keeps what is there);
// status[r] = iostat of READ r; reading stops at the first READ that fails.  -> READs completed
int gfio_read_counts(const char* path, const int* counts, int nreads, double* out, int* status) {
  static dt_t d;
  int done = 0;
  open_unit(64, path, "old");
  for (int r = 0; r < nreads; ++r) {
    int32_t st = 0;
    dt_begin(&d, DT_LIST_FORMAT | HAS_IOSTAT, 64, 0, 0);
    d.common.iostat = &st;
    st_read(&d);
    for (int k = 0; k < counts[r]; ++k) xfer_real(&d, &out[k], 8);
    st_read_done(&d);
    status[r] = st;
    if (st != 0) break;
    out += counts[r];
    ++done;
  }
  close_unit(64);
  return done;
}

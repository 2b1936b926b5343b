// oracle/ref_stubs/matstub.cpp -- TEST INFRASTRUCTURE ONLY.
// The handful of libmat / libmx calls the reference's damping code makes (damping/src/psd_model.cpp), served from the
// flat CRRES table of tools/gen_crres_table.py instead of a MATLAB run time.
//   matOpen(name)          opens $SRT_CRRES_TABLE, else `name` with its extension replaced by ".bin"
//   matGetVariable("crres") the 1 x 12 struct array; fields L, MLT, J_int (70 x 24, column-major), E, AE (1 x 1)
// Decision on a reference defect: the .mat stores AE as uint8 while psd_model.cpp:173 reads it with
// mxGetDoubles; here AE is handed out as the double the code evidently expects.  L and MLT are rebuilt as full
// arrays from their vectors (they are those vectors repeated; gen_crres_table.py checks it).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "matrix.h"
#include "mat.h"

struct mxArray_tag {
  std::vector<double> data;
  mwSize dims[2] = {1, 1};
  std::vector<mxArray_tag> entries;  // struct array: per element the fields in the order L, MLT, J_int, E, AE
  bool is_struct = false;
};
class MATFile {
 public:
  mxArray_tag root;
};

static mxArray_tag numeric(mwSize r, mwSize c) {
  mxArray_tag a;
  a.dims[0] = r;
  a.dims[1] = c;
  a.data.assign(r * c, 0.0);
  return a;
}

extern "C" {

MATFile* matOpen(const char* filename, const char* mode) {
  (void)mode;
  std::string path;
  if (const char* env = getenv("SRT_CRRES_TABLE")) path = env;
  else {
    path = filename;
    size_t dot = path.rfind('.');
    if (dot != std::string::npos) path.erase(dot);
    path += ".bin";
  }
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return NULL;
  int hdr[4];
  if (fread(hdr, sizeof(int), 4, f) != 4 || hdr[0] != 0x53525443) { fclose(f); return NULL; }
  const int n = hdr[1], nL = hdr[2], nM = hdr[3];
  MATFile* m = new MATFile();
  m->root.is_struct = true;
  m->root.dims[0] = 1;
  m->root.dims[1] = (mwSize)n;
  for (int k = 0; k < n; ++k) {
    double ea[2];
    std::vector<double> L(nL), M(nM);
    mxArray_tag J = numeric(nL, nM);
    if (fread(ea, 8, 2, f) != 2 || fread(L.data(), 8, nL, f) != (size_t)nL || fread(M.data(), 8, nM, f) != (size_t)nM ||
        fread(J.data.data(), 8, (size_t)nL * nM, f) != (size_t)nL * nM) { fclose(f); delete m; return NULL; }
    mxArray_tag La = numeric(nL, nM), Ma = numeric(nL, nM), E = numeric(1, 1), AE = numeric(1, 1);
    for (int c = 0; c < nM; ++c)
      for (int r = 0; r < nL; ++r) { La.data[r + (size_t)nL * c] = L[r]; Ma.data[r + (size_t)nL * c] = M[c]; }
    E.data[0] = ea[0];
    AE.data[0] = ea[1];
    mxArray_tag el;
    el.entries = {La, Ma, J, E, AE};
    m->root.entries.push_back(el);
  }
  fclose(f);
  return m;
}
matError matClose(MATFile* m) { delete m; return 0; }
mxArray* matGetVariable(MATFile* m, const char* name) { return (m && !strcmp(name, "crres")) ? &m->root : NULL; }

size_t mxGetNumberOfElements(const mxArray* a) { return a->is_struct ? a->entries.size() : a->data.size(); }
mwSize mxGetNumberOfDimensions(const mxArray*) { return 2; }
const mwSize* mxGetDimensions(const mxArray* a) { return a->dims; }
double* mxGetDoubles(const mxArray* a) { return const_cast<double*>(a->data.data()); }
mxArray* mxGetField(const mxArray* a, mwIndex index, const char* field) {
  static const char* names[5] = {"L", "MLT", "J_int", "E", "AE"};
  if (!a->is_struct || index >= a->entries.size()) return NULL;
  for (int i = 0; i < 5; ++i)
    if (!strcmp(field, names[i])) return const_cast<mxArray*>(&a->entries[index].entries[i]);
  return NULL;
}
bool mxIsNaN(double v) { return v != v; }

}  // extern "C"

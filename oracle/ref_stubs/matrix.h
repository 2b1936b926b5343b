/* oracle/ref_stubs/matrix.h -- TEST INFRASTRUCTURE ONLY.
 * Stand-in for MATLAB's <matrix.h> (not vendored by the reference: damping/include/mat.h:33 includes it), just
 * enough for the reference's own damping sources to compile unchanged (damping/src/psd_model.cpp:14-33,51-111,
 * 158-215).  The MAT-file behind it is the flat table tools/gen_crres_table.py writes from
 * damping/data/crres_clean.mat; matstub.cpp implements the calls. */
#ifndef SRT_REF_STUB_MATRIX_H
#define SRT_REF_STUB_MATRIX_H
#include <stddef.h>
typedef size_t mwSize;
typedef size_t mwIndex;
typedef struct mxArray_tag mxArray;
typedef struct { double real, imag; } mxComplexDouble;
#ifdef __cplusplus
extern "C" {
#endif
size_t mxGetNumberOfElements(const mxArray* a);
mwSize mxGetNumberOfDimensions(const mxArray* a);
const mwSize* mxGetDimensions(const mxArray* a);
double* mxGetDoubles(const mxArray* a);
mxArray* mxGetField(const mxArray* a, mwIndex index, const char* field);
bool mxIsNaN(double v);
#ifdef __cplusplus
}
#endif
#endif

/* srt.h -- C ABI of libsrt.so: batched Haselgrove ray integration on the B200s of one box.
 *
 * The reference has no C ABI; what this library replaces is the Fortran call
 *   call raytracer_run(pos,time,vprel,vgrel,n,B0,qs,ms,Ns,nus,stopcond, pos0,dir0,w,
 *        dt0,dtmax,maxerr,maxsteps,minalt,root,tmax,fixedstep,del,funcPlasmaParams,data,
 *        raytracer_stopconditions)
 * made once per input ray by fortran/raytracer_driver.f95:1154-1194 (signature
 * fortran/raytracer.f95:609-642), together with the model set-up that precedes it
 * (ngosetup, driver:256-369; interpsetup, driver:497-615).  The per-ray call becomes
 * one srt_trace_batch() over all rays of the input file.  `funcPlasmaParams` cannot
 * cross to the device as a callback: the two adapters on the hot path are built-in
 * models (srt_model_ngo = modelnum 1, srt_model_grid = modelnum 3), and `del` is
 * implied by the model (delSP = 1e-4 / delDP = 1e-6, driver:251-252,1158,1170).
 *
 * Conventions: plain C types; every call returns 0 or a negative SRT_E_* code (never
 * exits); input arrays are caller-owned; the result object is library-owned until
 * srt_result_free().  srt_trace_batch() spreads the rays of a batch over every device
 * selected by srt_init() / srt_init_devices() (ray i on device i mod n, one host thread
 * and one model replica per device, no exchange between devices: SURVEY.md 8b "blocking;
 * internally multi-GPU"); the other compute calls use the calling thread's current
 * device.  Under a one-process-per-GPU launcher each process selects its own device
 * and rays shard across processes by index (stanford_raytracer_b200/shard.py).  There
 * is no CPU fallback: without a CUDA device every compute call fails with
 * SRT_E_NODEVICE.
 *
 * INTEGRATION.md shows the ISO_C_BINDING interface block and the patched driver loop.
 */
#ifndef SRT_H
#define SRT_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define SRT_OK 0
#define SRT_E_NODEVICE -1   /* no usable CUDA device */
#define SRT_E_IO -2         /* cannot open / parse a model file */
#define SRT_E_ARG -3        /* invalid argument */
#define SRT_E_UNSUPPORTED -4 /* e.g. use_tsyganenko != 0 */
#define SRT_E_CUDA -5       /* CUDA runtime error (srt_last_cuda_error()) */
#define SRT_E_NOMEM -6      /* trajectory log cannot hold even one ray */
#define SRT_E_TOOBIG -7     /* more points than the static scratch of srt_model_eval_noalloc holds */

/* one recorded trajectory row: the non-constant columns that driver:1197-1217 writes */
#define SRT_ROW_DOUBLES 20
#define SRT_ROW_T 0      /* time            */
#define SRT_ROW_POS 1    /* pos(3)          */
#define SRT_ROW_VPREL 4  /* vprel(3)        */
#define SRT_ROW_VGREL 7  /* vgrel(3)        */
#define SRT_ROW_N 10     /* n(3)            */
#define SRT_ROW_B0 13    /* B0(3)           */
#define SRT_ROW_NS 16    /* Ns(nspec<=4)    */

/* per-ray flag bits (srt_result_flags) */
#define SRT_FLAG_FIXED_CONE_EXIT 1 /* fixed-step run left the resonance cone and returned with
                                      stopcond 0 (raytracer.f95:900-905) */

typedef struct srt_model srt_model;
typedef struct srt_result srt_result;

/* the scalar arguments of raytracer_run (raytracer.f95:620-623) + the driver's outputper (driver:223-228) */
typedef struct srt_run_params {
  double dt0, dtmax, maxerr, minalt, tmax;
  int32_t maxsteps, root, fixedstep, outputper;
} srt_run_params;

/* Selects the CUDA device of this process.  device < 0: keep the current one. */
int srt_init(int device);
/* Selects the devices srt_trace_batch() spreads a batch over (the reference's one-ray-at-a-time loop,
 * driver:1143-1232, becomes one call that uses the whole box).  ndevices <= 0: every visible device.
 * The first one becomes the calling thread's current device. */
int srt_init_devices(const int* devices, int ndevices);
int srt_device_count(void); /* devices selected by the last srt_init / srt_init_devices */
const char* srt_strerror(int code);
const char* srt_last_cuda_error(void);

/* modelnum 1 -- ngosetup + ngoStateData (driver:256-369; ngo_dens_model_adapter.f95:15-29,42-51) */
int srt_model_ngo(srt_model** m, const char* ngo_configfile, int yearday, int milliseconds_day, int use_igrf, int use_tsyganenko);
/* modelnum 3 -- interpsetup + interpStateData (driver:497-615; interp_dens_model_adapter.f95:16-39,52-134) */
int srt_model_grid(srt_model** m, const char* interp_interpfile, int yearday, int milliseconds_day, int use_igrf, int use_tsyganenko);
/* modelnum 4 -- scatteredinterpsetup + scatteredinterpStateData (driver:617-770; fortran/scattered_interp_dens_model_adapter.f95:17-48,
 * 63-212): density by moving least squares (fortran/lsinterp_mod.f95:255-449) over the samples of `interp_interpfile` within
 * window_scale * (largest nearest-neighbour distance of the file); order 0..3 = degree of the monomials; exact, scaled,
 * local_window_scale as the driver's --scattered_interp_* options (driver:704-729; the driver always passes scaled = 0).
 * nspec <= 4; order > 3 returns SRT_E_UNSUPPORTED.  The reference's kd-tree (fortran/kdtree_mod.f95) is replaced by a cell
 * list with the same neighbour sets; what that means for the last bits is stated in csrc/srt_scattered.h. */
int srt_model_scattered(srt_model** m, const char* interp_interpfile, int yearday, int milliseconds_day, int use_igrf, int use_tsyganenko,
                        double window_scale, int order, int exact, int scaled, double local_window_scale);
/* what set-up found: samples kept, the largest nearest-neighbour distance (adapter :160-192), the search radius, the number
 * of monomials, the dimensions of the cell list (any pointer may be NULL) */
int srt_model_scattered_info(const srt_model* m, int64_t* npoints, double* maxnearest, double* radius, int* nmono, int* cells /*[3]*/);
/* the samples set-up kept (repeated positions dropped, adapter :143-155), in the order a query walks them:
 * recs[npoints][8] = x, y, z, log-density of species 1..4, distance to the nearest other sample (adapter :160-191; 1.0 for a
 * sample inside the Earth).  capacity in doubles; recs = NULL only counts.  Returns npoints, or an error code < 0. */
int64_t srt_model_scattered_samples(const srt_model* m, double* recs, int64_t capacity);
/* use_tsyganenko = 1 (fortran/ngo_dens_model_adapter.f95:157-166,183-187): the TS05 external field T04_s
 * (tsyganenko/TS05_aka_TS04.for:5-116) is added to the base field.  parmod[10] = Pdyn, Dst, ByIMF, BzIMF, W1..W6, the
 * driver's --tsyganenko_* options (fortran/raytracer_driver.f95:262-291).  Also switches the term on for a model that
 * was set up with use_tsyganenko = 0.  The tilt argument is what the adapters pass (word 1 of Geopack's /GEOPACK1/).
 * Batches of such a model run the thread-per-ray kernel. */
int srt_model_tsyganenko(srt_model* m, const double* parmod /*[10]*/);
void srt_model_free(srt_model* m);
/* qs, ms, nus as the adapter returns them on every call (constant per model) */
int srt_model_species(const srt_model* m, int* nspec, double* qs, double* ms, double* nus);
/* funcPlasmaParams at n points (host arrays): out[n][19] = qs4, Ns4, ms4, nus4, B0(3).
 * The parity surface of fortran/dumpmodel.f95:374-378,649-653. */
int srt_model_eval(const srt_model* m, int64_t n, const double* x /*[n][3]*/, double* out /*[n][19]*/);
/* srt_model_eval for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.  Same arguments
 * and results as srt_model_eval; calls no cudaMalloc / cudaFree / allocating routine: points and results are staged in
 * a static device array of srt_model_eval_noalloc_max_points() = 16384 points (2.75 MiB, held by every GPU that loads
 * the library until exit).  n beyond that returns SRT_E_TOOBIG (text in srt_last_cuda_error()).  A gridded model must
 * already be on the device (any earlier ordinary call with it), else SRT_E_UNSUPPORTED.  All calls on a device share
 * the array, which keeps the last call's contents; the call is synchronous and serialised per device inside this
 * library, but calls from several PROCESSES do not share the array, and device work of this call overlapping other work
 * that used the same array would race. */
int srt_model_eval_noalloc(const srt_model* m, int64_t n, const double* x /*[n][3]*/, double* out /*[n][19]*/);
int64_t srt_model_eval_noalloc_max_points(void);

/* raytracer_run for nrays rays, host buffers.  pos0[nrays][3], dir0[nrays][3] (inout: an all-zero
 * direction is replaced by the field-aligned one, raytracer.f95:661-674), w0[nrays]. */
int srt_trace_batch(const srt_model* m, int64_t nrays, const double* pos0, double* dir0, const double* w0, const srt_run_params* p,
                    srt_result** out);

int64_t srt_result_nrays(const srt_result* r);
int64_t srt_result_total_rows(const srt_result* r);     /* rows held (after outputper decimation) */
int64_t srt_result_accepted_steps(const srt_result* r); /* sum over rays of (rows the reference would hold - 1) */
int64_t srt_result_attempts(const srt_result* r);       /* accepted + rejected + cone-refined attempts */
/* Of the accepted steps, those of rays whose state had become all-NaN: the reference steps such a ray on to
 * maxsteps / tmax, every row equal to the last but for t = t + dt (tests/test_nan_rays_cpu.py); the kernels
 * write those rows without integrating (srt_nan_fastforward).  Rates quoted "without fast-forwarded steps"
 * subtract this count. */
int64_t srt_result_fastforwarded_steps(const srt_result* r);
int srt_result_devices(const srt_result* r);            /* devices the batch ran on */
/* per ray */
int64_t srt_result_nrows(const srt_result* r, int64_t ray);       /* rows held for this ray = ceil(nsteps/outputper) */
int64_t srt_result_nsteps(const srt_result* r, int64_t ray);      /* size(time) in the reference */
int srt_result_stopcond(const srt_result* r, int64_t ray);
int srt_result_flags(const srt_result* r, int64_t ray);
int srt_result_copy_ray(const srt_result* r, int64_t ray, double* rows /*[nrows][SRT_ROW_DOUBLES]*/);
/* zero-copy view: *rows points at the rows of `ray` (row-major [nrows][20]) */
int srt_result_ray_view(const srt_result* r, int64_t ray, const double** rows, int64_t* nrows);
/* endpoint summary for all rays (any pointer may be NULL) */
int srt_result_summary(const srt_result* r, double* last_row /*[nrays][20]*/, int64_t* nsteps, int64_t* nrejected, int32_t* stopcond);
/* Where the rows live: a batch returns one row buffer ("chunk") per device and wave.  With srt_set_rows_shm() the
 * chunks are POSIX shared-memory segments (page-locked for the D2H copy), so that the gathering process of a
 * one-process-per-GPU run maps the rows of the other ranks instead of receiving a copy (shard.py). */
int srt_set_rows_shm(const char* prefix /* NULL or "": ordinary pinned memory */);
int srt_result_chunk_count(const srt_result* r);
int srt_result_chunk_info(const srt_result* r, int chunk, char* shm_name, int name_capacity, int64_t* nrows, const double** rows);
/* per ray: chunk, first row inside the chunk, rows held (arrays of nrays entries; any may be NULL) */
int srt_result_ray_location(const srt_result* r, int64_t nrays, int32_t* chunk, int64_t* first_row, int64_t* nrows);
void srt_result_free(srt_result* r);
/* frees every buffer the library caches between calls (device scratch, staging, pooled row buffers) */
int srt_release_scratch(void);

/* Device-resident variant (pipelines, bench `value`): inputs already in HBM, outputs stay there.
 *   d_rays      [7][nrays] structure-of-arrays: pos0 x,y,z; dir0 x,y,z; w0   (dir0 rows are updated in place)
 *   d_log       [log_capacity][21] append log: word 0 = (ray | row_index << 32) bit pattern, words 1..20 = the row
 *   d_summary   [nrays][24]: words 0..19 last row, 20 nsteps, 21 stopcond, 22 rejected+cone attempts, 23 flags
 *               (integers stored as doubles)
 *   d_counters  [8] uint64: log rows used, accepted steps, attempts, rays dropped because the log was full,
 *               ray-queue head, fast-forwarded NaN steps (included in accepted and attempts), attempts taken from a shadow team
 *               (team kernel, included in attempts), most block rounds of a block
 *               (zeroed by the call)
 * stream: a cudaStream_t (NULL = default stream).  Asynchronous: returns after the launch. */
int srt_trace_batch_device(const srt_model* m, int64_t nrays, double* d_rays, const srt_run_params* p, double* d_log, int64_t log_capacity,
                           double* d_summary, uint64_t* d_counters, void* stream);
/* kernels launched by this library since srt_init (for bench.py's gpu_launches) */
int64_t srt_launch_count(void);
/* which trace kernel a batch of nrays rays runs (mode 0: one thread per ray, 1: team of 8 lanes per ray;
 * SRT_KERNEL=thread|team overrides), its registers/thread, local bytes/thread, blocks/SM, persistent grid, block */
int srt_kernel_info(const srt_model* m, int64_t nrays, int* mode, int* regs, int* local_bytes, int* blocks_per_sm, int* grid, int* block);
/* Self-test: the interleaved IEEE division of the kernels (srt_divn, csrc/srt_math.h) against the compiler's `/` on
 * 4 * ngroups pseudo-random operand pairs incl. zeros, infinities, NaNs and subnormals; *mismatches = results that differ. */
int srt_selftest_div(int64_t ngroups, uint64_t seed, uint64_t* mismatches);
/* FP64 pipe micro-benchmark used as the roofline denominator: runs `iters` dependent-free DFMA (fma=1)
 * or DMUL+DADD pairs (fma=0) per thread on a full grid; returns achieved GFLOP/s in *gflops. */
int srt_fp64_peak(int fma, int iters, double* gflops);

/* ---- damping along the ray (BASELINE config 5): the reference's post-processor `damping` in its default mode 1
 * (damping/src/damping_main.cpp:131-143 -> damping_foust, damping/src/damping_foust.cpp:15-236), for all rays of a run.
 *   rows       [nrows][SRT_ROW_DOUBLES], the recorded rows of ray 0, then ray 1, ... (what damping reads back from the
 *              output file, damping/src/wipp_fileutils.cpp:6-170); first_row[nrays] = index of each ray's first row
 *   w[nrays]   wave angular frequency; qs, ms [nspec]: the species constants of the run
 *   Kp, AE_level, yearday, msec, include_geom_factor: the options -k -a -t -u -v of damping_main.cpp:44-98
 *   crres_table: path of the CRRES flux table (stanford_raytracer_b200/data/crres_clean.bin, tools/gen_crres_table.py)
 *   power[nrows]: relative wave power at every row (1 at each ray's first row) -- the third column of the file
 *              write_damping emits (wipp_fileutils.cpp:234-252). */
int srt_damping_foust(int64_t nrays, const int64_t* first_row, int64_t nrows, const double* rows, const double* w, int nspec, const double* qs,
                      const double* ms, double Kp, double AE_level, int yearday, int msec, int include_geom_factor, const char* crres_table,
                      double* power);
/* the same for every ray of a batch result, rows taken in ray order; first_row[nrays] and power[total rows] are the caller's */
int srt_result_damping(const srt_result* r, const srt_model* m, const double* w0 /*[nrays]*/, double Kp, double AE_level, int yearday, int msec,
                       int include_geom_factor, const char* crres_table, int64_t* first_row, double* power);
const char* srt_damping_last_error(void);

/* ---- the driver's two text formats (host side, no GPU; srt_io.cpp) ------------------------------------------------
 * Text of n doubles under the ES24.15E3 edit descriptor (driver:1197-1217), 24 characters each, no terminator. */
int srt_format_es24(const double* v, int64_t n, char* out /*[24*n]*/);
/* bytes of one output line for nspec species: 2 I10, 17 ES24, I10, 4*nspec ES24, newline */
int64_t srt_row_text_bytes(int nspec);
/* The lines the driver writes for one ray (driver:1197-1217): raynum (1-based), stopcond, the 16 leading columns of
 * each row, w, nspec, qs, ms, Ns (row columns 16..), nus.  Returns the bytes written (nrows * srt_row_text_bytes). */
int64_t srt_format_rows(int64_t raynum, int stopcond, const double* rows /*[nrows][SRT_ROW_DOUBLES]*/, int64_t nrows, double w, int nspec,
                        const double* qs, const double* ms, const double* nus, char* out);
/* The whole output file of a batch, byte for byte what the driver's loop writes (rays in input order, raynum counted
 * from first_raynum).  nthreads workers format, the caller's thread writes; nthreads <= 0: all host cores. */
int srt_result_write_text(const srt_result* r, const srt_model* m, const double* w0 /*[nrays]*/, const char* path, int64_t first_raynum,
                          int append, int nthreads);
/* inputraysfile (fortran/raytracer_driver.f95:1140-1150): the loop `read(infile,*,iostat=status) pos0, dir0, w` under gfortran's
 * list-directed rules (blank / comma / semicolon separators, records continuing, null values keeping the previous READ's value,
 * r*c and r*, `/` ending a READ, D / Q exponents, 1.5+3, INF / NAN); stops at the first token that is not a number or at the end
 * of the file.  The three arrays are allocated by the library: release each with srt_free. */
int srt_read_inputrays(const char* path, double** pos0 /*[n][3]*/, double** dir0 /*[n][3]*/, double** w0 /*[n]*/, int64_t* nrays);
void srt_free(void* p);

#ifdef __cplusplus
}
#endif
#endif

// srt_cli.cpp -- `raytracer`: the reference's command line (fortran/raytracer_driver.f95) over libsrt.so.
//
// Same `--name=value` options (driver:181-228 common, :263-337 Ngo, :502-575 interp), same input ray file
// (driver:1146), same output file byte for byte (driver:1197-1217).  The per-ray loop of the reference
// (read a ray, raytracer_run, write its rows; driver:1144-1232) becomes: read all rays, srt_trace_batch on the GPU,
// write all rows.  Models: --modelnum=1 (Ngo), 3 (gridded) and 4 (scattered samples); the others are not on the GPU path and are refused.
//
// Option parsing follows getopt_named (fortran/util.f95:53-84): an argument counts if it contains '=', its name is
// the text between the first two characters and the first '=', the first match wins.  Integers are read as reals and
// floored, as the driver does.  Options the reference leaves uninitialised when absent are required here.
//
// Additions (not in the reference): --device=N (one CUDA device; default: every GPU of the box), --threads=N (formatter threads,
// default all cores), --batch=N (rays per srt_trace_batch call; default: as many as keep the worst-case
// trajectory rows, maxsteps/outputper per ray, inside a quarter of the host's memory), --verbose=1 (timing to stderr).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include <unistd.h>

#include "srt.h"

namespace {

int g_argc;
char** g_argv;

bool getopt_named(const char* name, std::string& out) {
  for (int i = 1; i < g_argc; ++i) {
    const char* a = g_argv[i];
    const char* eq = strchr(a, '=');
    if (!eq || eq - a < 2) continue;
    if (std::string(a + 2, eq) == name) { out = eq + 1; return true; }
  }
  return false;
}

bool read_real(const std::string& s, double& v) {
  std::string t;
  for (char c : s) t.push_back((c == 'd' || c == 'D') ? 'e' : c);
  char* end = nullptr;
  v = strtod(t.c_str(), &end);
  while (end && (*end == ' ' || *end == ',')) ++end;
  return end != t.c_str();
}

[[noreturn]] void die(int code, const std::string& msg) {
  fprintf(stderr, "raytracer: %s\n", msg.c_str());
  exit(code);
}

double need_real(const char* name) {
  std::string b;
  double v;
  if (!getopt_named(name, b)) die(2, std::string("--") + name + " is required");
  if (!read_real(b, v)) die(2, std::string("--") + name + "=" + b + " is not a number");
  return v;
}

double opt_real(const char* name, double dflt) {
  std::string b;
  double v;
  if (!getopt_named(name, b)) return dflt;
  if (!read_real(b, v)) die(2, std::string("--") + name + "=" + b + " is not a number");
  return v;
}

void usage() {
  puts("Usage:\n"
       "  raytracer --param1=value1 --param2=value2 ...\n"
       "\n"
       "--dt0           initial timestep in seconds\n"
       "--dtmax         maximum timestep in seconds\n"
       "--tmax          maximum time for simulation\n"
       "--root          root number (2=whistler at VLF frequencies)\n"
       "--fixedstep     fixed timesteps (1) or adaptive (0)\n"
       "--maxerr        maximum error for adaptive timestepping\n"
       "--maxsteps      maximum number of timesteps\n"
       "--minalt        minimum altitude\n"
       "--inputraysfile ray input filename\n"
       "--outputfile    output filename\n"
       "--outputper     output fields only every \"outputper\" steps\n"
       "--modelnum      (1) Ngo model\n"
       "                (3) Interpolated model (gridded data)\n"
       "                (4) Interpolated model (scattered data)\n"
       "                (2, 5-7 are not available on the GPU path)\n"
       "\n"
       " Ngo parameters (model 1):\n"
       "   --ngo_configfile     newray input filename\n"
       " Interp parameters (model 3):\n"
       "   --interp_interpfile  grid filename\n"
       " Scattered interp parameters (model 4):\n"
       "   --interp_interpfile  data filename\n"
       "   --scattered_interp_window_scale        window radius as a factor of the maximum sample spacing\n"
       "   --scattered_interp_order               monomial order (0..3)\n"
       "   --scattered_interp_exact               exact (1) or inexact (0) weighting window\n"
       "   --scattered_interp_local_window_scale  window width above the local average sample spacing\n"
       " All:\n"
       "   --yearday            year and day, e.g., 1999098\n"
       "   --milliseconds_day   milliseconds of day\n"
       "   --use_igrf           (1=use, 0=do not use)\n"
       "   --use_tsyganenko     (1=use, 0=do not use)\n"
       "   --tsyganenko_Pdyn    between 0.5 and 10 nPa\n"
       "   --tsyganenko_Dst     between -100 and +20 in nT\n"
       "   --tsyganenko_ByIMF   between -10 and +10 nT\n"
       "   --tsyganenko_BzIMF   between -10 and +10 nT\n"
       "   --tsyganenko_W1 .. --tsyganenko_W6   TS04_s W1 .. W6\n"
       "\n"
       " Not in the reference:\n"
       "   --device=N (default: all GPUs)  --threads=N  --batch=N  --verbose=1");
}

}  // namespace

int main(int argc, char** argv) {
  g_argc = argc;
  g_argv = argv;
  if (argc <= 1) {  // driver:75-177: usage, stop
    usage();
    return 0;
  }
  srt_run_params p;
  p.dt0 = need_real("dt0");
  p.dtmax = need_real("dtmax");
  p.tmax = need_real("tmax");
  p.root = (int32_t)std::floor(need_real("root"));
  p.fixedstep = (int32_t)std::floor(need_real("fixedstep"));
  p.maxerr = need_real("maxerr");
  p.maxsteps = (int32_t)std::floor(need_real("maxsteps"));
  p.minalt = need_real("minalt");
  p.outputper = (int32_t)std::floor(opt_real("outputper", 1.0));  // driver:223-228
  if (p.outputper < 1) die(2, "--outputper must be >= 1");
  std::string inputraysfile, outputfile, modelfile;
  if (!getopt_named("inputraysfile", inputraysfile)) die(2, "--inputraysfile is required");
  if (!getopt_named("outputfile", outputfile)) die(2, "--outputfile is required");
  const int modelnum = (int)std::floor(need_real("modelnum"));
  const int yearday = (int)std::floor(opt_real("yearday", 0.0));
  const int msec = (int)std::floor(opt_real("milliseconds_day", 0.0));
  const int use_tsy = (int)std::floor(opt_real("use_tsyganenko", 0.0));
  const int use_igrf = (int)std::floor(opt_real("use_igrf", 0.0));
  const int device = (int)std::floor(opt_real("device", -1.0));
  const int threads = (int)std::floor(opt_real("threads", 0.0));
  long long batch = (long long)std::floor(opt_real("batch", 0.0));
  if (batch == 0) {
    // rows come back to host memory before they are written: bound the worst case (every ray runs to maxsteps)
    const double ram = (double)sysconf(_SC_PHYS_PAGES) * (double)sysconf(_SC_PAGE_SIZE);
    const double per_ray = ((double)p.maxsteps / p.outputper + 2.0) * SRT_ROW_DOUBLES * sizeof(double);
    batch = (long long)(0.25 * ram / per_ray);
    if (batch < 1024) batch = 1024;
    if (batch > 4000000) batch = 4000000;
  }
  const bool verbose = opt_real("verbose", 0.0) != 0.0;
  if (batch < 1) die(2, "--batch must be >= 1");

  // one call per batch of the ray file uses every GPU of the box unless --device names one
  int rc = device >= 0 ? srt_init(device) : srt_init_devices(nullptr, 0);
  if (rc) die(1, std::string("srt_init: ") + srt_strerror(rc));
  srt_model* model = nullptr;
  if (modelnum == 1) {
    if (!getopt_named("ngo_configfile", modelfile)) die(2, "--ngo_configfile is required for --modelnum=1");
    rc = srt_model_ngo(&model, modelfile.c_str(), yearday, msec, use_igrf, use_tsy);
  } else if (modelnum == 3) {
    if (!getopt_named("interp_interpfile", modelfile)) die(2, "--interp_interpfile is required for --modelnum=3");
    rc = srt_model_grid(&model, modelfile.c_str(), yearday, msec, use_igrf, use_tsy);
  } else if (modelnum == 4) {  // driver:617-770 (scaled is fixed at 0 there, :722)
    if (!getopt_named("interp_interpfile", modelfile)) die(2, "--interp_interpfile is required for --modelnum=4");
    const double window_scale = need_real("scattered_interp_window_scale");
    const int order = (int)std::floor(need_real("scattered_interp_order"));
    const int exact = (int)std::floor(need_real("scattered_interp_exact"));
    const double local_window_scale = need_real("scattered_interp_local_window_scale");
    rc = srt_model_scattered(&model, modelfile.c_str(), yearday, msec, use_igrf, use_tsy, window_scale, order, exact, 0, local_window_scale);
  } else {
    die(3, "--modelnum=" + std::to_string(modelnum) + " is not available on the GPU path (1 = Ngo, 3 = gridded, 4 = scattered samples)");
  }
  if (!rc && use_tsy == 1) {  // the parameters the driver reads into the adapter's state (raytracer_driver.f95:291-340)
    static const char* names[10] = {"tsyganenko_Pdyn", "tsyganenko_Dst", "tsyganenko_ByIMF", "tsyganenko_BzIMF", "tsyganenko_W1",
                                    "tsyganenko_W2",   "tsyganenko_W3",  "tsyganenko_W4",    "tsyganenko_W5",    "tsyganenko_W6"};
    double parmod[10];
    for (int i = 0; i < 10; ++i) parmod[i] = opt_real(names[i], 0.0);
    rc = srt_model_tsyganenko(model, parmod);
  }
  if (rc) die(1, std::string("model set-up (") + modelfile + "): " + srt_strerror(rc) + (rc == SRT_E_CUDA ? std::string(": ") + srt_last_cuda_error() : ""));

  double *pos0 = nullptr, *dir0 = nullptr, *w0 = nullptr;
  int64_t nrays = 0;
  rc = srt_read_inputrays(inputraysfile.c_str(), &pos0, &dir0, &w0, &nrays);  // open(..., status="old"): a missing file is fatal
  if (rc) die(1, "cannot read " + inputraysfile + ": " + srt_strerror(rc));
  {  // open(unit=outfile, status="replace")
    FILE* f = fopen(outputfile.c_str(), "wb");
    if (!f) die(1, "cannot create " + outputfile);
    fclose(f);
  }
  using clk = std::chrono::steady_clock;
  double t_trace = 0.0, t_write = 0.0;
  long long steps = 0, rows = 0;
  for (int64_t first = 0; first < nrays; first += batch) {
    const int64_t n = (nrays - first < batch) ? nrays - first : batch;
    srt_result* res = nullptr;
    auto t0 = clk::now();
    rc = srt_trace_batch(model, n, pos0 + 3 * first, dir0 + 3 * first, w0 + first, &p, &res);
    if (rc) die(1, std::string("srt_trace_batch: ") + srt_strerror(rc) + (rc == SRT_E_CUDA ? std::string(": ") + srt_last_cuda_error() : ""));
    auto t1 = clk::now();
    rc = srt_result_write_text(res, model, w0 + first, outputfile.c_str(), first + 1, 1, threads);
    if (rc) die(1, "cannot write " + outputfile + ": " + srt_strerror(rc));
    auto t2 = clk::now();
    t_trace += std::chrono::duration<double>(t1 - t0).count();
    t_write += std::chrono::duration<double>(t2 - t1).count();
    steps += srt_result_accepted_steps(res);
    rows += srt_result_total_rows(res);
    srt_result_free(res);
  }
  if (verbose)
    fprintf(stderr, "raytracer: %lld rays, %lld accepted steps, %lld rows written; trace %.3f s, write %.3f s\n", (long long)nrays, steps, rows,
            t_trace, t_write);
  srt_free(pos0);
  srt_free(dir0);
  srt_free(w0);
  srt_model_free(model);
  return 0;
}

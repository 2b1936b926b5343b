// tools/shadow_predict.cpp -- offline evaluation of the team kernel's shadow predictor (srt_team.cuh) on one config-2 ray: the sequence of attempts from the
// host build of the device physics, which failures a (primary, shadow) pair covers, and the attempt-times left on the critical path.
//   g++ -O2 -std=c++17 -ffp-contract=off -Wno-attributes -o tools/_bin/shadow_predict tools/shadow_predict.cpp
//   tools/_bin/shadow_predict tests/data/newray_default.in x y z omega
#include "../stanford_raytracer_b200/csrc/srt_physics.h"
#include "../stanford_raytracer_b200/csrc/srt_igrf.h"
#include "../stanford_raytracer_b200/csrc/srt_tricubic.h"
#include "../stanford_raytracer_b200/csrc/srt_host_setup.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
int main(int argc, char** argv) {
  SrtModel* M = new SrtModel(); memset(M, 0, sizeof *M);
  if (srt_host::load_ngo_config(argv[1], *M)) return 1;
  srt_host::set_time(*M, 2010001, 0);
  double pos0[3] = {atof(argv[2]), atof(argv[3]), atof(argv[4])}, dir0[3] = {0, 0, 0}, w0 = atof(argv[5]);
  SrtRunParams P = {1e-3, 0.1, 5e-4, 6371.2e3 + 100e3, 10.0, 20000, 2, 0, 1};
  SrtRayState st; double row[SRT_ROW_DOUBLES];
  srt_ray_init<0, 0>(SrtModelRef{M}, P, pos0, dir0, w0, st, row);
  long acc = 0, fail_err = 0, fail_cone = 0, hit = 0, seq_len = 0;
  int last_fail = 0; long miss_kind = 0, miss_dt = 0, chain = 0;
  double prev_dt = st.dt; int prev_lrd = st.lastrefinedown; bool prev_failed = false; double guess_dt = 0; int guess_kind = 0;
  std::vector<char> outcome;  // per attempt: 0 accepted, 1 error rejection, 2 cone refinement
  long critical = 0; long chains[8] = {0}; int run = 0; long first_miss_dt = 0; long growfail = 0;  // attempt-times on the critical path with shadows
  for (;;) {
    const double dt0 = st.dt; const int lrd0 = st.lastrefinedown; const long nrej0 = st.nrej, ncone0 = st.ncone;
    // the shadow's preparation for THIS attempt's failure (made from the state at its start)
    guess_kind = last_fail == 1 ? 1 : 2;
    if (guess_kind == 1) guess_dt = 0.8 * dt0; else { bool grow = lrd0 == 0 && dt0 * 1.25 < P.dtmax; guess_dt = grow ? (dt0 * 1.25) / 2.0 : dt0 / 2.0; }
    int r = srt_ray_attempt<0, 0>(SrtModelRef{M}, P, st, row);
    if (r == SRT_ATTEMPT_DONE) break;
    if (prev_failed && chain == 1) { /* this attempt was the shadow's: it cost no time */ } else critical++;
    if (r == SRT_ATTEMPT_ACCEPTED) { outcome.push_back(0); acc++; prev_failed = false; chain = 0; chains[run < 7 ? run : 7]++; run = 0; continue; }
    run++;
    const int kind = st.nrej > nrej0 ? 1 : 2;
    outcome.push_back((char)kind);
    (kind == 1 ? fail_err : fail_cone)++;
    // a failure of the primary is covered if the shadow prepared this kind and this dt; the attempt after it then is the shadow's.
    // A failure inside a shadow's own attempt (chain >= 1) is not covered: the pair re-synchronises first.
    if (run == 1 && guess_dt != st.dt && (st.nrej > nrej0 ? 1 : 2) == guess_kind) first_miss_dt++;
    bool covered = (!prev_failed || chain != 1) && kind == guess_kind && guess_dt == st.dt;
    if (covered) { hit++; chain = 1; } else { if (kind != guess_kind) miss_kind++; else miss_dt++; chain = 0; }
    prev_failed = covered;
    last_fail = kind;
  }
  printf("failure runs before an accept: 0:%ld 1:%ld 2:%ld 3:%ld 4:%ld 5+:%ld ; first failure of a run with the kind right and dt wrong: %ld\n", chains[0], chains[1], chains[2], chains[3], chains[4], chains[5]+chains[6]+chains[7], first_miss_dt);
  printf("accepted %ld err-fail %ld cone-fail %ld | hits %ld miss(kind) %ld miss(dt) %ld | attempt-times on the critical path %ld (without shadows %ld)\n", acc, fail_err, fail_cone, hit, miss_kind, miss_dt, critical, acc + fail_err + fail_cone);
  // deeper speculation, idealised: k teams per ray run the attempt and the k-1 retries that would follow a run of failures of the
  // kind last seen (0.8^j dt after error rejections, dt / 2^j after cone refinements: both exact); a window ends at the first
  // accepted attempt, at a failure of the other kind, or after k attempts.  Attempt-times on the critical path = windows.
  printf("attempt-times on the critical path with k teams per ray (idealised):");
  for (int k : {1, 2, 3, 4, 6, 8}) {
    long windows = 0; int guess = 1;
    for (size_t i = 0; i < outcome.size();) {
      ++windows;
      int j = 0;
      for (;;) {
        const int o = outcome[i + j];
        ++j;
        if (o == 0) break;
        if (o != guess) { guess = o; break; }
        if (j == k || i + j >= outcome.size()) break;
      }
      i += j;
    }
    printf(" k=%d: %ld", k, windows);
  }
  printf("\n");
  return 0;
}

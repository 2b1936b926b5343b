        // ---- srt_team_est.inl: error estimate, re-projection, accept/reject (raytracer.f95:775-922).
        // Included inside srt_trace_team_kernel; lane 6 holds the plasma at est1, lane 7 at est2 (fixed step:
        // both at est2), lanes 0-5 at est2 +- d e_i (the next attempt's stage-1 stencil, speculative).
        const double dt_used = dt;
        SrtPlasmaDerived pc1, pc2;
        team_bcast_derived(T, pd, 6, pc1);
        team_bcast_derived(T, pd, 7, pc2);
        bool retry = false;
        if (adaptive) {
          double sa = 0.0, sb = 0.0;
#pragma unroll
          for (int i = 0; i < 3; ++i) sa = sa + srt_fabs(e1[3 + i] - e2[3 + i]);
#pragma unroll
          for (int i = 0; i < 3; ++i) sb = sb + srt_fabs(e2[3 + i]);
          double err = srt_div(sa, sb);
          if (wl != 0.0) {
            const int set = r >> 2, q = r & 3;
            const double* kset = set ? (e2 + 3) : (e1 + 3);
            double n1[3], n2[3], dk1, dk2;
            team_dfdk_n(kset, wl, q >> 1, q & 1, n1, dk1);  // axes 0 and 1 on the four lanes of the set
            SrtPlasmaDerived ps = pc1;
            if (set) ps = pc2;
            SrtStix sw;
            srt_stix(ps, wl, sw);
            const double F1 = srt_F(n1, wl, ps, sw);
            team_dfdk_n(kset, wl, 2, q & 1, n2, dk2);       // axis 2 on the first two lanes of the set
            const double F2 = srt_F(n2, wl, ps, sw);
            double g[2][3];
#pragma unroll
            for (int s2 = 0; s2 < 2; ++s2) {
              const double* ks = s2 ? (e2 + 3) : (e1 + 3);
              g[s2][0] = srt_div(team_get(T, F1, s2 * 4 + 0) - team_get(T, F1, s2 * 4 + 1), srt_fd_step(1.0e-8, ks[0])) / 2.0;
              g[s2][1] = srt_div(team_get(T, F1, s2 * 4 + 2) - team_get(T, F1, s2 * 4 + 3), srt_fd_step(1.0e-8, ks[1])) / 2.0;
              g[s2][2] = srt_div(team_get(T, F2, s2 * 4 + 0) - team_get(T, F2, s2 * 4 + 1), srt_fd_step(1.0e-8, ks[2])) / 2.0;
            }
            double sc2 = 0.0, sd2 = 0.0;
#pragma unroll
            for (int i = 0; i < 3; ++i) sc2 = sc2 + srt_fabs(g[0][i] - g[1][i]);
#pragma unroll
            for (int i = 0; i < 3; ++i) sd2 = sd2 + srt_fabs(g[1][i]);
            const double eg = srt_div(sc2, sd2);
            if (eg > err || err != err) err = eg;  // MAX() as gfortran <= 8 expands it
          } else {
            if (err != err) err = srt_nan();  // Q1
          }
          err_small = (err < P.maxerr / 100.0) ? 1 : 0;
          if (err > P.maxerr) {
            dt = 0.8 * dt;
            lastrefinedown = 1;
            nrej++;
            retry = true;
          } else if (lastrefinedown == 0 && err < P.maxerr / 100.0 && dt * 1.25 < P.dtmax) {
            dt = dt * 1.25;
            lastrefinedown = 0;
          }
        }
        int outcome = retry ? 1 : 0;  // 0: goes on to the accept side, 1: rejected by the error test, 2: cone-refined, 3: cone exit (fixed step)
        double knew[3] = {0.0, 0.0, 0.0};
        if (!retry) {
          wl = w;  // raytracer.f95:821, before the cone test (see srt_ray_attempt)
          // re-project |k| at est2 (lane 7 holds the plasma there and its Stix set at w)
          SrtCplx ksel = srt_solve_root(e2 + 3, w, pd, sx, P.root);
          ksel.re = team_get(T, ksel.re, 7);
          ksel.im = team_get(T, ksel.im, 7);
          double kk = 0.0;
#pragma unroll
          for (int i = 0; i < 3; ++i) kk = kk + (e2[3 + i] * e2[3 + i] - (-0.0) * 0.0);
          SrtCplx nrm;
          srt_csqrt(kk, 0.0, nrm.re, nrm.im);
          double imsum = 0.0;
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            SrtCplx ki = {e2[3 + i], 0.0};
            SrtCplx kn = srt_cmul(ksel, srt_cdiv(ki, nrm));
            knew[i] = kn.re;
            imsum = imsum + kn.im * kn.im;
          }
          if (imsum > 0.0) {
            if (adaptive) {
              dt = dt / 2.0;
              lastrefinedown = 1;
              ncone++;
              retry = true;
              outcome = 2;
            } else {
              cone_exit = 1;
              stopcond = 0;
              finish_now = true;
              phase = SRT_PH_STAGE;
              attempt_start = true;
              outcome = 3;
            }
          }
        }
        // ---- shadow groups (srt_team.cuh): are the retries the ray now needs the attempts its shadows have just finished?  The shadow
        // of depth q is valid if every team before it failed the way the shadows prepared for and arrived at exactly its dt.
        bool i_live = true, valid_me = true;
        live_is_S = 0;
#if defined(SRT_TEAM_BLOCK_SYNC)
        if (outcome == 1 || outcome == 2) last_fail = outcome;
        if (group > 1 && shadow_ready) {
          const int prepared = __shfl_sync(gmask, shadow_kind, gbase + 8);
          int live_q = 0;
          bool chain = true;
          for (int q = 1; q < group; ++q) {
            const int out_prev = __shfl_sync(gmask, outcome, gbase + 8 * (q - 1));
            const long long dt_next_prev = __shfl_sync(gmask, __double_as_longlong(dt), gbase + 8 * (q - 1));
            const long long dt_used_q = __shfl_sync(gmask, __double_as_longlong(dt_used), gbase + 8 * q);
            chain = chain && out_prev == prepared && dt_next_prev == dt_used_q && !(__longlong_as_double(dt_used_q) < (double)1e-14f);
            if (chain) live_q = q;
          }
          live_is_S = live_q;
          i_live = gq == live_q;
          valid_me = gq <= live_q;
          if (gq >= 1 && valid_me && r == 0) ++my_spec;
          // the live attempt's failure, else the one before it in the chain, is the ray's most recent one
          if (i_live && gq >= 1 && outcome != 1 && outcome != 2) last_fail = prepared;
        }
#endif
        {
          if (outcome == 0 && i_live) {
            // ---- accept (raytracer.f95:909-922)
            xc = (comp == 0) ? e2[0] : (comp == 1) ? e2[1] : (comp == 2) ? e2[2] : (comp == 3) ? knew[0] : (comp == 4) ? knew[1] : knew[2];
            lastrefinedown = 0;
            t = t + dt_used;
            nstep = nstep + 1;
            // dF/dk and dF/dw at the new point: est2's plasma, spread over the 8 lanes
            const SrtPlasmaDerived& pc = pc2;
            const double wr2 = (r == 6) ? w + dw : (r == 7 ? w - dw : w);
            SrtStix s2;
            srt_stix(pc, wr2, s2);
            double nT[3], dkT;
            team_dfdk_n(knew, w, axis, minus, nT, dkT);
            if (r >= 6) {
#pragma unroll
              for (int j = 0; j < 3; ++j) nT[j] = srt_div(knew[j] * C, wr2);
            }
            const double FT = srt_F(nT, wr2, pc, s2);
            dfdw_rec = (team_get(T, FT, 6) - team_get(T, FT, 7)) / dw / 2.0;
            double dfdk_new[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) dfdk_new[i] = srt_div(team_get(T, FT, 2 * i) - team_get(T, FT, 2 * i + 1), srt_fd_step(1.0e-8, knew[i])) / 2.0;
            dfdk_c = (comp % 3 == 0) ? dfdk_new[0] : (comp % 3 == 1) ? dfdk_new[1] : dfdk_new[2];
            // the recorded row (raytracer.f95:937-986)
            double row[SRT_ROW_DOUBLES];
            double nn[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) nn[i] = srt_div(knew[i] * C, w);
            const double n2 = srt_dot3(nn, nn);
            row[0] = t;
#pragma unroll
            for (int i = 0; i < 3; ++i) {
              row[1 + i] = e2[i];
              row[4 + i] = srt_div(nn[i], n2);
              row[7 + i] = -srt_div(srt_div(dfdk_new[i], dfdw_rec), C);
              row[10 + i] = nn[i];
              row[13 + i] = team_get(T, pl.B0[i], 7);
            }
#pragma unroll
            for (int i = 0; i < SRT_NSPEC; ++i) row[16 + i] = team_get(T, pl.Ns[i], 7);
            vg[0] = row[7]; vg[1] = row[8]; vg[2] = row[9];
            if (r == 0) {
              ++my_accepted;
              ++my_attempts;
              double* sm = B.summary + slot * SRT_SUMMARY_WORDS;
#pragma unroll
              for (int i = 0; i < SRT_ROW_DOUBLES; ++i) sm[i] = row[i];
              const int rowidx = nstep - 1;
              if (rowidx % outputper == 0) {
                srt_log_row(B, slot, rowidx / outputper, row, dropped);
                if (!dropped) ++nlog;
              }
            }
            dropped = __shfl_sync(T.tmask, (int)dropped, T.tbase) != 0;
            // an all-NaN state is stepped on to maxsteps / tmax by the reference: srt_nan_fastforward (srt_kernels.cuh)
            if (srt_all_nan6(e2, knew)) {
              if (r == 0) {
                const unsigned nff = srt_nan_fastforward(P, B, slot, outputper, row, t, dt, nstep, nlog, dropped);
                my_accepted += nff; my_attempts += nff; my_ff += nff;
                B.summary[slot * SRT_SUMMARY_WORDS] = t;  // the last recorded row lives in the summary
              }
              t = __shfl_sync(T.tmask, t, T.tbase);
              nstep = __shfl_sync(T.tmask, nstep, T.tbase);
              dropped = __shfl_sync(T.tmask, (int)dropped, T.tbase) != 0;
            }
            // stage 1 of the next attempt (raytracer_evalrhs at the new point): dF/dk, dF/dw just recorded,
            // dF/dx from the six displaced points lanes 0-5 evaluated in this round
            {
              const double FA = srt_F(nn, w, pd, sx);
              const int j2 = 2 * (comp % 3);
              const double Fp = team_get(T, FA, j2), Fn = team_get(T, FA, j2 + 1);
              const double ddx = team_get(T, dx_axis, j2);
              if (comp < 3) rhs1c = -srt_div(dfdk_c, dfdw_rec);
              else rhs1c = srt_div(srt_div(Fp - Fn, ddx) / 2.0, dfdw_rec);
            }
            phase = SRT_PH_STAGE;  // the attempt-start block sets kk1, tmpc and stage = 2 from rhs1c
            attempt_start = true;
          }
        }
        if (retry) {
          if (r == 0 && valid_me) ++my_attempts;  // a primary's failed attempt counts; a shadow's only when it is one of the ray's attempts
          phase = SRT_PH_STAGE;
          attempt_start = true;
        }
        did_est = 1;

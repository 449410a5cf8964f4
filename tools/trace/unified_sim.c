/* tools/trace/unified_sim.c — DESIGN ANALYSIS TOOL: one persistent warp = 8 register slots (a 4-lane sub-group per patch,
 * template + tap window in registers, asynchronous progress, lock-step sample phases) + 32 lane slots (one lane per
 * patch, memoised replay).  Fresh patches enter through a register slot, move to a lane slot when they have not
 * missed for G passes, and come back to a register slot for their later misses.  Counts warp instructions per patch.
 * gcc -O2 -o /tmp/unified_sim tools/trace/unified_sim.c */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#define main p1main
#include "p1_sim.c"
#undef main

typedef struct { int pid, ip, lastmiss, nm, state, lane; cache_t c; } slot_t;   /* state: 0 free 1 run 2 miss(pending sample) 3 wait(lane slot waiting for a register slot) 4 reserved */

int main(int argc, char** argv) {
  load(argv[1]);
  int MINOLD = argc > 2 ? atoi(argv[2]) : 24;
  const double C_SAMPLE = 310, C_RSTEP = 30, C_RPASS = 13, C_BIND = 260, C_MOVE = 40, C_CHUNK = 6 * 11 + 22, C_OLDCTL = 40;
  int Gs[] = {4, 8, 16}; int Rs[] = {4, 8}; int K0s[] = {12, 16, 24};
  for (int gi = 0; gi < 3; ++gi) for (int ri = 1; ri < 2; ++ri) for (int ki = 0; ki < 3; ki += 2) {
    int G = Gs[gi], R = Rs[ri], K0 = K0s[ki];
    double cost = 0; long nsample = 0, nused = 0, nbind_ev = 0, nbind = 0, nchunk = 0, nrstep = 0; double lanes_active = 0;
    long fresh_done = 0; double c_sample = 0, c_bind = 0, c_replay = 0, c_old = 0;
    int nwarps = 32; long per = (NP + nwarps - 1) / nwarps;
    for (int wi = 0; wi < nwarps; ++wi) {
      long next = wi * per, q1 = next + per > NP ? NP : next + per;
      static slot_t reg[8], lane[32];
      for (int j = 0; j < 8; ++j) reg[j].state = 0;
      for (int j = 0; j < 32; ++j) lane[j].state = 0;
      for (;;) {
        /* A. bind: waiting lane patches first, then fresh ones (a fresh patch needs a lane slot to retire to) */
        int nfree = 0; for (int j = 0; j < 8; ++j) if (!reg[j].state) ++nfree;
        int nwait = 0, lfree = 0; for (int j = 0; j < 32; ++j) { if (lane[j].state == 3) ++nwait; else if (!lane[j].state) ++lfree; }
        int nreg_fresh = 0; for (int j = 0; j < 8; ++j) if (reg[j].state && reg[j].lane < 0) ++nreg_fresh;
        int avail_fresh = (int)(q1 - next); if (avail_fresh > lfree - nreg_fresh) avail_fresh = lfree - nreg_fresh; if (avail_fresh < 0) avail_fresh = 0;
        int nreg_busy = 8 - nfree;
        int cand = nwait + avail_fresh;
        if (nfree > 0 && cand > 0 && (nfree >= 2 || nreg_busy == 0 || cand >= nfree)) {
          int b = 0, nfresh_b = 0;
          for (int j = 0; j < 8; ++j) if (!reg[j].state) {
            int l = -1; for (int q = 0; q < 32; ++q) if (lane[q].state == 3) { l = q; break; }
            if (l >= 0) { reg[j] = lane[l]; reg[j].state = 2; reg[j].lane = l; lane[l].state = 4; ++b; }
            else if (avail_fresh > 0) { --avail_fresh; reg[j].pid = (int)next++; reg[j].ip = 0; reg[j].lastmiss = 0; reg[j].nm = 0; reg[j].state = 1; reg[j].lane = -1; cache_clear(&reg[j].c); ++b; ++nfresh_b; }
          }
          if (b) { double cb = nfresh_b ? C_BIND : 130; cost += cb; c_bind += cb; ++nbind_ev; nbind += b; }
        }
        int any_reg = 0; for (int j = 0; j < 8; ++j) if (reg[j].state) any_reg = 1;
        int any_lane = 0; for (int j = 0; j < 32; ++j) if (lane[j].state == 1) any_lane = 1;
        if (!any_reg && !any_lane) { int w = 0; for (int j = 0; j < 32; ++j) if (lane[j].state == 3) w = 1; if (!w && next >= q1) break; if (!w && lfree == 0) break; continue; }
        /* B. young step */
        if (any_reg) {
          int maxrun = 0;
          for (int j = 0; j < 8; ++j) if (reg[j].state == 1) {
            const patch_t* p = &P[reg[j].pid]; int run = 0;
            for (;;) {
              if (reg[j].ip >= p->n) { if (reg[j].lane >= 0) lane[reg[j].lane].state = 0; reg[j].state = 0; ++fresh_done; break; }
              if (reg[j].ip >= K0 && reg[j].ip - reg[j].lastmiss >= G) {   /* mature: to a lane slot */
                int l = reg[j].lane; if (l < 0) for (int q = 0; q < 32; ++q) if (!lane[q].state) { l = q; break; }
                if (l >= 0) { lane[l] = reg[j]; lane[l].state = 1; lane[l].lane = l; reg[j].state = 0; cost += C_MOVE / 2.0; break; }
              }
              if (!cache_lookup(&reg[j].c, p->tr[reg[j].ip])) { reg[j].state = 2; break; }
              if (run == R) break;
              ++run; ++reg[j].ip;
            }
            if (run > maxrun) maxrun = run;
          }
          cost += C_RSTEP + C_RPASS * maxrun; c_replay += C_RSTEP + C_RPASS * maxrun; ++nrstep;
          int npend = 0; for (int j = 0; j < 8; ++j) if (reg[j].state == 2) ++npend;
          if (npend) {
            cost += C_SAMPLE; c_sample += C_SAMPLE; ++nsample;
            for (int j = 0; j < 8; ++j) if (reg[j].state == 2) { cache_insert(&reg[j].c, P[reg[j].pid].tr[reg[j].ip]); reg[j].lastmiss = reg[j].ip; ++reg[j].nm; ++reg[j].ip; reg[j].state = 1; ++nused; }
          }
        }
        /* C. old chunks: up to 4 chunks of 6 passes per loop turn */
        int nrun0 = 0; for (int j = 0; j < 32; ++j) if (lane[j].state == 1) ++nrun0;
        if (any_lane && (nrun0 >= MINOLD || !any_reg)) {
          for (int ch = 0; ch < 4; ++ch) {
            int nr = 0;
            for (int j = 0; j < 32; ++j) if (lane[j].state == 1) {
              const patch_t* p = &P[lane[j].pid];
              for (int it = 0; it < 6; ++it) {
                if (lane[j].ip >= p->n) { lane[j].state = 0; ++fresh_done; break; }
                if (!cache_lookup(&lane[j].c, p->tr[lane[j].ip])) { lane[j].state = 3; break; }
                ++lane[j].ip;
              }
              if (lane[j].state == 1) ++nr;
            }
            cost += C_CHUNK; c_old += C_CHUNK; ++nchunk; lanes_active += nr;
            if (nr < MINOLD) break;
          }
          cost += C_OLDCTL; c_old += C_OLDCTL;
        }
      }
    }
    printf("G %2d R %d K0 %2d: %.0f wi/patch = sample %.0f (phases/patch %.2f util %.2f) + bind %.0f (%.2f binds/patch, %.2f per event) + young replay %.0f + old %.0f (lanes running/chunk %.1f) | samples/patch %.2f done %ld\n",
           G, R, K0, cost / NP, c_sample / NP, (double)nsample / NP, (double)nused / (8.0 * nsample), c_bind / NP, (double)nbind / NP, (double)nbind / nbind_ev, c_replay / NP, c_old / NP, lanes_active / nchunk, (double)nused / NP, fresh_done);
  }
  return 0;
}

/* tools/trace/p1_sim.c — DESIGN ANALYSIS TOOL: replays the position traces written by orbit_stats.c through
 * (a) the round-1 structure (lock-step k_converge batches of 8 + memoised walk) and
 * (b) an asynchronous phase 1: 8 sub-groups per warp, each with its own patch, pass counter and 16-entry cache;
 *     a warp step = replay phase (every sub-group advances through cache hits, at most R passes) + one sample
 *     phase for the sub-groups with a pending first-seen position; a patch leaves for the walk when mature.
 * Costs are warp instructions (measured/estimated constants below).  gcc -O2 -o /tmp/p1_sim tools/trace/p1_sim.c */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
typedef struct { unsigned x, y; } pos_t;
typedef struct { int n; int reset; pos_t* tr; } patch_t;

typedef struct { pos_t e[16]; unsigned fifo; } cache_t;
static int slot_of(unsigned bx, unsigned by) { return (int)((5u * bx + by) & 15u); }
static int way_xor(int w) { return ((w & 1) << 1) | ((w >> 1) << 3); }
static void cache_clear(cache_t* c) { memset(c->e, 0xff, sizeof c->e); c->fifo = 0; }
/* returns 0 miss, 1 hit in predicted way, 2 hit in another way */
static int cache_lookup(const cache_t* c, pos_t p) {
  int s = slot_of(p.x, p.y);
  if (c->e[s].x == p.x && c->e[s].y == p.y) return 1;
  for (int w = 1; w < 4; ++w) { int q = s ^ way_xor(w); if (c->e[q].x == p.x && c->e[q].y == p.y) return 2; }
  return 0;
}
static void cache_insert(cache_t* c, pos_t p) {
  int s = slot_of(p.x, p.y);
  pos_t old = c->e[s];
  if (old.x != 0xffffffffu) {
    int set = ((s >> 1) & 2) | (s & 1);
    int way = (c->fifo >> (2 * set)) & 3;
    c->fifo = (c->fifo & ~(3u << (2 * set))) | (((way + 1) & 3) << (2 * set));
    int q = (s & 5) | way_xor(way);
    if (q != s) c->e[q] = old;
  }
  c->e[s] = p;
}

static patch_t* P; static int NP;
static void load(const char* fn) {
  FILE* f = fopen(fn, "rb"); if (!f) { perror(fn); exit(1); }
  int cap = 1 << 16; P = malloc(sizeof(patch_t) * cap); NP = 0; int n;
  while (fread(&n, 4, 1, f) == 1) {
    if (NP == cap) { cap *= 2; P = realloc(P, sizeof(patch_t) * cap); }
    int m = n < 0 ? -n : n;
    P[NP].n = m; P[NP].reset = n <= 0; P[NP].tr = malloc(sizeof(pos_t) * (m ? m : 1));
    if (m && fread(P[NP].tr, sizeof(pos_t), m, f) != (size_t)m) exit(2);
    ++NP;
  }
  fclose(f);
}

/* walk model: from pass i0 with cache c: count misses / non-predicted hits to the end; straggler rule: handed off after
 * `handoff` misses (remaining passes all sampled in k_stragglers) */
typedef struct { long replay, miss, probe, strag_pass, strag_n; } walk_t;
static void walk(const patch_t* p, int i0, cache_t* c, int handoff, walk_t* w) {
  int misses = 0, streak = 0;
  for (int i = i0; i < p->n; ++i) {
    int h = cache_lookup(c, p->tr[i]);
    if (h == 0) {
      ++misses; ++streak;
      if (misses >= handoff || streak >= 64) { w->strag_pass += p->n - i; w->strag_n++; return; }
      cache_insert(c, p->tr[i]); w->miss++;
    } else { streak = 0; w->replay++; if (h == 2) w->probe++; }
  }
}

int main(int argc, char** argv) {
  if (argc < 2) return 1;
  load(argv[1]);
  int handoff = argc > 2 ? atoi(argv[2]) : 384;
  printf("%d patches\n", NP);
  /* ---- (a) round-1 structure ------------------------------------------------------------------------------- */
  {
    long samples1 = 0, phases = 0; walk_t w = {0};
    for (int b = 0; b + 8 <= NP; b += 8) {
      cache_t c[8]; int settled[8] = {0}, act[8]; int pass;
      for (int j = 0; j < 8; ++j) { cache_clear(&c[j]); act[j] = P[b + j].n > 0; }
      for (pass = 0; pass < 24; ++pass) {
        int any = 0, unsettled = 0;
        for (int j = 0; j < 8; ++j) { if (act[j] && pass >= P[b + j].n) act[j] = 0; any |= act[j]; if (act[j] && !settled[j]) ++unsettled; }
        if (!any) break;
        if (pass >= 12 && unsettled <= 2) break;
        ++phases;
        for (int j = 0; j < 8; ++j) if (act[j]) {
          ++samples1;
          if (pass >= 8) { if (cache_lookup(&c[j], P[b + j].tr[pass])) settled[j] = 1; else cache_insert(&c[j], P[b + j].tr[pass]); }
        }
      }
      for (int j = 0; j < 8; ++j) if (act[j]) walk(&P[b + j], pass, &c[j], handoff, &w);
    }
    printf("(a) round 1: k_converge phases/batch %.2f, samples/patch %.2f | walk: misses/patch %.2f, other-way hits %.2f, replays %.1f | stragglers %.3f %%, their passes/patch(avg over all) %.2f | total sampled/patch %.2f\n",
           8.0 * phases / NP, (double)samples1 / NP, (double)w.miss / NP, (double)w.probe / NP, (double)w.replay / NP, 100.0 * w.strag_n / NP, (double)w.strag_pass / NP,
           (double)(samples1 + w.miss + w.strag_pass) / NP);
  }
  /* ---- (b) asynchronous phase 1 ---------------------------------------------------------------------------- */
  const double C_SAMPLE = 300, C_REPLAY = 18, C_REBIND = 380, C_STEP = 24;
  int Rs[] = {4, 8, 16, 32}; int Ks[] = {24, 32, 48, 64}; int smins[] = {1, 3, 5}; int rbs[] = {1, 2, 3, 4};
  for (int ki = 0; ki < 4; ++ki) for (int ri = 0; ri < 4; ++ri) for (int si = 0; si < 3; ++si) for (int bi = 0; bi < 4; ++bi) {
    int K = Ks[ki], R = Rs[ri], smin = smins[si], rbmin = rbs[bi];
    /* maturity rule: pass >= K and no miss during the last G passes */
    int G = 8;
    long next = 0; double cost = 0; long nsample_phase = 0, nsample_used = 0, nrebind = 0, nreplay_pass = 0, early_strag = 0; walk_t w = {0};
    long s1 = 0;
    int nwarps = 64;   /* independent warps sharing the queue: simulate sequentially in chunks */
    long per = (NP + nwarps - 1) / nwarps;
    for (int wi = 0; wi < nwarps; ++wi) {
      long q0 = wi * per, q1 = q0 + per > NP ? NP : q0 + per; next = q0;
      int pid[8], ip[8], pend[8], lastmiss[8], nm[8]; cache_t c[8];
      for (int j = 0; j < 8; ++j) pid[j] = -1;
      for (;;) {
        int nfree = 0, nact = 0;
        for (int j = 0; j < 8; ++j) if (pid[j] < 0) ++nfree; else ++nact;
        if (next < q1 && (nfree >= rbmin || nact == 0)) {
          for (int j = 0; j < 8 && next < q1; ++j) if (pid[j] < 0) { pid[j] = (int)next++; ip[j] = 0; pend[j] = 0; lastmiss[j] = 0; nm[j] = 0; cache_clear(&c[j]); }
          cost += C_REBIND; ++nrebind;
        } else if (nact == 0) break;
        /* replay phase */
        int maxrun = 0;
        for (int j = 0; j < 8; ++j) if (pid[j] >= 0 && !pend[j]) {
          const patch_t* p = &P[pid[j]]; int run = 0;
          for (;;) {
            if (ip[j] >= p->n) { pid[j] = -1; break; }                                  /* finished (reset or budget) */
            if (ip[j] >= K && ip[j] - lastmiss[j] >= G) { walk(p, ip[j], &c[j], handoff, &w); pid[j] = -1; break; }   /* mature: to the walk */
            if (nm[j] >= 48 && ip[j] - nm[j] < 16) { w.strag_pass += p->n - ip[j]; w.strag_n++; ++early_strag; pid[j] = -1; break; }   /* misses on (nearly) every pass: straggler */
            if (!cache_lookup(&c[j], p->tr[ip[j]])) { pend[j] = 1; break; }
            if (run == R) break;
            ++run; ++ip[j];
          }
          if (run > maxrun) maxrun = run;
        }
        cost += C_STEP + C_REPLAY * maxrun; nreplay_pass += maxrun;
        /* sample phase */
        int npend = 0, canrun = 0;
        for (int j = 0; j < 8; ++j) if (pid[j] >= 0) { if (pend[j]) ++npend; else ++canrun; }
        if (npend >= smin || (npend > 0 && canrun == 0)) {
          cost += C_SAMPLE; ++nsample_phase;
          for (int j = 0; j < 8; ++j) if (pid[j] >= 0 && pend[j]) { cache_insert(&c[j], P[pid[j]].tr[ip[j]]); lastmiss[j] = ip[j]; ++nm[j]; ++ip[j]; pend[j] = 0; ++nsample_used; ++s1; }
        }
      }
    }
    /* walk cost model (warp instructions per patch): fast loop + service + control, calibrated on the round-1 ncu numbers */
    double walk_cost = ((double)w.replay * 11 / 32 / 0.85 + (double)w.miss * 90 + (double)w.probe * 12) / NP + 120;
    double strag_cost = (double)w.strag_pass * C_SAMPLE / 8 / NP;
    printf("(b) K %2d R %2d smin %d rb %d: P1 %.0f wi/patch (sample phases/patch %.2f, util %.2f, rebinds/patch %.2f, replay passes/patch %.1f) | P1 samples %.2f walk misses %.2f probes %.2f strag %.3f%% strag passes %.2f | walk %.0f strag %.0f | TOTAL %.0f\n",
           K, R, smin, rbmin, cost / NP, (double)nsample_phase / NP, (double)nsample_used / (8.0 * nsample_phase), (double)nrebind / NP, (double)nreplay_pass / NP,
           (double)s1 / NP, (double)w.miss / NP, (double)w.probe / NP, 100.0 * w.strag_n / NP, (double)w.strag_pass / NP, walk_cost, strag_cost, cost / NP + walk_cost + strag_cost);
  }
  return 0;
}

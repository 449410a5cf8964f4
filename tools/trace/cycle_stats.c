/* tools/trace/cycle_stats.c — DESIGN ANALYSIS TOOL (not product, not oracle): is the per-patch iteration of the reference
 * EXACTLY periodic in its state u = (ux, uy)?  The update u <- u - delta(pos(u)) is a pure function of u's bit pattern,
 * so once a u repeats the orbit is periodic for good and the result after pass 1000 follows without running the passes.
 * Prints, per level dump: how many patches enter a cycle, when (mu + lambda = first repeat), the period lambda, and
 * the passes a walk would have to execute with (a) perfect detection, (b) Brent's checkpoint detection (two compares a pass).
 * Build: gcc -O2 -ffp-contract=off -o gpurun_out/trace/cycle_stats tools/trace/cycle_stats.c -lm
 * Input: a raw dump written by tools/trace/dump_level.py. */
#define main orbit_main
#include "orbit_stats.c"
#undef main

typedef struct { unsigned x, y; } ubits_t;

static int cmp_int(const void* a, const void* b) { return *(const int*)a - *(const int*)b; }

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: cycle_stats dump.bin [stride] [first_pass]\n"); return 1; }
  FILE* f = fopen(argv[1], "rb");
  int hdr[8];
  if (fread(hdr, 4, 8, f) != 8) return 1;
  int stride = argc > 2 ? atoi(argv[2]) : 1;
  int P0 = argc > 3 ? atoi(argv[3]) : 24;          /* the walk starts at this pass (k_converge ran the earlier ones) */
  level_t L; level_setup(&L, hdr[0], hdr[1], hdr[2], hdr[3], hdr[4], hdr[5], hdr[6], hdr[7]);
  size_t pn = (size_t)L.tmp_w * (L.height + 2 * L.pad);
  float *gx = malloc(4 * pn), *gy = malloc(4 * pn), *im = malloc(4 * pn);
  int has_prev; if (fread(&has_prev, 4, 1, f) != 1) return 1;
  float* prev = NULL;
  if (fread(gx, 4, pn, f) != pn || fread(gy, 4, pn, f) != pn || fread(im, 4, pn, f) != pn) return 1;
  if (has_prev) { size_t fn = (size_t)(L.width / 2) * (L.height / 2) * 2; prev = malloc(4 * fn); if (fread(prev, 4, fn, f) != fn) return 1; }
  fclose(f);
  int NI = L.iterations + 2;
  pos_t* tr = malloc(sizeof(pos_t) * NI); float* dtr = malloc(8 * NI);
  ubits_t* us = malloc(sizeof(ubits_t) * NI);
  float* buf = malloc(sizeof(float) * 4 * L.ps * L.ps);
  long nfull = 0, ncyc = 0;
  int* first_rep = malloc(sizeof(int) * 400000); int* lam = malloc(sizeof(int) * 400000); int nrec = 0;
  double exec_perfect = 0, exec_brent = 0, exec_now = 0, exec_brent_lin = 0;
  long lam1 = 0, lam_le4 = 0, lam_le16 = 0;
  enum { HS = 4096 };
  static int htab[HS];
  for (int x = 0, ip = 0; x < L.npw; ++x)
    for (int y = 0; y < L.nph; ++y, ++ip) {
      if (ip % stride) continue;
      float refx = (float)(x * L.steps + L.offw), refy = (float)(y * L.steps + L.offh), ix = 0, iy = 0;
      if (prev) { int fx = (int)floorf(refx / 2), fy = (int)floorf(refy / 2); int i = fy * (L.width / 2) + fx; ix = prev[2 * i] * 2; iy = prev[2 * i + 1] * 2; }
      int n = trace_patch(&L, gx, gy, im, refx, refy, ix, iy, tr, dtr, buf);
      if (n <= 0) continue;
      ++nfull;
      /* the state before pass k: u_0 = init, u_{k+1} = u_k - delta_k (fp32, as the loop computes it) */
      float ux = ix, uy = iy;
      for (int k = 0; k < n; ++k) { memcpy(&us[k].x, &ux, 4); memcpy(&us[k].y, &uy, 4); ux = ux - dtr[2 * k]; uy = uy - dtr[2 * k + 1]; }
      /* first exact repeat of a state */
      memset(htab, -1, sizeof htab);
      int j0 = -1, i0 = -1;
      for (int k = 0; k < n && j0 < 0; ++k) {
        unsigned h = (us[k].x * 2654435761u ^ us[k].y * 40503u) >> 8;
        for (int probe = 0; probe < HS; ++probe) {
          int* slot = &htab[(h + probe) & (HS - 1)];
          if (*slot < 0) { *slot = k; break; }
          if (us[*slot].x == us[k].x && us[*slot].y == us[k].y) { j0 = k; i0 = *slot; break; }
        }
      }
      const int total = n;   /* 1001 passes */
      exec_now += total - P0;
      if (j0 < 0) { exec_perfect += total - P0; exec_brent += total - P0; exec_brent_lin += total - P0; first_rep[nrec] = 100000; lam[nrec++] = 100000; continue; }
      ++ncyc;
      const int la = j0 - i0;
      first_rep[nrec] = j0; lam[nrec++] = la;
      lam1 += la == 1; lam_le4 += la <= 4; lam_le16 += la <= 16;
      /* (a) perfect: run to the first repeat (at least from P0), then (remaining mod lambda) more passes */
      { int det = j0 < P0 ? P0 : j0; int rem = total - det; exec_perfect += (det - P0) + rem % la; }
      /* (b) Brent from pass P0: checkpoint at k = P0, P0 + 1, P0 + 3, P0 + 7, ... ; detects when u_k == checkpoint */
      {
        int cp = P0, next = 1, det = -1, per = 0;
        for (int k = P0 + 1; k < total; ++k) {
          if (us[k].x == us[cp].x && us[k].y == us[cp].y) { det = k; per = k - cp; break; }
          if (k - cp == next) { cp = k; next *= 2; }
        }
        if (det < 0) exec_brent += total - P0; else exec_brent += (det - P0) + (total - det) % per;
      }
      /* (c) checkpoint renewed every 32 passes (detects periods <= 32 only) */
      {
        int cp = P0, det = -1, per = 0;
        for (int k = P0 + 1; k < total; ++k) {
          if (us[k].x == us[cp].x && us[k].y == us[cp].y) { det = k; per = k - cp; break; }
          if (k - cp == 32) cp = k;
        }
        if (det < 0) exec_brent_lin += total - P0; else exec_brent_lin += (det - P0) + (total - det) % per;
      }
    }
  printf("level scale %d: full patches %ld, exactly periodic in u within the budget: %ld (%.2f %%)\n", hdr[2], nfull, ncyc, 100.0 * ncyc / nfull);
  printf("  period lambda: ==1 %.1f %%, <=4 %.1f %%, <=16 %.1f %% of the periodic ones\n", 100.0 * lam1 / ncyc, 100.0 * lam_le4 / ncyc, 100.0 * lam_le16 / ncyc);
  qsort(first_rep, nrec, sizeof(int), cmp_int); qsort(lam, nrec, sizeof(int), cmp_int);
  double qs[] = {0.1, 0.25, 0.5, 0.75, 0.9, 0.95, 0.99};
  printf("  first repeat (pass):"); for (int i = 0; i < 7; ++i) printf(" p%.0f=%d", 100 * qs[i], first_rep[(int)(qs[i] * (nrec - 1))]); printf("\n");
  printf("  period:             "); for (int i = 0; i < 7; ++i) printf(" p%.0f=%d", 100 * qs[i], lam[(int)(qs[i] * (nrec - 1))]); printf("\n");
  printf("  passes a walk from pass %d executes per patch: today %.1f | perfect detection %.1f | Brent %.1f | checkpoint every 32: %.1f\n",
         P0, exec_now / nfull, exec_perfect / nfull, exec_brent / nfull, exec_brent_lin / nfull);
  return 0;
}

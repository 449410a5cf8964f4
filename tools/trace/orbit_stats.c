/* tools/trace/orbit_stats.c — DESIGN ANALYSIS TOOL (not product, not oracle): runs the oracle's per-patch loop on one
 * level of one synthetic pair and records, per patch, the bit patterns of the sampling positions of every pass.
 * Prints: distinct positions per patch, when first visits happen, how an LRU / the engine's set-assoc cache would do.
 * Build: gcc -O2 -ffp-contract=off -o /tmp/orbit_stats tools/trace/orbit_stats.c -lm
 * Input: a raw dump written by tools/trace/dump_level.py (header ints + planes). */
#include "../../oracle/dis_oracle.c"
#include <stdio.h>

typedef struct { unsigned x, y; } pos_t;

static int trace_patch(const level_t* L, const float* gx1, const float* gy1, const float* img2, float refx, float refy,
                       float initx, float inity, pos_t* tr, float* dtr, float* buf) {
  const int n = L->ps * L->ps;
  float *gxp = buf, *gyp = buf + n, *r = buf + 2 * n, *tmp = buf + 3 * n;
  const int lb = -L->ps / 2, ub = L->ps / 2 - 1;
  int cx = (int)roundf(refx) + L->pad, cy = (int)roundf(refy) + L->pad, it = 0;
  for (int j = lb; j <= ub; ++j) for (int i = lb; i <= ub; ++i, ++it) { int idx = (cx + i) + (cy + j) * L->tmp_w; gxp[it] = gx1[idx]; gyp[it] = gy1[idx]; }
  float h00 = dot_sum(gxp, gxp, tmp, n), h01 = dot_sum(gxp, gyp, tmp, n), h11 = dot_sum(gyp, gyp, tmp, n), h10 = h01;
  if (h00 * h11 - h10 * h01 == 0) { h00 = (float)(h00 + 1e-10); h11 = (float)(h11 + 1e-10); }
  float ux = initx, uy = inity, px = refx + ux, py = refy + uy, sx = px, sy = py;
  if (px < L->lb || py < L->lb || px > L->ub_w || py > L->ub_h) return 0;
  sample_second(L, img2, px, py, r, n);
  int swap = fabsf(h10) > fabsf(h00);
  float p = swap ? h10 : h00, q = swap ? h11 : h01, s = swap ? h00 : h10, t = swap ? h01 : h11;
  float l = s / p, u11 = t - l * q;
  int counter = 0;
  for (;;) {
    memcpy(&tr[counter].x, &px, 4); memcpy(&tr[counter].y, &py, 4);
    counter++;
    float g0 = dot_sum(gxp, r, tmp, n), g1 = dot_sum(gyp, r, tmp, n);
    float r0 = swap ? g1 : g0, r1 = swap ? g0 : g1;
    float y1 = r1 - l * r0, d1 = y1 / u11, d0 = (r0 - q * d1) / p;
    dtr[2 * (counter - 1)] = d0; dtr[2 * (counter - 1) + 1] = d1;
    ux = ux - d0; uy = uy - d1; px = refx + ux; py = refy + uy;
    float ddx = sx - px, ddy = sy - py;
    if (sqrtf(ddx * ddx + ddy * ddy) > L->outlier || px < L->lb || py < L->lb || px > L->ub_w || py > L->ub_h) return -counter;
    if (px != px || py != py) return -counter;
    sample_second(L, img2, px, py, r, n);
    if (counter > L->iterations) break;
  }
  return counter;
}

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: orbit_stats dump.bin [stride]\n"); return 1; }
  FILE* f = fopen(argv[1], "rb");
  int hdr[8];   /* W0 H0 scale pad ps steps iterations patnorm */
  if (fread(hdr, 4, 8, f) != 8) return 1;
  int stride = argc > 2 ? atoi(argv[2]) : 1;
  FILE* dump = argc > 3 ? fopen(argv[3], "wb") : NULL;   /* per traced patch: int n (passes; <0: ended by a reset), n x {x bits, y bits} */
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
  float* buf = malloc(sizeof(float) * 4 * L.ps * L.ps);
  /* accumulators */
  long npatch = 0, nfull = 0, nreset = 0;
  enum { MAXD = 1100 };
  static long hist_distinct[MAXD + 1];
  static long first_by_pass[MAXD + 1];      /* first visits happening at pass p (all full patches) */
  static long first_by_pass_ns[MAXD + 1];   /* ... excluding stragglers (distinct > 200) */
  static long nonstrag = 0;
  int caps[6] = {4, 8, 12, 16, 24, 32};
  static long lru_miss[6];                  /* misses of a fully-assoc LRU of given capacity (non-stragglers), passes >= 0 */
  static long settle_hist[MAXD + 1];        /* first pass at which a position is re-visited */
  static long distinct_after[8];            /* distinct first visits at pass >= {8,12,16,20,24,32,48,64} (non-stragglers) */
  int after[8] = {8, 12, 16, 20, 24, 32, 48, 64};
  static long ucycle_found = 0; static double ucycle_onset = 0, ucycle_period = 0;
  for (int x = 0, ip = 0; x < L.npw; ++x)
    for (int y = 0; y < L.nph; ++y, ++ip) {
      if (ip % stride) continue;
      float refx = (float)(x * L.steps + L.offw), refy = (float)(y * L.steps + L.offh), ix = 0, iy = 0;
      if (prev) { int fx = (int)floorf(refx / 2), fy = (int)floorf(refy / 2); int i = fy * (L.width / 2) + fx; ix = prev[2 * i] * 2; iy = prev[2 * i + 1] * 2; }
      int n = trace_patch(&L, gx, gy, im, refx, refy, ix, iy, tr, dtr, buf);
      ++npatch;
      if (dump) { fwrite(&n, 4, 1, dump); fwrite(tr, sizeof(pos_t), n < 0 ? -n : n, dump); }
      if (n <= 0) { ++nreset; continue; }
      ++nfull;
      /* distinct positions + first visit pass */
      static int firsts[MAXD + 2]; int nd = 0; int settle = -1;
      for (int i = 0; i < n; ++i) {
        int seen = 0;
        for (int j = nd - 1; j >= 0; --j) if (tr[firsts[j]].x == tr[i].x && tr[firsts[j]].y == tr[i].y) { seen = 1; break; }
        if (!seen) firsts[nd++] = i; else if (settle < 0) settle = i;
      }
      hist_distinct[nd > MAXD ? MAXD : nd]++;
      settle_hist[settle < 0 ? MAXD : settle]++;
      for (int j = 0; j < nd; ++j) first_by_pass[firsts[j]]++;
      if (nd <= 200) {
        ++nonstrag;
        for (int j = 0; j < nd; ++j) { first_by_pass_ns[firsts[j]]++; for (int a = 0; a < 8; ++a) if (firsts[j] >= after[a]) distinct_after[a]++; }
        for (int c = 0; c < 6; ++c) {   /* LRU simulation */
          int cap = caps[c]; static pos_t lr[64]; int nl = 0; long miss = 0;
          for (int i = 0; i < n; ++i) {
            int hit = -1;
            for (int j = 0; j < nl; ++j) if (lr[j].x == tr[i].x && lr[j].y == tr[i].y) { hit = j; break; }
            pos_t cur = tr[i];
            if (hit < 0) { ++miss; if (nl < cap) ++nl; hit = nl - 1; }
            for (int j = hit; j > 0; --j) lr[j] = lr[j - 1];
            lr[0] = cur;
          }
          lru_miss[c] += miss;
        }
      }
    }
  printf("level scale %d: %dx%d ps %d, patches traced %ld, full %ld, reset/oob %ld\n", hdr[2], L.width, L.height, L.ps, npatch, nfull, nreset);
  long cum = 0; double mean = 0;
  for (int d = 0; d <= MAXD; ++d) mean += (double)d * hist_distinct[d];
  printf("distinct positions per full patch: mean %.2f\n", mean / nfull);
  int qs[] = {10, 25, 50, 75, 90, 95, 99, 995, 999}; int qi = 0;
  for (int d = 0; d <= MAXD && qi < 9; ++d) { cum += hist_distinct[d]; while (qi < 9 && cum * (qs[qi] >= 100 ? 1000.0 : 100.0) >= (double)qs[qi] * nfull) { printf("  p%d = %d\n", qs[qi], d); ++qi; } }
  long strag = nfull - nonstrag; double ms = 0; for (int d = 201; d <= MAXD; ++d) ms += (double)d * hist_distinct[d];
  printf("stragglers (>200 distinct): %ld (%.3f %%), mean distinct %.1f -> %.2f samples per patch averaged over all\n", strag, 100.0 * strag / nfull, strag ? ms / strag : 0, ms / nfull);
  double mns = 0; for (int d = 0; d <= 200; ++d) mns += (double)d * hist_distinct[d];
  printf("non-stragglers: mean distinct %.2f\n", mns / nonstrag);
  printf("non-stragglers: first visits at pass >= P (per patch):"); for (int a = 0; a < 8; ++a) printf("  P=%d: %.2f", after[a], (double)distinct_after[a] / nonstrag); printf("\n");
  printf("non-stragglers: LRU misses per patch:"); for (int c = 0; c < 6; ++c) printf("  cap %d: %.2f", caps[c], (double)lru_miss[c] / nonstrag); printf("\n");
  printf("first re-visit pass (settle): "); cum = 0; qi = 0; for (int d = 0; d <= MAXD && qi < 9; ++d) { cum += settle_hist[d]; while (qi < 9 && cum * (qs[qi] >= 100 ? 1000.0 : 100.0) >= (double)qs[qi] * nfull) { printf(" p%d=%d", qs[qi], d); ++qi; } } printf("\n");
  printf("non-straggler first visits by pass bucket (per patch): ");
  int edges[] = {0, 4, 8, 12, 16, 20, 24, 32, 48, 64, 96, 128, 192, 256, 384, 512, 768, 1002};
  for (int b = 0; b < 17; ++b) { long s = 0; for (int p = edges[b]; p < edges[b + 1] && p <= MAXD; ++p) s += first_by_pass_ns[p]; printf(" [%d,%d):%.2f", edges[b], edges[b + 1], (double)s / nonstrag); }
  printf("\n");
  return 0;
}

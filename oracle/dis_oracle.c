/*
 * oracle/dis_oracle.c — CPU restatement of the reference's Dense-Inverse-Search hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
 * The product path (the CUDA library behind include/dis_b200.h) never calls it and has no
 * CPU fallback.
 *
 * Parity status: the reference has no tests, fixtures or golden vectors of its own
 *  — "parity unpinned" by the reference's tests.  This restatement is
 * instead pinned against the reference ITSELF: oracle/ref_driver.cpp compiles the reference's
 * unmodified engine sources (src/patch.cpp, src/patch_grid.cpp, src/optical_flow.cpp) in place,
 * and tests/test_oracle.py + tests/golden/ check that this file reproduces their output
 * bit for bit.  The pre-processing half (OpenCV calls in src/main.cpp:12-50,139-160) is checked
 * against cv2 4.13 where that is importable.
 *
 * Plain C, fp32 scalar arithmetic written in the reference's association.  Must be compiled
 * with -ffp-contract=off (no FMA contraction): see oracle/Makefile.
 *
 * Every function cites the reference lines it restates (paths relative to /root/reference).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------
 * Pre-processing (src/main.cpp:12-50, 139-160; OpenCV semantics per SURVEY.md App. A.1)
 * ---------------------------------------------------------------------------------------- */

/* src/main.cpp:139-149 — padding needed to make a dimension divisible by 2^coarsest. */
ORACLE_API int oracle_div_pad(int dim, int coarsest) {
  int sf = 1 << coarsest;
  int r = dim % sf;
  return r > 0 ? sf - r : 0;
}

/* src/main.cpp:150-160 — copyMakeBorder(BORDER_REPLICATE) with the floor/ceil split, then
 * convertTo(CV_32F).  out is (h+padh) x (w+padw) floats. */
ORACLE_API void oracle_pad_convert(const uint8_t* img, int w, int h, int stride, int padw, int padh,
                                   float* out) {
  int left = padw / 2, top = padh / 2; /* floor(pad/2.0f); right/bottom get the ceil */
  int W = w + padw, H = h + padh;
  for (int y = 0; y < H; ++y) {
    int sy = y - top; if (sy < 0) sy = 0; if (sy > h - 1) sy = h - 1;
    for (int x = 0; x < W; ++x) {
      int sx = x - left; if (sx < 0) sx = 0; if (sx > w - 1) sx = w - 1;
      out[(size_t)y * W + x] = (float)img[(size_t)sy * stride + sx];
    }
  }
}

static inline int reflect101(int i, int n) {
  if (n == 1) return 0;
  if (i < 0) i = -i;
  if (i >= n) i = 2 * n - 2 - i;
  return i;
}

/* cv::Sobel(src, CV_32F, 1,0 / 0,1, ksize 3, scale 1/8, BORDER_DEFAULT) — src/main.cpp:19-20,34-35.
 * OpenCV folds the scale into the smoothing kernel (1,2,1)/8 = (.125,.25,.125); the derivative
 * kernel stays (-1,0,1).  Row pass first, then column pass. */
ORACLE_API void oracle_sobel(const float* p, int w, int h, float* gx, float* gy) {
  for (int y = 0; y < h; ++y) {
    int ym = reflect101(y - 1, h), yp = reflect101(y + 1, h);
    for (int x = 0; x < w; ++x) {
      int xm = reflect101(x - 1, w), xp = reflect101(x + 1, w);
      /* dx: horizontal difference, vertical smoothing */
      float d0 = p[(size_t)ym * w + xp] - p[(size_t)ym * w + xm];
      float d1 = p[(size_t)y * w + xp] - p[(size_t)y * w + xm];
      float d2 = p[(size_t)yp * w + xp] - p[(size_t)yp * w + xm];
      gx[(size_t)y * w + x] = (d0 + d2) * 0.125f + d1 * 0.25f;
      /* dy: horizontal smoothing, vertical difference */
      float s0 = (p[(size_t)ym * w + xm] + p[(size_t)ym * w + xp]) * 0.125f + p[(size_t)ym * w + x] * 0.25f;
      float s2 = (p[(size_t)yp * w + xm] + p[(size_t)yp * w + xp]) * 0.125f + p[(size_t)yp * w + x] * 0.25f;
      gy[(size_t)y * w + x] = s2 - s0;
    }
  }
}

/* src/main.cpp:18-26 — level 0 of the pyramid is the Sobel gradient magnitude. */
ORACLE_API void oracle_gradmag(const float* img, int w, int h, float* out) {
  float* gx = (float*)malloc(sizeof(float) * (size_t)w * h);
  float* gy = (float*)malloc(sizeof(float) * (size_t)w * h);
  oracle_sobel(img, w, h, gx, gy);
  for (size_t i = 0; i < (size_t)w * h; ++i) {
    float a = gx[i] * gx[i];
    float b = gy[i] * gy[i];
    out[i] = sqrtf(a + b);
  }
  free(gx); free(gy);
}

/* src/main.cpp:29 — cv::resize(fx=fy=.5, INTER_LINEAR) on even sizes = 2x2 mean in this order. */
ORACLE_API void oracle_halve(const float* p, int w, int h, float* out) {
  int w2 = w / 2, h2 = h / 2;
  for (int y = 0; y < h2; ++y)
    for (int x = 0; x < w2; ++x) {
      const float* r0 = p + (size_t)(2 * y) * w + 2 * x;
      const float* r1 = r0 + w;
      out[(size_t)y * w2 + x] = ((r0[0] + r0[1]) + (r1[0] + r1[1])) * 0.25f;
    }
}

/* src/main.cpp:43-47 — copyMakeBorder by `pad` on all sides; replicate (image) or constant 0. */
ORACLE_API void oracle_border(const float* p, int w, int h, int pad, int replicate, float* out) {
  int W = w + 2 * pad, H = h + 2 * pad;
  for (int y = 0; y < H; ++y) {
    int sy = y - pad;
    for (int x = 0; x < W; ++x) {
      int sx = x - pad;
      float v;
      if (replicate) {
        int cy = sy < 0 ? 0 : (sy > h - 1 ? h - 1 : sy);
        int cx = sx < 0 ? 0 : (sx > w - 1 ? w - 1 : sx);
        v = p[(size_t)cy * w + cx];
      } else {
        v = (sy < 0 || sy >= h || sx < 0 || sx >= w) ? 0.0f : p[(size_t)sy * w + sx];
      }
      out[(size_t)y * W + x] = v;
    }
  }
}

/* construct_pyramide, src/main.cpp:12-50.  img: fp32 frame (already divisibility padded), w x h.
 * For level i in 0..levels-1 writes padded planes of (h>>i + 2pad) x (w>>i + 2pad) floats into
 * pyr[i], pyr_dx[i], pyr_dy[i] (caller allocated). */
ORACLE_API void oracle_construct_pyramid(const float* img, int w, int h, int levels, int pad,
                                         float** pyr, float** pyr_dx, float** pyr_dy) {
  float* cur = (float*)malloc(sizeof(float) * (size_t)w * h);
  float* gx = (float*)malloc(sizeof(float) * (size_t)w * h);
  float* gy = (float*)malloc(sizeof(float) * (size_t)w * h);
  int cw = w, ch = h;
  oracle_gradmag(img, w, h, cur);
  for (int i = 0; i < levels; ++i) {
    if (i > 0) {
      float* nxt = (float*)malloc(sizeof(float) * (size_t)(cw / 2) * (ch / 2));
      oracle_halve(cur, cw, ch, nxt);
      free(cur); cur = nxt; cw /= 2; ch /= 2;
    }
    oracle_sobel(cur, cw, ch, gx, gy);
    oracle_border(cur, cw, ch, pad, 1, pyr[i]);
    oracle_border(gx, cw, ch, pad, 0, pyr_dx[i]);
    oracle_border(gy, cw, ch, pad, 0, pyr_dy[i]);
  }
  free(cur); free(gx); free(gy);
}

/* ------------------------------------------------------------------------------------------
 * Engine (src/optical_flow.cpp, src/patch_grid.cpp, src/patch.cpp)
 * ---------------------------------------------------------------------------------------- */

typedef struct {
  int width, height, pad, tmp_w;         /* include/optical_flow.hpp:14-24 */
  float lb, ub_w, ub_h;
  int ps, steps, iterations, patnorm;    /* include/optical_flow.hpp:26-37 */
  float outlier;
  int npw, nph, offw, offh;              /* src/patch_grid.cpp:20-23 */
} level_t;

/* src/optical_flow.cpp:38 — steps = max(1, (int)floor(ps*(1-overlap))) in fp32. */
ORACLE_API int oracle_steps(int patch_size, float overlap) {
  float one_minus = 1 - overlap;
  int s = (int)floorf((float)patch_size * one_minus);
  return s > 1 ? s : 1;
}

/* src/optical_flow.cpp:49-59 + src/patch_grid.cpp:20-25. */
static void level_setup(level_t* L, int width0, int height0, int scale, int pad, int ps, int steps,
                        int iterations, int patnorm) {
  float sf = (float)pow(2, -scale);
  L->height = (int)(height0 * sf);
  L->width = (int)(width0 * sf);
  L->pad = pad;
  L->lb = -(float)ps / 2;
  L->ub_w = (float)(L->width + ps / 2 - 2);
  L->ub_h = (float)(L->height + ps / 2 - 2);
  L->tmp_w = L->width + 2 * pad;
  L->ps = ps; L->steps = steps; L->iterations = iterations; L->patnorm = patnorm;
  L->outlier = (float)ps / 2;
  L->npw = (int)ceilf((float)L->width / (float)steps);
  L->nph = (int)ceilf((float)L->height / (float)steps);
  L->offw = (L->width - (L->npw - 1) * steps) / 2;
  L->offh = (L->height - (L->nph - 1) * steps) / 2;
}

/* Introspection used by the tests: geometry of one level. out = {w,h,npw,nph,offw,offh,steps,tmp_w}. */
ORACLE_API void oracle_level_info(int width0, int height0, int scale, int pad, int ps, float overlap,
                                  int* out) {
  level_t L;
  level_setup(&L, width0, height0, scale, pad, ps, oracle_steps(ps, overlap), 0, 0);
  out[0] = L.width; out[1] = L.height; out[2] = L.npw; out[3] = L.nph;
  out[4] = L.offw; out[5] = L.offh; out[6] = L.steps; out[7] = L.tmp_w;
}

/* Eigen 3.3 redux for n floats (used at src/patch.cpp:82-84,171-172,265). */
static float tree_sum(const float* v, int n) {
  if (n < 4) { float r = n > 0 ? v[0] : 0.0f; for (int i = 1; i < n; ++i) r += v[i]; return r; }
  int n4 = (n / 4) * 4, n8 = (n / 8) * 8;
  float a[4], b[4];
  for (int j = 0; j < 4; ++j) a[j] = v[j];
  if (n4 >= 8) {
    for (int j = 0; j < 4; ++j) b[j] = v[4 + j];
    for (int e = 8; e < n8; e += 8)
      for (int j = 0; j < 4; ++j) { a[j] += v[e + j]; b[j] += v[e + 4 + j]; }
    for (int j = 0; j < 4; ++j) a[j] = a[j] + b[j];
    if (n4 > n8) for (int j = 0; j < 4; ++j) a[j] += v[n8 + j];
  }
  float r = (a[0] + a[2]) + (a[1] + a[3]);
  for (int i = n4; i < n; ++i) r += v[i];
  return r;
}

static float dot_sum(const float* a, const float* b, float* tmp, int n) {
  for (int i = 0; i < n; ++i) tmp[i] = a[i] * b[i];
  return tree_sum(tmp, n);
}

/* Patch::get_patch_second_image, src/patch.cpp:207-267. */
static void sample_second(const level_t* L, const float* img2, float px, float py, float* out,
                          int n) {
  float l = floorf(px), k = floorf(py);
  float a = px - l, b = py - k;
  float w0 = (1 - a) * (1 - b), w1 = a * (1 - b), w2 = b * (1 - a), w3 = a * b;
  int posx = (int)ceilf(px + .00001f) + L->pad;
  int posy = (int)ceilf(py + .00001f) + L->pad;
  int lb = -L->ps / 2, ub = L->ps / 2 - 1;
  int ind_e = posx - L->ps / 2;
  int it = 0;
  for (int y = posy + lb; y <= posy + ub; ++y) {
    int ia = ind_e + y * L->tmp_w;
    int ic = ind_e + (y - 1) * L->tmp_w;
    for (int x = lb; x <= ub; ++x, ++ia, ++ic, ++it)
      out[it] = w3 * img2[ia] + w2 * img2[ia - 1] + w1 * img2[ic] + w0 * img2[ic - 1];
  }
  if (L->patnorm) {
    float m = tree_sum(out, n) / n;
    for (int i = 0; i < n; ++i) out[i] -= m;
  }
}

/* One patch, start to finish: Patch::init_patch (src/patch.cpp:31-91), inverse_search_start
 * (:119-154) and inverse_search (:156-203).  Returns the number of passes of the loop. */
static int run_patch(const level_t* L, const float* gx1, const float* gy1, const float* img2,
                     float refx, float refy, float initx, float inity, float* u_out,
                     float* gxp, float* gyp, float* r, float* tmp) {
  const int n = L->ps * L->ps;
  const int lb = -L->ps / 2, ub = L->ps / 2 - 1;
  /* get_gradients_on_patch, :47-73 */
  int cx = (int)roundf(refx) + L->pad, cy = (int)roundf(refy) + L->pad;
  int it = 0;
  for (int j = lb; j <= ub; ++j)
    for (int i = lb; i <= ub; ++i, ++it) {
      int idx = (cx + i) + (cy + j) * L->tmp_w;
      gxp[it] = gx1[idx];
      gyp[it] = gy1[idx];
    }
  /* compute_hessian_matrix, :75-91 */
  float h00 = dot_sum(gxp, gxp, tmp, n);
  float h01 = dot_sum(gxp, gyp, tmp, n);
  float h11 = dot_sum(gyp, gyp, tmp, n);
  float h10 = h01;
  if (h00 * h11 - h10 * h01 == 0) {
    h00 = (float)(h00 + 1e-10);   /* `+= 1e-10` is a double addition rounded back to fp32 */
    h11 = (float)(h11 + 1e-10);
  }
  /* inverse_search_start, :119-154 */
  float ux = initx, uy = inity;
  float px = refx + ux, py = refy + uy;
  float sx = px, sy = py;
  if (px < L->lb || py < L->lb || px > L->ub_w || py > L->ub_h) {
    u_out[0] = ux; u_out[1] = uy;
    return 0;
  }
  sample_second(L, img2, px, py, r, n);
  /* 2x2 PartialPivLU of the constant Hessian (src/patch.cpp:176; App. A.3) */
  int swap = fabsf(h10) > fabsf(h00);
  float p = swap ? h10 : h00, q = swap ? h11 : h01;
  float s = swap ? h00 : h10, t = swap ? h01 : h11;
  float l = s / p;
  float u11 = t - l * q;
  int counter = 0;
  for (;;) {
    counter++;
    float g0 = dot_sum(gxp, r, tmp, n);       /* :171 */
    float g1 = dot_sum(gyp, r, tmp, n);       /* :172 */
    float r0 = swap ? g1 : g0, r1 = swap ? g0 : g1;
    float y1 = r1 - l * r0;
    float d1 = y1 / u11;
    float d0 = (r0 - q * d1) / p;
    ux = ux - d0; uy = uy - d1;               /* :179 */
    px = refx + ux; py = refy + uy;           /* :182 */
    float ddx = sx - px, ddy = sy - py;
    int stop = 0;
    if (sqrtf(ddx * ddx + ddy * ddy) > L->outlier || px < L->lb || py < L->lb || px > L->ub_w ||
        py > L->ub_h) {                       /* :185-194 */
      ux = initx; uy = inity;
      px = refx + ux; py = refy + uy;
      stop = 1;
    }
    /* NaN convention: the reference would index out of bounds; reset. */
    if (!stop && (px != px || py != py)) { ux = initx; uy = inity; break; }
    sample_second(L, img2, px, py, r, n);     /* :196 */
    if (counter > L->iterations) stop = 1;    /* :199-201 */
    if (stop) break;
  }
  u_out[0] = ux; u_out[1] = uy;
  return counter;
}

/* PatchGrid::densification_and_create_dance_flow, src/patch_grid.cpp:121-182, with `weight`
 * zero-initialised. */
static void densify(const level_t* L, const float* u, float* dense) {
  size_t npx = (size_t)L->width * L->height;
  float* weight = (float*)calloc(npx, sizeof(float));
  memset(dense, 0, sizeof(float) * 2 * npx);
  int lb = -L->ps / 2, ub = L->ps / 2 - 1;
  int ip = 0;
  for (int gxi = 0; gxi < L->npw; ++gxi)
    for (int gyi = 0; gyi < L->nph; ++gyi, ++ip) {
      int rx = gxi * L->steps + L->offw, ry = gyi * L->steps + L->offh;
      float nu0 = u[2 * ip] * 0.5f, nu1 = u[2 * ip + 1] * 0.5f;
      for (int y = lb; y <= ub; ++y)
        for (int x = lb; x <= ub; ++x) {
          int xt = x + rx, yt = y + ry;
          if (xt >= 0 && yt >= 0 && xt < L->width && yt < L->height) {
            size_t i = (size_t)yt * L->width + xt;
            weight[i] += 0.5f;
            dense[2 * i] += nu0;
            dense[2 * i + 1] += nu1;
          }
        }
    }
  for (size_t i = 0; i < npx; ++i)
    if (weight[i] > 0) { dense[2 * i] /= weight[i]; dense[2 * i + 1] /= weight[i]; }
  free(weight);
}

/* One level: PatchGrid ctor + init_grid + set_target_image + patch_init_from_prev_flow +
 * inverse_search + densification (src/optical_flow.cpp:67-90, src/patch_grid.cpp).
 * flow_prev may be NULL (coarsest level).  patch_u / patch_iters may be NULL. */
static void run_level(const level_t* L, const float* gx1, const float* gy1, const float* img2,
                      const float* flow_prev, float* dense, float* patch_ref, float* patch_u,
                      int* patch_iters) {
  int n = L->ps * L->ps;
  int np = L->npw * L->nph;
  float* u = (float*)malloc(sizeof(float) * 2 * (size_t)np);
  float* buf = (float*)malloc(sizeof(float) * 4 * (size_t)n);
  int ip = 0;
  for (int x = 0; x < L->npw; ++x)
    for (int y = 0; y < L->nph; ++y, ++ip) {
      float refx = (float)(x * L->steps + L->offw);    /* src/patch_grid.cpp:43-44 */
      float refy = (float)(y * L->steps + L->offh);
      float ix = 0.0f, iy = 0.0f;
      if (flow_prev) {                                  /* src/patch_grid.cpp:108-119 */
        int fx = (int)floorf(refx / 2), fy = (int)floorf(refy / 2);
        int i = fy * (L->width / 2) + fx;
        ix = flow_prev[2 * i] * 2;
        iy = flow_prev[2 * i + 1] * 2;
      }
      int iters = run_patch(L, gx1, gy1, img2, refx, refy, ix, iy, u + 2 * ip, buf, buf + n,
                            buf + 2 * n, buf + 3 * n);
      if (patch_ref) { patch_ref[2 * ip] = refx; patch_ref[2 * ip + 1] = refy; }
      if (patch_iters) patch_iters[ip] = iters;
    }
  if (patch_u) memcpy(patch_u, u, sizeof(float) * 2 * (size_t)np);
  densify(L, u, dense);
  free(u); free(buf);
}

/* OpticalFlowClass::OpticalFlowClass, src/optical_flow.cpp:19-132 (draw_grid = false).
 * The six pyramids are arrays indexed by ABSOLUTE scale, planes padded by img_padding.
 * Only img_first_dx/dy and img_second are read; the other three pointers
 * are accepted for signature fidelity and may be NULL.
 *
 * Optional dumps for the parity harness (any may be NULL): for level index s = scale - finest,
 * lvl_flow[s] (2*w*h), lvl_ref[s] / lvl_u[s] (2*npatches), lvl_iters[s] (npatches). */
ORACLE_API void oracle_optical_flow(float** img_first, float** img_first_dx, float** img_first_dy,
                                    float** img_second, float** img_second_dx, float** img_second_dy,
                                    int img_padding, float* outflow, int width, int height,
                                    int coarsest, int finest, int iterations, int patch_size,
                                    float overlap, int patnorm, float** lvl_flow, float** lvl_ref,
                                    float** lvl_u, int** lvl_iters) {
  (void)img_first; (void)img_second_dx; (void)img_second_dy;
  int steps = oracle_steps(patch_size, overlap);
  int nl = coarsest - finest + 1;
  float** flows = (float**)calloc((size_t)nl, sizeof(float*));
  level_t* Ls = (level_t*)calloc((size_t)nl, sizeof(level_t));
  for (int scale = coarsest; scale >= finest; --scale) {
    int s = scale - finest;
    level_setup(&Ls[s], width, height, scale, img_padding, patch_size, steps, iterations, patnorm);
    flows[s] = (float*)malloc(sizeof(float) * 2 * (size_t)Ls[s].width * Ls[s].height);
  }
  for (int scale = coarsest; scale >= finest; --scale) {
    int s = scale - finest;
    float* dst = (scale == finest) ? outflow : flows[s];
    run_level(&Ls[s], img_first_dx[scale], img_first_dy[scale], img_second[scale],
              scale < coarsest ? flows[s + 1] : NULL, dst, lvl_ref ? lvl_ref[s] : NULL,
              lvl_u ? lvl_u[s] : NULL, lvl_iters ? lvl_iters[s] : NULL);
    if (lvl_flow && lvl_flow[s])
      memcpy(lvl_flow[s], dst, sizeof(float) * 2 * (size_t)Ls[s].width * Ls[s].height);
  }
  for (int s = 0; s < nl; ++s) free(flows[s]);
  free(flows); free(Ls);
}

/* The whole per-pair path of src/main.cpp:135-198 for finest_scale == 0 on two u8 grey frames:
 * divisibility pad -> float -> two pyramids -> engine -> crop.  flow_out is h x w x 2. */
ORACLE_API int oracle_run_pair_u8(const uint8_t* a, const uint8_t* b, int w, int h, int stride,
                                  int iterations, int patch_size, int coarsest, int finest,
                                  float overlap, int patnorm, float* flow_out) {
  if (finest != 0) return -1;
  int padw = oracle_div_pad(w, coarsest), padh = oracle_div_pad(h, coarsest);
  int W = w + padw, H = h + padh, nl = coarsest + 1, pad = patch_size;
  float* fa = (float*)malloc(sizeof(float) * (size_t)W * H);
  float* fb = (float*)malloc(sizeof(float) * (size_t)W * H);
  oracle_pad_convert(a, w, h, stride, padw, padh, fa);
  oracle_pad_convert(b, w, h, stride, padw, padh, fb);
  float** P[6];
  for (int k = 0; k < 6; ++k) {
    P[k] = (float**)calloc((size_t)nl, sizeof(float*));
    for (int i = 0; i < nl; ++i)
      P[k][i] = (float*)malloc(sizeof(float) * (size_t)((W >> i) + 2 * pad) * ((H >> i) + 2 * pad));
  }
  oracle_construct_pyramid(fa, W, H, nl, pad, P[0], P[1], P[2]);
  oracle_construct_pyramid(fb, W, H, nl, pad, P[3], P[4], P[5]);
  float* full = (float*)malloc(sizeof(float) * 2 * (size_t)W * H);
  oracle_optical_flow(P[0], P[1], P[2], P[3], P[4], P[5], pad, full, W, H, coarsest, finest,
                      iterations, patch_size, overlap, patnorm, NULL, NULL, NULL, NULL);
  int left = padw / 2, top = padh / 2;                   /* src/main.cpp:198 */
  for (int y = 0; y < h; ++y)
    memcpy(flow_out + (size_t)y * w * 2, full + ((size_t)(y + top) * W + left) * 2,
           sizeof(float) * 2 * (size_t)w);
  for (int k = 0; k < 6; ++k) { for (int i = 0; i < nl; ++i) free(P[k][i]); free(P[k]); }
  free(fa); free(fb); free(full);
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * Colour coding: draw_optical_flow / compute_color / is_flow_correct,
 * src/color_coding.cpp:8-117.  bgr is h*w*3 bytes, B,G,R per pixel (pix[2 - b], :77).
 * ---------------------------------------------------------------------------------------- */
static int flow_ok(float x, float y) {                       /* :8-11 */
  return !isnan(x) && !isnan(y) && fabsf(x) < 1e9f && fabsf(y) < 1e9f;
}

ORACLE_API void oracle_color_code(const float* flow, int w, int h, float maxmotion, uint8_t* bgr) {
  enum { RY = 15, YG = 6, GC = 4, CB = 11, BM = 13, MR = 6, NCOLS = RY + YG + GC + CB + BM + MR };
  int wheel[NCOLS][3];
  int k = 0;                                                 /* :30-50 */
  for (int i = 0; i < RY; ++i, ++k) { wheel[k][0] = 255; wheel[k][1] = 255 * i / RY; wheel[k][2] = 0; }
  for (int i = 0; i < YG; ++i, ++k) { wheel[k][0] = 255 - 255 * i / YG; wheel[k][1] = 255; wheel[k][2] = 0; }
  for (int i = 0; i < GC; ++i, ++k) { wheel[k][0] = 0; wheel[k][1] = 255; wheel[k][2] = 255 * i / GC; }
  for (int i = 0; i < CB; ++i, ++k) { wheel[k][0] = 0; wheel[k][1] = 255 - 255 * i / CB; wheel[k][2] = 255; }
  for (int i = 0; i < BM; ++i, ++k) { wheel[k][0] = 255 * i / BM; wheel[k][1] = 0; wheel[k][2] = 255; }
  for (int i = 0; i < MR; ++i, ++k) { wheel[k][0] = 255; wheel[k][1] = 0; wheel[k][2] = 255 - 255 * i / MR; }
  size_t n = (size_t)w * h;
  float maxrad = maxmotion;                                  /* :89-106 */
  if (maxmotion <= 0) {
    maxrad = 1;
    for (size_t i = 0; i < n; ++i) {
      float x = flow[2 * i], y = flow[2 * i + 1];
      if (!flow_ok(x, y)) continue;
      float r = sqrtf(x * x + y * y);
      if (r > maxrad) maxrad = r;
    }
  }
  memset(bgr, 0, 3 * n);                                     /* dst.setTo(0), :86 */
  for (size_t i = 0; i < n; ++i) {
    float x = flow[2 * i], y = flow[2 * i + 1];
    if (!flow_ok(x, y)) continue;
    float fx = x / maxrad, fy = y / maxrad;                  /* :113 */
    float rad = sqrtf(fx * fx + fy * fy);                    /* :55-60 */
    float a = atan2f(-fy, -fx) / (float)3.1415926535897932384626433832795;
    float fk = (a + 1.0f) / 2.0f * (NCOLS - 1);
    int k0 = (int)fk;
    int k1 = (k0 + 1) % NCOLS;
    float f = fk - k0;
    for (int b = 0; b < 3; ++b) {                            /* :64-78 */
      float col0 = wheel[k0][b] / 255.f;
      float col1 = wheel[k1][b] / 255.f;
      float col = (1 - f) * col0 + f * col1;
      if (rad <= 1) col = 1 - rad * (1 - col);
      else col *= .75f;
      bgr[3 * i + (2 - b)] = (uint8_t)(255.f * col);
    }
  }
}

// dis_search.cu — the per-level inverse search on sm_100a.
//
// Replaces, for every patch of a level and every pair of a batch:
//   PatchGrid::PatchGrid          src/patch_grid.cpp:17-51   grid geometry, x-major patch ids
//   Patch::init_patch             src/patch.cpp:31-91        template gradients + 2x2 Hessian
//   PatchGrid::patch_init_from_prev_flow  src/patch_grid.cpp:108-119
//   Patch::inverse_search(_start) src/patch.cpp:119-203      the iteration loop
//   Patch::get_patch_second_image src/patch.cpp:207-267      4-tap resample + mean normalisation
//
// Design (B200-first, not a translation):
//  * One LANE per patch carries the scalar state (u, pos, LU factors).  A block is a TILE of the
//    patch grid: warp w owns grid column (8*tile_x + w), lanes are 32 consecutive grid rows, so the
//    block's taps and templates share L1 lines.
//  * The expensive part of a pass (ps^2 4-tap samples + three tree reductions) is done by an
//    8-lane sub-warp per patch, four patches per warp round.  Lane j of a sub-warp owns the
//    elements e == j (mod 8), which makes Eigen's SSE reduction tree (App. A.3) exactly
//    "serial over the lane's elements, then xor-shuffle 4, 2, 1".
//  * EXACT memoisation: within a patch, delta_u is a pure function of the bit pattern of the
//    sampling position (H, the template and the image are constant), and so is the outcome of the
//    reset test.  Each patch keeps a 16-entry, 4-way cache {pos bits -> delta_u} in shared memory;
//    a pass whose position is cached costs a handful of scalar instructions instead of ~900.  The
//    fp32 orbit of a patch revisits a few dozen positions (it jitters at the ulp of the
//    coordinate), so ~98 % of the reference's 1001 passes per patch hit.  Results are bit-identical
//    to running every pass: nothing is approximated, only recomputation is skipped.
//  * Lanes that miss wait; the warp services misses in rounds of four so sub-warps stay busy.
//  * Patches whose orbit never settles (~0.15 %) would keep one lane of a warp busy for ~1000
//    rounds; they are handed to k_stragglers, which packs four of them per warp.
#include <stdlib.h>

#include "dis_kernels.cuh"

namespace dis {
namespace {

constexpr int kWarpsPerBlock = 8;     // = grid columns per tile
constexpr int kCacheEntries = 16;
constexpr int kWays = 4;
constexpr unsigned kFull = 0xffffffffu;
constexpr int kLonelyLimit = 8;       // consecutive under-filled forced rounds before a hand-off
constexpr int kHandoffMinIter = 48;   // never hand off a patch that is still in its convergence phase
constexpr int kCacheBytesPerWarp = kCacheEntries * 32 * 16;

template <int PS>
struct Geom {
  static constexpr int NE = PS * PS;
  static constexpr int K = NE / 8;
  static constexpr bool EXTRA = (NE % 8) >= 4;
  static constexpr int KX = K + (EXTRA ? 1 : 0);
};

// Eigen-3.3 redux of the per-lane partial sums.  `acc` is the lane's serial
// sum over its elements c, c+8, ...; `extra` the element 8*floor(n/8) + (lane&3) when n%8 >= 4.
template <bool EXTRA>
__device__ __forceinline__ float tree_finish(float acc, float extra) {
  acc = acc + __shfl_xor_sync(kFull, acc, 4);
  if (EXTRA) acc = acc + extra;
  acc = acc + __shfl_xor_sync(kFull, acc, 2);
  acc = acc + __shfl_xor_sync(kFull, acc, 1);
  return acc;
}

struct Straggler { int pair, pid, k; float ux, uy; };

// 2x2 partial-pivot LU of the constant Hessian (src/patch.cpp:176; App. A.3)
struct LU2 {
  float p, q, l, u11;
  bool swap;
  __device__ __forceinline__ void factor(float h00, float h01, float h11) {
    swap = fabsf(h01) > fabsf(h00);
    p = swap ? h01 : h00; q = swap ? h11 : h01;
    const float s = swap ? h00 : h01, t = swap ? h01 : h11;
    l = s / p;
    u11 = t - l * q;
  }
  __device__ __forceinline__ float2 solve(float g0, float g1) const {
    const float r0 = swap ? g1 : g0, r1 = swap ? g0 : g1;
    const float y1 = r1 - l * r0;
    const float d1 = y1 / u11;
    const float d0 = (r0 - q * d1) / p;
    return make_float2(d0, d1);
  }
};

__device__ __forceinline__ void fix_singular(float& h00, float h01, float& h11) {
  if (h00 * h11 - h01 * h01 == 0.f) {   // `+= 1e-10` is a double addition (src/patch.cpp:86-90)
    h00 = (float)((double)h00 + 1e-10);
    h11 = (float)((double)h11 + 1e-10);
  }
}

// Element offsets of sub-lane `sl` inside a patch window, relative to the patch centre.
template <int PS>
__device__ __forceinline__ void element_offsets(int sl, int tw, int (&eoff)[Geom<PS>::KX]) {
  using G = Geom<PS>;
#pragma unroll
  for (int k = 0; k < G::KX; ++k) {
    const int idx = (k < G::K) ? (sl + 8 * k) : (8 * G::K + (sl & 3));
    const int row = idx / PS, col = idx - row * PS;
    eoff[k] = (row - PS / 2) * tw + (col - PS / 2);
  }
}

// Patch::get_patch_second_image (src/patch.cpp:207-267) for the sub-lane's elements, followed by
// the two products-and-sums of src/patch.cpp:171-172.  gxv/gyv: the lane's template gradients.
// Returns (sum gx*r, sum gy*r), identical in all 8 lanes of the sub-warp.
template <int PS>
__device__ __forceinline__ float2 sample_and_project(const float* __restrict__ imB, int tw, int pad, int wmax,
                                                     int hmax, float opx, float opy, int sl, int patnorm,
                                                     const int (&eoff)[Geom<PS>::KX],
                                                     const float (&gxv)[Geom<PS>::KX],
                                                     const float (&gyv)[Geom<PS>::KX]) {
  using G = Geom<PS>;
  const float fl = floorf(opx), fk = floorf(opy);
  const float a = opx - fl, b = opy - fk;
  const float w0 = (1 - a) * (1 - b), w1 = a * (1 - b), w2 = b * (1 - a), w3 = a * b;
  // tap base; clamped so that a position about to be rejected by the reset test (out of bounds,
  // NaN) cannot address outside the plane — its result is discarded by the owner
  int bxi = (int)ceilf(opx + .00001f) + pad;
  int byi = (int)ceilf(opy + .00001f) + pad;
  bxi = min(max(bxi, PS / 2 + 1), wmax);
  byi = min(max(byi, PS / 2 + 1), hmax);
  const float* __restrict__ ib = imB + byi * tw + bxi;
  float v[G::KX];
  if (PS == 8) {
    // lane sl owns column sl: rows share taps, 9 x 2 loads instead of 8 x 4
    const float* col = ib + (sl - 4) - 5 * tw;
    float t0[9], t1[9];
#pragma unroll
    for (int r = 0; r < 9; ++r) { t0[r] = __ldg(col + r * tw - 1); t1[r] = __ldg(col + r * tw); }
#pragma unroll
    for (int r = 0; r < 8; ++r) v[r] = ((w3 * t1[r + 1] + w2 * t0[r + 1]) + w1 * t1[r]) + w0 * t0[r];
  } else {
#pragma unroll
    for (int q = 0; q < G::KX; ++q) {
      const float* t = ib + eoff[q];
      v[q] = ((w3 * __ldg(t) + w2 * __ldg(t - 1)) + w1 * __ldg(t - tw)) + w0 * __ldg(t - tw - 1);
    }
  }
  if (patnorm) {   // src/patch.cpp:264-266
    float acc = v[0];
#pragma unroll
    for (int q = 1; q < G::K; ++q) acc += v[q];
    const float sum = tree_finish<G::EXTRA>(acc, G::EXTRA ? v[G::KX - 1] : 0.f);
    const float mean = sum / (float)G::NE;
#pragma unroll
    for (int q = 0; q < G::KX; ++q) v[q] -= mean;
  }
  float sgx = 0.f, sgy = 0.f, egx = 0.f, egy = 0.f;
#pragma unroll
  for (int q = 0; q < G::KX; ++q) {
    const float tx = gxv[q] * v[q], ty = gyv[q] * v[q];
    if (q == 0) { sgx = tx; sgy = ty; }
    else if (q < G::K) { sgx += tx; sgy += ty; }
    else { egx = tx; egy = ty; }
  }
  sgx = tree_finish<G::EXTRA>(sgx, egx);
  sgy = tree_finish<G::EXTRA>(sgy, egy);
  return make_float2(sgx, sgy);
}

// The reset test of src/patch.cpp:185-194 (plus NaN => reset).
__device__ __forceinline__ bool must_reset(float px, float py, float stx, float sty, const SearchArgs& A) {
  const float ddx = stx - px, ddy = sty - py;
  return sqrtf(ddx * ddx + ddy * ddy) > A.outlier || px < A.lb || py < A.lb || px > A.ub_w || py > A.ub_h ||
         px != px || py != py;
}

enum : int { ST_RUN = 0, ST_WAIT = 1, ST_DONE = 2 };

template <int PS, bool FAST>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
k_inverse_search(const SearchArgs A) {
  using G = Geom<PS>;
  extern __shared__ uint4 smem_raw[];
  const int lane = threadIdx.x & 31;
  const int warp_in_block = threadIdx.x >> 5;
  // tile decomposition: blockIdx.x -> (pair, tile_x, tile_y); warp = grid column inside the tile
  const int tiles_x = (A.npw + kWarpsPerBlock - 1) / kWarpsPerBlock;
  const int tiles_y = (A.nph + 31) / 32;
  const int tiles = tiles_x * tiles_y;
  const int pair = blockIdx.x / tiles;
  const int t = blockIdx.x - pair * tiles;
  const int tile_x = t / tiles_y, tile_y = t - tile_x * tiles_y;
  const int gx_i = tile_x * kWarpsPerBlock + warp_in_block;
  if (gx_i >= A.npw) return;
  const int gy_raw = tile_y * 32 + lane;
  const bool valid = gy_raw < A.nph;
  const int gy_i = valid ? gy_raw : A.nph - 1;
  const int pid = gx_i * A.nph + gy_i;           // src/patch_grid.cpp:39-49: id = x*nph + y

  // per-warp cache: keys x [4 sets][32 lanes] uint4 (4 ways), keys y likewise, deltas [16][32] float2
  uint4* ckx = smem_raw + (size_t)warp_in_block * (kCacheBytesPerWarp / 16);
  uint4* cky = ckx + 4 * 32;
  float2* cdel = reinterpret_cast<float2*>(cky + 4 * 32);
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    ckx[s * 32 + lane] = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);   // NaN: never a key
    cky[s * 32 + lane] = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
  }

  const int2 fr = A.pairs[pair];
  const float* __restrict__ gxA = A.planes.base + (size_t)fr.x * A.planes.frame_stride + A.planes.plane_stride;
  const float* __restrict__ gyA = gxA + A.planes.plane_stride;
  const float* __restrict__ imB = A.planes.base + (size_t)fr.y * A.planes.frame_stride;
  const int tw = A.tmp_w;
  const int wmax = tw - PS / 2, hmax = A.height + 2 * A.pad - PS / 2;

  const int rx = gx_i * A.steps + A.offw, ry = gy_i * A.steps + A.offh;
  const float refx = (float)rx, refy = (float)ry;
  const int sub = lane >> 3, sl = lane & 7;
  int eoff[G::KX];
  element_offsets<PS>(sl, tw, eoff);

  // ---- template gradients and Hessian (src/patch.cpp:47-91), 4 patches per round ----------------
  float h00 = 0.f, h01 = 0.f, h11 = 0.f;
#pragma unroll 1
  for (int r = 0; r < 8; ++r) {
    const int ory = __shfl_sync(kFull, ry, r * 4 + sub);
    const int c = (ory + A.pad) * tw + (rx + A.pad);
    float sxx = 0.f, sxy = 0.f, syy = 0.f, exx = 0.f, exy = 0.f, eyy = 0.f;
#pragma unroll
    for (int k = 0; k < G::KX; ++k) {
      const float gx = __ldg(gxA + c + eoff[k]), gy = __ldg(gyA + c + eoff[k]);
      const float pxx = gx * gx, pxy = gx * gy, pyy = gy * gy;
      if (k == 0) { sxx = pxx; sxy = pxy; syy = pyy; }
      else if (k < G::K) { sxx += pxx; sxy += pxy; syy += pyy; }
      else { exx = pxx; exy = pxy; eyy = pyy; }
    }
    sxx = tree_finish<G::EXTRA>(sxx, exx);
    sxy = tree_finish<G::EXTRA>(sxy, exy);
    syy = tree_finish<G::EXTRA>(syy, eyy);
    const float t00 = __shfl_sync(kFull, sxx, (lane & 3) * 8);
    const float t01 = __shfl_sync(kFull, sxy, (lane & 3) * 8);
    const float t11 = __shfl_sync(kFull, syy, (lane & 3) * 8);
    if ((lane >> 2) == r) { h00 = t00; h01 = t01; h11 = t11; }
  }
  fix_singular(h00, h01, h11);
  LU2 lu;
  lu.factor(h00, h01, h11);

  // ---- initial displacement (src/patch_grid.cpp:108-119) ----------------------------------------
  float inx = 0.f, iny = 0.f;
  if (A.init_override) {
    const float2 v = reinterpret_cast<const float2*>(A.init_override)[(size_t)pair * A.np + pid];
    inx = v.x; iny = v.y;
  } else if (A.flow_prev) {
    const int fx = (int)floorf(refx / 2), fy = (int)floorf(refy / 2);
    const float2 v = reinterpret_cast<const float2*>(A.flow_prev + (size_t)pair * A.flow_prev_stride)[fy * (A.width / 2) + fx];
    inx = v.x * 2; iny = v.y * 2;
  }

  // ---- Patch::inverse_search_start (src/patch.cpp:119-154) --------------------------------------
  float ux = inx, uy = iny;
  float px = refx + ux, py = refy + uy;
  const float stx = px, sty = py;
  int k = 0;                       // counter_iter
  int state = ST_RUN;
  unsigned fifo = 0;               // 2-bit round-robin pointer per set
  unsigned n_sampled = 0, was_reset = 0, oob = 0;
  if (!valid) state = ST_DONE;
  else if (px < A.lb || py < A.lb || px > A.ub_w || py > A.ub_h || px != px || py != py) {
    state = ST_DONE; oob = 1;
  }

  unsigned trip = 0;
  int lonely = 0;
#pragma unroll 1
  for (;;) {
    // ---- scalar phase: consult the cache for the current position -----------------------------
    {
      const unsigned bx = __float_as_uint(px), by = __float_as_uint(py);
      const int set = ((bx & 1) << 1) | (by & 1);
      const uint4 kx = ckx[set * 32 + lane], ky = cky[set * 32 + lane];
      const bool m0 = (kx.x == bx) & (ky.x == by), m1 = (kx.y == bx) & (ky.y == by);
      const bool m2 = (kx.z == bx) & (ky.z == by), m3 = (kx.w == bx) & (ky.w == by);
      const bool hit = m0 | m1 | m2 | m3;
      const int way = m1 ? 1 : m2 ? 2 : m3 ? 3 : 0;
      const float2 d = cdel[((set << 2) + way) * 32 + lane];
      const bool run = state == ST_RUN;
      const bool over = k > A.iterations;                 // src/patch.cpp:199-201
      if (run & hit & !over) {
        ++k;
        ux = ux - d.x; uy = uy - d.y;                     // src/patch.cpp:179
        px = refx + ux; py = refy + uy;                   // src/patch.cpp:182
        if (FAST) { if (sqrtf(d.x * d.x + d.y * d.y) < A.conv_eps) state = ST_DONE; }
      } else if (run) {
        state = hit ? ST_DONE : ST_WAIT;   // a first-seen position still owes its reset test (service round)
      }
    }
    unsigned wait_mask = __ballot_sync(kFull, state == ST_WAIT);
    const unsigned run_mask = __ballot_sync(kFull, state == ST_RUN);
    if ((wait_mask | run_mask) == 0) break;
    ++trip;
    const int nwait = __popc(wait_mask);
    const bool force = (run_mask == 0) || ((trip & 7) == 0);
    if (run_mask != 0) lonely = 0;
    if (nwait == 0 || (nwait < 4 && !force)) continue;

    // ---- hand-off: the warp is down to a few lanes that miss on every pass ------------------------
    if (run_mask == 0 && nwait < 4) ++lonely;
    if (lonely > kLonelyLimit && A.stragglers) {
      if (state == ST_WAIT && k >= kHandoffMinIter) {
        const bool r = must_reset(px, py, stx, sty, A);
        if (r) { ux = inx; uy = iny; was_reset = 1; state = ST_DONE; }
        else if (k > A.iterations) state = ST_DONE;
        else {
          const int slot = atomicAdd(A.straggler_count, 1);
          if (slot < A.straggler_cap) {
            Straggler s; s.pair = pair; s.pid = pid; s.k = k; s.ux = ux; s.uy = uy;
            reinterpret_cast<Straggler*>(A.stragglers)[slot] = s;
            state = ST_DONE;
            k = -1;          // results and counters come from k_stragglers
          } else atomicSub(A.straggler_count, 1);
        }
      }
      wait_mask = __ballot_sync(kFull, state == ST_WAIT);
      if (wait_mask == 0) continue;
    }

    // ---- service rounds: four waiting patches at a time, one per 8-lane sub-warp ----------------
#pragma unroll 1
    while (wait_mask != 0 && (__popc(wait_mask) >= 4 || force)) {
      unsigned m = wait_mask;
      const int s0 = __ffs(m) - 1; m &= m - 1;
      const int s1 = m ? __ffs(m) - 1 : s0; m &= m - 1;
      const int s2 = m ? __ffs(m) - 1 : s0; m &= m - 1;
      const int s3 = m ? __ffs(m) - 1 : s0; m &= m - 1;
      wait_mask = m;
      const int src = sub == 0 ? s0 : sub == 1 ? s1 : sub == 2 ? s2 : s3;
      const float opx = __shfl_sync(kFull, px, src), opy = __shfl_sync(kFull, py, src);
      const int ory = __shfl_sync(kFull, ry, src);
      const int c = (ory + A.pad) * tw + (rx + A.pad);
      float gxv[G::KX], gyv[G::KX];
#pragma unroll
      for (int q = 0; q < G::KX; ++q) { gxv[q] = __ldg(gxA + c + eoff[q]); gyv[q] = __ldg(gyA + c + eoff[q]); }
      const float2 sg = sample_and_project<PS>(imB, tw, A.pad, wmax, hmax, opx, opy, sl, A.patnorm, eoff, gxv, gyv);

      // hand the two sums to the owning lanes
      const int slot = lane == s0 ? 0 : lane == s1 ? 1 : lane == s2 ? 2 : 3;
      const float g0 = __shfl_sync(kFull, sg.x, slot * 8), g1 = __shfl_sync(kFull, sg.y, slot * 8);
      const bool mine = (state == ST_WAIT) && (lane == s0 || lane == s1 || lane == s2 || lane == s3);
      if (mine) {
        if (must_reset(px, py, stx, sty, A)) {             // src/patch.cpp:185-194
          ux = inx; uy = iny; was_reset = 1; state = ST_DONE;
        } else if (k > A.iterations) {                     // src/patch.cpp:199-201
          state = ST_DONE;
        } else {
          const float2 d = lu.solve(g0, g1);               // src/patch.cpp:176
          const unsigned bx = __float_as_uint(px), by = __float_as_uint(py);
          const int set = ((bx & 1) << 1) | (by & 1);
          const int way = (fifo >> (2 * set)) & 3;
          fifo = (fifo & ~(3u << (2 * set))) | (((way + 1) & 3) << (2 * set));
          reinterpret_cast<unsigned*>(ckx)[(set * 32 + lane) * 4 + way] = bx;
          reinterpret_cast<unsigned*>(cky)[(set * 32 + lane) * 4 + way] = by;
          cdel[((set << 2) + way) * 32 + lane] = d;
          ++n_sampled;
          ++k;
          ux = ux - d.x; uy = uy - d.y;
          px = refx + ux; py = refy + uy;
          state = ST_RUN;
          if (FAST) { if (sqrtf(d.x * d.x + d.y * d.y) < A.conv_eps) state = ST_DONE; }
        }
      }
    }
  }

  // ---- results ----------------------------------------------------------------------------------
  const bool mine_out = valid && k >= 0;
  if (mine_out) {
    reinterpret_cast<float2*>(A.u_out)[(size_t)pair * A.np + pid] = make_float2(ux, uy);
    if (A.iters_out) A.iters_out[(size_t)pair * A.np + pid] = k;
  }
  if (A.counters) {
    unsigned long long it = mine_out ? (unsigned long long)k : 0ull;
    unsigned ns = valid ? n_sampled : 0u, nr = was_reset, no = oob;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      it += __shfl_xor_sync(kFull, it, o);
      ns += __shfl_xor_sync(kFull, ns, o);
      nr += __shfl_xor_sync(kFull, nr, o);
      no += __shfl_xor_sync(kFull, no, o);
    }
    if (lane == 0) {
      atomicAdd(&A.counters[0], it);
      atomicAdd(&A.counters[1], (unsigned long long)ns);
      atomicAdd(&A.counters[2], (unsigned long long)nr);
      atomicAdd(&A.counters[3], (unsigned long long)no);
    }
  }
}

// Patches handed off by k_inverse_search: every pass samples, four patches per warp (one per
// 8-lane sub-warp, template gradients in registers).  Same arithmetic, same order of tests.
template <int PS, bool FAST>
__global__ void __launch_bounds__(128) k_stragglers(const SearchArgs A) {
  using G = Geom<PS>;
  const int lane = threadIdx.x & 31, sub = lane >> 3, sl = lane & 7;
  const int gwarp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nwarps = gridDim.x * (blockDim.x >> 5);
  const int count = min(*A.straggler_count, A.straggler_cap);
  const Straggler* list = reinterpret_cast<const Straggler*>(A.stragglers);
  const int tw = A.tmp_w;
  const int wmax = tw - PS / 2, hmax = A.height + 2 * A.pad - PS / 2;
  int eoff[G::KX];
  element_offsets<PS>(sl, tw, eoff);
  unsigned long long it_total = 0;
  unsigned ns_total = 0, nr_total = 0;
#pragma unroll 1
  for (int base = gwarp * 4; base < count; base += nwarps * 4) {
    const bool have = base + sub < count;
    const Straggler e = list[have ? base + sub : base];
    const int2 fr = A.pairs[e.pair];
    const float* __restrict__ gxA = A.planes.base + (size_t)fr.x * A.planes.frame_stride + A.planes.plane_stride;
    const float* __restrict__ gyA = gxA + A.planes.plane_stride;
    const float* __restrict__ imB = A.planes.base + (size_t)fr.y * A.planes.frame_stride;
    const int gx_i = e.pid / A.nph, gy_i = e.pid - gx_i * A.nph;
    const int rx = gx_i * A.steps + A.offw, ry = gy_i * A.steps + A.offh;
    const float refx = (float)rx, refy = (float)ry;
    const int c = (ry + A.pad) * tw + (rx + A.pad);
    float gxv[G::KX], gyv[G::KX];
    float sxx = 0.f, sxy = 0.f, syy = 0.f, exx = 0.f, exy = 0.f, eyy = 0.f;
#pragma unroll
    for (int k = 0; k < G::KX; ++k) {
      const float gx = __ldg(gxA + c + eoff[k]), gy = __ldg(gyA + c + eoff[k]);
      gxv[k] = gx; gyv[k] = gy;
      const float pxx = gx * gx, pxy = gx * gy, pyy = gy * gy;
      if (k == 0) { sxx = pxx; sxy = pxy; syy = pyy; }
      else if (k < G::K) { sxx += pxx; sxy += pxy; syy += pyy; }
      else { exx = pxx; exy = pxy; eyy = pyy; }
    }
    float h00 = tree_finish<G::EXTRA>(sxx, exx);
    const float h01 = tree_finish<G::EXTRA>(sxy, exy);
    float h11 = tree_finish<G::EXTRA>(syy, eyy);
    fix_singular(h00, h01, h11);
    LU2 lu;
    lu.factor(h00, h01, h11);
    float inx = 0.f, iny = 0.f;
    if (A.init_override) {
      const float2 v = reinterpret_cast<const float2*>(A.init_override)[(size_t)e.pair * A.np + e.pid];
      inx = v.x; iny = v.y;
    } else if (A.flow_prev) {
      const int fx = (int)floorf(refx / 2), fy = (int)floorf(refy / 2);
      const float2 v = reinterpret_cast<const float2*>(A.flow_prev + (size_t)e.pair * A.flow_prev_stride)[fy * (A.width / 2) + fx];
      inx = v.x * 2; iny = v.y * 2;
    }
    const float stx = refx + inx, sty = refy + iny;
    float ux = e.ux, uy = e.uy;
    float px = refx + ux, py = refy + uy;
    int k = e.k;
    bool active = have;
    unsigned was_reset = 0, ns = 0;
#pragma unroll 1
    while (__any_sync(kFull, active)) {
      const float2 sg = sample_and_project<PS>(imB, tw, A.pad, wmax, hmax, px, py, sl, A.patnorm, eoff, gxv, gyv);
      if (active) {
        if (must_reset(px, py, stx, sty, A)) { ux = inx; uy = iny; was_reset = 1; active = false; }
        else if (k > A.iterations) active = false;
        else {
          const float2 d = lu.solve(sg.x, sg.y);
          ++k; ++ns;
          ux = ux - d.x; uy = uy - d.y;
          px = refx + ux; py = refy + uy;
          if (FAST) { if (sqrtf(d.x * d.x + d.y * d.y) < A.conv_eps) active = false; }
        }
      }
    }
    if (have && sl == 0) {
      reinterpret_cast<float2*>(A.u_out)[(size_t)e.pair * A.np + e.pid] = make_float2(ux, uy);
      if (A.iters_out) A.iters_out[(size_t)e.pair * A.np + e.pid] = k;
      it_total += (unsigned long long)k; ns_total += ns; nr_total += was_reset;
    }
  }
  if (A.counters) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      it_total += __shfl_xor_sync(kFull, it_total, o);
      ns_total += __shfl_xor_sync(kFull, ns_total, o);
      nr_total += __shfl_xor_sync(kFull, nr_total, o);
    }
    if (lane == 0 && (it_total | ns_total | nr_total)) {
      atomicAdd(&A.counters[0], it_total);
      atomicAdd(&A.counters[1], (unsigned long long)ns_total);
      atomicAdd(&A.counters[2], (unsigned long long)nr_total);
    }
  }
}

template <int PS, bool FAST>
int launch_ps(const SearchArgs& a, cudaStream_t s, int smem_pad) {
  const size_t smem = (size_t)kWarpsPerBlock * kCacheBytesPerWarp + (size_t)smem_pad;
  cudaFuncSetAttribute(k_inverse_search<PS, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int tiles_x = (a.npw + kWarpsPerBlock - 1) / kWarpsPerBlock, tiles_y = (a.nph + 31) / 32;
  const unsigned blocks = (unsigned)(tiles_x * tiles_y) * (unsigned)a.n_pairs;
  int n = 1;
  if (a.stragglers) cudaMemsetAsync(a.straggler_count, 0, sizeof(int), s);
  k_inverse_search<PS, FAST><<<blocks, kWarpsPerBlock * 32, smem, s>>>(a);
  if (a.stragglers) { k_stragglers<PS, FAST><<<148 * 4, 128, 0, s>>>(a); ++n; }
  return n;
}

}  // namespace

int launch_inverse_search(const SearchArgs& a, int patch_size, cudaStream_t s) {
  static int smem_pad = -1;
  if (smem_pad < 0) { const char* e = getenv("DIS_SMEM_PAD_KB"); smem_pad = e ? atoi(e) * 1024 : 0; }
  const bool fast = a.conv_eps > 0.f;
#define DIS_CASE(PS) case PS: return fast ? launch_ps<PS, true>(a, s, smem_pad) : launch_ps<PS, false>(a, s, smem_pad);
  switch (patch_size) {
    DIS_CASE(4) DIS_CASE(6) DIS_CASE(8) DIS_CASE(10) DIS_CASE(12) DIS_CASE(16)
    default: return -1;
  }
#undef DIS_CASE
}

}  // namespace dis

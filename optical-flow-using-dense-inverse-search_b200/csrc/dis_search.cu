// dis_search.cu — the per-level inverse search on sm_100a.
//
// Replaces, for every patch of a level and every pair of a batch:
//   PatchGrid::PatchGrid          src/patch_grid.cpp:17-51   grid geometry, x-major patch ids
//   Patch::init_patch             src/patch.cpp:31-91        template gradients + 2x2 Hessian
//   PatchGrid::patch_init_from_prev_flow  src/patch_grid.cpp:108-119
//   Patch::inverse_search(_start) src/patch.cpp:119-203      the iteration loop
//   Patch::get_patch_second_image src/patch.cpp:207-267      4-tap resample + mean normalisation
//
// Design (B200-first, not a translation).  A level runs as three kernels on one stream:
//  1. k_converge — the first kN1 passes of every patch visit almost only NEW positions, so they are
//     computed without memoisation by a sub-group of SG lanes per patch (SG = 4 for ps <= 8, else 8),
//     32/SG patches per warp in lock step.  Lane j of a sub-group owns the elements e with e mod 8 in
//     {j, j+4} (SG 4) or {j} (SG 8), which makes Eigen's SSE reduction tree (App. A.3) exactly "serial
//     over the lane's elements, then [in-lane add | xor-shuffle 4], xor-shuffle 2, 1".  Template
//     gradients, LU factors and (ps 8) the 9x9 tap window live in registers — a patch keeps the same
//     integer tap base for nearly all its passes — so a pass is pure fp32 arithmetic.  The kernel leaves
//     {u, k, LU} and a ready-made 16-entry cache image per patch in HBM (288 B per patch).
//  2. k_inverse_search (the walk) — EXACT memoisation: within a patch, delta_u is a pure function of
//     the bit pattern of the sampling position (H, the template and the image are constant), and so
//     is the outcome of the reset test.  After convergence the fp32 orbit of a patch jitters among a
//     few positions, so ~99 % of the remaining ~980 passes are cache hits.  PERSISTENT warps, one LANE
//     per patch: a lane carries the scalar state of its patch and a 16-entry cache {pos bits ->
//     delta_u} in shared memory (4 sets x 4 ways, one uint4 per entry); a hit costs ~14 scalar
//     instructions instead of ~1000.  Way prediction: an entry is inserted at the way named by bit 1 of
//     x and y, so the hot loop probes ONE slot (a single LDS.128); the other three ways are probed once
//     per chunk of kChunk passes, for the lanes that fell out of the fast loop.  Lanes that miss wait;
//     the warp serves them in rounds (one patch per sub-group; ps 8 uses 8-lane sub-groups here, which
//     fit 80 registers and so 24 warps per SM).  A finished lane is re-bound to the next patch index of
//     a global counter, so a warp never idles behind its slowest patch.  Results are bit-identical to
//     running every pass: nothing is approximated, only recomputation is skipped.
//  3. k_stragglers — patches whose orbit never settles (~0.6 %) sample on every pass; after
//     kStreakLimit such passes the walk hands them over, and they finish here with template and tap
//     window in registers, 32/SG per warp.
//  (tools/micro/ keeps the measured alternatives: single-kernel designs with tile-bound warps, and
//  warp-specialised walker/sampler blocks with a shared-memory request queue; DESIGN.md has the numbers.)
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <type_traits>

#include "dis_kernels.cuh"

namespace dis {
namespace {

constexpr int kSlots = kWalkSeeds;    // cache entries per patch: 4 sets (bit 0 of x, bit 0 of y) x kSlots/4 ways
static_assert(kSlots == 16 || kSlots == 32, "cache geometry");
// Ways per set.  4: 4 sets x 4 ways, the predicted way is probed in the fast loop and the other three once per chunk.
// 1 (-DDIS_WAYS=1, direct-mapped): no second probe and a one-store insertion, but two hot positions of a patch that share a
// slot evict each other on every visit: 41.2 instead of 32.4 sampled passes per patch and 35.9 instead of 31.4 ms per
// bench step (profiles/r2_direct_mapped_ab.log; tools/trace predicted it once hand-offs were counted).  Kept as an experiment.
#ifndef DIS_WAYS
#define DIS_WAYS 4
#endif
constexpr int kWays = DIS_WAYS == 1 ? 1 : kSlots / 4;
constexpr int kWayBits = kSlots == 32 ? 3 : 2;
// Slot predicted for a position.  16 slots: (5x + y) mod 16 — the best of the lattice hashes (ax + by) mod 16 on the
// position traces of the bench workload: half as many hits in a non-predicted way as "low 2 bits of x above low 2 bits of y" (the orbit of a
// patch spreads over more ulps in y, whose coordinates are smaller, than in x).  32 slots: low 3 bits of x above low 2 of y.
__device__ __forceinline__ int slot_of(unsigned bx, unsigned by) {
  return kSlots == 16 ? (int)((5u * bx + by) & 15u) : (int)(((bx & (kSlots / 4 - 1)) << 2) | (by & 3));
}
// &cache[slot_of(bx, by) * 32] for the walk's layout (slot s of a lane at cache[s * 32]) in three instructions:
// the slot number is formed in the top four bits, where the wrap-around of the shifts is the "mod 16".
__device__ __forceinline__ const uint4* slot_ptr32(const uint4* cache, unsigned bx, unsigned by) {
  static_assert(kSlots == 16 || kSlots == 32, "");
  if (kSlots != 16) return cache + slot_of(bx, by) * 32;
  const unsigned top = bx * 0x50000000u + (by << 28);
  return reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(cache) + (top >> 19));
}
// XOR distance, in slot index space, between way 0 and way w of a set (way bit 0 = bit 1 of y, bits 1.. = bits 1.. of x)
__host__ __device__ constexpr int way_xor(int w) { return ((w & 1) << 1) | ((w >> 1) << 3); }
#ifndef DIS_PACKED
#define DIS_PACKED 1      // packed fp32x2 arithmetic in the sampling code of k_converge / k_stragglers (ps 8)
#endif
#ifndef DIS_MOVED_ROLLING
#define DIS_MOVED_ROLLING 1   // the early hand-off rule of k_converge looks at the last three executed passes (0: only passes 21-23, round 1)
#endif
#ifndef DIS_CHUNK
#define DIS_CHUNK 6
#endif
#ifndef DIS_AGE
#define DIS_AGE 4
#endif
#ifndef DIS_HEAVY
#define DIS_HEAVY 100000
#endif
#ifndef DIS_EARLY
#define DIS_EARLY 0
#endif
#ifndef DIS_STRAG_SG8
#define DIS_STRAG_SG8 0
#endif
#ifndef DIS_STRAG_BLOCKS
#define DIS_STRAG_BLOCKS 4
#endif
#ifndef DIS_N1
#define DIS_N1 24
#endif
#ifndef DIS_REBIND_MIN
#define DIS_REBIND_MIN 4
#endif
constexpr int kChunk = DIS_CHUNK;           // fast-probe passes between two warp-level decisions
constexpr unsigned kFull = 0xffffffffu;
constexpr int kHandoffMinIter = 48;   // never hand off a patch that is still in its convergence phase

template <int PS, int SG>
struct Geom {
  static constexpr int NE = PS * PS;
  static constexpr int K = NE / 8;                       // full rounds of 8 (two SSE packets)
  static constexpr bool EXTRA = (NE % 8) >= 4;           // one more packet of 4
  static constexpr int H = 8 / SG;                       // residues mod 8 owned by a lane (1 or 2)
  static constexpr int NPL = K * H + (EXTRA ? 1 : 0);    // elements per lane
  static constexpr int G = 32 / SG;                      // patches per warp round
  static constexpr bool POW2 = (NE & (NE - 1)) == 0;
};

// Element index of the lane's i-th element (sl = lane within the sub-group).
template <int PS, int SG>
__device__ __forceinline__ int element_of(int i, int sl) {
  using G = Geom<PS, SG>;
  if (i < G::K) return 8 * i + sl;
  if (SG == 4 && i < 2 * G::K) return 8 * (i - G::K) + sl + 4;
  return 8 * G::K + (sl & 3);
}

// Eigen-3.3 redux of the sub-group's ps^2 values, v = the lane's elements in
// element_of order.  Returns the sum in every lane of the sub-group.
template <int PS, int SG>
__device__ __forceinline__ float tree_sum(const float (&v)[Geom<PS, SG>::NPL]) {
  using G = Geom<PS, SG>;
  float acc = v[0];
#pragma unroll
  for (int i = 1; i < G::K; ++i) acc += v[i];
  if (SG == 4) {
    float hi = v[G::K];
#pragma unroll
    for (int i = 1; i < G::K; ++i) hi += v[G::K + i];
    acc = acc + hi;                                      // packet 0 + packet 1
  } else {
    acc = acc + __shfl_xor_sync(kFull, acc, 4);
  }
  if (G::EXTRA) acc = acc + v[G::NPL - 1];
  acc = acc + __shfl_xor_sync(kFull, acc, 2);
  acc = acc + __shfl_xor_sync(kFull, acc, 1);
  return acc;
}

// Hand-off record walk -> k_stragglers, 32 bytes.  k doubles as the "published" marker: the list is cleared to
// 0xff (k = -1) before a launch and k is stored last, after a fence, because k_stragglers may run CONCURRENTLY
// with the walk and consume entries as they appear.
struct Straggler { int pair, pid; float ux, uy; int k; int pad[3]; };
static_assert(sizeof(Straggler) == kStragglerBytes, "hand-off list layout");
__device__ __forceinline__ int ld_volatile(const int* p) { return *reinterpret_cast<const volatile int*>(p); }

// Hand-off record between the two phases of a level (k_converge -> k_inverse_search), 32 bytes.
// k < 0: the patch finished in the first phase (its u is already in u_out).
struct WalkState { float ux, uy; int k, swap; float p, q, l, u11; };
constexpr int kN1 = DIS_N1;              // passes of the convergence phase (every one samples)
static_assert(kSlots == kWalkSeeds, "hand-off layout");
constexpr int kSeed = 16;             // the last kSeed passes of the convergence phase seed the walk phase's cache

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

// Offsets (in pixels of a padded plane of width tw) of the lane's elements relative to the patch
// centre; kept in registers for patch sizes other than 8 (ps 8 uses its column structure instead).
template <int PS, int SG>
__device__ __forceinline__ void element_offsets(int sl, int tw, int (&eoff)[Geom<PS, SG>::NPL]) {
  using G = Geom<PS, SG>;
#pragma unroll
  for (int i = 0; i < G::NPL; ++i) {
    const int idx = element_of<PS, SG>(i, sl);
    const int row = idx / PS, col = idx - row * PS;
    eoff[i] = (row - PS / 2) * tw + (col - PS / 2);
  }
}

// The lane's template gradients (gx, gy interleaved) of the patch centred at plane offset c
// (Patch::get_gradients_on_patch, src/patch.cpp:47-73; the intensity template is never used, F2).
template <int PS, int SG>
__device__ __forceinline__ void load_template(const float2* __restrict__ gxy, int c, int tw, int sl,
                                              const int (&eoff)[Geom<PS, SG>::NPL],
                                              float2 (&T)[Geom<PS, SG>::NPL]) {
  using G = Geom<PS, SG>;
  if (PS == 8 && SG == 4 && DIS_PACKED) {
    // 32-bit loads: the packed arithmetic pairs gx of the lane's two columns (Template8P), and a 64-bit load would tie
    // gx to gy in a register pair, which ptxas then re-packs with two moves per use on every pass
    const float* col = reinterpret_cast<const float*>(gxy + c + (sl - 4) - 4 * tw);
#pragma unroll
    for (int h = 0; h < G::H; ++h)
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        T[8 * h + r].x = __ldg(col + 2 * (4 * h + r * tw));
        T[8 * h + r].y = __ldg(col + 2 * (4 * h + r * tw) + 1);
      }
  } else if (PS == 8) {
    const float2* col = gxy + c + (sl - 4) - 4 * tw;
#pragma unroll
    for (int h = 0; h < G::H; ++h)
#pragma unroll
      for (int r = 0; r < 8; ++r) T[8 * h + r] = __ldg(col + 4 * h + r * tw);
  } else {
#pragma unroll
    for (int i = 0; i < G::NPL; ++i) T[i] = __ldg(gxy + c + eoff[i]);
  }
}

// compute_hessian_matrix (src/patch.cpp:75-91): the three products-and-sums, in every lane.
template <int PS, int SG>
__device__ __forceinline__ float3 hessian_sums(const float2 (&T)[Geom<PS, SG>::NPL]) {
  using G = Geom<PS, SG>;
  float a[G::NPL], b[G::NPL], c[G::NPL];
#pragma unroll
  for (int i = 0; i < G::NPL; ++i) { a[i] = T[i].x * T[i].x; b[i] = T[i].x * T[i].y; c[i] = T[i].y * T[i].y; }
  return make_float3(tree_sum<PS, SG>(a), tree_sum<PS, SG>(b), tree_sum<PS, SG>(c));
}

// Tap base of Patch::get_patch_second_image (src/patch.cpp:222-234), clamped into the padded plane so
// that a position the reset test is about to reject cannot address outside it.
template <int PS>
__device__ __forceinline__ int2 tap_base(float opx, float opy, int pad, int wmax, int hmax) {
  int bxi = (int)ceilf(opx + .00001f) + pad;
  int byi = (int)ceilf(opy + .00001f) + pad;
  bxi = min(max(bxi, PS / 2 + 1), wmax);
  byi = min(max(byi, PS / 2 + 1), hmax);
  return make_int2(bxi, byi);
}

// Mean normalisation (src/patch.cpp:264-266) of the lane's samples v, then the two
// products-and-sums of src/patch.cpp:171-172 with the template gradients T.
template <int PS, int SG>
__device__ __forceinline__ float2 normalise_and_project(float (&v)[Geom<PS, SG>::NPL], int patnorm,
                                                        const float2 (&T)[Geom<PS, SG>::NPL]) {
  using G = Geom<PS, SG>;
  if (patnorm) {   // x / 2^n == x * 2^-n bit for bit
    const float sum = tree_sum<PS, SG>(v);
    const float mean = G::POW2 ? sum * (1.0f / (float)G::NE) : sum / (float)G::NE;
#pragma unroll
    for (int i = 0; i < G::NPL; ++i) v[i] -= mean;
  }
  float px[G::NPL], py[G::NPL];
#pragma unroll
  for (int i = 0; i < G::NPL; ++i) { px[i] = T[i].x * v[i]; py[i] = T[i].y * v[i]; }
  return make_float2(tree_sum<PS, SG>(px), tree_sum<PS, SG>(py));
}

// Patch::get_patch_second_image (src/patch.cpp:207-267) for the lane's elements, followed by
// the two products-and-sums of src/patch.cpp:171-172.  T: the lane's template gradients.
// Returns (sum gx*r, sum gy*r), identical in all lanes of the sub-group.
template <int PS, int SG>
__device__ __forceinline__ float2 sample_and_project(const float* __restrict__ imB, int tw, int pad, int wmax,
                                                     int hmax, float opx, float opy, int sl, int patnorm,
                                                     const int (&eoff)[Geom<PS, SG>::NPL],
                                                     const float2 (&T)[Geom<PS, SG>::NPL]) {
  using G = Geom<PS, SG>;
  const float fl = floorf(opx), fk = floorf(opy);
  const float a = opx - fl, b = opy - fk;
  const float w0 = (1 - a) * (1 - b), w1 = a * (1 - b), w2 = b * (1 - a), w3 = a * b;
  // tap base; clamped so that a position about to be rejected by the reset test (out of bounds,
  // NaN) cannot address outside the plane — its result is discarded by the owner
  const int2 tb = tap_base<PS>(opx, opy, pad, wmax, hmax);
  const float* __restrict__ ib = imB + tb.y * tw + tb.x;
  float v[G::NPL];
  if (PS == 8) {
    // the lane owns whole columns: consecutive rows share taps, 9 x 2 loads per column instead of 8 x 4
#pragma unroll
    for (int h = 0; h < G::H; ++h) {
      const float* col = ib + (sl + 4 * h - 4) - 5 * tw;
      float t0[9], t1[9];
#pragma unroll
      for (int r = 0; r < 9; ++r) { t0[r] = __ldg(col + r * tw - 1); t1[r] = __ldg(col + r * tw); }
#pragma unroll
      for (int r = 0; r < 8; ++r) v[8 * h + r] = ((w3 * t1[r + 1] + w2 * t0[r + 1]) + w1 * t1[r]) + w0 * t0[r];
    }
  } else {
#pragma unroll
    for (int i = 0; i < G::NPL; ++i) {
      const float* t = ib + eoff[i];
      v[i] = ((w3 * __ldg(t) + w2 * __ldg(t - 1)) + w1 * __ldg(t - tw)) + w0 * __ldg(t - tw - 1);
    }
  }
  return normalise_and_project<PS, SG>(v, patnorm, T);
}

// ps == 8 with the tap window of the patch held in REGISTERS (k_converge): a patch keeps the same
// integer tap base for almost all of its passes (1.3 distinct bases per patch on the Sintel-shape
// workload), so the 9 x 2 x H taps of the lane's columns are loaded once and only the weights change.
template <int SG>
struct Window8 {
  static constexpr int H = 8 / SG;
  float t0[H][9], t1[H][9];
  int bx, by;                                  // the (clamped) tap base the registers hold; -1: nothing loaded
  // same taps from a shared-memory tile; w0 = address of tap (bxi - 5, byi - 5), rows `stride` floats apart
  __device__ __forceinline__ void load_tile(const float* w0, int stride, int sl, int bxi, int byi) {
    bx = bxi; by = byi;
#pragma unroll
    for (int h = 0; h < H; ++h)
#pragma unroll
      for (int r = 0; r < 9; ++r) { t0[h][r] = w0[r * stride + sl + 4 * h]; t1[h][r] = w0[r * stride + sl + 4 * h + 1]; }
  }
  __device__ __forceinline__ void load(const float* __restrict__ imB, int tw, int sl, int bxi, int byi) {
    bx = bxi; by = byi;
#pragma unroll
    for (int h = 0; h < H; ++h) {
      const float* col = imB + byi * tw + bxi + (sl + 4 * h - 4) - 5 * tw;
#pragma unroll
      for (int r = 0; r < 9; ++r) { t0[h][r] = __ldg(col + r * tw - 1); t1[h][r] = __ldg(col + r * tw); }
    }
  }
};

template <int SG>
__device__ __forceinline__ float2 sample_window8(const Window8<SG>& Wn, float opx, float opy, int patnorm,
                                                 const float2 (&T)[Geom<8, SG>::NPL]) {
  using G = Geom<8, SG>;
  const float fl = floorf(opx), fk = floorf(opy);
  const float a = opx - fl, b = opy - fk;
  const float w0 = (1 - a) * (1 - b), w1 = a * (1 - b), w2 = b * (1 - a), w3 = a * b;
  float v[G::NPL];
#pragma unroll
  for (int h = 0; h < G::H; ++h)
#pragma unroll
    for (int r = 0; r < 8; ++r)
      v[8 * h + r] = ((w3 * Wn.t1[h][r + 1] + w2 * Wn.t0[h][r + 1]) + w1 * Wn.t1[h][r]) + w0 * Wn.t0[h][r];
  return normalise_and_project<8, SG>(v, patnorm, T);
}

// Patch sizes other than 8: the four taps of each of the lane's elements, held in REGISTERS while the integer tap base
// stays the same (it does for ~99 % of a patch's passes), so that a pass of k_converge / k_stragglers is arithmetic only.
// The lane <-> element mapping stays Eigen's (e mod 8), which scatters a lane's elements over the patch: no taps are
// shared between them, so this is 4 x NPL registers (72 at ps 12) — affordable up to ps 12.
template <int PS, int SG>
struct TapWindow {
  static constexpr int NPL = Geom<PS, SG>::NPL;
  float t[NPL][4];
  int bx, by;
  __device__ __forceinline__ void load(const float* __restrict__ imB, int tw, const int (&eoff)[NPL], int bxi, int byi) {
    bx = bxi; by = byi;
    const float* __restrict__ ib = imB + byi * tw + bxi;
#pragma unroll
    for (int i = 0; i < NPL; ++i) {
      const float* p = ib + eoff[i];
      t[i][0] = __ldg(p); t[i][1] = __ldg(p - 1); t[i][2] = __ldg(p - tw); t[i][3] = __ldg(p - tw - 1);
    }
  }
};
template <int PS, int SG>
constexpr bool kUseTapWindow = PS != 8 && 4 * Geom<PS, SG>::NPL <= 72;

template <int PS, int SG>
__device__ __forceinline__ float2 sample_tapwindow(const TapWindow<PS, SG>& Wn, float opx, float opy, int patnorm,
                                                   const float2 (&T)[Geom<PS, SG>::NPL]) {
  using G = Geom<PS, SG>;
  const float fl = floorf(opx), fk = floorf(opy);
  const float a = opx - fl, b = opy - fk;
  const float w0 = (1 - a) * (1 - b), w1 = a * (1 - b), w2 = b * (1 - a), w3 = a * b;
  float v[G::NPL];
#pragma unroll
  for (int i = 0; i < G::NPL; ++i) v[i] = ((w3 * Wn.t[i][0] + w2 * Wn.t[i][1]) + w1 * Wn.t[i][2]) + w0 * Wn.t[i][3];
  return normalise_and_project<PS, SG>(v, patnorm, T);
}

// ---- packed fp32 arithmetic -------------------------------------------------------------------------------------
// sm_100a has packed fp32 instructions (PTX mul/fma.rn.f32x2, SASS FMUL2 / FFMA2): two IEEE round-to-nearest fp32
// operations on a 64-bit register pair in ONE issue slot (the FP32 pipe is busy two cycles, so the flop rate is that
// of the scalar forms; what is saved is instruction issue, which is what bounds these kernels:
// tools/micro/f32x2_probe.cu).  Each half is bit-identical to the scalar operation — PROVIDED nothing is contracted:
// ptxas 12.9 fuses mul.rn.f32x2 + add.rn.f32x2 into one FFMA2 in spite of the explicit rounding modifiers and of
// --fmad=false (it honours both for scalar code), and it also sees through fma(a, 1.0, b) and fma(a, b, -0.0).  So
// every packed addition is written fma(a, ONE, b) with ONE = {1.0f, 1.0f} read from the kernel arguments (opaque to
// ptxas): a * 1 is exact, the sum is rounded once, and an FFMA2 can be contracted with nothing.  The SASS of a build
// must show FMUL2 : FFMA2 = 48 : 53 per sample (checked by tests/test_abi.py::test_packed_arithmetic_is_not_contracted).
// ps 8 with 4 lanes per patch pairs the lane's two columns (elements r and 8 + r of the lane), which is also the
// split of Eigen's two SSE packets in tree_sum.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ float lo2(f32x2 a) { float l, h; asm("mov.b64 {%0, %1}, %2;" : "=f"(l), "=f"(h) : "l"(a)); return l; }
__device__ __forceinline__ float hi2(f32x2 a) { float l, h; asm("mov.b64 {%0, %1}, %2;" : "=f"(l), "=f"(h) : "l"(a)); return h; }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b, f32x2 one) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(one), "l"(b)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

// Eigen redux of the 64 values held as 8 pairs {v[r], v[8 + r]} by the 4 lanes of a sub-group: identical to
// tree_sum<8, 4> (serial over each packet's 8 values, packet 0 + packet 1, then xor-shuffles 2, 1).
__device__ __forceinline__ float tree_sum_p(const f32x2 (&v)[8], f32x2 one) {
  f32x2 s = v[0];
#pragma unroll
  for (int r = 1; r < 8; ++r) s = add2(s, v[r], one);
  float acc = lo2(s) + hi2(s);
  acc = acc + __shfl_xor_sync(kFull, acc, 2);
  acc = acc + __shfl_xor_sync(kFull, acc, 1);
  return acc;
}

// Template gradients of a ps-8 patch in packed form: x[r] = {gx[r], gx[8 + r]}, y[r] likewise.
struct Template8P {
  f32x2 x[8], y[8];
  __device__ __forceinline__ void set(const float2 (&T)[16]) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      x[r] = pk2(T[r].x, T[8 + r].x); y[r] = pk2(T[r].y, T[8 + r].y);
      // pin the pairs in registers: without this the compiler keeps the (gx, gy) layout of the loads alive and
      // re-packs the operands with two moves per multiplication on every pass
      asm volatile("" : "+l"(x[r]), "+l"(y[r]));
    }
  }
};

// Window8<4> with the taps of the lane's two columns paired: t0[r] = {tap(col - 1, r) of column sl, of column sl + 4}.
struct Window8P {
  f32x2 t0[9], t1[9];
  int bx, by;
  __device__ __forceinline__ void load_tile(const float* w0, int stride, int sl, int bxi, int byi) {
    bx = bxi; by = byi;
#pragma unroll
    for (int r = 0; r < 9; ++r) {
      t0[r] = pk2(w0[r * stride + sl], w0[r * stride + sl + 4]);
      t1[r] = pk2(w0[r * stride + sl + 1], w0[r * stride + sl + 5]);
    }
  }
  __device__ __forceinline__ void load(const float* __restrict__ imB, int tw, int sl, int bxi, int byi) {
    bx = bxi; by = byi;
    const float* col = imB + byi * tw + bxi + (sl - 4) - 5 * tw;
#pragma unroll
    for (int r = 0; r < 9; ++r) {
      t0[r] = pk2(__ldg(col + r * tw - 1), __ldg(col + r * tw + 3));
      t1[r] = pk2(__ldg(col + r * tw), __ldg(col + r * tw + 4));
    }
  }
};

// sample_window8<4> + normalise_and_project<8, 4> on pairs: same operations, same order, half the instructions.
// The packed template gradients of a ps-8 patch in shared memory: element r of the lane at base[r * 32] (x) and
// base[(8 + r) * 32] (y); base already points at the lane's column.  Holding them there instead of in 32 registers
// lets k_converge run 6 warps per scheduler instead of 4 — the kernel is bound by dependent-issue latency, not by
// the issue rate or the FP32 pipe.
struct Template8S {
  const f32x2* base;
  __device__ __forceinline__ f32x2 x(int r) const { return base[r * 32]; }
  __device__ __forceinline__ f32x2 y(int r) const { return base[(8 + r) * 32]; }
};
struct Template8R {   // register-resident variant with the same interface
  const Template8P* t;
  __device__ __forceinline__ f32x2 x(int r) const { return t->x[r]; }
  __device__ __forceinline__ f32x2 y(int r) const { return t->y[r]; }
};
template <typename TA>
__device__ __forceinline__ float2 sample_window8p(const Window8P& Wn, float opx, float opy, int patnorm, const TA T, float onef) {
  const float fl = floorf(opx), fk = floorf(opy);
  const float a = opx - fl, b = opy - fk;
  const float w0 = (1 - a) * (1 - b), w1 = a * (1 - b), w2 = b * (1 - a), w3 = a * b;
  const f32x2 W0 = pk2(w0, w0), W1 = pk2(w1, w1), W2 = pk2(w2, w2), W3 = pk2(w3, w3), one = pk2(onef, onef);
  f32x2 v[8];
#pragma unroll
  for (int r = 0; r < 8; ++r)
    v[r] = add2(add2(add2(mul2(W3, Wn.t1[r + 1]), mul2(W2, Wn.t0[r + 1]), one), mul2(W1, Wn.t1[r]), one), mul2(W0, Wn.t0[r]), one);
  if (patnorm) {
    const float nm = -(tree_sum_p(v, one) * (1.0f / 64.0f));   // v - mean == v + (-mean) bit for bit
    const f32x2 NM = pk2(nm, nm);
#pragma unroll
    for (int r = 0; r < 8; ++r) v[r] = add2(v[r], NM, one);
  }
  f32x2 px[8], py[8];
#pragma unroll
  for (int r = 0; r < 8; ++r) { px[r] = mul2(T.x(r), v[r]); py[r] = mul2(T.y(r), v[r]); }
  return make_float2(tree_sum_p(px, one), tree_sum_p(py, one));
}

// The reset test of src/patch.cpp:185-194 (plus NaN => reset).
// sqrtf is correctly rounded and monotone, so  sqrtf(s) > outlier  <=>  s > out2max  where out2max
// is the largest float whose root does not exceed outlier (found on the host, dis_api.cu).
__device__ __forceinline__ bool must_reset(float px, float py, float stx, float sty, const SearchArgs& A) {
  const float ddx = stx - px, ddy = sty - py;
  const float s = ddx * ddx + ddy * ddy;
  return !(s <= A.out2max) || px < A.lb || py < A.lb || px > A.ub_w || py > A.ub_h || px != px || py != py;
}

// (checked builds) is the ps x ps gather around plane offset c, or the (ps+1) x (ps+1) tap window ending at it, inside the padded plane?
template <int PS>
__device__ __forceinline__ bool gather_in_plane(const SearchArgs& A, int c) {
  const int th = A.height + 2 * A.pad, y = c / A.tmp_w, x = c - y * A.tmp_w;
  return c >= 0 && x - PS / 2 - 1 >= 0 && x + PS / 2 - 1 < A.tmp_w && y - PS / 2 - 1 >= 0 && y + PS / 2 - 1 < th;
}

// Insert {position bits -> delta_u} into a patch's 16-slot cache image (slot s at cache[s * STRIDE]).
// The entry goes to the way predicted by bit 1 of x and y; an entry already there moves to the set's
// FIFO victim way.  DEDUP: skip the insertion when the position is already present in its set.
// Returns true when (DEDUP) the position was already there.
template <int STRIDE, bool DEDUP>
__device__ __forceinline__ bool cache_insert(uint4* cache, unsigned& fifo, const uint4 e) {
  const int p = slot_of(e.x, e.y);
  if (kWays == 1) {                                             // direct-mapped: the newest position owns the slot
    if (DEDUP) {
      const uint2 old = *reinterpret_cast<const uint2*>(cache + p * STRIDE);
      if (old.x == e.x && old.y == e.y) return true;
    }
    cache[p * STRIDE] = e;
    return false;
  }
  const uint4 old = cache[p * STRIDE];
  if (DEDUP) {
    if (old.x == e.x && old.y == e.y) return true;
    bool found = false;
#pragma unroll
    for (int w = 1; w < kWays; ++w) {
      const uint4 o = cache[(p ^ way_xor(w)) * STRIDE];
      found = found || (o.x == e.x && o.y == e.y);
    }
    if (found) return true;
  }
  if (old.x != 0xffffffffu) {
    const int set = ((p >> 1) & 2) | (p & 1);
    const int way = (fifo >> (kWayBits * set)) & (kWays - 1);
    fifo = (fifo & ~((unsigned)(kWays - 1) << (kWayBits * set))) | (((way + 1) & (kWays - 1)) << (kWayBits * set));
    const int q = (p & 5) | way_xor(way);
    if (q != p) cache[q * STRIDE] = old;                      // the displaced entry moves there
  }
  cache[p * STRIDE] = e;
  return false;
}

enum : int { ST_RUN = 0, ST_WAIT = 1, ST_FREE = 3 };

#ifndef DIS_WARPS
#define DIS_WARPS (DIS_SLOTS == 32 ? 4 : 8)
#endif
constexpr int kWarps = DIS_WARPS;    // warps per block (independent of one another)
#ifndef DIS_WALK_SG8
#define DIS_WALK_SG8 1
#endif
// Resident blocks per SM of the walk kernel.  ps 8 serves its (rare) misses with 8-lane sub-groups: 8 elements
// per lane fit 85 registers, so three blocks (24 warps) hide the LDS -> compare -> add chain of the fast loop;
// the other patch sizes need 128 registers for their service rounds.
#ifdef DIS_WALK_BLOCKS
constexpr int walk_blocks(int, int) { return DIS_WALK_BLOCKS; }
#else
constexpr int walk_blocks(int ps, int sg) { return (kSlots == 32 || (ps == 8 && sg == 8)) ? 3 : 2; }
#endif
constexpr int kStreakLimit = 64;     // consecutive first-seen positions before a patch is handed off
constexpr int kRebindMin = DIS_REBIND_MIN;       // free lanes a warp gathers before it fetches new patches

// per-warp shared memory
struct WarpSmem {
  uint4 cache[kSlots][32];           // per lane: kSlots x {bx, by, dx, dy}
  float4 res[32];                    // service results, per lane
  int list[32];                      // lanes to serve, in rank order
};

// Geometry of work item `w` of the launch.  A launch covers the grid rows [band_y0, band_y0 + band_h) of every
// pair (the whole grid unless a large frame is row-tiled over several GPUs): w = (pair * npw + x) * band_h + y'.
// pid = x*nph + y is the reference's patch id (src/patch_grid.cpp:39-49); sidx = pair * np + pid indexes u / iters.
struct PatchGeom { int pair, pid, rx, ry, sidx; };
__device__ __forceinline__ PatchGeom patch_geom(const SearchArgs& A, int w) {
  PatchGeom g;
  const int npb = A.npw * A.band_h;
  g.pair = w / npb;
  const int r = w - g.pair * npb;
  const int gx = r / A.band_h, gy = A.band_y0 + (r - gx * A.band_h);
  g.pid = gx * A.nph + gy;
  g.sidx = g.pair * A.np + g.pid;
  g.rx = gx * A.steps + A.offw;
  g.ry = gy * A.steps + A.offh;
  return g;
}

// Initial displacement of a patch (PatchGrid::patch_init_from_prev_flow, src/patch_grid.cpp:108-119;
// or the explicit per-patch init of the PatchGrid-level API).
__device__ __forceinline__ float2 patch_init(const SearchArgs& A, int idx, int pair, float refx, float refy) {
  if (A.init_override) return reinterpret_cast<const float2*>(A.init_override)[idx];
  if (A.flow_prev) {
    const int fx = (int)floorf(refx / 2), fy = (int)floorf(refy / 2);
    const float2 v = reinterpret_cast<const float2*>(A.flow_prev + (size_t)pair * A.flow_prev_stride)[fy * (A.width / 2) + fx];
    return make_float2(v.x * 2, v.y * 2);
  }
  return make_float2(0.f, 0.f);
}

// ---- phase 1: convergence ------------------------------------------------------------------------
// The first kN1 passes of every patch visit (almost only) new positions, so they are run without any
// memoisation by one sub-group of SG lanes per patch, G patches per warp, in lock step: template
// gradients, LU factors and (ps 8) the tap window stay in registers, a pass is pure fp32 arithmetic.
// Patches that end here (start out of bounds, reset, max_iter < kN1, conv_eps) are written out;
// the others leave {u, k, LU} and their last kSeed {position -> delta_u} pairs for the walk phase.
#ifndef DIS_CONV_TSMEM
#define DIS_CONV_TSMEM 0     // ps 8: 1 = packed template gradients in shared memory (80 registers, 24 warps per SM): measured SLOWER (9.2 vs 7.5 ms per step), kept as an experiment
#endif
#ifndef DIS_CONV_WARPS
#define DIS_CONV_WARPS 8
#endif
#ifndef DIS_CONV_BLOCKS
#define DIS_CONV_BLOCKS (DIS_CONV_TSMEM && DIS_PACKED ? 3 : 2)
#endif
// ps <= 8: 8 warps x 2 blocks at <= 128 registers; larger patches keep more elements per lane and get the full register file
__host__ __device__ constexpr int conv_warps(int ps) { return ps <= 8 ? DIS_CONV_WARPS : 4; }
// ps 10 / 12 hold their tap window in registers (TapWindow): ~160 registers, three blocks of four warps per SM
#ifndef DIS_BIG_BLOCKS
#define DIS_BIG_BLOCKS 3
#endif
__host__ __device__ constexpr int conv_blocks(int ps) { return ps <= 8 ? DIS_CONV_BLOCKS : (ps <= 12 ? DIS_BIG_BLOCKS : 1); }
template <int PS, int SG, bool FAST>
__global__ void __launch_bounds__(conv_warps(PS) * 32, conv_blocks(PS)) k_converge(const SearchArgs A) {
  using G = Geom<PS, SG>;
  constexpr int kConvWarps = conv_warps(PS);
  const int lane = threadIdx.x & 31, sub = lane / SG, sl = lane % SG;
  const int gwarp = blockIdx.x * kConvWarps + (threadIdx.x >> 5);
  const int nwarps = gridDim.x * kConvWarps;
  const int tw = A.tmp_w;
  const int total = A.n_pairs * A.npw * A.band_h;
  const int wmax = tw - PS / 2, hmax = A.height + 2 * A.pad - PS / 2;
  const int iters = A.iterations;
  int eoff[G::NPL];
  element_offsets<PS, SG>(sl, tw, eoff);
  WalkState* __restrict__ states = reinterpret_cast<WalkState*>(A.walk_state);
  uint4* __restrict__ seeds = reinterpret_cast<uint4*>(A.walk_seeds);
  __shared__ uint4 images[kConvWarps][G::G * kSlots];         // the walk phase's cache image of each patch
  constexpr bool kPacked = DIS_PACKED && PS == 8 && SG == 4;
  constexpr bool kTsmem = kPacked && DIS_CONV_TSMEM;
  __shared__ f32x2 tsm[kTsmem ? kConvWarps : 1][16][32];      // packed template gradients, [element][lane]
  uint4* img_warp = images[threadIdx.x >> 5];
  uint4* img = img_warp + sub * kSlots;
  unsigned long long c_iters = 0;
  unsigned c_sampled = 0, c_resets = 0, c_oob = 0;
#pragma unroll 1
  for (int base = gwarp * G::G; base < total; base += nwarps * G::G) {
    const bool have = base + sub < total;
    const int idx = have ? base + sub : total - 1;
    const PatchGeom g = patch_geom(A, idx);
    const float refx = (float)g.rx, refy = (float)g.ry;
    const float2 in = patch_init(A, g.sidx, g.pair, refx, refy);
    float ux = in.x, uy = in.y;                               // Patch::inverse_search_start, src/patch.cpp:119-154
    float px = refx + ux, py = refy + uy;
    const float stx = px, sty = py;
    const bool oob = px < A.lb || py < A.lb || px > A.ub_w || py > A.ub_h || px != px || py != py;
    bool active = have && !oob;
    bool was_reset = false;
    const int2 fr = A.pairs[g.pair];
    const float2* __restrict__ gxyA =
        reinterpret_cast<const float2*>(A.planes.base + (size_t)fr.x * A.planes.frame_stride + A.planes.plane_stride);
    const float* __restrict__ imB = A.planes.base + (size_t)fr.y * A.planes.frame_stride;
    float2 T[G::NPL];
    DIS_CHECK(A, idx >= 0 && idx < total && (long long)idx < A.walk_capacity, 101);
    DIS_CHECK(A, g.pair >= 0 && g.pair < A.n_pairs && g.pid >= 0 && g.pid < A.np, 102);
    DIS_CHECK(A, gather_in_plane<PS>(A, (g.ry + A.pad) * tw + (g.rx + A.pad) + tw + 1), 103);   // template gather: rows/cols -ps/2 .. ps/2-1
    load_template<PS, SG>(gxyA, (g.ry + A.pad) * tw + (g.rx + A.pad), tw, sl, eoff, T);
    LU2 lu;
    {
      const float3 h = hessian_sums<PS, SG>(T);               // compute_hessian_matrix, src/patch.cpp:75-91
      float h00 = h.x, h11 = h.z;
      fix_singular(h00, h.y, h11);
      lu.factor(h00, h.y, h11);
    }
    typename std::conditional<kPacked, Window8P, Window8<SG>>::type Wn;
    Wn.bx = -1; Wn.by = -1;
    TapWindow<PS, kUseTapWindow<PS, SG> ? SG : 32> Wt;       // (SG 32: a one-element dummy for the sizes that do not use it)
    Wt.bx = -1; Wt.by = -1;
    Template8P Tp;
    if constexpr (kPacked && !kTsmem) Tp.set(T);
    if constexpr (kTsmem) {
      __syncwarp();
      f32x2* t = &tsm[threadIdx.x >> 5][0][lane];
#pragma unroll
      for (int r = 0; r < 8; ++r) { t[r * 32] = pk2(T[r].x, T[8 + r].x); t[(8 + r) * 32] = pk2(T[r].y, T[8 + r].y); }
      __syncwarp();
    }
    int k = 0;
    unsigned ns = 0, fifo = 0;
    float mv0 = 0.f, mv1 = 0.f, mv2 = 0.f;                    // |delta u| (largest component) of the last three EXECUTED passes
    __syncwarp();
#pragma unroll
    for (int i = lane; i < G::G * kSlots; i += 32) img_warp[i] = make_uint4(0xffffffffu, 0xffffffffu, 0u, 0u);
    __syncwarp();
#pragma unroll 1
    for (int pass = 0;; ++pass) {
      if (active) {
        if (must_reset(px, py, stx, sty, A)) { ux = in.x; uy = in.y; was_reset = true; active = false; }   // src/patch.cpp:185-194
        else if (k > iters) active = false;                                                              // src/patch.cpp:199-201
      }
      if (pass == kN1 || !__any_sync(kFull, active)) break;
      // (nearly) every patch of the batch has come back to a position it sampled before: the rest is cheaper in the walk
      if (pass >= A.conv_min && __popc(__ballot_sync(kFull, sl == 0 && active && (fifo >> 31) == 0u)) <= A.conv_slack) break;
      float2 sg;
      if constexpr (PS == 8) {
        const int2 tb = tap_base<PS>(px, py, A.pad, wmax, hmax);
        DIS_CHECK(A, gather_in_plane<PS>(A, tb.y * tw + tb.x), 104);   // tap window: rows/cols base-ps/2-1 .. base+ps/2-1
        if (__any_sync(kFull, active && (tb.x != Wn.bx || tb.y != Wn.by))) Wn.load(imB, tw, sl, tb.x, tb.y);
        if constexpr (kTsmem) sg = sample_window8p(Wn, px, py, A.patnorm, Template8S{&tsm[threadIdx.x >> 5][0][lane]}, A.one);
        else if constexpr (kPacked) sg = sample_window8p(Wn, px, py, A.patnorm, Template8R{&Tp}, A.one);
        else sg = sample_window8<SG>(Wn, px, py, A.patnorm, T);
      } else if constexpr (kUseTapWindow<PS, SG>) {
        const int2 tb = tap_base<PS>(px, py, A.pad, wmax, hmax);
        DIS_CHECK(A, gather_in_plane<PS>(A, tb.y * tw + tb.x), 104);
        if (__any_sync(kFull, active && (tb.x != Wt.bx || tb.y != Wt.by))) Wt.load(imB, tw, eoff, tb.x, tb.y);
        sg = sample_tapwindow<PS, SG>(Wt, px, py, A.patnorm, T);
      } else {
        sg = sample_and_project<PS, SG>(imB, tw, A.pad, wmax, hmax, px, py, sl, A.patnorm, eoff, T);
      }
      if (active) {
        const float2 d = lu.solve(sg.x, sg.y);                // src/patch.cpp:176
        if (sl == 0 && pass >= kN1 - kSeed)
          if (cache_insert<1, true>(img, fifo, make_uint4(__float_as_uint(px), __float_as_uint(py), __float_as_uint(d.x), __float_as_uint(d.y))))
            fifo |= 0x80000000u;                                // "settled" rides in the unused top bit
#if DIS_MOVED_ROLLING
        mv2 = mv1; mv1 = mv0; mv0 = fmaxf(fabsf(d.x), fabsf(d.y));   // a batch may leave at any pass from conv_min on
#else
        if (pass >= kN1 - 3) mv0 = fmaxf(mv0, fmaxf(fabsf(d.x), fabsf(d.y)));
#endif
        ++k; ++ns;
        ux = ux - d.x; uy = uy - d.y;                         // src/patch.cpp:179
        px = refx + ux; py = refy + uy;                       // src/patch.cpp:182
        if (FAST) { if (d.x * d.x + d.y * d.y < A.conv_eps2) active = false; }
      }
    }
    // A patch that still moves by a sizeable step after kN1 passes almost never settles (its walk would miss on
    // most passes): it goes to k_stragglers right away instead of after handoff_misses expensive walk misses.
    bool handed = false;
    const float moved = fmaxf(mv0, fmaxf(mv1, mv2));
    if (have && active && sl == 0 && A.stragglers && A.early_step > 0.f && moved >= A.early_step) {
      const int slot = atomicAdd(A.straggler_count, 1);
      if (slot < A.straggler_cap) {
        Straggler* rec = reinterpret_cast<Straggler*>(A.stragglers) + slot;
        *reinterpret_cast<float4*>(rec) = make_float4(__int_as_float(g.pair), __int_as_float(g.pid), ux, uy);
        __threadfence();
        *reinterpret_cast<volatile int*>(&rec->k) = k;        // publish
        handed = true;
      } else atomicSub(A.straggler_count, 1);
    }
    __syncwarp();
    if (__any_sync(kFull, active && !handed)) {                          // cache images of the batch's patches, coalesced
#pragma unroll
      for (int i = lane; i < G::G * kSlots; i += 32)
        if (base + i / kSlots < total) seeds[(size_t)base * kSlots + i] = img_warp[i];
    }
    if (have && sl == 0) {
      WalkState ws;
      ws.ux = ux; ws.uy = uy; ws.k = active && !handed ? k : -1; ws.swap = (lu.swap ? 1 : 0) | (int)((fifo & 0xffffu) << 8);
      ws.p = lu.p; ws.q = lu.q; ws.l = lu.l; ws.u11 = lu.u11;
      reinterpret_cast<float4*>(states)[2 * (size_t)idx] = make_float4(ws.ux, ws.uy, __int_as_float(ws.k), __int_as_float(ws.swap));
      reinterpret_cast<float4*>(states)[2 * (size_t)idx + 1] = make_float4(ws.p, ws.q, ws.l, ws.u11);
      c_sampled += ns;
      if (!active) {                                          // finished here: Patch::get_u_vector (handed: k_stragglers reports)
        reinterpret_cast<float2*>(A.u_out)[g.sidx] = make_float2(ux, uy);
        if (A.iters_out) A.iters_out[g.sidx] = k;
        c_iters += (unsigned long long)k;
        c_resets += was_reset ? 1u : 0u;
        c_oob += oob ? 1u : 0u;
      }
    }
  }
  if (A.counters) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      c_iters += __shfl_xor_sync(kFull, c_iters, o);
      c_sampled += __shfl_xor_sync(kFull, c_sampled, o);
      c_resets += __shfl_xor_sync(kFull, c_resets, o);
      c_oob += __shfl_xor_sync(kFull, c_oob, o);
    }
    if (lane == 0 && (c_iters | c_sampled | c_resets | c_oob)) {
      atomicAdd(&A.counters[0], c_iters);
      atomicAdd(&A.counters[1], (unsigned long long)c_sampled);
      atomicAdd(&A.counters[16], (unsigned long long)c_sampled);   // sampled passes of k_converge
      atomicAdd(&A.counters[2], (unsigned long long)c_resets);
      atomicAdd(&A.counters[3], (unsigned long long)c_oob);
    }
  }
}

// ---- phase 2: the walk ---------------------------------------------------------------------------
template <int PS, int SG, bool FAST>
__global__ void __launch_bounds__(kWarps * 32, walk_blocks(PS, SG))
k_inverse_search(const SearchArgs A) {
  using G = Geom<PS, SG>;
  extern __shared__ uint4 smem_raw[];
  const int lane = threadIdx.x & 31;
  WarpSmem& W = reinterpret_cast<WarpSmem*>(smem_raw)[threadIdx.x >> 5];
  uint4* cache = &W.cache[0][lane];                           // slot s at cache[s * 32]
  if (threadIdx.x == 0 && A.stragglers) atomicAdd(A.straggler_count + 3, 1);   // "this block of the walk is resident"
  const int tw = A.tmp_w;
  const int total = A.n_pairs * A.npw * A.band_h;
  const int wmax = tw - PS / 2, hmax = A.height + 2 * A.pad - PS / 2;
  const int sub = lane / SG, sl = lane % SG;
  int eoff[G::NPL];
  element_offsets<PS, SG>(sl, tw, eoff);
  const int iters = A.iterations;

  // per-lane patch state (a lane is re-bound to a new patch whenever it finishes one)
  int state = ST_FREE, idx = -1, k = 0, pair = 0, coff = 0;
  float refx = 0.f, refy = 0.f, inx = 0.f, iny = 0.f, stx = 0.f, sty = 0.f;
  float ux = 0.f, uy = 0.f, px = 0.f, py = 0.f;
  unsigned fifo = 0;
  int streak = 0, k_posted = -2, waited = 0, misses = 0;
  bool was_reset = false;
  LU2 lu;
  lu.p = lu.q = lu.l = lu.u11 = 0.f; lu.swap = false;
  unsigned long long c_iters = 0;
  unsigned c_sampled = 0, c_resets = 0, c_oob = 0;
  bool exhausted = false;

#pragma unroll 1
  for (;;) {
    // ---- free lanes fetch new patches (a few lanes at a time) --------------------------------------
    const unsigned free_m = __ballot_sync(kFull, state == ST_FREE);
    if (free_m != 0 && !exhausted &&
        (__popc(free_m) >= kRebindMin || __ballot_sync(kFull, state == ST_RUN) == 0)) {
      const int leader = __ffs(free_m) - 1;
      int base = 0;
      if (lane == leader) base = atomicAdd(A.work_counter, __popc(free_m));
      base = __shfl_sync(kFull, base, leader);
      if (base + __popc(free_m) >= total) exhausted = true;
#ifdef DIS_DEBUG_STATS
      if (lane == leader) { atomicAdd(&A.counters[10], (unsigned long long)__popc(free_m)); atomicAdd(&A.counters[11], 1ull); }
#endif
      if (state == ST_FREE) {
        idx = base + __popc(free_m & ((1u << lane) - 1u));
        DIS_CHECK(A, idx >= 0 && (idx >= total || (long long)idx < A.walk_capacity), 201);
        if (idx < total) {
          const float4 w0 = reinterpret_cast<const float4*>(A.walk_state)[2 * (size_t)idx];
          k = __float_as_int(w0.z);
          if (k >= 0) {                                        // else: finished in k_converge
            const float4 w1 = reinterpret_cast<const float4*>(A.walk_state)[2 * (size_t)idx + 1];
            const PatchGeom g = patch_geom(A, idx);
            pair = g.pair;
            coff = (g.ry + A.pad) * tw + (g.rx + A.pad);
            refx = (float)g.rx; refy = (float)g.ry;
            const float2 in = patch_init(A, g.sidx, g.pair, refx, refy);
            inx = in.x; iny = in.y;
            stx = refx + inx; sty = refy + iny;                // Patch::inverse_search_start, src/patch.cpp:119-154
            ux = w0.x; uy = w0.y;
            px = refx + ux; py = refy + uy;
            lu.p = w1.x; lu.q = w1.y; lu.l = w1.z; lu.u11 = w1.w;
            lu.swap = (__float_as_int(w0.w) & 1) != 0;
            fifo = (unsigned)__float_as_int(w0.w) >> 8;
            streak = 0; k_posted = -2; was_reset = false; waited = 0; misses = 0;
            const uint4* __restrict__ seed = reinterpret_cast<const uint4*>(A.walk_seeds) + (size_t)idx * kSlots;
#pragma unroll
            for (int s = 0; s < kSlots; ++s) cache[s * 32] = __ldg(seed + s);   // the cache image k_converge built
            state = ST_RUN;
          }
        }
      }
    }

    // ---- fast phase: kChunk passes probing only the predicted way; a lane leaves it at its first
    //      miss (or when it is done) and idles until the chunk ends -----------------------------------
    bool run = state == ST_RUN;
    bool fin = false;                                        // the lane's patch is finished in this round
#ifdef DIS_DEBUG_STATS
    const int dbg_k0 = k;
    if (lane == 0) { atomicAdd(&A.counters[4], 1ull); }
    { const unsigned m = __ballot_sync(kFull, run), f = __ballot_sync(kFull, state == ST_FREE);
      if (lane == 0) { atomicAdd(&A.counters[5], (unsigned long long)__popc(m)); atomicAdd(&A.counters[13], (unsigned long long)__popc(f)); } }
#endif
    // A round is up to walk_chunks times {kChunk passes probing only the predicted way; lanes that left the chunk
    // probe all four ways of the set and, on a hit, step one pass and rejoin}; it ends early once fewer than
    // walk_minrun lanes are running.  Lanes with a first-seen position then wait for the service below.
    // (A lane closer than kChunk passes to its budget steps through the four-way probe, one pass per chunk,
    // which keeps the budget test out of the inner loop.)
    bool missed = false;
    int chunk = 0;
#pragma unroll 1
    do {
      const bool budget = k + kChunk <= iters + 1;            // enough passes left for a whole chunk
      run = run && budget;
#pragma unroll
      for (int it = 0; it < kChunk; ++it) {
        if (DIS_EARLY && it == DIS_EARLY && !__any_sync(kFull, run)) break;
        const unsigned bx = __float_as_uint(px), by = __float_as_uint(py);
        const uint4 e = *slot_ptr32(cache, bx, by);
        run = run && e.x == bx && e.y == by;
        if (run) {
          ++k;
          const float dx = __uint_as_float(e.z), dy = __uint_as_float(e.w);
          ux = ux - dx; uy = uy - dy;                         // src/patch.cpp:179
          px = refx + ux; py = refy + uy;                     // src/patch.cpp:182
          if (FAST) { if (dx * dx + dy * dy < A.conv_eps2) { run = false; fin = true; } }
        }
      }
      bool probe = state == ST_RUN && !run && !fin && !missed;
      if (kWays == 1) {
        // direct-mapped: a lane that fell out of a chunk it had entered met a first-seen position; only a lane near the
        // end of its budget still has to step, one pass per chunk, through the slot test below
        if (probe && budget) { missed = true; probe = false; }
      }
      if (__any_sync(kFull, probe)) {
        if (probe) {
          const unsigned bx = __float_as_uint(px), by = __float_as_uint(py);
          const int p = slot_of(bx, by);
          uint4 e = cache[p * 32];
          bool hit = e.x == bx && e.y == by;
#pragma unroll
          for (int w = 1; w < kWays; ++w) {
            const uint4 o = cache[(p ^ way_xor(w)) * 32];
            if (o.x == bx && o.y == by) { e = o; hit = true; }
          }
          if (hit) {
#ifdef DIS_DEBUG_STATS
            atomicAdd(&A.counters[9], 1ull);
#endif
            if (k > iters) fin = true;                        // src/patch.cpp:199-201 (position already vetted)
            else {
              ++k;
              const float dx = __uint_as_float(e.z), dy = __uint_as_float(e.w);
              ux = ux - dx; uy = uy - dy;
              px = refx + ux; py = refy + uy;
              run = true;
              if (FAST) { if (dx * dx + dy * dy < A.conv_eps2) { run = false; fin = true; } }
            }
          } else missed = true;
        }
      }
    } while (++chunk < A.walk_chunks && __popc(__ballot_sync(kFull, run)) >= A.walk_minrun);
#ifdef DIS_DEBUG_STATS
    { int adv = k - dbg_k0;
      for (int o = 16; o > 0; o >>= 1) adv += __shfl_xor_sync(kFull, adv, o);
      if (lane == 0) atomicAdd(&A.counters[6], (unsigned long long)adv); }
#endif
    if (missed) {
      // a first-seen position: it owes its reset test and a sample
      streak = (k == k_posted + 1) ? streak + 1 : 1;
      k_posted = k;
      state = ST_WAIT;
      if (((streak >= kStreakLimit && k >= kHandoffMinIter) || misses >= A.handoff_misses) && A.stragglers) {
        // the orbit does not settle: hand the patch to k_stragglers
        if (must_reset(px, py, stx, sty, A)) { ux = inx; uy = iny; was_reset = true; fin = true; }
        else if (k > iters) fin = true;
        else {
          const int slot = atomicAdd(A.straggler_count, 1);
          DIS_CHECK(A, slot >= 0, 205);
          if (slot < A.straggler_cap) {
            Straggler* rec = reinterpret_cast<Straggler*>(A.stragglers) + slot;
            *reinterpret_cast<float4*>(rec) = make_float4(__int_as_float(pair), __int_as_float(patch_geom(A, idx).pid), ux, uy);
            __threadfence();
            *reinterpret_cast<volatile int*>(&rec->k) = k;     // publish
            state = ST_FREE;     // results and counters come from k_stragglers
          } else { atomicSub(A.straggler_count, 1); streak = 0; }
        }
      }
    }

    // ---- service: waiting patches (first-seen positions, new templates), one per sub-group and round.
    //      Whole rounds; an under-filled last round only if it is at least half full, nothing else can
    //      run, or one of its patches has already waited a chunk. -------------------------------------
    const bool needs = state == ST_WAIT && !fin;
    const unsigned wait_mask = __ballot_sync(kFull, needs);
    bool served = false;
    if (wait_mask != 0) {
      const int nwait = __popc(wait_mask);
      const int rem = nwait % G::G;
      const bool others = __ballot_sync(kFull, state == ST_RUN && !fin) != 0;
      const bool aged = __any_sync(kFull, needs && (waited >= A.walk_age || misses >= DIS_HEAVY));   // frequent missers are not made to wait
      const int nserv = (!others || 2 * rem >= G::G || aged) ? nwait : nwait - rem;
      const int rank = __popc(wait_mask & ((1u << lane) - 1u));
      served = needs && rank < nserv;
      if (served) W.list[rank] = lane;
#ifdef DIS_DEBUG_STATS
      if (lane == 0) { atomicAdd(&A.counters[7], (unsigned long long)((nserv + G::G - 1) / G::G)); atomicAdd(&A.counters[8], (unsigned long long)nserv);
                       atomicAdd(&A.counters[12], (unsigned long long)nwait); }
#endif
      __syncwarp();
#pragma unroll 1
      for (int base = 0; base < nserv; base += G::G) {
        const int src = W.list[min(base + sub, nserv - 1)];
        DIS_CHECK(A, src >= 0 && src < 32, 202);
        const float opx = __shfl_sync(kFull, px, src), opy = __shfl_sync(kFull, py, src);
        const int oc = __shfl_sync(kFull, coff, src), opair = __shfl_sync(kFull, pair, src);
        DIS_CHECK(A, opair >= 0 && opair < A.n_pairs && gather_in_plane<PS>(A, oc + tw + 1), 203);
        { const int2 tbc = tap_base<PS>(opx, opy, A.pad, wmax, hmax); DIS_CHECK(A, gather_in_plane<PS>(A, tbc.y * tw + tbc.x), 204); }
        const int2 fr = A.pairs[opair];
        const float2* __restrict__ gxyA =
            reinterpret_cast<const float2*>(A.planes.base + (size_t)fr.x * A.planes.frame_stride + A.planes.plane_stride);
        const float* __restrict__ imB = A.planes.base + (size_t)fr.y * A.planes.frame_stride;
        float2 T[G::NPL];
        load_template<PS, SG>(gxyA, oc, tw, sl, eoff, T);
        const float2 sg = sample_and_project<PS, SG>(imB, tw, A.pad, wmax, hmax, opx, opy, sl, A.patnorm, eoff, T);
        const float4 out = make_float4(sg.x, sg.y, 0.f, 0.f);
        if (sl == 0) W.res[src] = out;
      }
      __syncwarp();
    }
    if (served) {
      const float4 g = W.res[lane];
      waited = 0;
      if (must_reset(px, py, stx, sty, A)) {                 // src/patch.cpp:185-194
        ux = inx; uy = iny; was_reset = true; fin = true;
      } else if (k > iters) {                                // src/patch.cpp:199-201
        fin = true;
      } else {
        const float2 d = lu.solve(g.x, g.y);                 // src/patch.cpp:176
        cache_insert<32, false>(cache, fifo, make_uint4(__float_as_uint(px), __float_as_uint(py), __float_as_uint(d.x), __float_as_uint(d.y)));
        ++c_sampled; ++misses;
        ++k;
        ux = ux - d.x; uy = uy - d.y;
        px = refx + ux; py = refy + uy;
        state = ST_RUN;
        if (FAST) { if (d.x * d.x + d.y * d.y < A.conv_eps2) fin = true; }
      }
    } else if (needs) ++waited;
    if (fin) {                                               // Patch::get_u_vector of a finished patch
      const int sidx = patch_geom(A, idx).sidx;
      reinterpret_cast<float2*>(A.u_out)[sidx] = make_float2(ux, uy);
      if (A.iters_out) A.iters_out[sidx] = k;
      c_iters += (unsigned long long)k;
      c_resets += was_reset ? 1u : 0u;
      state = ST_FREE;
    }
    if (exhausted && __ballot_sync(kFull, state != ST_FREE) == 0) break;
  }
  if (A.counters) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      c_iters += __shfl_xor_sync(kFull, c_iters, o);
      c_sampled += __shfl_xor_sync(kFull, c_sampled, o);
      c_resets += __shfl_xor_sync(kFull, c_resets, o);
      c_oob += __shfl_xor_sync(kFull, c_oob, o);
    }
    if (lane == 0 && (c_iters | c_sampled | c_resets | c_oob)) {
      atomicAdd(&A.counters[0], c_iters);
      atomicAdd(&A.counters[1], (unsigned long long)c_sampled);
      atomicAdd(&A.counters[17], (unsigned long long)c_sampled);   // sampled passes of the walk
      atomicAdd(&A.counters[2], (unsigned long long)c_resets);
      atomicAdd(&A.counters[3], (unsigned long long)c_oob);
    }
  }
}

// Patches handed off by k_inverse_search: every pass samples, one patch per sub-group with its
// template gradients in registers.  Same arithmetic, same order of tests.
// Everything a straggler can touch lies within outlierthresh of its start position (beyond it the patch is
// reset, src/patch.cpp:185), so ps 8 stages that neighbourhood of the second image — kTile x kTile taps
// around the start — in shared memory once per patch: the ~950-pass chain then never waits on L2.
constexpr long long kStragWideBelow = 0;   // patches per launch
constexpr int kConvMin = 12;      // passes; >= kN1 = never
constexpr float kEarlyStep = 0.3f;   // px; measured on the bench workload: 0.1 ... 1 equivalent (+1 %), below 0.03 the stragglers take over
constexpr int kTile = 19, kTileStride = 20, kTileMargin = 5;   // 9-tap window + 2 * (ceil(outlierthresh) + 1)
constexpr int kStragWarps = 4;
constexpr long long kYieldGraceCycles = 20000;   // ~10 us: the walk's blocks announce themselves within a few us of their launch
// ps 8 is built twice: unconstrained (170 registers, two blocks per SM: the shortest chain per pass, which is what bounds a
// launch with few stragglers) and DENSE (three blocks per SM: ptxas then finds 136 registers without spilling; +50 %
// resident patches, which is what bounds a launch of many pairs: 31.4 -> 30.4 ms per 49-pair step, but 3.5 -> 3.8 ms
// for a single pair, so launch_ps picks by the number of pairs).
constexpr int kStragDenseBlocks = 3;
constexpr int kStragDensePairs = 16;   // DIS_STRAG_DENSE overrides (pairs per launch from which the dense build is used)
__host__ __device__ constexpr int strag_min_blocks(int ps, bool dense) {
  return (ps == 10 || ps == 12) ? DIS_BIG_BLOCKS : (ps == 8 && dense ? kStragDenseBlocks : 1);
}
template <int PS, int SG, bool FAST, bool DENSE = false>
__global__ void __launch_bounds__(kStragWarps * 32, strag_min_blocks(PS, DENSE)) k_stragglers(const SearchArgs A) {
  using G = Geom<PS, SG>;
  extern __shared__ float tiles[];
  const int th = A.height + 2 * A.pad;
  const bool use_tile = PS == 8 && A.tile_margin <= kTileMargin && A.tmp_w >= kTile && th >= kTile;
  float* tile = tiles + ((threadIdx.x >> 5) * G::G + (threadIdx.x & 31) / SG) * (kTile * kTileStride);
  const int lane = threadIdx.x & 31, sub = lane / SG, sl = lane % SG;
  const Straggler* list = reinterpret_cast<const Straggler*>(A.stragglers);
  int* const ctl = A.straggler_count;                        // [0] appended, [1] claimed, [2] walk finished
  const int tw = A.tmp_w;
  const int wmax = tw - PS / 2, hmax = A.height + 2 * A.pad - PS / 2;
  int eoff[G::NPL];
  element_offsets<PS, SG>(sl, tw, eoff);
  unsigned long long it_total = 0;
  unsigned ns_total = 0, nr_total = 0;
  const long long t_start = clock64();
#pragma unroll 1
  for (;;) {
    // ---- claim the next batch of G entries.  The kernel may run concurrently with the walk (ctl[2] == 0):
    //      entries are then consumed as they are published; a claimed slot that is still empty is waited for
    //      until the walk has finished, after which ctl[0] is final. ------------------------------------------
    int base = -1, nclaim = 0;
    if (lane == 0) {
      for (;;) {
        const int done = ld_volatile(ctl + 2);
        const int cnt = min(ld_volatile(ctl), A.straggler_cap);
        const int claimed = ld_volatile(ctl + 1);
        // no more than what has been appended: a warp does not sit on empty slots while hand-offs trickle in
        if (claimed < cnt) { nclaim = min(G::G, cnt - claimed); base = atomicAdd(ctl + 1, nclaim); break; }
        if (done) break;                                     // nothing left and nothing more to come
        // Never hold a block slot the walk still needs: had this kernel been scheduled ahead of the walk (the order
        // of two streams is not guaranteed) it would wait for hand-offs of a walk that cannot become resident.  An
        // idle warp therefore leaves while the walk is not fully resident; what it leaves behind is picked up by
        // the sweep launch that follows the walk in stream order (launch_ps).
        if (ld_volatile(ctl + 3) < A.walk_grid && clock64() - t_start > kYieldGraceCycles) break;
        __nanosleep(400);
        if (clock64() - t_start > 20000000000ll) {           // safety net (~10 s): never spin forever; the host reports it
          if (A.counters) atomicExch(&A.counters[15], 1ull);
          break;
        }
      }
    }
    base = __shfl_sync(kFull, base, 0);
    nclaim = __shfl_sync(kFull, nclaim, 0);
    if (base < 0) break;
    bool have = false;
    if (sl == 0 && sub < nclaim) {
      for (;;) {
        if (ld_volatile(&list[base + sub < A.straggler_cap ? base + sub : 0].k) >= 0 && base + sub < A.straggler_cap) { have = true; break; }
        if (ld_volatile(ctl + 2) && base + sub >= min(ld_volatile(ctl), A.straggler_cap)) break;   // walk finished: the slot stays empty
        __nanosleep(200);
        if (clock64() - t_start > 20000000000ll) {           // same safety net: the patch would be lost, so the host must know
          if (A.counters) atomicExch(&A.counters[15], 1ull);
          break;
        }
      }
      __threadfence();
    }
    have = __shfl_sync(kFull, have ? 1 : 0, sub * SG) != 0;
    if (!__any_sync(kFull, have)) continue;
    Straggler e;
    {
      const int src = __shfl_sync(kFull, have ? base + sub : -1, __ffs(__ballot_sync(kFull, have)) - 1);   // a valid entry for idle sub-groups
      const Straggler* rec = list + (have ? base + sub : src);
      const float4 f = __ldcg(reinterpret_cast<const float4*>(rec));
      e.pair = __float_as_int(f.x); e.pid = __float_as_int(f.y); e.ux = f.z; e.uy = f.w; e.k = ld_volatile(&rec->k);
    }
    DIS_CHECK(A, e.pair >= 0 && e.pair < A.n_pairs && e.pid >= 0 && e.pid < A.np && e.k >= 0 && e.k <= A.iterations + 1, 301);
    DIS_CHECK(A, base >= 0 && base + nclaim <= A.straggler_cap, 302);
    const int2 fr = A.pairs[e.pair];
    const float2* __restrict__ gxyA =
        reinterpret_cast<const float2*>(A.planes.base + (size_t)fr.x * A.planes.frame_stride + A.planes.plane_stride);
    const float* __restrict__ imB = A.planes.base + (size_t)fr.y * A.planes.frame_stride;
    const int gx_i = e.pid / A.nph, gy_i = e.pid - gx_i * A.nph;
    const int rx = gx_i * A.steps + A.offw, ry = gy_i * A.steps + A.offh;
    const float refx = (float)rx, refy = (float)ry;
    float2 T[G::NPL];
    load_template<PS, SG>(gxyA, (ry + A.pad) * tw + (rx + A.pad), tw, sl, eoff, T);
    const float3 h = hessian_sums<PS, SG>(T);
    float h00 = h.x, h11 = h.z;
    const float h01 = h.y;
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
    constexpr bool kPacked = DIS_PACKED && PS == 8 && SG == 4;
    typename std::conditional<kPacked, Window8P, Window8<SG>>::type Wn;
    Wn.bx = -1; Wn.by = -1;
    Template8P Tp;
    if constexpr (kPacked) Tp.set(T);
    TapWindow<PS, kUseTapWindow<PS, SG> ? SG : 32> Wt;
    Wt.bx = -1; Wt.by = -1;
    int ox = 0, oy = 0;
    if (PS == 8 && use_tile) {
      const int2 tb0 = tap_base<PS>(stx, sty, A.pad, wmax, hmax);
      ox = min(max(tb0.x - 5 - kTileMargin, 0), tw - kTile);
      oy = min(max(tb0.y - 5 - kTileMargin, 0), th - kTile);
      DIS_CHECK(A, ox >= 0 && oy >= 0 && ox + kTile <= tw && oy + kTile <= th, 303);
      __syncwarp();
      for (int i = sl; i < kTile * kTile; i += SG) {
        const int r = i / kTile, c = i - r * kTile;
        tile[r * kTileStride + c] = __ldg(imB + (oy + r) * tw + ox + c);
      }
      __syncwarp();
    }
#pragma unroll 1
    while (__any_sync(kFull, active)) {
      float2 sg;
      if constexpr (PS == 8) {                               // tap window in registers, reloaded when the base moves
        const int2 tb = tap_base<PS>(px, py, A.pad, wmax, hmax);
        if (__any_sync(kFull, active && (tb.x != Wn.bx || tb.y != Wn.by))) {
          if (use_tile) {   // a base outside the tile belongs to a position the reset test rejects: clamp, result unused
            const int rx = min(max(tb.x - 5 - ox, 0), kTile - 9), ry = min(max(tb.y - 5 - oy, 0), kTile - 9);
            // an ACTIVE patch (one the reset test has not rejected) must find its true window inside the tile
            DIS_CHECK(A, !(active && !must_reset(px, py, stx, sty, A)) || (rx == tb.x - 5 - ox && ry == tb.y - 5 - oy), 304);
            Wn.load_tile(tile + ry * kTileStride + rx, kTileStride, sl, tb.x, tb.y);
          } else Wn.load(imB, tw, sl, tb.x, tb.y);
        }
        if constexpr (kPacked) sg = sample_window8p(Wn, px, py, A.patnorm, Template8R{&Tp}, A.one);
        else sg = sample_window8<SG>(Wn, px, py, A.patnorm, T);
      } else if constexpr (kUseTapWindow<PS, SG>) {
        const int2 tb = tap_base<PS>(px, py, A.pad, wmax, hmax);
        if (__any_sync(kFull, active && (tb.x != Wt.bx || tb.y != Wt.by))) Wt.load(imB, tw, eoff, tb.x, tb.y);
        sg = sample_tapwindow<PS, SG>(Wt, px, py, A.patnorm, T);
      } else {
        sg = sample_and_project<PS, SG>(imB, tw, A.pad, wmax, hmax, px, py, sl, A.patnorm, eoff, T);
      }
      if (active) {
        if (must_reset(px, py, stx, sty, A)) { ux = inx; uy = iny; was_reset = 1; active = false; }
        else if (k > A.iterations) active = false;
        else {
          const float2 d = lu.solve(sg.x, sg.y);
          ++k; ++ns;
          ux = ux - d.x; uy = uy - d.y;
          px = refx + ux; py = refy + uy;
          if (FAST) { if (d.x * d.x + d.y * d.y < A.conv_eps2) active = false; }
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

template <int PS, int SG, bool FAST>
int launch_ps(const SearchArgs& a_in, cudaStream_t s, cudaEvent_t* ev, const SearchSide* side) {
  SearchArgs a = a_in;
  a.one = 1.0f;
  if (a.walk_capacity <= 0) a.walk_capacity = 1ll << 60;
  constexpr int WSG = (DIS_WALK_SG8 && PS == 8) ? 8 : SG;   // sub-group size of the walk phase's (rare) service rounds
  const size_t smem = (size_t)kWarps * sizeof(WarpSmem);
  static bool attr_done[64] = {false};                        // per template instantiation and device
  const int dev = current_device();
  const bool set_attr = !attr_done[dev];
  attr_done[dev] = true;
  if (set_attr) {
    cudaFuncSetAttribute(k_inverse_search<PS, WSG, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k_inverse_search<PS, WSG, FAST>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    constexpr int WS = PS == 8 ? 8 : SG;
    const size_t tsm_w = PS == 8 ? sizeof(float) * kStragWarps * Geom<PS, WS>::G * kTile * kTileStride : 0;
    cudaFuncSetAttribute(k_stragglers<PS, WS, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm_w);
    cudaFuncSetAttribute(k_stragglers<PS, WS, FAST>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    const size_t tsm_n = PS == 8 ? sizeof(float) * kStragWarps * Geom<PS, SG>::G * kTile * kTileStride : 0;
    cudaFuncSetAttribute(k_stragglers<PS, SG, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm_n);
    cudaFuncSetAttribute(k_stragglers<PS, SG, FAST>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    if constexpr (PS == 8) {
      cudaFuncSetAttribute(k_stragglers<PS, SG, FAST, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm_n);
      cudaFuncSetAttribute(k_stragglers<PS, SG, FAST, true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    }
  }
  // persistent warps: lanes fetch patches from a global counter until the level is done
  if (a.band_h <= 0) { a.band_y0 = 0; a.band_h = a.nph; }
  const long long total = (long long)a.n_pairs * a.npw * a.band_h;
  const long long want = (total + kWarps * 32 - 1) / (kWarps * 32);
  const int n_sm = sm_count();
  const Tuning& tn = tuning();
  const long long slots = ((long long)n_sm * walk_blocks(PS, WSG) + (a.share > 1 ? a.share : 1) - 1) / (a.share > 1 ? a.share : 1);
  const unsigned blocks = (unsigned)(want < slots ? want : slots);
  a.walk_grid = (int)blocks;
  int n = 2;
  {   // Small launches are bound by the serial miss chains of their few restless patches, not by throughput:
      // hand those to k_stragglers (one pass every ~0.6 us, template and taps on chip) after fewer misses.
    const int c = total < 200000 ? 0 : total < 600000 ? 1 : total < 2000000 ? 2 : 3;   // size class of the launch
    a.handoff_misses = tn.handoff[c];
    a.walk_chunks = tn.walk[3 * c]; a.walk_minrun = tn.walk[3 * c + 1]; a.walk_age = tn.walk[3 * c + 2];
    a.conv_min = tn.conv_min;
    a.conv_slack = tn.conv_slack;
    a.early_step = tn.early_step;
  }
  // k_stragglers: 4 lanes per patch (ps 8) have the fewest instructions per sample; 8 lanes per patch have the
  // shorter dependent chain per pass, which is what bounds a small launch (its last hand-offs run alone).
  const long long wide_below = tn.strag_wide;
  const bool wide = PS == 8 && SG == 4 && (DIS_STRAG_SG8 || total < wide_below);
  auto strag = [&](unsigned nblocks, cudaStream_t st) {
    if (wide) {
      constexpr int WS = PS == 8 ? 8 : SG;
      const size_t tsm = PS == 8 ? sizeof(float) * kStragWarps * Geom<PS, WS>::G * kTile * kTileStride : 0;
      k_stragglers<PS, WS, FAST><<<nblocks, kStragWarps * 32, tsm, st>>>(a);
    } else {
      const size_t tsm = PS == 8 ? sizeof(float) * kStragWarps * Geom<PS, SG>::G * kTile * kTileStride : 0;
      if constexpr (PS == 8) {
        if (a.n_pairs >= tn.strag_dense) { k_stragglers<PS, SG, FAST, true><<<nblocks, kStragWarps * 32, tsm, st>>>(a); return; }
      }
      k_stragglers<PS, SG, FAST><<<nblocks, kStragWarps * 32, tsm, st>>>(a);
    }
  };
  const bool overlap = a.stragglers && side && side->stream;   // k_stragglers runs beside the walk and eats its hand-offs as they come
  cudaMemsetAsync(a.work_counter, 0, sizeof(int), s);
  if (a.stragglers) {
    const long long cap = total / 4 + 64 < a.straggler_cap ? total / 4 + 64 : a.straggler_cap;
    a.straggler_cap = (int)cap;
    cudaMemsetAsync(a.straggler_count, 0, 4 * sizeof(int), s);
    cudaMemsetAsync(a.stragglers, 0xff, (size_t)cap * kStragglerBytes, s);
  }
  if (ev) cudaEventRecord(ev[0], s);
  {   // phase 1: G patches per warp, grid-stride over a few waves of resident blocks
    constexpr int kConvWarps = conv_warps(PS);
    const long long cwant = (total + kConvWarps * Geom<PS, SG>::G - 1) / (kConvWarps * Geom<PS, SG>::G);
    const long long cslots = (long long)n_sm * (conv_blocks(PS) > 1 ? conv_blocks(PS) * 4 : 12) / (a.share > 1 ? a.share : 1);
    k_converge<PS, SG, FAST><<<(unsigned)(cwant < cslots ? cwant : cslots), kConvWarps * 32, 0, s>>>(a);
  }
  if (ev) cudaEventRecord(ev[1], s);
  if (overlap) cudaEventRecord(side->fork, s);               // "convergence phase done"
  k_inverse_search<PS, WSG, FAST><<<blocks, kWarps * 32, smem, s>>>(a);
  if (a.stragglers) cudaMemsetAsync(a.straggler_count + 2, 1, sizeof(int), s);   // "walk finished": the entry count is final
  if (overlap) {
    // k_stragglers is launched behind the walk on the side stream: its blocks become resident as the walk's
    // persistent blocks retire, so the hand-offs are worked on during the walk's tail instead of after it.
    // (Enqueued after the flag's memset so that a host that serialises launches — CUDA_LAUNCH_BLOCKING, ncu —
    // still terminates: the kernel then simply starts with the flag already set.)
    cudaStreamWaitEvent(side->stream, side->fork, 0);
    strag(side->blocks, side->stream);
    cudaEventRecord(side->join, side->stream);
  }
  if (ev) cudaEventRecord(ev[2], s);
  if (a.stragglers) {
    if (overlap) {
      // Sweep: in stream order behind the walk and its "finished" flag, so it cannot wait on anything.  It finds
      // nothing to claim unless the side launch yielded (see k_stragglers) and then finishes the list.
      cudaStreamWaitEvent(s, side->join, 0);
      strag((unsigned)n_sm * DIS_STRAG_BLOCKS, s);
      ++n;
    } else strag((unsigned)n_sm * DIS_STRAG_BLOCKS, s);
    ++n;
  }
  if (ev) cudaEventRecord(ev[3], s);
  return n;
}

}  // namespace

namespace {
Tuning parse_tuning() {
    Tuning t;
    const int h[4] = {24, 48, 128, 384};
    for (int i = 0; i < 4; ++i) t.handoff[i] = h[i];
    if (const char* e = getenv("DIS_HANDOFF")) sscanf(e, "%d,%d,%d,%d", t.handoff, t.handoff + 1, t.handoff + 2, t.handoff + 3);
    for (int c = 0; c < 4; ++c) { t.walk[3 * c] = 10; t.walk[3 * c + 1] = 24; t.walk[3 * c + 2] = DIS_AGE; }   // measured: tools/sweep2.sh, DESIGN.md 3.2
    if (const char* e = getenv("DIS_WALK_TUNE"))
      sscanf(e, "%d,%d,%d:%d,%d,%d:%d,%d,%d:%d,%d,%d", t.walk, t.walk + 1, t.walk + 2, t.walk + 3, t.walk + 4, t.walk + 5, t.walk + 6,
             t.walk + 7, t.walk + 8, t.walk + 9, t.walk + 10, t.walk + 11);
    const char* e;
    t.conv_min = (e = getenv("DIS_CONV_MIN")) ? atoi(e) : kConvMin;
    t.conv_slack = (e = getenv("DIS_CONV_SLACK")) ? atoi(e) : 2;
    t.early_step = (e = getenv("DIS_EARLY_STEP")) ? (float)atof(e) : kEarlyStep;
    t.strag_wide = (e = getenv("DIS_STRAG_WIDE")) ? atoll(e) : kStragWideBelow;
    t.strag_dense = (e = getenv("DIS_STRAG_DENSE")) ? atoi(e) : kStragDensePairs;
    t.share = getenv("DIS_SHARE") ? 1 : 0;
    t.groups = (e = getenv("DIS_GROUPS")) ? (atoi(e) > 1 ? atoi(e) : 1) : 1;
    t.side_blocks = (e = getenv("DIS_SIDE_BLOCKS")) ? (atoi(e) > 0 ? atoi(e) : 0) : 4;
    t.out_slices_n = -1; t.out_slices_nw = 0;
    if ((e = getenv("DIS_OUT_SLICES"))) {
      if (strchr(e, ',')) {
        for (const char* q = e; q && *q && t.out_slices_nw < 16; q = strchr(q, ',') ? strchr(q, ',') + 1 : nullptr) {
          const double w = atof(q);
          t.out_slices_w[t.out_slices_nw++] = w > 0 ? w : 0;
        }
      } else t.out_slices_n = atoi(e);
    }
    return t;
}
Tuning g_tuning;
bool g_tuning_ready = false;
}  // namespace

const Tuning& tuning() {
  if (!g_tuning_ready) { g_tuning = parse_tuning(); g_tuning_ready = true; }
  return g_tuning;
}
void reparse_tuning() { g_tuning = parse_tuning(); g_tuning_ready = true; }

int current_device() {
  int dev = 0;
  return (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < 64) ? dev : 0;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

int launch_inverse_search(const SearchArgs& a, int patch_size, cudaStream_t s, cudaEvent_t* ev, const SearchSide* side) {
  const bool fast = a.conv_eps2 > 0.f;
#define DIS_CASE(PS, SG) case PS: return fast ? launch_ps<PS, SG, true>(a, s, ev, side) : launch_ps<PS, SG, false>(a, s, ev, side);
  switch (patch_size) {
    DIS_CASE(4, 4) DIS_CASE(6, 4) DIS_CASE(8, 4) DIS_CASE(10, 8) DIS_CASE(12, 8) DIS_CASE(16, 8)
    default: return -1;
  }
#undef DIS_CASE
}

}  // namespace dis

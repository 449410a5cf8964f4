// dis_kernels.cuh — shared declarations of the sm_100a kernels behind include/dis_b200.h.
//
// Arithmetic contract: fp32, IEEE round-to-nearest, expressions in the
// reference's association, NO FMA contraction.  The whole library is compiled with
// -fmad=false -prec-div=true -prec-sqrt=true -ftz=false; `+ - * / sqrtf floorf ceilf fabsf`
// are then bit-identical to the SSE2 scalar operations of the CPU reference.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dis {

// Planes of one pyramid level for all frames of a run, in HBM:
//   image:     base + frame*frame_stride + y*tmp_w + x                          (fp32)
//   gradients: base + frame*frame_stride + plane_stride + 2*(y*tmp_w + x) + {0: dx, 1: dy}
// i.e. the reference's padded planes (src/main.cpp:41-49): (h+2*pad) x (w+2*pad), replicate border
// for the image, zero border for the gradients — with dx and dy INTERLEAVED as float2 so that a
// template element is one 8-byte load.  The planar layout of the reference's float** API exists only
// at the C-ABI boundary (launch_interleave / launch_deinterleave).
struct LevelPlanes {
  float* base;
  size_t frame_stride;   // floats
  size_t plane_stride;   // floats
};

// Everything one level of the inverse search needs (image_parameters + fix_parameters of
// include/optical_flow.hpp:14-37 and the grid geometry of src/patch_grid.cpp:20-25).
struct SearchArgs {
  LevelPlanes planes;
  const int2* pairs;        // (first frame, second frame) per pair
  const float* flow_prev;   // coarser dense flow [pair][h/2][w/2][2], or nullptr at the coarsest level
  size_t flow_prev_stride;  // floats per pair
  const float* init_override; // optional explicit per-patch init [pair][np][2] (PatchGrid API); else nullptr
  float* u_out;             // [pair][np][2]
  int* iters_out;           // [pair][np] or nullptr
  unsigned long long* counters;  // [kCounters]: patch_iterations, sampled_iterations, resets, start_oob, (4-13: DIS_DEBUG_STATS), [15] = internal error flag, [16]/[17] sampled passes of k_converge / of the walk (or nullptr)
  void* stragglers;         // hand-off list for patches that miss on every pass (or nullptr: no hand-off)
  int* straggler_count;     // [0] entries appended, [1] entries claimed by k_stragglers, [2] "walk finished" flag,
                            // [3] blocks of the walk that have started (device)
  int* work_counter;        // next patch index of the launch (device; zeroed by the launcher)
  void* walk_state;         // [n_pairs*np] 32-byte hand-off records of the convergence phase
  void* walk_seeds;         // [n_pairs*np][kWalkSeeds] uint4 {pos bits x, y, delta_u x, y}
  int straggler_cap;
  int n_pairs;
  int width, height, pad, tmp_w;
  int npw, nph, offw, offh, steps, np;
  int band_y0, band_h;      // grid rows covered by this launch (band_h <= 0: all of them)
  int warps_per_pair;
  int share;                // this launch may use 1/share of the GPU's resident block slots (concurrent groups)
  int iterations, patnorm;
  float lb, ub_w, ub_h;
  float out2max;            // largest fp32 s with sqrtf(s) <= outlierthresh: |start-pos| > thr <=> s > out2max
  float conv_eps2;          // conv_eps^2; 0 = off (reference behaviour)
  int handoff_misses;       // walk -> k_stragglers after this many misses of a patch (set by the launcher from the launch size)
  int walk_chunks;          // the walk's fast phase runs up to this many chunks of kChunk passes per round ...
  int walk_minrun;          // ... while at least this many lanes of the warp are still running
  int walk_age;             // rounds a waiting patch may be held back for a fuller service round
  int walk_grid;            // blocks of the walk's launch (k_stragglers yields its slot while fewer have started)
  int conv_slack;           // ... all but this many of its patches
  int conv_min;             // k_converge: a warp's batch may leave after this many passes once each of its patches has re-sampled a seen position
  float early_step;         // k_converge -> k_stragglers when the last passes still move by this much (px); 0 = never (set by the launcher)
  int tile_margin;          // ceil(outlierthresh) + 1: how far (px) a tap base can move from the start position's before a reset
  long long walk_capacity;  // patches the hand-off buffers (walk_state / walk_seeds) of this launch can hold (checked builds)
  float one;                // 1.0f, set by the launcher: a run-time constant that keeps ptxas from contracting packed mul + add (dis_search.cu)
};

struct DensifyArgs {
  const float* u;           // [pair][np][2]
  float* dense;             // [pair][h][w][2]
  size_t dense_stride;      // floats per pair
  int n_pairs, width, height, npw, nph, offw, offh, steps, np, ps;
  int y0, y1;               // pixel rows [y0, y1) to densify (y1 <= y0: all rows) — row-banded runs over several GPUs
};

// Development knobs, read from the environment ONCE per process (first use) and never in launch code.  The defaults
// are the measured optimum for the bench workload and within a few per cent of it for the other content types of
// tools/bench_configs.py; none of them changes results.
struct Tuning {
  int handoff[4];        // DIS_HANDOFF=a,b,c,d   walk -> k_stragglers after this many misses, by launch size class
  int walk[12];          // DIS_WALK_TUNE=c,m,a:...  per size class: chunks per round, lanes that must still run, service ageing
  int conv_min;          // DIS_CONV_MIN          k_converge: earliest pass at which a settled batch may leave
  int conv_slack;        // DIS_CONV_SLACK        ... with at most this many unsettled patches
  float early_step;      // DIS_EARLY_STEP        k_converge -> k_stragglers when the last passes still move by this much (px)
  long long strag_wide;  // DIS_STRAG_WIDE        launches below this many patches use 8 lanes per straggler
  int strag_dense;       // DIS_STRAG_DENSE       launches of at least this many pairs use the three-blocks-per-SM build of k_stragglers (ps 8)
  int share;             // DIS_SHARE             concurrent pair groups size their persistent grids to 1/groups of the GPU
  int groups;            // DIS_GROUPS            pair groups on concurrent streams (default 1)
  int side_blocks;       // DIS_SIDE_BLOCKS       blocks per SM of the k_stragglers launch that runs beside the walk (0: after it)
  int out_slices_n;      // DIS_OUT_SLICES=n or w1,w2,...   slices of the finest level in host-buffer calls (-1: default)
  double out_slices_w[16];
  int out_slices_nw;
};
const Tuning& tuning();
void reparse_tuning();   // after the environment changed (dis_tune, a development entry point)
int current_device();
// multiProcessorCount of the current device (cached per device)
int sm_count();

// Checked build (-DDIS_CHECKED, tools/variants.sh): device-side bounds assertions on every class of address the search
// kernels form (plane taps, template gathers, hand-off records, straggler list slots, shared-memory tiles).  A violated
// check records its code in counters[14]; the host turns it into DIS_ERR_CUDA.  compute-sanitizer is closed on the
// B200 pool this was developed on, so the GPU test-suite run against the checked build is the evidence instead
// (profiles/r2_checked_build_pytest.log).
#ifdef DIS_CHECKED
#define DIS_CHECK(A, cond, code) do { if (!(cond) && (A).counters) atomicMax((A).counters + 14, (unsigned long long)(code)); } while (0)
#else
#define DIS_CHECK(A, cond, code) ((void)0)
#endif

#ifndef DIS_SLOTS
#define DIS_SLOTS 16
#endif
constexpr int kCounters = 24;
constexpr int kStragglerBytes = 32;      // one entry of the walk -> k_stragglers hand-off list
constexpr int kWalkSeeds = DIS_SLOTS;   // entries of a patch's {position -> delta_u} cache (16 or 32)
constexpr size_t kWalkBytesPerPatch = 32 + 16 * kWalkSeeds;   // walk_state + walk_seeds

// host-callable launchers (each returns the number of kernels launched)
// Optional second stream on which k_stragglers runs CONCURRENTLY with the walk (fork/join: timing-disabled events).
struct SearchSide { cudaStream_t stream; cudaEvent_t fork, join; unsigned blocks; };
// ev: optional 4 events recorded before k_converge, before the walk, after the walk and after k_stragglers has joined
int launch_inverse_search(const SearchArgs& a, int patch_size, cudaStream_t s, cudaEvent_t* ev = nullptr,
                          const SearchSide* side = nullptr);
int launch_densify(const DensifyArgs& a, cudaStream_t s);
// grid rows [row0, row0 + nrows) of a level's per-patch flows (x-major: u[gx * nph + gy]) from src (this or a peer
// device, read over NVLink by the copying SMs) into dst
int launch_copy_u_rows(const float* src, float* dst, int npw, int nph, int row0, int nrows, cudaStream_t s);

struct PyramidArgs {
  const uint8_t* frames;    // [frame][h0][stride0] u8 (un-padded input frames), or nullptr if f32 input
  const float* frames_f32;  // [frame][H][W] fp32 (already divisibility padded) when frames == nullptr
  size_t frame_stride_bytes;
  int stride0, w0, h0;      // input geometry (u8 path)
  int W, H;                 // level-0 size after divisibility padding
  int left, top;            // divisibility padding split (src/main.cpp:152)
  int n_frames, levels, pad;
  LevelPlanes planes[16];
};
int launch_pyramid(const PyramidArgs& a, cudaStream_t s);
// planar dx, dy (n floats each) <-> interleaved float2 plane
int launch_interleave(const float* dx, const float* dy, float* gxy, size_t n, cudaStream_t s);
int launch_deinterleave(const float* gxy, float* dx, float* dy, size_t n, cudaStream_t s);

// crop the divisibility padding of the finest dense flow into the output layout
int launch_crop(const float* dense, size_t dense_stride, int W, int H, int left, int top, float* out,
                int w, int h, int n_pairs, cudaStream_t s);
// finest_scale > 0 epilogue (src/main.cpp:192-198): flow * 2^f, bilinear x2^f upsample, crop
int launch_upsample_crop(const float* dense, size_t dense_stride, int Wf, int Hf, int finest, int left,
                         int top, float* out, int w, int h, int n_pairs, cudaStream_t s);

// Middlebury colour coding (src/color_coding.cpp): flow [n_pixels][2] -> bgr [n_pixels][3]; scratch: one unsigned
int set_last_error(int code, const char* msg);   // dis_api.cu: records msg for dis_last_error(), returns code
int launch_color_code(const float* flow, size_t n_pixels, float maxmotion, unsigned* scratch, uint8_t* bgr, cudaStream_t s);

}  // namespace dis

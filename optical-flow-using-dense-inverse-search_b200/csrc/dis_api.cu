// dis_api.cu — the C ABI of include/dis_b200.h: handle, device memory, level loop.
//
// Host-side orchestration of the path the reference runs inside the OpticalFlowClass constructor
// (src/optical_flow.cpp:19-132) and around it in main() (src/main.cpp:135-198).  Everything that
// computes is a kernel in dis_pyramid.cu / dis_search.cu / dis_densify.cu; there is no CPU path.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>   // header-only: ranges cost nothing unless a tool (nsys, ncu --nvtx) is attached

#include "../../include/dis_b200.h"
#include "dis_kernels.cuh"

namespace {

// NVTX ranges around the stages of a run (the reference's only trace is `cout << "scale: N"`, src/optical_flow.cpp:69):
// "dis.pyramids", "dis.level<scale>.search" (k_converge + walk + k_stragglers), "dis.level<scale>.densify",
// "dis.epilogue", "dis.upload", "dis.download".  Host-side ranges around asynchronous launches: a tool's CUDA
// projection maps them onto the kernels.
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  NvtxRange(const char* fmt, int a) { char b[64]; snprintf(b, sizeof b, fmt, a); nvtxRangePushA(b); }
  ~NvtxRange() { nvtxRangePop(); }
};

thread_local std::string g_err;
int fail(int code, const std::string& msg) { g_err = msg; return code; }

#define CU(call)                                                                          \
  do {                                                                                    \
    cudaError_t _e = (call);                                                              \
    if (_e != cudaSuccess)                                                                \
      return fail(DIS_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(_e));      \
  } while (0)

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return -1; }
    cap = bytes;
    return 0;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

constexpr int kMaxGroups = 4;

struct Level {            // one entry per scale finest..coarsest, index s = scale - finest
  dis_level_info info;
  float lb, ub_w, ub_h;
  size_t flow_stride;     // floats per pair
  DevBuf flow, u, iters;
};

}  // namespace

// the error slot of dis_last_error() for the other translation units of the library (dis_noalloc.cu)
namespace dis { int set_last_error(int code, const char* msg) { return fail(code, msg ? msg : ""); } }

struct dis_engine {
  dis_params prm;
  int device = 0;
  int steps = 1;
  float outlier = 4.f;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // Pairs are independent: a run is split into up to kMaxGroups groups of pairs, each with its own
  // stream, so that the latency-bound tails of one group's level (few patches left, straggler chains)
  // overlap with the bulk of another's.
  std::vector<cudaStream_t> gstream;    // group streams (group 0 runs on `stream` itself)
  std::vector<cudaEvent_t> gdone;
  cudaEvent_t fork_ev = nullptr;
  cudaStream_t copy_stream = nullptr;   // host-buffer calls: flows of a slice of pairs go home while the next slice is searched
  std::vector<cudaEvent_t> slice_ev;
  dis::SearchSide side = {nullptr, nullptr, nullptr, 0};   // k_stragglers beside the walk (single-group runs)
  int n_groups = 1;                     // groups of the last run
  std::vector<cudaEvent_t> search_ev;   // [group][level][4]: before k_converge | before the walk | before k_stragglers | after
  int search_launches = 0;              // per group
  size_t straggler_stride = 0;          // bytes per group in `stragglers`

  // geometry of the current run
  int w0 = 0, h0 = 0, W = 0, H = 0, left = 0, top = 0, pad = 0;
  int n_frames = 0, n_pairs = 0, stride_u8 = 0;
  bool have_run = false;
  bool tiled = false;                   // between dis_tiled_begin and dis_tiled_finish
  std::vector<cudaEvent_t> tiled_ev;    // per level: "this engine's band of the level has been searched" (row-banded runs)
  int2 tiled_pr = {0, 1};               // the pair record of a tiled run (outlives the asynchronous upload)
  unsigned long long tiled_patches = 0; // patches this engine searched in the current tiled run
  cudaEvent_t tiled_t0 = nullptr, tiled_t1 = nullptr;   // device time of a row-banded run: after the upload .. before the download
  float tiled_ms = 0.f;
  bool stats_fetched = false;
  bool out_sent = false;                // the last run already copied its flows to the caller's buffer
  bool keep_iters = true;
  bool host_slices = true;              // dis_set_host_slices
  int max_groups = 1;                   // DIS_GROUPS overrides; measured: no gain on B200 (levels are work-bound, not tail-bound)

  DevBuf frames_u8, frames_f32, pairs, planes_all, out, counters, stragglers, stage, walk;
  size_t walk_patches = 0;              // capacity of `walk` in patches (largest level x pairs)
  size_t walk_pair_patches = 0;         // patches of the largest level of one pair (rounded up to 8)
  int straggler_cap = 0;
  dis::LevelPlanes planes[16];          // by absolute scale
  size_t plane_elems[16];
  std::vector<Level> levels;
  dis_stats stats;
  unsigned long long launches = 0;
};

namespace {

int check_params(const dis_params& p) {
  if (p.patch_size < 4 || (p.patch_size & 1) || p.patch_size > 16 || p.patch_size == 14)
    return fail(DIS_ERR_PARAM, "patch_size must be one of 4, 6, 8, 10, 12, 16 (odd sizes are broken in the reference, SURVEY.md F9)");
  if (p.coarsest_scale < p.finest_scale || p.finest_scale < 0 || p.coarsest_scale > 15)
    return fail(DIS_ERR_PARAM, "need 0 <= finest_scale <= coarsest_scale <= 15");
  if (p.max_iter < 0) return fail(DIS_ERR_PARAM, "max_iter must be >= 0");
  if (!(p.patch_overlap >= 0.f && p.patch_overlap < 1.f)) return fail(DIS_ERR_PARAM, "patch_overlap must be in [0,1)");
  if (p.steps < 0 || p.steps > p.patch_size) return fail(DIS_ERR_PARAM, "steps must be in [1, patch_size] (0 = derive from patch_overlap)");
  if (!(p.outlier_thresh >= 0.f)) return fail(DIS_ERR_PARAM, "outlier_thresh must be >= 0 (0 = patch_size/2)");
  return DIS_OK;
}

// Largest fp32 s with sqrtf(s) <= thr.  sqrtf is correctly rounded and monotone, so the reference's
// `norm() > outlierthresh` (src/patch.cpp:185) is exactly `dx*dx + dy*dy > out2max`.
float sq_threshold(float thr) {
  float s = thr * thr;
  while (sqrtf(s) > thr) s = nextafterf(s, 0.f);
  while (sqrtf(nextafterf(s, INFINITY)) <= thr) s = nextafterf(s, INFINITY);
  return s;
}

// src/optical_flow.cpp:38
int steps_of(int ps, float overlap) {
  const float one_minus = 1 - overlap;
  const int s = (int)floorf((float)ps * one_minus);
  return std::max(1, s);
}

// (Re)define the geometry of a run and size every device buffer.  W,H: level-0 size after the
// divisibility padding; pad: img_padding.
int setup_run(dis_engine* e, int W, int H, int pad, int n_frames, int n_pairs) {
  const dis_params& p = e->prm;
  const int sf = 1 << p.coarsest_scale;
  if (W <= 0 || H <= 0 || (W % sf) || (H % sf))
    return fail(DIS_ERR_PARAM, "width/height must be positive multiples of 2^coarsest_scale");
  if (pad < p.patch_size) return fail(DIS_ERR_PARAM, "img_padding must be >= patch_size");
  if ((W >> p.coarsest_scale) < 2 || (H >> p.coarsest_scale) < 2)
    return fail(DIS_ERR_PARAM, "coarsest level would be smaller than 2x2");
  if (n_frames < 1 || n_pairs < 1) return fail(DIS_ERR_PARAM, "need at least one frame and one pair");
  e->W = W; e->H = H; e->pad = pad; e->n_frames = n_frames; e->n_pairs = n_pairs;
  // planes for ALL scales 0..coarsest (construct_pyramide builds every level, src/main.cpp:177)
  size_t total = 0;
  size_t off[16];
  for (int s = 0; s <= p.coarsest_scale; ++s) {
    const size_t tw = (W >> s) + 2 * pad, th = (H >> s) + 2 * pad;
    e->plane_elems[s] = (tw * th + 3) & ~(size_t)3;
    off[s] = total;
    total += 3 * e->plane_elems[s] * (size_t)n_frames;
  }
  if (e->planes_all.ensure(total * sizeof(float))) return fail(DIS_ERR_ALLOC, "cudaMalloc(planes) failed");
  for (int s = 0; s <= p.coarsest_scale; ++s) {
    e->planes[s].base = e->planes_all.as<float>() + off[s];
    e->planes[s].plane_stride = e->plane_elems[s];
    e->planes[s].frame_stride = 3 * e->plane_elems[s];
  }
  const int nl = p.coarsest_scale - p.finest_scale + 1;
  e->levels.resize(nl);
  for (int s = 0; s < nl; ++s) {
    Level& L = e->levels[s];
    const int scale = p.finest_scale + s;
    dis_level_info& I = L.info;
    const float sc = (float)pow(2, -scale);            // src/optical_flow.cpp:51-59
    I.scale = scale;
    I.height = (int)(H * sc);
    I.width = (int)(W * sc);
    I.tmp_w = I.width + 2 * pad;
    I.tmp_h = I.height + 2 * pad;
    I.steps = e->steps;
    L.lb = -(float)p.patch_size / 2;
    L.ub_w = (float)(I.width + p.patch_size / 2 - 2);
    L.ub_h = (float)(I.height + p.patch_size / 2 - 2);
    I.npw = (int)ceilf((float)I.width / (float)I.steps);   // src/patch_grid.cpp:20-23
    I.nph = (int)ceilf((float)I.height / (float)I.steps);
    I.offw = (I.width - (I.npw - 1) * I.steps) / 2;
    I.offh = (I.height - (I.nph - 1) * I.steps) / 2;
    I.num_patches = I.npw * I.nph;
    L.flow_stride = 2 * (size_t)I.width * I.height;
    if (L.flow.ensure(L.flow_stride * n_pairs * sizeof(float)) ||
        L.u.ensure(2 * (size_t)I.num_patches * n_pairs * sizeof(float)) ||
        L.iters.ensure((size_t)I.num_patches * n_pairs * sizeof(int)))
      return fail(DIS_ERR_ALLOC, "cudaMalloc(level buffers) failed");
  }
  if (e->pairs.ensure(sizeof(int2) * (size_t)n_pairs)) return fail(DIS_ERR_ALLOC, "cudaMalloc(pairs) failed");
  if (e->counters.ensure(sizeof(unsigned long long) * dis::kCounters)) return fail(DIS_ERR_ALLOC, "cudaMalloc(counters) failed");
  // hand-off list of the inverse search (20 bytes per entry + the counter in front)
  {
    size_t maxp = 0;
    for (const Level& L : e->levels) maxp = std::max(maxp, (size_t)L.info.num_patches * n_pairs);
    e->straggler_cap = (int)std::min<size_t>(maxp / 8 + 4096, (size_t)1 << 28);
    e->straggler_stride = (64 + (size_t)e->straggler_cap * dis::kStragglerBytes + 255) & ~(size_t)255;
    if (e->stragglers.ensure(kMaxGroups * e->straggler_stride)) return fail(DIS_ERR_ALLOC, "cudaMalloc(stragglers) failed");
    // hand-off between the two phases of a level: 32-byte records first, then the seed entries
    size_t maxnp = 0;
    for (const Level& L : e->levels) maxnp = std::max(maxnp, (size_t)L.info.num_patches);
    e->walk_pair_patches = (maxnp + 7) & ~(size_t)7;
    e->walk_patches = e->walk_pair_patches * n_pairs;
    if (e->walk.ensure(e->walk_patches * dis::kWalkBytesPerPatch)) return fail(DIS_ERR_ALLOC, "cudaMalloc(walk hand-off) failed");
  }
  return DIS_OK;
}

}  // namespace

namespace {

// Everything the search kernels need for level index s (= scale - finest) and pairs [p0, p1) of the run.
dis::SearchArgs level_search_args(dis_engine* e, int g, int s, int p0, int p1) {
  const dis_params& p = e->prm;
  const int nl = (int)e->levels.size();
  char* sbase = e->stragglers.as<char>() + (size_t)g * e->straggler_stride;
  Level& L = e->levels[s];
  const dis_level_info& I = L.info;
  dis::SearchArgs a;
  memset(&a, 0, sizeof a);
  a.planes = e->planes[I.scale];
  a.pairs = e->pairs.as<int2>() + p0;
  a.flow_prev = (s < nl - 1) ? e->levels[s + 1].flow.as<float>() + e->levels[s + 1].flow_stride * p0 : nullptr;
  a.flow_prev_stride = (s < nl - 1) ? e->levels[s + 1].flow_stride : 0;
  a.init_override = nullptr;
  a.u_out = L.u.as<float>() + 2 * (size_t)I.num_patches * p0;
  a.iters_out = e->keep_iters ? L.iters.as<int>() + (size_t)I.num_patches * p0 : nullptr;
  a.counters = e->counters.as<unsigned long long>();
  a.straggler_count = reinterpret_cast<int*>(sbase);
  a.work_counter = reinterpret_cast<int*>(sbase) + 8;
  a.stragglers = sbase + 64;
  a.straggler_cap = e->straggler_cap;
  // Regions are per PAIR at the size of the largest level, so that groups of pairs that work on different levels at
  // the same time (concurrent streams) never share hand-off records.
  a.walk_state = e->walk.as<char>() + 32 * e->walk_pair_patches * p0;
  a.walk_seeds = e->walk.as<char>() + 32 * e->walk_patches + 16 * (size_t)dis::kWalkSeeds * e->walk_pair_patches * p0;
  a.n_pairs = p1 - p0;
  a.walk_capacity = (long long)e->walk_pair_patches * (e->n_pairs - p0);
  a.share = dis::tuning().share ? e->n_groups : 1;   // full-size grids: a later group fills the SMs an earlier group's tail leaves idle
  a.width = I.width; a.height = I.height; a.pad = e->pad; a.tmp_w = I.tmp_w;
  a.npw = I.npw; a.nph = I.nph; a.offw = I.offw; a.offh = I.offh; a.steps = I.steps; a.np = I.num_patches;
  a.warps_per_pair = (I.num_patches + 31) / 32;
  a.iterations = p.max_iter; a.patnorm = p.patch_norm ? 1 : 0;
  a.lb = L.lb; a.ub_w = L.ub_w; a.ub_h = L.ub_h;
  a.out2max = sq_threshold(e->outlier); a.tile_margin = (int)ceilf(e->outlier) + 1;
  a.conv_eps2 = p.conv_eps * p.conv_eps;
  return a;
}

// PatchGrid::densification_and_create_dance_flow for level s, pairs [p0, p1)
void level_densify(dis_engine* e, int s, int p0, int p1, cudaStream_t st, int y0 = 0, int y1 = 0) {
  Level& L = e->levels[s];
  const dis_level_info& I = L.info;
  dis::DensifyArgs d;
  d.u = L.u.as<float>() + 2 * (size_t)I.num_patches * p0;
  d.dense = L.flow.as<float>() + L.flow_stride * p0; d.dense_stride = L.flow_stride;
  d.n_pairs = p1 - p0; d.width = I.width; d.height = I.height; d.npw = I.npw; d.nph = I.nph;
  d.offw = I.offw; d.offh = I.offh; d.steps = I.steps; d.np = I.num_patches; d.ps = e->prm.patch_size;
  d.y0 = y0; d.y1 = y1;
  e->launches += dis::launch_densify(d, st);
}

int run_epilogue(dis_engine* e, int p0, int p1);

// The coarse-to-fine loop of src/optical_flow.cpp:67-90 for pairs [p0, p1) of the run on stream `st`.
// host_out != null (host-buffer calls, one group): the finest level runs in slices of pairs, and each slice's
// cropped flows are copied to the caller's buffer on `copy_stream` while the next slice is being searched; the
// epilogue and the download are then done when this returns (the engine's stream waits for the last copy).
int run_levels_group(dis_engine* e, int g, int p0, int p1, cudaStream_t st, float* host_out = nullptr) {
  const int nl = (int)e->levels.size();
  int launch = 0;
  const dis::SearchSide* side = (e->n_groups == 1 && e->side.stream) ? &e->side : nullptr;
  for (int s = nl - 1; s >= 0; --s, ++launch) {
    cudaEvent_t* ev = &e->search_ev[4 * (g * nl + launch)];
    int slices = 1;
    std::vector<int> cut{p0, p1};         // slice c = pairs [cut[c], cut[c + 1])
    if (s == 0 && host_out) {
      // DIS_OUT_SLICES: a count (equal slices) or a comma list of weights, e.g. "4,3,2,1": the last slice's
      // copy is the only one nothing hides, so a tapering split shortens the exposed tail.
      const dis::Tuning& tn = dis::tuning();
      std::vector<double> wt;
      if (tn.out_slices_nw > 0) {
        wt.assign(tn.out_slices_w, tn.out_slices_w + tn.out_slices_nw);
      } else if (tn.out_slices_n < 0 && p1 - p0 >= 32) {
        wt = {3.0, 1.0};   // measured on the 49-pair sequence (profiles/r1k_e2e_slices.log): every further slice costs a launch tail
      } else {
        const int n = tn.out_slices_n >= 0 ? tn.out_slices_n : (p1 - p0 >= 8 ? 2 : 1);
        wt.assign(std::max(1, std::min(n, p1 - p0)), 1.0);
      }
      double tot = 0, run = 0;
      for (double x : wt) tot += x;
      cut.assign(1, p0);
      for (double x : wt) {
        run += x;
        const int q = tot > 0 ? p0 + (int)std::llround((p1 - p0) * run / tot) : p1;
        if (q > cut.back()) cut.push_back(std::min(q, p1));
      }
      if (cut.back() != p1) cut.push_back(p1);
      slices = (int)cut.size() - 1;
      while ((int)e->slice_ev.size() < slices + 1) {
        cudaEvent_t x; CU(cudaEventCreateWithFlags(&x, cudaEventDisableTiming)); e->slice_ev.push_back(x);
      }
    }
    for (int c = 0; c < slices; ++c) {
      const int q0 = cut[c], q1 = cut[c + 1];
      const dis::SearchArgs a = level_search_args(e, g, s, q0, q1);
      int n;
      {
        NvtxRange r("dis.level%d.search", e->levels[s].info.scale);
        n = dis::launch_inverse_search(a, e->prm.patch_size, st, c == 0 ? ev : nullptr, side);   // timings: first slice only
      }
      if (n < 0) return fail(DIS_ERR_PARAM, "unsupported patch_size");
      e->launches += n;
      {
        NvtxRange r("dis.level%d.densify", e->levels[s].info.scale);
        level_densify(e, s, q0, q1, st);
      }
      if (s == 0 && host_out) {
        if (int rc = run_epilogue(e, q0, q1)) return rc;
        const size_t off = 2 * (size_t)e->w0 * e->h0 * q0, cnt = 2 * (size_t)e->w0 * e->h0 * (q1 - q0);
        CU(cudaEventRecord(e->slice_ev[c], st));
        CU(cudaStreamWaitEvent(e->copy_stream, e->slice_ev[c], 0));
        CU(cudaMemcpyAsync(host_out + off, e->out.as<float>() + off, sizeof(float) * cnt, cudaMemcpyDeviceToHost, e->copy_stream));
      }
    }
    if (s == 0 && host_out) {
      CU(cudaEventRecord(e->slice_ev[slices], e->copy_stream));
      CU(cudaStreamWaitEvent(st, e->slice_ev[slices], 0));
    }
  }
  return DIS_OK;
}

int run_levels(dis_engine* e, float* host_out = nullptr) {
  const int nl = (int)e->levels.size();
  CU(cudaMemsetAsync(e->counters.p, 0, dis::kCounters * sizeof(unsigned long long), e->stream));
  const int G = std::max(1, std::min(std::min(kMaxGroups, e->max_groups), e->n_pairs));
  while ((int)e->search_ev.size() < 4 * nl * G) {
    cudaEvent_t ev; CU(cudaEventCreate(&ev)); e->search_ev.push_back(ev);
  }
  while ((int)e->gstream.size() < G - 1) {
    cudaStream_t st; cudaEvent_t ev;
    CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    e->gstream.push_back(st); e->gdone.push_back(ev);
  }
  if (!e->fork_ev) CU(cudaEventCreateWithFlags(&e->fork_ev, cudaEventDisableTiming));
  e->n_groups = G;
  e->search_launches = nl;
  if (host_out && (G > 1 || e->n_pairs < 8 || (dis::tuning().out_slices_nw == 0 && dis::tuning().out_slices_n >= 0 && dis::tuning().out_slices_n <= 1))) host_out = nullptr;
  if (!e->host_slices) host_out = nullptr;
  if (host_out && !e->copy_stream) CU(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
  if (host_out && e->out.ensure(sizeof(float) * 2 * (size_t)e->w0 * e->h0 * e->n_pairs)) return fail(DIS_ERR_ALLOC, "cudaMalloc(out) failed");
  e->out_sent = host_out != nullptr;
  if (G > 1) CU(cudaEventRecord(e->fork_ev, e->stream));
  for (int g = G - 1; g >= 0; --g) {          // group 0 last: it runs on the engine's own stream
    const int p0 = (int)((long long)e->n_pairs * g / G), p1 = (int)((long long)e->n_pairs * (g + 1) / G);
    cudaStream_t st = g == 0 ? e->stream : e->gstream[g - 1];
    if (g > 0) CU(cudaStreamWaitEvent(st, e->fork_ev, 0));
    if (int rc = run_levels_group(e, g, p0, p1, st, host_out)) return rc;
    if (g > 0) CU(cudaEventRecord(e->gdone[g - 1], st));
  }
  for (int g = 1; g < G; ++g) CU(cudaStreamWaitEvent(e->stream, e->gdone[g - 1], 0));
  CU(cudaGetLastError());
  unsigned long long patches = 0;
  for (const Level& L : e->levels) patches += (unsigned long long)L.info.num_patches * e->n_pairs;
  e->stats.patches = patches;
  e->have_run = true;
  e->stats_fetched = false;
  return DIS_OK;
}

int fetch_counters(dis_engine* e) {
  unsigned long long c[dis::kCounters];
  CU(cudaMemcpyAsync(c, e->counters.p, sizeof c, cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
#ifdef DIS_DEBUG_STATS
  fprintf(stderr, "[dis debug] walk: warp-iterations %llu | running lanes/iter %.2f free lanes/iter %.2f | passes advanced per running lane-chunk %.2f | "
                  "service: rounds/iter %.3f, patches/round %.2f, waiting at decision/iter %.2f | slow-path hits %llu | binds %llu in %llu events (%.1f lanes/event)\n",
          c[4], (double)c[5] / c[4], (double)c[13] / c[4], (double)c[6] / (c[5] ? c[5] : 1), (double)c[7] / c[4], (double)c[8] / (c[7] ? c[7] : 1),
          (double)c[12] / c[4], c[9], c[10], c[11], (double)c[10] / (c[11] ? c[11] : 1));
#endif
  if (c[14]) {
    static char msg[96];
    snprintf(msg, sizeof msg, "checked build: device assertion %llu failed (csrc/dis_search.cu DIS_CHECK)", c[14]);
    return fail(DIS_ERR_CUDA, msg);
  }
  if (c[15]) return fail(DIS_ERR_CUDA, "internal: k_stragglers timed out waiting for the walk's hand-offs; results are incomplete");
  e->stats.patch_iterations = c[0]; e->stats.sampled_iterations = c[1];
  e->stats.resets = c[2]; e->stats.start_oob = c[3];
  e->stats.sampled_converge = c[16]; e->stats.sampled_walk = c[17];
  e->stats.kernel_launches = e->launches;
  e->stats_fetched = true;
  return DIS_OK;
}

int run_pyramid_u8(dis_engine* e, int stride) {
  dis::PyramidArgs a;
  memset(&a, 0, sizeof a);
  a.frames = e->frames_u8.as<uint8_t>();
  a.frame_stride_bytes = (size_t)stride * e->h0;
  a.stride0 = stride; a.w0 = e->w0; a.h0 = e->h0;
  a.W = e->W; a.H = e->H; a.left = e->left; a.top = e->top;
  a.n_frames = e->n_frames; a.levels = e->prm.coarsest_scale + 1; a.pad = e->pad;
  for (int s = 0; s <= e->prm.coarsest_scale; ++s) a.planes[s] = e->planes[s];
  NvtxRange r("dis.pyramids");
  e->launches += dis::launch_pyramid(a, e->stream);
  CU(cudaGetLastError());
  return DIS_OK;
}

// finest dense flow -> cropped output (src/main.cpp:192-198), device to device, pairs [p0, p1)
int run_epilogue(dis_engine* e, int p0, int p1) {
  NvtxRange r("dis.epilogue");
  Level& L = e->levels[0];
  if (e->out.ensure(sizeof(float) * 2 * (size_t)e->w0 * e->h0 * e->n_pairs)) return fail(DIS_ERR_ALLOC, "cudaMalloc(out) failed");
  const float* dense = L.flow.as<float>() + L.flow_stride * p0;
  float* out = e->out.as<float>() + 2 * (size_t)e->w0 * e->h0 * p0;
  if (e->prm.finest_scale == 0)
    e->launches += dis::launch_crop(dense, L.flow_stride, L.info.width, L.info.height, e->left, e->top,
                                    out, e->w0, e->h0, p1 - p0, e->stream);
  else
    e->launches += dis::launch_upsample_crop(dense, L.flow_stride, L.info.width, L.info.height,
                                             e->prm.finest_scale, e->left, e->top, out, e->w0, e->h0,
                                             p1 - p0, e->stream);
  CU(cudaGetLastError());
  return DIS_OK;
}
int run_epilogue(dis_engine* e) { return run_epilogue(e, 0, e->n_pairs); }

int div_pad(int dim, int coarsest) {   // src/main.cpp:139-149
  const int sf = 1 << coarsest, r = dim % sf;
  return r > 0 ? sf - r : 0;
}

}  // namespace

extern "C" {

void dis_default_params(dis_params* p) {   // src/main.cpp:63-72
  if (!p) return;
  memset(p, 0, sizeof *p);
  p->max_iter = 1000; p->patch_size = 8; p->coarsest_scale = 3; p->finest_scale = 0;
  p->patch_overlap = 0.7f; p->patch_norm = 1; p->conv_eps = 0.f;
}

const char* dis_last_error(void) { return g_err.c_str(); }
const char* dis_version(void) { return "dis_b200 0.2 (sm_100a)"; }

int dis_tune(const char* name, const char* value) {
  if (!name || strncmp(name, "DIS_", 4) != 0) return fail(DIS_ERR_PARAM, "dis_tune: knob names start with DIS_");
  if (value) setenv(name, value, 1); else unsetenv(name);
  dis::reparse_tuning();
  return DIS_OK;
}

int dis_create(const dis_params* p, int device, dis_engine** out) {
  if (!p || !out) return fail(DIS_ERR_PARAM, "null argument");
  *out = nullptr;
  if (int rc = check_params(*p)) return rc;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    return fail(DIS_ERR_CUDA, "no CUDA device: this engine has no CPU path");
  }
  if (device < 0 || device >= n) return fail(DIS_ERR_PARAM, "device ordinal out of range");
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return fail(DIS_ERR_CUDA, "kernels are built for sm_100a only");
  CU(cudaSetDevice(device));
  dis_engine* e = new dis_engine();
  e->prm = *p; e->device = device;
  e->steps = p->steps > 0 ? p->steps : steps_of(p->patch_size, p->patch_overlap);
  e->outlier = p->outlier_thresh > 0.f ? p->outlier_thresh : (float)p->patch_size / 2;   // src/optical_flow.cpp:34
  e->max_groups = dis::tuning().groups;
  {
    // k_stragglers beside the walk: side_blocks blocks per SM (0 = after the walk instead)
    e->side.blocks = (unsigned)(dis::tuning().side_blocks * prop.multiProcessorCount);
    if (e->side.blocks > 0 &&
        (cudaStreamCreateWithFlags(&e->side.stream, cudaStreamNonBlocking) != cudaSuccess ||
         cudaEventCreateWithFlags(&e->side.fork, cudaEventDisableTiming) != cudaSuccess ||
         cudaEventCreateWithFlags(&e->side.join, cudaEventDisableTiming) != cudaSuccess)) {
      cudaGetLastError();
      e->side.stream = nullptr;
    }
  }
  memset(&e->stats, 0, sizeof e->stats);
  if (cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&e->ev0) != cudaSuccess || cudaEventCreate(&e->ev1) != cudaSuccess) {
    delete e;
    return fail(DIS_ERR_CUDA, "stream/event creation failed");
  }
  *out = e;
  return DIS_OK;
}

void dis_destroy(dis_engine* e) {
  if (!e) return;
  cudaSetDevice(e->device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  for (Level& L : e->levels) { L.flow.release(); L.u.release(); L.iters.release(); }
  e->frames_u8.release(); e->frames_f32.release(); e->pairs.release(); e->planes_all.release(); e->stragglers.release(); e->walk.release();
  e->stage.release();
  e->out.release(); e->counters.release();
  for (cudaEvent_t ev : e->search_ev) cudaEventDestroy(ev);
  if (e->side.stream) { cudaStreamDestroy(e->side.stream); cudaEventDestroy(e->side.fork); cudaEventDestroy(e->side.join); }
  for (cudaEvent_t ev : e->gdone) cudaEventDestroy(ev);
  for (cudaStream_t st : e->gstream) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
  if (e->fork_ev) cudaEventDestroy(e->fork_ev);
  for (cudaEvent_t ev : e->slice_ev) cudaEventDestroy(ev);
  if (e->copy_stream) { cudaStreamSynchronize(e->copy_stream); cudaStreamDestroy(e->copy_stream); }
  if (e->ev0) cudaEventDestroy(e->ev0);
  if (e->ev1) cudaEventDestroy(e->ev1);
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
}

int dis_set_host_slices(dis_engine* e, int enabled) {
  if (!e) return fail(DIS_ERR_PARAM, "null engine");
  e->host_slices = enabled != 0;
  return DIS_OK;
}

int dis_upload_frames_u8(dis_engine* e, const uint8_t* frames, int n_frames, int w, int h, int stride,
                         size_t frame_stride, const int* first, const int* second, int n_pairs) {
  if (!e || !frames || !first || !second) return fail(DIS_ERR_PARAM, "null argument");
  if (w <= 0 || h <= 0 || stride < w || frame_stride < (size_t)stride * h) return fail(DIS_ERR_PARAM, "bad frame geometry");
  CU(cudaSetDevice(e->device));
  const int padw = div_pad(w, e->prm.coarsest_scale), padh = div_pad(h, e->prm.coarsest_scale);
  e->w0 = w; e->h0 = h; e->left = padw / 2; e->top = padh / 2;
  if (int rc = setup_run(e, w + padw, h + padh, e->prm.patch_size, n_frames, n_pairs)) return rc;
  std::vector<int2> pr(n_pairs);
  for (int i = 0; i < n_pairs; ++i) {
    if (first[i] < 0 || first[i] >= n_frames || second[i] < 0 || second[i] >= n_frames)
      return fail(DIS_ERR_PARAM, "pair index out of range");
    pr[i] = make_int2(first[i], second[i]);
  }
  const size_t fbytes = (size_t)stride * h;
  if (e->frames_u8.ensure(fbytes * n_frames)) return fail(DIS_ERR_ALLOC, "cudaMalloc(frames) failed");
  if (frame_stride == fbytes) {
    CU(cudaMemcpyAsync(e->frames_u8.p, frames, fbytes * n_frames, cudaMemcpyHostToDevice, e->stream));
  } else {
    CU(cudaMemcpy2DAsync(e->frames_u8.p, fbytes, frames, frame_stride, fbytes, n_frames, cudaMemcpyHostToDevice, e->stream));
  }
  CU(cudaMemcpyAsync(e->pairs.p, pr.data(), sizeof(int2) * n_pairs, cudaMemcpyHostToDevice, e->stream));
  CU(cudaStreamSynchronize(e->stream));   // pr is a host temporary
  e->stride_u8 = stride;
  e->have_run = false;
  return DIS_OK;
}

int dis_run_resident(dis_engine* e) {
  if (!e) return fail(DIS_ERR_PARAM, "null engine");
  if (!e->frames_u8.p || e->n_pairs <= 0) return fail(DIS_ERR_STATE, "dis_upload_frames_u8 first");
  CU(cudaSetDevice(e->device));
  e->launches = 0;
  if (int rc = run_pyramid_u8(e, e->stride_u8)) return rc;
  if (int rc = run_levels(e)) return rc;
  return run_epilogue(e);
}

int dis_download_flows(dis_engine* e, float* flow_out) {
  if (!e || !flow_out) return fail(DIS_ERR_PARAM, "null argument");
  if (!e->have_run) return fail(DIS_ERR_STATE, "no run to download");
  CU(cudaSetDevice(e->device));
  CU(cudaMemcpyAsync(flow_out, e->out.p, sizeof(float) * 2 * (size_t)e->w0 * e->h0 * e->n_pairs,
                     cudaMemcpyDeviceToHost, e->stream));
  return fetch_counters(e);   // synchronises; also surfaces the kernels' internal error flag
}

int dis_synchronize(dis_engine* e) {
  if (!e) return fail(DIS_ERR_PARAM, "null engine");
  CU(cudaStreamSynchronize(e->stream));
  return DIS_OK;
}

int dis_timer_start(dis_engine* e) {
  if (!e) return fail(DIS_ERR_PARAM, "null engine");
  CU(cudaEventRecord(e->ev0, e->stream));
  return DIS_OK;
}

int dis_timer_stop_ms(dis_engine* e, float* ms) {
  if (!e || !ms) return fail(DIS_ERR_PARAM, "null argument");
  CU(cudaEventRecord(e->ev1, e->stream));
  CU(cudaEventSynchronize(e->ev1));
  CU(cudaEventElapsedTime(ms, e->ev0, e->ev1));
  return DIS_OK;
}

int dis_last_search_level_ms(dis_engine* e, int level, float* ms) {
  if (!e || !ms) return fail(DIS_ERR_PARAM, "null argument");
  if (!e->have_run) return fail(DIS_ERR_STATE, "no run yet");
  const int nl = (int)e->levels.size();
  if (level < 0 || level >= nl || e->search_launches != nl) return fail(DIS_ERR_PARAM, "level out of range");
  CU(cudaStreamSynchronize(e->stream));
  const int i = nl - 1 - level;   // launches go coarsest -> finest
  *ms = 0.f;
  for (int g = 0; g < e->n_groups; ++g) {   // groups run concurrently: this is the sum of their device times
    float t = 0.f;
    CU(cudaEventElapsedTime(&t, e->search_ev[4 * (g * nl + i)], e->search_ev[4 * (g * nl + i) + 3]));
    *ms += t;
  }
  return DIS_OK;
}

int dis_last_search_phase_ms(dis_engine* e, int level, int phase, float* ms) {
  if (!e || !ms) return fail(DIS_ERR_PARAM, "null argument");
  if (!e->have_run) return fail(DIS_ERR_STATE, "no run yet");
  const int nl = (int)e->levels.size();
  if (level < 0 || level >= nl || e->search_launches != nl || phase < 0 || phase > 2) return fail(DIS_ERR_PARAM, "level or phase out of range");
  CU(cudaStreamSynchronize(e->stream));
  const int i = nl - 1 - level;
  *ms = 0.f;
  for (int g = 0; g < e->n_groups; ++g) {
    float t = 0.f;
    CU(cudaEventElapsedTime(&t, e->search_ev[4 * (g * nl + i) + phase], e->search_ev[4 * (g * nl + i) + phase + 1]));
    *ms += t;
  }
  return DIS_OK;
}

int dis_last_search_ms(dis_engine* e, float* ms, int* launches) {
  if (!e || !ms) return fail(DIS_ERR_PARAM, "null argument");
  if (!e->have_run) return fail(DIS_ERR_STATE, "no run yet");
  CU(cudaStreamSynchronize(e->stream));
  float total = 0.f;
  for (int i = 0; i < e->search_launches * e->n_groups; ++i) {
    float t = 0.f;
    CU(cudaEventElapsedTime(&t, e->search_ev[4 * i], e->search_ev[4 * i + 3]));
    total += t;
  }
  *ms = total;
  if (launches) *launches = e->search_launches * e->n_groups;
  return DIS_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// host-buffer entry points
// ---------------------------------------------------------------------------------------------
namespace {

constexpr int kMaxChunkPairs = 128;   // bounds device memory per chunk (~18 MB per Sintel pair)

// Upload `n_frames` separately addressed frames, run everything, download the cropped flows.
int run_frames(dis_engine* e, const uint8_t* const* fptr, int n_frames, int w, int h, int stride,
               const int* first, const int* second, int n_pairs, float* flow_out) {
  CU(cudaSetDevice(e->device));
  const int padw = div_pad(w, e->prm.coarsest_scale), padh = div_pad(h, e->prm.coarsest_scale);
  e->w0 = w; e->h0 = h; e->left = padw / 2; e->top = padh / 2;
  if (int rc = setup_run(e, w + padw, h + padh, e->prm.patch_size, n_frames, n_pairs)) return rc;
  const size_t fbytes = (size_t)w * h;   // frames are stored densely on the device
  if (e->frames_u8.ensure(fbytes * n_frames)) return fail(DIS_ERR_ALLOC, "cudaMalloc(frames) failed");
  bool dense = stride == w;                                   // one copy when the frames are back to back in host memory
  for (int f = 1; f < n_frames && dense; ++f) dense = fptr[f] == fptr[f - 1] + fbytes;
  if (dense) CU(cudaMemcpyAsync(e->frames_u8.p, fptr[0], fbytes * n_frames, cudaMemcpyHostToDevice, e->stream));
  else
    for (int f = 0; f < n_frames; ++f)
      CU(cudaMemcpy2DAsync(e->frames_u8.as<uint8_t>() + fbytes * f, w, fptr[f], stride, w, h,
                           cudaMemcpyHostToDevice, e->stream));
  std::vector<int2> pr(n_pairs);
  for (int i = 0; i < n_pairs; ++i) pr[i] = make_int2(first[i], second[i]);
  CU(cudaMemcpyAsync(e->pairs.p, pr.data(), sizeof(int2) * n_pairs, cudaMemcpyHostToDevice, e->stream));
  e->stride_u8 = w;
  e->launches = 0;
  if (int rc = run_pyramid_u8(e, w)) return rc;
  if (int rc = run_levels(e, flow_out)) return rc;
  if (!e->out_sent) {           // else: cropped and copied slice by slice inside the finest level
    if (int rc = run_epilogue(e)) return rc;
    CU(cudaMemcpyAsync(flow_out, e->out.p, sizeof(float) * 2 * fbytes * n_pairs, cudaMemcpyDeviceToHost, e->stream));
  }
  return fetch_counters(e);   // synchronises; pr stays alive until here
}

void add_stats(dis_stats* acc, const dis_stats& s) {
  acc->patches += s.patches; acc->patch_iterations += s.patch_iterations;
  acc->sampled_iterations += s.sampled_iterations; acc->resets += s.resets;
  acc->start_oob += s.start_oob; acc->kernel_launches += s.kernel_launches;
  acc->sampled_converge += s.sampled_converge; acc->sampled_walk += s.sampled_walk;
}

}  // namespace

extern "C" {

int dis_run_pair_u8(dis_engine* e, const uint8_t* a, const uint8_t* b, int w, int h, int stride, float* flow_out) {
  if (!e || !a || !b || !flow_out) return fail(DIS_ERR_PARAM, "null argument");
  if (w <= 0 || h <= 0 || stride < w) return fail(DIS_ERR_PARAM, "bad frame geometry");
  const uint8_t* f[2] = {a, b};
  const int first = 0, second = 1;
  return run_frames(e, f, 2, w, h, stride, &first, &second, 1, flow_out);
}

int dis_run_batch_u8(dis_engine* e, const uint8_t* const* a, const uint8_t* const* b, int n_pairs, int w, int h,
                     int stride, float* flow_out) {
  if (!e || !a || !b || !flow_out) return fail(DIS_ERR_PARAM, "null argument");
  if (n_pairs <= 0 || w <= 0 || h <= 0 || stride < w) return fail(DIS_ERR_PARAM, "bad batch geometry");
  dis_stats acc; memset(&acc, 0, sizeof acc);
  std::vector<const uint8_t*> f;
  std::vector<int> fi, se;
  for (int base = 0; base < n_pairs; base += kMaxChunkPairs) {
    const int n = std::min(kMaxChunkPairs, n_pairs - base);
    f.resize(2 * n); fi.resize(n); se.resize(n);
    for (int i = 0; i < n; ++i) {
      if (!a[base + i] || !b[base + i]) return fail(DIS_ERR_PARAM, "null frame pointer");
      f[2 * i] = a[base + i]; f[2 * i + 1] = b[base + i]; fi[i] = 2 * i; se[i] = 2 * i + 1;
    }
    if (int rc = run_frames(e, f.data(), 2 * n, w, h, stride, fi.data(), se.data(), n,
                            flow_out + (size_t)base * 2 * w * h)) return rc;
    add_stats(&acc, e->stats);
  }
  if (n_pairs > kMaxChunkPairs) e->stats = acc;
  return DIS_OK;
}

int dis_run_sequence_u8(dis_engine* e, const uint8_t* frames, int n_frames, int w, int h, int stride,
                        size_t frame_stride, float* flow_out) {
  if (!e || !frames || !flow_out) return fail(DIS_ERR_PARAM, "null argument");
  if (n_frames < 2 || w <= 0 || h <= 0 || stride < w || frame_stride < (size_t)stride * h)
    return fail(DIS_ERR_PARAM, "bad sequence geometry");
  dis_stats acc; memset(&acc, 0, sizeof acc);
  std::vector<const uint8_t*> f;
  std::vector<int> fi, se;
  const int n_pairs = n_frames - 1;
  for (int base = 0; base < n_pairs; base += kMaxChunkPairs) {
    const int n = std::min(kMaxChunkPairs, n_pairs - base);
    f.resize(n + 1); fi.resize(n); se.resize(n);
    for (int i = 0; i <= n; ++i) f[i] = frames + frame_stride * (size_t)(base + i);
    for (int i = 0; i < n; ++i) { fi[i] = i; se[i] = i + 1; }
    if (int rc = run_frames(e, f.data(), n + 1, w, h, stride, fi.data(), se.data(), n,
                            flow_out + (size_t)base * 2 * w * h)) return rc;
    add_stats(&acc, e->stats);
  }
  if (n_pairs > kMaxChunkPairs) e->stats = acc;
  return DIS_OK;
}

int dis_run_pyramids(dis_engine* e, const float* const* img_first, const float* const* img_first_dx,
                     const float* const* img_first_dy, const float* const* img_second,
                     const float* const* img_second_dx, const float* const* img_second_dy, int img_padding,
                     float* outflow, int width, int height) {
  (void)img_first; (void)img_second_dx; (void)img_second_dy;   // never read on this path
  if (!e || !img_first_dx || !img_first_dy || !img_second || !outflow) return fail(DIS_ERR_PARAM, "null argument");
  CU(cudaSetDevice(e->device));
  e->w0 = width; e->h0 = height; e->left = 0; e->top = 0;
  if (int rc = setup_run(e, width, height, img_padding, 2, 1)) return rc;
  e->tiled_pr = make_int2(0, 1);
  CU(cudaMemcpyAsync(e->pairs.p, &e->tiled_pr, sizeof e->tiled_pr, cudaMemcpyHostToDevice, e->stream));
  for (const Level& L : e->levels) {
    const int s = L.info.scale;
    if (!img_first_dx[s] || !img_first_dy[s] || !img_second[s]) return fail(DIS_ERR_PARAM, "null pyramid plane");
    const size_t n = (size_t)L.info.tmp_w * L.info.tmp_h, bytes = sizeof(float) * n;
    const dis::LevelPlanes& P = e->planes[s];
    if (e->stage.ensure(2 * bytes)) return fail(DIS_ERR_ALLOC, "cudaMalloc(stage) failed");
    float* st = e->stage.as<float>();   // planar dx, dy of the reference's API -> the engine's float2 plane
    CU(cudaMemcpyAsync(st, img_first_dx[s], bytes, cudaMemcpyHostToDevice, e->stream));
    CU(cudaMemcpyAsync(st + n, img_first_dy[s], bytes, cudaMemcpyHostToDevice, e->stream));
    dis::launch_interleave(st, st + n, P.base + P.plane_stride, n, e->stream);
    CU(cudaMemcpyAsync(P.base + P.frame_stride, img_second[s], bytes, cudaMemcpyHostToDevice, e->stream));
  }
  e->launches = 0;
  if (int rc = run_levels(e)) return rc;
  const Level& F = e->levels[0];
  CU(cudaMemcpyAsync(outflow, F.flow.p, sizeof(float) * F.flow_stride, cudaMemcpyDeviceToHost, e->stream));
  return fetch_counters(e);
}

int dis_construct_pyramid(dis_engine* e, const float* img, int width, int height, int lv_f, int img_padding,
                          float* const* img_pyr, float* const* img_dx_pyr, float* const* img_dy_pyr) {
  if (!e || !img || !img_pyr || !img_dx_pyr || !img_dy_pyr) return fail(DIS_ERR_PARAM, "null argument");
  if (lv_f < 0 || lv_f > 15) return fail(DIS_ERR_PARAM, "lv_f out of range");
  CU(cudaSetDevice(e->device));
  const dis_params saved = e->prm;
  e->prm.coarsest_scale = lv_f; e->prm.finest_scale = 0;
  e->w0 = width; e->h0 = height; e->left = 0; e->top = 0;
  int rc = setup_run(e, width, height, img_padding, 1, 1);
  if (rc == DIS_OK && e->frames_f32.ensure(sizeof(float) * (size_t)width * height)) rc = fail(DIS_ERR_ALLOC, "cudaMalloc failed");
  if (rc != DIS_OK) { e->prm = saved; e->levels.clear(); return rc; }
  auto body = [&]() -> int {
    CU(cudaMemcpyAsync(e->frames_f32.p, img, sizeof(float) * (size_t)width * height, cudaMemcpyHostToDevice, e->stream));
    dis::PyramidArgs a;
    memset(&a, 0, sizeof a);
    a.frames = nullptr; a.frames_f32 = e->frames_f32.as<float>();
    a.W = width; a.H = height; a.n_frames = 1; a.levels = lv_f + 1; a.pad = img_padding;
    for (int s = 0; s <= lv_f; ++s) a.planes[s] = e->planes[s];
    e->launches = dis::launch_pyramid(a, e->stream);
    CU(cudaGetLastError());
    for (int s = 0; s <= lv_f; ++s) {
      const size_t n = (size_t)((width >> s) + 2 * img_padding) * ((height >> s) + 2 * img_padding), bytes = sizeof(float) * n;
      const dis::LevelPlanes& P = e->planes[s];
      if (img_pyr[s]) CU(cudaMemcpyAsync(img_pyr[s], P.base, bytes, cudaMemcpyDeviceToHost, e->stream));
      if (img_dx_pyr[s] || img_dy_pyr[s]) {
        if (e->stage.ensure(2 * bytes)) return fail(DIS_ERR_ALLOC, "cudaMalloc(stage) failed");
        float* st = e->stage.as<float>();
        dis::launch_deinterleave(P.base + P.plane_stride, st, st + n, n, e->stream);
        if (img_dx_pyr[s]) CU(cudaMemcpyAsync(img_dx_pyr[s], st, bytes, cudaMemcpyDeviceToHost, e->stream));
        if (img_dy_pyr[s]) CU(cudaMemcpyAsync(img_dy_pyr[s], st + n, bytes, cudaMemcpyDeviceToHost, e->stream));
      }
    }
    CU(cudaStreamSynchronize(e->stream));
    return DIS_OK;
  };
  rc = body();
  e->prm = saved;
  e->levels.clear();       // geometry no longer matches the engine's own parameters
  e->have_run = false;
  return rc;
}

// ---- introspection ---------------------------------------------------------------------------
int dis_num_levels(dis_engine* e) { return e ? (int)e->levels.size() : 0; }

int dis_get_level_info(dis_engine* e, int level, dis_level_info* out) {
  if (!e || !out) return fail(DIS_ERR_PARAM, "null argument");
  if (level < 0 || level >= (int)e->levels.size()) return fail(DIS_ERR_PARAM, "level out of range");
  *out = e->levels[level].info;
  return DIS_OK;
}

int dis_get_patches(dis_engine* e, int pair, int level, float* ref_xy, float* u_xy, int* iters) {
  if (!e) return fail(DIS_ERR_PARAM, "null engine");
  if (!e->have_run) return fail(DIS_ERR_STATE, "no run yet");
  if (level < 0 || level >= (int)e->levels.size() || pair < 0 || pair >= e->n_pairs) return fail(DIS_ERR_PARAM, "index out of range");
  const Level& L = e->levels[level];
  const dis_level_info& I = L.info;
  CU(cudaSetDevice(e->device));
  if (ref_xy) {   // src/patch_grid.cpp:39-49
    int id = 0;
    for (int x = 0; x < I.npw; ++x)
      for (int y = 0; y < I.nph; ++y, ++id) {
        ref_xy[2 * id] = (float)(x * I.steps + I.offw);
        ref_xy[2 * id + 1] = (float)(y * I.steps + I.offh);
      }
  }
  if (u_xy) CU(cudaMemcpyAsync(u_xy, L.u.as<float>() + 2 * (size_t)I.num_patches * pair, sizeof(float) * 2 * I.num_patches, cudaMemcpyDeviceToHost, e->stream));
  if (iters) {
    if (!e->keep_iters) return fail(DIS_ERR_STATE, "iteration counts were not kept");
    CU(cudaMemcpyAsync(iters, L.iters.as<int>() + (size_t)I.num_patches * pair, sizeof(int) * I.num_patches, cudaMemcpyDeviceToHost, e->stream));
  }
  CU(cudaStreamSynchronize(e->stream));
  return DIS_OK;
}

int dis_get_level_flow(dis_engine* e, int pair, int level, float* flow) {
  if (!e || !flow) return fail(DIS_ERR_PARAM, "null argument");
  if (!e->have_run) return fail(DIS_ERR_STATE, "no run yet");
  if (level < 0 || level >= (int)e->levels.size() || pair < 0 || pair >= e->n_pairs) return fail(DIS_ERR_PARAM, "index out of range");
  const Level& L = e->levels[level];
  CU(cudaSetDevice(e->device));
  CU(cudaMemcpyAsync(flow, L.flow.as<float>() + L.flow_stride * pair, sizeof(float) * L.flow_stride, cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return DIS_OK;
}

int dis_get_plane(dis_engine* e, int frame, int scale, int which, float* out) {
  if (!e || !out) return fail(DIS_ERR_PARAM, "null argument");
  if (!e->have_run) return fail(DIS_ERR_STATE, "no run yet");
  if (frame < 0 || frame >= e->n_frames || scale < 0 || scale > e->prm.coarsest_scale || which < 0 || which > 2)
    return fail(DIS_ERR_PARAM, "index out of range");
  const dis::LevelPlanes& P = e->planes[scale];
  const size_t n = (size_t)((e->W >> scale) + 2 * e->pad) * ((e->H >> scale) + 2 * e->pad);
  CU(cudaSetDevice(e->device));
  const float* src = P.base + P.frame_stride * frame;
  if (which > 0) {
    if (e->stage.ensure(2 * sizeof(float) * n)) return fail(DIS_ERR_ALLOC, "cudaMalloc(stage) failed");
    float* st = e->stage.as<float>();
    dis::launch_deinterleave(src + P.plane_stride, st, st + n, n, e->stream);
    src = st + (which - 1) * n;
  }
  CU(cudaMemcpyAsync(out, src, sizeof(float) * n, cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return DIS_OK;
}

int dis_get_stats(dis_engine* e, dis_stats* out) {
  if (!e || !out) return fail(DIS_ERR_PARAM, "null argument");
  if (!e->have_run) return fail(DIS_ERR_STATE, "no run yet");
  CU(cudaSetDevice(e->device));
  if (!e->stats_fetched)
    if (int rc = fetch_counters(e)) return rc;
  *out = e->stats;
  return DIS_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// PatchGrid-level API (include/patch_grid.hpp:16-30): one level of one pair, host pointers
// ---------------------------------------------------------------------------------------------
struct dis_grid {
  dis_engine* e;
  dis_level_info info;
  int pad;
  float lb, ub_w, ub_h;
  size_t plane_elems;
  DevBuf planes, flow_prev, init, u, dense, pair, stage, walk;
  std::vector<float> init_host;
  bool have_first = false, have_second = false, use_prev = false, use_override = false, searched = false;
  std::vector<float> u_host;
};

extern "C" {

static int grid_create(dis_engine* e, int width, int height, int img_padding, bool single, int ref_x, int ref_y, dis_grid** out) {
  if (!e || !out) return fail(DIS_ERR_PARAM, "null argument");
  *out = nullptr;
  if (width < 2 || height < 2) return fail(DIS_ERR_PARAM, "level too small");
  if (img_padding < e->prm.patch_size) return fail(DIS_ERR_PARAM, "img_padding must be >= patch_size");
  CU(cudaSetDevice(e->device));
  dis_grid* g = new dis_grid();
  g->e = e; g->pad = img_padding;
  dis_level_info& I = g->info;
  const int ps = e->prm.patch_size;
  I.scale = 0; I.width = width; I.height = height;
  I.tmp_w = width + 2 * img_padding; I.tmp_h = height + 2 * img_padding;
  I.steps = e->steps;
  I.npw = (int)ceilf((float)width / (float)I.steps);
  I.nph = (int)ceilf((float)height / (float)I.steps);
  I.offw = (width - (I.npw - 1) * I.steps) / 2;
  I.offh = (height - (I.nph - 1) * I.steps) / 2;
  if (single) { I.npw = 1; I.nph = 1; I.offw = ref_x; I.offh = ref_y; }
  I.num_patches = I.npw * I.nph;
  g->lb = -(float)ps / 2; g->ub_w = (float)(width + ps / 2 - 2); g->ub_h = (float)(height + ps / 2 - 2);
  g->plane_elems = ((size_t)I.tmp_w * I.tmp_h + 3) & ~(size_t)3;
  const int2 pr = make_int2(0, 1);
  if (g->planes.ensure(sizeof(float) * 6 * g->plane_elems) || g->stage.ensure(sizeof(float) * 2 * g->plane_elems) ||
      g->u.ensure(sizeof(float) * 2 * I.num_patches) ||
      g->dense.ensure(sizeof(float) * 2 * (size_t)width * height) || g->pair.ensure(64) ||
      g->init.ensure(sizeof(float) * 2 * I.num_patches) ||
      g->walk.ensure(dis::kWalkBytesPerPatch * (size_t)I.num_patches) ||
      g->flow_prev.ensure(sizeof(float) * 2 * (size_t)(width / 2 + 1) * (height / 2 + 1))) {
    dis_grid_destroy(g);
    return fail(DIS_ERR_ALLOC, "cudaMalloc(grid) failed");
  }
  cudaMemcpy(g->pair.p, &pr, sizeof pr, cudaMemcpyHostToDevice);
  g->init_host.assign(2 * (size_t)I.num_patches, 0.f);
  *out = g;
  return DIS_OK;
}

int dis_grid_create(dis_engine* e, int width, int height, int img_padding, dis_grid** out) {
  return grid_create(e, width, height, img_padding, false, 0, 0, out);
}

int dis_grid_create_single(dis_engine* e, int width, int height, int img_padding, float ref_x, float ref_y, dis_grid** out) {
  if (!(ref_x == floorf(ref_x) && ref_y == floorf(ref_y))) return fail(DIS_ERR_PARAM, "patch reference point must be integer valued");
  const int half = e ? e->prm.patch_size / 2 : 0;
  // the template gather (src/patch.cpp:47-73) must stay inside the padded plane
  if (e && (ref_x < (float)(half - img_padding) || ref_y < (float)(half - img_padding) ||
            ref_x > (float)(width + img_padding - half) || ref_y > (float)(height + img_padding - half)))
    return fail(DIS_ERR_PARAM, "patch reference point outside the padded image");
  return grid_create(e, width, height, img_padding, true, (int)ref_x, (int)ref_y, out);
}

void dis_grid_destroy(dis_grid* g) {
  if (!g) return;
  cudaSetDevice(g->e->device);
  cudaStreamSynchronize(g->e->stream);
  g->planes.release(); g->flow_prev.release(); g->init.release(); g->u.release(); g->dense.release(); g->pair.release();
  g->stage.release(); g->walk.release();
  delete g;
}

int dis_grid_init(dis_grid* g, const float* img_first, const float* dx, const float* dy) {
  (void)img_first;   // gathered by the reference but never used
  if (!g || !dx || !dy) return fail(DIS_ERR_PARAM, "null argument");
  CU(cudaSetDevice(g->e->device));
  const size_t n = (size_t)g->info.tmp_w * g->info.tmp_h, bytes = sizeof(float) * n;
  float* base = g->planes.as<float>();
  float* st = g->stage.as<float>();
  CU(cudaMemcpyAsync(st, dx, bytes, cudaMemcpyHostToDevice, g->e->stream));
  CU(cudaMemcpyAsync(st + n, dy, bytes, cudaMemcpyHostToDevice, g->e->stream));
  dis::launch_interleave(st, st + n, base + g->plane_elems, n, g->e->stream);
  CU(cudaStreamSynchronize(g->e->stream));
  std::fill(g->init_host.begin(), g->init_host.end(), 0.f);   // src/patch_grid.cpp:80
  g->use_prev = g->use_override = false;
  g->have_first = true; g->searched = false;
  return DIS_OK;
}

int dis_grid_set_target(dis_grid* g, const float* img_second, const float* dx, const float* dy) {
  (void)dx; (void)dy;   // never read
  if (!g || !img_second) return fail(DIS_ERR_PARAM, "null argument");
  CU(cudaSetDevice(g->e->device));
  const size_t bytes = sizeof(float) * (size_t)g->info.tmp_w * g->info.tmp_h;
  CU(cudaMemcpyAsync(g->planes.as<float>() + 3 * g->plane_elems, img_second, bytes, cudaMemcpyHostToDevice, g->e->stream));
  CU(cudaStreamSynchronize(g->e->stream));
  g->have_second = true; g->searched = false;
  return DIS_OK;
}

int dis_grid_init_from_prev_flow(dis_grid* g, const float* flow_prev) {
  if (!g || !flow_prev) return fail(DIS_ERR_PARAM, "null argument");
  CU(cudaSetDevice(g->e->device));
  const size_t n = 2 * (size_t)(g->info.width / 2) * (g->info.height / 2);
  CU(cudaMemcpyAsync(g->flow_prev.p, flow_prev, sizeof(float) * n, cudaMemcpyHostToDevice, g->e->stream));
  CU(cudaStreamSynchronize(g->e->stream));
  g->use_prev = true; g->use_override = false;
  return DIS_OK;
}

int dis_grid_set_patch_init(dis_grid* g, int i, float ux, float uy) {
  if (!g) return fail(DIS_ERR_PARAM, "null grid");
  if (i < 0 || i >= g->info.num_patches) return fail(DIS_ERR_PARAM, "patch index out of range");
  if (g->use_prev) return fail(DIS_ERR_STATE, "explicit inits cannot be mixed with init_from_prev_flow");
  g->init_host[2 * i] = ux; g->init_host[2 * i + 1] = uy;
  g->use_override = true;
  return DIS_OK;
}

int dis_grid_inverse_search(dis_grid* g) {
  if (!g) return fail(DIS_ERR_PARAM, "null grid");
  if (!g->have_first || !g->have_second) return fail(DIS_ERR_STATE, "init_grid and set_target_image first");
  dis_engine* e = g->e;
  CU(cudaSetDevice(e->device));
  const dis_level_info& I = g->info;
  if (g->use_override)
    CU(cudaMemcpyAsync(g->init.p, g->init_host.data(), sizeof(float) * 2 * I.num_patches, cudaMemcpyHostToDevice, e->stream));
  dis::SearchArgs a;
  memset(&a, 0, sizeof a);
  a.planes.base = g->planes.as<float>(); a.planes.plane_stride = g->plane_elems; a.planes.frame_stride = 3 * g->plane_elems;
  a.pairs = g->pair.as<int2>();
  a.work_counter = g->pair.as<int>() + 8;
  a.flow_prev = g->use_prev ? g->flow_prev.as<float>() : nullptr;
  a.init_override = g->use_override ? g->init.as<float>() : nullptr;
  a.u_out = g->u.as<float>();
  a.walk_state = g->walk.p;
  a.walk_seeds = g->walk.as<char>() + 32 * (size_t)I.num_patches;
  a.n_pairs = 1;
  a.width = I.width; a.height = I.height; a.pad = g->pad; a.tmp_w = I.tmp_w;
  a.npw = I.npw; a.nph = I.nph; a.offw = I.offw; a.offh = I.offh; a.steps = I.steps; a.np = I.num_patches;
  a.warps_per_pair = (I.num_patches + 31) / 32;
  a.iterations = e->prm.max_iter; a.patnorm = e->prm.patch_norm ? 1 : 0;
  a.lb = g->lb; a.ub_w = g->ub_w; a.ub_h = g->ub_h; a.out2max = sq_threshold(e->outlier); a.tile_margin = (int)ceilf(e->outlier) + 1;
  a.conv_eps2 = e->prm.conv_eps * e->prm.conv_eps;
  if (dis::launch_inverse_search(a, e->prm.patch_size, e->stream) < 0) return fail(DIS_ERR_PARAM, "unsupported patch_size");
  CU(cudaGetLastError());
  g->u_host.resize(2 * (size_t)I.num_patches);
  CU(cudaMemcpyAsync(g->u_host.data(), g->u.p, sizeof(float) * 2 * I.num_patches, cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  g->searched = true;
  return DIS_OK;
}

int dis_grid_densify(dis_grid* g, float* dense_flow) {
  if (!g || !dense_flow) return fail(DIS_ERR_PARAM, "null argument");
  if (!g->searched) return fail(DIS_ERR_STATE, "inverse_search first");
  dis_engine* e = g->e;
  CU(cudaSetDevice(e->device));
  const dis_level_info& I = g->info;
  dis::DensifyArgs d;
  d.u = g->u.as<float>(); d.dense = g->dense.as<float>(); d.dense_stride = 2 * (size_t)I.width * I.height;
  d.n_pairs = 1; d.width = I.width; d.height = I.height; d.npw = I.npw; d.nph = I.nph;
  d.offw = I.offw; d.offh = I.offh; d.steps = I.steps; d.np = I.num_patches; d.ps = e->prm.patch_size;
  dis::launch_densify(d, e->stream);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(dense_flow, g->dense.p, sizeof(float) * d.dense_stride, cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return DIS_OK;
}

int dis_grid_num_patches(dis_grid* g) { return g ? g->info.num_patches : 0; }

int dis_grid_get_patch(dis_grid* g, int i, float* ref_xy, float* u_xy) {
  if (!g) return fail(DIS_ERR_PARAM, "null grid");
  if (i < 0 || i >= g->info.num_patches) return fail(DIS_ERR_PARAM, "patch index out of range");
  const dis_level_info& I = g->info;
  if (ref_xy) {
    ref_xy[0] = (float)((i / I.nph) * I.steps + I.offw);
    ref_xy[1] = (float)((i % I.nph) * I.steps + I.offh);
  }
  if (u_xy) {
    if (g->searched) { u_xy[0] = g->u_host[2 * i]; u_xy[1] = g->u_host[2 * i + 1]; }
    else { u_xy[0] = 0.f; u_xy[1] = 0.f; }   // Patch::reset_patch, src/patch.cpp:102-116
  }
  return DIS_OK;
}

// ---- one large frame tiled by grid rows over several GPUs (BASELINE.json config 5) ---------
// Within a level the patches are independent and densification is a local gather, so every GPU searches the
// patches of its band of grid rows, the per-patch flows u are exchanged (a few MB per level; the caller does it,
// e.g. an NCCL all-gather over NVLink), and every GPU densifies the whole level redundantly (< 1 % of the work).
// The result is bit-identical to the single-GPU run.
int dis_tiled_begin(dis_engine* e, const uint8_t* a, const uint8_t* b, int w, int h, int stride) {
  if (!e || !a || !b) return fail(DIS_ERR_PARAM, "null argument");
  if (w <= 0 || h <= 0 || stride < w) return fail(DIS_ERR_PARAM, "bad frame geometry");
  CU(cudaSetDevice(e->device));
  const int padw = div_pad(w, e->prm.coarsest_scale), padh = div_pad(h, e->prm.coarsest_scale);
  e->w0 = w; e->h0 = h; e->left = padw / 2; e->top = padh / 2;
  if (int rc = setup_run(e, w + padw, h + padh, e->prm.patch_size, 2, 1)) return rc;
  const size_t fbytes = (size_t)w * h;
  if (e->frames_u8.ensure(fbytes * 2)) return fail(DIS_ERR_ALLOC, "cudaMalloc(frames) failed");
  CU(cudaMemcpy2DAsync(e->frames_u8.as<uint8_t>(), w, a, stride, w, h, cudaMemcpyHostToDevice, e->stream));
  CU(cudaMemcpy2DAsync(e->frames_u8.as<uint8_t>() + fbytes, w, b, stride, w, h, cudaMemcpyHostToDevice, e->stream));
  e->tiled_pr = make_int2(0, 1);
  CU(cudaMemcpyAsync(e->pairs.p, &e->tiled_pr, sizeof e->tiled_pr, cudaMemcpyHostToDevice, e->stream));
  e->stride_u8 = w;
  e->launches = 0;
  if (!e->tiled_t0) { CU(cudaEventCreate(&e->tiled_t0)); CU(cudaEventCreate(&e->tiled_t1)); }
  CU(cudaEventRecord(e->tiled_t0, e->stream));
  CU(cudaEventRecord(e->tiled_t1, e->stream));
  if (int rc = run_pyramid_u8(e, w)) return rc;
  CU(cudaMemsetAsync(e->counters.p, 0, dis::kCounters * sizeof(unsigned long long), e->stream));
  for (Level& L : e->levels)   // patches outside this GPU's bands are filled in by dis_tiled_set_u
    CU(cudaMemsetAsync(L.u.p, 0, 2 * sizeof(float) * (size_t)L.info.num_patches, e->stream));
  const int nl = (int)e->levels.size();
  while ((int)e->search_ev.size() < 4 * nl) { cudaEvent_t ev; CU(cudaEventCreate(&ev)); e->search_ev.push_back(ev); }
  while ((int)e->tiled_ev.size() < nl) { cudaEvent_t ev; CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)); e->tiled_ev.push_back(ev); }
  e->n_groups = 1; e->search_launches = nl;
  e->tiled_patches = 0;
  e->have_run = false; e->tiled = true;
  return DIS_OK;
}

int dis_tiled_search(dis_engine* e, int level, int row0, int row1) {
  if (!e) return fail(DIS_ERR_PARAM, "null engine");
  if (!e->tiled) return fail(DIS_ERR_STATE, "dis_tiled_begin first");
  const int nl = (int)e->levels.size();
  if (level < 0 || level >= nl) return fail(DIS_ERR_PARAM, "level out of range");
  const dis_level_info& I = e->levels[level].info;
  if (row0 < 0 || row1 > I.nph || row0 > row1) return fail(DIS_ERR_PARAM, "grid rows out of range");
  CU(cudaSetDevice(e->device));
  if (row0 == row1) {                // an empty band (more GPUs than grid rows)
    CU(cudaEventRecord(e->tiled_ev[level], e->stream));
    return DIS_OK;
  }
  dis::SearchArgs a = level_search_args(e, 0, level, 0, 1);
  a.band_y0 = row0; a.band_h = row1 - row0;
  e->tiled_patches += (unsigned long long)I.npw * (row1 - row0);
  cudaEvent_t* ev = &e->search_ev[4 * (nl - 1 - level)];
  const int n = dis::launch_inverse_search(a, e->prm.patch_size, e->stream, ev, e->side.stream ? &e->side : nullptr);
  if (n < 0) return fail(DIS_ERR_PARAM, "unsupported patch_size");
  e->launches += n;
  CU(cudaGetLastError());
  CU(cudaEventRecord(e->tiled_ev[level], e->stream));
  return DIS_OK;
}

// ---- halo exchange of a row-banded run, device to device (src/patch_grid.cpp:108-119, 121-182 define
//      what a band needs: the covering patches of the pixel rows it densifies, and the coarser-flow rows its patches
//      are initialised from).  All calls are asynchronous: e's stream waits for src's search of the level by event.
int dis_tiled_fetch_u_rows(dis_engine* e, int level, dis_engine* src, int row0, int row1) {
  if (!e || !src) return fail(DIS_ERR_PARAM, "null engine");
  if (!e->tiled || !src->tiled) return fail(DIS_ERR_STATE, "dis_tiled_begin first (both engines)");
  const int nl = (int)e->levels.size();
  if (level < 0 || level >= nl || (int)src->levels.size() != nl) return fail(DIS_ERR_PARAM, "level out of range");
  const dis_level_info& I = e->levels[level].info;
  if (src->levels[level].info.num_patches != I.num_patches) return fail(DIS_ERR_PARAM, "engines run different geometries");
  if (row0 < 0 || row1 > I.nph || row0 > row1) return fail(DIS_ERR_PARAM, "grid rows out of range");
  if (row0 == row1 || e == src) return DIS_OK;
  CU(cudaSetDevice(e->device));
  CU(cudaStreamWaitEvent(e->stream, src->tiled_ev[level], 0));
  e->launches += dis::launch_copy_u_rows(src->levels[level].u.as<float>(), e->levels[level].u.as<float>(), I.npw, I.nph,
                                         row0, row1 - row0, e->stream);
  CU(cudaGetLastError());
  return DIS_OK;
}

int dis_tiled_densify_rows(dis_engine* e, int level, int y0, int y1) {
  if (!e) return fail(DIS_ERR_PARAM, "null engine");
  if (!e->tiled) return fail(DIS_ERR_STATE, "dis_tiled_begin first");
  if (level < 0 || level >= (int)e->levels.size()) return fail(DIS_ERR_PARAM, "level out of range");
  const dis_level_info& I = e->levels[level].info;
  if (y0 < 0 || y1 > I.height || y0 > y1) return fail(DIS_ERR_PARAM, "pixel rows out of range");
  if (y0 == y1) return DIS_OK;
  CU(cudaSetDevice(e->device));
  level_densify(e, level, 0, 1, e->stream, y0, y1);
  CU(cudaGetLastError());
  return DIS_OK;
}

// crop + download of output rows [y0, y1) only (rows of the un-padded frame); asynchronous — dis_tiled_end waits
int dis_tiled_finish_rows(dis_engine* e, float* flow_out, int y0, int y1) {
  if (!e || !flow_out) return fail(DIS_ERR_PARAM, "null argument");
  if (!e->tiled) return fail(DIS_ERR_STATE, "dis_tiled_begin first");
  if (e->prm.finest_scale != 0) return fail(DIS_ERR_PARAM, "row-banded download needs finest_scale == 0");
  if (y0 < 0 || y1 > e->h0 || y0 > y1) return fail(DIS_ERR_PARAM, "output rows out of range");
  CU(cudaSetDevice(e->device));
  CU(cudaEventRecord(e->tiled_t1, e->stream));
  if (y0 == y1) return DIS_OK;
  Level& L = e->levels[0];
  const size_t row = 2 * (size_t)e->w0;
  if (e->out.ensure(sizeof(float) * row * e->h0)) return fail(DIS_ERR_ALLOC, "cudaMalloc(out) failed");
  e->launches += dis::launch_crop(L.flow.as<float>(), L.flow_stride, L.info.width, L.info.height, e->left, e->top + y0,
                                  e->out.as<float>() + row * y0, e->w0, y1 - y0, 1, e->stream);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(flow_out + row * y0, e->out.as<float>() + row * y0, sizeof(float) * row * (y1 - y0), cudaMemcpyDeviceToHost, e->stream));
  return DIS_OK;
}

int dis_tiled_end(dis_engine* e) {
  if (!e) return fail(DIS_ERR_PARAM, "null engine");
  if (!e->tiled) return fail(DIS_ERR_STATE, "dis_tiled_begin first");
  CU(cudaSetDevice(e->device));
  e->stats.patches = e->tiled_patches;   // this engine's bands; the counters below are its share too
  e->have_run = true; e->stats_fetched = false; e->tiled = false;
  const int rc = fetch_counters(e);      // synchronises the stream
  e->tiled_ms = 0.f;
  if (!rc && e->tiled_t0) { if (cudaEventElapsedTime(&e->tiled_ms, e->tiled_t0, e->tiled_t1) != cudaSuccess) { cudaGetLastError(); e->tiled_ms = 0.f; } }
  return rc;
}

int dis_tiled_last_ms(dis_engine* e, float* ms) {
  if (!e || !ms) return fail(DIS_ERR_PARAM, "null argument");
  *ms = e->tiled_ms;
  return DIS_OK;
}

static int tiled_copy_u(dis_engine* e, int level, void* dst, const void* src, bool get) {
  if (!e || (get ? !dst : !src)) return fail(DIS_ERR_PARAM, "null argument");
  if (!e->tiled) return fail(DIS_ERR_STATE, "dis_tiled_begin first");
  if (level < 0 || level >= (int)e->levels.size()) return fail(DIS_ERR_PARAM, "level out of range");
  CU(cudaSetDevice(e->device));
  Level& L = e->levels[level];
  const size_t bytes = 2 * sizeof(float) * (size_t)L.info.num_patches;
  CU(cudaMemcpyAsync(get ? dst : L.u.p, get ? L.u.p : src, bytes, cudaMemcpyDeviceToDevice, e->stream));
  CU(cudaStreamSynchronize(e->stream));   // the caller's buffers live on other streams (torch / NCCL)
  return DIS_OK;
}
int dis_tiled_get_u(dis_engine* e, int level, void* device_dst) { return tiled_copy_u(e, level, device_dst, nullptr, true); }
int dis_tiled_set_u(dis_engine* e, int level, const void* device_src) { return tiled_copy_u(e, level, nullptr, device_src, false); }

int dis_tiled_densify(dis_engine* e, int level) {
  if (!e) return fail(DIS_ERR_PARAM, "null engine");
  if (!e->tiled) return fail(DIS_ERR_STATE, "dis_tiled_begin first");
  if (level < 0 || level >= (int)e->levels.size()) return fail(DIS_ERR_PARAM, "level out of range");
  CU(cudaSetDevice(e->device));
  level_densify(e, level, 0, 1, e->stream);
  CU(cudaGetLastError());
  return DIS_OK;
}

int dis_tiled_finish(dis_engine* e, float* flow_out) {
  if (!e || !flow_out) return fail(DIS_ERR_PARAM, "null argument");
  if (!e->tiled) return fail(DIS_ERR_STATE, "dis_tiled_begin first");
  CU(cudaSetDevice(e->device));
  if (int rc = run_epilogue(e)) return rc;
  CU(cudaMemcpyAsync(flow_out, e->out.p, sizeof(float) * 2 * (size_t)e->w0 * e->h0, cudaMemcpyDeviceToHost, e->stream));
  e->stats.patches = 0;
  for (const Level& L : e->levels) e->stats.patches += (unsigned long long)L.info.num_patches;
  e->have_run = true; e->stats_fetched = false; e->tiled = false;
  return fetch_counters(e);
}

// ---- colour coding (src/color_coding.cpp:83-117) -------------------------------------------------
static int color_device(dis_engine* e, const float* dflow, size_t npx, float maxmotion, uint8_t* bgr_out) {
  const size_t off = (2 * sizeof(float) * npx + 255) & ~(size_t)255;   // stage: [flow | scratch | bgr]
  if (e->stage.ensure(off + 256 + 3 * npx)) return fail(DIS_ERR_ALLOC, "cudaMalloc(colour stage) failed");
  char* st = e->stage.as<char>();
  e->launches += dis::launch_color_code(dflow, npx, maxmotion, reinterpret_cast<unsigned*>(st + off),
                                        reinterpret_cast<uint8_t*>(st + off + 256), e->stream);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(bgr_out, st + off + 256, 3 * npx, cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return DIS_OK;
}

int dis_color_code(dis_engine* e, const float* flow, int w, int h, float maxmotion, uint8_t* bgr_out) {
  if (!e || !flow || !bgr_out) return fail(DIS_ERR_PARAM, "null argument");
  if (w <= 0 || h <= 0) return fail(DIS_ERR_PARAM, "empty flow field");
  CU(cudaSetDevice(e->device));
  const size_t npx = (size_t)w * h;
  const size_t off = (2 * sizeof(float) * npx + 255) & ~(size_t)255;
  if (e->stage.ensure(off + 256 + 3 * npx)) return fail(DIS_ERR_ALLOC, "cudaMalloc(colour stage) failed");
  CU(cudaMemcpyAsync(e->stage.p, flow, 2 * sizeof(float) * npx, cudaMemcpyHostToDevice, e->stream));
  return color_device(e, e->stage.as<float>(), npx, maxmotion, bgr_out);
}

int dis_color_code_pair(dis_engine* e, int pair, float maxmotion, uint8_t* bgr_out) {
  if (!e || !bgr_out) return fail(DIS_ERR_PARAM, "null argument");
  if (!e->have_run || !e->out.p) return fail(DIS_ERR_STATE, "no run yet");
  if (pair < 0 || pair >= e->n_pairs) return fail(DIS_ERR_PARAM, "pair index out of range");
  CU(cudaSetDevice(e->device));
  const size_t npx = (size_t)e->w0 * e->h0;
  return color_device(e, e->out.as<float>() + 2 * npx * pair, npx, maxmotion, bgr_out);
}

// ---- Middlebury .flo (src/IO_flow.cpp:10-98) ----------------------------------------------------
int dis_write_flo(const char* path, const float* flow, int w, int h) {
  if (!path || !flow || w <= 0 || h <= 0) return fail(DIS_ERR_PARAM, "bad argument");
  FILE* f = fopen(path, "wb");
  if (!f) return fail(DIS_ERR_PARAM, std::string("cannot open ") + path);
  const int32_t wh[2] = {w, h};
  bool ok = fwrite("PIEH", 1, 4, f) == 4 && fwrite(wh, sizeof(int32_t), 2, f) == 2 &&
            fwrite(flow, sizeof(float), 2 * (size_t)w * h, f) == 2 * (size_t)w * h;
  fclose(f);
  return ok ? DIS_OK : fail(DIS_ERR_PARAM, "short write");
}

int dis_read_flo_size(const char* path, int* w, int* h) {
  if (!path || !w || !h) return fail(DIS_ERR_PARAM, "bad argument");
  FILE* f = fopen(path, "rb");
  if (!f) return fail(DIS_ERR_PARAM, std::string("cannot open ") + path);
  char tag[4]; int32_t wh[2];
  bool ok = fread(tag, 1, 4, f) == 4 && fread(wh, sizeof(int32_t), 2, f) == 2 && memcmp(tag, "PIEH", 4) == 0;
  fclose(f);
  if (!ok) return fail(DIS_ERR_PARAM, "not a .flo file");
  *w = wh[0]; *h = wh[1];
  return DIS_OK;
}

int dis_read_flo(const char* path, float* flow, int w, int h) {
  int fw = 0, fh = 0;
  if (int rc = dis_read_flo_size(path, &fw, &fh)) return rc;
  if (!flow || fw != w || fh != h) return fail(DIS_ERR_PARAM, "size mismatch");
  FILE* f = fopen(path, "rb");
  if (!f) return fail(DIS_ERR_PARAM, std::string("cannot open ") + path);
  fseek(f, 12, SEEK_SET);
  const size_t n = 2 * (size_t)w * h;
  const bool ok = fread(flow, sizeof(float), n, f) == n;
  fclose(f);
  return ok ? DIS_OK : fail(DIS_ERR_PARAM, "file too short");
}

}  // extern "C"

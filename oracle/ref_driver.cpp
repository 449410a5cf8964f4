// oracle/ref_driver.cpp — drives the UNMODIFIED reference engine. TEST INFRASTRUCTURE ONLY.
//
// Built by oracle/Makefile together with /root/reference/src/{patch,patch_grid,optical_flow}.cpp
// (compiled where they lie; nothing is copied) against the stub headers in oracle/stubs/ into
// oracle/_ref/libdis_ref.so.  Used to (1) validate oracle/dis_oracle.c, (2) generate the
// fixtures in tests/golden/, (3) serve as the "reference" CPU baseline of bench.py.
//
// The reference allocates its densification `weight` array with new[] and never clears it
// (src/patch_grid.cpp:123-126).  Per SURVEY.md App. C-1 the oracle defines that
// as zero-initialised: this TU replaces global operator new[] with a zero-filling one (the
// library is linked -Bsymbolic so the replacement binds inside the .so only).
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <new>
#include <vector>

#include "optical_flow.hpp"
#include "patch_grid.hpp"

void* operator new[](std::size_t n) {
  void* p = std::calloc(n ? n : 1, 1);
  if (!p) throw std::bad_alloc();
  return p;
}
void operator delete[](void* p) noexcept { std::free(p); }
void operator delete[](void* p, std::size_t) noexcept { std::free(p); }

namespace {
struct QuietCout {   // the ctor prints "scale: N" per level (src/optical_flow.cpp:69)
  std::streambuf* old;
  QuietCout() : old(std::cout.rdbuf(nullptr)) {}
  ~QuietCout() { std::cout.rdbuf(old); std::cout.clear(); }
};
}  // namespace

extern "C" {

// 1 when array-new hands out zeroed memory inside this library.
__attribute__((visibility("default"))) int ref_new_is_zeroing() {
  for (int rep = 0; rep < 8; ++rep) {
    float* a = new float[4096];
    for (int i = 0; i < 4096; ++i) if (a[i] != 0.0f) { delete[] a; return 0; }
    std::memset(a, 0xff, sizeof(float) * 4096);
    delete[] a;
  }
  return 1;
}

// The reference's own entry point for the path: the OpticalFlowClass constructor
// (include/optical_flow.hpp:43-54), called exactly as src/main.cpp:184-189 does.
__attribute__((visibility("default"))) void ref_optical_flow(
    float** img_first, float** img_first_dx, float** img_first_dy, float** img_second,
    float** img_second_dx, float** img_second_dy, int img_padding, float* outflow, int width,
    int height, int coarsest, int finest, int iterations, int patch_size, float overlap,
    int patnorm) {
  QuietCout q;
  OpticalFlow::OpticalFlowClass ofc(img_first, img_first_dx, img_first_dy, img_second,
                                    img_second_dx, img_second_dy, img_padding, outflow, width,
                                    height, coarsest, finest, iterations, patch_size, overlap,
                                    patnorm != 0, false);
}

// Same computation, but driving PatchGrid through its public API (include/patch_grid.hpp:16-30)
// level by level so that per-level artefacts can be dumped: the sequence of calls repeats
// src/optical_flow.cpp:33-90.  For level index s = scale - finest: lvl_flow[s] (2*w*h),
// lvl_ref[s] and lvl_u[s] (2*npatches, u = query_pos - ref_pos is NOT used: u is read back as
// get_patch_query_pos - get_patch_ref_pos only for the position dump; the exact u comes from
// the dense flow test). Any dump pointer may be null.
__attribute__((visibility("default"))) void ref_optical_flow_levels(
    float** img_first, float** img_first_dx, float** img_first_dy, float** img_second,
    float** img_second_dx, float** img_second_dy, int img_padding, float* outflow, int width,
    int height, int coarsest, int finest, int iterations, int patch_size, float overlap,
    int patnorm, float** lvl_flow, float** lvl_ref, float** lvl_query, int* lvl_npatch) {
  using namespace OpticalFlow;
  fix_parameters fp;
  fp.patch_size = patch_size;
  fp.outlierthresh = (float)patch_size / 2;
  fp.coarsest_scale = coarsest;
  fp.finest_scale = finest;
  fp.iterations = iterations;
  fp.steps = std::max(1, (int)std::floor(patch_size * (1 - overlap)));
  fp.num_points_patch = patch_size * patch_size;
  fp.patch_normalization = patnorm != 0;
  const int nl = coarsest - finest + 1;
  std::vector<image_parameters> ip(nl);
  std::vector<PatchGrid*> grids(nl);
  std::vector<float*> flows(nl);
  for (int scale = coarsest; scale >= finest; --scale) {
    const int s = scale - finest;
    const float sf = std::pow(2, -scale);
    ip[s].height = height * sf;
    ip[s].width = width * sf;
    ip[s].img_padding = img_padding;
    ip[s].tmp_lb = -(float)patch_size / 2;
    ip[s].tmp_ub_w = (float)(ip[s].width + patch_size / 2 - 2);
    ip[s].tmp_ub_h = (float)(ip[s].height + patch_size / 2 - 2);
    ip[s].tmp_w = ip[s].width + 2 * img_padding;
    ip[s].tmp_h = ip[s].height + 2 * img_padding;
    flows[s] = new float[2 * ip[s].width * ip[s].height];
    grids[s] = new PatchGrid(&ip[s], &fp);
  }
  for (int scale = coarsest; scale >= finest; --scale) {
    const int s = scale - finest;
    grids[s]->init_grid(img_first[scale], img_first_dx[scale], img_first_dy[scale]);
    grids[s]->set_target_image(img_second[scale], img_second_dx[scale], img_second_dy[scale]);
    if (scale < coarsest) grids[s]->patch_init_from_prev_flow(flows[s + 1]);
    grids[s]->inverse_search();
    float* dst = (scale == finest) ? outflow : flows[s];
    grids[s]->densification_and_create_dance_flow(dst);
    const int np = grids[s]->get_num_all_patch();
    if (lvl_npatch) lvl_npatch[s] = np;
    if (lvl_flow && lvl_flow[s])
      std::memcpy(lvl_flow[s], dst, sizeof(float) * 2 * ip[s].width * ip[s].height);
    for (int i = 0; i < np; ++i) {
      if (lvl_ref && lvl_ref[s]) {
        lvl_ref[s][2 * i] = grids[s]->get_patch_ref_pos(i)[0];
        lvl_ref[s][2 * i + 1] = grids[s]->get_patch_ref_pos(i)[1];
      }
      if (lvl_query && lvl_query[s]) {
        lvl_query[s][2 * i] = grids[s]->get_patch_query_pos(i)[0];
        lvl_query[s][2 * i + 1] = grids[s]->get_patch_query_pos(i)[1];
      }
    }
  }
  for (int s = 0; s < nl; ++s) { delete[] flows[s]; delete grids[s]; }
}

}  // extern "C"

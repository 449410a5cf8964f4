// include/dis_b200.hpp — the reference's C++ class API on top of the C ABI of dis_b200.h.
//
// Header-only host C++ (no CUDA, no Eigen, no OpenCV needed to compile it; link libdis_b200.so).
// It re-creates, with the same names, parameter order and call sequence, the three classes a user
// of the reference touches (paths relative to /root/reference):
//
//   OpticalFlow::OpticalFlowClass   include/optical_flow.hpp:40-66, src/optical_flow.cpp:19-132
//   OpticalFlow::PatchGrid          include/patch_grid.hpp:13-52,   src/patch_grid.cpp:17-182
//   OpticalFlow::Patch              include/patch.hpp:30-73,        src/patch.cpp:11-267
//   OpticalFlow::image_parameters / fix_parameters   include/optical_flow.hpp:14-37
//
// Differences, all forced by the toolchain or by defects documented in SURVEY.md App. C:
//   * Eigen::Vector2f -> OpticalFlow::Vector2f (layout-compatible POD of two floats);
//     Eigen::Map<MatrixXf> -> OpticalFlow::PlaneMap (pointer + dims; only data() is ever used by
//     the reference, src/patch.cpp:47-73, 207-262).
//   * All arithmetic runs on the GPU; errors (no device, bad sizes — the reference would read out of
//     bounds) surface as std::runtime_error carrying dis_last_error().
//   * draw_grid (an OpenCV GUI path, src/optical_flow.cpp:92-125) is accepted and ignored.
#ifndef DIS_B200_HPP
#define DIS_B200_HPP

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "dis_b200.h"

namespace OpticalFlow {

struct Vector2f {                       // stands in for Eigen::Vector2f
  float v[2];
  Vector2f() : v{0.f, 0.f} {}
  Vector2f(float x, float y) : v{x, y} {}
  float& operator[](int i) { return v[i]; }
  float operator[](int i) const { return v[i]; }
  float x() const { return v[0]; }
  float y() const { return v[1]; }
  float* data() { return v; }
};

struct PlaneMap {                       // stands in for Eigen::Map<MatrixXf> over a padded plane
  float* ptr;
  int rows_, cols_;
  PlaneMap(float* p, int rows, int cols) : ptr(p), rows_(rows), cols_(cols) {}
  float* data() const { return ptr; }
  int rows() const { return rows_; }
  int cols() const { return cols_; }
};

typedef struct {                        // include/optical_flow.hpp:14-24
  int width, height, img_padding;
  float tmp_lb, tmp_ub_w, tmp_ub_h;
  int tmp_w, tmp_h;
} image_parameters;

typedef struct {                        // include/optical_flow.hpp:26-37
  int coarsest_scale, finest_scale, patch_size, iterations;
  bool patch_normalization;
  float outlierthresh;
  int steps;
  int num_points_patch;
} fix_parameters;

namespace detail {
inline int& device_ordinal() {
  static int dev = [] { const char* s = std::getenv("DIS_DEVICE"); return s ? std::atoi(s) : 0; }();
  return dev;
}
inline void check(int rc) {
  if (rc != DIS_OK) throw std::runtime_error(std::string("dis_b200: ") + dis_last_error());
}
// An engine for ONE level with explicitly given fix_parameters (what PatchGrid / Patch receive).
inline dis_engine* level_engine(const fix_parameters* f) {
  dis_params p;
  dis_default_params(&p);
  p.max_iter = f->iterations;
  p.patch_size = f->patch_size;
  p.coarsest_scale = 0; p.finest_scale = 0;
  p.patch_norm = f->patch_normalization ? 1 : 0;
  p.steps = f->steps;
  p.outlier_thresh = f->outlierthresh;
  dis_engine* e = nullptr;
  check(dis_create(&p, device_ordinal(), &e));
  return e;
}
}  // namespace detail

// CUDA device used by objects created afterwards (default: $DIS_DEVICE or 0).
inline void set_device(int ordinal) { detail::device_ordinal() = ordinal; }

// ---- Patch: one patch of one level (include/patch.hpp:30-73) ------------------------------------
class Patch {
 public:
  Patch(image_parameters* image_param_in, fix_parameters* fix_param_in)
      : image_param(image_param_in), fix_param(fix_param_in) {}
  ~Patch() { release(); }
  Patch(const Patch&) = delete;
  Patch& operator=(const Patch&) = delete;

  // src/patch.cpp:31-45: remember the first image's planes and the reference point, gather the
  // template gradients and build the Hessian (here: on the device, at inverse_search time).
  void init_patch(PlaneMap* img_first_in, PlaneMap* img_first_dx_in, PlaneMap* img_first_dy_in, Vector2f pt_ref_in) {
    release();
    engine = detail::level_engine(fix_param);
    detail::check(dis_grid_create_single(engine, image_param->width, image_param->height, image_param->img_padding,
                                         pt_ref_in[0], pt_ref_in[1], &grid));
    detail::check(dis_grid_init(grid, img_first_in ? img_first_in->data() : nullptr, img_first_dx_in->data(),
                                img_first_dy_in->data()));
    patch_first_pos = pt_ref_in;
    reset_patch();
  }
  // src/patch.cpp:93-100
  void set_target_image(PlaneMap* img_second_in, PlaneMap* img_second_dx_in, PlaneMap* img_second_dy_in) {
    need_grid();
    detail::check(dis_grid_set_target(grid, img_second_in->data(), img_second_dx_in ? img_second_dx_in->data() : nullptr,
                                      img_second_dy_in ? img_second_dy_in->data() : nullptr));
    reset_patch();
  }
  // src/patch.cpp:156-203 (with inverse_search_start, :119-154): iterate from the initial displacement.
  void inverse_search(Vector2f patch_in) {
    need_grid();
    detail::check(dis_grid_set_patch_init(grid, 0, patch_in[0], patch_in[1]));
    detail::check(dis_grid_inverse_search(grid));
    float ref[2], uu[2];
    detail::check(dis_grid_get_patch(grid, 0, ref, uu));
    u = Vector2f(uu[0], uu[1]);
    pos = Vector2f(ref[0] + uu[0], ref[1] + uu[1]);
  }
  Vector2f get_patch_pos() { return pos; }          // include/patch.hpp:41
  bool is_valid() { return true; }                  // `invalid` is never set by the reference
  Vector2f* get_u_vector() { return &u; }           // include/patch.hpp:43

 private:
  void reset_patch() { u = Vector2f(); pos = patch_first_pos; }   // src/patch.cpp:102-116
  void need_grid() const { if (!grid) throw std::runtime_error("dis_b200: Patch::init_patch must be called first"); }
  void release() {
    if (grid) dis_grid_destroy(grid);
    if (engine) dis_destroy(engine);
    grid = nullptr; engine = nullptr;
  }
  image_parameters* image_param;
  fix_parameters* fix_param;
  dis_engine* engine = nullptr;
  dis_grid* grid = nullptr;
  Vector2f patch_first_pos, u, pos;
};

// ---- PatchGrid: all patches of one level (include/patch_grid.hpp:13-52) ---------------------------
class PatchGrid {
 public:
  PatchGrid(image_parameters* image_param_in, fix_parameters* fix_param_in)   // src/patch_grid.cpp:17-51
      : image_param(image_param_in), fix_param(fix_param_in) {
    engine = detail::level_engine(fix_param);
    int rc = dis_grid_create(engine, image_param->width, image_param->height, image_param->img_padding, &grid);
    if (rc != DIS_OK) { dis_destroy(engine); engine = nullptr; detail::check(rc); }
    num_all_patch = dis_grid_num_patches(grid);
  }
  ~PatchGrid() {
    if (grid) dis_grid_destroy(grid);
    if (engine) dis_destroy(engine);
  }
  PatchGrid(const PatchGrid&) = delete;
  PatchGrid& operator=(const PatchGrid&) = delete;

  void init_grid(float* img_first_in, float* img_first_dx_in, float* img_first_dy_in) {          // :68-82
    detail::check(dis_grid_init(grid, img_first_in, img_first_dx_in, img_first_dy_in));
    searched = false;
  }
  void set_target_image(float* img_second_in, float* img_second_dx_in, float* img_second_dy_in) { // :84-97
    detail::check(dis_grid_set_target(grid, img_second_in, img_second_dx_in, img_second_dy_in));
    searched = false;
  }
  void patch_init_from_prev_flow(float* flow_prev) {                                              // :108-119
    detail::check(dis_grid_init_from_prev_flow(grid, flow_prev));
  }
  void inverse_search() {                                                                         // :99-106
    detail::check(dis_grid_inverse_search(grid));
    searched = true;
  }
  void densification_and_create_dance_flow(float* dense_flow) {                                   // :121-182
    if (!searched) {   // every u is still 0 (Patch::reset_patch): the weighted mean of zeros
      std::memset(dense_flow, 0, sizeof(float) * 2 * (size_t)image_param->width * image_param->height);
      return;
    }
    detail::check(dis_grid_densify(grid, dense_flow));
  }
  int get_num_all_patch() { return num_all_patch; }
  Vector2f get_patch_ref_pos(int i) {                                                             // hpp:29
    float r[2]; detail::check(dis_grid_get_patch(grid, i, r, nullptr)); return Vector2f(r[0], r[1]);
  }
  Vector2f get_patch_query_pos(int i) {                                                           // hpp:30
    float r[2], uu[2]; detail::check(dis_grid_get_patch(grid, i, r, uu)); return Vector2f(r[0] + uu[0], r[1] + uu[1]);
  }

 private:
  image_parameters* image_param;
  fix_parameters* fix_param;
  dis_engine* engine = nullptr;
  dis_grid* grid = nullptr;
  int num_all_patch = 0;
  bool searched = false;
};

// ---- OpticalFlowClass: the whole coarse-to-fine run happens in the constructor --------------------
class OpticalFlowClass {
 public:
  OpticalFlowClass(float** img_first_in, float** img_first_dx_in, float** img_first_dy_in,
                   float** img_second_in, float** img_second_dx_in, float** img_second_dy_in,
                   int img_padding_in, float* outflow, int width_in, int height_in,
                   int coarsest_scale, int finest_scale, int iterations, int patch_size,
                   float patch_overlap, bool patnorm_in, bool draw_grid) {
    (void)draw_grid;
    dis_params p;
    dis_default_params(&p);
    p.max_iter = iterations; p.patch_size = patch_size;
    p.coarsest_scale = coarsest_scale; p.finest_scale = finest_scale;
    p.patch_overlap = patch_overlap; p.patch_norm = patnorm_in ? 1 : 0;
    dis_engine* e = nullptr;
    detail::check(dis_create(&p, detail::device_ordinal(), &e));
    for (int s = coarsest_scale; s >= finest_scale; --s) std::printf("scale: %d\n", s);   // src/optical_flow.cpp:69
    int rc = dis_run_pyramids(e, img_first_in, img_first_dx_in, img_first_dy_in, img_second_in, img_second_dx_in,
                              img_second_dy_in, img_padding_in, outflow, width_in, height_in);
    dis_destroy(e);
    detail::check(rc);
  }
};

}  // namespace OpticalFlow
#endif  // DIS_B200_HPP

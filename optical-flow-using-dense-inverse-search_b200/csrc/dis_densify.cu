// dis_densify.cu — densification of per-patch displacements into a dense flow field, the
// finest-level crop and the finest_scale > 0 epilogue.
//
// Replaces PatchGrid::densification_and_create_dance_flow (src/patch_grid.cpp:121-182) and
// src/main.cpp:192-198.  The reference scatters patch by patch (ascending patch id, x-major) into
// `dense_flow` and a weight image; here every pixel GATHERS the patches that cover it in the same
// ascending-id order, so the fp32 sums are bit-identical and no atomics are needed.  The weight is
// the reference's constant 0.5 per patch, with `weight` taken as zero-initialised
// (SURVEY.md F5 / App. C-1).
#include "dis_kernels.cuh"

namespace dis {
namespace {

__device__ __forceinline__ int ceil_div_pos(int a, int b) { return a >= 0 ? (a + b - 1) / b : -((-a) / b); }
__device__ __forceinline__ int floor_div(int a, int b) { return a >= 0 ? a / b : -((-a + b - 1) / b); }

__global__ void __launch_bounds__(256) k_densify(const DensifyArgs A) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = A.y0 + blockIdx.y * blockDim.y + threadIdx.y;
  const int pair = blockIdx.z;
  if (x >= A.width || y >= A.y1) return;
  const int half = A.ps / 2;
  // patch (gx,gy) with centre c covers pixels c-half .. c+half-1  <=>  x-half+1 <= c <= x+half
  int gx_lo = ceil_div_pos(x - half + 1 - A.offw, A.steps), gx_hi = floor_div(x + half - A.offw, A.steps);
  int gy_lo = ceil_div_pos(y - half + 1 - A.offh, A.steps), gy_hi = floor_div(y + half - A.offh, A.steps);
  gx_lo = max(gx_lo, 0); gy_lo = max(gy_lo, 0);
  gx_hi = min(gx_hi, A.npw - 1); gy_hi = min(gy_hi, A.nph - 1);
  const float2* __restrict__ u = reinterpret_cast<const float2*>(A.u) + (size_t)pair * A.np;
  float sx = 0.f, sy = 0.f, wsum = 0.f;
  for (int gx = gx_lo; gx <= gx_hi; ++gx)          // ascending patch id = x-major (src/patch_grid.cpp:39-49)
    for (int gy = gy_lo; gy <= gy_hi; ++gy) {
      const float2 v = __ldg(u + gx * A.nph + gy);
      sx += v.x * 0.5f;                             // new_u = u * 0.5f (src/patch_grid.cpp:156)
      sy += v.y * 0.5f;
      wsum += 0.5f;                                 // src/patch_grid.cpp:157
    }
  if (wsum > 0.f) { sx /= wsum; sy /= wsum; }       // src/patch_grid.cpp:173-177
  reinterpret_cast<float2*>(A.dense + (size_t)pair * A.dense_stride)[(size_t)y * A.width + x] = make_float2(sx, sy);
}

// Halo exchange of a row-banded run: grid rows [row0, row0 + nrows) of the per-patch flows.  `src` may
// live on a PEER GPU (peer access enabled): the loads then travel over NVLink, issued by this GPU's SMs.
__global__ void __launch_bounds__(256) k_copy_u_rows(const float2* __restrict__ src, float2* __restrict__ dst, int npw,
                                                     int nph, int row0, int nrows) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npw * nrows) return;
  const int gx = i / nrows, gy = row0 + (i - gx * nrows);
  dst[(size_t)gx * nph + gy] = src[(size_t)gx * nph + gy];
}

__global__ void __launch_bounds__(256) k_crop(const float2* __restrict__ dense, size_t dense_stride2, int W,
                                              int left, int top, float2* __restrict__ out, int w, int h) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  const int pair = blockIdx.z;
  if (x >= w || y >= h) return;
  out[((size_t)pair * h + y) * w + x] = dense[(size_t)pair * dense_stride2 + (size_t)(y + top) * W + (x + left)];
}

// src/main.cpp:192-198: flowout *= sc; cv::resize(sc, sc, INTER_LINEAR); crop.  OpenCV's INTER_LINEAR for fp32 is
// separable and runs in fp32 without FMA: destination x -> source fx = (x + 0.5) / sc - 0.5, sx = floor(fx), weights
// (1 - (fx - sx), fx - sx); HORIZONTALLY a tap index outside the row is clamped and its weight set to 0 (resize.cpp,
// xofs/alpha set-up), VERTICALLY only the row indices are clamped and the weights stay (so a border row is
// S * b0 + S * b1, not S); rows are interpolated horizontally first, then S0 * b0 + S1 * b1.  Restated in exactly that
// order; bit-identical to cv2.resize (4.13, setUseOptimized(False)) for scales 2, 4, 8 (tests/test_gpu_api.py).
__global__ void __launch_bounds__(256) k_upsample_crop(const float2* __restrict__ dense, size_t dense_stride2,
                                                       int Wf, int Hf, int finest, int left, int top,
                                                       float2* __restrict__ out, int w, int h) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  const int pair = blockIdx.z;
  if (x >= w || y >= h) return;
  const float sc = (float)(1 << finest), inv = 1.0f / sc;
  const float fx = ((float)(x + left) + 0.5f) * inv - 0.5f, fy = ((float)(y + top) + 0.5f) * inv - 0.5f;
  int x0 = (int)floorf(fx), sy = (int)floorf(fy);
  float ax = fx - (float)x0;
  const float ay = fy - (float)sy;
  if (x0 < 0) { x0 = 0; ax = 0.f; }
  if (x0 >= Wf - 1) { x0 = Wf - 1; ax = 0.f; }
  const int x1 = min(x0 + 1, Wf - 1);
  const int y0 = min(max(sy, 0), Hf - 1), y1 = min(max(sy + 1, 0), Hf - 1);
  const float a0 = 1.f - ax, a1 = ax, b0 = 1.f - ay, b1 = ay;
  const float2* d = dense + (size_t)pair * dense_stride2;
  const float2 p00 = d[(size_t)y0 * Wf + x0], p01 = d[(size_t)y0 * Wf + x1];
  const float2 p10 = d[(size_t)y1 * Wf + x0], p11 = d[(size_t)y1 * Wf + x1];
  const float hx0 = (p00.x * sc) * a0 + (p01.x * sc) * a1, hx1 = (p10.x * sc) * a0 + (p11.x * sc) * a1;
  const float hy0 = (p00.y * sc) * a0 + (p01.y * sc) * a1, hy1 = (p10.y * sc) * a0 + (p11.y * sc) * a1;
  out[((size_t)pair * h + y) * w + x] = make_float2(hx0 * b0 + hx1 * b1, hy0 * b0 + hy1 * b1);
}

}  // namespace

int launch_densify(const DensifyArgs& a_in, cudaStream_t s) {
  DensifyArgs a = a_in;
  if (a.y1 <= a.y0) { a.y0 = 0; a.y1 = a.height; }
  a.y0 = a.y0 < 0 ? 0 : a.y0;
  a.y1 = a.y1 > a.height ? a.height : a.y1;
  if (a.y1 <= a.y0) return 0;
  dim3 b(32, 8), g((a.width + 31) / 32, (a.y1 - a.y0 + 7) / 8, a.n_pairs);
  k_densify<<<g, b, 0, s>>>(a);
  return 1;
}

int launch_copy_u_rows(const float* src, float* dst, int npw, int nph, int row0, int nrows, cudaStream_t s) {
  if (nrows <= 0) return 0;
  const int n = npw * nrows;
  k_copy_u_rows<<<(n + 255) / 256, 256, 0, s>>>(reinterpret_cast<const float2*>(src), reinterpret_cast<float2*>(dst), npw, nph, row0, nrows);
  return 1;
}

int launch_crop(const float* dense, size_t dense_stride, int W, int H, int left, int top, float* out, int w,
                int h, int n_pairs, cudaStream_t s) {
  (void)H;
  dim3 b(32, 8), g((w + 31) / 32, (h + 7) / 8, n_pairs);
  k_crop<<<g, b, 0, s>>>(reinterpret_cast<const float2*>(dense), dense_stride / 2, W, left, top,
                         reinterpret_cast<float2*>(out), w, h);
  return 1;
}

int launch_upsample_crop(const float* dense, size_t dense_stride, int Wf, int Hf, int finest, int left, int top,
                         float* out, int w, int h, int n_pairs, cudaStream_t s) {
  dim3 b(32, 8), g((w + 31) / 32, (h + 7) / 8, n_pairs);
  k_upsample_crop<<<g, b, 0, s>>>(reinterpret_cast<const float2*>(dense), dense_stride / 2, Wf, Hf, finest, left,
                                  top, reinterpret_cast<float2*>(out), w, h);
  return 1;
}

}  // namespace dis

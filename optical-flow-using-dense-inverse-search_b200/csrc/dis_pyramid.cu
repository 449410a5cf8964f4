// dis_pyramid.cu — pre-processing on the device.
//
// Replaces, for every frame of a run:
//   divisibility pad + u8->f32         src/main.cpp:139-160
//   construct_pyramide                 src/main.cpp:12-50
//     level 0 = sqrt(Sobel_x^2 + Sobel_y^2) of the frame        :18-26
//     level i = 0.5x INTER_LINEAR resize of level i-1            :29   (2x2 mean, App. A.1 item 4)
//     per level Sobel dx, dy (scale 1/8, BORDER_REFLECT_101)     :34-35
//     copyMakeBorder by img_padding: replicate / zero            :41-49
// The padded planes are written directly (no separate copyMakeBorder pass): interior by the
// producing kernel, borders by k_sobel_border.  All arithmetic in the association of App. A.1.
#include "dis_kernels.cuh"

namespace dis {
namespace {

__device__ __forceinline__ int reflect101(int i, int n) {
  if (n == 1) return 0;
  if (i < 0) i = -i;
  if (i >= n) i = 2 * n - 2 - i;
  return i;
}

// cv::Sobel ksize 3, scale 1/8 (the scale is folded into the smoothing taps .125,.25,.125)
struct Sobel { float gx, gy; };
template <typename F>
__device__ __forceinline__ Sobel sobel_at(F at, int x, int y, int w, int h) {
  const int xm = reflect101(x - 1, w), xp = reflect101(x + 1, w);
  const int ym = reflect101(y - 1, h), yp = reflect101(y + 1, h);
  const float a = at(xm, ym), b = at(x, ym), c = at(xp, ym);
  const float d = at(xm, y), f = at(xp, y);
  const float g = at(xm, yp), hh = at(x, yp), i = at(xp, yp);
  Sobel r;
  const float d0 = c - a, d1 = f - d, d2 = i - g;
  r.gx = (d0 + d2) * 0.125f + d1 * 0.25f;
  const float s0 = (a + c) * 0.125f + b * 0.25f;
  const float s2 = (g + i) * 0.125f + hh * 0.25f;
  r.gy = s2 - s0;
  return r;
}

// level 0 image plane interior = gradient magnitude of the (divisibility padded) frame
__global__ void __launch_bounds__(256) k_gradmag(const PyramidArgs A) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  const int f = blockIdx.z;
  if (x >= A.W || y >= A.H) return;
  Sobel s;
  if (A.frames) {
    const uint8_t* img = A.frames + (size_t)f * A.frame_stride_bytes;
    auto at = [&](int xx, int yy) {   // copyMakeBorder(BORDER_REPLICATE) of src/main.cpp:152 + convertTo
      const int sx = min(max(xx - A.left, 0), A.w0 - 1), sy = min(max(yy - A.top, 0), A.h0 - 1);
      return (float)__ldg(img + (size_t)sy * A.stride0 + sx);
    };
    s = sobel_at(at, x, y, A.W, A.H);
  } else {
    const float* img = A.frames_f32 + (size_t)f * A.W * A.H;
    auto at = [&](int xx, int yy) { return __ldg(img + (size_t)yy * A.W + xx); };
    s = sobel_at(at, x, y, A.W, A.H);
  }
  const float m = sqrtf(s.gx * s.gx + s.gy * s.gy);
  const LevelPlanes& P = A.planes[0];
  const int tw = A.W + 2 * A.pad;
  P.base[(size_t)f * P.frame_stride + (size_t)(y + A.pad) * tw + (x + A.pad)] = m;
}

// level `lvl` image plane interior from level lvl-1
__global__ void __launch_bounds__(256) k_halve(const PyramidArgs A, int lvl) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  const int f = blockIdx.z;
  const int w = A.W >> lvl, h = A.H >> lvl;
  if (x >= w || y >= h) return;
  const LevelPlanes& S = A.planes[lvl - 1];
  const LevelPlanes& D = A.planes[lvl];
  const int stw = (A.W >> (lvl - 1)) + 2 * A.pad, dtw = w + 2 * A.pad;
  const float* r0 = S.base + (size_t)f * S.frame_stride + (size_t)(2 * y + A.pad) * stw + (2 * x + A.pad);
  const float* r1 = r0 + stw;
  D.base[(size_t)f * D.frame_stride + (size_t)(y + A.pad) * dtw + (x + A.pad)] =
      ((r0[0] + r0[1]) + (r1[0] + r1[1])) * 0.25f;
}

// one thread per padded-plane pixel of level `lvl`: interior -> Sobel dx, dy of the image plane;
// border -> replicate the image plane, zero the gradient planes
__global__ void __launch_bounds__(256) k_sobel_border(const PyramidArgs A, int lvl) {
  const int X = blockIdx.x * blockDim.x + threadIdx.x;
  const int Y = blockIdx.y * blockDim.y + threadIdx.y;
  const int f = blockIdx.z;
  const int w = A.W >> lvl, h = A.H >> lvl, pad = A.pad;
  const int tw = w + 2 * pad, th = h + 2 * pad;
  if (X >= tw || Y >= th) return;
  const LevelPlanes& P = A.planes[lvl];
  float* img = P.base + (size_t)f * P.frame_stride;
  float2* gxy = reinterpret_cast<float2*>(img + P.plane_stride);
  const int x = X - pad, y = Y - pad;
  const size_t o = (size_t)Y * tw + X;
  if (x >= 0 && x < w && y >= 0 && y < h) {
    auto at = [&](int xx, int yy) { return img[(size_t)(yy + pad) * tw + (xx + pad)]; };
    const Sobel s = sobel_at(at, x, y, w, h);
    gxy[o] = make_float2(s.gx, s.gy);
  } else {
    const int cx = min(max(x, 0), w - 1), cy = min(max(y, 0), h - 1);
    img[o] = img[(size_t)(cy + pad) * tw + (cx + pad)];
    gxy[o] = make_float2(0.f, 0.f);
  }
}

__global__ void __launch_bounds__(256) k_interleave(const float* __restrict__ dx, const float* __restrict__ dy,
                                                    float2* __restrict__ gxy, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    gxy[i] = make_float2(dx[i], dy[i]);
}

__global__ void __launch_bounds__(256) k_deinterleave(const float2* __restrict__ gxy, float* __restrict__ dx,
                                                      float* __restrict__ dy, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float2 v = gxy[i];
    dx[i] = v.x; dy[i] = v.y;
  }
}

}  // namespace

int launch_interleave(const float* dx, const float* dy, float* gxy, size_t n, cudaStream_t s) {
  const unsigned blocks = (unsigned)((n + 255) / 256 < (size_t)sm_count() * 8 ? (n + 255) / 256 : (size_t)sm_count() * 8);
  k_interleave<<<blocks, 256, 0, s>>>(dx, dy, reinterpret_cast<float2*>(gxy), n);
  return 1;
}

int launch_deinterleave(const float* gxy, float* dx, float* dy, size_t n, cudaStream_t s) {
  const unsigned blocks = (unsigned)((n + 255) / 256 < (size_t)sm_count() * 8 ? (n + 255) / 256 : (size_t)sm_count() * 8);
  k_deinterleave<<<blocks, 256, 0, s>>>(reinterpret_cast<const float2*>(gxy), dx, dy, n);
  return 1;
}

int launch_pyramid(const PyramidArgs& a, cudaStream_t s) {
  int n = 0;
  dim3 b(32, 8);
  {
    dim3 g((a.W + 31) / 32, (a.H + 7) / 8, a.n_frames);
    k_gradmag<<<g, b, 0, s>>>(a); ++n;
  }
  for (int l = 1; l < a.levels; ++l) {
    dim3 g(((a.W >> l) + 31) / 32, ((a.H >> l) + 7) / 8, a.n_frames);
    k_halve<<<g, b, 0, s>>>(a, l); ++n;
  }
  for (int l = 0; l < a.levels; ++l) {
    dim3 g(((a.W >> l) + 2 * a.pad + 31) / 32, ((a.H >> l) + 2 * a.pad + 7) / 8, a.n_frames);
    k_sobel_border<<<g, b, 0, s>>>(a, l); ++n;
  }
  return n;
}

}  // namespace dis

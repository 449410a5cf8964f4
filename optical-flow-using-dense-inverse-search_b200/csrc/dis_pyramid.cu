// dis_pyramid.cu — pre-processing on the device.
//
// Replaces, for every frame of a run:
//   divisibility pad + u8->f32         src/main.cpp:139-160
//   construct_pyramide                 src/main.cpp:12-50
//     level 0 = sqrt(Sobel_x^2 + Sobel_y^2) of the frame        :18-26
//     level i = 0.5x INTER_LINEAR resize of level i-1            :29   (2x2 mean, App. A.1 item 4)
//     per level Sobel dx, dy (scale 1/8, BORDER_REFLECT_101)     :34-35
//     copyMakeBorder by img_padding: replicate / zero            :41-49
// The padded planes are written directly (no separate copyMakeBorder pass).  All arithmetic in the
// association of App. A.1.
#include "dis_kernels.cuh"

namespace dis {
namespace {

__device__ __forceinline__ int reflect101(int i, int n) {
  if (n == 1) return 0;
  if (i < 0) i = -i;
  if (i >= n) i = 2 * n - 2 - i;
  return i;
}

// BORDER_REFLECT_101 for any offset (cv::borderInterpolate loops the same way); tiles that stick out of a small or
// partial image ask for positions more than one period away — their values feed nothing that is stored
__device__ __forceinline__ int reflect_any(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
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

// ---- one kernel per level, shared-memory tiles, 128-bit stores --------------------------------------------------------
// A block of 256 threads produces a 64 x 16 tile of a level: it stages the tile of the level's image plane P with a
// 1-pixel halo in shared memory (level 0: computes it there as the Sobel gradient magnitude of a u8 tile with a
// 2-pixel halo, loaded with uchar4 where the group lies inside the frame), takes the Sobel gradients from the tile,
// and writes
//   * Gx,Gy interleaved: 4 pixels per thread = two float4 stores,
//   * level 0 only: P itself, one float4 per thread,
//   * the NEXT level's P (2x2 mean of the tile, one value per thread, coalesced rows),
//   * for tiles on the rim of the image: the img_padding border of both planes (replicate / zero, src/main.cpp:41-49).
// So a frame's pyramid is `levels` launches (was 2 * levels), every plane is written once and read once.
// REFLECT_101 is applied when the tile is STAGED: a halo value outside the image is the value at the mirrored position;
// for level 0 it is computed from the mirrored u8 neighbourhood, which gives the same magnitude bit for bit (gx changes
// sign, (a + c) is commutative).  All arithmetic in the association of App. A.1.
constexpr int kTW = 64, kTH = 16;          // tile of output pixels; thread (tx, ty) owns pixels 4 tx .. 4 tx + 3 of row ty
constexpr int kPW = kTW + 2, kPH = kTH + 2;      // P tile with halo 1
constexpr int kUW = kTW + 4, kUH = kTH + 4;      // u8 tile with halo 2 (level 0)

template <int MODE>   // 0: level 0 from u8 frames, 1: level 0 from fp32 frames, 2: level >= 1 from its own image plane
__global__ void __launch_bounds__(256) k_level(const PyramidArgs A, int lvl) {
  __shared__ float Ps[kPH][kPW + 2];
  __shared__ float Us[MODE == 2 ? 1 : kUH][MODE == 2 ? 1 : kUW];
  const int f = blockIdx.z;
  const int w = A.W >> lvl, h = A.H >> lvl, pad = A.pad;
  const int tw = w + 2 * pad;
  const int x0 = blockIdx.x * kTW, y0 = blockIdx.y * kTH;
  const int tid = threadIdx.x;
  const LevelPlanes& P = A.planes[lvl];
  float* img = P.base + (size_t)f * P.frame_stride;
  float2* gxy = reinterpret_cast<float2*>(img + P.plane_stride);

  if (MODE == 2) {
    // stage P (66 x 18) from the level's own plane, mirrored at the image rim
    for (int i = tid; i < kPH * kPW; i += 256) {
      const int r = i / kPW, c = i - r * kPW;
      const int yy = reflect_any(y0 + r - 1, h), xx = reflect_any(x0 + c - 1, w);
      Ps[r][c] = img[(size_t)(yy + pad) * tw + (xx + pad)];
    }
  } else {
    // stage the source frame (68 x 20) as float; (X, Y) are coordinates of the divisibility-padded frame
    if (MODE == 0) {
      // 18 aligned groups of 4 bytes [x0 - 4 + 4 g, +4) cover the 68 tile columns X = x0 - 2 .. x0 + 65 (tile column
      // c = X - (x0 - 2) = 4 g - 2 + k); a group that lies inside the frame and is 4-aligned in memory is ONE uchar4 load
      const uint8_t* src = A.frames + (size_t)f * A.frame_stride_bytes;
      constexpr int kG = 18;
      for (int i = tid; i < kUH * kG; i += 256) {
        const int r = i / kG, g = i - r * kG;
        const int Y = reflect_any(y0 + r - 2, A.H);
        const uint8_t* row = src + (size_t)min(max(Y - A.top, 0), A.h0 - 1) * A.stride0;   // copyMakeBorder(REPLICATE), src/main.cpp:152
        const int Xa = x0 - 4 + 4 * g, sxa = Xa - A.left;
        float v[4];
        if (Xa >= 0 && Xa + 3 < A.W && sxa >= 0 && sxa + 3 < A.w0 && (reinterpret_cast<size_t>(row + sxa) & 3) == 0) {
          const uchar4 q = __ldg(reinterpret_cast<const uchar4*>(row + sxa));
          v[0] = (float)q.x; v[1] = (float)q.y; v[2] = (float)q.z; v[3] = (float)q.w;
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) v[k] = (float)__ldg(row + min(max(reflect_any(Xa + k, A.W) - A.left, 0), A.w0 - 1));
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int c = 4 * g - 2 + k;
          if (c >= 0 && c < kUW) Us[r][c] = v[k];
        }
      }
    } else {
      const float* src = A.frames_f32 + (size_t)f * A.W * A.H;
      for (int i = tid; i < kUH * kUW; i += 256) {
        const int r = i / kUW, c = i - r * kUW;
        const int Y = reflect_any(y0 + r - 2, A.H), X = reflect_any(x0 + c - 2, A.W);
        Us[r][c] = __ldg(src + (size_t)Y * A.W + X);
      }
    }
    __syncthreads();
    // P = sqrt(gx^2 + gy^2) of the frame (src/main.cpp:18-26) for the tile and its halo
    for (int i = tid; i < kPH * kPW; i += 256) {
      const int r = i / kPW, c = i - r * kPW;              // P tile (r, c) <-> source tile (r + 1, c + 1)
      const float a = Us[r][c], b = Us[r][c + 1], cc = Us[r][c + 2];
      const float d = Us[r + 1][c], ff = Us[r + 1][c + 2];
      const float g = Us[r + 2][c], hh = Us[r + 2][c + 1], ii = Us[r + 2][c + 2];
      const float d0 = cc - a, d1 = ff - d, d2 = ii - g;
      const float gx = (d0 + d2) * 0.125f + d1 * 0.25f;
      const float s0 = (a + cc) * 0.125f + b * 0.25f, s2 = (g + ii) * 0.125f + hh * 0.25f;
      const float gy = s2 - s0;
      Ps[r][c] = sqrtf(gx * gx + gy * gy);
    }
  }
  __syncthreads();

  const int tx = tid & 15, ty = tid >> 4;
  const int x = x0 + 4 * tx, y = y0 + ty;
  const bool vec = ((tw & 3) == 0) && ((pad & 3) == 0) && x + 3 < w;
  if (y < h && x < w) {
    float4 pv;
    float2 gv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int r = ty + 1, c = 4 * tx + k + 1;
      const float a = Ps[r - 1][c - 1], b = Ps[r - 1][c], cc = Ps[r - 1][c + 1];
      const float d = Ps[r][c - 1], ff = Ps[r][c + 1];
      const float g = Ps[r + 1][c - 1], hh = Ps[r + 1][c], ii = Ps[r + 1][c + 1];
      const float d0 = cc - a, d1 = ff - d, d2 = ii - g;
      gv[k].x = (d0 + d2) * 0.125f + d1 * 0.25f;                       // cv::Sobel(1, 0), scale 1/8 (src/main.cpp:34)
      const float s0 = (a + cc) * 0.125f + b * 0.25f, s2 = (g + ii) * 0.125f + hh * 0.25f;
      gv[k].y = s2 - s0;                                               // cv::Sobel(0, 1) (src/main.cpp:35)
      (&pv.x)[k] = Ps[r][c];
    }
    const size_t o = (size_t)(y + pad) * tw + (x + pad);
    if (vec) {
      reinterpret_cast<float4*>(gxy + o)[0] = make_float4(gv[0].x, gv[0].y, gv[1].x, gv[1].y);
      reinterpret_cast<float4*>(gxy + o)[1] = make_float4(gv[2].x, gv[2].y, gv[3].x, gv[3].y);
      if (MODE != 2) *reinterpret_cast<float4*>(img + o) = pv;
    } else {
      for (int k = 0; k < 4 && x + k < w; ++k) {
        gxy[o + k] = gv[k];
        if (MODE != 2) img[o + k] = (&pv.x)[k];
      }
    }
  }
  // next level's image plane: 2x2 mean of this tile (src/main.cpp:29; App. A.1 item 4), one value per thread
  if (lvl + 1 < A.levels) {
    const int hx = tid & 31, hy = tid >> 5;                            // 32 x 8 outputs
    const int X = (x0 >> 1) + hx, Y = (y0 >> 1) + hy;
    const int w2 = w >> 1, h2 = h >> 1;
    if (X < w2 && Y < h2) {
      const LevelPlanes& D = A.planes[lvl + 1];
      const int r = 2 * hy + 1, c = 2 * hx + 1;
      D.base[(size_t)f * D.frame_stride + (size_t)(Y + pad) * (w2 + 2 * pad) + (X + pad)] =
          ((Ps[r][c] + Ps[r][c + 1]) + (Ps[r + 1][c] + Ps[r + 1][c + 1])) * 0.25f;
    }
  }
  // the img_padding border next to rim tiles: image plane replicated, gradient planes zero (src/main.cpp:41-49).
  // A rim tile owns the strips beside its own rows / columns; corner tiles also the corners.
  const int tile_w = min(kTW, w - x0), tile_h = min(kTH, h - y0);
  const bool left = x0 == 0, right = x0 + kTW >= w, top = y0 == 0, bottom = y0 + kTH >= h;
  if (left || right) {
    for (int i = tid; i < tile_h * pad; i += 256) {
      const int r = i / pad, k = i - r * pad;
      const size_t row = (size_t)(y0 + r + pad) * tw;
      if (left) { img[row + k] = Ps[r + 1][1]; gxy[row + k] = make_float2(0.f, 0.f); }
      if (right) { img[row + pad + w + k] = Ps[r + 1][tile_w]; gxy[row + pad + w + k] = make_float2(0.f, 0.f); }
    }
  }
  if (top || bottom) {
    const int cx0 = left ? -pad : 0, cx1 = tile_w + (right ? pad : 0);     // columns relative to x0, corners included
    const int span = cx1 - cx0;
    for (int i = tid; i < span * pad; i += 256) {
      const int k = i / span, c = cx0 + (i - k * span);
      const int cc = min(max(c, 0), tile_w - 1);                           // replicate sideways into the corners
      const size_t col = (size_t)(x0 + c + pad);
      if (top) { img[(size_t)k * tw + col] = Ps[1][cc + 1]; gxy[(size_t)k * tw + col] = make_float2(0.f, 0.f); }
      if (bottom) { img[(size_t)(pad + h + k) * tw + col] = Ps[tile_h][cc + 1]; gxy[(size_t)(pad + h + k) * tw + col] = make_float2(0.f, 0.f); }
    }
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
  for (int l = 0; l < a.levels; ++l) {
    dim3 g(((a.W >> l) + kTW - 1) / kTW, ((a.H >> l) + kTH - 1) / kTH, a.n_frames);
    if (l > 0) k_level<2><<<g, 256, 0, s>>>(a, l);
    else if (a.frames) k_level<0><<<g, 256, 0, s>>>(a, l);
    else k_level<1><<<g, 256, 0, s>>>(a, l);
    ++n;
  }
  return n;
}

}  // namespace dis

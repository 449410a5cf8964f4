// dis_color.cu — Middlebury colour coding of a flow field on the device.
//
// Replaces draw_optical_flow / compute_color / is_flow_correct of src/color_coding.cpp:8-117:
//   * maxrad = maxmotion if > 0, else max(1, max over valid pixels of |u|)       :89-106
//   * valid pixel: no NaN and |u.x|, |u.y| < 1e9; invalid pixels stay (0,0,0)     :8-11, :113
//   * colour wheel of 55 integer entries (RY 15, YG 6, GC 4, CB 11, BM 13, MR 6)  :21-53
//   * a = atan2(-fy,-fx)/pi; fk = (a+1)/2*54; linear blend of wheel[k0], wheel[k0+1 mod 55];
//     saturation rises with radius, x0.75 beyond radius 1; stored B,G,R            :55-80
// fp32 in the reference's association (the library is built without FMA contraction).  atan2 is
// evaluated in double and rounded to fp32, i.e. correctly rounded like glibc's atan2f.
#include "dis_kernels.cuh"

namespace dis {
namespace {

constexpr int kNCols = 55;
__constant__ unsigned char c_wheel[kNCols][3];

__device__ __forceinline__ bool flow_ok(float2 u) {
  return !(u.x != u.x) && !(u.y != u.y) && fabsf(u.x) < 1e9f && fabsf(u.y) < 1e9f;
}

// max over valid pixels of u.x^2 + u.y^2 (sqrtf is monotone, so max |u| = sqrtf of this), as the
// bit pattern of a non-negative float (ordered like an unsigned integer).
__global__ void __launch_bounds__(256) k_flow_maxsq(const float2* __restrict__ flow, size_t n, unsigned* __restrict__ out) {
  float m = 0.f;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float2 u = flow[i];
    if (flow_ok(u)) m = fmaxf(m, u.x * u.x + u.y * u.y);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(out, __float_as_uint(m));
}

__global__ void __launch_bounds__(256) k_color_code(const float2* __restrict__ flow, size_t n, float maxmotion,
                                                    const unsigned* __restrict__ maxsq, uchar3* __restrict__ bgr) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float maxrad = maxmotion;
  if (!(maxmotion > 0.f)) maxrad = fmaxf(1.f, sqrtf(__uint_as_float(*maxsq)));
  const float2 u = flow[i];
  uchar3 pix = make_uchar3(0, 0, 0);
  if (flow_ok(u)) {
    const float fx = u.x / maxrad, fy = u.y / maxrad;
    const float rad = sqrtf(fx * fx + fy * fy);
    const float a = (float)atan2((double)-fy, (double)-fx) / 3.14159265358979323846f;
    const float fk = (a + 1.0f) / 2.0f * (float)(kNCols - 1);
    const int k0 = (int)fk;
    const int k1 = (k0 + 1) % kNCols;
    const float f = fk - (float)k0;
    unsigned char c[3];
#pragma unroll
    for (int b = 0; b < 3; ++b) {
      const float col0 = (float)c_wheel[k0][b] / 255.f;
      const float col1 = (float)c_wheel[k1][b] / 255.f;
      float col = (1 - f) * col0 + f * col1;
      if (rad <= 1) col = 1 - rad * (1 - col);
      else col *= .75f;
      c[b] = (unsigned char)(int)(255.f * col);
    }
    pix = make_uchar3(c[2], c[1], c[0]);   // pix[2 - b]: B, G, R
  }
  bgr[i] = pix;
}

}  // namespace

int launch_color_code(const float* flow, size_t n_pixels, float maxmotion, unsigned* scratch, uint8_t* bgr, cudaStream_t s) {
  unsigned char w[kNCols][3];
  int k = 0;
  const int RY = 15, YG = 6, GC = 4, CB = 11, BM = 13, MR = 6;
  for (int i = 0; i < RY; ++i, ++k) { w[k][0] = 255; w[k][1] = (unsigned char)(255 * i / RY); w[k][2] = 0; }
  for (int i = 0; i < YG; ++i, ++k) { w[k][0] = (unsigned char)(255 - 255 * i / YG); w[k][1] = 255; w[k][2] = 0; }
  for (int i = 0; i < GC; ++i, ++k) { w[k][0] = 0; w[k][1] = 255; w[k][2] = (unsigned char)(255 * i / GC); }
  for (int i = 0; i < CB; ++i, ++k) { w[k][0] = 0; w[k][1] = (unsigned char)(255 - 255 * i / CB); w[k][2] = 255; }
  for (int i = 0; i < BM; ++i, ++k) { w[k][0] = (unsigned char)(255 * i / BM); w[k][1] = 0; w[k][2] = 255; }
  for (int i = 0; i < MR; ++i, ++k) { w[k][0] = 255; w[k][1] = 0; w[k][2] = (unsigned char)(255 - 255 * i / MR); }
  cudaMemcpyToSymbolAsync(c_wheel, w, sizeof w, 0, cudaMemcpyHostToDevice, s);
  int n = 0;
  if (!(maxmotion > 0.f)) {
    cudaMemsetAsync(scratch, 0, sizeof(unsigned), s);
    const unsigned blocks = (unsigned)((n_pixels + 255) / 256 < (size_t)sm_count() * 8 ? (n_pixels + 255) / 256 : (size_t)sm_count() * 8);
    k_flow_maxsq<<<blocks ? blocks : 1, 256, 0, s>>>(reinterpret_cast<const float2*>(flow), n_pixels, scratch);
    ++n;
  }
  k_color_code<<<(unsigned)((n_pixels + 255) / 256), 256, 0, s>>>(reinterpret_cast<const float2*>(flow), n_pixels, maxmotion, scratch,
                                                                 reinterpret_cast<uchar3*>(bgr));
  return n + 1;
}

}  // namespace dis

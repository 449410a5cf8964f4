// dis_noalloc.cu — dis_color_code_device_noalloc: the Middlebury colour coding of dis_color.cu for callers that
// cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// NO COUNTERPART IN THE REFERENCE.  The arithmetic is that of
// draw_optical_flow / compute_color (src/color_coding.cpp:8-117) exactly as dis_color.cu restates it; results are
// byte-identical to dis_color_code.  What differs is memory: inputs and outputs are the CALLER's device buffers, and
// the one piece of scratch the path needs — the reduction of max |u| over the field — lives in a file-scope
// __device__ array of fixed size instead of the engine's cudaMalloc'ed stage.  The entry point calls no cudaMalloc /
// cudaFree / allocating routine: it resolves the array's device address once per device (cudaGetSymbolAddress) and
// launches three kernels on the caller's stream.
//
// The array is one per GPU per process and is shared by every call: calls that overlap on one GPU (two streams, two
// threads) race on it.  Nothing in it has to start at zero: every partial a call reads was written by that call.
#include <cuda_runtime.h>
#include <stdio.h>

#include <mutex>
#include <string>

#include "../../include/dis_b200.h"
#include "dis_kernels.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kPxPerBlock = DIS_NOALLOC_PIXELS_PER_PARTIAL;                            // pixels one block of the reduction covers
constexpr int kMaxPartials = DIS_NOALLOC_MAX_PIXELS / DIS_NOALLOC_PIXELS_PER_PARTIAL;  // 8192
constexpr int kHeaderWords = 64;                                                       // [0] = max |u|^2 of the field (bit pattern)

// scratch: 256-byte header + one partial maximum per reduction block (33 KB, held by every GPU that loads the library)
static __device__ __align__(256) unsigned g_noalloc_scratch[kHeaderWords + kMaxPartials];

__device__ __forceinline__ bool flow_ok(float2 u) {   // is_flow_correct, src/color_coding.cpp:8-11
  return !(u.x != u.x) && !(u.y != u.y) && fabsf(u.x) < 1e9f && fabsf(u.y) < 1e9f;
}

// colour wheel entry k, channel b (0 R, 1 G, 2 B), src/color_coding.cpp:21-53 (RY 15, YG 6, GC 4, CB 11, BM 13, MR 6)
__device__ __forceinline__ int wheel(int k, int b) {
  int r, g, bl;
  if (k < 15) { r = 255; g = 255 * k / 15; bl = 0; }
  else if (k < 21) { r = 255 - 255 * (k - 15) / 6; g = 255; bl = 0; }
  else if (k < 25) { r = 0; g = 255; bl = 255 * (k - 21) / 4; }
  else if (k < 36) { r = 0; g = 255 - 255 * (k - 25) / 11; bl = 255; }
  else if (k < 49) { r = 255 * (k - 36) / 13; g = 0; bl = 255; }
  else { r = 255; g = 0; bl = 255 - 255 * (k - 49) / 6; }
  return b == 0 ? r : b == 1 ? g : bl;
}

__device__ __forceinline__ float block_max(float m) {
  __shared__ float s[kThreads / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = m;
  __syncthreads();
  m = threadIdx.x < kThreads / 32 ? s[threadIdx.x] : 0.f;
#pragma unroll
  for (int o = kThreads / 64; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  return m;   // valid in thread 0
}

// stage 1: block b writes max over its kPxPerBlock pixels of u.x^2 + u.y^2 (valid pixels only) to partial[b]
__global__ void __launch_bounds__(kThreads) k_na_partials(const float2* __restrict__ flow, size_t n, unsigned* __restrict__ scratch) {
  const size_t base = (size_t)blockIdx.x * kPxPerBlock;
  float m = 0.f;
#pragma unroll 4
  for (int j = threadIdx.x; j < kPxPerBlock; j += kThreads) {
    const size_t i = base + j;
    if (i < n) {
      const float2 u = flow[i];
      if (flow_ok(u)) m = fmaxf(m, u.x * u.x + u.y * u.y);
    }
  }
  m = block_max(m);
  if (threadIdx.x == 0) scratch[kHeaderWords + blockIdx.x] = __float_as_uint(m);
}

// stage 2: one block folds the partials of THIS call into scratch[0]
__global__ void __launch_bounds__(kThreads) k_na_fold(unsigned* __restrict__ scratch, int n_partials) {
  float m = 0.f;
  for (int j = threadIdx.x; j < n_partials; j += kThreads) m = fmaxf(m, __uint_as_float(scratch[kHeaderWords + j]));
  m = block_max(m);
  if (threadIdx.x == 0) scratch[0] = __float_as_uint(m);
}

// stage 3: compute_color per pixel (same expression order as dis_color.cu's k_color_code)
__global__ void __launch_bounds__(kThreads) k_na_color(const float2* __restrict__ flow, size_t n, float maxmotion,
                                                       const unsigned* __restrict__ scratch, uchar3* __restrict__ bgr) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float maxrad = maxmotion;
  if (!(maxmotion > 0.f)) maxrad = fmaxf(1.f, sqrtf(__uint_as_float(scratch[0])));
  const float2 u = flow[i];
  uchar3 pix = make_uchar3(0, 0, 0);
  if (flow_ok(u)) {
    const float fx = u.x / maxrad, fy = u.y / maxrad;
    const float rad = sqrtf(fx * fx + fy * fy);
    const float a = (float)atan2((double)-fy, (double)-fx) / 3.14159265358979323846f;
    const float fk = (a + 1.0f) / 2.0f * 54.f;
    const int k0 = (int)fk;
    const int k1 = (k0 + 1) % 55;
    const float f = fk - (float)k0;
    unsigned char c[3];
#pragma unroll
    for (int b = 0; b < 3; ++b) {
      const float col0 = (float)wheel(k0, b) / 255.f;
      const float col1 = (float)wheel(k1, b) / 255.f;
      float col = (1 - f) * col0 + f * col1;
      if (rad <= 1) col = 1 - rad * (1 - col);
      else col *= .75f;
      c[b] = (unsigned char)(int)(255.f * col);
    }
    pix = make_uchar3(c[2], c[1], c[0]);
  }
  bgr[i] = pix;
}

// device address of the scratch array, resolved once per device with that device current.  In host code the name
// g_noalloc_scratch is only a stand-in for the symbol, never a device address: kernels get the resolved pointer.
std::mutex g_na_mutex;
unsigned* g_na_ptr[64] = {nullptr};

int scratch_for(int device, unsigned** out) {
  std::lock_guard<std::mutex> g(g_na_mutex);
  if (!g_na_ptr[device]) {
    unsigned* p = nullptr;
    const cudaError_t rc = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (rc != cudaSuccess || !p) {
      cudaGetLastError();
      return dis::set_last_error(DIS_ERR_CUDA, (std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(rc)).c_str());
    }
    g_na_ptr[device] = p;
  }
  *out = g_na_ptr[device];
  return DIS_OK;
}

}  // namespace

extern "C" int dis_color_code_device_noalloc(int device, const float* d_flow, int w, int h, float maxmotion,
                                             uint8_t* d_bgr_out, void* cuda_stream) {
  if (!d_flow || !d_bgr_out) return dis::set_last_error(DIS_ERR_PARAM, "dis_color_code_device_noalloc: null buffer");
  if (w <= 0 || h <= 0) return dis::set_last_error(DIS_ERR_PARAM, "dis_color_code_device_noalloc: empty flow field");
  if (device < 0 || device >= 64) return dis::set_last_error(DIS_ERR_PARAM, "dis_color_code_device_noalloc: device ordinal out of range");
  const size_t npx = (size_t)w * (size_t)h;
  if (npx > (size_t)DIS_NOALLOC_MAX_PIXELS) {
    char msg[192];
    snprintf(msg, sizeof msg, "dis_color_code_device_noalloc: %d x %d = %zu pixels exceeds the fixed scratch (DIS_NOALLOC_MAX_PIXELS = %d); use dis_color_code",
             w, h, npx, (int)DIS_NOALLOC_MAX_PIXELS);
    return dis::set_last_error(DIS_ERR_PARAM, msg);
  }
  int prev = -1;
  cudaError_t rc = cudaGetDevice(&prev);
  if (rc == cudaSuccess && prev != device) rc = cudaSetDevice(device);
  if (rc != cudaSuccess) {
    cudaGetLastError();
    return dis::set_last_error(DIS_ERR_CUDA, (std::string("dis_color_code_device_noalloc: cudaSetDevice: ") + cudaGetErrorString(rc)).c_str());
  }
  unsigned* scratch = nullptr;
  int st = scratch_for(device, &scratch);
  if (st == DIS_OK) {
    cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
    const float2* flow = reinterpret_cast<const float2*>(d_flow);
    if (!(maxmotion > 0.f)) {
      const int n_partials = (int)((npx + kPxPerBlock - 1) / kPxPerBlock);
      k_na_partials<<<n_partials, kThreads, 0, s>>>(flow, npx, scratch);
      k_na_fold<<<1, kThreads, 0, s>>>(scratch, n_partials);
    }
    k_na_color<<<(unsigned)((npx + kThreads - 1) / kThreads), kThreads, 0, s>>>(flow, npx, maxmotion, scratch,
                                                                                 reinterpret_cast<uchar3*>(d_bgr_out));
    rc = cudaGetLastError();
    if (rc != cudaSuccess)
      st = dis::set_last_error(DIS_ERR_CUDA, (std::string("dis_color_code_device_noalloc: launch: ") + cudaGetErrorString(rc)).c_str());
  }
  if (prev >= 0 && prev != device) cudaSetDevice(prev);
  return st;
}

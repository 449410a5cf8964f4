// Micro-probe (development aid): issue rate of packed FADD2/FMUL2 (add.rn.f32x2 / mul.rn.f32x2) against
// scalar FADD/FMUL on sm_100a.  Prints cycles per instruction per SM sub-partition for 1..8 warps per SMSP.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float add1(float a, float b) { float r; asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float mul1(float a, float b) { float r; asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
constexpr int N = 2048, C = 8;
template <int MODE>
__global__ void k(float* out, long long* cyc, float seed) {
  float s[C]; u64 p[C];
  for (int i = 0; i < C; ++i) { s[i] = seed + i + threadIdx.x; float2 f = make_float2(s[i], s[i] + 1); p[i] = *reinterpret_cast<u64*>(&f); }
  const u64 inc2 = p[0]; const float inc = seed;
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < N; ++it) {
#pragma unroll
    for (int i = 0; i < C; ++i) {
      if (MODE == 0) s[i] = add1(s[i], inc);
      if (MODE == 1) p[i] = add2(p[i], inc2);
      if (MODE == 2) s[i] = mul1(s[i], inc);
      if (MODE == 3) p[i] = mul2(p[i], inc2);
      if (MODE == 4) { s[i] = add1(s[i], inc); p[i] = add2(p[i], inc2); }
    }
  }
  long long t1 = clock64();
  float acc = 0; for (int i = 0; i < C; ++i) { float2 f = *reinterpret_cast<float2*>(&p[i]); acc += s[i] + f.x + f.y; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 1 << 20); cudaMallocManaged(&cyc, 8);
  const char* names[5] = {"FADD", "FADD2", "FMUL", "FMUL2", "FADD+FADD2"};
  for (int mode = 0; mode < 5; ++mode)
    for (int wps = 1; wps <= 8; wps *= 2) {   // warps per SMSP (block = 4*wps warps on one SM)
      const int threads = 128 * wps;
      for (int rep = 0; rep < 2; ++rep) {
        if (mode == 0) k<0><<<1, threads>>>(out, cyc, 1.0f);
        if (mode == 1) k<1><<<1, threads>>>(out, cyc, 1.0f);
        if (mode == 2) k<2><<<1, threads>>>(out, cyc, 1.0f);
        if (mode == 3) k<3><<<1, threads>>>(out, cyc, 1.0f);
        if (mode == 4) k<4><<<1, threads>>>(out, cyc, 1.0f);
        cudaDeviceSynchronize();
      }
      const double insts = (double)N * C * (mode == 4 ? 2 : 1) * wps;   // per SMSP
      printf("%-11s warps/SMSP %d: %lld cycles, %.3f cycles per warp-instruction per SMSP\n", names[mode], wps, *cyc, *cyc / insts);
    }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}

// Micro-probe (development aid): how long do __nanosleep(t) and a timed-out mbarrier.try_wait take on
// this GPU, and what is the shared-memory flag ping-pong latency between two warps of a block?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_sleep(long long* out) {
  const unsigned ts[6] = {20, 40, 100, 200, 1000, 4000};
  for (int i = 0; i < 6; ++i) {
    long long best = 1 << 30, sum = 0;
    for (int r = 0; r < 64; ++r) {
      long long t0 = clock64();
      __nanosleep(ts[i]);
      long long t1 = clock64();
      best = min(best, t1 - t0); sum += t1 - t0;
    }
    if (threadIdx.x == 0) { out[2 * i] = best; out[2 * i + 1] = sum / 64; }
  }
}
__global__ void k_trywait(long long* out) {
  __shared__ unsigned long long bar;
  if (threadIdx.x == 0) asm volatile("mbarrier.init.shared.b64 [%0], 1;" ::"r"((unsigned)__cvta_generic_to_shared(&bar)));
  __syncthreads();
  const unsigned hints[4] = {100, 500, 2000, 10000};
  for (int i = 0; i < 4; ++i) {
    long long sum = 0;
    for (int r = 0; r < 32; ++r) {
      long long t0 = clock64();
      unsigned ok;
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared.b64 p, [%1], %2, %3; selp.u32 %0, 1, 0, p; }"
                   : "=r"(ok) : "r"((unsigned)__cvta_generic_to_shared(&bar)), "r"(0u), "r"(hints[i]) : "memory");
      long long t1 = clock64();
      sum += t1 - t0;
    }
    if (threadIdx.x == 0) out[i] = sum / 32;
  }
}
__global__ void k_pingpong(long long* out) {
  __shared__ volatile int f[2];
  if (threadIdx.x == 0) { f[0] = 0; f[1] = 0; }
  __syncthreads();
  const int w = threadIdx.x >> 5;
  long long t0 = clock64();
  for (int i = 1; i <= 1000; ++i) {
    if (w == 0) { if ((threadIdx.x & 31) == 0) f[0] = i; while (f[1] != i) {} }
    else { while (f[0] != i) {} if ((threadIdx.x & 31) == 0) f[1] = i; }
    __syncwarp();
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) out[0] = (t1 - t0) / 1000;
}
int main() {
  long long* d; cudaMalloc(&d, 256); long long h[16];
  k_sleep<<<1, 32>>>(d); cudaMemcpy(h, d, 96, cudaMemcpyDeviceToHost);
  const int ts[6] = {20, 40, 100, 200, 1000, 4000};
  for (int i = 0; i < 6; ++i) printf("nanosleep(%d): min %lld avg %lld cycles\n", ts[i], h[2 * i], h[2 * i + 1]);
  k_trywait<<<1, 32>>>(d); cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
  const int hs[4] = {100, 500, 2000, 10000};
  for (int i = 0; i < 4; ++i) printf("try_wait timeout hint %d ns: avg %lld cycles\n", hs[i], h[i]);
  k_pingpong<<<1, 64>>>(d); cudaMemcpy(h, d, 8, cudaMemcpyDeviceToHost);
  printf("smem flag ping-pong round trip: %lld cycles\n", h[0]);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}

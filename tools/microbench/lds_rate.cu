// Shared-memory load throughput per SM by access width and pattern (one CTA per SM, 8 warps).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 lds_rate.cu -o lds_rate
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256, 1) lds_kernel(int iters, long long* cycles, float* sink) {
  __shared__ __align__(1024) unsigned char buf[32768];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < 32768 / 4; i += 256) reinterpret_cast<float*>(buf)[i] = 1.0f;
  __syncthreads();
  unsigned base = static_cast<unsigned>(__cvta_generic_to_shared(buf)) + warp * 4096;
  unsigned addr;
  if (MODE == 0) addr = base + lane * 16;                                   // LDS.128, 512 contiguous bytes
  if (MODE == 1) addr = base + (lane >> 3) * 256 + (((lane & 7) ^ (lane >> 3)) << 4);  // LDS.128, one 128 B row per 8 lanes, XOR-permuted chunks
  if (MODE == 2) addr = base + lane * 8;                                    // LDS.64 contiguous
  if (MODE == 3) addr = base + lane * 4;                                    // LDS.32 contiguous
  if (MODE == 4) addr = base;                                               // LDS.128 broadcast
  if (MODE == 5) addr = base + (lane & 7) * 16 + (lane >> 3) * 128 * 3;     // LDS.128 rows 3 apart
  float acc = 0.f;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const unsigned a = addr + ((it + u) & 3) * 512;
      if (MODE == 2) { float x, y; asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(x), "=f"(y) : "r"(a)); acc += x + y; }
      else if (MODE == 3) { float x; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(a)); acc += x; }
      else { float x, y, z, w; asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x), "=f"(y), "=f"(z), "=f"(w) : "r"(a)); acc += x + y + z + w; }
    }
  }
  __syncthreads();
  long long t1 = clock64();
  if (tid == 0 && blockIdx.x == 0) cycles[0] = t1 - t0;
  if (acc == 12345.f) sink[0] = acc;
}

int main() {
  long long* d_c; float* d_s; cudaMalloc(&d_c, 8); cudaMalloc(&d_s, 4);
  const int iters = 4000;
  const char* names[] = {"LDS.128 contiguous 512 B/warp", "LDS.128 row per 8 lanes, XOR chunks", "LDS.64 contiguous", "LDS.32 contiguous", "LDS.128 broadcast", "LDS.128 rows 384 B apart"};
  const int bytes[] = {512, 512, 256, 128, 512, 512};
  for (int m = 0; m < 6; ++m) {
    for (int rep = 0; rep < 2; ++rep) {
      switch (m) {
        case 0: lds_kernel<0><<<148, 256>>>(iters, d_c, d_s); break;
        case 1: lds_kernel<1><<<148, 256>>>(iters, d_c, d_s); break;
        case 2: lds_kernel<2><<<148, 256>>>(iters, d_c, d_s); break;
        case 3: lds_kernel<3><<<148, 256>>>(iters, d_c, d_s); break;
        case 4: lds_kernel<4><<<148, 256>>>(iters, d_c, d_s); break;
        case 5: lds_kernel<5><<<148, 256>>>(iters, d_c, d_s); break;
      }
      cudaError_t e = cudaDeviceSynchronize();
      long long c = 0; cudaMemcpy(&c, d_c, 8, cudaMemcpyDeviceToHost);
      const double per = (double)c / (iters * 8.0 * 8.0);   // cycles per warp-instruction per SM
      if (rep) printf("%-40s: %.2f cycles / warp-instr -> %.1f B/clk/SM [%s]\n", names[m], per, bytes[m] / per, cudaGetErrorString(e));
    }
  }
  return 0;
}

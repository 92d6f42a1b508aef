// Where does a stem CTA spend a tile?  Builds csrc/stem.cu with -DB2I_STEM_TRACE (clock64 stamps of
// CTA 0's loader / MMA / drain roles) and prints the per-tile timeline of a batch-256 launch.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DB2I_STEM_TRACE -I../../distributed-machine-learning-cluster_b200/csrc \
//        stem_trace.cu ../../distributed-machine-learning-cluster_b200/csrc/stem.cu ../../distributed-machine-learning-cluster_b200/csrc/conv_umma.cu -o stem_trace -lcuda
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "conv.h"
namespace b2i { cudaError_t stem_trace_read(long long* host); cudaError_t stem_cta_read(long long* host); }
using namespace b2i;

int main() {
  const int N = 256;
  uint8_t* d_in; void* d_out; void* d_w; float* d_bias;
  cudaMalloc(&d_in, (size_t)N * 150528); cudaMalloc(&d_out, (size_t)N * 56 * 56 * 64 * 2);
  cudaMalloc(&d_w, 64 * 256 * 2); cudaMalloc(&d_bias, 256);
  unsigned int* d_sched; cudaMalloc(&d_sched, 8); cudaMemset(d_sched, 0, 8);
  std::vector<uint8_t> h((size_t)N * 150528);
  for (auto& v : h) v = rand() & 255;
  cudaMemcpy(d_in, h.data(), h.size(), cudaMemcpyHostToDevice);
  cudaMemset(d_w, 0, 64 * 256 * 2); cudaMemset(d_bias, 0, 256);
  static float h_bias[64];
  CUtensorMap tm; const char* err = nullptr;
  if (make_tmap_weights(&tm, d_w, 64, 256, 64, &err)) { printf("tmap: %s\n", err); return 1; }
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int rep = 0; rep < 200; ++rep) {
    cudaEventRecord(e0);
    cudaError_t e = launch_stem(d_in, d_out, &tm, h_bias, N, d_sched, /*f16=*/0, 0);
    cudaEventRecord(e1);
    cudaError_t s = cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    if (rep < 3 || rep >= 197) printf("rep %d: %.4f ms [%s / %s]\n", rep, ms, cudaGetErrorString(e), cudaGetErrorString(s));
  }
#ifdef B2I_STEM_TRACE
  static long long t[64][12];
  stem_trace_read(&t[0][0]);
  const long long z = t[0][0];
  printf("tile | loader: start  empty-ok stored | mma: ready  issued | drain: accfull tmem-read staged bar1 pooled bar2   (cycles since CTA 0's first stamp)\n");
  for (int i = 0; i < 56; ++i) {
    printf("%4d |", i);
    for (int k = 0; k < 11; ++k) { if (k == 3 || k == 5) printf(" |"); printf(" %7lld", t[i][k] - z); }
    printf("\n");
  }
  printf("tiles 0..55 of CTA 0: %lld cycles in %lld ns -> %.0f MHz\n", t[55][10] - t[0][10], t[55][11] - t[0][11], 1e3 * (t[55][10] - t[0][10]) / (double)(t[55][11] - t[0][11]));
  static long long c[512][2];
  stem_cta_read(&c[0][0]);
  long long t_first = c[0][0], t_last_start = 0, t_first_end = 1LL << 62, t_end = 0;
  for (int i = 0; i < 296; ++i) {
    if (c[i][0] < t_first) t_first = c[i][0];
    if (c[i][0] > t_last_start) t_last_start = c[i][0];
    if (c[i][1] < t_first_end) t_first_end = c[i][1];
    if (c[i][1] > t_end) t_end = c[i][1];
  }
  printf("CTAs: first start 0, last start %+lld ns, first exit %+lld ns, last exit %+lld ns; CTA 0: %lld..%lld; CTA 0 first tile stamp %+lld ns\n",
         t_last_start - t_first, t_first_end - t_first, t_end - t_first, c[0][0] - t_first, c[0][1] - t_first, t[0][11] - t_first);
  printf("exit time per CTA (us):");
  for (int i = 0; i < 296; ++i) { if (i % 16 == 0) printf("\n%3d:", i); printf(" %5.1f", (c[i][1] - t_first) * 1e-3); }
  printf("\n");
#endif
  return 0;
}

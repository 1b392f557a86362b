// exp_check.cu — exp_ref (gsplat_internal.h: glibc's expf algorithm restated) on the GPU against the
// host libm's expf, bit for bit, over N pseudo-random arguments in the range K5 feeds it.
//   nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -O3 -o exp_check exp_check.cu && ./exp_check
#include <cmath>
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <vector>
#include "../../gaussian_splat_ipu_b200/csrc/gsplat_internal.h"
using namespace gsplat;
__global__ void k(const float *x, float *y, int n) {
  __shared__ unsigned long long tab[32];
  if (threadIdx.x < 32) tab[threadIdx.x] = EXPF_TAB[threadIdx.x];
  __syncthreads();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = exp_ref(x[i], tab);
}
int main() {
  const int N = 1 << 28;
  std::vector<float> x(N), y(N);
  uint64_t st = 88172645463325252ull;
  for (int i = 0; i < N; ++i) {
    st ^= st << 13; st ^= st >> 7; st ^= st << 17;
    float u = (float)((st >> 40) * (1.0 / 16777216.0));
    x[i] = (i & 3) == 0 ? -u * 1e-3f : (i & 3) == 1 ? -u : -12.0f * u;
  }
  float *dx, *dy;
  cudaMalloc(&dx, N * 4); cudaMalloc(&dy, N * 4);
  cudaMemcpy(dx, x.data(), N * 4, cudaMemcpyHostToDevice);
  k<<<(N + 255) / 256, 256>>>(dx, dy, N);
  cudaMemcpy(y.data(), dy, N * 4, cudaMemcpyDeviceToHost);
  long bad = 0;
  for (int i = 0; i < N; ++i) {
    float h = expf(x[i]);
    uint32_t a, b; memcpy(&a, &h, 4); memcpy(&b, &y[i], 4);
    if (a != b) { if (bad < 8) printf("x %a (%g): host %a gpu %a\n", x[i], x[i], h, y[i]); ++bad; }
  }
  printf("exp_ref vs host expf: %ld of %d differ (%s)\n", bad, N, cudaGetErrorString(cudaGetLastError()));
  return 0;
}

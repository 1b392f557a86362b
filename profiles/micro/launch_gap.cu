// launch-gap microbenchmark: dependent tiny kernels in one stream, plain vs programmatic dependent launch
#include <cstdio>
#include <cuda_runtime.h>
__global__ void tiny(int *p) { if (threadIdx.x == 0 && blockIdx.x == 0) p[0] += 1; }
__global__ void tiny_pdl(int *p) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (threadIdx.x == 0 && blockIdx.x == 0) p[0] += 1;
}
__global__ void tiny_pdl2(int *p) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (threadIdx.x == 0 && blockIdx.x == 0) p[0] += 1;
}
int main() {
  int *d; cudaMalloc(&d, 4); cudaMemset(d, 0, 4);
  cudaStream_t st; cudaStreamCreate(&st);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  const int K = 2000;
  for (int grid : {1, 148, 444}) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(a, st);
      for (int i = 0; i < K; ++i) tiny<<<grid, 256, 0, st>>>(d);
      cudaEventRecord(b, st); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b);
      if (rep) printf("plain grid %d: %.2f us per launch\n", grid, ms * 1e3 / K);
    }
    for (int rep = 0; rep < 2; ++rep) {
      cudaLaunchConfig_t cfg = {}; cfg.gridDim = grid; cfg.blockDim = 256; cfg.stream = st;
      cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[0].val.programmaticStreamSerializationAllowed = 1; cfg.attrs = at; cfg.numAttrs = 1;
      cudaEventRecord(a, st);
      for (int i = 0; i < K; ++i) cudaLaunchKernelEx(&cfg, tiny_pdl, d);
      cudaEventRecord(b, st); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b);
      if (rep) printf("pdl   grid %d: %.2f us per launch\n", grid, ms * 1e3 / K);
    }
    for (int rep = 0; rep < 2; ++rep) {
      cudaLaunchConfig_t cfg = {}; cfg.gridDim = grid; cfg.blockDim = 256; cfg.stream = st;
      cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[0].val.programmaticStreamSerializationAllowed = 1; cfg.attrs = at; cfg.numAttrs = 1;
      cudaEventRecord(a, st);
      for (int i = 0; i < K; ++i) cudaLaunchKernelEx(&cfg, tiny_pdl2, d);
      cudaEventRecord(b, st); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b);
      if (rep) printf("pdl+early trigger grid %d: %.2f us per launch\n", grid, ms * 1e3 / K);
    }
  }
  int h; cudaMemcpy(&h, d, 4, cudaMemcpyDeviceToHost); printf("count %d err %s\n", h, cudaGetErrorString(cudaGetLastError()));
}

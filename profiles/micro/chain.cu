// Microbenchmark: where does a chained-scan compaction kernel (select.cu / duplicate.cu pattern) spend
// its time?   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I gaussian_splat_ipu_b200/csrc -o /tmp/chain profiles/micro/chain.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "gsplat_internal.h"
using namespace gsplat;
#ifndef CT
#define CT 256
#endif
#ifndef CI
#define CI 16
#endif
constexpr int T = CT, ITEMS = CI, CHUNK = T * ITEMS, NWC = ITEMS * (T / 32), PER = NWC / 32;
// MODE 0: ticket + chain; 1: ticket, no chain; 2: blockIdx order + chain; 3: no ticket, no chain
template <int MODE>
__global__ void __launch_bounds__(T) compact(const uint32_t *__restrict__ key, uint32_t n, uint32_t tau,
                                             uint32_t *__restrict__ out, unsigned long long *lookback, uint32_t *ticket,
                                             uint32_t *total_out) {
  __shared__ uint32_t wc[NWC + 1];
  __shared__ uint32_t chain_s;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (MODE == 0 || MODE == 1) {
    if (tid == 0) chain_s = atomicAdd(ticket, 1u);
    __syncthreads();
  }
  const uint32_t b = (MODE == 0 || MODE == 1) ? chain_s : blockIdx.x;
  const uint32_t base = b * CHUNK;
  uint32_t k[ITEMS]; unsigned long long sel = 0;
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) { const uint32_t j = base + i * T + tid; k[i] = j < n ? key[j] : 0xffffffffu; }
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const bool take = k[i] < tau;
    const uint32_t m = __ballot_sync(0xffffffffu, take);
    if (take) sel |= 1ull << i;
    if (lane == 0) wc[i * (T / 32) + warp] = __popc(m);
  }
  __syncthreads();
  if (warp == 0) {
    uint32_t v[PER], sum = 0;
#pragma unroll
    for (int q = 0; q < PER; ++q) { v[q] = wc[lane * PER + q]; sum += v[q]; }
    uint32_t inc = sum;
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    uint32_t run = inc - sum;
#pragma unroll
    for (int q = 0; q < PER; ++q) { wc[lane * PER + q] = run; run += v[q]; }
    if (lane == 31) wc[NWC] = inc;
  }
  __syncthreads();
  const uint32_t total = wc[NWC];
  if (warp == 0) {
    unsigned long long excl;
    if (MODE == 0 || MODE == 2) excl = lb_exclusive_prefix(lookback, b, total, lane);
    else excl = (unsigned long long)b * 330;
    if (lane == 0) { chain_s = (uint32_t)excl; if (b == gridDim.x - 1) *total_out = (uint32_t)(excl + total); }
  }
  __syncthreads();
  const uint32_t cbase = chain_s;
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const bool take = sel >> i & 1ull;
    const uint32_t m = __ballot_sync(0xffffffffu, take);
    if (take) out[cbase + wc[i * (T / 32) + warp] + __popc(m & ((1u << lane) - 1u))] = k[i];
  }
}
template <int MODE> float run(const uint32_t *key, uint32_t n, uint32_t tau, uint32_t *out, unsigned long long *lb, uint32_t *sc) {
  const int grid = (n + CHUNK - 1) / CHUNK;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9f;
  for (int r = 0; r < 12; ++r) {
    cudaMemsetAsync(lb, 0, sizeof(unsigned long long) * (grid + 1));
    cudaMemsetAsync(sc, 0, 64);
    cudaEventRecord(e0);
    compact<MODE><<<grid, T>>>(key, n, tau, out, lb, sc, sc + 1);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (r > 1 && ms < best) best = ms;
  }
  uint32_t tot; cudaMemcpy(&tot, sc + 1, 4, cudaMemcpyDeviceToHost);
  printf("  mode %d: %.1f us (selected %u)%s\n", MODE, best * 1e3f, tot, cudaGetLastError() ? " ERR" : "");
  return best;
}
int main(int argc, char **argv) {
  const uint32_t n = argc > 1 ? atoi(argv[1]) : 6000000;
  const double frac = argc > 2 ? atof(argv[2]) : 0.08;
  std::vector<uint32_t> h(n);
  uint64_t s = 88172645463325252ull;
  for (auto &x : h) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; x = (uint32_t)(s >> 33); }
  uint32_t *key, *out, *sc; unsigned long long *lb;
  cudaMalloc(&key, 4ull * n); cudaMalloc(&out, 4ull * n + 4096 * 400); cudaMalloc(&sc, 64); cudaMalloc(&lb, 8ull * (n / CHUNK + 2));
  cudaMemcpy(key, h.data(), 4ull * n, cudaMemcpyHostToDevice);
  const uint32_t tau = (uint32_t)(frac * 2147483648.0);
  printf("n %u frac %.2f window %d T %d items %d links %d\n", n, frac, 32 * LB_WINDOW, T, ITEMS, (n + CHUNK - 1) / CHUNK);
  run<0>(key, n, tau, out, lb, sc); run<1>(key, n, tau, out, lb, sc);
  return 0;
}

// warp_op_probe.cu — issue rates of the warp-wide operations the radix ranking is made of (match.any, ballot, shuffle),
// measured on the GPU at hand: warp-instructions per cycle per SM sub-partition, for 1 .. 8 warps per sub-partition and,
// for match.any, for 1 / 8 / 32 distinct values per warp.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/warp_op_probe.cu -o tools/bin/warp_op_probe
#include <cstdio>
#include <cuda_runtime.h>

#define ITER 2048
// OP 0: match.any   1: eight ballots + eight and/not (what replaces one match.any on an 8-bit digit)   2: shfl.idx   3: one ballot
template <int OP>
__global__ void probe(unsigned *out, unsigned distinct_mask, unsigned one, long long *cycles) {
  unsigned x[4];
#pragma unroll
  for (int i = 0; i < 4; i++) x[i] = (threadIdx.x * 2654435761u + i) & distinct_mask;
  long long t0 = clock64();
  for (int it = 0; it < ITER; it++) {
#pragma unroll
    for (int i = 0; i < 4; i++) {
      if (OP == 0) x[i] = (__match_any_sync(0xffffffffu, x[i]) * one + threadIdx.x) & distinct_mask;
      if (OP == 1) {
        unsigned m = 0xffffffffu;
#pragma unroll
        for (int b = 0; b < 8; b++) { const bool bit = (x[i] >> b) & 1u; const unsigned v = __ballot_sync(0xffffffffu, bit); m &= bit ? v : ~v; }
        x[i] = (m * one + threadIdx.x) & distinct_mask;
      }
      if (OP == 2) x[i] = __shfl_sync(0xffffffffu, x[i] + one, (int)(x[i] & 31u));
      if (OP == 3) x[i] = (__ballot_sync(0xffffffffu, x[i] & 1u) * one + threadIdx.x) & distinct_mask;
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = x[0] + x[1] + x[2] + x[3];
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int OP>
void run(const char *name, unsigned distinct_mask, int warps_per_smsp) {
  unsigned *out; long long *cyc, h;
  const int threads = 32 * 4 * warps_per_smsp;
  cudaMalloc(&out, 148 * threads * 4); cudaMalloc(&cyc, 8);
  probe<OP><<<148, threads>>>(out, distinct_mask, 1u, cyc);
  probe<OP><<<148, threads>>>(out, distinct_mask, 1u, cyc);
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  const double ops = (double)ITER * 4 * warps_per_smsp;                  // probed operations per SMSP
  printf("%-44s warps/SMSP %d: %8.2f cycles per operation per SMSP\n", name, warps_per_smsp, h / ops);
  cudaFree(out); cudaFree(cyc);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  printf("# %s, %d SMs; cycles are clock64 ticks of one SM, a chain of 4 independent operations per warp\n", p.name, p.multiProcessorCount);
  for (int w : {1, 2, 4, 8}) {
    run<0>("match.any, 1 distinct value per warp", 0x0u, w);
    run<0>("match.any, <= 8 distinct values per warp", 0x7u, w);
    run<0>("match.any, <= 32 distinct values (8-bit digit)", 0xffu, w);
    run<1>("8 x (ballot + and) on an 8-bit digit", 0xffu, w);
    run<3>("one ballot", 0xffu, w);
    run<2>("shfl.idx", 0xffu, w);
  }
  return 0;
}

// pipe_probe.cu — issue rates of the integer instructions the DP kernels are made of, measured on
// the GPU at hand (B200): warp-instructions per cycle per SM sub-partition for each op alone and
// for the ALU/FMA mixes.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/pipe_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

#define ITER 4096
template <int OP>
__global__ void probe(int *out, int a0, int b0, int c0, int one, long long *cycles) {
  int x[8];
#pragma unroll
  for (int i = 0; i < 8; i++) x[i] = a0 + i + threadIdx.x;
  int b = b0 + threadIdx.x, c = c0;
  long long t0 = clock64();
  for (int it = 0; it < ITER; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (OP == 0) x[i] = __viaddmax_s32(x[i], b, c);                 // VIADDMNMX
      if (OP == 1) x[i] = __vimax3_s32(x[i], b, c + i);               // VIMNMX3
      if (OP == 2) x[i] = (x[i] & b) ^ c;                             // LOP3
      if (OP == 3) x[i] = x[i] * one + b;                             // IMAD
      if (OP == 4) x[i] = max(x[i], b + i);                           // VIMNMX / IMNMX
      if (OP == 5) x[i] = __viaddmax_s32_relu(x[i], b, c);            // VIADDMNMX.RELU
      if (OP == 6) { x[i] = __viaddmax_s32(x[i], b, c); x[i] = x[i] * one + b; }           // 1 ALU + 1 FMA
      if (OP == 7) { x[i] = __viaddmax_s32(x[i], b, c); x[i] = (x[i] & b) ^ c; }           // 2 ALU
      if (OP == 8) { x[i] = __vimax3_s32(x[i], b, c + i); x[i] = x[i] * one + b; x[i] = x[i] * one + c; }   // 1 ALU + 2 FMA
      if (OP == 9) x[i] = __funnelshift_r(x[i], b, 2);                // SHF
      if (OP == 10) x[i] = __byte_perm(x[i], b, 0x4441);              // PRMT
      if (OP == 11) x[i] = __vadd2(x[i], b);                          // VIADD.16x2
      if (OP == 12) x[i] = __vimax3_s16x2(x[i], b, c + i);            // VIMNMX3.S16x2
      if (OP == 13) x[i] = __viaddmax_s16x2_relu(x[i], b, c);         // VIADDMNMX.S16x2.RELU
      if (OP == 14) x[i] = __vmaxs2(x[i], b + i);                     // VIMNMX.S16x2
      if (OP == 15) { x[i] = __vimax3_s16x2(x[i], b, c + i); x[i] = x[i] * one + b; }      // packed max3 + IMAD
      if (OP == 16) { x[i] = __vimax3_s16x2(x[i], b, c + i); x[i] = __vadd2(x[i], b); }    // packed max3 + packed add
      if (OP == 17) { x[i] = __vadd2(x[i], b); x[i] = x[i] * one + b; }                    // packed add + IMAD
      if (OP == 18) x[i] = x[i] + b + (c + i);                         // IADD3
      if (OP == 19) x[i] = __vimax3_u16x2(x[i], b, c + i);             // VIMNMX3.U16x2
      if (OP == 20) x[i] = __viaddmax_u16x2(x[i], b, c);               // VIADDMNMX.U16x2
      if (OP == 21) x[i] = __vmaxu2(x[i], b + i);                      // VIMNMX.U16x2
      if (OP == 22) { x[i] = __vimax3_s16x2(x[i], b, c + i); x[i] = x[i] * one + b; x[i] = x[i] * one + c; }   // packed max3 + 2 IMAD
      if (OP == 23) { x[i] = __viaddmax_s16x2_relu(x[i], b, c); x[i] = __vimax3_s16x2(x[i], b, c + i);
                      x[i] = x[i] * one + b; x[i] = x[i] * one + c; }                       // 2 ALU + 2 FMA
      if (OP == 24) { x[i] = __viaddmax_s16x2_relu(x[i], b, c); x[i] = x[i] + b + (c + i); }  // packed add-max + IADD3
    }
  }
  long long t1 = clock64();
  int s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int OP>
void run(const char *name, int per_iter, int warps_per_smsp) {
  int *out; long long *cyc, h;
  const int threads = 32 * 4 * warps_per_smsp;
  cudaMalloc(&out, 148 * threads * 4); cudaMalloc(&cyc, 8);
  probe<OP><<<148, threads>>>(out, 1, 3, 5, 1, cyc);
  probe<OP><<<148, threads>>>(out, 1, 3, 5, 1, cyc);
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  const double inst = (double)ITER * 8 * per_iter * warps_per_smsp;      // warp-instructions per SMSP
  printf("%-34s warps/SMSP %d: %.3f warp-inst/cycle/SMSP (%.2f cycles per inst)\n", name, warps_per_smsp, inst / h, h / inst);
  cudaFree(out); cudaFree(cyc);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  printf("# %s, %d SMs, max SM clock %d MHz; rates are per cycle (clock64), so independent of the clock\n", p.name, p.multiProcessorCount, khz / 1000);
  for (int w : {1, 2, 4, 8}) {
    run<0>("VIADDMNMX", 1, w); run<1>("VIMNMX3", 1, w); run<2>("LOP3", 1, w); run<3>("IMAD", 1, w);
    run<4>("VIMNMX", 1, w); run<5>("VIADDMNMX.RELU", 1, w); run<9>("SHF", 1, w); run<10>("PRMT", 1, w);
    run<11>("VIADD.16x2", 1, w); run<12>("VIMNMX3.S16x2", 1, w); run<13>("VIADDMNMX.S16x2.RELU", 1, w);
    run<14>("VIMNMX.S16x2", 1, w); run<15>("VIMNMX3.S16x2 + IMAD", 2, w); run<16>("VIMNMX3.S16x2 + VIADD.16x2", 2, w);
    run<17>("VIADD.16x2 + IMAD", 2, w);
    run<18>("IADD3", 1, w); run<19>("VIMNMX3.U16x2", 1, w); run<20>("VIADDMNMX.U16x2", 1, w); run<21>("VIMNMX.U16x2", 1, w);
    run<22>("VIMNMX3.S16x2 + 2 IMAD", 3, w); run<23>("2 packed ALU + 2 IMAD", 4, w); run<24>("VIADDMNMX.S16x2.RELU + IADD3", 2, w);
    run<6>("VIADDMNMX + IMAD", 2, w); run<7>("VIADDMNMX + LOP3", 2, w); run<8>("VIMNMX3 + 2 IMAD", 3, w);
  }
  return 0;
}

// fp64_latency.cu -- dependent-issue latency and single-warp throughput of the FP64 pipe on sm_100a.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o tools/_bin/fp64_latency tools/fp64_latency.cu
// One warp per SM sub-partition (128-thread blocks, one block per SM), ILP independent chains per thread, each chain
// a loop of dependent DADD / DMUL / DFMA / (DMUL+DADD).  Prints cycles per instruction per warp: with ILP = 1 this is
// the dependent-issue latency; as ILP grows it falls to the issue interval of the pipe (2 cycles per warp instruction
// if the sub-partition has 16 FP64 lanes).  Also: MUFU.RCP64H + Newton (IEEE division) and DSETP+SEL chains.
#include <cstdio>
#include <cuda_runtime.h>
template <int OP, int ILP>
__global__ void chain(int iters, double seed, double* out, long long* cyc) {
  double a[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) a[i] = seed + threadIdx.x + i;
  const double m = 0.999999999, c = 1e-9;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      if (OP == 0) a[i] = __dadd_rn(a[i], c);
      else if (OP == 1) a[i] = __dmul_rn(a[i], m);
      else if (OP == 2) a[i] = fma(a[i], m, c);
      else if (OP == 3) a[i] = __dadd_rn(__dmul_rn(a[i], m), c);
      else if (OP == 4) a[i] = c / a[i] + 1.5;          // IEEE division + add
      else if (OP == 5) a[i] = sqrt(a[i]) + 1.5;        // IEEE sqrt + add
      else if (OP == 6) a[i] = (a[i] > 1.0) ? __dadd_rn(a[i], -c) : __dadd_rn(a[i], c);  // compare + select + add
    }
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += a[i];
  if (s == 123.456) out[0] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
template <int OP, int ILP>
void run(const char* name, int nops, int warps_per_smsp) {
  double* d; long long* c; cudaMalloc(&d, 8); cudaMalloc(&c, 8);
  const int iters = 20000;
  chain<OP, ILP><<<148, 128 * warps_per_smsp>>>(iters, 1.0, d, c);
  chain<OP, ILP><<<148, 128 * warps_per_smsp>>>(iters, 1.0, d, c);
  long long h = 0; cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
  printf("%-14s ILP %d warps/SMSP %d: %7.2f cycles per op per chain-step, %6.2f cycles per warp-instruction-group\n", name, ILP, warps_per_smsp,
         (double)h / iters, (double)h / iters / (ILP * nops));
  cudaFree(d); cudaFree(c);
}
int main() {
  run<0, 1>("DADD", 1, 1); run<0, 2>("DADD", 1, 1); run<0, 4>("DADD", 1, 1); run<0, 8>("DADD", 1, 1);
  run<1, 1>("DMUL", 1, 1); run<1, 4>("DMUL", 1, 1);
  run<2, 1>("DFMA", 1, 1); run<2, 2>("DFMA", 1, 1); run<2, 4>("DFMA", 1, 1); run<2, 8>("DFMA", 1, 1);
  run<3, 1>("DMUL+DADD", 2, 1); run<3, 4>("DMUL+DADD", 2, 1); run<3, 8>("DMUL+DADD", 2, 1);
  run<4, 1>("DIV+DADD", 1, 1); run<4, 2>("DIV+DADD", 1, 1); run<4, 4>("DIV+DADD", 1, 1);
  run<5, 1>("SQRT+DADD", 1, 1); run<5, 4>("SQRT+DADD", 1, 1);
  run<6, 1>("DSETP+SEL+DADD", 1, 1); run<6, 4>("DSETP+SEL+DADD", 1, 1);
  run<2, 1>("DFMA", 1, 2); run<2, 1>("DFMA", 1, 4); run<2, 1>("DFMA", 1, 8); run<2, 4>("DFMA", 1, 4);
  run<4, 1>("DIV+DADD", 1, 4); run<4, 1>("DIV+DADD", 1, 8);
  return 0;
}

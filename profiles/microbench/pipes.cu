// pipes.cu — which pipe limits a compare-exchange on sm_100a?  Warp-instructions per clock per SM
// for streams of min/max and multiply-add flavours, 8 independent chains per thread.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu ; run on a B200.
#include <cstdint>
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define ITER 4096

template <int OP>
__global__ void __launch_bounds__(256) probe(float* out, float seed, long long* clocks) {
  float f[8]; int n[8]; __half2 h[8]; double d[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    f[i] = seed + threadIdx.x * 0.001f + i; n[i] = (int)(threadIdx.x * 977u + i * 131u + (int)seed);
    h[i] = __floats2half2_rn(f[i], f[i] * 0.5f); d[i] = f[i];
  }
  float g = seed * 0.999f; int m = (int)seed | 1; __half2 hg = __floats2half2_rn(g, -g); double dg = g;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) { f[i] = fminf(f[i], g); f[i] = fmaxf(f[i], -g); }                       // 2 FMNMX
      else if (OP == 1) { f[i] = fminf(f[i], g); f[i] = f[i] * g; }                          // FMNMX + FMUL
      else if (OP == 2) { n[i] = min(n[i], m); n[i] = max(n[i], -m); }                       // 2 VIMNMX
      else if (OP == 3) { const int lo = min(n[i], m); const int s = n[i] * 1 + m;           // int CE: VIMNMX + 2 IMAD
                          asm volatile("mad.lo.s32 %0, %1, -1, %2;" : "=r"(n[i]) : "r"(lo), "r"(s)); m ^= lo; }
      else if (OP == 4) { asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(n[i]) : "r"(m), "r"(i)); asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(n[i]) : "r"(m), "r"(i)); }  // 2 IMAD
      else if (OP == 5) { f[i] = fmaf(f[i], g, g); f[i] = fmaf(f[i], g, -g); }               // 2 FFMA
      else if (OP == 6) { h[i] = __hmin2(h[i], hg); h[i] = __hmax2(h[i], __hneg2(hg)); }     // 2 HMNMX2
      else if (OP == 7) { f[i] = fminf(f[i], g); h[i] = __hmin2(h[i], hg); }                 // FMNMX + HMNMX2
      else if (OP == 8) { const bool p = f[i] < g; f[i] = p ? f[i] : -g; }                   // FSETP + FSEL
      else if (OP == 9) { d[i] = fmin(d[i], dg); d[i] = fmax(d[i], -dg); }                   // fp64 min/max
      else if (OP == 10) { f[i] = fminf(f[i], g); f[i] = fmaxf(f[i], -g); f[i] = f[i] * g; f[i] = fmaf(f[i], g, g); } // 2 FMNMX + 2 FMA-pipe
      else if (OP == 11) { f[i] = fminf(fminf(f[i], g), (float)i); f[i] = fmaxf(fmaxf(f[i], -g), -1.0f); }           // FMNMX3?
    }
  }
  const long long t1 = clock64();
  float acc = 0; int an = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) { acc += f[i] + __low2float(h[i]) + (float)d[i]; an += n[i]; }
  out[blockIdx.x * 256 + threadIdx.x] = acc + an + m;
  if (threadIdx.x == 0) clocks[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int per_chain, int ctas_per_sm, float* out, long long* clk) {
  const int grid = 148 * ctas_per_sm;
  probe<OP><<<grid, 256>>>(out, 1.5f, clk);
  cudaDeviceSynchronize();
  probe<OP><<<grid, 256>>>(out, 2.5f, clk);
  cudaDeviceSynchronize();
  static long long h[148 * 8];
  cudaMemcpy(h, clk, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < grid; ++i) avg += (double)h[i]; avg /= grid;
  const double warp_instr = (double)ITER * 8 * per_chain * 8 /*warps per CTA*/ * ctas_per_sm;
  printf("%-28s warps/SM %2d  %6.2f warp-instr/clk/SM  err=%s\n", name, 8 * ctas_per_sm, warp_instr / avg,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  float* out; long long* clk;
  cudaMalloc(&out, 148 * 8 * 256 * 4); cudaMalloc(&clk, 148 * 8 * 8);
  for (int c : {2, 4}) {
    run<0>("fmnmx x2", 2, c, out, clk);
    run<1>("fmnmx + fmul", 2, c, out, clk);
    run<2>("vimnmx x2", 2, c, out, clk);
    run<3>("vimnmx + 2 imad (int CE)", 3, c, out, clk);
    run<4>("imad x2", 2, c, out, clk);
    run<5>("ffma x2", 2, c, out, clk);
    run<6>("hmnmx2 x2", 2, c, out, clk);
    run<7>("fmnmx + hmnmx2", 2, c, out, clk);
    run<8>("fsetp + fsel", 2, c, out, clk);
    run<9>("fp64 min + max", 2, c, out, clk);
    run<10>("2 fmnmx + fmul + ffma", 4, c, out, clk);
    run<11>("3-input min + 3-input max", 2, c, out, clk);
  }
  return 0;
}

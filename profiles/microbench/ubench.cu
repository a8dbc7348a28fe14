// ubench.cu — instruction-throughput probes that decide the shape of the rank stage.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench ubench.cu ; run on a B200.
// Each probe runs 148*4 CTAs x 256 threads, ITER iterations of an unrolled body, and prints
// warp-instructions per SM per clock (or cycles per warp-instruction) from the elapsed SM clocks.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define ITER 2048

template <int OP>
__global__ void __launch_bounds__(256) probe(uint32_t* out, uint32_t seed, long long* clocks) {
  __shared__ uint32_t sm[4096];
  const uint32_t tid = threadIdx.x, lane = tid & 31;
  for (int i = tid; i < 4096; i += 256) sm[i] = i * 2654435761u;
  __syncthreads();
  uint32_t a = seed + tid * 2654435761u, b = a ^ 0x9e3779b9u, c = 0, d = 1;
  float fa = __uint_as_float((a & 0x007fffff) | 0x3f800000), fb = __uint_as_float((b & 0x007fffff) | 0x3f800000);
  double da = (double)fa, db = (double)fb, dc = 0.0;
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (OP == 0) {           // MATCH.ANY, all lanes distinct
        c += __match_any_sync(0xffffffffu, a + u);
        a += c;
      } else if (OP == 1) {    // MATCH.ANY, 4 distinct values per warp
        c += __match_any_sync(0xffffffffu, (a + u) & 3u);
        a += (c & 1);
      } else if (OP == 2) {    // SHFL.IDX with a data-dependent source lane
        a = __shfl_sync(0xffffffffu, a, (a + u) & 31);
      } else if (OP == 3) {    // LDS.32 random bank
        a = sm[(a + u) & 4095];
      } else if (OP == 4) {    // LDS.32 conflict-free
        a = sm[((a & 127u) * 32u + lane + u) & 4095];
      } else if (OP == 5) {    // FSETP + select chain (ALU pipe)
        c += (fa < fb) ? 1u : 0u; fa += 1.0f; fb = fb * 1.0001f;
      } else if (OP == 6) {    // FMNMX pair (compare-exchange)
        const float lo = fminf(fa, fb), hi = fmaxf(fa, fb); fa = lo + 1.0f; fb = hi;
      } else if (OP == 7) {    // DFMA
        dc = fma(da, db, dc); da += 1.0;
      } else if (OP == 8) {    // F2F.F64.F32 + DADD
        dc += (double)fa; fa += 1.0f;
      } else if (OP == 9) {    // ATOMS.ADD random address
        atomicAdd(&sm[(a + u * 977) & 4095], 1u); a = a * 1664525u + 1013904223u;
      } else if (OP == 10) {   // STS.U16 + LDS.U16 random (hash election)
        reinterpret_cast<uint16_t*>(sm)[(a + u) & 8191] = (uint16_t)tid;
        a += reinterpret_cast<uint16_t*>(sm)[(b + u) & 8191]; b = b * 1664525u + 1013904223u;
      } else if (OP == 11) {   // VOTE.BALLOT + POPC
        c += __popc(__ballot_sync(0xffffffffu, (a >> u) & 1)); a += c;
      } else if (OP == 12) {   // IMAD hash + shift
        a = (a * 0x9E3779B1u) >> 3; a += u;
      } else if (OP == 13) {   // LDS.128 random 16-byte node
        const uint4 v = reinterpret_cast<const uint4*>(sm)[(a + u) & 1023]; a = v.x + v.w;
      } else if (OP == 14) {   // REDUX (warp reduce add)
        a += __reduce_add_sync(0xffffffffu, a + u);
      } else if (OP == 15) {   // ATOMS.ADD, the 32 lanes of a warp on ONE address (heavy ties)
        atomicAdd(&sm[((tid >> 5) * 64 + u) & 4095], 1u);
      } else if (OP == 16) {   // ATOMS.ADD, 4 addresses per warp
        atomicAdd(&sm[((tid >> 5) * 64 + u * 4 + (lane & 3)) & 4095], 1u);
      } else if (OP == 17) {   // ATOMS.CAS, one address per warp
        a += atomicCAS(&sm[((tid >> 5) * 64 + u) & 4095], a, b);
      } else if (OP == 18) {   // ATOMS.CAS random address
        a += atomicCAS(&sm[(a + u * 977) & 4095], a, b); a = a * 1664525u + 1013904223u;
      } else if (OP == 19) {   // STS of one value to one address from every lane, then LDS of it (election)
        sm[((tid >> 5) * 64 + u) & 4095] = a; a += sm[((tid >> 5) * 64 + u) & 4095];
      } else if (OP == 20) {   // ATOMS.ADD returning, one address per warp
        a += atomicAdd(&sm[((tid >> 5) * 64 + u) & 4095], 1u);
      }
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * 256 + tid] = a + b + c + d + __float_as_uint(fa) + __float_as_uint(fb) + (uint32_t)(dc + da + db);
  if (tid == 0) clocks[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int instr_per_body, uint32_t* out, long long* clk) {
  const int grid = 148 * 4;
  probe<OP><<<grid, 256>>>(out, 12345u, clk);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  probe<OP><<<grid, 256>>>(out, 777u, clk);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  static long long h[148 * 4];
  cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < grid; ++i) avg += (double)h[i]; avg /= grid;
  // per SM: 4 CTAs x 8 warps resident, each did ITER*8 bodies
  const double warp_bodies_per_sm = 4.0 * 8.0 * ITER * 8.0;
  printf("%-34s  %8.3f ms  %10.0f clk/CTA  %7.2f clk per warp-body per SM  (x%d instr)  err=%s\n", name, ms, avg,
         avg / warp_bodies_per_sm, instr_per_body, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  uint32_t* out; long long* clk;
  cudaMalloc(&out, 148 * 4 * 256 * 4); cudaMalloc(&clk, 148 * 4 * 8);
  run<0>("match.any distinct", 1, out, clk);
  run<1>("match.any 4 values", 1, out, clk);
  run<2>("shfl.idx dependent", 1, out, clk);
  run<3>("lds.32 random", 1, out, clk);
  run<4>("lds.32 conflict-free", 1, out, clk);
  run<5>("fsetp+sel+fadd+fmul", 4, out, clk);
  run<6>("fmnmx pair + fadd", 3, out, clk);
  run<7>("dfma + dadd", 2, out, clk);
  run<8>("f2f.f64.f32 + dadd + fadd", 3, out, clk);
  run<9>("atoms.add random", 1, out, clk);
  run<10>("sts.u16 + lds.u16 random", 2, out, clk);
  run<11>("ballot + popc", 2, out, clk);
  run<12>("imad + shf", 2, out, clk);
  run<13>("lds.128 random", 1, out, clk);
  run<14>("redux.add", 1, out, clk);
  run<15>("atoms.add one address per warp", 1, out, clk);
  run<16>("atoms.add 4 addresses per warp", 1, out, clk);
  run<17>("atoms.cas one address per warp", 1, out, clk);
  run<18>("atoms.cas random", 1, out, clk);
  run<19>("sts + lds one address per warp", 2, out, clk);
  run<20>("atoms.add (returning) one address", 1, out, clk);
  return 0;
}

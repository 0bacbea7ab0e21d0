// Micro-benchmark: issue rate of integer instructions on sm_100a (cycles per warp instruction
// per SM sub-partition), to balance k_scan_wt's per-op work between the ALU and FMA pipes.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/pipe_bench tools/pipe_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITER 4096
#define NACC 8

template <int OP>
__global__ void __launch_bounds__(256) k(uint32_t* out, uint32_t a, uint32_t b, long long* cyc) {
  uint32_t x[NACC];
  unsigned long long y[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { x[i] = threadIdx.x * 7 + i + a; y[i] = x[i]; }
  long long t0 = clock64();
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) {
      if (OP == 0) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
      if (OP == 1) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(a));
      if (OP == 2) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
      if (OP == 3) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(y[i]) : "r"(x[i]), "r"(a));
      if (OP == 4) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
      if (OP == 5) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(a), "r"(b));
      if (OP == 6) asm volatile("shr.u32 %0, %0, 1;" : "+r"(x[i]));
      if (OP == 7) asm volatile("max.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(a));
      if (OP == 8) {  // mix: 1 prmt + 1 mad.lo
        asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[(i + 1) % NACC]) : "r"(a), "r"(b));
      }
      if (OP == 9) {  // mix: 1 prmt + 1 mul.hi
        asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
        asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(x[(i + 1) % NACC]) : "r"(a));
      }
      if (OP == 10) asm volatile("bfe.u32 %0, %0, 4, 28;" : "+r"(x[i]));
      if (OP == 11) asm volatile("shf.r.wrap.b32 %0, %0, %0, 4;" : "+r"(x[i]));
      if (OP == 12) asm volatile("add.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(a));
      if (OP == 13) asm volatile("dp4a.u32.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
      if (OP == 14) {  // mix: 1 prmt + 1 dp4a
        asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
        asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(x[(i + 1) % NACC]) : "r"(a), "r"(b));
      }
      if (OP == 15) {  // the scan kernel's per-op work today: and, shr, mad(sel), 3 prmt, 3 mad, mad + min, max  (12)
        uint32_t w = x[i], code, len, sel, tr, tq, tc, t;
        asm volatile("and.b32 %0, %1, 15;" : "=r"(code) : "r"(w));
        asm volatile("shr.u32 %0, %1, 4;" : "=r"(len) : "r"(w));
        asm volatile("mul.lo.u32 %0, %1, 0x1111;" : "=r"(sel) : "r"(code));
        asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(tr) : "r"(a), "r"(b), "r"(sel));
        asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(tq) : "r"(b), "r"(a), "r"(sel));
        asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(tc) : "r"(a), "r"(a), "r"(sel));
        asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(x[(i + 1) % NACC]) : "r"(len), "r"(tr));
        asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(x[(i + 2) % NACC]) : "r"(len), "r"(tq));
        asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(x[(i + 3) % NACC]) : "r"(tc), "r"(a));
        asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(t) : "r"(len), "r"(tc));
        asm volatile("min.s32 %0, %0, %1;" : "+r"(x[(i + 4) % NACC]) : "r"(t));
        asm volatile("max.u32 %0, %0, %1;" : "+r"(x[(i + 5) % NACC]) : "r"(code));
      }
      if (OP == 16) {  // candidate per-op work: shr, 3 prmt (selector = the word), 3 dp4a, min  (8)
        uint32_t w = x[i], len, tr, tq, tc, t;
        asm volatile("shr.u32 %0, %1, 4;" : "=r"(len) : "r"(w));
        asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(tr) : "r"(a), "r"(b), "r"(w));
        asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(tq) : "r"(b), "r"(a), "r"(w));
        asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(tc) : "r"(a), "r"(a), "r"(w));
        asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(x[(i + 1) % NACC]) : "r"(len), "r"(tr));
        asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(x[(i + 2) % NACC]) : "r"(len), "r"(tq));
        asm volatile("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(t) : "r"(len), "r"(tc), "r"(0));
        asm volatile("min.s32 %0, %0, %1;" : "+r"(x[(i + 4) % NACC]) : "r"(t));
      }
    }
  }
  long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += x[i] + (uint32_t)y[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int OP>
void run(const char* name, int per_iter) {
  uint32_t* out; long long* cyc; long long h;
  cudaMalloc(&out, 148 * 8 * 256 * 4); cudaMalloc(&cyc, 8);
  for (int warps = 8; warps >= 1; warps /= 2) {  // warps per SM sub-partition: blockDim 256 = 2 warps/SMSP; use 1..4 blocks/SM
    int blocks_per_sm = warps / 2 > 0 ? warps / 2 : 1;
    int threads = warps >= 2 ? 256 : 128;
    k<OP><<<148 * blocks_per_sm, threads>>>(out, 3, 5, cyc);
    k<OP><<<148 * blocks_per_sm, threads>>>(out, 3, 5, cyc);
    cudaDeviceSynchronize();
    if (cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("%s: failed\n", name); return; }
    double wi = (double)ITER * NACC * per_iter * (threads / 32) * blocks_per_sm / 4.0;  // warp instr per SMSP
    printf("%-22s warps/SMSP=%d  cycles/warp-instr/SMSP = %.3f\n", name, (threads / 32) * blocks_per_sm / 4, h / wi);
  }
}

int main() {
  run<0>("mad.lo (IMAD)", 1);
  run<1>("mul.hi (IMAD.HI)", 1);
  run<2>("mad.hi", 1);
  run<3>("mad.wide (IMAD.WIDE)", 1);
  run<4>("prmt", 1);
  run<5>("lop3", 1);
  run<6>("shr", 1);
  run<7>("max.u32", 1);
  run<8>("prmt+mad.lo pair", 2);
  run<9>("prmt+mul.hi pair", 2);
  run<10>("bfe", 1);
  run<11>("shf.r.wrap", 1);
  run<12>("add", 1);
  run<13>("dp4a", 1);
  run<14>("prmt+dp4a pair", 2);
  run<15>("scan per-op, today (x12)", 1);
  run<16>("scan per-op, dp4a (x8)", 1);
  cudaError_t e = cudaGetLastError();
  printf("status: %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}

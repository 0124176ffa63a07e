// Issue-rate microbenchmarks for the instructions the cost kernel's warp-cooperative scans rely on
// (SHFL, REDUX, uniform LDS.128, FFMA, FFMA2, FMNMX).  Prints warp-instructions per cycle per SM.
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
template <int OP> __global__ void k(float* out, long long* cyc) {
  __shared__ float4 sm[64];
  if (threadIdx.x < 64) sm[threadIdx.x] = make_float4(threadIdx.x, 1.f, 2.f, 3.f);
  __syncthreads();
  float a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  int i0 = threadIdx.x, i1 = i0 + 1, i2 = i0 + 2, i3 = i0 + 3;
  unsigned long long p0 = threadIdx.x, p1 = p0 + 1, p2 = p0 + 2, p3 = p0 + 3;
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
    if (OP == 0) { a0 = fmaf(a0, a1, a2); a1 = fmaf(a1, a2, a3); a2 = fmaf(a2, a3, a4); a3 = fmaf(a3, a4, a5); a4 = fmaf(a4, a5, a6); a5 = fmaf(a5, a6, a7); a6 = fmaf(a6, a7, a0); a7 = fmaf(a7, a0, a1); }
    if (OP == 1) { a0 = __shfl_xor_sync(~0u, a0, 1); a1 = __shfl_xor_sync(~0u, a1, 2); a2 = __shfl_xor_sync(~0u, a2, 4); a3 = __shfl_xor_sync(~0u, a3, 8); a4 = __shfl_xor_sync(~0u, a4, 16); a5 = __shfl_xor_sync(~0u, a5, 1); a6 = __shfl_xor_sync(~0u, a6, 2); a7 = __shfl_xor_sync(~0u, a7, 4); }
    if (OP == 2) { i0 = __reduce_max_sync(~0u, i0 + it); i1 = __reduce_max_sync(~0u, i1 + it); i2 = __reduce_max_sync(~0u, i2 + it); i3 = __reduce_max_sync(~0u, i3 + it);
                   i0 = __reduce_min_sync(~0u, i0 ^ it); i1 = __reduce_min_sync(~0u, i1 ^ it); i2 = __reduce_min_sync(~0u, i2 ^ it); i3 = __reduce_min_sync(~0u, i3 ^ it); }
    if (OP == 3) { float4 v0 = sm[(it) & 63], v1 = sm[(it + 1) & 63], v2 = sm[(it + 2) & 63], v3 = sm[(it + 3) & 63], v4 = sm[(it + 4) & 63], v5 = sm[(it + 5) & 63], v6 = sm[(it + 6) & 63], v7 = sm[(it + 7) & 63];
                   a0 += v0.x; a1 += v1.y; a2 += v2.z; a3 += v3.w; a4 += v4.x; a5 += v5.y; a6 += v6.z; a7 += v7.w; }
    if (OP == 4) { asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p0) : "l"(p1), "l"(p2)); asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p1) : "l"(p2), "l"(p3));
                   asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p2) : "l"(p3), "l"(p0)); asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p3) : "l"(p0), "l"(p1));
                   asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p0) : "l"(p1), "l"(p3)); asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p1) : "l"(p2), "l"(p0));
                   asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p2) : "l"(p3), "l"(p1)); asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p3) : "l"(p0), "l"(p2)); }
    if (OP == 5) { a0 = fmaxf(a0, a1); a1 = fmaxf(a1, a2); a2 = fmaxf(a2, a3); a3 = fmaxf(a3, a4); a4 = fmaxf(a4, a5); a5 = fmaxf(a5, a6); a6 = fmaxf(a6, a7); a7 = fmaxf(a7, a0 + 1.f); }
    if (OP == 6) { unsigned b0 = __ballot_sync(~0u, a0 > it), b1 = __ballot_sync(~0u, a1 > it), b2 = __ballot_sync(~0u, a2 > it), b3 = __ballot_sync(~0u, a3 > it);
                   unsigned b4 = __ballot_sync(~0u, a4 > it), b5 = __ballot_sync(~0u, a5 > it), b6 = __ballot_sync(~0u, a6 > it), b7 = __ballot_sync(~0u, a7 > it);
                   i0 += b0 ^ b1; i1 += b2 ^ b3; i2 += b4 ^ b5; i3 += b6 ^ b7; a0 += 1.f; a1 += 1.f; a2 += 1.f; a3 += 1.f; a4 += 1.f; a5 += 1.f; a6 += 1.f; a7 += 1.f; }
    if (OP == 7) { double d0 = a0, d1 = a1, d2 = a2, d3 = a3; d0 = fma(d0, d1, d2); d1 = fma(d1, d2, d3); d2 = fma(d2, d3, d0); d3 = fma(d3, d0, d1); d0 = fma(d0, d1, d2); d1 = fma(d1, d2, d3); d2 = fma(d2, d3, d0); d3 = fma(d3, d0, d1);
                   a0 = d0; a1 = d1; a2 = d2; a3 = d3; }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 + i0 + i1 + i2 + i3 + (float)(p0 + p1 + p2 + p3);
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int OP> void run(const char* name, int ops_per_iter, int warps_per_sm) {
  int nsm = 148; float* out; long long* cyc;
  int threads = 256, blocks_per_sm = warps_per_sm * 32 / threads;
  cudaMalloc(&out, sizeof(float) * nsm * blocks_per_sm * threads); cudaMalloc(&cyc, sizeof(long long) * nsm * blocks_per_sm);
  k<OP><<<nsm * blocks_per_sm, threads>>>(out, cyc); cudaDeviceSynchronize();
  k<OP><<<nsm * blocks_per_sm, threads>>>(out, cyc); cudaDeviceSynchronize();
  long long h[4096]; cudaMemcpy(h, cyc, sizeof(long long) * nsm * blocks_per_sm, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < nsm * blocks_per_sm; ++i) avg += h[i]; avg /= nsm * blocks_per_sm;
  printf("%-28s warps/SM %2d  %.3f warp-instr/cycle/SM (named op only; loop overhead ignored)\n", name, warps_per_sm, (double)ITERS * ops_per_iter * warps_per_sm / avg);
  cudaFree(out); cudaFree(cyc);
}
int main() {
  for (int w : {8, 16, 32}) {
    run<0>("FFMA", 8, w); run<4>("FFMA2", 8, w); run<5>("FMNMX", 8, w); run<1>("SHFL.BFLY", 8, w); run<2>("REDUX.MAX/MIN", 8, w);
    run<3>("LDS.128 uniform", 8, w); run<6>("VOTE.ballot (+FSETP)", 8, w); run<7>("DFMA (+cvt)", 8, w);
  }
  return 0;
}

// Gather-rate microbenchmark: the roofline the north_star's grid design (8 manual fp32 corner loads per
// trilinear sample) is measured against.  MEASURED_PEAKS.json holds an HBM copy number only; this
// program measures, on the same GPU,
//   random4   independent random 4-byte loads over an array of S bytes
//   trilerp8  the access pattern of grid_sd(): a random cell of an [nz][ny][nx] fp32 grid, its 8 corners
//             (4 rows of 2 adjacent floats)
// for S = 48 MiB (the baked grid pool of the bench scene: L2-resident on a 126 MB L2) and S = 4 GiB
// (HBM-resident).  Reported: loads/s, "algorithmic" GB/s (4 B per load) and "sector" GB/s (32 B per
// distinct sector touched, what ncu's lts__t_bytes counts).  Output: one JSON line.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o l2_gather l2_gather.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned mix(unsigned x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

template <int ILP>
__global__ void __launch_bounds__(256) random4(const float* __restrict__ a, unsigned long long n, int iters, float* out) {
  unsigned s = mix(blockIdx.x * 256u + threadIdx.x + 1u);
  float acc = 0.f;
  for (int it = 0; it < iters; ++it) {
    float v[ILP];
#pragma unroll
    for (int k = 0; k < ILP; ++k) { s = mix(s + 0x9e3779b9u); v[k] = __ldg(a + (unsigned long long)s % n); }
#pragma unroll
    for (int k = 0; k < ILP; ++k) acc += v[k];
  }
  out[blockIdx.x * 256 + threadIdx.x] = acc;
}

// coherent = 1: the 32 lanes of a warp sample cells within a small neighbourhood (what Morton-ordered points do)
template <int COHERENT>
__global__ void __launch_bounds__(256) trilerp8(const float* __restrict__ g, int nx, int ny, int nz, int iters, float* out) {
  unsigned s = mix(blockIdx.x * 256u + threadIdx.x + 1u);
  unsigned sw = mix((blockIdx.x * 256u + threadIdx.x) / 32u + 77u);
  const long long sy = nx, sz = (long long)nx * ny;
  float acc = 0.f;
  for (int it = 0; it < iters; ++it) {
    s = mix(s + 0x9e3779b9u);
    int ix, iy, iz;
    if (COHERENT) {
      sw = mix(sw + 0x9e3779b9u);
      ix = (int)(sw % (unsigned)(nx - 5)) + (int)(s & 3u); iy = (int)((sw >> 10) % (unsigned)(ny - 5)) + (int)((s >> 2) & 3u); iz = (int)((sw >> 20) % (unsigned)(nz - 5)) + (int)((s >> 4) & 3u);
    } else {
      ix = (int)(s % (unsigned)(nx - 1)); iy = (int)(mix(s) % (unsigned)(ny - 1)); iz = (int)(mix(s + 1u) % (unsigned)(nz - 1));
    }
    const float* p = g + iz * sz + iy * sy + ix;
    const float c000 = __ldg(p), c100 = __ldg(p + 1), c010 = __ldg(p + sy), c110 = __ldg(p + sy + 1);
    const float c001 = __ldg(p + sz), c101 = __ldg(p + sz + 1), c011 = __ldg(p + sz + sy), c111 = __ldg(p + sz + sy + 1);
    acc += ((c000 + c100) + (c010 + c110)) + ((c001 + c101) + (c011 + c111));
  }
  out[blockIdx.x * 256 + threadIdx.x] = acc;
}

template <class F> static double time_ms(F launch) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  launch(); launch(); cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) { cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms; }
  return best;
}

int main() {
  cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
  const int blocks = pr.multiProcessorCount * 8, threads = 256;
  float* out; cudaMalloc(&out, sizeof(float) * blocks * threads);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"l2_bytes\": %d", pr.name, pr.multiProcessorCount, pr.l2CacheSize);
  const size_t sizes[2] = {48ull << 20, 4ull << 30};
  const char* names[2] = {"l2_48MiB", "hbm_4GiB"};
  for (int si = 0; si < 2; ++si) {
    const size_t S = sizes[si];
    float* a; if (cudaMalloc(&a, S) != cudaSuccess) { printf(", \"%s\": null", names[si]); continue; }
    cudaMemset(a, 0, S);
    const unsigned long long n = S / 4;
    const int iters = 512;
    const double loads = (double)blocks * threads * iters * 8;
    double ms = time_ms([&] { random4<8><<<blocks, threads>>>(a, n, iters, out); });
    printf(", \"%s\": {\"random4\": {\"loads_per_s\": %.4g, \"alg_gbs\": %.1f, \"sector_gbs\": %.1f}", names[si], loads / ms * 1e3, loads * 4 / ms / 1e6, loads * 32 / ms / 1e6);
    // grid of S bytes, roughly cubic
    int nx = 1; while ((size_t)(nx + 1) * (nx + 1) * (nx + 1) * 4 <= S) ++nx;
    for (int coh = 0; coh < 2; ++coh) {
      ms = coh ? time_ms([&] { trilerp8<1><<<blocks, threads>>>(a, nx, nx, nx, iters, out); }) : time_ms([&] { trilerp8<0><<<blocks, threads>>>(a, nx, nx, nx, iters, out); });
      const double samples = (double)blocks * threads * iters;
      printf(", \"trilerp8_%s\": {\"samples_per_s\": %.4g, \"alg_gbs\": %.1f}", coh ? "warp_coherent" : "random", samples / ms * 1e3, samples * 32 / ms / 1e6);
    }
    printf("}");
    cudaFree(a);
  }
  printf("}\n");
  return 0;
}

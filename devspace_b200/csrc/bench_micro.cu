// bench_micro.cu -- calibration microbenchmarks for the design of the LZ77 kernel and the e2e roofline.
// Stand-alone binary (not part of libdsgpack.so). Measures on one B200:
//   * shared-memory atomicExch (32-bit) / atomicCAS (16-bit) / random LDS / __match_any_sync rates,
//     expressed in SM cycles per warp instruction with all warps of a full-occupancy CTA issuing;
//   * pinned host<->device copy bandwidth (PCIe Gen5): the denominator of the end-to-end roofline.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e = (x);                                                           \
    if (e != cudaSuccess) {                                                        \
      fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e));                      \
      exit(1);                                                                     \
    }                                                                              \
  } while (0)

constexpr int kIters = 2048;
constexpr int kTable = 8192;  // words (32 KiB static shared memory)

__device__ __forceinline__ uint32_t rng(uint32_t &s) {
  s ^= s << 13;
  s ^= s >> 17;
  s ^= s << 5;
  return s;
}

template <int MODE>
__global__ void __launch_bounds__(1024) smem_kernel(uint32_t *out, long long *cycles) {
  __shared__ uint32_t tab[kTable];
  for (int i = threadIdx.x; i < kTable; i += blockDim.x) tab[i] = i * 2654435761u;
  __syncthreads();
  uint32_t s = threadIdx.x * 747796405u + blockIdx.x * 2891336453u + 1u;
  uint32_t acc = 0;
  long long t0 = clock64();
#pragma unroll 4
  for (int i = 0; i < kIters; i++) {
    uint32_t r = rng(s);
    if (MODE == 0) {  // random 32-bit LDS
      acc += tab[r & (kTable - 1)];
    } else if (MODE == 1) {  // random atomicExch 32-bit
      acc += atomicExch(&tab[r & (kTable - 1)], r);
    } else if (MODE == 2) {  // random atomicCAS 16-bit (exchange emulation, one attempt)
      unsigned short *p = reinterpret_cast<unsigned short *>(tab) + (r & (2 * kTable - 1));
      unsigned short old = *p;
      acc += atomicCAS(p, old, (unsigned short)r);
    } else if (MODE == 3) {  // match.any on 15-bit keys (mostly distinct)
      acc += __match_any_sync(0xffffffffu, r & 0x7fff);
    } else if (MODE == 4) {  // match.any on 3-bit keys (heavy duplication)
      acc += __match_any_sync(0xffffffffu, r & 7);
    } else if (MODE == 5) {  // random 16-bit LDS
      acc += reinterpret_cast<unsigned short *>(tab)[r & (2 * kTable - 1)];
    } else if (MODE == 6) {  // dependent pointer chase through 16-bit links (latency)
      acc = reinterpret_cast<unsigned short *>(tab)[(acc + r) & (2 * kTable - 1)];
    } else if (MODE == 7) {  // random plain 16-bit store
      reinterpret_cast<unsigned short *>(tab)[r & (2 * kTable - 1)] = (unsigned short)r;
    } else if (MODE == 9) {  // random atomicCAS 32-bit (exchange emulation, one attempt)
      uint32_t *p = &tab[r & (kTable - 1)];
      uint32_t old = *p;
      acc += atomicCAS(p, old, r);
    } else if (MODE == 8) {  // shared atomicAdd histogram on 286 bins
      atomicAdd(&tab[r % 286u], 1u);
    }
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc + tab[threadIdx.x];
}

template <int MODE>
void run_smem(const char *name, int threads) {
  int nblk = 148;
  uint32_t *out;
  long long *cyc;
  CK(cudaMalloc(&out, sizeof(uint32_t) * nblk * 1024));
  CK(cudaMalloc(&cyc, sizeof(long long) * nblk));
  smem_kernel<MODE><<<nblk, threads>>>(out, cyc);
  CK(cudaDeviceSynchronize());
  smem_kernel<MODE><<<nblk, threads>>>(out, cyc);
  CK(cudaDeviceSynchronize());
  long long h[148];
  CK(cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost));
  double avg = 0;
  for (int i = 0; i < nblk; i++) avg += (double)h[i];
  avg /= nblk;
  int warps = threads / 32;
  // cycles per warp-instruction at SM level = elapsed / (iters * warps)
  printf("{\"micro\":\"%s\",\"threads\":%d,\"cycles_total\":%.0f,\"cyc_per_warp_instr\":%.2f,\"cyc_per_lane_op\":%.3f}\n",
         name, threads, avg, avg / ((double)kIters * warps), avg / ((double)kIters * warps * 32));
  cudaFree(out);
  cudaFree(cyc);
}

void run_pcie() {
  size_t bytes = 1ull << 30;
  void *h, *d;
  CK(cudaMallocHost(&h, bytes));
  CK(cudaMalloc(&d, bytes));
  memset(h, 1, bytes);
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  cudaStream_t s1, s2;
  CK(cudaStreamCreate(&s1));
  CK(cudaStreamCreate(&s2));
  for (int dir = 0; dir < 3; dir++) {
    float best = 1e30f;
    for (int it = 0; it < 4; it++) {
      CK(cudaDeviceSynchronize());
      CK(cudaEventRecord(a, s1));
      if (dir == 0) CK(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, s1));
      if (dir == 1) CK(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, s1));
      if (dir == 2) {  // both directions at once, half a GiB each way
        CK(cudaMemcpyAsync(d, h, bytes / 2, cudaMemcpyHostToDevice, s1));
        CK(cudaMemcpyAsync((char *)h + bytes / 2, (char *)d + bytes / 2, bytes / 2, cudaMemcpyDeviceToHost, s2));
        CK(cudaStreamSynchronize(s2));
      }
      CK(cudaEventRecord(b, s1));
      CK(cudaEventSynchronize(b));
      float ms;
      CK(cudaEventElapsedTime(&ms, a, b));
      if (ms < best) best = ms;
    }
    const char *nm = dir == 0 ? "h2d_pinned" : dir == 1 ? "d2h_pinned" : "bidir_pinned_each_half";
    printf("{\"micro\":\"%s\",\"bytes\":%zu,\"ms\":%.3f,\"GBps\":%.2f}\n", nm, bytes, best, bytes / 1e6 / best);
  }
  cudaFreeHost(h);
  cudaFree(d);
}

int main() {
  cudaDeviceProp p;
  CK(cudaGetDeviceProperties(&p, 0));
  printf("{\"device\":\"%s\",\"sms\":%d,\"smem_per_block_optin\":%zu,\"clock_khz\":%d}\n", p.name, p.multiProcessorCount,
         p.sharedMemPerBlockOptin, p.clockRate);
  for (int threads : {512, 1024}) {
    run_smem<0>("lds32_random", threads);
    run_smem<5>("lds16_random", threads);
    run_smem<7>("sts16_random", threads);
    run_smem<6>("lds16_dependent_chase", threads);
    run_smem<1>("atoms_exch32_random", threads);
    run_smem<2>("atoms_cas16_random", threads);
    run_smem<9>("atoms_cas32_random", threads);
    run_smem<8>("atoms_add_hist286", threads);
    run_smem<3>("match_any_15bit", threads);
    run_smem<4>("match_any_3bit", threads);
  }
  run_pcie();
  return 0;
}

// tools/micro_match.cu -- DEVELOPMENT AID (not built by build(), not linked into the product): cycles of the
// building blocks of the ordered link step of deflate_lz.cuh on sm_100a -- MATCH.ANY with distinct / equal values,
// an indexed shuffle, a shared-memory round trip with __syncwarp, a shared-memory atomic.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/micro_match tools/micro_match.cu && /tmp/micro_match
#include <cstdint>
#include <cstdio>

__global__ void k(unsigned long long *out, int iters) {
  __shared__ uint32_t head[8192];
  __shared__ uint16_t prev[8192];
  const uint32_t lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) {
    head[i] = i;
    prev[i] = (uint16_t)i;
  }
  __syncthreads();
  uint32_t acc = 0;
  long long t0, t1;
  // 1. match.any, all distinct
  uint32_t v = lane * 2654435761u;
  __syncwarp();
  t0 = clock64();
  for (int i = 0; i < iters; i++) {
    const uint32_t m = __match_any_sync(0xffffffffu, v);
    acc += m;
    v = v * 1664525u + 1013904223u + m;
  }
  t1 = clock64();
  if (threadIdx.x == 0) out[0] = (unsigned long long)(t1 - t0);
  // 2. match.any, all equal
  v = 7;
  t0 = clock64();
  for (int i = 0; i < iters; i++) {
    const uint32_t m = __match_any_sync(0xffffffffu, v);
    acc += m;
    v = (v + m) & 0xffu ? 7u : 7u;
  }
  t1 = clock64();
  if (threadIdx.x == 0) out[1] = (unsigned long long)(t1 - t0);
  // 3. match.any, 4 groups of 8
  v = lane & 3;
  t0 = clock64();
  for (int i = 0; i < iters; i++) {
    const uint32_t m = __match_any_sync(0xffffffffu, v);
    acc += m;
    v = (v + (m & 4u)) & 3u;
  }
  t1 = clock64();
  if (threadIdx.x == 0) out[2] = (unsigned long long)(t1 - t0);
  // 4. indexed shuffle chain
  v = lane;
  t0 = clock64();
  for (int i = 0; i < iters; i++) v = __shfl_sync(0xffffffffu, v * 5u + 1u, (int)(v & 31u));
  t1 = clock64();
  acc += v;
  if (threadIdx.x == 0) out[3] = (unsigned long long)(t1 - t0);
  // 5. shared round trip: read head, syncwarp, write head, syncwarp (dependent through memory only)
  uint32_t h = (lane * 2654435761u) >> 19;
  t0 = clock64();
  for (int i = 0; i < iters; i++) {
    const uint32_t w = head[h];
    __syncwarp();
    head[h] = (uint32_t)i;
    prev[(h + i) & 8191] = (uint16_t)w;
    __syncwarp();
    h = (h + 977u) & 8191u;
  }
  t1 = clock64();
  if (threadIdx.x == 0) out[4] = (unsigned long long)(t1 - t0);
  // 6. the same with the address depending on the value read (a true chain)
  t0 = clock64();
  for (int i = 0; i < iters; i++) {
    const uint32_t w = head[h];
    __syncwarp();
    head[h] = (uint32_t)i + lane;
    __syncwarp();
    h = (h + w + 977u) & 8191u;
  }
  t1 = clock64();
  if (threadIdx.x == 0) out[5] = (unsigned long long)(t1 - t0);
  // 7. shared atomicCAS 32-bit, distinct addresses
  t0 = clock64();
  for (int i = 0; i < iters; i++) {
    const uint32_t w = atomicCAS(&head[h], (uint32_t)i + lane, 5u);
    h = (h + w + 977u) & 8191u;
  }
  t1 = clock64();
  if (threadIdx.x == 0) out[6] = (unsigned long long)(t1 - t0);
  // 8. a full link step as in insert_link (distinct buckets)
  v = lane * 2654435761u;
  t0 = clock64();
  for (int i = 0; i < iters; i++) {
    const uint32_t hv = (v >> 19);
    const uint32_t mask = __match_any_sync(0xffffffffu, hv);
    const uint32_t lower = mask & ((1u << lane) - 1u);
    const uint32_t below = __shfl_sync(0xffffffffu, v, lower ? 31 - __clz((int)lower) : (int)lane);
    const uint32_t word = !lower ? head[hv] : 0u;
    __syncwarp();
    if ((mask >> lane) == 1u) head[hv] = v;
    prev[(hv * 3u + i) & 8191u] = (uint16_t)(lower ? below : word);
    __syncwarp();
    v = v * 1664525u + 1013904223u;
  }
  t1 = clock64();
  if (threadIdx.x == 0) out[7] = (unsigned long long)(t1 - t0);
  if (acc == 0x12345678u) out[15] = acc;
}

int main() {
  unsigned long long *d, h[16];
  cudaMalloc(&d, sizeof h);
  const int iters = 4096;
  const char *names[8] = {"match.any distinct", "match.any equal", "match.any 4 groups", "shfl idx chain", "lds/sync/sts/sync",
                          "lds->addr chain", "atomicCAS smem", "link step"};
  for (int threads : {32, 256, 1024}) {
    cudaMemset(d, 0, sizeof h);
    k<<<1, threads>>>(d, iters);
    k<<<1, threads>>>(d, iters);
    cudaDeviceSynchronize();
    cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
    printf("threads=%d\n", threads);
    for (int i = 0; i < 8; i++) printf("  %-22s %8.1f cycles/iter\n", names[i], (double)h[i] / iters);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}

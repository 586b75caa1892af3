// Probe: how many DRAM bytes does a random 32-byte read cost on B200, and which knob
// changes it?  Each thread reads one random 32-byte-aligned 16-byte word of a large
// array.  Run under `ncu --metrics dram__bytes_read.sum` (one line per variant).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t rnd(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x;
}
template <int MODE>
__global__ void k_probe(const uint4* __restrict__ a, uint64_t n_sectors, uint32_t* out, uint32_t salt) {
  uint32_t acc = 0;
  const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
  for (int it = 0; it < 4; ++it) {
    const uint64_t s = ((uint64_t)rnd(gid * 4 + it + salt) * n_sectors) >> 32;
    const uint4* p = a + s * 2;  // 32-byte sectors, 16-byte elements
    uint4 v;
    if (MODE == 0) v = *p;
    else if (MODE == 1) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    else if (MODE == 2) asm volatile("ld.global.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    else if (MODE == 3) asm volatile("ld.global.cv.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    else if (MODE == 4) asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    else if (MODE == 5) asm volatile("ld.global.L1::evict_first.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    acc += v.x ^ v.y ^ v.z ^ v.w;
  }
  if (acc == 0x12345678u) out[0] = acc;
}
// cp.async variants (what the row gather uses)
template <int MODE>
__global__ void k_probe_cpasync(const uint4* __restrict__ a, uint64_t n_sectors, uint32_t* out, uint32_t salt) {
  __shared__ uint4 sh[256 * 4];
  const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
  for (int it = 0; it < 4; ++it) {
    const uint64_t s = ((uint64_t)rnd(gid * 4 + it + salt) * n_sectors) >> 32;
    const uint4* p = a + s * 2;
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(&sh[threadIdx.x * 4 + it]);
    if (MODE == 0) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(p) : "memory");
    else if (MODE == 1) asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(p) : "memory");
    else if (MODE == 2) asm volatile("cp.async.cg.shared.global.L2::64B [%0], [%1], 16;" ::"r"(sa), "l"(p) : "memory");
  }
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  uint4 v = sh[(threadIdx.x * 4 + 1) % 1024];
  if ((v.x ^ v.y) == 0x12345678u) out[0] = v.x;
}

int main(int argc, char** argv) {
  int want = argc > 1 ? atoi(argv[1]) : -1;
  size_t lim = 0;
  cudaDeviceGetLimit(&lim, cudaLimitMaxL2FetchGranularity);
  printf("default cudaLimitMaxL2FetchGranularity = %zu\n", lim);
  if (want >= 0) {
    cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)want);
    cudaDeviceGetLimit(&lim, cudaLimitMaxL2FetchGranularity);
    printf("set %d -> %s, now %zu\n", want, cudaGetErrorString(e), lim);
  }
  const uint64_t bytes = 8ull << 30;
  uint4* a; uint32_t* out;
  cudaMalloc(&a, bytes); cudaMalloc(&out, 4);
  cudaMemset(a, 1, bytes);
  const uint64_t n_sectors = bytes / 32;
  const int threads = 256, blocks = 148 * 64;   // 2.4M threads x 4 reads = 9.7M sector reads = 310 MB useful
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
#define RUN(name, call)                                                    \
  { call; cudaDeviceSynchronize(); cudaEventRecord(e0); call; cudaEventRecord(e1); cudaEventSynchronize(e1); \
    float ms; cudaEventElapsedTime(&ms, e0, e1);                          \
    printf("%-28s %.3f ms  (%.1f GB/s of useful 32 B sectors)\n", name, ms, blocks * threads * 4.0 * 32 / ms / 1e6); }
  RUN("ld default", (k_probe<0><<<blocks, threads>>>(a, n_sectors, out, 1)));
  RUN("ld.nc.L1::no_allocate", (k_probe<1><<<blocks, threads>>>(a, n_sectors, out, 2)));
  RUN("ld.L2::64B", (k_probe<2><<<blocks, threads>>>(a, n_sectors, out, 3)));
  RUN("ld.cv", (k_probe<3><<<blocks, threads>>>(a, n_sectors, out, 4)));
  RUN("ld.cg", (k_probe<4><<<blocks, threads>>>(a, n_sectors, out, 5)));
  RUN("ld evict_first", (k_probe<5><<<blocks, threads>>>(a, n_sectors, out, 6)));
  RUN("cp.async.cg", (k_probe_cpasync<0><<<blocks, threads>>>(a, n_sectors, out, 7)));
  RUN("cp.async.ca", (k_probe_cpasync<1><<<blocks, threads>>>(a, n_sectors, out, 8)));
  RUN("cp.async.cg.L2::64B", (k_probe_cpasync<2><<<blocks, threads>>>(a, n_sectors, out, 9)));
  printf("last error: %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}

// microbench_gather.cu — how many DRAM bytes does one random 16-byte gather cost on B200, and
// what streaming read bandwidth does the sweep's load pattern reach?  (development aid)
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
#include <algorithm>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint4 ld_nc_na(const uint4* p) { uint4 r; asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p)); return r; }
__device__ __forceinline__ uint4 ld_plain(const uint4* p) { uint4 r; asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p)); return r; }
__device__ __forceinline__ uint4 ld_nc_64(const uint4* p) { uint4 r; asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p)); return r; }
__device__ __forceinline__ uint4 ld_ef(const uint4* p) { uint4 r; asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p)); return r; }
struct u8x { uint32_t a[8]; };
__device__ __forceinline__ u8x ld_v8(const void* p) { u8x r; asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r.a[0]), "=r"(r.a[1]), "=r"(r.a[2]), "=r"(r.a[3]), "=r"(r.a[4]), "=r"(r.a[5]), "=r"(r.a[6]), "=r"(r.a[7]) : "l"(p)); return r; }
__global__ void __launch_bounds__(256) stream256_kernel(const uint4* src, int64_t n_tiles, unsigned long long* sink) {
  uint32_t acc = 0;
  for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const char* p = (const char*)(src + t * 2048 + (threadIdx.x >> 5) * 256) + (threadIdx.x & 31) * 32;
    u8x v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = ld_v8(p + j * 1024);
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int k = 0; k < 8; ++k) acc += (v[j].a[k] == 12345u);
  }
  if (acc == 0xFFFFFFFFu) *sink = acc;
}
__device__ __forceinline__ uint4 ld_cv(const uint4* p) { uint4 r; asm volatile("ld.global.cv.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p)); return r; }

template <int V>
__global__ void gather_kernel(const uint4* ids, const uint32_t* rows, uint4* out, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4* p = ids + rows[i];
  uint4 v;
  if (V == 0) v = ld_nc_na(p);
  else if (V == 1) v = ld_plain(p);
  else if (V == 2) v = ld_nc_64(p);
  else if (V == 3) v = ld_ef(p);
  else v = ld_cv(p);
  out[i] = v;
}

// 8-byte gather of two u64 halves stored in separate columns? no: 32-byte aligned pair read test
__global__ void gather32_kernel(const uint4* ids, const uint32_t* rows, uint4* out, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = ld_nc_na(ids + rows[i]);
}

// streaming read in the sweep's pattern: each warp 8 x LDG.128 coalesced, grid-stride tiles
template <int V>
__global__ void __launch_bounds__(256) stream_kernel(const uint4* src, int64_t n_tiles, unsigned long long* sink) {
  uint32_t acc = 0;
  for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const uint4* p = src + t * 2048 + (threadIdx.x >> 5) * 256 + (threadIdx.x & 31);
    uint4 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = V == 0 ? ld_nc_na(p + j * 32) : ld_plain(p + j * 32);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += (v[j].x == 12345u) + (v[j].y == 12345u) + (v[j].z == 12345u) + (v[j].w == 12345u);
  }
  if (acc == 0xFFFFFFFFu) *sink = acc;
}

int main(int argc, char** argv) {
  const int64_t N = 256ll << 20;  // rows
  const int64_t H = N / 40;       // 2.5 % hits
  uint4* ids; uint32_t* rows; uint4* out; uint32_t* due; unsigned long long* sink;
  CK(cudaMalloc(&ids, N * 16)); CK(cudaMalloc(&rows, H * 4)); CK(cudaMalloc(&out, H * 16));
  CK(cudaMalloc(&due, N * 4)); CK(cudaMalloc(&sink, 8));
  CK(cudaMemset(ids, 1, N * 16)); CK(cudaMemset(due, 0, N * 4));
  std::vector<uint32_t> h(H);
  uint64_t s = 88172645463325252ull;
  for (int64_t i = 0; i < H; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (uint32_t)(i * 40 + s % 40); }
  CK(cudaMemcpy(rows, h.data(), H * 4, cudaMemcpyHostToDevice));
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  size_t gran = 0; cudaDeviceGetLimit(&gran, cudaLimitMaxL2FetchGranularity);
  printf("default cudaLimitMaxL2FetchGranularity = %zu\n", gran);
  auto run = [&](const char* name, auto launch) {
    float best = 1e9f;
    for (int r = 0; r < 4; ++r) {
      cudaMemset(due, r, 512 << 20);  // flush L2
      cudaEventRecord(a); launch(); cudaEventRecord(b); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b); best = std::min(best, ms);
    }
    printf("%-28s %8.3f ms\n", name, best);
  };
  unsigned g = (unsigned)((H + 255) / 256);
  for (int pass = 0; pass < 2; ++pass) {
    if (pass == 1) {
      cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, 32);
      cudaDeviceGetLimit(&gran, cudaLimitMaxL2FetchGranularity);
      printf("set granularity 32 -> %s, now %zu\n", cudaGetErrorString(e), gran);
    }
    run("gather nc.no_allocate", [&] { gather_kernel<0><<<g, 256>>>(ids, rows, out, H); });
    run("gather plain", [&] { gather_kernel<1><<<g, 256>>>(ids, rows, out, H); });
    run("gather nc L2::64B", [&] { gather_kernel<2><<<g, 256>>>(ids, rows, out, H); });
    run("gather nc L2::128B", [&] { gather_kernel<3><<<g, 256>>>(ids, rows, out, H); });
    run("gather cv", [&] { gather_kernel<4><<<g, 256>>>(ids, rows, out, H); });
  }
  const int64_t tiles = N / 8192;
  for (int mult : {2, 4, 8}) {
    char nm[64];
    snprintf(nm, sizeof nm, "stream nc x%d CTAs/SM", mult);
    run(nm, [&] { stream_kernel<0><<<148 * mult, 256>>>((const uint4*)due, tiles, sink); });
    float ms = 0; (void)ms;
  }
  run("stream v8(256-bit) x8", [&] { stream256_kernel<<<148 * 8, 256>>>((const uint4*)due, tiles, sink); });
  run("stream v8(256-bit) x4", [&] { stream256_kernel<<<148 * 4, 256>>>((const uint4*)due, tiles, sink); });
  run("stream plain x8", [&] { stream_kernel<1><<<148 * 8, 256>>>((const uint4*)due, tiles, sink); });
  printf("stream bytes = %.3f GB ; gather hits = %lld\n", N * 4 / 1e9, (long long)H);
  CK(cudaDeviceSynchronize());
  return 0;
}

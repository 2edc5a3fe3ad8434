// sweep_lab.cu — development harness: ordering strategies for the sweep's compaction, timed on
// synthetic columns (5 % matches, 50 % completed).  Not part of the product.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
#include <algorithm>
#include "../aca-dotnet-workshop_b200/csrc/sweep.cuh"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
using namespace tt;

__global__ void fill(uint32_t* due, uint32_t* comp, uint32_t* over, int64_t n, uint32_t yday) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t h = tt_mix64(i * 0x9E3779B97F4A7C15ull + 12345);
  due[i] = (h % 100 < 5) ? yday : yday + 1 + (uint32_t)(h >> 40) % 300;
  uint32_t c = __ballot_sync(0xFFFFFFFFu, (h >> 20) & 1);
  if ((threadIdx.x & 31) == 0) { comp[i >> 5] = c; over[i >> 5] = 0; }
}

struct LabCtl { unsigned int tile_counter; unsigned int pad; unsigned long long cursor; };

int main(int argc, char** argv) {
  const int64_t N = argc > 1 ? atoll(argv[1]) : (1ll << 30);
  const int64_t NP = (N + 32767) / 32768 * 32768;
  const uint32_t yday = 739924;
  uint32_t *due, *comp, *over, *rows, *status, *ubase, *tcount, *gsum; SweepCtl* ctl; LabCtl* lctl;
  CK(cudaMalloc(&due, NP * 4)); CK(cudaMalloc(&comp, NP / 8)); CK(cudaMalloc(&over, NP / 8));
  CK(cudaMalloc(&rows, NP / 16 * 4)); CK(cudaMalloc(&status, (NP / 8192 + 4) * 4));
  CK(cudaMalloc(&ubase, NP / 8192 * 4)); CK(cudaMalloc(&tcount, NP / 8192 * 4));
  CK(cudaMalloc(&ctl, sizeof(SweepCtl))); CK(cudaMalloc(&lctl, sizeof(LabCtl)));
  CK(cudaMemset(due, 0xFF, NP * 4));
  fill<<<(unsigned)((NP + 255) / 256), 256>>>(due, comp, over, N, yday);
  CK(cudaDeviceSynchronize());
  CK(cudaMalloc(&gsum, 4096));
  SweepParams p; p.due_code = due; p.completed = comp; p.overdue = over; p.list = rows; p.tile_slot = ubase; p.tile_count = tcount; p.group_sum = gsum;
  p.ctl = ctl; p.cap = NP / 16; p.first_tile = 0; p.num_tiles = (uint32_t)(NP / 8192); p.mask = TT_MASK_STORE; p.yesterday = yday; p.mark_enable = 1;
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  auto run = [&](const char* name, int grid_mult, auto launch) {
    float best = 1e9f; unsigned long long total = 0;
    for (int r = 0; r < 4; ++r) {
      cudaMemset(over, 0, NP / 8);
      SweepCtl init; init.tile_counter = 0; init.pad = 0; init.cursor = 0; init.stored = ~0ull;
      cudaMemset(gsum, 0, 4096);
      cudaMemcpy(ctl, &init, sizeof init, cudaMemcpyHostToDevice);
      cudaMemset(lctl, 0, sizeof(LabCtl));
      cudaMemset(status, 0, (NP / 8192 + 4) * 4);
      cudaEventRecord(a); launch(148 * grid_mult); cudaEventRecord(b); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b); best = std::min(best, ms);
      SweepCtl h; cudaMemcpy(&h, ctl, sizeof h, cudaMemcpyDeviceToHost);
      LabCtl lh; cudaMemcpy(&lh, lctl, sizeof lh, cudaMemcpyDeviceToHost);
      total = h.cursor; (void)lh;
    }
    printf("%-40s %8.3f ms  %7.1f GB/s  hits=%llu  err=%s\n", name, best, N * 4.25 / best / 1e6, total, cudaGetErrorString(cudaGetLastError()));
  };
  p.num_tiles = (uint32_t)(NP / 8192);
  run("library sweep 256thr x4/SM", 4, [&](int g) { sweep_kernel<true, true, 256, 4><<<g, 256>>>(p); });
  run("library sweep 256thr x3/SM", 3, [&](int g) { sweep_kernel<true, true, 256, 4><<<g, 256>>>(p); });
  run("library sweep 256thr x2/SM", 2, [&](int g) { sweep_kernel<true, true, 256, 4><<<g, 256>>>(p); });
  p.list = nullptr;
  run("library sweep count-only x4/SM", 4, [&](int g) { sweep_kernel<true, false, 256, 4><<<g, 256>>>(p); });
  return 0;
}

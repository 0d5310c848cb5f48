// Microbenchmark: HBM write bandwidth of the epilogue's store pattern.  148 CTAs x 8 warps; per tile a warp
// writes NCH pieces of 512 B (32 lanes x 16 B).  "planes": piece j goes to chunk plane j (planes are tens
// of MB apart, the layout the kernels use); "blocked": the tile's pieces are adjacent (16 KB per tile).
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(uint4* out, long long plane_elems, int tiles_per_cta, int nch) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int t = 0; t < tiles_per_cta; ++t) {
    const long long tile = (long long)t * gridDim.x + blockIdx.x;
    for (int j = 0; j < nch; ++j) {
      long long idx;
      if (MODE == 0) idx = (long long)j * plane_elems + tile * 256 + w * 32 + lane;
      else idx = (tile * nch + j) * 256 + w * 32 + lane;
      out[idx] = make_uint4(t, j, w, lane);
    }
  }
}

int main() {
  const int nch = 4, tiles = 110, grid = 148;
  const long long plane = (long long)grid * tiles * 256;   // uint4 elements per plane
  uint4* d;
  cudaMalloc(&d, plane * nch * 16);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int mode = 0; mode < 2; ++mode) {
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) k<0><<<grid, 256>>>(d, plane, tiles, nch);
      else k<1><<<grid, 256>>>(d, plane, tiles, nch);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      printf("%s: %.1f us  %.0f GB/s\n", mode == 0 ? "planes " : "blocked", ms * 1e3, plane * nch * 16 / (ms * 1e-3) / 1e9);
    }
  }
  return 0;
}

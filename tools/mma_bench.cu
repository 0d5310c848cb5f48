// Microbenchmarks behind the design of gemm_tc.cuh: what one tcgen05.mma costs as a function of
// N, the shared-memory layout (no swizzle vs 128-byte swizzle), the A source (shared / tensor memory), the number of
// accumulators it rotates over and the CTA group (1 SM / an SM pair), and how fast an SM can pull a weight set that
// every SM reads out of L2.  Data are zeros: only the issue / operand-fetch timing is of interest.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/mma_bench tools/mma_bench.cu
//   tools/mma_bench            (prints one line per case; every wait is bounded)
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../open_universe_b200/csrc/gemm_tc.cuh"

using namespace uv;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

struct BenchParams {
  int N;        // UMMA N
  int layout;   // 0 = no swizzle, 2 = 128-byte swizzle (descriptor layout type)
  int ts;       // 1: A operand from tensor memory
  int iters;    // groups of 16 MMAs
  int n_acc;    // accumulators the MMAs rotate over (1, 2, 4)
  int accum;    // accumulate flag (0: every MMA overwrites: no dependency between them)
  long long* out;   // [grid][2]: cycles, MMAs
  int* err;
};

__device__ __forceinline__ uint64_t make_desc_l(uint32_t saddr, uint32_t lbo, uint32_t sbo, int layout) {
  return make_desc(saddr, lbo, sbo) | ((uint64_t)layout << 61);
}
__device__ __forceinline__ void tc_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
               "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc)
               : "memory");
}

// ---------------------------------------------------------------- one SM
__global__ void __launch_bounds__(128, 1) mma_bench_kernel(const BenchParams P) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 200 * 1024 / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_ptr)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_ptr;
  if (threadIdx.x == 0) {
    const int N = P.N;
    const uint32_t idesc = make_idesc(128, N);
    const uint32_t a_base = smem_u32(smem), b_base = smem_u32(smem + 96 * 1024);
    // no swizzle: A slab [K chunk][264 rows][16 B] (LBO = 264 * 16), B image [K chunk][N][16 B]; a K step = 2 chunks
    // 128-byte swizzle: A [rows][128 B] atoms of 8 rows (SBO = 1024), a K step = +32 B inside the row
    uint64_t ad[8], bd[8];
    for (int k = 0; k < 8; ++k) {
      if (P.layout == 0) {
        ad[k] = make_desc_l(a_base + k * 2 * 264 * 16, 264 * 16, 128, 0);
        bd[k] = make_desc_l(b_base + k * 2 * N * 16, N * 16, 128, 0);
      } else {
        ad[k] = make_desc_l(a_base + (k >> 2) * 256 * 128 + (k & 3) * 32, 16, 1024, 2);
        bd[k] = make_desc_l(b_base + (k >> 2) * N * 128 + (k & 3) * 32, 16, 1024, 2);
      }
    }
    const uint32_t a_t = tm + 480;   // TS mode: 8 columns per K step, 4 steps
    const uint32_t acc_stride = accum_stride_cols(N);
    const uint32_t amask = (uint32_t)P.n_acc - 1;
    const uint32_t accum = (uint32_t)P.accum;
    const uint64_t a_sub = (uint64_t)((128 * (P.layout == 0 ? 16 : 128)) >> 4);
    const long long t0 = clock64();
    uint32_t j = 0;
    for (int it = 0; it < P.iters; ++it) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t d0 = tm + (j & amask) * acc_stride, d1 = tm + ((j + 1) & amask) * acc_stride;
        j += 2;
        if (P.ts) {
          tc_mma_ts(d0, a_t + (k & 3) * 8, bd[k], idesc, accum);
          tc_mma_ts(d1, a_t + (k & 3) * 8, bd[k], idesc, accum);
        } else {
          tc_mma_bf16(d0, ad[k], bd[k], idesc, accum);
          tc_mma_bf16(d1, ad[k] + a_sub, bd[k], idesc, accum);
        }
      }
    }
    tc_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0, P.err, 1);
    const long long t1 = clock64();
    P.out[blockIdx.x * 2] = t1 - t0;
    P.out[blockIdx.x * 2 + 1] = (long long)P.iters * 16;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(512) : "memory");
  }
}

// ---------------------------------------------------------------- an SM pair (cta_group::2): M = 256 across the pair,
// every CTA holds its 128 rows of A and its N / 2 columns of B
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) mma_bench2_kernel(const BenchParams P) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5;
  const uint32_t rank = cta_rank_in_cluster();
  for (int i = threadIdx.x; i < 200 * 1024 / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  cluster_barrier();
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_ptr)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_barrier();
  tc_fence_after();
  const uint32_t tm = tmem_ptr;
  if (threadIdx.x == 0 && rank == 0) {
    const int N = P.N, Nh = N / 2;
    const uint32_t idesc = make_idesc(256, N);
    const uint32_t a_base = smem_u32(smem), b_base = smem_u32(smem + 96 * 1024);
    uint64_t ad[8], bd[8];
    for (int k = 0; k < 8; ++k) {
      if (P.layout == 0) {
        ad[k] = make_desc_l(a_base + k * 2 * 264 * 16, 264 * 16, 128, 0);
        bd[k] = make_desc_l(b_base + k * 2 * Nh * 16, Nh * 16, 128, 0);
      } else {
        ad[k] = make_desc_l(a_base + (k >> 2) * 256 * 128 + (k & 3) * 32, 16, 1024, 2);
        bd[k] = make_desc_l(b_base + (k >> 2) * Nh * 128 + (k & 3) * 32, 16, 1024, 2);
      }
    }
    const uint32_t acc_stride = accum_stride_cols(N);
    const uint32_t amask = (uint32_t)P.n_acc - 1;
    const uint32_t accum = (uint32_t)P.accum;
    const long long t0 = clock64();
    uint32_t j = 0;
    for (int it = 0; it < P.iters; ++it) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        tc_mma2_bf16(tm + (j & amask) * acc_stride, ad[k], bd[k], idesc, accum);
        tc_mma2_bf16(tm + ((j + 1) & amask) * acc_stride, ad[k], bd[k], idesc, accum);
        j += 2;
      }
    }
    tc_commit2(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0, P.err, 2);
    const long long t1 = clock64();
    P.out[blockIdx.x * 2] = t1 - t0;
    P.out[blockIdx.x * 2 + 1] = (long long)P.iters * 16;
  } else if (threadIdx.x == 0) {
    mbar_wait(smem_u32(&bar), 0, P.err, 3);   // the peer's shared / tensor memory is in use until the commit lands
    P.out[blockIdx.x * 2] = 0;
    P.out[blockIdx.x * 2 + 1] = 0;
  }
  tc_fence_before();
  __syncthreads();
  cluster_barrier();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(512) : "memory");
  }
}

// ---------------------------------------------------------------- L2 / HBM -> shared-memory feed: every CTA streams the
// same `set_bytes` (a weight set every SM needs: L2 hits after the first touch) or its own slice of a large buffer (HBM)
__global__ void __launch_bounds__(64, 1) feed_kernel(const char* src, long long set_bytes, long long cta_stride, int rounds, int piece,
                                                   long long* out, int* err) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t full[8];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&full[i]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const char* base = src + cta_stride * blockIdx.x;
    const int n_piece = (int)(set_bytes / piece);
    const long long total = (long long)rounds * n_piece;
    const long long t0 = clock64();
    // 6 pieces in flight through a ring of 6 buffers; a buffer is reused as soon as its copy has landed (no consumer)
    long long issued = 0, waited = 0;
    uint32_t ph[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    while (waited < total) {
      while (issued < total && issued - waited < 6) {
        const int slot = (int)(issued % 6);
        mbar_expect_tx(smem_u32(&full[slot]), piece);
        bulk_g2s(smem_u32(smem + (size_t)slot * piece), base + (issued % n_piece) * (long long)piece, piece, smem_u32(&full[slot]));
        ++issued;
      }
      const int slot = (int)(waited % 6);
      mbar_wait(smem_u32(&full[slot]), ph[slot], err, 4);
      ph[slot] ^= 1;
      ++waited;
    }
    const long long t1 = clock64();
    out[blockIdx.x * 2] = t1 - t0;
    out[blockIdx.x * 2 + 1] = total * piece;
  }
}

static void report(const char* name, const std::vector<long long>& h, int grid, double per_unit_macs) {
  double worst = 0, sum = 0;
  int n = 0;
  long long units = 0;
  for (int i = 0; i < grid; ++i) {
    if (h[2 * i + 1] == 0) continue;
    const double c = (double)h[2 * i] / (double)h[2 * i + 1];
    worst = c > worst ? c : worst;
    sum += c;
    units = h[2 * i + 1];
    ++n;
  }
  const double mean = sum / (n ? n : 1);
  if (per_unit_macs > 0)
    printf("%-44s  clk/MMA mean %7.1f worst %7.1f   -> %6.0f MAC/clk/issuing-CTA (%lld MMAs, %d CTAs)\n", name, mean, worst, per_unit_macs / mean,
           units, n);
  else
    printf("%-44s  B/clk/SM mean %6.1f (slowest CTA %6.1f), %d CTAs, %lld bytes each\n", name, 1.0 / mean, 1.0 / worst, n, units);
}

int main() {
  int dev_sms = 0;
  CK(cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, 0));
  long long* d_out;
  int* d_err;
  CK(cudaMalloc(&d_out, 4096 * sizeof(long long)));
  CK(cudaMalloc(&d_err, 4));
  CK(cudaMemset(d_err, 0, 4));
  CK(cudaFuncSetAttribute(mma_bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  CK(cudaFuncSetAttribute(mma_bench2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  CK(cudaFuncSetAttribute(feed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  std::vector<long long> h(4096);
  const int iters = 512;
  char name[128];
  for (int ts = 0; ts < 2; ++ts)
    for (int layout : {0, 2})
      for (int N : {32, 64, 128, 256})
        for (int n_acc : {1, 2, 4})
          for (int accum : {1, 0}) {
            if ((int)accum_stride_cols(N) * n_acc > 448) continue;
            if ((layout == 2 || n_acc != 2) && accum != 1) continue;
            if (layout == 2 && n_acc != 2) continue;
            BenchParams P{N, layout, ts, iters, n_acc, accum, d_out, d_err};
            mma_bench_kernel<<<dev_sms, 128, 200 * 1024>>>(P);
            CK(cudaDeviceSynchronize());
            CK(cudaMemcpy(h.data(), d_out, 2 * dev_sms * sizeof(long long), cudaMemcpyDeviceToHost));
            snprintf(name, sizeof(name), "1-SM M128 N%-3d %s A=%s acc x%d %s", N, layout ? "sw128" : "nosw ", ts ? "tmem" : "smem", n_acc,
                     accum ? "accumulate" : "overwrite ");
            report(name, h, dev_sms, 128.0 * N * 16);
          }
  const int grid2 = dev_sms & ~1;
  for (int layout : {0, 2})
    for (int N : {32, 64, 128, 256})
      for (int n_acc : {1, 2, 4}) {
        if ((int)accum_stride_cols(N) * n_acc > 512) continue;
        if (layout == 2 && n_acc != 2) continue;
        BenchParams P{N, layout, 0, iters, n_acc, 1, d_out, d_err};
        mma_bench2_kernel<<<grid2, 128, 200 * 1024>>>(P);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(h.data(), d_out, 2 * grid2 * sizeof(long long), cudaMemcpyDeviceToHost));
        snprintf(name, sizeof(name), "2-SM M256 N%-3d %s A=smem acc x%d accumulate", N, layout ? "sw128" : "nosw ", n_acc);
        report(name, h, grid2, 256.0 * N * 16);
      }
  char* buf;
  const long long big = 4ll << 30;
  CK(cudaMalloc(&buf, big));
  CK(cudaMemset(buf, 0, big));
  for (int piece : {8192, 32768}) {
    for (long long set : {327680ll, 655360ll}) {
      feed_kernel<<<dev_sms, 64, 200 * 1024>>>(buf, set, 0, 64, piece, d_out, d_err);   // same set for every CTA: L2
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(h.data(), d_out, 2 * dev_sms * sizeof(long long), cudaMemcpyDeviceToHost));
      snprintf(name, sizeof(name), "feed: shared %lld KB set, %d B pieces (L2)", set / 1024, piece);
      report(name, h, dev_sms, 0);
    }
    const long long per = (big / dev_sms) & ~((long long)piece * 64 - 1);
    feed_kernel<<<dev_sms, 64, 200 * 1024>>>(buf, per, per, 1, piece, d_out, d_err);   // disjoint slices: HBM
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(h.data(), d_out, 2 * dev_sms * sizeof(long long), cudaMemcpyDeviceToHost));
    snprintf(name, sizeof(name), "feed: disjoint slices, %d B pieces (HBM)", piece);
    report(name, h, dev_sms, 0);
  }
  int err = 0;
  CK(cudaMemcpy(&err, d_err, 4, cudaMemcpyDeviceToHost));
  printf("watchdog flag: %d\n", err);
  return 0;
}

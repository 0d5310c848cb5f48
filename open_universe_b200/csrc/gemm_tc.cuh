// Multi-tap implicit-GEMM conv1d on the sm_100a tensor cores (tcgen05 + TMEM).
//
//   D[row, n] = sum_{src s} sum_{shift q} sum_{c}  A_s[row + q, c] * W[s][q][n][c]
//
// covers every dense layer of the reference's ConvBlock
// (open_universe/networks/universe/blocks.py:331-416): "same" k=3/5 convs
// (one source, shifts -pad..pad), anti-aliased strided down convs (r phase
// sources, shifts -1..1; blocks.py:217-221 folded), transposed up convs (one
// source, N = r phases x C_out; blocks.py:223-225 folded), the conditioner's
// st_convs (condition.py:53-59) and the GRU input projection.
//
// A tiles are halo'd [KC/8][128 + smax - smin][8] bf16 slabs fetched with 1-D
// bulk async copies straight out of the planes layout (uv_common.cuh); that is
// the canonical no-swizzle K-major UMMA layout, in which a tap shift is
// +16 bytes on the descriptor start address, so one slab serves all taps.
// Weight tiles are pre-packed images [KC/8][N_TILE][8].
//
// Warp roles (192 threads): warp0 = bulk-copy producer, warp1 = MMA issuer and
// TMEM owner, warps 2-5 = epilogue (TMEM -> registers -> fused bias / add /
// FiLM / PReLU -> coalesced 16-byte stores into up to three planes tensors).
// Accumulators are double-buffered in TMEM so the epilogue of tile i overlaps
// the MMAs of tile i+1.  Persistent grid, static round-robin tile schedule.
#pragma once
#include "uv_common.cuh"

namespace uv {

constexpr int kMaxAStages = 4;
constexpr int kMaxBStages = 8;
constexpr int kGemmThreads = 192;

// TMEM columns per accumulator stage (power of two >= 32); two stages are allocated.
__host__ __device__ inline uint32_t accum_stride_cols(int n_tile) {
  uint32_t c = 32;
  while ((int)c < n_tile) c <<= 1;
  return c;
}

// ------------------------------------------------------------------ PTX glue
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must trap, never hang the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int* err_flag, int code) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 26)) {
      if (err_flag) atomicExch(err_flag, code);
      __threadfence_system();
      __trap();
    }
  }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
               : "memory");
}
__device__ __forceinline__ void tc_mma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// K-major, no-swizzle shared-memory matrix descriptor (sm_100 format, version 1).
//   lbo: byte stride between the two 8x16B core matrices of one K=16 step
//   sbo: byte stride between 8-row groups along M / N
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}

// kind::f16 instruction descriptor: D = f32, A = B = bf16, both K-major.
__device__ __forceinline__ uint32_t make_idesc(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ------------------------------------------------------------------ kernel
__global__ void __launch_bounds__(kGemmThreads, 1) gemm_tc_kernel(const __grid_constant__ GemmParams P) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // smem carve-up: A ring | B ring | barriers | tmem ptr
  uint8_t* a_ring = smem;
  uint8_t* b_ring = a_ring + (size_t)P.na_stages * P.a_stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(b_ring + (size_t)P.nb_stages * P.b_stage_bytes);
  uint64_t* a_full = bars;
  uint64_t* a_empty = a_full + kMaxAStages;
  uint64_t* b_full = a_empty + kMaxAStages;
  uint64_t* b_empty = b_full + kMaxBStages;
  uint64_t* t_full = b_empty + kMaxBStages;
  uint64_t* t_empty = t_full + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(t_empty + 2);

  if (threadIdx.x == 0) {
    for (int i = 0; i < P.na_stages; ++i) { mbar_init(smem_u32(&a_full[i]), 1); mbar_init(smem_u32(&a_empty[i]), 1); }
    for (int i = 0; i < P.nb_stages; ++i) { mbar_init(smem_u32(&b_full[i]), 1); mbar_init(smem_u32(&b_empty[i]), 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(smem_u32(&t_full[i]), 1); mbar_init(smem_u32(&t_empty[i]), 128); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)),
                 "r"(2u * accum_stride_cols(P.N_TILE))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t kAccumStride = accum_stride_cols(P.N_TILE);

  const int n_shift = P.smax - P.smin + 1;
  const int rows_a = kTileM + P.smax - P.smin;  // halo'd slab rows
  const int kc_chunks = P.KC >> 3;
  const int total_tiles = P.n_tiles_m * P.n_tiles_n;
  const uint32_t a_slab_chunk_bytes = (uint32_t)rows_a * 16u;
  const uint32_t a_bytes = a_slab_chunk_bytes * kc_chunks;
  const uint32_t b_bytes = (uint32_t)P.N_TILE * P.KC * 2u;

  if (warp == 0) {
    // ============================== producer
    if (lane == 0) {
      int ia = 0, ib = 0;
      uint32_t pa = 0, pb = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int n_tile = tile % P.n_tiles_n;
        const int m_tile = tile / P.n_tiles_n;
        const long long row0 = (long long)P.a_G + (long long)m_tile * kTileM + P.smin;
        for (int s = 0; s < P.n_src; ++s) {
          for (int g = 0; g < P.n_groups; ++g) {
            mbar_wait(smem_u32(&a_empty[ia]), pa ^ 1, P.err_flag, 1);
            const uint32_t fb = smem_u32(&a_full[ia]);
            mbar_expect_tx(fb, a_bytes);
            const uint32_t dst = smem_u32(a_ring + (size_t)ia * P.a_stage_bytes);
            const long long plane0 = (long long)s * P.a_nchunks + (long long)g * kc_chunks;
            for (int jc = 0; jc < kc_chunks; ++jc) {
              const char* src = P.a_base + (((plane0 + jc) * P.a_P + row0) << 4);
              bulk_g2s(dst + jc * a_slab_chunk_bytes, src, a_slab_chunk_bytes, fb);
            }
            if (++ia == P.na_stages) { ia = 0; pa ^= 1; }
            for (int q = 0; q < n_shift; ++q) {
              mbar_wait(smem_u32(&b_empty[ib]), pb ^ 1, P.err_flag, 2);
              const uint32_t fbb = smem_u32(&b_full[ib]);
              mbar_expect_tx(fbb, b_bytes);
              const long long img = (((long long)n_tile * P.n_src + s) * P.n_groups + g) * n_shift + q;
              bulk_g2s(smem_u32(b_ring + (size_t)ib * P.b_stage_bytes), P.w + img * b_bytes, b_bytes, fbb);
              if (++ib == P.nb_stages) { ib = 0; pb ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer
    if (lane == 0) {
      const uint32_t idesc = make_idesc(kTileM, P.N_TILE);
      uint32_t lbo_a = a_slab_chunk_bytes, sbo_a = 128;
      uint32_t lbo_b = (uint32_t)P.N_TILE * 16u, sbo_b = 128;
      if (P.dbg_swap_lbo) {
        uint32_t t = lbo_a; lbo_a = sbo_a; sbo_a = t;
        t = lbo_b; lbo_b = sbo_b; sbo_b = t;
      }
      int ia = 0, ib = 0, it = 0;
      uint32_t pa = 0, pb = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1;
        mbar_wait(smem_u32(&t_empty[as]), ((it >> 1) & 1) ^ 1, P.err_flag, 3);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * kAccumStride;
        uint32_t accumulate = 0;
        for (int s = 0; s < P.n_src; ++s) {
          for (int g = 0; g < P.n_groups; ++g) {
            mbar_wait(smem_u32(&a_full[ia]), pa, P.err_flag, 4);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(a_ring + (size_t)ia * P.a_stage_bytes);
            for (int q = 0; q < n_shift; ++q) {
              mbar_wait(smem_u32(&b_full[ib]), pb, P.err_flag, 5);
              tc_fence_after();
              const uint32_t b_addr = smem_u32(b_ring + (size_t)ib * P.b_stage_bytes);
              for (int kk = 0; kk < (P.KC >> 4); ++kk) {
                const uint64_t ad = make_desc(a_addr + q * 16 + kk * 2 * a_slab_chunk_bytes, lbo_a, sbo_a);
                const uint64_t bd = make_desc(b_addr + kk * 2 * (uint32_t)P.N_TILE * 16u, lbo_b, sbo_b);
                tc_mma_bf16(d_tmem, ad, bd, idesc, accumulate);
                accumulate = 1;
              }
              tc_commit(smem_u32(&b_empty[ib]));
              if (++ib == P.nb_stages) { ib = 0; pb ^= 1; }
            }
            tc_commit(smem_u32(&a_empty[ia]));
            if (++ia == P.na_stages) { ia = 0; pa ^= 1; }
          }
        }
        tc_commit(smem_u32(&t_full[as]));
      }
    }
  } else {
    // ============================== epilogue (warps 2..5; TMEM lane quarter = warp % 4)
    const int quarter = warp & 3;
    const int row_in_tile = quarter * 32 + lane;
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int as = it & 1;
      const int n_tile = tile % P.n_tiles_n;
      const int m_tile = tile / P.n_tiles_n;
      const int n0 = n_tile * P.N_TILE;
      const int phase = n0 / P.C_out;
      const int co0 = n0 - phase * P.C_out;

      const int g = m_tile * kTileM + row_in_tile;
      const int b = g / P.Lh_A;
      const int m = g - b * P.Lh_A;
      const bool valid = (g < P.m_rows) && (m < P.M_valid);
      const int l_out = P.up * m + phase;

      RowAddr ra1 = row_addr(P.add1, b, l_out);
      RowAddr ra2 = row_addr(P.add2, b, l_out);
      RowAddr ro[3];
#pragma unroll
      for (int o = 0; o < 3; ++o) {
        if (o < P.n_out) ro[o] = row_addr(P.out[o].t, b, l_out);
        else { ro[o].p = nullptr; ro[o].pstride = 0; }
      }

      mbar_wait(smem_u32(&t_full[as]), (it >> 1) & 1, P.err_flag, 6);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + as * kAccumStride + ((uint32_t)(quarter * 32) << 16);

      for (int c0 = 0; c0 < P.N_TILE; c0 += 16) {
        float acc[16];
        tc_ld16(t_addr + c0, acc);
        if (valid) {
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            float* v = acc + 8 * h;
            const int co = co0 + c0 + 8 * h;  // channel of v[0]
            const int chunk = co >> 3;
            if (P.bias) {
              const float4 b0 = *reinterpret_cast<const float4*>(P.bias + co);
              const float4 b1 = *reinterpret_cast<const float4*>(P.bias + co + 4);
              v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
              v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
            }
            if (ra1.p) {
              float t[8];
              load8(ra1, P.add1.esize_log2, chunk, t);
#pragma unroll
              for (int i = 0; i < 8; ++i) v[i] += t[i];
            }
            if (ra2.p) {
              float t[8];
              load8(ra2, P.add2.esize_log2, chunk, t);
#pragma unroll
              for (int i = 0; i < 8; ++i) v[i] += t[i];
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] *= P.scale;
            if (P.film) {
#pragma unroll
              for (int i = 0; i < 8; ++i) v[i] = P.film[co + i] * v[i] + P.film[P.C_out + co + i];
            }
            if (P.rm_out) {
              float* dst = P.rm_out + (long long)g * P.rm_ld + n0 + c0 + 8 * h;
              *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
              *reinterpret_cast<float4*>(dst + 4) = make_float4(v[4], v[5], v[6], v[7]);
            }
#pragma unroll
            for (int o = 0; o < 3; ++o) {
              if (ro[o].p) {
                float w[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) w[i] = apply_act(v[i], P.out[o].act, P.out[o].a1, P.out[o].a2);
                store8(ro[o], P.out[o].t.esize_log2, chunk, w);
              }
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&t_empty[as]));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2u * kAccumStride) : "memory");
  }
}

// ------------------------------------------------------------------ host side
inline size_t gemm_smem_bytes(const GemmParams& p) {
  return (size_t)p.na_stages * p.a_stage_bytes + (size_t)p.nb_stages * p.b_stage_bytes +
         (2 * kMaxAStages + 2 * kMaxBStages + 4) * sizeof(uint64_t) + 16;
}

// Fills the pipeline fields of `p` (stage sizes / counts) for a 227 KB budget.
inline void gemm_plan_pipeline(GemmParams& p) {
  const int rows_a = kTileM + p.smax - p.smin;
  int a_bytes = (p.KC / 8) * rows_a * 16;
  int b_bytes = p.N_TILE * p.KC * 2;
  p.a_stage_bytes = (a_bytes + 127) & ~127;
  p.b_stage_bytes = (b_bytes + 127) & ~127;
  const int budget = 200 * 1024;
  int na = 3, nb = 6;
  while (na * p.a_stage_bytes + nb * p.b_stage_bytes > budget && nb > 2) --nb;
  while (na * p.a_stage_bytes + nb * p.b_stage_bytes > budget && na > 2) --na;
  p.na_stages = na;
  p.nb_stages = nb;
  p.n_tiles_m = (p.m_rows + kTileM - 1) / kTileM;
}

inline cudaError_t gemm_launch(GemmParams p, int num_sms, cudaStream_t stream) {
  gemm_plan_pipeline(p);
  const size_t smem = gemm_smem_bytes(p);
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  const int total = p.n_tiles_m * p.n_tiles_n;
  int grid = total < num_sms ? total : num_sms;
  if (grid < 1) grid = 1;
  gemm_tc_kernel<<<grid, kGemmThreads, smem, stream>>>(p);
  return cudaGetLastError();
}

}  // namespace uv

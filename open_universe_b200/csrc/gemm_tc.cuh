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
// Warp roles (320 threads): warp0 = bulk-copy producer, warp1 = MMA issuer and
// TMEM owner, warps 2-9 = two epilogue groups that take alternate tiles (TMEM ->
// registers -> fused bias / add / FiLM / PReLU -> coalesced 16-byte stores into up
// to three planes tensors; the add operands of the next 16-column group are
// prefetched while the current one is processed).  Up to four accumulator stages
// live in TMEM so the MMAs run ahead of the epilogues.  Small layers keep their
// whole weight set resident in shared memory and give the rest to a deep A ring
// (the memory latency of these HBM-bound layers needs many tiles in flight).
// Persistent grid, static round-robin tile schedule.
#pragma once
#include "uv_common.cuh"

namespace uv {

constexpr int kMaxAStages = 16;
constexpr int kMaxBStages = 8;
constexpr int kGemmThreads = 320;   // producer warp, MMA warp, 2 x 4 epilogue warps
constexpr int kMaxAcc = 4;          // TMEM accumulator stages

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
struct AddBuf {
  uint4 a1[4];  // 16 columns of add1: 2 x uint4 (bf16) or 4 x uint4 (f32)
  uint4 a2[4];
};

__device__ __forceinline__ void add_prefetch(const GemmParams& P, bool valid, int b, int l_out, int chunk0, AddBuf& buf) {
  if (!valid) return;
  if (P.add1.base) {
    const RowAddr r = row_addr(P.add1, b, l_out);
    if (P.add1.esize_log2 == 1) {
      buf.a1[0] = *reinterpret_cast<const uint4*>(r.p + (long long)chunk0 * r.pstride);
      buf.a1[1] = *reinterpret_cast<const uint4*>(r.p + (long long)(chunk0 + 1) * r.pstride);
    } else {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const char* p = r.p + (long long)(chunk0 + h) * r.pstride;
        buf.a1[2 * h] = *reinterpret_cast<const uint4*>(p);
        buf.a1[2 * h + 1] = *reinterpret_cast<const uint4*>(p + 16);
      }
    }
  }
  if (P.add2.base) {
    const RowAddr r = row_addr(P.add2, b, l_out);
    if (P.add2.esize_log2 == 1) {
      buf.a2[0] = *reinterpret_cast<const uint4*>(r.p + (long long)chunk0 * r.pstride);
      buf.a2[1] = *reinterpret_cast<const uint4*>(r.p + (long long)(chunk0 + 1) * r.pstride);
    } else {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const char* p = r.p + (long long)(chunk0 + h) * r.pstride;
        buf.a2[2 * h] = *reinterpret_cast<const uint4*>(p);
        buf.a2[2 * h + 1] = *reinterpret_cast<const uint4*>(p + 16);
      }
    }
  }
}

// adds the 8 channels of half `h` (0/1) of a prefetched 16-column group to v
__device__ __forceinline__ void add_apply(const uint4* a, int esize_log2, int h, float* v) {
  if (esize_log2 == 1) {
    const __nv_bfloat162* q = reinterpret_cast<const __nv_bfloat162*>(&a[h]);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __bfloat1622float2(q[i]);
      v[2 * i] += f.x;
      v[2 * i + 1] += f.y;
    }
  } else {
    const float* f = reinterpret_cast<const float*>(&a[2 * h]);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] += f[i];
  }
}

struct TileCtx {
  int b, l_out, g, n0, co0;
  bool valid;
};

__device__ __forceinline__ TileCtx tile_ctx(const GemmParams& P, int tile, int row_in_tile) {
  TileCtx c;
  const int n_tile = tile % P.n_tiles_n;
  const int m_tile = tile / P.n_tiles_n;
  c.n0 = n_tile * P.N_TILE;
  const int phase = c.n0 / P.C_out;
  c.co0 = c.n0 - phase * P.C_out;
  c.g = m_tile * kTileM + row_in_tile;
  c.b = c.g / P.Lh_A;
  const int m = c.g - c.b * P.Lh_A;
  c.valid = (c.g < P.m_rows) && (m < P.M_valid);
  c.l_out = P.up * m + phase;
  return c;
}

// One 16-column group of the epilogue: TMEM -> registers -> fused math -> stores.
__device__ __forceinline__ void epilogue_group(const GemmParams& P, const TileCtx& c, uint32_t t_addr, int c0,
                                               const AddBuf& buf, const RowAddr* ro) {
  float acc[16];
  tc_ld16(t_addr + c0, acc);
  if (!c.valid) return;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    float* v = acc + 8 * h;
    const int co = c.co0 + c0 + 8 * h;
    const int chunk = co >> 3;
    if (P.bias) {
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(P.bias + co));
      const float4 b1 = __ldg(reinterpret_cast<const float4*>(P.bias + co + 4));
      v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
      v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
    }
    if (P.add1.base) add_apply(buf.a1, P.add1.esize_log2, h, v);
    if (P.add2.base) add_apply(buf.a2, P.add2.esize_log2, h, v);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] *= P.scale;
    if (P.film) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(P.film + co));
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(P.film + co + 4));
      const float4 e0 = __ldg(reinterpret_cast<const float4*>(P.film + P.C_out + co));
      const float4 e1 = __ldg(reinterpret_cast<const float4*>(P.film + P.C_out + co + 4));
      v[0] = fmaf(g0.x, v[0], e0.x); v[1] = fmaf(g0.y, v[1], e0.y); v[2] = fmaf(g0.z, v[2], e0.z); v[3] = fmaf(g0.w, v[3], e0.w);
      v[4] = fmaf(g1.x, v[4], e1.x); v[5] = fmaf(g1.y, v[5], e1.y); v[6] = fmaf(g1.z, v[6], e1.z); v[7] = fmaf(g1.w, v[7], e1.w);
    }
    if (P.rm_out) {
      float* dst = P.rm_out + (long long)c.g * P.rm_ld + c.n0 + c0 + 8 * h;
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

__global__ void __launch_bounds__(kGemmThreads, 1) gemm_tc_kernel(const __grid_constant__ GemmParams P) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // smem carve-up: A ring | B ring (or resident weight set) | barriers | tmem ptr
  uint8_t* a_ring = smem;
  uint8_t* b_ring = a_ring + (size_t)P.na_stages * P.a_stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(b_ring + (size_t)P.nb_stages * P.b_stage_bytes);
  uint64_t* a_full = bars;
  uint64_t* a_empty = a_full + kMaxAStages;
  uint64_t* b_full = a_empty + kMaxAStages;
  uint64_t* b_empty = b_full + kMaxBStages;
  uint64_t* t_full = b_empty + kMaxBStages;
  uint64_t* t_empty = t_full + kMaxAcc;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(t_empty + kMaxAcc);

  const uint32_t acc_stride = accum_stride_cols(P.N_TILE);
  const int n_acc = P.n_acc;
  if (threadIdx.x == 0) {
    for (int i = 0; i < P.na_stages; ++i) { mbar_init(smem_u32(&a_full[i]), 1); mbar_init(smem_u32(&a_empty[i]), 1); }
    for (int i = 0; i < kMaxBStages; ++i) { mbar_init(smem_u32(&b_full[i]), 1); mbar_init(smem_u32(&b_empty[i]), 1); }
    for (int i = 0; i < kMaxAcc; ++i) { mbar_init(smem_u32(&t_full[i]), 1); mbar_init(smem_u32(&t_empty[i]), 128); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)),
                 "r"((uint32_t)n_acc * acc_stride)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  const int n_shift = P.smax - P.smin + 1;
  const int rows_a = kTileM + P.smax - P.smin;  // halo'd slab rows
  const int kc_chunks = P.KC >> 3;
  const int total_tiles = P.n_tiles_m * P.n_tiles_n;
  const uint32_t a_slab_chunk_bytes = (uint32_t)rows_a * 16u;
  const uint32_t a_bytes = a_slab_chunk_bytes * kc_chunks;
  const uint32_t b_bytes = (uint32_t)P.N_TILE * P.KC * 2u;
  const int n_img = P.n_src * P.n_groups * n_shift;  // weight images per N tile

  if (warp == 0) {
    // ============================== producer
    if (lane == 0) {
      if (P.b_resident) {
        const uint32_t fbb = smem_u32(&b_full[0]);
        const uint32_t total = (uint32_t)n_img * b_bytes;
        mbar_expect_tx(fbb, total);
        for (uint32_t off = 0; off < total; off += 32768u) {
          const uint32_t n = total - off < 32768u ? total - off : 32768u;
          bulk_g2s(smem_u32(b_ring + off), P.w + off, n, fbb);
        }
      }
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
            if (!P.b_resident) {
              for (int q = 0; q < n_shift; ++q) {
                mbar_wait(smem_u32(&b_empty[ib]), pb ^ 1, P.err_flag, 2);
                const uint32_t fbb = smem_u32(&b_full[ib]);
                mbar_expect_tx(fbb, b_bytes);
                const long long img = ((long long)n_tile * P.n_src + s) * P.n_groups * n_shift + (long long)g * n_shift + q;
                bulk_g2s(smem_u32(b_ring + (size_t)ib * P.b_stage_bytes), P.w + img * b_bytes, b_bytes, fbb);
                if (++ib == P.nb_stages) { ib = 0; pb ^= 1; }
              }
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
      if (P.b_resident) {
        mbar_wait(smem_u32(&b_full[0]), 0, P.err_flag, 7);
        tc_fence_after();
      }
      int ia = 0, ib = 0, it = 0;
      uint32_t pa = 0, pb = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const int as = it % n_acc;
        mbar_wait(smem_u32(&t_empty[as]), ((it / n_acc) & 1) ^ 1, P.err_flag, 3);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * acc_stride;
        uint32_t accumulate = 0;
        for (int s = 0; s < P.n_src; ++s) {
          for (int g = 0; g < P.n_groups; ++g) {
            mbar_wait(smem_u32(&a_full[ia]), pa, P.err_flag, 4);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(a_ring + (size_t)ia * P.a_stage_bytes);
            for (int q = 0; q < n_shift; ++q) {
              uint32_t b_addr;
              if (P.b_resident) {
                b_addr = smem_u32(b_ring) + (uint32_t)((s * P.n_groups + g) * n_shift + q) * b_bytes;
              } else {
                mbar_wait(smem_u32(&b_full[ib]), pb, P.err_flag, 5);
                tc_fence_after();
                b_addr = smem_u32(b_ring + (size_t)ib * P.b_stage_bytes);
              }
              for (int kk = 0; kk < (P.KC >> 4); ++kk) {
                const uint64_t ad = make_desc(a_addr + q * 16 + kk * 2 * a_slab_chunk_bytes, lbo_a, sbo_a);
                const uint64_t bd = make_desc(b_addr + kk * 2 * (uint32_t)P.N_TILE * 16u, lbo_b, sbo_b);
                tc_mma_bf16(d_tmem, ad, bd, idesc, accumulate);
                accumulate = 1;
              }
              if (!P.b_resident) {
                tc_commit(smem_u32(&b_empty[ib]));
                if (++ib == P.nb_stages) { ib = 0; pb ^= 1; }
              }
            }
            tc_commit(smem_u32(&a_empty[ia]));
            if (++ia == P.na_stages) { ia = 0; pa ^= 1; }
          }
        }
        tc_commit(smem_u32(&t_full[as]));
      }
    }
  } else {
    // ============================== epilogue: 2 groups x 4 warps, alternate tiles; TMEM lane quarter = warp % 4
    const int egroup = (warp - 2) >> 2;
    const int quarter = warp & 3;
    const int row_in_tile = quarter * 32 + lane;
    const int ngroups = P.N_TILE >> 4;
    int it = egroup;
    int tile = blockIdx.x + it * gridDim.x;
    int cg = 0;
    TileCtx cur;
    RowAddr ro[3];
    uint32_t t_addr = 0;
    AddBuf bufA, bufB;
    bool have = tile < total_tiles;
    if (have) {
      cur = tile_ctx(P, tile, row_in_tile);
      add_prefetch(P, cur.valid, cur.b, cur.l_out, cur.co0 >> 3, bufA);
    }
    // processes the current (tile, column group) with `bc`, prefetching the next one into `bn`
    auto step = [&](const AddBuf& bc, AddBuf& bn) -> bool {
      int n_it = it, n_cg = cg + 1, n_tile = tile;
      if (n_cg == ngroups) {
        n_cg = 0;
        n_it = it + 2;
        n_tile = blockIdx.x + n_it * gridDim.x;
      }
      const bool n_have = n_tile < total_tiles;
      TileCtx nxt = cur;
      if (n_have) {
        if (n_cg == 0) nxt = tile_ctx(P, n_tile, row_in_tile);
        add_prefetch(P, nxt.valid, nxt.b, nxt.l_out, (nxt.co0 + 16 * n_cg) >> 3, bn);
      }
      const int as = it % n_acc;
      if (cg == 0) {
#pragma unroll
        for (int o = 0; o < 3; ++o) {
          if (o < P.n_out && cur.valid) ro[o] = row_addr(P.out[o].t, cur.b, cur.l_out);
          else { ro[o].p = nullptr; ro[o].pstride = 0; }
        }
        mbar_wait(smem_u32(&t_full[as]), (it / n_acc) & 1, P.err_flag, 6);
        tc_fence_after();
        t_addr = tmem_base + as * acc_stride + ((uint32_t)(quarter * 32) << 16);
      }
      epilogue_group(P, cur, t_addr, cg * 16, bc, ro);
      if (cg == ngroups - 1) {
        tc_fence_before();
        mbar_arrive(smem_u32(&t_empty[as]));
      }
      it = n_it; cg = n_cg; tile = n_tile; cur = nxt;
      return n_have;
    };
    while (have) {
      have = step(bufA, bufB);
      if (!have) break;
      have = step(bufB, bufA);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)n_acc * acc_stride)
                 : "memory");
  }
}

// ------------------------------------------------------------------ host side
inline size_t gemm_smem_bytes(const GemmParams& p) {
  return (size_t)p.na_stages * p.a_stage_bytes + (size_t)p.nb_stages * p.b_stage_bytes +
         (2 * kMaxAStages + 2 * kMaxBStages + 2 * kMaxAcc) * sizeof(uint64_t) + 16;
}

// Fills the pipeline fields of `p` (stage sizes / counts, resident-weight mode) for the 227 KB smem of an SM.
inline void gemm_plan_pipeline(GemmParams& p) {
  const int rows_a = kTileM + p.smax - p.smin;
  const int a_bytes = (p.KC / 8) * rows_a * 16;
  const int b_bytes = p.N_TILE * p.KC * 2;
  p.a_stage_bytes = (a_bytes + 127) & ~127;
  p.n_tiles_m = (p.m_rows + kTileM - 1) / kTileM;
  p.n_acc = accum_stride_cols(p.N_TILE) <= 128 ? 4 : 2;
  const int budget = 216 * 1024;
  const int n_img = p.n_src * p.n_groups * (p.smax - p.smin + 1);
  const long long total_b = (long long)n_img * b_bytes;
  if (p.n_tiles_n == 1 && total_b <= 112 * 1024 && (budget - total_b) / p.a_stage_bytes >= 3) {
    p.b_resident = 1;
    p.b_stage_bytes = b_bytes;
    p.nb_stages = n_img;
    int na = (int)((budget - total_b) / p.a_stage_bytes);
    p.na_stages = na > kMaxAStages ? kMaxAStages : na;
    return;
  }
  p.b_resident = 0;
  p.b_stage_bytes = (b_bytes + 127) & ~127;
  int na = 3, nb = 6;
  while (na * p.a_stage_bytes + nb * p.b_stage_bytes > budget && nb > 2) --nb;
  while (na * p.a_stage_bytes + nb * p.b_stage_bytes > budget && na > 2) --na;
  while (na < 6 && (na + 1) * p.a_stage_bytes + nb * p.b_stage_bytes <= budget) ++na;
  p.na_stages = na;
  p.nb_stages = nb;
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

// Fused ConvBlock for the wide, narrow-channel levels (C <= 64): conv1 (k5) -> [+cond]/sqrt2 -> [FiLM] ->
// PReLU -> conv2 (k3) -> PReLU -> conv3 (k3) -> (h + c)/sqrt2 in ONE kernel
// (reference: open_universe/networks/universe/blocks.py:387-403).
//
// Run as three GEMM launches these layers are HBM bound: each conv reads and writes a full
// [rows][C] activation (263 MB at levels 0/1 of the benchmark), 3 reads + 3-4 writes per block.  Here a
// tile of 252 output rows keeps the two intermediate activations in shared memory: HBM sees h_act, h (the
// residual), the optional condition, and the block outputs only.
//
// Tile geometry (flat rows of the level's row space, r0 = 252 * tile):
//   buffer rows j = 0..259  <- h_act rows r0-4+j                 (TMA bulk copies, one per channel chunk)
//   conv1 row i = 0..255 = flat r0-2+i, taps read buffer rows i..i+4;  c1 written back to buffer row i
//   conv2 row i          = flat r0-1+i, taps read buffer rows i..i+2;  c2 written back to buffer row i
//   conv3 row i          = flat r0+i,   taps read buffer rows i..i+2;  rows i < 252 are this tile's output
// Every stage is two M=128 tcgen05 MMAs chains (rows 0-127 / 128-255) with the accumulators in tensor
// memory; rows that are padding (between utterances, before 0, past the end) are written back as zeros,
// which is exactly the "same" padding each conv of the reference sees.
//
// Roles (320 threads): warp 0 = TMA producer, warp 1 = MMA issuer, warps 2-9 = epilogue (TMEM lane quarter x
// half).  A tile is a chain MMA -> epilogue -> MMA -> epilogue -> MMA -> epilogue with a barrier hand-off
// (several hundred cycles) at every arrow, so n_slots tiles are in flight and both roles walk them stage
// by stage: the hand-off latency of one tile is covered by the work on the others.  A slot is one
// accumulator pair in tensor memory (2C columns); row buffers rotate independently (n_bufs >= n_slots).
#pragma once
#include "gemm_tc.cuh"

namespace uv {

constexpr int kCbEpiGroups = 1;      // groups of 8 epilogue warps; group g serves slots g, g + 2, ...
constexpr int kCbThreads = 64 + 256 * kCbEpiGroups;
constexpr int kCbRowsOut = 252;       // output rows per tile
constexpr int kCbBufRows = 260;       // buffer rows (256 + 4 halo)
constexpr int kCbMaxSlots = 6;
constexpr int kCbMaxBufs = 8;
constexpr int kCbMaxC = 64;

struct CbParams {
  const char* x_base;   // h_act: bf16 planes, split 1, C/8 chunks
  int x_P, x_G;
  int C;                // channels in = out of all three convs (multiple of 16, <= 64)
  const char* w[3];     // packed images [tap][C/8][C][8] bf16, then (has_bias) the bias image [2][C][8]
  int has_bias[3];
  int m_rows, Lh, M_valid, n_tiles;
  int n_slots, n_bufs;  // tiles in flight / row buffers (host: cb_plan)
  TensDesc cond;        // optional bf16 tensor added to conv1's output
  float scale1;         // applied after the condition add
  const float* film;    // optional [2C] gamma|beta
  OutDesc cond_out;     // optional output of stage 1 (after add / FiLM, before the activation)
  float slope2, slope3; // PReLU slopes in front of conv2 / conv3
  TensDesc h_raw;       // residual of the block output
  float scale3;
  int n_out;
  OutDesc out[3];
  int* err_flag;
};

struct CbSmem {
  uint32_t buf_bytes, w_off[3], w_bytes[3], film_off, bufs_off, bars_off, total;
};
__host__ __device__ inline CbSmem cb_smem_layout(int C, const int* has_bias, int n_bufs) {
  CbSmem L;
  uint32_t off = kOnesBytes;
  L.film_off = off;
  off += 2 * kCbMaxC * 4;
  for (int s = 0; s < 3; ++s) {
    const int taps = s == 0 ? 5 : 3;
    L.w_off[s] = off;
    L.w_bytes[s] = (uint32_t)taps * C * C * 2 + (has_bias[s] ? (uint32_t)C * 32u : 0u);
    off += (L.w_bytes[s] + 127u) & ~127u;
  }
  L.buf_bytes = (uint32_t)kCbBufRows * (C / 8) * 16u;
  L.bufs_off = (off + 1023u) & ~1023u;
  L.bars_off = L.bufs_off + (uint32_t)n_bufs * L.buf_bytes;
  L.total = L.bars_off + 256;
  return L;
}

// stage-s epilogue of one warp on one tile: TMEM -> registers -> (stage math) -> shared / global
// Row of the bf16 operand a (tile, stage) item adds in its epilogue (stage 1: the condition, stage 3: the
// residual h), fetched one work item ahead into registers: a DRAM round trip is longer than the item.
__device__ __forceinline__ void cb_prefetch(const CbParams& P, int st, long long r0, int i, float inv_Lh, uint4* pre) {
  const TensDesc& t = st == 0 ? P.cond : P.h_raw;
  if (st == 1 || t.base == nullptr) return;
  const long long g = r0 + i - (2 - st);
  if (g < 0 || g >= P.m_rows || (st == 2 && i >= kCbRowsOut)) return;
  const int b = fast_div((int)g, P.Lh, inv_Lh);
  if ((int)g - b * P.Lh >= P.M_valid) return;
  const char* p = t.base + ((long long)t.G + g) * 16;
  const long long stride = (long long)t.P * 16;
#pragma unroll
  for (int c = 0; c < kCbMaxC / 8; ++c)
    if (c < (P.C >> 3)) pre[c] = *reinterpret_cast<const uint4*>(p + c * stride);
}

template <int ST, bool LE1>
__device__ __forceinline__ void cb_epilogue_stage(const CbParams& P, uint32_t taddr, uint8_t* buf, uint32_t plane_bytes,
                                                  const float* film_s, int i, long long r0, float inv_Lh, const uint4* pre) {
  const int C = P.C, ngroups = C >> 4;
  const long long g = r0 + i - (2 - ST);          // flat row of accumulator row i
  bool valid = g >= 0 && g < P.m_rows;
  int b = 0, m = 0;
  if (valid) {
    b = fast_div((int)g, P.Lh, inv_Lh);
    m = (int)g - b * P.Lh;
    valid = m < P.M_valid;
  }
  uint32_t r[2][16];
  tc_ld16_issue(taddr, r[0]);
  if (ST < 2) {
    // ---- stage 1 / 2: next conv's A operand (bf16, zero on padding rows) back into the row buffer
    const bool has_cond = ST == 0 && P.cond.base != nullptr;
    const bool has_film = ST == 0 && P.film != nullptr;
    const bool own = valid && i >= 2 && i < 2 + kCbRowsOut;   // cond_out rows written by this tile
    OutPtr po;
    po.p = nullptr; po.pstride = 0;
    if (ST == 0 && P.cond_out.t.base && own) po = out_ptr(P.cond_out.t, b, m, 0, 1.f);
    const float slope = ST == 0 ? P.slope2 : P.slope3;
    uint8_t* dst = buf + (size_t)i * 16;
#pragma unroll
    for (int cg = 0; cg < kCbMaxC / 16; ++cg) {   // fully unrolled: the register double buffer needs static indices
      if (cg >= ngroups) break;
      tc_ld_wait();
      if (cg + 1 < ngroups) tc_ld16_issue(taddr + 16u * (cg + 1), r[(cg + 1) & 1]);
      float v[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) v[k] = __uint_as_float(r[cg & 1][k]);
      if (has_cond) {
        const uint4 cc[2] = {pre[2 * cg], pre[2 * cg + 1]};   // (rows that are padding are zeroed below)
        add_unpack(cc, v);
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] *= P.scale1;
      }
      if (has_film) {
        const float4* ga = reinterpret_cast<const float4*>(film_s + 16 * cg);
        const float4* be = reinterpret_cast<const float4*>(film_s + C + 16 * cg);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 gq = ga[q], bq = be[q];
          v[4 * q] = fmaf(gq.x, v[4 * q], bq.x);
          v[4 * q + 1] = fmaf(gq.y, v[4 * q + 1], bq.y);
          v[4 * q + 2] = fmaf(gq.z, v[4 * q + 2], bq.z);
          v[4 * q + 3] = fmaf(gq.w, v[4 * q + 3], bq.w);
        }
      }
      if (po.p) {
        store16_act<LE1>(po.p, po.pstride, P.cond_out.t.esize_log2 == 2, 1.f, false, 1.f, v);
        po.p += 2 * po.pstride;
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint4 u;
        __nv_bfloat162* q = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          q[k] = __floats2bfloat162_rn(prelu_t<LE1>(v[8 * h + 2 * k], slope), prelu_t<LE1>(v[8 * h + 2 * k + 1], slope));
        if (!valid) u = make_uint4(0, 0, 0, 0);
        *reinterpret_cast<uint4*>(dst + (size_t)(2 * cg + h) * plane_bytes) = u;
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic stores -> tensor-core reads
  } else {
    // ---- stage 3: residual, scale, block outputs
    const bool out_row = valid && i < kCbRowsOut;
    OutPtr op[3];
#pragma unroll
    for (int o = 0; o < 3; ++o) {
      op[o].p = nullptr; op[o].pstride = 0;
      if (o < P.n_out && out_row) op[o] = out_ptr(P.out[o].t, b, m, 0, 1.f / (float)(P.out[o].t.split > 0 ? P.out[o].t.split : 1));
    }
#pragma unroll
    for (int cg = 0; cg < kCbMaxC / 16; ++cg) {
      if (cg >= ngroups) break;
      tc_ld_wait();
      if (cg + 1 < ngroups) tc_ld16_issue(taddr + 16u * (cg + 1), r[(cg + 1) & 1]);
      float v[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) v[k] = __uint_as_float(r[cg & 1][k]);
      const uint4 hh[2] = {pre[2 * cg], pre[2 * cg + 1]};
      if (out_row) {
        if (P.h_raw.base) add_unpack(hh, v);
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] *= P.scale3;
#pragma unroll
        for (int o = 0; o < 3; ++o) {
          if (op[o].p) {
            store16_act<LE1>(op[o].p, op[o].pstride, P.out[o].t.esize_log2 == 2, P.out[o].act >= 1 ? P.out[o].a1 : 1.f,
                             P.out[o].act == 2, P.out[o].a2, v);
            op[o].p += 2 * op[o].pstride;
          }
        }
      }
    }
  }
}

template <bool LE1>
__device__ __forceinline__ void cb_epilogue_role(const CbParams& P, const CbSmem& L, uint8_t* smem, uint32_t tmem_base,
                                                 uint64_t* mma_done, uint64_t* epi_done, int my_tiles, int warp, int lane) {
  const int sub = ((warp - 2) >> 2) & 1, quarter = warp & 3;
  const int i = sub * kSubM + quarter * 32 + lane;
  const uint32_t plane_bytes = (uint32_t)kCbBufRows * 16u;
  const float inv_Lh = 1.f / (float)P.Lh;
  const float* film_s = reinterpret_cast<const float*>(smem + L.film_off);
  const uint32_t C = (uint32_t)P.C;
  const int NS = P.n_slots, NB = P.n_bufs;
  const int grp = (warp - 2) >> 3;   // this group of 8 warps serves slots grp, grp + kCbEpiGroups, ...
  auto tile_r0 = [&](int k) { return (long long)(blockIdx.x + (long long)k * gridDim.x) * kCbRowsOut; };
  for (int k0 = 0, j = 0; k0 < my_tiles; k0 += NS, ++j) {   // round j: one tile per slot
    const int n2 = my_tiles - k0 < NS ? my_tiles - k0 : NS;
#pragma unroll
    for (int st = 0; st < 3; ++st) {
      for (int s = grp; s < n2; s += kCbEpiGroups) {
        const int k = k0 + s;
        const long long r0 = tile_r0(k);
        uint4 pre[kCbMaxC / 8];
#pragma unroll
        for (int c = 0; c < kCbMaxC / 8; ++c) pre[c] = make_uint4(0, 0, 0, 0);
        cb_prefetch(P, st, r0, i, inv_Lh, pre);   // in flight across the barrier wait and the first TMEM load
        uint8_t* buf = smem + L.bufs_off + (size_t)(k % NB) * L.buf_bytes;
        mbar_wait(smem_u32(&mma_done[s]), (uint32_t)(3 * j + st) & 1u, P.err_flag, 12);
        tc_fence_after();
        const uint32_t taddr = tmem_base + (uint32_t)(s * 2 + sub) * C + ((uint32_t)(quarter * 32) << 16);
        if (st == 0) cb_epilogue_stage<0, LE1>(P, taddr, buf, plane_bytes, film_s, i, r0, inv_Lh, pre);
        else if (st == 1) cb_epilogue_stage<1, LE1>(P, taddr, buf, plane_bytes, film_s, i, r0, inv_Lh, pre);
        else cb_epilogue_stage<2, LE1>(P, taddr, buf, plane_bytes, film_s, i, r0, inv_Lh, pre);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&epi_done[s]));
      }
    }
  }
}

__global__ void __launch_bounds__(kCbThreads, 1) convblock_kernel(const __grid_constant__ CbParams P) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const CbSmem L = cb_smem_layout(P.C, P.has_bias, P.n_bufs);
  const int NS = P.n_slots, NB = P.n_bufs;
  uint8_t* ones = smem;
  float* film_s = reinterpret_cast<float*>(smem + L.film_off);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bars_off);
  uint64_t* x_full = bars;                      // [bufs] tile rows landed in buffer b
  uint64_t* x_empty = bars + kCbMaxBufs;        // [bufs] conv3 of the tile in buffer b has read it
  uint64_t* mma_done = bars + 2 * kCbMaxBufs;   // [slots] completes once per stage
  uint64_t* epi_done = mma_done + kCbMaxSlots;  // [slots] 8 epilogue warps, once per stage
  uint64_t* w_full = epi_done + kCbMaxSlots;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(w_full + 1);
  const int C = P.C;
  uint32_t tmem_cols = 32;
  while ((int)tmem_cols < NS * 2 * C) tmem_cols <<= 1;

  if (threadIdx.x == 0) {
    for (int b = 0; b < NB; ++b) { mbar_init(smem_u32(&x_full[b]), 1); mbar_init(smem_u32(&x_empty[b]), 1); }
    for (int s = 0; s < NS; ++s) { mbar_init(smem_u32(&mma_done[s]), 1); mbar_init(smem_u32(&epi_done[s]), 8); }
    mbar_init(smem_u32(w_full), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 256) {   // ones tile: K columns 0 / 1 meet the hi / lo halves of a bias image
    uint4 u = make_uint4(0u, 0u, 0u, 0u);
    if (threadIdx.x < kSubM) u.x = 0x3F803F80u;
    reinterpret_cast<uint4*>(ones)[threadIdx.x] = u;
  }
  if (P.film)
    for (int k = threadIdx.x; k < 2 * C; k += kCbThreads) film_s[k] = P.film[k];
  // the buffers' slack rows are read by the last taps of discarded accumulator rows: keep them finite
  for (uint32_t k = threadIdx.x; k < (uint32_t)NB * L.buf_bytes / 16; k += kCbThreads)
    reinterpret_cast<uint4*>(smem + L.bufs_off)[k] = make_uint4(0, 0, 0, 0);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  const int my_tiles = (int)blockIdx.x < P.n_tiles ? (P.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const uint32_t plane_bytes = (uint32_t)kCbBufRows * 16u;
  const int nch = C >> 3;

  if (warp == 0) {
    // ============================== producer: resident weights once, then one row buffer per tile
    if (elect_one()) {
      uint32_t wtot = 0;
      for (int s = 0; s < 3; ++s) wtot += L.w_bytes[s];
      mbar_expect_tx(smem_u32(w_full), wtot);
      for (int s = 0; s < 3; ++s) {
        for (uint32_t off = 0; off < L.w_bytes[s]; off += 32768u) {
          const uint32_t n = L.w_bytes[s] - off < 32768u ? L.w_bytes[s] - off : 32768u;
          bulk_g2s(smem_u32(smem + L.w_off[s] + off), P.w[s] + off, n, smem_u32(w_full));
        }
      }
      for (int k = 0; k < my_tiles; ++k) {
        const int b = k % NB;
        const uint32_t u = (uint32_t)(k / NB);
        mbar_wait(smem_u32(&x_empty[b]), (u & 1u) ^ 1u, P.err_flag, 10);
        const long long row = (long long)P.x_G + (long long)(blockIdx.x + (long long)k * gridDim.x) * kCbRowsOut - 4;
        long long nrows = (long long)P.x_P - row;     // never read past the plane (rows past m_rows are masked)
        if (nrows > kCbBufRows) nrows = kCbBufRows;
        const uint32_t bytes = (uint32_t)nrows * 16u;
        const uint32_t fb = smem_u32(&x_full[b]);
        mbar_expect_tx(fb, bytes * (uint32_t)nch);
        const uint32_t dst = smem_u32(smem + L.bufs_off + (size_t)b * L.buf_bytes);
        for (int c = 0; c < nch; ++c) bulk_g2s(dst + c * plane_bytes, P.x_base + (((long long)c * P.x_P + row) << 4), bytes, fb);
      }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer
    if (elect_one()) {
      const uint32_t idesc = make_idesc(kSubM, C);
      const uint32_t desc_hi = (128u >> 4) | (1u << 14);
      const uint32_t lbo_b = (uint32_t)C * 16u;
      const uint32_t a_kstep = (2u * plane_bytes) >> 4, b_kstep = (2u * lbo_b) >> 4;
      const uint32_t tap16 = ((uint32_t)C * C * 2u) >> 4;
      const uint64_t ones_desc = make_desc(smem_u32(ones), (uint32_t)kSubM * 16u, 128);
      auto dsc = [&](uint32_t lo) { return ((uint64_t)desc_hi << 32) | (uint64_t)lo; };
      const int ksteps = C >> 4;
      mbar_wait(smem_u32(w_full), 0, P.err_flag, 11);
      tc_fence_after();
      for (int k0 = 0, j = 0; k0 < my_tiles; k0 += NS, ++j) {
        const int n2 = my_tiles - k0 < NS ? my_tiles - k0 : NS;
#pragma unroll
        for (int st = 0; st < 3; ++st) {
          for (int s = 0; s < n2; ++s) {
            const int k = k0 + s, b = k % NB;
            if (st == 0) {
              mbar_wait(smem_u32(&x_full[b]), (uint32_t)(k / NB) & 1u, P.err_flag, 13);
              if (j > 0) mbar_wait(smem_u32(&epi_done[s]), (uint32_t)(3 * j - 1) & 1u, P.err_flag, 14);
            } else {
              mbar_wait(smem_u32(&epi_done[s]), (uint32_t)(3 * j + st - 1) & 1u, P.err_flag, 14);
            }
            tc_fence_after();
            const uint32_t d0 = tmem_base + (uint32_t)(s * 2) * (uint32_t)C, d1 = d0 + (uint32_t)C;
            const uint32_t a_lo = (smem_u32(smem + L.bufs_off + (size_t)b * L.buf_bytes) >> 4) | ((plane_bytes >> 4) << 16);
            uint32_t w_lo = (smem_u32(smem + L.w_off[st]) >> 4) | ((lbo_b >> 4) << 16);
            const int taps = st == 0 ? 5 : 3;
            uint32_t acc = 0;
            if (P.has_bias[st]) {
              const uint32_t bias_lo = w_lo + (uint32_t)taps * tap16;
              tc_mma_bf16(d0, ones_desc, dsc(bias_lo), idesc, 0);
              tc_mma_bf16(d1, ones_desc, dsc(bias_lo), idesc, 0);
              acc = 1;
            }
            for (int q = 0; q < taps; ++q) {
              uint32_t al = a_lo + (uint32_t)q, bl = w_lo;
              for (int kk = 0; kk < ksteps; ++kk) {
                tc_mma_bf16(d0, dsc(al), dsc(bl), idesc, acc);
                tc_mma_bf16(d1, dsc(al + (uint32_t)(kSubM * 16 >> 4)), dsc(bl), idesc, acc);
                acc = 1;
                al += a_kstep;
                bl += b_kstep;
              }
              w_lo += tap16;
            }
            tc_commit(smem_u32(&mma_done[s]));
            if (st == 2) tc_commit(smem_u32(&x_empty[b]));
          }
        }
      }
    }
  } else {
    bool le1 = P.slope2 <= 1.f && P.slope3 <= 1.f;
    for (int o = 0; o < 3; ++o)
      if (o < P.n_out && ((P.out[o].act >= 1 && !(P.out[o].a1 <= 1.f)) || (P.out[o].act == 2 && !(P.out[o].a2 <= 1.f)))) le1 = false;
    if (le1) cb_epilogue_role<true>(P, L, smem, tmem_base, mma_done, epi_done, my_tiles, warp, lane);
    else cb_epilogue_role<false>(P, L, smem, tmem_base, mma_done, epi_done, my_tiles, warp, lane);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
  }
}

// Tiles in flight: one accumulator pair (2C columns of 512) per slot, at most kCbMaxSlots; row buffers: one per
// slot plus one being loaded when shared memory allows.
inline bool cb_plan(CbParams& p) {
  if (p.C % 16 || p.C < 16 || p.C > kCbMaxC || p.x_G < 4) return false;
  int ns = 512 / (2 * p.C);
  if (ns > kCbMaxSlots) ns = kCbMaxSlots;
  for (; ns >= 1; --ns) {
    for (int nb = ns + 1 <= kCbMaxBufs ? ns + 1 : ns; nb >= ns; --nb) {
      if (cb_smem_layout(p.C, p.has_bias, nb).total <= 227 * 1024) {
        p.n_slots = ns;
        p.n_bufs = nb;
        return true;
      }
    }
  }
  return false;
}

inline cudaError_t convblock_set_attributes() {   // per device (uv_model_create)
  return cudaFuncSetAttribute(convblock_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
}

inline cudaError_t convblock_launch(CbParams p, int num_sms, cudaStream_t s) {
  if (!cb_plan(p)) return cudaErrorInvalidConfiguration;
  const CbSmem L = cb_smem_layout(p.C, p.has_bias, p.n_bufs);
  const int grid = p.n_tiles < num_sms ? p.n_tiles : num_sms;
  convblock_kernel<<<grid, kCbThreads, L.total, s>>>(p);
  return cudaGetLastError();
}

}  // namespace uv

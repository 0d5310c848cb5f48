// Fused ConvBlock for the HBM-bound levels (64 GEMM columns: C = 64, or C = 32 stored phase-split by 2):
//   x -> PReLU -> conv1 -> [+cond] * s1 -> [FiLM] -> PReLU -> conv2 -> PReLU -> conv3 -> (. + x) * s3 -> outputs
// (reference: open_universe/networks/universe/blocks.py:387-403) in ONE kernel on the SM pairs of gemm_tc.cuh.
//
// Run as three GEMM launches these layers move every activation through HBM 7-8 times per block (read x, write c1,
// read c1, write c2, read c2, read x again as the residual, write the outputs; + the condition) and each launch reaches
// 50-60 % of the HBM bandwidth.  Here a CTA keeps its 256-row slab of x (raw, for the residual) and a working buffer
// (PReLU(x) -> c1 -> c2, each written in place by the stage epilogues) in shared memory: HBM sees x, the optional
// condition and the block outputs only.
//
// Tile geometry (flat rows of the block's row space; every CTA of a pair owns 252 output rows, R0 = 252 (2 tile + rank)):
//   buffer row j = 0..259    <- x rows R0 - H + j          (H = h1 + h2 + h3, h_k = taps of conv k on either side)
//   conv1 row i = 0..255     = flat R0 - H + h1 + i,       reads buffer rows i .. i + 2 h1;  c1 -> working buffer row i
//   conv2 row i              = flat R0 - H + h1 + h2 + i,  reads rows i .. i + 2 h2;         c2 -> working buffer row i
//   conv3 row i              = flat R0 + i,                reads rows i .. i + 2 h3;  rows i < 252 are the tile's output
// Every stage is two M = 256 tcgen05.mma.cta_group::2 chains (sub-tiles of 128 rows per CTA) with the accumulators in
// tensor memory; rows that are padding (between utterances, before 0, past the end) are written back as zeros, which
// is exactly the "same" padding each conv of the reference sees.  The residual enters through an identity MMA on the
// resident raw slab, the bias through the constant ones tile (as in gemm_tc.cuh).
//
// Roles per CTA (384 threads, as gemm_tc.cuh): warp 0 = loads (x slabs; weights once), warp 1 = MMA issuer (leader CTA)
// / relay (peer), warps 2-9 = stage epilogues, warps 10-11 = PReLU(x) into the working buffer.  Two tiles are in
// flight per CTA (slots), their stages interleaved (t0 s1, t1 s1, t0 s2, t1 s2, ...) so that the tensor core works on
// one slot while the epilogue warps drain the other.  Tiles are claimed dynamically by the leader (same ring as
// gemm_tc.cuh).
#pragma once
#include "gemm_tc.cuh"

namespace uv {

constexpr int kCbN = 64;             // GEMM columns = channels of every tensor of the block (as 8 chunk planes)
constexpr int kCbPlanes = kCbN / 8;
constexpr int kCbRowsOut = 252;      // output rows per CTA and tile
constexpr int kCbCtaRows = 256;      // accumulator rows per CTA (two sub-tiles)
constexpr int kCbBufRows = 264;      // buffer rows (256 + halo, padded)
constexpr int kCbLoadRows = 260;
constexpr int kCbSlots = 2;          // tiles in flight per CTA: working buffer + accumulator pair each
constexpr int kCbXBufs = 3;          // raw x slabs: the loads run one tile ahead of the slots
constexpr int kCbXfWarps = 4;        // warps applying PReLU(x) (the GEMM kernel has 2: its register budget is tighter)
constexpr int kCbThreads = 64 + 32 * kEpiWarps + 32 * kCbXfWarps;
constexpr uint32_t kCbPlaneBytes = kCbBufRows * 16;
constexpr uint32_t kCbBufBytes = kCbPlanes * kCbPlaneBytes;     // 33 792
constexpr int kCbIdentBytes = kCbN * (kCbN / 2) * 2;            // one identity image (add_kc = 64), this CTA's half

struct CbParams {
  const char* x_base;   // input (raw): bf16 planes, 8 chunk planes (a plain 64-channel tensor, or 32 channels split by 2)
  int x_P, x_G;
  float a_slope;        // PReLU in front of conv1
  const char* w[3];     // packed like a GEMM layer: per half [src][shift][kc/8][32][8] images, then the bias image
  int has_bias[3];
  int n_src[3], n_shift[3];   // K sources (phase planes of the input: kc = 64 / n_src channels each) and taps per stage
  int m_rows, Lh, M_valid, n_tiles;
  const int* mv;        // optional per-utterance valid rows (variable-length calls), else M_valid
  TensDesc cond;        // optional bf16 tensor (same geometry as x) added to conv1's output
  float scale1;         // applied after the condition add
  const float* film;    // optional [2 film_C] gamma | beta; channel of column n = n % film_C
  int film_C;
  OutDesc cond_out;     // optional output of stage 1 (after add / FiLM, before the activation)
  float slope2, slope3; // PReLU slopes in front of conv2 / conv3
  float scale3;         // out = (conv3 + x) * scale3
  int n_out;
  OutDesc out[3];
  int reverse;
  int bulk_out;         // host: the single output is a plain tensor of the block's geometry: stage 3 stages its rows in the
                        // tile's (by then consumed) raw slab and one thread writes them with bulk shared -> global copies
  int* err_flag;
  int* sched;
  long long* clk;       // debug (UV_CB_CLK=1): [16] clock sums of pair 0's leader: MMA thread 0-5, epilogue warp 2 8-13
};

struct CbSmem {
  uint32_t film_off, mv_off, ident_off, w_off[3], w_bytes[3], bufs_off, bars_off, total;
};
__host__ __device__ inline uint32_t cb_stage_bytes(int n_src, int n_shift, int has_bias) {
  return (uint32_t)(n_src * n_shift) * (uint32_t)((kCbN / 2) * (kCbN / n_src) * 2) + (has_bias ? (kCbN / 2) * 32u : 0u);
}
__host__ __device__ inline CbSmem cb_smem_layout(const CbParams& p) {
  CbSmem L;
  uint32_t off = kOnesBytes;
  L.film_off = off;
  off += 2 * kCbN * 4;
  L.mv_off = off;       // per-utterance valid rows of a variable-length call (first kMvEntries utterances)
  off += kMvBytes;
  L.ident_off = off;
  off += kCbIdentBytes;
  for (int s = 0; s < 3; ++s) {
    L.w_off[s] = off;
    L.w_bytes[s] = cb_stage_bytes(p.n_src[s], p.n_shift[s], p.has_bias[s]);
    off += (L.w_bytes[s] + 127u) & ~127u;
  }
  L.bufs_off = (off + 1023u) & ~1023u;
  L.bars_off = L.bufs_off + (uint32_t)(kCbXBufs + kCbSlots) * kCbBufBytes;   // raw x slabs, then the working buffers
  L.total = L.bars_off + 1024;
  return L;
}

// barriers (same offsets in both CTAs)
struct CbBars {
  uint64_t x_full[kCbXBufs], x_free[kCbXBufs], w_free[kCbSlots], xa_ready[kCbSlots], mma_done[kCbSlots], epi_done[kCbSlots];
  uint64_t w_full;
  uint64_t s_full[kSchedSlots], s_empty[kSchedSlots];
  int s_ring[kSchedSlots];
  uint32_t tmem_ptr;
  int st_cnt[kCbSlots];   // bulk_out: epilogue warps that have staged their rows of the slot's tile
};
static_assert(sizeof(CbBars) <= 1024, "barrier block");

// One stage epilogue of one warp on one tile: TMEM -> registers -> stage math -> working buffer (stages 0 / 1) or
// global outputs (stage 2).  lane = accumulator row.
// valid rows of utterance b
__device__ __forceinline__ int cb_valid_rows(const CbParams& P, const int* mv_s, int b) {
  return P.mv ? (b < kMvEntries ? mv_s[b] : P.mv[b]) : P.M_valid;
}

template <int ST, bool LE1>
__device__ __forceinline__ void cb_stage_epilogue(const CbParams& P, uint32_t taddr, uint8_t* wbuf, const float* film_s, const int* mv_s,
                                                  int i, long long g, float inv_Lh, int own_lo, const uint4* pre) {
  bool valid = g >= 0 && g < P.m_rows;
  int b = 0, m = 0;
  if (valid) {
    b = fast_div((int)g, P.Lh, inv_Lh);
    m = (int)g - b * P.Lh;
    valid = m < cb_valid_rows(P, mv_s, b);
  }
  uint32_t r[2][16];
  tc_ld16_issue(taddr, r[0]);
  if (ST < 2) {
    const bool has_cond = ST == 0 && P.cond.base != nullptr;
    const bool has_film = ST == 0 && P.film != nullptr;
    const bool own = valid && i >= own_lo && i < own_lo + kCbRowsOut;   // stage-1 output rows written by this tile
    OutPtr po;
    po.p = nullptr; po.pstride = 0;
    if (ST == 0 && P.cond_out.t.base && own)
      po = out_ptr(P.cond_out.t, b, m, 0, 1.f / (float)(P.cond_out.t.split > 0 ? P.cond_out.t.split : 1));
    const float slope = ST == 0 ? P.slope2 : P.slope3;
    uint8_t* dst = wbuf + (size_t)i * 16;
#pragma unroll
    for (int cg = 0; cg < kCbN / 16; ++cg) {
      tc_ld_wait();
      if (cg + 1 < kCbN / 16) tc_ld16_issue(taddr + 16u * (cg + 1), r[(cg + 1) & 1]);
      float v[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) v[k] = __uint_as_float(r[cg & 1][k]);
      if (has_cond) {
        const uint4 cc[2] = {pre[2 * cg], pre[2 * cg + 1]};   // zeros on rows that are padding
        add_unpack(cc, v);
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] *= P.scale1;
      }
      if (has_film) {
        const float4* ga = reinterpret_cast<const float4*>(film_s + 16 * cg);
        const float4* be = reinterpret_cast<const float4*>(film_s + kCbN + 16 * cg);
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
        *reinterpret_cast<uint4*>(dst + (size_t)(2 * cg + h) * kCbPlaneBytes) = u;
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic stores -> tensor-core reads
  } else if (P.bulk_out) {
    // stage 3, staged: the finished rows go to a shared-memory buffer (the caller passes the tile's raw slab; zeros on
    // padding rows: those rows of the output tensor are zero and stay zero), like c1 / c2; cb_store_rows() writes them out
    const float a1 = P.out[0].act >= 1 ? P.out[0].a1 : 1.f, a2 = P.out[0].a2;
    const bool dbl = P.out[0].act == 2;
    uint8_t* dst = wbuf + (size_t)i * 16;
#pragma unroll
    for (int cg = 0; cg < kCbN / 16; ++cg) {
      tc_ld_wait();
      if (cg + 1 < kCbN / 16) tc_ld16_issue(taddr + 16u * (cg + 1), r[(cg + 1) & 1]);
      float v[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) v[k] = __uint_as_float(r[cg & 1][k]) * P.scale3;
      if (a1 != 1.f) {
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = prelu_t<LE1>(v[k], a1);
      }
      if (dbl) {
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = prelu_t<LE1>(v[k], a2);
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint4 u;
        __nv_bfloat162* q = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
        for (int k = 0; k < 4; ++k) q[k] = __floats2bfloat162_rn(v[8 * h + 2 * k], v[8 * h + 2 * k + 1]);
        if (!valid) u = make_uint4(0, 0, 0, 0);
        *reinterpret_cast<uint4*>(dst + (size_t)(2 * cg + h) * kCbPlaneBytes) = u;
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic stores -> bulk-copy reads
  } else {
    const bool out_row = valid && i < kCbRowsOut;
    OutPtr op[3];
#pragma unroll
    for (int o = 0; o < 3; ++o) {
      op[o].p = nullptr; op[o].pstride = 0;
      if (o < P.n_out && out_row) op[o] = out_ptr(P.out[o].t, b, m, 0, 1.f / (float)(P.out[o].t.split > 0 ? P.out[o].t.split : 1));
    }
#pragma unroll
    for (int cg = 0; cg < kCbN / 16; ++cg) {
      tc_ld_wait();
      if (cg + 1 < kCbN / 16) tc_ld16_issue(taddr + 16u * (cg + 1), r[(cg + 1) & 1]);
      if (out_row) {
        float v[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = __uint_as_float(r[cg & 1][k]) * P.scale3;
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

// bulk_out: rows [0, 252) of the staging buffer -> the output tensor (one bulk copy per chunk plane); returns when
// the copies have READ shared memory, i.e. the buffer may be overwritten.  One thread.
__device__ __forceinline__ void cb_store_rows(const CbParams& P, const uint8_t* wbuf, long long r0) {
  long long nrows = (long long)P.m_rows - r0;
  if (nrows > kCbRowsOut) nrows = kCbRowsOut;
  if (nrows > 0) {
    const TensDesc& t = P.out[0].t;
    const uint32_t bytes = (uint32_t)nrows * 16u;
    char* dst = t.base + (((long long)t.G + r0) << 4);
    const long long pstride = (long long)t.P << 4;
    for (int c = 0; c < kCbPlanes; ++c)
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + c * pstride),
                   "r"(smem_u32(wbuf + (size_t)c * kCbPlaneBytes)), "r"(bytes)
                   : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
}

// Row of the condition tensor that stage 1 adds, fetched before the wait for the accumulator (8 x 16 bytes per row).
__device__ __forceinline__ void cb_prefetch_cond(const CbParams& P, const int* mv_s, long long g, float inv_Lh, uint4* pre) {
#pragma unroll
  for (int c = 0; c < kCbPlanes; ++c) pre[c] = make_uint4(0, 0, 0, 0);
  if (P.cond.base == nullptr || g < 0 || g >= P.m_rows) return;
  const int b = fast_div((int)g, P.Lh, inv_Lh);
  if ((int)g - b * P.Lh >= cb_valid_rows(P, mv_s, b)) return;
  const char* p = P.cond.base + ((long long)P.cond.G + g) * 16;
  const long long stride = (long long)P.cond.P * 16;
#pragma unroll
  for (int c = 0; c < kCbPlanes; ++c) pre[c] = *reinterpret_cast<const uint4*>(p + c * stride);
}

__global__ void __launch_bounds__(kCbThreads, 1) convblock_pair_kernel(const __grid_constant__ CbParams P) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cta_rank_in_cluster();
  const bool leader = rank == 0;
  pdl_launch_dependents();
  int claim0 = 0;   // the pair's first two tile claims: one atomic, issued before the prologue (as in gemm_tc.cuh)
  if (threadIdx.x == 0 && leader) claim0 = atomicAdd(P.sched, 2);
  const CbSmem L = cb_smem_layout(P);
  uint8_t* ones = smem;
  float* film_s = reinterpret_cast<float*>(smem + L.film_off);
  int* mv_s = reinterpret_cast<int*>(smem + L.mv_off);
  uint8_t* ident = smem + L.ident_off;
  CbBars* B = reinterpret_cast<CbBars*>(smem + L.bars_off);
  auto xbuf = [&](int x) { return smem + L.bufs_off + (size_t)x * kCbBufBytes; };
  auto wbuf = [&](int s) { return smem + L.bufs_off + (size_t)(kCbXBufs + s) * kCbBufBytes; };
  constexpr int n_half = kCbN / 2;
  constexpr uint32_t tmem_cols = kCbSlots * 2 * kCbN;   // per slot: two sub-tile accumulators of 64 columns

  {
    // barrier set-up spread over threads (as in gemm_tc.cuh)
    const uint32_t pr = leader ? 1u : 0u;
    const int t = threadIdx.x;
    if (t < kSchedSlots) {
      const uint32_t ring_readers = (1 + kEpiWarps + kCbXfWarps) + (2 + kEpiWarps + kCbXfWarps);
      mbar_init(smem_u32(&B->s_full[t]), 1);
      mbar_init(smem_u32(&B->s_empty[t]), ring_readers);
    } else if (t < kSchedSlots + kCbXBufs) {
      const int x = t - kSchedSlots;
      mbar_init(smem_u32(&B->x_full[x]), 1);
      mbar_init(smem_u32(&B->x_free[x]), 1);
    } else if (t < kSchedSlots + kCbXBufs + kCbSlots) {
      const int s = t - kSchedSlots - kCbXBufs;
      mbar_init(smem_u32(&B->w_free[s]), 1);
      mbar_init(smem_u32(&B->xa_ready[s]), kCbXfWarps + pr);   // leader: + the peer's relay
      mbar_init(smem_u32(&B->mma_done[s]), 1);
      mbar_init(smem_u32(&B->epi_done[s]), 2 * kEpiWarps);   // (leader's is used) both CTAs' epilogue warps
      B->st_cnt[s] = 0;
    } else if (t == kSchedSlots + kCbXBufs + kCbSlots) {
      mbar_init(smem_u32(&B->w_full), 1 + pr);
    }
    if (t <= kSchedSlots + kCbXBufs + kCbSlots) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 256) {   // ones tile: K columns 0 / 1 meet the hi / lo halves of a bias image
    uint4 u = make_uint4(0u, 0u, 0u, 0u);
    if (threadIdx.x < kSubM) u.x = 0x3F803F80u;
    reinterpret_cast<uint4*>(ones)[threadIdx.x] = u;
  }
  // FiLM table / per-utterance valid rows: loaded to registers here, stored behind the cluster barrier below so that
  // the load latency overlaps it (both are written in front of the call, not by a predecessor: safe before pdl_wait)
  static_assert(2 * kCbN <= kCbThreads && kMvEntries <= kCbThreads, "one entry per thread");
  float film_r = 0.f;
  int mv_r = 0;
  const bool film_on = P.film != nullptr && threadIdx.x < 2 * kCbN;
  const bool mv_on = P.mv != nullptr && (int)threadIdx.x < kMvEntries && (long long)threadIdx.x * P.Lh < P.m_rows;
  if (film_on) { const int k = threadIdx.x; film_r = P.film[(k < kCbN ? 0 : P.film_C) + (k % kCbN) % P.film_C]; }
  if (mv_on) mv_r = P.mv[threadIdx.x];
  {  // identity image [8 chunks][n_half][8]: residual channel k -> column k; local column n stands for rank * n_half + n
    for (int i = threadIdx.x; i < kCbPlanes * n_half; i += kCbThreads) {
      const int c = i / n_half, n = i - c * n_half + (int)rank * n_half;
      const int e = n - 8 * c;
      uint32_t w[4] = {0u, 0u, 0u, 0u};
      if (e >= 0 && e < 8) w[e >> 1] = (e & 1) ? 0x3F800000u : 0x00003F80u;
      reinterpret_cast<uint4*>(ident)[i] = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
  // slack rows of the buffers are read by the taps of discarded accumulator rows: keep them finite
  for (uint32_t k = threadIdx.x; k < (uint32_t)(kCbXBufs + kCbSlots) * kCbBufBytes / 16; k += kCbThreads)
    reinterpret_cast<uint4*>(smem + L.bufs_off)[k] = make_uint4(0, 0, 0, 0);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  cluster_barrier();
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&B->tmem_ptr)), "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  if (film_on) film_s[threadIdx.x] = film_r;
  if (mv_on) mv_s[threadIdx.x] = mv_r;
  tc_fence_before();
  __syncthreads();
  cluster_barrier();
  tc_fence_after();
  const uint32_t tmem_base = B->tmem_ptr;

  int h[3], H = 0;
  for (int s = 0; s < 3; ++s) { h[s] = (P.n_shift[s] - 1) >> 1; H += h[s]; }
  const uint32_t sempty_cl0 = map_shared_rank(smem_u32(B->s_empty), 0);
  // first flat row of this CTA's part of tile t
  auto tile_r0 = [&](int t) { return ((long long)t * kPairCtas + rank) * kCbRowsOut; };

  if (warp == 0) {
    // ============================== loads
    claim0 = __shfl_sync(0xffffffffu, claim0, 0);
    if (elect_one()) {
      {  // resident weights of the three stages (this CTA's half of the columns)
        uint32_t tot = 0;
        for (int s = 0; s < 3; ++s) tot += L.w_bytes[s];
        mbar_expect_tx(smem_u32(&B->w_full), tot);
        for (int s = 0; s < 3; ++s) bulk_g2s(smem_u32(smem + L.w_off[s]), P.w[s] + (size_t)rank * L.w_bytes[s], L.w_bytes[s], smem_u32(&B->w_full));
      }
      int ss = 0;
      uint32_t sp = 0;
      int tile = 0, tile_nxt = 0;
      const uint32_t ring_peer = map_shared_rank(smem_u32(B->s_ring), 1), sfull_peer = map_shared_rank(smem_u32(B->s_full), 1);
      if (leader) {
        tile = claim0;
        tile_nxt = claim0 + 1;
      }
      pdl_wait();
      uint32_t pf[kCbXBufs] = {0, 0, 0};
      int xb = 0;
      // variable-length calls: a tile whose output rows all lie past their utterances' ends is dropped by the leader
      // (nothing to compute: those rows are zero padding and stay zero)
      auto tile_has_rows = [&](int claim) {
        const int t = P.reverse ? P.n_tiles - 1 - claim : claim;
        const long long g0 = (long long)t * kPairCtas * kCbRowsOut;
        long long g1 = g0 + kPairCtas * kCbRowsOut;
        if (g1 > P.m_rows) g1 = P.m_rows;
        for (int b = (int)(g0 / P.Lh); (long long)b * P.Lh < g1; ++b) {
          const long long m_lo = g0 - (long long)b * P.Lh;
          if ((m_lo < 0 ? 0 : m_lo) < cb_valid_rows(P, mv_s, b)) return true;
        }
        return false;
      };
      for (int k = 0;; ++k) {
        int tile_id;
        if (leader) {
          if (P.mv) {
            while (tile < P.n_tiles && !tile_has_rows(tile)) {
              tile = tile_nxt;
              if (tile_nxt < P.n_tiles) tile_nxt = atomicAdd(P.sched, 1);
            }
          }
          const bool live = tile < P.n_tiles;
          mbar_wait(smem_u32(&B->s_empty[ss]), sp ^ 1, P.err_flag, 9);
          tile_id = live ? (P.reverse ? P.n_tiles - 1 - tile : tile) : -1;
          B->s_ring[ss] = tile_id;
          mbar_arrive(smem_u32(&B->s_full[ss]));
          mbar_expect_tx_cl(sfull_peer + 8u * ss, 4u);
          st_async_u32(ring_peer + 4u * ss, (uint32_t)tile_id, sfull_peer + 8u * ss);
        } else {
          mbar_wait(smem_u32(&B->s_full[ss]), sp, P.err_flag, 8);
          tile_id = *reinterpret_cast<const volatile int*>(B->s_ring + ss);
          mbar_arrive_cl(sempty_cl0 + 8u * ss);
        }
        if (++ss == kSchedSlots) { ss = 0; sp ^= 1; }
        if (tile_id < 0) break;
        int tile_nn = tile_nxt;
        if (leader && tile_nxt < P.n_tiles) tile_nn = atomicAdd(P.sched, 1);
        if (k >= kCbXBufs) {   // the tile that used this slab three tiles ago has been consumed (its stage-3 MMAs are complete)
          mbar_wait(smem_u32(&B->x_free[xb]), pf[xb], P.err_flag, 10);
          pf[xb] ^= 1;
        }
        const long long row = (long long)P.x_G + tile_r0(tile_id) - H;
        long long nrows = (long long)P.x_P - row;     // never read past the plane (rows past m_rows are masked)
        if (nrows > kCbLoadRows) nrows = kCbLoadRows;
        const uint32_t fb = smem_u32(&B->x_full[xb]);
        if (nrows <= 0) {
          mbar_arrive(fb);   // this CTA's part of the last tile lies wholly past the plane: every row is masked
        } else {
          const uint32_t bytes = (uint32_t)nrows * 16u;
          mbar_expect_tx(fb, bytes * kCbPlanes);
          const uint32_t dst = smem_u32(xbuf(xb));
          for (int c = 0; c < kCbPlanes; ++c) bulk_g2s(dst + c * kCbPlaneBytes, P.x_base + (((long long)c * P.x_P + row) << 4), bytes, fb);
        }
        if (++xb == kCbXBufs) xb = 0;
        tile = tile_nxt;
        tile_nxt = tile_nn;
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      int ss = 0;
      uint32_t sp = 0;
      auto next_tile = [&]() {
        mbar_wait(smem_u32(&B->s_full[ss]), sp, P.err_flag, 8);
        const int t = *reinterpret_cast<const volatile int*>(B->s_ring + ss);
        mbar_arrive_cl(sempty_cl0 + 8u * ss);
        if (++ss == kSchedSlots) { ss = 0; sp ^= 1; }
        return t;
      };
      if (leader) {
        // ============================== MMA issuer: rounds of up to two tiles, stage by stage
        const uint32_t idesc = make_idesc(kTileM, kCbN);
        const uint32_t desc_hi = (128u >> 4) | (1u << 14);
        const uint32_t lbo_a = kCbPlaneBytes, lbo_b = (uint32_t)n_half * 16u;
        const uint32_t a_kstep = (2u * lbo_a) >> 4, b_kstep = (2u * lbo_b) >> 4;
        const uint32_t sub16 = (uint32_t)(kSubM * 16) >> 4;
        const uint64_t ones_desc = make_desc(smem_u32(ones), (uint32_t)kSubM * 16u, 128);
        auto dsc = [&](uint32_t lo) { return ((uint64_t)desc_hi << 32) | (uint64_t)lo; };
        mbar_wait(smem_u32(&B->w_full), 0, P.err_flag, 11);   // both CTAs' weights
        tc_fence_after();
        uint32_t px[kCbSlots] = {0, 0}, pe[kCbSlots] = {0, 0};
        bool used[kCbSlots] = {false, false};
        const bool ck = P.clk != nullptr && blockIdx.x == 0;
        long long c_wait_epi = 0, c_wait_xa = 0, c_issue = 0, c_items = 0, t_prev = ck ? clock64() : 0, c_ring = 0;
        int xb0 = 0;   // x slab of the round's first tile (the second one uses the next)
        for (;;) {
          int tl[kCbSlots];
          tl[0] = next_tile();
          if (tl[0] < 0) break;
          tl[1] = next_tile();
          if (ck) { const long long t = clock64(); c_ring += t - t_prev; t_prev = t; }
          const int xbs[kCbSlots] = {xb0, xb0 + 1 == kCbXBufs ? 0 : xb0 + 1};
          xb0 = (xb0 + 2) % kCbXBufs;
#pragma unroll
          for (int st = 0; st < 3; ++st) {
            for (int s = 0; s < kCbSlots; ++s) {
              if (tl[s] < 0) continue;
              if (used[s]) {   // the slot's previous item has been drained / written back by both CTAs' epilogue warps
                mbar_wait(smem_u32(&B->epi_done[s]), pe[s], P.err_flag, 14);
                pe[s] ^= 1;
              }
              if (ck) { const long long t = clock64(); c_wait_epi += t - t_prev; t_prev = t; }
              used[s] = true;
              if (st == 0) {   // PReLU(x) is in the working buffer of both CTAs
                mbar_wait(smem_u32(&B->xa_ready[s]), px[s], P.err_flag, 13);
                px[s] ^= 1;
              }
              if (ck) { const long long t = clock64(); c_wait_xa += t - t_prev; t_prev = t; }
              tc_fence_after();
              const uint32_t d0 = tmem_base + (uint32_t)(s * 2) * kCbN, d1 = d0 + kCbN;
              const uint32_t a_lo0 = (smem_u32(wbuf(s)) >> 4) | ((lbo_a >> 4) << 16);
              uint32_t w_lo = (smem_u32(smem + L.w_off[st]) >> 4) | ((lbo_b >> 4) << 16);
              const int kc = kCbN / P.n_src[st];
              const uint32_t img16 = ((uint32_t)n_half * kc * 2u) >> 4;
              uint32_t acc = 0;
              if (P.has_bias[st]) {
                const uint32_t bias_lo = w_lo + (uint32_t)(P.n_src[st] * P.n_shift[st]) * img16;
                tc_mma2_bf16(d0, ones_desc, dsc(bias_lo), idesc, 0);
                tc_mma2_bf16(d1, ones_desc, dsc(bias_lo), idesc, 0);
                acc = 1;
              }
              for (int src = 0; src < P.n_src[st]; ++src) {
                const uint32_t a_src = a_lo0 + (uint32_t)src * (uint32_t)(kc >> 3) * (lbo_a >> 4);
                for (int q = 0; q < P.n_shift[st]; ++q) {
                  uint32_t al = a_src + (uint32_t)q, bl = w_lo;
                  for (int kk = 0; kk < (kc >> 4); ++kk) {
                    tc_mma2_bf16(d0, dsc(al), dsc(bl), idesc, acc);
                    tc_mma2_bf16(d1, dsc(al + sub16), dsc(bl), idesc, acc);
                    acc = 1;
                    al += a_kstep;
                    bl += b_kstep;
                  }
                  w_lo += img16;
                }
              }
              if (st == 2) {
                // residual: out row i = flat R0 + i = raw slab row i + H, through the identity image
                uint32_t al = ((smem_u32(xbuf(xbs[s])) >> 4) | ((lbo_a >> 4) << 16)) + (uint32_t)H;
                uint32_t bl = (smem_u32(ident) >> 4) | ((lbo_b >> 4) << 16);
                for (int kk = 0; kk < (kCbN >> 4); ++kk) {
                  tc_mma2_bf16(d0, dsc(al), dsc(bl), idesc, 1);
                  tc_mma2_bf16(d1, dsc(al + sub16), dsc(bl), idesc, 1);
                  al += a_kstep;
                  bl += b_kstep;
                }
              }
              tc_commit2(smem_u32(&B->mma_done[s]));
              if (st == 2) {   // the tile's working buffer and raw slab may be refilled
                tc_commit2(smem_u32(&B->w_free[s]));
                if (!P.bulk_out) tc_commit2(smem_u32(&B->x_free[xbs[s]]));   // (bulk_out: the slab then stages the output rows and
                                                                             // is freed by the thread that stores them)
              }
              if (ck) { const long long t = clock64(); c_issue += t - t_prev; t_prev = t; ++c_items; }
            }
          }
          if (tl[1] < 0) break;
        }
        if (ck) { P.clk[0] = c_wait_epi; P.clk[1] = c_wait_xa; P.clk[2] = c_issue; P.clk[3] = c_items; P.clk[4] = c_ring; }
      } else {
        // ============================== relay: this CTA's weights / activated slabs are in place
        mbar_wait(smem_u32(&B->w_full), 0, P.err_flag, 11);
        mbar_arrive_cl(map_shared_rank(smem_u32(&B->w_full), 0));
        const uint32_t xa_rem = map_shared_rank(smem_u32(B->xa_ready), 0);
        uint32_t px[kCbSlots] = {0, 0};
        for (int k = 0;; ++k) {
          if (next_tile() < 0) break;
          const int s = k & (kCbSlots - 1);
          mbar_wait(smem_u32(&B->xa_ready[s]), px[s], P.err_flag, 13);
          px[s] ^= 1;
          mbar_arrive_cl(xa_rem + 8u * s);
        }
      }
    }
  } else if (warp >= 2 + kEpiWarps) {
    // ============================== PReLU(x): raw slab -> working buffer
    const int w = warp - (2 + kEpiWarps);
    const bool le1 = P.a_slope <= 1.f;
    int ss = 0;
    uint32_t sp = 0, pxf[kCbXBufs] = {0, 0, 0}, pwf[kCbSlots] = {0, 0};
    int xb = 0;
    for (int k = 0;; ++k) {
      mbar_wait(smem_u32(&B->s_full[ss]), sp, P.err_flag, 8);
      const int tile = *reinterpret_cast<const volatile int*>(B->s_ring + ss);
      __syncwarp();
      if (lane == 0) mbar_arrive_cl(sempty_cl0 + 8u * ss);
      if (++ss == kSchedSlots) { ss = 0; sp ^= 1; }
      if (tile < 0) break;
      const int s = k & (kCbSlots - 1);
      mbar_wait(smem_u32(&B->x_full[xb]), pxf[xb], P.err_flag, 12);
      pxf[xb] ^= 1;
      if (k >= kCbSlots) {   // the slot's previous tile no longer reads the working buffer
        mbar_wait(smem_u32(&B->w_free[s]), pwf[s], P.err_flag, 15);
        pwf[s] ^= 1;
      }
      const uint4* src = reinterpret_cast<const uint4*>(xbuf(xb));
      uint4* dst = reinterpret_cast<uint4*>(wbuf(s));
      if (++xb == kCbXBufs) xb = 0;
      constexpr int kStep = 32 * kCbXfWarps;
      constexpr int nvec = kCbPlanes * kCbBufRows;
      constexpr int kPer = (nvec + kStep - 1) / kStep;   // 17 vectors per thread, every load in flight before the first
      const int i0 = w * 32 + lane;                      // store: the pass is latency bound (the MMAs keep shared memory busy)
      uint4 u[kPer];
#pragma unroll
      for (int j = 0; j < kPer; ++j)
        if (i0 + j * kStep < nvec) u[j] = src[i0 + j * kStep];
#pragma unroll
      for (int j = 0; j < kPer; ++j) {
        if (i0 + j * kStep < nvec) {
          u[j].x = prelu_bf16x2(u[j].x, P.a_slope, le1);
          u[j].y = prelu_bf16x2(u[j].y, P.a_slope, le1);
          u[j].z = prelu_bf16x2(u[j].z, P.a_slope, le1);
          u[j].w = prelu_bf16x2(u[j].w, P.a_slope, le1);
          dst[i0 + j * kStep] = u[j];
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&B->xa_ready[s]));
    }
  } else {
    // ============================== stage epilogues
    // (every PReLU slope of the block is <= 1: cb_params_ok; PReLU(v) = max(v, a v))
    const int sub = (warp - 2) >> 2, quarter = warp & 3;
    const int i = sub * kSubM + quarter * 32 + lane;
    const float inv_Lh = 1.f / (float)P.Lh;
    const uint32_t epi_cl0 = map_shared_rank(smem_u32(B->epi_done), 0);
    int ss = 0;
    uint32_t sp = 0, pm[kCbSlots] = {0, 0};
    auto next_tile = [&]() {
      mbar_wait(smem_u32(&B->s_full[ss]), sp, P.err_flag, 8);
      const int t = *reinterpret_cast<const volatile int*>(B->s_ring + ss);
      __syncwarp();
      if (lane == 0) mbar_arrive_cl(sempty_cl0 + 8u * ss);
      if (++ss == kSchedSlots) { ss = 0; sp ^= 1; }
      return t;
    };
    bool first = true;
    const bool ck = P.clk != nullptr && blockIdx.x == 0 && warp == 2 && lane == 0;
    long long c_wait = 0, c_work[3] = {0, 0, 0}, c_ring = 0, t_prev = ck ? clock64() : 0;
    int xb0 = 0;   // x slab of the round's first tile, as in the MMA issuer
    for (;;) {
      int tl[kCbSlots];
      tl[0] = next_tile();
      if (tl[0] < 0) break;
      tl[1] = next_tile();
      const int xbs[kCbSlots] = {xb0, xb0 + 1 == kCbXBufs ? 0 : xb0 + 1};
      xb0 = (xb0 + 2) % kCbXBufs;
      if (first) { pdl_wait(); first = false; }
      if (ck) { const long long t = clock64(); c_ring += t - t_prev; t_prev = t; }   // condition loads / output stores touch the predecessors' tensors
#pragma unroll
      for (int st = 0; st < 3; ++st) {
        for (int s = 0; s < kCbSlots; ++s) {
          if (tl[s] < 0) continue;
          const long long r0 = tile_r0(tl[s]);
          // flat row of accumulator row i at this stage
          const long long g = r0 - H + (st >= 0 ? h[0] : 0) + (st >= 1 ? h[1] : 0) + (st >= 2 ? h[2] : 0) + i;
          uint4 pre[kCbPlanes];
          if (st == 0) cb_prefetch_cond(P, mv_s, g, inv_Lh, pre);
          mbar_wait(smem_u32(&B->mma_done[s]), pm[s], P.err_flag, 12);
          pm[s] ^= 1;
          if (ck) { const long long t = clock64(); c_wait += t - t_prev; t_prev = t; }
          tc_fence_after();
          const uint32_t taddr = tmem_base + (uint32_t)(s * 2 + sub) * kCbN + ((uint32_t)(quarter * 32) << 16);
          // stages 1 / 2 write the working buffer; a staged stage 3 writes the tile's raw slab (its last reader, the
          // residual MMA, has completed)
          uint8_t* wb = st == 2 && P.bulk_out ? xbuf(xbs[s]) : wbuf(s);
          if (st == 0) cb_stage_epilogue<0, true>(P, taddr, wb, film_s, mv_s, i, g, inv_Lh, H - h[0], pre);
          else if (st == 1) cb_stage_epilogue<1, true>(P, taddr, wb, film_s, mv_s, i, g, inv_Lh, 0, pre);
          else cb_stage_epilogue<2, true>(P, taddr, wb, film_s, mv_s, i, g, inv_Lh, 0, pre);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cl(epi_cl0 + 8u * s);
          if (st == 2 && P.bulk_out && lane == 0) {
            // the last warp to have staged its rows writes the tile out and hands the slab back to the loads
            __threadfence_block();
            if (atomicAdd(&B->st_cnt[s], 1) == kEpiWarps - 1) {
              B->st_cnt[s] = 0;
              __threadfence_block();
              asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
              cb_store_rows(P, wb, r0);
              mbar_arrive(smem_u32(&B->x_free[xbs[s]]));
            }
          }
          __syncwarp();
          if (ck) { const long long t = clock64(); c_work[st] += t - t_prev; t_prev = t; }
        }
      }
      if (tl[1] < 0) break;
    }
    if (ck) { P.clk[8] = c_wait; P.clk[9] = c_work[0]; P.clk[10] = c_work[1]; P.clk[11] = c_work[2]; P.clk[12] = c_ring; }
  }

  tc_fence_before();
  __syncthreads();
  cluster_barrier();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
  }
  if (threadIdx.x == 0 && leader) {
    __threadfence();
    if (atomicAdd(P.sched + 1, 1) == (int)(gridDim.x >> 1) - 1) {
      P.sched[0] = 0;
      P.sched[1] = 0;
    }
  }
}

inline bool cb_params_ok(const CbParams& p) {
  int H = 0;
  for (int s = 0; s < 3; ++s) {
    if (p.n_src[s] < 1 || kCbN % (16 * p.n_src[s]) || p.n_shift[s] % 2 == 0 || p.n_shift[s] > 5) return false;
    H += (p.n_shift[s] - 1) / 2;
  }
  if (H > 4 || p.x_G < H || kCbCtaRows - kCbRowsOut < 2 * ((p.n_shift[1] - 1) / 2 + (p.n_shift[2] - 1) / 2)) return false;
  if (p.film && (p.film_C <= 0 || kCbN % p.film_C)) return false;
  // the epilogues compute PReLU(v) as max(v, a v): slopes <= 1 only (the host compiler keeps other blocks unfused)
  if (!(p.slope2 <= 1.f) || !(p.slope3 <= 1.f)) return false;
  for (int o = 0; o < p.n_out && o < 3; ++o)
    if ((p.out[o].act >= 1 && !(p.out[o].a1 <= 1.f)) || (p.out[o].act == 2 && !(p.out[o].a2 <= 1.f))) return false;
  return cb_smem_layout(p).total <= 227 * 1024;
}

inline cudaError_t convblock_pair_set_attributes() {   // per device (uv_model_create)
  return cudaFuncSetAttribute(convblock_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
}

inline cudaError_t convblock_pair_launch(CbParams p, int num_sms, cudaStream_t stream) {
  if (!cb_params_ok(p) || !p.sched) return cudaErrorInvalidConfiguration;
  p.n_tiles = (p.m_rows + kPairCtas * kCbRowsOut - 1) / (kPairCtas * kCbRowsOut);
  int pairs = num_sms / kPairCtas;
  if (p.n_tiles < pairs) pairs = p.n_tiles;
  if (pairs < 1) pairs = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(pairs * kPairCtas);
  cfg.blockDim = dim3(kCbThreads);
  cfg.dynamicSmemBytes = 227 * 1024;   // the whole SM, like the GEMM kernel (never co-resident with a GRU CTA)
  cfg.stream = stream;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = kPairCtas;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, convblock_pair_kernel, p);
}

}  // namespace uv

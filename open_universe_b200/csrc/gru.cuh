// Bidirectional GRU recurrence (torch.nn.GRU semantics; reference score.py:84-90,117 and
// condition.py:173-179,212).  The input projection x W_ih^T + b_ih is a tensor-core GEMM
// (gemm_tc.cuh, f32 row-major output `gi`); this kernel runs the serial part:
//   r = sig(gi_r + W_hr h + b_hr), z = sig(gi_z + W_hz h + b_hz),
//   n = tanh(gi_n + r * (W_hn h + b_hn)),  h' = (1 - z) n + z h
//
// The chain of T dependent steps is latency bound, so the design minimises the per-step
// critical path instead of chasing throughput:
//   * one thread-block CLUSTER of 8 CTAs per (direction, group of <= 64 utterances); CTA `rank`
//     owns hidden units [rank*U, (rank+1)*U), U = H/8, i.e. 3U rows of W_hh, which stay resident
//     in its shared memory for the whole sequence (bf16, K-major UMMA layout);
//   * per step one elected thread issues  gh[3U x NB] = W_slice[3U x H] . h^T[H x NB]  as
//     tcgen05 MMAs (M = 128 row blocks, N = utterances, accumulators in TMEM);
//   * the gate math runs on all 8 warps with fp32 state in registers (thread = 8 hidden units of
//     one utterance), and the new h slice is written as bf16 straight into the h operand buffer of
//     every CTA of the cluster (distributed shared memory), followed by one release-arrive per warp
//     on each peer's mbarrier -- no cluster-wide barrier on the critical path;
//   * every per-step global access is a bulk copy or a full 128-byte line (scattered 16-byte accesses made
//     the first version L1-wavefront bound): the gate inputs gi of step t+1 arrive as ONE bulk copy into
//     shared memory (the input-projection GEMM writes them as [t][dir][rank][b][3U+4] blocks), and the new
//     h is staged (bf16) for 8 steps, then flushed as 128-byte runs of the planes outputs together with the
//     fused epilogue of the reference ((out + res)/sqrt2, PReLU of the consumers: condition.py:215,
//     blocks.py:380).
#pragma once
#include "gemm_tc.cuh"

namespace uv {

constexpr int kGruNC = 8;          // CTAs per cluster (portable maximum)
constexpr int kGruThreads = 288;   // 8 compute warps + 1 MMA-issue warp
constexpr int kGruCompute = 256;
constexpr int kGruMaxNB = 64;      // utterances per cluster (UMMA N)
constexpr int kGruMaxH = 384;

struct GruParams {
  const float* gi;          // [T][2 dirs][8 ranks][Bp][3U + 4] f32: gate inputs, one contiguous block per (t, dir, rank)
  int Bp;                   // batch pitch of gi (B rounded up to a multiple of NB)
  const char* whh;          // [2 dirs][8 ranks][H/8][3U][8] bf16 (K-major UMMA images of the W_hh slices)
  const float* bhh;         // [2][3H]
  int B, T, H, Lh;
  int NB;                   // utterances per cluster, multiple of 16
  TensDesc add1;
  float scale;
  int n_out;
  OutDesc out[3];
  int* err_flag;
  long long* clk;           // optional [16] phase-time accumulators (debug, UV_GRU_CLK=1)
  int dbg_flags;            // UV_GRU_FLAGS: spare debug switches (unused)
  const RowGeo* rows;       // per-utterance lengths (null: every utterance has T steps): an utterance of F < T frames
                            // is a sequence of F steps (forward: t = 0 .. F-1, backward: F-1 .. 0), the rest is zero
};

struct GruSmem {
  int w_bytes, h_bytes, gh_stride, gh_bytes, gi_bytes, st_bytes, union_bytes, total;
};
constexpr int kGruStagePitch = 144;   // bytes per (8 units, utterance) pair in the output staging: 8 steps x 16 B + 16 B skew
__host__ __device__ inline GruSmem gru_smem_layout(int H, int NB) {
  GruSmem s;
  const int U = H / kGruNC, rows = 3 * U;
  s.w_bytes = rows * H * 2 + 128 * 16;          // + one row block of slack: the last M=128 block may over-read
  s.h_bytes = NB * H * 2;                       // one h operand buffer [H/8][NB][8] bf16
  s.gh_stride = NB + 1;
  s.gh_bytes = ((rows * s.gh_stride * 4) + 127) & ~127;
  s.gi_bytes = NB * (rows + 4) * 4;             // gate inputs of one step [NB][3U + 4] f32 (row pitch 16 B off a bank multiple)
  s.st_bytes = (U / 8) * NB * kGruStagePitch;   // 8 steps of bf16 h per (chunk, utterance) pair (two such buffers)
  // W is only staged in shared memory until it has been copied into tensor memory: the staging area is then
  // reused for the per-step buffers (gh, gi, output staging)
  const int step_bytes = s.gh_bytes + s.gi_bytes + 2 * s.st_bytes;
  s.union_bytes = s.w_bytes > step_bytes ? s.w_bytes : step_bytes;
  s.total = 2 * s.h_bytes + s.union_bytes + 256;
  return s;
}

// gh += A (tensor memory: lane = row, 32-bit column = two consecutive K elements) . B (shared memory descriptor)
__device__ __forceinline__ void tc_mma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 16 registers -> 32 lanes x 16 consecutive 32-bit columns of tensor memory
__device__ __forceinline__ void tc_st16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
      "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
// Slices of h are consumed in three groups, each behind one barrier per operand buffer: the CTA's own
// slice, the next three ranks', the last four ranks' (a barrier wait costs ~90 clk of latency on the
// issuing thread even when the phase is complete, so not one per slice).
constexpr int kGruGroups = 3;
__host__ __device__ constexpr int gru_group_of(int c) { return c == 0 ? 0 : (c <= 3 ? 1 : 2); }
__host__ __device__ constexpr int gru_group_slices(int g) { return g == 0 ? 1 : (g == 1 ? 3 : 4); }
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t map_to_rank(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, const uint4& v) {
  asm volatile("st.shared::cluster.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
// bulk copy of this CTA's shared memory into a peer CTA's; completes `bytes` on the peer's mbarrier
__device__ __forceinline__ void bulk_s2peer(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t bar_cluster) {
  asm volatile(
      "cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_cluster),
      "r"(src_cta), "r"(bytes), "r"(bar_cluster)
      : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity, int* err_flag, int code) {
  uint32_t spins = 0;
  while (!mbar_try_wait_cluster(bar, parity)) {
    if (++spins > (1u << 22)) {
      if (err_flag) *reinterpret_cast<volatile int*>(err_flag) = code;
      __threadfence_system();
      __trap();
    }
  }
}
__device__ __forceinline__ void compute_sync() { asm volatile("bar.sync 1, %0;" ::"n"(256) : "memory"); }
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// sigmoid / tanh on the two SFU approximations (ex2, rcp: ~1e-7 relative error each, no slow path)
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// The gate phase is SFU bound (16 MUFU/clk/SM): one MUFU.TANH per gate instead of ex2 + rcp.
// tanh.approx.f32 has a maximum relative error of 2^-11, below the bf16 rounding (2^-9) h takes anyway
// as the MMA operand of the next step.
__device__ __forceinline__ float fast_tanh(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_sigmoid(float x) { return fmaf(0.5f, fast_tanh(0.5f * x), 0.5f); }


// The MMA issuer (one thread).  W_hh's slice is resident in tensor memory (A operand), so an MMA only
// reads the 16 x NB slab of h from shared memory.  Per step it waits for the three slice groups in the
// order they are sent and issues their K steps; everything that does not depend on the step sits in
// registers, because this thread's instruction latency is on the critical path of the recurrence.
template <int KPR, int NBLK>
__device__ __forceinline__ void gru_mma_role(const GruParams& P, const GruSmem& L, uint32_t h_addr, uint32_t hfull_addr,
                                             uint32_t done_addr, uint32_t tmem_base, uint32_t w_tmem, uint32_t rank, int kpr_rt,
                                             int nblk_rt, uint32_t slice_bytes, bool mclk, volatile long long* pub_clk) {
  const int kpr = KPR ? KPR : kpr_rt, nblk = NBLK ? NBLK : nblk_rt;
  const int NB = P.NB, T = P.T;
  const uint32_t idesc = make_idesc(128, NB);
  const uint32_t lbo_h = (uint32_t)NB * 16u;
  const uint64_t hks = (uint64_t)((2u * lbo_h) >> 4);   // descriptor step per K=16
  const uint64_t hbuf = (uint64_t)((uint32_t)L.h_bytes >> 4);
  const uint64_t hdesc0 = make_desc(h_addr, lbo_h, 128);
  const uint32_t wblk = (uint32_t)P.H >> 1;             // tensor-memory columns of one 128-row block of W
  uint32_t wa[kGruNC];
  uint64_t hd[kGruNC];
#pragma unroll
  for (int c = 0; c < kGruNC; ++c) {
    const uint32_t src = (rank + c) & (kGruNC - 1);
    wa[c] = w_tmem + src * kpr * 8u;
    hd[c] = hdesc0 + (uint64_t)(src * kpr) * hks;
  }
  long long mk[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) mk[i] = 0;
  for (int step = 0; step < T; ++step) {
    const uint32_t boff = (step & 1) ? 8u * kGruGroups : 0u;
    const uint64_t hb = (step & 1) ? hbuf : 0;
    const uint32_t par = ((step - 1) >> 1) & 1;
#pragma unroll
    for (int c = 0; c < kGruNC; ++c) {
      if (c == 0 || gru_group_of(c) != gru_group_of(c - 1)) {
        if (step > 0) mbar_wait(hfull_addr + boff + 8u * gru_group_of(c), par, P.err_flag, 22);
        tc_fence_after();
        if (mclk && step > 0) mk[gru_group_of(c)] += clock64() - *pub_clk;
      }
#pragma unroll
      for (int blk = 0; blk < (NBLK ? NBLK : 2); ++blk) {
        if (blk < nblk) {
#pragma unroll
          for (int k2 = 0; k2 < (KPR ? KPR : 3); ++k2)
            if (k2 < kpr) tc_mma_bf16_ts(tmem_base + blk * NB, wa[c] + blk * wblk + k2 * 8u, hd[c] + hb + k2 * hks, idesc, (c | k2) ? 1u : 0u);
        }
      }
    }
    tc_commit(done_addr);
    // the peer groups of this buffer are filled again at the end of step+1 (for step+2): arm those
    // phases now (a phase cannot complete before its arming arrive)
    if (step > 0 && step + 2 < T) {
#pragma unroll
      for (int g = 1; g < kGruGroups; ++g) mbar_expect_tx(hfull_addr + boff + 8u * g, slice_bytes * gru_group_slices(g));
    }
    if (mclk && step > 0) mk[8] += clock64() - *pub_clk;
  }
  if (mclk) {
#pragma unroll
    for (int i = 0; i < 9; ++i) P.clk[16 + i] = mk[i];
  }
}

__global__ void __launch_bounds__(kGruThreads, 1) gru_kernel(const __grid_constant__ GruParams P) {
  extern __shared__ __align__(1024) uint8_t gsm[];
  const int H = P.H, U = H / kGruNC, rows = 3 * U, NB = P.NB;
  const GruSmem L = gru_smem_layout(H, NB);
  uint8_t* h_s = gsm;                                   // two operand buffers
  uint8_t* w_s = gsm + 2 * L.h_bytes;                   // W staging (start-up only), then:
  float* gh_s = reinterpret_cast<float*>(w_s);
  float* gi_s = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(gh_s) + L.gh_bytes);
  uint8_t* st_s = reinterpret_cast<uint8_t*>(gi_s) + L.gi_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(w_s + L.union_bytes);
  uint64_t* h_full = bars;         // [2 buffers][3 groups]: that group's slices of h have landed in buffer i
  uint64_t* mma_done = bars + 16;  // [1]
  uint64_t* w_full = bars + 17;    // [1]
  uint64_t* gi_full = bars + 18;   // [1]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 19);
  volatile long long* pub_clk = reinterpret_cast<volatile long long*>(bars + 20);   // debug: clock of the last publish

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_rank();
  const int dir = blockIdx.z;
  const int b0 = blockIdx.y * NB;
  const int T = P.T;
  const int n_blocks = (rows + 127) / 128;
  const int S = rows + 4;                                // gi row pitch (floats)
  // tensor memory: accumulators [n_blocks][NB columns], then the W_hh slice [n_blocks][H/2 columns]
  const uint32_t w_col0 = (uint32_t)(n_blocks * NB);
  uint32_t tmem_cols = 32;
  while (tmem_cols < w_col0 + (uint32_t)(n_blocks * (H >> 1))) tmem_cols <<= 1;
  const uint32_t slice_bytes = (uint32_t)(U >> 3) * NB * 16u;
  const uint32_t gi_bytes = (uint32_t)NB * S * 4u;
  // gi block of time t for this CTA: ((t*2 + dir)*8 + rank) * Bp + b0 rows of S floats
  const long long gi_tstride = 16ll * P.Bp * S;
  const float* gi_blk0 = P.gi + ((long long)(dir * 8 + (int)rank) * P.Bp + b0) * S;
  const int t0 = dir == 0 ? 0 : T - 1;
  const int tstep = dir == 0 ? 1 : -1;

  if (tid == 0) {
    // one barrier per (buffer, slice group), count 1: the peer slices complete the bytes armed by the MMA
    // thread's arrive.expect_tx, the local slice is a plain arrive
    for (int i = 0; i < 2 * kGruGroups; ++i) mbar_init(smem_u32(&h_full[i]), 1);
    mbar_init(smem_u32(mma_done), 1);
    mbar_init(smem_u32(w_full), 1);
    mbar_init(smem_u32(gi_full), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    for (int g = 1; g < kGruGroups; ++g) {
      if (T > 1) mbar_expect_tx(smem_u32(&h_full[kGruGroups + g]), slice_bytes * gru_group_slices(g));   // buffer 1: filled at the end of step 0
      if (T > 2) mbar_expect_tx(smem_u32(&h_full[g]), slice_bytes * gru_group_slices(g));                // buffer 0: filled at the end of step 1
    }
    // resident W_hh slice
    const uint32_t wb = (uint32_t)rows * H * 2;
    const char* src = P.whh + ((size_t)dir * kGruNC + rank) * wb;
    mbar_expect_tx(smem_u32(w_full), wb);
    for (uint32_t off = 0; off < wb; off += 32768u) {
      const uint32_t n = wb - off < 32768u ? wb - off : 32768u;
      bulk_g2s(smem_u32(w_s + off), src + off, n, smem_u32(w_full));
    }
  }
  // h_0 = 0 (both buffers), slack rows of W zero
  for (int i = tid; i < (2 * L.h_bytes) / 16; i += kGruThreads) reinterpret_cast<uint4*>(h_s)[i] = make_uint4(0, 0, 0, 0);
  for (int i = tid; i < 128; i += kGruThreads) reinterpret_cast<uint4*>(w_s + (size_t)rows * H * 2)[i] = make_uint4(0, 0, 0, 0);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)),
                 "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  // Everything above (barriers, tensor-memory allocation, the W_hh bulk load) is independent of the predecessor.
  // No early pdl_launch_dependents() here: the dependents' CTAs would sit on the free SMs for the whole recurrence
  // and keep the other utterance lane's kernels out.
  pdl_wait();
  const uint32_t tmem_base = *tmem_ptr_smem;
  if (warp < 4) {
    // W_hh slice: staging image in shared memory ([K chunk][row][8]) -> tensor memory (lane = row,
    // column = K pair); rows past 3U are zero.  The staging area is dead afterwards.
    mbar_wait(smem_u32(w_full), 0, P.err_flag, 21);
    for (int blk = 0; blk < n_blocks; ++blk) {
      const int r = blk * 128 + warp * 32 + lane;
      const uint32_t ta = tmem_base + w_col0 + (uint32_t)(blk * (H >> 1)) + ((uint32_t)(warp * 32) << 16);
      for (int kc = 0; kc < (H >> 3); kc += 4) {   // 4 chunks of 8 K values = 16 columns per store
        uint32_t v[16];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 x = make_uint4(0, 0, 0, 0);
          if (r < rows) x = *reinterpret_cast<const uint4*>(w_s + ((size_t)(kc + j) * rows + r) * 16);
          v[4 * j] = x.x; v[4 * j + 1] = x.y; v[4 * j + 2] = x.z; v[4 * j + 3] = x.w;
        }
        tc_st16(ta + (uint32_t)(kc * 4), v);
      }
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (tid == 0) {
    // gate inputs of the first step (their buffer overlays the W staging area, free only now)
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    mbar_expect_tx(smem_u32(gi_full), gi_bytes);
    bulk_g2s(smem_u32(gi_s), gi_blk0 + (long long)t0 * gi_tstride, gi_bytes, smem_u32(gi_full));
  }
  cluster_sync_all();   // every CTA's barriers and zeroed buffers exist before any remote access

  const bool clk_on = P.clk != nullptr && tid == 0 && rank == 0 && blockIdx.y == 0 && dir == 0;
  long long ck[10];
#pragma unroll
  for (int i = 0; i < 10; ++i) ck[i] = 0;
#define GRU_STAMP(i) do { if (clk_on) { const long long now_ = clock64(); ck[i] += now_ - last_; last_ = now_; } } while (0)
  long long last_ = clk_on ? clock64() : 0;

  if (warp == 8) {
    // ============================== MMA warp: gh = W_slice . h^T on the tensor core.  It only waits for the
    // h operand and issues, so the compute warps' flush / bookkeeping of step t overlaps the MMAs of t+1.
    // One elected thread runs the whole loop (no per-step warp reconvergence).
    if (elect_one()) {
      const bool mclk = P.clk != nullptr && rank == 0 && blockIdx.y == 0 && dir == 0;
      const int kpr = U >> 4, sel = kpr * 4 + n_blocks;
      if (sel == 2 * 4 + 1)
        gru_mma_role<2, 1>(P, L, smem_u32(h_s), smem_u32(h_full), smem_u32(mma_done), tmem_base, tmem_base + w_col0, rank, kpr, n_blocks,
                           slice_bytes, mclk, pub_clk);
      else if (sel == 3 * 4 + 2)
        gru_mma_role<3, 2>(P, L, smem_u32(h_s), smem_u32(h_full), smem_u32(mma_done), tmem_base, tmem_base + w_col0, rank, kpr, n_blocks,
                           slice_bytes, mclk, pub_clk);
      else
        gru_mma_role<0, 0>(P, L, smem_u32(h_s), smem_u32(h_full), smem_u32(mma_done), tmem_base, tmem_base + w_col0, rank, kpr, n_blocks,
                           slice_bytes, mclk, pub_clk);
    }
    __syncwarp();
  } else {
    // ============================== compute warps
    // (a) TMEM drain: lane = accumulator row; warps 0-3 take the first columns, warps 4-7 the rest
    // (b) gate math: pair = (utterance gn, chunk gc of 8 hidden units), chunk fastest
    const int cpu = U >> 3;
    const int n_pairs = cpu * NB;
    const bool gate_thread = tid < n_pairs;
    const int gn = tid / cpu, gc = tid - gn * cpu;
    const bool gvalid = gate_thread && (b0 + gn) < P.B;
    const int t_len = (gvalid && P.rows) ? P.rows[b0 + gn].F : T;
    const int ju = (int)rank * U + gc * 8;             // first hidden unit of this thread
    const int q4 = warp & 3;
    const int c_split = ((NB / 16 + 1) / 2) * 16;
    const int dc0 = warp < 4 ? 0 : c_split, dc1 = warp < 4 ? c_split : NB;
    float bias_r0 = 0.f, bias_r1 = 0.f;               // b_hh of the (at most two) accumulator rows this thread drains
    {
      const int r0 = q4 * 32 + lane, r1 = 128 + r0;
      if (r0 < rows) bias_r0 = P.bhh[dir * 3 * H + (r0 / U) * H + (int)rank * U + r0 % U];
      if (r1 < rows) bias_r1 = P.bhh[dir * 3 * H + (r1 / U) * H + (int)rank * U + r1 % U];
    }
    float hprev[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) hprev[i] = 0.f;

    // flush-time constants (pieces of 16 B = 8 units of one utterance at one time step)
    const int hc = H >> 3;
    const bool has_add = P.add1.base != nullptr;
    const float scale = P.scale;
    const int n_out = P.n_out;
    const float* gi_row = gi_s + (size_t)gn * S + gc * 8;
    uint8_t* st_mine = st_s + (size_t)tid * kGruStagePitch;
    const uint32_t h_off = (uint32_t)(((ju >> 3) * NB + gn) * 16);

    // Output flush, spread over time: the octet of 8 time steps staged in buffer (t8/8)&1 is written out one
    // eighth per step (one 16-byte piece per thread: lanes 0-7 of a group = 8 consecutive time rows of one
    // chunk = one 128-byte line) at the top of the following steps, while this warp would otherwise idle
    // waiting for the recurrent MMA.  Fused epilogue of the reference: (h + add1) * scale -> act -> planes.
    // Split in two so that no load latency is exposed: flush_load issues the reads of a piece (staging +
    // residual) at the top of a step, flush_store consumes them after that step's publish.
    uint4 f_hv = make_uint4(0, 0, 0, 0), f_av = make_uint4(0, 0, 0, 0);
    long long f_chunk = 0, f_row = 0;
    bool f_ok = false;
    auto flush_load = [&](int t8, int k) {
      f_ok = false;
      if (!gate_thread) return;
      const int idx = k * n_pairs + tid;
      const int pair = idx >> 3, tt = idx & 7;
      const int time = t8 + tt;
      const int pn = pair / cpu, pc = pair - pn * cpu;
      const int b = b0 + pn;
      if (time >= T || b >= P.B) return;
      f_ok = true;
      f_hv = *reinterpret_cast<const uint4*>(st_s + ((t8 >> 3) & 1) * L.st_bytes + (size_t)pair * kGruStagePitch + tt * 16);
      f_chunk = dir * hc + (int)rank * cpu + pc;
      f_row = (long long)b * P.Lh + time;
      if (has_add) f_av = *reinterpret_cast<const uint4*>(P.add1.base + ((f_chunk * P.add1.P + P.add1.G + f_row) << 4));
    };
    auto flush_store = [&]() {
      if (!f_ok) return;
      f_ok = false;
      // opaque redefinition: keeps ptxas from hoisting the bf16 unpack (and with it the wait for the loads
      // issued a whole step ago) up to the load site
      asm volatile("" : "+r"(f_av.x), "+r"(f_av.y), "+r"(f_av.z), "+r"(f_av.w), "+r"(f_hv.x), "+r"(f_hv.y), "+r"(f_hv.z), "+r"(f_hv.w));
      float v[8];
      const __nv_bfloat162* q = reinterpret_cast<const __nv_bfloat162*>(&f_hv);
      const __nv_bfloat162* qa = reinterpret_cast<const __nv_bfloat162*>(&f_av);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 f = __bfloat1622float2(q[i]);
        const float2 a = __bfloat1622float2(qa[i]);   // zeros when there is no residual
        v[2 * i] = (f.x + a.x) * scale;
        v[2 * i + 1] = (f.y + a.y) * scale;
      }
#pragma unroll
      for (int o = 0; o < 3; ++o) {
        if (o < n_out) {
          const TensDesc& td = P.out[o].t;
          char* op = td.base + (((f_chunk * td.P + td.G + f_row) * 8) << td.esize_log2);
          float w[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) w[i] = apply_act(v[i], P.out[o].act, P.out[o].a1, P.out[o].a2);
          if (td.esize_log2 == 2) {
            *reinterpret_cast<float4*>(op) = make_float4(w[0], w[1], w[2], w[3]);
            *reinterpret_cast<float4*>(op + 16) = make_float4(w[4], w[5], w[6], w[7]);
          } else {
            uint4 u;
            __nv_bfloat162* hq = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
            for (int i = 0; i < 4; ++i) hq[i] = __floats2bfloat162_rn(w[2 * i], w[2 * i + 1]);
            *reinterpret_cast<uint4*>(op) = u;
          }
        }
      }
    };
    auto flush_part = [&](int t8, int k) { flush_load(t8, k); flush_store(); };
    int cur_t8 = t0 & ~7, drain_t8 = -1, drain_k = 8;

    for (int step = 0; step < T; ++step) {
      const int buf = step & 1;
      const int t = t0 + tstep * step;
      GRU_STAMP(0);
      if ((t & ~7) != cur_t8) {
        // a new octet begins: the one just completed starts draining (the previous drain has had 8 steps)
        while (drain_k < 8) flush_part(drain_t8, drain_k++);
        drain_t8 = cur_t8; drain_k = 0; cur_t8 = t & ~7;
      }
      if (drain_k < 8) flush_load(drain_t8, drain_k++);
      GRU_STAMP(8);
      // ---- drain TMEM -> gh_s[row][utterance] (+ b_hh)
      mbar_wait(smem_u32(mma_done), step & 1, P.err_flag, 23);
      tc_fence_after();
      GRU_STAMP(1);
      for (int blk = 0; blk < n_blocks; ++blk) {
        const int r = blk * 128 + q4 * 32 + lane;
        const float bb = blk == 0 ? bias_r0 : bias_r1;
        const uint32_t ta = tmem_base + blk * NB + ((uint32_t)(q4 * 32) << 16);
        float* dst = gh_s + (size_t)r * L.gh_stride;
        uint32_t v0[16], v1[16];
        if (dc0 < dc1) tc_ld16_issue(ta + dc0, v0);
        for (int c0 = dc0; c0 < dc1; c0 += 32) {
          tc_ld_wait();
          if (c0 + 16 < dc1) tc_ld16_issue(ta + c0 + 16, v1);
          if (r < rows) {
#pragma unroll
            for (int i = 0; i < 16; ++i) dst[c0 + i] = __uint_as_float(v0[i]) + bb;
          }
          if (c0 + 16 < dc1) {
            tc_ld_wait();
            if (c0 + 32 < dc1) tc_ld16_issue(ta + c0 + 32, v0);
            if (r < rows) {
#pragma unroll
              for (int i = 0; i < 16; ++i) dst[c0 + 16 + i] = __uint_as_float(v1[i]) + bb;
            }
          }
        }
      }
      tc_fence_before();
      GRU_STAMP(2);
      mbar_wait(smem_u32(gi_full), step & 1, P.err_flag, 24);   // gate inputs of this step have landed
      compute_sync();
      GRU_STAMP(3);
      // ---- gates: thread = (8 hidden units, one utterance)
      if (gate_thread) {
        float hn[8], gif[24];
#pragma unroll
        for (int g3 = 0; g3 < 3; ++g3) {
          const float4 x = *reinterpret_cast<const float4*>(gi_row + g3 * U);
          const float4 y = *reinterpret_cast<const float4*>(gi_row + g3 * U + 4);
          gif[8 * g3] = x.x; gif[8 * g3 + 1] = x.y; gif[8 * g3 + 2] = x.z; gif[8 * g3 + 3] = x.w;
          gif[8 * g3 + 4] = y.x; gif[8 * g3 + 5] = y.y; gif[8 * g3 + 6] = y.z; gif[8 * g3 + 7] = y.w;
        }
        const float* gr = gh_s + (size_t)(gc * 8) * L.gh_stride + gn;
        const float* gz = gr + (size_t)U * L.gh_stride;
        const float* gq = gz + (size_t)U * L.gh_stride;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float r = fast_sigmoid(gif[i] + gr[i * L.gh_stride]);
          const float z = fast_sigmoid(gif[8 + i] + gz[i * L.gh_stride]);
          const float n = fast_tanh(gif[16 + i] + r * gq[i * L.gh_stride]);
          // (steps past the utterance's own length: the state is zero -- the backward direction enters its last frame
          // with h = 0, as in the reference -- and so are the outputs)
          hn[i] = (gvalid && t < t_len) ? (1.f - z) * n + z * hprev[i] : 0.f;
          hprev[i] = hn[i];
        }
        uint4 hv;
        __nv_bfloat162* hq = reinterpret_cast<__nv_bfloat162*>(&hv);
#pragma unroll
        for (int i = 0; i < 4; ++i) hq[i] = __floats2bfloat162_rn(hn[2 * i], hn[2 * i + 1]);
        // -> this CTA's own operand buffer (step+1)&1 (the bulk copies below fan it out) and the output staging
        if (step + 1 < T) {
          uint8_t* hdst = h_s + (size_t)(buf ^ 1) * L.h_bytes + h_off;
          *reinterpret_cast<uint4*>(hdst) = hv;
        }
        *reinterpret_cast<uint4*>(st_mine + ((t >> 3) & 1) * L.st_bytes + (t & 7) * 16) = hv;
      }
      GRU_STAMP(4);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      compute_sync();
      GRU_STAMP(5);
      // ---- publish h (one bulk copy per peer, completes bytes on the peer's barrier) and fetch the next gi block
      if (step + 1 < T && warp == 1 && elect_one()) {
        if (P.clk) *pub_clk = clock64();
        const uint32_t bar0 = smem_u32(&h_full[(buf ^ 1) * kGruGroups]);   // group barriers of the buffer, same offsets in every CTA
        mbar_arrive(bar0);   // local slice is in place (fenced + barrier above)
        const uint32_t local = smem_u32(h_s + (size_t)(buf ^ 1) * L.h_bytes) + rank * slice_bytes;
        // (per-thread st.async stores with complete_tx land no earlier than these copies and lengthen the gate phase)
#pragma unroll
        for (uint32_t c = 1; c < kGruNC; ++c) {
          const uint32_t peer = (rank + kGruNC - c) & (kGruNC - 1);    // peer consumes us as its c-th slice
          bulk_s2peer(map_to_rank(local, peer), local, slice_bytes, map_to_rank(bar0 + 8u * gru_group_of((int)c), peer));
        }
        mbar_expect_tx(smem_u32(gi_full), gi_bytes);
        bulk_g2s(smem_u32(gi_s), gi_blk0 + (long long)(t + tstep) * gi_tstride, gi_bytes, smem_u32(gi_full));
      }
      GRU_STAMP(6);
      flush_store();
      GRU_STAMP(7);
      // gh_s / gi_s are rewritten only after the next step's compute_sync, and every warp finishes the
      // reads above before it arrives there; a staging buffer is rewritten 8 steps after its drain began.
    }
    while (drain_k < 8) flush_part(drain_t8, drain_k++);
    for (int k = 0; k < 8; ++k) flush_part(cur_t8, k);   // last octet (its last slot was published by the barrier above)
  }

  if (clk_on) {
#pragma unroll
    for (int i = 0; i < 10; ++i) P.clk[i] = ck[i];
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // no CTA may exit while peers can still write into its shared memory
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
  }
}

// Host-side check of a (H, NB) choice made by the program builder (program.py gru_geometry()).
inline bool gru_config_ok(int H, int NB) {
  if (H % (16 * kGruNC) || H > kGruMaxH || NB % 16 || NB < 16 || NB > kGruMaxNB) return false;
  const int U = H / kGruNC;
  return (U / 8) * NB <= kGruCompute && gru_smem_layout(H, NB).total <= 227 * 1024;
}

inline cudaError_t gru_launch(const GruParams& g, cudaStream_t s) {
  if (!gru_config_ok(g.H, g.NB)) return cudaErrorInvalidConfiguration;
  const GruSmem L = gru_smem_layout(g.H, g.NB);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(kGruNC, (g.B + g.NB - 1) / g.NB, 2);
  cfg.blockDim = dim3(kGruThreads);
  cfg.dynamicSmemBytes = L.total;
  cfg.stream = s;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = kGruNC;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, gru_kernel, g);
}

}  // namespace uv

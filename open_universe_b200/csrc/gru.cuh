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
//   * gi and the residual operand of step t+1 are prefetched while step t computes.
// The output epilogue fuses the residual add / scale / PReLU that follows the GRU in the
// reference ((out + res)/sqrt2: condition.py:215, blocks.py:380) and writes planes tensors.
#pragma once
#include "gemm_tc.cuh"

namespace uv {

constexpr int kGruNC = 8;          // CTAs per cluster (portable maximum)
constexpr int kGruThreads = 288;   // 8 compute warps + 1 MMA-issue warp
constexpr int kGruCompute = 256;
constexpr int kGruMaxNB = 64;      // utterances per cluster (UMMA N)
constexpr int kGruMaxH = 384;

struct GruParams {
  const float* gi;          // [rows][6H], rows = b*Lh + t, cols = dir*3H + gate*H + j
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
};

struct GruSmem {
  int w_bytes, h_bytes, gh_stride, gh_bytes, total;
};
__host__ __device__ inline GruSmem gru_smem_layout(int H, int NB) {
  GruSmem s;
  const int U = H / kGruNC, rows = 3 * U;
  s.w_bytes = rows * H * 2 + 128 * 16;          // + one row block of slack: the last M=128 block may over-read
  s.h_bytes = NB * H * 2;                       // one h operand buffer [H/8][NB][8] bf16
  s.gh_stride = NB + 1;
  s.gh_bytes = ((rows * s.gh_stride * 4) + 127) & ~127;
  s.total = s.w_bytes + 2 * s.h_bytes + s.gh_bytes + 64;
  return s;
}

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
    __nanosleep(40);   // the waiter shares its scheduler with two compute warps: do not burn their issue slots
    if (++spins > (1u << 22)) {
      if (err_flag) atomicExch(err_flag, code);
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
__device__ __forceinline__ float fast_sigmoid(float x) { return rcp_approx(1.f + ex2_approx(-1.4426950408889634f * x)); }
__device__ __forceinline__ float fast_tanh(float x) { return fmaf(2.f, rcp_approx(1.f + ex2_approx(-2.8853900817779268f * x)), -1.f); }

__global__ void __launch_bounds__(kGruThreads, 1) gru_kernel(const __grid_constant__ GruParams P) {
  extern __shared__ __align__(1024) uint8_t gsm[];
  const int H = P.H, U = H / kGruNC, rows = 3 * U, NB = P.NB;
  const GruSmem L = gru_smem_layout(H, NB);
  uint8_t* w_s = gsm;
  uint8_t* h_s = gsm + L.w_bytes;                       // two operand buffers
  float* gh_s = reinterpret_cast<float*>(h_s + 2 * L.h_bytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(gh_s) + L.gh_bytes);
  uint64_t* h_full = bars;        // [2]: all 8 CTAs have written their slice of h into buffer i
  uint64_t* mma_done = bars + 2;  // [1]
  uint64_t* w_full = bars + 3;    // [1]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 4);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_rank();
  const int dir = blockIdx.z;
  const int b0 = blockIdx.y * NB;
  const int n_blocks = (rows + 127) / 128;
  uint32_t tmem_cols = 32;
  while ((int)tmem_cols < n_blocks * NB) tmem_cols <<= 1;

  if (tid == 0) {
    // per fill: one arming arrive (expect_tx of the 7 peer slices) + one arrive for the local slice
    mbar_init(smem_u32(&h_full[0]), 2);
    mbar_init(smem_u32(&h_full[1]), 2);
    mbar_init(smem_u32(mma_done), 1);
    mbar_init(smem_u32(w_full), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const uint32_t peer_bytes = (uint32_t)(kGruNC - 1) * (uint32_t)(U >> 3) * NB * 16u;
    if (P.T > 1) mbar_expect_tx(smem_u32(&h_full[1]), peer_bytes);   // filled at the end of step 0
    if (P.T > 2) mbar_expect_tx(smem_u32(&h_full[0]), peer_bytes);   // filled at the end of step 1
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
  cluster_sync_all();   // every CTA's barriers and zeroed buffers exist before any remote access
  const uint32_t tmem_base = *tmem_ptr_smem;

  // ---- static thread roles
  // (a) TMEM drain: warps 0-3, lane = accumulator row within a 128-row block
  // (b) gate math: pair index = (chunk c of 8 hidden units, utterance n)
  const int n_pairs = (U >> 3) * NB;
  const bool gate_thread = tid < n_pairs;
  // chunk index fastest: the U/8 lanes of one utterance read one contiguous 4*U-byte run of each gi gate row
  const int cpu = U >> 3;
  const int gn = tid / cpu, gc = tid - gn * cpu;    // utterance within the group, chunk within the slice
  const int gb = b0 + gn;
  const bool gvalid = gate_thread && gb < P.B;
  const int ju = (int)rank * U + gc * 8;             // first hidden unit of this thread
  float bias_r0 = 0.f, bias_r1 = 0.f;               // b_hh of the (at most two) accumulator rows this thread drains
  if (warp < 4) {
    const int r0 = warp * 32 + lane, r1 = 128 + r0;
    if (r0 < rows) bias_r0 = P.bhh[dir * 3 * H + (r0 / U) * H + (int)rank * U + r0 % U];
    if (r1 < rows) bias_r1 = P.bhh[dir * 3 * H + (r1 / U) * H + (int)rank * U + r1 % U];
  }
  float hprev[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) hprev[i] = 0.f;

  // Everything the loop touches per step is a pointer that advances by a constant (time runs forwards
  // or backwards): gi row, residual row, output rows.  No parameter-space reads inside the loop.
  const int hc = H >> 3;                               // chunks per direction in the output tensors
  const int out_chunk = dir * hc + (ju >> 3);
  const int t0 = dir == 0 ? 0 : P.T - 1;
  const long long tstep = dir == 0 ? 1 : -1;
  const bool has_add = P.add1.base != nullptr;
  const bool add_f32 = has_add && P.add1.esize_log2 == 2;
  const float scale = P.scale;
  const int n_out = P.n_out;
  const float* gi_p = nullptr;
  const char* add_p = nullptr;
  long long add_inc = 0;
  char* out_p[3] = {nullptr, nullptr, nullptr};
  long long out_inc[3] = {0, 0, 0};
  bool out_f32[3] = {false, false, false};
  int o_act[3];
  float o_a1[3], o_a2[3];
#pragma unroll
  for (int o = 0; o < 3; ++o) { o_act[o] = P.out[o].act; o_a1[o] = P.out[o].a1; o_a2[o] = P.out[o].a2; }
  if (gvalid) {
    gi_p = P.gi + ((size_t)gb * P.Lh + t0) * (6 * H) + dir * 3 * H + ju;
    if (has_add) {
      const RowAddr r = row_addr(P.add1, gb, t0);    // GRU operands are plain (split == 1): row t+1 is +16/32 bytes
      add_p = r.p + (long long)out_chunk * r.pstride;
      add_inc = tstep * (8ll << P.add1.esize_log2);
    }
#pragma unroll
    for (int o = 0; o < 3; ++o) {
      if (o < n_out) {
        const RowAddr r = row_addr(P.out[o].t, gb, t0);
        out_p[o] = r.p + (long long)out_chunk * r.pstride;
        out_inc[o] = tstep * (8ll << P.out[o].t.esize_log2);
        out_f32[o] = P.out[o].t.esize_log2 == 2;
      }
    }
  }
  const long long gi_inc = tstep * 6 * H;

  // register double buffer for the step-ahead prefetch of gi (3 gates x 8 units) and the residual
  float4 gcur[6], gnxt[6];
  uint4 acur[2], anxt[2];
#pragma unroll
  for (int i = 0; i < 6; ++i) gcur[i] = gnxt[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  acur[0] = acur[1] = anxt[0] = anxt[1] = make_uint4(0, 0, 0, 0);
  if (gvalid) {
#pragma unroll
    for (int gate = 0; gate < 3; ++gate) {
      gnxt[2 * gate] = __ldg(reinterpret_cast<const float4*>(gi_p + gate * H));
      gnxt[2 * gate + 1] = __ldg(reinterpret_cast<const float4*>(gi_p + gate * H + 4));
    }
    if (has_add) {
      anxt[0] = *reinterpret_cast<const uint4*>(add_p);
      if (add_f32) anxt[1] = *reinterpret_cast<const uint4*>(add_p + 16);
    }
  }

  const uint32_t idesc = make_idesc(128, NB);
  const uint32_t lbo_w = (uint32_t)rows * 16u, lbo_h = (uint32_t)NB * 16u;
  const uint32_t peer_bytes = (uint32_t)(kGruNC - 1) * (uint32_t)(U >> 3) * NB * 16u;
  const uint32_t slice_bytes = (uint32_t)(U >> 3) * NB * 16u;
  const uint64_t wdesc0 = make_desc(smem_u32(w_s), lbo_w, 128);
  const uint64_t hdesc0 = make_desc(smem_u32(h_s), lbo_h, 128);
  const uint64_t wdesc_kstep = (uint64_t)((2u * lbo_w) >> 4), hdesc_kstep = (uint64_t)((2u * lbo_h) >> 4);
  const uint64_t hdesc_buf = (uint64_t)((uint32_t)L.h_bytes >> 4);
  const int ksteps = H >> 4;
  const bool clk_on = P.clk != nullptr && tid == 0 && rank == 0 && blockIdx.y == 0 && dir == 0;
  long long ck[10];
#pragma unroll
  for (int i = 0; i < 10; ++i) ck[i] = 0;
#define GRU_STAMP(i) do { if (clk_on) { const long long now_ = clock64(); ck[i] += now_ - last_; last_ = now_; } } while (0)
  long long last_ = clk_on ? clock64() : 0;
  const int T = P.T;

  if (warp == 8) {
    // ============================== MMA warp: gh = W_slice . h^T on the tensor core, one step ahead of nothing
    // else: it only waits for the h operand and issues; the compute warps' epilogue of step t overlaps it.
    mbar_wait(smem_u32(w_full), 0, P.err_flag, 21);
    for (int step = 0; step < T; ++step) {
      const int buf = step & 1;
      if (step > 0) mbar_wait_cluster(smem_u32(&h_full[buf]), ((step - 1) >> 1) & 1, P.err_flag, 22);
      tc_fence_after();
      if (elect_one()) {
        // this buffer is filled again at the end of step+1 (for step+2): arm that phase now
        if (step > 0 && step + 2 < T) mbar_expect_tx(smem_u32(&h_full[buf]), peer_bytes);
        const uint64_t hd = hdesc0 + (buf ? hdesc_buf : 0);
        for (int blk = 0; blk < n_blocks; ++blk) {
          const uint64_t wd = wdesc0 + (uint64_t)(blk * 128);   // 128 rows x 16 B >> 4
          const uint32_t d = tmem_base + blk * NB;
#pragma unroll 4
          for (int kk = 0; kk < ksteps; ++kk) tc_mma_bf16(d, wd + kk * wdesc_kstep, hd + kk * hdesc_kstep, idesc, kk > 0 ? 1u : 0u);
        }
        tc_commit(smem_u32(mma_done));
      }
      __syncwarp();
    }
  } else {
  // ============================== compute warps
  for (int step = 0; step < T; ++step) {
    const int buf = step & 1;
    GRU_STAMP(0);
    // ---- rotate the prefetch registers (loads issued one whole step ago) and issue the next step's loads
#pragma unroll
    for (int i = 0; i < 6; ++i) gcur[i] = gnxt[i];
    acur[0] = anxt[0];
    acur[1] = anxt[1];
    if (gvalid && step + 1 < T) {
      gi_p += gi_inc;
#pragma unroll
      for (int gate = 0; gate < 3; ++gate) {
        gnxt[2 * gate] = __ldg(reinterpret_cast<const float4*>(gi_p + gate * H));
        gnxt[2 * gate + 1] = __ldg(reinterpret_cast<const float4*>(gi_p + gate * H + 4));
      }
      if (has_add) {
        add_p += add_inc;
        anxt[0] = *reinterpret_cast<const uint4*>(add_p);
        if (add_f32) anxt[1] = *reinterpret_cast<const uint4*>(add_p + 16);
      }
    }
    GRU_STAMP(3);
    // ---- drain TMEM -> gh_s[row][utterance] (+ b_hh)
    if (warp < 4) {
      mbar_wait(smem_u32(mma_done), step & 1, P.err_flag, 23);
      tc_fence_after();
      GRU_STAMP(4);
      for (int blk = 0; blk < n_blocks; ++blk) {
        const int r = blk * 128 + warp * 32 + lane;
        const float bb = blk == 0 ? bias_r0 : bias_r1;
        const uint32_t ta = tmem_base + blk * NB + ((uint32_t)(warp * 32) << 16);
        float* dst = gh_s + (size_t)r * L.gh_stride;
        uint32_t v0[16], v1[16];
        tc_ld16_issue(ta, v0);
        for (int c0 = 0; c0 < NB; c0 += 32) {
          tc_ld_wait();
          if (c0 + 16 < NB) tc_ld16_issue(ta + c0 + 16, v1);
          if (r < rows) {
#pragma unroll
            for (int i = 0; i < 16; ++i) dst[c0 + i] = __uint_as_float(v0[i]) + bb;
          }
          if (c0 + 16 < NB) {
            tc_ld_wait();
            if (c0 + 32 < NB) tc_ld16_issue(ta + c0 + 32, v0);
            if (r < rows) {
#pragma unroll
              for (int i = 0; i < 16; ++i) dst[c0 + 16 + i] = __uint_as_float(v1[i]) + bb;
            }
          }
        }
      }
      tc_fence_before();
    }
    GRU_STAMP(5);
    compute_sync();
    GRU_STAMP(6);
    // ---- gates: thread = (8 hidden units, one utterance)
    if (gate_thread) {
      float hn[8];
      const float* gr = gh_s + (size_t)(gc * 8) * L.gh_stride + gn;
      const float* gz = gr + (size_t)U * L.gh_stride;
      const float* gq = gz + (size_t)U * L.gh_stride;
      float gif[24];
#pragma unroll
      for (int i = 0; i < 6; ++i) { gif[4 * i] = gcur[i].x; gif[4 * i + 1] = gcur[i].y; gif[4 * i + 2] = gcur[i].z; gif[4 * i + 3] = gcur[i].w; }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float r = fast_sigmoid(gif[i] + gr[i * L.gh_stride]);
        const float z = fast_sigmoid(gif[8 + i] + gz[i * L.gh_stride]);
        const float n = fast_tanh(gif[16 + i] + r * gq[i * L.gh_stride]);
        hn[i] = gvalid ? (1.f - z) * n + z * hprev[i] : 0.f;
        hprev[i] = hn[i];
      }
      // new h slice -> this CTA's own operand buffer (step+1)&1; the bulk copies below fan it out
      if (step + 1 < T) {
        uint4 hv;
        __nv_bfloat162* hq = reinterpret_cast<__nv_bfloat162*>(&hv);
#pragma unroll
        for (int i = 0; i < 4; ++i) hq[i] = __floats2bfloat162_rn(hn[2 * i], hn[2 * i + 1]);
        *reinterpret_cast<uint4*>(h_s + (size_t)(buf ^ 1) * L.h_bytes + (size_t)(((ju >> 3) * NB + gn) * 16)) = hv;
      }
    }
    GRU_STAMP(7);
    // ---- publish: the slice [U/8 chunks][NB][8] is contiguous in every CTA's operand buffer, so one bulk
    // copy per peer moves it through the DSMEM fabric and completes its bytes on that peer's barrier
    if (step + 1 < T) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      compute_sync();
      if (warp == 1 && elect_one()) {
        const uint32_t local = smem_u32(h_s + (size_t)(buf ^ 1) * L.h_bytes) + rank * slice_bytes;
        const uint32_t bar = smem_u32(&h_full[buf ^ 1]);
        for (uint32_t c = 1; c < kGruNC; ++c) {
          const uint32_t peer = (rank + c) & (kGruNC - 1);
          bulk_s2peer(map_to_rank(local, peer), local, slice_bytes, map_to_rank(bar, peer));
        }
        mbar_arrive(bar);
      }
    }
    GRU_STAMP(8);
    // ---- output epilogue (off the critical path): (h + add1) * scale -> act -> planes
    if (gvalid) {
      float v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = hprev[i];
      if (has_add) {
        if (!add_f32) {
          const __nv_bfloat162* q = reinterpret_cast<const __nv_bfloat162*>(&acur[0]);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float2 f = __bfloat1622float2(q[i]);
            v[2 * i] += f.x;
            v[2 * i + 1] += f.y;
          }
        } else {
          v[0] += __uint_as_float(acur[0].x); v[1] += __uint_as_float(acur[0].y);
          v[2] += __uint_as_float(acur[0].z); v[3] += __uint_as_float(acur[0].w);
          v[4] += __uint_as_float(acur[1].x); v[5] += __uint_as_float(acur[1].y);
          v[6] += __uint_as_float(acur[1].z); v[7] += __uint_as_float(acur[1].w);
        }
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] *= scale;
#pragma unroll
      for (int o = 0; o < 3; ++o) {
        if (o < n_out) {
          float w[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) w[i] = apply_act(v[i], o_act[o], o_a1[o], o_a2[o]);
          if (out_f32[o]) {
            *reinterpret_cast<float4*>(out_p[o]) = make_float4(w[0], w[1], w[2], w[3]);
            *reinterpret_cast<float4*>(out_p[o] + 16) = make_float4(w[4], w[5], w[6], w[7]);
          } else {
            uint4 u;
            __nv_bfloat162* hq = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
            for (int i = 0; i < 4; ++i) hq[i] = __floats2bfloat162_rn(w[2 * i], w[2 * i + 1]);
            *reinterpret_cast<uint4*>(out_p[o]) = u;
          }
          out_p[o] += out_inc[o];
        }
      }
    }
    GRU_STAMP(9);
    // gh_s is rewritten only after the next step's MMAs complete, which need this CTA's own slice,
    // published after the barrier above: every warp is past its gh_s reads by then.
  }
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

// Utterances per cluster for hidden size H: limited by the UMMA N (64), the gate-thread mapping
// (U/8 * NB <= 256 threads) and shared memory.
inline int gru_pick_nb(int H, int B) {
  const int U = H / kGruNC;
  int nb = kGruMaxNB;
  while (nb > 16 && ((U / 8) * nb > kGruCompute || gru_smem_layout(H, nb).total > 227 * 1024)) nb -= 16;
  const int need = (B + 15) / 16 * 16;
  return need < nb ? need : nb;
}

inline cudaError_t gru_launch(GruParams g, cudaStream_t s) {
  g.NB = gru_pick_nb(g.H, g.B);
  const GruSmem L = gru_smem_layout(g.H, g.NB);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(kGruNC, (g.B + g.NB - 1) / g.NB, 2);
  cfg.blockDim = dim3(kGruThreads);
  cfg.dynamicSmemBytes = L.total;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = kGruNC;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, gru_kernel, g);
}

}  // namespace uv

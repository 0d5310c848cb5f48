// Multi-tap implicit-GEMM conv1d on the sm_100a tensor cores (tcgen05 + TMEM).
//
//   D[row, n] = bias[n] + sum_{src s} sum_{shift q} sum_{c}  A_s[row + q, c] * W[s][q][n][c]
//
// covers every dense layer of the reference's ConvBlock
// (open_universe/networks/universe/blocks.py:331-416): "same" k=3/5 convs
// (one source, shifts -pad..pad), anti-aliased strided down convs (r phase
// sources, shifts -1..1; blocks.py:217-221 folded), transposed up convs (one
// source, N = r phases x C_out; blocks.py:223-225 folded), the conditioner's
// st_convs (condition.py:53-59) and the GRU input projection.
//
// Tile = 256 rows x N_TILE columns, issued as two M=128 UMMAs per K step that share
// every weight image (halves the L2->smem weight traffic per FLOP; the 8 epilogue
// warps split into one group per 128-row half).  A tiles are halo'd
// [KC/8][256 + smax - smin][8] bf16 slabs fetched with 1-D bulk async copies straight
// out of the planes layout (uv_common.cuh); that is the canonical no-swizzle K-major
// UMMA layout, in which a tap shift is +16 bytes on the descriptor start address, so
// one slab serves all taps.  Weight tiles are pre-packed images [KC/8][N_TILE][8].
// The bias enters through the tensor core as well: a constant "ones" A tile times a
// [hi | lo] bf16 split of the fp32 bias initialises the accumulator, so the epilogue
// carries no per-column bias loads.
//
// Warp roles (320 threads): warp0 = bulk-copy producer, warp1 = MMA issuer and TMEM
// owner, warps 2-9 = epilogue (TMEM -> registers -> fused add / scale / FiLM / PReLU ->
// coalesced 16-byte stores into up to three planes tensors).  The epilogue is
// specialised at compile time on (has adds, has FiLM, single plain output) and
// software-pipelined: the TMEM load and the residual loads of column group g+1 are in
// flight while group g is processed.  2-4 accumulator stages live in TMEM so the MMAs
// run ahead of the epilogues.  Layers whose whole weight set fits keep it resident in
// shared memory and give the rest to a deep A ring.  A tensor consumed through a PReLU may be
// stored RAW (a_act): two transform warps then apply the activation to every A slab in shared
// memory between its arrival and the MMAs (fp32 math, one bf16 rounding), so the producer of
// that tensor writes ONE copy instead of a raw and an activated one -- the layers around it are
// HBM bound.  Persistent grid with a DYNAMIC tile
// schedule: the producer thread claims tile indices from a global counter (two claims in
// flight, so the atomic's latency hides behind the current tile's copies) and publishes them
// through a small shared-memory ring to the MMA issuer and the epilogue warps; -1 ends the
// kernel.  A CTA that becomes resident late (its SM was busy with another stream's kernel,
// e.g. the GRU clusters of the other utterance lane) finds the counter exhausted and exits at
// once instead of holding a statically assigned share of the tiles.  The last CTA to leave
// resets the counter pair, so the launch is replayable without a memset.
#pragma once
#include "uv_common.cuh"

namespace uv {

constexpr int kSubM = 128;          // rows of a tile owned by one CTA of the pair (= its TMEM lanes)
constexpr int kTileM = 256;         // rows per tile = UMMA M of the SM pair (tcgen05 cta_group::2)
constexpr int kMaxAStages = 16;
constexpr int kMaxBStages = 12;
constexpr int kEpiWarps = 8;        // per CTA: 4 TMEM lane quarters x 2 halves of the tile's column groups
#ifndef UV_XF_WARPS
#define UV_XF_WARPS 2   // 4 removes the transform cost at level 1 (k5: 104 vs 125 us) but costs 128 registers + spills everywhere else: 45.3 vs 45.2 ms per pass
#endif
constexpr int kXfWarps = UV_XF_WARPS;         // A-operand transform warps (PReLU of a raw tensor in shared memory)
constexpr int kGemmThreads = 64 + 32 * kEpiWarps + 32 * kXfWarps;   // producer warp, MMA warp, epilogue warps, transform warps
constexpr int kMaxAcc = 4;          // TMEM accumulator stages
constexpr int kSchedSlots = 32;     // ring of claimed tile indices (producer -> MMA issuer + epilogue warps)
constexpr int kSchedBytes = kSchedSlots * (4 + 2 * 8);
constexpr int kPairCtas = 2;        // CTAs per cluster: one tcgen05 CTA pair
constexpr int kOnesBytes = 2 * kSubM * 16;   // [2 chunks][128 rows][8] bf16
constexpr int kFilmMaxC = 768;
constexpr int kFilmBytes = 2 * kFilmMaxC * 4;
constexpr int kGemmSmemBudget = 227 * 1024;
constexpr int kRmPitch = 20;        // floats per row of the GRU-layout staging buffer (16 + 4: conflict-free 16-byte accesses)
constexpr int kRmScratchBytes = kEpiWarps * 32 * kRmPitch * 4;
constexpr int kMvEntries = 128;              // per-utterance valid-row table kept in shared memory (variable-length calls)
constexpr int kMvBytes = kMvEntries * 4;
constexpr int kAddRowsBytes = kSubM * 16;    // one 8-channel chunk of a residual slab (no halo), this CTA's rows

// TMEM columns per accumulator (power of two >= 32).
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
// Non-blocking probe (try_wait may suspend the thread for a while when the phase is not complete).
__device__ __forceinline__ bool mbar_test_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
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
      // err_flag is mapped pinned host memory (capi.cu): a plain store + system fence reaches the host even though
      // the trap below kills the context
      if (err_flag) *reinterpret_cast<volatile int*>(err_flag) = code;
      __threadfence_system();
      __trap();
    }
  }
}
// ---- CTA-pair plumbing.  Remote arrives and the waits on barriers that receive them use the DEFAULT (cta-scope)
// semantics, like every cluster barrier of CUTLASS: what they order lives in shared / tensor memory and is moved by
// the async proxy or fenced by tcgen05.fence / fence.proxy.async before the arrive.  Cluster-scope release / acquire
// compile to MEMBAR.ALL.GPU + CGAERRBAR / CCTL.IVALL on every arrive / wait, which made the epilogue warps wait for
// all their outstanding global stores at every tile (level-0 convs 3.5x slower, measured).  The one datum that crosses
// CTAs through a generic store -- the claimed tile index -- is sent with st.async, whose completion is part of the
// receiving barrier's transaction count.
__device__ __forceinline__ uint32_t cta_rank_in_cluster() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t map_shared_rank(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
// arrive on a barrier given by its shared::cluster address (own or peer CTA)
__device__ __forceinline__ void mbar_arrive_cl(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// 4-byte store into the peer's shared memory that completes 4 bytes of the transaction count of the peer's barrier
// (armed by mbar_expect_tx_cl): the value is visible to whoever observes that barrier phase complete.
__device__ __forceinline__ void mbar_expect_tx_cl(uint32_t cluster_bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cluster.b64 _, [%0], %1;" ::"r"(cluster_bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void st_async_u32(uint32_t cluster_addr, uint32_t v, uint32_t cluster_bar) {
  asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.u32 [%0], %1, [%2];" ::"r"(cluster_addr), "r"(v), "r"(cluster_bar)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cl(uint32_t bar, uint32_t parity) {
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
__device__ __forceinline__ bool mbar_test_wait_cl(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_cl(uint32_t bar, uint32_t parity, int* err_flag, int code) {
  uint32_t spins = 0;
  while (!mbar_try_wait_cl(bar, parity)) {
    if (++spins > (1u << 26)) {
      if (err_flag) *reinterpret_cast<volatile int*>(err_flag) = code;
      __threadfence_system();
      __trap();
    }
  }
}
__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}
// One lane of the (converged) warp; lets ptxas see that the role body runs single-threaded, so
// descriptors / addresses go through uniform registers without a per-lane serialisation loop.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// phase stamps of one CTA (debug; P.clk == nullptr in production): 0 entry, 1 prologue done, 2 first tile's loads
// issued, 3 resident weights landed, 4 first A slab landed, 5 first accumulator complete, 6 last tile stored, 7 exit
#define UV_STAMP(i) do { if (P.clk) P.clk[(size_t)blockIdx.x * 8 + (i)] = gtimer(); } while (0)
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
// The SM pair: M = 256 (128 rows per CTA), each CTA supplies N / 2 columns of B, issued by the leader CTA only;
// the descriptors name the same shared-memory offsets in both CTAs.
__device__ __forceinline__ void tc_mma2_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrives, once all earlier MMAs of the pair have completed, on the barrier at this offset in BOTH CTAs of the pair.
__device__ __forceinline__ void tc_commit2(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"((uint16_t)3)
               : "memory");
}
// 32 lanes x 16 consecutive fp32 columns -> 16 registers; completes at tc_ld_wait().
__device__ __forceinline__ void tc_ld16_issue(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tc_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

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

// ------------------------------------------------------------------ epilogue helpers
// PReLU: slopes <= 1 (every trained / initialised slope of these networks) take the two-instruction
// form max(v, a v); the LE1 = false instantiation is the general compare-free form.
template <bool LE1>
__device__ __forceinline__ float prelu_t(float v, float a) {
  if (LE1) return fmaxf(v, a * v);
  return fmaf(a, fminf(v, 0.f), fmaxf(v, 0.f));
}

// PReLU of a bf16 pair: the products a x in fp32 (packed multiply), ONE rounding to bf16, then a packed max / min
// against x itself: for a <= 1, max(x, rn(a x)) is x for x >= 0 and rn(a x) for x < 0 (rounding is monotone and x
// is a bf16 value); for a > 1 the same holds with min.
__device__ __forceinline__ uint32_t prelu_bf16x2(uint32_t x, float a, bool le1) {
  const float lo = __uint_as_float(x << 16), hi = __uint_as_float(x & 0xffff0000u);
  uint64_t s2, a2, p2;
  asm("mov.b64 %0, {%1, %2};" : "=l"(s2) : "f"(lo), "f"(hi));
  asm("mov.b64 %0, {%1, %1};" : "=l"(a2) : "f"(a));
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(p2) : "l"(s2), "l"(a2));
  float plo, phi;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(plo), "=f"(phi) : "l"(p2));
  uint32_t p, r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(p) : "f"(phi), "f"(plo));
  if (le1) asm("max.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(p));
  else asm("min.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(p));
  return r;
}

// floor(g / d) for 0 <= g < 2^24-ish without an integer division (per tile, per thread)
__device__ __forceinline__ int fast_div(int g, int d, float inv_d) {
  int q = __float2int_rz(((float)g + 0.5f) * inv_d);
  const int r = g - q * d;
  if (r < 0) --q;
  else if (r >= d) ++q;
  return q;
}

// One output tensor of the epilogue, reduced to what the column loop needs.
struct OutPtr {
  char* p;             // (row, first chunk of this tile); nullptr = nothing to store for this thread
  long long pstride;   // bytes between chunk planes
};

__device__ __forceinline__ OutPtr out_ptr(const TensDesc& t, int b, int l, int chunk0, float inv_split) {
  OutPtr o;
  int ph = 0, ll = l;
  if (t.split > 1) {
    ll = fast_div(l, t.split, inv_split);
    ph = l - ll * t.split;
  }
  const long long row = (long long)ph * t.nchunks * t.P + t.G + (long long)b * t.Lh + ll;
  o.pstride = ((long long)t.P * 8) << t.esize_log2;
  o.p = t.base + ((row * 8) << t.esize_log2) + (long long)chunk0 * o.pstride;
  return o;
}

// 16 columns -> one output tensor.  a1 = 1 stands for "no activation"; both tests are warp uniform.
template <bool LE1>
__device__ __forceinline__ void store16_act(char* p, long long pstride, bool f32, float a1, bool dbl, float a2, const float* v) {
  float w[16];
  if (a1 != 1.f) {
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = prelu_t<LE1>(v[i], a1);
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = v[i];
  }
  if (dbl) {
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = prelu_t<LE1>(w[i], a2);
  }
  if (!f32) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint4 u;
      __nv_bfloat162* q = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
      for (int i = 0; i < 4; ++i) q[i] = __floats2bfloat162_rn(w[8 * h + 2 * i], w[8 * h + 2 * i + 1]);
      *reinterpret_cast<uint4*>(p + h * pstride) = u;
    }
  } else {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      *reinterpret_cast<float4*>(p + h * pstride) = make_float4(w[8 * h], w[8 * h + 1], w[8 * h + 2], w[8 * h + 3]);
      *reinterpret_cast<float4*>(p + h * pstride + 16) = make_float4(w[8 * h + 4], w[8 * h + 5], w[8 * h + 6], w[8 * h + 7]);
    }
  }
}

struct AddRegs {
  uint4 a1[2], a2[2];  // 16 bf16 columns of each residual operand
};

__device__ __forceinline__ void add_unpack(const uint4* a, float* v) {
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const __nv_bfloat162* q = reinterpret_cast<const __nv_bfloat162*>(&a[h]);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __bfloat1622float2(q[i]);
      v[8 * h + 2 * i] += f.x;
      v[8 * h + 2 * i + 1] += f.y;
    }
  }
}
__device__ __forceinline__ void add_f32(const char* p, long long pstride, float* v) {
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const float4 x = *reinterpret_cast<const float4*>(p + h * pstride);
    const float4 y = *reinterpret_cast<const float4*>(p + h * pstride + 16);
    v[8 * h] += x.x; v[8 * h + 1] += x.y; v[8 * h + 2] += x.z; v[8 * h + 3] += x.w;
    v[8 * h + 4] += y.x; v[8 * h + 5] += y.y; v[8 * h + 6] += y.z; v[8 * h + 7] += y.w;
  }
}

// Epilogue of one warp: lane = TMEM lane = output row among this CTA's 128 rows of every tile (rows 128 rank ...);
// two warps share a lane quarter and take one half of the tile's 16-column groups each.
// The work is a flat sequence of units (tile, 16-column group) run as a two-buffer software pipeline
// that crosses tile boundaries: while unit u is processed, the TMEM load of unit u+1 is in flight, also
// when u+1 is the first group of the next tile (its accumulator barrier is probed without blocking).
// Per-tile address work is done once (no integer divisions; 64-bit pointers advance by a constant per
// group) and overlaps that load.
template <bool ADD, bool FILM, bool MULTI, bool LE1>
__device__ __forceinline__ void epilogue_role(const GemmParams& P, uint32_t tmem_base, uint64_t* t_full, uint64_t* t_empty,
                                              const float* film_s, const int* s_ring, uint64_t* s_full, uint64_t* s_empty,
                                              float* rm_scratch, const int* mv_s, int warp, int lane, uint32_t rank) {
  // msub = 1: the two warps of a lane quarter split the tile's column groups; msub = 2 (narrow tiles): they take one
  // 128-row sub-tile of this CTA's 256 rows each
  const int msub = P.msub;
  const int wh = (warp - 2) >> 2;               // 0 / 1: which of the two warps of this lane quarter
  const int sub = msub == 2 ? wh : 0;
  const int chalf = msub == 2 ? 0 : wh;
  const int quarter = warp & 3;                 // a warp may only touch TMEM lanes 32*(warp%4)..+31
  const int tile_rows = kTileM * msub;
  const int row_in_tile = (int)rank * kSubM * msub + sub * kSubM + quarter * 32 + lane;
  const int ngroups = P.N_TILE >> 4;
  const int g_begin = chalf ? (ngroups + 1) >> 1 : 0;
  const int g_end = (msub == 2 || chalf) ? ngroups : (ngroups + 1) >> 1;
  const uint32_t acc_stride = accum_stride_cols(P.N_TILE);
  const int n_acc = P.n_acc;
  const int ntn = P.n_tiles_n;
  const float inv_ntn = 1.f / (float)ntn;
  const bool a1_bf16 = ADD && P.add1.esize_log2 == 1;
  const bool a2_on = ADD && P.add2.base != nullptr;
  const bool a2_bf16 = a2_on && P.add2.esize_log2 == 1;
  const float scale = P.scale;
  const bool do_scale = scale != 1.f;
  const int n_out = P.n_out;
  const float inv_Lh = 1.f / (float)P.Lh_A;
  const float inv_Cout = 1.f / (float)P.C_out;
  float inv_split[3], o_a1[3];
  bool o_f32[3], o_dbl[3];
#pragma unroll
  for (int o = 0; o < 3; ++o) {
    inv_split[o] = 1.f / (float)(P.out[o].t.split > 0 ? P.out[o].t.split : 1);
    o_f32[o] = P.out[o].t.esize_log2 == 2;
    o_a1[o] = P.out[o].act >= 1 ? P.out[o].a1 : 1.f;
    o_dbl[o] = P.out[o].act == 2;
  }
  const float inv_split_a1 = 1.f / (float)(P.add1.split > 0 ? P.add1.split : 1);
  const float inv_split_a2 = 1.f / (float)(P.add2.split > 0 ? P.add2.split : 1);
  const uint32_t t_lane = tmem_base + (uint32_t)sub * acc_stride + ((uint32_t)(quarter * 32) << 16);
  const uint32_t t_stage = (uint32_t)msub * acc_stride;
  const uint32_t full0 = smem_u32(t_full);
  const uint32_t empty_cl0 = map_shared_rank(smem_u32(t_empty), 0);   // the MMA issuer (leader CTA) waits for both CTAs' epilogues

  // claimed-tile ring (filled by the producer thread): every consumer reads every slot once
  const uint32_t sfull0 = smem_u32(s_full);
  const uint32_t sempty_cl0 = map_shared_rank(smem_u32(s_empty), 0);   // the scheduler lives in the leader CTA
  int s_slot = 0;
  uint32_t s_par = 0;
  auto sched_take = [&]() {   // slot s_slot is known to be full
    const int t = *reinterpret_cast<const volatile int*>(s_ring + s_slot);
    __syncwarp();
    if (lane == 0) mbar_arrive_cl(sempty_cl0 + 8u * s_slot);
    if (++s_slot == kSchedSlots) { s_slot = 0; s_par ^= 1u; }
    return t;
  };
  auto sched_next = [&]() {
    mbar_wait_cl(sfull0 + 8u * s_slot, s_par, P.err_flag, 8);
    return sched_take();
  };
  int cur_tile = sched_next();
  if (cur_tile < 0) return;
  pdl_wait();   // residual loads and output stores below touch the predecessors' tensors
  int n_tile = 0, m_tile = 0;

  // state of the tile being processed
  int stage = 0;
  uint32_t phase = 0;
  bool valid = false;
  int n0 = 0, co0 = 0, t_b = 0, t_l = 0;
  const int gpp = P.N_TILE > P.C_out ? (P.C_out >> 4) : 0x7fffffff;   // column groups per phase of a multi-phase tile
  OutPtr op[3], pa1, pa2;
  float* rm_row = nullptr;
  // GRU gate-input layout: a (t, b) row's 16 columns are 64 contiguous bytes, rows are ~400 KB apart.  Stored straight
  // from the accumulator layout (lane = row) a warp store touches 32 lines for 16 bytes each; staged through shared
  // memory, four lanes write one row's 64 bytes and a store instruction covers 8 rows (8 lines).
  const bool rm_x = MULTI && P.rm_xpose;
  float* rm_ptr[4] = {nullptr, nullptr, nullptr, nullptr};   // rows 8 j + lane / 4 of this warp's 32 rows
  float* rm_sc = rm_scratch + (warp - 2) * 32 * kRmPitch;
#pragma unroll
  for (int o = 0; o < 3; ++o) { op[o].p = nullptr; op[o].pstride = 0; }
  pa1.p = pa2.p = nullptr; pa1.pstride = pa2.pstride = 0;

  auto tile_setup = [&]() {
    m_tile = cur_tile; n_tile = 0;
    if (ntn > 1) {
      m_tile = fast_div(cur_tile, ntn, inv_ntn);
      n_tile = cur_tile - m_tile * ntn;
    }
    n0 = n_tile * P.N_TILE;
    int ph = 0;
    co0 = n0;
    if (P.up > 1) {
      ph = fast_div(n0, P.C_out, inv_Cout);
      co0 = n0 - ph * P.C_out;
    }
    const int g = m_tile * tile_rows + row_in_tile;
    const int b = fast_div(g, P.Lh_A, inv_Lh);
    const int m = g - b * P.Lh_A;
    // valid rows of utterance b: the table of a variable-length call (first kMvEntries in shared memory), else M_valid
    valid = (g < P.m_rows) && (m < (P.mv ? (b < kMvEntries ? mv_s[b] : P.mv[b]) : P.M_valid));
    // first column group of this warp: in a multi-phase tile it may lie in a later phase (= a later output row)
    int ph_b = 0, gin = g_begin;
    if (g_begin >= gpp) { ph_b = g_begin / gpp; gin = g_begin - ph_b * gpp; }
    const int l_out = P.up * m + ph + ph_b;
    t_b = b; t_l = l_out;
    const int chunk_base = co0 >> 3;
#pragma unroll
    for (int o = 0; o < 3; ++o) {
      if (o < n_out && valid && (MULTI || o == 0)) op[o] = out_ptr(P.out[o].t, b, l_out, chunk_base + 2 * gin, inv_split[o]);
      else op[o].p = nullptr;
    }
    if (ADD) {
      pa1.p = pa2.p = nullptr;
      if (valid) {
        pa1 = out_ptr(P.add1, b, l_out, chunk_base, inv_split_a1);
        if (a2_on) pa2 = out_ptr(P.add2, b, l_out, chunk_base, inv_split_a2);
      }
    }
    if (rm_x) {
      const int S = 3 * (P.rm_H >> 3) + 4;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int gj = g - lane + 8 * j + (lane >> 2);
        const int bj = fast_div(gj, P.Lh_A, inv_Lh);
        const int mj = gj - bj * P.Lh_A;
        rm_ptr[j] = (gj < P.m_rows && mj < (P.mv ? (bj < kMvEntries ? mv_s[bj] : P.mv[bj]) : P.M_valid)) ? P.rm_out + ((long long)mj * 16 * P.rm_Bp + bj) * S : nullptr;
      }
    } else if (MULTI) {
      rm_row = nullptr;
      if (P.rm_out && valid) {
        if (P.rm_H == 0) rm_row = P.rm_out + (long long)g * P.rm_ld + n0;
        else rm_row = P.rm_out + ((long long)m * 16 * P.rm_Bp + b) * (3 * (P.rm_H >> 3) + 4);   // (t, b) corner of the GRU layout
      }
    }
  };

  struct AddRegs ad0, ad1;
  auto add_issue = [&](int g16, AddRegs& ad) {
    if (!ADD) return;
    if (a1_bf16 && pa1.p) {
      const char* q = pa1.p + (long long)(2 * g16) * pa1.pstride;
      ad.a1[0] = *reinterpret_cast<const uint4*>(q);
      ad.a1[1] = *reinterpret_cast<const uint4*>(q + pa1.pstride);
    }
    if (a2_bf16 && pa2.p) {
      const char* q = pa2.p + (long long)(2 * g16) * pa2.pstride;
      ad.a2[0] = *reinterpret_cast<const uint4*>(q);
      ad.a2[1] = *reinterpret_cast<const uint4*>(q + pa2.pstride);
    }
  };

  auto process = [&](int g16, const uint32_t* r, const AddRegs& ad) {
    if (rm_x) {
      // (no other output, add or FiLM on this path: host-checked)  every lane takes part, rows are masked per store
      const int H = P.rm_H, U = H >> 3, S = 3 * U + 4;
      const int n = n0 + 16 * g16;
      const int dirn = n / (3 * H), rem = n - dirn * 3 * H;
      const int gate = rem / H, jj = rem - gate * H;
      const int rank = jj / U, u = jj - rank * U;
      const long long coloff = (long long)(dirn * 8 + rank) * P.rm_Bp * S + gate * U + u + 4 * (lane & 3);
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<uint4*>(rm_sc + lane * kRmPitch + 4 * i) = make_uint4(r[4 * i], r[4 * i + 1], r[4 * i + 2], r[4 * i + 3]);
      __syncwarp();
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint4 x = *reinterpret_cast<const uint4*>(rm_sc + (8 * j + (lane >> 2)) * kRmPitch + 4 * (lane & 3));
        if (rm_ptr[j]) *reinterpret_cast<uint4*>(rm_ptr[j] + coloff) = x;
      }
      return;
    }
    if (!valid) return;
    if (g16 != g_begin && g16 >= gpp && g16 % gpp == 0) {
      // next phase of a multi-phase tile (transposed conv): one output row further, channel chunks from 0 again
      ++t_l;
#pragma unroll
      for (int o = 0; o < 3; ++o)
        if ((MULTI || o == 0) && op[o].p) op[o] = out_ptr(P.out[o].t, t_b, t_l, 0, inv_split[o]);
    }
    float v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
    if (ADD) {
      if (a1_bf16) add_unpack(ad.a1, v);
      else add_f32(pa1.p + (long long)(2 * g16) * pa1.pstride, pa1.pstride, v);
      if (a2_on) {
        if (a2_bf16) add_unpack(ad.a2, v);
        else add_f32(pa2.p + (long long)(2 * g16) * pa2.pstride, pa2.pstride, v);
      }
    }
    if (do_scale) {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] *= scale;
    }
    if (FILM) {
      const int gc = g16 >= gpp ? g16 - (g16 / gpp) * gpp : g16;   // column group within its phase (multi-phase tiles)
      const float4* ga = reinterpret_cast<const float4*>(film_s + co0 + 16 * gc);
      const float4* be = reinterpret_cast<const float4*>(film_s + P.C_out + co0 + 16 * gc);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 gq = ga[i], bq = be[i];
        v[4 * i] = fmaf(gq.x, v[4 * i], bq.x);
        v[4 * i + 1] = fmaf(gq.y, v[4 * i + 1], bq.y);
        v[4 * i + 2] = fmaf(gq.z, v[4 * i + 2], bq.z);
        v[4 * i + 3] = fmaf(gq.w, v[4 * i + 3], bq.w);
      }
    }
    if (MULTI && rm_row) {
      float4* dst = reinterpret_cast<float4*>(rm_row + 16 * g16);
      if (P.rm_H) {
        // column n = dir*3H + gate*H + rank*U + u  ->  block (dir, rank), row b, offset gate*U + u (16 columns never straddle: U % 16 == 0)
        const int H = P.rm_H, U = H >> 3, S = 3 * U + 4;
        const int n = n0 + 16 * g16;
        const int dirn = n / (3 * H), rem = n - dirn * 3 * H;
        const int gate = rem / H, j = rem - gate * H;
        const int rank = j / U, u = j - rank * U;
        dst = reinterpret_cast<float4*>(rm_row + (long long)(dirn * 8 + rank) * P.rm_Bp * S + gate * U + u);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
#pragma unroll
    for (int o = 0; o < 3; ++o) {
      if ((MULTI || o == 0) && op[o].p) {
        store16_act<LE1>(op[o].p, op[o].pstride, o_f32[o], o_a1[o], o_dbl[o], P.out[o].a2, v);
        op[o].p += 2 * op[o].pstride;
      }
    }
  };

  // Two-buffer software pipeline over this warp's 16-column groups of a tile (group g+1 loads while g is
  // processed).  It crosses the tile boundary when the group count is even: the first group of the next
  // tile is loaded into r0 while the last group (in r1) is processed, provided that accumulator is
  // already complete.
  uint32_t r0[16], r1[16];
  const bool have = g_begin < g_end;
  mbar_wait(full0, 0, P.err_flag, 6);
  tc_fence_after();
  if (warp == 2 && lane == 0) UV_STAMP(5);
  if (have) tc_ld16_issue(t_lane + 16u * g_begin, r0);
  for (;;) {
    tile_setup();
    add_issue(g_begin, ad0);
    const uint32_t taddr = t_lane + (uint32_t)stage * t_stage;
    int nstage = stage + 1;
    uint32_t nphase = phase;
    if (nstage == n_acc) { nstage = 0; nphase ^= 1u; }
    int next_tile = -1;
    bool next_known = false;
    bool prefetched = false;
    auto release = [&]() {   // this warp's part of the accumulator is in registers: the MMA warp may reuse it
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cl(empty_cl0 + 8u * stage);
    };
    if (!have) release();
    for (int g2 = g_begin; g2 < g_end; g2 += 2) {
      tc_ld_wait();
      const bool more1 = g2 + 1 < g_end;
      if (more1) {
        tc_ld16_issue(taddr + 16u * (g2 + 1), r1);
        add_issue(g2 + 1, ad1);
      } else {
        release();
      }
      process(g2, r0, ad0);
      if (more1) {
        tc_ld_wait();
        if (g2 + 2 < g_end) {
          tc_ld16_issue(taddr + 16u * (g2 + 2), r0);
          add_issue(g2 + 2, ad0);
        } else {
          release();
          // never block here: this group is still to be processed
          if (__all_sync(0xffffffffu, mbar_test_wait_cl(sfull0 + 8u * s_slot, s_par))) {
            next_tile = sched_take();
            next_known = true;
          }
          if (next_known && next_tile >= 0 && __all_sync(0xffffffffu, mbar_test_wait(full0 + 8u * nstage, nphase))) {
            tc_fence_after();
            tc_ld16_issue(t_lane + (uint32_t)nstage * t_stage + 16u * g_begin, r0);
            prefetched = true;
          }
        }
        process(g2 + 1, r1, ad1);
      }
    }
    if (!next_known) next_tile = sched_next();
    if (next_tile < 0) {
      if (warp == 2 && lane == 0) UV_STAMP(6);
      break;
    }
    if (!prefetched) {
      mbar_wait(full0 + 8u * nstage, nphase, P.err_flag, 6);
      tc_fence_after();
      if (have) tc_ld16_issue(t_lane + (uint32_t)nstage * t_stage + 16u * g_begin, r0);
    }
    cur_tile = next_tile;
    stage = nstage;
    phase = nphase;
  }
}


// ------------------------------------------------------------------ kernel
// Launched as clusters of two CTAs (one tcgen05 CTA pair on two SMs of a TPC).  A tile is 256 rows x N_TILE columns:
// CTA `rank` holds rows [128 rank, 128 rank + 128) of every A slab and columns [N_TILE/2 rank, ...) of every weight
// image in its own shared memory and its rows of the accumulator in its own tensor memory; the leader (rank 0)
// issues ONE M = 256 tcgen05.mma.cta_group::2 per K step for the pair.  Measured on B200 (tools/mma_bench.cu): a
// 1-SM M128 MMA costs ~97 clk whatever N <= 128 is (66 % of the tensor peak at N = 128, 17-33 % at N = 32-64), the
// pair's M256 MMA costs 64 clk at N = 128 and 128 clk at N = 256 (100 %), 51 clk at N = 64.
// Division of labour in a pair: both CTAs load, transform and drain their own halves; tile claims, MMA issue and all
// "stage consumed" / "accumulator complete" signals (multicast tcgen05.commit) come from the leader; the peer's
// warp 1 relays "my half of this stage has landed" to the leader's barriers.
__global__ void __launch_bounds__(kGemmThreads, 1) gemm_tc_kernel(const __grid_constant__ GemmParams P) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cta_rank_in_cluster();
  const bool leader = rank == 0;
  if (threadIdx.x == 0 && leader) UV_STAMP(0);
  pdl_launch_dependents();   // the next kernel's CTAs take over this SM when this CTA exits and start their own prologue
  // the pair's first two tile claims, in one atomic issued before the prologue so that its round trip (~0.6 us) hides
  // behind the barrier set-up and the tensor-memory allocation (the counter was rewound by this op's previous launch)
  int claim0 = 0;
  if (threadIdx.x == 0 && leader) claim0 = atomicAdd(P.sched, 2);

  // smem carve-up: ones tile | FiLM table | identity images | A ring | B ring (or resident weight set) | barriers | tmem ptr
  uint8_t* ones = smem;
  float* film_s = reinterpret_cast<float*>(smem + kOnesBytes);
  uint8_t* ident = smem + kOnesBytes + kFilmBytes;
  uint8_t* a_ring = ident + P.ident_bytes;
  uint8_t* b_ring = a_ring + (size_t)P.na_stages * P.a_stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(b_ring + (size_t)P.nb_stages * P.b_stage_bytes);
  uint64_t* a_full = bars;
  uint64_t* a_empty = a_full + kMaxAStages;
  uint64_t* b_full = a_empty + kMaxAStages;
  uint64_t* b_empty = b_full + kMaxBStages;
  uint64_t* t_full = b_empty + kMaxBStages;
  uint64_t* t_empty = t_full + kMaxAcc;
  uint64_t* s_full = t_empty + kMaxAcc;
  uint64_t* s_empty = s_full + kSchedSlots;
  uint64_t* a_ready = s_empty + kSchedSlots;
  int* s_ring = reinterpret_cast<int*>(a_ready + kMaxAStages);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(s_ring + kSchedSlots);
  int* mv_s = reinterpret_cast<int*>(tmem_ptr_smem + 4);             // kMvBytes: P.mv[0 .. kMvEntries)
  float* rm_scratch = reinterpret_cast<float*>(mv_s + kMvEntries);   // kRmScratchBytes, only with rm_xpose

  const int n_half = P.N_TILE >> 1;                // weight / identity columns held by this CTA
  const uint32_t acc_stride = accum_stride_cols(P.N_TILE);
  const int n_acc = P.n_acc;
  const int msub = P.msub;                         // 128-row sub-tiles per CTA (2 for narrow tiles)
  const uint32_t tmem_cols = (uint32_t)n_acc * msub * acc_stride;
  const uint32_t xf = P.a_act ? (uint32_t)kXfWarps : 0u;
  {
    // barrier set-up, one barrier group per thread (144 inits by one thread were ~1 us of every launch's prologue).
    // arrival counts: the leader's "full" barriers also take the peer's relay arrive; the leader's "empty" barriers
    // of the accumulators and of the tile ring take the readers of both CTAs
    const uint32_t pr = leader ? 1u : 0u;
    int i = threadIdx.x;
    if (i < kMaxAStages) {
      mbar_init(smem_u32(&a_full[i]), 1 + (P.a_act ? 0u : pr));
      mbar_init(smem_u32(&a_ready[i]), kXfWarps + pr);
      mbar_init(smem_u32(&a_empty[i]), 1);
    } else if ((i -= kMaxAStages) < kMaxBStages) {
      mbar_init(smem_u32(&b_full[i]), 1 + pr);
      mbar_init(smem_u32(&b_empty[i]), 1);
    } else if ((i -= kMaxBStages) < kMaxAcc) {
      mbar_init(smem_u32(&t_full[i]), 1);
      mbar_init(smem_u32(&t_empty[i]), 2 * kEpiWarps);
    } else if ((i -= kMaxAcc) < kSchedSlots) {
      // ring readers: leader = MMA issuer + epilogue + transform warps, peer = producer + relay + epilogue + transform warps
      const uint32_t ring_readers = (1 + kEpiWarps + xf) + (2 + kEpiWarps + xf);
      mbar_init(smem_u32(&s_full[i]), 1);
      mbar_init(smem_u32(&s_empty[i]), ring_readers);
    }
    if (threadIdx.x < kMaxAStages + kMaxBStages + kMaxAcc + kSchedSlots) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // ones tile: K columns 0 and 1 are 1.0 (they meet the hi / lo halves of the bias image), the rest 0
  if (threadIdx.x < 256) {
    uint4 u = make_uint4(0u, 0u, 0u, 0u);
    if (threadIdx.x < kSubM) u.x = 0x3F803F80u;
    reinterpret_cast<uint4*>(ones)[threadIdx.x] = u;
  }
  // FiLM table and per-utterance valid rows: loaded into registers here, stored to shared memory behind the cluster
  // barrier / tensor-memory allocation below, so the global-load latency (~0.7 us) overlaps them instead of preceding
  // them.  (Both are written in front of the call -- plan-time FiLM kernel, host tables --, not by a predecessor
  // kernel: safe before pdl_wait.)
  constexpr int kFilmPerThread = (kFilmBytes / 4 + kGemmThreads - 1) / kGemmThreads;
  float film_r[kFilmPerThread];
  int mv_r = 0;
  const bool mv_on = P.mv != nullptr && (int)threadIdx.x < kMvEntries && (int)threadIdx.x * P.Lh_A < P.m_rows;
  if (P.film) {
#pragma unroll
    for (int j = 0; j < kFilmPerThread; ++j) {
      const int i = threadIdx.x + j * kGemmThreads;
      film_r[j] = i < 2 * P.C_out ? P.film[i] : 0.f;
    }
  }
  if (mv_on) mv_r = P.mv[threadIdx.x];
  if (P.add_kc) {
    // this CTA's half of the identity images [N_TILE / add_kc][add_kc / 8][n_half][8]: image j maps residual channel
    // j*add_kc + k to column j*add_kc + k; local column n stands for column rank*n_half + n
    const int per_img = (P.add_kc >> 3) * n_half;
    const int n_ent = (P.N_TILE / P.add_kc) * per_img;
    for (int i = threadIdx.x; i < n_ent; i += kGemmThreads) {
      const int j = i / per_img, rem = i - j * per_img;
      const int c = rem / n_half, n = rem - c * n_half + (int)rank * n_half;
      const int e = n - j * P.add_kc - 8 * c;   // K element of this 16-byte entry that is 1.0
      uint32_t w[4] = {0u, 0u, 0u, 0u};
      if (e >= 0 && e < 8) w[e >> 1] = (e & 1) ? 0x3F800000u : 0x00003F80u;
      reinterpret_cast<uint4*>(ident)[i] = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  cluster_barrier();   // both CTAs' barriers exist before any remote arrive / multicast commit
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)),
                 "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  if (P.film) {
#pragma unroll
    for (int j = 0; j < kFilmPerThread; ++j) {
      const int i = threadIdx.x + j * kGemmThreads;
      if (i < 2 * P.C_out) film_s[i] = film_r[j];
    }
  }
  if (mv_on) mv_s[threadIdx.x] = mv_r;
  tc_fence_before();
  __syncthreads();
  cluster_barrier();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  if (threadIdx.x == 0 && leader) UV_STAMP(1);

  const int n_shift = P.smax - P.smin + 1;
  const int cta_rows = kSubM * msub;            // rows of a tile held by this CTA (contiguous)
  const int rows_a = cta_rows + P.smax - P.smin;   // halo'd slab rows of this CTA
  const uint32_t add_rows_bytes = (uint32_t)cta_rows * 16u;   // one 8-channel chunk of a residual slab (no halo)
  const int kc_chunks = P.KC >> 3;
  const int total_tiles = P.n_tiles_m * P.n_tiles_n;
  const uint32_t a_slab_chunk_bytes = (uint32_t)rows_a * 16u;
  const uint32_t a_bytes = a_slab_chunk_bytes * kc_chunks;
  const uint32_t b_bytes = (uint32_t)n_half * P.KC * 2u;            // this CTA's half of a weight image
  const uint32_t bias_bytes = P.has_bias ? (uint32_t)n_half * 32u : 0u;
  const int n_img = P.n_src * P.n_groups * n_shift;  // weight images per N tile
  const long long w_half_stride = (long long)n_img * b_bytes + bias_bytes;   // packed per (N tile, half): images, bias image
  const int n_conv = P.n_src * P.n_groups;
  const int n_res = P.add_kc ? P.N_TILE / P.add_kc : 0;

  if (warp == 0) {
    // ============================== producer (both CTAs: own rows of A, own columns of the weights)
    claim0 = __shfl_sync(0xffffffffu, claim0, 0);
    if (elect_one()) {
      if (P.b_resident) {
        const uint32_t fbb = smem_u32(&b_full[0]);
        const uint32_t total = (uint32_t)w_half_stride;
        const char* wsrc = P.w + (long long)rank * w_half_stride;
        mbar_expect_tx(fbb, total);
        for (uint32_t off = 0; off < total; off += 32768u) {
          const uint32_t n = total - off < 32768u ? total - off : 32768u;
          bulk_g2s(smem_u32(b_ring + off), wsrc + off, n, fbb);
        }
      }
      int ia = 0, ib = 0;
      uint32_t pa = 0, pb = 0;
      int ss = 0;
      uint32_t sp = 0;
      // leader: dynamic schedule, two claims in flight, each published to BOTH CTAs' rings (-1 = no more tiles) before
      // its loads are issued.  peer: takes the tile indices from its ring like every other reader.
      int tile = 0, tile_nxt = 0;
      const uint32_t ring_peer = map_shared_rank(smem_u32(s_ring), 1), sfull_peer = map_shared_rank(smem_u32(s_full), 1);
      const uint32_t sempty_cl0 = map_shared_rank(smem_u32(s_empty), 0);
      if (leader) {
        tile = claim0;
        tile_nxt = claim0 + 1;
      }
      pdl_wait();   // weights and the tile counter do not depend on the predecessor; the activations do
      // variable-length calls: a tile whose rows all lie past their utterances' ends has nothing to compute or store
      // (those rows are zero padding and stay zero): the leader drops its claim and takes the next one
      auto tile_has_rows = [&](int claim) {
        const int t = P.reverse ? total_tiles - 1 - claim : claim;
        const int g0 = (t / P.n_tiles_n) * kPairCtas * cta_rows;
        int g1 = g0 + kPairCtas * cta_rows;
        if (g1 > P.m_rows) g1 = P.m_rows;
        for (int b = g0 / P.Lh_A; (long long)b * P.Lh_A < g1; ++b) {
          const int m_lo = g0 - b * P.Lh_A;   // first row of the tile inside utterance b (<= 0: the tile starts before it)
          if ((m_lo < 0 ? 0 : m_lo) < (b < kMvEntries ? mv_s[b] : P.mv[b])) return true;
        }
        return false;
      };
      for (;;) {
        int tile_id;
        if (leader) {
          if (P.mv) {
            while (tile < total_tiles && !tile_has_rows(tile)) {
              tile = tile_nxt;
              if (tile_nxt < total_tiles) tile_nxt = atomicAdd(P.sched, 1);
            }
          }
          const bool live = tile < total_tiles;
          mbar_wait_cl(smem_u32(&s_empty[ss]), sp ^ 1, P.err_flag, 9);
          // reverse: claim k is tile total-1-k.  A kernel's output (263 MB at the finest levels) exceeds the 126 MB L2, so
          // what is still cached when the next kernel starts is the END of the tensor; a consumer that starts from the
          // end reads that part from L2 instead of evicting it on its way there (the builder alternates the direction
          // along each chain of kernels).
          tile_id = live ? (P.reverse ? total_tiles - 1 - tile : tile) : -1;
          s_ring[ss] = tile_id;
          mbar_arrive(smem_u32(&s_full[ss]));
          mbar_expect_tx_cl(sfull_peer + 8u * ss, 4u);
          st_async_u32(ring_peer + 4u * ss, (uint32_t)tile_id, sfull_peer + 8u * ss);
        } else {
          mbar_wait_cl(smem_u32(&s_full[ss]), sp, P.err_flag, 8);
          tile_id = *reinterpret_cast<const volatile int*>(s_ring + ss);
          mbar_arrive_cl(sempty_cl0 + 8u * ss);
        }
        if (++ss == kSchedSlots) { ss = 0; sp ^= 1; }
        if (tile_id < 0) break;
        int tile_nn = tile_nxt;
        if (leader && tile_nxt < total_tiles) tile_nn = atomicAdd(P.sched, 1);
        const int n_tile = tile_id % P.n_tiles_n;
        const int m_tile = tile_id / P.n_tiles_n;
        const long long row0 = (long long)P.a_G + ((long long)m_tile * kPairCtas + rank) * cta_rows + P.smin;
        const char* w_tile = P.w + (long long)(2 * n_tile + (int)rank) * w_half_stride;
        if (!P.b_resident && bias_bytes) {
          mbar_wait(smem_u32(&b_empty[ib]), pb ^ 1, P.err_flag, 2);
          const uint32_t fbb = smem_u32(&b_full[ib]);
          mbar_expect_tx(fbb, bias_bytes);
          bulk_g2s(smem_u32(b_ring + (size_t)ib * P.b_stage_bytes), w_tile + (long long)n_img * b_bytes, bias_bytes, fbb);
          if (++ib == P.nb_stages) { ib = 0; pb ^= 1; }
        }
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
                const long long img = ((long long)s * P.n_groups + g) * n_shift + q;
                bulk_g2s(smem_u32(b_ring + (size_t)ib * P.b_stage_bytes), w_tile + img * b_bytes, b_bytes, fbb);
                if (++ib == P.nb_stages) { ib = 0; pb ^= 1; }
              }
            }
          }
        }
        if (P.add_kc) {
          // residual operand (same row space, phase plane = output phase): one un-halo'd slab per identity image
          const int cpi = P.add_kc >> 3;
          for (int j = 0; j < n_res; ++j) {
            const int col = n_tile * P.N_TILE + j * P.add_kc;   // a tile may span several phases; a slab never does
            const int phase = col / P.C_out;
            const int co0 = col - phase * P.C_out;
            mbar_wait(smem_u32(&a_empty[ia]), pa ^ 1, P.err_flag, 1);
            const uint32_t fb = smem_u32(&a_full[ia]);
            mbar_expect_tx(fb, (uint32_t)cpi * add_rows_bytes);
            const uint32_t dst = smem_u32(a_ring + (size_t)ia * P.a_stage_bytes);
            const long long plane0 = (long long)phase * P.add1.nchunks + (co0 >> 3);
            const long long rowa = (long long)P.add1.G + ((long long)m_tile * kPairCtas + rank) * cta_rows;
            for (int jc = 0; jc < cpi; ++jc) {
              const char* src = P.add1.base + (((plane0 + jc) * P.add1.P + rowa) << 4);
              bulk_g2s(dst + jc * add_rows_bytes, src, add_rows_bytes, fb);
            }
            if (++ia == P.na_stages) { ia = 0; pa ^= 1; }
          }
        }
        if (leader && ss == 1 && sp == 0) UV_STAMP(2);
        tile = tile_nxt;
        tile_nxt = tile_nn;
      }
    }
  } else if (warp == 1) {
    if (leader) {
      // ============================== MMA issuer (leader CTA)
      // One thread; its instruction latency paces small-K layers, so descriptors are split into a constant
      // high word and a low word (address >> 4 | LBO) that advances by plain 32-bit adds.
      if (elect_one()) {
        const uint32_t idesc = make_idesc(kTileM, P.N_TILE);
        const uint32_t lbo_a = a_slab_chunk_bytes, lbo_b = (uint32_t)n_half * 16u;
        const uint32_t desc_hi = (128u >> 4) | (1u << 14);                       // SBO = 128 B, descriptor version 1
        const uint32_t a_lo0 = (smem_u32(a_ring) >> 4) | ((lbo_a >> 4) << 16);
        const uint32_t r_lo0 = (smem_u32(a_ring) >> 4) | ((add_rows_bytes >> 4) << 16);   // residual slabs (no halo)
        const uint32_t b_lo0 = (smem_u32(b_ring) >> 4) | ((lbo_b >> 4) << 16);
        const uint32_t i_lo0 = (smem_u32(ident) >> 4) | ((lbo_b >> 4) << 16);
        const uint32_t a_stage16 = (uint32_t)P.a_stage_bytes >> 4, b_stage16 = (uint32_t)P.b_stage_bytes >> 4;
        const uint32_t a_kstep = (2u * lbo_a) >> 4, r_kstep = (2u * add_rows_bytes) >> 4, b_kstep = (2u * lbo_b) >> 4;
        const uint32_t sub16 = (uint32_t)(kSubM * 16) >> 4;   // second sub-tile: 128 rows further down the slab
        const bool two = msub == 2;
        const uint32_t b_img16 = b_bytes >> 4;
        const uint64_t ones_desc = make_desc(smem_u32(ones), (uint32_t)kSubM * 16u, 128);
        const int ksteps = P.KC >> 4;
        auto dsc = [&](uint32_t lo) { return ((uint64_t)desc_hi << 32) | (uint64_t)lo; };
        if (P.b_resident) {
          mbar_wait_cl(smem_u32(&b_full[0]), 0, P.err_flag, 7);
          tc_fence_after();
        }
        UV_STAMP(3);
        bool first_slab = true;
        const uint32_t a_wait0 = P.a_act ? smem_u32(a_ready) : smem_u32(a_full);
        const uint32_t sempty0 = smem_u32(s_empty);
        int ia = 0, ib = 0, as = 0, ss = 0;
        uint32_t pa = 0, pb = 0, pt = 1, sp = 0;
        for (;;) {
          mbar_wait(smem_u32(&s_full[ss]), sp, P.err_flag, 8);
          const int tile = s_ring[ss];
          mbar_arrive(sempty0 + 8u * ss);
          if (++ss == kSchedSlots) { ss = 0; sp ^= 1; }
          if (tile < 0) break;
          mbar_wait_cl(smem_u32(&t_empty[as]), pt, P.err_flag, 3);   // both CTAs' epilogues have drained this accumulator
          tc_fence_after();
          const uint32_t d0 = tmem_base + (uint32_t)(as * msub) * acc_stride;
          const uint32_t d1 = d0 + acc_stride;
          uint32_t accumulate = 0;
          if (bias_bytes) {
            uint32_t bias_lo;
            if (P.b_resident) {
              bias_lo = b_lo0 + (uint32_t)n_img * b_img16;
            } else {
              mbar_wait_cl(smem_u32(&b_full[ib]), pb, P.err_flag, 5);
              tc_fence_after();
              bias_lo = b_lo0 + (uint32_t)ib * b_stage16;
            }
            tc_mma2_bf16(d0, ones_desc, dsc(bias_lo), idesc, 0);
            if (two) tc_mma2_bf16(d1, ones_desc, dsc(bias_lo), idesc, 0);
            accumulate = 1;
            if (!P.b_resident) {
              tc_commit2(smem_u32(&b_empty[ib]));
              if (++ib == P.nb_stages) { ib = 0; pb ^= 1; }
            }
          }
          uint32_t img_lo = b_lo0;   // resident weights: images are consumed in storage order
          for (int sg = 0; sg < n_conv; ++sg) {
            mbar_wait_cl(a_wait0 + 8u * ia, pa, P.err_flag, 4);   // landed (and, for a raw A tensor, activated) in both CTAs
            tc_fence_after();
            if (first_slab) { UV_STAMP(4); first_slab = false; }
            const uint32_t a_lo = a_lo0 + (uint32_t)ia * a_stage16;
            for (int q = 0; q < n_shift; ++q) {
              uint32_t bl;
              if (P.b_resident) {
                bl = img_lo;
                img_lo += b_img16;
              } else {
                mbar_wait_cl(smem_u32(&b_full[ib]), pb, P.err_flag, 5);
                tc_fence_after();
                bl = b_lo0 + (uint32_t)ib * b_stage16;
              }
              uint32_t al = a_lo + (uint32_t)q;   // tap q = slab shifted by q rows of 16 B
              for (int kk = 0; kk < ksteps; ++kk) {
                tc_mma2_bf16(d0, dsc(al), dsc(bl), idesc, accumulate);
                if (two) tc_mma2_bf16(d1, dsc(al + sub16), dsc(bl), idesc, accumulate);
                accumulate = 1;
                al += a_kstep;
                bl += b_kstep;
              }
              if (!P.b_resident) {
                tc_commit2(smem_u32(&b_empty[ib]));
                if (++ib == P.nb_stages) { ib = 0; pb ^= 1; }
              }
            }
            tc_commit2(smem_u32(&a_empty[ia]));
            if (++ia == P.na_stages) { ia = 0; pa ^= 1; }
          }
          if (P.add_kc) {
            const uint32_t img16 = ((uint32_t)n_half * P.add_kc * 2u) >> 4;
            for (int j = 0; j < n_res; ++j) {
              mbar_wait_cl(a_wait0 + 8u * ia, pa, P.err_flag, 4);
              tc_fence_after();
              uint32_t al = r_lo0 + (uint32_t)ia * a_stage16;
              uint32_t bl = i_lo0 + (uint32_t)j * img16;
              for (int kk = 0; kk < (P.add_kc >> 4); ++kk) {
                tc_mma2_bf16(d0, dsc(al), dsc(bl), idesc, accumulate);
                if (two) tc_mma2_bf16(d1, dsc(al + sub16), dsc(bl), idesc, accumulate);
                accumulate = 1;
                al += r_kstep;
                bl += b_kstep;
              }
              tc_commit2(smem_u32(&a_empty[ia]));
              if (++ia == P.na_stages) { ia = 0; pa ^= 1; }
            }
          }
          tc_commit2(smem_u32(&t_full[as]));
          if (++as == n_acc) { as = 0; pt ^= 1; }
        }
      }
    } else {
      // ============================== relay (peer CTA): walks the stages in the issuer's order and tells the leader's
      // barriers when this CTA's half of a stage has landed (and, for a raw A tensor, has been activated)
      if (elect_one()) {
        const uint32_t a_loc0 = P.a_act ? smem_u32(a_ready) : smem_u32(a_full);
        const uint32_t a_rem0 = map_shared_rank(a_loc0, 0);
        const uint32_t b_loc0 = smem_u32(b_full), b_rem0 = map_shared_rank(b_loc0, 0);
        const uint32_t sempty_cl0 = map_shared_rank(smem_u32(s_empty), 0);
        if (P.b_resident) {
          mbar_wait(b_loc0, 0, P.err_flag, 7);
          mbar_arrive_cl(b_rem0);
        }
        int ia = 0, ib = 0, ss = 0;
        uint32_t pa = 0, pb = 0, sp = 0;
        auto relay_b = [&]() {
          mbar_wait(b_loc0 + 8u * ib, pb, P.err_flag, 5);
          mbar_arrive_cl(b_rem0 + 8u * ib);
          if (++ib == P.nb_stages) { ib = 0; pb ^= 1; }
        };
        for (;;) {
          mbar_wait_cl(smem_u32(&s_full[ss]), sp, P.err_flag, 8);
          const int tile = *reinterpret_cast<const volatile int*>(s_ring + ss);
          mbar_arrive_cl(sempty_cl0 + 8u * ss);
          if (++ss == kSchedSlots) { ss = 0; sp ^= 1; }
          if (tile < 0) break;
          if (!P.b_resident && bias_bytes) relay_b();
          for (int sl = 0; sl < n_conv + n_res; ++sl) {
            mbar_wait(a_loc0 + 8u * ia, pa, P.err_flag, 4);
            mbar_arrive_cl(a_rem0 + 8u * ia);
            if (++ia == P.na_stages) { ia = 0; pa ^= 1; }
            if (!P.b_resident && sl < n_conv)
              for (int q = 0; q < n_shift; ++q) relay_b();
          }
        }
      }
    }
  } else if (warp >= 2 + kEpiWarps) {
    // ============================== A-operand transform (only for a raw A tensor): every CTA activates its own slabs
    if (P.a_act) {
      const int w = warp - (2 + kEpiWarps);
      const float slope = P.a_slope;
      const bool le1 = slope <= 1.f;
      const int nvec = (int)(a_bytes >> 4);
      const uint32_t sempty_cl0 = map_shared_rank(smem_u32(s_empty), 0);
      int ia = 0, ss = 0;
      uint32_t pa = 0, sp = 0;
      for (;;) {
        mbar_wait_cl(smem_u32(&s_full[ss]), sp, P.err_flag, 8);
        const int tile = *reinterpret_cast<const volatile int*>(s_ring + ss);
        __syncwarp();
        if (lane == 0) mbar_arrive_cl(sempty_cl0 + 8u * ss);
        if (++ss == kSchedSlots) { ss = 0; sp ^= 1; }
        if (tile < 0) break;
        for (int sl = 0; sl < n_conv + n_res; ++sl) {   // residual slabs pass through untouched (but are handed on: stage roles rotate)
          mbar_wait(smem_u32(&a_full[ia]), pa, P.err_flag, 4);
          uint4* v = reinterpret_cast<uint4*>(a_ring + (size_t)ia * P.a_stage_bytes);
          if (sl < n_conv) {
            // four 16-byte vectors per lane in flight (the loop is shared-memory latency bound otherwise)
            constexpr int kStep = 32 * kXfWarps;
            for (int i0 = w * 32 + lane; i0 < nvec; i0 += 4 * kStep) {
              uint4 u[4];
#pragma unroll
              for (int k = 0; k < 4; ++k)
                if (i0 + k * kStep < nvec) u[k] = v[i0 + k * kStep];
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                if (i0 + k * kStep < nvec) {
                  u[k].x = prelu_bf16x2(u[k].x, slope, le1);
                  u[k].y = prelu_bf16x2(u[k].y, slope, le1);
                  u[k].z = prelu_bf16x2(u[k].z, slope, le1);
                  u[k].w = prelu_bf16x2(u[k].w, slope, le1);
                  v[i0 + k * kStep] = u[k];
                }
              }
            }
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic stores -> tensor-core reads
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&a_ready[ia]));
          if (++ia == P.na_stages) { ia = 0; pa ^= 1; }
        }
      }
    }
  } else {
    // ============================== epilogue (specialised once, outside the tile loop)
    const bool has_add = P.add1.base != nullptr && P.add_kc == 0;   // add_kc: the residual went through the tensor core
    const bool has_film = P.film != nullptr;
    const bool multi = P.n_out != 1 || P.rm_out != nullptr;
    bool le1 = true;   // every PReLU slope of every output <= 1: two-instruction PReLU
    for (int o = 0; o < 3; ++o)
      if (o < P.n_out && ((P.out[o].act >= 1 && !(P.out[o].a1 <= 1.f)) || (P.out[o].act == 2 && !(P.out[o].a2 <= 1.f)))) le1 = false;
    const int kind = (has_add ? 1 : 0) | (has_film ? 2 : 0) | (multi ? 4 : 0) | (le1 ? 8 : 0);
#define UV_EPI(A, F, M, L) epilogue_role<A, F, M, L>(P, tmem_base, t_full, t_empty, film_s, s_ring, s_full, s_empty, rm_scratch, mv_s, warp, lane, rank)
    switch (kind) {
      case 0: UV_EPI(false, false, false, false); break;
      case 1: UV_EPI(true, false, false, false); break;
      case 2: UV_EPI(false, true, false, false); break;
      case 3: UV_EPI(true, true, false, false); break;
      case 4: UV_EPI(false, false, true, false); break;
      case 5: UV_EPI(true, false, true, false); break;
      case 6: UV_EPI(false, true, true, false); break;
      case 7: UV_EPI(true, true, true, false); break;
      case 8: UV_EPI(false, false, false, true); break;
      case 9: UV_EPI(true, false, false, true); break;
      case 10: UV_EPI(false, true, false, true); break;
      case 11: UV_EPI(true, true, false, true); break;
      case 12: UV_EPI(false, false, true, true); break;
      case 13: UV_EPI(true, false, true, true); break;
      case 14: UV_EPI(false, true, true, true); break;
      default: UV_EPI(true, true, true, true); break;
    }
#undef UV_EPI
  }

  tc_fence_before();
  __syncthreads();
  cluster_barrier();   // neither CTA leaves (or frees tensor memory) while the other can still signal / be signalled
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
  }
  // all claims of this pair are done: the last pair to leave rewinds the counter pair for the next launch
  if (threadIdx.x == 0 && leader) {
    UV_STAMP(7);
    __threadfence();
    if (atomicAdd(P.sched + 1, 1) == (int)(gridDim.x >> 1) - 1) {
      P.sched[0] = 0;
      P.sched[1] = 0;
    }
  }
}

// ------------------------------------------------------------------ host side
inline size_t gemm_smem_bytes(const GemmParams& p) {
  return (size_t)kOnesBytes + kFilmBytes + p.ident_bytes + (size_t)p.na_stages * p.a_stage_bytes + (size_t)p.nb_stages * p.b_stage_bytes +
         (3 * kMaxAStages + 2 * kMaxBStages + 2 * kMaxAcc) * sizeof(uint64_t) + kSchedBytes + 16 + kMvBytes;
}

// 128-row sub-tiles per CTA of the pair: a tile is 256 * gemm_msub(N_TILE) rows (planes are padded to 512 rows).
#ifndef UV_MSUB_MAXN
#define UV_MSUB_MAXN 128
#endif
inline int gemm_msub(int n_tile) { return n_tile <= UV_MSUB_MAXN ? 2 : 1; }
constexpr int kMaxTileRows = 2 * kTileM;

// Fills the pipeline fields of `p` (stage sizes / counts, resident-weight mode) for the 227 KB smem of an SM.
// All sizes are per CTA of the pair: 128 (+ halo) rows of A, N_TILE / 2 columns of the weights.
// Returns false if the layer cannot be scheduled.
inline bool gemm_plan_pipeline(GemmParams& p) {
  if (p.N_TILE % 16 || p.N_TILE > 256 || p.N_TILE < 16) return false;
  const int n_half = p.N_TILE / 2;
  // N tiles up to UV_MSUB_MAXN columns: 256 contiguous rows per CTA (two sub-tiles, two MMAs per K step), so the
  // per-tile costs of every role (claims, barrier round trips, address set-up) are paid half as often; measured with
  // 128 rows per CTA the HBM-bound levels 0 / 1 ran 2x slower than with 256 (test_gemm perf), and on the whole model
  // 64 / 128 as the limit gave 52.4 / 47.0 ms per eager pass (tools/ab_msub.sh).
  p.msub = gemm_msub(p.N_TILE);
  const int rows_a = kSubM * p.msub + p.smax - p.smin;
  const int a_bytes = (p.KC / 8) * rows_a * 16;
  const int b_bytes = n_half * p.KC * 2;
  const int bias_bytes = p.has_bias ? n_half * 32 : 0;
  p.a_stage_bytes = (a_bytes + 127) & ~127;
  // Residual through the tensor core (identity MMA): a bf16 add1 laid out in the GEMM's own row space
  // (phase-split like the output of an up conv), memory-bound layers only (C_out <= 128).
  p.add_kc = 0;
  p.ident_bytes = 0;
  if (p.add1.base && !p.add2.base && p.add1.esize_log2 == 1 && p.add1.Lh == p.Lh_A && p.add1.split == p.up &&
      p.add1.G == p.a_G && p.C_out <= 128 && !p.no_add_mma) {
    for (int kc = 64; kc >= 16; kc -= 16)
      if (p.N_TILE % kc == 0 && p.C_out % kc == 0) { p.add_kc = kc; break; }
    if (p.add_kc) {
      p.ident_bytes = p.N_TILE * n_half * 2;
      const int add_bytes = (p.add_kc / 8) * kAddRowsBytes * p.msub;
      if (add_bytes > p.a_stage_bytes) p.a_stage_bytes = add_bytes;
    }
  }
  if (p.N_TILE > p.C_out) {
    // tile spanning several phases of a transposed conv: whole phases only, adds through the tensor core only
    if (p.up <= 1 || p.N_TILE % p.C_out || p.C_out % 16 || p.rm_out || p.add2.base || (p.add1.base && !p.add_kc)) return false;
  }
  p.n_tiles_m = (p.m_rows + kTileM * p.msub - 1) / (kTileM * p.msub);
  const int stride = (int)accum_stride_cols(p.N_TILE) * p.msub;
  p.n_acc = 512 / stride;
  if (p.n_acc > kMaxAcc) p.n_acc = kMaxAcc;
  if (p.n_acc < 1) return false;
  p.rm_xpose = (p.rm_out && p.rm_H > 0 && p.n_out == 0 && !p.add1.base && !p.add2.base && !p.film && p.scale == 1.f) ? 1 : 0;
  const long long fixed = kOnesBytes + kFilmBytes + p.ident_bytes + (3 * kMaxAStages + 2 * kMaxBStages + 2 * kMaxAcc) * 8 + kSchedBytes + 16 + kMvBytes + 1024 +
                          (p.rm_xpose ? kRmScratchBytes : 0);
  const long long budget = kGemmSmemBudget - fixed;
  const int n_img = p.n_src * p.n_groups * (p.smax - p.smin + 1);
  const long long total_b = (((long long)n_img * b_bytes + bias_bytes) + 127) & ~127LL;
  if (p.n_tiles_n == 1 && total_b + 3LL * p.a_stage_bytes <= budget) {
    p.b_resident = 1;
    p.b_stage_bytes = 128;
    p.nb_stages = (int)(total_b / 128);
    int na = (int)((budget - total_b) / p.a_stage_bytes);
    p.na_stages = na > kMaxAStages ? kMaxAStages : na;
    return true;
  }
  p.b_resident = 0;
  p.b_stage_bytes = (b_bytes + 127) & ~127;
  int na = 2, nb = 3;
  if ((long long)na * p.a_stage_bytes + (long long)nb * p.b_stage_bytes > budget) return false;
  // grow both rings round-robin, B (one image per tap) twice as eagerly as A (one slab per K group)
  for (;;) {
    bool grew = false;
    for (int k = 0; k < 2; ++k)
      if (nb < kMaxBStages && (long long)na * p.a_stage_bytes + (long long)(nb + 1) * p.b_stage_bytes <= budget) { ++nb; grew = true; }
    if (na < 6 && (long long)(na + 1) * p.a_stage_bytes + (long long)nb * p.b_stage_bytes <= budget) { ++na; grew = true; }
    if (!grew) break;
  }
  p.na_stages = na;
  p.nb_stages = nb;
  return true;
}

// Function attributes are per device: call once for every device the kernel is launched on (uv_model_create does).
inline cudaError_t gemm_set_attributes() {
  return cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmemBudget);
}

inline cudaError_t gemm_launch(GemmParams p, int num_sms, cudaStream_t stream) {
  if (!gemm_plan_pipeline(p)) return cudaErrorInvalidConfiguration;
  if (gemm_smem_bytes(p) + (p.rm_xpose ? kRmScratchBytes : 0) > (size_t)kGemmSmemBudget) return cudaErrorInvalidConfiguration;
  if (!p.sched) return cudaErrorInvalidValue;
  const int total = p.n_tiles_m * p.n_tiles_n;
  int pairs = num_sms / kPairCtas;
  if (total < pairs) pairs = total;
  if (pairs < 1) pairs = 1;
  // always the whole shared memory of the SM: one CTA per SM, and never co-resident with a GRU CTA of the other
  // lane (whose tensor-memory allocation would stall this CTA's tcgen05.alloc for the length of the recurrence)
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(pairs * kPairCtas);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = kGemmSmemBudget;
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
  return cudaLaunchKernelEx(&cfg, gemm_tc_kernel, p);
}

}  // namespace uv

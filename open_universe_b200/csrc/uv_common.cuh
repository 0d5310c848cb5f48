// Shared device-side definitions for the open-universe B200 kernels.
//
// Activation layout ("planes"): every activation tensor of C channels lives in
// HBM as  [phase t < split][chunk j < C/8][row < P][8 channels]  (bf16 or f32),
// where a row is  G + b*Lh + (l / split)  for utterance b, sample l (phase
// t = l % split).  Rows G + b*Lh + [Lvalid, Lh) between utterances and the
// guard rows [0,G) / tail are zero and are NEVER written: they are the zero
// padding of every "same" convolution (reference: padding="same" in
// open_universe/networks/universe/blocks.py:293-316), so a conv tap is just a
// row shift of the A operand and needs no bounds handling.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <utility>

namespace uv {

constexpr int kGuard = 8;        // guard rows at the start/end of each plane

struct TensDesc {
  char* base;        // nullptr => unused
  int esize_log2;    // 1 = bf16, 2 = f32
  int nchunks;       // C / 8
  int split;         // phase split r (1 = plain)
  int Lh;            // utterance pitch in rows
  int P;             // plane pitch in rows
  int G;             // guard rows
  int pad_;
};

struct OutDesc {
  TensDesc t;
  int act;           // 0 none, 1 prelu(a1), 2 prelu(prelu(.,a1),a2)
  float a1, a2;
  int pad_;
};

// Per-utterance geometry of a call whose rows have different lengths (uv_set_row_tables): the reference enhances file
// by file (bin/enhance.py:168-192), each with its own padding (universe.py:219-223), statistics, mel frames and GRU
// length; a batch row b of a variable-length call behaves exactly like that file alone.  An op's `rows` table holds
// one record per utterance of its lane; absent = every row has the plan's geometry.
struct RowGeo {
  int T;         // valid samples of the input row
  int pad_left;  // (tot_ds - T % tot_ds) / 2
  int Lp;        // padded length, a multiple of tot_ds (<= the plan's Lp)
  int F;         // frames: Lp / tot_ds
};

struct GemmParams {
  // ---- A operand: bf16 planes tensor, sources = its phases
  const char* a_base;
  int a_nchunks;     // chunks per source (C_in / 8)
  int a_P;           // plane pitch (rows)
  int a_G;
  int n_src;         // number of sources (phases of A)
  int smin, smax;    // row shifts applied to every source
  int KC;            // channels per K group (multiple of 16)
  int n_groups;      // C_in / KC
  // ---- B operand: packed weight images
  //  per N tile: [src][group][shift][KC/8][N_TILE][8] bf16, then (has_bias) the bias image
  //  [2][N_TILE][8] bf16 whose K columns 0 / 1 hold the hi / lo bf16 halves of the fp32 bias
  const char* w;
  int has_bias;
  int N_TILE;        // multiple of 16, <= 256
  int n_tiles_n;
  // ---- row space
  int m_rows;        // B * Lh_A flat rows
  int Lh_A;          // utterance pitch of the row space
  int M_valid;       // valid rows per utterance
  int n_tiles_m;
  // ---- epilogue
  int up;            // l_out = up*m + phase, phase = n_glob / C_out
  int C_out;
  const float* film; // [2*C_out] gamma|beta or null
  TensDesc add1, add2;
  float scale;
  int n_out;
  OutDesc out[3];
  float* rm_out;     // optional f32 row-major output [m_rows][rm_ld], column n_glob
  int rm_ld;
  int rm_H, rm_Bp;   // rm_H > 0: rm_out is the GRU gate-input layout [t][dir][rank 8][b < rm_Bp][3U + 4] (gru.cuh), H = rm_H
  // ---- pipeline (filled by host)
  int na_stages, nb_stages;
  int a_stage_bytes, b_stage_bytes;
  int add_kc;        // > 0: add1 is accumulated by identity MMAs in K slabs of add_kc channels (host decides)
  int ident_bytes;   // smem bytes of the identity images
  int no_add_mma;    // test hook: keep add1 on the register path
  int b_resident;    // 1: all weight images of the (single) N tile stay in smem for the whole kernel
  int n_acc;         // TMEM accumulator stages (1, 2 or 4)
  int msub;          // host: 128-row sub-tiles per CTA of the pair (2 for narrow N tiles: a tile is then 512 rows)
  int rm_xpose;      // host: rm_out in the GRU layout is stored through the shared-memory transpose (gemm_tc.cuh)
  int reverse;       // 1: tiles are taken from the last row tile down (see gemm_tc.cuh: L2 reuse between consecutive kernels)
  int a_act;         // 1: A is stored RAW; PReLU(a_slope) is applied to every slab in shared memory before the MMAs read it
  float a_slope;
  const int* mv;     // optional per-utterance valid rows (replaces M_valid for that utterance: variable-length calls)
  int* err_flag;     // set on watchdog timeout
  int* sched;        // [2] zero before the first launch: tile counter, finished-CTA counter (the kernel rewinds both)
  unsigned long long* clk;   // debug (test_gemm phases): [gridDim.x][8] globaltimer stamps of the CTA's phases, or null
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// Programmatic dependent launch (every kernel of the per-step chain is launched with the "programmatic stream
// serialization" attribute): the next kernel's CTAs may become resident and run their prologue (barrier init,
// tensor-memory allocation, weight loads) as soon as every CTA of this grid has called pdl_launch_dependents();
// they must call pdl_wait() before touching memory that a predecessor writes.  Without the launch attribute both
// are no-ops.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Host side: <<<>>> with the optional programmatic-serialization attribute.
#ifdef __CUDACC__
inline bool pdl_enabled() {
  static const bool on = !(getenv("UV_PDL") && atoi(getenv("UV_PDL")) == 0);
  return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}
#endif

__device__ __forceinline__ float prelu1(float v, float a) { return v >= 0.f ? v : a * v; }

__device__ __forceinline__ float apply_act(float v, int act, float a1, float a2) {
  if (act >= 1) v = prelu1(v, a1);
  if (act == 2) v = prelu1(v, a2);
  return v;
}

struct RowAddr {
  char* p;            // address of (row, chunk 0) in phase plane; nullptr if unused
  long long pstride;  // bytes between chunk planes
};

__device__ __forceinline__ RowAddr row_addr(const TensDesc& t, int b, int l) {
  RowAddr r;
  if (t.base == nullptr) {
    r.p = nullptr;
    r.pstride = 0;
    return r;
  }
  int ph = 0, ll = l;
  if (t.split == 2) {   // the space-to-depth tensors of the finest level: no division
    ll = l >> 1;
    ph = l & 1;
  } else if (t.split > 1) {
    ll = l / t.split;
    ph = l - ll * t.split;
  }
  long long row = (long long)ph * t.nchunks * t.P + t.G + (long long)b * t.Lh + ll;
  r.p = t.base + ((row * 8) << t.esize_log2);
  r.pstride = ((long long)t.P * 8) << t.esize_log2;
  return r;
}

__device__ __forceinline__ void load8(const RowAddr& r, int esize_log2, int chunk, float* v) {
  const char* p = r.p + (long long)chunk * r.pstride;
  if (esize_log2 == 1) {
    uint4 u = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f = __bfloat1622float2(h[i]);
      v[2 * i] = f.x;
      v[2 * i + 1] = f.y;
    }
  } else {
    float4 a = *reinterpret_cast<const float4*>(p);
    float4 c = *reinterpret_cast<const float4*>(p + 16);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
    v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
  }
}

__device__ __forceinline__ void store8(const RowAddr& r, int esize_log2, int chunk, const float* v) {
  char* p = r.p + (long long)chunk * r.pstride;
  if (esize_log2 == 1) {
    uint4 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = u;
  } else {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 16) = make_float4(v[4], v[5], v[6], v[7]);
  }
}

}  // namespace uv

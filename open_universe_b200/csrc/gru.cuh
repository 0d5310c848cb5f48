// Bidirectional GRU recurrence (torch.nn.GRU semantics; reference score.py:84-90,117 and
// condition.py:173-179,212).  The input projection x W_ih^T + b_ih is a tensor-core GEMM
// (gemm_tc.cuh, f32 row-major output `gi`); this kernel runs the serial part:
//   r = sig(gi_r + W_hr h + b_hr), z = sig(gi_z + W_hz h + b_hz),
//   n = tanh(gi_n + r * (W_hn h + b_hn)),  h' = (1 - z) n + z h
// One CTA = one direction x 8 utterances.  Per step the [3H x H] x [H x 8] product runs on
// mma.sync m16n8k16 (bf16 operands, f32 accumulate; h is kept in f32, rounded to bf16 only as
// the MMA operand); W_hh A-fragments are pre-packed in fragment order and streamed from L2.
// The output epilogue fuses the residual add / scale / PReLU that follows the GRU in the
// reference ((out + res)/sqrt2: condition.py:215, blocks.py:380) and writes planes tensors.
#pragma once
#include "uv_common.cuh"

namespace uv {

constexpr int kGruNB = 8;        // utterances per CTA (= MMA n)
constexpr int kGruThreads = 256;
constexpr int kGruMaxH = 512;

struct GruParams {
  const float* gi;          // [rows][6H], rows = b*Lh + t, cols = dir*3H + gate*H + j
  const uint4* whh;         // [2][3H/16][H/16][32] A fragments
  const float* bhh;         // [2][3H]
  int B, T, H, Lh;
  TensDesc add1;
  float scale;
  int n_out;
  OutDesc out[3];
};

__device__ __forceinline__ void mma_bf16_16816(float* c, const uint4& a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b0), "r"(b1));
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

__global__ void __launch_bounds__(kGruThreads, 1) gru_kernel(const __grid_constant__ GruParams P) {
  extern __shared__ __align__(16) uint8_t gsm[];
  const int H = P.H, H3 = 3 * P.H;
  const int hp = H + 8;                                   // bf16 row pitch (bank-conflict free B fragments)
  __nv_bfloat16* hb = reinterpret_cast<__nv_bfloat16*>(gsm);             // [8][H+8]
  float* gh = reinterpret_cast<float*>(gsm + kGruNB * hp * 2);           // [3H][8]
  float* hout = gh + H3 * kGruNB;                                        // [8][H]

  const int dir = blockIdx.y;
  const int b0 = blockIdx.x * kGruNB;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int MT = H3 >> 4, KT = H >> 4;
  const uint4* whh = P.whh + (size_t)dir * MT * KT * 32;
  const float* bhh = P.bhh + dir * H3;

  for (int i = tid; i < kGruNB * hp; i += kGruThreads) hb[i] = __float2bfloat16(0.f);
  float hreg[2][kGruNB];
#pragma unroll
  for (int u = 0; u < 2; ++u)
#pragma unroll
    for (int n = 0; n < kGruNB; ++n) hreg[u][n] = 0.f;
  float br[2], bz[2], bn[2];
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const int j = tid + u * kGruThreads;
    br[u] = j < H ? bhh[j] : 0.f;
    bz[u] = j < H ? bhh[H + j] : 0.f;
    bn[u] = j < H ? bhh[2 * H + j] : 0.f;
  }
  __syncthreads();

  for (int step = 0; step < P.T; ++step) {
    const int t = dir == 0 ? step : P.T - 1 - step;
    // ---- issue the gi loads of this step early (consumed after the MMA phase)
    float gir[2][kGruNB], giz[2][kGruNB], gin[2][kGruNB];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int j = tid + u * kGruThreads;
#pragma unroll
      for (int n = 0; n < kGruNB; ++n) {
        const int b = b0 + n;
        if (j < H && b < P.B) {
          const float* g = P.gi + ((size_t)b * P.Lh + t) * (6 * H) + dir * H3 + j;
          gir[u][n] = g[0]; giz[u][n] = g[H]; gin[u][n] = g[2 * H];
        } else {
          gir[u][n] = giz[u][n] = gin[u][n] = 0.f;
        }
      }
    }
    // ---- gh = W_hh h   (mma.sync; warp w owns m-tiles w, w+8, ...)
    {
      const int nq = lane >> 2, kq = (lane & 3) * 2;
      const __nv_bfloat16* hrow = hb + nq * hp + kq;
      for (int mt = warp; mt < MT; mt += 8) {
        float c[4] = {0.f, 0.f, 0.f, 0.f};
        const uint4* wa = whh + (size_t)mt * KT * 32 + lane;
#pragma unroll 8
        for (int kt = 0; kt < KT; ++kt) {
          const uint4 a = __ldg(wa + kt * 32);
          const uint32_t bb0 = *reinterpret_cast<const uint32_t*>(hrow + kt * 16);
          const uint32_t bb1 = *reinterpret_cast<const uint32_t*>(hrow + kt * 16 + 8);
          mma_bf16_16816(c, a, bb0, bb1);
        }
        float* d = gh + (mt * 16 + nq) * kGruNB + kq;
        *reinterpret_cast<float2*>(d) = make_float2(c[0], c[1]);
        *reinterpret_cast<float2*>(d + 8 * kGruNB) = make_float2(c[2], c[3]);
      }
    }
    __syncthreads();
    // ---- gates (thread = hidden unit j, all 8 utterances)
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int j = tid + u * kGruThreads;
      if (j < H) {
#pragma unroll
        for (int n = 0; n < kGruNB; ++n) {
          const float r = sigmoidf_(gir[u][n] + gh[j * kGruNB + n] + br[u]);
          const float z = sigmoidf_(giz[u][n] + gh[(H + j) * kGruNB + n] + bz[u]);
          const float nn = tanhf(gin[u][n] + r * (gh[(2 * H + j) * kGruNB + n] + bn[u]));
          const float h = (1.f - z) * nn + z * hreg[u][n];
          hreg[u][n] = h;
          hb[n * hp + j] = __float2bfloat16(h);
          hout[n * H + j] = h;
        }
      }
    }
    __syncthreads();
    // ---- output epilogue: (h + add1) * scale -> act -> planes (thread = (utterance, 8-channel chunk))
    const int hc = H >> 3;
    for (int idx = tid; idx < kGruNB * hc; idx += kGruThreads) {
      const int n = idx / hc, ch = idx - n * hc;
      const int b = b0 + n;
      if (b >= P.B) continue;
      float v[8];
      const float4 v0 = *reinterpret_cast<const float4*>(hout + n * H + ch * 8);
      const float4 v1 = *reinterpret_cast<const float4*>(hout + n * H + ch * 8 + 4);
      v[0] = v0.x; v[1] = v0.y; v[2] = v0.z; v[3] = v0.w; v[4] = v1.x; v[5] = v1.y; v[6] = v1.z; v[7] = v1.w;
      const int chunk = dir * hc + ch;
      if (P.add1.base) {
        float a[8];
        load8(row_addr(P.add1, b, t), P.add1.esize_log2, chunk, a);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] += a[i];
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] *= P.scale;
#pragma unroll
      for (int o = 0; o < 3; ++o) {
        if (o < P.n_out) {
          float w[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) w[i] = apply_act(v[i], P.out[o].act, P.out[o].a1, P.out[o].a2);
          store8(row_addr(P.out[o].t, b, t), P.out[o].t.esize_log2, chunk, w);
        }
      }
    }
    // the next step's gate phase rewrites hout only after its own __syncthreads
  }
}

inline size_t gru_smem_bytes(int H) {
  return (size_t)kGruNB * (H + 8) * 2 + (size_t)3 * H * kGruNB * 4 + (size_t)kGruNB * H * 4;
}

}  // namespace uv

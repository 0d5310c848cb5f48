// Kernels of the keyword tail of enhance(): aux_to_wav (use_aux_signal / warm_start) and the signal_median ensemble
// statistic.  Off the hot path (one launch per call), plain CUDA-core code, fp32 throughout.
#pragma once
#include "small_ops.cuh"

namespace uv {

// ---------------------------------------------------------------- AUX   signal_decoupling_layer
// reference universe_gan.py:119-128,147-151 = PReLU_Conv(C -> 1, k3, act = AliasFreeSnake) (blocks.py:209-231):
//   act = bigvgan alias_free_act.py:8-30: torchaudio Resample(1 -> 2) . Snake . Resample(2 -> 1)
//         up   u[2 i + p] = sum_t kup[p][t] x[i + t - wu],        wu = (KU - 1) / 2   (x = 0 outside [0, L))
//         snake s[n] = u[n] + sin^2(a u[n]) / (a + 1e-9),         a = exp(alpha[c])   (snake.py:53-64, alpha_logscale)
//         down d[i] = sum_t kdn[t] s[2 i + t - wd],               wd = (KD - 2) / 2   (s = 0 outside [0, 2 L))
//   or act = PReLU(slope) (mode 1);  then y[l] = bias + sum_c sum_k w[c][k] d_c[l + k - 1]  (d = 0 outside [0, L)).
// Fused with the state initialisation that follows it in enhance() (universe.py:317-331):
//   use_aux_signal: x = y;    warm_start n: x = y + sigma_n z_n
// One CTA = kAuxTile output samples of one utterance, all channels; x tile, s tile and d tile live in shared memory.
constexpr int kAuxTile = 128;
constexpr int kAuxMaxKU = 16, kAuxMaxKD = 32;

struct AuxParams {
  TensDesc a;            // y_hat: bf16 or f32 planes, C channels, plain (split 1)
  int B, Lp, C;
  int mode;              // 0 = snake, 1 = PReLU
  int KU, KD;            // taps per phase of the up kernel / taps of the down kernel
  const float* wts;      // f32: alpha or slope [C] | kup [2][KU] | kdn [KD] | w [C][3] | bias [1]
  float* x;              // [B][Lp]
  float sigma;           // 0: no noise
  int step;              // noise step n (warm_start)
  int b_off, B_total;
  const IoTable* io;
  const RowGeo* rows;    // per-utterance lengths (null: every row is Lp long)
};

__host__ __device__ inline size_t aux_smem_bytes(int C, int KU, int KD) {
  const int nx = kAuxTile + 2 + KD / 2 + KU + 2;   // x samples per channel
  const int ns = 2 * (kAuxTile + 2) + KD;          // s samples per channel
  const int nd = kAuxTile + 2;
  return (size_t)C * (nx + ns + nd) * sizeof(float);
}

__global__ void __launch_bounds__(256) aux_to_wav_kernel(const AuxParams P) {
  extern __shared__ __align__(16) float sm[];
  const int C = P.C, KU = P.KU, KD = P.KD;
  const int wu = (KU - 1) / 2, wd = (KD - 2) / 2;
  const int nd = kAuxTile + 2;                    // d[j], j = l0 - 1 .. l0 + tile
  const int ns = 2 * nd + KD;                     // s[n], n = s_lo .. (covers 2 j + t - wd)
  const int nx = kAuxTile + 2 + KD / 2 + KU + 2;  // x[i], i = x_lo ..
  float* xs = sm;                  // [C][nx]
  float* ss = xs + (size_t)C * nx; // [C][ns]
  float* ds = ss + (size_t)C * ns; // [C][nd]
  const int tiles = (P.Lp + kAuxTile - 1) / kAuxTile;
  const int b = blockIdx.x / tiles, l0 = (blockIdx.x - b * tiles) * kAuxTile;
  const int Lpb = P.rows ? P.rows[b].Lp : P.Lp;   // the utterance's own length: nothing exists past it
  const float* alpha = P.wts;
  const float* kup = alpha + C;
  const float* kdn = kup + 2 * KU;
  const float* w = kdn + KD;
  const float* bias = w + 3 * C;
  const int j_lo = l0 - 1;                       // first d index
  const int s_lo = 2 * j_lo - wd;                // first s index
  const int x_lo = (s_lo >> 1) - wu - 1;         // first x index (floor division: s_lo may be negative)

  if (P.mode == 0) {
    // ---- x tile (zero outside [0, Lp))
    for (int e = threadIdx.x; e < (C >> 3) * nx; e += blockDim.x) {
      const int ch = e / nx, i = e - ch * nx;
      const int l = x_lo + i;
      float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (l >= 0 && l < Lpb) load8(row_addr(P.a, b, l), P.a.esize_log2, ch, v);
#pragma unroll
      for (int k = 0; k < 8; ++k) xs[(size_t)(ch * 8 + k) * nx + i] = v[k];
    }
    __syncthreads();
    // ---- s = snake(up(x)) (zero outside [0, 2 Lp))
    for (int e = threadIdx.x; e < C * ns; e += blockDim.x) {
      const int c = e / ns, k = e - c * ns;
      const int n = s_lo + k;
      float sv = 0.f;
      if (n >= 0 && n < 2 * Lpb) {
        const int i = n >> 1, ph = n & 1;
        const float* xr = xs + (size_t)c * nx + (i - wu - x_lo);
        const float* kr = kup + ph * KU;
        float u = 0.f;
        for (int t = 0; t < KU; ++t) u = fmaf(kr[t], xr[t], u);
        const float a = expf(alpha[c]);
        const float sn = sinf(u * a);
        sv = u + (1.f / (a + 1e-9f)) * sn * sn;
      }
      ss[e] = sv;
    }
    __syncthreads();
    // ---- d = down(s) (zero outside [0, Lp): the conv's "same" padding)
    for (int e = threadIdx.x; e < C * nd; e += blockDim.x) {
      const int c = e / nd, k = e - c * nd;
      const int j = j_lo + k;
      float dv = 0.f;
      if (j >= 0 && j < Lpb) {
        const float* sr = ss + (size_t)c * ns + 2 * k;   // s index 2 j - wd + t = s_lo + 2 k + t
        for (int t = 0; t < KD; ++t) dv = fmaf(kdn[t], sr[t], dv);
      }
      ds[e] = dv;
    }
  } else {
    // ---- d = PReLU(x)
    for (int e = threadIdx.x; e < (C >> 3) * nd; e += blockDim.x) {
      const int ch = e / nd, k = e - ch * nd;
      const int l = j_lo + k;
      float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (l >= 0 && l < Lpb) load8(row_addr(P.a, b, l), P.a.esize_log2, ch, v);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float sl = alpha[ch * 8 + q];
        ds[(size_t)(ch * 8 + q) * nd + k] = v[q] >= 0.f ? v[q] : sl * v[q];
      }
    }
  }
  __syncthreads();
  // ---- C -> 1 conv + state initialisation
  for (int k = threadIdx.x; k < kAuxTile; k += blockDim.x) {
    const int l = l0 + k;
    if (l >= P.Lp) break;
    if (l >= Lpb) { P.x[(size_t)b * P.Lp + l] = 0.f; continue; }
    float acc = bias[0];
    for (int c = 0; c < C; ++c) {
      const float* dr = ds + (size_t)c * nd + k;   // d[l - 1], d[l], d[l + 1]
      acc = fmaf(w[3 * c], dr[0], acc);
      acc = fmaf(w[3 * c + 1], dr[1], acc);
      acc = fmaf(w[3 * c + 2], dr[2], acc);
    }
    if (P.sigma != 0.f) {
      const float z = P.io->noise ? P.io->noise[((long long)P.step * P.B_total + P.b_off + b) * P.Lp + l]
                                  : unit_normal(P.io->seed, P.step, P.b_off + b, l);
      acc = fmaf(P.sigma, z, acc);
    }
    P.x[(size_t)b * P.Lp + l] = acc;
  }
}

// ---------------------------------------------------------------- signal_median (reference utils/stats.py:22-66)
// signal [E][B][S] -> out [B][S]: per (b, sample) sort the E members, find the sorted position holding the member
// whose INDEX is closest to E / 2, count how often every position wins over the samples of b, and return whole the
// member whose index equals the most frequent position.  Two launches: votes (shared-memory histogram per CTA, one
// atomicAdd per bin), then pick + gather.  Ties: first position / lowest index (stable sort), like the CPU oracle.
constexpr int kMedMaxE = 32;

__global__ void __launch_bounds__(256) signal_median_votes_kernel(const float* __restrict__ sig, int E, int B, long long S, int* __restrict__ votes) {
  __shared__ int hist[kMedMaxE];
  const int b = blockIdx.y;
  if (threadIdx.x < kMedMaxE) hist[threadIdx.x] = 0;
  __syncthreads();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < S; i += (long long)gridDim.x * blockDim.x) {
    float v[kMedMaxE];
    int id[kMedMaxE];
#pragma unroll 1
    for (int e = 0; e < E; ++e) {   // stable insertion sort of (value, member)
      const float x = sig[((long long)e * B + b) * S + i];
      int k = e;
      while (k > 0 && v[k - 1] > x) { v[k] = v[k - 1]; id[k] = id[k - 1]; --k; }
      v[k] = x; id[k] = e;
    }
    int best = 0;
    float bd = fabsf((float)id[0] - 0.5f * (float)E);
    for (int k = 1; k < E; ++k) {
      const float d = fabsf((float)id[k] - 0.5f * (float)E);
      if (d < bd) { bd = d; best = k; }
    }
    atomicAdd(&hist[best], 1);
  }
  __syncthreads();
  if (threadIdx.x < E && hist[threadIdx.x]) atomicAdd(&votes[b * kMedMaxE + threadIdx.x], hist[threadIdx.x]);
}

__global__ void __launch_bounds__(256) signal_median_gather_kernel(const float* __restrict__ sig, int E, int B, long long S,
                                                                   const int* __restrict__ votes, float* __restrict__ out) {
  const int b = blockIdx.y;
  int sel = 0, best = votes[b * kMedMaxE];
  for (int e = 1; e < E; ++e) {
    const int c = votes[b * kMedMaxE + e];
    if (c > best) { best = c; sel = e; }
  }
  const float* src = sig + ((long long)sel * B + b) * S;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < S; i += (long long)gridDim.x * blockDim.x) out[(long long)b * S + i] = src[i];
}

}  // namespace uv

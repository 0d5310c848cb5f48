// Memory-bound kernels of the enhance() path: normalise/pad, mel front-end,
// 1->C input conv, C->1 output conv fused with the sampler update, post-processing
// and the FiLM table.  All coalesced + vectorised; reductions by warp shuffle.
#pragma once
#include "uv_common.cuh"

namespace uv {

struct IoTable {
  const float* mix;    // [B][T]
  const float* noise;  // [n_steps][B][Lp]; null = the sampler draws are generated in the kernels (unit_normal below)
  float* out;          // [B][T]
  float* dbg;          // [n_steps][B][Lp] or null
  unsigned long long seed;   // key of the in-kernel draws (uv_set_noise_seed)
  unsigned long long pad_;
};

// ---------------------------------------------------------------- in-kernel sampler noise ("throughput mode")
// z[step][utterance][sample] ~ N(0, 1), a pure function of (seed, step, utterance, sample): Philox4x32-10 keyed by the
// seed on the counter (sample / 4, utterance, step, tag), four normals per block by Box-Muller.  Independent of the
// launch geometry, of the lane split and of which rank owns the utterance.  Stands for the reference's
// torch.randn draws (universe.py:39-41) when the caller does not need torch's own stream; oracle/philox.py is the
// CPU restatement the GPU test checks it against.
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0;
    c1 = lo1;
    c2 = hi0 ^ c3 ^ k1;
    c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
constexpr uint32_t kNoiseTag = 0x554E4956u;   // "UNIV"
__device__ __forceinline__ float unit_normal(unsigned long long seed, int step, int utt, int l) {
  uint32_t r[4];
  philox4x32_10((uint32_t)l >> 2, (uint32_t)utt, (uint32_t)step, kNoiseTag, (uint32_t)seed, (uint32_t)(seed >> 32), r);
  const int pair = (l >> 1) & 1;
  // 24-bit uniforms in (0, 1): (r >> 8) + 0.5 scaled by 2^-24
  const float u1 = ((float)(r[2 * pair] >> 8) + 0.5f) * 5.9604644775390625e-8f;
  const float u2 = ((float)(r[2 * pair + 1] >> 8) + 0.5f) * 5.9604644775390625e-8f;
  const float rad = sqrtf(-2.f * logf(u1));
  float sn, cs;
  sincospif(2.f * u2, &sn, &cs);
  return rad * ((l & 1) ? sn : cs);
}

constexpr int kEdgeMaxC = 128;   // channels of the 1 <-> C edge convs (32 / 48 in the reference configs)

struct StepCoef {
  float a, b, c, pad_;
};

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum of up to 3 doubles (blockDim.x multiple of 32, <= 1024). Result valid in all threads.
__device__ __forceinline__ void block_sum3(double& a, double& b, double& c, double* sh /* [3*32] */) {
  a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) { sh[w] = a; sh[32 + w] = b; sh[64 + w] = c; }
  __syncthreads();
  a = (l < nw) ? sh[l] : 0.0; b = (l < nw) ? sh[32 + l] : 0.0; c = (l < nw) ? sh[64 + l] : 0.0;
  a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
}

// ---------------------------------------------------------------- PREP
// Universe.pad (universe.py:219-223) + utils.normalize_batch norm=2, zero_mean
// (utils/norm.py:47-87): statistics over the PADDED signal, unbiased std, clamp 1e-5.
// Also mix_rms over the unpadded input (universe.py:259).  One CTA per utterance.
// b_off = first utterance of this lane in the caller's batch (the io tensors cover all lanes).
__global__ void __launch_bounds__(1024) prep_kernel(const IoTable* io, int b_off, int T, int Lp, int pad_left, float level,
                                                    float* sig, float* mix_rms, const RowGeo* __restrict__ rows) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ double sh[96];
  const int b = blockIdx.x;
  const float* mix = io->mix + (size_t)(b_off + b) * T;   // (T = row pitch of the caller's tensor)
  int Tb = T, Lpb = Lp;
  if (rows) { const RowGeo r = rows[b]; Tb = r.T; pad_left = r.pad_left; Lpb = r.Lp; }
  double s = 0, ss = 0, dummy = 0;
  for (int i = threadIdx.x; i < Tb; i += blockDim.x) {
    const double v = mix[i];
    s += v; ss += v * v;
  }
  block_sum3(s, ss, dummy, sh);
  const double mean = s / Lpb;
  double var = (ss - (double)Lpb * mean * mean) / (double)(Lpb - 1);
  if (var < 0) var = 0;
  const float meanf = (float)mean;
  const float gain = level / fmaxf((float)sqrt(var), 1e-5f);
  if (threadIdx.x == 0) mix_rms[b] = (float)sqrt(ss / Tb);
  float* o = sig + (size_t)b * Lp;
  for (int i = threadIdx.x; i < Lp; i += blockDim.x) {
    const int j = i - pad_left;
    const float v = (j >= 0 && j < Tb) ? mix[j] : 0.f;
    o[i] = i < Lpb ? (v - meanf) * gain : 0.f;   // nothing exists past the row's own padded length
  }
}

// ---------------------------------------------------------------- MEL
// MelAdapter.compute_mel_spec (condition.py:92-108): explicit (n_fft-hop)/2 padding,
// Hann-windowed power spectrum (torchaudio MelSpectrogram, center=False, power=2),
// mel filterbank.  The windowed real-DFT basis [n_fft][2*nbp] and fb come from the
// checkpoint buffers.  One CTA = 32 frames of one utterance, one thread per frequency bin (block = bins rounded
// up to a warp); fp32 throughout.  The 1.7 MB basis is streamed from L2 once per CTA, so the frame tile
// sets the L2 traffic (8 frames per CTA were L2-bandwidth bound at 3.5 TB/s), and one bin per thread avoids a
// second, mostly idle pass over the bins (321 bins on 256 threads).
constexpr int kMelFT = 32;
constexpr int kMelFbRows = 16;            // filterbank rows (bins) per staged chunk
constexpr int kMelPwPitch = kMelFT + 4;   // power spectrum rows [bin][frame], padded: conflict-free 16-byte accesses
// Folded DFT: the window is symmetric about n_fft / 2 (Hann; checked by the host when the basis is packed), so
//   re[k] = sum_{n=0}^{N/2} (x[n] + x[N-n]) C[n][k],   im[k] = sum_{n=1}^{N/2-1} (x[n] - x[N-n]) S[n][k]
// (n = 0 and n = N/2 unpaired): the folded frames a = x[n] + x[N-n], b = x[n] - x[N-n] are built once per CTA in shared
// memory and the multiply-adds per bin are halved.  Only basis rows 0 .. N/2 are read.
__host__ __device__ inline int mel_fold_pitch(int n_fft) { return n_fft / 2 + 4; }   // floats per folded frame (n = 0 .. N/2, padded)
__host__ __device__ inline size_t mel_smem_bytes(int n_fft, int hop) {
  const int span = (kMelFT - 1) * hop + n_fft;
  return ((size_t)((span + 3) & ~3) + 2 * (size_t)kMelFT * mel_fold_pitch(n_fft)) * sizeof(float);   // (the power rows overlay the folded frames)
}
__global__ void __launch_bounds__(512) mel_kernel(const float* __restrict__ sig, int Lp, int F, int n_fft, int hop,
                                                  int nb, int nbp, int n_mels, int pad_left,
                                                  const float* __restrict__ basis, const float* __restrict__ fb,
                                                  float* __restrict__ melraw, float* __restrict__ energy, int nft,
                                                  const RowGeo* __restrict__ rows) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ float smem_f[];
  const int span = (kMelFT - 1) * hop + n_fft;
  const int nh = n_fft >> 1, fp = mel_fold_pitch(n_fft);
  float* xs = smem_f;                           // [span] (later: filterbank staging)
  float* fa = smem_f + ((span + 3) & ~3);       // [kMelFT][fp]  x[n] + x[N - n]
  float* fd = fa + (size_t)kMelFT * fp;         // [kMelFT][fp]  x[n] - x[N - n]
  float* pw = fa;                               // [nb][kMelPwPitch]: overlays the folded frames once the DFT is done
  __shared__ double sh[96];
  const int b = blockIdx.y, f0 = blockIdx.x * kMelFT;
  const int Fb = rows ? rows[b].F : F;
  const float* x = sig + (size_t)b * Lp;
  for (int i = threadIdx.x; i < span; i += blockDim.x) {
    const int j = f0 * hop + i - pad_left;
    xs[i] = (j >= 0 && j < Lp) ? x[j] : 0.f;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < kMelFT * fp; e += blockDim.x) {
    const int f = e / fp, n = e - f * fp;
    const float* xf = xs + f * hop;
    float va = 0.f, vd = 0.f;
    if (n == 0 || n == nh) {
      va = xf[n];
    } else if (n < nh) {
      va = xf[n] + xf[n_fft - n];
      vd = xf[n] - xf[n_fft - n];
    }
    fa[e] = va;
    fd[e] = vd;
  }
  __syncthreads();
  // Thread = (half h of the tile's frames, two adjacent frequency bins): blockDim = 2 x (threads covering ceil(nb / 2)
  // bin pairs).  A 16-byte load of a folded frame is a broadcast that costs the shared-memory pipe as much as a full
  // 512-byte access, so what counts is FMAs per load: 16 here (2 bins x 4 samples x re / im per pair of loads) against
  // 8 with one bin x 32 frames per thread, in the same 64 accumulator registers.
  constexpr int kHalf = kMelFT / 2;
  const int nth = blockDim.x >> 1;
  const int h = threadIdx.x >= nth ? 1 : 0;
  const int k0 = 2 * (threadIdx.x - h * nth);
  const bool bin_on = k0 < nb;             // (bin k0 + 1 may be the padding column nb: its basis entries are zero)
  float re0[kHalf], im0[kHalf], re1[kHalf], im1[kHalf];
#pragma unroll
  for (int f = 0; f < kHalf; ++f) re0[f] = im0[f] = re1[f] = im1[f] = 0.f;
  if (bin_on) {
    const size_t bst = (size_t)2 * nbp;
    const float* bc = basis + k0;          // cos | sin halves of a basis row; two adjacent bins = one 8-byte load
    const float* bs = basis + nbp + k0;
    const float* fah = fa + (size_t)h * kHalf * fp;
    const float* fdh = fd + (size_t)h * kHalf * fp;
    float2 c[4], s4[4], cn[4], sn[4];      // basis of this / the next step (nothing else hides the L2 latency)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      cn[j] = *reinterpret_cast<const float2*>(bc + (size_t)j * bst);
      sn[j] = *reinterpret_cast<const float2*>(bs + (size_t)j * bst);
    }
    for (int n = 0; n < nh; n += 4) {
#pragma unroll
      for (int j = 0; j < 4; ++j) { c[j] = cn[j]; s4[j] = sn[j]; }
      if (n + 4 < nh) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          cn[j] = *reinterpret_cast<const float2*>(bc + (size_t)(n + 4 + j) * bst);
          sn[j] = *reinterpret_cast<const float2*>(bs + (size_t)(n + 4 + j) * bst);
        }
      }
#pragma unroll
      for (int f = 0; f < kHalf; ++f) {
        const float4 va = *reinterpret_cast<const float4*>(fah + f * fp + n);
        const float4 vd = *reinterpret_cast<const float4*>(fdh + f * fp + n);
        re0[f] = fmaf(va.x, c[0].x, re0[f]); re1[f] = fmaf(va.x, c[0].y, re1[f]);
        im0[f] = fmaf(vd.x, s4[0].x, im0[f]); im1[f] = fmaf(vd.x, s4[0].y, im1[f]);
        re0[f] = fmaf(va.y, c[1].x, re0[f]); re1[f] = fmaf(va.y, c[1].y, re1[f]);
        im0[f] = fmaf(vd.y, s4[1].x, im0[f]); im1[f] = fmaf(vd.y, s4[1].y, im1[f]);
        re0[f] = fmaf(va.z, c[2].x, re0[f]); re1[f] = fmaf(va.z, c[2].y, re1[f]);
        im0[f] = fmaf(vd.z, s4[2].x, im0[f]); im1[f] = fmaf(vd.z, s4[2].y, im1[f]);
        re0[f] = fmaf(va.w, c[3].x, re0[f]); re1[f] = fmaf(va.w, c[3].y, re1[f]);
        im0[f] = fmaf(vd.w, s4[3].x, im0[f]); im1[f] = fmaf(vd.w, s4[3].y, im1[f]);
      }
    }
    const float2 cm = *reinterpret_cast<const float2*>(bc + (size_t)nh * bst);   // the unpaired middle sample (its sine term is zero)
#pragma unroll
    for (int f = 0; f < kHalf; ++f) {
      const float xm = fah[f * fp + nh];
      re0[f] = fmaf(xm, cm.x, re0[f]);
      re1[f] = fmaf(xm, cm.y, re1[f]);
    }
  }
  __syncthreads();   // every thread has read the folded frames: the power rows may overwrite them
  if (bin_on) {
#pragma unroll
    for (int f = 0; f < kHalf; f += 4) {
      *reinterpret_cast<float4*>(pw + (size_t)k0 * kMelPwPitch + h * kHalf + f) =
          make_float4(re0[f] * re0[f] + im0[f] * im0[f], re0[f + 1] * re0[f + 1] + im0[f + 1] * im0[f + 1],
                      re0[f + 2] * re0[f + 2] + im0[f + 2] * im0[f + 2], re0[f + 3] * re0[f + 3] + im0[f + 3] * im0[f + 3]);
      if (k0 + 1 < nb)
        *reinterpret_cast<float4*>(pw + (size_t)(k0 + 1) * kMelPwPitch + h * kHalf + f) =
            make_float4(re1[f] * re1[f] + im1[f] * im1[f], re1[f + 1] * re1[f + 1] + im1[f + 1] * im1[f + 1],
                        re1[f + 2] * re1[f + 2] + im1[f + 2] * im1[f + 2], re1[f + 3] * re1[f + 3] + im1[f + 3] * im1[f + 3]);
    }
  }
  __syncthreads();
  // mel filterbank: thread = (mel band m, group g of the tile's frames); per bin one coalesced fb load and
  // broadcast 16-byte loads of the power row feed 8-32 FMAs (a thread per (frame, band) item with a scalar load
  // pair per FMA was latency bound and took as long as the DFT)
  double e = 0, d1 = 0, d2 = 0;
  int ng = blockDim.x / n_mels;
  ng = ng >= 4 ? 4 : (ng >= 2 ? 2 : 1);
  const int fpt = kMelFT / ng;   // frames per thread: 8, 16 or 32
  // fb rows are staged through the (now dead) signal buffer in chunks of kMelFbRows bins, double buffered
  const int tpg = blockDim.x / ng;                 // threads per frame group
  const int g = threadIdx.x / tpg, m = threadIdx.x % tpg;
  const bool active = g < ng && m < n_mels;        // (n_mels <= tpg is checked by the host)
  float acc[kMelFT];
#pragma unroll
  for (int f = 0; f < kMelFT; ++f) acc[f] = 0.f;
  const int chunk = kMelFbRows * n_mels;
  float* fbs = xs;                                 // [2][kMelFbRows][n_mels]
  const int n_chunks = (nb + kMelFbRows - 1) / kMelFbRows;
  auto load_chunk = [&](int ci, float* dst) {
    const int k0 = ci * kMelFbRows;
    const int cnt = (nb - k0 < kMelFbRows ? nb - k0 : kMelFbRows) * n_mels;
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) dst[i] = fb[(size_t)k0 * n_mels + i];
  };
  load_chunk(0, fbs);
  __syncthreads();
  for (int ci = 0; ci < n_chunks; ++ci) {
    float* cur = fbs + (ci & 1) * chunk;
    if (ci + 1 < n_chunks) load_chunk(ci + 1, fbs + ((ci + 1) & 1) * chunk);
    if (active) {
      const int k0 = ci * kMelFbRows;
      const int kn = nb - k0 < kMelFbRows ? nb - k0 : kMelFbRows;
      const float* prow = pw + (size_t)k0 * kMelPwPitch + g * fpt;
      for (int kk = 0; kk < kn; ++kk) {
        const float w = cur[kk * n_mels + m];
#pragma unroll
        for (int f = 0; f < kMelFT; f += 4) {
          if (f < fpt) {
            const float4 v = *reinterpret_cast<const float4*>(prow + (size_t)kk * kMelPwPitch + f);
            acc[f] = fmaf(v.x, w, acc[f]); acc[f + 1] = fmaf(v.y, w, acc[f + 1]);
            acc[f + 2] = fmaf(v.z, w, acc[f + 2]); acc[f + 3] = fmaf(v.w, w, acc[f + 3]);
          }
        }
      }
    }
    __syncthreads();
  }
  if (active) {
#pragma unroll
    for (int f = 0; f < kMelFT; ++f) {
      const int fr = f0 + g * fpt + f;
      if (f < fpt && fr < F) {
        const float a = fr < Fb ? acc[f] : 0.f;   // frames past the row's own length do not exist
        melraw[((size_t)b * F + fr) * n_mels + m] = a;
        e += (double)a * a;
      }
    }
  }
  block_sum3(e, d1, d2, sh);
  if (threadIdx.x == 0) energy[(size_t)b * nft + blockIdx.x] = (float)e;
}

// ---------------------------------------------------------------- mel front end around the tensor-core DFT
// The framed DFT is one multi-tap GEMM (weights.w4_dft, run by gemm_hp_tc_kernel: split-bf16 products, fp32
// accumulation) on the padded signal stored as f32 planes in rows of hop / 2 samples, phase-split by 2; its output is
// f32 row-major [b Lh + f][ld] with the cos bins at column 0 and the sin bins at column nbq.
// SIGP: sig [B][Lp] -> that planes tensor (thread = one 8-sample piece; reads are contiguous, writes 32-byte pieces).
__global__ void __launch_bounds__(256) sig_planes_kernel(const float* __restrict__ sig, int Lp, int F, int half, OutDesc out,
                                                         const RowGeo* __restrict__ rows) {
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.y, nch = half >> 3;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;   // (half-hop h, chunk), chunk fastest = signal order
  if (idx >= 2 * F * nch) return;
  const int h = idx / nch, ch = idx - h * nch;
  if (rows && h >= 2 * rows[b].F) return;   // past the row's own frames: the planes stay zero (as the signal is)
  const float* src = sig + (size_t)b * Lp + (size_t)idx * 8;
  const float4 a = *reinterpret_cast<const float4*>(src), c = *reinterpret_cast<const float4*>(src + 4);
  const float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
  store8(row_addr(out.t, b, h), out.t.esize_log2, ch, v);
}

// MEL_POW: power spectrum re^2 + im^2 of 32 frames -> shared memory [bin][frame] -> mel filterbank -> melraw [B][F][n_mels]
// and the partial energy sums mel_fin_kernel reduces.  Thread = (mel band m, group g of the tile's frames); a band only
// walks the bins [lo, hi) its filter is non-zero on (rng, from the host: triangular filters see a few dozen of the
// n_fft / 2 + 1 bins), reading its weights straight from the (cached) filterbank.
__host__ __device__ inline size_t mel_pow_smem_bytes(int nb) { return (size_t)nb * kMelPwPitch * sizeof(float); }
__global__ void __launch_bounds__(512) mel_pow_kernel(const float* __restrict__ D, int ld, int Lh, int F, int nb, int nbq,
                                                      int n_mels, const float* __restrict__ fb, const int2* __restrict__ rng,
                                                      float* __restrict__ melraw, float* __restrict__ energy, int nft,
                                                      const RowGeo* __restrict__ rows) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ float smem_f[];
  float* pw = smem_f;                                   // [nb][kMelPwPitch]
  __shared__ double sh[96];
  const int b = blockIdx.y, f0 = blockIdx.x * kMelFT;
  const int Fb = rows ? rows[b].F : F;
  for (int e = threadIdx.x; e < kMelFT * nb; e += blockDim.x) {
    const int f = e / nb, k = e - f * nb;
    float p = 0.f;
    if (f0 + f < Fb) {   // (rows of frames past the utterance's length are never written by the GEMM)
      const float* row = D + ((size_t)b * Lh + f0 + f) * ld;
      const float re = row[k], im = row[nbq + k];
      p = re * re + im * im;
    }
    pw[(size_t)k * kMelPwPitch + f] = p;
  }
  __syncthreads();
  double e = 0, d1 = 0, d2 = 0;
  int ng = blockDim.x / n_mels;
  ng = ng >= 4 ? 4 : (ng >= 2 ? 2 : 1);
  const int fpt = kMelFT / ng;   // frames per thread: 8, 16 or 32
  const int tpg = blockDim.x / ng;                 // threads per frame group
  const int g = threadIdx.x / tpg, m = threadIdx.x % tpg;
  const bool active = g < ng && m < n_mels;        // (n_mels <= tpg is checked by the host)
  if (active) {
    float acc[kMelFT];
#pragma unroll
    for (int f = 0; f < kMelFT; ++f) acc[f] = 0.f;
    const int2 r = rng[m];
    const float* prow = pw + g * fpt;
    for (int k = r.x; k < r.y; ++k) {
      const float w = fb[(size_t)k * n_mels + m];
#pragma unroll
      for (int f = 0; f < kMelFT; f += 4) {
        if (f < fpt) {
          const float4 v = *reinterpret_cast<const float4*>(prow + (size_t)k * kMelPwPitch + f);
          acc[f] = fmaf(v.x, w, acc[f]); acc[f + 1] = fmaf(v.y, w, acc[f + 1]);
          acc[f + 2] = fmaf(v.z, w, acc[f + 2]); acc[f + 3] = fmaf(v.w, w, acc[f + 3]);
        }
      }
    }
#pragma unroll
    for (int f = 0; f < kMelFT; ++f) {
      const int fr = f0 + g * fpt + f;
      if (f < fpt && fr < F) {
        const float a = fr < Fb ? acc[f] : 0.f;   // frames past the row's own length do not exist
        melraw[((size_t)b * F + fr) * n_mels + m] = a;
        e += (double)a * a;
      }
    }
  }
  block_sum3(e, d1, d2, sh);
  if (threadIdx.x == 0) energy[(size_t)b * nft + blockIdx.x] = (float)e;
}

// Global mel normalisation (condition.py:105-106): x / clamp(sqrt(mean_t sum_m x^2), 1e-5) -> bf16 planes.
__global__ void __launch_bounds__(256) mel_fin_kernel(const float* __restrict__ melraw, const float* __restrict__ energy,
                                                      int F, int n_mels, int nft, OutDesc out, const RowGeo* __restrict__ rows) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float s_inv;
  const int b = blockIdx.y;
  const int Fb = rows ? rows[b].F : F;
  if (threadIdx.x == 0) {
    float e = 0.f;
    for (int i = 0; i < nft; ++i) e += energy[(size_t)b * nft + i];
    s_inv = 1.f / fmaxf(sqrtf(e / (float)Fb), 1e-5f);
  }
  __syncthreads();
  const float inv = s_inv;
  const int nch = n_mels >> 3;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;  // (frame, chunk), frame fastest
  if (idx >= F * nch) return;
  const int ch = idx / F, f = idx - ch * F;
  if (f >= Fb) return;   // (rows past the utterance's length stay zero)
  const float* src = melraw + ((size_t)b * F + f) * n_mels + ch * 8;
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = apply_act(src[i] * inv, out.act, out.a1, out.a2);
  RowAddr r = row_addr(out.t, b, f);
  store8(r, out.t.esize_log2, ch, v);
}

// ---------------------------------------------------------------- CONV_IN
// 1 -> C, k=3, "same" input convolution on an f32 signal (score.py:285 on w_in*x,
// condition.py:360), writing the raw tensor and / or PReLU-activated copies.
// Both edge convs were instruction bound (ncu: sm 77 % / 60 %, dram 33 % / 31 %: ~550 / ~640 instructions per sample):
// the grid is 2-D (block row = utterance: no 64-bit division per thread), activations that are not asked for are not
// computed (a uniform branch instead of per-element selects), plane pointers advance by additions.
__device__ __forceinline__ void store8_act(char* p, int esize_log2, int act, float a1, float a2, const float* v) {
  float u[8];
  if (act == 0) {
#pragma unroll
    for (int i = 0; i < 8; ++i) u[i] = v[i];
  } else if (act == 1) {
#pragma unroll
    for (int i = 0; i < 8; ++i) u[i] = prelu1(v[i], a1);
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) u[i] = prelu1(prelu1(v[i], a1), a2);
  }
  if (esize_log2 == 1) {
    uint4 q;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(u[2 * i], u[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = q;
  } else {
    *reinterpret_cast<float4*>(p) = make_float4(u[0], u[1], u[2], u[3]);
    *reinterpret_cast<float4*>(p + 16) = make_float4(u[4], u[5], u[6], u[7]);
  }
}

__global__ void __launch_bounds__(256) conv_in_kernel(const float* __restrict__ x, int B, int Lp, int C, float in_scale,
                                                      const float* __restrict__ w /*[3][C]*/, const float* __restrict__ bias,
                                                      int n_out, OutDesc o0, OutDesc o1, int reverse, const RowGeo* __restrict__ rows) {
  pdl_launch_dependents();
  // the three taps and the bias of every channel, staged once per block: one 16-byte broadcast load per channel
  // instead of four global loads per channel and thread
  __shared__ float4 ws[kEdgeMaxC];
  for (int c = threadIdx.x; c < C; c += blockDim.x) ws[c] = make_float4(w[c], w[C + c], w[2 * C + c], bias[c]);
  __syncthreads();
  pdl_wait();
  // grid (blocks per utterance, utterances); reverse: blocks walk the signal from its end (the consumer of these
  // tensors starts where this kernel finished)
  const int bx = reverse ? (int)(gridDim.x - 1 - blockIdx.x) : (int)blockIdx.x;
  const int b = reverse ? (int)(gridDim.y - 1 - blockIdx.y) : (int)blockIdx.y;
  const int l = bx * (int)blockDim.x + (int)threadIdx.x;
  if (l >= Lp) return;
  const int Lpb = rows ? rows[b].Lp : Lp;
  const float* xr = x + (size_t)b * Lp;
  const float xm = (l > 0 ? xr[l - 1] : 0.f) * in_scale;
  const float x0 = xr[l] * in_scale;
  const float xp = (l + 1 < Lp ? xr[l + 1] : 0.f) * in_scale;
  if (l >= Lpb) return;   // rows past the utterance's own length stay zero (x is zero there as well)
  const RowAddr r0 = row_addr(o0.t, b, l), r1 = row_addr(o1.t, b, l);
  char* p0 = r0.p;
  char* p1 = r1.p;
  const int nch = C >> 3;
  for (int ch = 0; ch < nch; ++ch) {
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4 q = ws[ch * 8 + i];
      v[i] = fmaf(q.x, xm, fmaf(q.y, x0, fmaf(q.z, xp, q.w)));
    }
    if (n_out > 0) {
      store8_act(p0, o0.t.esize_log2, o0.act, o0.a1, o0.a2, v);
      p0 += r0.pstride;
    }
    if (n_out > 1) {
      store8_act(p1, o1.t.esize_log2, o1.act, o1.a1, o1.a2, v);
      p1 += r1.pstride;
    }
  }
}

// ---------------------------------------------------------------- INIT_X   x = sigma_0 * z_0 (universe.py:326)
__global__ void __launch_bounds__(256) init_x_kernel(const IoTable* io, long long n, int step, int b_off, int B_total, int Lp, float sigma0,
                                                     float* x, const RowGeo* __restrict__ rows) {
  pdl_launch_dependents();
  pdl_wait();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int b = (int)(i / Lp), l = (int)(i - (long long)b * Lp);
  if (rows && l >= rows[b].Lp) { x[i] = 0.f; return; }
  const float z = io->noise ? io->noise[((long long)step * B_total + b_off) * Lp + i] : unit_normal(io->seed, step, b_off + b, l);
  x[i] = sigma0 * z;
}

// ---------------------------------------------------------------- CONV_OUT
// output_conv (C -> 1, k=3; score.py:290) on the already doubly-activated planes, fused with
// the EDM wrapper and the sampler update (universe.py:197-209, 334-343):  x <- a x + b net + c z.
template <bool BF16>
__device__ __forceinline__ void unpack8(const char* p, float* v) {
  if (BF16) {
    const uint4 u = *reinterpret_cast<const uint4*>(p);
    const uint32_t q[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {   // bf16 -> f32 is a 16-bit shift
      v[2 * i] = __uint_as_float(q[i] << 16);
      v[2 * i + 1] = __uint_as_float(q[i] & 0xFFFF0000u);
    }
  } else {
    const float4 a = *reinterpret_cast<const float4*>(p), c = *reinterpret_cast<const float4*>(p + 16);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
  }
}
__device__ __forceinline__ float fma8(const float* v, const float* wk, float acc) {
  const float4 w0 = *reinterpret_cast<const float4*>(wk), w1 = *reinterpret_cast<const float4*>(wk + 4);
  acc = fmaf(v[0], w0.x, acc); acc = fmaf(v[1], w0.y, acc); acc = fmaf(v[2], w0.z, acc); acc = fmaf(v[3], w0.w, acc);
  acc = fmaf(v[4], w1.x, acc); acc = fmaf(v[5], w1.y, acc); acc = fmaf(v[6], w1.z, acc); acc = fmaf(v[7], w1.w, acc);
  return acc;
}

// Phase-split-by-2 input (the space-to-depth form of the finest level): thread = row m of the two phase planes =
// samples 2m and 2m + 1, whose taps are P1[m-1], P0[m], P1[m] and P0[m], P1[m], P0[m+1]: four coalesced 16-byte
// loads per chunk feed two samples (a thread per sample needs six, alternating between the planes).  Rows -1 and
// Lp / 2 are the never-written zero rows of the layout.  Per sample the sum runs chunk by chunk, tap by tap, channel
// by channel, exactly like the general path below.
template <bool BF16>
__device__ __forceinline__ void conv_out_pair(const TensDesc& a, int b, int m, int C, const float* ws, float& acc0, float& acc1) {
  const long long es8 = 8ll << a.esize_log2;
  const long long pstride = (long long)a.P * es8;
  const char* p0 = a.base + ((long long)a.G + (long long)b * a.Lh + m) * es8;   // phase 0, chunk 0, row m
  const char* p1 = p0 + (long long)a.nchunks * pstride;                          // phase 1
  const int nch = C >> 3;
  for (int ch = 0; ch < nch; ++ch) {
    float va[8], vb[8], vc[8], vd[8];
    unpack8<BF16>(p1 - es8, va);   // sample 2m - 1
    unpack8<BF16>(p0, vb);         // 2m
    unpack8<BF16>(p1, vc);         // 2m + 1
    unpack8<BF16>(p0 + es8, vd);   // 2m + 2
    const float* w0 = ws + ch * 8;
    acc0 = fma8(va, w0, acc0); acc0 = fma8(vb, w0 + C, acc0); acc0 = fma8(vc, w0 + 2 * C, acc0);
    acc1 = fma8(vb, w0, acc1); acc1 = fma8(vc, w0 + C, acc1); acc1 = fma8(vd, w0 + 2 * C, acc1);
    p0 += pstride;
    p1 += pstride;
  }
}

__global__ void __launch_bounds__(256) conv_out_kernel(TensDesc a, int B, int Lp, int C, const float* __restrict__ w /*[3][C]*/,
                                                       const float* __restrict__ bias, float* __restrict__ x,
                                                       const StepCoef* __restrict__ coef, int step, const IoTable* io,
                                                       int b_off, int B_total, int reverse, const RowGeo* __restrict__ rows) {
  pdl_launch_dependents();
  __shared__ __align__(16) float ws[3 * kEdgeMaxC];   // [tap][C]: 16-byte broadcast loads instead of one global load per weight
  for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) ws[i] = w[i];
  __syncthreads();
  pdl_wait();
  // grid (blocks per utterance, utterances): no 64-bit division per thread
  const int bx = reverse ? (int)(gridDim.x - 1 - blockIdx.x) : (int)blockIdx.x;
  const int b = reverse ? (int)(gridDim.y - 1 - blockIdx.y) : (int)blockIdx.y;
  const int Lpb = rows ? rows[b].Lp : Lp;
  const StepCoef cf = coef[step];
  const long long nbase = ((long long)step * B_total + b_off) * Lp;   // this lane's rows of step `step` in noise / dbg
  if (a.split == 2 && !(Lp & 1)) {
    const int m = bx * (int)blockDim.x + (int)threadIdx.x;
    const int l = 2 * m;
    if (l >= Lp) return;
    float acc0 = bias[0], acc1 = acc0;
    if (a.esize_log2 == 1) conv_out_pair<true>(a, b, m, C, ws, acc0, acc1);
    else conv_out_pair<false>(a, b, m, C, ws, acc0, acc1);
    const long long idx = (long long)b * Lp + l;
    float2* xp = reinterpret_cast<float2*>(x + idx);
    if (l >= Lpb) {   // past the utterance's own length (even, like Lp): the state stays zero
      *xp = make_float2(0.f, 0.f);
      if (io->dbg) { io->dbg[nbase + idx] = 0.f; io->dbg[nbase + idx + 1] = 0.f; }   // (caller memory: no alignment assumed)
      return;
    }
    const float2 xo = *xp;
    float2 xn = make_float2(cf.a * xo.x + cf.b * acc0, cf.a * xo.y + cf.b * acc1);
    if (cf.c != 0.f) {
      float2 z;
      if (io->noise) z = make_float2(io->noise[nbase + (long long)B_total * Lp + idx], io->noise[nbase + (long long)B_total * Lp + idx + 1]);
      else z = make_float2(unit_normal(io->seed, step + 1, b_off + b, l), unit_normal(io->seed, step + 1, b_off + b, l + 1));
      xn.x = fmaf(cf.c, z.x, xn.x);
      xn.y = fmaf(cf.c, z.y, xn.y);
    }
    *xp = xn;
    if (io->dbg) { io->dbg[nbase + idx] = acc0; io->dbg[nbase + idx + 1] = acc1; }
    return;
  }
  const int l = bx * (int)blockDim.x + (int)threadIdx.x;
  if (l >= Lp) return;
  const long long idx = (long long)b * Lp + l;
  // rows l-1, l, l+1 of the planes tensor (plain: neighbouring rows, rows -1 and Lp are the zero gap; phase-split:
  // the neighbours live in other phase planes, samples outside [0, Lp) are skipped)
  RowAddr rt[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int lk = l + k - 1;
    if (a.split > 1) {
      rt[k] = row_addr(a, b, lk < 0 ? 0 : lk);
      if (lk < 0 || lk >= Lp) rt[k].p = nullptr;
    } else {
      rt[k] = row_addr(a, b, l);
      rt[k].p += (long long)(k - 1) * (8 << a.esize_log2);
    }
  }
  float acc = bias[0];
  for (int ch = 0; ch < (C >> 3); ++ch) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const RowAddr rk = rt[k];
      if (rk.p == nullptr) continue;
      float v[8];
      load8(rk, a.esize_log2, ch, v);
      acc = fma8(v, ws + k * C + ch * 8, acc);
    }
  }
  if (l >= Lpb) {   // past the utterance's own length: the state stays zero (the rows read above are zero padding)
    x[idx] = 0.f;
    if (io->dbg) io->dbg[nbase + idx] = 0.f;
    return;
  }
  float xn = cf.a * x[idx] + cf.b * acc;
  if (cf.c != 0.f) {
    const float z = io->noise ? io->noise[nbase + (long long)B_total * Lp + idx] : unit_normal(io->seed, step + 1, b_off + b, l);
    xn = fmaf(cf.c, z, xn);
  }
  x[idx] = xn;
  if (io->dbg) io->dbg[nbase + idx] = acc;
}

// ---------------------------------------------------------------- POST
// unpad, optional keep_rms, per-utterance peak normalisation (universe.py:349-357). One CTA per utterance.
__global__ void __launch_bounds__(1024) post_kernel(const float* __restrict__ x, int T, int Lp, int pad_left,
                                                    const float* __restrict__ mix_rms, const int* __restrict__ keep_rms,
                                                    const IoTable* io, int b_off, const RowGeo* __restrict__ rows) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ double sh[96];
  __shared__ float s_max[32];
  const int b = blockIdx.x;
  const int Tp = T;   // row pitch of the caller's output tensor
  if (rows) { T = rows[b].T; pad_left = rows[b].pad_left; }
  const float* xr = x + (size_t)b * Lp + pad_left;
  double ss = 0, d1 = 0, d2 = 0;
  float mx = 0.f;
  for (int i = threadIdx.x; i < T; i += blockDim.x) {
    const float v = xr[i];
    ss += (double)v * v;
    mx = fmaxf(mx, fabsf(v));
  }
  block_sum3(ss, d1, d2, sh);
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = (threadIdx.x & 31) < (blockDim.x >> 5) ? s_max[threadIdx.x & 31] : 0.f;
  mx = warp_max(mx);
  float g = 1.f;
  if (*keep_rms) {
    const float x_rms = fmaxf((float)sqrt(ss / T), 1e-5f);
    g = mix_rms[b] / x_rms;
  }
  const float peak = mx * g;
  float* o = io->out + (size_t)(b_off + b) * Tp;
  for (int i = threadIdx.x; i < Tp; i += blockDim.x) {
    float v = i < T ? xr[i] * g : 0.f;
    if (peak > 1.f) v = v / peak;
    o[i] = v;
  }
}

// ---------------------------------------------------------------- FILM
// Noise-level embedding (SimpleTimeEmbedding sigma_block.py:73-78 / SigmaBlock :50-57) followed by
// every FiLM projection Linear(D -> 2C) of the score network (score.py:109,208) for one sampler step.
__global__ void __launch_bounds__(512) film_kernel(const float* __restrict__ log10s, int kind, int D, int n_rff, int R,
                                                   const float* __restrict__ emb, const float* __restrict__ W,
                                                   const float* __restrict__ bias, float* __restrict__ film) {
  extern __shared__ float smem_f[];
  float* g = smem_f;        // [D]
  float* tmp = smem_f + D;  // [D]
  const int step = blockIdx.x;
  const float ls = log10s[step];
  const float two_pi = 6.283185307179586f;
  if (kind == 0) {
    const float f = 0.5f * (1.f / (1.f + expf(-(emb[0] * ls + emb[1]))));
    for (int i = threadIdx.x; i < D / 2; i += blockDim.x) {
      const float p = two_pi * f * (float)i;
      g[i] = sinf(p);
      g[D / 2 + i] = cosf(p);
    }
    __syncthreads();
  } else {
    for (int i = threadIdx.x; i < n_rff; i += blockDim.x) {
      const float p = two_pi * emb[i] * ls;
      g[i] = sinf(p);
      g[n_rff + i] = cosf(p);
    }
    __syncthreads();
    const float* p = emb + n_rff;
    int in = 2 * n_rff;
    float* src = g;
    float* dst = tmp;
    for (int layer = 0; layer < 3; ++layer) {
      const int out = layer == 2 ? D : 2 * in;
      const float slope = p[0];
      const float* Wl = p + 1;
      const float* bl = Wl + (size_t)in * out;
      for (int o = threadIdx.x; o < out; o += blockDim.x) {
        float acc = bl[o];
        for (int k = 0; k < in; ++k) acc = fmaf(Wl[(size_t)o * in + k], src[k], acc);
        dst[o] = prelu1(acc, slope);
      }
      __syncthreads();
      p = bl + out;
      in = out;
      float* t = src; src = dst; dst = t;
    }
    if (src != g) {
      for (int i = threadIdx.x; i < D; i += blockDim.x) g[i] = src[i];
      __syncthreads();
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int r = warp; r < R; r += nw) {
    float acc = 0.f;
    for (int k = lane; k < D; k += 32) acc = fmaf(W[(size_t)r * D + k], g[k], acc);
    acc = warp_sum(acc);
    if (lane == 0) film[(size_t)step * R + r] = acc + bias[r];
  }
}

}  // namespace uv

// Standalone GPU unit test for gemm_tc_kernel (no torch): random planes tensors,
// CPU fp64 reference computed from the same bf16-rounded data.
//   ./test_gemm <case> [swap_lbo]      case = name or "perf"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <string>
#include <algorithm>
#include <vector>

#include "gemm_tc.cuh"

using namespace uv;

#define CK(x)                                                                        \
  do {                                                                               \
    cudaError_t e_ = (x);                                                            \
    if (e_ != cudaSuccess) {                                                         \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(2);                                                                       \
    }                                                                                \
  } while (0)

static uint32_t rng_state = 12345;
static float frand() {
  rng_state = rng_state * 1664525u + 1013904223u;
  return ((rng_state >> 8) & 0xFFFF) / 32768.0f - 1.0f;
}
static uint16_t f2bf(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  u += 0x7FFF + ((u >> 16) & 1);
  return (uint16_t)(u >> 16);
}
static float bf2f(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

// Host mirror of a planes tensor.
struct HT {
  int C, split, L, B, Lh, P, G, esz;  // L = valid unsplit length; Lh in split rows
  std::vector<uint8_t> data;
  char* dev = nullptr;
  size_t idx(int b, int l, int c) const {
    int ph = l % split, ll = l / split;
    size_t row = (size_t)(ph * (C / 8) + c / 8) * P + G + (size_t)b * Lh + ll;
    return (row * 8 + c % 8);
  }
  float get(int b, int l, int c) const {
    size_t i = idx(b, l, c);
    if (esz == 2) return bf2f(((const uint16_t*)data.data())[i]);
    return ((const float*)data.data())[i];
  }
  void set(int b, int l, int c, float v) {
    size_t i = idx(b, l, c);
    if (esz == 2) ((uint16_t*)data.data())[i] = f2bf(v);
    else ((float*)data.data())[i] = v;
  }
  // raw row access in phase plane `ph` at row-space index m (may be outside [0, L/split))
  float get_row(int ph, int b, int m, int c) const {
    size_t row = (size_t)(ph * (C / 8) + c / 8) * P + G + (size_t)b * Lh + m;
    size_t i = row * 8 + c % 8;
    return bf2f(((const uint16_t*)data.data())[i]);
  }
};

static HT make_tensor(int B, int L, int C, int split, int esz, bool randomize) {
  HT t;
  t.B = B; t.L = L; t.C = C; t.split = split; t.esz = esz; t.G = kGuard;
  int rows = (L + split - 1) / split;
  t.Lh = ((rows + 2 + 7) / 8) * 8;
  t.P = t.G + ((B * t.Lh + kMaxTileRows - 1) / kMaxTileRows) * kMaxTileRows + t.G;
  t.data.assign((size_t)split * (C / 8) * t.P * 8 * esz, 0);
  if (randomize)
    for (int b = 0; b < B; ++b)
      for (int l = 0; l < L; ++l)
        for (int c = 0; c < C; ++c) t.set(b, l, c, frand());
  CK(cudaMalloc(&t.dev, t.data.size()));
  CK(cudaMemcpy(t.dev, t.data.data(), t.data.size(), cudaMemcpyHostToDevice));
  return t;
}
static TensDesc desc_of(const HT& t) {
  TensDesc d;
  d.base = t.dev; d.esize_log2 = t.esz == 2 ? 1 : 2; d.nchunks = t.C / 8; d.split = t.split;
  d.Lh = t.Lh; d.P = t.P; d.G = t.G; d.pad_ = 0;
  return d;
}
static void fetch(HT& t) { CK(cudaMemcpy(t.data.data(), t.dev, t.data.size(), cudaMemcpyDeviceToHost)); }

struct Case {
  const char* name;
  int B, L_out_space;  // rows per utterance in A row space (M_valid)
  int C_in, n_src, smin, smax, KC;
  int C_out, up, N_TILE;
  bool bias, film, add1, add2f32;
  float scale;
  int out1_split;   // 0 = no second output
  bool rm;
};

static const Case kCases[] = {
    {"k1", 2, 200, 64, 1, 0, 0, 64, 64, 1, 64, false, false, false, false, 1.f, 0, false},
    {"k5", 2, 200, 64, 1, -2, 2, 64, 64, 1, 64, true, false, false, false, 1.f, 0, false},
    {"c32", 3, 301, 32, 1, -2, 2, 32, 32, 1, 32, true, false, false, false, 1.f, 0, false},
    {"c48", 2, 150, 48, 1, -1, 1, 48, 48, 1, 48, true, false, false, false, 1.f, 0, false},
    {"big", 2, 401, 256, 1, -1, 1, 64, 512, 1, 256, true, false, false, false, 1.f, 0, false},
    {"down", 2, 100, 64, 4, -1, 1, 64, 128, 1, 128, true, false, false, false, 1.f, 0, false},
    {"up", 2, 100, 128, 1, -1, 1, 64, 64, 4, 64, true, false, true, false, 0.70710678f, 0, false},
    {"epi", 2, 200, 64, 1, -2, 2, 64, 64, 1, 64, true, true, true, true, 0.5f, 2, true},
    {"mel", 2, 201, 80, 1, -1, 1, 80, 512, 1, 128, true, false, false, false, 1.f, 0, false},
    {"big128", 3, 401, 256, 1, -2, 2, 64, 512, 1, 128, true, false, true, false, 0.70710678f, 0, false},
    {"film", 2, 700, 64, 1, -2, 2, 64, 64, 1, 64, true, true, true, false, 0.70710678f, 0, false},
    {"nobias", 2, 300, 32, 1, -1, 1, 32, 32, 1, 32, false, false, false, false, 1.f, 0, false},
    {"c96", 2, 333, 96, 1, -1, 1, 48, 96, 1, 96, true, false, true, true, 1.f, 0, false},
    {"add32", 3, 777, 32, 1, -1, 1, 32, 32, 1, 32, true, false, true, false, 0.70710678f, 0, false},
    {"add64x2", 2, 500, 64, 1, -2, 2, 64, 64, 1, 64, true, true, true, false, 0.70710678f, 2, false},
    {"add128", 2, 400, 128, 1, -1, 1, 64, 128, 1, 128, true, false, true, false, 0.70710678f, 0, false},
    {"add96", 2, 300, 96, 1, -1, 1, 48, 96, 1, 96, true, false, true, false, 1.f, 0, false},
    {"upsplit", 2, 300, 128, 1, -1, 1, 64, 64, 4, 64, true, false, true, false, 0.70710678f, 0, false},
    {"add256reg", 2, 300, 256, 1, -1, 1, 64, 256, 1, 128, true, false, true, false, 0.70710678f, 0, false},
    {"rmonly", 2, 130, 64, 1, 0, 0, 64, 256, 1, 128, true, false, false, false, 1.f, -1, true},
    // raw A tensor, PReLU applied in shared memory by the transform warps (names starting with "xf")
    {"xf32", 3, 777, 32, 1, -2, 2, 32, 32, 1, 32, true, false, false, false, 1.f, 0, false},
    {"xf64add", 2, 500, 64, 1, -2, 2, 64, 64, 1, 64, true, true, true, false, 0.70710678f, 2, false},
    {"xfdown", 2, 100, 64, 4, -1, 1, 64, 128, 1, 128, true, false, false, false, 1.f, 0, false},
    {"xfbig", 2, 401, 256, 1, -1, 1, 64, 512, 1, 128, true, false, false, false, 1.f, 0, false},
    {"xfupsplit", 2, 300, 128, 1, -1, 1, 64, 64, 4, 64, true, false, true, false, 0.70710678f, 0, false},
    // transposed convs whose N tile spans two whole phases (N_TILE = 2 C_out)
    {"mp2upsplit", 3, 300, 64, 1, -1, 1, 64, 32, 2, 64, true, false, true, false, 0.70710678f, 0, false},
    {"mp4upsplit", 2, 300, 128, 1, -1, 1, 64, 64, 4, 128, true, false, true, false, 0.70710678f, 0, false},
    {"mp2up", 2, 301, 64, 1, -1, 1, 64, 32, 2, 64, true, false, false, false, 1.f, 2, false},
    {"xfmp3up", 2, 200, 96, 1, -1, 1, 48, 32, 3, 96, true, false, false, false, 1.f, 0, false},
    // N tiles of 256 columns (two accumulator stages of the SM pair)
    {"n256", 3, 401, 256, 1, -1, 1, 64, 512, 1, 256, true, false, false, false, 1.f, 0, false},
    {"n256add", 2, 300, 256, 1, -2, 2, 64, 256, 1, 256, true, true, true, false, 0.70710678f, 2, false},
    {"n256upsplit", 2, 100, 128, 1, -1, 1, 64, 64, 4, 256, true, false, true, false, 0.70710678f, 0, false},
    {"rm256", 2, 130, 64, 1, 0, 0, 64, 512, 1, 256, true, false, false, false, 1.f, -1, true},
    {"tiny", 1, 10, 32, 1, -1, 1, 32, 32, 1, 32, true, false, false, false, 1.f, 0, false},
};

static int run_case(const Case& c, int swap_lbo, int num_sms) {
  const bool a_act = strncmp(c.name, "xf", 2) == 0;
  const float a_slope = 0.3f;
  auto xf = [&](float v) { return a_act ? bf2f(f2bf(v >= 0.f ? v : a_slope * v)) : v; };
  const int n_shift = c.smax - c.smin + 1;
  const int N_total = c.up * c.C_out;
  // A: split tensor with n_src phases; unsplit valid length = M_valid * n_src
  HT A = make_tensor(c.B, c.L_out_space * c.n_src, c.C_in, c.n_src, 2, true);
  const int L_out = c.L_out_space * c.up;
  // weights [n][s][q][ci]
  std::vector<float> W((size_t)N_total * c.n_src * n_shift * c.C_in);
  for (auto& w : W) w = bf2f(f2bf(frand() * 0.1f));
  const int n_groups = c.C_in / c.KC, n_tiles_n = N_total / c.N_TILE, kcc = c.KC / 8;
  std::vector<float> bias(c.C_out), film(2 * c.C_out);
  for (auto& v : bias) v = frand() * 0.5f;
  for (auto& v : film) v = frand();
  // packed per (N tile, half of its columns): the CTA of rank h of a pair holds columns [h N_TILE/2, (h+1) N_TILE/2)
  const int NH = c.N_TILE / 2;
  const size_t n_img_elems = (size_t)c.n_src * n_groups * n_shift * NH * c.KC;
  const size_t bias_elems = c.bias ? (size_t)NH * 16 : 0;
  const size_t nt_stride = n_img_elems + bias_elems;   // per half
  std::vector<uint16_t> img(nt_stride * 2 * n_tiles_n, 0);
  for (int nt = 0; nt < n_tiles_n; ++nt)
    for (int h = 0; h < 2; ++h) {
      const size_t hb = (size_t)(2 * nt + h) * nt_stride;
      for (int s = 0; s < c.n_src; ++s)
        for (int g = 0; g < n_groups; ++g)
          for (int q = 0; q < n_shift; ++q) {
            size_t base = hb + (((size_t)s * n_groups + g) * n_shift + q) * (size_t)NH * c.KC;
            for (int j = 0; j < kcc; ++j)
              for (int n = 0; n < NH; ++n)
                for (int e = 0; e < 8; ++e) {
                  int ci = g * c.KC + j * 8 + e;
                  int ng = nt * c.N_TILE + h * NH + n;
                  img[base + ((size_t)j * NH + n) * 8 + e] =
                      f2bf(W[(((size_t)ng * c.n_src + s) * n_shift + q) * c.C_in + ci]);
                }
          }
      if (c.bias)
        for (int n = 0; n < NH; ++n) {
          const float bv = bias[(nt * c.N_TILE + h * NH + n) % c.C_out];
          const uint16_t hi = f2bf(bv);
          img[hb + n_img_elems + (size_t)n * 8 + 0] = hi;
          img[hb + n_img_elems + (size_t)n * 8 + 1] = f2bf(bv - bf2f(hi));
        }
    }
  char* d_w;
  CK(cudaMalloc(&d_w, img.size() * 2));
  CK(cudaMemcpy(d_w, img.data(), img.size() * 2, cudaMemcpyHostToDevice));
  float *d_bias, *d_film;
  CK(cudaMalloc(&d_bias, bias.size() * 4));
  CK(cudaMalloc(&d_film, film.size() * 4));
  CK(cudaMemcpy(d_bias, bias.data(), bias.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_film, film.data(), film.size() * 4, cudaMemcpyHostToDevice));

  HT add1 = make_tensor(c.B, L_out, c.C_out, strstr(c.name, "split") ? c.up : 1, 2, true);
  HT add2 = make_tensor(c.B, L_out, c.C_out, 1, 4, true);
  HT out0 = make_tensor(c.B, L_out, c.C_out, 1, 2, false);
  HT out1 = make_tensor(c.B, L_out, c.C_out, c.out1_split > 0 ? c.out1_split : 1, 2, false);
  float* d_rm = nullptr;
  const int m_rows = c.B * A.Lh;
  if (c.rm) {
    CK(cudaMalloc(&d_rm, (size_t)m_rows * N_total * 4));
    CK(cudaMemset(d_rm, 0, (size_t)m_rows * N_total * 4));
  }
  int* d_err;
  CK(cudaMalloc(&d_err, 4));
  CK(cudaMemset(d_err, 0, 4));

  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.a_base = A.dev; p.a_nchunks = c.C_in / 8; p.a_P = A.P; p.a_G = A.G; p.n_src = c.n_src;
  p.smin = c.smin; p.smax = c.smax; p.KC = c.KC; p.n_groups = n_groups;
  p.w = d_w; p.N_TILE = c.N_TILE; p.n_tiles_n = n_tiles_n;
  p.m_rows = m_rows; p.Lh_A = A.Lh; p.M_valid = c.L_out_space;
  p.up = c.up; p.C_out = c.C_out;
  p.has_bias = c.bias ? 1 : 0;
  p.film = c.film ? d_film : nullptr;
  if (c.add1) p.add1 = desc_of(add1);
  if (c.add2f32) p.add2 = desc_of(add2);
  p.scale = c.scale;
  p.n_out = c.out1_split < 0 ? 0 : (c.out1_split ? 2 : 1);
  p.out[0].t = desc_of(out0); p.out[0].act = 0;
  p.out[1].t = desc_of(out1); p.out[1].act = 2; p.out[1].a1 = 0.25f; p.out[1].a2 = -0.5f;
  p.rm_out = d_rm; p.rm_ld = N_total;
  p.err_flag = d_err;
  p.a_act = a_act ? 1 : 0;
  p.a_slope = a_slope;
  int* d_sched;
  CK(cudaMalloc(&d_sched, 8));
  CK(cudaMemset(d_sched, 0, 8));
  p.sched = d_sched;

  CK(gemm_launch(p, num_sms, 0));
  CK(cudaDeviceSynchronize());      // (one counter pair per op: back-to-back launches of the SAME op must not overlap under PDL)
  CK(gemm_launch(p, num_sms, 0));   // relaunch: the kernel rewinds its own tile counter
  cudaError_t e = cudaDeviceSynchronize();
  int err = 0;
  if (e != cudaSuccess) {
    printf("[%s] kernel failed: %s\n", c.name, cudaGetErrorString(e));
    return 1;
  }
  CK(cudaMemcpy(&err, d_err, 4, cudaMemcpyDeviceToHost));
  fetch(out0);
  fetch(out1);
  std::vector<float> rm;
  if (c.rm) {
    rm.resize((size_t)m_rows * N_total);
    CK(cudaMemcpy(rm.data(), d_rm, rm.size() * 4, cudaMemcpyDeviceToHost));
  }

  double max_err = 0, max_err1 = 0, max_rm = 0, max_ref = 0;
  for (int b = 0; b < c.B; ++b)
    for (int m = 0; m < c.L_out_space; ++m)
      for (int n = 0; n < N_total; ++n) {
        double acc = 0;
        for (int s = 0; s < c.n_src; ++s)
          for (int q = 0; q < n_shift; ++q)
            for (int ci = 0; ci < c.C_in; ++ci)
              acc += (double)xf(A.get_row(s, b, m + c.smin + q, ci)) *
                     W[(((size_t)n * c.n_src + s) * n_shift + q) * c.C_in + ci];
        int phase = n / c.C_out, co = n % c.C_out, l = c.up * m + phase;
        double v = acc;
        if (c.bias) v += bias[co];
        if (c.add1) v += add1.get(b, l, co);
        if (c.add2f32) v += add2.get(b, l, co);
        v *= c.scale;
        if (c.film) v = film[co] * v + film[c.C_out + co];
        max_ref = fmax(max_ref, fabs(v));
        if (c.out1_split >= 0) max_err = fmax(max_err, fabs(v - out0.get(b, l, co)));
        if (c.out1_split > 0) {
          double w = v >= 0 ? v : 0.25 * v;
          w = w >= 0 ? w : -0.5 * w;
          max_err1 = fmax(max_err1, fabs(w - out1.get(b, l, co)));
        }
        if (c.rm) max_rm = fmax(max_rm, fabs(v - rm[(size_t)(b * A.Lh + m) * N_total + n]));
      }
  // gap rows must stay zero
  double gap = 0;
  for (int b = 0; b < c.B; ++b)
    for (int co = 0; co < c.C_out; ++co) {
      size_t row = (size_t)(co / 8) * out0.P + out0.G + (size_t)b * out0.Lh + L_out;
      gap = fmax(gap, fabs(bf2f(((uint16_t*)out0.data.data())[row * 8 + co % 8])));
    }
  bool ok = max_err < 0.02 * fmax(1.0, max_ref) && max_err1 < 0.02 * fmax(1.0, max_ref) && max_rm < 1e-3 * fmax(1.0, max_ref) &&
            gap == 0 && err == 0;
  printf("[%s] swap=%d max_ref=%.4f err_out0=%.5f err_out1=%.5f err_rm=%.6f gap=%.3g flag=%d -> %s\n", c.name, swap_lbo,
         max_ref, max_err, max_err1, max_rm, gap, err, ok ? "PASS" : "FAIL");
  return ok ? 0 : 1;
}

// Throughput of layer-shaped problems (cfg 2 sizes: B=64, 4 s clips).
// phases = true: a few layers at 1/4 and full size with the per-CTA phase stamps of the kernel (where a launch's
// fixed cost goes); times in us relative to the earliest CTA entry, min / median / max over the CTAs.
static void run_perf(int num_sms, bool phases = false) {
  struct PC { const char* name; int B, Lv, C, k, add, nout; };
  std::vector<PC> pcs;
  if (phases) pcs = {{"L0 C32 k3 B16", 16, 64160, 32, 3, 0, 1}, {"L0 C32 k3 B64", 64, 64160, 32, 3, 0, 1},
                     {"L1 C64 k5 B16", 16, 32080, 64, 5, 0, 1}, {"L1 C64 k5 B64", 64, 32080, 64, 5, 0, 1},
                     {"L2 C128 k3 B16", 16, 8020, 128, 3, 0, 1}, {"L2 C128 k3 B64", 64, 8020, 128, 3, 0, 1},
                     {"L3 C256 k3 B16", 16, 2005, 256, 3, 0, 1}, {"L3 C256 k3 B64", 64, 2005, 256, 3, 0, 1},
                     {"L4 C512 k3 B64", 64, 401, 512, 3, 0, 1}};
  else pcs = {
      {"L0 C32 k5", 64, 64160, 32, 5, 0, 1}, {"L0 C32 k3", 64, 64160, 32, 3, 0, 1}, {"L0 C32 k3 +add 2out", 64, 64160, 32, 3, 1, 2},
      {"L0 C32 k3 +add 1out", 64, 64160, 32, 3, 1, 1}, {"L0 C32 k3 2out", 64, 64160, 32, 3, 0, 2}, {"L0 C32 k3 3out", 64, 64160, 32, 3, 0, 3},
      {"L1 C64 k3 +add 1out", 64, 32080, 64, 3, 1, 1}, {"L1 C64 k3 2out", 64, 32080, 64, 3, 0, 2},
      {"L1 C64 k5", 64, 32080, 64, 5, 0, 1}, {"L1 C64 k3", 64, 32080, 64, 3, 0, 1}, {"L1 C64 k3 +add 2out", 64, 32080, 64, 3, 1, 2},
      {"L2 C128 k5", 64, 8020, 128, 5, 0, 1}, {"L2 C128 k3", 64, 8020, 128, 3, 0, 1}, {"L2 C128 k3 +add 2out", 64, 8020, 128, 3, 1, 2},
      {"L3 C256 k5", 64, 2005, 256, 5, 0, 1}, {"L3 C256 k3", 64, 2005, 256, 3, 0, 1},
      {"L4 C512 k5", 64, 401, 512, 5, 0, 1}, {"L4 C512 k3", 64, 401, 512, 3, 0, 1},
      // raw A tensor + in-kernel PReLU ("xf"): against the same layer above
      {"L0 C32 k5 xf", 64, 64160, 32, 5, 0, 1}, {"L1 C64 k5 xf", 64, 32080, 64, 5, 0, 1}, {"L2 C128 k5 xf", 64, 8020, 128, 5, 0, 1},
      {"L3 C256 k5 xf", 64, 2005, 256, 5, 0, 1}, {"L4 C512 k5 xf", 64, 401, 512, 5, 0, 1},
      // the same level-0 / level-1 layers on tensors stored phase-split by 2 ("space to depth": 2 sources x 3 shifts,
      // N = 2 C columns): what a k5 / k3 "same" conv would cost in that formulation (dense weight images)
      {"L0 C32 s2d", 64, 64160, 32, 3, 0, 1}, {"L0 C32 s2d xf", 64, 64160, 32, 3, 0, 1}, {"L0 C32 s2d +add", 64, 64160, 32, 3, 1, 1},
      {"L1 C64 s2d", 64, 32080, 64, 3, 0, 1}, {"L1 C64 s2d xf", 64, 32080, 64, 3, 0, 1}};
  unsigned long long* d_clk = nullptr;
  if (phases) CK(cudaMalloc(&d_clk, (size_t)num_sms * 8 * 8));
  for (auto& pc : pcs) {
    const bool s2d = strstr(pc.name, "s2d") != nullptr;
    const int sp = s2d ? 2 : 1;
    HT A = make_tensor(pc.B, pc.Lv, pc.C, sp, 2, false);
    CK(cudaMemset(A.dev, 0x3c, A.data.size()));  // small finite bf16 values everywhere
    HT O = make_tensor(pc.B, pc.Lv, pc.C, sp, 2, false);
    HT O2 = make_tensor(pc.B, pc.Lv, pc.C, 1, 2, false);
    const int ntmax = getenv("UV_PERF_NT") ? atoi(getenv("UV_PERF_NT")) : 256;
    const int KC = pc.C < 64 ? pc.C : 64, NT = s2d ? 2 * pc.C : (pc.C < ntmax ? pc.C : ntmax);
    size_t wbytes = (size_t)pc.C * pc.C * pc.k * 2 * sp * sp + (size_t)pc.C * 32 * sp;
    char* d_w;
    CK(cudaMalloc(&d_w, wbytes));
    CK(cudaMemset(d_w, 0x3c, wbytes));
    GemmParams p;
    memset(&p, 0, sizeof(p));
    p.a_base = A.dev; p.a_nchunks = pc.C / 8; p.a_P = A.P; p.a_G = A.G; p.n_src = 1;
    p.smin = -(pc.k / 2); p.smax = pc.k / 2; p.KC = KC; p.n_groups = pc.C / KC;
    p.w = d_w; p.N_TILE = NT; p.n_tiles_n = pc.C / NT;
    p.m_rows = pc.B * A.Lh; p.Lh_A = A.Lh; p.M_valid = pc.Lv / sp; p.up = sp; p.C_out = pc.C;
    if (s2d) { p.n_src = 2; p.smin = -1; p.smax = 1; p.n_tiles_n = 1; }
    p.has_bias = 1; p.scale = 1.f; p.n_out = pc.nout;
    p.out[0].t = desc_of(O); p.out[1].t = desc_of(O2); p.out[1].act = 1; p.out[1].a1 = 0.25f;
    HT O3 = make_tensor(pc.B, pc.Lv, pc.C, 1, 2, false);
    p.out[2].t = desc_of(O3); p.out[2].act = 1; p.out[2].a1 = 0.5f;
    HT AD = make_tensor(pc.B, pc.Lv, pc.C, sp, 2, false);
    if (pc.add) p.add1 = desc_of(AD);
    if (strstr(pc.name, " xf")) { p.a_act = 1; p.a_slope = 0.25f; }
    int* d_err; CK(cudaMalloc(&d_err, 4)); CK(cudaMemset(d_err, 0, 4));
    p.err_flag = d_err;
    int* d_sched; CK(cudaMalloc(&d_sched, 64)); CK(cudaMemset(d_sched, 0, 64));
    p.sched = d_sched;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    // one counter pair per op in the product; here the same op is relaunched back to back, so rotate pairs
    for (int i = 0; i < 3; ++i) { p.sched = d_sched + 2 * (i & 7); CK(gemm_launch(p, num_sms, 0)); }
    CK(cudaDeviceSynchronize());
    const int iters = 10;
    CK(cudaEventRecord(e0));
    for (int i = 0; i < iters; ++i) { p.sched = d_sched + 2 * (i & 7); CK(gemm_launch(p, num_sms, 0)); }
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= iters;
    double flop = 2.0 * pc.B * pc.Lv * (double)pc.C * pc.C * pc.k;
    double bytes = (double)pc.B * pc.Lv * pc.C * 2 * (1 + pc.nout + pc.add);
    printf("[perf %s] %.3f ms  %.1f TFLOP/s  %.0f GB/s (algorithmic reads + writes)\n", pc.name, ms, flop / ms * 1e-9, bytes / ms * 1e-6);
    if (phases) {
      CK(cudaMemset(d_clk, 0, (size_t)num_sms * 64));
      p.clk = d_clk;
      CK(cudaEventRecord(e0));
      CK(gemm_launch(p, num_sms, 0));
      CK(cudaEventRecord(e1));
      CK(cudaDeviceSynchronize());
      CK(cudaEventElapsedTime(&ms, e0, e1));
      std::vector<unsigned long long> h((size_t)num_sms * 8);
      CK(cudaMemcpy(h.data(), d_clk, h.size() * 8, cudaMemcpyDeviceToHost));
      unsigned long long t0 = ~0ull;
      int nc = 0;
      for (int c = 0; c < num_sms; ++c)
        if (h[(size_t)c * 8]) { t0 = std::min(t0, h[(size_t)c * 8]); ++nc; }
      static const char* nm[8] = {"entry", "prologue done", "1st tile loads issued", "weights landed", "1st A slab landed",
                                  "1st accumulator done", "last tile stored", "exit"};
      printf("   single launch %.1f us (events), %d CTAs; phase: min / median / max us after the first CTA entry\n", ms * 1e3, nc);
      for (int i = 0; i < 8; ++i) {
        std::vector<double> v;
        for (int c = 0; c < num_sms; ++c)
          if (h[(size_t)c * 8 + i]) v.push_back((double)(h[(size_t)c * 8 + i] - t0) * 1e-3);
        if (v.empty()) continue;
        std::sort(v.begin(), v.end());
        printf("   %-22s %8.2f %8.2f %8.2f\n", nm[i], v.front(), v[v.size() / 2], v.back());
      }
    }
    cudaFree(O3.dev); cudaFree(AD.dev); cudaFree(A.dev); cudaFree(O.dev); cudaFree(O2.dev); cudaFree(d_w); cudaFree(d_err);
  }
}

int main(int argc, char** argv) {
  if (gemm_set_attributes() != cudaSuccess) { fprintf(stderr, "cudaFuncSetAttribute failed (no sm_100 device?)\n"); }
  if (argc < 2) { printf("usage: test_gemm <case|perf|list> [swap]\n"); return 2; }
  int dev_sms = 0;
  CK(cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, 0));
  std::string name = argv[1];
  int swap = argc > 2 ? atoi(argv[2]) : 0;
  if (name == "list") { for (auto& c : kCases) printf("%s\n", c.name); return 0; }
  if (name == "perf") { run_perf(dev_sms); return 0; }
  if (name == "phases") { run_perf(dev_sms, true); return 0; }
  for (auto& c : kCases)
    if (name == c.name) return run_case(c, swap, dev_sms);
  printf("unknown case %s\n", name.c_str());
  return 2;
}

// C ABI (include/universe_b200.h): model = device weight blob, plan = op program +
// arena + CUDA graph.  Translates uv_op records into kernel launches.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/universe_b200.h"
#include "gemm_tc.cuh"
#include "gru.cuh"
#include "convblock_pair.cuh"
#include "small_ops.cuh"
#include "aux_ops.cuh"
#include "hp_ops.cuh"
#include "wav_io.cuh"

using namespace uv;

static constexpr int kMelMaxSmem = 200 * 1024;   // dynamic shared memory the mel kernel may ask for (set per device in uv_model_create)

static thread_local std::string g_err;
static int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
#define CUDA_TRY(x)                                                                              \
  do {                                                                                           \
    cudaError_t e_ = (x);                                                                        \
    if (e_ != cudaSuccess)                                                                       \
      return fail(UV_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_) + " (" + __FILE__ + ":" + \
                                   std::to_string(__LINE__) + ")");                              \
  } while (0)

// Entry points select the handle's device and restore the caller's current device on return.
struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

struct uv_model {
  int device = 0;
  char* blob = nullptr;
  size_t nbytes = 0;
  int num_sms = 0;
};

struct uv_plan {
  uv_model* model = nullptr;
  std::vector<uv_op> ops;
  char* arena = nullptr;
  size_t arena_bytes = 0;
  IoTable* io_dev = nullptr;
  // pinned staging of the per-call io table and sampler coefficients: a ring of kStage slots, each guarded by an
  // event recorded behind its H2D copies, so that a call never waits for the previous call's device work
  static constexpr int kStage = 4;
  char* stage_host = nullptr;     // kStage x stage_bytes: IoTable | uv_step_coef[n_steps] | int keep_rms (+pad)
  size_t stage_bytes = 0, coef_bytes = 0;
  cudaEvent_t stage_ev[kStage] = {nullptr, nullptr, nullptr, nullptr};
  bool stage_used[kStage] = {false, false, false, false};
  int stage_next = 0;
  std::vector<int64_t> coef_offs;   // one per utterance lane
  int n_steps = 0;
  int* err_host = nullptr;        // mapped pinned host int written by a tripped device watchdog (mbar_wait)
  int* err_flag = nullptr;        // its device alias
  std::mutex mu;                  // calls on one plan are serialised (one arena, one io table)
  int* sched = nullptr;           // [n_ops][2] tile / finished-CTA counters of the GEMM launches (self-rewinding)
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  cudaStream_t cap_stream = nullptr;
  cudaStream_t lane_stream = nullptr;   // second capture stream: ops of lane 1
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_lane[2] = {nullptr, nullptr};
  int n_launches = 0;
  bool use_graph = false;
  bool lazy_graph = false;        // use_graph == 2: the first uv_enhance() runs eagerly, the second captures the graph
  long long n_calls = 0;
  // uv_set_row_tables: per-utterance geometry of variable-length calls, staged like the sampler coefficients
  std::vector<char> row_blob;
  size_t row_off = 0, clear_off = 0, clear_bytes = 0;
  bool row_dirty = false;         // copy the tables (and clear the activation planes) in front of the next call
  char* row_host = nullptr;       // pinned ring: kStage x row_cap
  size_t row_cap = 0;
  unsigned long long noise_seed = 0;   // uv_set_noise_seed: key of the in-kernel draws of calls made with noise == NULL
  bool seeded = false;
};

extern "C" int uv_abi_version(void) { return UV_ABI_VERSION; }
extern "C" size_t uv_sizeof_op(void) { return sizeof(uv_op); }
extern "C" const char* uv_last_error(void) { return g_err.c_str(); }

extern "C" int uv_model_create(uv_model** out, const void* weights, size_t nbytes, int device) {
  if (!out || !weights || nbytes == 0) return fail(UV_ERR_INVALID, "uv_model_create: null argument");
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device >= n) return fail(UV_ERR_NO_DEVICE, "no such CUDA device");
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return fail(UV_ERR_NO_DEVICE, "device is not sm_100 (B200): found sm_" + std::to_string(prop.major * 10 + prop.minor));
  DeviceGuard guard(device);
  // function attributes are per device: every model (= device) sets them, no process-wide "done" flags
  CUDA_TRY(gemm_set_attributes());
  CUDA_TRY(convblock_pair_set_attributes());
  CUDA_TRY(gemm_hp_tc_set_attributes());
  CUDA_TRY(gru_hp_set_attributes());
  CUDA_TRY(cudaFuncSetAttribute(aux_to_wav_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  CUDA_TRY(cudaFuncSetAttribute(gru_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  CUDA_TRY(cudaFuncSetAttribute(mel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMelMaxSmem));
  CUDA_TRY(cudaFuncSetAttribute(mel_pow_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMelMaxSmem));
  char* blob = nullptr;
  CUDA_TRY(cudaMalloc(&blob, nbytes));
  cudaError_t e = cudaMemcpy(blob, weights, nbytes, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    cudaFree(blob);
    return fail(UV_ERR_CUDA, std::string("weight upload: ") + cudaGetErrorString(e));
  }
  uv_model* m = new uv_model();
  m->device = device;
  m->nbytes = nbytes;
  m->num_sms = prop.multiProcessorCount;
  m->blob = blob;
  *out = m;
  return UV_OK;
}

extern "C" void uv_model_destroy(uv_model* m) {
  if (!m) return;
  DeviceGuard guard(m->device);
  cudaFree(m->blob);
  delete m;
}

// ------------------------------------------------------------------ op translation
static TensDesc to_desc(const uv_plan* p, const uv_tens& t) {
  TensDesc d;
  memset(&d, 0, sizeof(d));
  if (t.off < 0) return d;
  d.base = p->arena + t.off;
  d.esize_log2 = t.dtype == 1 ? 1 : 2;
  d.nchunks = t.nch; d.split = t.split; d.Lh = t.Lh; d.P = t.P; d.G = t.G;
  return d;
}
static OutDesc to_out(const uv_plan* p, const uv_out& o) {
  OutDesc d;
  memset(&d, 0, sizeof(d));
  d.t = to_desc(p, o.t);
  d.act = o.act; d.a1 = o.a1; d.a2 = o.a2;
  return d;
}
template <typename T>
static T* wptr(const uv_plan* p, int64_t off) { return off < 0 ? nullptr : reinterpret_cast<T*>(p->model->blob + off); }
template <typename T>
static T* aptr(const uv_plan* p, int64_t off) { return off < 0 ? nullptr : reinterpret_cast<T*>(p->arena + off); }

static const RowGeo* rowgeo(const uv_plan* p, const uv_op& op) { return aptr<const RowGeo>(p, op.rows); }

static int check_tens(const uv_plan* p, const uv_tens& t, const char* what) {
  if (t.off < 0) return UV_OK;
  const size_t es = t.dtype == 1 ? 2 : 4;
  const size_t end = (size_t)t.off + (size_t)t.split * t.nch * t.P * 8 * es;
  if (end > p->arena_bytes || (t.off & 15)) return fail(UV_ERR_INVALID, std::string("tensor out of arena: ") + what);
  return UV_OK;
}

static int launch_op(uv_plan* p, int op_index, cudaStream_t s) {
  const uv_op& op = p->ops[op_index];
  switch (op.type) {
    case UV_OP_GEMM: {
      if (op.a.dtype == 2) {
        // fp32 parity tier (csrc/hp_ops.cuh): f32 planes in, the fp32 image of the same multi-tap GEMM at p5
        if (op.p5 < 0) return fail(UV_ERR_INVALID, "GEMM on f32 planes needs the fp32 weight image (p5; pack_model(hp=True))");
        if (op.rm_out >= 0 && op.H > 0) return fail(UV_ERR_INVALID, "fp32 GEMM: row-major outputs only (H == 0)");
        HpGemmParams h;
        memset(&h, 0, sizeof(h));
        h.a_base = reinterpret_cast<const float*>(p->arena + op.a.off);
        h.a_nch = op.a.nch; h.a_P = op.a.P; h.a_G = op.a.G;
        h.n_src = op.n_src; h.smin = op.smin; h.n_shift = op.smax - op.smin + 1;
        h.w = wptr<const float>(p, op.p5);
        h.bias = wptr<const float>(p, op.bias_w);
        h.N_total = op.N_total; h.C_out = op.C_out; h.up = op.up;
        h.m_rows = op.m_rows; h.Lh_A = op.Lh_A; h.M_valid = op.M_valid;
        h.mv = op.rows >= 0 ? aptr<const int>(p, op.rows) : nullptr;
        h.film = aptr<const float>(p, op.film);
        h.add1 = to_desc(p, op.add1); h.add2 = to_desc(p, op.add2);
        h.scale = op.scale; h.n_out = op.n_out;
        for (int i = 0; i < 3; ++i) h.out[i] = to_out(p, op.out[i]);
        h.rm_out = aptr<float>(p, op.rm_out); h.rm_ld = op.rm_ld;
        h.a_act = op.a_act; h.a_slope = op.a_slope;
        if (h.N_total % 8 || h.C_out % 8 || h.a_P < h.a_G + ((h.m_rows + kHpBM - 1) / kHpBM) * kHpBM + op.smax)
          return fail(UV_ERR_INVALID, "fp32 GEMM: geometry unsupported");
        // default: split-bf16 ("bf16x3") on the tensor cores; UV_HP_SIMT=1: the fp32 FMA kernel (same results to ~1e-5)
        static const bool hp_simt = getenv("UV_HP_SIMT") && atoi(getenv("UV_HP_SIMT")) != 0;
        h.w_tc = wptr<const char>(p, op.p4); h.KC = op.i0; h.N_TILE = op.i2;
        h.err_flag = p->err_flag;
        if (!hp_simt && gemm_hp_tc_ok(h)) CUDA_TRY(gemm_hp_tc_launch(h, s));
        else CUDA_TRY(gemm_hp_launch(h, s));
        break;
      }
      GemmParams g;
      memset(&g, 0, sizeof(g));
      g.a_base = p->arena + op.a.off;
      g.a_nchunks = op.a.nch; g.a_P = op.a.P; g.a_G = op.a.G;
      g.n_src = op.n_src; g.smin = op.smin; g.smax = op.smax; g.KC = op.KC;
      g.n_groups = op.a.nch * 8 / op.KC;
      g.w = wptr<const char>(p, op.w_w);
      g.N_TILE = op.N_TILE; g.n_tiles_n = op.N_total / op.N_TILE;
      g.m_rows = op.m_rows; g.Lh_A = op.Lh_A; g.M_valid = op.M_valid;
      g.up = op.up; g.C_out = op.C_out;
      g.has_bias = op.bias_w >= 0 ? 1 : 0;
      g.film = aptr<const float>(p, op.film);
      g.add1 = to_desc(p, op.add1); g.add2 = to_desc(p, op.add2);
      g.scale = op.scale; g.n_out = op.n_out;
      for (int i = 0; i < 3; ++i) g.out[i] = to_out(p, op.out[i]);
      g.rm_out = aptr<float>(p, op.rm_out); g.rm_ld = op.rm_ld;
      g.rm_H = op.rm_out >= 0 ? op.H : 0; g.rm_Bp = op.i1;   // GEMM ops with a row-major output feed the GRU (H, batch pitch)
      g.mv = op.rows >= 0 ? aptr<const int>(p, op.rows) : nullptr;
      g.err_flag = p->err_flag;
      g.a_act = op.a_act; g.a_slope = op.a_slope;
      g.reverse = op.i3 & 1;   // tile order (L2 reuse with the producer of A)
      g.sched = p->sched + 2 * op_index;
      CUDA_TRY(gemm_launch(g, p->model->num_sms, s));
      break;
    }
    case UV_OP_CONVBLOCK: {
      // every tensor of the block as 8 chunk planes of 64 "channels": a plain 64-channel tensor, or 32 channels stored
      // phase-split by 2 in this row space (the space-to-depth form of the finest level)
      auto as64 = [](TensDesc d) {
        if (d.base && d.split == 2 && d.nchunks == 4) { d.split = 1; d.nchunks = 8; }
        return d;
      };
      CbParams c;
      memset(&c, 0, sizeof(c));
      const TensDesc x = as64(to_desc(p, op.a));
      if (!x.base || x.split != 1 || x.nchunks != kCbPlanes || x.esize_log2 != 1 || !op.a_act)
        return fail(UV_ERR_INVALID, "fused ConvBlock: input must be a raw 64-column bf16 tensor");
      c.x_base = x.base; c.x_P = x.P; c.x_G = x.G;
      c.a_slope = op.a_slope;
      c.w[0] = wptr<const char>(p, op.w_w); c.w[1] = wptr<const char>(p, op.p0); c.w[2] = wptr<const char>(p, op.p1);
      for (int k = 0; k < 3; ++k) {
        c.has_bias[k] = (op.i0 >> k) & 1;
        c.n_src[k] = (op.i1 >> (8 * k)) & 15;
        c.n_shift[k] = (op.i1 >> (8 * k + 4)) & 15;
      }
      c.m_rows = op.m_rows; c.Lh = op.Lh_A; c.M_valid = op.M_valid;
      c.cond = as64(to_desc(p, op.add1)); c.scale1 = op.f0;
      if (c.cond.base && (c.cond.split != 1 || c.cond.nchunks != kCbPlanes || c.cond.esize_log2 != 1 || c.cond.Lh != x.Lh))
        return fail(UV_ERR_INVALID, "fused ConvBlock: condition must be a 64-column bf16 tensor of the block's geometry");
      c.film = aptr<const float>(p, op.film); c.film_C = op.C;
      c.slope2 = op.f1; c.slope3 = op.f2; c.scale3 = op.scale;
      c.n_out = op.n_out;
      for (int i = 0; i < 3; ++i) { c.out[i] = to_out(p, op.out[i]); c.out[i].t = as64(c.out[i].t); }
      if (op.i2) {
        if (op.n_out > 2) return fail(UV_ERR_INVALID, "fused ConvBlock: at most two block outputs next to a stage-1 output");
        c.cond_out = c.out[2];
        memset(&c.out[2], 0, sizeof(c.out[2]));
      }
      for (int i = 0; i < 3; ++i)
        if (c.out[i].t.base && c.out[i].t.nchunks != kCbPlanes) return fail(UV_ERR_INVALID, "fused ConvBlock: outputs must have 64 columns");
      c.reverse = op.i3 & 1;
      {
        static const bool bulk_on = !(getenv("UV_CB_BULK") && atoi(getenv("UV_CB_BULK")) == 0);
        const TensDesc& o = c.out[0].t;
        c.bulk_out = (bulk_on && c.n_out == 1 && o.base && o.split == 1 && o.nchunks == kCbPlanes && o.esize_log2 == 1 && o.Lh == x.Lh) ? 1 : 0;
      }
      c.mv = op.rows >= 0 ? aptr<const int>(p, op.rows) : nullptr;
      c.err_flag = p->err_flag;
      c.sched = p->sched + 2 * op_index;
      if (!cb_params_ok(c)) return fail(UV_ERR_INVALID, "fused ConvBlock: unsupported stage geometry");
      static const bool cb_clk = getenv("UV_CB_CLK") != nullptr;   // debug: where pair 0's MMA thread / epilogue warp spend their clocks
      if (cb_clk) {
        long long* d = nullptr;
        CUDA_TRY(cudaMalloc(&d, 16 * sizeof(long long)));
        CUDA_TRY(cudaMemset(d, 0, 16 * sizeof(long long)));
        c.clk = d;
        CUDA_TRY(convblock_pair_launch(c, p->model->num_sms, s));
        CUDA_TRY(cudaStreamSynchronize(s));
        long long h[16];
        CUDA_TRY(cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost));
        cudaFree(d);
        const double it = h[3] > 0 ? (double)h[3] : 1.0;
        fprintf(stderr, "[cb clk] items %lld (stage launches of pair 0): MMA thread per item: wait epi %.0f, wait xa %.0f, issue+commit %.0f, ring %.0f | "
                        "epilogue warp per item: wait mma %.0f, work s1 %.0f s2 %.0f s3 %.0f (per tile), ring %.0f\n",
                h[3], h[0] / it, h[1] / it, h[2] / it, h[4] / it, h[8] / it, 3 * h[9] / it, 3 * h[10] / it, 3 * h[11] / it, h[12] / it);
        break;
      }
      CUDA_TRY(convblock_pair_launch(c, p->model->num_sms, s));
      break;
    }
    case UV_OP_PREP:
      CUDA_TRY(launch_pdl(prep_kernel, dim3(op.B), dim3(1024), 0, s, p->io_dev, op.i1, op.T, op.Lp, op.i0, op.f0, aptr<float>(p, op.p0), aptr<float>(p, op.p1), rowgeo(p, op)));
      break;
    case UV_OP_MEL: {
      if (op.T != kMelFT) return fail(UV_ERR_INVALID, "MEL op built for another frame tile than the kernel's");
      const int nft = (op.F + kMelFT - 1) / kMelFT;
      const size_t sm = mel_smem_bytes(op.i0, op.i1);
      const int threads = std::max(2 * (((op.i2 + 1) / 2 + 31) / 32 * 32), (op.C + 31) / 32 * 32);   // (half of the frames, two bins) per thread; >= one thread per mel band
      // (the filterbank stage keeps two chunks of kMelFbRows x n_mels floats in the signal buffer, one band per thread;
      // the folded DFT needs n_fft / 2 to be a multiple of 4 and all n_fft / 2 + 1 bins)
      if (threads > 512 || sm > (size_t)kMelMaxSmem || op.i0 % 8 || op.i1 % 4 || op.C > threads || op.i2 != op.i0 / 2 + 1 ||
          2 * kMelFbRows * op.C > (kMelFT - 1) * op.i1 + op.i0)
        return fail(UV_ERR_INVALID, "MEL front end: n_fft / hop / n_mels unsupported");
      CUDA_TRY(launch_pdl(mel_kernel, dim3(nft, op.B), dim3(threads), sm, s, aptr<const float>(p, op.p0), op.Lp, op.F, op.i0, op.i1, op.i2, op.H, op.C,
                                                      op.i3, wptr<const float>(p, op.p3), wptr<const float>(p, op.p4),
                                                      aptr<float>(p, op.p1), aptr<float>(p, op.p2), nft, rowgeo(p, op)));
      break;
    }
    case UV_OP_SIGP: {
      const OutDesc o = to_out(p, op.out[0]);
      if (op.C % 8 || !o.t.base || o.t.split != 2 || o.t.nchunks * 8 < op.C || op.Lp < 2 * op.F * op.C)
        return fail(UV_ERR_INVALID, "SIGP: needs half-hop rows of a multiple of 8 samples and a phase-split-by-2 output");
      const int items = 2 * op.F * (op.C / 8);
      CUDA_TRY(launch_pdl(sig_planes_kernel, dim3((items + 255) / 256, op.B), dim3(256), 0, s, aptr<const float>(p, op.p0), op.Lp, op.F, op.C, o,
                          rowgeo(p, op)));
      break;
    }
    case UV_OP_MEL_POW: {
      if (op.T != kMelFT) return fail(UV_ERR_INVALID, "MEL_POW op built for another frame tile than the kernel's");
      const int nft = (op.F + kMelFT - 1) / kMelFT;
      const int threads = std::max(2 * (((op.i0 + 1) / 2 + 31) / 32 * 32), (op.C + 31) / 32 * 32);   // as the MEL op (same filterbank stage)
      const size_t sm = mel_pow_smem_bytes(op.i0);
      if (threads > 512 || sm > (size_t)kMelMaxSmem || op.C > threads || op.i1 < op.i0 || op.i2 < op.i1 + op.i0 || op.p4 < 0 || op.p5 < 0)
        return fail(UV_ERR_INVALID, "MEL_POW: bins / bands unsupported");
      CUDA_TRY(launch_pdl(mel_pow_kernel, dim3(nft, op.B), dim3(threads), sm, s, aptr<const float>(p, op.p0), op.i2, op.Lh_A, op.F, op.i0, op.i1, op.C,
                          wptr<const float>(p, op.p4), wptr<const int2>(p, op.p5), aptr<float>(p, op.p1), aptr<float>(p, op.p2), nft, rowgeo(p, op)));
      break;
    }
    case UV_OP_MEL_FIN: {
      const int items = op.F * (op.C / 8);
      CUDA_TRY(launch_pdl(mel_fin_kernel, dim3((items + 255) / 256, op.B), dim3(256), 0, s, aptr<const float>(p, op.p0), aptr<const float>(p, op.p1),
                                                                     op.F, op.C, op.i0, to_out(p, op.out[0]), rowgeo(p, op)));
      break;
    }
    case UV_OP_CONV_IN: {
      if (op.C > kEdgeMaxC || op.C % 8) return fail(UV_ERR_INVALID, "input conv: channel count unsupported");
      if (op.B > 65535) return fail(UV_ERR_INVALID, "input conv: more than 65535 utterances per lane");
      CUDA_TRY(launch_pdl(conv_in_kernel, dim3((unsigned)((op.Lp + 255) / 256), (unsigned)op.B), dim3(256), 0, s, aptr<const float>(p, op.p0), op.B, op.Lp, op.C, op.f0,
                                                                 wptr<const float>(p, op.p3), wptr<const float>(p, op.p4),
                                                                 op.n_out, to_out(p, op.out[0]), to_out(p, op.out[1]), op.i3 & 1, rowgeo(p, op)));
      break;
    }
    case UV_OP_INIT_X: {
      const long long n = (long long)op.B * op.Lp;
      const int B_total = op.i2 > 0 ? op.i2 : op.B;   // i1 / i2: first utterance of this lane / utterances of the whole call
      CUDA_TRY(launch_pdl(init_x_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, s, p->io_dev, n, op.i0, op.i1, B_total, op.Lp, op.f0,
                                                                aptr<float>(p, op.p0), rowgeo(p, op)));
      break;
    }
    case UV_OP_CONV_OUT: {
      if (op.C > kEdgeMaxC || op.C % 8) return fail(UV_ERR_INVALID, "output conv: channel count unsupported");
      if (op.B > 65535) return fail(UV_ERR_INVALID, "output conv: more than 65535 utterances per lane");
      // (a phase-split-by-2 input is walked in sample pairs: half the threads)
      const int per_row = (op.a.split == 2 && !(op.Lp & 1)) ? op.Lp / 2 : op.Lp;
      CUDA_TRY(launch_pdl(conv_out_kernel, dim3((unsigned)((per_row + 255) / 256), (unsigned)op.B), dim3(256), 0, s, to_desc(p, op.a), op.B, op.Lp, op.C, wptr<const float>(p, op.p3),
                                                                  wptr<const float>(p, op.p4), aptr<float>(p, op.p0),
                                                                  aptr<const StepCoef>(p, op.p1), op.i0, p->io_dev, op.i1,
                                                                  op.i2 > 0 ? op.i2 : op.B, op.i3 & 1, rowgeo(p, op)));
      break;
    }
    case UV_OP_AUX: {
      AuxParams a;
      a.a = to_desc(p, op.a);
      a.B = op.B; a.Lp = op.Lp; a.C = op.C;
      a.mode = op.i3; a.KU = op.H; a.KD = op.F;
      a.wts = wptr<const float>(p, op.p3);
      a.x = aptr<float>(p, op.p0);
      a.sigma = op.f0; a.step = op.i0;
      a.b_off = op.i1; a.B_total = op.i2 > 0 ? op.i2 : op.B;
      a.io = p->io_dev;
      a.rows = rowgeo(p, op);
      const size_t smem = aux_smem_bytes(a.C, a.KU, a.KD);
      if (a.C % 8 || a.C > kEdgeMaxC || a.a.split > 1 || a.mode < 0 || a.mode > 1 || a.KU > kAuxMaxKU || a.KD > kAuxMaxKD ||
          (a.mode == 0 && (a.KU < 1 || a.KD < 2)) || smem > 200 * 1024)
        return fail(UV_ERR_INVALID, "aux_to_wav: unsupported geometry");
      const int tiles = (a.Lp + kAuxTile - 1) / kAuxTile;
      aux_to_wav_kernel<<<dim3((unsigned)(a.B * tiles)), 256, smem, s>>>(a);
      CUDA_TRY(cudaGetLastError());
      break;
    }
    case UV_OP_POST:
      CUDA_TRY(launch_pdl(post_kernel, dim3(op.B), dim3(1024), 0, s, aptr<const float>(p, op.p0), op.T, op.Lp, op.i0, aptr<const float>(p, op.p1),
                                        aptr<const int>(p, op.p2), p->io_dev, op.i1, rowgeo(p, op)));
      break;
    case UV_OP_FILM:
      film_kernel<<<op.i0, 512, 2 * op.C * sizeof(float), s>>>(aptr<const float>(p, op.p0), op.i1, op.C, op.H, op.i2,
                                                               wptr<const float>(p, op.p3), wptr<const float>(p, op.p4),
                                                               wptr<const float>(p, op.p5), aptr<float>(p, op.p1));
      break;
    case UV_OP_GRU: {
      if (op.p5 >= 0) {
        // fp32 parity tier: f32 recurrence on row-major gate inputs (csrc/hp_ops.cuh)
        if (op.H > 384 || op.H % 32) return fail(UV_ERR_INVALID, "fp32 GRU: hidden size unsupported (needs H % 32 == 0, H <= 384)");
        HpGruParams h;
        memset(&h, 0, sizeof(h));
        h.gi = aptr<const float>(p, op.p0);
        h.whh = wptr<const float>(p, op.p5);
        h.bhh = wptr<const float>(p, op.p4);
        h.B = op.B; h.T = op.T; h.H = op.H; h.Lh = op.Lh_A;
        h.add1 = to_desc(p, op.add1); h.scale = op.scale; h.n_out = op.n_out;
        for (int i = 0; i < 3; ++i) h.out[i] = to_out(p, op.out[i]);
        if (h.add1.base && h.add1.split != 1) return fail(UV_ERR_INVALID, "GRU residual must be a plain tensor");
        for (int i = 0; i < h.n_out; ++i)
          if (h.out[i].t.split != 1) return fail(UV_ERR_INVALID, "GRU outputs must be plain tensors");
        h.rows = op.rows >= 0 ? aptr<const RowGeo>(p, op.rows) : nullptr;
        h.err_flag = p->err_flag;
        if (h.H % hp_gru_kch(h.H) || hp_gru_smem(h.H) > 200 * 1024) return fail(UV_ERR_INVALID, "fp32 GRU: hidden size unsupported");
        CUDA_TRY(gru_hp_launch(h, s));
        break;
      }
      if (op.H > kGruMaxH || op.H % (8 * kGruNC)) return fail(UV_ERR_INVALID, "GRU hidden size unsupported (needs H % 64 == 0, H <= 384)");
      GruParams g;
      memset(&g, 0, sizeof(g));
      g.gi = aptr<const float>(p, op.p0);
      g.whh = wptr<const char>(p, op.p3);
      g.bhh = wptr<const float>(p, op.p4);
      g.B = op.B; g.T = op.T; g.H = op.H; g.Lh = op.Lh_A;
      g.NB = op.i0; g.Bp = op.i1;
      if (!gru_config_ok(g.H, g.NB) || g.Bp % g.NB || g.Bp < g.B) return fail(UV_ERR_INVALID, "GRU batch geometry (NB, Bp) invalid");
      g.add1 = to_desc(p, op.add1); g.scale = op.scale; g.n_out = op.n_out;
      for (int i = 0; i < 3; ++i) g.out[i] = to_out(p, op.out[i]);
      if (g.add1.base && (g.add1.split != 1 || g.add1.esize_log2 != 1)) return fail(UV_ERR_INVALID, "GRU residual must be a plain bf16 tensor");
      for (int i = 0; i < g.n_out; ++i)
        if (g.out[i].t.split != 1) return fail(UV_ERR_INVALID, "GRU outputs must be plain tensors");
      g.rows = op.rows >= 0 ? aptr<const RowGeo>(p, op.rows) : nullptr;
      g.err_flag = p->err_flag;
      static const int gru_flags = getenv("UV_GRU_FLAGS") ? atoi(getenv("UV_GRU_FLAGS")) : 1;
      g.dbg_flags = gru_flags;
      static const bool clk_dbg = getenv("UV_GRU_CLK") != nullptr;   // debug: per-phase clocks of one CTA, printed to stderr
      if (clk_dbg) {
        long long* d = nullptr;
        CUDA_TRY(cudaMalloc(&d, 32 * sizeof(long long)));
        CUDA_TRY(cudaMemset(d, 0, 32 * sizeof(long long)));
        g.clk = d;
        CUDA_TRY(gru_launch(g, s));
        long long h[32];
        CUDA_TRY(cudaStreamSynchronize(s));
        CUDA_TRY(cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost));
        cudaFree(d);
        fprintf(stderr, "[gru clk/step] T=%d:", op.T);
        for (int i = 0; i < 10; ++i) fprintf(stderr, " p%d=%lld", i, h[i] / (op.T > 0 ? op.T : 1));
        fprintf(stderr, " | publish->slice group landed (MMA thread; last = committed):");
        for (int i = 16; i < 25; ++i) fprintf(stderr, " %lld", h[i] / (op.T > 1 ? op.T - 1 : 1));
        fprintf(stderr, "\n");
        break;
      }
      CUDA_TRY(gru_launch(g, s));
      break;
    }
    default:
      return fail(UV_ERR_INVALID, "unknown op type " + std::to_string(op.type));
  }
  CUDA_TRY(cudaGetLastError());
  return UV_OK;
}

static int validate_ops(uv_plan* p) {
  for (size_t i = 0; i < p->ops.size(); ++i) {
    const uv_op& op = p->ops[i];
    int rc;
    if ((rc = check_tens(p, op.a, "a"))) return rc;
    if ((rc = check_tens(p, op.add1, "add1"))) return rc;
    if ((rc = check_tens(p, op.add2, "add2"))) return rc;
    for (int k = 0; k < 3; ++k)
      if ((rc = check_tens(p, op.out[k].t, "out"))) return rc;
    if (op.type == UV_OP_GEMM) {
      if (op.a.off < 0 || op.KC % 16 || op.N_TILE % 16 || op.N_TILE > 256 || op.N_total % op.N_TILE ||
          (op.a.nch * 8) % op.KC || (op.C_out % op.N_TILE && (op.up <= 1 || op.N_TILE % op.C_out)) || op.smin < -kGuard || op.smax > kGuard || op.a.split != op.n_src)
        return fail(UV_ERR_INVALID, "malformed GEMM op #" + std::to_string(i));
      const int tile_rows = kTileM * gemm_msub(op.N_TILE);
      if (op.a.P < op.a.G + ((op.m_rows + tile_rows - 1) / tile_rows) * tile_rows + op.smax)
        return fail(UV_ERR_INVALID, "GEMM A plane too short, op #" + std::to_string(i));
    }
  }
  return UV_OK;
}

static int enqueue_range(uv_plan* p, int first, int last, cudaStream_t s, bool skip_setup) {
  for (int i = first; i < last; ++i) {
    if (skip_setup && p->ops[i].step == -2) continue;
    int rc = launch_op(p, i, s);
    if (rc) return rc;
  }
  return UV_OK;
}

// Graph capture of the whole program.  Ops of the two utterance lanes are independent (each lane owns its
// tensors): lane 1 is captured on a second stream forked from / joined to the capture stream, so its kernels
// run next to lane 0's -- in particular next to lane 0's GRU recurrence, which occupies 16 of the 148 SMs for
// hundreds of dependent steps.  An op flagged sync & 1 records its lane's event after the launch, an op
// flagged sync & 2 first waits for the other lane's event: the builder uses one such edge to start lane 1
// behind lane 0, so that the two lanes do not reach their recurrences together.
static int ensure_lane_objects(uv_plan* p, bool two) {
  if (!p->cap_stream) CUDA_TRY(cudaStreamCreateWithFlags(&p->cap_stream, cudaStreamNonBlocking));
  if (two && !p->lane_stream) {
    CUDA_TRY(cudaStreamCreateWithFlags(&p->lane_stream, cudaStreamNonBlocking));
    CUDA_TRY(cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming));
    for (int l = 0; l < 2; ++l) CUDA_TRY(cudaEventCreateWithFlags(&p->ev_lane[l], cudaEventDisableTiming));
  }
  return UV_OK;
}

static bool has_two_lanes(const uv_plan* p) {
  for (const uv_op& op : p->ops)
    if (op.lane == 1) return true;
  return false;
}

// Enqueues the program with lane 0 on s0 and lane 1 on s1 (fork at the start, join at the end, cross-lane edges
// as flagged).  Works both under stream capture and eagerly; `ev` (optional, 2 per op) brackets every launch with
// timing events on its own stream (uv_trace_ops).
static int enqueue_lanes(uv_plan* p, cudaStream_t s0, cudaStream_t s1, bool two, cudaEvent_t* ev) {
  const int n_ops = (int)p->ops.size();
  bool recorded[2] = {false, false};
  if (two) {
    CUDA_TRY(cudaEventRecord(p->ev_fork, s0));
    CUDA_TRY(cudaStreamWaitEvent(s1, p->ev_fork, 0));
  }
  for (int i = 0; i < n_ops; ++i) {
    const uv_op& op = p->ops[i];
    if (op.step == -2) continue;
    const int lane = two && op.lane == 1 ? 1 : 0;
    cudaStream_t s = lane ? s1 : s0;
    if (two && (op.sync & 2) && recorded[lane ^ 1]) CUDA_TRY(cudaStreamWaitEvent(s, p->ev_lane[lane ^ 1], 0));
    if (ev) CUDA_TRY(cudaEventRecord(ev[2 * i], s));
    int rc = launch_op(p, i, s);
    if (rc) return rc;
    if (ev) CUDA_TRY(cudaEventRecord(ev[2 * i + 1], s));
    if (two && (op.sync & 1)) {
      CUDA_TRY(cudaEventRecord(p->ev_lane[lane], s));
      recorded[lane] = true;
    }
  }
  if (two) {
    CUDA_TRY(cudaEventRecord(p->ev_join, s1));
    CUDA_TRY(cudaStreamWaitEvent(s0, p->ev_join, 0));
  }
  return UV_OK;
}

static int capture_program(uv_plan* p) {
  const bool two = has_two_lanes(p);
  int rc = ensure_lane_objects(p, two);
  if (rc) return rc;
  CUDA_TRY(cudaStreamBeginCapture(p->cap_stream, cudaStreamCaptureModeThreadLocal));
  rc = enqueue_lanes(p, p->cap_stream, p->lane_stream, two, nullptr);
  cudaError_t e = cudaStreamEndCapture(p->cap_stream, &p->graph);
  if (rc) return rc;
  if (e != cudaSuccess) return fail(UV_ERR_CUDA, std::string("graph capture: ") + cudaGetErrorString(e));
  CUDA_TRY(cudaGraphInstantiate(&p->exec, p->graph, 0));
  return UV_OK;
}

static int plan_build(uv_plan* p, const void* init, size_t nbytes_init, size_t init_off) {
  const int n_ops = (int)p->ops.size();
  CUDA_TRY(cudaMalloc(&p->arena, p->arena_bytes));
  CUDA_TRY(cudaMemset(p->arena, 0, p->arena_bytes));
  if (init && nbytes_init) {
    if (init_off + nbytes_init > p->arena_bytes) return fail(UV_ERR_INVALID, "init blob out of arena");
    CUDA_TRY(cudaMemcpy(p->arena + init_off, init, nbytes_init, cudaMemcpyHostToDevice));
  }
  CUDA_TRY(cudaMalloc(&p->io_dev, sizeof(IoTable)));
  p->coef_bytes = (size_t)p->n_steps * 16 + 16;
  p->stage_bytes = (sizeof(IoTable) + p->coef_bytes + 63) & ~(size_t)63;
  CUDA_TRY(cudaMallocHost(&p->stage_host, p->stage_bytes * uv_plan::kStage));
  for (int i = 0; i < uv_plan::kStage; ++i) CUDA_TRY(cudaEventCreateWithFlags(&p->stage_ev[i], cudaEventDisableTiming));
  CUDA_TRY(cudaHostAlloc(&p->err_host, sizeof(int), cudaHostAllocMapped));
  *p->err_host = 0;
  CUDA_TRY(cudaHostGetDevicePointer(&p->err_flag, p->err_host, 0));
  CUDA_TRY(cudaMalloc(&p->sched, (size_t)n_ops * 2 * sizeof(int)));
  CUDA_TRY(cudaMemset(p->sched, 0, (size_t)n_ops * 2 * sizeof(int)));
  int rc = validate_ops(p);
  if (rc) return rc;
  for (const uv_op& op : p->ops) {
    if (op.type == UV_OP_CONV_OUT) {
      bool seen = false;
      for (int64_t o : p->coef_offs) seen = seen || o == op.p1;
      if (!seen) p->coef_offs.push_back(op.p1);
    }
    if (op.step != -2) ++p->n_launches;
  }
  // plan-time setup ops (FiLM table)
  for (int i = 0; i < n_ops; ++i)
    if (p->ops[i].step == -2 && (rc = launch_op(p, i, 0))) return rc;
  CUDA_TRY(cudaDeviceSynchronize());
  if (p->use_graph && !p->lazy_graph && (rc = capture_program(p))) return rc;
  return UV_OK;
}

extern "C" int uv_plan_create(uv_plan** out, uv_model* m, const uv_op* ops, int n_ops, size_t arena_bytes,
                              const void* init, size_t nbytes_init, size_t init_off, int n_steps, int use_graph) {
  if (!out || !m || !ops || n_ops <= 0) return fail(UV_ERR_INVALID, "uv_plan_create: null argument");
  DeviceGuard guard(m->device);
  uv_plan* p = new uv_plan();
  p->model = m;
  p->ops.assign(ops, ops + n_ops);
  p->arena_bytes = arena_bytes;
  p->n_steps = n_steps;
  p->use_graph = use_graph != 0;
  p->lazy_graph = use_graph == 2;
  const int rc = plan_build(p, init, nbytes_init, init_off);
  if (rc) {   // every early return above lands here: nothing of a half-built plan is leaked
    const std::string msg = g_err;
    uv_plan_destroy(p);
    return fail(rc, msg);
  }
  *out = p;
  return UV_OK;
}

extern "C" void uv_plan_destroy(uv_plan* p) {
  if (!p) return;
  DeviceGuard guard(p->model->device);
  if (p->exec) cudaGraphExecDestroy(p->exec);
  if (p->graph) cudaGraphDestroy(p->graph);
  if (p->cap_stream) cudaStreamDestroy(p->cap_stream);
  if (p->lane_stream) cudaStreamDestroy(p->lane_stream);
  if (p->ev_fork) cudaEventDestroy(p->ev_fork);
  if (p->ev_join) cudaEventDestroy(p->ev_join);
  for (int l = 0; l < 2; ++l)
    if (p->ev_lane[l]) cudaEventDestroy(p->ev_lane[l]);
  for (int i = 0; i < uv_plan::kStage; ++i)
    if (p->stage_ev[i]) cudaEventDestroy(p->stage_ev[i]);
  cudaFree(p->sched);
  cudaFree(p->arena);
  cudaFree(p->io_dev);
  if (p->err_host) cudaFreeHost(p->err_host);
  if (p->stage_host) cudaFreeHost(p->stage_host);
  if (p->row_host) cudaFreeHost(p->row_host);
  delete p;
}

extern "C" size_t uv_plan_arena_bytes(const uv_plan* p) { return p ? p->arena_bytes : 0; }
extern "C" int uv_plan_num_launches(const uv_plan* p) { return p ? p->n_launches : 0; }

// A tripped device watchdog (mbar_wait -> trap) leaves its code in mapped host memory: visible without any
// device call, also after the trap has made the context unusable.
static int check_err_flag(uv_plan* p) {
  const int flag = *reinterpret_cast<volatile int*>(p->err_host);
  if (flag) return fail(UV_ERR_KERNEL, "device watchdog fired, code " + std::to_string(flag));
  return UV_OK;
}

// Stages the io table and the sampler coefficients of one call through the next slot of the pinned ring and
// enqueues their H2D copies on `s`.  The host only waits if that slot's copies of kStage calls ago are still
// pending, so batch k+1 can be queued behind batch k.
static int stage_io(uv_plan* p, const float* mix, const float* noise, float* out, float* dbg,
                    const uv_step_coef* coef, int keep_rms, cudaStream_t s) {
  if (!mix || !out || !coef) return fail(UV_ERR_INVALID, "uv_enhance: null pointer");
  if (!noise && !p->seeded) return fail(UV_ERR_INVALID, "uv_enhance: noise == NULL needs uv_set_noise_seed() first");
  int rc = check_err_flag(p);   // outcome of the previous calls
  if (rc) return rc;
  const int slot = p->stage_next;
  p->stage_next = (slot + 1) % uv_plan::kStage;
  if (p->stage_used[slot]) CUDA_TRY(cudaEventSynchronize(p->stage_ev[slot]));
  char* h = p->stage_host + (size_t)slot * p->stage_bytes;
  IoTable* io = reinterpret_cast<IoTable*>(h);
  io->mix = mix; io->noise = noise; io->out = out; io->dbg = dbg;
  io->seed = p->noise_seed; io->pad_ = 0;
  CUDA_TRY(cudaMemcpyAsync(p->io_dev, io, sizeof(IoTable), cudaMemcpyHostToDevice, s));
  if (p->row_dirty) {
    // new row lengths: rows past an utterance's end are never written by the kernels (they are the zero padding of
    // the convolutions), so what a longer utterance of an earlier call left there has to go first
    if (p->clear_bytes) CUDA_TRY(cudaMemsetAsync(p->arena + p->clear_off, 0, p->clear_bytes, s));
    char* rh = p->row_host + (size_t)slot * p->row_cap;
    memcpy(rh, p->row_blob.data(), p->row_blob.size());
    CUDA_TRY(cudaMemcpyAsync(p->arena + p->row_off, rh, p->row_blob.size(), cudaMemcpyHostToDevice, s));
    p->row_dirty = false;
  }
  if (!p->coef_offs.empty()) {
    char* c = h + sizeof(IoTable);
    memcpy(c, coef, (size_t)p->n_steps * 16);
    int kr[4] = {keep_rms, 0, 0, 0};
    memcpy(c + (size_t)p->n_steps * 16, kr, 16);
    for (int64_t off : p->coef_offs)
      CUDA_TRY(cudaMemcpyAsync(p->arena + off, c, p->coef_bytes, cudaMemcpyHostToDevice, s));
  }
  CUDA_TRY(cudaEventRecord(p->stage_ev[slot], s));
  p->stage_used[slot] = true;
  return UV_OK;
}

extern "C" int uv_enhance(uv_plan* p, const float* mix, const float* noise, float* out, float* dbg_net,
                          const uv_step_coef* coef, int keep_rms, void* stream) {
  if (!p) return fail(UV_ERR_INVALID, "uv_enhance: null plan");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  DeviceGuard guard(p->model->device);
  std::lock_guard<std::mutex> lock(p->mu);
  int rc = stage_io(p, mix, noise, out, dbg_net, coef, keep_rms, s);
  if (rc) return rc;
  const long long call = p->n_calls++;
  if (p->use_graph && !p->exec) {
    // lazy capture: a shape that is used once (a corpus where every file has its own length) never pays for the
    // capture + instantiation of several hundred kernel nodes; the first call runs the same two-lane schedule eagerly
    if (call == 0) {
      const bool two = has_two_lanes(p);
      if ((rc = ensure_lane_objects(p, two))) return rc;
      return enqueue_lanes(p, s, p->lane_stream, two, nullptr);
    }
    if ((rc = capture_program(p))) return rc;
  }
  if (p->use_graph) {
    cudaError_t e = cudaGraphLaunch(p->exec, s);
    if (e != cudaSuccess) {
      if (check_err_flag(p)) return UV_ERR_KERNEL;   // a sticky error left by a tripped watchdog
      return fail(UV_ERR_CUDA, std::string("cudaGraphLaunch: ") + cudaGetErrorString(e));
    }
  } else if ((rc = enqueue_range(p, 0, (int)p->ops.size(), s, true))) {
    return rc;
  }
  return UV_OK;
}

extern "C" int uv_set_row_tables(uv_plan* p, const void* tables, size_t nbytes, size_t arena_off, size_t clear_off, size_t clear_bytes) {
  if (!p || !tables || nbytes == 0) return fail(UV_ERR_INVALID, "uv_set_row_tables: null argument");
  if (arena_off + nbytes > p->arena_bytes || (arena_off & 15) || clear_off + clear_bytes > p->arena_bytes)
    return fail(UV_ERR_INVALID, "uv_set_row_tables: range outside the arena");
  if (arena_off < clear_off + clear_bytes && clear_off < arena_off + nbytes)
    return fail(UV_ERR_INVALID, "uv_set_row_tables: the tables lie inside the cleared range");
  DeviceGuard guard(p->model->device);
  std::lock_guard<std::mutex> lock(p->mu);
  if (nbytes > p->row_cap) {   // (first call, or a larger table: outside the steady state)
    if (p->row_host) {
      for (int i = 0; i < uv_plan::kStage; ++i)
        if (p->stage_used[i]) CUDA_TRY(cudaEventSynchronize(p->stage_ev[i]));
      cudaFreeHost(p->row_host);
      p->row_host = nullptr;
    }
    p->row_cap = (nbytes + 4095) & ~(size_t)4095;
    CUDA_TRY(cudaMallocHost(&p->row_host, p->row_cap * uv_plan::kStage));
  }
  p->row_blob.assign(static_cast<const char*>(tables), static_cast<const char*>(tables) + nbytes);
  p->row_off = arena_off; p->clear_off = clear_off; p->clear_bytes = clear_bytes;
  p->row_dirty = true;
  return UV_OK;
}

extern "C" int uv_set_noise_seed(uv_plan* p, unsigned long long seed) {
  if (!p) return fail(UV_ERR_INVALID, "uv_set_noise_seed: null plan");
  std::lock_guard<std::mutex> lock(p->mu);
  p->noise_seed = seed;
  p->seeded = true;
  return UV_OK;
}

extern "C" int uv_signal_median(const float* signal, int n_members, int batch, long long samples, float* out, int* scratch, void* stream) {
  if (!signal || !out || !scratch) return fail(UV_ERR_INVALID, "uv_signal_median: null pointer");
  if (n_members < 1 || n_members > kMedMaxE || batch < 1 || batch > 65535 || samples < 1)
    return fail(UV_ERR_INVALID, "uv_signal_median: 1..32 members, 1..65535 rows");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  CUDA_TRY(cudaMemsetAsync(scratch, 0, (size_t)batch * kMedMaxE * sizeof(int), s));
  long long nb = (samples + 255) / 256;
  if (nb > 1184) nb = 1184;   // 8 CTAs per SM
  signal_median_votes_kernel<<<dim3((unsigned)nb, (unsigned)batch), 256, 0, s>>>(signal, n_members, batch, samples, scratch);
  signal_median_gather_kernel<<<dim3((unsigned)nb, (unsigned)batch), 256, 0, s>>>(signal, n_members, batch, samples, scratch, out);
  CUDA_TRY(cudaGetLastError());
  return UV_OK;
}

extern "C" int uv_plan_status(uv_plan* p, void* stream) {
  if (!p) return fail(UV_ERR_INVALID, "uv_plan_status: null plan");
  DeviceGuard guard(p->model->device);
  cudaError_t e = cudaStreamSynchronize(reinterpret_cast<cudaStream_t>(stream));
  int rc = check_err_flag(p);
  if (rc) return rc;
  if (e != cudaSuccess) return fail(UV_ERR_CUDA, std::string("cudaStreamSynchronize: ") + cudaGetErrorString(e));
  return UV_OK;
}

extern "C" int uv_run_ops(uv_plan* p, int first, int last, const float* mix, const float* noise, float* out,
                          float* dbg_net, const uv_step_coef* coef, int keep_rms, void* stream) {
  if (!p || first < 0 || last > (int)p->ops.size()) return fail(UV_ERR_INVALID, "uv_run_ops: bad range");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  DeviceGuard guard(p->model->device);
  std::lock_guard<std::mutex> lock(p->mu);
  int rc = stage_io(p, mix, noise, out, dbg_net, coef, keep_rms, s);
  if (rc) return rc;
  if ((rc = enqueue_range(p, first, last, s, false))) return rc;
  CUDA_TRY(cudaStreamSynchronize(s));
  return check_err_flag(p);
}

extern "C" int uv_profile_ops(uv_plan* p, const float* mix, const float* noise, float* out, float* dbg_net,
                              const uv_step_coef* coef, int keep_rms, void* stream, float* ms_per_op) {
  if (!p || !ms_per_op) return fail(UV_ERR_INVALID, "uv_profile_ops: null argument");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  DeviceGuard guard(p->model->device);
  std::lock_guard<std::mutex> lock(p->mu);
  int rc = stage_io(p, mix, noise, out, dbg_net, coef, keep_rms, s);
  if (rc) return rc;
  const int n = (int)p->ops.size();
  std::vector<cudaEvent_t> ev(2 * n);
  for (auto& e : ev) CUDA_TRY(cudaEventCreate(&e));
  for (int i = 0; i < n; ++i) {
    if (p->ops[i].step == -2) continue;
    CUDA_TRY(cudaEventRecord(ev[2 * i], s));
    if ((rc = launch_op(p, i, s))) return rc;
    CUDA_TRY(cudaEventRecord(ev[2 * i + 1], s));
  }
  CUDA_TRY(cudaStreamSynchronize(s));
  for (int i = 0; i < n; ++i) {
    ms_per_op[i] = 0.f;
    if (p->ops[i].step != -2) CUDA_TRY(cudaEventElapsedTime(&ms_per_op[i], ev[2 * i], ev[2 * i + 1]));
  }
  for (auto& e : ev) cudaEventDestroy(e);
  return check_err_flag(p);
}

extern "C" int uv_trace_ops(uv_plan* p, const float* mix, const float* noise, float* out, float* dbg_net,
                            const uv_step_coef* coef, int keep_rms, float* t_start_ms, float* t_end_ms) {
  if (!p || !t_start_ms || !t_end_ms) return fail(UV_ERR_INVALID, "uv_trace_ops: null argument");
  DeviceGuard guard(p->model->device);
  std::lock_guard<std::mutex> lock(p->mu);
  const bool two = has_two_lanes(p);
  int rc = ensure_lane_objects(p, two);
  if (rc) return rc;
  cudaStream_t s0 = p->cap_stream;
  if ((rc = stage_io(p, mix, noise, out, dbg_net, coef, keep_rms, s0))) return rc;
  const int n = (int)p->ops.size();
  std::vector<cudaEvent_t> ev(2 * n);
  for (auto& e : ev) CUDA_TRY(cudaEventCreate(&e));
  cudaEvent_t base;
  CUDA_TRY(cudaEventCreate(&base));
  CUDA_TRY(cudaEventRecord(base, s0));
  if ((rc = enqueue_lanes(p, s0, p->lane_stream, two, ev.data()))) return rc;
  CUDA_TRY(cudaStreamSynchronize(s0));
  for (int i = 0; i < n; ++i) {
    t_start_ms[i] = t_end_ms[i] = 0.f;
    if (p->ops[i].step == -2) continue;
    CUDA_TRY(cudaEventElapsedTime(&t_start_ms[i], base, ev[2 * i]));
    CUDA_TRY(cudaEventElapsedTime(&t_end_ms[i], base, ev[2 * i + 1]));
  }
  for (auto& e : ev) cudaEventDestroy(e);
  cudaEventDestroy(base);
  return check_err_flag(p);
}

extern "C" int uv_arena_read(uv_plan* p, size_t off, void* host_dst, size_t nbytes) {
  if (!p || off + nbytes > p->arena_bytes) return fail(UV_ERR_INVALID, "uv_arena_read: out of range");
  DeviceGuard guard(p->model->device);
  CUDA_TRY(cudaDeviceSynchronize());
  CUDA_TRY(cudaMemcpy(host_dst, p->arena + off, nbytes, cudaMemcpyDeviceToHost));
  return UV_OK;
}

extern "C" void* uv_plan_arena_ptr(uv_plan* p) { return p ? (void*)p->arena : nullptr; }

extern "C" int uv_arena_write(uv_plan* p, size_t off, const void* host_src, size_t nbytes) {
  if (!p || off + nbytes > p->arena_bytes) return fail(UV_ERR_INVALID, "uv_arena_write: out of range");
  DeviceGuard guard(p->model->device);
  CUDA_TRY(cudaMemcpy(p->arena + off, host_src, nbytes, cudaMemcpyHostToDevice));
  return UV_OK;
}

// ------------------------------------------------------------------ allocation-free mel front end
// uv_mel_spec_static (include/universe_b200.h): the mel front end of the conditioner for callers that cannot
// allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  The scratch the two kernels exchange (raw mel bands,
// partial energy sums) lives in a file-scope device array of UV_STATIC_SCRATCH_BYTES that every GPU which loads
// the library holds until exit.  In host code `g_static_scratch` is only a stand-in for the symbol: the device
// address is fetched with cudaGetSymbolAddress once per device, with that device current, and handed to the kernels.
static __device__ __align__(256) float g_static_scratch[UV_STATIC_SCRATCH_BYTES / sizeof(float)];

static std::mutex g_static_mu;
static float* g_static_ptr[64] = {nullptr};   // per device

static int static_scratch_ptr(int device, float** out) {
  if (device < 0 || device >= 64) return fail(UV_ERR_INVALID, "uv_mel_spec_static: device index out of range");
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ptr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);   // the current device is `device` (DeviceGuard of the caller)
    if (e != cudaSuccess || !p)
      return fail(UV_ERR_CUDA, std::string("cudaGetSymbolAddress(g_static_scratch): ") + cudaGetErrorString(e));
    g_static_ptr[device] = static_cast<float*>(p);
  }
  *out = g_static_ptr[device];
  return UV_OK;
}

// Variant of mel_fin_kernel: same normalisation (condition.py:105-106), plain f32 output [B][n_mels][F] instead of
// bf16 planes.  One thread per (band, frame), frame fastest: coalesced stores.
__global__ void __launch_bounds__(256) mel_fin_plain_kernel(const float* __restrict__ melraw, const float* __restrict__ energy,
                                                            int F, int n_mels, int nft, float* __restrict__ out) {
  __shared__ float s_inv;
  const int b = blockIdx.y;
  if (threadIdx.x == 0) {
    float e = 0.f;
    for (int i = 0; i < nft; ++i) e += energy[(size_t)b * nft + i];
    s_inv = 1.f / fmaxf(sqrtf(e / (float)F), 1e-5f);
  }
  __syncthreads();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= F * n_mels) return;
  const int m = idx / F, f = idx - m * F;
  out[((size_t)b * n_mels + m) * F + f] = melraw[((size_t)b * F + f) * n_mels + m] * s_inv;
}

extern "C" int uv_mel_spec_static(uv_model* m, const uv_mel_desc* d, const float* sig, int B, int Lp, float* mel_out,
                                  void* stream) {
  if (!m || !d || !sig || !mel_out) return fail(UV_ERR_INVALID, "uv_mel_spec_static: null argument");
  if (B <= 0 || Lp <= 0 || d->hop <= 0 || Lp % d->hop) return fail(UV_ERR_INVALID, "uv_mel_spec_static: Lp must be a positive multiple of hop");
  const int F = Lp / d->hop;
  const int nft = (F + kMelFT - 1) / kMelFT;
  const int threads = std::max(2 * (((d->n_bins + 1) / 2 + 31) / 32 * 32), (d->n_mels + 31) / 32 * 32);
  const size_t sm = mel_smem_bytes(d->n_fft, d->hop);
  if (threads > 512 || sm > (size_t)kMelMaxSmem || d->n_fft % 8 || d->hop % 4 || d->n_mels > threads || d->n_mels <= 0 ||
      d->n_bins != d->n_fft / 2 + 1 ||
      d->nbp < d->n_bins || 2 * kMelFbRows * d->n_mels > (kMelFT - 1) * d->hop + d->n_fft)
    return fail(UV_ERR_INVALID, "uv_mel_spec_static: n_fft / hop / n_mels unsupported");
  const size_t need = ((size_t)B * F * d->n_mels + (size_t)B * nft) * sizeof(float);
  if (need > (size_t)UV_STATIC_SCRATCH_BYTES)
    return fail(UV_ERR_INVALID, "uv_mel_spec_static: B x frames x n_mels needs " + std::to_string(need) +
                                    " bytes of scratch, the static array holds " + std::to_string((size_t)UV_STATIC_SCRATCH_BYTES));
  const size_t basis_bytes = (size_t)d->n_fft * 2 * d->nbp * 4, fb_bytes = (size_t)d->n_bins * d->n_mels * 4;
  if (d->basis_w < 0 || d->fb_w < 0 || (size_t)d->basis_w + basis_bytes > m->nbytes || (size_t)d->fb_w + fb_bytes > m->nbytes)
    return fail(UV_ERR_INVALID, "uv_mel_spec_static: weight offsets outside the model blob");
  DeviceGuard guard(m->device);
  float* scratch = nullptr;
  int rc = static_scratch_ptr(m->device, &scratch);
  if (rc) return rc;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  float* melraw = scratch;                                  // [B][F][n_mels], fully rewritten by every call
  float* energy = scratch + (size_t)B * F * d->n_mels;      // [B][nft], fully rewritten by every call
  mel_kernel<<<dim3(nft, B), dim3(threads), sm, s>>>(sig, Lp, F, d->n_fft, d->hop, d->n_bins, d->nbp, d->n_mels,
                                                     (d->n_fft - d->hop) / 2,
                                                     reinterpret_cast<const float*>(m->blob + d->basis_w),
                                                     reinterpret_cast<const float*>(m->blob + d->fb_w), melraw, energy, nft, nullptr);
  CUDA_TRY(cudaGetLastError());
  mel_fin_plain_kernel<<<dim3((F * d->n_mels + 255) / 256, B), dim3(256), 0, s>>>(melraw, energy, F, d->n_mels, nft, mel_out);
  CUDA_TRY(cudaGetLastError());
  return UV_OK;
}

/* universe_b200.h -- C ABI of the B200-native UNIVERSE / UNIVERSE++ enhance() path.
 *
 * The reference (kolyangg/open-universe) is 100 % Python and has no FFI of its
 * own: its boundary for this path is
 *     open_universe/inference_utils/model_loader.py:65-140   load_model()
 *     open_universe/networks/universe/universe.py:231-375    Universe.enhance()
 * which today bottom out in torch / cuDNN library calls.  This header is the
 * native boundary that replaces those library calls: plain pointers and sizes,
 * no torch types.  The Python drop-in (open_universe_b200/) binds it with
 * ctypes; INTEGRATION.md shows the stub a reference maintainer would add.
 *
 * Model = a read-only weight blob (folded + packed by the host compiler).
 * Plan  = an "op program" for one (batch, length, n_steps) shape: a flat array
 *         of uv_op records, each one kernel launch of the hot path, plus the
 *         workspace ("arena") size.  uv_plan_create() captures the whole program
 *         (conditioner once + n_steps score evaluations + sampler updates) into
 *         ONE CUDA graph; uv_enhance() replays it.
 *
 * All functions return 0 on success, a negative uv_status otherwise;
 * uv_last_error() gives a human readable message (thread local).
 * No exceptions cross the boundary.  Calls on one plan are serialised by the
 * library (a plan owns one arena and one io table); use one plan per stream /
 * thread for concurrency.  All device work is enqueued on the cudaStream_t
 * passed in (as void*; NULL = legacy default stream); uv_enhance() does not wait
 * for the device: batch k+1 can be queued behind batch k.  Every entry point
 * selects the handle's device and restores the caller's current device.
 */
#ifndef UNIVERSE_B200_H_
#define UNIVERSE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UV_ABI_VERSION 10

typedef enum {
  UV_OK = 0,
  UV_ERR_INVALID = -1,   /* bad argument / malformed program            */
  UV_ERR_CUDA = -2,      /* CUDA runtime error (message has the detail) */
  UV_ERR_NO_DEVICE = -3, /* no sm_100 device                            */
  UV_ERR_KERNEL = -4     /* device-side watchdog fired                  */
} uv_status;

/* ---- tensor addressing ("planes" layout, see csrc/uv_common.cuh) ---------
 * element (utterance b, sample l, channel c) lives at
 *   off + (((l % split) * nch + c / 8) * P + G + b * Lh + l / split) * 8 + c % 8   [elements]
 */
typedef struct {
  int64_t off;    /* byte offset into the arena; < 0 => unused            */
  int32_t dtype;  /* 1 = bf16, 2 = f32                                    */
  int32_t nch;    /* channels / 8                                         */
  int32_t split;  /* phase split r (1 = plain)                            */
  int32_t Lh;     /* utterance pitch (rows)                               */
  int32_t P;      /* plane pitch (rows)                                   */
  int32_t G;      /* guard rows at the start of each plane                */
} uv_tens;

typedef struct {
  uv_tens t;
  int32_t act;    /* 0 none, 1 prelu(a1), 2 prelu(prelu(., a1), a2)       */
  float a1, a2;
  int32_t pad_;
} uv_out;

typedef enum {
  UV_OP_PREP = 1,     /* pad + per-utterance normalise   (universe.py:219-223, utils/norm.py:47-87)   */
  UV_OP_MEL = 2,      /* framed DFT power -> mel         (condition.py:92-108)                        */
  UV_OP_MEL_FIN = 3,  /* global mel normalisation -> bf16 planes                                      */
  UV_OP_CONV_IN = 4,  /* 1 -> C k3 input conv on a signal (score.py:285, condition.py:360)            */
  UV_OP_GEMM = 5,     /* multi-tap tensor-core conv GEMM (blocks.py:331-416, every dense layer)       */
  UV_OP_GRU = 6,      /* bidirectional GRU recurrence    (score.py:117, condition.py:212)             */
  UV_OP_CONV_OUT = 7, /* C -> 1 k3 output conv + EDM wrapper + sampler update (universe.py:197-209,334-343) */
  UV_OP_INIT_X = 8,   /* x = sigma_0 * z_0               (universe.py:326)                            */
  UV_OP_POST = 9,     /* crop, keep_rms, peak normalise  (universe.py:349-357)                        */
  UV_OP_FILM = 10,    /* noise-level embedding + FiLM table (sigma_block.py, score.py:109,208)        */
  UV_OP_CONVBLOCK = 11, /* fused ConvBlock of a 64-column level (C = 64, or C = 32 stored phase-split by 2): PReLU -> conv1 ->
                        * [+cond] * f0 -> [FiLM] -> PReLU -> conv2 -> PReLU -> conv3 -> (. + x) * scale  (blocks.py:387-403).
                        * a = x (raw), a_slope = slope in front of conv1, w_w / p0 / p1 = the three convs' packed weights
                        * (i0 bit s: conv s has a bias; i1 byte s: n_src | n_shift << 4 of conv s), add1 = condition,
                        * film (C = its channel count), f1 / f2 = slopes in front of conv2 / conv3, out[0..n_out) = block
                        * outputs, i2 = 1: out[2] is the stage-1 output (the conditioner's conditioning signal)          */
  UV_OP_AUX = 12,      /* aux_to_wav fused with the state initialisation of use_aux_signal / warm_start (universe_gan.py:
                        * 119-128,147-151; bigvgan/alias_free_act.py:8-30, snake.py:53-64; universe.py:317-331):
                        * a = y_hat (bf16 planes, C channels), p0 = x [B][Lp], p3 = weights f32 (alpha or slope [C] |
                        * up kernel [2][H] | down kernel [F] | conv [C][3] | bias), i3 = 0 snake / 1 PReLU, f0 = sigma_n
                        * (0: x = aux_to_wav(y_hat)), i0 = n (the draw added: x += sigma_n z_n), i1 / i2 as INIT_X   */
  UV_OP_SIGP = 13,     /* padded signal p0 f32 [B][Lp] -> out[0]: f32 planes tensor in rows of C = hop / 2 samples, phase-split by 2
                        * (row m of phase s = samples [(2m + s) C, (2m + s + 1) C)), F frames: the A operand of the framed DFT of
                        * the mel front end run as ONE multi-tap GEMM on the tensor cores (condition.py:92-108; a UV_OP_GEMM on
                        * f32 planes with a row-major output, see the fp32 notes below) */
  UV_OP_MEL_POW = 14   /* p0 = that GEMM's output f32 [b Lh_A + f][i2] (cos bins at column 0, sin bins at column i1, i0 bins)
                        * -> power spectrum -> mel filterbank p4 (f32 [i0][C]; p5 = int32 [C][2]: the bins [lo, hi) of each band) -> p1 / p2 as UV_OP_MEL (raw mel bands, partial
                        * energy sums per T frames) */
} uv_op_type;

/* One record per kernel launch.  Offsets named *_w index the model's weight
 * blob, all others index the plan's arena; < 0 means "absent".            */
typedef struct {
  int32_t type;
  int32_t step;          /* sampler step this op belongs to (-1 = conditioner / setup) */
  /* ---- UV_OP_GEMM (also GRU epilogue fields: add1, scale, n_out, out) */
  uv_tens a;             /* A operand (bf16); its phases are the K sources */
  int32_t n_src, smin, smax, KC;
  int64_t w_w;           /* packed weight images                           */
  int32_t N_TILE, N_total;
  int32_t m_rows, Lh_A, M_valid, up, C_out;
  int64_t bias_w;        /* f32 [C_out]                                    */
  int64_t film;          /* arena: f32 [2*C_out] gamma|beta                */
  uv_tens add1, add2;
  float scale;
  int32_t n_out;
  uv_out out[3];
  int64_t rm_out;        /* arena: f32 row-major [m_rows][rm_ld]           */
  int32_t rm_ld;
  /* ---- generic fields used by the small ops (meaning per type: launch_op() in open_universe_b200/csrc/capi.cu) */
  int32_t B, T, Lp, C, H, F;
  int32_t i0, i1, i2, i3;
  int64_t p0, p1, p2, p3, p4, p5;
  float f0, f1, f2, f3;
  /* ---- utterance lanes: a program may process two independent halves of the batch ("lanes", each with its
   * own tensors); uv_plan_create captures lane 1 on a second stream so that its kernels overlap lane 0's
   * GRU recurrence.  PREP / INIT_X / CONV_OUT / POST carry i1 = first utterance of their lane and
   * i2 = utterances of the whole call (the io tensors of uv_enhance always cover all lanes).            */
  int32_t lane;          /* 0 or 1                                                                */
  int32_t sync;          /* bit 0: record this lane's event after the op; bit 1: wait for the other lane's event first */
  /* ---- UV_OP_GEMM: A stored raw.  a_act = 1: the kernel applies PReLU(a_slope) to A in shared memory, so the
   * producer of A writes one tensor instead of a raw and an activated copy (blocks.py:209-231: PReLU -> conv). */
  int32_t a_act;
  float a_slope;
  /* ---- variable-length calls (uv_set_row_tables): arena offset of this op's per-utterance table, < 0 = every
   * utterance has the plan's geometry.  GEMM / CONVBLOCK: int32 [B] valid rows of each utterance in the op's row
   * space (replaces M_valid); every other op: uv_row_geo [B].                                                    */
  int64_t rows;
  /* ---- fp32 parity tier (north_star: rel-L2 <= 1e-4 with fp32 operands; the reference's own arithmetic is fp32,
   * bin/enhance.py:83-192 sets no precision flag).  The precision of a plan is a property of its program: every
   * planes tensor has dtype 2 (f32), a UV_OP_GEMM whose A is f32 names in p5 the fp32 image of the same multi-tap GEMM
   * (f32 [n_src][n_shift][C_in][N_total]; bias_w = f32 [C_out]) and writes an optional row-major output
   * [m_rows][rm_ld] (H = 0), a UV_OP_GRU with p5 >= 0 (f32 W_hh [dir][k < H][gate 3][u < H]) reads row-major gate inputs
   * [b Lh + t][6H] at p0 (csrc/hp_ops.cuh).  The GEMM runs on the tensor cores as a 3-term split-bf16 product (p4 = images
   * [N tile][src][K group][shift][hi | lo][i0 / 8][i2][8] bf16, i0 = K group, i2 = N tile; ~1e-5 per layer) or, with
   * UV_HP_SIMT=1 / when p4 < 0, as an fp32 FMA kernel from p5; the recurrence is an fp32 FMA kernel.  UV_OP_CONVBLOCK is not used by such programs.
   * The host compiler builds them with build_program(..., precision="fp32") (Universe.precision = "fp32").        */
  /* UV_OP_GEMM / CONV_IN / CONV_OUT: i3 bit 0 = walk the tensor from its end (tiles / blocks in descending order).
   * The builder alternates it along each chain of kernels so that a consumer starts where its producer finished,
   * i.e. on the part of the tensor the L2 still holds.                                                          */
} uv_op;

/* Geometry of one utterance of a variable-length call: the reference enhances file by file (bin/enhance.py:168-192),
 * each file with its own padding (universe.py:219-223), normalisation statistics, mel frames and GRU length. */
typedef struct {
  int32_t T;        /* valid samples of the row                      */
  int32_t pad_left; /* (tot_ds - T % tot_ds) / 2                     */
  int32_t Lp;       /* T + tot_ds - T % tot_ds (<= the plan's Lp)    */
  int32_t F;        /* Lp / tot_ds: mel frames = GRU steps           */
} uv_row_geo;

typedef struct uv_model uv_model;
typedef struct uv_plan uv_plan;

/* Sampler coefficients of one step (host computed from sigma, eta, beta and the
 * EDM weights; universe.py:175-209,301-343):  x <- a*x + b*net + c*z          */
typedef struct {
  float a, b, c, pad_;
} uv_step_coef;

int uv_abi_version(void);
size_t uv_sizeof_op(void); /* sizeof(uv_op), for binding self-checks */
const char* uv_last_error(void);

/* Uploads the packed weight blob to `device`. */
int uv_model_create(uv_model** out, const void* weights, size_t nbytes, int device);
void uv_model_destroy(uv_model* m);

/* Builds a plan: allocates + zeroes the arena, copies `init` (nbytes_init bytes)
 * to arena offset init_off, runs the ops flagged step == -2 once (plan-time
 * setup such as the FiLM table) and captures the remaining ops in a CUDA graph.
 * use_graph: 0 = eager launches, 1 = capture now, 2 = capture lazily (the first uv_enhance() of the plan runs
 * eagerly, the second captures: a shape that is used once never pays for the capture). */
int uv_plan_create(uv_plan** out, uv_model* m, const uv_op* ops, int n_ops, size_t arena_bytes,
                   const void* init, size_t nbytes_init, size_t init_off, int n_steps, int use_graph);
void uv_plan_destroy(uv_plan* p);
size_t uv_plan_arena_bytes(const uv_plan* p);

/* One enhance() pass.  Device pointers: mix [B][T] f32, noise [n_steps][B][Lp]
 * f32, out [B][T] f32, dbg_net (optional) [n_steps][B][Lp] f32 per-step network
 * outputs.  coef: host array of n_steps uv_step_coef. keep_rms as in enhance().
 * noise = the sampler's draws in the reference's order (randn, universe.py:39-41, called :326,338): parity mode,
 * the caller draws them with torch's generator.  noise == NULL (after uv_set_noise_seed): throughput mode, the
 * kernels draw z[step][utterance][sample] themselves (Philox4x32-10 keyed by the seed, Box-Muller; a pure function
 * of (seed, step, utterance, sample), restated on the CPU in oracle/philox.py): same distribution, another stream. */
int uv_enhance(uv_plan* p, const float* mix, const float* noise, float* out, float* dbg_net,
               const uv_step_coef* coef, int keep_rms, void* stream);

/* Variable-length batches: the rows of the next calls have their own lengths.  `tables` (host, nbytes) is the block
 * of per-utterance tables the ops' `rows` offsets point into, to be placed at arena_off; [clear_off, clear_off +
 * clear_bytes) = the activation planes, cleared once in front of the next call (rows past an utterance's end are the
 * zero padding of the convolutions and are never written).  Both happen on the stream of the next uv_enhance /
 * uv_run_ops; call again only when the lengths change.  The host side (open_universe_b200/program.py row_tables())
 * builds the block. */
int uv_set_row_tables(uv_plan* p, const void* tables, size_t nbytes, size_t arena_off, size_t clear_off,
                      size_t clear_bytes);

/* Key of the in-kernel draws of the following calls made with noise == NULL (uv_enhance, uv_run_ops, ...). */
int uv_set_noise_seed(uv_plan* p, unsigned long long seed);

/* Waits for `stream` and reports what the calls queued so far left behind: UV_ERR_KERNEL if a device-side
 * watchdog fired (a pipeline barrier that never completed; the code is in uv_last_error()), UV_ERR_CUDA for
 * a CUDA error, UV_OK otherwise.  uv_enhance() itself checks the flag of the PREVIOUS calls on entry. */
int uv_plan_status(uv_plan* p, void* stream);

/* Debug / parity: run ops [first, last) eagerly (no graph) on `stream` with the
 * same io pointers, and copy a byte range of the arena back / forth.       */
int uv_run_ops(uv_plan* p, int first, int last, const float* mix, const float* noise, float* out,
               float* dbg_net, const uv_step_coef* coef, int keep_rms, void* stream);
int uv_arena_read(uv_plan* p, size_t off, void* host_dst, size_t nbytes);
/* Device address of the plan's arena (for zero-copy views of intermediate tensors, e.g. the conditioner's
 * y_hat consumed by aux_to_wav, reference universe_gan.py:147-151; the caller must not free it).          */
void* uv_plan_arena_ptr(uv_plan* p);
int uv_arena_write(uv_plan* p, size_t off, const void* host_src, size_t nbytes);

/* Measurement: runs the program eagerly with a cudaEvent pair around every launch (on `stream`)
 * and returns the device time of each op in ms_per_op[n_ops] (0 for plan-time setup ops).  */
int uv_profile_ops(uv_plan* p, const float* mix, const float* noise, float* out, float* dbg_net,
                   const uv_step_coef* coef, int keep_rms, void* stream, float* ms_per_op);

/* Measurement: runs the program eagerly the way the graph runs it (lane 0 and lane 1 on two internal streams,
 * same fork / join / stagger edges) with a cudaEvent pair around every launch, and returns each op's start and
 * end time in ms relative to the start of the pass: the timeline that shows what overlaps the GRU recurrences. */
int uv_trace_ops(uv_plan* p, const float* mix, const float* noise, float* out, float* dbg_net,
                 const uv_step_coef* coef, int keep_rms, float* t_start_ms, float* t_end_ms);

/* Ensemble statistic "signal_median" (reference utils/stats.py:22-66, used by universe.py:359-367): signal
 * [n_members][batch][samples] f32 -> out [batch][samples]: the member whose index equals the sorted position that
 * most often holds the member closest to n_members / 2.  Device pointers; scratch = batch * 32 ints (the call
 * clears it); n_members <= 32; allocates nothing. */
int uv_signal_median(const float* signal, int n_members, int batch, long long samples, float* out, int* scratch,
                     void* stream);

/* Number of kernel launches one uv_enhance() replays (for bench accounting). */
int uv_plan_num_launches(const uv_plan* p);

/* ---- allocation-free entry point -------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 *
 * The conditioner's mel front end (MelAdapter.compute_mel_spec: power mel spectrogram, then division by
 * sqrt(mean_t sum_m S^2)) on a padded, normalised signal, with the scratch the two kernels exchange (raw mel
 * bands + partial energy sums: (B * F * n_mels + B * ceil(F / 32)) floats, F = Lp / hop) in a file-scope static
 * device array of UV_STATIC_SCRATCH_BYTES.  Every GPU that loads the library holds that array until exit.
 * The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine); a problem whose scratch
 * exceeds the array is rejected with UV_ERR_INVALID.  Nothing in the array has to start at zero (both regions are
 * rewritten in full by every call), but ALL calls on a device share it: calls that overlap on the GPU (different
 * streams, or a second call before the first has finished on another stream) race; serialise them.
 * sig: device f32 [B][Lp] (Lp a multiple of hop), mel_out: device f32 [B][n_mels][Lp / hop]; weights come from
 * the model blob (offsets as in the packed layer "c.mel").                                                  */
#define UV_STATIC_SCRATCH_BYTES (16u << 20)
typedef struct {
  int32_t n_fft, hop, n_bins, nbp; /* window length, hop, n_fft / 2 + 1, bins padded to a multiple of 32 */
  int32_t n_mels, pad_;
  int64_t basis_w;                 /* model blob: f32 [n_fft][2 * nbp] windowed DFT basis (cos | sin)   */
  int64_t fb_w;                    /* model blob: f32 [n_bins][n_mels] mel filterbank                   */
} uv_mel_desc;
int uv_mel_spec_static(uv_model* m, const uv_mel_desc* d, const float* sig, int B, int Lp, float* mel_out, void* stream);

/* ---- batch WAV IO of the file pipeline (host side; replaces the per-file torchaudio.load / torchaudio.save of
 * open_universe/bin/enhance.py:183,192 for RIFF/WAVE files).  A window of files is decoded / encoded by `threads` host
 * threads without the interpreter lock, into / out of ONE caller-owned float32 staging buffer (pinned, so that the
 * window crosses PCIe as a single copy).  File i occupies [offs[i], offs[i] + channels * frames) floats, planar
 * [channels][frames], scaled like torchaudio.load: int16 / 2^15, int32 / 2^31, (uint8 - 128) / 128, float32 as is.
 * uv_wav_probe fills info[i] (frames = -1: not a PCM 8/16/32-bit or float32 RIFF/WAVE file -- the caller falls back
 * to its own decoder for that file); all three return the number of files that failed (status[i] != 0), < 0 on a
 * bad argument.  uv_wav_write writes IEEE-float WAV files (what torchaudio.save writes for a float tensor).     */
typedef struct {
  int32_t channels, fs;
  int32_t format;   /* 1 = integer PCM, 3 = IEEE float */
  int32_t bits;
  int64_t frames;   /* < 0: unsupported / unreadable */
  int64_t data_off; /* byte offset of the samples in the file */
} uv_wav_info;
int uv_wav_probe(const char* const* paths, int n, uv_wav_info* info, int threads);
int uv_wav_read(const char* const* paths, int n, const uv_wav_info* info, const int64_t* offs, float* dst, int threads,
                int32_t* status);
int uv_wav_write(const char* const* paths, int n, const int32_t* channels, const int64_t* frames, const int32_t* fs,
                 const int64_t* offs, const float* src, int threads, int32_t* status);

#ifdef __cplusplus
}
#endif
#endif /* UNIVERSE_B200_H_ */

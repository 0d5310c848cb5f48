// Native batch WAV IO of the file pipeline (host code only; declared in include/universe_b200.h).
//
// The reference reads and writes one file per enhance() call with torchaudio (open_universe/bin/enhance.py:183,192).
// At ~6000 audio-seconds per second a GPU goes through ~1500 four-second files per second, and the Python side of
// file IO (scipy's pure-Python RIFF parser, one tensor and one copy per file, all under the GIL) took longer per
// 64-file window than the window's batch on the device.  These entry points decode / encode a whole window with a
// few host threads, without the GIL, straight into / out of ONE (pinned) staging buffer that goes to the device as
// a single copy.  Planar float32 [channels][frames] per file, scaled like torchaudio.load(normalize=True).
#pragma once
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/universe_b200.h"

namespace uvwav {

static inline uint32_t rd32(const unsigned char* p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
static inline uint16_t rd16(const unsigned char* p) { return (uint16_t)(p[0] | (p[1] << 8)); }

template <typename F>
static void parallel_for(int n, int threads, F f) {
  if (threads < 1) threads = 1;
  if (threads > n) threads = n;
  if (threads <= 1) {
    for (int i = 0; i < n; ++i) f(i);
    return;
  }
  std::atomic<int> next(0);
  std::vector<std::thread> pool;
  pool.reserve(threads);
  for (int t = 0; t < threads; ++t)
    pool.emplace_back([&]() {
      for (;;) {
        const int i = next.fetch_add(1);
        if (i >= n) break;
        f(i);
      }
    });
  for (auto& t : pool) t.join();
}

// Walks the RIFF chunks up to 'data'.  Returns false (info.frames = -1) for anything this reader does not decode.
static bool probe_one(const char* path, uv_wav_info* info) {
  memset(info, 0, sizeof(*info));
  info->frames = -1;
  FILE* f = fopen(path, "rb");
  if (!f) return false;
  unsigned char h[12];
  bool ok = fread(h, 1, 12, f) == 12 && !memcmp(h, "RIFF", 4) && !memcmp(h + 8, "WAVE", 4);
  bool have_fmt = false;
  int block_align = 0;
  long long pos = 12;
  while (ok) {
    unsigned char c[8];
    if (fread(c, 1, 8, f) != 8) { ok = false; break; }
    const uint32_t size = rd32(c + 4);
    pos += 8;
    if (!memcmp(c, "fmt ", 4)) {
      unsigned char b[40];
      const size_t n = size < sizeof(b) ? size : sizeof(b);
      if (size < 16 || fread(b, 1, n, f) != n) { ok = false; break; }
      int tag = rd16(b);
      info->channels = rd16(b + 2);
      info->fs = (int32_t)rd32(b + 4);
      block_align = rd16(b + 12);
      info->bits = rd16(b + 14);
      if (tag == 0xFFFE && n >= 26) tag = rd16(b + 24);   // WAVE_FORMAT_EXTENSIBLE: the sub-format's first two bytes
      info->format = tag;
      have_fmt = true;
      if (fseek(f, (long)(size - n + (size & 1)), SEEK_CUR)) { ok = false; break; }
      pos += size + (size & 1);
    } else if (!memcmp(c, "data", 4)) {
      if (!have_fmt) { ok = false; break; }
      const bool pcm = info->format == 1 && (info->bits == 8 || info->bits == 16 || info->bits == 32);
      const bool flt = info->format == 3 && info->bits == 32;
      if (!(pcm || flt) || info->channels < 1 || block_align != info->channels * info->bits / 8) { ok = false; break; }
      info->data_off = pos;
      info->frames = (long long)size / block_align;
      break;
    } else {
      if (fseek(f, (long)(size + (size & 1)), SEEK_CUR)) { ok = false; break; }
      pos += size + (size & 1);
    }
  }
  fclose(f);
  if (!ok) info->frames = -1;
  return ok && info->frames >= 0;
}

static bool read_one(const char* path, const uv_wav_info& in, float* dst) {
  if (in.frames < 0) return false;
  FILE* f = fopen(path, "rb");
  if (!f) return false;
  const int C = in.channels, bytes = in.bits / 8;
  const long long F = in.frames;
  bool ok = fseek(f, (long)in.data_off, SEEK_SET) == 0;
  std::vector<unsigned char> buf;
  const long long chunk = 1 << 16;   // frames per read
  if (ok) buf.resize((size_t)(F < chunk ? F : chunk) * C * bytes);
  for (long long f0 = 0; ok && f0 < F; f0 += chunk) {
    const long long nf = F - f0 < chunk ? F - f0 : chunk;
    if (C == 1 && in.format == 3) {   // mono float: the file's bytes are the result
      ok = fread(dst + f0, 4, (size_t)nf, f) == (size_t)nf;
      continue;
    }
    if (fread(buf.data(), (size_t)C * bytes, (size_t)nf, f) != (size_t)nf) { ok = false; break; }
    for (int c = 0; c < C; ++c) {
      float* o = dst + (long long)c * F + f0;
      const unsigned char* p = buf.data() + (size_t)c * bytes;
      const size_t stride = (size_t)C * bytes;
      if (in.format == 3) {
        for (long long i = 0; i < nf; ++i) memcpy(o + i, p + i * stride, 4);
      } else if (in.bits == 16) {
        for (long long i = 0; i < nf; ++i) o[i] = (float)(int16_t)rd16(p + i * stride) / 32768.0f;
      } else if (in.bits == 32) {
        for (long long i = 0; i < nf; ++i) o[i] = (float)(int32_t)rd32(p + i * stride) / 2147483648.0f;
      } else {
        for (long long i = 0; i < nf; ++i) o[i] = ((float)p[i * stride] - 128.0f) / 128.0f;
      }
    }
  }
  fclose(f);
  return ok;
}

static inline void wr32(unsigned char* p, uint32_t v) { p[0] = v & 255; p[1] = (v >> 8) & 255; p[2] = (v >> 16) & 255; p[3] = (v >> 24) & 255; }
static inline void wr16(unsigned char* p, uint16_t v) { p[0] = v & 255; p[1] = (v >> 8) & 255; }

// IEEE-float WAV (format tag 3, 'fact' chunk), what torchaudio.save / scipy write for a float32 tensor.
static bool write_one(const char* path, int C, long long F, int fs, const float* src) {
  if (C < 1 || F < 0 || (unsigned long long)F * C * 4 > 0xFFFFFF00ull) return false;
  FILE* f = fopen(path, "wb");
  if (!f) return false;
  const uint32_t data_bytes = (uint32_t)(F * C * 4);
  unsigned char h[58];
  memcpy(h, "RIFF", 4); wr32(h + 4, 50 + data_bytes); memcpy(h + 8, "WAVE", 4);
  memcpy(h + 12, "fmt ", 4); wr32(h + 16, 18); wr16(h + 20, 3); wr16(h + 22, (uint16_t)C); wr32(h + 24, (uint32_t)fs);
  wr32(h + 28, (uint32_t)fs * C * 4); wr16(h + 32, (uint16_t)(C * 4)); wr16(h + 34, 32); wr16(h + 36, 0);
  memcpy(h + 38, "fact", 4); wr32(h + 42, 4); wr32(h + 46, (uint32_t)F);
  memcpy(h + 50, "data", 4); wr32(h + 54, data_bytes);
  bool ok = fwrite(h, 1, 58, f) == 58;
  if (ok && C == 1) {
    ok = fwrite(src, 4, (size_t)F, f) == (size_t)F;
  } else if (ok) {
    const long long chunk = 1 << 15;
    std::vector<float> buf((size_t)(F < chunk ? F : chunk) * C);
    for (long long f0 = 0; ok && f0 < F; f0 += chunk) {
      const long long nf = F - f0 < chunk ? F - f0 : chunk;
      for (int c = 0; c < C; ++c)
        for (long long i = 0; i < nf; ++i) buf[(size_t)i * C + c] = src[(long long)c * F + f0 + i];
      ok = fwrite(buf.data(), 4, (size_t)(nf * C), f) == (size_t)(nf * C);
    }
  }
  ok = (fclose(f) == 0) && ok;
  return ok;
}

}  // namespace uvwav

extern "C" int uv_wav_probe(const char* const* paths, int n, uv_wav_info* info, int threads) {
  if (!paths || !info || n < 0) return UV_ERR_INVALID;
  std::atomic<int> bad(0);
  uvwav::parallel_for(n, threads, [&](int i) {
    if (!paths[i] || !uvwav::probe_one(paths[i], &info[i])) ++bad;
  });
  return bad.load();
}

extern "C" int uv_wav_read(const char* const* paths, int n, const uv_wav_info* info, const int64_t* offs, float* dst,
                           int threads, int32_t* status) {
  if (!paths || !info || !offs || !dst || n < 0) return UV_ERR_INVALID;
  std::atomic<int> bad(0);
  uvwav::parallel_for(n, threads, [&](int i) {
    const bool ok = paths[i] && uvwav::read_one(paths[i], info[i], dst + offs[i]);
    if (status) status[i] = ok ? 0 : -1;
    if (!ok) ++bad;
  });
  return bad.load();
}

extern "C" int uv_wav_write(const char* const* paths, int n, const int32_t* channels, const int64_t* frames, const int32_t* fs,
                            const int64_t* offs, const float* src, int threads, int32_t* status) {
  if (!paths || !channels || !frames || !fs || !offs || !src || n < 0) return UV_ERR_INVALID;
  std::atomic<int> bad(0);
  uvwav::parallel_for(n, threads, [&](int i) {
    const bool ok = paths[i] && uvwav::write_one(paths[i], channels[i], frames[i], fs[i], src + offs[i]);
    if (status) status[i] = ok ? 0 : -1;
    if (!ok) ++bad;
  });
  return bad.load();
}

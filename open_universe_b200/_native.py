"""ctypes binding of libuniverse_b200.so (include/universe_b200.h).

There is NO fallback: if the shared library is missing or does not match the header,
loading raises -- the product path never routes through PyTorch ops or the CPU oracle.
"""
import ctypes as C
import os
import subprocess

from . import _abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("UV_LIB") or os.path.join(_HERE, "libuniverse_b200.so")   # UV_LIB: A/B runs of two builds
CSRC = os.path.join(_HERE, "csrc")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17"]

EXPORTS = [
    "uv_abi_version", "uv_sizeof_op", "uv_last_error", "uv_model_create", "uv_model_destroy", "uv_plan_create",
    "uv_plan_destroy", "uv_plan_arena_bytes", "uv_enhance", "uv_run_ops", "uv_arena_read", "uv_arena_write",
    "uv_plan_num_launches", "uv_profile_ops", "uv_plan_arena_ptr", "uv_trace_ops", "uv_plan_status",
    "uv_mel_spec_static", "uv_set_noise_seed", "uv_signal_median", "uv_set_row_tables",
    "uv_wav_probe", "uv_wav_read", "uv_wav_write",
]


def build_library(verbose=False):
    """Compiles csrc/capi.cu for sm_100a into the in-tree shared library (nvcc cross-compiles without a GPU)."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + ["-shared", "-Xcompiler", "-fPIC", "-o", LIB_PATH, os.path.join(CSRC, "capi.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    return LIB_PATH, res.stdout + res.stderr


def _needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh"))]
    srcs.append(os.path.join(_HERE, "..", "include", "universe_b200.h"))
    return any(os.path.exists(s) and os.path.getmtime(s) > t for s in srcs)


_lib = None


def lib():
    """Loads (once) and returns the native library; raises if it is absent or inconsistent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"native library {LIB_PATH} not found. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(needs nvcc); open_universe_b200 has no PyTorch / CPU fallback."
        )
    L = C.CDLL(LIB_PATH)
    for name in EXPORTS:
        if not hasattr(L, name):
            raise RuntimeError(f"{LIB_PATH} does not export {name}")
    L.uv_abi_version.restype = C.c_int
    L.uv_sizeof_op.restype = C.c_size_t
    L.uv_last_error.restype = C.c_char_p
    L.uv_model_create.argtypes = [C.POINTER(C.c_void_p), C.c_void_p, C.c_size_t, C.c_int]
    L.uv_model_destroy.argtypes = [C.c_void_p]
    L.uv_model_destroy.restype = None
    L.uv_plan_create.argtypes = [C.POINTER(C.c_void_p), C.c_void_p, C.POINTER(_abi.UvOp), C.c_int, C.c_size_t,
                                 C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, C.c_int]
    L.uv_plan_destroy.argtypes = [C.c_void_p]
    L.uv_plan_destroy.restype = None
    L.uv_plan_arena_bytes.argtypes = [C.c_void_p]
    L.uv_plan_arena_bytes.restype = C.c_size_t
    L.uv_plan_num_launches.argtypes = [C.c_void_p]
    io = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(_abi.UvStepCoef), C.c_int, C.c_void_p]
    L.uv_enhance.argtypes = [C.c_void_p] + io
    L.uv_run_ops.argtypes = [C.c_void_p, C.c_int, C.c_int] + io
    L.uv_profile_ops.argtypes = [C.c_void_p] + io + [C.POINTER(C.c_float)]
    L.uv_trace_ops.argtypes = [C.c_void_p] + io[:-1] + [C.POINTER(C.c_float), C.POINTER(C.c_float)]
    L.uv_arena_read.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
    L.uv_arena_write.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
    L.uv_plan_status.argtypes = [C.c_void_p, C.c_void_p]
    L.uv_mel_spec_static.argtypes = [C.c_void_p, C.POINTER(_abi.UvMelDesc), C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    L.uv_set_noise_seed.argtypes = [C.c_void_p, C.c_ulonglong]
    L.uv_set_row_tables.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t]
    L.uv_signal_median.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p]
    L.uv_wav_probe.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.POINTER(_abi.UvWavInfo), C.c_int]
    L.uv_wav_read.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.POINTER(_abi.UvWavInfo), C.POINTER(C.c_int64), C.c_void_p, C.c_int,
                              C.POINTER(C.c_int32)]
    L.uv_wav_write.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_int32),
                               C.POINTER(C.c_int64), C.c_void_p, C.c_int, C.POINTER(C.c_int32)]
    L.uv_plan_arena_ptr.argtypes = [C.c_void_p]
    L.uv_plan_arena_ptr.restype = C.c_void_p
    if L.uv_abi_version() != _abi.UV_ABI_VERSION:
        raise RuntimeError(f"ABI version mismatch: library {L.uv_abi_version()}, binding {_abi.UV_ABI_VERSION}")
    if L.uv_sizeof_op() != C.sizeof(_abi.UvOp):
        raise RuntimeError(f"uv_op size mismatch: library {L.uv_sizeof_op()}, binding {C.sizeof(_abi.UvOp)}")
    _lib = L
    return L


def _check(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {lib().uv_last_error().decode()}")


class NativeModel:
    """Device-resident packed weights."""

    def __init__(self, blob_bytes, device_index):
        self.handle = C.c_void_p()
        self.device_index = device_index
        buf = C.create_string_buffer(blob_bytes, len(blob_bytes))
        _check(lib().uv_model_create(C.byref(self.handle), buf, len(blob_bytes), device_index), "uv_model_create")

    def mel_spec_static(self, mel_layer, sig_ptr, B, Lp, out_ptr, stream_ptr=None):
        """Allocation-free mel front end (uv_mel_spec_static); `mel_layer` = the packed layer record "c.mel".
        Returns the status code (0 = ok) instead of raising, so that callers can test the rejection path."""
        d = _abi.UvMelDesc(mel_layer["n_fft"], mel_layer["hop"], mel_layer["n_bins"], mel_layer["nbp"], mel_layer["n_mels"], 0,
                           mel_layer["basis_w"], mel_layer["fb_w"])
        return lib().uv_mel_spec_static(self.handle, C.byref(d), sig_ptr, B, Lp, out_ptr, stream_ptr)

    def close(self):
        if self.handle:
            lib().uv_model_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class NativePlan:
    """One (B, T, n_steps) plan: arena + captured CUDA graph."""

    def __init__(self, model, pg, use_graph=True):
        """use_graph: False / 0 eager launches, True / 1 capture at creation, 2 capture at the second enhance()."""
        self.model = model  # keep alive
        self.pg = pg
        self.handle = C.c_void_p()
        n = len(pg.ops)
        arr = (_abi.UvOp * n)(*pg.ops)
        init = C.create_string_buffer(pg.init_blob, len(pg.init_blob))
        _check(lib().uv_plan_create(C.byref(self.handle), model.handle, arr, n, pg.arena, init, len(pg.init_blob),
                                    pg.init_off, pg.n_steps, int(use_graph)), "uv_plan_create")
        self.arena_bytes = int(pg.arena)
        self.num_launches = lib().uv_plan_num_launches(self.handle)
        self._lengths = None   # the plan starts with the tables of a full-length batch (program._finish_init)

    def set_lengths(self, lengths):
        """Per-row valid lengths of the following calls (None: every row is T long): a row then behaves like a file of
        that length enhanced on its own (uv_set_row_tables).  A no-op while the lengths do not change."""
        from . import program

        key = None if lengths is None else tuple(int(v) for v in lengths)
        if key is not None and all(v == self.pg.T for v in key):
            key = None
        if key == self._lengths:
            return
        blob, off, lo, n = program.row_tables(self.pg, key)
        buf = C.create_string_buffer(blob, len(blob))
        _check(lib().uv_set_row_tables(self.handle, buf, len(blob), off, lo, n), "uv_set_row_tables")
        self._lengths = key

    def _coefs(self, coefs):
        arr = (_abi.UvStepCoef * len(coefs))()
        for i, (a, b, c) in enumerate(coefs):
            arr[i].a, arr[i].b, arr[i].c = a, b, c
        return arr

    def enhance(self, mix_ptr, noise_ptr, out_ptr, dbg_ptr, coefs, keep_rms, stream_ptr):
        _check(lib().uv_enhance(self.handle, mix_ptr, noise_ptr, out_ptr, dbg_ptr, self._coefs(coefs),
                                1 if keep_rms else 0, stream_ptr), "uv_enhance")

    def set_noise_seed(self, seed):
        """Key of the in-kernel sampler draws of the following calls that pass noise_ptr = None."""
        _check(lib().uv_set_noise_seed(self.handle, int(seed) & 0xFFFFFFFFFFFFFFFF), "uv_set_noise_seed")

    def status(self, stream_ptr=None):
        """Waits for the stream; raises if a device watchdog fired or a CUDA error is pending."""
        _check(lib().uv_plan_status(self.handle, stream_ptr), "uv_plan_status")

    def run_ops(self, first, last, mix_ptr, noise_ptr, out_ptr, dbg_ptr, coefs, keep_rms, stream_ptr):
        _check(lib().uv_run_ops(self.handle, first, last, mix_ptr, noise_ptr, out_ptr, dbg_ptr, self._coefs(coefs),
                                1 if keep_rms else 0, stream_ptr), "uv_run_ops")

    def profile_ops(self, mix_ptr, noise_ptr, out_ptr, dbg_ptr, coefs, keep_rms, stream_ptr):
        """Per-op device times (ms) of one eager pass, measured with CUDA events."""
        ms = (C.c_float * len(self.pg.ops))()
        _check(lib().uv_profile_ops(self.handle, mix_ptr, noise_ptr, out_ptr, dbg_ptr, self._coefs(coefs),
                                    1 if keep_rms else 0, stream_ptr, ms), "uv_profile_ops")
        return list(ms)

    def trace_ops(self, mix_ptr, noise_ptr, out_ptr, dbg_ptr, coefs, keep_rms):
        """(start, end) device times in ms of every op of one eager pass run on the plan's own lane streams."""
        n = len(self.pg.ops)
        t0, t1 = (C.c_float * n)(), (C.c_float * n)()
        _check(lib().uv_trace_ops(self.handle, mix_ptr, noise_ptr, out_ptr, dbg_ptr, self._coefs(coefs),
                                  1 if keep_rms else 0, t0, t1), "uv_trace_ops")
        return list(t0), list(t1)

    def arena_view(self, off, n_floats, device):
        """Zero-copy f32 torch view of arena[off : off + 4 n_floats] (the plan keeps owning the memory)."""
        import torch

        class _Ptr:
            pass

        h = _Ptr()
        h.__cuda_array_interface__ = dict(shape=(n_floats,), typestr="<f4", data=(lib().uv_plan_arena_ptr(self.handle) + off, False),
                                          version=2)
        t = torch.as_tensor(h, device=device)
        t._uv_plan = self   # keep the arena alive as long as the view
        return t

    def arena_read(self, off, nbytes):
        buf = C.create_string_buffer(nbytes)
        _check(lib().uv_arena_read(self.handle, off, buf, nbytes), "uv_arena_read")
        return buf.raw

    def close(self):
        if self.handle:
            lib().uv_plan_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def signal_median(signal):
    """Ensemble statistic on the device (uv_signal_median; reference utils/stats.py:22-66).
    signal: CUDA float tensor (ensemble, batch, ...) -> (batch, ...)."""
    import torch

    if not signal.is_cuda:
        raise RuntimeError("open_universe_b200 runs only on a CUDA device")
    shape = signal.shape
    E, B = int(shape[0]), int(shape[1])
    flat = signal.reshape(E, B, -1).to(torch.float32).contiguous()
    S = int(flat.shape[2])
    out = torch.empty((B, S), dtype=torch.float32, device=signal.device)
    scratch = torch.empty((B, 32), dtype=torch.int32, device=signal.device)
    with torch.cuda.device(signal.device):
        _check(lib().uv_signal_median(flat.data_ptr(), E, B, S, out.data_ptr(), scratch.data_ptr(),
                                      torch.cuda.current_stream(signal.device).cuda_stream), "uv_signal_median")
    return out.reshape(shape[1:]).to(signal.dtype)


# ---------------------------------------------------------------- batch WAV IO (uv_wav_probe / uv_wav_read / uv_wav_write)
def _c_paths(paths):
    arr = (C.c_char_p * len(paths))()
    arr[:] = [os.fsencode(str(p)) for p in paths]
    return arr


def wav_probe(paths, threads=4):
    """-> list of (channels, frames, fs) per file, None where the native reader does not decode the file."""
    n = len(paths)
    info = (_abi.UvWavInfo * max(n, 1))()
    rc = lib().uv_wav_probe(_c_paths(paths), n, info, threads)
    if rc < 0:
        raise RuntimeError("uv_wav_probe: bad argument")
    return info, [(i.channels, i.frames, i.fs) if i.frames >= 0 else None for i in info[:n]]


def wav_read(paths, info, offsets, dst, threads=4):
    """Decodes the probed files into the float32 CPU tensor `dst` (file i at offsets[i], planar [channels][frames]).
    -> list of booleans (decoded?)."""
    n = len(paths)
    offs = (C.c_int64 * max(n, 1))(*offsets)
    status = (C.c_int32 * max(n, 1))()
    assert dst.dtype.itemsize == 4 and dst.is_contiguous() and not dst.is_cuda
    rc = lib().uv_wav_read(_c_paths(paths), n, info, offs, dst.data_ptr(), threads, status)
    if rc < 0:
        raise RuntimeError("uv_wav_read: bad argument")
    return [status[i] == 0 for i in range(n)]


def wav_write(paths, shapes, rates, offsets, src, threads=4):
    """Writes float32 WAV files from the float32 CPU tensor `src` (file i = shapes[i] = (channels, frames) at
    offsets[i]).  -> list of booleans (written?)."""
    n = len(paths)
    ch = (C.c_int32 * max(n, 1))(*[s[0] for s in shapes])
    fr = (C.c_int64 * max(n, 1))(*[s[1] for s in shapes])
    fs = (C.c_int32 * max(n, 1))(*rates)
    offs = (C.c_int64 * max(n, 1))(*offsets)
    status = (C.c_int32 * max(n, 1))()
    assert src.dtype.itemsize == 4 and src.is_contiguous() and not src.is_cuda
    rc = lib().uv_wav_write(_c_paths(paths), n, ch, fr, fs, offs, src.data_ptr(), threads, status)
    if rc < 0:
        raise RuntimeError("uv_wav_write: bad argument")
    return [status[i] == 0 for i in range(n)]

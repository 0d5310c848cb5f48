"""ctypes mirror of include/universe_b200.h (records only; the loaded library
lives in `_native.py`)."""
import ctypes as C

UV_ABI_VERSION = 10

OP_PREP, OP_MEL, OP_MEL_FIN, OP_CONV_IN, OP_GEMM, OP_GRU, OP_CONV_OUT, OP_INIT_X, OP_POST, OP_FILM, OP_CONVBLOCK, OP_AUX, OP_SIGP, OP_MEL_POW = range(1, 15)
OP_NAMES = {
    1: "PREP", 2: "MEL", 3: "MEL_FIN", 4: "CONV_IN", 5: "GEMM", 6: "GRU", 7: "CONV_OUT", 8: "INIT_X", 9: "POST", 10: "FILM",
    11: "CONVBLOCK", 12: "AUX", 13: "SIGP", 14: "MEL_POW",
}

DT_BF16, DT_F32 = 1, 2


class UvTens(C.Structure):
    _fields_ = [
        ("off", C.c_int64),
        ("dtype", C.c_int32),
        ("nch", C.c_int32),
        ("split", C.c_int32),
        ("Lh", C.c_int32),
        ("P", C.c_int32),
        ("G", C.c_int32),
    ]


class UvOut(C.Structure):
    _fields_ = [("t", UvTens), ("act", C.c_int32), ("a1", C.c_float), ("a2", C.c_float), ("pad_", C.c_int32)]


class UvOp(C.Structure):
    _fields_ = [
        ("type", C.c_int32),
        ("step", C.c_int32),
        ("a", UvTens),
        ("n_src", C.c_int32),
        ("smin", C.c_int32),
        ("smax", C.c_int32),
        ("KC", C.c_int32),
        ("w_w", C.c_int64),
        ("N_TILE", C.c_int32),
        ("N_total", C.c_int32),
        ("m_rows", C.c_int32),
        ("Lh_A", C.c_int32),
        ("M_valid", C.c_int32),
        ("up", C.c_int32),
        ("C_out", C.c_int32),
        ("bias_w", C.c_int64),
        ("film", C.c_int64),
        ("add1", UvTens),
        ("add2", UvTens),
        ("scale", C.c_float),
        ("n_out", C.c_int32),
        ("out", UvOut * 3),
        ("rm_out", C.c_int64),
        ("rm_ld", C.c_int32),
        ("B", C.c_int32),
        ("T", C.c_int32),
        ("Lp", C.c_int32),
        ("C", C.c_int32),
        ("H", C.c_int32),
        ("F", C.c_int32),
        ("i0", C.c_int32),
        ("i1", C.c_int32),
        ("i2", C.c_int32),
        ("i3", C.c_int32),
        ("p0", C.c_int64),
        ("p1", C.c_int64),
        ("p2", C.c_int64),
        ("p3", C.c_int64),
        ("p4", C.c_int64),
        ("p5", C.c_int64),
        ("f0", C.c_float),
        ("f1", C.c_float),
        ("f2", C.c_float),
        ("f3", C.c_float),
        ("lane", C.c_int32),
        ("sync", C.c_int32),
        ("a_act", C.c_int32),
        ("a_slope", C.c_float),
        ("rows", C.c_int64),
    ]


class UvMelDesc(C.Structure):
    _fields_ = [("n_fft", C.c_int32), ("hop", C.c_int32), ("n_bins", C.c_int32), ("nbp", C.c_int32), ("n_mels", C.c_int32),
                ("pad_", C.c_int32), ("basis_w", C.c_int64), ("fb_w", C.c_int64)]


class UvWavInfo(C.Structure):
    _fields_ = [("channels", C.c_int32), ("fs", C.c_int32), ("format", C.c_int32), ("bits", C.c_int32),
                ("frames", C.c_int64), ("data_off", C.c_int64)]


UV_STATIC_SCRATCH_BYTES = 16 << 20


class UvStepCoef(C.Structure):
    _fields_ = [("a", C.c_float), ("b", C.c_float), ("c", C.c_float), ("pad_", C.c_float)]


def no_tens():
    return UvTens(-1, 0, 0, 1, 0, 0, 0)


def new_op(type_, step=-1):
    op = UvOp()
    op.type = type_
    op.step = step
    op.a = no_tens()
    op.add1 = no_tens()
    op.add2 = no_tens()
    for i in range(3):
        op.out[i].t = no_tens()
    op.w_w = op.bias_w = op.film = op.rm_out = -1
    op.p0 = op.p1 = op.p2 = op.p3 = op.p4 = op.p5 = -1
    op.rows = -1
    op.scale = 1.0
    op.up = 1
    op.n_src = 1
    return op

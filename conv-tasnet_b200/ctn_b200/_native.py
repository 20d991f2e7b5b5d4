"""ctypes binding of libctn_b200.so (C ABI declared in include/ctn_b200.h).

There is no fallback: if the library is missing or a call fails, a RuntimeError
(or ValueError for the reference's ValueError cases) is raised.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# Precision of the tensor-core GEMM operands, chosen once per process (CTN_PRECISION):
#   "bf16x3" (default)  bf16 (hi, lo) pairs, three MMAs per product: 1.5e-5 rel-L2 at the paper configuration,
#                       the only mode training supports
#   "fp16"              fp16 operands, one MMA per product: inference only, 4.9e-4 rel-L2 (tolerance 1e-3),
#                       +14-19 % throughput.  Same sources compiled with
#                       -DCTN_OPERAND_FP16=1 into a second library (build_native.py --fp16).
_FORMATS = {"bf16x3": ("libctn_b200.so", 3), "fp16": ("libctn_b200_fp16.so", 101)}
PRECISION = os.environ.get("CTN_PRECISION", "bf16x3")
if PRECISION not in _FORMATS:
    raise RuntimeError("CTN_PRECISION=%r: expected one of %s" % (PRECISION, sorted(_FORMATS)))
_LIB_PATH = os.path.join(os.path.dirname(_HERE), _FORMATS[PRECISION][0])
if os.environ.get("CTN_ALLOW_EXPERIMENT_BUILD") == "1" and os.environ.get("CTN_LIB_FILE"):
    # A/B runs of experiment builds (build_native.py with CTN_LIB_NAME / NVCC_FLAGS): never the product path
    _LIB_PATH = os.path.join(os.path.dirname(_HERE), os.environ["CTN_LIB_FILE"])

NORM_NONE, NORM_GLN, NORM_CLN = 0, 1, 2
MASK_RELU, MASK_SOFTMAX = 0, 1
ROW_TILE, COL_TILE, DW_TILE = 128, 128, 16

_vp = C.c_void_p


class GemmArgs(C.Structure):
    """struct ctn_gemm_args"""
    _fields_ = [
        ("M", C.c_int), ("K", C.c_int), ("Cin", C.c_int), ("Cout", C.c_int),
        ("a_hi", _vp), ("a_lo", _vp), ("w_hi", _vp), ("w_lo", _vp),
        ("out_f32", _vp), ("out_hi", _vp), ("out_lo", _vp),
        ("resid", _vp), ("col_g", _vp), ("col_bias", _vp),
        ("in_norm", C.c_int), ("in_stat", _vp), ("in_P", C.c_int), ("in_count", C.c_double),
        ("out_stat", C.c_int), ("out_stat_buf", _vp), ("stat_prelu", _vp),
        ("force_simt", C.c_int),
        ("bw_pre", _vp), ("bw_gamma", _vp), ("bw_slope", _vp), ("bw_stat", _vp), ("bw_P", C.c_int),
        ("resid_hi", _vp), ("resid_lo", _vp), ("store_prelu", C.c_int),
    ]


class DwconvArgs(C.Structure):
    """struct ctn_dwconv_args"""
    _fields_ = [
        ("M", C.c_int), ("K", C.c_int), ("H", C.c_int), ("P", C.c_int),
        ("dilation", C.c_int), ("causal", C.c_int),
        ("inp", _vp), ("prelu_in", _vp), ("gamma_in", _vp), ("beta_in", _vp),
        ("norm", C.c_int), ("in_stat", _vp), ("in_P", C.c_int),
        ("wd", _vp), ("prelu_out", _vp), ("out_pre", _vp), ("out_hi", _vp), ("out_lo", _vp),
        ("out_stat", _vp),
    ]


class DwpwArgs(C.Structure):
    """struct ctn_dwpw_args"""
    _fields_ = [
        ("M", C.c_int), ("K", C.c_int), ("H", C.c_int), ("B", C.c_int), ("dilation", C.c_int), ("causal", C.c_int),
        ("a", _vp), ("prelu_in", _vp), ("gamma_in", _vp), ("beta_in", _vp), ("in_stat", _vp), ("in_P", C.c_int),
        ("wd", _vp), ("prelu_out", _vp), ("w_hi", _vp), ("w_lo", _vp), ("y", _vp), ("out_stat", _vp), ("d_out", _vp),
    ]


class Updc1Args(C.Structure):
    """struct ctn_updc1_args"""
    _fields_ = [
        ("M", C.c_int), ("K", C.c_int), ("B", C.c_int), ("H", C.c_int),
        ("x_hi", _vp), ("x_lo", _vp), ("y", _vp), ("in_stat", _vp), ("in_P", C.c_int), ("in_count", C.c_double),
        ("col_g", _vp), ("col_bias", _vp), ("w_hi", _vp), ("w_lo", _vp), ("a", _vp), ("out_stat", _vp), ("slope", _vp),
        ("x_out_hi", _vp), ("x_out_lo", _vp), ("store_raw", C.c_int),
    ]


class OptChunk(C.Structure):
    """struct ctn_opt_chunk"""
    _fields_ = [("param", _vp), ("grad", _vp), ("exp_avg", _vp), ("exp_avg_sq", _vp), ("n", C.c_int), ("pad_", C.c_int)]


class BwdDwArgs(C.Structure):
    """struct ctn_bwd_dw_args"""
    _fields_ = [
        ("M", C.c_int), ("K", C.c_int), ("H", C.c_int), ("P", C.c_int), ("dilation", C.c_int), ("causal", C.c_int),
        ("P1", C.c_int), ("P2", C.c_int), ("Pb", C.c_int),
        ("du", _vp), ("d", _vp), ("a", _vp), ("stat1", _vp), ("stat2", _vp), ("bstat", _vp),
        ("g1", _vp), ("b1", _vp), ("g2", _vp), ("wd", _vp), ("slope1", _vp), ("slope2", _vp),
        ("dn1", _vp), ("chan_part", _vp), ("scal_part", _vp), ("norm", C.c_int),
        ("b2", _vp), ("u_hi", _vp), ("u_lo", _vp),
    ]


class CopyItem(C.Structure):
    """struct ctn_copy_item"""
    _fields_ = [("dst", _vp), ("src", _vp), ("bytes", C.c_longlong)]


class CopyTable:
    """Device-resident table of ctn_copy_item for ctn_copy_table (one launch per run)."""

    def __init__(self, device, pairs):
        """pairs: list of (dst tensor view, src tensor view), both contiguous, same number of bytes."""
        items = []
        self.max_bytes = 0
        for dst, src in pairs:
            assert dst.is_contiguous() and src.is_contiguous()
            n = dst.numel() * dst.element_size()
            assert n == src.numel() * src.element_size() and n % 4 == 0
            it = CopyItem()
            it.dst, it.src, it.bytes = dst.data_ptr(), src.data_ptr(), n
            items.append(it)
            self.max_bytes = max(self.max_bytes, n)
        self.n = len(items)
        arr = (CopyItem * self.n)(*items)
        self.dev = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(device)

    def run(self):
        check(lib().ctn_copy_table(self.dev.data_ptr(), self.n, self.max_bytes, stream()))


class PrepItem(C.Structure):
    """struct ctn_prep_item"""
    _fields_ = [("W", _vp), ("gamma", _vp), ("beta", _vp), ("w_hi", _vp), ("w_lo", _vp), ("col_g", _vp),
                ("col_bias", _vp), ("rows", C.c_int), ("cols", C.c_int), ("transpose", C.c_int),
                ("reserved", C.c_int)]


class PrepTable:
    """Device-resident table of ctn_prep_item for ctn_weight_prep_batch (one launch per refresh)."""

    def __init__(self, device):
        self.device, self.items, self.dev, self.max_rows = device, [], None, 0

    def add(self, W, rows, cols, gamma, beta, transpose, out):
        it = PrepItem()
        it.W, it.gamma, it.beta = W.data_ptr(), ptr(gamma), ptr(beta)
        it.w_hi, it.w_lo = out[0].data_ptr(), out[1].data_ptr()
        it.col_g = out[2].data_ptr() if len(out) > 2 else None
        it.col_bias = out[3].data_ptr() if len(out) > 3 else None
        it.rows, it.cols, it.transpose = rows, cols, int(transpose)
        self.items.append(it)
        self.max_rows = max(self.max_rows, rows)

    def run(self):
        if self.dev is None:
            arr = (PrepItem * len(self.items))(*self.items)
            self.dev = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(self.device)
        check(lib().ctn_weight_prep_batch(self.dev.data_ptr(), len(self.items), self.max_rows, stream()))


MAX_SEG = 12


class ReduceArgs(C.Structure):
    """struct ctn_reduce_args"""
    _fields_ = [
        ("part", _vp), ("n_part", C.c_int), ("part_stride", C.c_longlong), ("nseg", C.c_int),
        ("accumulate", C.c_int),
        ("src_off", C.c_longlong * MAX_SEG), ("len", C.c_int * MAX_SEG), ("src_stride", C.c_int * MAX_SEG),
        ("dst", _vp * MAX_SEG), ("dst_stride", C.c_int * MAX_SEG),
        ("seg_part", _vp * MAX_SEG), ("seg_n_part", C.c_int * MAX_SEG), ("seg_part_stride", C.c_longlong * MAX_SEG),
    ]


_i, _ll = C.c_int, C.c_longlong

_SIGNATURES = {
    "ctn_version": (C.c_int, []),
    "ctn_operand_format": (C.c_int, []),
    "ctn_last_error": (C.c_char_p, []),
    "ctn_gemm_gln_partials": (C.c_int, [C.c_int, C.c_int]),
    "ctn_dwconv_gln_partials": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "ctn_gemm_cln_groups": (C.c_int, [C.c_int]),
    "ctn_weight_prep_batch": (C.c_int, [_vp, C.c_int, C.c_int, _vp]),
    "ctn_weight_prep": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp]),
    "ctn_encoder_fwd": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp]),
    "ctn_gemm1x1": (C.c_int, [C.POINTER(GemmArgs), _vp]),
    "ctn_dwconv_fwd": (C.c_int, [C.POINTER(DwconvArgs), _vp]),
    "ctn_clip_adam": (C.c_int, [_vp, C.c_int, _vp, C.c_float, C.c_float, C.c_double, C.c_double, C.c_float, C.c_float,
                                C.c_float, C.c_int, _vp, _vp]),
    "ctn_updc1_supported": (C.c_int, [C.c_int, C.c_int, C.c_int]),
    "ctn_updc1_fwd": (C.c_int, [C.POINTER(Updc1Args), _vp]),
    "ctn_dwpw_supported": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "ctn_dwpw_partials": (C.c_int, [C.c_int, C.c_int]),
    "ctn_dwpw_fwd": (C.c_int, [C.POINTER(DwpwArgs), _vp]),
    "ctn_resid_update": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int, C.c_double, _vp, _vp, C.c_int, C.c_int, C.c_int, _vp]),
    "ctn_decoder_fwd": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                  _vp, _vp]),
    "ctn_overlap_add": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp, _vp]),
    "ctn_norm_split": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _vp, _vp, _vp]),
    "ctn_wgrad_splits": (_i, [_ll, _i, _i]),
    "ctn_wgrad": (_i, [_vp, _vp, _vp, _vp, _ll, _i, _i, _vp, _i, _vp]),
    "ctn_bwd_dw_layout": (_i, [_i, _i, _i, _i, _i, _i, C.POINTER(_i), C.POINTER(_i)]),
    "ctn_bwd_dwconv": (_i, [C.POINTER(BwdDwArgs), _vp]),
    "ctn_bwd_dw_emits_u": (_i, [_i, _i, _i]),
    "ctn_bwd_prelu_norm_tiles": (_i, [_i]),
    "ctn_bwd_prelu_norm": (_i, [_vp, _vp, _vp, _i, _vp, _i, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "ctn_bwd_prelu_norm_cln": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "ctn_reduce_partials": (_i, [C.POINTER(ReduceArgs), _vp]),
    "ctn_decoder_bwd_tiles": (_i, [_i]),
    "ctn_decoder_bwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "ctn_cln_bwd_tiles": (_i, [_i]),
    "ctn_cln_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp]),
    "ctn_encoder_wgrad_tiles": (_i, [_i]),
    "ctn_encoder_wgrad": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "ctn_pit_workspace_doubles": (C.c_int, [C.c_int, C.c_int, C.c_int]),
    "ctn_pit_sisnr_fwd": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int, _vp, C.c_int, _vp, _vp, _vp, _vp,
                                    _vp, _vp, _vp]),
    "ctn_pit_sisnr_fwd_static": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int, _vp, C.c_int, _vp, _vp, _vp, _vp,
                                           _vp, _vp]),
    "ctn_pit_reorder": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int, _vp, _vp]),
    "ctn_bn_fold": (C.c_int, [_vp, _vp, _vp, _vp, C.c_float, C.c_int, _vp, _vp, _vp]),
    "ctn_bn_tiles": (_i, [_ll]),
    "ctn_bn_colsums": (_i, [_vp, _vp, _vp, _vp, _vp, _ll, _i, C.c_double, _vp, _vp, _vp]),
    "ctn_bn_train_fold": (_i, [_vp, C.c_double, _vp, _vp, C.c_float, C.c_float, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp,
                               _vp, _vp]),
    "ctn_bn_bwd_center": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, C.c_double, _i, _ll, _i, _vp, _vp, _vp]),
    "ctn_copy_table": (C.c_int, [_vp, C.c_int, C.c_longlong, _vp]),
    "ctn_pad_batch": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int, C.c_float, _vp, _vp, _vp]),
    "ctn_remove_pad": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int, _vp, _vp]),
    "ctn_sisnri": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp]),
    "ctn_pit_sisnr_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int, C.c_int, C.c_int, _vp, _vp]),
}

_lib = None


def exported_symbols():
    """Names include/ctn_b200.h declares (used by the CPU-side ABI test)."""
    return sorted(_SIGNATURES)


def lib():
    """Loads the shared library once; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise RuntimeError(
                "%s not found: build it with `python conv-tasnet_b200/build_native.py%s` (there is no CPU or "
                "PyTorch fallback)" % (_LIB_PATH, "" if PRECISION == "bf16x3" else " --fp16"))
        handle = C.CDLL(_LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        fmt = handle.ctn_operand_format()
        if fmt != _FORMATS[PRECISION][1] and os.environ.get("CTN_ALLOW_EXPERIMENT_BUILD") != "1":
            raise RuntimeError(
                "%s was built with operand format %d, CTN_PRECISION=%s needs %d (see DESIGN.md section 8); rebuild "
                "with `python conv-tasnet_b200/build_native.py --all --force` or set CTN_ALLOW_EXPERIMENT_BUILD=1"
                % (_LIB_PATH, fmt, PRECISION, _FORMATS[PRECISION][1]))
        _lib = handle
    return _lib


def check(rc):
    """Translates a C return code into the reference's Python error conventions."""
    if rc == 0:
        return
    msg = lib().ctn_last_error().decode("utf-8", "replace")
    if "Unsupported mask non-linear function" in msg:
        raise ValueError(msg)  # conv_tasnet.py:208
    raise RuntimeError("libctn_b200: %s (code %d)" % (msg, rc))


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream


_capture_streams = {}


def capture_stream(device):
    """A capture stream that belongs to `device`.  torch.cuda.graph's default capture stream is created
    once per process on whatever device was current then, which breaks captures on any other GPU of a
    process that drives several (torch.nn.DataParallel)."""
    st = _capture_streams.get(device.index)
    if st is None:
        # high priority: kernel nodes inherit it, so in a captured training backward the dgrad chain's CTAs are placed
        # before those of the weight-gradient kernels that share the graph from a default-priority side stream
        # (backward 3.89 -> 3.61 ms at batch 3)
        st = _capture_streams[device.index] = torch.cuda.Stream(device=device, priority=-1)
    return st


def require_cuda(t, what):
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise RuntimeError("%s must be a CUDA tensor: this implementation has no CPU path" % what)

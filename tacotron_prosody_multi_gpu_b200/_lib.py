"""ctypes binding of libtaco_b200.so (include/taco_b200.h).  There is NO fallback: if the CUDA
extension is missing the import of this module's ``lib()`` raises."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TACO_B200_LIB") or os.path.join(HERE, "libtaco_b200.so")   # env: A/B builds of the same library (development)

TACO_OK = 0
TACO_MODE_FREE = 0
TACO_MODE_TEACHER = 1
TACO_PRECISION_FP32 = 0
TACO_PRECISION_BF16_CONV = 1
TACO_PRECISION_BF16 = 2
TACO_STAGE_ENCODE, TACO_STAGE_DECODER_KERNEL, TACO_STAGE_POSTNET = 0, 1, 2

ERROR_NAMES = {-1: "TACO_ERR_INVALID", -2: "TACO_ERR_UNSUPPORTED", -3: "TACO_ERR_CUDA",
               -4: "TACO_ERR_MISSING_WEIGHT", -5: "TACO_ERR_STATE"}

# every symbol include/taco_b200.h declares
EXPORTED_SYMBOLS = [
    "taco_create", "taco_destroy", "taco_last_error", "taco_version", "taco_set_weight",
    "taco_finalize_weights", "taco_set_precision", "taco_set_encoder_ctas", "taco_encode", "taco_decode",
    "taco_transpose_alignments", "taco_postnet", "taco_launch_count", "taco_get_decoder_state",
    "taco_stage_ms", "taco_debug_read_workspace",
]


class TacoHParams(C.Structure):
    _fields_ = [
        ("n_symbols", C.c_int32), ("embedding_dim", C.c_int32), ("conv_num_layers", C.c_int32),
        ("conv_kernel_size", C.c_int32), ("conv_channels", C.c_int32), ("enc_lstm_hidden_size", C.c_int32),
        ("use_gst", C.c_int32), ("num_gst", C.c_int32), ("num_heads", C.c_int32),
        ("style_embed_depth", C.c_int32), ("style_att_dim", C.c_int32), ("reference_depth", C.c_int32),
        ("n_reference_filters", C.c_int32), ("reference_filters", C.c_int32 * 8),
        ("attention_dim", C.c_int32), ("attention_filters", C.c_int32), ("attention_kernel", C.c_int32),
        ("prenet_layers", C.c_int32 * 2), ("decoder_lstm_units", C.c_int32), ("num_mels", C.c_int32),
        ("outputs_per_step", C.c_int32), ("postnet_num_layers", C.c_int32), ("postnet_kernel_size", C.c_int32),
        ("postnet_channels", C.c_int32), ("mask_encoder", C.c_int32), ("smoothing", C.c_int32),
        ("cumulative_weights", C.c_int32), ("stop_at_any", C.c_int32), ("clip_outputs", C.c_int32),
        ("symmetric_mels", C.c_int32), ("zoneout_rate", C.c_float), ("dropout_rate", C.c_float),
        ("max_abs_value", C.c_float), ("lower_bound_decay", C.c_float),
    ]


_lib = None


class TacoError(RuntimeError):
    pass


def lib():
    """Load the shared library once; fail loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise TacoError(
            "libtaco_b200.so is missing (%s). Build it with `python -m tacotron_prosody_multi_gpu_b200.build` "
            "(needs nvcc); this package has no CPU or PyTorch fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, u64 = C.c_void_p, C.c_int, C.c_int64, C.c_uint64
    L.taco_create.argtypes = [C.POINTER(TacoHParams), i32, C.POINTER(vp)]
    L.taco_create.restype = i32
    L.taco_destroy.argtypes = [vp]
    L.taco_destroy.restype = None
    L.taco_last_error.argtypes = [vp]
    L.taco_last_error.restype = C.c_char_p
    L.taco_version.argtypes = []
    L.taco_version.restype = C.c_char_p
    L.taco_set_weight.argtypes = [vp, C.c_char_p, vp, C.POINTER(i64), i32, i32]
    L.taco_set_weight.restype = i32
    L.taco_finalize_weights.argtypes = [vp]
    L.taco_finalize_weights.restype = i32
    L.taco_set_precision.argtypes = [vp, i32]
    L.taco_set_precision.restype = i32
    L.taco_set_encoder_ctas.argtypes = [vp, i32]
    L.taco_set_encoder_ctas.restype = i32
    L.taco_encode.argtypes = [vp, vp, vp, i32, i32, vp, i32, vp, vp, vp]
    L.taco_encode.restype = i32
    L.taco_decode.argtypes = [vp, i32, vp, vp, i32, i32, vp, i32, i32, vp, u64, i32, i32, vp, vp, vp, vp, vp]
    L.taco_decode.restype = i32
    L.taco_transpose_alignments.argtypes = [vp, vp, i32, i32, i32, vp, vp]
    L.taco_transpose_alignments.restype = i32
    L.taco_postnet.argtypes = [vp, vp, i32, i32, i32, vp, vp, vp]
    L.taco_postnet.restype = i32
    L.taco_launch_count.argtypes = [vp]
    L.taco_launch_count.restype = i64
    L.taco_get_decoder_state.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp, vp, vp]
    L.taco_get_decoder_state.restype = i32
    L.taco_stage_ms.argtypes = [vp, i32, C.POINTER(C.c_float)]
    L.taco_stage_ms.restype = i32
    L.taco_debug_read_workspace.argtypes = [vp, C.c_char_p, vp, C.c_size_t]
    L.taco_debug_read_workspace.restype = i32
    _lib = L
    return L


def make_hparams(hp, n_symbols):
    s = TacoHParams()
    s.n_symbols = n_symbols
    s.embedding_dim = hp.embedding_dim
    s.conv_num_layers = hp.conv_num_layers
    s.conv_kernel_size = int(hp.conv_kernel_size[0]) if isinstance(hp.conv_kernel_size, (tuple, list)) else int(hp.conv_kernel_size)
    s.conv_channels = hp.conv_channels
    s.enc_lstm_hidden_size = hp.enc_lstm_hidden_size
    s.use_gst = int(bool(hp.use_gst))
    s.num_gst = hp.num_gst
    s.num_heads = hp.num_heads
    s.style_embed_depth = hp.style_embed_depth
    s.style_att_dim = hp.style_att_dim
    s.reference_depth = hp.reference_depth
    filt = list(hp.reference_filters)
    s.n_reference_filters = len(filt)
    for i, f in enumerate(filt[:8]):
        s.reference_filters[i] = int(f)
    s.attention_dim = hp.attention_dim
    s.attention_filters = hp.attention_filters
    s.attention_kernel = int(hp.attention_kernel[0]) if isinstance(hp.attention_kernel, (tuple, list)) else int(hp.attention_kernel)
    s.prenet_layers[0], s.prenet_layers[1] = int(hp.prenet_layers[0]), int(hp.prenet_layers[1])
    s.decoder_lstm_units = hp.decoder_lstm_units
    s.num_mels = hp.num_mels
    s.outputs_per_step = hp.outputs_per_step
    s.postnet_num_layers = hp.postnet_num_layers
    s.postnet_kernel_size = int(hp.postnet_kernel_size[0]) if isinstance(hp.postnet_kernel_size, (tuple, list)) else int(hp.postnet_kernel_size)
    s.postnet_channels = hp.postnet_channels
    s.mask_encoder = int(bool(hp.mask_encoder))
    s.smoothing = int(bool(hp.smoothing))
    s.cumulative_weights = int(bool(hp.cumulative_weights))
    s.stop_at_any = int(bool(hp.stop_at_any))
    s.clip_outputs = int(bool(hp.clip_outputs))
    s.symmetric_mels = int(bool(hp.symmetric_mels))
    s.zoneout_rate = float(hp.zoneout_rate)
    s.dropout_rate = float(hp.dropout_rate)
    s.max_abs_value = float(hp.max_abs_value)
    s.lower_bound_decay = float(hp.lower_bound_decay)
    return s

"""ctypes binding of libdyibf.so (include/dyibf.h).  Loading fails loudly when the CUDA library is missing:
there is no CPU implementation behind this package."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libdyibf.so")

OK, ERR_ARG, ERR_IO, ERR_CUDA, ERR_FORMAT, ERR_SIZE_MISMATCH, ERR_CAPACITY, ERR_NOMEM, ERR_BUSY = 0, -1, -2, -3, -4, -5, -6, -7, -8
DEVERR_BIN_RANGE, DEVERR_SEQ_TOO_LONG = 1, 2
GROUP_READ_SHARDED, GROUP_BIN_SHARDED = 0, 1
ENC_DNA5, ENC_DNA5Q, ENC_ASCII = 0, 1, 2
MEM_HOST, MEM_DEVICE = 0, 1
FMT_FASTA, FMT_FASTQ = 0, 1
ALPHA_DNA5, ALPHA_IUPAC = 0, 1
MAX_HASH = 8
OPT_EARLY_EXIT = 1
OPT_BUILD_BUCKET_MIB = 2
OPT_BUILD_CHUNK_MIB = 3
OPT_QUERY_CHUNK_MIB = 4
OPT_BUILD_MODE = 5
OPT_LAST_BUILD_MODE = 6
BUILD_AUTO, BUILD_DIRECT, BUILD_BUCKETED, BUILD_SLICED = 0, 1, 2, 3


class Params(C.Structure):
    _fields_ = [("bins", C.c_uint32), ("n_hash", C.c_uint32), ("kmer", C.c_uint32), ("words_per_block", C.c_uint32),
                ("nbits", C.c_uint64), ("n_blocks", C.c_uint64), ("n_words", C.c_uint64), ("device", C.c_int)]


class Profile(C.Structure):
    _fields_ = [("pack_ms", C.c_double), ("query_ms", C.c_double), ("compact_ms", C.c_double),
                ("build_ms", C.c_double), ("clear_ms", C.c_double),
                ("pack_launches", C.c_uint64), ("query_launches", C.c_uint64), ("compact_launches", C.c_uint64),
                ("build_launches", C.c_uint64), ("clear_launches", C.c_uint64), ("query_lookups", C.c_uint64),
                ("build_apply_ms", C.c_double), ("build_apply_launches", C.c_uint64),
                ("build_split2_ms", C.c_double), ("build_split2_launches", C.c_uint64)]


class IoStats(C.Structure):
    _fields_ = [("bytes", C.c_uint64), ("seconds", C.c_double), ("file_seconds", C.c_double),
                ("copy_wait_seconds", C.c_double), ("threads", C.c_int), ("devices", C.c_int),
                ("submit_seconds", C.c_double), ("starve_seconds", C.c_double), ("pool_seconds", C.c_double),
                ("chunk_bytes", C.c_uint64)]


# every symbol include/dyibf.h declares: name -> (restype, argtypes)
_P = C.c_void_p
_U8P, _U16P, _U32P, _U64P = _P, _P, _P, _P      # raw addresses (numpy .ctypes.data or device pointers)
SYMBOLS = {
    "dyibf_last_error": (C.c_char_p, []),
    "dyibf_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "dyibf_create": (C.c_int, [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, C.c_int, C.POINTER(_P)]),
    "dyibf_attach": (C.c_int, [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, C.c_int, _P, C.POINTER(_P)]),
    "dyibf_load_file": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(_P)]),
    "dyibf_load_file_expect": (C.c_int, [C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, C.c_int, C.POINTER(_P)]),
    "dyibf_peek_file": (C.c_int, [C.c_char_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)]),
    "dyibf_destroy": (None, [_P]),
    "dyibf_get_params": (C.c_int, [_P, C.POINTER(Params)]),
    "dyibf_size_mb": (C.c_double, [_P]),
    "dyibf_device_words": (_P, [_P]),
    "dyibf_upload_words": (C.c_int, [_P, _U64P, C.c_uint64, C.c_uint64]),
    "dyibf_download_words": (C.c_int, [_P, _U64P, C.c_uint64, C.c_uint64]),
    "dyibf_save_file": (C.c_int, [_P, C.c_char_p]),
    "dyibf_add_sequences": (C.c_int, [_P, _U8P, _U64P, C.c_uint32, _U32P, C.c_int, C.c_int, _P]),
    "dyibf_clear_bins": (C.c_int, [_P, _U32P, C.c_uint32]),
    "dyibf_threshold": (C.c_uint16, [C.c_int, C.c_uint32, C.c_uint16]),
    "dyibf_threshold_rate": (C.c_uint16, [C.c_float, C.c_uint32, C.c_uint16]),
    "dyibf_query_batch": (C.c_int, [_P, _U8P, _U64P, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_uint32,
                                    _U64P, _U16P, C.c_int, _P]),
    "dyibf_compact": (C.c_int, [_P, _U64P, C.c_uint32, _U64P, _U32P, C.c_uint64, C.POINTER(C.c_uint64), C.c_int, _P]),
    "dyibf_compact_static": (C.c_int, [C.c_int, C.c_uint32, _U64P, C.c_uint32, _U64P, _U32P, C.c_uint64, C.POINTER(C.c_uint64), _P]),
    "dyibf_classify": (C.c_int, [_P, _U8P, _U64P, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_uint32,
                                 _U64P, _U32P, C.c_uint64, C.POINTER(C.c_uint64)]),
    "dyibf_classify_fetch": (C.c_int, [_P, _U64P, _U32P, C.c_uint64, C.POINTER(C.c_uint64)]),
    "dyibf_sorted_bins": (C.c_int, [_U64P, C.c_uint32, _U32P]),
    "dyibf_gather_reads": (C.c_int, [_P, C.c_uint32, _U8P, C.c_uint64, _U64P, _U32P, C.c_uint64,
                                     C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "dyibf_parse_records": (C.c_int, [C.c_int, _P, C.c_uint64, C.c_int, C.c_int, _U8P, C.c_uint64, _U64P, C.c_uint64, C.c_uint64,
                                      C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_int, _P]),
    "dyibf_classify_text": (C.c_int, [_P, _P, C.c_uint64, _P, C.c_uint64, C.c_int, C.c_int, _U64P, _U32P, C.c_uint64,
                                      C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "dyibf_classify_text_begin": (C.c_int, [_P, _P, C.c_uint64, _P, C.c_uint64, C.c_int, C.c_int, C.POINTER(C.c_uint32),
                                            C.POINTER(C.c_uint32), C.POINTER(C.c_void_p)]),
    "dyibf_add_fasta_texts": (C.c_int, [_P, _P, _P, _P, _P, C.c_uint32, C.POINTER(C.c_uint32)]),
    "dyibf_set_option": (C.c_int, [_P, C.c_int, C.c_int64]),
    "dyibf_get_option": (C.c_int, [_P, C.c_int, C.POINTER(C.c_int64)]),
    "dyibf_profile_enable": (C.c_int, [_P, C.c_int]),
    "dyibf_profile_get": (C.c_int, [_P, C.POINTER(Profile), C.c_int]),
    "dyibf_launch_count": (C.c_uint64, [_P]),
    "dyibf_device_errors": (C.c_int, [_P, C.POINTER(C.c_uint32), C.c_int]),
    "dyibf_set_io_threads": (C.c_int, [C.c_int]),
    "dyibf_set_io_chunk_mib": (C.c_int, [C.c_int]),
    "dyibf_last_io_stats": (C.c_int, [C.POINTER(IoStats)]),
    "dyibf_classify_begin": (C.c_int, [_P, _U8P, _U64P, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_uint32, C.POINTER(_P)]),
    "dyibf_classify_poll": (C.c_int, [_P, C.POINTER(C.c_int)]),
    "dyibf_classify_end": (C.c_int, [_P, _U64P, _U32P, C.c_uint64, C.POINTER(C.c_uint64)]),
    "dyibf_classify_cancel": (None, [_P]),
    "dyibf_shard_bins": (C.c_int, [C.c_uint32, C.c_int, C.c_int, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "dyibf_shard_reads": (C.c_int, [C.c_uint32, C.c_int, C.c_int, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "dyibf_group_create": (C.c_int, [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, C.POINTER(C.c_int), C.c_int, C.c_int, C.POINTER(_P)]),
    "dyibf_group_load_file": (C.c_int, [C.c_char_p, C.POINTER(C.c_int), C.c_int, C.c_int, C.POINTER(_P)]),
    "dyibf_group_load_file_expect": (C.c_int, [C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, C.POINTER(C.c_int), C.c_int, C.c_int, C.POINTER(_P)]),
    "dyibf_group_destroy": (None, [_P]),
    "dyibf_group_get_params": (C.c_int, [_P, C.POINTER(Params), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "dyibf_group_member": (C.c_int, [_P, C.c_int, C.POINTER(_P), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "dyibf_group_set_option": (C.c_int, [_P, C.c_int, C.c_int64]),
    "dyibf_group_save_file": (C.c_int, [_P, C.c_char_p]),
    "dyibf_broadcast_filter": (C.c_int, [_P, C.c_int]),
    "dyibf_group_classify": (C.c_int, [_P, _U8P, _U64P, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_uint32,
                                       _U64P, _U32P, C.c_uint64, C.POINTER(C.c_uint64)]),
    "dyibf_group_classify_fetch": (C.c_int, [_P, _U64P, _U32P, C.c_uint64, C.POINTER(C.c_uint64)]),
    "dyibf_group_classify_begin": (C.c_int, [_P, _U8P, _U64P, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_uint32, C.POINTER(_P)]),
    "dyibf_group_add_sequences": (C.c_int, [_P, _U8P, _U64P, C.c_uint32, _U32P, C.c_int]),
    "dyibf_group_clear_bins": (C.c_int, [_P, _U32P, C.c_uint32]),
    "dyibf_host_alloc": (C.c_int, [C.POINTER(_P), C.c_size_t]),
    "dyibf_host_free": (None, [_P]),
}

_lib = None


class DyibfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("dyibf error %d: %s" % (code, msg))
        self.code = code


def lib():
    """The loaded library. Raises if libdyibf.so has not been built (python -m dream_yara_b200.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("%s is missing: build it with `python -m dream_yara_b200.build` "
                               "(the IBF path is CUDA-only; there is no CPU fallback)" % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc):
    if rc != OK:
        raise DyibfError(rc, lib().dyibf_last_error().decode("utf-8", "replace"))
    return rc

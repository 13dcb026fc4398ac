"""ctypes binding of the C-ABI library (`include/fls_b200.h`) and its in-tree build.

The library is plain `nvcc`-built code (no torch types cross the boundary): torch is used by the
host code above it only for device memory and streams.  There is no fallback: if the shared
object is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(CSRC, "libfls_b200.so")
SOURCES = ["api.cu", "gemm_sm100.cu", "gemm_wide_sm100.cu", "elementwise.cu", "attention.cu", "attention_sm100.cu", "host_cache.cpp",
           "noalloc.cu"]
HEADERS = ["common.cuh", "ptx.cuh", "gemm_tile.cuh", "host_cache.h", os.path.join("..", "..", "include", "fls_b200.h")]
NVCC_FLAGS = ["-shared", "-Xcompiler", "-fPIC", "-std=c++17", "-O3", "-lineinfo",
              "-gencode", "arch=compute_100a,code=sm_100a"]


class FlsError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"fls_b200 error {code}: {msg}")
        self.code = code


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into csrc/libfls_b200.so (cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, s) for s in SOURCES]
    deps = srcs + [os.path.join(CSRC, h) for h in HEADERS]
    if not force and os.path.exists(LIB_PATH):
        if os.path.getmtime(LIB_PATH) >= max(os.path.getmtime(d) for d in deps):
            return LIB_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if not os.path.exists(nvcc):
        nvcc = "nvcc"
    cmd = [nvcc, *NVCC_FLAGS, "-o", LIB_PATH + ".tmp", *srcs]
    if verbose:
        print(" ".join(cmd))
    subprocess.run(cmd, check=True, cwd=CSRC)
    os.replace(LIB_PATH + ".tmp", LIB_PATH)
    return LIB_PATH


class ModelDesc(C.Structure):
    _fields_ = [("hidden_size", C.c_int32), ("intermediate_size", C.c_int32),
                ("num_attention_heads", C.c_int32), ("num_key_value_heads", C.c_int32),
                ("head_dim", C.c_int32), ("vocab_size", C.c_int32), ("num_hidden_layers", C.c_int32),
                ("rms_norm_eps", C.c_float)]


class Stats(C.Structure):
    _fields_ = [("kernel_launches", C.c_uint64), ("h2d_bytes", C.c_uint64), ("h2d_ms", C.c_double),
                ("file_bytes_read", C.c_uint64), ("pinned_bytes", C.c_uint64), ("workspace_bytes", C.c_uint64),
                ("p2p_bytes", C.c_uint64), ("h2d_copies", C.c_uint64), ("mapped_bytes", C.c_uint64)]


class KernelProfile(C.Structure):
    _fields_ = [("ms", C.c_double), ("work", C.c_double), ("launches", C.c_uint64)]


PROF_CLASSES = ["gemm_qkv", "gemm_o", "gemm_gate_up", "gemm_down", "gemm_head", "attention", "rmsnorm", "rope",
                "embed", "other"]
LAYER_EMBED, LAYER_DECODER, LAYER_NORM, LAYER_HEAD = 0, 1, 2, 3
EPI_STORE, EPI_RESIDUAL, EPI_SWIGLU = 0, 1, 2
KV_NONE, KV_SAVE, KV_LOAD = 0, 1, 2
ATTN_TILE_INTS = 8  # sizeof(fls_attn_tile) / 4
IPC_HANDLE_BYTES = 64  # FLS_IPC_HANDLE_BYTES

_P = C.c_void_p
_I = C.c_int
_SIGNATURES = {
    # name: (restype, argtypes)            -- must list every symbol include/fls_b200.h declares
    "fls_version": (_I, []),
    "fls_last_error": (C.c_char_p, [_P]),
    "fls_create": (_I, [C.POINTER(ModelDesc), _I, C.POINTER(_P)]),
    "fls_destroy": (_I, [_P]),
    "fls_configure": (_I, [_P, _I, C.c_size_t, _I, _I]),
    "fls_set_rope_table": (_I, [_P, _P, _P, _I]),
    "fls_register_layer_file": (_I, [_P, _I, _I, C.c_char_p, _I]),
    "fls_layer_slot_bytes": (_I, [_P, _I, C.POINTER(C.c_size_t)]),
    "fls_prefetch_shard": (_I, [_P, _I, C.POINTER(C.c_int32), _I]),
    "fls_prefetch_pump": (_I, [_P, _I, C.c_size_t, _P]),
    "fls_acquire_shard": (_I, [_P, _I, _P]),
    "fls_release_shard": (_I, [_P, _I, _P]),
    "fls_preload_layer_file": (_I, [C.c_char_p]),
    "fls_set_host_cache_limit": (_I, [C.c_size_t]),
    "fls_trim_host_cache": (_I, [C.c_size_t]),
    "fls_set_host_image_mode": (_I, [_I]),
    "fls_set_stage_dir": (_I, [C.c_char_p]),
    "fls_peer_copy": (_I, [_P, _P, _I, _P, C.c_size_t, _P]),
    "fls_ipc_alloc": (_I, [_P, C.c_size_t, C.POINTER(_P), _P]),
    "fls_ipc_free": (_I, [_P, _P]),
    "fls_ipc_open": (_I, [_P, _P, C.POINTER(_P)]),
    "fls_ipc_close": (_I, [_P, _P]),
    "fls_flag_signal": (_I, [_P, _P, C.c_uint32, _P]),
    "fls_flag_wait": (_I, [_P, _P, C.c_uint32, C.c_uint32, _P]),
    "fls_flag_timeouts": (_I, [_P, C.POINTER(C.c_uint32)]),
    "fls_embed": (_I, [_P, _I, _I, _P, _I, _P, _P]),
    "fls_decoder_layer": (_I, [_P, _I, _I, _P, _I, _P, _P, _I, _P]),
    "fls_decoder_layer_kv": (_I, [_P, _I, _I, _P, _I, _P, _P, _I, _I, _P, _I, _P, _P]),
    "fls_final_norm": (_I, [_P, _I, _I, _P, _P, _I, _P, _P]),
    "fls_lm_head": (_I, [_P, _I, _I, _P, _I, _P, _P, _I, _P, _P, _P]),
    "fls_op_gemm": (_I, [_P, _I, _P, _I, _P, _I, _I, _I, _I, _I, _P, _I, _P, _I, _I, _P]),
    "fls_op_rmsnorm": (_I, [_P, _P, _P, _P, _P, C.c_float, _I, _I, _P, _P, _P]),
    "fls_op_rope": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, _P]),
    "fls_op_attention": (_I, [_P, _P, _I, _I, _I, _I, _I, _P, _I, _P, _I, _P]),
    "fls_op_swiglu": (_I, [_P, _P, _P, _I, _I, _I, _P, _I, _P]),
    "fls_op_embed": (_I, [_P, _P, _P, _I, _I, _P, _P]),
    "fls_op_head": (_I, [_P, _P, _P, _I, _I, _I, _P, _P, _I, _P, _P, _P]),
    "fls_op_head_noalloc": (_I, [_P, _P, _I, _I, _I, _P, _P, _I, _P, _P, _P]),
    "fls_set_option": (_I, [_P, C.c_char_p, _I]),
    "fls_set_profiling": (_I, [_P, _I]),
    "fls_get_profile": (_I, [_P, _P]),
    "fls_get_stats": (_I, [_P, C.POINTER(Stats)]),
    "fls_reset_stats": (_I, [_P]),
}

_lib = None
_lib_lock = threading.Lock()


def load() -> C.CDLL:
    """dlopen the library and type every entry point.  Raises if the library is missing."""
    global _lib
    with _lib_lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise FileNotFoundError(
                    f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                    "(there is no non-CUDA fallback)")
            lib = C.CDLL(LIB_PATH)
            for name, (res, args) in _SIGNATURES.items():
                fn = getattr(lib, name)  # AttributeError if the symbol is not exported
                fn.restype, fn.argtypes = res, args
            _lib = lib
    return _lib


def exported_symbols() -> list[str]:
    return list(_SIGNATURES)


class Context:
    """Thin owner of one `fls_ctx*` (one per GPU)."""

    def __init__(self, desc: ModelDesc, device: int):
        self.lib = load()
        self._h = _P()
        self.device = device
        rc = self.lib.fls_create(C.byref(desc), device, C.byref(self._h))
        if rc != 0:
            raise FlsError(rc, self.lib.fls_last_error(None).decode())

    def call(self, name: str, *args):
        rc = getattr(self.lib, name)(self._h, *args)
        if rc != 0:
            raise FlsError(rc, self.lib.fls_last_error(self._h).decode())

    def stats(self) -> Stats:
        s = Stats()
        self.call("fls_get_stats", C.byref(s))
        return s

    def profile(self) -> dict:
        arr = (KernelProfile * len(PROF_CLASSES))()
        self.call("fls_get_profile", C.cast(arr, _P))
        return {n: dict(ms=arr[i].ms, work=arr[i].work, launches=int(arr[i].launches))
                for i, n in enumerate(PROF_CLASSES)}

    def close(self):
        if self._h:
            self.lib.fls_destroy(self._h)
            self._h = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

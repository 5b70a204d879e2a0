"""ctypes binding of libpa_b200.so - the C ABI declared in include/pa_b200.h.

There is NO fallback: if the library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import (POINTER, Structure, c_char_p, c_int, c_int32, c_int64, c_size_t, c_void_p)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpa_b200.so")

ABI_VERSION = 3          # PA_B200_ABI_VERSION of include/pa_b200.h
PA_BF16 = 0
PA_F32 = 1


class PaKvCache(Structure):
    """Mirror of `struct PaKvCache` (include/pa_b200.h)."""
    _fields_ = [
        ("k", c_void_p),
        ("v", c_void_p),
        ("stride_block", c_int64),
        ("stride_head", c_int64),
        ("stride_slot", c_int64),
        ("num_blocks", c_int32),
        ("num_kv_heads", c_int32),
        ("block_size", c_int32),
        ("head_dim", c_int32),
        ("dtype", c_int32),
        ("reserved", c_int32),
    ]


class PaPeerGather(Structure):
    """Mirror of `struct PaPeerGather` (include/pa_b200.h)."""
    _fields_ = [
        ("num_remote", c_int32),
        ("remote_out", c_void_p * 7),
        ("remote_sig", c_void_p * 7),
        ("local_sig", c_void_p),
        ("ll_epoch", c_void_p),
    ]


# name -> (restype, argtypes); one entry per symbol declared in include/pa_b200.h (the drop-in ABI)
SIGNATURES = {
    "pa_abi_version": (c_int, []),
    "pa_status_string": (c_char_p, [c_int]),
    "pa_decode_num_workers": (c_int, []),
    "pa_decode_workspace_bytes": (c_size_t, [c_int32, c_int32, c_int32, c_int32]),
    "pa_decode_workspace_init": (c_int, [c_void_p, c_size_t, c_void_p]),
    "pa_decode": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, POINTER(PaKvCache),
                          c_void_p, c_int32, c_void_p, c_int32, c_int32, c_void_p, c_size_t, c_void_p]),
    # no workspace argument: the library's static device array (PA_DECODE_STATIC_BYTES), see include/pa_b200.h
    "pa_decode_static": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, POINTER(PaKvCache),
                                 c_void_p, c_int32, c_void_p, c_int32, c_int32, c_void_p]),
    "pa_decode_gather": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, POINTER(PaKvCache),
                                 c_void_p, c_int32, c_void_p, c_int32, c_int32, c_void_p, c_size_t,
                                 POINTER(PaPeerGather), c_void_p]),
    "pa_peer_wait": (c_int, [c_void_p, c_int32, ctypes.c_uint32, c_int64, c_void_p, c_void_p]),
    "pa_peer_wait_next": (c_int, [c_void_p, c_int32, c_void_p, ctypes.c_uint32, c_int64, c_void_p, c_void_p]),
    "pa_peer_unpack": (c_int, [c_void_p, c_void_p, c_int32, c_int32, c_int32, c_void_p, c_void_p, c_int64,
                               c_void_p, c_void_p]),
    "pa_peer_alloc": (c_int, [c_size_t, POINTER(c_void_p), c_void_p]),
    "pa_peer_open": (c_int, [c_void_p, POINTER(c_void_p)]),
    "pa_peer_close": (c_int, [c_void_p]),
    "pa_peer_free": (c_int, [c_void_p]),
    "pa_slot_mapping": (c_int, [c_void_p, c_int32, c_int32, c_int32, c_void_p, c_void_p]),
    "pa_block_table_step": (c_int, [c_void_p, c_int32, c_int32, c_void_p, c_void_p, c_int32, c_void_p, c_void_p,
                                    c_int32, c_int32, c_void_p, c_int32, c_void_p]),
    "pa_reshape_and_cache": (c_int, [c_void_p, c_void_p, c_int64, c_int64, POINTER(PaKvCache), c_void_p,
                                     c_int32, c_void_p]),
    "pa_rope_and_cache": (c_int, [c_void_p, c_int64, c_int32, c_void_p, c_void_p, c_int64, c_int32, c_void_p,
                                  c_void_p, c_int32, c_void_p, c_int64, c_int64, POINTER(PaKvCache), c_void_p]),
    "pa_copy_blocks": (c_int, [c_void_p, c_int64, c_int32, c_int64, c_int64, c_void_p, c_int32, c_void_p]),
    "pa_swap_blocks": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, c_int32, c_int64,
                               c_void_p, c_int32, c_int32, c_void_p]),
    "pa_swap_blocks_staged": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, c_int32, c_int64,
                                      c_void_p, c_void_p, c_int32, c_int32, c_void_p, c_int64, c_void_p]),
    "pa_prefill_workspace_bytes": (c_size_t, [c_int32, c_int32, c_int32, c_int32]),
    "pa_prefill": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, POINTER(PaKvCache),
                           c_void_p, c_int32, c_int32, c_int32, c_void_p, c_size_t, c_void_p]),
    "pa_prefill_varlen": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, POINTER(PaKvCache),
                                  c_void_p, c_int32, c_int32, c_void_p, c_void_p, c_int32, c_int32, c_int32, c_void_p]),
    "pa_prefill_tc_available": (c_int, [POINTER(PaKvCache), c_int32]),
}

# test / bring-up / measurement hooks declared in include/pa_b200_debug.h (process-global state;
# not part of the drop-in ABI)
DEBUG_SIGNATURES = {
    "pa_decode_force_path": (c_int, [c_int]),
    "pa_prefill_force_path": (c_int, [c_int]),
    "pa_decode_set_tuning": (c_int, [c_int, c_int, c_int, c_int]),
    "pa_decode_set_pdl": (c_int, [c_int]),
    "pa_decode_debug_times": (c_int, [c_void_p]),
    "pa_prefill_debug_times": (c_int, [c_void_p]),
    "pa_profile_begin": (c_int, [c_int32]),
    "pa_profile_end": (c_int, [c_void_p, c_void_p, c_int32]),
}

_lib = None


def load() -> ctypes.CDLL:
    """Load the library once.  Fails loudly - the product path has no CPU fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m vllm_from_scratch_b200.build` "
            "(nvcc, sm_100a).  There is no CPU / PyTorch fallback for the paged-attention path.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in list(SIGNATURES.items()) + list(DEBUG_SIGNATURES.items()):
        fn = getattr(lib, name)           # AttributeError if a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.pa_abi_version() != ABI_VERSION:
        raise RuntimeError(f"libpa_b200.so ABI version {lib.pa_abi_version()} != {ABI_VERSION}: rebuild it "
                           "(python -m vllm_from_scratch_b200.build --force)")
    _lib = lib
    return lib


_ext = None


def load_ext():
    """The thin torch C++ extension over the same C ABI (csrc/torch_shim.cpp -> _pa_torch.so): tensors in, one
    extern "C" call out.  The hot per-layer entry points go through it (3 us of host time instead of ~20 us through
    ctypes); setup, peer plumbing, block copy / swap and the debug hooks stay on ctypes.  Fails loudly when missing."""
    global _ext
    if _ext is not None:
        return _ext
    load()
    try:
        from . import _pa_torch
    except ImportError as e:
        raise RuntimeError(f"vllm_from_scratch_b200/_pa_torch.so is missing or does not load ({e}): build it with "
                           "`python -m vllm_from_scratch_b200.build`.  There is no fallback binding.") from e
    if _pa_torch.abi_version() != ABI_VERSION:
        raise RuntimeError(f"_pa_torch.so was built against ABI {_pa_torch.abi_version()} != {ABI_VERSION}: rebuild it")
    _ext = _pa_torch
    return _ext


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().pa_status_string(rc).decode()
        raise RuntimeError(f"{what} failed: {msg} (status {rc})")

"""The C-ABI boundary without a GPU: libpa_b200.so builds / loads, exports every function
include/pa_b200.h declares, the ctypes table binds exactly that set, and argument validation
returns status codes (no compute call is made here)."""
import ctypes
import os
import re
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vllm_from_scratch_b200 import _lib  # noqa: E402
from vllm_from_scratch_b200.build import build_library  # noqa: E402

HEADER = os.path.join(ROOT, "include", "pa_b200.h")
DEBUG_HEADER = os.path.join(ROOT, "include", "pa_b200_debug.h")


def declared_functions(header=HEADER):
    src = open(header).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"^\s*(?:int|size_t|const char\*)\s+(pa_[a-z0-9_]+)\s*\(", src, flags=re.M)))


@pytest.fixture(scope="module")
def lib():
    return ctypes.CDLL(build_library())


def test_header_declares_the_expected_entry_points():
    names = declared_functions()
    for must in ("pa_decode", "pa_decode_gather", "pa_reshape_and_cache", "pa_slot_mapping", "pa_copy_blocks",
                 "pa_swap_blocks", "pa_swap_blocks_staged", "pa_prefill", "pa_prefill_varlen", "pa_rope_and_cache",
                 "pa_peer_wait", "pa_peer_wait_next", "pa_peer_unpack", "pa_decode_workspace_bytes"):
        assert must in names, must


def test_library_exports_every_declared_symbol(lib):
    for name in declared_functions() + declared_functions(DEBUG_HEADER):
        assert hasattr(lib, name), f"{name} is declared in include/ but not exported"


def test_ctypes_table_binds_exactly_the_declared_set():
    assert sorted(_lib.SIGNATURES) == declared_functions()
    assert sorted(_lib.DEBUG_SIGNATURES) == declared_functions(DEBUG_HEADER)


def test_process_global_hooks_are_not_in_the_public_abi():
    public = declared_functions()
    for name in ("pa_decode_force_path", "pa_prefill_force_path", "pa_decode_set_tuning", "pa_decode_debug_times",
                 "pa_prefill_debug_times", "pa_profile_begin", "pa_profile_end"):
        assert name not in public and name in _lib.DEBUG_SIGNATURES


def test_struct_mirrors_have_the_c_layout():
    assert ctypes.sizeof(_lib.PaKvCache) == 2 * 8 + 3 * 8 + 6 * 4
    assert ctypes.sizeof(_lib.PaPeerGather) == 8 + 7 * 8 + 7 * 8 + 8 + 8        # int32 + padding, 2 x 7 pointers, 2 pointers


def test_status_strings_and_argument_validation_without_a_gpu(lib):
    lib.pa_status_string.restype = ctypes.c_char_p
    assert lib.pa_abi_version() >= 1
    assert lib.pa_status_string(0) == b"ok" or lib.pa_status_string(0)
    # null / malformed arguments come back as negative status codes, never as a crash or an exception
    assert lib.pa_copy_blocks(None, 0, 1, 0, 16, None, 1, None) < 0
    assert lib.pa_swap_blocks(None, 0, 0, None, 0, 0, 1, 16, None, 1, 0, None) < 0
    assert lib.pa_slot_mapping(None, 0, 4, 16, None, None) < 0
    assert lib.pa_peer_wait(None, 2, 0, 10, None, None) < 0
    assert lib.pa_decode_gather(None, 0, 0, None, 0, 0, None, None, 0, None, 1, 1, None, 0, None, None) < 0
    # empty work is a no-op success
    assert lib.pa_copy_blocks(None, 0, 0, 0, 0, None, 0, None) == 0


def test_static_scratch_decode_validates_before_it_touches_a_device(lib):
    """pa_decode_static (no counterpart in the reference): the size check against PA_DECODE_STATIC_BYTES and the
    argument checks come before any CUDA call, so they answer without a GPU."""
    from vllm_from_scratch_b200._lib import PaKvCache, SIGNATURES
    res, args = SIGNATURES["pa_decode_static"]
    lib.pa_decode_static.restype, lib.pa_decode_static.argtypes = res, args
    lib.pa_decode_workspace_bytes.restype = ctypes.c_size_t
    lib.pa_decode_workspace_bytes.argtypes = [ctypes.c_int32] * 4
    lib.pa_status_string.restype = ctypes.c_char_p
    header = open(HEADER).read()
    limit = int(re.search(r"#define PA_DECODE_STATIC_BYTES \((\d+)u << (\d+)\)", header).group(1)) << 20
    assert limit == 96 << 20
    c = PaKvCache(16, 16, 8 * 16 * 128, 16 * 128, 128, 64, 8, 16, 128, 0, 0)
    assert lib.pa_decode_workspace_bytes(512, 64, 8, 128) > limit
    rc = lib.pa_decode_static(16, 64 * 128, 128, 16, 64 * 128, 128, ctypes.byref(c), 16, 1, 16, 512, 64, None)
    assert rc == -7 and b"static" in lib.pa_status_string(rc)
    assert lib.pa_decode_static(16, 1, 1, 16, 1, 1, None, 16, 1, 16, 1, 4, None) == -1          # no cache descriptor
    assert lib.pa_decode_static(None, 0, 0, None, 0, 0, ctypes.byref(c), None, 0, None, 0, 32, None) == 0   # empty batch
    assert lib.pa_decode_static(16, 1, 1, 16, 1, 1, ctypes.byref(c), 16, 1, 16, -3, 32, None) == -1


def test_torch_shim_loads_and_mirrors_the_error_contract():
    """The thin torch C++ extension (csrc/torch_shim.cpp) is built in-tree, reports the same ABI version and raises
    the host mirror's exception types without touching a device (CPU tensors: RuntimeError, no CPU fallback)."""
    import pytest
    import torch
    from vllm_from_scratch_b200 import _lib, paged_attention as P
    ext = _lib.load_ext()
    assert ext.abi_version() == _lib.ABI_VERSION
    for name in ("reshape_and_cache", "rope_and_cache", "decode", "prefill_varlen", "prefill_chunk", "block_table_step",
                 "prefill_tc_available", "decode_workspace_bytes"):
        assert callable(getattr(ext, name))
    assert ext.decode_workspace_bytes(64, 32, 8, 128) == _lib.load().pa_decode_workspace_bytes(64, 32, 8, 128)
    k = torch.zeros(4, 16, 2, 8)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        P.reshape_and_cache(torch.zeros(3, 2, 8), torch.zeros(3, 2, 8), (k, k.clone()), torch.arange(3))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        P.paged_attention_decode_batched(torch.zeros(1, 2, 8), (k, k.clone()), torch.zeros(1, 1, dtype=torch.int32),
                                         torch.ones(1, dtype=torch.int32), workspace=torch.zeros(8))
    with pytest.raises(TypeError):
        P.reshape_and_cache("not a tensor", torch.zeros(3, 2, 8), (k, k.clone()), torch.arange(3))
    # c10 errors map to the host mirror's exception types (and never carry a C++ backtrace)
    for kind, exc in ((0, RuntimeError), (1, ValueError), (2, TypeError), (3, IndexError), (4, RuntimeError)):
        with pytest.raises(exc) as info:
            ext._raise(kind, 7)
        assert "frame #" not in str(info.value)

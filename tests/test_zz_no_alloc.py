"""``pa_decode_static`` (include/pa_b200.h): pa_decode for callers that may not allocate device memory.

No counterpart in the reference.  The entry point takes its
split-K scratch from a file-scope device array inside libpa_b200.so instead of a caller workspace; everything
else is pa_decode.  The checks, all through the C ABI (ctypes, raw pointers):
  * two calls in a row, different inputs, each equal to ``pa_decode`` on the same inputs - on the fast path with
    split-K partials actually in play (long ragged contexts), and on the generic fp32 path;
  * free device memory does not move across a call (nothing is allocated);
  * a shape beyond PA_DECODE_STATIC_BYTES is refused with its own status and launches nothing.
Collected last on purpose (file name).
"""
import ctypes
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu

DEV = "cuda"
TOL = {torch.float32: 1e-5, torch.bfloat16: 2e-2}          # the tolerances of tests/test_gpu_parity.py
STATIC_BYTES = 96 << 20                                      # PA_DECODE_STATIC_BYTES


@pytest.fixture(scope="module")
def lib():
    from vllm_from_scratch_b200 import _lib
    return _lib.load()


def _case(seed, dtype, batch, hq, hkv, d, bs, lens):
    """A 'hbsd' cache with every sequence's blocks scattered over the pool, q, the table and the lengths."""
    g = torch.Generator().manual_seed(seed)
    blocks = [(n + bs - 1) // bs for n in lens]
    nb = sum(blocks) + 3
    kv = (torch.randn(2, nb, hkv, bs, d, generator=g) * 0.5).to(dtype).to(DEV)
    perm = torch.randperm(nb, generator=g).tolist()
    width = max(blocks)
    bt = torch.full((batch, width), -1, dtype=torch.int32)
    at = 0
    for i, n in enumerate(blocks):
        bt[i, :n] = torch.tensor(perm[at:at + n], dtype=torch.int32)
        at += n
    q = torch.randn(batch, hq, d, generator=g).to(dtype).to(DEV)
    return kv, q, bt.to(DEV), torch.tensor(lens, dtype=torch.int32, device=DEV)


def _desc(kv):
    from vllm_from_scratch_b200._lib import PaKvCache
    _, nb, h, bs, d = kv.shape
    return PaKvCache(kv[0].data_ptr(), kv[1].data_ptr(), kv.stride(1), kv.stride(2), kv.stride(3), nb, h, bs, d,
                     0 if kv.dtype == torch.bfloat16 else 1, 0)


def _with_workspace(lib, kv, q, bt, lens, ws):
    out = torch.empty_like(q)
    c = _desc(kv)
    rc = lib.pa_decode(q.data_ptr(), q.stride(0), q.stride(1), out.data_ptr(), out.stride(0), out.stride(1),
                       ctypes.byref(c), bt.data_ptr(), bt.stride(0), lens.data_ptr(), q.shape[0], q.shape[1],
                       ws.data_ptr(), ws.numel(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.pa_status_string(rc)
    return out


def _static(lib, kv, q, bt, lens, out):
    c = _desc(kv)
    return lib.pa_decode_static(q.data_ptr(), q.stride(0), q.stride(1), out.data_ptr(), out.stride(0), out.stride(1),
                                ctypes.byref(c), bt.data_ptr(), bt.stride(0), lens.data_ptr(), q.shape[0], q.shape[1],
                                torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("dtype,hq,hkv,d,bs,lens_a,lens_b", [
    # fast path (bf16, BS 16, D 128): few long ragged sequences -> every segment is split, partials + folds in scratch
    (torch.bfloat16, 32, 8, 128, 16, [9000, 1, 4097, 777], [33, 12001, 16, 2500, 5000]),
    # fast path at head_dim 64, GQA 8
    (torch.bfloat16, 16, 2, 64, 16, [6000, 300], [15, 7000, 4096]),
    # generic streaming path (fp32, the reference's dtype): per-worker partials + the reduce kernel
    (torch.float32, 8, 4, 32, 16, [3000, 5, 640], [1, 2049, 100, 4000]),
], ids=["bf16-d128", "bf16-d64", "f32-generic"])
def test_static_scratch_decode_matches_pa_decode_on_two_calls_in_a_row(lib, dtype, hq, hkv, d, bs, lens_a, lens_b):
    cases = [_case(11, dtype, len(lens_a), hq, hkv, d, bs, lens_a), _case(12, dtype, len(lens_b), hq, hkv, d, bs, lens_b)]
    # expected: the existing entry point with a caller-owned, initialised workspace
    need = max(lib.pa_decode_workspace_bytes(c[1].shape[0], hq, hkv, d) for c in cases)
    assert need <= STATIC_BYTES
    ws = torch.zeros(need, dtype=torch.uint8, device=DEV)
    want = [_with_workspace(lib, *c, ws) for c in cases]
    outs = [torch.full_like(c[1], float("nan")) for c in cases]
    torch.cuda.synchronize()

    # two calls back to back on one stream, nothing in between: the second runs on what the first left behind
    free0 = torch.cuda.mem_get_info()[0]
    rcs = [_static(lib, *c, o) for c, o in zip(cases, outs)]
    torch.cuda.synchronize()
    assert rcs == [0, 0], [lib.pa_status_string(r) for r in rcs]
    for got, ref in zip(outs, want):
        assert torch.isfinite(got.float()).all()
        assert (got.float() - ref.float()).abs().max().item() <= TOL[dtype]
    # ... and a third time over the first inputs: still the same answer
    again = torch.full_like(outs[0], float("nan"))
    assert _static(lib, *cases[0], again) == 0
    torch.cuda.synchronize()
    assert (again.float() - want[0].float()).abs().max().item() <= TOL[dtype]
    # the module (and with it the array) is resident after the first call: later calls must not move free memory
    free1 = torch.cuda.mem_get_info()[0]
    assert _static(lib, *cases[1], outs[1]) == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free1
    assert free1 <= free0


def test_static_scratch_decode_refuses_shapes_beyond_its_array(lib):
    hq, hkv, d, batch = 64, 8, 128, 512
    assert lib.pa_decode_workspace_bytes(batch, hq, hkv, d) > STATIC_BYTES
    kv, q, bt, lens = _case(3, torch.bfloat16, 2, hq, hkv, d, 16, [40, 17])
    c = _desc(kv)
    out = torch.full((batch, hq, d), 7.0, dtype=torch.bfloat16, device=DEV)
    # pointers are never dereferenced: the size check comes before any launch
    rc = lib.pa_decode_static(q.data_ptr(), hq * d, d, out.data_ptr(), hq * d, d, ctypes.byref(c), bt.data_ptr(),
                              bt.stride(0), lens.data_ptr(), batch, hq, None)
    torch.cuda.synchronize()
    assert rc == -7 and b"static" in lib.pa_status_string(rc)
    assert (out == 7.0).all()
    # argument validation is pa_decode's
    assert lib.pa_decode_static(q.data_ptr(), hq * d, d, out.data_ptr(), hq * d, d, None, bt.data_ptr(), 1,
                                lens.data_ptr(), 2, hq, None) == -1
    assert lib.pa_decode_static(None, 0, 0, None, 0, 0, ctypes.byref(c), None, 0, None, 0, hq, None) == 0   # empty batch

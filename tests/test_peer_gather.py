"""Fused output gather of KV-head-sharded decode (pa_decode_gather / pa_peer_wait).

One GPU is enough to pin the kernel side: with world 1 the decode kernel still counts every
completed (sequence, KV head) segment into the arrival counter through all three code paths
(whole-segment units, split segments merged by the finisher warps, empty sequences), the wait
kernel orders the consumer behind it, and the outputs must be the plain decode's bits.  The
multi-rank form (NVLink peer stores) runs under torchrun: tools/peer_gather_check.py.
"""
import ctypes
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _case(lens, hq=32, hkv=8, d=128, bs=16, seed=0):
    g = torch.Generator().manual_seed(seed)
    lens = torch.tensor(lens, dtype=torch.int32)
    nblk = (lens + bs - 1) // bs
    nb = int(nblk.sum()) + 8
    kv = torch.randn(2, nb, hkv, bs, d, generator=g).to(torch.bfloat16).to(DEV)
    q = torch.randn(len(lens), hq, d, generator=g).to(torch.bfloat16).to(DEV)
    perm = torch.randperm(nb, generator=g).to(torch.int32)
    bts = torch.zeros(len(lens), max(int(nblk.max()), 1), dtype=torch.int32)
    at = 0
    for i, n in enumerate(nblk.tolist()):
        bts[i, :n] = perm[at:at + n]
        at += n
    return q, kv, bts.to(DEV), lens.to(DEV)


@pytest.mark.parametrize("protocol", ["counter", "ll"])
@pytest.mark.parametrize("lens", [[2048] * 16, [0, 1, 17, 5000, 0, 333, 40000, 16], [100] * 3], ids=["whole", "mixed", "short"])
def test_world1_fused_gather_counts_segments_and_matches_plain_decode(lens, protocol):
    from vllm_from_scratch_b200 import paged_attention as P
    from vllm_from_scratch_b200 import sharding as S
    q, kv, bts, sl = _case(lens)
    want = P.paged_attention_decode_batched(q, kv, bts, sl)
    pg = S.PeerGather(len(lens), 32, 8, 128, DEV, world=1, rank=0, timeout_us=500_000, protocol=protocol)
    try:
        for step in range(1, 4):
            got = pg.decode(q, kv, bts, sl)
            assert got.shape == (1, len(lens), 32, 128)
            assert torch.equal(pg.full(got), want), f"step {step}"
            pg.check()                                        # no timeout: every segment was counted
            if protocol == "counter":
                assert int(pg.sig[0].item()) == step * len(lens) * 8
            assert int(pg.epoch.item()) == step                # the step number advanced on the device
        assert pg.local_out().data_ptr() != got[0].data_ptr()   # the two buffer sets alternate
    finally:
        pg.close()


def test_peer_wait_times_out_instead_of_hanging():
    from vllm_from_scratch_b200 import _lib
    lib = _lib.load()
    sig = torch.zeros(8, dtype=torch.int32, device=DEV)
    status = torch.zeros(1, dtype=torch.int32, device=DEV)
    sig[0] = 5
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.pa_peer_wait(sig.data_ptr(), 2, 5, 2000, status.data_ptr(), st), "pa_peer_wait")
    torch.cuda.synchronize()
    assert int(status.item()) == 2                            # rank 1 never arrived -> 1 + 1
    # once failed, later waits return at once
    _lib.check(lib.pa_peer_wait(sig.data_ptr(), 2, 5, 5_000_000, status.data_ptr(), st), "pa_peer_wait")
    torch.cuda.synchronize()
    status.zero_()
    sig[1] = 7                                                # counters are monotonic: >= expected passes
    _lib.check(lib.pa_peer_wait(sig.data_ptr(), 2, 5, 2000, status.data_ptr(), st), "pa_peer_wait")
    torch.cuda.synchronize()
    assert int(status.item()) == 0


def test_gather_refuses_shapes_outside_the_fast_path():
    from vllm_from_scratch_b200 import _lib
    from vllm_from_scratch_b200 import paged_attention as P
    q = torch.randn(2, 4, 32, device=DEV)
    kv = torch.randn(2, 8, 4, 16, 32, device=DEV)             # fp32, D 32: generic path only
    bts = torch.zeros(2, 2, dtype=torch.int32, device=DEV)
    sl = torch.tensor([5, 20], dtype=torch.int32, device=DEV)
    sig = torch.zeros(1, dtype=torch.int32, device=DEV)
    pg = _lib.PaPeerGather()
    pg.num_remote = 0
    pg.local_sig = sig.data_ptr()
    with pytest.raises(RuntimeError, match="pa_decode_gather"):
        P.paged_attention_decode_batched(q, kv, bts, sl, peers=pg)


def test_ll_unpack_polls_flags_and_times_out():
    """pa_peer_unpack on hand-made LL words: a 'remote' slot whose words carry this step's flag is
    unpacked; a word with a stale flag makes the kernel give up after the timeout and name the rank."""
    from vllm_from_scratch_b200 import _lib
    lib = _lib.load()
    world, rank, words = 3, 1, 1000
    g = torch.Generator().manual_seed(3)
    data = torch.randint(-2**31, 2**31 - 1, (world, words), generator=g, dtype=torch.int64).to(torch.int32)
    ll = torch.zeros(world, words, 2, dtype=torch.int32)
    ll[:, :, 0] = data
    ll[:, :, 1] = 5                                           # flag of step 5 = epoch 4 + 1
    ll[rank] = -1                                             # the local LL slot is never read
    ll_d = ll.to(DEV)
    gathered = torch.full((world, words), 77, dtype=torch.int32, device=DEV)
    epoch = torch.tensor([4], dtype=torch.int32, device=DEV)
    done = torch.zeros(1, dtype=torch.int32, device=DEV)
    status = torch.zeros(1, dtype=torch.int32, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.pa_peer_unpack(ll_d.data_ptr(), gathered.data_ptr(), world, rank, words, epoch.data_ptr(),
                                  done.data_ptr(), 200_000, status.data_ptr(), st), "pa_peer_unpack")
    torch.cuda.synchronize()
    assert int(status.item()) == 0 and int(epoch.item()) == 5 and int(done.item()) == 0
    got = gathered.cpu()
    assert torch.equal(got[0], data[0]) and torch.equal(got[2], data[2])
    assert (got[rank] == 77).all()                            # local slot untouched
    # next step: rank 2's word 10 still carries the old flag -> bounded wait, status names rank 2
    ll_d[:, :, 1] = 6
    ll_d[2, 10, 1] = 5
    _lib.check(lib.pa_peer_unpack(ll_d.data_ptr(), gathered.data_ptr(), world, rank, words, epoch.data_ptr(),
                                  done.data_ptr(), 3000, status.data_ptr(), st), "pa_peer_unpack")
    torch.cuda.synchronize()
    assert int(status.item()) == 3

"""Worker of tests/test_multi_gpu.py: multi-rank parity of KV-head-sharded decode.

Launched as  python -m torch.distributed.run --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 \
             tests/multi_gpu_worker.py [--steps N] [--out PATH]
one rank per GPU.  Every rank builds the SAME full-head batch (same seed), keeps its KV-head slice
``[2, NB, Hkv / W, BS, D]`` and the matching query heads, and runs the three gather implementations:

  * ``PeerGather(protocol='ll')``      decode kernel stores {data, step} words into every peer over NVLink
  * ``PeerGather(protocol='counter')`` plain peer stores + fence + arrival counters
  * ``HeadGather``                     decode + one NCCL all_gather_into_tensor   (the comparison arm)

Checked every step, on every rank:
  (1) TRANSPORT IS BIT-EXACT: gathered[r] == the bits rank r's own plain ``pa_decode`` produced for its head
      slice (collected independently with an NCCL all-gather of the plain outputs) - the fusion, the LL
      packing, the buffer-set alternation and the receive side may not change one bit;
  (2) every rank holds the same gathered buffer (follows from (1));
  (3) the gathered ``[B, Hq, D]`` equals the single-GPU decode of the FULL-head batch within the
      scale-aware bf16 bound of tests/test_decode_strict.py and vs the fp64 oracle on rank 0.  (Bit-equality
      with the full-head launch is not a property of split-K: the unit sizes are chosen from the local block
      count, so a rank with Hkv / W heads merges its partials at other boundaries than the 8-head launch;
      world 1 IS bit-equal and is asserted.)
Cases: ragged lengths incl. ``seq_len == 0``, 1, non-multiples of 16 and a 32k sequence; the C3 per-rank shape
(8 x 32768); ``--steps`` consecutive steps (default 50) with rank-dependent host delays so the ranks drift a
step apart (two buffer sets alternate; the LL step number wraps its parity bit every step); CUDA-graph replay
of the captured (decode + receive) pair as bench.py uses it; finally a DEAD PEER: one rank skips a step and the
others must report it through ``PeerGather.check()`` within the timeout instead of hanging.
"""
import argparse
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import paged_attention_oracle as O  # noqa: E402
from vllm_from_scratch_b200 import paged_attention as P  # noqa: E402
from vllm_from_scratch_b200 import sharding as S  # noqa: E402

HQ, HKV, D, BS = 32, 8, 128, 16


def build(lens, seed, dev):
    g = torch.Generator().manual_seed(seed)
    sl = torch.tensor(lens, dtype=torch.int32)
    nblk = (sl + BS - 1) // BS
    nb = int(nblk.sum()) + 8
    gd = torch.Generator(device=dev).manual_seed(seed)
    kv = torch.randn(2, nb, HKV, BS, D, generator=gd, device=dev, dtype=torch.float32).to(torch.bfloat16)
    perm = torch.randperm(nb, generator=g).to(torch.int32)
    bts = torch.zeros(len(lens), max(int(nblk.max()), 1), dtype=torch.int32)
    at = 0
    for i, n in enumerate(nblk.tolist()):
        bts[i, :n] = perm[at:at + n]
        at += n
    return kv, bts.to(dev), sl.to(dev), g


def strict_ok(got, ref):
    return bool(((got.double() - ref.double()).abs() <= 2e-3 + 2.0 ** -7 * ref.double().abs()).all())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    lines, ok = [], True

    def note(msg):
        lines.append(f"[rank {rank}/{world}] {msg}")

    kvh, qh = S.shard_kv_heads(HKV, world, rank), S.shard_q_heads(HQ, HKV, world, rank)
    cases = {"ragged": [0, 1, 17, 5000, 333, 32768, 16, 2048, 0, 4095], "c3_8x32k": [32768] * 8}
    for cname, lens in cases.items():
        b = len(lens)
        kv, bts, sl, g = build(lens, 7 + len(lens), dev)
        kv_local = kv[:, :, kvh.start:kvh.stop].contiguous()
        steps = args.steps if cname == "ragged" else max(6, args.steps // 8)
        arms = [("peer-ll", lambda: S.PeerGather(b, HQ, HKV, D, dev, world=world, rank=rank, timeout_us=10_000_000, protocol="ll")),
                ("peer-counter", lambda: S.PeerGather(b, HQ, HKV, D, dev, world=world, rank=rank, timeout_us=10_000_000, protocol="counter")),
                ("nccl", lambda: S.HeadGather(b, HQ, D, torch.bfloat16, dev, world=world, rank=rank))]
        for aname, make in arms:
            pg = make()
            bad_bits = bad_tol = 0
            worst = 0.0
            for step in range(steps):
                q = torch.randn(b, HQ, D, generator=g).to(torch.bfloat16).to(dev)      # same on every rank
                q_local = q[:, qh.start:qh.stop].contiguous()
                plain = P.paged_attention_decode_batched(q_local, kv_local, bts, sl)    # this rank's own bits
                want = torch.empty(world, b, HQ // world, D, dtype=torch.bfloat16, device=dev)
                dist.all_gather_into_tensor(want.view(-1), plain.view(-1))
                torch.cuda.synchronize()
                if step % 5 == 1:
                    time.sleep(0.003 * ((rank + step) % world))                        # let the ranks drift apart
                if aname == "nccl":
                    P.paged_attention_decode_batched(q_local, kv_local, bts, sl, out=pg.local_out())
                    pg.all_gather()
                    got = pg.buf
                    full = pg.full()
                else:
                    got = pg.decode(q_local, kv_local, bts, sl)
                    pg.check()
                    full = pg.full(got)
                if not torch.equal(got, want):
                    bad_bits += 1
                ref_full = P.paged_attention_decode_batched(q, kv, bts, sl)             # full-head launch on this GPU
                worst = max(worst, float((full.float() - ref_full.float()).abs().max()))
                if world == 1 and not torch.equal(full, ref_full):
                    bad_tol += 1
                if not strict_ok(full, ref_full):
                    bad_tol += 1
                if step == 0 and rank == 0 and cname == "ragged":
                    ref64 = O.decode_batched(q.cpu(), kv.cpu(), bts.cpu(), sl.cpu(), BS, compute_dtype=torch.float64)
                    if not strict_ok(full.cpu(), ref64):
                        bad_tol += 1
                for i, n in enumerate(lens):
                    if n == 0 and float(full[i].abs().sum()) != 0.0:
                        bad_tol += 1
            if aname != "nccl":
                pg.close()
            good = bad_bits == 0 and bad_tol == 0
            ok &= good
            note(f"{cname:9s} {aname:12s} {steps:3d} steps: transport bit-exact on {steps - bad_bits}/{steps}, "
                 f"vs full-head single-GPU decode max-abs {worst:.2e} (bound violations {bad_tol}) -> {'OK' if good else 'MISMATCH'}")

    # ---- CUDA-graph replay of (decode + receive), as bench.py --shard head runs it ---------------------
    lens = cases["ragged"]
    b = len(lens)
    kv, bts, sl, g = build(lens, 99, dev)
    kv_local = kv[:, :, kvh.start:kvh.stop].contiguous()
    pg = S.PeerGather(b, HQ, HKV, D, dev, world=world, rank=rank, timeout_us=10_000_000, protocol="ll")
    q_local = torch.zeros(b, HQ // world, D, dtype=torch.bfloat16, device=dev)
    ws = P.new_decode_workspace(b, HQ // world, HKV // world, D, dev)
    for i in range(2):                                                                   # warm both buffer sets eagerly
        pg.decode(q_local, kv_local, bts, sl, workspace=ws, parity=i)
    torch.cuda.synchronize()
    graphs = []
    for i in range(2):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            pg.decode(q_local, kv_local, bts, sl, workspace=ws, parity=i)
        graphs.append(gr)
    bad = 0
    for step in range(20):
        q = torch.randn(b, HQ, D, generator=g).to(torch.bfloat16).to(dev)
        q_local.copy_(q[:, qh.start:qh.stop])
        plain = P.paged_attention_decode_batched(q_local, kv_local, bts, sl)
        want = torch.empty(world, b, HQ // world, D, dtype=torch.bfloat16, device=dev)
        dist.all_gather_into_tensor(want.view(-1), plain.view(-1))
        graphs[step % 2].replay()
        torch.cuda.synchronize()
        bad += int(not torch.equal(pg.sets[step % 2], want))
    pg.check()
    ok &= bad == 0
    note(f"graph replay (decode + LL receive, 2 alternating graphs) 20 steps: bit-exact on {20 - bad}/20 -> {'OK' if bad == 0 else 'MISMATCH'}")
    pg.close()

    # ---- a dead peer must be reported, not hang -----------------------------------------------------------
    if world > 1:
        for protocol in ("ll", "counter"):
            pg = S.PeerGather(b, HQ, HKV, D, dev, world=world, rank=rank, timeout_us=300_000, protocol=protocol)
            pg.decode(q_local, kv_local, bts, sl)
            pg.check()                                                                   # a healthy step first
            dead = world - 1
            t0 = time.perf_counter()
            reported = None
            if rank != dead:
                pg.decode(q_local, kv_local, bts, sl)                                    # rank `dead` never sends this step
                try:
                    pg.check()
                except RuntimeError as e:
                    reported = str(e)
            dt = time.perf_counter() - t0
            good = rank == dead or (reported is not None and f"rank {dead}" in reported and dt < 5.0)
            ok &= good
            note(f"dead peer ({protocol}): " + ("I am the silent rank" if rank == dead else
                                                f"reported after {dt * 1e3:.0f} ms: {reported!r}") + f" -> {'OK' if good else 'FAIL'}")
            torch.cuda.synchronize()
            dist.barrier()
            pg.close()

    flag = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    gathered = [None] * world
    dist.all_gather_object(gathered, lines)
    if rank == 0:
        text = "\n".join(ln for rl in gathered for ln in rl) + f"\nmulti_gpu_worker world {world}: {'ALL OK' if int(flag) else 'FAILED'}\n"
        print(text, flush=True)
        if args.out:
            os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
            with open(args.out, "w") as f:
                f.write(text)
    dist.destroy_process_group()
    sys.exit(0 if int(flag) else 1)


if __name__ == "__main__":
    main()

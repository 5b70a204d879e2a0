"""Golden vectors of the chained block hash, produced by the UNMODIFIED reference
(/root/reference/06_global_prefix_cache/global_kv_pool.py:471-500) in the build container:

    python tests/golden/make_hash_golden.py        # writes tests/golden/block_hash_cases.json

hash() of nested tuples of ints is deterministic (no string hashing involved), so the values
travel to the GPU box (same CPython 3.12)."""
import json
import os
import random
import sys

sys.path.insert(0, "/root/reference/06_global_prefix_cache")
import global_kv_pool as ref  # noqa: E402

rng = random.Random(0)
cases = []
for name, n, bs, extra in [("kat_32_bs16", 32, 16, None), ("ragged_70_bs16", 70, 16, None), ("short_15_bs16", 15, 16, None),
                           ("bs4_23", 23, 4, None), ("extra_key", 48, 16, 7), ("long_2048_bs16", 2048, 16, None),
                           ("bs8_64", 64, 8, None)]:
    toks = [rng.randrange(0, 32000) for _ in range(n)]
    cases.append({"name": name, "token_ids": toks, "block_size": bs, "extra_key": extra,
                  "hashes": ref.compute_block_hashes(toks, bs, extra)})
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "block_hash_cases.json")
json.dump({"source": "06_global_prefix_cache/global_kv_pool.py:471-500", "python": sys.version.split()[0], "cases": cases},
          open(out, "w"))
print(out, len(cases))

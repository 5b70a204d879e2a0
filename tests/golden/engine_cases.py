"""Engine-loop golden cases (config C4) shared by make_engine_golden.py (which runs the UNMODIFIED
reference ``05_mini_vllm/mini_vllm.py`` in the build container) and tests/test_engine_golden.py
(which replays the recorded steps on the B200 through the CUDA kernels).

Everything both sides must agree on lives here and is derived from seeds: the engine
configuration, the prompts, the arrival schedule and the model weights.  The weights are NOT the
module's construction-order initialisation (that would tie the replica to torch's RNG consumption
order inside the reference's constructors): every parameter is drawn from its own generator seeded
by the crc32 of its state_dict name, so the real ``MiniVLLM`` (weights loaded with
``load_state_dict``) and the replica model hold bit-identical tensors; the golden file pins their
sha256.
"""
import random
import zlib

import torch

# name -> engine configuration.  ``c4_f32`` is the reference's own benchmark_throughput shape
# (mini_vllm.py:943-954: vocab 1000, hidden 256, 4 layers, 8 heads -> head_dim 32, inter 1024,
# block_size 16, max_num_seqs 16) with the token budget cut to 64 to force chunked prefill and
# a 28-block pool to force preemption by swap.  ``c4_bf16_d128`` widens the
# heads to head_dim 128 in bf16 so the same loop runs on the TMA decode / tcgen05 prefill kernels.
ENGINE_CASES = {
    "c4_f32": dict(vocab_size=1000, hidden_size=256, num_layers=4, num_heads=8, intermediate_size=1024,
                   num_kv_blocks=28, block_size=16, max_num_seqs=16, max_num_batched_tokens=64,
                   dtype="f32", num_host_blocks=64, max_new_tokens=16, seed=0),
    "c4_bf16_d128": dict(vocab_size=1000, hidden_size=512, num_layers=2, num_heads=4, intermediate_size=1024,
                         num_kv_blocks=28, block_size=16, max_num_seqs=16, max_num_batched_tokens=64,
                         dtype="bf16", num_host_blocks=64, max_new_tokens=16, seed=1),
}

ENGINE_KEYS = ("vocab_size", "hidden_size", "num_layers", "num_heads", "intermediate_size", "num_kv_blocks",
               "block_size", "max_num_seqs", "max_num_batched_tokens")


def engine_kwargs(cfg):
    return {k: cfg[k] for k in ENGINE_KEYS}


def make_prompts(cfg):
    """16 prompts of 10-50 tokens (mini_vllm.py:957-960), 8 of which start with the same 32-token
    (2-block) prefix; returns (prompts, arrival_step) - 4 requests at step 0, then one per step, so
    the prefix blocks are in the cache when the sharing prompts arrive."""
    rng = random.Random(cfg["seed"])
    v = cfg["vocab_size"]
    prefix = [rng.randint(1, v - 1) for _ in range(32)]
    prompts = []
    for i in range(16):
        body = [rng.randint(1, v - 1) for _ in range(rng.randint(10, 50))]
        prompts.append(prefix + body[:18] if i % 2 == 0 else body)
    arrival = [0 if i < 4 else i - 3 for i in range(16)]
    return prompts, arrival


def param_shapes(cfg):
    """state_dict names and shapes of the reference MiniTransformer's parameters
    (mini_vllm.py:344-352, :390-391, :494-496, :536-542)."""
    h, inter, v = cfg["hidden_size"], cfg["intermediate_size"], cfg["vocab_size"]
    shapes = {"embedding.weight": (v, h), "norm.weight": (h,), "lm_head.weight": (v, h)}
    for i in range(cfg["num_layers"]):
        p = f"layers.{i}."
        shapes[p + "attn.qkv.weight"] = (3 * h, h)
        shapes[p + "attn.out_proj.weight"] = (h, h)
        shapes[p + "ffn.gate.weight"] = (inter, h)
        shapes[p + "ffn.up.weight"] = (inter, h)
        shapes[p + "ffn.down.weight"] = (h, inter)
        shapes[p + "norm1.weight"] = (h,)
        shapes[p + "norm2.weight"] = (h,)
    return shapes


def det_weights(cfg):
    """name -> fp32 tensor, each from its own generator seeded by crc32(name) ^ case seed."""
    out = {}
    for name, shape in param_shapes(cfg).items():
        g = torch.Generator().manual_seed((zlib.crc32(name.encode()) ^ (cfg["seed"] * 7919)) & 0x7fffffff)
        if len(shape) == 1:
            w = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif name == "embedding.weight":
            w = torch.randn(shape, generator=g)
        else:
            # q / k projections a little hot so attention is not uniform and a wrong block shows in the logits
            scale = (2.0 if name.endswith("attn.qkv.weight") else 1.0) / shape[1] ** 0.5
            w = scale * torch.randn(shape, generator=g)
        out[name] = w
    return out


def torch_dtype(cfg):
    return torch.float32 if cfg["dtype"] == "f32" else torch.bfloat16

"""TEST INFRASTRUCTURE: tiny deterministic HuggingFace-layout checkpoints (config.json, tokenizer.json, sharded
.safetensors) for the converter tests — three architectures the reference's convert.py accepts (convert.py:15-19)."""
import json
import os

import numpy as np

SP = "▁"          # sentencepiece's visible space


def _tok_sentencepiece(vocab_size):
    vocab = {"<unk>": 0, "<s>": 1, "</s>": 2}
    for b in range(256):
        vocab["<0x%02X>" % b] = 3 + b
    i = len(vocab)
    for p in [SP, SP + SP, SP + "t", "in", "er", SP + "the", "é", SP + "caf", "日本"]:
        vocab[p] = i
        i += 1
    while i < vocab_size - 2:
        vocab[f"{SP}w{i}"] = i
        i += 1
    return {"model": {"type": "BPE", "byte_fallback": True, "vocab": vocab},
            "added_tokens": [{"id": vocab_size - 2, "content": "<|eot_id|>"}, {"id": vocab_size - 1, "content": "a b"}]}


def _tok_bytelevel(vocab_size):
    # GPT-2 style: pieces are written in the printable-unicode byte alphabet (U+0120 = space, U+010A = newline,
    # U+00C3 U+00A9 = the two UTF-8 bytes of e-acute)
    vocab = {"!": 0, "Ġthe": 1, "Ċ": 2, "ĠcafÃ©": 3, "hello": 4}
    i = len(vocab)
    while i < vocab_size - 1:
        vocab[f"Ġw{i}"] = i
        i += 1
    return {"model": {"type": "BPE", "vocab": vocab}, "added_tokens": [{"id": vocab_size - 1, "content": "<|im_end|>"}]}


CASES = {
    "llama-tied": dict(architectures=["LlamaForCausalLM"], hidden_size=64, intermediate_size=96, num_attention_heads=4,
                       num_key_value_heads=2, num_hidden_layers=2, vocab_size=300, max_position_embeddings=2048, bos_token_id=1,
                       eos_token_id=2, rope_theta=500000.0, rms_norm_eps=1e-05, hidden_act="silu", tie_word_embeddings=True,
                       partial_rotary_factor=0.5, _src="float32", _tok="sp"),
    "mistral": dict(architectures=["MistralForCausalLM"], hidden_size=64, intermediate_size=128, num_attention_heads=4,
                    num_key_value_heads=2, num_hidden_layers=2, vocab_size=280, max_position_embeddings=32768, bos_token_id=1,
                    eos_token_id=2, rope_theta=1000000.0, rms_norm_eps=1e-05, hidden_act="silu", tie_word_embeddings=False,
                    _src="bfloat16", _tok="sp"),
    "mixtral": dict(architectures=["MixtralForCausalLM"], hidden_size=64, intermediate_size=96, num_attention_heads=4,
                    num_key_value_heads=4, num_hidden_layers=2, vocab_size=270, max_position_embeddings=4096, bos_token_id=1,
                    eos_token_id=2, rms_norm_eps=1e-06, hidden_act="gelu", tie_word_embeddings=False, num_local_experts=4,
                    num_experts_per_tok=2, head_dim=16, _src="bfloat16", _tok="bytelevel"),
}


def write_hf_dir(path, case):
    """Writes the checkpoint with the real safetensors library (two shards), returns the config."""
    import torch
    from safetensors.torch import save_file
    cfg = dict(CASES[case])
    src, tok = cfg.pop("_src"), cfg.pop("_tok")
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(cfg, f)
    with open(os.path.join(path, "tokenizer.json"), "w") as f:
        json.dump(_tok_sentencepiece(cfg["vocab_size"]) if tok == "sp" else _tok_bytelevel(cfg["vocab_size"]), f)
    rng = np.random.default_rng(sorted(CASES).index(case) + 100)
    dim, hid, H, KV = cfg["hidden_size"], cfg["intermediate_size"], cfg["num_attention_heads"], cfg["num_key_value_heads"]
    hd = cfg.get("head_dim", dim // H)
    dt = getattr(torch, src)

    def t(*shape):
        return torch.from_numpy(rng.standard_normal(shape).astype(np.float32)).to(dt)

    shards = [{}, {}]
    shards[0]["model.embed_tokens.weight"] = t(cfg["vocab_size"], dim)
    for l in range(cfg["num_hidden_layers"]):
        p, d = f"model.layers.{l}.", shards[l % 2]
        d[p + "input_layernorm.weight"] = t(dim)
        d[p + "post_attention_layernorm.weight"] = t(dim)
        d[p + "self_attn.q_proj.weight"] = t(H * hd, dim)
        d[p + "self_attn.k_proj.weight"] = t(KV * hd, dim)
        d[p + "self_attn.v_proj.weight"] = t(KV * hd, dim)
        d[p + "self_attn.o_proj.weight"] = t(dim, H * hd)
        if "num_local_experts" in cfg:
            d[p + "block_sparse_moe.gate.weight"] = t(cfg["num_local_experts"], dim)
            for e in range(cfg["num_local_experts"]):
                d[p + f"block_sparse_moe.experts.{e}.w1.weight"] = t(hid, dim)
                d[p + f"block_sparse_moe.experts.{e}.w2.weight"] = t(dim, hid)
                d[p + f"block_sparse_moe.experts.{e}.w3.weight"] = t(hid, dim)
        else:
            d[p + "mlp.gate_proj.weight"] = t(hid, dim)
            d[p + "mlp.down_proj.weight"] = t(dim, hid)
            d[p + "mlp.up_proj.weight"] = t(hid, dim)
    shards[1]["model.norm.weight"] = t(dim)
    if not cfg.get("tie_word_embeddings"):
        shards[1]["lm_head.weight"] = t(cfg["vocab_size"], dim)
    for i, d in enumerate(shards):
        save_file(d, os.path.join(path, f"model-0000{i + 1}-of-00002.safetensors"))
    return cfg

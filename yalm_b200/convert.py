#!/usr/bin/env python
"""HuggingFace checkpoint directory -> `.yalm`, compatible with the reference's convert.py:

    python -m yalm_b200.convert out.yalm hf_dir/ [--dtype fp32|fp16|fp8]

Same container, tensor names, shapes, metadata strings and tokenizer blob as the reference writes (convert.py:22-81
metadata, :98-125 tokens, :127-207 weights), so that a file produced here loads in the reference and vice versa:
  * wq / wk rows are un-permuted from HF's two-halves rotary layout to interleaved pairs (convert.py:145-158),
  * Mixtral experts are stacked into (n_experts, ...) tensors (convert.py:188-193),
  * norm weights stay fp32, everything else is cast to --dtype.
Differences, all deliberate:
  * `--dtype fp8` writes F8_E4M3 codes + one fp32 scale per output row (`<name>.scale`) — the format this library
    computes with.  The reference casts to unscaled F8_E5M2, for which it has no arithmetic.
  * streaming and torch-free: shards are mmap'ed and converted tensor by tensor with numpy; peak memory is one tensor,
    not the whole model (the reference materialises every shard, then every converted tensor).
"""
from __future__ import annotations

import argparse
import json
import os
import struct

import numpy as np

from . import yalm_file as yf

SUPPORTED_ARCHITECTURES = ["LlamaForCausalLM", "MistralForCausalLM", "MixtralForCausalLM"]
SUPPORTED_DTYPES = ["fp32", "fp16", "fp8"]


def metadata_of(config: dict, dtype: str) -> dict:
    """All values strings, formatted as Python's str() formats them (the reference's Metadata.to_dict)."""
    arch = config["architectures"][0]
    if arch not in SUPPORTED_ARCHITECTURES:
        raise ValueError(f"Architecture {arch} is not supported, must be one of {SUPPORTED_ARCHITECTURES}")
    if dtype not in SUPPORTED_DTYPES:
        raise ValueError(f"Data type {dtype} is not supported, must be one of {SUPPORTED_DTYPES}")
    if config.get("attention_bias", False) or config.get("mlp_bias", False):
        raise ValueError("biases are not supported")
    if config["hidden_act"] not in ("gelu", "silu"):
        raise ValueError(f"hidden_act {config['hidden_act']} is not supported")
    head_dim = config.get("head_dim", config["hidden_size"] // config["num_attention_heads"])
    md = {
        "arch": arch, "dtype": dtype,
        "dim": config["hidden_size"], "hidden_dim": config["intermediate_size"], "head_dim": head_dim,
        "n_layers": config["num_hidden_layers"], "n_heads": config["num_attention_heads"],
        "n_kv_heads": config.get("num_key_value_heads", config["num_attention_heads"]),
        "vocab_size": config["vocab_size"], "max_seq_len": config["max_position_embeddings"],
        "bos_token_id": config["bos_token_id"], "eos_token_id": config["eos_token_id"],
        "rope_theta": config.get("rope_theta", 10000.0),
        "rotary_dim": int(head_dim * config.get("partial_rotary_factor", 1)),
        "norm_eps": config["rms_norm_eps"], "norm_type": "rmsnorm", "act_type": config["hidden_act"],
    }
    if arch == "MixtralForCausalLM":
        md["n_experts"] = config["num_local_experts"]
        md["n_experts_active"] = config["num_experts_per_tok"]
    return {k: str(v) for k, v in md.items()}


def _gpt2_unicode_to_byte() -> dict:
    """Inverse of GPT-2's printable-unicode byte alphabet (the ByteLevel pre-tokenizer of HF tokenizers)."""
    keep = list(range(ord("!"), ord("~") + 1)) + list(range(ord("¡"), ord("¬") + 1)) + list(range(ord("®"), ord("ÿ") + 1))
    table, extra = {}, 0
    for b in range(256):
        if b in keep:
            table[chr(b)] = b
        else:
            table[chr(256 + extra)] = b
            extra += 1
    return table


def token_blob(tokenizer: dict, vocab_size: int) -> np.ndarray:
    """`tokenizer.tokens`: the pieces as UTF-8, NUL-terminated, in id order (convert.py:98-125, :245)."""
    pieces = [""] * vocab_size
    vocab = tokenizer["model"]["vocab"]
    if len(vocab) > vocab_size:
        raise ValueError("tokenizer vocabulary larger than config vocab_size")
    for t, i in vocab.items():
        pieces[i] = t
    for added in tokenizer.get("added_tokens", []):
        pieces[added["id"]] = added["content"]
    byte_level = not tokenizer["model"].get("byte_fallback", False)
    inv = _gpt2_unicode_to_byte()
    out = bytearray()
    for t in pieces:
        b = bytes(inv.get(ch, 0) for ch in t) if byte_level else t.replace("▁", " ").encode("utf-8")
        out += b.replace(b"\0", b"\7") + b"\0"
    return np.frombuffer(bytes(out), np.uint8)


# ---- safetensors shards (the same container as .yalm) -------------------------------------------------------------
class _Shards:
    def __init__(self, paths):
        self.index = {}
        for p in paths:
            buf = np.memmap(p, dtype=np.uint8, mode="r")
            (hlen,) = struct.unpack("<Q", bytes(buf[:8]))
            header = json.loads(bytes(buf[8:8 + hlen]))
            header.pop("__metadata__", None)
            for name, info in header.items():
                if name in self.index:
                    raise ValueError(f"tensor {name} appears in two shards")
                self.index[name] = (buf, 8 + hlen, info)

    def f32(self, name: str) -> np.ndarray:
        buf, base, info = self.index[name]
        s, e = info["data_offsets"]
        raw, shape, dt = buf[base + s: base + e], info["shape"], info["dtype"]
        if dt == "F32":
            return np.array(raw.view(np.float32)).reshape(shape)
        if dt == "F16":
            return raw.view(np.float16).astype(np.float32).reshape(shape)
        if dt == "BF16":   # bf16 = the upper half of an fp32
            return (raw.view(np.uint16).astype(np.uint32) << 16).view(np.float32).reshape(shape)
        raise ValueError(f"{name}: unsupported source dtype {dt}")


def permute_reverse(w: np.ndarray, heads: int, rotary_dim: int) -> np.ndarray:
    """HF stores the rotary rows of every head as two halves (all even-pair members, then all odd ones); the forward pass
    rotates interleaved pairs (src/infer.cpp:200-213).  Rows [h, r] -> [h, interleave(r)] for the rotary part, rest kept."""
    head_dim = w.shape[0] // heads
    if rotary_dim > head_dim:
        raise ValueError("rotary_dim > head_dim")
    v = w.reshape(heads, head_dim, -1)
    rot = v[:, :rotary_dim].reshape(heads, 2, rotary_dim // 2, -1).transpose(0, 2, 1, 3).reshape(heads, rotary_dim, -1)
    return np.concatenate([rot, v[:, rotary_dim:]], axis=1).reshape(w.shape)


def plan(config: dict, md: dict, tied: bool):
    """[(yalm name, how)] in the reference's order; how = ("norm", hf) | ("w", hf) | ("perm", hf, heads) | ("stack", [hf...])."""
    L, moe = config["num_hidden_layers"], config["architectures"][0] == "MixtralForCausalLM"
    out = [("model.embed.weight", ("w", "model.embed_tokens.weight"))]
    for l in range(L):
        hf, y = f"model.layers.{l}.", f"model.layers.{l}."
        out += [
            (y + "attn.norm.weight", ("norm", hf + "input_layernorm.weight")),
            (y + "attn.wq.weight", ("perm", hf + "self_attn.q_proj.weight", int(md["n_heads"]))),
            (y + "attn.wk.weight", ("perm", hf + "self_attn.k_proj.weight", int(md["n_kv_heads"]))),
            (y + "attn.wv.weight", ("w", hf + "self_attn.v_proj.weight")),
            (y + "attn.wo.weight", ("w", hf + "self_attn.o_proj.weight")),
            (y + "mlp.norm.weight", ("norm", hf + "post_attention_layernorm.weight")),
        ]
        if moe:
            E = config["num_local_experts"]
            out.append((y + "moegate.weight", ("w", hf + "block_sparse_moe.gate.weight")))
            for w in ("w1", "w2", "w3"):
                out.append((y + f"mlp.{w}.weight", ("stack", [hf + f"block_sparse_moe.experts.{e}.{w}.weight" for e in range(E)])))
        else:
            out += [(y + "mlp.w1.weight", ("w", hf + "mlp.gate_proj.weight")), (y + "mlp.w2.weight", ("w", hf + "mlp.down_proj.weight")),
                    (y + "mlp.w3.weight", ("w", hf + "mlp.up_proj.weight"))]
    out.append(("model.norm.weight", ("norm", "model.norm.weight")))
    if not tied:
        out.append(("model.output.weight", ("w", "lm_head.weight")))
    return out


def _materialise(shards: _Shards, how, rotary_dim: int) -> np.ndarray:
    if how[0] in ("norm", "w"):
        return shards.f32(how[1])
    if how[0] == "perm":
        return permute_reverse(shards.f32(how[1]), how[2], rotary_dim)
    return np.stack([shards.f32(n) for n in how[1]])


def convert(input_dir: str, output: str, dtype: str = "fp16") -> dict:
    with open(os.path.join(input_dir, "config.json")) as f:
        config = json.load(f)
    with open(os.path.join(input_dir, "tokenizer.json")) as f:
        tokenizer = json.load(f)
    files = sorted(os.path.join(input_dir, n) for n in os.listdir(input_dir) if n.endswith(".safetensors"))
    if not files:
        raise FileNotFoundError(f"no .safetensors files found in {input_dir}")
    md = metadata_of(config, dtype)
    shards = _Shards(files)
    tied = config.get("tie_word_embeddings", None) != False  # noqa: E712  (the reference tests `== False`: None means tied)
    steps = plan(config, md, tied)
    rotary_dim, st = int(md["rotary_dim"]), {"fp32": "F32", "fp16": "F16", "fp8": "F8_E4M3"}[dtype]
    esz = {"F32": 4, "F16": 2, "F8_E4M3": 1}[st]

    # pass 1: the header (shapes from the shard headers: nothing is read yet)
    def shape_of(how):
        if how[0] == "stack":
            return [len(how[1])] + list(shards.index[how[1][0]][2]["shape"])
        return list(shards.index[how[1]][2]["shape"])

    tokens = token_blob(tokenizer, int(md["vocab_size"]))
    entries = []
    for name, how in steps:
        shape = shape_of(how)
        n = int(np.prod(shape))
        if how[0] == "norm":
            entries.append((name, "F32", shape, 4 * n, how))
        else:
            entries.append((name, st, shape, esz * n, how))
            if dtype == "fp8":
                entries.append((name + ".scale", "F32", shape[:-1], 4 * int(np.prod(shape[:-1])), None))
    entries.append(("tokenizer.tokens", "U8", [tokens.size], tokens.size, None))
    header, off = {"__metadata__": md}, 0
    for name, sdt, shape, nb, _ in entries:
        off = (off + 63) // 64 * 64
        header[name] = {"dtype": sdt, "shape": shape, "data_offsets": [off, off + nb]}
        off += nb
    hj = json.dumps(header, separators=(",", ":")).encode()
    hj += b" " * ((-(8 + len(hj))) % 64)
    # pass 2: one tensor at a time
    with open(output, "wb") as f:
        f.write(struct.pack("<Q", len(hj)) + hj)
        base = 8 + len(hj)
        f.truncate(base + off)
        for name, sdt, shape, nb, how in entries:
            if how is None:
                continue
            w = _materialise(shards, how, rotary_dim)
            pos = base + header[name]["data_offsets"][0]
            if how[0] == "norm" or dtype == "fp32":
                f.seek(pos)
                f.write(np.ascontiguousarray(w, np.float32).tobytes())
            elif dtype == "fp16":
                f.seek(pos)
                f.write(w.astype(np.float16).tobytes())
            else:
                codes, scale = yf.quantize_rows_e4m3(w)
                f.seek(pos)
                f.write(codes.tobytes())
                f.seek(base + header[name + ".scale"]["data_offsets"][0])
                f.write(scale.astype(np.float32).tobytes())
        f.seek(base + header["tokenizer.tokens"]["data_offsets"][0])
        f.write(tokens.tobytes())
    return md


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("output")
    ap.add_argument("input")
    ap.add_argument("--dtype", default="fp16", choices=SUPPORTED_DTYPES)
    args = ap.parse_args()
    for need in ("config.json", "tokenizer.json"):
        if not os.path.exists(os.path.join(args.input, need)):
            ap.error(f"{need} not found in {args.input}")
    md = convert(args.input, args.output, args.dtype)
    print(f"wrote {args.output}: {md['arch']} {md['n_layers']} layers, dtype {args.dtype}")


if __name__ == "__main__":
    main()

"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

ctypes front-ends for the two checkers:
  * `OracleModel` — our plain-C restatement (oracle/yalm_oracle.c -> oracle/liboracle.so)
  * `RefModel`    — the unmodified reference built into oracle/_ref/libyalm_ref.so (oracle/Makefile)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys
import tempfile

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))
from yalm_b200 import yalm_file as yf  # noqa: E402

DT = {"F32": 0, "F16": 1, "F8_E4M3": 4}
_fp = C.POINTER(C.c_float)


def build(ref: bool = True) -> None:
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])
    if ref and os.path.isdir("/root/reference/src"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "ref"])


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class _Cfg(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("dim", "hidden_dim", "head_dim", "n_layers", "n_heads", "n_kv_heads",
                                        "vocab_size", "max_seq_len")] + \
               [("rope_theta", C.c_float), ("rotary_dim", C.c_int), ("norm_eps", C.c_float), ("act", C.c_int),
                ("qkv_clip", C.c_float), ("n_experts", C.c_int), ("n_experts_active", C.c_int),
                ("weight_dtype", C.c_int)]


class _Layer(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("rms_att", "rms_ffn", "wq", "wk", "wv", "wo", "w1", "w2", "w3", "moegate",
                                           "sq", "sk", "sv", "so", "s1", "s2", "s3", "sgate",
                                           "key_cache", "value_cache")]


class _Model(C.Structure):
    _fields_ = [("c", _Cfg), ("embed", C.c_void_p), ("embed_scale", C.c_void_p), ("rms_final", C.c_void_p),
                ("wcls", C.c_void_p), ("wcls_scale", C.c_void_p), ("layers", C.POINTER(_Layer))] + \
               [(n, C.c_void_p) for n in ("x", "xb", "xb2", "hb", "hb2", "q", "k", "v", "att", "logits",
                                           "moe_weights", "active_w", "active_e")]


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build(ref=False)
        _lib = C.CDLL(path)
        _lib.yo_forward.argtypes = [C.POINTER(_Model), C.c_int, C.c_int, C.c_int]
        _lib.yo_argmax.argtypes = [C.c_void_p, C.c_int]
        _lib.yo_matmul.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        _lib.yo_mha.argtypes = [C.c_void_p] * 5 + [C.c_int] * 5
        _lib.yo_attn.argtypes = [C.c_void_p] * 5 + [C.c_int] * 3
        _lib.yo_ffn.argtypes = [C.c_void_p] * 8 + [C.c_int] * 4
        _lib.yo_moe_gate.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        _lib.yo_rope.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int]
        _lib.yo_rmsnorm.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float]
    return _lib


# ---- free functions (numpy in / numpy out) ---------------------------------------------------
def matmul(x, w, scale=None, dtype="F32"):
    d, n = w.shape
    out = np.empty(d, np.float32)
    lib().yo_matmul(_ptr(out), _ptr(np.ascontiguousarray(x, np.float32)), _ptr(w), _ptr(scale), n, d, DT[dtype])
    return out


def mha(kb, vb, q, head_dim, kv_len, max_seq_len, n_heads, n_kv_heads):
    xout = np.zeros(n_heads * head_dim, np.float32)
    att = np.zeros(n_heads * max_seq_len, np.float32)
    lib().yo_mha(_ptr(xout), _ptr(att), _ptr(kb), _ptr(vb), _ptr(q), head_dim, kv_len, max_seq_len, n_heads,
                 n_kv_heads)
    return xout, att.reshape(n_heads, max_seq_len)


def ffn(x, w1, w2, w3, act="gelu", dtype="F32", s1=None, s2=None, s3=None):
    hidden, dim = w1.shape
    out = np.empty(dim, np.float32)
    lib().yo_ffn(_ptr(out), _ptr(x), _ptr(w1), _ptr(w2), _ptr(w3), _ptr(s1), _ptr(s2), _ptr(s3), hidden, dim,
                 0 if act == "gelu" else 1, DT[dtype])
    return out


def moe_gate(logits, n_active):
    logits = np.ascontiguousarray(logits, np.float32)
    w = np.zeros(n_active, np.float32)
    e = np.zeros(n_active, np.int32)
    lib().yo_moe_gate(_ptr(w), _ptr(e), _ptr(logits), len(logits), n_active)
    return w, e


def sample(logits, temperature: float, r: float) -> int:
    """Sampler::sample (src/sampler.cpp:40-65) with the std::rand() draw replaced by `r`: fp32 running sum in index
    order, first index whose sum reaches r, vocab - 1 if none.  temperature 0 = first-max arg-max (:27-38)."""
    l = np.ascontiguousarray(logits, np.float32)
    if temperature == 0:
        return int(np.argmax(l))
    m = l.max()
    e = np.exp(((l - m) / np.float32(temperature)).astype(np.float32)).astype(np.float32)
    s = np.float32(0)
    for v in e:                       # sequential fp32 sum, as the reference accumulates it
        s = np.float32(s + v)
    c = np.cumsum((e / s).astype(np.float32), dtype=np.float32)   # numpy's cumsum is sequential in index order
    hit = np.nonzero(c >= np.float32(r))[0]
    return int(hit[0]) if len(hit) else len(l) - 1


def sample_prob(logits, index: int) -> float:
    """Sampler::sample_prob (src/sampler.cpp:11-25)."""
    l = np.ascontiguousarray(logits, np.float32)
    e = np.exp((l - l.max()).astype(np.float32)).astype(np.float32)
    s = np.float32(0)
    for v in e:
        s = np.float32(s + v)
    return float(e[index] / s)


def tokenize(vocab, text: bytes, bos_id: int, add_bos: bool = True):
    """Tokenizer::encode (src/tokenizer.cpp:57-94): greedy longest match over the pieces, byte fallback through
    "<0x00>".."<0xFF>"; later duplicates of a piece win (trie insertion order, :32-41).  vocab: list of bytes."""
    ids = {}
    for i, p in enumerate(vocab):
        ids[p] = i
    maxlen = max((len(p) for p in vocab), default=0)
    prefixes = {p[:k] for p in vocab for k in range(1, len(p) + 1)}
    fb = ids.get(b"<0x00>", -1)
    out = [bos_id] if add_bos else []
    i = 0
    while i < len(text):
        best, l = None, 0
        while i + l < len(text) and l < maxlen and text[i:i + l + 1] in prefixes:
            l += 1
            if text[i:i + l] in ids:
                best = l
        if best is None:
            if fb >= 0:
                out.append(text[i] + fb)
            i += 1
        else:
            out.append(ids[text[i:i + best]])
            i += best
    return out


def detokenize(vocab, ids, bos_id: int) -> bytes:
    """Tokenizer::decode_one chained (src/tokenizer.cpp:44-55)."""
    fb = {p: i for i, p in enumerate(vocab)}.get(b"<0x00>", -1)
    out, prev = b"", -1
    for t in ids:
        p = vocab[t]
        if prev == bos_id and p[:1] == b" ":
            out += p[1:]
        elif fb >= 0 and fb <= t < fb + 256:
            out += bytes([t - fb])
        else:
            out += p
        prev = t
    return out


class OracleModel:
    """Our C restatement driven over a parsed .yalm (path or in-memory image)."""

    def __init__(self, yalm, context: int = 0):
        md, t, kinds = yf.load(yalm)
        self.md, self.t = md, t
        g = lambda k, d=None: md.get(k, d)
        c = _Cfg()
        for k in ("dim", "hidden_dim", "head_dim", "n_layers", "n_heads", "n_kv_heads", "vocab_size", "rotary_dim"):
            setattr(c, k, int(md[k]))
        c.max_seq_len = context if context else min(int(md["max_seq_len"]), 4096)   # src/model.cpp:31-36
        c.rope_theta = float(md["rope_theta"])
        c.norm_eps = float(g("norm_eps", "1e-5"))
        c.act = 1 if g("act_type", "gelu") == "silu" else 0
        c.qkv_clip = float(g("qkv_clip", str(np.finfo(np.float32).max)))
        c.n_experts = int(g("n_experts", "0"))
        c.n_experts_active = int(g("n_experts_active", "0"))
        wkind = {"fp32": "F32", "fp16": "F16", "fp8": "F8_E4M3"}[md["dtype"]]
        c.weight_dtype = DT[wkind]
        self.cfg = c
        kv_dim = c.n_kv_heads * c.head_dim
        self._keep = []
        self.kc = [np.zeros((c.max_seq_len, kv_dim), np.uint16) for _ in range(c.n_layers)]
        self.vc = [np.zeros((c.max_seq_len, kv_dim), np.uint16) for _ in range(c.n_layers)]
        layers = (_Layer * c.n_layers)()

        def P(name):
            a = t.get(name)
            return None if a is None else a.ctypes.data

        for l in range(c.n_layers):
            p = f"model.layers.{l}."
            L = layers[l]
            L.rms_att, L.rms_ffn = P(p + "attn.norm.weight"), P(p + "mlp.norm.weight")
            for f, n in (("wq", "attn.wq"), ("wk", "attn.wk"), ("wv", "attn.wv"), ("wo", "attn.wo"),
                         ("w1", "mlp.w1"), ("w2", "mlp.w2"), ("w3", "mlp.w3"), ("moegate", "moegate")):
                setattr(L, f, P(p + n + ".weight"))
                setattr(L, "s" + (f[1:] if f != "moegate" else "gate"), P(p + n + ".weight.scale"))
            L.key_cache, L.value_cache = self.kc[l].ctypes.data, self.vc[l].ctypes.data
        m = _Model()
        m.c, m.layers = c, layers
        m.embed, m.embed_scale = P("model.embed.weight"), P("model.embed.weight.scale")
        m.rms_final = P("model.norm.weight")
        tied = "model.output.weight" not in t
        m.wcls = m.embed if tied else P("model.output.weight")
        m.wcls_scale = m.embed_scale if tied else P("model.output.weight.scale")
        sizes = dict(x=c.dim, xb=c.dim, xb2=max(c.dim, c.n_heads * c.head_dim), hb=max(c.hidden_dim, c.dim), hb2=c.hidden_dim,  # hb doubles as the wo output (src/infer.cpp:335): dim floats
                     q=c.n_heads * c.head_dim, k=kv_dim, v=kv_dim, att=c.n_heads * c.max_seq_len,
                     logits=c.vocab_size, moe_weights=max(1, c.n_experts), active_w=max(1, c.n_experts_active))
        self.buf = {k: np.zeros(n, np.float32) for k, n in sizes.items()}
        self.buf["active_e"] = np.zeros(max(1, c.n_experts_active), np.int32)
        for k, a in self.buf.items():
            setattr(m, k, a.ctypes.data)
        self.m, self._layers = m, layers

    def forward(self, token: int, pos: int, mode: int = 1):
        lib().yo_forward(C.byref(self.m), token, pos, mode)
        return self.buf["logits"]

    def argmax(self) -> int:
        return lib().yo_argmax(_ptr(self.buf["logits"]), self.cfg.vocab_size)

    def generate(self, prompt, n_gen, keep_logits=False):
        """Same loop as run_completion (src/main.cpp:91-112), greedy."""
        for pos, tok in enumerate(prompt):
            self.forward(tok, pos, 1 if pos + 1 == len(prompt) else 0)
        toks, logits = [], []
        for i in range(n_gen):
            tok = self.argmax()
            toks.append(tok)
            if keep_logits:
                logits.append(self.buf["logits"].copy())
            if i + 1 < n_gen:
                self.forward(tok, len(prompt) + i, 1)
        return (toks, np.stack(logits)) if keep_logits else toks


_ref = None


def ref_available() -> bool:
    return os.path.exists(os.path.join(_HERE, "_ref", "libyalm_ref.so"))


def ref_lib():
    global _ref
    if _ref is None:
        L = C.CDLL(os.path.join(_HERE, "_ref", "libyalm_ref.so"))
        L.ref_open.restype = C.c_void_p
        L.ref_open.argtypes = [C.c_char_p, C.c_int]
        L.ref_close.argtypes = [C.c_void_p]
        L.ref_cuda.argtypes = [C.c_void_p]
        L.ref_config.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        L.ref_forward.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.ref_logits.restype = _fp
        L.ref_logits.argtypes = [C.c_void_p]
        L.ref_argmax.argtypes = [C.c_void_p]
        L.ref_state_buffer.restype = C.c_void_p
        L.ref_state_buffer.argtypes = [C.c_void_p, C.c_int]
        L.ref_kv_cache.restype = C.c_void_p
        L.ref_kv_cache.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.ref_generate.restype = C.c_double
        L.ref_generate.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.ref_attn.argtypes = [C.c_void_p] * 5 + [C.c_int] * 3
        L.ref_mha_cpu.argtypes = [C.c_void_p] * 5 + [C.c_int] * 5
        L.ref_mha_cuda.argtypes = [C.c_void_p] * 5 + [C.c_int] * 5
        for n in ("ref_matmul_cpu_f32", "ref_matmul_cpu_f16", "ref_matmul_cuda_f32", "ref_matmul_cuda_f16"):
            getattr(L, n).argtypes = [C.c_void_p] * 3 + [C.c_int] * 2
        L.ref_ffn_cpu.argtypes = [C.c_void_p] * 5 + [C.c_int] * 3
        L.ref_set_threads.argtypes = [C.c_int]
        L.ref_tokenize.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_void_p, C.c_int]
        L.ref_detokenize.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_char_p, C.c_int]
        L.ref_sample_prob.restype = C.c_float
        L.ref_sample_prob.argtypes = [C.c_void_p, C.c_int]
        _ref = L
    return _ref


class RefModel:
    """The unmodified reference (Model::forward on Device::CPU, or ::CUDA after .cuda())."""

    def __init__(self, yalm, context: int = 0):
        self._tmp = None
        if not isinstance(yalm, str):
            self._tmp = tempfile.NamedTemporaryFile(suffix=".yalm", delete=False)
            self._tmp.write(yalm)
            self._tmp.close()
            yalm = self._tmp.name
        self.h = ref_lib().ref_open(yalm.encode(), context)
        if not self.h:
            raise RuntimeError("reference failed to open " + yalm)
        cfg = (C.c_int * 16)()
        ref_lib().ref_config(self.h, cfg)
        (self.dim, self.hidden_dim, self.head_dim, self.n_layers, self.n_heads, self.n_kv_heads, self.vocab_size,
         self.max_seq_len, self.rotary_dim, self.n_experts, self.n_experts_active, self.weight_dtype,
         self.act) = list(cfg)[:13]

    def cuda(self):
        ref_lib().ref_cuda(self.h)

    def forward(self, token, pos, mode=1):
        ref_lib().ref_forward(self.h, token, pos, mode)
        return self.logits()

    def logits(self):
        return np.ctypeslib.as_array(ref_lib().ref_logits(self.h), (self.vocab_size,))

    def argmax(self):
        return ref_lib().ref_argmax(self.h)

    def kv(self, layer):
        n = self.max_seq_len * self.n_kv_heads * self.head_dim
        out = []
        for which in (0, 1):
            p = C.cast(ref_lib().ref_kv_cache(self.h, layer, which), C.POINTER(C.c_uint16))
            out.append(np.ctypeslib.as_array(p, (n,)).reshape(self.max_seq_len, -1))
        return out

    def generate(self, prompt, n_gen, keep_logits=False):
        prompt = np.ascontiguousarray(prompt, np.int32)
        toks = np.zeros(n_gen, np.int32)
        lg = np.zeros((n_gen, self.vocab_size), np.float32) if keep_logits else None
        secs = ref_lib().ref_generate(self.h, _ptr(prompt), len(prompt), n_gen, _ptr(toks), _ptr(lg))
        self.decode_seconds = secs
        return (toks.tolist(), lg) if keep_logits else toks.tolist()

    def tokenize(self, text: bytes, add_bos: bool = True):
        out = np.zeros(2 * len(text) + 2, np.int32)
        n = ref_lib().ref_tokenize(self.h, text, int(add_bos), _ptr(out), out.size)
        return out[:n].tolist()

    def detokenize(self, ids) -> bytes:
        ids = np.ascontiguousarray(ids, np.int32)
        buf = C.create_string_buffer(64 * len(ids) + 16)
        n = ref_lib().ref_detokenize(self.h, _ptr(ids), len(ids), buf, len(buf))
        return buf.raw[:n]

    def sample_prob(self, index: int) -> float:
        return float(ref_lib().ref_sample_prob(self.h, index))

    def close(self):
        if self.h:
            ref_lib().ref_close(self.h)
            self.h = None
        if self._tmp:
            os.unlink(self._tmp.name)
            self._tmp = None

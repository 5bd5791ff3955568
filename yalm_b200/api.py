"""ctypes binding of include/yalm_b200.h — the Python-side mirror of yalm's Model / InferenceState /
test entry points (reference src/model.h:84-327, 353-384).

There is no Python or CPU implementation behind these calls: if libyalm_b200.so is missing the import
of the library fails loudly, and every compute entry point returns YALM_ENOGPU without a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("YALM_B200_LIB") or os.path.join(_HERE, "libyalm_b200.so")  # override: A/B builds

F32, F16, F8E4M3 = 0, 1, 4
ACT_GELU, ACT_SILU = 0, 1
MODE_HYDRATE, MODE_LOGITS = 0, 1
DTYPE_OF = {"fp32": F32, "fp16": F16, "fp8": F8E4M3}

# every symbol include/yalm_b200.h declares (tests check the library exports all of them)
SYMBOLS = [
    "upload_cuda", "download_cuda_raw", "register_cuda_host", "free_cuda", "unregister_cuda_host",
    "set_cuda_device", "init_cuda_stream",
    "yalm_embed", "yalm_qkv_rope_append", "yalm_attn_workspace_bytes", "yalm_attn_decode", "yalm_gemv_residual",
    "yalm_ffn_glu", "yalm_lm_head", "yalm_argmax_advance", "yalm_gemv", "yalm_gemv_residual_ep", "yalm_ffn_glu_ep",
    "yalm_attn_decode_static", "yalm_attn_decode_static_max_bytes",
    "yalm_matmul_host", "yalm_mha_host", "yalm_ffn_host", "yalm_bench_op",
    "yalm_model_open_file", "yalm_model_open_memory", "yalm_model_cuda", "yalm_model_close", "yalm_model_config",
    "yalm_model_device_bytes", "yalm_model_load_info", "yalm_state_new", "yalm_state_free", "yalm_forward", "yalm_logits",
    "yalm_generate_greedy", "yalm_state_read", "yalm_forward_launches", "yalm_model_uses_megakernel", "yalm_model_mega_info", "yalm_mega_profile",
    "yalm_profile_ffn", "yalm_sample_advance", "yalm_sample_device1", "yalm_token_probs", "yalm_sample_host",
    "yalm_token_probs_host", "yalm_gemm_f16", "yalm_gemm_glu_f16", "yalm_attn_prefill", "yalm_gemm_host",
    "yalm_model_batch_capacity", "yalm_forward_batch", "yalm_generate", "yalm_state_debug_set_launch_seq", "yalm_tokenize", "yalm_detokenize",
    "yalm_tp_unique_id", "yalm_tp_init", "yalm_tp_finalize", "yalm_tp_shard_plan", "yalm_mega_tile_plan", "yalm_mega_attn_split", "yalm_mega_mma_plan", "yalm_mega_hb_share", "yalm_error_string", "yalm_version",
]


class YalmError(RuntimeError):
    def __init__(self, code, what):
        self.code = code
        super().__init__(f"{what}: {error_string(code)} ({code})")


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} not built (run `make lib` or __graft_entry__.build()); "
                          "yalm_b200 has no fallback implementation")
    L = C.CDLL(LIB_PATH)
    vp, i, f, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    L.yalm_error_string.restype = C.c_char_p
    L.yalm_error_string.argtypes = [i]
    L.yalm_version.restype = C.c_char_p
    L.yalm_matmul_host.argtypes = [vp, vp, vp, vp, i, i, i]
    L.yalm_mha_host.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, i]
    L.yalm_ffn_host.argtypes = [vp] * 8 + [i] * 4
    L.yalm_bench_op.argtypes = [C.c_char_p, i, i, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.yalm_model_open_file.argtypes = [C.c_char_p, i, C.POINTER(vp)]
    L.yalm_model_open_memory.argtypes = [vp, sz, i, C.POINTER(vp)]
    L.yalm_model_cuda.argtypes = [vp, i, i, i]
    L.yalm_model_close.argtypes = [vp]
    L.yalm_model_close.restype = None
    L.yalm_model_config.argtypes = [vp, C.POINTER(i)]
    L.yalm_model_config.restype = None
    L.yalm_model_load_info.argtypes = [vp, C.POINTER(i)]
    L.yalm_model_load_info.restype = None
    L.yalm_model_device_bytes.argtypes = [vp]
    L.yalm_model_device_bytes.restype = sz
    L.yalm_state_new.argtypes = [vp, C.POINTER(vp)]
    L.yalm_state_free.argtypes = [vp]
    L.yalm_state_free.restype = None
    L.yalm_forward.argtypes = [vp, vp, i, i, i]
    L.yalm_logits.argtypes = [vp]
    L.yalm_logits.restype = C.POINTER(f)
    L.yalm_generate_greedy.argtypes = [vp, vp, i, i, i, vp, vp, C.POINTER(f)]
    L.yalm_generate.argtypes = [vp, vp, i, i, i, f, C.c_ulonglong, vp, vp, C.POINTER(f)]
    L.yalm_state_debug_set_launch_seq.argtypes = [vp, C.c_uint]
    L.yalm_state_debug_set_launch_seq.restype = None
    L.yalm_sample_host.argtypes = [vp, i, f, f, C.POINTER(i)]
    L.yalm_token_probs_host.argtypes = [vp, i, i, vp, vp]
    L.yalm_gemm_host.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, i]
    L.yalm_model_batch_capacity.argtypes = [vp]
    L.yalm_forward_batch.argtypes = [vp, vp, vp, i, i, vp, vp]
    L.yalm_tokenize.argtypes = [vp, sz, C.c_char_p, i, vp, i]
    L.yalm_detokenize.argtypes = [vp, sz, vp, i, C.c_char_p, i]
    L.yalm_state_read.argtypes = [vp, vp, i, i, vp, sz]
    L.yalm_forward_launches.argtypes = [vp, i]
    L.yalm_model_uses_megakernel.argtypes = [vp]
    L.yalm_model_mega_info.argtypes = [vp, C.POINTER(i)]
    L.yalm_mega_profile.argtypes = [vp, vp, i, i, vp, sz, C.POINTER(i)]
    L.yalm_profile_ffn.argtypes = [vp, vp, i, i, i, C.POINTER(C.c_double)]
    L.yalm_tp_unique_id.argtypes = [vp]
    L.yalm_tp_init.argtypes = [vp, i, i]
    L.yalm_tp_finalize.restype = None
    L.yalm_attn_workspace_bytes.argtypes = [i, i, i]
    L.yalm_attn_workspace_bytes.restype = sz
    L.yalm_attn_decode.argtypes = [vp, vp, vp, vp, vp, i, i, i, i, vp, vp]
    L.yalm_attn_decode_static.argtypes = [vp, vp, vp, vp, i, i, i, i, vp, vp]
    L.yalm_attn_decode_static_max_bytes.restype = sz
    L.upload_cuda.argtypes = [vp, sz]
    L.upload_cuda.restype = vp
    L.free_cuda.argtypes = [vp]
    L.free_cuda.restype = None
    _lib = L
    return L


def error_string(code: int) -> str:
    return lib().yalm_error_string(code).decode()


def _check(rc, what):
    if rc != 0:
        raise YalmError(rc, what)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


# ---- test entry points (reference: matmul_cuda / mha_cuda / ffn_cuda, host arrays in and out) ----------
def matmul_cuda(x: np.ndarray, w: np.ndarray, dtype: int = F32, scale: np.ndarray | None = None) -> np.ndarray:
    """xout = W x with W (d, n) row-major; reference argument order is (xout, x, w, n, d)."""
    d, n = w.shape
    x = np.ascontiguousarray(x, np.float32)
    w = np.ascontiguousarray(w)
    out = np.empty(d, np.float32)
    _check(lib().yalm_matmul_host(_p(out), _p(x), _p(w), _p(scale), n, d, dtype), "matmul_cuda")
    return out


def mha_cuda(kb, vb, q, head_dim, kv_len, max_seq_len, n_heads, n_kv_heads, want_att=True):
    kb, vb = np.ascontiguousarray(kb, np.uint16), np.ascontiguousarray(vb, np.uint16)
    q = np.ascontiguousarray(q, np.float32)
    xout = np.zeros(n_heads * head_dim, np.float32)
    att = np.zeros((n_heads, max_seq_len), np.float32) if want_att else None
    _check(lib().yalm_mha_host(_p(xout), _p(att), _p(kb), _p(vb), _p(q), head_dim, kv_len, max_seq_len, n_heads,
                               n_kv_heads), "mha_cuda")
    return xout, att


def ffn_cuda(x, w1, w2, w3, act=ACT_GELU, dtype=F32, s1=None, s2=None, s3=None):
    hidden, dim = w1.shape
    out = np.empty(dim, np.float32)
    x = np.ascontiguousarray(x, np.float32)
    _check(lib().yalm_ffn_host(_p(out), _p(x), _p(np.ascontiguousarray(w1)), _p(np.ascontiguousarray(w2)),
                               _p(np.ascontiguousarray(w3)), _p(s1), _p(s2), _p(s3), hidden, dim, act, dtype),
           "ffn_cuda")
    return out


def sample_cuda(logits: np.ndarray, temperature: float, r: float) -> int:
    """Sampler::sample (src/sampler.cpp:40-65) on the device with the uniform draw supplied by the caller."""
    logits = np.ascontiguousarray(logits, np.float32)
    tok = C.c_int()
    _check(lib().yalm_sample_host(_p(logits), logits.size, temperature, r, C.byref(tok)), "sample_cuda")
    return tok.value


def token_probs_cuda(logits: np.ndarray, targets) -> np.ndarray:
    """Sampler::sample_prob per row (src/sampler.cpp:11-25) on the device."""
    logits = np.ascontiguousarray(logits, np.float32)
    targets = np.ascontiguousarray(targets, np.int32)
    out = np.empty(len(targets), np.float32)
    _check(lib().yalm_token_probs_host(_p(logits), logits.shape[1], logits.shape[0], _p(targets), _p(out)), "token_probs_cuda")
    return out


def gemm_cuda(a: np.ndarray, w: np.ndarray, c: np.ndarray | None = None, w3: np.ndarray | None = None, act: int = ACT_SILU):
    """Prefill GEMM test entry: a (M, K) fp16, w (N, K) fp16 -> a w^T as fp32 (+ c), or with w3 the fused GLU as fp16."""
    a, w = np.ascontiguousarray(a, np.float16), np.ascontiguousarray(w, np.float16)
    M, K = a.shape
    N = w.shape[0]
    if w3 is not None:
        h = np.zeros((M, N), np.float16)
        _check(lib().yalm_gemm_host(None, _p(h), _p(a), _p(w), _p(np.ascontiguousarray(w3, np.float16)), M, N, K, 0, act), "gemm_cuda")
        return h
    out = np.zeros((M, N), np.float32) if c is None else np.ascontiguousarray(c, np.float32).copy()
    _check(lib().yalm_gemm_host(_p(out), None, _p(a), _p(w), None, M, N, K, int(c is not None), act), "gemm_cuda")
    return out


def tokenize(image, text: bytes, add_bos: bool = True):
    img = np.frombuffer(image, dtype=np.uint8)
    out = np.zeros(max(16, 2 * len(text) + 2), np.int32)
    n = lib().yalm_tokenize(_p(img), img.nbytes, text, int(add_bos), _p(out), out.size)
    if n < 0:
        raise YalmError(n, "tokenize")
    return out[:n].tolist()


def detokenize(image, ids) -> bytes:
    img = np.frombuffer(image, dtype=np.uint8)
    ids = np.ascontiguousarray(ids, np.int32)
    buf = C.create_string_buffer(64 * len(ids) + 16)
    n = lib().yalm_detokenize(_p(img), img.nbytes, _p(ids), len(ids), buf, len(buf))
    if n < 0:
        raise YalmError(n, "detokenize")
    return buf.raw[:n]


def bench_op(name: str, iters: int = 20, warmup: int = 3):
    us, by = C.c_double(), C.c_double()
    _check(lib().yalm_bench_op(name.encode(), iters, warmup, C.byref(us), C.byref(by)), f"bench_op({name})")
    return us.value, by.value


# ---- Model / InferenceState --------------------------------------------------------------------------
class Model:
    """yalm's `Model` (src/model.h:303-327) behind an opaque handle."""

    def __init__(self, yalm, context: int = 0):
        h = C.c_void_p()
        self._image = None
        if isinstance(yalm, str):
            _check(lib().yalm_model_open_file(yalm.encode(), context, C.byref(h)), "Model(file)")
        else:
            self._image = np.frombuffer(yalm, dtype=np.uint8)   # keep alive until cuda()
            _check(lib().yalm_model_open_memory(self._image.ctypes.data_as(C.c_void_p), self._image.nbytes, context,
                                                C.byref(h)), "Model(memory)")
        self.h = h
        self._read_config()

    def _read_config(self):
        v = (C.c_int * 16)()
        lib().yalm_model_config(self.h, v)
        (self.dim, self.hidden_dim, self.head_dim, self.n_layers, self.n_heads, self.n_kv_heads, self.vocab_size,
         self.max_seq_len, self.rotary_dim, self.n_experts, self.n_experts_active, self.weight_dtype, self.act,
         self.tp_rank, self.tp_size, _) = list(v)

    def cuda(self, device: int = 0, tp_rank: int = 0, tp_size: int = 1):
        _check(lib().yalm_model_cuda(self.h, device, tp_rank, tp_size), "Model.cuda")
        self._image = None
        self._read_config()
        return self

    def device_bytes(self) -> int:
        return lib().yalm_model_device_bytes(self.h)

    def load_info(self):
        v = (C.c_int * 2)()
        lib().yalm_model_load_info(self.h, v)
        return {"device_allocations": v[0], "pinned_staging": bool(v[1])}

    def batch_capacity(self) -> int:
        return lib().yalm_model_batch_capacity(self.h)

    def uses_megakernel(self) -> bool:
        return bool(lib().yalm_model_uses_megakernel(self.h))

    def mega_info(self):
        v = (C.c_int * 8)()
        _check(lib().yalm_model_mega_info(self.h, v), "mega_info")
        return {"stages": v[0], "vec_floats": v[1], "w2_pipe": bool(v[2]), "w2_k_splits": v[3], "mma": bool(v[4])}

    def launches(self, mode=MODE_LOGITS) -> int:
        return lib().yalm_forward_launches(self.h, mode)

    def close(self):
        if self.h:
            lib().yalm_model_close(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class InferenceState:
    """yalm's `InferenceState` (src/model.h:84-190), created on the model's GPU."""

    def __init__(self, model: Model):
        self.model = model
        h = C.c_void_p()
        _check(lib().yalm_state_new(model.h, C.byref(h)), "InferenceState")
        self.h = h

    def forward(self, token: int, pos: int, mode: int = MODE_LOGITS):
        """Model::forward (src/model.cpp:396-407). Returns the host logits view in MODE_LOGITS."""
        _check(lib().yalm_forward(self.model.h, self.h, token, pos, mode), "forward")
        return self.logits() if mode == MODE_LOGITS else None

    def logits(self) -> np.ndarray:
        return np.ctypeslib.as_array(lib().yalm_logits(self.h), (self.model.vocab_size,))

    def generate_greedy(self, token: int, pos: int, n_gen: int, keep_logits: bool = False):
        toks = np.zeros(n_gen, np.int32)
        lg = np.zeros((n_gen, self.model.vocab_size), np.float32) if keep_logits else None
        ms = C.c_float()
        _check(lib().yalm_generate_greedy(self.model.h, self.h, token, pos, n_gen, _p(toks), _p(lg), C.byref(ms)),
               "generate_greedy")
        self.device_ms = ms.value
        return (toks.tolist(), lg) if keep_logits else toks.tolist()

    def forward_batch(self, tokens, pos0: int, targets=None):
        """Positions pos0 .. pos0 + len(tokens) - 1 in one pass (Model::forward_batch); returns the per-position probabilities
        of `targets` (or None).  logits() afterwards = the last position's."""
        tokens = np.ascontiguousarray(tokens, np.int32)
        tg = None if targets is None else np.ascontiguousarray(targets, np.int32)
        probs = None if targets is None else np.zeros(len(tokens), np.float32)
        _check(lib().yalm_forward_batch(self.model.h, self.h, _p(tokens), len(tokens), pos0, _p(tg), _p(probs)), "forward_batch")
        return probs

    def generate_sampled(self, token: int, pos: int, n_gen: int, temperature: float, seed: int = 0):
        """n_gen steps with Sampler::sample on the device (temperature 0 = greedy); only token ids come back."""
        toks = np.zeros(n_gen, np.int32)
        ms = C.c_float()
        _check(lib().yalm_generate(self.model.h, self.h, token, pos, n_gen, temperature, seed, _p(toks), None, C.byref(ms)),
               "generate")
        self.device_ms = ms.value
        return toks.tolist()

    def set_launch_seq(self, seq: int):
        lib().yalm_state_debug_set_launch_seq(self.h, seq)

    def mega_profile(self, token: int, pos: int) -> np.ndarray:
        """(grid, n_layers, 6 phases, 8 slots) clock64 stamps of one forward pass on the megakernel."""
        g = C.c_int()
        _check(lib().yalm_mega_profile(self.model.h, self.h, token, pos, None, 0, C.byref(g)), "mega_profile")
        out = np.zeros((g.value, self.model.n_layers, 6, 8), np.int64)
        _check(lib().yalm_mega_profile(self.model.h, self.h, token, pos, _p(out), out.size, C.byref(g)), "mega_profile")
        return out

    def profile_ffn(self, token: int, pos: int, n_steps: int) -> float:
        us = C.c_double()
        _check(lib().yalm_profile_ffn(self.model.h, self.h, token, pos, n_steps, C.byref(us)), "profile_ffn")
        return us.value

    def generate(self, prompt, n_gen: int, keep_logits: bool = False):
        """run_completion's loop (src/main.cpp:91-112) with host-side first-max argmax, via forward()."""
        for pos, tok in enumerate(prompt):
            self.forward(tok, pos, MODE_LOGITS if pos + 1 == len(prompt) else MODE_HYDRATE)
        toks, lgs = [], []
        for i in range(n_gen):
            lg = self.logits()
            tok = int(np.argmax(lg))          # numpy argmax = first maximum = Sampler::sample_argmax
            toks.append(tok)
            if keep_logits:
                lgs.append(lg.copy())
            if i + 1 < n_gen:
                self.forward(tok, len(prompt) + i, MODE_LOGITS)
        return (toks, np.stack(lgs)) if keep_logits else toks

    def read(self, which: int, count: int, dtype=np.float32, layer: int = 0) -> np.ndarray:
        out = np.empty(count, dtype)
        _check(lib().yalm_state_read(self.model.h, self.h, which, layer, _p(out), out.nbytes), "state_read")
        return out

    def close(self):
        if self.h:
            lib().yalm_state_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

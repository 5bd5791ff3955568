#!/usr/bin/env python
"""Batch-1 decode benchmark (BASELINE.json metric): tok/s and fraction of the HBM roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

N = 1 workload = BASELINE.json configs[1]: Mistral-7B-v0.2 architecture, fp16 random-init weights,
4k context, batch-1 greedy decode on one B200.  A "step" is one decoded token = one pass of the forward
path.  The K timed steps are the LAST K positions of the 4096-token context (kv_len 4096-K+1 .. 4096), after
hydrating the KV cache through the same forward path.
  value    device-timed (CUDA events) tok/s of the on-device greedy loop, inputs resident in HBM
  e2e      same metric through the public per-token API Model::forward(token, pos) driven from the host:
           token/pos travel host->device every step and the logits come back device->host (host arg-max)
  roofline the whole-token kernel (one launch per token): algorithmic bytes (weights + KV rows) / CUDA-event duration
  parity   N = 1: the first tokens + logits of the SAME full-size image against the unmodified reference's CPU backend;
           N > 1: the K timed greedy tokens identical on every rank, and a 2-layer slim model of the same architecture,
           sharded the same way, against the reference
  cpu_baseline  the unmodified reference CPU backend (oracle/_ref) timed on this box's host cores
  gpu_reference the unmodified reference's own `-d cuda` path on this GPU (reported beside the CPU baseline)
N > 1 (torchrun, one process per GPU): the same model tensor-parallel over N ranks (strong scaling); the
all-reduce behind the row-parallel wo / w2 GEMVs happens inside the kernel over NVLink peer memory.

--impl reference: the reference's own CPU implementation of the path (oracle/_ref; the oracle port if the
reference library is absent) on the host cores, same metric / unit / config; rank 0 only.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from yalm_b200 import yalm_file as yf  # noqa: E402

WORKLOADS = {
    # name: (arch, dtype, context(-T), description)
    "mistral-7b-fp16-4k": ("mistral-7b", "fp16", 4096, "Mistral-7B-v0.2 fp16 random-init, 4k context, batch-1 decode"),
    "mistral-7b-fp8-32k": ("mistral-7b", "fp8", 32768, "Mistral-7B-v0.2 fp8(e4m3) weights, 32k sliding-window (-T 32768) decode"),
    "mixtral-8x7b-fp16-4k": ("mixtral-8x7b", "fp16", 4096, "Mixtral-8x7B fp16 top-2 MoE, 4k context, batch-1 decode"),
    "llama-3-70b-fp16-4k": ("llama-3-70b", "fp16", 4096, "Llama-3-70B-arch fp16 random-init, 4k context, batch-1 decode"),
    "llama-3.2-1b-fp32": ("llama-3.2-1b", "fp32", 4096, "Llama-3.2-1B fp32 random-init, batch-1 decode"),
    "tiny": ("tiny-moe", "fp16", 64, "tiny MoE test model (debug only)"),
    # batched prefill: a different metric (prompt tokens hydrated per second), its own line
    "mistral-7b-fp16-prefill": ("mistral-7b", "fp16", 4096, "Mistral-7B-v0.2 fp16 random-init, 4096-token prompt hydrated in 512-token chunks (tcgen05 GEMMs)"),
}
HEADLINE = "mistral-7b-fp16-4k"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE timed-shape launch of the whole-token kernel on the headline
    workload — logits mode at pos ~4090, captured by tools/gpu/ncu_full_logits.sh with `ncu --set full` and summarised
    into profiles/*logits*_full_summary.txt (tools/ncu_summary.py).  None if no such capture is committed."""
    import glob
    import re
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*logits*full_summary.txt")))
    if not files:
        return None, None
    txt = open(files[-1]).read()
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    tot = 0.0
    for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        m = re.search(rf"^{re.escape(key)}\s+([0-9.]+)\s+(\w+)\s*$", txt, re.M)
        if not m or m.group(2) not in unit:
            return None, None
        tot += float(m.group(1)) * unit[m.group(2)]
    return tot, os.path.relpath(files[-1], ROOT) + " (one logits-mode launch at pos ~4090, ncu --set full)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def shm_path(name):
    d = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else "/tmp"
    return os.path.join(d, name)


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def write_model(path, arch, dtype, part=None, create=True):
    """Seeded random-init image of `arch` straight into `path` (large tensors drawn on the GPU when there is one).
    part = (rank, world): this process draws every world-th tensor (all ranks of a node fill one file together)."""
    import torch
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    yf.random_model_streamed(arch, dtype, seed=0, device=dev, path=path, part=part, create=create)


def run_reference(args, path, context, start=(1, 0), want_logits=False):
    """Reference arm / cpu_baseline: the reference's CPU backend on the host cores over the image at `path`.  One step =
    one decoded token (weights streamed once) at a short context: the CPU path is weight-bound (KV at 4k context adds
    3.8 % to the bytes of the headline model: stated in BASELINE.md).  Returns (record, seconds, tokens, last logits)."""
    # torchrun exports OMP_NUM_THREADS=1 to its workers: the reference arm uses every host thread it can get
    n_thr = host_threads()
    os.environ["OMP_NUM_THREADS"] = str(n_thr)
    from oracle import oracle as O
    kind = "reference" if O.ref_available() else "port"
    if kind == "reference":
        O.ref_lib().ref_set_threads(n_thr)
        m = O.RefModel(path, context)
        cores = O.ref_lib().ref_omp_threads()
    else:
        m = O.OracleModel(path, context)
        cores = n_thr
    tok, pos = start
    toks, lg = [], None
    for _ in range(args.warmup):
        lg = m.forward(tok, pos, 1)
        tok = m.argmax()
        toks.append(tok)
        pos += 1
    t0 = time.perf_counter()
    for _ in range(args.steps):
        lg = m.forward(tok, pos, 1)
        tok = m.argmax()
        toks.append(tok)
        pos += 1
    dt = time.perf_counter() - t0
    last = np.array(lg, np.float32, copy=True) if want_logits else None
    if kind == "reference":
        m.close()
    v = args.steps / dt
    rec = {"value": v, "unit": "tok/s", "cores": cores, "kind": kind,
           "sample": f"{args.steps} greedy decode steps of the same model after {args.warmup} warm-up steps, "
                     f"positions {start[1] + args.warmup}..{start[1] + args.warmup + args.steps - 1} (short context), "
                     f"{'oracle/_ref (unmodified reference, OpenMP)' if kind == 'reference' else 'oracle C port'}"}
    return rec, dt, toks, last


GPU_REF_SNIPPET = r"""
import json, sys, time
sys.path.insert(0, {root!r})
from oracle import oracle as O
m = O.RefModel({path!r}, {context})
m.cuda()
prompt = [1, 17, 300, 42, 5, 9, 77, 4]
m.generate(prompt, {warm})                 # captures + instantiates the reference's CUDA graph, warms up
t = m.generate([3], {steps})                 # positions restart at 0: decode-only seconds of {steps} greedy steps
print("GPUREF " + json.dumps({{"tok_s": ({steps} - 1) / m.decode_seconds}}))   # its timed loop holds steps - 1 forwards
"""


def run_gpu_reference(path, context, steps=32, warm=8):
    """The unmodified reference's own `-d cuda` path (src/infer.cu, compiled for sm_100 by oracle/Makefile) on this GPU, in
    a subprocess (its error convention is abort()).  Short context, decode-only seconds from its own greedy loop."""
    from oracle import oracle as O
    if not O.ref_available():
        return {"value": None, "unit": "tok/s", "note": "oracle/_ref not built"}
    code = GPU_REF_SNIPPET.format(root=ROOT, path=path, context=context, steps=steps, warm=warm)
    try:
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    except subprocess.TimeoutExpired:
        return {"value": None, "unit": "tok/s", "note": "reference -d cuda timed out (300 s)"}
    for ln in r.stdout.splitlines():
        if ln.startswith("GPUREF "):
            v = json.loads(ln[7:])["tok_s"]
            return {"value": v, "unit": "tok/s",
                    "note": f"unmodified reference, Device::CUDA (src/infer.cu built -gencode arch=compute_100,code=sm_100), same image, "
                            f"{steps} greedy decode steps at positions 1..{steps} after warm-up (short context: favours the reference), "
                            "host-timed around its own loop incl. its per-token stream sync; reference: unsupported for MoE / fp8 / "
                            "multi-GPU / positions past the window (BASELINE.md 4.5)"}
    return {"value": None, "unit": "tok/s", "note": f"reference -d cuda failed (rc {r.returncode}): {(r.stderr or r.stdout)[-300:]}"}


def run_prefill(args, arch, dtype, context, desc, base, path):
    """--workload *-prefill: hydrate a `context`-token prompt in chunks through Model::forward_batch.  One step = one chunk.
    value = prompt tokens per second (CUDA events around the K timed chunks = the last K chunks of the context, after W
    untimed ones); roofline bound = tensor: 2 * weight parameters * tokens (+ attention) FLOP against the measured bf16
    tensor throughput.  The token loop it replaces (the reference's hydrate, src/main.cpp:91-97) is timed beside it."""
    import torch
    from yalm_b200 import api
    write_model(path, arch, dtype)
    try:
        model = api.Model(path, context).cuda(0)
    finally:
        os.unlink(path)
    state = api.InferenceState(model)
    cap, V = model.batch_capacity(), model.vocab_size
    if not cap:
        raise SystemExit("bench.py: this model has no batched prefill path")
    ids = [(i * 7919 + 1) % V for i in range(context)]
    chunks = [(p, min(cap, context - p)) for p in range(0, context, cap)]
    K = min(args.steps, len(chunks) - 3)
    W = len(chunks) - K
    clocks = ClockSampler(0)
    clocks.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    t_host = 0.0
    for i, (p, n) in enumerate(chunks):
        if i == W:
            torch.cuda.synchronize()
            ev[0].record(torch.cuda.current_stream())   # forward_batch synchronises its own stream before returning
            t_host = time.perf_counter()
        state.forward_batch(ids[p:p + n], p)
    torch.cuda.synchronize()
    t_host = time.perf_counter() - t_host
    clk = clocks.stop()
    toks = sum(n for _, n in chunks[W:])
    # token-loop hydrate of the same model for comparison (64 positions at the end of the context)
    st2 = api.InferenceState(model)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for pos in range(context - 64, context):
        st2.forward(ids[pos], pos, api.MODE_HYDRATE)
    torch.cuda.synchronize()
    loop_tok_s = 64 / (time.perf_counter() - t0)
    a = arch
    q_dim, kv_dim = a.n_heads * a.head_dim, a.n_kv_heads * a.head_dim
    params = a.n_layers * (2 * q_dim * a.dim + 2 * kv_dim * a.dim + 3 * a.dim * a.hidden_dim)
    mean_kv = sum(p + n / 2 for p, n in chunks[W:]) / K
    flop = 2.0 * params * toks + 4.0 * a.n_layers * q_dim * mean_kv * toks
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak = float(json.load(open(pk))["bf16_tflops_sustained"]) if os.path.exists(pk) else 1374.5
    line = dict(base, impl="ours", metric="prefill tok/s (prompt tokens hydrated per second)", value=toks / t_host, steps=K, warmup=W,
                ms_per_step=1000 * t_host / K, clocks=clk, gpu_launches=K * (2 + 12 * a.n_layers),
                e2e={"value": toks / t_host, "unit": "tok/s", "h2d_bytes_per_step": 4 * cap, "d2h_bytes_per_step": 4 * V,
                     "note": "host-timed around Model::forward_batch calls with host token buffers; each call copies the chunk's ids in and the last logits out"},
                roofline={"bound": "tensor", "kernel": "gemm_f16_kernel (tcgen05.mma kind::f16, TMA operands, TMEM accumulator) + attn_prefill_kernel",
                          "achieved": flop / t_host / 1e12, "peak": peak, "unit": "TFLOP/s", "frac": flop / t_host / 1e12 / peak,
                          "traffic": None, "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained"},
                token_loop_hydrate_tok_s=loop_tok_s, speedup_vs_token_loop=toks / t_host / loop_tok_s,
                cpu_baseline=None)
    _emit(line)
    state.close()
    st2.close()
    model.close()


def _emit(line):
    """The one JSON line of the contract, on the real stdout (see main: fd 1 is pointed at stderr meanwhile)."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def greedy_host_loop(state, tok, pos, n):
    toks, lg = [], None
    for _ in range(n):
        lg = state.forward(tok, pos)
        tok = int(np.argmax(lg))
        toks.append(tok)
        pos += 1
    return toks, np.array(lg, np.float32, copy=True)


def main():
    # Native libraries (the loader's log line, NCCL's version banner, the reference's spdlog) write to fd 1:
    # keep a private copy of stdout for the JSON line and send everything else to stderr.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=HEADLINE, choices=sorted(WORKLOADS))
    ap.add_argument("--layers", type=int, default=0, help="debug: truncate the model to this many layers")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-reference", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity legs (ncu captures: keeps launch indices simple)")
    ap.add_argument("--moe-ep", action="store_true", help="MoE: expert parallelism (whole experts per rank) instead of slicing every expert")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    if args.moe_ep:
        os.environ["YALM_MOE_EP"] = "1"
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    arch_name, dtype, context, desc = WORKLOADS[args.workload]
    arch = yf.ARCHS[arch_name]
    if args.layers:
        arch = yf.slim(arch, args.layers)
    base = {"metric": "batch-1 decode tok/s", "unit": "tok/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "strong" if world > 1 else "weak",
            "vs_baseline": None, "dtype": dtype, "data": "synthetic (seeded random-init weights, synthetic token ids)",
            "config": {"workload": desc, "arch": arch.name, "weights": dtype, "context": context,
                       "parallelism": (f"ep{world} (whole experts per rank) + tp{world} attention / LM head" if args.moe_ep else f"tp{world}") if world > 1 else "single-gpu",
                       "l2": "inputs larger than L2 (weights streamed per token >> 126 MB)"}}
    tag = f"{arch.name}_{dtype}_{os.environ.get('MASTER_PORT', '0')}_{os.getppid() if world > 1 else os.getpid()}"
    path = shm_path(f"yalm_bench_{tag}.yalm")

    if args.impl == "reference":
        if rank != 0:
            return
        write_model(path, arch, dtype)
        try:
            cb, dt, _, _ = run_reference(args, path, context)
        finally:
            os.unlink(path)
        line = dict(base, impl="reference", n_gpus=world, value=cb["value"], ms_per_step=1000 * dt / args.steps,
                    cpu_baseline=cb, gpu_launches=0,
                    e2e={"value": cb["value"], "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
        _emit(line)
        return

    import torch
    import torch.distributed as dist
    from yalm_b200 import api
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; yalm_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if args.workload.endswith("-prefill"):
        if world > 1:
            raise SystemExit("bench.py: the prefill workload is single-GPU")
        return run_prefill(args, arch, dtype, context, desc, base, path)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            buf = np.zeros(128, np.uint8)
            api._check(api.lib().yalm_tp_unique_id(buf.ctypes.data), "tp_unique_id")
            idt.copy_(torch.from_numpy(buf))
        dist.broadcast(idt, 0)
        buf = idt.cpu().numpy()
        api._check(api.lib().yalm_tp_init(buf.ctypes.data, rank, world), "tp_init")

    def barrier():
        # Drain this rank's own launches BEFORE the collective: hydrate launches are asynchronous, and an NCCL kernel that
        # slips between two queued whole-token launches on one rank (it waits for its peers) while a peer's NCCL kernel is
        # still queued behind a whole-token kernel (which waits for this rank's next launch) is a cross-GPU deadlock — the
        # cooperative launch needs every SM, so it cannot start beside the spinning NCCL kernel.
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def load(p, ctx):
        m = api.Model(p, ctx).cuda(local_rank, rank, world)
        return m, api.InferenceState(m)

    # ---- the model image: ONE copy per node (rank 0 writes it under /dev/shm, every rank maps it and uploads its slices)
    t_gen = time.perf_counter()
    if rank == 0:
        write_model(path, arch, dtype, part=(0, world))     # sizes the file, writes the header and its share of the tensors
    barrier()
    if rank != 0:
        write_model(path, arch, dtype, part=(rank, world), create=False)
    torch.cuda.empty_cache()
    barrier()
    t_gen = time.perf_counter() - t_gen

    cpu_baseline = gpu_reference = parity = None
    ref_toks = ref_last = None
    try:
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            # bounded sample of the same workload on the host cores (reported baseline, not the target); its tokens and last
            # logits are what the GPU arm is checked against below
            sub = argparse.Namespace(steps=12, warmup=2)
            try:
                cpu_baseline, _, ref_toks, ref_last = run_reference(sub, path, context, want_logits=True)
            except Exception as e:  # the checker failing must not hide the GPU number
                cpu_baseline = {"value": None, "unit": "tok/s", "cores": host_threads(), "kind": "reference",
                                "sample": f"failed: {e}"}
        if rank == 0 and world == 1 and not args.no_gpu_reference and not arch.n_experts and dtype != "fp8":
            gpu_reference = run_gpu_reference(path, context)

        t_load = time.perf_counter()
        model, state = load(path, context)
        t_load = time.perf_counter() - t_load
        load_info = model.load_info()
        barrier()                       # every rank has uploaded its slices: the image can go
    finally:
        if rank == 0 and os.path.exists(path):
            os.unlink(path)

    K, W, V = args.steps, args.warmup, model.vocab_size
    # ---- parity, N = 1: the full-size image, first tokens from (token 1, pos 0), vs the unmodified reference (CPU) -----
    if ref_toks is not None and not args.no_parity:
        sp = api.InferenceState(model)
        got, last = greedy_host_loop(sp, 1, 0, len(ref_toks))
        sp.close()
        parity = {"tokens_identical": got == ref_toks, "max_abs_dlogit": float(np.abs(last - ref_last).max()),
                  "n_tokens": len(ref_toks), "tolerance": 2e-2,
                  "vs": f"oracle/_ref (unmodified reference, CPU backend) on the same {arch.n_layers}-layer image, positions 0..{len(ref_toks) - 1}"}
    # ---- parity, N > 1: a 2-layer slim model of the same architecture, sharded the same way, vs the reference on rank 0 -----
    slim_par = None
    if world > 1 and not args.no_parity:
        sarch = yf.slim(arch, 2)
        spath = shm_path(f"yalm_bench_slim_{tag}.yalm")
        want = None
        if rank == 0:
            write_model(spath, sarch, dtype)
            try:    # the CPU reference first: the other ranks wait at the barrier, not inside a kernel
                _, _, want, want_last = run_reference(argparse.Namespace(steps=6, warmup=2), spath, context, want_logits=True)
            except Exception as e:
                slim_par = {"tokens_identical": None, "note": f"reference failed: {e}"}
        barrier()
        try:
            sm, ss = load(spath, context)
            got, last = greedy_host_loop(ss, 1, 0, 8)
            ss.close()
            sm.close()
            barrier()
        finally:
            if rank == 0 and os.path.exists(spath):
                os.unlink(spath)
        if rank == 0 and want is not None:
            slim_par = {"tokens_identical": got == want, "max_abs_dlogit": float(np.abs(last - want_last).max()), "n_tokens": 8,
                        "tolerance": 2e-2, "vs": f"oracle/_ref (CPU) on a 2-layer slim {arch.name}, tp{world} vs unsharded reference"}

    p0 = max(1, context - 2 * (K + W))
    for pos in range(p0):                                   # prompt hydrate, same forward path, untimed
        state.forward((pos * 7919 + 1) % V, pos, api.MODE_HYDRATE)
    barrier()

    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    # ---- e2e: host-driven Model::forward(token,pos) -> host logits -> host first-max argmax ------------------
    tok, pos = 1, p0
    for _ in range(W):
        tok = int(np.argmax(state.forward(tok, pos)))
        pos += 1
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        tok = int(np.argmax(state.forward(tok, pos)))
        pos += 1
    barrier()
    e2e_s = time.perf_counter() - t0
    # ---- value: on-device greedy loop, CUDA events ---------------------------------------------------------------
    toks = state.generate_greedy(tok, pos, W)
    tok, pos = toks[-1], pos + W
    barrier()
    toks = state.generate_greedy(tok, pos, K)
    dev_ms = state.device_ms
    barrier()
    tok, pos = toks[-1], pos + K
    clk = clocks.stop() if rank == 0 else None
    # ---- dominant kernel, CUDA events around each launch (separate pass, does not perturb the numbers above) -----
    mega = model.uses_megakernel()
    ffn_us = 0.0 if mega else state.profile_ffn(tok, pos - 1, min(K, 8))

    ranks_agree = None
    if world > 1:
        t = torch.tensor([dev_ms, e2e_s, ffn_us], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, e2e_s, ffn_us = t.tolist()
        mine = torch.tensor(toks, dtype=torch.int32, device="cuda")
        allt = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allt, mine)
        ranks_agree = all(bool((a == allt[0]).all().item()) for a in allt)
    if rank == 0:
        if world > 1:
            parity = {"timed_tokens_identical_on_all_ranks": ranks_agree, "n_tokens": K, "slim": slim_par}
        peak, peak_src = peaks()
        wsize = {"fp32": 4, "fp16": 2, "fp8": 1}[dtype]
        k_act = max(1, arch.n_experts_active)
        ffn_bytes = k_act * 2 * arch.hidden_dim * arch.dim * wsize / world + 4 * arch.dim * 2
        wbytes = yf.weight_bytes_per_token(arch, dtype)
        kvbytes = yf.kv_bytes_per_token(arch, min(context, pos))
        tok_s = K / (dev_ms * 1e-3)
        if mega:
            # one persistent kernel per token IS the step: algorithmic bytes = weights + KV rows it reads
            kname = "decode_mega_kernel (whole token: all GEMVs + flash-decode + LM head, TMA ring)"
            ffn_bytes = (wbytes + yf.kv_bytes_per_token(arch, min(context, pos - K // 2))) / world   # this rank's shard
            if world > 1:
                kname += f", tp{world}: all-reduce fused into the wo/w2 epilogues over NVLink peer memory"
            ffn_us = 1000.0 * dev_ms / K
        else:
            kname = "ffn_glu_kernel (rmsnorm + w1/w3 GEMV + GLU)"
        achieved = ffn_bytes / (ffn_us * 1e-6) / 1e9
        # DRAM bytes of one launch from the committed ncu capture: only for the workload it was taken on
        traffic, traffic_src = ncu_traffic() if (mega and world == 1 and args.workload == HEADLINE and not args.layers) else (None, None)
        line = dict(base, impl="ours", value=tok_s, ms_per_step=dev_ms / K, clocks=clk,
                    e2e={"value": K / e2e_s, "unit": "tok/s", "h2d_bytes_per_step": 8, "d2h_bytes_per_step": 4 * V},
                    gpu_launches=(model.launches(api.MODE_LOGITS) + (0 if mega else 1)) * K,
                    roofline={"bound": "hbm", "kernel": kname,
                              "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                              "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                              "bytes_per_launch": ffn_bytes,
                              "launch_us": ffn_us},
                    parity=parity, cpu_baseline=cpu_baseline, gpu_reference=gpu_reference,
                    whole_step={"weight_bytes_per_token": wbytes, "kv_bytes_per_token": kvbytes,
                                "gbs_weights": wbytes * tok_s / 1e9 / world, "frac_of_measured_peak_weights": wbytes * tok_s / 1e9 / world / peak,
                                "frac_of_measured_peak_weights_kv": (wbytes + kvbytes) * tok_s / 1e9 / world / peak,
                                "frac_of_nominal_8TBs_weights": wbytes * tok_s / 8e12 / world,
                                "kv_len_range": [pos - K, pos]},
                    setup_s={"generate": round(t_gen, 1), "load_upload": round(t_load, 1), "loader": load_info})
        _emit(line)
    state.close()
    model.close()
    if world > 1:
        api.lib().yalm_tp_finalize()
        dist.destroy_process_group()
    if parity and (parity.get("tokens_identical") is False or parity.get("timed_tokens_identical_on_all_ranks") is False):
        raise SystemExit("bench.py: PARITY FAILURE (see the parity key of the line above)")


if __name__ == "__main__":
    main()

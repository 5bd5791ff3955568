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
  roofline dominant kernel (fused rmsnorm + w1/w3 + SwiGLU GEMV), algorithmic bytes / CUDA-event duration
  cpu_baseline  the unmodified reference CPU backend (oracle/_ref) timed on this box's host cores
N > 1 (torchrun, one process per GPU): the same model tensor-parallel over N ranks (strong scaling),
NCCL all-reduce after the row-parallel wo / w2 GEMVs.

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
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one megakernel launch on the headline workload, from the newest
    committed `ncu --set full` summary (profiles/*_full_summary.txt, written by tools/ncu_summary.py).  None if absent."""
    import glob
    import re
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*mega*full_summary.txt")))
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
    return tot, os.path.relpath(files[-1], ROOT) + " (captured launch: hydrate step at pos 3 = no LM head, no KV: 13.96 GB algorithmic)"


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


def run_reference(args, arch, dtype, context, desc):
    """Reference arm / cpu_baseline: the reference's CPU backend on the host cores.  One step = one decoded
    token (weights streamed once) at a short context: the CPU path is weight-bound, KV at 4k adds < 4 %."""
    from oracle import oracle as O
    import torch
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    img = yf.random_model_streamed(arch, dtype, seed=0, device=dev)
    path = shm_path(f"yalm_bench_{arch.name}_{dtype}.yalm")
    with open(path, "wb") as f:
        f.write(memoryview(img))
    del img
    kind = "reference" if O.ref_available() else "port"
    try:
        if kind == "reference":
            m = O.RefModel(path, context)
            cores = O.ref_lib().ref_omp_threads()
            step = lambda tok, pos: (m.forward(tok, pos, 1), m.argmax())[1]
        else:
            m = O.OracleModel(path, context)
            cores = os.cpu_count()
            step = lambda tok, pos: (m.forward(tok, pos, 1), m.argmax())[1]
        tok, pos = 1, 0
        for _ in range(args.warmup):
            tok = step(tok, pos)
            pos += 1
        t0 = time.perf_counter()
        for _ in range(args.steps):
            tok = step(tok, pos)
            pos += 1
        dt = time.perf_counter() - t0
    finally:
        os.unlink(path)
    v = args.steps / dt
    return {"value": v, "unit": "tok/s", "cores": cores, "kind": kind,
            "sample": f"{args.steps} greedy decode steps of the same model after {args.warmup} warm-up steps, "
                      f"positions {args.warmup}..{args.warmup + args.steps - 1} (short context), "
                      f"{'oracle/_ref (unmodified reference, OpenMP)' if kind == 'reference' else 'oracle C port'}"}, dt


def _emit(line):
    """The one JSON line of the contract, on the real stdout (see main: fd 1 is pointed at stderr meanwhile)."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


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
    ap.add_argument("--workload", default="mistral-7b-fp16-4k", choices=sorted(WORKLOADS))
    ap.add_argument("--layers", type=int, default=0, help="debug: truncate the model to this many layers")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

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
                       "parallelism": f"tp{world}" if world > 1 else "single-gpu",
                       "l2": "inputs larger than L2 (weights streamed per token >> 126 MB)"}}

    if args.impl == "reference":
        if rank != 0:
            return
        cb, dt = run_reference(args, arch, dtype, context, desc)
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
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    t_gen = time.perf_counter()
    img = yf.random_model_streamed(arch, dtype, seed=0, device="cuda")
    torch.cuda.empty_cache()
    t_gen = time.perf_counter() - t_gen

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # bounded sample of the same workload on the host cores (reported baseline, not the target)
        sub = argparse.Namespace(steps=12, warmup=2)
        try:
            cpu_baseline, _ = run_reference(sub, arch, dtype, context, desc)
        except Exception as e:  # the checker failing must not hide the GPU number
            cpu_baseline = {"value": None, "unit": "tok/s", "cores": os.cpu_count(), "kind": "reference",
                            "sample": f"failed: {e}"}

    t_load = time.perf_counter()
    model = api.Model(img, context).cuda(local_rank, rank, world)
    del img
    state = api.InferenceState(model)
    t_load = time.perf_counter() - t_load

    K, W, V = args.steps, args.warmup, model.vocab_size
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

    if world > 1:
        t = torch.tensor([dev_ms, e2e_s, ffn_us], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, e2e_s, ffn_us = t.tolist()
    if rank == 0:
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
        traffic, traffic_src = ncu_traffic() if (mega and world == 1 and args.workload == "mistral-7b-fp16-4k" and not args.layers) else (None, None)
        line = dict(base, impl="ours", value=tok_s, ms_per_step=dev_ms / K, clocks=clk,
                    e2e={"value": K / e2e_s, "unit": "tok/s", "h2d_bytes_per_step": 8, "d2h_bytes_per_step": 4 * V},
                    gpu_launches=(model.launches(api.MODE_LOGITS) + (0 if mega else 1)) * K,
                    roofline={"bound": "hbm", "kernel": kname,
                              "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                              "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                              "bytes_per_launch": ffn_bytes,
                              "launch_us": ffn_us},
                    cpu_baseline=cpu_baseline,
                    whole_step={"weight_bytes_per_token": wbytes, "kv_bytes_per_token": kvbytes,
                                "gbs_weights": wbytes * tok_s / 1e9 / world, "frac_of_measured_peak_weights": wbytes * tok_s / 1e9 / world / peak,
                                "frac_of_measured_peak_weights_kv": (wbytes + kvbytes) * tok_s / 1e9 / world / peak,
                                "frac_of_nominal_8TBs_weights": wbytes * tok_s / 8e12 / world,
                                "kv_len_range": [pos - K, pos]},
                    setup_s={"generate": round(t_gen, 1), "load_upload": round(t_load, 1)})
        _emit(line)
    state.close()
    model.close()
    if world > 1:
        api.lib().yalm_tp_finalize()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""A/B of plan-time settings of the whole-token kernel on one model image (the model is re-created per setting, the
environment is read when the plan is made): tok/s of the on-device greedy loop at ~4k context and token identity.
    python tools/mega_ab.py [--workload NAME] [--pos P] "YALM_MEGA_W2_PIPE=0" "YALM_MEGA_STAGES=20" ..."""
import argparse
import os
import sys

os.environ["YALM_MEGA_TUNE"] = "1"  # launch-time knobs (YALM_MEGA_DEBUG, ...) are re-read at every launch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from yalm_b200 import api, yalm_file as yf  # noqa: E402
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="mistral-7b-fp16-4k")
    ap.add_argument("--pos", type=int, default=3600)
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("settings", nargs="*")
    args = ap.parse_args()
    arch_name, dtype, context, desc = bench.WORKLOADS[args.workload]
    arch = yf.ARCHS[arch_name]
    if args.layers:
        arch = yf.slim(arch, args.layers)
    img = yf.random_model_streamed(arch, dtype, seed=0, device="cuda")
    ref = None
    for st in [""] + list(args.settings):
        keys = []
        for kv in st.split():
            k, v = kv.split("=")
            os.environ[k] = v
            keys.append(k)
        m = api.Model(img, context).cuda()
        s = api.InferenceState(m)
        for p in range(args.pos):
            s.forward((p * 7919 + 1) % m.vocab_size, p, api.MODE_HYDRATE)
        best = 0.0
        for rep in range(3):
            s.generate_greedy(1, args.pos, 8)
            toks = s.generate_greedy(1, args.pos, 48)
            best = max(best, 48 / (s.device_ms * 1e-3))
        if ref is None:
            ref = toks
        print(f"{st or 'default':50s} {best:7.1f} tok/s  {'tokens identical' if toks == ref else 'TOKENS DIFFER'}", flush=True)
        s.close()
        m.close()
        for k in keys:
            os.environ.pop(k, None)


if __name__ == "__main__":
    main()

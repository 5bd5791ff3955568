#!/usr/bin/env python
"""Sweep of the megakernel's run-time knobs on one loaded model (with YALM_MEGA_TUNE=1 the knobs are read from the
environment at every launch): tok/s of the on-device greedy loop at 4k context per setting.
    python tools/mega_sweep.py "YALM_MEGA_L2_AHEAD=-16" "YALM_MEGA_L2_AHEAD=-16 YALM_MEGA_DEBUG=16" ..."""
import os
import sys

os.environ["YALM_MEGA_TUNE"] = "1"  # the library re-reads its knobs at every launch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from yalm_b200 import api, yalm_file as yf  # noqa: E402
import bench  # noqa: E402

KNOBS = ["YALM_MEGA_L2_AHEAD", "YALM_MEGA_DEBUG", "YALM_MEGA_STAGES"]


def main():
    settings = sys.argv[1:] or [""]
    arch_name, dtype, context, desc = bench.WORKLOADS["mistral-7b-fp16-4k"]
    img = yf.random_model_streamed(yf.ARCHS[arch_name], dtype, seed=0, device="cuda")
    m = api.Model(img, context).cuda()
    del img
    s = api.InferenceState(m)
    pos = 3600
    for p in range(pos):
        s.forward((p * 7919 + 1) % m.vocab_size, p, api.MODE_HYDRATE)
    ref = None
    for rep in range(2):
        for st in [""] + settings:
            for k in KNOBS:
                os.environ.pop(k, None)
            for kv in st.split():
                k, v = kv.split("=")
                os.environ[k] = v
            s.generate_greedy(1, pos, 8)
            toks = s.generate_greedy(1, pos, 48)
            if ref is None:
                ref = toks
            print(f"{st or 'default':50s} {48 / (s.device_ms * 1e-3):7.1f} tok/s  {'tokens identical' if toks == ref else 'TOKENS DIFFER'}", flush=True)


if __name__ == "__main__":
    main()

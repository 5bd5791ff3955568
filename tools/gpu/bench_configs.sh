#!/bin/bash
# bench lines for the headline config and the fp8 / 32k sliding-window config
cd /root/repo
timeout 600 python bench.py --steps 64 --warmup 4 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/bench_fp16_4k.json; python -c "import json; d=json.load(open('gpurun_out/bench_fp16_4k.json')); print('fp16-4k', round(d['value'],1), d['roofline']['frac'])"
timeout 1200 python bench.py --workload mistral-7b-fp8-32k --steps 32 --warmup 3 --no-cpu-baseline 2> gpurun_out/bench_fp8_32k.err | tail -1 > gpurun_out/bench_fp8_32k.json; tail -3 gpurun_out/bench_fp8_32k.err; python -c "import json; d=json.load(open('gpurun_out/bench_fp8_32k.json')); print('fp8-32k', round(d['value'],1), d['roofline'], d['whole_step'])"

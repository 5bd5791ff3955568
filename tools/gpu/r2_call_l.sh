#!/bin/bash
# Round-2 GPU call L (1 GPU): kernel variants as separate instantiations — whole-token kernel tests (all three variants), then the
# headline model on the default variant (tools/mega_ab.py) and the bench line.
cd /root/repo
mkdir -p gpurun_out
YALM_MEGA_TIMEOUT_S=6 timeout 900 python -m pytest tests/test_gpu_mega.py tests/test_gpu_model.py -x -q > gpurun_out/l_mega_tests.log 2>&1; echo "mega+model tests rc=$?"; tail -4 gpurun_out/l_mega_tests.log
timeout 600 python tools/mega_ab.py "YALM_MEGA_W2_PIPE=1" "YALM_MEGA_MMA=1" > gpurun_out/l_ab.log 2>&1; echo "ab rc=$?"; grep -v "loading model" gpurun_out/l_ab.log | tail -5
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/r2b_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['parity']['tokens_identical'], d['clocks'])"

#!/bin/bash
# Round-2 final GPU call (1 GPU): the full GPU suite, smoke, the bench line (parity + cpu_baseline + gpu_reference), the ncu launch
# list of the timed launches, one --set full capture of a LOGITS-mode launch at pos ~4089 (the shape the bench times), the phase
# profile at 4k and the fp8 / 32k phase profile.
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r2b_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/r2b_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2b_smoke.log 2>&1; tail -3 gpurun_out/r2b_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench rc=$?"; tail -c 1800 gpurun_out/r2b_bench.json
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-gpu-reference --no-parity"
# whole-token kernel launches: 4086 hydrate, then e2e 3 + 2 (logits mode, pos 4086..4090), then device loop 3 + 2 (greedy mode)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:decode_mega --launch-skip 4080 --launch-count 16 --csv --log-file gpurun_out/r2b_launches_timed.csv $B > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"; tail -3 gpurun_out/r2b_launches_timed.csv
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:decode_mega --launch-skip 4089 --launch-count 1 -o gpurun_out/r2b_mega_logits4089_full -f $B > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -2 gpurun_out/ncu_full.log
timeout 600 python tools/mega_profile.py --pos 4000 --out gpurun_out/r2b_mega_phase_profile.txt > gpurun_out/phase.log 2>&1; echo "phase rc=$?"; tail -12 gpurun_out/r2b_mega_phase_profile.txt
timeout 600 python tools/mega_profile.py --workload mistral-7b-fp8-32k --pos 32700 --out gpurun_out/r2b_mega_phase_profile_fp8_32k.txt > gpurun_out/phase32k.log 2>&1; echo "phase 32k rc=$?"; tail -12 gpurun_out/r2b_mega_phase_profile_fp8_32k.txt

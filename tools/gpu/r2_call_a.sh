#!/bin/bash
# Round-2 GPU call A (1 GPU): system facts, the full GPU suite, smoke, the bench (both arms), the ncu launch list of the
# timed launches and one --set full capture of a LOGITS-mode launch at pos ~4089 (the shape the bench times).
cd /root/repo
mkdir -p gpurun_out
{ nvidia-smi -L; nproc; free -g | head -2; df -h /dev/shm | tail -1; } > gpurun_out/sysinfo.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -5 gpurun_out/gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r2a.json 2> gpurun_out/bench_r2a.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench_r2a.json
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_r2a_reference.json 2> gpurun_out/bench_r2a_reference.err; echo "reference rc=$?"
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-gpu-reference --no-parity"
# whole-token kernel launches: 4086 hydrate, then e2e 3 + 2 (logits mode, pos 4086..4090), then device loop 3 + 2 (greedy mode)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:decode_mega --launch-skip 4080 --launch-count 16 --csv --log-file gpurun_out/launches_r2a.csv $B > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"; tail -4 gpurun_out/launches_r2a.csv
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:decode_mega --launch-skip 4089 --launch-count 1 -o gpurun_out/mega_logits4089_full -f $B > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -2 gpurun_out/ncu_full.log
timeout 600 python tools/mega_profile.py --pos 4000 --out gpurun_out/r2a_mega_phase_profile.txt > gpurun_out/phase.log 2>&1; echo "phase rc=$?"; cat gpurun_out/r2a_mega_phase_profile.txt

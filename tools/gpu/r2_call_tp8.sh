#!/bin/bash
# Round-2 8-GPU call: tensor-parallel parity at world 8 (whole-token kernel path), then the Llama-3-70B-arch TP8 bench line
# (BASELINE config 5's decode half).  Extra workloads as arguments.
N=8
WL=${@:-llama-3-70b-fp16-4k}
cd /root/repo
mkdir -p gpurun_out
nvidia-smi -L | head -8; free -g | head -2 | tail -1
YALM_TP_MEGA=1 OMP_NUM_THREADS=4 timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 tests/tp_worker.py > gpurun_out/tp${N}_parity_mega1.log 2>&1
echo "tp$N parity mega=1 rc=$?"; grep -E "PASS|FAIL|skipped|all-reduce over" gpurun_out/tp${N}_parity_mega1.log | grep -E "rank 0|FAIL|skipped|all-reduce" | head -12
P=29720
for W in $WL; do
  P=$((P+1))
  S=$(echo $W | sed 's/-fp16-4k//; s/-8x7b//; s/-3-70b/70b/; s/mistral-7b/mistral/')
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --workload $W --steps 32 --warmup 4 > gpurun_out/r2_bench_${S}_tp${N}.json 2> gpurun_out/r2_bench_${S}_tp${N}.err
  echo "$W tp$N rc=$?"; tail -c 900 gpurun_out/r2_bench_${S}_tp${N}.json | head -c 900; echo; grep -iE "error|Traceback|PARITY|timed out" gpurun_out/r2_bench_${S}_tp${N}.err | head -5
done

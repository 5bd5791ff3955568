#!/bin/bash
# Round-2 multi-GPU call (N = $1 GPUs): tensor-parallel parity at world N (both paths), then bench lines at N.
# usage: r2_call_tp.sh N [workloads...]      (default workloads: mistral-7b-fp16-4k)
N=${1:-2}; shift
WL=${@:-mistral-7b-fp16-4k}
cd /root/repo
mkdir -p gpurun_out
nvidia-smi -L | head -8; free -g | head -2 | tail -1
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
for MEGA in ${PARITY_PATHS-1 0}; do
  YALM_TP_MEGA=$MEGA OMP_NUM_THREADS=4 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700+MEGA)) tests/tp_worker.py > gpurun_out/tp${N}_parity_mega${MEGA}.log 2>&1
  echo "tp$N parity mega=$MEGA rc=$?"; grep -E "PASS|FAIL|skipped|all-reduce over" gpurun_out/tp${N}_parity_mega${MEGA}.log | grep -E "rank 0|FAIL|skipped|all-reduce" | head -12
done
P=29720
for W in $WL; do
  P=$((P+1))
  S=$(echo $W | sed 's/-fp16-4k//; s/-8x7b//; s/-3-70b/70b/; s/mistral-7b/mistral/')
  timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --workload $W --steps 64 --warmup 8 > gpurun_out/r2_bench_${S}_tp${N}.json 2> gpurun_out/r2_bench_${S}_tp${N}.err
  echo "$W tp$N rc=$?"; tail -c 900 gpurun_out/r2_bench_${S}_tp${N}.json | head -c 900; echo; grep -iE "error|Traceback|PARITY|timed out" gpurun_out/r2_bench_${S}_tp${N}.err | head -5
done
if [ -n "$EP" ]; then   # expert parallelism as the measured alternative (per-op path)
  P=$((P+1))
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --workload mixtral-8x7b-fp16-4k --moe-ep --steps 64 --warmup 8 > gpurun_out/r2_bench_mixtral_ep${N}.json 2> gpurun_out/r2_bench_mixtral_ep${N}.err
  echo "mixtral ep$N rc=$?"; tail -c 600 gpurun_out/r2_bench_mixtral_ep${N}.json; echo; grep -iE "error|Traceback|PARITY|timed out" gpurun_out/r2_bench_mixtral_ep${N}.err | head -5
fi
if [ -f build/lib_tp_barrier.so ]; then
  P=$((P+1))
  YALM_B200_LIB=build/lib_tp_barrier.so timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --steps 64 --warmup 8 --no-parity > gpurun_out/r2_bench_mistral_tp${N}_barrier_variant.json 2> gpurun_out/r2_bench_mistral_tp${N}_barrier_variant.err
  echo "barrier variant tp$N rc=$?"; python -c "import json,sys; d=json.load(open('gpurun_out/r2_bench_mistral_tp${N}_barrier_variant.json')); print('barrier variant', d['value'], d['e2e']['value'])"
fi

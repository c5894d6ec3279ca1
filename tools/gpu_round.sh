#!/usr/bin/env bash
# One GPU-box call: parity sweep, bench, ncu launch list + full capture of the step kernel.
# usage (here): gpurun --timeout 1500 -- 'bash tools/gpu_round.sh [tag]'
set -u
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > $OUT/clocks_$TAG.csv &
SMI=$!
python tools/gpu_check.py 2048 > $OUT/parity_$TAG.log 2>&1
echo "parity rc=$?"; tail -12 $OUT/parity_$TAG.log
python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?"; tail -5 $OUT/pytest_gpu_$TAG.log
python bench.py > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; cat $OUT/bench_$TAG.json; tail -3 $OUT/bench_$TAG.err
python bench.py --envs 4096 --no-cpu-baseline > $OUT/bench4096_$TAG.json 2>> $OUT/bench_$TAG.err
cat $OUT/bench4096_$TAG.json
kill $SMI
CMD="python bench.py --steps 8 --warmup 3 --no-cpu-baseline"
$CMD > $OUT/plain_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launch_$TAG.log 2>&1
echo "ncu launches rc=$?"
$CMD > $OUT/plain2_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lgk_step_kernel -s 12 -c 2 -o $OUT/prof_step_$TAG $CMD > $OUT/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"

# ANYmal LSTM actuator kernel: timing + one full capture; the reference's op-by-op path on CUDA torch
python tools/bench_lstm.py 4096 16384 65536 > $OUT/bench_lstm_$TAG.json 2> $OUT/bench_lstm_$TAG.err && \
ncu --set full --clock-control none --import-source on -k regex:lgk_lstm_kernel -s 6 -c 1 -o $OUT/prof_lstm_$TAG python tools/bench_lstm.py 65536 > $OUT/ncu_lstm_$TAG.log 2>&1
echo "ncu lstm rc=$?"
python tests/bench_torch_cuda_reference.py 4096 65536 > $OUT/torch_cuda_reference_$TAG.json 2> $OUT/torch_cuda_reference_$TAG.err
echo "torch cuda reference rc=$?"; cat $OUT/torch_cuda_reference_$TAG.json
# Go1 actuator MLP: the three kernels, config 3 as a whole step, one full capture of the tcgen05 kernel
for k in umma mma ffma; do for n in 4096 16384 65536; do LGK_MLP_KERNEL=$k python tools/bench_actuator.py $n | tail -3; done; done > $OUT/bench_actuator_$TAG.txt 2>&1
python bench.py --scenario go1_rough_actuator --envs 16384 --no-cpu-baseline 2>/dev/null | tail -1 > $OUT/bench_config3_$TAG.json
ncu --set full --clock-control none --import-source on -k regex:lgk_act_mlp_umma -s 6 -c 1 -o $OUT/prof_umma_$TAG python tools/bench_actuator.py 65536 > $OUT/ncu_umma_$TAG.log 2>&1
echo "ncu umma rc=$?"
ls -la $OUT

#!/bin/bash
# Whole-box pass (gpurun --gpus N): BASELINE configs[4] (scenes-per-GPU sweep at N GPUs) and configs[3] (data generation with rendered
# observations, NCCL gather of the demonstrations to rank 0).   gpurun --gpus 8 -- 'bash tools/gpu_n8_round.sh r02 8'
TAG=${1:-r02}; N=${2:-8}
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > $O/${TAG}_n${N}_smi.txt 2>&1
: > $O/${TAG}_sweep_c3_n${N}.jsonl
PORT=29600
for E in 1024 4096 16384 65536; do
  PORT=$((PORT+1))
  timeout 400 $TR --master-port $PORT bench.py --gpus $N --envs $E --steps 4 --no-cpu >> $O/${TAG}_sweep_c3_n${N}.jsonl 2>> $O/${TAG}_sweep_n${N}.err; echo "envs $E x $N GPUs rc=$?"
done
python - <<P
import json
for l in open("$O/${TAG}_sweep_c3_n${N}.jsonl"):
    d = json.loads(l); print(d["n_gpus"], d["config"]["envs_per_gpu"], "%.2f M env-steps/s" % (d["value"] / 1e6), "e2e %.2f" % (d["e2e"]["value"] / 1e6))
P
timeout 900 $TR --master-port 29650 -m mujoco_robot_environments_b200.data_generation --envs 1024 --max-steps 3 --render --out $O/${TAG}_demos_n${N} > $O/${TAG}_datagen_n${N}.json 2> $O/${TAG}_datagen_n${N}.err; echo "datagen rc=$?"; cat $O/${TAG}_datagen_n${N}.json; tail -3 $O/${TAG}_datagen_n${N}.err
rm -f $O/${TAG}_demos_n${N}.npz
timeout 900 $TR --master-port 29651 -m mujoco_robot_environments_b200.data_generation --envs 8192 --recycle --episodes 2 > $O/${TAG}_datagen_recycle_n${N}.json 2> $O/${TAG}_datagen_recycle_n${N}.err; echo "datagen recycle rc=$?"; cat $O/${TAG}_datagen_recycle_n${N}.json

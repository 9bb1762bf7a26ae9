#!/bin/bash
# 2-GPU pass: weak-scaling bench lines (N=1,2) + batched data generation with rendered observations and NCCL gather to rank 0.
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 python bench.py --workload render --no-cpu > $O/${TAG}_bench_render.json 2> $O/${TAG}_bench_render.err; echo "render rc=$?"; cut -c1-330 $O/${TAG}_bench_render.json
timeout 600 python bench.py --gpus 1 --steps 5 --no-cpu > $O/${TAG}_bench_c3_n1.json 2> $O/${TAG}_bench_c3_n1.err; echo "n1 rc=$?"; cut -c1-200 $O/${TAG}_bench_c3_n1.json
timeout 600 $TR --nproc-per-node 2 --master-port 29511 bench.py --gpus 2 --steps 5 > $O/${TAG}_bench_c3_n2.json 2> $O/${TAG}_bench_c3_n2.err; echo "n2 rc=$?"; cut -c1-200 $O/${TAG}_bench_c3_n2.json; tail -3 $O/${TAG}_bench_c3_n2.err
timeout 600 $TR --nproc-per-node 2 --master-port 29512 bench.py --gpus 2 --steps 5 --workload render > $O/${TAG}_bench_render_n2.json 2> $O/${TAG}_bench_render_n2.err; echo "render n2 rc=$?"; cut -c1-200 $O/${TAG}_bench_render_n2.json
timeout 900 $TR --nproc-per-node 2 --master-port 29513 -m mujoco_robot_environments_b200.data_generation --envs 512 --max-steps 3 --render --out $O/${TAG}_demos > $O/${TAG}_datagen.json 2> $O/${TAG}_datagen.err; echo "datagen rc=$?"; cat $O/${TAG}_datagen.json; tail -5 $O/${TAG}_datagen.err
rm -f $O/${TAG}_demos.npz

#!/bin/bash
# multi-GPU evidence run (under gpurun --gpus N): config 3 headline, config 4, config 5, pinned-H2D ceiling
N=${1:-8}; tag=${2:-r02}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29601 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/${tag}_bench_${N}gpu.json 2> gpurun_out/${tag}_bench_${N}gpu.err; tail -c 400 gpurun_out/${tag}_bench_${N}gpu.json; tail -3 gpurun_out/${tag}_bench_${N}gpu.err
$TR --master-port 29602 bench.py --gpus $N --workload h2d > gpurun_out/${tag}_h2d_${N}gpu.json 2>/dev/null; cat gpurun_out/${tag}_h2d_${N}gpu.json
$TR --master-port 29603 bench.py --gpus $N --workload fusion_train --no-cpu-baseline > gpurun_out/${tag}_bench_fusion_train_${N}gpu.json 2> gpurun_out/${tag}_ft.err; tail -c 300 gpurun_out/${tag}_bench_fusion_train_${N}gpu.json; tail -2 gpurun_out/${tag}_ft.err
$TR --master-port 29604 bench.py --gpus $N --workload stress --no-cpu-baseline --steps 5 > gpurun_out/${tag}_bench_stress_${N}gpu.json 2> gpurun_out/${tag}_st.err; tail -c 300 gpurun_out/${tag}_bench_stress_${N}gpu.json; tail -2 gpurun_out/${tag}_st.err
$TR --master-port 29605 bench.py --gpus $N --exchange nccl --steps 5 --warmup 3 > gpurun_out/${tag}_bench_${N}gpu_nccl.json 2>/dev/null; python -c "
import json; a=json.load(open('gpurun_out/${tag}_bench_${N}gpu.json')); b=json.load(open('gpurun_out/${tag}_bench_${N}gpu_nccl.json'))
print('peer', a['value'], a['ms_per_step'], a['multi_gpu_parity'], '| nccl', b['value'], b['ms_per_step'], b['multi_gpu_parity'])"
nvidia-smi topo -m > gpurun_out/${tag}_topo_${N}gpu.txt; grep MemTotal /proc/meminfo; nproc

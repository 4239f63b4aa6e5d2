set -x
python -m pytest tests -x -q -m gpu > gpurun_out/r01h_pytest_gpu.log 2>&1; tail -3 gpurun_out/r01h_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r01h_smoke.log 2>&1; tail -2 gpurun_out/r01h_smoke.log
python bench.py > gpurun_out/r01h_bench.json 2> gpurun_out/r01h_bench.err; tail -c 600 gpurun_out/r01h_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01h_bench_reference.json 2> gpurun_out/r01h_bench_reference.err
python bench.py --workload stress --no-cpu-baseline > gpurun_out/r01h_bench_stress.json 2>/dev/null
python bench.py --workload avm_train --no-cpu-baseline > gpurun_out/r01h_bench_avm_train.json 2>/dev/null
python bench.py --workload fusion_train --no-cpu-baseline > gpurun_out/r01h_bench_fusion_train.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01h_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r01h_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -c 24 -o gpurun_out/r01h_full -f python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/r01h_ncu_full.log 2>&1
ls -la gpurun_out/ | tail -15

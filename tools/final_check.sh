# end-of-round evidence run on ONE B200 (under gpurun): GPU suite, smoke, bench lines, ncu launch list of the bench command
tag=${1:-r02z}
python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_pytest_gpu.log 2>&1; tail -3 gpurun_out/${tag}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; tail -2 gpurun_out/${tag}_smoke.log
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -c 300 gpurun_out/${tag}_bench.json; tail -2 gpurun_out/${tag}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err; tail -c 300 gpurun_out/${tag}_bench_reference.json
python bench.py --workload stress --no-cpu-baseline > gpurun_out/${tag}_bench_stress.json 2>/dev/null
python bench.py --workload avm_train --no-cpu-baseline > gpurun_out/${tag}_bench_avm_train.json 2>/dev/null
python bench.py --workload fusion_train --no-cpu-baseline > gpurun_out/${tag}_bench_fusion_train.json 2>/dev/null
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-share > gpurun_out/${tag}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 200 --csv --log-file gpurun_out/${tag}_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-share > gpurun_out/${tag}_ncu_launch.log 2>&1
tail -2 gpurun_out/${tag}_ncu_launch.log
ls gpurun_out/ | grep ${tag}

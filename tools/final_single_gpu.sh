set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > gpurun_out/final_gpu_tests.txt
python bench.py > gpurun_out/r02_bench_c3_1gpu.json 2> gpurun_out/bench_c3.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_c3_reference.json 2>/dev/null
for c in C1 C2 C4; do python bench.py --config $c --no-strong > gpurun_out/r02_bench_${c}_1gpu.json 2>/dev/null; done
timeout 600 ncu --set full --clock-control none --profile-from-start off -c 8 -o /tmp/step -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-strong > gpurun_out/ncu_step.log 2>&1
ncu -i /tmp/step.ncu-rep --page raw --csv > gpurun_out/r02_ncu_full_step.csv 2>/dev/null
mkdir -p profiles
python tools/make_traffic_json.py gpurun_out/r02_ncu_full_step.csv C3 512 && cp profiles/r02_traffic.json gpurun_out/r02_traffic.json
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 8 --csv --log-file gpurun_out/r02_launches_c3_step.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-strong > /dev/null 2>&1
python bench.py --no-cpu-baseline --no-strong --steps 20 > gpurun_out/b_after.json 2>/dev/null; python tools/_pj.py gpurun_out/b_after.json
cat gpurun_out/final_gpu_tests.txt
python tools/_pj.py gpurun_out/r02_bench_c3_1gpu.json
# C4 step: full counter set for the per-kernel table
timeout 600 ncu --set full --clock-control none --profile-from-start off -c 10 -o /tmp/step_C4 -f python bench.py --config C4 --steps 1 --warmup 3 --no-cpu-baseline --no-strong > gpurun_out/ncu_C4.log 2>&1
ncu -i /tmp/step_C4.ncu-rep --page raw --csv > gpurun_out/r02_ncu_full_c4_step.csv 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 10 --csv --log-file gpurun_out/r02_launches_C4_step.csv python bench.py --config C4 --steps 2 --warmup 3 --no-cpu-baseline --no-strong > /dev/null 2>&1
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')"

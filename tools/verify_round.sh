# Round-end verification on a GPU box (run from the repo root, e.g. `gpurun --timeout 1800 -- 'bash tools/verify_round.sh'`).
# Everything lands under gpurun_out/verify_*; nothing here reads /root/reference.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/verify_gpu_tests.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/verify_smoke.log 2>&1
python bench.py > gpurun_out/verify_bench.json 2> gpurun_out/verify_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/verify_bench_reference.json 2> gpurun_out/verify_bench_reference.err
python tools/optimize_c3.py 50 --fit > gpurun_out/verify_optimize_fit.json 2> gpurun_out/verify_optimize_fit.err
python tools/bench_constrained.py > gpurun_out/verify_constrained.json 2> gpurun_out/verify_constrained.err
python tools/saas_bench.py > gpurun_out/verify_saas.json 2> gpurun_out/verify_saas.err

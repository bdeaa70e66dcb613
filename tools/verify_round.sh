set -x
python tools/optimize_c3.py 50 --fit > gpurun_out/s38_optimize_fit.json 2> gpurun_out/s38_optimize_fit.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s38_smoke.log 2>&1
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/s38_bench2.json 2> gpurun_out/s38_bench2.err

set -x
N=${1:-8}
nvidia-smi topo -m > gpurun_out/r02_topo_8gpu.txt 2>&1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 tools/scaling.py --what strong,c5,shapes --out gpurun_out/r02_scaling_8gpu.json > gpurun_out/scaling8.log 2>&1
echo rc=$?
tail -c 600 gpurun_out/scaling8.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02_bench_c3_8gpu.json 2> gpurun_out/bench8.err
echo rc=$?
python tools/_pj.py gpurun_out/r02_bench_c3_8gpu.json

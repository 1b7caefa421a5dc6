# 8 GPUs: the driver's SCALE command at N=8 (strong object: cfg4 p2p strips + shared host frame, cfg5)
set -x
N=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29601 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2_s49_bench_n$N.json 2> gpurun_out/r2_s49_bench_n$N.err; echo "rc=$?"
tail -c 1200 gpurun_out/r2_s49_bench_n$N.err
cut -c1-400 gpurun_out/r2_s49_bench_n$N.json

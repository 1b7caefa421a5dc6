# 2 GPUs: the multi-rank bench paths (strong-scaling object, shared host frame, p2p strips)
set -x
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29591 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2_s43_bench_n$N.json 2> gpurun_out/r2_s43_bench_n$N.err; echo "rc=$?"
tail -c 1500 gpurun_out/r2_s43_bench_n$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29592 bench.py --gpus $N --steps 5 --warmup 3 --impl reference > gpurun_out/r2_s43_bench_ref_n$N.json 2> gpurun_out/r2_s43_bench_ref_n$N.err; echo "rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29593 bench.py --gpus $N --steps 5 --warmup 3 --workload cfg4 --gather nccl > gpurun_out/r2_s43_bench_cfg4_nccl_n$N.json 2> gpurun_out/r2_s43_bench_cfg4_nccl_n$N.err; echo "rc=$?"
tail -c 600 gpurun_out/r2_s43_bench_cfg4_nccl_n$N.err
cut -c1-600 gpurun_out/r2_s43_bench_n$N.json

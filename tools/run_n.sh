N=$1
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r01m_bench_c3_n$N.json 2> gpurun_out/r01m_bench_c3_n$N.err; echo "n$N rc=$?"; cut -c1-250 gpurun_out/r01m_bench_c3_n$N.json

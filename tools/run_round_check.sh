# Round check on one GPU box: GPU tests, smoke, both bench arms, the reference's CUDA path beside ours, ncu launch list
# and full captures. TAG names the output files under gpurun_out/ (copy what is to be judged into profiles/).
#   gpurun --timeout 2400 -- 'TAG=r02m bash tools/run_round_check.sh'
TAG=${TAG:-r02x}
O=gpurun_out
set -x
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
timeout 1200 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 600 python bench.py > $O/${TAG}_bench_c3.json 2> $O/${TAG}_bench_c3.err; echo "bench rc=$?"; cut -c1-400 $O/${TAG}_bench_c3.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err; echo "ref rc=$?"; cut -c1-300 $O/${TAG}_bench_ref.json
timeout 200 python tools/ref_gpu_run.py c1 > $O/${TAG}_ref_gpu_c1.json 2> $O/${TAG}_ref_gpu_c1.err; echo "refgpu c1 rc=$?"; cut -c1-300 $O/${TAG}_ref_gpu_c1.json
timeout 300 python tools/ref_gpu_run.py c3 > $O/${TAG}_ref_gpu_c3.json 2> $O/${TAG}_ref_gpu_c3.err; echo "refgpu c3 rc=$?"; cut -c1-300 $O/${TAG}_ref_gpu_c3.json
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $O/${TAG}_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras > $O/${TAG}_ncu_list.log 2>&1; echo "ncu list rc=$?"
QP_SPP=4 timeout 120 python tools/ncu_target.py 2>&1 | tail -3
QP_SPP=4 timeout 500 ncu --set full --clock-control none --import-source on -k regex:wf_ -s 8 -c 6 -o $O/prof_wf_${TAG} -f python tools/ncu_target.py > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
QP_SPP=4 timeout 500 ncu --set full --clock-control none --import-source on -k regex:trace_batch -c 3 -o $O/prof_batch_${TAG} -f python tools/ncu_target.py > $O/${TAG}_ncu_batch.log 2>&1; echo "ncu batch rc=$?"
ls -la $O

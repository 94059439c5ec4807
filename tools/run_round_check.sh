set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01m_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r01m_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 400 python bench.py > gpurun_out/r01m_bench_c3.json 2> gpurun_out/r01m_bench_c3.err; echo "bench rc=$?"; cat gpurun_out/r01m_bench_c3.json | cut -c1-400
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01m_bench_ref.json 2> gpurun_out/r01m_bench_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/r01m_bench_ref.json
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r01m_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras > gpurun_out/r01m_ncu_list.log 2>&1; echo "ncu list rc=$?"
QP_SPP=4 timeout 120 python tools/ncu_target.py 2>&1 | tail -3
QP_SPP=4 timeout 500 ncu --set full --clock-control none --import-source on -k regex:wf_ -s 8 -c 6 -o gpurun_out/prof_wf_r01m -f python tools/ncu_target.py > gpurun_out/r01m_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
QP_SPP=16 timeout 500 ncu --set full --clock-control none --import-source on -k regex:wf_finish -c 2 -o gpurun_out/prof_finish_r01m -f python tools/ncu_target.py > gpurun_out/r01m_ncu_finish.log 2>&1; echo "ncu finish rc=$?"
for w in c1 c5 c4 c2; do timeout 600 python bench.py --workload $w --no-cpu-baseline > gpurun_out/r01m_bench_$w.json 2> gpurun_out/r01m_bench_$w.err; echo "$w rc=$?"; cut -c1-200 gpurun_out/r01m_bench_$w.json; done

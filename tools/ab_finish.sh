# A/B of the tail-finish threshold (AKB_FINISH_DIV=0: wavefront to the end) on one GPU
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for w in c3 c1 c5; do
  for fd in 0 2 4 8 16 64; do
    AKB_FINISH_DIV=$fd timeout 300 python bench.py --workload $w --no-cpu-baseline --no-extras > gpurun_out/abf_${w}_${fd}.json 2> gpurun_out/abf_${w}_${fd}.err
    python -c "import json,sys; d=json.load(open('gpurun_out/abf_${w}_${fd}.json')); print('$w finish_div=$fd', round(d['value'],1), round(d['ms_per_step'],3), round(d['e2e']['value'],1), d['gpu_launches'], d['rays'])"
  done
done
QP_SPP=64 AKB_TIMELINE=1 timeout 300 python tools/ncu_target.py > gpurun_out/timeline_c3_finish.log 2>&1
tail -2 gpurun_out/timeline_c3_finish.log

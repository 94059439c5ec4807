mkdir -p gpurun_out
for w in c3 c1 c5; do
  for fl in 32 16 8 4; do
    AKB_FINISH_LANES=$fl timeout 300 python bench.py --workload $w --no-cpu-baseline --no-extras > gpurun_out/abl_${w}_${fl}.json 2> gpurun_out/abl_${w}_${fl}.err
    python -c "import json,sys; d=json.load(open('gpurun_out/abl_${w}_${fl}.json')); print('$w finish_lanes=$fl', round(d['value'],1), round(d['ms_per_step'],3))"
  done
done
AKB_FINISH_LANES=8 timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3

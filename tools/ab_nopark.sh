mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
AKB_NO_PARK=0 timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run() { # label, workload
  timeout 300 python bench.py --workload $2 --no-cpu-baseline --no-extras > gpurun_out/abp_$2_$1.json 2> gpurun_out/abp.err
  python -c "import json,sys; d=json.load(open('gpurun_out/abp_$2_$1.json')); print('$2 $1', round(d['value'],1), round(d['ms_per_step'],3), d['gpu_launches'], d['rays'])"
}
for w in c3 c1 c5; do
  AKB_NO_PARK=0 run park $w
  AKB_NO_PARK=1 run nopark $w
done

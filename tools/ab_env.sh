mkdir -p gpurun_out
run() { # label, workload
  timeout 300 python bench.py --workload $2 --no-cpu-baseline --no-extras > gpurun_out/abe.json 2> gpurun_out/abe.err
  python -c "import json,sys; d=json.load(open('gpurun_out/abe.json')); print('$2 $1', round(d['value'],1), round(d['ms_per_step'],3))"
}
for w in c3 c1; do
  run base $w
  AKB_CHECK_EVERY=2 run check2 $w
  AKB_CHECK_EVERY=8 run check8 $w
  AKB_FINISH_MAX=32768 run max32k $w
  AKB_FINISH_MAX=98304 AKB_FINISH_DIV=6 run max96k_div6 $w
  AKB_FINISH_DIV=4 run div4_max64k $w
done

# env-only sweeps of the scheduling switches on one GPU (development aid)
mkdir -p gpurun_out
run() { # label, workload
  timeout 300 python bench.py --workload $2 --no-cpu-baseline --no-extras > gpurun_out/abe.json 2> gpurun_out/abe.err
  python -c "import json,sys; d=json.load(open('gpurun_out/abe.json')); print('$2 $1', round(d['value'],1), round(d['ms_per_step'],3))"
}
for w in c1; do
  run base $w
  AKB_PIPELINES=1 run pipes1 $w
  AKB_CHECK_EVERY=2 run check2 $w
  AKB_CHECK_EVERY=8 run check8 $w
  AKB_FINISH_DIV=2 run div2 $w
  AKB_FINISH_DIV=8 run div8 $w
  AKB_REFILL_BELOW=20 run refill20 $w
  run base_again $w
done

# A/B of the small-queue fetch rule (AKB_FETCH_CHUNK=-1 = the previous rule) on one GPU
set -x
mkdir -p gpurun_out
for w in c3 c1 c5; do
  for fc in -1 0; do
    AKB_FETCH_CHUNK=$fc timeout 300 python bench.py --workload $w --no-cpu-baseline --no-extras > gpurun_out/ab_${w}_fc${fc}.json 2> gpurun_out/ab_${w}_fc${fc}.err
    python -c "import json,sys; d=json.load(open('gpurun_out/ab_${w}_fc${fc}.json')); print('$w fc=$fc', d['value'], d['ms_per_step'], d['e2e']['value'])"
  done
done
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3

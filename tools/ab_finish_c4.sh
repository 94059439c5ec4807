mkdir -p gpurun_out
for cfg in "0 131072" "8 131072" "8 1000000" "8 32768"; do
  set -- $cfg
  AKB_FINISH_DIV=$1 AKB_FINISH_MAX=$2 timeout 600 python bench.py --workload c4 --spp 64 --steps 3 --no-cpu-baseline --no-extras > gpurun_out/abf_c4_$1_$2.json 2> gpurun_out/abf_c4_$1_$2.err
  python -c "import json,sys; d=json.load(open('gpurun_out/abf_c4_$1_$2.json')); print('c4 64spp finish_div=$1 max=$2', round(d['value'],1), round(d['ms_per_step'],3), d['gpu_launches'])"
done

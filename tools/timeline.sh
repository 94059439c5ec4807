mkdir -p gpurun_out
QP_SPP=64 AKB_TIMELINE=1 timeout 300 python tools/ncu_target.py > gpurun_out/timeline_c3.log 2>&1
tail -3 gpurun_out/timeline_c3.log

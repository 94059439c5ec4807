# ncu --set full captures of the kernels beside the wavefront pair (judge's N2 row): the megakernel, the AO kernel, the
# device BVH builder's kernels, the streaming kernel. Each command runs once without ncu first.
#   gpurun --timeout 1800 -- 'TAG=r02p bash tools/ncu_other_kernels.sh'
TAG=${TAG:-r02x}
O=gpurun_out
mkdir -p $O
NCU="ncu --set full --clock-control none --import-source on -f"
QP_MODE=mega QP_SPP=4 timeout 200 python tools/ncu_target.py 2>&1 | tail -2
QP_MODE=mega QP_SPP=4 timeout 600 $NCU -k regex:megakernel -c 1 -o $O/prof_mega_$TAG python tools/ncu_target.py > $O/${TAG}_ncu_mega.log 2>&1; echo "mega rc=$?"
QP_SPP=4 timeout 200 python tools/ao_perf.py 2>&1 | tail -2
QP_SPP=4 timeout 600 $NCU -k regex:ao_kernel -c 1 -o $O/prof_ao_$TAG python tools/ao_perf.py > $O/${TAG}_ncu_ao.log 2>&1; echo "ao rc=$?"
AKB_STREAM=1 QP_SPP=4 timeout 200 python tools/ncu_target.py 2>&1 | tail -2
AKB_STREAM=1 QP_SPP=4 timeout 600 $NCU -k regex:wf_stream -c 1 -o $O/prof_stream_$TAG python tools/ncu_target.py > $O/${TAG}_ncu_stream.log 2>&1; echo "stream rc=$?"
QP_SPP=1 timeout 300 python tools/bvh_build_perf.py c3 2>&1 | tail -3
QP_SPP=1 timeout 900 $NCU -k regex:"ploc_|fit_kernel|level_|morton_kernel|centroid_bounds|leaf_box|hierarchy" -c 14 -o $O/prof_bvh_$TAG python tools/bvh_build_perf.py c3 > $O/${TAG}_ncu_bvh.log 2>&1; echo "bvh rc=$?"
ls -la $O | grep prof_

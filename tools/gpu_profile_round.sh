#!/bin/bash
# GPU box only: ncu full-set capture of EVERY kernel of the hot path (tools/profile_all.py), reduced on the box to small
# text files (the .ncu-rep of ~100 MB stays in /tmp; gpurun copies back at most 64 MiB).
#   usage: bash tools/gpu_profile_round.sh <tag>      -> gpurun_out/<tag>_all_kernels.md, <tag>_hot_<kernel>.txt,
#                                                        <tag>_sass_<kernel>.txt, <tag>_launches.csv
set -u
tag=${1:-r2}
out=gpurun_out
mkdir -p $out
rep=/tmp/${tag}_all.ncu-rep
python tools/profile_all.py > $out/${tag}_profall.log 2>&1 || { echo "profile_all failed"; tail -20 $out/${tag}_profall.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -f -o ${rep%.ncu-rep} \
    python tools/profile_all.py > $out/${tag}_ncu.log 2>&1
ls -la $rep
ALL='k_level0 k_resize_rows_tma k_fast_band k_octree k_blur_tma k_describe k_bruteforce$ k_bruteforce_tc k_expand_pm1 k_merge_partials k_merge_exchange k_bow_scatter12 k_proj_search k_stereo_rows k_stereo_refine k_stereo_medians k_stereo_reject k_bow_transform k_bow_nodes k_bow_rotation k_search_best k_init_search k_triang_nodes k_distinctive k_undistort k_assign_grid'
python tools/ncu_report.py $rep --sass $ALL > $out/${tag}_all_kernels.md 2> $out/${tag}_report.err
for k in k_fast_band k_resize_rows_tma k_describe k_blur_tma k_octree k_bruteforce_tc; do
    python tools/ncu_hot.py $rep $k 60 > $out/${tag}_hot_$k.txt 2>&1
done
python tools/ncu_sass.py $rep k_fast_band > $out/${tag}_sass_k_fast_band.txt 2>&1
python tools/ncu_sass.py $rep k_resize_rows > $out/${tag}_sass_k_resize_rows.txt 2>&1
python tools/ncu_sass.py $rep k_describe > $out/${tag}_sass_k_describe.txt 2>&1
# launch list of one bench run (not replayed; cold-cache, serialised)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 1 --warmup 3 --inner 2 --e2e-inner 1 --no-cpu-baseline --no-extra > $out/${tag}_launches.log 2>&1
du -sh $out

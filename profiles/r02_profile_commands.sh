# Round-2 evidence, run as one gpurun call on one B200 (after the plain run exits 0):
#   bash profiles/r02_profile_commands.sh
set -x
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-wisdm --no-config5 --no-config4"
$CMD > gpurun_out/r02_prof_plain.json 2> gpurun_out/r02_prof_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r02_launches.csv $CMD > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_window_features_v4|k_seg_tiles|k_seg_scan|k_seg_emit|k_predict_walk" -s 15 -c 5 -o gpurun_out/r02_full_window_predict $CMD > gpurun_out/r02_full_a.log 2>&1
# fit kernels: the root level (-s 0 of the timed steps' first set) and a deep level of the same step
ncu --set full --clock-control none --import-source on -k regex:"k_node_hist_lanes|k_best_split|k_part_count|k_part_scatter" -s 220 -c 12 -o gpurun_out/r02_full_fit $CMD > gpurun_out/r02_full_b.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_node_hist_lanes" -s 60 -c 2 -o gpurun_out/r02_full_hist_root $CMD > gpurun_out/r02_full_c.log 2>&1
tail -n 2 gpurun_out/r02_full_a.log gpurun_out/r02_full_b.log gpurun_out/r02_full_c.log

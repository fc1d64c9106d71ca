#!/bin/bash
# Round-2 profiler captures, run on the GPU box AFTER the same commands have exited 0 without ncu.  The .ncu-rep files are
# summarised ON the box (scripts/ncu_summary.py: raw page -> the CSVs committed under profiles/; scripts/sass_lines.py:
# instructions per source line) and deleted: gpurun_out/ only carries 64 MiB back.
set -x
mkdir -p gpurun_out
O=gpurun_out
B="python bench.py --no-cpu-baseline --no-also --frames-in-flight 1"
# 1. launch lists (durations only; per-launch times are cold-cache and serialised: shares, not absolutes)
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/r02_launches_ball_1080p_256spp.csv $B --steps 2 --warmup 1 > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/r02_launches_ball_1080p_32spp.csv $B --spp 32 --steps 2 --warmup 1 > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 60 --csv --log-file $O/r02_launches_soup_build.csv python bench.py --workload soup --steps 1 --warmup 1 > /dev/null 2>&1
# 2. full captures of the two render kernels of a ball frame (first launch of each after the warm-up frame)
ncu --set full --clock-control none --import-source on -k regex:k_primary_coop --launch-skip 1 -c 1 -f -o /tmp/r02_k_primary_ball $B --steps 1 --warmup 1 > /dev/null 2>&1
python scripts/ncu_summary.py /tmp/r02_k_primary_ball.ncu-rep $O/r02_ncu_full_k_primary_ball.csv
python scripts/sass_lines.py /tmp/r02_k_primary_ball.ncu-rep k_primary_coop tinypathtracer_b200/lib/libtpt_b200.so 30 > $O/r02_hot_lines_k_primary_ball.txt 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wave --launch-skip 8 -c 1 -f -o /tmp/r02_k_wave_ball $B --steps 1 --warmup 1 > /dev/null 2>&1
python scripts/ncu_summary.py /tmp/r02_k_wave_ball.ncu-rep $O/r02_ncu_full_k_wave_ball.csv
python scripts/sass_lines.py /tmp/r02_k_wave_ball.ncu-rep k_wave tinypathtracer_b200/lib/libtpt_b200.so 30 > $O/r02_hot_lines_k_wave_ball.txt 2>&1
# 3. the 10 M-triangle soup: the trace launches after the warm-up (primary, secondary) and one whole build
ncu --set full --clock-control none -k regex:k_trace_rays_persistent --launch-skip 2 -c 2 -f -o /tmp/r02_k_trace_rays_soup python bench.py --workload soup --steps 1 --warmup 1 > /dev/null 2>&1
python scripts/ncu_summary.py /tmp/r02_k_trace_rays_soup.ncu-rep $O/r02_ncu_full_k_trace_rays_soup.csv
ncu --set full --clock-control none -k 'regex:k_transform|k_morton_hist|k_onesweep|k_leaves_build|k_build_upper' --launch-skip 8 -c 8 -f -o /tmp/r02_bvh_build_soup python bench.py --workload soup --steps 1 --warmup 1 > /dev/null 2>&1
python scripts/ncu_summary.py /tmp/r02_bvh_build_soup.ncu-rep $O/r02_ncu_full_bvh_build_soup.csv
ls -la $O/r02_*

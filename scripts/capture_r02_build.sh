set -x
mkdir -p gpurun_out
O=gpurun_out
B="python bench.py --no-cpu-baseline --no-also --frames-in-flight 1"
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/r02_launches_ball_1080p_256spp.csv $B --steps 2 --warmup 1 > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/r02_launches_ball_1080p_32spp.csv $B --spp 32 --steps 2 --warmup 1 > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 60 --csv --log-file $O/r02_launches_soup_build.csv python bench.py --workload soup --steps 1 --warmup 1 > /dev/null 2>&1
ncu --set full --clock-control none -k 'regex:k_transform|k_morton_hist|k_onesweep|k_leaves_build|k_build_upper' --launch-skip 8 -c 8 -f -o /tmp/r02_bvh_build_soup python bench.py --workload soup --steps 1 --warmup 1 > /dev/null 2>&1
python scripts/ncu_summary.py /tmp/r02_bvh_build_soup.ncu-rep $O/r02_ncu_full_bvh_build_soup.csv
ls -la $O/r02_*

set -x
python bench.py > gpurun_out/r01e_bench.json 2> gpurun_out/r01e_bench.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01e_bench_reference.json 2> gpurun_out/r01e_bench_reference.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r01e_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r01e_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:paf_parse_kernel\|walk_mc_kernel\|rank_small_kernel -s 6 -c 5 -o gpurun_out/r01e_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r01e_ncu2.log 2>&1
tail -2 gpurun_out/r01e_ncu2.log

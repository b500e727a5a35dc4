# One round of measurements on the GPU box: TAG=r01f bash tools/profile_round.sh
TAG=${TAG:-r01f}
set -x
mkdir -p gpurun_out
(free -g | head -2; df -h /tmp | tail -1; nproc; nvidia-smi --query-gpu=name,memory.total --format=csv,noheader) > gpurun_out/${TAG}_host.txt 2>&1
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:paf_parse_kernel\|walk_mc_kernel\|rank_small_kernel -s 6 -c 5 -o gpurun_out/${TAG}_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu2.log 2>&1
tail -2 gpurun_out/${TAG}_ncu2.log

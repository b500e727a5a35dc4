# One round of ncu evidence on the GPU box (one gpurun call, ncu only):  TAG=r02w bash tools/profile_round.sh
# Every ncu run follows a plain run of the same command that exited 0 (B200_PROFILING.md).
TAG=${TAG:-r02w}
set -x
mkdir -p gpurun_out
(free -g | head -2; df -h /tmp | tail -1; nproc; nvidia-smi --query-gpu=name,memory.total --format=csv,noheader) > gpurun_out/${TAG}_host.txt 2>&1
# 1. launch list of the bench command (configs[3]): every launch with its device time (cold-cache, serialised: compare SHARES)
BENCH="python bench.py --steps 1 --warmup 3 --no-secondary --no-cpu-baseline"
$BENCH > gpurun_out/${TAG}_bench_plain.json 2> gpurun_out/${TAG}_bench_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${TAG}_launches.csv $BENCH > gpurun_out/${TAG}_ncu1.log 2>&1
# 2. --set full of EVERY kernel of one step at configs[1] size (small: 40 replays per launch); the report is turned into
#    the raw csv here because it exceeds what gpurun copies back
export HERA_BENCH_WORKLOAD=ecoli
python tools/prof_step.py 2 > gpurun_out/${TAG}_step_ecoli_plain.log 2>&1 &&
ncu --set full --clock-control none -c 400 -o /tmp/${TAG}_full_ecoli -f python tools/prof_step.py 2 > gpurun_out/${TAG}_ncu2.log 2>&1
ncu -i /tmp/${TAG}_full_ecoli.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_ecoli_raw.csv 2>/dev/null
# 3. the parse kernel at configs[3] size with source correlation (roofline.traffic comes from this capture)
export HERA_BENCH_WORKLOAD=dmel
python tools/prof_step.py 1 > gpurun_out/${TAG}_step_dmel_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:paf_parse_kernel -c 1 -o gpurun_out/${TAG}_parse_dmel -f python tools/prof_step.py 1 > gpurun_out/${TAG}_ncu3.log 2>&1
tail -2 gpurun_out/${TAG}_ncu1.log gpurun_out/${TAG}_ncu2.log gpurun_out/${TAG}_ncu3.log
ls -la gpurun_out

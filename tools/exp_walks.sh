set -x
python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -x -q 2>&1 | tail -3
for v in base e1; do
  if [ $v = base ]; then export HERA_B200_LIB=$PWD/telomeri_b200/build/variants/walks_base.so; else unset HERA_B200_LIB; fi
  echo "== $v"; python tools/prof_walks.py 64 2>&1 | grep warps/SM; python tools/prof_step.py 2>&1 | tail -3
done

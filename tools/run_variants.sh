for v in telomeri_b200/build/variants/*.so; do echo "== $v"; HERA_B200_LIB=$PWD/$v python tools/prof_parse.py 4 2>&1 | tail -2; done

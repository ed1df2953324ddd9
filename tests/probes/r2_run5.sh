for v in a; do echo $v; TDOD_LIB=build_exp/$v/libtdod.so python tests/probes/r2_corun.py 2>&1 | tail -3;  TDOD_LIB=build_exp/$v/libtdod.so python tests/probes/r2_step.py 2>&1 | tail -1; done

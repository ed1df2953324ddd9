python tests/probes/r2_probe3_kb.py > gpurun_out/r2p_new.txt 2>&1; tail -4 gpurun_out/r2p_new.txt
for v in v33 v43; do echo $v; TDOD_LIB=build_exp/$v/libtdod.so python tests/probes/r2_probe3_kb.py 2>&1 | tail -3; done
for v in prof prof33; do echo $v; TDOD_LIB=build_exp/$v/libtdod.so python tests/probes/r2_kb_prof.py 2>&1 | grep -v " 0.0 us"; done

python tests/probes/r2_probe3_kb.py > gpurun_out/r2p_new.txt 2>&1; tail -4 gpurun_out/r2p_new.txt
TDOD_LIB=build_exp/old/libtdod.so python tests/probes/r2_probe3_kb.py > gpurun_out/r2p_old.txt 2>&1; tail -1 gpurun_out/r2p_old.txt
TDOD_LIB=build_exp/prof/libtdod.so python tests/probes/r2_kb_prof.py > gpurun_out/r2p_prof.txt 2>&1; cat gpurun_out/r2p_prof.txt | grep -v " 0.0 us"
python tests/probes/kb_sweep.py 600 > gpurun_out/r2d_kb_sweep.txt 2>&1; echo rc=$?
tail -6 gpurun_out/r2d_kb_sweep.txt

set -x
python tests/probes/r2_probe3_kb.py > gpurun_out/r2d_kb.txt 2>&1; echo rc=$?
tail -12 gpurun_out/r2d_kb.txt
python tests/probes/kb_sweep.py 600 > gpurun_out/r2d_kb_sweep.txt 2>&1; echo rc=$?
tail -6 gpurun_out/r2d_kb_sweep.txt
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mask_rect or pipeline or selection" > gpurun_out/r2d_pytest.log 2>&1; echo rc=$?; tail -3 gpurun_out/r2d_pytest.log

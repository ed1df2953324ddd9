# the round-1 stress sweeps once more with the end-of-round-2 kernels (outputs -> gpurun_out/, copied to profiles/r02_*)
( time python tests/probes/select_sweep.py 120 ) > gpurun_out/r02_select_sweep.txt 2>&1; tail -3 gpurun_out/r02_select_sweep.txt
( time python tests/probes/plane_sweep.py 60 ) > gpurun_out/r02_plane_sweep.txt 2>&1; tail -3 gpurun_out/r02_plane_sweep.txt
( time python tests/probes/variants_sweep.py 15 ) > gpurun_out/r02_variants_sweep.txt 2>&1; tail -3 gpurun_out/r02_variants_sweep.txt
( time python tests/probes/modeb_sweep.py ) > gpurun_out/r02_modeb_sweep.txt 2>&1; tail -2 gpurun_out/r02_modeb_sweep.txt

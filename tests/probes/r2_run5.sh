for o in 1 3; do echo over $o; TDOD_CF_OVER=$o python tests/probes/r2_step.py 2>&1 | tail -4; done

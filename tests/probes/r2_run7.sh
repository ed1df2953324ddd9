for c in 1 2 4 8; do echo cluster $c; TDOD_PSS_CLUSTER=$c python tests/probes/r2_c1.py C1 2>&1 | tail -2; done

for v in main mb3; do
  if [ $v = main ]; then unset TDOD_LIB; else export TDOD_LIB=build_exp/$v/libtdod.so; fi
  echo $v; python tests/probes/r2_modeb_phases.py 64 2>&1 | grep -E "select mode_b"; python tests/probes/r2_modeb_phases.py 8 2>&1 | grep -E "select mode_b"
done

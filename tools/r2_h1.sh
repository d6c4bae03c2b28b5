# the documented A/B switches still give a green suite
for e in "FINDB200_PRESORT=0" "FINDB200_DW_SPLIT=0" "FINDB200_CHAIN=1" "FINDB200_PDL=0" "FINDB200_KNN_MODE=2" "FINDB200_SORT_CTA_MAX=32768"; do
  env $e timeout 600 python -m pytest tests/test_field_gpu.py tests/test_step_gpu.py tests/test_parity_full_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/h1_$e.log 2>&1; echo "$e rc=$?"; tail -1 gpurun_out/h1_$e.log
done

mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout=120 > gpurun_out/h_noalloc_alone.txt 2>&1; echo "alone rc=$?" >> gpurun_out/h_noalloc_alone.txt
tail -15 gpurun_out/h_noalloc_alone.txt
timeout 500 python -m pytest tests -m gpu -q -x --timeout=120 > gpurun_out/h_tests_all.txt 2>&1; echo "all rc=$?" >> gpurun_out/h_tests_all.txt
tail -6 gpurun_out/h_tests_all.txt

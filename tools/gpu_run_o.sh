set -u
T=${TAG:-r2o}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu --timeout 300 > gpurun_out/${T}_pytest_noalloc_alone.log 2>&1; echo "alone rc=$?"
tail -5 gpurun_out/${T}_pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/${T}_pytest_gpu_all.log 2>&1; echo "all rc=$?"
tail -6 gpurun_out/${T}_pytest_gpu_all.log

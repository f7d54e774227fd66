set -u
T=${TAG:-r2v}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q --timeout 900 -k "prove_batch or quotient or proof" > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/${T}_pytest.log
timeout 900 python tools/gpu_probe_batch_roots.py > gpurun_out/${T}_batch_roots.json 2> gpurun_out/${T}_batch_roots.err; echo "probe rc=$?"; tail -3 gpurun_out/${T}_batch_roots.err
cat gpurun_out/${T}_batch_roots.json

set -u
T=${TAG:-r2lad}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tail_and_group.py -x -q --timeout 600 -k "ladder or same_bytes" > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/${T}_pytest.log
timeout 600 python tools/gpu_probe_ladder.py > gpurun_out/${T}_probe_ladder.json 2> gpurun_out/${T}_probe_ladder.err; echo "probe rc=$?"; tail -3 gpurun_out/${T}_probe_ladder.err
cat gpurun_out/${T}_probe_ladder.json

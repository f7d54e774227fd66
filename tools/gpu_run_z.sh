set -u
T=${TAG:-r2z}
mkdir -p gpurun_out
cat > /tmp/lat.py <<'PY'
import os, sys
sys.path.insert(0, "kzg-commitments_b200"); sys.path.insert(0, ".")
import kzg_b200
ctx = kzg_b200.Context("bls12381", 0)
kind = int(sys.argv[1])
print(kind, 1e6 / ctx.microbench(kind))
ctx.close()
PY
for K in 13 10; do
  python /tmp/lat.py $K
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:bench_latency_chain -c 1 -o /tmp/lat$K python /tmp/lat.py $K > gpurun_out/${T}_ncu$K.log 2>&1; echo "ncu rc=$?"
  ncu -i /tmp/lat$K.ncu-rep --page raw --csv > gpurun_out/${T}_lat$K.raw.csv 2>/dev/null
  ncu -i /tmp/lat$K.ncu-rep --page source --csv > /tmp/lat$K.source.csv 2>/dev/null
  python tools/ncu_source_top.py /tmp/lat$K.source.csv gpurun_out/${T}_lat$K.source_top.txt; echo "top rc=$?"
  head -c 3000 /tmp/lat$K.source.csv > gpurun_out/${T}_lat$K.source_head.txt
done
ls -la gpurun_out/ | tail -8

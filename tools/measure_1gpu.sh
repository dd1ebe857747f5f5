#!/bin/bash
# Single-GPU record of a round: the full GPU test suite, one bench line per configuration, the batch-size sweep of
# config 2 and the launch list of the default bench command.  Run on the GPU box:
#   gpurun --timeout 1500 -- bash tools/measure_1gpu.sh      (outputs in gpurun_out/, copied to profiles/ by hand)
R=gpurun_out
python -m pytest tests -q -m gpu 2>&1 | tail -3 > $R/r02_gputests.txt; cat $R/r02_gputests.txt
for c in 2 2het 3 4 4h 5; do
  python bench.py --config $c 2>> $R/r02_bench.err | tail -1 > $R/r02_bench_$c.json
  python - <<PY
import json
d = json.load(open("$R/r02_bench_$c.json"))
print("$c", round(d["value"]), round(d["e2e"]["value"]), d["roofline"].get("kernel_ms"), d.get("cpu_baseline", {}).get("value"))
PY
done
: > $R/r02_batch_sweep.json
for n in 1024 2048 3000 4096 5000 6144 8192; do
  python bench.py --config 2 --streams $n --steps 30 --warmup 5 --no-cpu-baseline 2>> $R/r02_bench.err | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(json.dumps({'streams': $n, 'value': d['value'], 'e2e': d['e2e']['value'], 'isolated': d['isolated_steps']['value'], 'kernel_ms': d['roofline']['kernel_ms']}))" >> $R/r02_batch_sweep.json
done
cat $R/r02_batch_sweep.json
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $R/r02_launches_decode.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > /dev/null 2>&1
python tools/launch_times.py $R/r02_launches_decode.csv | head -12
tail -3 $R/r02_bench.err

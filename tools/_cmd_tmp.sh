python -m pytest tests/test_gpu_decode.py tests/test_gpu_robustness.py -x -q -m gpu 2>&1 | tail -2
for s in 0 8; do OPUSB_PARSE_SPW=$s python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/b2.json; python - <<PY
import json
d=json.load(open("gpurun_out/b2.json"))
print($s, d["value"], d["e2e"]["value"], d["e2e"]["synchronous_call_value"], d["isolated_steps"]["value"], d["roofline"]["kernel_ms"])
PY
done

python -m pytest tests/test_gpu_decode.py tests/test_gpu_robustness.py -x -q -m gpu 2>&1 | tail -2
python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/b2.json; python - <<PY
import json
d=json.load(open("gpurun_out/b2.json"))
print(d["value"], d["e2e"]["value"], d["isolated_steps"]["value"], d["roofline"]["kernel_ms"])
PY

#!/bin/bash
# 1/2/4/8-GPU lines of configs 2 and 5 only (the decode kernels changed after tools/scale_run.sh was last run):
#   gpurun --gpus 8 -- bash tools/scale_run2.sh
R=gpurun_out
run() {  # config n
  if [ "$2" = 1 ]; then
    python bench.py --gpus 1 --config $1 --steps 60 --warmup 5 --no-cpu-baseline
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29533 \
        bench.py --gpus $2 --config $1 --steps 60 --warmup 5 --no-cpu-baseline
  fi 2>> $R/r02_scale.err | tail -1
}
for c in ${SCALE_CONFIGS:-2 5}; do
  : > $R/r02_scale_$c.json
  for n in 1 2 4 8; do run $c $n >> $R/r02_scale_$c.json; done
done
python - <<'PY'
import json, os
for c in os.environ.get("SCALE_CONFIGS", "2 5").split():
    for l in open("gpurun_out/r02_scale_%s.json" % c):
        try:
            d = json.loads(l)
            print(c, d["n_gpus"], round(d["value"]), round(d["e2e"]["value"]))
        except Exception as e:
            print(c, "bad line", l[:200])
PY
tail -3 $R/r02_scale.err

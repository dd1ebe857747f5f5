#!/bin/bash
# 1/2/4/8-GPU lines of configs 2, 3 and 5 (one process per GPU, the driver's launch line), the copy-only probe and
# the one-process thread mode, on an 8-GPU box:  gpurun --gpus 8 -- bash tools/scale_run.sh
R=gpurun_out
run() {  # config n
  if [ "$2" = 1 ]; then
    python bench.py --gpus 1 --config $1 --steps 60 --warmup 5 --no-cpu-baseline
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29533 \
        bench.py --gpus $2 --config $1 --steps 60 --warmup 5 --no-cpu-baseline
  fi 2>> $R/r02_scale.err | tail -1
}
for c in 2 3 5; do
  : > $R/r02_scale_$c.json
  for n in 1 2 4 8; do run $c $n >> $R/r02_scale_$c.json; done
done
: > $R/r02_copy_probe.json
for n in 1 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29534 tools/copy_probe.py \
      2>> $R/r02_scale.err | tail -1 >> $R/r02_copy_probe.json
done
: > $R/r02_threads.json
for n in 2 8; do python bench.py --threads-per-process $n --config 2 --steps 60 --warmup 5 2>> $R/r02_scale.err | tail -1 >> $R/r02_threads.json; done
python -m pytest tests/test_gpu_threads.py -q 2>&1 | tail -2
nproc; tail -5 $R/r02_scale.err

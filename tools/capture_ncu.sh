#!/bin/bash
# One `ncu --set full` capture per kernel of the decode / encode paths (run on the GPU box through gpurun; the
# .ncu-rep files land in gpurun_out/ and are summarised into profiles/ with tools/ncu_summary.py and
# tools/ncu_by_function.py).  Each capture profiles ONE launch of the named kernel in a short bench run.
set -u
R=gpurun_out
SRC="--import-source on"
cap() {  # name kernel-regex config skip
  ncu --set full --clock-control none $SRC -k "regex:$2" -s "$4" -c 1 -f -o $R/r02_$1 \
      python bench.py --config "$3" --steps 4 --warmup 3 --no-cpu-baseline ${5:-} > /dev/null 2> $R/r02_$1.err
  ls -la $R/r02_$1.ncu-rep 2>/dev/null || tail -3 $R/r02_$1.err
}
# (gpurun brings back at most 64 MiB: the captures go in two batches, "main" and "rest")
if [ "${1:-main}" = main ]; then
  cap parse   celt_parse_kernel   2 4
  cap rebuild celt_rebuild_kernel 2 4
  cap synth   celt_synth_kernel   2 4
  cap silk    plan_silk_kernel    4 4
  cap encode  opus_encode_kernel  3 4
else
  SRC=""
  cap general opus_general_kernel 5 4
  cap msplan  opus_ms_plan_kernel 5 4
  cap scatter opus_ms_scatter_kernel 5 4
fi

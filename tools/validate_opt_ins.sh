#!/usr/bin/env bash
# First GPU call of the next round: run the parity tests of everything that was written without hardware access, then
# measure each opt-in against the default.  One GPU, about two minutes.  Results land in gpurun_out/opt_ins.log.
#   /usr/local/graft/bin/gpurun --timeout 400 -- 'bash tools/validate_opt_ins.sh'
set -u
mkdir -p gpurun_out
log=gpurun_out/opt_ins.log
: > "$log"
echo "== parity of the opt-in paths" | tee -a "$log"
DLS_TEST_UNVERIFIED=1 timeout 200 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_graph.py -k "skinny or fused_single" -q 2>&1 | tail -15 | tee -a "$log"
for wl in c3 c1; do
  for flags in "" "DLS_FUSED_SGD=1" "DLS_SKINNY_GEMM=1" "DLS_FUSED_SGD=1 DLS_SKINNY_GEMM=1"; do
    echo "== $wl [$flags]" | tee -a "$log"
    env $flags timeout 120 python bench.py --workload $wl --steps 200 --warmup 10 2>>"$log" \
      | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['launches_per_step'], d['loss'])" | tee -a "$log"
  done
done

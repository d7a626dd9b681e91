#!/bin/bash
# One `ncu --set full` capture per hot kernel family (run under gpurun, one GPU, each only after the same command
# exited 0 without ncu).  Summaries are made ON the box (gpurun_out/ is capped at 64 MiB): <tag>_<name>_summary.txt
# (tools/ncu_summary.py) and <tag>_<name>_lines.txt (stall samples per source line); reports above 12 MB are dropped.
#   bash tools/ncu_round.sh r1k
tag=${1:-rX}
out=gpurun_out
mkdir -p $out
NCU="ncu --set full --clock-control none --import-source on -f"
run() { # name, kernel regex, skip, count, command...
  name=$1; k=$2; s=$3; c=$4; shift 4
  "$@" > $out/${tag}_${name}_plain.log 2>&1 || { echo "$name: plain run failed"; tail -5 $out/${tag}_${name}_plain.log; return; }
  tail -2 $out/${tag}_${name}_plain.log
  timeout 600 $NCU -k "regex:$k" -s $s -c $c -o $out/${tag}_${name} "$@" > $out/${tag}_${name}_ncu.log 2>&1
  echo "$name: ncu rc=$?"
  rep=$out/${tag}_${name}.ncu-rep
  [ -f $rep ] || return
  { echo "# $NCU -k regex:$k -s $s -c $c  $*"; python tools/ncu_summary.py $rep; } > $out/${tag}_${name}_summary.txt 2>&1
  ncu -i $rep --page source --csv --print-source cuda,sass > /tmp/${name}_src.csv 2>/dev/null
  python tools/ncu_lines.py /tmp/${name}_src.csv 25 > $out/${tag}_${name}_lines.txt 2>&1
  [ $(stat -c %s $rep) -gt 12000000 ] && rm -f $rep
}
# kbench attn --iters 3 launches: 7 forwards (1 + 3 warm-up + 3 timed), then 6 backwards
run attn128 'attn2_' 6 2 python tools/kbench.py attn --iters 3
# multi-tile: [prerotate, forward] x 7, then [delta, dQ pass, dK|dV pass] x 6: skip 12 -> the 7th forward pair + one backward
# (S = 512: the multi-step kernels; 192 x 196: the wide-key forward and dQ pass, the multi-step dK|dV pass)
run attn512 'attn2_fwd_ms|attn2_bwd_|attn_delta|attn_tc_|attn_prerotate' 12 5 python tools/kbench.py attn --n 128 --s 512 --iters 3
run attn196 'attn2_fwd_wk|attn2_fwd_ms|attn2_bwd_|attn_delta|attn_tc_|attn_prerotate' 12 5 python tools/kbench.py attn --n 192 --s 196 --causal 0 --iters 3
run norm 'norm_' 5 3 python tools/kbench.py norm --iters 3
run gemm 'gemm_tn|nvjet' 5 2 python tools/kbench.py gemm --iters 3
# general forms: after the 6 + 6 forward launches come 6 dgrad launches, then 6 wgrad launches (gemm_gen_kernel)
run gemm2 'gemm_gen' 5 2 python tools/kbench.py gemm --iters 3
ls -la $out/

#!/bin/bash
# One `ncu --set full` capture per hot kernel family (run under gpurun, one GPU, after the same commands exited 0
# without ncu).  Reports land in gpurun_out/<tag>_*.ncu-rep; summarise here with tools/ncu_summary.py.
#   bash tools/ncu_round.sh r1j
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
}
run attn128 'attn2_' 6 2 python tools/kbench.py attn --iters 3
run attn512 'attn_tc_|attn_prerotate' 0 40 python tools/kbench.py attn --n 128 --s 512 --iters 3
run attn196 'attn_tc_|attn_prerotate' 0 40 python tools/kbench.py attn --n 192 --s 196 --causal 0 --iters 3
run norm 'norm_' 5 3 python tools/kbench.py norm --iters 3
run gemm 'gemm_tn|nvjet' 5 2 python tools/kbench.py gemm --iters 3
ls -la $out/${tag}_*.ncu-rep

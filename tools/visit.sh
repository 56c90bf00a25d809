mkdir -p gpurun_out
( for b in 256 32; do for k in 0 1 2; do MSTK_WG_KSMAX=$k timeout 120 python tools/wgrad_tune.py $b 2>&1 | tail -11; done; done ) > gpurun_out/wg.log 2>&1
cat gpurun_out/wg.log

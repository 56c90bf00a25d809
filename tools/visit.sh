mkdir -p gpurun_out
( for n in 96 192 384; do for m in 1 2 4 100; do echo "n=$n min=$m"; MSTK_HALO_MIN=$m timeout 120 python tools/halo_tune.py $n 2>&1 | tail -6 | cut -c 1-60; done; done ) > gpurun_out/halo6.log 2>&1
cat gpurun_out/halo6.log

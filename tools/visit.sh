mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_conv_halo.py -q -m gpu -x --timeout 300 2>&1 | tail -2
for g in 8 4; do echo "CTG $g"; MSTK_HALO_CTG=$g timeout 120 python tools/tcp_tune.py 2>&1 | tail -2 | head -1; done ) > gpurun_out/halo5.log 2>&1
cat gpurun_out/halo5.log | cut -c 1-200

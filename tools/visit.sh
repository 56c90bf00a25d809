mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_audio_inverse.py tests/test_gpu_mel.py -q -m gpu -x --timeout 600 2>&1 | tail -3
timeout 300 python tools/bench_inverse.py 64 128 5 ) > gpurun_out/gl2.log 2>&1
cat gpurun_out/gl2.log | cut -c 1-400

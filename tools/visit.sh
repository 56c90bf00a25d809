mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_audio_inverse.py tests/test_gpu_mel.py tests/test_zz_no_alloc.py -q -m gpu -x -s --timeout 600 > gpurun_out/t_inv.log 2>&1; echo "pytest rc=$?"; tail -n 40 gpurun_out/t_inv.log
timeout 300 python tools/bench_inverse.py > gpurun_out/inv_bench.json 2>gpurun_out/inv_bench.err; echo "inv bench rc=$?"; cat gpurun_out/inv_bench.json; tail -3 gpurun_out/inv_bench.err
timeout 300 python tools/profile_train.py 256 5 > gpurun_out/pt_plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/train_launches.csv python tools/profile_train.py 256 3 > gpurun_out/pt_ncu.log 2>&1; echo "ncu train rc=$?"
python tools/summarize_launches.py gpurun_out/train_launches.csv > gpurun_out/train_launches.md 2>&1; head -50 gpurun_out/train_launches.md

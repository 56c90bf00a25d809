mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 2400 python -m pytest tests -q -m gpu --timeout 900 > gpurun_out/t_gpu.log 2>&1; echo "pytest_gpu rc=$?" | tee gpurun_out/rc.txt
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/rc.txt
timeout 1200 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; echo "bench_ref rc=$?" | tee -a gpurun_out/rc.txt
tail -n 4 gpurun_out/t_gpu.log; tail -n 3 gpurun_out/smoke.log; tail -n 3 gpurun_out/bench.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench.log").read().strip().splitlines()[-1])
s = d["summary"]
print(json.dumps(s["sampling"])); print(json.dumps(s["train_strong"])); print(json.dumps(s["vae"])); print(json.dumps(s["config5"])); print(s["autoencoder_samples_per_s"], json.dumps(s["gpu_baseline"])[:600])
print(json.dumps(d["train"]["e2e"]), json.dumps(d["train_autoencoder"]["e2e"]))
PY
timeout 300 python tools/profile_train.py 256 3 > gpurun_out/pt_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/train_launches.csv python tools/profile_train.py 256 3 > gpurun_out/pt_ncu.log 2>&1; echo "ncu_train rc=$?"
for m in eval train; do
  timeout 300 python tools/profile_vae.py $m 3 > /dev/null 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file gpurun_out/vae_${m}_launches.csv python tools/profile_vae.py $m 3 > /dev/null 2>&1; echo "ncu_vae_$m rc=$?"
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"conv_halo_kernel" -s 1 -c 1 -f -o gpurun_out/prof_halot python tools/profile_vae.py eval 3 > gpurun_out/prof_halot.log 2>&1; echo "ncu_full_halot rc=$?"
[ -f gpurun_out/prof_halot.ncu-rep ] && ncu -i gpurun_out/prof_halot.ncu-rep --page raw --csv > gpurun_out/prof_halot_raw.csv 2>/dev/null; rm -f gpurun_out/prof_halot.ncu-rep
du -sh gpurun_out

#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench (both arms), ncu launch list, ncu --set full of the tcgen05 kernel.
# Everything lands in gpurun_out/; numbers printed under ncu are never bench values.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 > gpurun_out/t_gpu.log 2>&1; echo "pytest_gpu rc=$?" | tee gpurun_out/rc.txt
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/rc.txt
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; echo "bench_ref rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/b_short.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1; echo "ncu_list rc=$?" | tee -a gpurun_out/rc.txt
timeout 300 python tools/profile_step.py 4 > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"conv_tc_kernel|attention" -s 36 -c 17 -o gpurun_out/prof_step python tools/profile_step.py 4 > gpurun_out/prof_ncu.log 2>&1; echo "ncu_full rc=$?" | tee -a gpurun_out/rc.txt
tail -n 5 gpurun_out/t_gpu.log; tail -n 4 gpurun_out/smoke.log; cut -c 1-1800 gpurun_out/bench.log; tail -n 5 gpurun_out/bench.err; cut -c 1-600 gpurun_out/bench_ref.log
timeout 300 python tools/profile_train.py 256 3 > gpurun_out/pt_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/train_launches.csv python tools/profile_train.py 256 3 > gpurun_out/pt_ncu.log 2>&1; echo "ncu_train rc=$?" | tee -a gpurun_out/rc.txt
timeout 900 ncu --set full --clock-control none -k regex:"conv_tc_kernel|wgrad_tc_kernel|feat_fused3_kernel|conv_cin1_row|wgrad_edge" --launch-skip 70 --launch-count 36 -o gpurun_out/prof_train python tools/profile_train.py 256 3 > gpurun_out/prof_train_ncu.log 2>&1; echo "ncu_full_train rc=$?" | tee -a gpurun_out/rc.txt
# keep the raw pages as CSV; drop reports that would push gpurun_out over the 64 MiB merge limit
for r in prof_step prof_train; do
  if [ -f gpurun_out/$r.ncu-rep ]; then
    ncu -i gpurun_out/$r.ncu-rep --page raw --csv > gpurun_out/${r}_raw.csv 2>/dev/null
    sz=$(stat -c %s gpurun_out/$r.ncu-rep); if [ "$sz" -gt 25000000 ]; then rm -f gpurun_out/$r.ncu-rep; fi
  fi
done
[ -f gpurun_out/prof_step_raw.csv ] && python tools/traffic_from_ncu.py gpurun_out/prof_step_raw.csv conv_tc_kernel gpurun_out/conv_tc_traffic.json \
  "ncu --set full --clock-control none -k regex:conv_tc_kernel|attention -s 36 -c 17 python tools/profile_step.py 4"
# config-2 VAE round trip: launch list with DRAM bytes (eval and train mode)
for m in eval train; do
  timeout 300 python tools/profile_vae.py $m 3 > /dev/null 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file gpurun_out/vae_${m}_launches.csv python tools/profile_vae.py $m 3 > /dev/null 2>&1
done
du -sh gpurun_out | tee -a gpurun_out/rc.txt

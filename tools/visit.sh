mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-train --no-vae --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/bl_plain.log 2> gpurun_out/bl_plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/bench_launches.csv $CMD > gpurun_out/bl_ncu.log 2> gpurun_out/bl_ncu.err; echo "ncu rc=$?"
wc -l gpurun_out/bench_launches.csv
python tools/summarize_launches.py gpurun_out/bench_launches.csv --md > gpurun_out/bench_launches.md; head -25 gpurun_out/bench_launches.md | cut -c 1-200

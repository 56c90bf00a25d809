mkdir -p gpurun_out
run() { MSTK_TC_FORCE="$1" timeout 300 python bench.py --steps 6 --no-train --no-vae --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline > gpurun_out/bk.log 2> gpurun_out/bk.err; echo "force=$1 rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/bk.log').read().strip().splitlines()[-1]); s=d['summary']['sampling']; print(s['unet_step_us'], s['sum_isolated_kernel_us'], s['pdl_overlap_frac'])"; }
run "0,0,0,0"; run "0,0,2,2"; run "0,0,3,1"; run "0,0,2,1"; run "0,0,4,1"; run "0,0,3,2"

#!/bin/bash
# training legs only (config 4 step + autoencoder step)
mkdir -p gpurun_out
python bench.py --steps 5 --no-vae --no-mel --no-fp32 --no-gpu-baseline --no-cpu-baseline > gpurun_out/bt.log 2>gpurun_out/bt.err; tail -2 gpurun_out/bt.err
python -c "
import json; d=json.loads(open('gpurun_out/bt.log').read().strip().splitlines()[-1]); print(json.dumps(d['summary']['train_strong'])); print(d['train']['gpu_launches_per_step'], d['summary']['autoencoder_samples_per_s'])"

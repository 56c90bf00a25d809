#!/bin/bash
# quick A/B visit: edge-conv + feature-loss + model tests, then two short bench runs
timeout 600 python -m pytest tests/test_gpu_conv_edge.py tests/test_gpu_feature_loss.py tests/test_gpu_model.py -q -m gpu -x 2>&1 | tail -3
for i in 1 2; do
timeout 500 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-vae > gpurun_out/bq$i.log 2> gpurun_out/bq$i.err; tail -3 gpurun_out/bq$i.err
python -c "
import json; d=json.loads(open('gpurun_out/bq$i.log').read().strip().splitlines()[-1]); print(d['value'], d['train']['value'], d['train']['ms_per_step'], d['train_autoencoder']['value'])"
done

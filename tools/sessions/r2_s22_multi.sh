#!/bin/bash
# round 2, GPU session 22 (2 GPUs): multi-device contexts created by one worker thread per device -- parity subset, headless set-up time,
# the torchrun arm at N = 2 (final library)
set -u
o=gpurun_out/s22; mkdir -p $o
N=$(nvidia-smi -L | wc -l); echo "GPUs: $N"
timeout 900 python -m pytest tests -m gpu -x -q -k "multi or bands or headless or no_alloc" > $o/pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -2 $o/pytest_multi.log
simplepathtracer_b200/host/spt_headless --scene random --seed 1 --width 1920 --height 1080 --spp 1024 --depth 50 --tracer task --gpus $N --repeat 2 --out $o/h_n.bmp | tail -5
rm -f $o/*.bmp
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29577 bench.py --gpus $N --steps 3 --warmup 3 > $o/bench_n$N.json 2>$o/bench_n$N.err; echo "n$N rc=$?"
python - <<P
import json
d=json.loads(open('$o/bench_n$N.json').read().strip().splitlines()[-1])
print('n$N value %.1f ms %.2f | e2e %.1f (%.2f ms) | frac %.4f step %.4f' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['roofline']['frac'], d['roofline']['step']['frac']))
P

#!/bin/bash
# round 2, GPU session 13 (8 GPUs of one box, lean: the call is charged eightfold): the torchrun arm at N = 8 with the end-to-end leg
# through ONE multi-device context over the eight GPUs, and the headless program with --gpus 8
set -u
o=gpurun_out/s13; mkdir -p $o
N=$(nvidia-smi -L | wc -l); echo "GPUs: $N" | tee $o/gpus.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29588 bench.py --gpus $N --steps 3 --warmup 3 > $o/bench_n$N.json 2>$o/bench_n$N.err; echo "n$N rc=$?"
python - <<P
import json
try:
    d=json.loads(open('$o/bench_n$N.json').read().strip().splitlines()[-1])
    print('n$N value %.1f ms %.2f | e2e %.1f (%.2f ms, %s launches) | frac %.4f step %.4f share %.4f' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e'].get('launches_per_step'), d['roofline']['frac'], d['roofline']['step']['frac'], d['roofline']['step']['kernel_share_of_step']))
    print(d['roofline']['kernels']['trace_secondary'].get('lanes_with_a_ray'), d['config']['staging'])
except Exception as e:
    print('FAILED', e); print(open('$o/bench_n$N.err').read()[-2500:])
P
simplepathtracer_b200/host/spt_headless --scene random --seed 1 --width 3840 --height 2160 --spp 1024 --depth 50 --tracer task --gpus $N --repeat 2 --out $o/h_n.bmp | tail -5
rm -f $o/*.bmp

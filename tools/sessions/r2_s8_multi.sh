#!/bin/bash
# round 2, GPU session 8 (N GPUs of one box; N = number visible): the multi-device context on real devices, the torchrun arm
# (device-timed value + e2e through spt_cuda_create_multi on rank 0), headless --gpus N
set -u
o=gpurun_out/s8; mkdir -p $o
N=$(nvidia-smi -L | wc -l); echo "GPUs: $N" | tee $o/gpus.txt
timeout 900 python -m pytest tests -m gpu -x -q -k "multi or bands or headless or no_alloc" > $o/pytest_multi.log 2>&1; echo "pytest rc=$?" | tee -a $o/pytest_multi.log; tail -4 $o/pytest_multi.log
timeout 600 python bench.py --steps 3 --warmup 2 --no-extra > $o/bench_n1.json 2>$o/bench_n1.err; echo "n1 rc=$?"
for n in 2 4 8; do
  [ $n -le $N ] || continue
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 295$n bench.py --gpus $n --steps 3 --warmup 2 > $o/bench_n$n.json 2>$o/bench_n$n.err; echo "n$n rc=$?"
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29599 bench.py --gpus $N --steps 3 --warmup 2 --partition bands > $o/bench_n${N}_bands.json 2>$o/bench_n${N}_bands.err; echo "n$N bands rc=$?"
python - <<P
import json,glob
for f in sorted(glob.glob('$o/bench_n*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1], 'value %.1f ms %.2f | e2e %.1f (%.2f ms, %s launches) | frac %.4f step %.4f share %.4f' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e'].get('launches_per_step'), d['roofline']['frac'], d['roofline']['step']['frac'], d['roofline']['step']['kernel_share_of_step']))
    except Exception as e:
        print(f,'FAILED',e); print(open(f.replace('.json','.err')).read()[-1500:])
P
simplepathtracer_b200/host/spt_headless --scene random --seed 1 --width 1920 --height 1080 --spp 1024 --depth 50 --tracer task --gpus $N --repeat 3 --out $o/h_n.bmp | tail -7
simplepathtracer_b200/host/spt_headless --scene random --seed 1 --width 1920 --height 1080 --spp 1024 --depth 50 --tracer task --repeat 3 --out $o/h_1.bmp | tail -7
rm -f $o/*.bmp

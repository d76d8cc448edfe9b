#!/bin/bash
# Round-2 scaling evidence on an N-GPU box: pre-training presets, the fine-tune workload and the 2-rank NCCL tests.
# usage: bash tools/scale_runs.sh N tag
N=${1:-8}; TAG=${2:-r2e}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --no-kernel-timings"
$T --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_${N}gpu.json 2> gpurun_out/${TAG}_${N}.err
$T --steps 5 --warmup 3 --config vit_l16 --micro-batch 512 > gpurun_out/${TAG}_vitl_${N}gpu.json 2>> gpurun_out/${TAG}_${N}.err
$T --steps 5 --warmup 3 --config vit_h14 --micro-batch 256 > gpurun_out/${TAG}_vith_${N}gpu.json 2>> gpurun_out/${TAG}_${N}.err
$T --steps 5 --warmup 3 --workload finetune > gpurun_out/${TAG}_finetune_${N}gpu.json 2>> gpurun_out/${TAG}_${N}.err
python -m pytest tests/test_dp_gpu.py -x -q 2>&1 | tail -3
for f in gpurun_out/${TAG}_*_${N}gpu.json; do python -c "
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], round(d['value']), round(d['e2e']['value']), round(d['roofline']['frac'],4), d['clocks']['sm_mhz'])
" $f; done
tail -3 gpurun_out/${TAG}_${N}.err

# NCCL channel-count sweep at 2 GPUs in the per-rank regime of the 8-GPU run (512 images per rank and step)
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --global-batch 1024 --micro-batch 512 --steps 20 --warmup 5 --no-kernel-timings --no-cpu-baseline"
for v in default min16 min32 default2; do
  case $v in
    min16) NCCL_MIN_NCHANNELS=16 NCCL_MIN_CTAS=16 $T > gpurun_out/nccl_$v.json 2>/dev/null;;
    min32) NCCL_MIN_NCHANNELS=32 NCCL_MIN_CTAS=32 $T > gpurun_out/nccl_$v.json 2>/dev/null;;
    *) $T > gpurun_out/nccl_$v.json 2>/dev/null;;
  esac
done
python - <<PY
import json
for n in ("default","min16","min32","default2"):
    try:
        d=json.loads(open(f"gpurun_out/nccl_{n}.json").read().strip().splitlines()[-1]); print(n, round(d["value"]), d["ms_per_step"], round(d["e2e"]["value"]), d["clocks"]["sm_mhz"])
    except Exception as e: print(n, "fail", e)
PY

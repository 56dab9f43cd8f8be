set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu > gpurun_out/scale_n1.json 2> gpurun_out/scale.err
for n in 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29600+n)) bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/scale_n$n.json 2>> gpurun_out/scale.err
done
$TR --nproc-per-node 8 --master-port 29620 bench.py --gpus 8 --steps 10 --warmup 3 --contiguous > gpurun_out/scale_n8_contig.json 2>> gpurun_out/scale.err
$TR --nproc-per-node 8 --master-port 29621 bench.py --gpus 8 --steps 10 --warmup 3 --gather nccl > gpurun_out/scale_n8_nccl.json 2>> gpurun_out/scale.err
for p in mixed fp16 bf16; do
  $TR --nproc-per-node 8 --master-port 29630 bench.py --gpus 8 --orbit 200 --res 1600 --precision $p > gpurun_out/orbit_$p.json 2>> gpurun_out/scale.err
done
tail -c 600 gpurun_out/scale.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/scale_n*.json"))+sorted(glob.glob("gpurun_out/orbit_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d["value"]/1e6,3), "Mrays/s", round(d["ms_per_step"],2), "ms", (d.get("e2e") or {}).get("value"), (d.get("parity") or {}))
    except Exception as e: print(f,"ERR",e)
PY

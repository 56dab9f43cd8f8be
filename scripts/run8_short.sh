set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu > gpurun_out/scale_n1.json 2> gpurun_out/scale.err
for n in 2 4 8; do
  $TR --nproc-per-node $n --master-port $((29600+n)) bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/scale_n$n.json 2>> gpurun_out/scale.err
done
$TR --nproc-per-node 8 --master-port 29630 bench.py --gpus 8 --orbit 200 --res 1600 --precision mixed > gpurun_out/orbit_mixed.json 2>> gpurun_out/scale.err
tail -c 300 gpurun_out/scale.err

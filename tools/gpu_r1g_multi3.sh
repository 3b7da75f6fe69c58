set -x
mkdir -p gpurun_out/r1g
for n in 8 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline --shape 22x1125 --layout btc --batch 56832 2>> gpurun_out/r1g/multi3.err | tail -1 > gpurun_out/r1g/scale_22x1125_btc_peer_n$n.json
cut -c1-200 gpurun_out/r1g/scale_22x1125_btc_peer_n$n.json
done
tail -3 gpurun_out/r1g/multi3.err

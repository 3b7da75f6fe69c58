set -x
mkdir -p gpurun_out/r1g
run() { # N tag args...
  n=$1; tag=$2; shift 2
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline "$@" 2>> gpurun_out/r1g/multi2.err | tail -1 > gpurun_out/r1g/scale_${tag}_n$n.json
  cut -c1-300 gpurun_out/r1g/scale_${tag}_n$n.json
}
run 8 stream --workload stream
run 4 stream --workload stream
run 2 stream --workload stream
run 8 22x1125_btc --shape 22x1125 --layout btc --batch 56832
run 2 64x480 --shape 64x480
tail -3 gpurun_out/r1g/multi2.err

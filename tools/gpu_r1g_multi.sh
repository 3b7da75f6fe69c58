set -x
mkdir -p gpurun_out/r1g
run() { # N shape extra
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29500 + $1)) bench.py --gpus $1 --steps 20 --warmup 5 --shape $2 --no-cpu-baseline $3 2>> gpurun_out/r1g/multi.err | tail -1 > gpurun_out/r1g/scale_$2_n$1.json
  cat gpurun_out/r1g/scale_$2_n$1.json | cut -c1-400
}
run 8 22x1125
run 4 22x1125
run 2 22x1125
run 4 64x480
run 8 64x480
tail -3 gpurun_out/r1g/multi.err

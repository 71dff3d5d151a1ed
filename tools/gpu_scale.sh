# strong-scaling bench at N GPUs of one box, launched as the driver does
N=$1
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 40 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "rc=$?"; tail -2 gpurun_out/bench_n$N.err; cat gpurun_out/bench_n$N.json | cut -c1-600

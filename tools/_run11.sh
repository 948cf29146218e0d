set -u
out=gpurun_out/r2j; mkdir -p $out
LYNX_EX_TRACE=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 100 --warmup 10 --no-extras > $out/bench_n2.json 2> $out/bench_n2.err; echo "bench n2 rc=$?"; grep "rank" $out/bench_n2.err | tail -12; head -c 300 $out/bench_n2.json

set -u
out=gpurun_out/r2k; mkdir -p $out
export LYNX_B200_LIB=lynx_b200/lib/liblynx_b200_exp.so
for m in push pull; do
echo "== $m"
if [ $m = pull ]; then export LYNX_EX_PULL=1; fi
LYNX_EX_TRACE=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 100 --warmup 10 --no-extras > $out/bench_n2_$m.json 2> $out/bench_n2.err; grep "rank 0" $out/bench_n2.err | grep "us after\|stamp kernel" | sed 's/.*\(us after\)/\1/'; python -c "
import json;s=open('$out/bench_n2_$m.json').read();d=json.loads(s[s.index('{'):]);print('step',d['ms_per_step']*1e3,'check',d.get('exchange_check'))"
done

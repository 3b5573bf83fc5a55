F="--gpus 2 --steps 20 --warmup 5 --strong none --no-c5 --no-parity --no-unpruned --no-exe-check --no-cpu-baseline"
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555"
for v in sharded replicated nogather; do
  case $v in
    sharded) $R bench.py $F > gpurun_out/n2_$v.json 2> gpurun_out/n2_$v.err;;
    replicated) $R bench.py $F --table replicated > gpurun_out/n2_$v.json 2> gpurun_out/n2_$v.err;;
    nogather) EREF_BENCH_NO_GATHER=1 $R bench.py $F --table replicated > gpurun_out/n2_$v.json 2> gpurun_out/n2_$v.err;;
  esac
  python -c "
import json; d=json.load(open('gpurun_out/n2_$v.json')); print('$v', d['ms_per_step'], sum(d['phases_ms'].values()), d['phases_ms'])"
done

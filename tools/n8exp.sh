# diagnostic: weak-scaled headline step at N GPUs, sharded vs replicated table build, with and without the clock sampler
N=${1:-8}
F="--gpus $N --steps 30 --warmup 5 --strong none --no-c5 --no-parity --no-unpruned --no-exe-check --no-cpu-baseline"
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555"
run() { name=$1; shift; "$@" > gpurun_out/n${N}_$name.json 2> gpurun_out/n${N}_$name.err; python -c "
import json; d=json.load(open('gpurun_out/n${N}_$name.json')); print('$name', d['ms_per_step'], round(sum(d['phases_ms'].values()),3), d['step_spread_ms_rank0'], d['phases_ms']['query_bucket_ms'], d['phases_ms']['query_count_ms'])"; }
run sharded $R bench.py $F
run replicated $R bench.py $F --table replicated
EREF_BENCH_NO_SAMPLER=1 run sharded_nosampler $R bench.py $F

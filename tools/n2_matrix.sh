export TK_BENCH_NO_DIAG=1
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --warmup 5 --e2e-steps 0"
pick='import json,sys; l=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(sys.argv[1], round(l["ms_per_step"]*1e3,3), "us/step", l["stats_allreduce"]["launched_in_timed_region"])'
for i in 1 2; do
$R --steps 20 2>/dev/null | python -c "$pick" n2_steps20
TK_BENCH_NO_HOOK=1 $R --steps 20 2>/dev/null | python -c "$pick" n2_steps20_nohook
$R --steps 100 2>/dev/null | python -c "$pick" n2_steps100
python bench.py --steps 20 --warmup 5 --e2e-steps 0 --no-extras --no-cpu-baseline 2>/dev/null | python -c "$pick" n1_steps20
TK_BENCH_NO_HOOK=1 python bench.py --steps 20 --warmup 5 --e2e-steps 0 --no-extras --no-cpu-baseline 2>/dev/null | python -c "$pick" n1_steps20_nohook
python bench.py --steps 100 --warmup 5 --e2e-steps 0 --no-extras --no-cpu-baseline 2>/dev/null | python -c "$pick" n1_steps100
done

#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out results
{
echo "=== p2p bench"; timeout 200 python -m ddlb_b200.launcher -np $N --timeout 150 -m ddlb_b200.parallel.p2p_bench --mib 128 --iters 3 2>&1 | tee gpurun_out/p2p_bench_$N.json | python -c "import sys,json; t=sys.stdin.read(); d=json.loads(t[t.index('{'):]); print({k:v for k,v in d['results'].items() if '148' in k or 'local' in k or '32' in k})"
echo "=== pytest multigpu"; timeout 600 python -m pytest tests/test_multigpu.py -x -q -k "$N" 2>&1 | tail -5
echo "=== bench ours N=$N"; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 10 2>gpurun_out/bench_ours_$N.err | tee gpurun_out/bench_ours_n$N.json | cut -c1-600; tail -3 gpurun_out/bench_ours_$N.err
echo "=== bench reference N=$N"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --impl reference --gpus $N --steps 30 --warmup 5 2>gpurun_out/bench_ref_$N.err | tee gpurun_out/bench_ref_n$N.json | cut -c1-600; tail -3 gpurun_out/bench_ref_$N.err
echo "=== harness sweeps"
timeout 900 python -m ddlb_b200.launcher -np $N --timeout 800 scripts/run_benchmark.py scripts/dev/dev_perf8_col.json 2>&1 | grep -E "Error|error|Traceback" | head -5
timeout 600 python -m ddlb_b200.launcher -np $N --timeout 500 scripts/run_benchmark.py scripts/dev/dev_perf8_row.json 2>&1 | grep -E "Error|error|Traceback" | head -5
python - <<'PY'
import pandas as pd, glob
pd.set_option("display.width", 250); pd.set_option("display.max_colwidth", 100)
for f in sorted(glob.glob("results/dev_perf_*.csv")):
    df = pd.read_csv(f)
    df["impl"] = df["implementation"].str.replace(", comm_sms=0, cta_group=0, block_n=0", "").str.replace(", cta_group=0, block_n=0, transport=p2p","")
    print(f); print(df[["impl", "option", "mean_time (ms)", "min_time", "std_time", "Throughput (TFLOPS)", "valid", "pct_roofline", "error"]].to_string())
PY
cp results/dev_perf_*.csv gpurun_out/
} 2>&1 | tee gpurun_out/session_$N.log

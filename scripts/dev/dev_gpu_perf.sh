#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out results
{
timeout 900 python -m ddlb_b200.launcher -np $N --timeout 800 scripts/run_benchmark.py scripts/dev/dev_perf_col.json 2>&1 | grep -E "Error|error|Traceback" | head
timeout 600 python -m ddlb_b200.launcher -np $N --timeout 500 scripts/run_benchmark.py scripts/dev/dev_perf_row.json 2>&1 | grep -E "Error|error|Traceback" | head
python - <<'PY'
import pandas as pd, glob
pd.set_option("display.width", 250); pd.set_option("display.max_colwidth", 90)
for f in sorted(glob.glob("results/dev_perf_*.csv")):
    df = pd.read_csv(f)
    df["impl"] = df["implementation"].str.replace(", cta_group=0, block_n=0, transport=p2p", "").str.replace(", block_n=0, transport=p2p","")
    print(f); print(df[["impl", "mean_time (ms)", "min_time", "std_time", "Throughput (TFLOPS)", "valid", "pct_roofline", "error"]].to_string())
PY
cp results/dev_perf_*.csv gpurun_out/
} 2>&1 | tee gpurun_out/perf_$N.log

set -x
mkdir -p gpurun_out
T=r2v
for K in 8 16 32 64 128; do
VPG_TOK_BLOCKS_PER_SM=$K timeout 300 python scripts/bench_token_ops.py > gpurun_out/${T}_token_ops_k$K.jsonl 2> gpurun_out/${T}_token_ops_k$K.err; echo "k=$K rc=$?"
done
python - <<'PY'
import json
tabs={}
for K in (8,16,32,64,128):
    for l in open(f"gpurun_out/r2v_token_ops_k{K}.jsonl"):
        d=json.loads(l)
        if "frac_of_hbm_peak" in d: tabs.setdefault(d["op"][:44],{})[K]=d["frac_of_hbm_peak"]
        elif d.get("config","").startswith("C3"): tabs.setdefault("C3us "+d["op"][:38],{})[K]=d["us"]
for op,v in tabs.items(): print(f"{op:46s}", v)
PY

# r02 call O: decode tests, instruction counts per kernel of one file job, rate with 24 slots
export ICB_POOL_CACHE=/tmp/icb_pool
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -3
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_decode_inst.csv python profiles/r02/decode_ncu.py > gpurun_out/ncu_o.log 2>&1
python - <<'PY'
import csv, collections
rows = list(csv.DictReader(l for l in open("gpurun_out/r02_decode_inst.csv") if l.startswith('"')))
agg = collections.OrderedDict()
for r in rows:
    k = r["Kernel Name"].split("(")[0][:28]
    d = agg.setdefault(k, {"n": 0, "ns": 0.0, "inst": 0.0})
    v = float(r["Metric Value"].replace(",", ""))
    if r["Metric Name"] == "gpu__time_duration.sum": d["n"] += 1; d["ns"] += v * (1000 if r["Metric Unit"] in ("us", "usecond") else 1)
    elif r["Metric Name"] == "smsp__inst_executed.sum": d["inst"] += v
jobs = 3
for k, d in agg.items():
    print("%-28s launches/job %5.1f  us/job %8.1f  Mwarp-inst/job %7.2f" % (k, d["n"] / jobs, d["ns"] / jobs / 1000, d["inst"] / jobs / 1e6))
PY
PROBE_SLOTS=24 timeout 600 python profiles/r02/decode_probe2.py 2>&1 | sed -n 2,4p

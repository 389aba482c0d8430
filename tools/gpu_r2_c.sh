#!/bin/bash
# round 2, GPU pass C: batched engine tests, then an ncu launch list of a few batched decode steps (no graph: every kernel listed)
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_t3_batch_gpu.py -q -s > gpurun_out/pytest_batch.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_batch.log
grep -E "max-abs|equal|passed|failed|Error|error|assert" gpurun_out/pytest_batch.log | tail -40
for u in 1 8 64; do timeout 300 python tools/t3b_time.py $u 2>&1 | tail -1; done | tee gpurun_out/t3b_sweep.log
CBX_T3B_GRAPH=0 timeout 300 python tools/t3b_time.py 64 6 > gpurun_out/plain_t3b.log 2>&1 &&
CBX_T3B_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_t3b.csv python tools/t3b_time.py 64 6 > gpurun_out/ncu_t3b.log 2>&1
tail -2 gpurun_out/ncu_t3b.log
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open('gpurun_out/launches_t3b.csv')) if len(r) > 5]
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); gi = hdr.index('Grid Size') if 'Grid Size' in hdr else None
agg = collections.OrderedDict()
for r in rows[1:]:
    try: v = float(r[vi].replace(',', ''))
    except: continue
    k = r[ki][:70] + (' ' + r[gi] if gi is not None else '')
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
    print(f"{k:100s} {a[0]:6d} {a[1] / a[0] / 1e3:9.2f} us {a[1] / 1e6:9.2f} ms {100 * a[1] / tot:5.1f}%")
PY

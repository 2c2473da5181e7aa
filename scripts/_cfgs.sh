python scripts/run_lattice.py --n 408 --structures 8 > gpurun_out/s10_c5_8x408.json 2>gpurun_out/s10.err; cut -c1-1200 gpurun_out/s10_c5_8x408.json
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --per-case-section > gpurun_out/s10_bench_c2b.json 2>>gpurun_out/s10.err; cut -c1-200 gpurun_out/s10_bench_c2b.json; python -c "
import json; d=json.load(open('gpurun_out/s10_bench_c2b.json')); print(d['value'], d['roofline']['kernel_ms'], d['e2e']['value'])"
python - <<'PY'
import sys,time,json
sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np, fixtures
from openbeam_b200 import obgpu
from openbeam_b200.problem import CStructureProblem, StaticSolveProblemInfo, StressInfo
ctx=obgpu.Context(0)
st=fixtures.demo1()
p=CStructureProblem(st,ctx=ctx)
ts=[]
for i in range(50):
    info,s=StaticSolveProblemInfo(),StressInfo()
    t=time.perf_counter(); p.solveStatic(info); p.postProcCalcStress(s,info); ts.append(time.perf_counter()-t)
print(json.dumps({"c1_demo1_latency_us_median": float(np.median(ts[5:])*1e6), "min": float(min(ts)*1e6), "U_f": info.U_f.tolist(), "F_b": info.F_b.tolist()}))
PY

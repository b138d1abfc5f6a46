#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ingest_gpu.py -q -m gpu -s 2>&1 | grep -E "device ingest|passed|failed" ; 
cat > /tmp/ing.py <<'PY'
import sys; sys.path.insert(0, '.')
import sparsecholesky_b200 as sc
from sparsecholesky_b200 import synth
dims=(1024,1024,1)
perm=synth.nd_perm(*dims); ti,tj,tx,n=synth.laplacian_triplets(*dims,perm=perm)
for _ in range(3):
    A,ms=sc.triplet_to_csc_matrix_gpu(ti,tj,tx,n,return_ms=True); print("ms",ms)
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2o_launches_ingest.csv python /tmp/ing.py > gpurun_out/r2o_ing.log 2>&1; tail -3 gpurun_out/r2o_ing.log
python tools/summarize_launches.py gpurun_out/r2o_launches_ingest.csv gpurun_out/r2o_launches_ingest.txt "ingest 1024^2 x3" | head -20
cat > /tmp/solve.py <<'PY'
import sys, numpy as np; sys.path.insert(0, '.')
import sparsecholesky_b200 as sc
from sparsecholesky_b200 import synth
dims=(64,64,64)
perm=synth.nd_perm(*dims); ti,tj,tx,n=synth.laplacian_triplets(*dims,perm=perm)
A=sc.triplet_to_csc_matrix(ti,tj,tx,n); S=sc.schol(A)
import ctypes as C
from sparsecholesky_b200._lib import lib, check, ptr
info=C.c_int(0); plan=S.plan(0)
lbuf=np.empty(S.plan_stats(0)["lbuf_entries"])
check(lib().spchol_factor_lbuf(plan.h, ptr(A.x), ptr(lbuf), C.byref(info)))
b=np.ones(n); 
for _ in range(2): x=sc.solve(S,b)
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2o_launches_solve.csv python /tmp/solve.py > gpurun_out/r2o_solve.log 2>&1; tail -2 gpurun_out/r2o_solve.log
python tools/summarize_launches.py gpurun_out/r2o_launches_solve.csv gpurun_out/r2o_launches_solve.txt "64^3 factor + 2 solves" | grep -E "solve|TOTAL|kernel"

#!/bin/bash
mkdir -p gpurun_out
timeout 60 python bench_ingest.py > gpurun_out/r2z_bench_ingest.json 2> gpurun_out/r2z_bench_ingest.err; echo "rc=$?"; cat gpurun_out/r2z_bench_ingest.json; tail -2 gpurun_out/r2z_bench_ingest.err

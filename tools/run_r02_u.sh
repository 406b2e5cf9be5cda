#!/bin/bash
# source-level captures of the three kernels that bound the bench: fast_kernel (batched ingest), klt_kernel (sync), pnp
mkdir -p gpurun_out
O=gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
ncu --set full --clock-control none --import-source on -k regex:"fast_kernel" -s 2 -c 1 -f -o $O/u_fast \
    python tools/ingest_probe.py 3 > $O/u_fast.log 2>&1; echo "fast rc $?"
ncu --set full --clock-control none --import-source on -k regex:"klt_kernel" -s 8 -c 1 -f -o $O/u_klt_sync \
    python tools/trace_sync.py > $O/u_klt_sync.log 2>&1; echo "klt sync rc $?"
for r in u_fast u_klt_sync; do
  ncu -i $O/$r.ncu-rep --page source --csv > $O/$r.source.csv 2>/dev/null
  ncu -i $O/$r.ncu-rep --page raw --csv > $O/$r.raw.csv 2>/dev/null
  ls -la $O/$r.ncu-rep | awk '{print $5, $9}'
done

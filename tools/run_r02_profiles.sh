#!/bin/bash
# Round-2 evidence pass (one GPU): tests, the bench line of both arms, the ncu launch list of the bench command,
# ncu --set full captures of the ingest / batched KLT / PnP kernels (raw pages exported here), the two timelines.
mkdir -p gpurun_out
O=gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -2
python bench.py > $O/r02_bench.json 2> $O/r02_bench.err; echo "bench rc $?"; tail -2 $O/r02_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/r02_bench_ref.json 2> $O/r02_bench_ref.err; echo "ref rc $?"
# launch list of the bench command (short run: ncu serialises every launch)
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/r02_launches.csv \
    python bench.py --frames 40 --steps 1 --warmup 1 --no-cpu-baseline > $O/r02_launches_run.log 2>&1; echo "launch list rc $?"
# full captures
ncu --set full --clock-control none --import-source on -k regex:"fast_kernel|pyrdown" -s 5 -c 5 -f -o $O/r02_ingest_full \
    python tools/ingest_probe.py 4 > $O/r02_ingest_full.log 2>&1; echo "ingest full rc $?"
ncu --set full --clock-control none --import-source on -k regex:"klt_kernel" -s 2 -c 2 -f -o $O/r02_klt_batch_full \
    python tools/ingest_probe.py 4 > $O/r02_klt_batch_full.log 2>&1; echo "klt full rc $?"
ncu --set full --clock-control none --import-source on -k regex:"pnp_" -s 8 -c 8 -f -o $O/r02_pnp_full \
    python tools/trace_sync.py > $O/r02_pnp_full.log 2>&1; echo "pnp full rc $?"
ncu --set full --clock-control none --import-source on -k regex:"klt_" -s 8 -c 6 -f -o $O/r02_klt_sync_full \
    python tools/trace_sync.py > $O/r02_klt_sync_full.log 2>&1; echo "klt sync full rc $?"
for r in r02_ingest_full r02_klt_batch_full r02_pnp_full r02_klt_sync_full; do
  ncu -i $O/$r.ncu-rep --page raw --csv > $O/$r.raw.csv 2>/dev/null
  ls -la $O/$r.ncu-rep | awk '{print $5, $9}'
done
# timelines (SVO_TRACE: per-kernel start/stop of the library's own event trace)
SVO_TRACE=$O/r02_timeline_stream.csv python tools/trace_run.py > $O/r02_timeline_stream.log 2>&1; echo "trace stream rc $?"
SVO_TRACE=$O/r02_timeline_sync.csv python tools/trace_sync.py > $O/r02_timeline_sync.log 2>&1; echo "trace sync rc $?"
(pkg-config --modversion opencv4; echo "pkg-config rc $?"; python -c "import cv2; print(cv2.__version__)"; ls /usr/include/opencv4 /usr/local/include/opencv4) > $O/r02_opencv_probe.txt 2>&1
compute-sanitizer --version > $O/r02_sanitizer_probe.txt 2>&1; echo "sanitizer rc $?" >> $O/r02_sanitizer_probe.txt
python -c "
import json; d=json.load(open('$O/r02_bench.json')); print(d['value'], d['e2e']['value'], d['e2e'].get('sync_per_frame_value'), d['batch64']); print(d['stages']); print(d['roofline'])"

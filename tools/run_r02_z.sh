#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
bash tools/quick_bench.sh new_defaults
SVO_LIB_PATH=tools/libsvo_w9.so bash tools/quick_bench.sh warps9
SVO_LIB_PATH=tools/libsvo_old.so bash tools/quick_bench.sh old_head
bash tools/quick_bench.sh new_defaults_again
for lib in stereo_visual_odometry_b200/libsvo_b200.so tools/libsvo_old.so; do
SVO_LIB_PATH=$lib ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:"fast_kernel|pyrdown" -s 5 -c 5 --csv --log-file $O/z_l.csv python tools/ingest_probe.py 3 > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("$O/z_l.csv")) if len(r)>5 and r[0].isdigit()]
print("$lib", [(r[4][:22], r[-3], r[-1]) for r in rows])
PY
done

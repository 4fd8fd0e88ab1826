#!/bin/bash
# Round-2 evidence for the bench's roofline block (run on the GPU box, after the programs ran clean without ncu):
#  1. launch list of one solve (gpu__time_duration per launch) -> gpurun_out/r2_solve_launches.csv + per-kernel summary
#  2. ncu --set full of the finest-level float tile kernel (Chebyshev step, mode 3, and the other modes that follow)
mkdir -p gpurun_out
python tools/prof_tile.py > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start on -c 6000 --csv \
    --log-file gpurun_out/r2_solve_launches_all.csv python tools/prof_tile.py > gpurun_out/r2_solve_launches.log 2>&1
python - <<'PY'
import csv
rows = [r for r in csv.reader(open("gpurun_out/r2_solve_launches_all.csv")) if len(r) > 10]
hdr, body = rows[0], rows[1:]
ki = hdr.index("Kernel Name")
# the launches of solve_cg: from the first k_reduce of the solve (after set-up, the two float matvecs and the double one)
first = [i for i, r in enumerate(body) if "solve_cg" in r[ki]]
start = first[0] if first else 0
w = csv.writer(open("gpurun_out/r2_solve_launches.csv", "w"))
w.writerow(hdr)
w.writerows(body[start:])
PY
python tools/launch_summary.py gpurun_out/r2_solve_launches.csv 30 > gpurun_out/r2_solve_launch_summary.txt
head -14 gpurun_out/r2_solve_launch_summary.txt
bash tools/prof_tile.sh k_apply_tile_f32x2 2178 6 r2_tile_f32x2 "<4, 3>"
ncu -i gpurun_out/r2_tile_f32x2.ncu-rep --page raw --csv > gpurun_out/r2_tile_f32x2_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/r2_tile_f32x2_raw.csv > gpurun_out/r2_tile_f32x2_ncu_full_summary.txt
grep -A6 "Kernel Name" gpurun_out/r2_tile_f32x2_ncu_full_summary.txt | head -40

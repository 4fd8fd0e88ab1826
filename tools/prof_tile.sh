#!/bin/bash
# usage (on the GPU box): tools/prof_tile.sh <kernel base name> <grid size of the launches wanted> <count> <output name> [<substring of the demangled name, e.g. '<4, 3>'>]
# pass 1 lists all matching launches (cheap), pass 2 captures `count` of them with --set full
set -e
RE=$1; GRID=$2; CNT=$3; OUT=$4; SUB=${5:-}
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:$RE -c 5000 --csv --log-file gpurun_out/${OUT}_list.csv python tools/prof_tile.py > /dev/null 2>&1
SKIP=$(python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/${OUT}_list.csv")) if len(r)>10]
h=rows[0]; gi=h.index("Grid Size"); ki=h.index("Kernel Name"); ids=[]
for n,r in enumerate(rows[1:]):
    if r[gi].replace(" ","").startswith("($GRID,") and "$SUB" in r[ki]:
        ids.append(n)
# the last launches are those of solve_cg; take the ones in the middle of the solve
print(ids[len(ids)//2] if ids else 0)
PY
)
echo "skip $SKIP"
ncu --set full --clock-control none --import-source on -k regex:$RE --launch-skip $SKIP -c $CNT -f -o gpurun_out/$OUT python tools/prof_tile.py > gpurun_out/${OUT}.log 2>&1
tail -2 gpurun_out/${OUT}.log

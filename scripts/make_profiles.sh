#!/bin/bash
# Turn one gpu_round.sh session (gpurun_out/<tag>) into the tracked summaries under profiles/ (round-prefixed).
# Usage: bash scripts/make_profiles.sh <tag> <round-prefix e.g. r01>
set -e
TAG=$1; R=$2; SRC=gpurun_out/$TAG; DST=profiles
mkdir -p $DST
cp $SRC/launches_bench.csv $DST/${R}_launches_bench.csv
python scripts/ncu_launches.py $SRC/launches_bench.csv > $DST/${R}_launches_summary.txt
for rep in $SRC/prof_*.ncu-rep; do
  name=$(basename $rep .ncu-rep | sed 's/^prof_//')
  { echo "# ncu --set full --clock-control none, $name (scripts/gpu_round.sh); summary by scripts/ncu_summary.py"; python scripts/ncu_summary.py $rep; } > $DST/${R}_ncu_$name.txt
  ncu -i $rep --page source --print-source cuda,sass --csv > /tmp/_src.csv 2>/dev/null || true
  { echo; echo "# hottest source lines (share of executed warp instructions / of stall samples)"; python scripts/ncu_lines.py /tmp/_src.csv 25; } >> $DST/${R}_ncu_$name.txt
done
for f in $SRC/bench_*.json; do cp $f $DST/${R}_$(basename $f); done
cp $SRC/parity.log $DST/${R}_parity.log
tail -3 $SRC/pytest_gpu.log > $DST/${R}_pytest_gpu.txt
python - <<PY
import csv, glob, io, json, os, subprocess
out = {}
for rep in glob.glob("$SRC/prof_*.ncu-rep"):
    name = os.path.basename(rep)[5:-8]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, u = rows[0], rows[1]
    tot = 0.0
    for v in rows[2:3]:
        for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = h.index(key); x = float(v[i]); unit = u[i]
            tot += x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
    key = name.replace("_paper", ".json:paper") if name.endswith("_paper") else name + ".json"
    out[key] = int(tot)
json.dump(out, open("$DST/traffic.json", "w"), indent=1, sort_keys=True)
print(out)
PY
ls $DST

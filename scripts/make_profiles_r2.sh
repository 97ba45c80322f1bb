#!/bin/bash
# Turn one gpu_final.sh session (gpurun_out/<tag>) into the tracked summaries under profiles/ (round-prefixed).
# Usage: bash scripts/make_profiles_r2.sh <tag> <prefix e.g. r02z>
set -e
TAG=$1; R=$2; SRC=gpurun_out/$TAG; DST=profiles
mkdir -p $DST
cp $SRC/launches_bench.csv $DST/${R}_launches_bench.csv
python scripts/ncu_launches.py $SRC/launches_bench.csv > $DST/${R}_launches_summary.txt
for rep in $SRC/prof_*.ncu-rep; do
  name=$(basename $rep .ncu-rep | sed 's/^prof_//')
  ncu -i $rep --page source --print-source cuda,sass --csv > /tmp/_src_$name.csv 2>/dev/null || true
  {
    echo "# ncu --set full --clock-control none --import-source on, $name (scripts/gpu_final.sh); summaries by scripts/ncu_*.py"
    n=$(ncu -i $rep --page raw --csv 2>/dev/null | tail -n +3 | wc -l)
    python scripts/ncu_summary.py $rep
    if [ "$n" -gt 1 ]; then echo; echo "# second kernel of the capture"; python scripts/ncu_summary.py $rep 1 || true; fi
    echo; echo "# hottest source lines (share of executed warp instructions / of stall samples)"; python scripts/ncu_lines.py /tmp/_src_$name.csv 25
    echo; echo "# executed warp instructions by region and class (scripts/ncu_regions.py)"; python scripts/ncu_regions.py /tmp/_src_$name.csv | grep -v " 0.0% |   0.0%   0.0%   0.0%   0.0%   0.0%   0.0%   0.0%"
    echo; echo "# opcode histogram (scripts/ncu_opcodes.py)"; python scripts/ncu_opcodes.py /tmp/_src_$name.csv | head -45
  } > $DST/${R}_ncu_$name.txt
done
for f in $SRC/bench_*.json; do grep '^{' $f | tail -1 > $DST/${R}_$(basename $f); done
tail -150 $SRC/pytest_gpu.log | grep -E "PASSED|FAILED|SKIPPED|passed|failed|rc=" > $DST/${R}_pytest_gpu.txt
cp $SRC/smoke.log $DST/${R}_smoke.txt
python - <<PY
import csv, glob, io, json, os, subprocess
out = {}
for rep in glob.glob("$SRC/prof_*.ncu-rep"):
    name = os.path.basename(rep)[5:-8]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, u = rows[0], rows[1]
    tot = 0.0
    for v in rows[2:]:                      # (paper capture: G-buffer kernel + post kernel = one frame)
        for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = h.index(key); x = float(v[i]); unit = u[i]
            tot += x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
    key = name.replace("_paper", ".json:paper") if name.endswith("_paper") else name + ".json"
    out[key] = int(tot)
json.dump(out, open("$DST/traffic.json", "w"), indent=1, sort_keys=True)
print(out)
PY
ls $DST | grep "^${R}_" | tr '\n' ' '

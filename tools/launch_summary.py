"""Per-kernel durations of the LAST of the three evaluations of every shape in an ncu launch list written by
tools/gpu_small_shapes.sh (one process, shapes in command-line order; an evaluation starts at its colstats kernel):
python tools/launch_summary.py gpurun_out/<tag>_launches.csv "N D prec" ..."""
import csv, sys

f, shapes = sys.argv[1], sys.argv[2:]
with open(f) as fh:
    lines = [l for l in fh if not l.startswith("==")]
evals = []
for row in csv.DictReader(lines):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(row["Metric Value"].replace(",", ""))
    u = row["Metric Unit"]
    v = v / 1e3 if u in ("nsecond", "ns") else (v * 1e3 if u in ("msecond", "ms") else v)
    name = row["Kernel Name"]
    if "colstats" in name:
        evals.append([])
    if evals:
        evals[-1].append((name[:64], v))
for k in range(2, len(evals), 3):
    label = shapes[k // 3] if k // 3 < len(shapes) else f"shape {k // 3}"
    print(f"{label}: {len(evals[k])} launches, {sum(v for _, v in evals[k]):.1f} us")
    for name, v in evals[k]:
        print(f"   {name:<64s} {v:9.1f} us")

"""Top source lines by warp-stall samples from an ncu source page.
usage: ncu -i X.ncu-rep --page source --csv --print-source sass,cuda > src.csv
       python tools/ncu_source_top.py src.csv [n]
Rows with a line number carry that line's aggregate; SASS rows under them are skipped."""
import collections, csv, glob, os, sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 24
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = {os.path.basename(p): open(p).read().splitlines() for p in glob.glob(os.path.join(root, "qwen3-tts-jetson_b200", "csrc", "*"))}
agg = collections.Counter()
cur = None
csv.field_size_limit(1 << 30)
for r in csv.reader(open(path)):
    if len(r) >= 2 and r[0] == "File Path":
        cur = os.path.basename(r[1])
        continue
    if len(r) < 7 or r[0] in ("", "Line No", "Function Name"):
        continue
    try:
        agg[(cur, int(r[0]))] += int(r[4])
    except ValueError:
        pass
total = sum(agg.values())
print(f"total samples {total}")
for (f, l), s in agg.most_common(top):
    text = src[f][l - 1].strip()[:100] if f in src and l <= len(src[f]) else ""
    print(f"{s:9d} {100.0 * s / max(total, 1):5.1f}%  {f}:{l}  {text}")

"""profiles/traffic.json from an `ncu --page raw --csv` dump: DRAM bytes per launch of the dominant kernel, tagged with the
hash of the kernel sources it was taken on (bench.py emits roofline.traffic only when that hash matches its own build).
python tools/make_traffic.py <workload> <kernel substring> <raw.csv> <capture name> [out.json]"""
import csv, json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from bench import source_hash
workload, kern, path, capture = sys.argv[1:5]
out = Path(sys.argv[5] if len(sys.argv) > 5 else "profiles/traffic.json")
rows = list(csv.reader(open(path)))
hdr, units = rows[0], rows[1]
k, r, w = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
best = None
for row in rows[2:]:
    if kern in row[k]:
        best = float(row[r].replace(",", "")) * scale[units[r]] + float(row[w].replace(",", "")) * scale[units[w]]
doc = json.loads(out.read_text()) if out.exists() else {}
doc[workload] = {"dram_bytes": best, "kernel": kern, "capture": capture, "source_hash": source_hash()}
out.write_text(json.dumps(doc, indent=1) + "\n")
print(doc[workload])

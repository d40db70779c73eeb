"""profiles/traffic.json from ncu --set full captures: dram__bytes_read.sum + dram__bytes_write.sum per launch of every
kernel in the given .ncu-rep files (the largest launch of a kernel wins: a step launches the selection twice, the second
pass over the flagged streams only).
   python tools/make_traffic.py out.json lk.ncu-rep sad.ncu-rep"""
import csv, json, subprocess, sys
out, reps = sys.argv[1], sys.argv[2:]
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
doc = {"_what": "dram__bytes_read.sum + dram__bytes_write.sum per launch from the ncu --set full captures named in 'source' (B200, the "
                "bench workloads: 512 LK streams of 640x480 / 4096 SAD pairs per launch); bench.py copies these into roofline.traffic",
       "source": [r.replace("gpurun_out/", "profiles/").replace("_full.ncu-rep", "_ncu_full_summary.txt") for r in reps]}
for rep in reps:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        name = r[ix["Kernel Name"]].split("(")[0].split("<")[0].replace("void ", "").replace("aof::", "").strip()
        tot = 0.0
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            tot += float(r[ix[m]].replace(",", "")) * UNIT.get(units[ix[m]], 1)
        doc[name] = max(int(tot), doc.get(name, 0))
json.dump(doc, open(out, "w"), indent=2)
print(json.dumps(doc, indent=1))

#!/usr/bin/env python
"""roofline.traffic from an ncu capture of THIS binary: reads an .ncu-rep taken with `ncu --set full -k regex:gemv_kernel`
over consecutive launches of the bench step (the three Llama-2-7B shapes in turn), takes dram__bytes_read.sum +
dram__bytes_write.sum per launch, and writes profiles/gemv_traffic.json with the hash of the kernel sources (the built .so is
not reproducible byte for byte) — bench.py reports the figure only while the sources it was built from have that hash.   python tools/measure_traffic.py gpurun_out/prof.ncu-rep [Q4_0]"""
import csv
import hashlib
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "llama.kotlin_b200", "libllamakotlin_cuda.so")


def to_bytes(v, unit):
    mul = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
    return float(v.replace(",", "")) * mul


def main():
    rep, qtype = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "Q4_0")
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ir, iw, it, ik = (hdr.index(k) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "Kernel Name"))
    per = [{"kernel": r[ik][:60], "bytes": to_bytes(r[ir], units[ir]) + to_bytes(r[iw], units[iw]), "duration": r[it] + " " + units[it]}
           for r in data if "gemv_kernel" in r[ik]]
    if not per:
        sys.exit("no gemv_kernel launch in the report")
    sys.path.insert(0, ROOT)
    from bench import kernel_sources_sha256
    rec = {qtype: sum(p["bytes"] for p in per) / len(per), "launches": per, "sources_sha256": kernel_sources_sha256(),
           "how": f"ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, mean over {len(per)} consecutive launches of the bench step ({os.path.basename(rep)})"}
    with open(os.path.join(ROOT, "profiles", "gemv_traffic.json"), "w") as f:
        json.dump(rec, f, indent=1)
    print(json.dumps(rec)[:400])


if __name__ == "__main__":
    main()

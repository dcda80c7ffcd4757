#!/usr/bin/env python
"""DRAM traffic of every kernel of ONE step, from an ncu report -> profiles/r02_traffic_lists.json
(what bench.py reports as roofline.traffic: measured, not typed in).

    gpurun -- 'python tools/list_prof.py --execs 3 > /dev/null &&
               ncu --set full --clock-control none -s <launches of 2 execs> -c <launches of 1 exec> \
                   -o gpurun_out/r02_list_step python tools/list_prof.py --execs 3'
    ncu -i gpurun_out/r02_list_step.ncu-rep --page raw --csv > gpurun_out/r02_list_step_raw.csv
    python tools/ncu_traffic.py gpurun_out/r02_list_step_raw.csv profiles/r02_traffic_lists.json \
        [atoms_warp_instr_per_s_peak]
"""
import csv
import json
import re
import sys


def num(x):
    return float(x.replace(",", "")) if x not in ("", "n/a") else 0.0


def scale(unit):
    return {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    kernels, total = {}, 0.0
    for r in rows[2:]:
        name = re.sub(r"<.*", "", r[col["Kernel Name"]].split("(")[0]).replace("void ", "").replace("pmb::", "").strip()
        rd = num(r[col["dram__bytes_read.sum"]]) * scale(units[col["dram__bytes_read.sum"]])
        wr = num(r[col["dram__bytes_write.sum"]]) * scale(units[col["dram__bytes_write.sum"]])
        k = kernels.setdefault(name, {"launches": 0, "dram_bytes": 0.0, "dram_read": 0.0, "dram_write": 0.0,
                                      "shared_atomic_instr": 0.0, "warp_instr": 0.0, "duration_us": 0.0})
        k["launches"] += 1
        k["dram_bytes"] += rd + wr
        k["dram_read"] += rd
        k["dram_write"] += wr
        k["shared_atomic_instr"] += num(r[col["smsp__inst_executed_op_shared_atom.sum"]])
        k["warp_instr"] += num(r[col["smsp__inst_executed.sum"]])
        k["duration_us"] += num(r[col["gpu__time_duration.sum"]]) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(
            units[col["gpu__time_duration.sum"]], 1.0)
        total += rd + wr
    out = {"source": "ncu --set full --clock-control none over every kernel of one step of tools/list_prof.py "
                     "(hg38, 20 M reads, NCC+MSCC, max_shift 1000, list path, 1 B200): dram__bytes_read.sum + "
                     "dram__bytes_write.sum per launch; %s" % sys.argv[1],
           "step_dram_bytes": total, "kernels": kernels}
    if len(sys.argv) > 3:
        out["atoms_warp_instr_per_s_peak"] = float(sys.argv[3])
        out["atoms_peak_source"] = ("profiles/r02_atom_probe.jsonl: shared-memory atomicAdd (ATOMS.POPC.INC), "
                                    "full warps on random bins, updates/s / 32")
    json.dump(out, open(sys.argv[2], "w"), indent=1)
    print(json.dumps({k: (v["dram_bytes"], v["launches"]) for k, v in kernels.items()}), total)


if __name__ == "__main__":
    main()

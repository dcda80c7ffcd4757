"""Measure the integer-pipe issue rates of the device (POPC, LOP3, SHF, IADD3, mixes).
Writes gpurun_out/pipe_probe.json.  The POPC rate is the roofline denominator of bench.py."""
import json
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pymasc_b200 import _shim  # noqa: E402

NAMES = {0: "popc", 1: "lop3", 2: "popc+lop3", 3: "shf", 4: "iadd", 5: "popc+imad",
         6: "shf+2lop3+popc+iadd"}


def sm_clock():
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm", "--format=csv,noheader,nounits"],
                             capture_output=True, text=True).stdout.split("\n")[0]
        return [int(x) for x in out.split(",")]
    except Exception:
        return None


def main():
    d = _shim.Device(0)
    res = {}
    for what, name in NAMES.items():
        ops, ms = d.pipe_probe(what)
        res[name] = {"ops_per_s": ops, "ms": ms, "per_clk_per_sm_at_1965MHz": ops / 148 / 1.965e9}
        print("%-22s %.3e thread-ops/s  (%.2f ms)  = %.1f /clk/SM @1965 MHz" %
              (name, ops, ms, ops / 148 / 1.965e9), flush=True)
    res["clocks_after"] = sm_clock()
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/pipe_probe.json", "w") as fh:
        json.dump(res, fh, indent=1)


if __name__ == "__main__":
    main()

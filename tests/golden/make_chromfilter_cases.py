#!/usr/bin/env python
"""tests/golden/chromfilter_cases.json: what the reference's filter_chroms (utils/calc.py:49-91)
selects for a list of include / exclude options.  Run in the build container (imports the
compiled reference from oracle/_ref); the fixture travels, the reference does not."""
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle", "_ref"))
from PyMaSC.utils.calc import filter_chroms  # noqa: E402

chroms = ["chr1", "chr2", "chr10", "chr11", "chrX", "chrY", "chrM", "chr1_random", "chrUn_gl000220", "scaffold_7"]
pats = ["chr*", "chr1*", "chr?", "*_random", "chrUn*", "chr[XY]", "scaffold_*", "chrM", "nomatch*", "*"]
rnd = random.Random(7)
cases = [None, [(True, ["chr?"])], [(False, ["*_random", "chrUn*"])], [(False, ["chr1*"]), (True, ["chr1_random"])],
         [(True, ["chr*"]), (False, ["chrM"])], [(True, ["chr1"]), (True, ["chr2"])], [(False, ["nomatch*"])],
         [(True, ["nomatch*"])]]
for _ in range(60):
    cases.append([(rnd.random() < 0.5, rnd.sample(pats, rnd.randint(1, 3))) for _ in range(rnd.randint(1, 4))])
out = [[f, sorted(filter_chroms(chroms, f))] for f in cases]
with open(os.path.join(HERE, "chromfilter_cases.json"), "w") as fh:
    json.dump({"chroms": chroms, "cases": out}, fh, indent=0)
print(len(out), "cases")

"""north_star: "the derived NCC/MSCC curves, NSC/RSC/VSN and estimated fragment/library length
must match within 1e-9 relative".  The drop-in calculator runs inside the reference package
(oracle/_ref, the compiled unmodified reference) and the reference's OWN stats.py consumes its
results; the statistics are compared with the reference's golden stats table
(tests/golden/ENCFF000RMB-test_stats.tab -> golden_encff000rmb.json)."""
import json
import os
import subprocess
import sys

import pytest

import helpers

RTOL = 1e-9
SCRIPT = os.path.join(helpers.ROOT, "tests", "stats_dropin_script.py")
HAVE_REF = os.path.isdir(os.path.join(helpers.ROOT, "oracle", "_ref", "PyMaSC"))


def run(mode):
    env = dict(os.environ)
    env.pop("PYMASC_B200_MIRROR_RESULTS", None)
    p = subprocess.run([sys.executable, SCRIPT, mode], capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 0, p.stderr[-2000:]
    return json.loads(p.stdout.strip().splitlines()[-1])


def check(got):
    g = helpers.load_golden_tables()["stats"]

    def close(x, name):
        want = float(g[name])
        assert abs(x - want) <= RTOL * abs(want), (name, x, want)

    assert got["forward_reads"] == int(g["Forward reads"]) and got["reverse_reads"] == int(g["Reverse reads"])
    n, m = got["ncc"], got["mscc"]
    # the NCC metrics are taken at the MSCC-estimated library length (stats.py:627-642)
    assert n["fragment_length"] == m["est_lib_len"] == int(g["Estimated library length"]) == 65
    assert n["fwhm"] == int(g["Estimated FWHM"]) and m["fwhm"] == int(g["Estimated MSCC FWHM"])
    close(n["cc_min"], "Minimum NCC")
    close(n["ccrl"], "NCC at read length")
    close(n["ccfl"], "NCC at estimated library length")
    close(n["nsc"], "Estimated NSC")
    close(n["rsc"], "Estimated RSC")
    close(n["vsn"], "Estimated VSN")
    close(m["cc_min"], "Minimum MSCC")
    close(m["ccrl"], "MSCC at read length")
    close(m["ccfl"], "MSCC at estimated library length")
    close(m["nsc"], "Estimated MSCC NSC")
    close(m["rsc"], "Estimated MSCC RSC")
    close(m["vsn"], "Estimated MSCC VSN")


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref (compiled reference) not built")
def test_reference_stats_on_dropin_results_match_golden_stats_tab():
    check(run("oracle"))


@pytest.mark.gpu
@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref (compiled reference) not built")
def test_reference_stats_on_device_results_match_golden_stats_tab():
    check(run("device"))

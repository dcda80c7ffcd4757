"""The drop-in at the reference's real seam: PyMaSC.handler.factory (oracle/_ref, unmodified)
with INTEGRATION.md section 1's swap applied builds the calculator through `create_calculator(config)`
(factory.py:108-136, :224-258) for NCC / BOTH / --skip-ncc; the `isinstance` assertion of the
reference's tests/unit/test_factory_architecture.py:216-217 holds, the results equal the vectors
generated from the compiled reference (tests/golden/ref_vectors.json), and the progress hook sees
the protocol of mscc.pyx:151-159, :317, :325."""
import json
import os
import subprocess
import sys

import pytest

import helpers

RTOL = 1e-9
SCRIPT = os.path.join(helpers.ROOT, "tests", "seam_script.py")
HAVE_REF = os.path.isdir(os.path.join(helpers.ROOT, "oracle", "_ref", "PyMaSC"))
L_CHR1 = 249250621


def run(mode, tmp_path):
    env = dict(os.environ)
    env.pop("PYMASC_B200_MIRROR_RESULTS", None)
    p = subprocess.run([sys.executable, SCRIPT, mode, str(tmp_path)], capture_output=True, text=True,
                       timeout=600, env=env)
    assert p.returncode == 0, p.stderr[-3000:]
    return json.loads(p.stdout.strip().splitlines()[-1])


def check(out):
    vec = helpers.load_ref_vectors()
    assert [out[k]["type"] for k in ("fixture_ncc", "fixture_both", "fixture_mscc_only")] == \
        ["NCCGenomeWideResult", "BothGenomeWideResult", "BothGenomeWideResult"]
    for label in ("fixture_ncc", "fixture_both", "fixture_mscc_only"):
        got, want = out[label]["whole"], vec[label]
        for k in ("chroms", "mappable_chroms"):       # the committed vectors keep chr1, chr2 and chrM of the 86 references
            if k in want:
                assert set(want[k]) <= set(got[k])
                got[k] = {c: v for c, v in got[k].items() if c in want[k]}
        helpers.assert_snapshots_equal(got, want, rtol=RTOL)
    # progress: a hook was passed, so (mscc.pyx:140-159) every chromosome whose track is loaded
    # gets the mappability bar -- enable, reset with the chromosome's name and length, filled,
    # cleaned -- and the per-shift bar is disabled; chr1 (the one with reads) comes first
    assert [c[0] for c in out["fixture_ncc"]["progress"]] == ["disable_bar", "update", "clean"]
    calls = out["fixture_both"]["progress"]
    assert calls[:4] == [["enable_bar"], ["reset_progress", "@@@@@^" * 12, "^", " ", "chr1", L_CHR1],
                         ["update", L_CHR1], ["clean"]]
    assert ["disable_bar"] in calls and ["update", 300] in calls
    assert not any(c[0] == "reset_progress" and c[2] == "X" for c in calls)      # no per-shift bar through a hook
    assert calls[-1] == ["clean"]


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref (compiled reference) not built")
def test_dropin_through_the_reference_factory(tmp_path):
    check(run("oracle", tmp_path))


@pytest.mark.gpu
@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref (compiled reference) not built")
def test_dropin_through_the_reference_factory_on_the_device(tmp_path):
    check(run("device", tmp_path))

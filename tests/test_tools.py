"""The measurement tools keep working: tools/bam_e2e.py (files -> result objects; here with the
CPU checker engine on a tiny genome) and tools/decode_bench.py (BAM decoder alone, own inflate vs
zlib) run as a user would run them and print their JSON line."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, tmp_path):
    r = subprocess.run([sys.executable] + args, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return json.loads(r.stdout.strip().splitlines()[-1]), r.stderr


def test_bam_e2e_runs_from_files_to_result_objects(tmp_path):
    line, _ = _run(["tools/bam_e2e.py", "--reads", "20000", "--chroms", "chr21,chr22", "--max-shift", "100",
                    "--repeat", "1", "--dir", str(tmp_path), "--oracle-engine"], tmp_path)
    assert line["unit"] == "s" and line["value"] > 0 and line["files"] == 1
    assert line["result"]["type"] == "BothGenomeWideResult" and line["result"]["chroms"] == 2
    assert line["result"]["forward_sum"] + line["result"]["reverse_sum"] <= line["reads_in_files"]
    assert line["read_len_estimated"] == 36
    assert set(line["stages_s"]) >= {"decode", "feed", "correlate", "wall"}


def test_decode_bench_times_both_inflate_routes(tmp_path):
    line, err = _run(["tools/decode_bench.py", "--reads", "200000", "--dir", str(tmp_path), "--repeat", "1"], tmp_path)
    assert line["accepted"] == line["records"] > 100000
    assert line["own_inflate_and_crc_s"] > 0 and line["zlib_s"] > 0
    assert err.count("[pmio bam]") == 2                 # the PMIO_TIMING stage line of each child


def test_inflate_bench_compares_the_decoders(tmp_path):
    line, _ = _run(["tools/inflate_bench.py", "--blocks", "4", "--reps", "1"], tmp_path)
    for kind in ("literal_heavy", "match_heavy"):
        assert all(line[kind][k] > 0 for k in ("own", "own_two_streams", "zlib", "crc32_clmul", "crc32_zlib"))
    assert line["literal_heavy"]["compression_ratio"] < 4 < line["match_heavy"]["compression_ratio"]

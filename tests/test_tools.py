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


def test_ncu_traffic_sums_dram_bytes_per_kernel(tmp_path):
    """tools/ncu_traffic.py turns `ncu --page raw --csv` into the measured-traffic file bench.py
    reports as roofline.traffic (no typed-in constants): units are honoured, launches summed."""
    hdr = ["ID", "Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "smsp__inst_executed_op_shared_atom.sum", "smsp__inst_executed.sum", "gpu__time_duration.sum"]
    units = ["", "", "Mbyte", "Kbyte", "inst", "inst", "us"]
    rows = [["0", "void pmb::list_fwd_kernel<1, 1>(pmb::ListParams)", "91.5", "3,684.0", "11063486", "279181269", "441.0"],
            ["1", "void pmb::list_rev_kernel<1>(pmb::ListParams)", "51.8", "149.76", "2540866", "113168051", "190.3"],
            ["2", "void pmb::list_fwd_kernel<1, 1>(pmb::ListParams)", "1.5", "0", "10", "100", "9.0"]]
    src = tmp_path / "raw.csv"
    with open(src, "w") as fh:
        for r in (hdr, units) + tuple(rows):
            fh.write(",".join('"%s"' % x for x in r) + "\n")
    out = tmp_path / "traffic.json"
    r = subprocess.run([sys.executable, "tools/ncu_traffic.py", str(src), str(out), "9.8e10"], cwd=ROOT,
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.load(open(out))
    fwd = d["kernels"]["list_fwd_kernel"]
    assert fwd["launches"] == 2 and abs(fwd["dram_bytes"] - (91.5e6 + 3684.0e3 + 1.5e6)) < 1
    assert fwd["shared_atomic_instr"] == 11063496 and abs(fwd["duration_us"] - 450.0) < 1e-9
    assert abs(d["step_dram_bytes"] - (fwd["dram_bytes"] + 51.8e6 + 149.76e3)) < 1
    assert d["atoms_warp_instr_per_s_peak"] == 9.8e10


def test_bench_blocks_generate_in_parallel_what_synth_generates(tmp_path):
    """tools/bench_blocks.generate (spawned workers, longest first) returns exactly the arrays
    pymasc_b200.synth produces in-process, in task order."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import numpy as np
    import bench_blocks
    from pymasc_b200 import synth
    tasks = [(20, 3000, 20260101, True), (21, 2500, 20260201, False), (23, 1000, 20260101, True)]
    got = bench_blocks.generate(tasks, procs=2)
    assert [g[0] for g in got] == [20, 21, 23]
    for (ci, n, seed, wt), (gi, gseed, fwd, rev, b, e) in zip(tasks, got):
        L = synth.HG38[ci][1]
        f2, r2 = synth.chip_reads(ci, L, n, 36, seed=seed)
        assert gseed == seed and np.array_equal(fwd, f2) and np.array_equal(rev, r2)
        if wt:
            b2, e2 = synth.mappability(ci, L)
            assert np.array_equal(b, b2) and np.array_equal(e, e2)
        else:
            assert b is None and e is None
    assert sum(bench_blocks.reads_per_chrom(20_000_000)) > 19_900_000


def test_ncu_summary_and_launch_list_tools(tmp_path):
    """tools/ncu_summary.py (selected metrics per launch of an `ncu --page raw --csv` export) and
    tools/ncu_launches.py (per-kernel totals and shares of a `--metrics gpu__time_duration.sum`
    launch list): the text summaries under profiles/ are their output."""
    hdr = ["ID", "Kernel Name", "gpu__time_duration.sum", "smsp__inst_executed.sum", "dram__bytes_read.sum"]
    units = ["", "", "us", "inst", "Mbyte"]
    rows = [["0", "void pmb::list_fwd_kernel<1, 1>(pmb::ListParams)", "300.0", "181017628", "92.0"],
            ["1", "pmb::list_finalize_kernel(pmb::ListParams)", "100.0", "347590", "1.6"]]
    raw = tmp_path / "raw.csv"
    with open(raw, "w") as fh:
        for r in (hdr, units) + tuple(rows):
            fh.write(",".join('"%s"' % x for x in r) + "\n")
    r = subprocess.run([sys.executable, "tools/ncu_summary.py", str(raw)], cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "== void list_fwd_kernel<1, 1>(ListParams)   (75.0 % of the captured kernels' time)" in r.stdout
    assert "smsp__inst_executed.sum" in r.stdout and "181017628 inst" in r.stdout
    assert "(25.0 %" in r.stdout

    launches = tmp_path / "launches.csv"
    with open(launches, "w") as fh:
        fh.write("==PROF== Connected to process 1\n")
        fh.write('"ID","Process ID","Process Name","Host Name","Kernel Name","Context","Stream","Block Size","Grid Size",'
                 '"Device","CC","Section Name","Metric Name","Metric Unit","Metric Value"\n')
        for i, (name, ns) in enumerate([("void list_fwd_kernel<1, 1>(ListParams)", "300000"),
                                        ("void list_rev_kernel<1>(ListParams)", "100,000"),
                                        ("void list_fwd_kernel<1, 1>(ListParams)", "100000"),
                                        ("void probe_kernel<0>(unsigned int *)", "999000")]):
            fh.write('"%d","1","python","h","%s","1","7","(512, 1, 1)","(296, 1, 1)","0","10.0","s",'
                     '"gpu__time_duration.sum","ns","%s"\n' % (i, name, ns))
    r = subprocess.run([sys.executable, "tools/ncu_launches.py", str(launches)], cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = {l.split()[0]: l.split() for l in r.stdout.splitlines() if not l.startswith("#")}
    assert lines["list_fwd_kernel<1,"][2:] == ["2", "400.0", "200.00", "0.800"]
    assert lines["list_rev_kernel<1>"][1:] == ["1", "100.0", "100.00", "0.200"]
    assert lines["probe_kernel<0>"][-1] == "-"


def test_list_sweep_caches_the_bench_workload_once(tmp_path):
    """tools/list_sweep.py: the workload every (library, options) child loads is the bench
    workload (synth.workload's seed), cut to a rank's share on request."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import numpy as np
    import list_sweep
    from pymasc_b200 import synth
    list_sweep.save_workload(str(tmp_path), 200000, ["chr21", "chr22"], 2)
    meta = json.load(open(tmp_path / "meta.json"))
    assert [m["name"] for m in meta] == ["chr21", "chr22"]
    total = sum(m["length"] for m in meta)
    assert meta[0]["hi"] == meta[0]["length"] and meta[0]["hi"] + meta[1]["hi"] == total // 2
    want = {w["name"]: w for w in synth.workload(n_reads=200000, chroms=["chr21", "chr22"])}
    # (read counts are allotted against the whole genome by the tool, as bench.py does)
    n21 = max(8, int(round(200000 * dict(synth.HG38)["chr21"] / float(sum(L for _, L in synth.HG38)))))
    fwd, rev = synth.chip_reads([n for n, _ in synth.HG38].index("chr21"), dict(synth.HG38)["chr21"], n21, 36)
    assert np.array_equal(np.load(tmp_path / "chr21_fwd.npy"), fwd)
    assert np.array_equal(np.load(tmp_path / "chr21_rev.npy"), rev)
    assert np.array_equal(np.load(tmp_path / "chr22_begins.npy"), want["chr22"]["begins"])

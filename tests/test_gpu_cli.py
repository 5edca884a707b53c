"""GPU tests of the two CLI drivers (drop-ins for the reference tools) and BASELINE config 5:
the GPU-projected db fed unchanged into the reference's pairwise_comp_optimized / query_pc_mat."""
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

from oracle_util import GOLDEN, load_kat, load_toy, load_toy_ref, ref_dir

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "metagenome_vector_sketches_b200", "bin")
PE = os.path.join(BIN, "project_everything")
SP = os.path.join(BIN, "standalone_projection")
REFV3 = os.path.join(ROOT, "oracle", "_ref", "v3")


@pytest.fixture(scope="module")
def toy_hash_file(tmp_path_factory):
    names, offsets, hashes = load_toy()
    p = tmp_path_factory.mktemp("toy") / "toy_hashes.txt"
    with open(p, "w") as f:
        for i, nme in enumerate(names):
            f.write(nme + ":" + "".join(" %d" % int(h) for h in hashes[int(offsets[i]):int(offsets[i + 1])]) + "\n")
    assert hashlib.sha256(open(p, "rb").read()).hexdigest().startswith("2eab37d3")    # SURVEY KAT-3 input
    return str(p)


def run(argv, **kw):
    return subprocess.run(argv, capture_output=True, text=True, timeout=300, **kw)


@pytest.mark.parametrize("tag,extra", [("d2048", ["-d", "2048"]), ("d2048", []), ("d100", ["-d", "100", "-t", "8"]),
                                       ("d2048_i16", ["--int16"]), ("d2048", ["--dimension", "2048", "--gpus", "1"])])
def test_sketch_cli_writes_reference_db(tmp_path, toy_hash_file, tag, extra):
    db = tmp_path / "db"
    db.mkdir()
    (db / "stale_file").write_text("must be wiped")                     # sketch() empties the folder (:244-249)
    r = run([PE, "sketch", toy_hash_file, str(db)] + extra)
    assert r.returncode == 0, r.stderr
    assert "Loaded 61 hash sets from" in r.stdout and "Time to compute all projected vectors:" in r.stdout
    assert r.stdout.count("Projected ") == 61 and "Projected DRR000001, vector dimension" in r.stdout
    ref = load_toy_ref(tag)
    assert sorted(os.listdir(db)) == ["dimension.txt", "dtype.txt", "vector_norms.txt", "vectors.bin"]
    assert hashlib.sha256((db / "vectors.bin").read_bytes()).hexdigest() == ref["vectors_sha256"]
    assert (db / "dimension.txt").read_text() == ref["dimension_txt"]
    assert (db / "dtype.txt").read_text() == ref["dtype_txt"]
    mine = (db / "vector_norms.txt").read_text().split("\n")[:-1]
    theirs = ref["norms_txt"].split("\n")[:-1]
    assert [l.rsplit(" ", 1)[0] for l in mine] == [l.rsplit(" ", 1)[0] for l in theirs]


def test_sketch_cli_readme_spelling_and_seed(tmp_path, toy_hash_file):
    db = tmp_path / "db2"
    r = run([PE, toy_hash_file, str(db) + "/", "-t", "8", "-d", "2048", "-s", "0", "--quiet"])
    assert r.returncode == 0, r.stderr
    assert "Projected " not in r.stdout
    ref = load_toy_ref("d2048")
    assert hashlib.sha256((db / "vectors.bin").read_bytes()).hexdigest() == ref["vectors_sha256"]
    r = run([PE, "sketch", toy_hash_file, str(db), "-s", "7", "--quiet"])
    assert r.returncode == 0
    assert hashlib.sha256((db / "vectors.bin").read_bytes()).hexdigest() != ref["vectors_sha256"]
    r = run([PE, "sketch", str(tmp_path / "missing.txt"), str(db)])
    assert r.returncode == 0 and "Error opening" in r.stderr            # reference returns 0 here


def test_standalone_cli_matches_reference_stdout(tmp_path):
    for case in load_kat()["standalone"]:
        p = tmp_path / "q.hashes"
        p.write_text("".join(l + "\n" for l in case["lines"]))
        r = run([SP, str(p), str(case["d"])])
        assert r.returncode == 0, r.stderr
        assert r.stdout == case["stdout"]
    r = run([SP, str(tmp_path / "nope"), "64"])
    assert r.returncode == 1 and "Error opening file" in r.stderr


def _neighbours(stdout):
    return [l for l in stdout.split("\n") if l.startswith(("Query:", "Top ")) or ". Neighbor:" in l]


def test_config5_topk_identical_through_reference_consumers(tmp_path, toy_hash_file):
    """BASELINE config 5: our db -> reference pairwise_comp_optimized -> reference query_pc_mat."""
    pw, q = os.path.join(REFV3, "pairwise_comp_optimized"), os.path.join(REFV3, "query_pc_mat")
    if not (os.path.exists(pw) and os.path.exists(q)):
        pytest.skip("oracle/_ref/v3 consumers not prebuilt")
    names, _, _ = load_toy()

    def pipeline(db, tag, query_lines):
        mat = tmp_path / ("mat_" + tag)
        mat.mkdir()
        r = run([pw, "--db", db, "--output_folder", str(mat) + "/", "--max_memory_gb", "12", "--num_threads", "8",
                 "--num_shards", "1", "--shard_idx", "0"])
        assert r.returncode == 0, r.stderr
        qf = tmp_path / ("q_%s.txt" % tag)
        qf.write_text("".join(x + "\n" for x in query_lines))
        r = run([q, "--matrix", str(mat) + "/", "--db", db, "--query_file", str(qf), "--top", "10", "--print"])
        assert r.returncode == 0, r.stderr
        return _neighbours(r.stdout)

    ours = tmp_path / "db_ours"
    assert run([PE, "sketch", toy_hash_file, str(ours), "-d", "2048", "--quiet"]).returncode == 0
    golden = json.load(open(os.path.join(GOLDEN, "pairwise_top10.json")))
    got = pipeline(str(ours) + "/", "ours2", ["DRR000821", "DRR000837"])
    assert got == _neighbours(golden["stdout"])                          # KAT-4
    if ref_dir() is not None:                                            # all 61 queries vs the reference's own db
        theirs = tmp_path / "db_ref"
        r = run([os.path.join(ref_dir(), "project_everything"), "sketch", toy_hash_file, str(theirs), "-d", "2048"])
        assert r.returncode == 0
        assert (theirs / "vectors.bin").read_bytes() == (ours / "vectors.bin").read_bytes()
        assert pipeline(str(ours) + "/", "oursall", names) == pipeline(str(theirs) + "/", "refall", names)


def test_reference_driver_linked_against_our_shim(tmp_path, toy_hash_file):
    """INTEGRATION.md option 2: the reference's unmodified project_everything.cpp + our
    transform_set_into_vector shim (prebuilt by oracle/Makefile) writes the reference's vectors.bin."""
    exe = os.path.join(REFV3, "project_everything_b200shim")
    if not os.path.exists(exe):
        pytest.skip("shim-linked reference driver not prebuilt")
    db = tmp_path / "db_shim"
    r = run([exe, "sketch", toy_hash_file, str(db), "-d", "2048"], env=dict(os.environ, OMP_NUM_THREADS="4"))
    assert r.returncode == 0, r.stderr
    ref = load_toy_ref("d2048")
    assert hashlib.sha256((db / "vectors.bin").read_bytes()).hexdigest() == ref["vectors_sha256"]
    assert (db / "vector_norms.txt").read_text() == ref["norms_txt"] or ref_dir() != REFV3

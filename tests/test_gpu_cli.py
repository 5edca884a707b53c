"""GPU tests of the two CLI drivers (drop-ins for the reference tools) and BASELINE config 5:
the GPU-projected db fed unchanged into the reference's pairwise_comp_optimized / query_pc_mat."""
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

from oracle_util import (GOLDEN, load_kat, load_toy, load_toy_ref, ref_dir, golden_host_has_same_rsqrt,
                         write_family_hash_file)

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
    return subprocess.run(argv, capture_output=True, text=True, timeout=600, **kw)


DB_FILES = ("vectors.bin", "vector_norms.txt", "dimension.txt", "dtype.txt")


def reference_sketch(hash_file, db, extra):
    """The reference binary on THIS host (v3 build travels with the repo), or None."""
    rd = ref_dir()
    if rd is None:
        return None
    r = run([os.path.join(rd, "project_everything"), "sketch", hash_file, str(db)] + extra)
    assert r.returncode == 0, r.stderr
    return db


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
    if golden_host_has_same_rsqrt(int(ref["dimension_txt"])):           # committed file, printed on the build host's CPU
        assert (db / "vector_norms.txt").read_text() == ref["norms_txt"]    # 61/61 lines byte for byte
    live = reference_sketch(toy_hash_file, tmp_path / "db_live", [a for a in extra if a not in ("--gpus", "1")])
    if live is not None:                                                 # the reference binary on this very host
        for f in DB_FILES:
            assert (db / f).read_bytes() == (live / f).read_bytes(), f
    else:
        assert golden_host_has_same_rsqrt(int(ref["dimension_txt"])), "no pin for vector_norms.txt on this host"


@pytest.mark.parametrize("extra", [["-d", "2048"], ["-d", "1000", "--devices", "0,0"], ["-d", "2048", "--int16", "--devices", "0,0,0"]])
def test_sketch_cli_db_identical_to_live_reference_10k_samples(tmp_path, extra):
    """>= 10 000 samples in families, ragged sizes: every file of db/ byte-identical to what the reference binary
    writes on this host -- including vector_norms.txt (reference :328-330) -- also when the batch is fanned out over
    several contexts (`--devices 0,0`: the --gpus N code path on a 1-GPU box)."""
    hf = tmp_path / "fam.txt"
    write_family_hash_file(str(hf), 10000, seed=21, mean_size=100)
    ref_extra = [a for i, a in enumerate(extra) if a != "--devices" and (i == 0 or extra[i - 1] != "--devices")]
    live = reference_sketch(str(hf), tmp_path / "db_ref", ref_extra)
    if live is None:
        pytest.skip("oracle/_ref not prebuilt for this host")
    db = tmp_path / "db"
    r = run([PE, "sketch", str(hf), str(db), "--quiet"] + extra)
    assert r.returncode == 0, r.stderr
    for f in DB_FILES:
        assert (db / f).read_bytes() == (live / f).read_bytes(), f
    assert len((db / "vector_norms.txt").read_text().split("\n")) == 10001
    # the other norms mode is a documented deviation: same names, values within print precision
    r = run([PE, "sketch", str(hf), str(db), "--quiet", "--norms", "exact"] + extra)
    assert r.returncode == 0
    assert (db / "vectors.bin").read_bytes() == (live / "vectors.bin").read_bytes()
    a, b = (db / "vector_norms.txt").read_text().split("\n"), (live / "vector_norms.txt").read_text().split("\n")
    differ = sum(x != y for x, y in zip(a, b))
    assert 0 < differ < 0.2 * len(a)                                    # which is why "reference" is the default


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


def test_sketch_cli_from_binary_hash_file_writes_the_same_db(tmp_path, toy_hash_file):
    """`pack` once, then sketch from the binary file (no text parse): the db is byte-identical to the one sketched from
    the text file, at two dimensions, int16 and over two shards."""
    packed = tmp_path / "toy.u64"
    r = run([PE, "pack", toy_hash_file, str(packed)])
    assert r.returncode == 0 and "Packed 61 samples, 261591 hashes" in r.stdout
    for extra in (["-d", "2048"], ["-d", "100", "--int16"], ["--devices", "0,0"]):
        a, b = tmp_path / "db_text", tmp_path / "db_bin"
        ra = run([PE, "sketch", toy_hash_file, str(a), "--quiet"] + extra)
        rb = run([PE, "sketch", str(packed), str(b), "--quiet"] + extra)
        assert ra.returncode == 0 and rb.returncode == 0, (ra.stderr, rb.stderr)
        for f in DB_FILES:
            assert (a / f).read_bytes() == (b / f).read_bytes(), (extra, f)
    ref = load_toy_ref("d2048")
    assert hashlib.sha256((tmp_path / "db_bin" / "vectors.bin").read_bytes()).hexdigest() == ref["vectors_sha256"]


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
        for f in DB_FILES:
            assert (theirs / f).read_bytes() == (ours / f).read_bytes(), f
        assert pipeline(str(ours) + "/", "oursall", names) == pipeline(str(theirs) + "/", "refall", names)


def test_config5_topk_identical_at_2500_samples(tmp_path):
    """BASELINE config 5 above toy size: 2 500 samples in families of 10 (several kept neighbours per row).  The
    reference's db and ours go through the reference's pairwise_comp_optimized and query_pc_mat unchanged; the
    top-10 list of EVERY row must be identical (north_star), which needs vectors.bin and the printed norms equal."""
    pw, q = os.path.join(REFV3, "pairwise_comp_optimized"), os.path.join(REFV3, "query_pc_mat")
    if not (os.path.exists(pw) and os.path.exists(q)) or ref_dir() is None:
        pytest.skip("oracle/_ref consumers not prebuilt")
    hf = tmp_path / "fam.txt"
    names, _, _ = write_family_hash_file(str(hf), 2500, seed=33, mean_size=400)
    theirs = reference_sketch(str(hf), tmp_path / "db_ref", ["-d", "2048"])
    ours = tmp_path / "db_ours"
    assert run([PE, "sketch", str(hf), str(ours), "-d", "2048", "--quiet"]).returncode == 0
    for f in DB_FILES:
        assert (theirs / f).read_bytes() == (ours / f).read_bytes(), f
    qf = tmp_path / "q.txt"
    qf.write_text("".join(x + "\n" for x in names))
    lists = {}
    for tag, db in (("ref", theirs), ("ours", ours)):
        mat = tmp_path / ("mat_" + tag)
        mat.mkdir()
        r = run([pw, "--db", str(db) + "/", "--output_folder", str(mat) + "/", "--max_memory_gb", "2", "--num_threads", "8",
                 "--num_shards", "1", "--shard_idx", "0"])
        assert r.returncode == 0, r.stderr
        r = run([q, "--matrix", str(mat) + "/", "--db", str(db) + "/", "--query_file", str(qf), "--top", "10", "--print"])
        assert r.returncode == 0, r.stderr
        lists[tag] = _neighbours(r.stdout)
    assert lists["ours"] == lists["ref"]
    assert sum(l.startswith("Query:") for l in lists["ours"]) == 2500
    assert sum(". Neighbor:" in l for l in lists["ours"]) > 3 * 2500        # families: several neighbours per row
    # and through OUR all-pairs tool on our db: same shard files, hence the same lists
    if os.path.exists(PWB):
        mat = tmp_path / "mat_b200"
        mat.mkdir()
        r = run([PWB, "--db", str(ours) + "/", "--output_folder", str(mat) + "/", "--max_memory_gb", "2", "--num_threads", "8",
                 "--num_shards", "1", "--shard_idx", "0"])
        assert r.returncode == 0, r.stderr
        for f in SHARD_FILES:
            assert (mat / "shard_0" / f).read_bytes() == (tmp_path / "mat_ref" / "shard_0" / f).read_bytes(), f


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
    # the norms here are printed by the reference's own main(): the file the reference binary writes on this host
    live = reference_sketch(toy_hash_file, tmp_path / "db_live", ["-d", "2048"])
    assert live is not None
    for f in DB_FILES:
        assert (db / f).read_bytes() == (live / f).read_bytes(), f


@pytest.mark.parametrize("threads,serial", [(1, False), (24, False), (8, True)])
def test_shim_called_concurrently_from_openmp_threads(threads, serial):
    """The a4 sample loop (project_everything.cpp:287-298) over in-memory sets: our replacement of
    transform_set_into_vector, one context per calling thread, against the reference's function."""
    ours, theirs = os.path.join(REFV3, "a4_loop_b200shim"), os.path.join(REFV3, "a4_loop_ref")
    if not (os.path.exists(ours) and os.path.exists(theirs)):
        pytest.skip("a4 loop harness not prebuilt")
    for n, h, d in ((70, 3000, 2048), (33, 1, 100), (9, 40000, 4096)):
        env = dict(os.environ, OMP_NUM_THREADS=str(threads))
        if serial:
            env["MVS_SHIM_SERIAL"] = "1"
        a = run([ours, str(n), str(h), str(d), "2"], env=env)
        b = run([theirs, str(n), str(h), str(d), "1"], env=dict(os.environ, OMP_NUM_THREADS="8"))
        assert a.returncode == 0 and b.returncode == 0, a.stderr + b.stderr
        ja, jb = json.loads(a.stdout.strip().splitlines()[-1]), json.loads(b.stdout.strip().splitlines()[-1])
        assert ja["checksum"] == jb["checksum"] and ja["hashes"] == jb["hashes"] == n * h


PWB = os.path.join(BIN, "pairwise_comp_b200")
SHARD_FILES = ("matrix.bin", "row_index.bin", "neighbor_start.bin")


def _synthetic_db(path, n=700, d=256, seed=5):
    """A db/ with families of similar vectors (so that most rows have several neighbours), written in the
    reference's layout with the norms printed like sketch() prints them."""
    rng = np.random.default_rng(seed)
    fam = np.rint(rng.normal(0, 120, ((n + 9) // 10, d)))
    v = (fam[np.arange(n) // 10] + np.rint(rng.normal(0, 70, (n, d)))).astype(np.int32)
    v[13] = 0                                                           # a zero vector: no neighbours at all, not even itself
    os.makedirs(path, exist_ok=True)
    v.tofile(os.path.join(path, "vectors.bin"))
    open(os.path.join(path, "dimension.txt"), "w").write("%d\n" % d)
    open(os.path.join(path, "dtype.txt"), "w").write("int32\n")
    norm = np.sqrt((v.astype(np.float64) ** 2).sum(axis=1) / d)
    with open(os.path.join(path, "vector_norms.txt"), "w") as f:
        for i, x in enumerate(norm):
            f.write("s%04d %g\n" % (i, x))
    return path + "/"


@pytest.mark.parametrize("case", ["toy_one_chunk", "toy_chunks_of_16", "synthetic_shard_1_of_3", "synthetic_small_chunks"])
def test_pairwise_cli_writes_the_reference_shard_files(tmp_path, toy_hash_file, case):
    """SURVEY 8f rank 4 end to end: our GPU pairwise CLI against the reference binary on the same db and the same
    arguments -- matrix.bin / row_index.bin / neighbor_start.bin byte for byte (the row order inside them is the
    iteration order of an unordered_map fed in the reference's block order, which depends on --max_memory_gb)."""
    ref = os.path.join(REFV3, "pairwise_comp_optimized")
    if not (os.path.exists(ref) and os.path.exists(PWB)):
        pytest.skip("reference pairwise binary / pairwise_comp_b200 not prebuilt")
    if case.startswith("toy"):
        db = tmp_path / "db"
        assert run([PE, "sketch", toy_hash_file, str(db), "-d", "2048", "--quiet"]).returncode == 0
        db = str(db) + "/"
        # chunk = max_bytes / (4 d)^2: 12 GB -> 192 vectors (one block); 1 GB -> 16 vectors (4 x 4 blocks of the 61)
        mem, shards, idx = ("12", "1", "0") if case == "toy_one_chunk" else ("1", "1", "0")
    else:
        db = _synthetic_db(str(tmp_path / "sdb"))
        # d = 256: (4 d)^2 = 2^20 -> 0.1 GB = 102 vectors per chunk, 0.02 GB = 20 per chunk
        mem, shards, idx = ("0.1", "3", "1") if case == "synthetic_shard_1_of_3" else ("0.02", "2", "0")
    outs = {}
    for tag, exe in (("ref", ref), ("ours", PWB)):
        out = tmp_path / ("mat_" + tag)
        out.mkdir()
        r = run([exe, "--db", db, "--output_folder", str(out) + "/", "--max_memory_gb", mem, "--num_threads", "4",
                 "--num_shards", shards, "--shard_idx", idx])
        assert r.returncode == 0, r.stdout + r.stderr
        outs[tag] = {f: (out / ("shard_" + idx) / f).read_bytes() for f in SHARD_FILES}
        assert "Total computation time" in r.stdout
    for f in SHARD_FILES:
        assert len(outs["ref"][f]) > 8 and outs["ours"][f] == outs["ref"][f], (case, f)


def test_pairwise_cli_feeds_query_pc_mat_kat4(tmp_path, toy_hash_file):
    """Our projection -> our pairwise -> the reference's query_pc_mat gives KAT-4's top-10 lists."""
    q = os.path.join(REFV3, "query_pc_mat")
    if not (os.path.exists(q) and os.path.exists(PWB)):
        pytest.skip("query_pc_mat / pairwise_comp_b200 not prebuilt")
    db = tmp_path / "db"
    assert run([PE, "sketch", toy_hash_file, str(db), "-d", "2048", "--quiet"]).returncode == 0
    mat = tmp_path / "mat"
    mat.mkdir()
    r = run([PWB, "--db", str(db) + "/", "--output_folder", str(mat) + "/", "--max_memory_gb", "12", "--num_threads", "1",
             "--num_shards", "1", "--shard_idx", "0"])
    assert r.returncode == 0, r.stderr
    qf = tmp_path / "q.txt"
    qf.write_text("DRR000821\nDRR000837\n")
    r = run([q, "--matrix", str(mat) + "/", "--db", str(db) + "/", "--query_file", str(qf), "--top", "10", "--print"])
    assert r.returncode == 0, r.stderr
    golden = json.load(open(os.path.join(GOLDEN, "pairwise_top10.json")))
    assert _neighbours(r.stdout) == _neighbours(golden["stdout"])


def test_pairwise_cli_accepts_an_int16_db(tmp_path, toy_hash_file):
    """`sketch --int16` writes a db our own all-pairs tool must be able to read (the reference routes such dbs to its
    older 16-bit tool): the rows are widened on load, so without saturation the shard files equal the int32 db's."""
    if not os.path.exists(PWB):
        pytest.skip("pairwise_comp_b200 not prebuilt")
    outs = {}
    for tag, extra in (("i32", []), ("i16", ["--int16"])):
        db = tmp_path / ("db_" + tag)
        assert run([PE, "sketch", toy_hash_file, str(db), "-d", "2048", "--quiet"] + extra).returncode == 0
        mat = tmp_path / ("mat_" + tag)
        mat.mkdir()
        r = run([PWB, "--db", str(db) + "/", "--output_folder", str(mat) + "/", "--max_memory_gb", "12", "--num_threads", "1",
                 "--num_shards", "1", "--shard_idx", "0"])
        assert r.returncode == 0, r.stderr
        outs[tag] = {f: (mat / "shard_0" / f).read_bytes() for f in SHARD_FILES}
    assert (tmp_path / "db_i16" / "dtype.txt").read_text() == "int16\n"
    assert outs["i16"] == outs["i32"] and len(outs["i32"]["matrix.bin"]) > 8


def test_pairwise_cli_grammar(tmp_path):
    if not os.path.exists(PWB):
        pytest.skip("pairwise_comp_b200 not prebuilt")
    assert run([PWB]).returncode == 1                                    # missing required options: usage, exit 1
    assert run([PWB, "--help"]).returncode == 0
    r = run([PWB, "--db", str(tmp_path) + "/", "--output_folder", str(tmp_path) + "/", "--max_memory_gb", "1", "--num_threads", "1",
             "--num_shards", "1", "--shard_idx", "0"])
    assert r.returncode == 1 and "vector_norms.txt" in r.stderr          # reference :858-861

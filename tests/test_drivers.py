"""CPU tests of the host-side drivers (parser, db writer, convert, partition, CLI grammar).
None of these needs a GPU: they exercise the C++ code above the projection call."""
import ctypes as C
import hashlib
import os
import subprocess

import numpy as np
import pytest

import metagenome_vector_sketches_b200 as mvs
from metagenome_vector_sketches_b200 import _lib
from oracle_util import (load_kat, load_toy, load_toy_ref, parse_line_like_reference, oracle_set, ref_dir,
                         golden_host_has_same_rsqrt, oracle_reference_norms, write_family_hash_file)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "metagenome_vector_sketches_b200", "bin")


def parse_file(path, named):
    lib = mvs.load_library()
    hb = C.POINTER(_lib.HashBatch)()
    rc = lib.mvs_parse_hash_file(path.encode(), int(named), C.byref(hb))
    if rc != 0:
        return rc, None, None
    b = hb.contents
    n = b.n_samples
    off = np.ctypeslib.as_array(b.offsets, shape=(n + 1,)).copy()
    hashes = np.ctypeslib.as_array(b.hashes, shape=(max(1, int(b.n_hashes)),))[: int(b.n_hashes)].copy()
    names = [b.names[i].decode() for i in range(n)] if named else None
    lib.mvs_free_hash_batch(hb)
    # the order inside a set is free (first occurrences in file order; the projection is a sum):
    # return every set sorted, after checking it holds no repeated value
    sets = [hashes[int(off[i]):int(off[i + 1])] for i in range(n)]
    for x in sets:
        assert len(np.unique(x)) == len(x)
    return 0, names, [np.sort(x) for x in sets]


def test_parser_matches_reference_edge_cases(tmp_path):
    kat = load_kat()["sketch_parser"]
    p = tmp_path / "edge.txt"
    p.write_bytes(kat["text"].encode())
    rc, names, sets = parse_file(str(p), True)
    assert rc == 0
    want_names, want_sets = [], []
    for line in kat["text"].split("\n"):
        if ":" not in line:
            continue
        want_names.append(line[: line.index(":")])
        want_sets.append(sorted(set(parse_line_like_reference(line[line.index(":") + 1:]))))
    assert names == want_names
    assert [s.tolist() for s in sets] == want_sets
    # ... and the sets reproduce the reference's vectors.bin for that file (golden, d=8)
    got = np.concatenate([oracle_set(s, kat["d"]) for s in sets])
    assert np.array_equal(got, np.array(kat["vectors"], dtype=np.int32))
    # norms file of the reference lists the same names in the same order
    assert [l.rsplit(" ", 1)[0] for l in kat["norms_txt"].split("\n")[:-1]] == want_names


def test_parser_standalone_mode_lines(tmp_path):
    p = tmp_path / "q.hashes"
    p.write_text("1 2 3\n\n5 5 5\nx 9\n7")
    rc, names, sets = parse_file(str(p), False)
    assert rc == 0 and names is None
    assert [s.tolist() for s in sets] == [[1, 2, 3], [], [5], [], [7]]
    p.write_text("")
    assert parse_file(str(p), False)[2] == []
    assert parse_file(str(tmp_path / "missing"), True)[0] == -6      # MVS_ERR_IO


def test_parser_swar_fuzz_against_stream_semantics(tmp_path):
    """The 8-digits-at-a-time path (mvs_driver.cpp parse_tokens) against the character-by-character
    model of `iss >> unsigned long`: every digit count 1..22 (overflow at 20+), leading zeros, signs,
    junk that ends a line, duplicates / unsorted lines, a last line without newline that ends exactly
    at the end of the mapping (no 8-byte load may cross it)."""
    rng = np.random.default_rng(7)
    lines = []
    for k in range(400):
        toks = []
        for _ in range(int(rng.integers(0, 40))):
            nd = int(rng.integers(1, 23))
            t = "".join(rng.choice(list("0123456789"), size=nd))
            r = rng.random()
            if r < 0.05:
                t = "-" + t
            elif r < 0.1:
                t = "+" + t
            elif r < 0.12:
                t = t + "x"
            elif r < 0.14:
                t = "18446744073709551615"
            elif r < 0.16:
                t = "18446744073709551616"
            elif r < 0.2 and toks:
                t = toks[int(rng.integers(0, len(toks)))]               # duplicate
            toks.append(t)
        sep = [" ", "  ", "\t", " \r"][k % 4]
        lines.append("s%d:" % k + sep.join([""] + toks))
    for tail in ("\n", "", " 12345678", " 1234567", " 123456789012345678901"):
        text = "\n".join(lines) + tail
        p = tmp_path / "fuzz.txt"
        p.write_bytes(text.encode())
        rc, names, sets = parse_file(str(p), True)
        assert rc == 0
        want = [l for l in text.split("\n") if ":" in l]
        assert names == [l[: l.index(":")] for l in want]
        for l, s in zip(want, sets):
            assert s.tolist() == sorted(set(parse_line_like_reference(l[l.index(":") + 1:]))), l


def test_parser_toy_roundtrip(tmp_path):
    names, offsets, hashes = load_toy()
    p = tmp_path / "toy.txt"
    with open(p, "w") as f:
        for i, nme in enumerate(names):
            f.write(nme + ":" + "".join(" %d" % int(h) for h in hashes[int(offsets[i]):int(offsets[i + 1])]) + "\n")
    rc, got_names, sets = parse_file(str(p), True)
    assert rc == 0 and got_names == names
    for i, s in enumerate(sets):
        assert np.array_equal(s, np.unique(hashes[int(offsets[i]):int(offsets[i + 1])]))


def _raw_batch(path, named):
    """mvs_parse_hash_file -> (names, offsets, hashes) exactly as returned (no sorting)."""
    lib = mvs.load_library()
    hb = C.POINTER(_lib.HashBatch)()
    rc = lib.mvs_parse_hash_file(str(path).encode(), int(named), C.byref(hb))
    if rc != 0:
        return rc, None
    b = hb.contents
    n = b.n_samples
    off = np.ctypeslib.as_array(b.offsets, shape=(n + 1,)).copy()
    hashes = np.ctypeslib.as_array(b.hashes, shape=(max(1, int(b.n_hashes)),))[: int(b.n_hashes)].copy()
    names = [b.names[i] for i in range(n)] if named else None
    lib.mvs_free_hash_batch(hb)
    return 0, (names, off, hashes)


def _write_text_case(path):
    """ragged text hash file: empty sets, names with spaces, duplicates, a line without ':' (skipped), a long sample"""
    rng = np.random.default_rng(31)
    with open(path, "w") as f:
        f.write("first sample: 5 3 5 9\n")
        f.write("no colon here 1 2 3\n")
        f.write("empty:\n")
        f.write(": 7\n")
        for i in range(40):
            k = int(rng.integers(0, 3000))
            f.write("s%d/x y.sig:" % i + "".join(" %d" % int(h) for h in rng.integers(0, 2 ** 64 - 1, k, dtype=np.uint64)) + "\n")
        f.write("last: 18446744073709551615 0")                                     # no trailing newline


def test_binary_hash_file_round_trip(tmp_path):
    """pack: text -> batch -> binary file -> the same batch (names, offsets, payload order), in both parser modes, through
    the library call and through `project_everything pack`."""
    lib = mvs.load_library()
    txt, binf = tmp_path / "h.txt", tmp_path / "h.u64"
    _write_text_case(txt)
    for named in (1, 0):
        rc, want = _raw_batch(txt, named)
        assert rc == 0
        hb = C.POINTER(_lib.HashBatch)()
        assert lib.mvs_parse_hash_file(str(txt).encode(), named, C.byref(hb)) == 0
        assert lib.mvs_write_hash_batch(hb, str(binf).encode()) == 0
        lib.mvs_free_hash_batch(hb)
        raw = open(binf, "rb").read()
        assert raw[:8] == b"MVSU64\x00\x01" and len(raw) % 8 == 0
        n, nh, nb, pay = np.frombuffer(raw[8:40], dtype=np.uint64)
        assert int(pay) % 4096 == 0 and len(raw) == int(pay) + 8 * int(nh) and int(n) == len(want[1]) - 1
        assert np.array_equal(np.frombuffer(raw[int(pay):], dtype=np.uint64), want[2])   # the payload, verbatim
        rc, got = _raw_batch(binf, named)
        assert rc == 0 and got[0] == want[0] and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
        if named:                                     # a named file also serves the unnamed parser; not the reverse
            rc, got0 = _raw_batch(binf, 0)
            assert rc == 0 and got0[0] is None and np.array_equal(got0[2], want[2])
        else:
            assert _raw_batch(binf, 1)[0] == -1
    pe = os.path.join(BIN, "project_everything")
    if os.path.exists(pe):
        out = tmp_path / "cli.u64"
        r = subprocess.run([pe, "pack", str(txt), str(out)], capture_output=True, text=True)
        assert r.returncode == 0 and "Packed 44 samples" in r.stdout
        rc, want = _raw_batch(txt, 1)
        rc2, got = _raw_batch(out, 1)
        assert rc == 0 and rc2 == 0 and got[0] == want[0] and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
        r = subprocess.run([pe, "pack", str(tmp_path / "missing.txt"), str(out)], capture_output=True, text=True)
        assert r.returncode == 2 and "pack:" in r.stderr


def test_binary_hash_file_rejects_damaged_files(tmp_path):
    lib = mvs.load_library()
    txt, binf = tmp_path / "h.txt", tmp_path / "h.u64"
    _write_text_case(txt)
    hb = C.POINTER(_lib.HashBatch)()
    assert lib.mvs_parse_hash_file(str(txt).encode(), 1, C.byref(hb)) == 0
    assert lib.mvs_write_hash_batch(hb, str(binf).encode()) == 0
    lib.mvs_free_hash_batch(hb)
    good = bytearray(open(binf, "rb").read())
    n = int(np.frombuffer(bytes(good[8:16]), dtype=np.uint64)[0])

    def load(data):
        p = tmp_path / "bad.u64"
        open(p, "wb").write(bytes(data))
        return _raw_batch(p, 1)[0]

    assert load(good) == 0
    assert load(good[:-8]) == -6                                                    # truncated payload
    assert load(good + b"\0" * 8) == -6                                            # trailing bytes
    assert load(good[:100]) == -6                                                   # header only
    bad = bytearray(good); bad[16:24] = (10 ** 15).to_bytes(8, "little"); assert load(bad) == -6      # n_hashes lies
    bad = bytearray(good); bad[8:16] = (n + 1).to_bytes(8, "little"); assert load(bad) == -6          # n_samples lies
    bad = bytearray(good); bad[32:40] = (4097).to_bytes(8, "little"); assert load(bad) == -6          # payload offset
    bad = bytearray(good)                                                           # offsets not monotone
    bad[64 + 8 * 5:64 + 8 * 6] = (2 ** 40).to_bytes(8, "little")
    assert load(bad) == -6
    bad = bytearray(good)                                                           # names without terminators
    nb = int(np.frombuffer(bytes(good[24:32]), dtype=np.uint64)[0])
    at = 64 + 8 * (n + 1)
    bad[at:at + nb] = b"x" * nb
    assert load(bad) == -6


@pytest.mark.parametrize("tag,i16", [("d2048", False), ("d2048_i16", True), ("d100", False)])
@pytest.mark.parametrize("norms_from", ["ref_norms", "vec32"])
def test_write_db_matches_reference_files(tmp_path, tag, i16, norms_from):
    """All four db files byte-identical to the reference binary's (golden: toy, 61 samples)."""
    lib = mvs.load_library()
    names, _, _ = load_toy()
    ref = load_toy_ref(tag)
    vec = ref["vectors"]
    n, d = vec.shape
    if not golden_host_has_same_rsqrt(d):
        pytest.skip("this CPU's rsqrtss differs from the host that printed the golden norms")
    v32 = load_toy_ref("d2048")["vectors"].astype(np.int32) if i16 else vec.astype(np.int32)   # norms come from the int32 rows (:328)
    arr = (C.c_char_p * n)(*[s.encode() for s in names])
    v16 = np.ascontiguousarray(vec.astype(np.int16)) if i16 else None
    out = tmp_path / "db"
    out.mkdir()
    if norms_from == "vec32" and i16:
        pytest.skip("the writer derives norms itself only from int32 rows")
    rn = mvs.reference_norms(v32) if norms_from == "ref_norms" else None
    rc = lib.mvs_write_db(str(out).encode(), n, d, arr, rn.ctypes.data if rn is not None else None, None,
                          None if i16 else v32.ctypes.data, v16.ctypes.data if i16 else None)
    assert rc == 0
    raw = (out / "vectors.bin").read_bytes()
    assert hashlib.sha256(raw).hexdigest() == ref["vectors_sha256"]          # byte-identical vectors.bin
    assert (out / "dimension.txt").read_text() == ref["dimension_txt"]
    assert (out / "dtype.txt").read_text() == ref["dtype_txt"]
    assert (out / "vector_norms.txt").read_text() == ref["norms_txt"]        # 61/61 lines, byte for byte


def test_write_db_exact_norms_mode_within_print_precision(tmp_path):
    """The caller's other choice: sqrt(exact sum of squares / d).  Tolerance parity only (half a unit of the
    6th digit + 2e-6 relative for the reference's float32 evaluation)."""
    lib = mvs.load_library()
    names, _, _ = load_toy()
    ref = load_toy_ref("d2048")
    v32 = ref["vectors"].astype(np.int32)
    n, d = v32.shape
    sumsq = (v32.astype(np.int64) ** 2).sum(axis=1).astype(np.uint64)
    arr = (C.c_char_p * n)(*[s.encode() for s in names])
    out = tmp_path / "db"
    out.mkdir()
    assert lib.mvs_write_db(str(out).encode(), n, d, arr, None, sumsq.ctypes.data, v32.ctypes.data, None) == 0
    for a, b in zip((out / "vector_norms.txt").read_text().split("\n")[:-1], ref["norms_txt"].split("\n")[:-1]):
        (na, va), (nb, vb) = a.rsplit(" ", 1), b.rsplit(" ", 1)
        exact = float(va)
        ulp6 = 10.0 ** (np.floor(np.log10(exact)) - 5) if exact > 0 else 0.0
        assert na == nb and abs(float(vb) - exact) <= 1.0 * ulp6 + 2e-6 * exact


def test_norms_file_identical_to_live_reference_binary_10k(tmp_path):
    """>= 10 000 samples: the reference binary sketches a family hash file on THIS host; our writer, fed with the
    binary's own vectors.bin, must print the same vector_norms.txt byte for byte (project_everything.cpp:328-330)."""
    rd = ref_dir()
    if rd is None:
        pytest.skip("oracle/_ref not built on this host")
    lib = mvs.load_library()
    hf = tmp_path / "fam.txt"
    names, offsets, hashes = write_family_hash_file(str(hf), 10000, seed=3, mean_size=120)
    db = tmp_path / "db_ref"
    r = subprocess.run([os.path.join(rd, "project_everything"), "sketch", str(hf), str(db), "-d", "1000"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    d = 1000
    v32 = np.fromfile(db / "vectors.bin", dtype=np.int32).reshape(-1, d)
    assert len(v32) == 10000
    out = tmp_path / "db_ours"
    out.mkdir()
    arr = (C.c_char_p * len(names))(*[s.encode() for s in names])
    assert lib.mvs_write_db(str(out).encode(), len(names), d, arr, None, None, v32.ctypes.data, None) == 0
    for f in ("vector_norms.txt", "vectors.bin", "dimension.txt", "dtype.txt"):
        assert (out / f).read_bytes() == (db / f).read_bytes(), f
    # the oracle's restatement agrees on every row as well
    assert np.array_equal(oracle_reference_norms(v32[:500]).view(np.uint32), mvs.reference_norms(v32[:500]).view(np.uint32))


def test_convert_reads_toy_signatures(tmp_path):
    toy = "/root/reference/test/toy"
    if not os.path.isdir(toy):
        pytest.skip("reference toy signatures not present on this host")
    lib = mvs.load_library()
    hf = tmp_path / "toy_hashes.txt"
    assert lib.mvs_convert_signatures(toy.encode(), str(hf).encode(), 4, 0) == 0
    rc, got_names, sets = parse_file(str(hf), True)
    names, offsets, hashes = load_toy()
    assert rc == 0 and got_names == names
    for i, s in enumerate(sets):
        assert np.array_equal(s, np.unique(hashes[int(offsets[i]):int(offsets[i + 1])]))


def test_partition_is_contiguous_balanced_cover():
    lib = mvs.load_library()
    sizes = mvs.synthetic.lognormal_sizes(5000, seed=3)
    off = np.zeros(len(sizes) + 1, dtype=np.uint64)
    off[1:] = np.cumsum(sizes)
    for parts in (1, 2, 4, 8):
        b = np.zeros(parts + 1, dtype=np.int64)
        lib.mvs_partition_samples(off.ctypes.data, len(sizes), parts, b.ctypes.data)
        assert b[0] == 0 and b[-1] == len(sizes) and np.all(np.diff(b) >= 0)
        loads = [int(off[b[k + 1]] - off[b[k]]) for k in range(parts)]
        assert sum(loads) == int(off[-1])
        assert max(loads) <= int(off[-1]) / parts + int(sizes.max()) + len(sizes)   # balanced within one sample
    # empty samples only: still a cover
    off0 = np.zeros(11, dtype=np.uint64)
    b = np.zeros(5, dtype=np.int64)
    lib.mvs_partition_samples(off0.ctypes.data, 10, 4, b.ctypes.data)
    assert b[0] == 0 and b[-1] == 10 and np.all(np.diff(b) >= 0)


def _plan(offsets, d, sm_count=148):
    lib = mvs.load_library()
    off = np.ascontiguousarray(offsets, dtype=np.uint64)
    n = len(off) - 1
    cap, nt, pr = C.c_uint64(), C.c_uint64(), C.c_uint64()
    assert lib.mvs_plan_tasks(off.ctypes.data, n, d, sm_count, C.byref(cap), C.byref(nt), C.byref(pr), None, None, None, None, 0) == 0
    t = int(nt.value)
    begin, count = np.zeros(t + 1, np.uint64), np.zeros(t + 1, np.uint32)
    sample, pas = np.zeros(t + 1, np.uint32), np.zeros(t + 1, np.uint32)
    assert lib.mvs_plan_tasks(off.ctypes.data, n, d, sm_count, C.byref(cap), C.byref(nt), C.byref(pr), begin.ctypes.data,
                              count.ctypes.data, sample.ctypes.data, pas.ctypes.data, t) == 0
    return int(cap.value), t, int(pr.value), begin[:t], count[:t], sample[:t], pas[:t]


@pytest.mark.parametrize("d", [1, 64, 200, 256, 512, 2048, 2049, 4096, 5000])
def test_planner_tasks_cover_every_hash_exactly_once_per_pass(d):
    """The host-side planner (no GPU): every (sample, column pass) is covered by contiguous chunks that partition the
    sample's hash range; chunk sizes respect the kernel's limits (multiple of the trip quantum except the last chunk of
    a sample, at most 65535 hashes per 64-column block and 16320 rounds per lane)."""
    rng = np.random.default_rng(d)
    sizes = np.concatenate([[0, 1, 63, 64, 65, 16320, 16321, 70000, 0, 1_300_000], rng.integers(0, 40000, 300)])
    off = np.zeros(len(sizes) + 1, dtype=np.uint64)
    off[1:] = np.cumsum(sizes)
    off += np.uint64(12345)                                           # offsets[0] need not be 0
    cap, t, prows, begin, count, sample, pas = _plan(off, d)
    nb = (d + 63) // 64
    passes = (nb + 31) // 32
    lph = 32 if nb >= 32 else 1 << max(0, (nb - 1).bit_length())
    S = 32 // lph
    max_rounds = min(16320, (65535 // S) // 64 * 64)
    quantum = min(64 * S, max(256, 16 * S))
    assert cap <= max_rounds * S and cap % quantum == 0
    assert count.max() <= cap
    covered = {}
    for b, c, s_, p_ in zip(begin, count, sample, pas):
        covered.setdefault((int(s_), int(p_)), []).append((int(b), int(c)))
    for s_ in range(len(sizes)):
        for p_ in range(passes):
            chunks = sorted(covered.get((s_, p_), []))
            assert chunks, (s_, p_)
            pos = int(off[s_] - off[0])
            for i, (b, c) in enumerate(chunks):
                assert b == pos and (c % quantum == 0 or i == len(chunks) - 1)
                pos += c
            assert pos == int(off[s_ + 1] - off[0])
    assert t == sum(len(v) for v in covered.values())
    assert prows == sum(len(v) for (s_, p_), v in covered.items() if p_ == 0 and len(v) > 1)


def test_planner_avoids_a_nearly_empty_last_wave():
    """BASELINE configs[2] (1024 x 100 000 hashes): round 1 cut 7168 tasks at d = 256 = 12.1 CTAs per SM on 148 SMs, the
    13th almost alone.  The planner's model (SM quantisation, fixed cost per task, latency-bound tail) picks 8192 tasks at
    d <= 1024 and 16 384 (two column passes) at d = 4096, the measured optima of profiles/r02z_ksweep_c3.jsonl.  Big
    launches keep the largest task the bit planes allow."""
    off = np.arange(1025, dtype=np.uint64) * np.uint64(100000)
    for d, want in ((256, 8192), (512, 8192), (1024, 8192), (4096, 16384)):
        cap, t, *_ = _plan(off, d)
        assert t == want, (d, t)
    assert _plan(off, 2048)[1] in (8192, 15360)                        # equal within 0.1 % on the GPU
    off2 = np.arange(10001, dtype=np.uint64) * np.uint64(50000)        # configs[1]
    cap2, t2, *_ = _plan(off2, 2048)
    assert cap2 == 16320 and t2 == 40000
    # a tiny batch is cut fine enough to occupy many SMs
    off3 = np.arange(4, dtype=np.uint64) * np.uint64(40000)
    cap3, t3, *_ = _plan(off3, 2048)
    assert cap3 == 512 and t3 >= 3 * 78


def test_cli_grammar_and_exit_codes(tmp_path):
    pe = os.path.join(BIN, "project_everything")
    sp = os.path.join(BIN, "standalone_projection")
    if not (os.path.exists(pe) and os.path.exists(sp)):
        pytest.skip("CLI drivers not built")
    for argv in ([pe], [pe, "sketch"], [pe, "sketch", "a"], [pe, "sketch", "a", "b", "--nope"],
                 [pe, "sketch", "a", "b", "-d"], [pe, "bogus"], [pe, "convert", "a"]):
        r = subprocess.run(argv, capture_output=True, text=True)
        assert r.returncode == 1 and "Usage:" in r.stderr             # reference: usage + exit 1
    r = subprocess.run([sp], capture_output=True, text=True)
    assert r.returncode == 1 and "Usage:" in r.stderr and "<hashes_file> <dimension>" in r.stderr
    r = subprocess.run([sp, "only_one"], capture_output=True, text=True)
    assert r.returncode == 1

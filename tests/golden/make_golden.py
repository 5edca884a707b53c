#!/usr/bin/env python
"""Regenerate the committed golden fixtures from the REFERENCE ITSELF.

Run in the build container only (needs /root/reference and `make -C oracle ref`):

    python tests/golden/make_golden.py

Everything written here is produced by the unmodified reference binaries in
oracle/_ref (compiled from /root/reference/src by oracle/Makefile):

  toy_hashes.npz     k=31 FracMinHash `mins` of /root/reference/test/toy/*.sig.zip,
                     files sorted by name, hashes in file order (what
                     `project_everything convert` would emit, project_everything.cpp:94-229;
                     `unzip` is absent here so the zip/gzip/JSON layers are read with
                     the Python standard library instead).
  toy_ref_d2048.npz  vectors.bin / vector_norms.txt / dimension.txt / dtype.txt written by
                     `oracle/_ref/v3/project_everything sketch toy_hashes.txt db/ -d 2048`
                     (project_everything.cpp:238-362)  == SURVEY.md KAT-3.
  toy_ref_d100.npz   same with -d 100 (d not a multiple of 64, random_projection.cpp:19).
  toy_ref_d2048_i16.npz  same with --int16.
  kat.json           `oracle/_ref/v3/standalone_projection` outputs for hand-written sets
                     (standalone_projection.cpp:11-46)  == SURVEY.md KAT-1/KAT-2, plus the
                     parser edge cases of SURVEY.md section 8(a5) pushed through `sketch`.
  pairwise_top10.json  KAT-4: reference pairwise_comp_optimized + query_pc_mat on the
                     reference toy db (top-10 neighbours of DRR000821 / DRR000837).
"""
import glob
import gzip
import hashlib
import json
import os
import subprocess
import sys
import tempfile
import zipfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
REFBIN = os.path.join(ROOT, "oracle", "_ref", "v3")


def extract_toy():
    names, sets = [], []
    for path in sorted(glob.glob(os.path.join(REF, "test", "toy", "*.sig.zip"))):
        base = os.path.basename(path)
        name = base[: base.index(".")]            # project_everything.cpp:210-211
        mins = []
        with zipfile.ZipFile(path) as z:
            for member in z.namelist():
                if not member.endswith(".gz"):
                    continue
                doc = json.loads(gzip.decompress(z.read(member)))
                for rec in doc:
                    for sig in rec["signatures"]:
                        if sig["ksize"] == 31:    # project_everything.cpp:116
                            mins.extend(int(m) for m in sig["mins"])
        names.append(name)
        sets.append(np.array(mins, dtype=np.uint64))
    return names, sets


def write_hash_file(path, names, sets):
    with open(path, "w") as f:
        for name, hs in zip(names, sets):
            f.write(name + ":" + "".join(" %d" % int(h) for h in hs) + "\n")


def run_sketch(hash_file, d, int16=False):
    with tempfile.TemporaryDirectory() as td:
        db = os.path.join(td, "db")
        cmd = [os.path.join(REFBIN, "project_everything"), "sketch", hash_file, db, "-d", str(d)]
        if int16:
            cmd.append("--int16")
        subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL)
        raw = open(os.path.join(db, "vectors.bin"), "rb").read()
        out = {
            "vectors": np.frombuffer(raw, dtype=np.int16 if int16 else np.int32).copy(),
            "vectors_sha256": hashlib.sha256(raw).hexdigest(),
            "norms_txt": open(os.path.join(db, "vector_norms.txt")).read(),
            "dimension_txt": open(os.path.join(db, "dimension.txt")).read(),
            "dtype_txt": open(os.path.join(db, "dtype.txt")).read(),
        }
        return out, db


def standalone(lines, d):
    with tempfile.NamedTemporaryFile("w", suffix=".hashes", delete=False) as f:
        f.write("".join(l + "\n" for l in lines))
        p = f.name
    r = subprocess.run([os.path.join(REFBIN, "standalone_projection"), p, str(d)],
                       check=True, capture_output=True, text=True)
    os.unlink(p)
    return r.stdout


def main():
    if not os.path.isdir(REF) or not os.path.exists(os.path.join(REFBIN, "project_everything")):
        sys.exit("needs /root/reference and `make -C oracle ref`")
    names, sets = extract_toy()
    offsets = np.zeros(len(sets) + 1, dtype=np.uint64)
    offsets[1:] = np.cumsum([len(s) for s in sets])
    hashes = np.concatenate(sets)
    np.savez_compressed(os.path.join(HERE, "toy_hashes.npz"), names=np.array(names),
                        offsets=offsets, hashes=hashes)
    print("toy: %d samples, %d hashes" % (len(names), len(hashes)))

    with tempfile.TemporaryDirectory() as td:
        hf = os.path.join(td, "toy_hashes.txt")
        write_hash_file(hf, names, sets)
        print("hash file sha256", hashlib.sha256(open(hf, "rb").read()).hexdigest())
        for tag, d, i16 in (("d2048", 2048, False), ("d100", 100, False), ("d2048_i16", 2048, True)):
            out, _ = run_sketch(hf, d, i16)
            v = out.pop("vectors")
            assert np.abs(v).max() < 32768
            np.savez_compressed(os.path.join(HERE, "toy_ref_%s.npz" % tag),
                                vectors=v.astype(np.int16).reshape(len(names), d),
                                **{k: np.array(val) for k, val in out.items()})
            print(tag, out["vectors_sha256"], out["norms_txt"].splitlines()[:2])

        # KAT-4: downstream top-10 on the reference db (config 5)
        db = os.path.join(td, "db/")
        subprocess.run([os.path.join(REFBIN, "project_everything"), "sketch", hf, db, "-d", "2048"],
                       check=True, stdout=subprocess.DEVNULL)
        mat = os.path.join(td, "mat/")
        os.makedirs(mat)
        subprocess.run([os.path.join(REFBIN, "pairwise_comp_optimized"), "--db", db,
                        "--output_folder", mat, "--max_memory_gb", "12", "--num_threads", "8",
                        "--num_shards", "1", "--shard_idx", "0"], check=True, stdout=subprocess.DEVNULL)
        q = subprocess.run([os.path.join(REFBIN, "query_pc_mat"), "--matrix", mat, "--db", db,
                            "--query_file", os.path.join(REF, "test", "query_strs.txt"),
                            "--top", "10", "--print"], capture_output=True, text=True)
        json.dump({"cmd": "pairwise_comp_optimized --max_memory_gb 12 --num_threads 8 --num_shards 1 "
                          "--shard_idx 0; query_pc_mat --query_file test/query_strs.txt --top 10 --print",
                   "returncode": q.returncode, "stdout": q.stdout},
                  open(os.path.join(HERE, "pairwise_top10.json"), "w"), indent=1)
        print(q.stdout[-600:], q.stderr[-300:])

    kat = {"standalone": [], "sketch_parser": {}}
    cases = [
        (64, ["0"]),
        (8, ["1", "0 1", "18446744073709551615", "5", "", "12345 67890 12345", "1 2 3"]),
        (100, ["7 8 9 1000000007", "42"]),
        (130, ["18446744073709551615 18446744073709551614 0"]),
        (2048, ["3 1 4 1 5 9 2 6 5 3 5 8 9 7 9"]),
        (4160, ["123456789012345 987654321"]),
    ]
    for d, lines in cases:
        kat["standalone"].append({"d": d, "lines": lines, "stdout": standalone(lines, d)})
    # parser edge cases of sketch() (project_everything.cpp:264-282)
    text = ("a: 1 2 3\n" "no colon line 5 6\n" "c: 5 x 6\n" "neg: -1\n"
            "big: 7 99999999999999999999999 8\n" "empty:\n" "sp ace: 10 10 11\n"
            "two:colons: 4\n" "tab:\t12\t13  14 \n" "plus: +15 016\n"
            "glued: 5-3 7+8\n" "hex: 0x10 3\n" "exp: 1e5 2\n" "cr: 21 22\r\n" "max: 18446744073709551615 18446744073709551616 9\n"
            "negbig: -18446744073709551615 -18446744073709551616 4\n" "last: 31 32")
    with tempfile.TemporaryDirectory() as td:
        hf = os.path.join(td, "edge.txt")
        open(hf, "w").write(text)
        out, _ = run_sketch(hf, 8)
        kat["sketch_parser"] = {"text": text, "d": 8, "vectors": out["vectors"].tolist(),
                                "norms_txt": out["norms_txt"]}
    json.dump(kat, open(os.path.join(HERE, "kat.json"), "w"), indent=1)
    print(kat["standalone"][1]["stdout"])
    print(kat["sketch_parser"]["norms_txt"])


if __name__ == "__main__":
    main()

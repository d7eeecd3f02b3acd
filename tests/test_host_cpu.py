"""CPU-only checks of the product's host side: the C-ABI library loads, exports every symbol
include/cmsbwt_b200.h declares, and its host loaders (reference reader, augmentation, collection
parser with -p) agree with the oracle on the fuzz families and the golden fixtures.  No compute
call is made (there is no GPU here and no CPU fallback)."""
import glob
import os
import re
import numpy as np
import pytest

import cms_bwt_b200 as P
import oracle_lib as O
import fuzz_cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "cmsbwt_b200.h")).read()
    declared = set(re.findall(r"\b(cmsbwt_[a-z_0-9]+)\s*\(", hdr))
    declared.discard("cmsbwt_ctx")
    L = P.lib()
    for name in sorted(declared):
        assert hasattr(L, name), name
    assert declared == set(P.EXPORTS)


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(P.CmsBwtError):
        P.Context(0)


def test_host_loaders_match_oracle_fuzz():
    for seed in range(300):
        c = fuzz_cases.make_case(seed)
        r1, r2 = P.read_reference(c["ref_file"]), O.read_reference(c["ref_file"])
        assert np.array_equal(r1, r2), c["desc"]
        assert np.array_equal(P.augment(r1), O.augment(r2)), c["desc"]
        t1, nd = P.parse_collection(c["coll_file"])
        t2, ds = O.parse_collection(c["coll_file"])
        assert np.array_equal(t1, t2), c["desc"]
        assert nd == len(ds) - 1


def test_host_parser_prefix_matches_oracle():
    for f in sorted(glob.glob(os.path.join(GOLDEN, "*.npz"))):
        z = np.load(f)
        coll = z["coll_file"].tobytes()
        for pfx in [int(z["prefix"]), 1, 2, 7, 100, 2001, 2002, 2003, len(coll) - 1, len(coll)]:
            t1, _ = P.parse_collection(coll, pfx)
            t2, _ = O.parse_collection(coll, pfx)
            assert np.array_equal(t1, t2), (f, pfx)


def test_augment_rejects_reserved_bytes():
    with pytest.raises(P.CmsBwtError):
        P.augment(np.frombuffer(b"AC\x02GT", dtype=np.uint8))


def _odd_files():
    """Collections around the line rules: unterminated last line, blank-line splits, no header,
    CR LF, header last, only headers, empty lines first."""
    return [b"", b"\n", b">a\n", b"ACGT", b"ACGT\n", b">a\nAC\nGT", b">a\nAC\n\nGT\n", b"AC\nGT\n>b\n",
            b"\n\n>a\n>b\nA\n", b">a\r\nAC\r\nGT\r\n", b">a\nACGT\n>b\nAC\n>c", b"A\n" * 50,
            b">x\n" + b"ACGTACGTAC\n" * 40 + b">y\n" + b"TTTT\n" * 33 + b"\n" + b"GG"]


def test_parallel_parser_matches_sequential():
    # streamed ingest: same bytes and document count as the sequential parse
    cases = [fuzz_cases.make_case(s)["coll_file"] for s in range(120)] + _odd_files()
    rng = np.random.default_rng(3)
    big = b"".join(b">s%d\n" % k + b"\n".join(bytes(rng.choice(list(b"ACGT"), int(rng.integers(0, 70))).astype(np.uint8))
                                             for _ in range(int(rng.integers(0, 400)))) + b"\n"
                   for k in range(60))
    cases.append(big)
    for coll in cases:
        want, nd = P.parse_collection(coll)
        for T in (1, 2, 3, 8, 0):
            got, nd2 = P.parse_collection_mt(coll, threads=T)
            assert np.array_equal(got, want) and nd2 == nd, (coll[:40], T)
        for piece in (1, 7, 64, 1000, 1 << 20):
            if len(coll) > 20000 and piece < 64:
                continue
            parts = list(P.parse_pieces(coll, piece))
            got = np.concatenate(parts) if parts else np.empty(0, np.uint8)
            assert np.array_equal(got, want), (coll[:40], piece)


def test_parallel_parser_prefix_falls_back_to_sequential():
    coll = b">a\n" + b"ACGTACGTAC\n" * 300 + b">b\n" + b"TTTTGGGG\n" * 200
    for pfx in (1, 2, 100, 2001, len(coll) - 1, len(coll), len(coll) + 5, P.UINT64_MAX):
        t1, n1 = P.parse_collection(coll, pfx)
        t2, n2 = P.parse_collection_mt(coll, pfx, threads=4)
        assert np.array_equal(t1, t2) and n1 == n2, pfx


def test_cli_exit_codes(tmp_path):
    """Process contract of main.cpp: usage -> exit(-1) (:23, :55-59), unreadable or empty inputs ->
    exit(1) (:91-115, CMS-BWT-functions.cpp:156-160); without a device the program fails, it does
    not fall back to the CPU."""
    import subprocess
    import torch
    exe = os.path.join(ROOT, "cms_bwt_b200", "cms_bwt")
    assert os.path.exists(exe), "build first"
    run = lambda *a: subprocess.run([exe, *a], cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    (tmp_path / "r.fa").write_bytes(b">ref\nACGTACGTTTGACA\n")
    (tmp_path / "c.fa").write_bytes(b">s1\nACGTACGATTGACA\n>s2\nACGTACGTTTGCA\n")
    (tmp_path / "e.fa").write_bytes(b"")
    (tmp_path / "ok.txt").write_text("r.fa\nc.fa\n")
    (tmp_path / "noref.txt").write_text("nope.fa\nc.fa\n")
    (tmp_path / "nocoll.txt").write_text("r.fa\nnope.fa\n")
    (tmp_path / "empty.txt").write_text("r.fa\ne.fa\n")
    (tmp_path / "short.txt").write_text("r.fa\n")
    assert run().returncode == 255
    assert run("-x", "ok.txt").returncode == 255
    assert run("a", "b").returncode == 255
    assert run("missing.txt").returncode == 1
    for lst in ("noref.txt", "nocoll.txt", "empty.txt", "short.txt"):
        r = run(lst)
        assert r.returncode == 1 and r.stderr, lst
    r = run("ok.txt")
    assert b"==== Command line:" in r.stdout and b"==== Parameters:" in r.stdout and b"==== CMS-BWT" in r.stdout
    if not torch.cuda.is_available():
        assert r.returncode == 1 and b"no CPU fallback" in r.stderr
        assert not (tmp_path / "ok.txt.bwt").exists()


def test_sharded_ingest_concatenates_to_the_whole_parse(tmp_path):
    """Multi-GPU command line: every rank parses only its byte range of the collection file
    (multi.file_shard_cuts / file_shard_text); the rank texts concatenated are the whole parse."""
    import mmap
    from cms_bwt_b200 import multi
    cases = [fuzz_cases.make_case(s)["coll_file"] for s in range(150)] + _odd_files()
    cases.append(b">a\nAC\n\n>b\n>c\nGT\nAA\n\n\n>d\nT\n>e\n" * 7 + b"ACG")
    used = 0
    for i, coll in enumerate(cases):
        want, nd = P.parse_collection(coll)
        path = tmp_path / ("c%d.fa" % i)
        path.write_bytes(coll)
        bufs = [coll]
        if len(coll) and i % 3 == 0:
            with open(path, "rb") as fh:
                bufs.append(mmap.mmap(fh.fileno(), 0, access=mmap.ACCESS_READ))
        for buf in bufs:
            for world in (1, 2, 3, 5, 8):
                cuts = multi.file_shard_cuts(buf, world)
                assert cuts[0] == 0 and cuts[-1] == len(coll) and all(x <= y for x, y in zip(cuts, cuts[1:]))
                if not multi.file_shards_usable(buf, cuts):
                    continue               # main() falls back to the whole-file parse
                parts = [multi.file_shard_text(buf, cuts, r, threads=1 + r % 2) for r in range(world)]
                got = np.concatenate([t for t, _ in parts])
                assert np.array_equal(got, want), (i, world)
                assert sum(n for _, n in parts) == nd, (i, world)
                for t, n in parts:
                    if world > 1:
                        assert len(t) > 0, (i, world)
                    assert len(t) == 0 or (t[-1] == 2 and n == int((t == 2).sum())), (i, world)
                used += world > 1
    assert used > 100

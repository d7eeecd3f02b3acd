#!/usr/bin/env python3
"""Generates tests/golden/*.npz by running the UNMODIFIED reference binary (oracle/_ref/cmsbwt,
built by oracle/Makefile from /root/reference) on small inputs.  Run in the build container:

    make -C oracle ref && python tests/golden/make_golden.py

Each fixture holds the two input files, the -p value, the reference's .bwt bytes and the
maximal-run RLE of those bytes in the reference's 9-byte record format (the .rl_bwt parity
target; SURVEY.md §0 fact 3 explains why the as-shipped -r output is not used).  The reference's
own -r output is stored too (`rl_bwt_as_shipped`) purely as a record of the defect.
"""
import os
import sys
import tempfile
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(HERE)), "tools"))
import oracle_lib as O        # noqa: E402
import fuzz_cases             # noqa: E402
import gen_synth              # noqa: E402


def rle_np(b):
    if len(b) == 0:
        return np.zeros(0, np.uint8)
    starts = np.flatnonzero(np.concatenate([[True], b[1:] != b[:-1]]))
    runs = np.diff(np.concatenate([starts, [len(b)]])).astype("<u8")
    rec = np.zeros((len(starts), 9), np.uint8)
    rec[:, :8] = runs.view(np.uint8).reshape(-1, 8)
    rec[:, 8] = b[starts]
    return rec.ravel()


def emit(name, ref_file, coll_file, prefix=2 ** 64 - 1):
    with tempfile.TemporaryDirectory() as wd:
        rp, cp = os.path.join(wd, "r"), os.path.join(wd, "c")
        open(rp, "wb").write(ref_file)
        open(cp, "wb").write(coll_file)
        p = None if prefix == 2 ** 64 - 1 else prefix
        bwt = O.run_reference(rp, cp, wd, prefix=p)
        shipped = O.run_reference(rp, cp, wd, rle_mode=True, prefix=p, name="rl")
    np.savez_compressed(os.path.join(HERE, name + ".npz"),
                        ref_file=np.frombuffer(ref_file, np.uint8),
                        coll_file=np.frombuffer(coll_file, np.uint8),
                        prefix=np.uint64(prefix), bwt=bwt, rl_bwt=rle_np(bwt),
                        rl_bwt_as_shipped=shipped)
    print(name, len(bwt), "bytes")


def main():
    assert O.have_ref(), "build oracle/_ref first"
    for seed in (3, 17, 42, 77, 101, 150, 199):
        c = fuzz_cases.make_case(seed)
        if len(O.parse_collection(c["coll_file"])[0]):
            emit("fuzz_%03d" % seed, c["ref_file"], c["coll_file"])
    # config-1 shape, scaled: 20 kbp reference + 8 copies at 0.1 % (large path needs n >= 1e6,
    # so this exercises the reference's small path) ...
    ref, copies = gen_synth.make_collection(20000, 8, 0.001, 42)
    emit("synth_20k_x8", gen_synth.fasta_bytes("ref", ref),
         b"".join(gen_synth.fasta_bytes("s%d" % k, t) for k, t in enumerate(copies)))
    # ... and 1 % divergence with block events (config-5 shape)
    ref, copies = gen_synth.make_collection(30000, 4, 0.01, 5, svs=3)
    emit("synth_30k_x4_sv", gen_synth.fasta_bytes("ref", ref),
         b"".join(gen_synth.fasta_bytes("s%d" % k, t) for k, t in enumerate(copies)))
    # many short documents (config-4 shape)
    ref, copies = gen_synth.make_collection(600, 200, 0.005, 9)
    emit("synth_600_x200", gen_synth.fasta_bytes("ref", ref),
         b"".join(gen_synth.fasta_bytes("s%d" % k, t) for k, t in enumerate(copies)))
    # -p cut inside a document and exactly on a boundary
    ref, copies = gen_synth.make_collection(2000, 5, 0.002, 3)
    coll = b"".join(gen_synth.fasta_bytes("s%d" % k, t) for k, t in enumerate(copies))
    emit("prefix_5000", gen_synth.fasta_bytes("ref", ref), coll, 5000)
    emit("prefix_boundary", gen_synth.fasta_bytes("ref", ref), coll, len(copies[0]) + 3)
    # edge cases of the readers
    emit("edge_raw_dollar", b"ACGTACGTTTGACCA$\n", b">a\nACGTTTGACC\n>b\n\n>c\nNNACGTn\nACG")
    emit("edge_headerless", b">r\nGATTACAGATTACA\n", b"GATTACA\n\nGATTTACA\nTTT\n\n\nACA\n")


if __name__ == "__main__":
    main()

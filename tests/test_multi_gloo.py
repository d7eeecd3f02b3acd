"""world_size-2 test of the multi-GPU host layer on the gloo backend (CPU): the orchestration of
cms_bwt_b200.multi.ShardedPath (shard plan, splitters, split sizes, all-to-all / all-gather order,
refinement loop, output ranges) with a numpy model of the device steps (tests/shard_model.py)
must reproduce the oracle's BWT of the whole collection.  The same orchestration drives the CUDA
steps on NCCL (tests/test_gpu_parity.py::test_multi_gpu_*, bench.py --gpus N)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import oracle_lib as O
import fuzz_cases
import gen_synth
import cms_bwt_b200 as P
from cms_bwt_b200 import multi

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run_world(x, text, workdir, name, world=2, coll_file=None, emit="pull"):
    case = os.path.join(workdir, name + ".npz")
    if coll_file is None:
        np.savez(case, x=x, text=text)
    else:
        np.savez(case, x=x, text=text, coll_file=np.frombuffer(coll_file, np.uint8))
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="1", CMSBWT_MODEL_EMIT=emit)
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "shard_worker.py"), case,
                                       os.path.join(workdir, name)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT))
    logs = [p.communicate(timeout=300)[0].decode() for p in procs]
    for p, lg in zip(procs, logs):
        assert p.returncode == 0, lg[-3000:]
    parts = [np.load(os.path.join(workdir, name + ".%d.npz" % r)) for r in range(world)]
    off = 0
    for z in parts:
        assert int(z["begin"]) == off or len(z["out"]) == 0
        off += len(z["out"])
    got = np.concatenate([z["out"] for z in parts])
    # -r across ranks: the concatenated records are the maximal-run RLE of the whole BWT
    recs = np.concatenate([z["rle"] for z in parts])
    assert np.array_equal(recs, O.rle(got)), "sharded RLE records"
    cnt = 0
    for z in parts:
        assert int(z["rle_first"]) == cnt
        cnt += len(z["rle"]) // 9
    return got, parts


def test_shard_documents_plan():
    assert multi.shard_documents([10, 10, 10, 10], 2) == [(0, 2), (2, 4)]
    assert multi.shard_documents([0, 100, 100], 2) == [(0, 2), (2, 3)]
    plan = multi.shard_documents([5] * 7, 8)
    assert plan[0][0] == 0 and plan[-1][1] == 7 and all(a <= b for a, b in plan)
    assert all(plan[i][1] == plan[i + 1][0] for i in range(7))
    assert multi.shard_documents([], 2) == [(0, 0), (0, 0)]


def test_plan_rle_borders():
    A, C_, G = 65, 67, 71
    # rank 1's single run and rank 2's first run continue rank 0's last run; rank 3 is empty
    info = [(3, 5, A, 2, C_), (1, 4, C_, 4, C_), (2, 1, C_, 7, G), (0, 0, 0, 0, 0), (2, 3, G, 1, A)]
    drop, add = multi.plan_rle_borders(info)
    assert drop == [False, True, True, False, True]
    assert add == [5, 0, 3, 0, 0]
    drop, add = multi.plan_rle_borders([(2, 1, A, 1, C_), (2, 1, A, 1, C_)])
    assert drop == [False, False] and add == [0, 0]


def test_merge_rle_borders():
    assert multi.merge_rle_borders([[(3, 65), (1, 67)], [(2, 67), (5, 71)], [], [(1, 71)]]) == \
        [(3, 65), (3, 67), (6, 71)]


@pytest.mark.parametrize("emit", ["pull", "bytes"])
@pytest.mark.parametrize("world", [2, 3])
def test_sharded_path_synthetic(workdir, world, emit):
    """emit = pull: stage 3 by SA-rank range as the C++ host layer does it (csrc/sharded.cu); bytes: by output
    byte range with head placements (the path kept for tiles too deep for k_tile_rank)"""
    ref, copies = gen_synth.make_collection(3000, 7, 0.01, 5)
    x = O.augment(ref)
    text = O.docs_to_text([np.zeros(0, np.uint8)] + copies)      # leading empty document (FASTA)
    got, parts = _run_world(x, text, workdir, "synth%d%s" % (world, emit), world, emit=emit)
    want = O.Oracle(x).bwt(text)
    assert np.array_equal(got, want)
    if emit == "pull":
        # slices are the buckets of contiguous SA-rank ranges: rank 0's starts with the terminators
        nd = int(np.sum(text == 2))
        assert np.array_equal(parts[0]["out"][:nd], want[:nd]) and len(parts[0]["out"]) >= nd


def test_sharded_path_duplicates_and_short_documents(workdir):
    rng = np.random.default_rng(3)
    ref = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, 500)]
    docs = [ref[a:a + 40].copy() for a in rng.integers(0, 460, 60)]
    mut = ref.copy()
    for q in (50, 120, 121, 300, 420):                                      # a few SNPs
        mut[q] = ord("A") if mut[q] != ord("A") else ord("C")
    # exact duplicates of a mutated copy: their tails tie on (ISA rank, cls) and need the rank of
    # the next tail, several rounds deep
    docs += [docs[3].copy(), docs[3].copy(), ref.copy(), mut.copy(), mut.copy(), mut[100:].copy(), mut.copy()]
    x = O.augment(ref)
    text = O.docs_to_text(docs)
    got, parts = _run_world(x, text, workdir, "dups")
    assert np.array_equal(got, O.Oracle(x).bwt(text))
    assert int(parts[0]["rounds"]) >= 1


def test_sharded_path_fuzz(workdir):
    done = 0
    for seed in (3, 17, 42, 77, 101, 150):
        c = fuzz_cases.make_case(seed)
        x = O.augment(O.read_reference(c["ref_file"]))
        text, docs = O.parse_collection(c["coll_file"])
        if len(text) == 0:
            continue
        got, _ = _run_world(x, text, workdir, "fuzz%d" % seed)
        assert np.array_equal(got, O.Oracle(x).bwt(text)), c["desc"]
        done += 1
    assert done >= 4


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_ingest_by_file_ranges(workdir, world):
    """The multi-GPU command line's ingest (multi.file_shard_cuts / file_shard_text): every rank
    parses its own byte range of the collection file; a range that holds only header and blank lines
    gives a rank nothing but empty documents."""
    ref, copies = gen_synth.make_collection(2500, 6, 0.01, 8)
    x = O.augment(ref)
    fa = b"".join(gen_synth.fasta_bytes("s%d" % k, c) for k, c in enumerate(copies))
    # blank lines and a header run in the middle (empty documents), unterminated last line
    mid = len(fa) // 2
    cutpos = fa.index(b"\n>", mid) + 1
    coll = fa[:cutpos] + b"\n>e1\n>e2\n\n" + fa[cutpos:] + b"ACGTAC"
    text, docs = O.parse_collection(coll)
    cuts = multi.file_shard_cuts(coll, world)
    assert multi.file_shards_usable(coll, cuts)
    got, parts = _run_world(x, text, workdir, "fileshard%d" % world, world, coll_file=coll)
    assert np.array_equal(got, O.Oracle(x).bwt(text))

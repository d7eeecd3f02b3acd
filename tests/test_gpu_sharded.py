"""GPU parity tests of the multi-GPU path (cms_bwt_b200.multi + the shard API of the C ABI).

On a single-GPU box the ranks are threads of this process, one context each on cuda:0, exchanging
through multi.ThreadComm -- every CUDA shard step and the whole orchestration run exactly as under
NCCL; only the transport differs.  With >= 2 GPUs the same cases also run as one process per GPU over
NCCL (torchrun-style environment, tests/shard_worker_gpu.py).  Bit-exact against the oracle."""
import glob
import os
import socket
import subprocess
import sys
import threading

import numpy as np
import pytest
import torch

import cms_bwt_b200 as P
from cms_bwt_b200 import multi
import oracle_lib as O
import fuzz_cases
import gen_synth

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")


def split_text(text, world):
    ends = np.flatnonzero(text == 2)
    starts = np.concatenate([[0], ends[:-1] + 1])
    plan = multi.shard_documents((ends - starts).tolist(), world)
    out = []
    for a, b in plan:
        lo = int(starts[a]) if a < len(ends) else len(text)
        hi = int(ends[b - 1]) + 1 if b > a else lo
        out.append((text[lo:hi], b - a))
    return out


def run_threads(x, text, world, params=None):
    dev = torch.device("cuda:0")
    shards = split_text(text, world)
    shared = multi.ThreadComm.Shared(world)
    res, err = [None] * world, [None] * world

    def work(r):
        ctx = None
        try:
            torch.cuda.set_device(0)
            ctx = P.Context(0)
            for k, v in (params or {}).items():
                ctx.set_param(k, v)
            ctx.index_upload(x)
            shard, nd = shards[r]
            pad = np.full(64, 2, np.uint8)
            t = torch.from_numpy(np.concatenate([shard, pad])).to(dev)
            if len(shard) == 0:
                raise RuntimeError("empty shard")
            sp, out = multi.run_sharded(ctx, dev, t.data_ptr(), len(shard), nd, r, world,
                                        comm=multi.ThreadComm(shared, r))
            plain = out.cpu().numpy().copy()
            rl, rl_first = sp.rle()
            res[r] = (sp.out_begin, plain, sp.rank_rounds, ctx.stats(), rl.cpu().numpy().copy(), rl_first)
        except BaseException as e:          # noqa: BLE001  (a dead rank must not leave the others waiting)
            err[r] = e
            shared.barrier.abort()
        finally:
            if ctx is not None:
                ctx.close()

    th = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for e in err:
        if e is not None and not isinstance(e, threading.BrokenBarrierError):
            raise e
    for e in err:
        if e is not None:
            raise e
    off = 0
    for r_ in res:
        assert r_[0] == off or len(r_[1]) == 0
        off += len(r_[1])
    return np.concatenate([r_[1] for r_ in res]), res


def _check(x, text, world, desc, params=None):
    want = O.Oracle(x).bwt(text)
    got, res = run_threads(x, text, world, params)
    if not np.array_equal(got, want):
        bad = np.flatnonzero(got != want) if len(got) == len(want) else []
        raise AssertionError("sharded BWT %s world=%d: len %d/%d, %d mismatches, first %s"
                             % (desc, world, len(got), len(want), len(bad), bad[:8]))
    # -r across ranks: concatenated records = maximal-run RLE of the whole BWT
    assert np.array_equal(np.concatenate([r_[4] for r_ in res]), O.rle(want)), "sharded RLE " + desc
    cnt = 0
    for r_ in res:
        assert r_[5] == cnt
        cnt += len(r_[4]) // 9
    return res


@pytest.mark.parametrize("world", [1, 2, 3, 4])
def test_sharded_synthetic(world):
    ref, copies = gen_synth.make_collection(30000, 9, 0.004, 21)
    x = P.augment(ref)
    text = O.docs_to_text([np.zeros(0, np.uint8)] + copies)
    res = _check(x, text, world, "synthetic 30k x 9")
    assert all(r[3]["fast_path"] == 1 for r in res)


@pytest.mark.parametrize("world", [1, 2, 3])
def test_sharded_deep_tile_ranges(world):
    """owners count with the sort-based sweep (forced here; taken by itself when a tile holds more
    entries than shared memory, e.g. config 4's 300,000 documents over 30 kbp)"""
    ref, copies = gen_synth.make_collection(30000, 9, 0.004, 21)
    x = P.augment(ref)
    text = O.docs_to_text([np.zeros(0, np.uint8)] + copies)
    res = _check(x, text, world, "synthetic 30k x 9, deep", params={"tile_budget": 0})
    assert all(r[3]["light_tiles"] == 0 for r in res)
    done = 0
    for seed in range(40, 80):
        c = fuzz_cases.make_case(seed)
        x = P.augment(P.read_reference(c["ref_file"]))
        text, nd = P.parse_collection(c["coll_file"])
        if nd < world + 1 or len(text) < 40 or any(len(s_) == 0 for s_, _ in split_text(text, world)):
            continue
        _check(x, text, world, "fuzz %d deep" % seed, params={"tile_budget": 0})
        done += 1
    assert done >= 10


@pytest.mark.parametrize("kind", ["units", "period7", "homopolymer"])
def test_sharded_repetitive_references(kind):
    """all-repeat references (long tie chains refined across ranks, skewed key ranges and tiles) on 3 ranks"""
    from test_gpu_parity import _repetitive
    ref, copies = _repetitive(kind, np.random.default_rng(6))
    x = P.augment(ref)
    text = O.docs_to_text([np.zeros(0, np.uint8)] + copies)
    _check(x, text, 3, "repetitive " + kind)


def test_sharded_golden_fixtures():
    n = 0
    for f in sorted(glob.glob(os.path.join(GOLDEN, "*.npz"))):
        z = np.load(f)
        x = P.augment(P.read_reference(z["ref_file"].tobytes()))
        text, nd = P.parse_collection(z["coll_file"].tobytes(), int(z["prefix"]))
        if nd < 2:
            continue
        got, _ = run_threads(x, text, 2)
        assert np.array_equal(got, z["bwt"]), f
        n += 1
    assert n >= 5


def test_sharded_fuzz_general_alphabets():
    done = 0
    for seed in range(40):
        c = fuzz_cases.make_case(seed)
        x = P.augment(P.read_reference(c["ref_file"]))
        text, nd = P.parse_collection(c["coll_file"])
        if nd < 3 or len(text) < 40:
            continue
        shards = split_text(text, 2)
        if any(len(s) == 0 for s, _ in shards):
            continue
        _check(x, text, 2, c["desc"])
        done += 1
    assert done >= 10


def test_sharded_duplicates_need_refinement_rounds():
    rng = np.random.default_rng(8)
    ref = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, 20000)]
    mut = ref.copy()
    for q in rng.integers(100, 19900, 40):
        mut[q] = ord("A") if mut[q] != ord("A") else ord("C")
    docs = [mut.copy(), ref.copy(), mut.copy(), mut[5000:].copy(), ref[:7000].copy(), mut.copy(), mut.copy()]
    x = P.augment(ref)
    text = O.docs_to_text(docs)
    res = _check(x, text, 3, "duplicates")
    assert max(r[2] for r in res) >= 1


def test_sharded_many_short_documents():
    rng = np.random.default_rng(4)
    ref = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, 3000)]
    docs = []
    for k in range(1500):
        a = int(rng.integers(0, 2900))
        d = ref[a:a + int(rng.integers(1, 100))].copy()
        if rng.random() < 0.3 and len(d) > 2:
            d[len(d) // 2] = ord("A") if d[len(d) // 2] != ord("A") else ord("G")
        docs.append(d)
    x = P.augment(ref)
    _check(x, O.docs_to_text(docs), 4, "1500 short documents")


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_sharded_nccl_all_gpus(workdir):
    """one process per visible GPU over NCCL (2, 4 or 8 ranks)"""
    world = min(8, torch.cuda.device_count())
    ref, copies = gen_synth.make_collection(200000, 19, 0.002, 3)
    x = P.augment(ref)
    text = O.docs_to_text([np.zeros(0, np.uint8)] + copies)
    case = os.path.join(workdir, "nccl_case.npz")
    np.savez(case, x=x, text=text)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "shard_worker_gpu.py"), case,
                                       os.path.join(workdir, "nccl_out")], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT))
    logs = [p.communicate(timeout=600)[0].decode() for p in procs]
    for p, lg in zip(procs, logs):
        assert p.returncode == 0, lg[-3000:]
    zs = [np.load(os.path.join(workdir, "nccl_out.%d.npz" % r)) for r in range(world)]
    want = O.Oracle(x).bwt(text)
    assert np.array_equal(np.concatenate([z["out"] for z in zs]), want)
    assert np.array_equal(np.concatenate([z["rle"] for z in zs]), O.rle(want))
    cnt, off = 0, 0
    for z in zs:
        assert int(z["begin"]) == off and int(z["rle_first"]) == cnt
        cnt += len(z["rle"]) // 9
        off += len(z["out"])


def test_native_host_layer_one_rank():
    """cmsbwt_run_sharded (C++ host layer, NCCL communicator of one rank) on a single GPU: every
    exchange runs as a self send/receive; BWT and RLBWT against the oracle"""
    dev = torch.device("cuda:0")
    cases = []
    ref, copies = gen_synth.make_collection(30000, 9, 0.004, 21)
    cases.append((P.augment(ref), O.docs_to_text([np.zeros(0, np.uint8)] + copies)))
    rng = np.random.default_rng(8)
    ref = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, 20000)]
    mut = ref.copy()
    for q in rng.integers(100, 19900, 40):
        mut[q] = ord("A") if mut[q] != ord("A") else ord("C")
    cases.append((P.augment(ref), O.docs_to_text([mut.copy(), ref.copy(), mut.copy(), mut[5000:].copy(), mut.copy()])))
    for seed in (3, 7, 11):
        c = fuzz_cases.make_case(seed)
        t, nd = P.parse_collection(c["coll_file"])
        if len(t) >= 40:
            cases.append((P.augment(P.read_reference(c["ref_file"])), t))
    ctx = P.Context(0)
    try:
        multi.comm_init(ctx, 0, 1, dev)
        for x, text in cases:
            want = O.Oracle(x).bwt(text)
            ctx.index_upload(x)
            t = torch.from_numpy(np.concatenate([text, np.full(64, 2, np.uint8)])).to(dev)
            nd = int((text == 2).sum())
            r = multi.run_sharded_native(ctx, dev, t.data_ptr(), len(text), nd, P.MODE_PLAIN)
            assert r.out_begin == 0 and r.out_len == len(text) == r.N_total
            assert np.array_equal(r.out.cpu().numpy(), want)
            r2 = multi.run_sharded_native(ctx, dev, t.data_ptr(), len(text), nd, P.MODE_RLE)
            assert r2.first_record == 0 and np.array_equal(r2.records.cpu().numpy(), O.rle(want))
            assert r.collectives >= 8
    finally:
        ctx.close()


@pytest.mark.parametrize("rle_mode", [False, True])
def test_cli_gpus_option_matches_golden(workdir, rle_mode):
    """cms_bwt --gpus W (C++ host: one thread per GPU, NCCL inside the library) on every visible GPU
    (W = 1 on a single-GPU box) against a golden .bwt of the reference binary"""
    z = np.load(os.path.join(GOLDEN, "synth_20k_x8.npz"))
    d = os.path.join(workdir, "cli_gpus_%d" % rle_mode)
    os.makedirs(d, exist_ok=True)
    z["ref_file"].tofile(os.path.join(d, "ref.fa"))
    z["coll_file"].tofile(os.path.join(d, "coll.fa"))
    with open(os.path.join(d, "list.txt"), "w") as f:
        f.write(os.path.join(d, "ref.fa") + "\n" + os.path.join(d, "coll.fa") + "\n")
    world = max(1, min(4, torch.cuda.device_count()))
    exe = os.path.join(os.path.dirname(HERE), "cms_bwt_b200", "cms_bwt")
    cmd = [exe, "--gpus", str(world), "-o", os.path.join(d, "out")] + (["-r"] if rle_mode else []) + [os.path.join(d, "list.txt")]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=600)
    assert r.returncode == 0 and b"==== Time elapsed:" in r.stdout, r.stdout.decode()[-3000:]
    got = np.fromfile(os.path.join(d, "out.rl_bwt" if rle_mode else "out.bwt"), dtype=np.uint8)
    want = O.rle(z["bwt"]) if rle_mode else z["bwt"]
    assert np.array_equal(got, want)
    assert "Multi-GPU path on %d GPUs" % world in open(os.path.join(d, "out.log")).read()


@pytest.mark.parametrize("rle_mode", [False, True])
def test_cli_multi_matches_golden(workdir, rle_mode):
    """tools/cms_bwt_multi.py (the cms_bwt command line over a torchrun launch) on every visible GPU
    (one rank on a single-GPU box) against a golden .bwt of the reference binary"""
    z = np.load(os.path.join(GOLDEN, "synth_20k_x8.npz"))
    d = os.path.join(workdir, "cli_multi_%d" % rle_mode)
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "ref.fa"), "wb") as f:
        f.write(z["ref_file"].tobytes())
    with open(os.path.join(d, "coll.fa"), "wb") as f:
        f.write(z["coll_file"].tobytes())
    with open(os.path.join(d, "list.txt"), "w") as f:
        f.write(os.path.join(d, "ref.fa") + "\n" + os.path.join(d, "coll.fa") + "\n")
    world = max(1, min(4, torch.cuda.device_count()))
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(os.path.dirname(HERE), "tools", "cms_bwt_multi.py"), "-o", os.path.join(d, "out")]
    if rle_mode:
        cmd.append("-r")
    cmd.append(os.path.join(d, "list.txt"))
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=600)
    assert r.returncode == 0, r.stdout.decode()[-3000:]
    got = np.fromfile(os.path.join(d, "out.rl_bwt" if rle_mode else "out.bwt"), dtype=np.uint8)
    want = O.rle(z["bwt"]) if rle_mode else z["bwt"]
    assert np.array_equal(got, want)

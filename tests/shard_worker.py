"""One rank of the world_size-2 gloo test of cms_bwt_b200.multi.ShardedPath (TEST INFRASTRUCTURE).
usage: shard_worker.py <case.npz> <outprefix>   (RANK / WORLD_SIZE / MASTER_* from the environment)"""
import os
import sys

import numpy as np
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
for p in (HERE, os.path.dirname(HERE)):
    if p not in sys.path:
        sys.path.insert(0, p)

import oracle_lib as O                      # noqa: E402
from shard_model import NumpyShardOps       # noqa: E402
from cms_bwt_b200 import multi              # noqa: E402


def main():
    case, outp = sys.argv[1], sys.argv[2]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo", rank=rank, world_size=world)
    z = np.load(case)
    x, text = z["x"], z["text"]
    if "coll_file" in z.files:
        # the multi-GPU command line's sharded ingest: this rank parses only its byte range of the file
        coll = z["coll_file"].tobytes()
        cuts = multi.file_shard_cuts(coll, world)
        assert multi.file_shards_usable(coll, cuts)
        shard, ndocs = multi.file_shard_text(coll, cuts, rank)
        a, b = 0, ndocs
    else:
        ends = np.flatnonzero(text == 2)
        starts = np.concatenate([[0], ends[:-1] + 1])
        lens = (ends - starts).tolist()
        a, b = multi.shard_documents(lens, world)[rank]
        lo = int(starts[a]) if a < len(lens) else len(text)
        hi = int(ends[b - 1]) + 1 if b > a else lo
        shard = text[lo:hi]
    orc = O.Oracle(x)
    ops = NumpyShardOps(orc, shard)
    sp = multi.ShardedPath(ops, rank, world)
    sp.pull = os.environ.get("CMSBWT_MODEL_EMIT", "pull") == "pull"      # stage 3 by SA-rank range / by output byte range
    out = sp.run(len(shard), b - a)
    rl, rl_first = sp.rle()
    np.savez(outp + ".%d.npz" % rank, out=out.numpy(), begin=sp.out_begin, rounds=sp.rank_rounds,
             collectives=sp.collectives, rle=rl.numpy(), rle_first=rl_first)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

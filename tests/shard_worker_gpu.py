"""One rank (one GPU) of the NCCL parity test of the multi-GPU path (TEST INFRASTRUCTURE).
usage: shard_worker_gpu.py <case.npz> <outprefix>   (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*)"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
for p in (HERE, os.path.dirname(HERE), os.path.join(os.path.dirname(HERE), "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

import cms_bwt_b200 as P                    # noqa: E402
from cms_bwt_b200 import multi              # noqa: E402
from test_gpu_sharded import split_text     # noqa: E402


def main():
    case, outp = sys.argv[1], sys.argv[2]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    z = np.load(case)
    x, text = z["x"], z["text"]
    shard, nd = split_text(text, world)[rank]
    ctx = P.Context(local)
    ctx.index_upload(x)
    t = torch.from_numpy(np.concatenate([shard, np.full(64, 2, np.uint8)])).to(dev)
    if os.environ.get("CMSBWT_TEST_PYTHON_HOST") == "1":        # the Python orchestration over torch.distributed
        sp, out = multi.run_sharded(ctx, dev, t.data_ptr(), len(shard), nd, rank, world)
        np.savez(outp + ".%d.npz" % rank, out=out.cpu().numpy(), begin=sp.out_begin)
    else:                                                        # the C++ host layer over the library's own NCCL
        multi.comm_init(ctx, rank, world, dev)
        r = multi.run_sharded_native(ctx, dev, t.data_ptr(), len(shard), nd, P.MODE_PLAIN)
        plain = r.out.cpu().numpy().copy()
        r2 = multi.run_sharded_native(ctx, dev, t.data_ptr(), len(shard), nd, P.MODE_RLE)
        np.savez(outp + ".%d.npz" % rank, out=plain, begin=r.out_begin, rle=r2.records.cpu().numpy(),
                 rle_first=r2.first_record, collectives=r.collectives)
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

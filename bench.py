#!/usr/bin/env python3
"""bench.py -- collection Gbp/s to BWT (MS + head sort + emit) on B200.

A "step" is one pass of the hot path (cmsbwt_ms_run_device + cmsbwt_heads_sort + cmsbwt_emit,
plain BWT) over the whole synthetic collection, text resident in HBM; `value` = sequence symbols /
time.  `e2e` is the same job through the host-buffer call cmsbwt_run (H2D of the text and D2H of
the BWT inside the timed region, pinned host memory).  `--impl reference` times the unmodified
reference binary (oracle/_ref/cmsbwt, 1 thread -- it has none to use) on a bounded sample.

Workloads (BASELINE.json configs): config3 = 59 Mbp reference + 100 haplotypes at 0.1 % (5.9 Gbp,
the configuration the metric is quoted on; default), config2 = 5 Mbp + 100 copies, config1 = 1 Mbp
+ 10 copies.  Synthetic data, generated on the device by tools/gen_device.py.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

WORKLOADS = {
    "config3": dict(ref_bp=59_000_000, copies=100, div=0.001, seed=13),
    "config2": dict(ref_bp=5_000_000, copies=100, div=0.001, seed=11),
    "config1": dict(ref_bp=1_000_000, copies=10, div=0.001, seed=42),
    # config 5 shape (248 Mbp reference, 1 % + 20 block SVs per copy): 4 copies fit the single-GPU
    # bench; the full 64 copies are for --gpus 8
    "config4_10": dict(ref_bp=30_000, copies=30_000, div=0.001, seed=9, many=True),     # config 4 at 1/10
    "config4": dict(ref_bp=30_000, copies=300_000, div=0.001, seed=9, many=True),
    "config5_4": dict(ref_bp=248_000_000, copies=4, div=0.01, seed=17, svs=20),
    "config5_16": dict(ref_bp=248_000_000, copies=16, div=0.01, seed=17, svs=20),
    "config5": dict(ref_bp=248_000_000, copies=64, div=0.01, seed=17, svs=20),
}
METRIC = "collection Gbp/s to BWT (MS+head sort+emit)"


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for k, nm in enumerate(names):
                    if r[2 + k].lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


def host_info():
    model = "unknown"
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
    except Exception:
        pass
    try:
        allowed = sorted(os.sched_getaffinity(0))
    except Exception:
        allowed = list(range(os.cpu_count() or 1))
    return {"cpu_model": model, "cores_total": os.cpu_count(), "cores_allowed": len(allowed)}, allowed


def _elapsed_ms(stdout):
    for line in stdout.splitlines():
        if line.startswith("==== Time elapsed:"):
            return float(line.split(":")[1].split()[0])
    return None


def _log_ms(path, key):
    try:
        for line in open(path):
            if line.startswith(key):
                return float(line[len(key):].split()[0])
    except Exception:
        pass
    return None


def run_reference_cli(list_file, out_prefix, cpu=None):
    """one run of the unmodified reference binary, pinned to one core; -> dict of its own timers"""
    exe = os.path.join(ROOT, "oracle", "_ref", "cmsbwt")
    cmd = [exe, "-o", out_prefix, list_file]
    if cpu is not None:
        cmd = ["taskset", "-c", str(cpu)] + cmd
    t = time.time()
    r = subprocess.run(cmd, check=True, stdout=subprocess.PIPE, text=True)
    wall = (time.time() - t) * 1e3
    el = _elapsed_ms(r.stdout) or wall
    pre = _log_ms(out_prefix + ".log", "Preprocessing done in") or 0.0
    return {"elapsed_ms": el, "wall_ms": wall, "preprocessing_ms": pre, "path_ms": el - pre}


def same_work(div):
    """CLI against CLI on the SAME files (BASELINE config 2 whole: 5 Mbp reference + 100 copies, plain
    BWT): the unmodified reference binary once on one pinned host core, cms_bwt_b200/cms_bwt three
    times (best).  Both numbers are the programs' own `==== Time elapsed` (main.cpp:126): file parsing,
    index build, the path and the output file included.  -> (same_work dict, cpu_baseline dict)"""
    import genpin
    exe_ref = os.path.join(ROOT, "oracle", "_ref", "cmsbwt")
    exe_b200 = os.path.join(ROOT, "cms_bwt_b200", "cms_bwt")
    if not (os.path.exists(exe_ref) and os.path.exists(exe_b200)):
        return None, None
    info, allowed = host_info()
    cpu = allowed[-1] if allowed else None
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
    wd = tempfile.mkdtemp(prefix="cmssame", dir=base)
    g = dict(ref_bp=5_000_000, copies=100, div=div, seed=11, svs=0)
    pre = os.path.join(wd, "c2")
    genpin.write_files(g["ref_bp"], g["copies"], g["div"], g["seed"], g["svs"], pre)
    nsym = None
    try:
        ref = run_reference_cli(pre + ".txt", pre + ".ref_out", cpu)
        N = os.path.getsize(pre + ".ref_out.bwt")
        nsym = N - (g["copies"] + 1)
        ours = []
        for _ in range(3):
            r = subprocess.run([exe_b200, "-o", pre + ".b200_out", pre + ".txt"], check=True, stdout=subprocess.PIPE,
                               text=True)
            ours.append({"elapsed_ms": _elapsed_ms(r.stdout),
                         "context_ms": _log_ms(pre + ".b200_out.log", "Device context created in"),
                         "path_ms": _log_ms(pre + ".b200_out.log", "Path (MS + head sort + emit + copies) wall:")})
        same = subprocess.run(["cmp", "-s", pre + ".ref_out.bwt", pre + ".b200_out.bwt"]).returncode == 0
        best = min(ours, key=lambda o: o["elapsed_ms"])
        desc = "config2 whole: %d bp reference + %d copies at %g (%.1f Mbp), plain BWT, files in %s" % (
            g["ref_bp"], g["copies"], g["div"], nsym / 1e6, base or "tmp")
        sw = {"same_config": True, "workload": desc, "what": "program wall time, `==== Time elapsed` of both CLIs on the same files",
              "reference_ms": ref["elapsed_ms"], "reference_path_ms": ref["path_ms"], "reference_runs": 1,
              "b200_ms": best["elapsed_ms"], "b200_context_ms": best["context_ms"], "b200_path_ms": best["path_ms"],
              "b200_runs": len(ours), "ratio": ref["elapsed_ms"] / best["elapsed_ms"], "identical_bytes": same,
              "host": info, "reference_pinned_to_cpu": cpu}
        cb = {"value": nsym / 1e9 / (ref["elapsed_ms"] / 1e3), "unit": "Gbp/s", "cores": 1, "kind": "reference",
              "sample": desc + "; whole program incl. index build and file I/O, %.1f s; without its preprocessing: %.4f Gbp/s"
                        % (ref["elapsed_ms"] / 1e3, nsym / 1e9 / (ref["path_ms"] / 1e3)),
              "host": info, "pinned_to_cpu": cpu}
        return sw, cb
    finally:
        import shutil
        shutil.rmtree(wd, ignore_errors=True)


def reference_arm(args, rank):
    """The reference's own CPU implementation on a bounded sample of the workload."""
    if rank != 0:
        return
    exe = os.path.join(ROOT, "oracle", "_ref", "cmsbwt")
    # bounded sample: about 1 s of index build + 0.4 s per 5 Mbp copy on one core; the whole
    # --steps/--warmup run is kept under ~150 s
    runs = max(1, args.steps + args.warmup)
    ncopies = int(max(2, min(40, (150.0 / runs - 1.0) / 0.42)))
    sample = dict(ref_bp=5_000_000, copies=ncopies, div=WORKLOADS[args.workload]["div"], seed=13)
    if not os.path.exists(exe):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/cmsbwt not built"}))
        return
    import genpin
    wd = tempfile.mkdtemp(prefix="cmsref", dir="/dev/shm" if os.access("/dev/shm", os.W_OK) else None)
    genpin.write_files(sample["ref_bp"], sample["copies"], sample["div"], sample["seed"], 0, os.path.join(wd, "s"))
    _, _, Ns, nds = genpin.make_text(sample["ref_bp"], sample["copies"], sample["div"], sample["seed"], 0)
    nsym = int(Ns - nds)
    info, allowed = host_info()
    cpu = allowed[-1] if allowed else None
    times, paths = [], []
    for it in range(args.warmup + args.steps):
        r = run_reference_cli(os.path.join(wd, "s.txt"), os.path.join(wd, "out"), cpu)
        if it >= args.warmup:
            times.append(r["wall_ms"] / 1e3)
            paths.append(r["path_ms"] / 1e3)
    dt = sum(times) / len(times)
    dp = sum(paths) / len(paths)
    # `value` = the path alone (the program's elapsed time minus its own "Preprocessing done in": the
    # MS, head-sort and writer stages with the file parsing and writing they contain), the same scope
    # as the B200 arm's `value`; `e2e` = the whole program (index build included)
    val, val_e2e = nsym / 1e9 / dp, nsym / 1e9 / dt
    desc = ("bounded sample of %s: %d bp reference + %d copies at %g (%.1f Mbp), unmodified reference binary on one "
            "pinned core; value = without its preprocessing (index build), e2e = whole program" % (
                args.workload, sample["ref_bp"], sample["copies"], sample["div"], nsym / 1e6))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "Gbp/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dp * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic", "config": {"workload": args.workload, "sample": desc, "host": info,
                                        "pinned_to_cpu": cpu, "program_ms_per_step": dt * 1e3},
        "cpu_baseline": {"value": val, "unit": "Gbp/s", "cores": 1, "kind": "reference",
                         "sample": desc},
        "e2e": {"value": val_e2e, "unit": "Gbp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_multi(args, rank, world, local, dev, dist):
    """N > 1: one process per GPU, the collection's documents sharded by rank (strong scaling: the
    workload is the same 100-copy collection at every N), cms_bwt_b200.multi over NCCL."""
    import numpy as np
    import torch
    import cms_bwt_b200 as P
    from cms_bwt_b200 import multi
    import gen_device
    w = WORKLOADS[args.workload]
    # pinned host buffers should live on the NUMA node of this rank's GPU: run this process on the
    # CPUs NVML reports as local to the device (first touch then places the pages there)
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = [64 * i + b for i, wd in enumerate(words) for b in range(64) if (wd >> b) & 1]
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass
    # contiguous ranges of copies per rank (all copies have nearly the same length)
    plan = multi.shard_documents([w["ref_bp"]] * w["copies"], world)
    a, b = plan[rank]
    if b <= a:
        raise SystemExit("bench.py: more ranks than documents")
    if w.get("many"):       # shards on batch borders of the generator
        a, b = (a // 2000) * 2000, (w["copies"] if rank == world - 1 else (b // 2000) * 2000)
        ref, text, N, ndocs = gen_device.make_text_many(w["ref_bp"], w["copies"], w["div"], w["seed"], dev,
                                                        first_copy=a, ncopies=b - a)
    else:
        ref, text, N, ndocs = gen_device.make_text(w["ref_bp"], w["copies"], w["div"], w["seed"], dev,
                                                   first_copy=a, ncopies=b - a, svs=w.get("svs", 0))
    torch.cuda.synchronize()
    x = P.augment(ref.cpu().numpy())
    ctx = P.Context(local)
    ctx.index_upload(x)                       # replicated index, built on every GPU, timed separately
    st0 = ctx.stats()
    state = {"phase": {}}
    # the shard plan (symbols and documents per rank) is exchanged once, outside the timed region
    pt = torch.tensor([N, ndocs], dtype=torch.int64, device=dev)
    allp = torch.empty(2 * world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(allp, pt)
    plan = allp.cpu().view(world, 2).tolist()

    # the C++ host layer (cms_bwt_b200/csrc/sharded.cu): the library's own NCCL communicator and exchanges
    multi.comm_init(ctx, rank, world, dev)
    ndocs_total = int(sum(p_[1] for p_ in plan))

    def step():
        sp = multi.run_sharded_native(ctx, dev, text.data_ptr(), N, ndocs, P.MODE_PLAIN)
        sp.ndocs_total = ndocs_total
        state["sp"], state["out"] = sp, sp.out
        for k, v in sp.phase_ms.items():
            state["phase"][k] = state["phase"].get(k, 0.0) + v

    def allmax(v):
        t = torch.tensor([float(v)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(v):
        t = torch.tensor([float(v)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    state["phase"] = {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage = {"ms_ms": 0.0, "sort_ms": 0.0, "emit_ms": 0.0, "ms_kernel_ms": 0.0, "pack_ms": 0.0}
    launches = 0
    t0 = time.time()
    e0.record()
    for _ in range(args.steps):
        step()
        s = ctx.stats()
        for k in stage:
            stage[k] += s[k]
        launches += s["kernel_launches"]
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    wall = time.time() - t0
    dev_ms_local = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = allmax(max(wall * 1e3, dev_ms_local)) / args.steps
    sp = state["sp"]
    nsym = sp.N_total - sp.ndocs_total
    s = ctx.stats()
    H_total = allsum(s["n_heads"])
    launches_total = allsum(launches)
    comm_bytes = allsum(sp.comm_bytes)
    lib_ms = (stage["ms_ms"] + stage["sort_ms"] + stage["emit_ms"]) / args.steps
    lib_ms_max = allmax(lib_ms)

    # e2e: every rank copies its documents H2D and its slice of the BWT D2H inside the timed region
    e2e = None
    if not args.no_e2e:
        host_text = torch.empty(N + 64, dtype=torch.uint8, pin_memory=True)
        host_text.copy_(text[:N + 64])
        dtext = torch.empty(N + 64, dtype=torch.uint8, device=dev)
        host_out = torch.empty(sp.out_len + 64, dtype=torch.uint8, pin_memory=True)

        def e2e_step():
            dtext.copy_(host_text, non_blocking=True)
            torch.cuda.synchronize()
            out = multi.run_sharded_native(ctx, dev, dtext.data_ptr(), N, ndocs, P.MODE_PLAIN).out
            host_out[:out.numel()].copy_(out, non_blocking=True)
            torch.cuda.synchronize()
        e2e_step()
        reps = max(1, min(args.steps, 3))
        dist.barrier()
        t1 = time.time()
        for _ in range(reps):
            e2e_step()
        dist.barrier()
        dt = allmax(time.time() - t1) / reps
        e2e = {"value": nsym / 1e9 / dt, "unit": "Gbp/s", "h2d_bytes_per_step": int(sp.N_total),
               "d2h_bytes_per_step": int(sp.N_total), "ms_per_step": dt * 1e3}
    if rank == 0:
        peak, peak_kind = measured_peak()
        kms = stage["ms_kernel_ms"] / args.steps
        alg_bytes = N + 16 * s["n_heads"]
        achieved = alg_bytes / 1e9 / (kms / 1e3)
        n = len(x)
        path_bytes = nsym + sp.N_total + 30 * n * world + 48 * H_total
        # per-stage roofline of rank 0 (library stage timers on its stream); algorithmic bytes of the whole
        # job split evenly over the ranks, the replicated index term 30 n counted once per rank
        st_ms = {"stage1_ms": stage["ms_ms"] / args.steps, "stage2_head_order": stage["sort_ms"] / args.steps,
                 "stage3_emit": stage["emit_ms"] / args.steps}
        st_alg = {"stage1_ms": nsym / world, "stage2_head_order": 48 * H_total / world,
                  "stage3_emit": sp.N_total / world + 30 * n}
        stages_m = {k: {"alg_bytes": int(st_alg[k]), "ms": st_ms[k],
                        "frac": st_alg[k] / 1e9 / (st_ms[k] / 1e3) / peak if st_ms[k] else None} for k in st_ms}
        dom = max(st_ms, key=lambda k: st_ms[k])
        roofline_multi = {"bound": "hbm", "kernel": "%s on rank 0 (kernel time of the stage, exchanges excluded)" % dom,
                          "achieved": st_alg[dom] / 1e9 / (st_ms[dom] / 1e3), "peak": peak, "peak_kind": peak_kind,
                          "unit": "GB/s", "frac": stages_m[dom]["frac"], "traffic": None,
                          "alg_bytes_per_launch": int(st_alg[dom]), "kernel_ms": st_ms[dom], "stages": stages_m,
                          "ms_kernel": {"ms": kms, "alg_bytes": int(alg_bytes)},
                          "path_alg_bytes": path_bytes,
                          "path_frac": path_bytes / 1e9 / (ms_per_step / 1e3) / (peak * world)}
        line = {
            "metric": METRIC, "value": nsym / 1e9 / (ms_per_step / 1e3), "unit": "Gbp/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "%s: %d bp random ACGT reference + %d mutated copies at %g (SNP/ins/del)%s, "
                                   "plain BWT, documents sharded over %d GPUs" %
                                   (args.workload, w["ref_bp"], w["copies"], w["div"],
                                    " + %d block SVs per copy" % w["svs"] if w.get("svs") else "", world),
                       "N_symbols": int(nsym), "documents": int(sp.ndocs_total), "n_ref_augmented": int(n),
                       "heads": int(H_total), "cache": "inputs (%.1f GB per GPU) larger than L2" % (N / 1e9),
                       "parallelism": "documents sharded (stage 1), tails by key range, tiles by reference-"
                                      "position range, output by SA-rank range (pull emit); index replicated; host layer = "
                                      "cmsbwt_run_sharded (C++, NCCL grouped send/recv + all-gather on the library's stream)",
                       "device_ms_per_step_rank0": dev_ms_local / args.steps,
                       "kernel_stage_ms_max": lib_ms_max,
                       "stage_ms_rank0": {k: v / args.steps for k, v in stage.items()},
                       "phase_ms_rank0": {k: round(v / args.steps, 3) for k, v in state["phase"].items()},
                       "exchange_bytes_per_step": comm_bytes, "collectives_per_step": sp.collectives,
                       "rank_rounds": sp.rank_rounds,
                       "index_ms": {"sa_build_gpu": st0["sa_build_ms"], "index_build": st0["index_build_ms"]}},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches_total),
            "roofline": roofline_multi,
            "cpu_baseline": None,
        }
        print(json.dumps(line))
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="config3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return
    if args.warmup < 3:
        args.warmup = 3

    import numpy as np
    import torch
    import cms_bwt_b200 as P
    import gen_device
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    if world > 1:
        run_multi(args, rank, world, local, dev, dist)
        return

    w = WORKLOADS[args.workload]
    if w.get("many"):
        ref, text, N, ndocs = gen_device.make_text_many(w["ref_bp"], w["copies"], w["div"], w["seed"], dev)
    else:
        ref, text, N, ndocs = gen_device.make_text(w["ref_bp"], w["copies"], w["div"], w["seed"], dev,
                                                   svs=w.get("svs", 0))
    torch.cuda.synchronize()
    nsym = N - ndocs
    x = P.augment(ref.cpu().numpy())
    ctx = P.Context(local)
    for kv in os.environ.get("CMSBWT_PARAMS", "").split():          # tuning experiments only
        k, v = kv.split("=")
        ctx.set_param(k, int(v))
    ctx.index_upload(x)                       # SA + index built on the GPU, timed separately
    st0 = ctx.stats()

    def step():
        ctx.ms_run_device(text.data_ptr(), N)
        ctx.heads_sort()
        ctx.emit(P.MODE_PLAIN)

    sampler = ClockSampler(local)
    sampler.start()                            # nvidia-smi needs ~0.1 s to come up: start it before
    for _ in range(args.warmup):               # the warm-up (same load as the timed steps)
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    profiling = os.environ.get("CMSBWT_PROFILE") == "1"      # ncu --profile-from-start off
    if profiling:
        torch.cuda.profiler.start()
    stage = {"ms_ms": 0.0, "sort_ms": 0.0, "emit_ms": 0.0, "ms_kernel_ms": 0.0, "pack_ms": 0.0}
    groups = {g: 0.0 for g in P.Stats.GROUPS}
    launches = 0
    t0 = time.time()
    e0.record()
    for _ in range(args.steps):
        step()                                 # every call synchronises the library's stream
        s = ctx.stats()
        for k in stage:
            stage[k] += s[k]
        for g in groups:
            groups[g] += s["group_ms"][g]
        launches += s["kernel_launches"]
    e1.record()
    torch.cuda.synchronize()
    wall = time.time() - t0
    if profiling:
        torch.cuda.profiler.stop()
    clocks = sampler.stop()
    s = ctx.stats()
    # device time of the timed region = sum of the library's CUDA-event stage times (its own
    # stream); wall clock over the same region is reported beside it.
    dev_ms = (stage["ms_ms"] + stage["sort_ms"] + stage["emit_ms"]) / args.steps
    ms_per_step = wall * 1e3 / args.steps
    value = nsym / 1e9 / (ms_per_step / 1e3)
    H = s["n_heads"]
    n = len(x)

    # ---- roofline: algorithmic bytes of the path = N_sym (every collection symbol read
    # once) + N (every BWT byte written once) + 30 n (index arrays read once, bucket counts written and
    # read) + 48 H (one 24-byte head record written once and read once), split by stage; time = the
    # CUDA-event time of the stage's kernel groups on the library's stream (stats.group_ms, live in
    # this run).  `roofline` describes the dominant stage (largest share of the step); `traffic` is
    # the DRAM read+write of its kernels per launch from the committed ncu --set full capture of the
    # same command (profiles/traffic.json, made by tools/ncu_traffic.py).
    peak, peak_kind = measured_peak()
    grp = {k: v / args.steps for k, v in groups.items()}
    stage_groups = {"stage1_ms": ("pack", "ms", "gather"), "stage2_head_order": ("rank", "tiles", "place"),
                    "stage3_emit": ("offsets", "fill", "rle")}
    stage_alg = {"stage1_ms": nsym, "stage2_head_order": 48 * H, "stage3_emit": N + 30 * n}
    stage_kernels = {"stage1_ms": "k_pack + k_ms_fast + k_msf_gather" if s.get("fast_path") else "k_ms + k_compact",
                     "stage2_head_order": "k_link, k_tail_label, tail radix sort + refinement, k_tile_hist/scatter, k_tile_rank",
                     "stage3_emit": "k_cnt_sa, offsets scan, k_fill_part, k_fill"}
    traffic_tab, traffic_capture = {}, None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
        traffic_tab = tj.get(args.workload, {})
        traffic_capture = tj.get("_capture")
    except Exception:
        pass
    stages = {}
    for name, gs in stage_groups.items():
        ms = sum(grp[g] for g in gs)
        tr = [traffic_tab.get(g) for g in gs if traffic_tab.get(g) is not None]
        stages[name] = {"alg_bytes": int(stage_alg[name]), "ms": ms, "GBps": stage_alg[name] / 1e9 / (ms / 1e3) if ms else None,
                        "frac": stage_alg[name] / 1e9 / (ms / 1e3) / peak if ms else None,
                        "groups_ms": {g: grp[g] for g in gs}, "traffic": int(sum(tr)) if tr else None}
    dom = max(stages, key=lambda k: stages[k]["ms"])
    path_alg = nsym + N + 30 * n + 48 * H
    roofline = {"bound": "hbm", "kernel": "%s (%s)" % (dom, stage_kernels[dom]),
                "achieved": stages[dom]["GBps"], "peak": peak, "peak_kind": peak_kind, "unit": "GB/s",
                "frac": stages[dom]["frac"], "traffic": stages[dom]["traffic"],
                "alg_bytes_per_launch": stages[dom]["alg_bytes"], "kernel_ms": stages[dom]["ms"],
                "share_of_step": stages[dom]["ms"] / dev_ms,
                "stages": stages, "path_alg_bytes": int(path_alg),
                "path_frac": path_alg / 1e9 / (dev_ms / 1e3) / peak,
                "path_frac_of_8TBps_nominal": path_alg / 1e9 / (dev_ms / 1e3) / 8000.0,
                "ms_kernel": {"name": "k_ms_fast" if s.get("fast_path") else "k_ms", "ms": grp["ms"],
                              "alg_bytes": int(N // 4 + 16 * H),
                              "note": "its own inputs/outputs: the 2-bit text (N/4) + one 16-byte record per head"}}
    if traffic_capture:
        roofline["traffic_capture"] = traffic_capture

    # ---- e2e: host buffers through cmsbwt_run (H2D text + D2H BWT inside the timed region)
    e2e = None
    if not args.no_e2e:
        host_text = torch.empty(N, dtype=torch.uint8, pin_memory=True)
        host_text.copy_(text[:N])
        host_out = torch.empty(N, dtype=torch.uint8, pin_memory=True)
        torch.cuda.synchronize()
        ht, ho = host_text.numpy(), host_out.numpy()
        ctx.run(ht, P.MODE_PLAIN, ho)
        reps = max(1, min(args.steps, 3))
        t1 = time.time()
        for _ in range(reps):
            ctx.run(ht, P.MODE_PLAIN, ho)
        dt = (time.time() - t1) / reps
        e2e = {"value": nsym / 1e9 / dt, "unit": "Gbp/s", "h2d_bytes_per_step": int(N),
               "d2h_bytes_per_step": int(N), "ms_per_step": dt * 1e3}
        # cheap self-check: same bytes as the device-resident run
        chk = torch.from_numpy(ho[:1 << 20]).to(dev)
    ctx.close()
    sw, cb = (None, None) if args.no_cpu_baseline else same_work(w["div"])
    line = {
        "metric": METRIC, "value": value, "unit": "Gbp/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "%s: %d bp random ACGT reference + %d mutated copies at %g "
                               "(SNP/ins/del)%s, plain BWT" % (args.workload, w["ref_bp"], w["copies"], w["div"],
                                                               " + %d block SVs per copy" % w["svs"] if w.get("svs") else ""),
                   "N_symbols": int(nsym), "documents": int(ndocs), "n_ref_augmented": int(n),
                   "heads": int(H), "tails": int(s["n_tails"]), "tails_tied_after_radix_sort": int(s["rank_tied"]),
                   "rank_rounds": int(s["rank_rounds"]), "tile_entries": int(s["n_items"]),
                   "escapes": int(s["n_escapes"]), "cache": "inputs (%.1f GB) larger than L2" % (N / 1e9),
                   "device_ms_per_step": dev_ms,
                   "stage_ms": {k: v / args.steps for k, v in stage.items()},
                   "index_ms": {"sa_build_gpu": st0["sa_build_ms"], "index_build": st0["index_build_ms"]}},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
        "cpu_baseline": cb, "same_work": sw,
    }
    print(json.dumps(line))


if __name__ == "__main__":
    main()

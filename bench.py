"""bench.py -- HyDia approach 5 encrypted match latency on B200 (BASELINE.json metric).

A step = one query through DiagonalSender.membershipScenario + indexScenario (/root/reference/src/main.cpp:343,361), each
recomputing the similarity as the reference does (src/sender/sender_diag.cpp:12-63).

Workloads:
  --gpus 1           BASELINE configs[1]: 2^14-vector synthetic DB = 1 block of 512 diagonal ciphertexts (N = 2^15, 12 + 4 limbs);
                     the line also carries `share_of_configs4`: the same step on 8 blocks, one GPU's share of configs[4]
  --gpus 2 | 4 | 8   BASELINE configs[4]: 2^20 vectors = 64 blocks sharded contiguously over the ranks (strong scaling: the
                     database is the same for every N >= 2); --total-blocks overrides
  --impl reference   the CPU arm: the oracle port of the same path on all host cores (oracle/cpu_path.py; OpenFHE is not
                     installable here), rank 0 only
Every field of the JSON line is described in DESIGN.md "Measurement".
"""
import argparse
import json
import os
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "encrypted_match_latency_membership_plus_index"
UNIT = "s"
SLOTS = 16384
# BASELINE.md "HyDia, approach 5": published Sender compute seconds (membership, index) per DB size on a 48-thread
# Xeon Gold 5412U (tools/figures/approach5.csv) -- other hardware, not re-measured; vs_baseline = value / (m + i)
PUBLISHED_APPROACH5 = {1 << 14: (4.518783, 4.175511), 1 << 15: (6.155975, 5.85484), 1 << 16: (9.560245, 9.011757),
                       1 << 17: (16.00675, 15.46935), 1 << 18: (28.72538, 27.62988), 1 << 19: (53.24866, 49.70151),
                       1 << 20: (102.411, 96.52172)}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--total-blocks", type=int, default=0, help="database blocks of 16384 vectors over all GPUs (0: 1 at --gpus 1, 64 = configs[4] otherwise)")
    ap.add_argument("--ring-dim", type=int, default=0, help="0 = the 128-bit ring of the reference (2^15); smaller rings are for quick checks only")
    ap.add_argument("--ndiag", type=int, default=512)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-share", action="store_true", help="--gpus 1: skip the extra 8-block measurement (share_of_configs4)")
    ap.add_argument("--cpu-budget-s", type=float, default=200.0, help="--impl reference: wall-clock budget that sizes the per-step sample")
    ap.add_argument("--rot-chunk", type=int, default=0, help="override the engine's hoisted-rotation batch size")
    ap.add_argument("--rot-streams", type=int, default=0, help="override the number of streams the hoisted-rotation batches use")
    ap.add_argument("--ks-moddown-w", type=int, default=-1, help="override the engine's warp-task ModDown kernel (0 = the r1 launch sequence)")
    ap.add_argument("--rot-lane", type=int, default=-1, help="override the engine's rotation lane (0 = all work ordered on the engine stream)")
    ap.add_argument("--shard", default="diagonals", choices=["diagonals", "blocks"],
                    help="N > 1: 'diagonals' = every rank holds its diagonal range of all blocks, computes only its rotations, partial sums "
                         "reduce-scattered; 'blocks' = round 1 (whole blocks per rank, all rotations on every rank)")
    ap.add_argument("--sync-exchange", action="store_true", help="N > 1: host-synchronous result exchange (no overlap of the scenarios)")
    ap.add_argument("--ncu-range", action="store_true", help="run one extra step inside cudaProfilerStart/Stop (for ncu --profile-from-start off)")
    a = ap.parse_args()
    if a.total_blocks <= 0:
        a.total_blocks = 1 if a.gpus == 1 else 64
    if a.total_blocks % a.gpus:
        raise SystemExit(f"--total-blocks {a.total_blocks} is not a multiple of --gpus {a.gpus}")
    return a


def slots_of(args):
    return (args.ring_dim // 2) if args.ring_dim else SLOTS


def vs_published(args, seconds):
    if args.ring_dim or args.ndiag != 512:
        return None
    pub = PUBLISHED_APPROACH5.get(args.total_blocks * SLOTS)
    return seconds / sum(pub) if pub else None


def synth_db(n, seed, match_rows):
    """Seeded restatement of tools/gen_dataset.sh: rows in [-99, 99], match rows in {1,2,3}, query all ones."""
    rng = np.random.default_rng(seed)
    db = rng.integers(-99, 100, size=(n, 512)).astype(np.float64)
    for r in match_rows:
        db[r] = rng.integers(1, 4, size=512)
    return db


def workload_config(args, total_blocks=None):
    tb = total_blocks or args.total_blocks
    full = not args.ring_dim and args.ndiag == 512
    name = " = BASELINE configs[1]" if (tb == 1 and full) else (" = BASELINE configs[4]" if (tb == 64 and full) else "")
    return {"workload": f"HyDia approach 5, membership + index scenarios, synthetic DB of {tb * slots_of(args)} vectors x 512 dims "
                        f"({tb} block(s) of {args.ndiag} diagonal ciphertexts, {tb // args.gpus} per GPU){name}",
            "approach": 5, "ring_dim": args.ring_dim or 32768, "limbs": "12+4", "total_blocks": tb, "blocks_per_gpu": tb // args.gpus,
            "ndiag": args.ndiag, "sharding": (args.shard if args.gpus > 1 else None),
            "scaling_note": "N = 1 runs configs[1] (1 block); every N >= 2 runs the same 64-block database (strong scaling)",
            "l2_policy": "inputs larger than L2: each step streams 13.3 GB of rotation keys and 3.2 GB of database per block (L2 = 126 MB)"}


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, device):
        super().__init__(daemon=True)
        self.device, self.samples, self.reasons, self.stop_flag, self.max_mhz = device, [], set(), False, None

    def run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.device)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            names = {"hw_slowdown": pynvml.nvmlClocksThrottleReasonHwSlowdown,
                     "hw_thermal_slowdown": pynvml.nvmlClocksThrottleReasonHwThermalSlowdown,
                     "sw_thermal_slowdown": pynvml.nvmlClocksThrottleReasonSwThermalSlowdown,
                     "sw_power_cap": pynvml.nvmlClocksThrottleReasonSwPowerCap}
            while not self.stop_flag:
                self.samples.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
                time.sleep(0.02)
        except Exception as e:   # clocks are reporting, not compute
            self.reasons.add(f"sampler_error:{type(e).__name__}")

    def result(self):
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# ---------------------------------------------------------------------------------------------------- CPU arm
def cpu_arm(args, total_blocks, steps, warmup, budget_s):
    """The oracle port of sender_diag.cpp:12-63 on all host cores (oracle/cpu_path.py restates the reference's OpenMP loops).
    A step runs the path IN FULL when the workload is one block (configs[1]); for more blocks the per-block part (512 products +
    511 adds + relinearise + rescale + compare) is run on a bounded sample of blocks and scaled -- the fixed part (511 hoisted
    rotations per scenario) always runs in full.  Two database modes: (ii) ciphertexts in RAM -- the
    reported value -- and (i) re-read from files inside the timed region as sender_diag.cpp:87-91 does, timed once.
    Returns (seconds per step [list], description dict)."""
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)   # torchrun exports 1 to its workers
    from oracle import ckks_oracle as O
    from oracle.cpu_path import CpuDiagPath, DiskDb, host_threads
    ndiag = args.ndiag
    threads = host_threads()
    P = O.gen_params(11, ring_dim=args.ring_dim or None)
    orc = O.Oracle(P)
    orc.keygen(1)
    orc.mult_keygen()
    orc.rot_keygen(list(range(1, ndiag)) + [1 << i for i in range(int(np.log2(P.slots)))])
    rng = np.random.default_rng(0)
    ndist = 64   # distinct database ciphertexts (encryption time; the arithmetic per diagonal does not care)
    base = [orc.encrypt(rng.uniform(-0.04, 0.04, P.slots)) for _ in range(ndist)]
    q = orc.encrypt(rng.uniform(-0.04, 0.04, P.slots))
    ram = lambda g, d: base[(d * 7 + g) % ndist]
    cp = CpuDiagPath(orc, threads)

    def timed_step(load, blocks):
        t0 = time.perf_counter()
        cp.membership(q, load, blocks, ndiag)
        cp.index(q, load, blocks, ndiag)
        return time.perf_counter() - t0

    # size the sample: the fixed part in full, as many blocks as the budget allows
    t_one = timed_step(ram, [0])                          # also the first warm-up step (pages the library and the keys in)
    sample_blocks, t_fixed = total_blocks, 0.0
    if total_blocks > 1:
        t0 = time.perf_counter()
        dg = orc.fast_rotation_precompute(q)
        list(cp.pool.map(lambda i: orc.fast_rotation(q, i, dg), range(1, ndiag)))
        t_fixed = 2 * (time.perf_counter() - t0)
        t_block = max(t_one - t_fixed, 1e-3)
        per_step_budget = budget_s / max(1, steps + warmup)
        sample_blocks = int(max(1, min(total_blocks, (per_step_budget - t_fixed) // t_block)))
    blocks = list(range(sample_blocks))
    scale = lambda t: t if sample_blocks == total_blocks else t_fixed + (t - t_fixed) * (total_blocks / sample_blocks)
    vals = []
    for i in range(max(0, warmup - 1) + steps):
        t = timed_step(ram, blocks)
        if i >= max(0, warmup - 1):
            vals.append(scale(t))
    # mode (i): database re-read from disk inside the timed region, one step
    disk_s = None
    try:
        with tempfile.TemporaryDirectory(prefix="hx_cpu_db_") as td:
            ddb = DiskDb(td)
            for k, c in enumerate(base):
                ddb.store(0, k, c)
            disk_s = scale(timed_step(lambda g, d: ddb.load(0, (d * 7 + g) % ndist), blocks))
    except OSError:
        pass
    cp.close()
    desc = {"kind": "port", "cores": threads,
            "sample": (f"oracle port (Python + C/OpenMP, oracle/cpu_path.py) of sender_diag.cpp:12-63, N={P.N}, {ndiag} diagonals: "
                       + ("the whole path per step (2 x 511 hoisted rotations, 2 x 512 products, relinearise, rescale, compare, EvalSum)"
                          if sample_blocks == total_blocks else
                          f"2 x 511 hoisted rotations in full + {sample_blocks} of {total_blocks} blocks (512 products, 511 adds, relinearise, "
                          f"rescale, compare each), block part scaled to {total_blocks}")
                       + f"; database ciphertexts in RAM ({ndist} distinct, tiled)"),
            "value_db_from_disk": disk_s, "blocks_run": sample_blocks}
    return vals, desc


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  OpenFHE v1.2.3 is not buildable here (un-vendored,
    DESIGN.md), so this is the oracle port on all host cores; rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    vals, desc = cpu_arm(args, args.total_blocks, args.steps, args.warmup, args.cpu_budget_s)
    v = float(np.mean(vals))
    n_total = args.total_blocks * slots_of(args)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": len(vals),
            "warmup": args.warmup, "ms_per_step": v * 1e3, "higher_is_better": False, "scaling": "strong" if args.gpus > 1 else "weak",
            "vs_baseline": vs_published(args, v), "dtype": "u64", "data": "synthetic", "config": workload_config(args),
            "vectors_per_s": n_total / v,
            "cpu_baseline": dict(desc, value=v, unit=UNIT),
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------- B200 arm
def stamped(path, stamp):
    """A committed profile summary (profiles/*.json) counts only while it was captured from the kernels this library holds."""
    try:
        d = json.load(open(path))
    except Exception:
        return None
    return d if d.get("kernel_stamp") == stamp else None


def run_b200(args):
    import torch
    import torch.distributed as dist

    import image_matching_b200 as hx

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1")   # the collectives' kernels must not queue behind the low-priority rotation lanes
        dist.init_process_group("nccl", device_id=dev)

    ndiag = args.ndiag
    cc = hx.CryptoContext.generate(depth=11, ring_dim=args.ring_dim, device=local)
    for name, v in (("rot_chunk", args.rot_chunk or -1), ("rot_streams", args.rot_streams or -1), ("ks_moddown_w", args.ks_moddown_w),
                    ("rot_lane", args.rot_lane)):
        if v >= 0:
            cc.set_option(name, v)
    for kv in filter(None, os.environ.get("HYDIA_ENGINE_OPTS", "").split(",")):   # diagnostic: any hx_ctx_set_option pair, name=value
        cc.set_option(kv.split("=")[0], int(kv.split("=")[1]))
    slots = cc.GetBatchSize()
    cl = hx.Client(cc, seed=1)                     # same seed on every rank: replicated key set
    cl.EvalMultKeyGen()
    cl.EvalSumKeyGen()
    by_diag = world > 1 and args.shard == "diagonals"
    if by_diag:
        from image_matching_b200.sharded import diag_range
        d0, d1 = diag_range(rank, world, ndiag)
        cc.set_option("db_diag0", d0)
        cc.set_option("db_diag_count", d1 - d0)
        cl.EvalRotateKeyGen(list(range(max(d0, 1), d1)))   # a rank needs the keys of its own rotations only
    else:
        d0, d1 = 0, ndiag
        cl.EvalRotateKeyGen(list(range(1, ndiag)))

    def enroll(total_blocks):
        """matches at rows 0 and n-1.  'blocks': global block g lives on rank g // bpg as local block g % bpg; 'diagonals': every
        rank enrolls its diagonal range of every block under the global block id"""
        bpg = total_blocks // world
        for g in (range(total_blocks) if by_diag else range(rank * bpg, (rank + 1) * bpg)):
            match = ([0] if g == 0 else []) + ([slots - 1] if g == total_blocks - 1 else [])
            db = synth_db(slots, seed=100 + g, match_rows=match)
            rows = hx.diagonal_pack(db, slots)
            for d in range(d0, d1):
                cl.enroll_row(g if by_diag else g - rank * bpg, d, rows[d])
            del db, rows
        return sorted({0, total_blocks * slots - 1})

    # query: encrypted once on rank 0, kept in pinned host memory (what a Receiver sends)
    ct_words = 2 * cc.L * cc.N
    q_host = torch.empty(ct_words, dtype=torch.int64).pin_memory()
    if rank == 0:
        q0 = cl.encryptQuery(np.ones(512))[0]
        q0.download_ptr(q_host.data_ptr())
        del q0
    if world > 1:
        q_dev = q_host.to(dev)
        dist.broadcast(q_dev, src=0)
        q_host.copy_(q_dev.cpu())
    res_words = 2 * cc.N

    def upload_query_from_host():
        if world == 1:
            return cc.upload_ptr(q_host.data_ptr(), 2, cc.L)
        q_dev = torch.empty(ct_words, dtype=torch.int64, device=dev)
        if rank == 0:
            q_dev.copy_(q_host, non_blocking=True)
        dist.broadcast(q_dev, src=0)
        torch.cuda.synchronize()
        return cc.upload_dev(q_dev.data_ptr(), 2, cc.L)

    def fence():
        cc.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def measure(total_blocks, steps, warmup, profile):
        """warm-up, then `steps` steps with the query resident in HBM and `steps` steps end to end (query from pinned host memory,
        results back to host), each arm timed in two halves that alternate (value, e2e, value, e2e) so that neither arm owns the
        warmer half of the run; device time from CUDA events on the engine stream with a host clock beside it"""
        n_total = total_blocks * slots
        if world > 1:
            from image_matching_b200.sharded import TorchTransport
            snd = hx.ShardedDiagonalSender(cc, None, n_total, device=dev, transport=TorchTransport(cc, dev, async_exchange=not args.sync_exchange),
                                           shard=args.shard)
        else:
            snd = hx.DiagonalSender(cc, None, n_total)
        idx_host = torch.empty((total_blocks, res_words), dtype=torch.int64).pin_memory()
        mem_host = torch.empty(res_words, dtype=torch.int64).pin_memory()

        e2e_parts = os.environ.get("HYDIA_BENCH_E2E_PARTS", "ud")   # diagnostic: "u" = upload only, "d" = download only
        trace2 = os.environ.get("HYDIA_BENCH_TRACE") == "2"
        trace = os.environ.get("HYDIA_BENCH_TRACE") == "1"   # host-clock phases of the end-to-end step on rank 0 (diagnostic)

        def step(q, to_host):
            t = [time.perf_counter()]
            m = snd.membershipScenario([q])
            t.append(time.perf_counter())
            idx = snd.indexScenario([q])
            t.append(time.perf_counter())
            if to_host and rank == 0:
                m.download_ptr(mem_host.data_ptr())
                t.append(time.perf_counter())
                cc.download_many_ptr(idx, idx_host.data_ptr())
                t.append(time.perf_counter())
            if trace and rank == 0:
                print("trace to_host=%d issue_membership %.2f issue_index %.2f" % (to_host, (t[1] - t[0]) * 1e3, (t[2] - t[1]) * 1e3)
                      + (" wait+download_membership %.2f download_index %.2f" % ((t[3] - t[2]) * 1e3, (t[4] - t[3]) * 1e3) if len(t) > 3 else ""),
                      file=sys.stderr, flush=True)
            return m, idx

        def timed(nsteps, e2e):
            fence()
            t0 = time.perf_counter()
            cc.timer_start()
            out = None
            for _ in range(nsteps):
                if trace2:
                    cc.trace_start()
                ta = time.perf_counter()
                q = upload_query_from_host() if (e2e and "u" in e2e_parts) else q_res
                tb = time.perf_counter()
                out = step(q, e2e and "d" in e2e_parts)
                tc = time.perf_counter()
                cc.sync()   # a step is one query's latency: the next query starts when this one's results are complete
                if trace2:
                    cc.trace_mark("step_end")
                    mk = cc.trace_dump()
                    print("timeline timed e2e=%d rank %d: %s" % (e2e, rank, json.dumps({k: [round(x, 1) for x in v] for k, v in mk.items()})), file=sys.stderr, flush=True)
                if trace and rank == 0:
                    print("trace e2e=%d upload %.2f step %.2f sync %.2f" % (e2e, (tb - ta) * 1e3, (tc - tb) * 1e3, (time.perf_counter() - tc) * 1e3),
                          file=sys.stderr, flush=True)
            cc.sync()
            torch.cuda.synchronize()
            ms = cc.timer_stop()
            wall_ms = (time.perf_counter() - t0) * 1e3
            if world > 1:
                t = torch.tensor([ms, wall_ms], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms, wall_ms = float(t[0].item()), float(t[1].item())
            return ms, wall_ms, out

        q_res = upload_query_from_host()
        prof_all, dominant = {}, None
        if profile:   # the first warm-up step runs with every kernel family timed to find the dominant one
            cc.prof_reset()
            cc.prof_enable(None)
            step(q_res, False)
            prof_all = cc.prof_read()
            dominant = max(prof_all, key=lambda k: prof_all[k]["ms"])
            cc.prof_enable([])
        nw = max(0, warmup - (1 if profile else 0))
        for i in range(nw):
            if i == nw - 1:   # the last warm-up step takes the end-to-end route (pinned upload, results to host) so that arm is warm too
                step(upload_query_from_host(), True)
            else:
                step(q_res, False)
            cc.sync()
        cc.sync()
        if profile:
            cc.prof_reset()
            cc.prof_enable([dominant])
        l0 = cc.launches()
        halves = [n for n in (steps // 2, steps - steps // 2) if n]
        remeasured = False
        while True:
            ms_v = ms_e = wall_v = wall_e = 0.0
            launches = 0
            out = None
            per_half = []
            prof = {dominant: {"launches": 0, "ms": 0.0, "bytes": 0.0}} if profile else {}
            for n in halves:
                l0 = cc.launches()
                p0 = cc.prof_read()[dominant] if profile else None
                a, b, out = timed(n, False)
                ms_v += a; wall_v += b
                launches += cc.launches() - l0
                if profile:
                    # launches and bytes count whether or not the family is being timed, its milliseconds only while it is:
                    # take all three as differences around the value half, so that they describe the same launches
                    p1 = cc.prof_read()[dominant]
                    for k in ("launches", "ms", "bytes"):
                        prof[dominant][k] += p1[k] - p0[k]
                    cc.prof_enable([])      # the dominant family is timed in the value halves only
                a2, b2, _ = timed(n, True)
                ms_e += a2; wall_e += b2
                per_half.append((a / n, a2 / n))
                if profile:
                    cc.prof_enable([dominant])
            # the two halves of an arm time the same work: a gap above 25 % is a disturbance (first-touch allocation, a clock
            # transition), not a property of the path -- measure once more and say so
            spread = max(max(h[i] for h in per_half) / min(h[i] for h in per_half) for i in (0, 1))
            if spread <= 1.25 or remeasured:
                break
            remeasured = True
            if profile:
                cc.prof_reset()
                cc.prof_enable([dominant])
        if profile:
            cc.prof_enable([])
        fence()
        return {"sec": ms_v / steps / 1e3, "sec_e2e": ms_e / steps / 1e3, "wall_sec": wall_v / steps / 1e3, "wall_sec_e2e": wall_e / steps / 1e3,
                "halves_ms": [[round(x, 3) for x in h] for h in per_half], "remeasured": remeasured,
                "launches": launches, "prof": prof, "prof_all": prof_all, "dominant": dominant, "out": out, "n_total": n_total, "snd": snd, "q_res": q_res, "mem_host": mem_host, "idx_host": idx_host}

    expected_index = enroll(args.total_blocks)
    sampler = ClockSampler(local)
    sampler.start()
    r = measure(args.total_blocks, args.steps, args.warmup, profile=True)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    m, idx = r["out"]

    if os.environ.get("HYDIA_BENCH_TRACE") == "1":
        # one more step with the engine's timeline marks on: where a step's time goes on this rank (ms since the step's start)
        for route in ("value", "e2e"):
            fence()
            cc.trace_start()
            qn = upload_query_from_host() if route == "e2e" else r["q_res"]
            tm, ti = r["snd"].membershipScenario([qn]), r["snd"].indexScenario([qn])
            if route == "e2e" and rank == 0:
                tm.download_ptr(r["mem_host"].data_ptr())
                cc.download_many_ptr(ti, r["idx_host"].data_ptr())
            cc.trace_mark("step_end")
            marks = cc.trace_dump()
            del tm, ti, qn
            print("timeline %s rank %d: %s" % (route, rank, json.dumps({k: [round(x, 2) for x in v] for k, v in marks.items()})),
                  file=sys.stderr, flush=True)
        fence()

    if args.ncu_range:
        fence()
        torch.cuda.cudart().cudaProfilerStart()
        qn = upload_query_from_host()
        r["snd"].membershipScenario([qn])
        r["snd"].indexScenario([qn])
        fence()
        torch.cuda.cudart().cudaProfilerStop()

    # correctness of what was timed (outside the timed region): decisions as receiver_hers.cpp:26-54
    ok = True
    if rank == 0:
        got_idx = cl.decryptIndex(idx)
        got_mem = cl.decryptMembership(m)
        ok = (got_idx == expected_index) and got_mem
        if not ok:
            print(f"bench.py: WRONG RESULT index={got_idx[:8]} (expected {expected_index}) membership={got_mem}", file=sys.stderr)
    launches = r["launches"]
    if world > 1:
        lt = torch.tensor([launches], dtype=torch.int64, device=dev)
        dist.all_reduce(lt)
        launches = int(lt.item())

    # --gpus 1: one GPU's share of configs[4] (8 blocks) as an extra measurement on the same context
    share = None
    if world == 1 and not args.no_share and not args.ring_dim and args.total_blocks == 1 and not args.ncu_range:
        del m, idx
        r["out"] = None
        exp8 = enroll(8)
        nshare = max(2, args.steps // 2)
        r8 = measure(8, nshare, 2, profile=False)
        m8, idx8 = r8["out"]
        ok8 = (cl.decryptIndex(idx8) == exp8) and cl.decryptMembership(m8)
        ok = ok and ok8
        share = {"workload": "8 blocks (2^17 vectors) on one GPU = one GPU's database share of BASELINE configs[4]; the 511 hoisted rotations are "
                             "computed in full here (at N > 1 they are sharded over the ranks)",
                 "value": r8["sec"], "e2e": r8["sec_e2e"], "unit": UNIT, "steps": nshare, "results_ok": bool(ok8),
                 "vectors_per_s": 8 * slots / r8["sec"]}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 0)) or 6500.0
        peak_src = "MEASURED_PEAKS.json hbm_gbs" if peaks.get("hbm_gbs") else "fallback 6500 GB/s (B200_PROFILING.md)"
        dominant, prof, prof_all = r["dominant"], r["prof"], r["prof_all"]
        d = prof[dominant]
        stamp = cc.kernel_stamp()
        tr = stamped(os.path.join(ROOT, "profiles", "r2_traffic.json"), stamp)
        traffic = (tr or {}).get(dominant, {}).get("traffic_bytes_per_launch")
        ls = stamped(os.path.join(ROOT, "profiles", "r2_launch_summary.json"), stamp)
        # a "launch" of a fused family is its group of kernels: P-limb kernel, ModDown middle (1 warp-task kernel, or 2 with
        # --ks-moddown-w 0), Q-limb kernel once for the FP64 limbs and once for the 60-bit limb; 4 for ModUp and rescale
        ks_group = 5 if args.ks_moddown_w == 0 else 4
        groups = max(1, d["launches"] // {"ks_hoisted": ks_group, "ks_fused": ks_group, "modup": 4, "rescale": 4}.get(dominant, 1))
        avg_ms = d["ms"] / groups
        achieved = (d["bytes"] / groups) / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
        sec, sec_e2e = r["sec"], r["sec_e2e"]
        line = {"metric": METRIC, "value": sec, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": sec * 1e3, "higher_is_better": False, "scaling": "strong" if world > 1 else "weak",
                "vs_baseline": vs_published(args, sec), "dtype": "u64", "data": "synthetic", "config": workload_config(args),
                "vectors_per_s": r["n_total"] / sec,
                "e2e": {"value": sec_e2e, "unit": UNIT, "h2d_bytes_per_step": ct_words * 8,
                        "d2h_bytes_per_step": (args.total_blocks + 1) * res_words * 8, "host_clock_value": r["wall_sec_e2e"]},
                "host_clock_value": r["wall_sec"], "halves_ms_per_step_value_e2e": r["halves_ms"], "remeasured_after_disturbance": r["remeasured"],
                "gpu_launches": launches, "clocks": sampler.result(), "results_ok": bool(ok),
                "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": traffic,
                             "traffic_source": ("profiles/r2_traffic.json (ncu --set full, same kernel_stamp)" if traffic else
                                                "null: no ncu capture committed for this library's kernel_stamp"),
                             "algorithmic_bytes_per_launch": d["bytes"] / groups, "peak_source": peak_src,
                             "launches": groups, "kernels": d["launches"], "avg_launch_ms": avg_ms, "share_of_step": d["ms"] / (sec * 1e3 * args.steps),
                             "kernel_stamp": stamp,
                             "warmup_profile_ms": {k: round(v["ms"], 3) for k, v in prof_all.items() if v["launches"]},
                             # the other kernel families: from the serialized ncu launch list of this library's kernels when one is
                             # committed (profiles/r2_launch_summary.json); per-launch events in a warm-up step overstate small kernels
                             "other_kernels": (ls or {}).get("families")}}
        if share:
            line["share_of_configs4"] = share
        if not args.no_cpu_baseline and world == 1:
            vals, desc = cpu_arm(args, args.total_blocks, 1, 1, 120.0)
            line["cpu_baseline"] = dict(desc, value=float(np.mean(vals)), unit=UNIT)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if not ok:
        sys.exit(1)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)

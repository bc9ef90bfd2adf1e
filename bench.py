"""bench.py -- HyDia approach 5 encrypted match latency on B200 (BASELINE.json metric).

A step = one query through DiagonalSender.membershipScenario + indexScenario
(/root/reference/src/main.cpp:343,361), each recomputing the similarity as the reference does.
Workload at N=1: configs[1] (2^14-vector synthetic DB = 1 block of 512 diagonal ciphertexts);
N>1: weak scaling, --blocks-per-gpu blocks on every rank (8 blocks x 8 GPUs = configs[4], 2^20).
See DESIGN.md "Measurement" for every field of the JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "encrypted_match_latency_membership_plus_index"
UNIT = "s"
# BASELINE.md "HyDia, approach 5": published Sender compute seconds (membership, index) per DB size on a 48-thread
# Xeon Gold 5412U (tools/figures/approach5.csv) -- other hardware, not re-measured; vs_baseline = value / (m + i)
PUBLISHED_APPROACH5 = {1 << 14: (4.518783, 4.175511), 1 << 15: (6.155975, 5.85484), 1 << 16: (9.560245, 9.011757),
                       1 << 17: (16.00675, 15.46935), 1 << 18: (28.72538, 27.62988), 1 << 19: (53.24866, 49.70151),
                       1 << 20: (102.411, 96.52172)}


def vs_published(args, seconds):
    if args.ring_dim or args.ndiag != 512:
        return None
    pub = PUBLISHED_APPROACH5.get(args.blocks_per_gpu * args.gpus * 16384)
    return seconds / sum(pub) if pub else None


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--blocks-per-gpu", type=int, default=1)
    ap.add_argument("--ring-dim", type=int, default=0, help="0 = the 128-bit ring of the reference (2^15)")
    ap.add_argument("--ndiag", type=int, default=512)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--rot-chunk", type=int, default=0, help="override the engine's hoisted-rotation batch size")
    ap.add_argument("--rot-streams", type=int, default=0, help="override the number of streams the hoisted-rotation batches use")
    ap.add_argument("--ks-moddown-w", type=int, default=-1, help="override the engine's warp-task ModDown kernel (0 = the r1 launch sequence)")
    ap.add_argument("--rot-lane", type=int, default=-1, help="override the engine's rotation lane (0 = all work ordered on the engine stream)")
    ap.add_argument("--sync-exchange", action="store_true", help="N > 1: host-synchronous result exchange (no overlap of the scenarios)")
    ap.add_argument("--ncu-range", action="store_true",
                    help="run one extra step inside cudaProfilerStart/Stop (for ncu --profile-from-start off)")
    return ap.parse_args()


def synth_db(n, seed, match_rows):
    """Seeded restatement of tools/gen_dataset.sh: rows in [-99, 99], match rows in {1,2,3}, query all ones."""
    rng = np.random.default_rng(seed)
    db = rng.integers(-99, 100, size=(n, 512)).astype(np.float64)
    for r in match_rows:
        db[r] = rng.integers(1, 4, size=512)
    return db


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


def cpu_port_latency(ring_dim, ndiag, blocks, n_rot_sample=96):
    """CPU arm: the oracle (oracle/ckks_oracle.py over oracle/ckks_ref.c, OpenMP on all host cores) runs the
    same op sequence as sender_diag.cpp on a bounded sample -- n_rot_sample hoisted rotations and diagonal
    products instead of ndiag-1 / ndiag, everything else in full -- and the per-op times are scaled to the
    full workload. Returns (estimated seconds for membership + index, description, cores)."""
    from oracle import ckks_oracle as O
    P = O.gen_params(11, ring_dim=ring_dim or None)
    orc = O.Oracle(P)
    orc.keygen(1)
    orc.mult_keygen()
    ns = min(n_rot_sample, ndiag - 1)
    orc.rot_keygen(list(range(1, ns + 1)) + [1 << i for i in range(2)])
    rng = np.random.default_rng(0)
    q = orc.encrypt(rng.uniform(-0.04, 0.04, P.slots))
    dbc = [orc.encrypt(rng.uniform(-0.04, 0.04, P.slots)) for _ in range(4)]
    t0 = time.perf_counter()
    dg = orc.fast_rotation_precompute(q)
    t_pre = time.perf_counter() - t0
    t0 = time.perf_counter()
    rot = [orc.fast_rotation(q, r, dg) for r in range(1, ns + 1)]
    t_rot = (time.perf_counter() - t0) / ns
    t0 = time.perf_counter()
    acc = orc.mult_norelin(q, dbc[0])
    for i, r in enumerate(rot):
        acc = orc.add(acc, orc.mult_norelin(r, dbc[(i + 1) % 4]))
    t_mac = (time.perf_counter() - t0) / (ns + 1)
    t0 = time.perf_counter()
    s = orc.rescale(orc.relinearize(acc))
    t_relin = time.perf_counter() - t0
    t0 = time.perf_counter()
    c = orc.chebyshev_compare(s, 0.44, 10)
    t_cmp = time.perf_counter() - t0
    t0 = time.perf_counter()
    m = orc.add(c, orc.automorphism(c, 5))
    m = orc.add(m, orc.automorphism(m, 25))
    t_sumrot = (time.perf_counter() - t0) / 2
    nsum = int(np.log2(P.slots))
    scen = t_pre + (ndiag - 1) * t_rot + blocks * (ndiag * t_mac + t_relin + t_cmp)
    total = 2 * scen + nsum * t_sumrot
    cores = os.cpu_count() or 1
    desc = (f"oracle port, N={P.N}: 1 precompute + {ns} hoisted rotations + {ns + 1} diagonal products (scaled to "
            f"{ndiag - 1}/{ndiag}) + relin + rescale + 1 chebyshevCompare + 2 EvalSum steps (scaled to {nsum}); "
            f"per-op s: rot {t_rot:.4f} mac {t_mac:.4f} relin+rescale {t_relin:.3f} compare {t_cmp:.3f}")
    return total, desc, cores


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path. OpenFHE v1.2.3 is not buildable here
    (un-vendored), so this is the oracle port on all host cores; rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    nsteps = max(1, min(args.steps, 3))
    vals = []
    for i in range(args.warmup + nsteps):
        v, desc, cores = cpu_port_latency(args.ring_dim, args.ndiag, args.blocks_per_gpu * args.gpus, n_rot_sample=48)
        if i >= args.warmup:
            vals.append(v)
        if i == 0 and args.warmup > 1:
            args.warmup = 1   # one untimed pass is enough to page the library in; keeps the run to minutes
    v = float(np.mean(vals))
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": len(vals),
            "warmup": args.warmup, "ms_per_step": v * 1e3, "higher_is_better": False, "scaling": "weak", "vs_baseline": vs_published(args, v),
            "dtype": "u64", "data": "synthetic", "config": workload_config(args),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))


def workload_config(args):
    blocks = args.blocks_per_gpu * args.gpus
    return {"workload": f"HyDia approach 5, membership + index scenarios, synthetic DB of {blocks * 16384 if not args.ring_dim else blocks * args.ring_dim // 2} "
                        f"vectors x 512 dims ({args.blocks_per_gpu} block(s) of {args.ndiag} diagonal ciphertexts per GPU)"
                        + (" = BASELINE configs[1]" if blocks == 1 and not args.ring_dim and args.ndiag == 512 else "")
                        + (" = BASELINE configs[4]" if blocks == 64 and not args.ring_dim and args.ndiag == 512 else ""),
            "approach": 5, "ring_dim": args.ring_dim or 32768, "limbs": "12+4", "blocks_per_gpu": args.blocks_per_gpu,
            "ndiag": args.ndiag, "l2_policy": "inputs larger than L2: each step streams 13.3 GB of rotation keys and "
            "3.2 GB of database per block (L2 = 126 MB)"}


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

    bpg, ndiag = args.blocks_per_gpu, args.ndiag
    cc = hx.CryptoContext.generate(depth=11, ring_dim=args.ring_dim, device=local)
    if args.rot_chunk:
        cc.set_option("rot_chunk", args.rot_chunk)
    if args.rot_streams:
        cc.set_option("rot_streams", args.rot_streams)
    if args.rot_lane >= 0:
        cc.set_option("rot_lane", args.rot_lane)
    if args.ks_moddown_w >= 0:
        cc.set_option("ks_moddown_w", args.ks_moddown_w)
    slots = cc.GetBatchSize()
    total_blocks = bpg * world
    n_total = total_blocks * slots
    cl = hx.Client(cc, seed=1)                     # same seed on every rank: replicated key set
    cl.EvalMultKeyGen()
    cl.EvalSumKeyGen()
    cl.EvalRotateKeyGen(list(range(1, ndiag)))
    # database: global block g lives on rank g // bpg as local block g % bpg; matches at rows 0 and n-1
    for b in range(bpg):
        g = rank * bpg + b
        match = ([0] if g == 0 else []) + ([slots - 1] if g == total_blocks - 1 else [])
        db = synth_db(slots, seed=100 + g, match_rows=match)
        rows = hx.diagonal_pack(db, slots)
        for d in range(ndiag):
            cl.enroll_row(b, d, rows[d])
        del db, rows
    expected_index = sorted({0, n_total - 1})

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
    idx_host = torch.empty((total_blocks, res_words), dtype=torch.int64).pin_memory()
    mem_host = torch.empty(res_words, dtype=torch.int64).pin_memory()

    if world > 1:
        from image_matching_b200.sharded import TorchTransport
        snd = hx.ShardedDiagonalSender(cc, None, n_total, device=dev, transport=TorchTransport(cc, dev, async_exchange=not args.sync_exchange))
    else:
        snd = hx.DiagonalSender(cc, None, n_total)

    def upload_query_from_host():
        if world == 1:
            return cc.upload_ptr(q_host.data_ptr(), 2, cc.L)
        q_dev = torch.empty(ct_words, dtype=torch.int64, device=dev)
        if rank == 0:
            q_dev.copy_(q_host, non_blocking=True)
        dist.broadcast(q_dev, src=0)
        torch.cuda.synchronize()
        return cc.upload_dev(q_dev.data_ptr(), 2, cc.L)

    def step(q, to_host):
        m = snd.membershipScenario([q])
        idx = snd.indexScenario([q])
        if to_host and rank == 0:
            m.download_ptr(mem_host.data_ptr())
            for i, c in enumerate(idx):
                c.download_ptr(idx_host[i].data_ptr())
        return m, idx

    def fence():
        cc.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(nsteps, e2e):
        fence()
        cc.timer_start()
        out = None
        for _ in range(nsteps):
            q = upload_query_from_host() if e2e else q_res
            out = step(q, e2e)
            cc.sync()   # a step is one query's latency: the next query starts when this one's results are complete
        cc.sync()
        torch.cuda.synchronize()
        ms = cc.timer_stop()
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, out

    q_res = upload_query_from_host()
    # warm-up; the first warm-up step runs with every kernel family timed to find the dominant one
    cc.prof_reset()
    cc.prof_enable(None)
    step(q_res, False)
    prof_all = cc.prof_read()
    dominant = max(prof_all, key=lambda k: prof_all[k]["ms"])
    cc.prof_enable([])
    for _ in range(max(0, args.warmup - 1)):
        step(q_res, False)

    sampler = ClockSampler(local)
    sampler.start()
    cc.prof_reset()
    cc.prof_enable([dominant])
    l0 = cc.launches()
    ms_total, (m, idx) = timed(args.steps, False)
    launches = cc.launches() - l0
    prof = cc.prof_read()
    cc.prof_enable([])
    ms_e2e, _ = timed(args.steps, True)
    sampler.stop_flag = True
    sampler.join(timeout=2)

    if args.ncu_range:
        fence()
        torch.cuda.cudart().cudaProfilerStart()
        step(q_res, False)
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
    if world > 1:
        lt = torch.tensor([launches], dtype=torch.int64, device=dev)
        dist.all_reduce(lt)
        launches = int(lt.item())

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 0)) or 6500.0
        peak_src = "MEASURED_PEAKS.json hbm_gbs" if peaks.get("hbm_gbs") else "fallback 6500 GB/s (B200_PROFILING.md)"
        d = prof[dominant]
        traffic = None
        try:   # ncu --set full DRAM bytes per launch of this kernel family (captured once per change, see profiles/)
            traffic = json.load(open(os.path.join(ROOT, "profiles", "r1_traffic.json"))).get(dominant, {}).get("traffic_bytes_per_launch")
        except Exception:
            pass
        # a "launch" of a fused family is its group of kernels: P-limb kernel, ModDown middle (1 warp-task kernel, or 2 with
        # --ks-moddown-w 0), Q-limb kernel once for the FP64 limbs and once for the 60-bit limb; 4 for ModUp and rescale
        ks_group = 5 if args.ks_moddown_w == 0 else 4
        groups = max(1, d["launches"] // {"ks_hoisted": ks_group, "ks_fused": ks_group, "modup": 4, "rescale": 4}.get(dominant, 1))
        avg_ms = d["ms"] / groups
        achieved = (d["bytes"] / groups) / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
        sec = ms_total / args.steps / 1e3
        sec_e2e = ms_e2e / args.steps / 1e3
        line = {"metric": METRIC, "value": sec, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": sec * 1e3, "higher_is_better": False, "scaling": "weak", "vs_baseline": vs_published(args, sec), "dtype": "u64",
                "data": "synthetic", "config": workload_config(args),
                "e2e": {"value": sec_e2e, "unit": UNIT, "h2d_bytes_per_step": ct_words * 8,
                        "d2h_bytes_per_step": (total_blocks + 1) * res_words * 8},
                "gpu_launches": launches, "clocks": sampler.result(), "results_ok": bool(ok),
                "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": traffic, "algorithmic_bytes_per_launch": d["bytes"] / groups, "peak_source": peak_src,
                             "launches": groups, "kernels": d["launches"], "avg_launch_ms": avg_ms, "share_of_step": d["ms"] / ms_total,
                             "warmup_profile_ms": {k: round(v["ms"], 3) for k, v in prof_all.items() if v["launches"]},
                             # the other kernel families of the step from the warm-up profile (each launch timed with events, so
                             # small kernels read high): GB/s on their algorithmic bytes against the same peak
                             "other_kernels": {k: {"algorithmic_gbs": round(v["bytes"] / (v["ms"] * 1e-3) / 1e9, 1),
                                                   "frac": round(v["bytes"] / (v["ms"] * 1e-3) / 1e9 / peak, 3), "launches": v["launches"]}
                                               for k, v in prof_all.items() if v["launches"] and v["ms"] > 0 and k != dominant}}}
        if not args.no_cpu_baseline and world == 1:
            v, desc, cores = cpu_port_latency(args.ring_dim, ndiag, total_blocks, n_rot_sample=96)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc}
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

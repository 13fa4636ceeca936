#!/usr/bin/env python3
"""bench.py -- pivots/sec of the dense-tableau simplex pivot path on B200.

Contract (see the task statement): `python bench.py --gpus N --steps K --warmup W` prints ONE
JSON line on rank 0.  Workload at every N: BASELINE.json config 4 -- one dense 8192 x 16384
fp64 tableau per GPU (random feasible LP, seed 1234, LP index = rank), a "step" = a window of
--pivots pivots (pricing + ratio test + pivot-row scaling + rank-1 elimination each).
  value  : pivots/s summed over ranks, tableau resident in HBM when the timed region starts
  e2e    : the same metric through lpgpu_solve_dense() with HOST (pinned) buffers: H2D copy of
           the 1.07 GB tableau, P pivots, D2H of x / basis / status inside the timed region
  roofline: the dominant kernel = the pass that applies a block of K pivots to the tableau
           (blk_flush_kernel; with --variant 0x800 the rank-1 elimination kernel), algorithmic bytes
           16*(m+1)*(n+1) per launch over its CUDA-event duration, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline: the CPU oracle (oracle/, kind "port": the reference has no pivot loop to time)
           on the host cores, bounded sample of the same workload
`--impl reference` times that CPU oracle alone, on the same config/metric/unit.
N > 1: one process per GPU (torchrun), independent LPs per rank, no data-path collective
("weak" scaling); timing = max over ranks.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

M = 8192
NS = 8192
N = M + NS            # total columns incl. slacks: tableau (M+1) x (N+1)
SEED = 1234
ALG_BYTES = 16 * (M + 1) * (N + 1)        # one 8-B read + one 8-B write of every element

BM, BNS, BN_LP = 32, 32, 100_000           # config 3 (extra): batched 100k LPs m=32 n=64


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--pivots", type=int, default=1000, help="pivots per step (window)")
    ap.add_argument("--variant", type=lambda s: int(s, 0), default=None, help="dense kernel variant override")
    ap.add_argument("--no-extra", action="store_true", help="skip the batched (config 3) extra measurement")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the sharded-vs-oracle check that precedes the timing")
    ap.add_argument("--m", type=int, default=M, help=argparse.SUPPRESS)     # debugging at reduced size only
    ap.add_argument("--shard-m", type=int, default=16384, help="rows of the column-sharded extra (N > 1); n = 4m")
    ap.add_argument("--shard-pivots", type=int, default=200)
    ap.add_argument("--batched5", type=int, default=1_000_000, help="LPs of the m=64 n=128 batched extra (whole job)")
    return ap.parse_args()


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.p = index, [], None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "20"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for nme, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                pass
        # "under load" = the upper half of the samples (idle samples at the edges drop out)
        sm_sorted = sorted(sm)
        load = sm_sorted[len(sm_sorted) // 2:] if sm_sorted else []
        return {"sm_mhz": statistics.median(load) if load else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ------------------------------------------------------------------ CPU legs (oracle = the checker, timed)
class CpuOracle:
    """The CPU oracle (oracle/lp_oracle.c, OpenMP over rows) on the same synthetic tableau."""

    def __init__(self, m, ns, threads, lp_index=0):
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import ctypes as C
        import numpy as np
        import oracle_lib as O
        self.C, self.O, self.L = C, O, O.lib()
        self.m, self.n, self.threads = m, m + ns, threads
        self.ns, self.lp_index = ns, lp_index
        self.T = np.zeros((m + 1, self.n + 1))
        self.basis = np.zeros(m, np.int32)
        self.synth()

    def synth(self):
        C = self.C
        self.L.lpo_synth_feasible(SEED, self.lp_index, self.m, self.ns, self.T.ctypes.data_as(C.POINTER(C.c_double)),
                                  self.n + 1, self.basis.ctypes.data_as(C.POINTER(C.c_int)))

    def run(self, pivots):
        """`pivots` more pivots from the current state; returns (pivots done, seconds).  If the LP reaches
        its optimum first, the tableau is regenerated (untimed) and the window continues."""
        C, O = self.C, self.O
        done, secs = 0, 0.0
        while done < pivots:
            res = O.Result()
            opts = O.Opts(1e-9, int(pivots - done), O.PLAIN, 0, self.threads)
            t0 = time.perf_counter()
            rc = self.L.lpo_solve_f64(self.m, self.n, 1, self.T.ctypes.data_as(C.POINTER(C.c_double)), self.n + 1,
                                      self.basis.ctypes.data_as(C.POINTER(C.c_int)), C.byref(opts), None, 0,
                                      C.byref(res))
            secs += time.perf_counter() - t0
            assert rc == 0
            done += res.iters
            if res.status != O.ITER_LIMIT:
                self.synth()
        return done, secs

    def calibrate(self, budget_s):
        self.run(1)                                   # page in, start the thread pool
        it, dt = self.run(2)
        return int(max(2, min(5000, budget_s / max(dt / max(it, 1), 1e-6))))


def cpu_baseline(m, ns, budget_s=12.0):
    threads = host_threads()
    o = CpuOracle(m, ns, threads)
    p = o.calibrate(budget_s)
    it, dt = o.run(p)
    return {"value": it / dt, "unit": "pivots/s", "cores": threads, "kind": "port",
            "sample": f"{it} pivots of the same {m}x{m + ns} tableau (seed {SEED}) in {dt:.1f} s, "
                      f"oracle/lp_oracle.c with OpenMP over rows; the reference has no pivot loop to time"}


def reference_arm(args):
    """`--impl reference`: the reference has NO implementation of this path (Source/simplex.c:57-63 are
    empty cases), so the timed CPU implementation is this repo's oracle port, all host threads."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    threads = host_threads()
    m = ns = args.m
    o = CpuOracle(m, ns, threads)
    p = o.calibrate(3.0)                              # each step: a bounded sample of ~3 s
    times, done = [], []
    for s in range(args.warmup + args.steps):
        it, dt = o.run(p)
        if s >= args.warmup:
            times.append(dt); done.append(it)
    value = sum(done) / sum(times)
    sample = f"{p} pivots per step of the {m}x{m + ns} tableau (seed {SEED}), OpenMP over rows, {threads} threads"
    line = {
        "impl": "reference", "metric": "pivots_per_sec", "value": value, "unit": "pivots/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / len(times),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"dense tableau {m}x{m + ns} fp64 random feasible LP per GPU (BASELINE config 4)",
                   "m": m, "n": m + ns, "tableau_bytes": (m + 1) * (m + ns + 1) * 8, "pivots_per_step": p, "seed": SEED,
                   "sample": sample},
        "cpu_baseline": {"value": value, "unit": "pivots/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "pivots/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "the reference has no pivot loop (Source/simplex.c:57-63 empty cases); timed = this repo's CPU oracle port",
    }
    emit(line)


# ------------------------------------------------------------------ GPU arm
def gpu_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import linearprogramming_b200 as lp

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    m = ns = args.m
    n = m + ns
    P, K, W = args.pivots, args.steps, max(args.warmup, 0)
    alg_bytes = 16 * (m + 1) * (n + 1)

    # ---- N > 1: bit-exact check of the column-sharded path against the oracle at THIS world size, before anything is
    # timed (the driver reaches more than one real GPU only through this script): small cases (plain, Big-M, two-phase,
    # Bland; blocks of 32 and 5, rank-1 pipeline) over the peer-memory transport, then the first 64 pivots of the
    # config-5b tableau against the single-tableau path.  Outcome -> config.sharded_parity; a failure fails the run.
    sharded_parity = None
    if world > 1 and not args.no_extra and not args.no_parity:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import sharded_parity as SP
        from linearprogramming_b200 import sharding as _shd

        class _HostDist:
            """object collectives over a gloo group (host plumbing of the checks; the data path is the library's own)"""
            def __init__(self):
                self.g = dist.new_group(backend="gloo")
            def get_rank(self): return rank
            def get_world_size(self): return world
            def barrier(self): dist.barrier(group=self.g)
            def all_gather_object(self, out, obj): dist.all_gather_object(out, obj, group=self.g)
            def broadcast_object_list(self, box, src=0): dist.broadcast_object_list(box, src=src, group=self.g)
            def gather_object(self, obj, parts, dst=0): dist.gather_object(obj, parts, dst=dst, group=self.g)

        hd = _HostDist()
        pctx = lp.Context(local)
        t0p = time.perf_counter()
        try:
            runs = SP.check_small(lp, _shd, hd, pctx, rank, world, transports=("p2p",))
            SP.check_large_prefix(lp, _shd, hd, pctx, rank, world, args.shard_m, 4 * args.shard_m, pivots=64)
            sharded_parity = f"ok ({runs} solves vs oracle + first 64 pivots of {args.shard_m}x{4 * args.shard_m} vs the single-tableau path, world {world}, {time.perf_counter() - t0p:.0f} s)"
        except AssertionError as e:
            sharded_parity = "FAILED: " + str(e)[:300]
        pctx.close()
        barrier()
    variant = args.variant if args.variant is not None else int(os.environ.get("LPGPU_VARIANT", "0"), 0)

    stream = torch.cuda.Stream()
    ctx = lp.Context(local, stream=stream.cuda_stream)

    def timed(fn):
        """CUDA events on the launching stream, synchronize on both sides; returns seconds."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e-3

    state = {"opts": {}}

    def start(d, **opts):
        state["opts"] = opts
        state["iters"] = 0
        d.synth(SEED, rank, ns)
        d.reset(variant=state.get("variant", variant), check_every=256, **opts)

    def window(d):
        """P pivots; if the LP reaches its optimum inside the window (about 17k pivots at the default
        size) it is regenerated and the window continues, so every step performs exactly P pivots."""
        done, upd_ms, sel_ms, passes = 0, 0.0, 0.0, 0
        while done < P:
            r = d.run(P - done)
            done += r["iters"] - state["iters"]
            state["iters"] = r["iters"]
            passes += r["update_launches"]; upd_ms += r["update_ms"]; sel_ms += r["select_ms"]
            if r["status"] != lp.RUNNING:
                start(d, **state["opts"])
        return {"pivots": done, "update_launches": passes, "update_ms": upd_ms, "select_ms": sel_ms}

    # ---- value: resident tableau, K windows of P pivots
    d = lp.DenseTableau(ctx, m, n, 1)
    start(d)
    for _ in range(W):
        window(d)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ctx.launch_count()
    t_value = timed(lambda: [window(d) for _ in range(K)])
    launches = ctx.launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    t_value = max_over_ranks(t_value)
    value = world * K * P / t_value

    # ---- roofline of the dominant kernel: same region, every pass over the tableau bracketed by events
    from linearprogramming_b200 import capi as _capi
    RANK1, INPLACE, PIPELINED_BIT = _capi.RANK1, _capi.INPLACE, _capi.PIPELINED
    blocked = not (variant & RANK1)
    # the default picks pipelined blocks from 0.8 GB of tableau on (csrc/dense.cu want_pipeline); 0x1000 forces them
    pipelined = blocked and not (variant & INPLACE) and bool((variant & PIPELINED_BIT) or (m + 1) * ((n + 1 + 15) // 16 * 16) * 8 >= (800 << 20))
    block_k = ((variant >> 16) & 0x3f) or (28 if pipelined else 32)      # csrc/dense.cu make_bufs

    def kernel_times():
        upd_ms = sel_ms = 0.0
        upd_n = piv = 0
        for _ in range(W):
            window(d)
        for _ in range(K):
            r = window(d)
            upd_ms += r["update_ms"]; sel_ms += r["select_ms"]; upd_n += r["update_launches"]; piv += r["pivots"]
        return upd_ms, sel_ms, upd_n, piv

    start(d, time_kernels=1)
    upd_ms, sel_ms, upd_n, piv_n = kernel_times()
    peak, peak_src = measured_peak()
    traffic, traffic_src = None, None
    pass_kernel = ("blk_flush_mma_kernel" if variant & _capi.DMMA_PASS
                   else "blk_flush_sp_kernel" if variant & _capi.DFMA_PASS else "blk_flush_ring_kernel")
    kname = pass_kernel if blocked else {0: "dense_update_dynamic", 1: "dense_update_tma_staged", 2: "dense_update_direct"}[variant & 0xf]
    try:        # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture
        tfile = {"blk_flush_sp_kernel": "r1_flush_sp_traffic.json",
                 "blk_flush_ring_kernel": "r2_flush_ring_traffic.json"}.get(pass_kernel, "none.json") if blocked else "r1_update_traffic.json"
        with open(os.path.join(ROOT, "profiles", tfile)) as f:
            tj = json.load(f)
        if m == M and (blocked or tj["variant"] == (variant & 0xf)):
            traffic, traffic_src = tj["traffic_bytes_per_launch"], tj["source"]
    except Exception:
        pass
    upd_avg_ms = upd_ms / upd_n
    achieved = alg_bytes / (upd_avg_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": kname,
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
                "avg_launch_ms": upd_avg_ms, "launches": upd_n, "pivots_per_launch": piv_n / upd_n,
                "decision_kernels_avg_ms_per_pivot": sel_ms / piv_n,
                "kernel_share_of_step": upd_ms / (upd_ms + sel_ms), "frac_of_nominal_8TBs": achieved / 8000.0}
    if pipelined:
        # the pass of a block runs on a second stream WHILE the next block is decided: the two kernel families overlap, so their
        # summed times exceed the step; kernel_share_of_step stays the share of summed kernel time (what an ncu launch list shows)
        roofline["overlap"] = ("pipelined blocks: the pass (stream 2 in a green context over every SM but the 16 of the step cluster) runs concurrently with "
                               "the decisions of the next block; avg_launch_ms is measured under that contention "
                               "(alone on the GPU the same launch takes 0.357 ms = 0.92 of the peak: tools/ring_lab.cu, profiles/r2_ring_lab.txt)")
        roofline["pass_time_over_step_time"] = (upd_ms / K) / (1e3 * t_value / K)
    if blocked:
        flops = 2.0 * (piv_n / upd_n) * (m + 1) * (n + 1)
        roofline["fp64_TFLOPs_in_kernel"] = flops / (upd_avg_ms * 1e-3) / 1e12
        # FP64 side, against the NOMINAL rate (64 FMA / clock / SM x 148 SMs x 1.965 GHz = 37.2 TFLOP/s).  Footnote: what this
        # pool's B200 delivers for the instruction mix of the kernel -- DMMA with both operands from shared memory 36.8 TFLOP/s
        # (tools/dmma_probe.cu, profiles/r2_dmma_probe.txt), scalar outer-product DFMAs 25.3 TFLOP/s (tools/dfma_peak.cu)
        sms_pass = 148 - 16 if pipelined else 148
        roofline["fp64_nominal_TFLOPs"] = 37.2
        roofline["fp64_frac_of_nominal"] = roofline["fp64_TFLOPs_in_kernel"] / 37.2
        measured_ceiling = 36.8 if pass_kernel in ("blk_flush_ring_kernel", "blk_flush_mma_kernel") else 25.3
        t_hbm_ms, t_fp64_ms = alg_bytes / (peak * 1e9) * 1e3, flops / (measured_ceiling * sms_pass / 148 * 1e12) * 1e3
        roofline["footnote_measured_fp64_ceiling_TFLOPs"] = measured_ceiling * sms_pass / 148
        roofline["footnote_hbm_ms_and_fp64_ms_at_measured_ceilings"] = [t_hbm_ms, t_fp64_ms]
        roofline["note"] = ("one launch applies pivots_per_launch pivots: every element is read and written once and runs through "
                            "that many chained FMAs (DMMA: four steps per instruction, in pivot order); with the FP64 side at "
                            "the matrix-instruction rate the launch is bounded by HBM alone")
    # the same blocks applied in place (steps, pass, steps on one stream), for comparison
    inplace = None
    if pipelined and not args.no_extra:
        state["variant"] = variant | INPLACE
        start(d)
        window(d)
        t_ip = timed(lambda: window(d))
        start(d, time_kernels=1)
        window(d)
        r = window(d)
        inplace = {"pivots_per_sec": world * P / max_over_ranks(t_ip), "kernel": kname,
                   "avg_launch_ms": r["update_ms"] / r["update_launches"],
                   "roofline_frac": alg_bytes / (r["update_ms"] / r["update_launches"] * 1e-3) / 1e9 / peak,
                   "decision_kernels_avg_ms_per_pivot": r["select_ms"] / r["pivots"]}
        state.pop("variant")
    # the rank-1 path (one pass per pivot) on the same workload, for comparison with earlier rounds
    rank1 = None
    if blocked and not args.no_extra:
        state["variant"] = variant | RANK1
        start(d)
        p_save, P = P, min(P, 300)
        window(d)
        t_r1 = timed(lambda: window(d))
        start(d, time_kernels=1)
        window(d)
        r = window(d)
        rank1 = {"pivots_per_sec": world * P / max_over_ranks(t_r1), "kernel": "dense_update_dynamic",
                 "avg_launch_ms": r["update_ms"] / r["update_launches"],
                 "roofline_frac": alg_bytes / (r["update_ms"] / r["update_launches"] * 1e-3) / 1e9 / peak}
        P = p_save
        state.pop("variant")
    d.close()

    # ---- e2e: host (pinned) buffers through the one-shot C-ABI call, copies inside the timed region
    Th = ctx.pinned((m + 1, n + 1))
    bh = ctx.pinned((m,), np.int32)
    T0, b0 = lp.synth_host(SEED, rank, m, ns)
    Th[...] = T0
    bh[...] = b0
    del T0

    def e2e_step():
        bh[...] = b0
        r = lp.solve_dense(ctx, Th, bh, max_iter=P, results_only=True, inplace=True, variant=variant)
        if r["iters"] != P:
            raise SystemExit("e2e step did not perform the requested pivots")
        return r

    for _ in range(min(W, 2)):
        e2e_step()
    t_e2e = max_over_ranks(timed(lambda: [e2e_step() for _ in range(K)]))

    # the same job as a stream of problems: two tableau objects, the upload of step i+1 (lpgpu_dense_upload_async, its own
    # copy stream) travels while step i pivots.  Every step still uploads ITS input from pinned host memory and reads ITS
    # result back inside the timed region; only the order of the calls differs from the one-shot lpgpu_solve_dense above.
    Th2 = ctx.pinned((m + 1, n + 1))
    bh2 = ctx.pinned((m,), np.int32)
    Th2[...] = Th
    objs = [lp.DenseTableau(ctx, m, n, 1), lp.DenseTableau(ctx, m, n, 1)]
    bufs = [(Th, bh), (Th2, bh2)]
    cnt = {"i": 0}
    bh[...] = b0
    objs[0].upload_async(*bufs[0])

    def e2e_stream_step():
        i = cnt["i"]; cnt["i"] += 1
        X, (Tn, bn) = objs[i % 2], bufs[(i + 1) % 2]
        bn[...] = b0                                    # the next step's input arrives in its pinned buffer ...
        objs[(i + 1) % 2].upload_async(Tn, bn)          # ... and starts travelling
        X.reset(max_iter=P, check_every=256, variant=variant)
        r = X.run(P)
        out = X.download(want_T=False, want_log=False)  # x and basis of this step back on the host
        if r["iters"] != P or out["x"] is None:
            raise SystemExit("e2e stream step did not perform the requested pivots")
        return r

    for _ in range(min(W, 2)):
        e2e_stream_step()
    t_e2e_stream = max_over_ranks(timed(lambda: [e2e_stream_step() for _ in range(K)]))
    for o in objs:
        o.close()
    # ---- e2e headline: what the reference's entry point does -- ONE lpgpu_solve_dense call per model (Source/simplex.c:27-73):
    # upload the 1.07 GB tableau from pinned host memory, pivot to OPTIMAL (about 17k pivots), read x / basis / status back
    golden = None
    try:
        with open(os.path.join(ROOT, "tests", "golden", "cfg4_full_solve.json")) as f:
            golden = json.load(f)
    except Exception:
        pass
    full = {"iters": 0, "status": None, "z": None}

    def e2e_full_step():
        bh[...] = b0
        r = lp.solve_dense(ctx, Th, bh, results_only=True, inplace=True, variant=variant, check_every=256)
        full.update(iters=r["iters"], status=r["status"], z=r["z"], basis=bh.copy())
        return r

    e2e_full_step()                                     # warm-up (allocations of the cached tableau object are behind us anyway)
    KF = max(1, min(K, 5))
    t_full = max_over_ranks(timed(lambda: [e2e_full_step() for _ in range(KF)]))
    full_parity = "no golden record"
    if golden is not None and m == golden["m"] and rank == 0:
        okg = (lp.STATUS_NAMES[full["status"]] == golden["status"] and full["iters"] == golden["iters"] and
               float(full["z"]).hex() == golden["z_hex"] and full["basis"].tolist() == golden["basis"])
        full_parity = "ok" if okg else f"FAILED: status/iters/z/basis {lp.STATUS_NAMES[full['status']]}/{full['iters']}/{full['z']!r} vs golden {golden['status']}/{golden['iters']}/{golden['z']!r}"
    e2e = {"value": world * KF * full["iters"] / t_full, "unit": "pivots/s",
           "h2d_bytes_per_step": int((m + 1) * (n + 1) * 8 + m * 4), "d2h_bytes_per_step": int(n * 8 + m * 4 + 64),
           "api": "ONE lpgpu_solve_dense per step on pinned host buffers: upload, pivot to OPTIMAL, read x / basis / status back",
           "pivots_per_step": full["iters"], "status": lp.STATUS_NAMES.get(full["status"], "?"), "ms_per_solve": 1e3 * t_full / KF,
           "steps": KF, "golden": full_parity,
           "window_one_shot": {"value": world * K * P / t_e2e, "unit": "pivots/s", "pivots_per_step": P,
                               "api": "lpgpu_solve_dense(max_iter = pivots per step): one upload per window of pivots, nothing overlapped"},
           "window_streamed": {"value": world * K * P / t_e2e_stream, "unit": "pivots/s", "pivots_per_step": P,
                               "api": "two tableau objects: lpgpu_dense_upload_async of window i+1 overlaps lpgpu_dense_run of window i"}}

    # ---- extra: batched path, BASELINE config 3 (100k LPs m=32 n=64 per GPU, one LP per CTA in smem)
    extra = {}
    if not args.no_extra:
        bn = BN_LP
        b = lp.BatchedLPs(ctx, bn, BM, BM + BNS, 1)
        b.synth(SEED, rank * bn, BNS)
        for _ in range(3):
            b.run()
        ms = [b.run() for _ in range(5)]
        out = b.download(want_x=False)
        b.close()
        t_b = max_over_ranks(statistics.median(ms) * 1e-3)
        lp_bytes = 8 * ((BM + 1) * (BM + BNS + 1) + (BM + BNS) + 4)
        extra["batched"] = {"metric": "LPs_per_sec", "value": world * bn / t_b, "unit": "LPs/s",
                            "workload": f"{bn} LPs m={BM} n={BM + BNS} fp64 per GPU, resident (BASELINE config 3)",
                            "kernel": "batched_reg_kernel<1,32,2> (one warp per LP, tableau in registers)",
                            "kernel_ms": t_b * 1e3, "mean_pivots_per_lp": float(out["iters"].mean()),
                            "all_optimal": bool((out["status"] == lp.OPTIMAL).all()),
                            "hbm_GBs_algorithmic": world * bn * lp_bytes / t_b / 1e9,
                            "frac_of_measured_peak_per_gpu": bn * lp_bytes / t_b / 1e9 / peak}
        # BASELINE config 5a: m=64 n=128, instances sharded over the ranks by contiguous index range
        from linearprogramming_b200 import sharding as _sh
        n5 = args.batched5                   # the whole 1 M-LP job at every N (67 GB of tableaux fit one B200)
        lo5, hi5 = _sh.lp_range(n5, world, rank)
        b5 = lp.BatchedLPs(ctx, hi5 - lo5, 64, 128, 1)
        b5.synth(SEED, lo5, 64)
        b5.run()
        ms5 = [b5.run() for _ in range(3)]
        out5 = b5.download(want_x=False)
        b5.close()
        t5 = max_over_ranks(statistics.median(ms5) * 1e-3)
        lp5_bytes = 8 * (65 * 129 + 128 + 4)
        extra["batched_m64_n128"] = {"metric": "LPs_per_sec", "value": n5 / t5, "unit": "LPs/s",
                                     "workload": f"{n5} LPs m=64 n=128 fp64 sharded by index range over {world} GPU(s), "
                                                 "no collective (BASELINE config 5a)",
                                     "kernel": "batched_reg_kernel<4,16,4>", "kernel_ms": t5 * 1e3,
                                     "mean_pivots_per_lp": float(out5["iters"].mean()),
                                     "all_optimal": bool((out5["status"] == lp.OPTIMAL).all()),
                                     "hbm_GBs_algorithmic": n5 * lp5_bytes / t5 / 1e9}

    # ---- extra (N = 1): BASELINE config 5b's tableau (16384 x 65536, 8.6 GB) on ONE GPU through the blocked path,
    # for comparison with the column-sharded rank-1 runs at N > 1
    big_n1 = None
    if rank == 0 and not args.no_extra and m == M:
        lm, ln = args.shard_m, 4 * args.shard_m
        big = lp.DenseTableau(ctx, lm, ln, 1)
        big.synth(SEED, 0, ln - lm)
        big.reset(variant=variant, check_every=256)
        big.run(64)
        lpv = 256
        torch.cuda.synchronize()
        eb0, eb1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        eb0.record(stream)
        big.run(lpv)
        eb1.record(stream)
        torch.cuda.synchronize()
        t_big = eb0.elapsed_time(eb1) * 1e-3          # rank 0 alone (no collective in this leg)
        r_big = big.run(0)
        big.close()
        big_n1 = lpv / t_big
        extra["large_tableau_single_gpu"] = {"metric": "pivots_per_sec", "value": lpv / t_big, "unit": "pivots/s",
                                             "workload": f"one dense tableau {lm}x{ln} fp64 ({(lm + 1) * (ln + 1) * 8 / 1e9:.2f} GB) resident on one GPU, default path",
                                             "pivots": lpv, "us_per_pivot": 1e6 * t_big / lpv, "iters_after": r_big["iters"]}

    # ---- extra (rank 0): the dual simplex (the reference's menu stub Source/router.c:32-34) on a 3000 x 5000
    # dual-feasible LP: max c.x with c <= 0, about half the rows '>=' rows turned into '<=' rows with a negative b
    if rank == 0 and not args.no_extra and m == M:
        rng = np.random.default_rng(77)
        dm, dns = 3000, 2000
        A = rng.uniform(0.0, 1.5, size=(dm, dns))
        A[rng.random((dm, dns)) < 0.2] = 0.0
        ge = rng.random(dm) < 0.5
        ax = A @ rng.uniform(0.0, 2.0, size=dns)
        bb = np.where(ge, np.maximum(ax - rng.uniform(0.1, 1.0, dm), 0.05), ax + rng.uniform(0.1, 2.0, dm))
        sgn = np.where(ge, -1.0, 1.0)
        Td = np.zeros((dm + 1, dns + dm + 1))
        Td[:dm, 0] = sgn * bb
        Td[:dm, 1:dns + 1] = sgn[:, None] * A
        Td[:dm, dns + 1:] = np.eye(dm)
        Td[dm, 1:dns + 1] = -rng.uniform(0.1, 1.0, size=dns)
        dd = lp.DenseTableau(ctx, dm, dns + dm, 1)
        dd.upload(Td, (dns + 1 + np.arange(dm)).astype(np.int32))
        dd.reset(method=lp.DUAL, check_every=256)
        dd.run(32)
        torch.cuda.synchronize()
        ed0, ed1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ed0.record(stream)
        r_d = dd.run(1 << 20)
        ed1.record(stream)
        torch.cuda.synchronize()
        dd.close()
        extra["dual_simplex"] = {"metric": "pivots_per_sec", "value": (r_d["iters"] - 32) / (ed0.elapsed_time(ed1) * 1e-3), "unit": "pivots/s",
                                 "workload": f"dual simplex, {dm} x {dns + dm} fp64 dual-feasible LP, blocked path in place", "iters": r_d["iters"],
                                 "status": lp.STATUS_NAMES[r_d["status"]]}

    # ---- extra (N > 1): BASELINE config 5b, ONE 16384 x 65536 tableau column-sharded over the ranks,
    # candidate + entering-column exchange per pivot over our own NCCL communicator
    if world > 1 and not args.no_extra:
        barrier()
        from linearprogramming_b200 import sharding
        sm_, sn_ = args.shard_m, 4 * args.shard_m
        sp = args.shard_pivots
        per_transport = {}
        for transport in ("p2p", "p2p_rank1", "nccl"):
            if transport == "nccl":
                ident = sharding.exchange_nccl_id(dist)
                sh = lp.ShardedTableau(ctx, sm_, sn_, 1, world=world, rank=rank, nccl_id=ident)
            else:
                sh = lp.ShardedTableau(ctx, sm_, sn_, 1, world=world, rank=rank, p2p=True)
                sharding.attach_p2p(dist, sh)
            sh.synth(SEED, 0, sn_ - sm_)
            sh.reset(variant=RANK1 if transport == "p2p_rank1" else 0, check_every=256)
            sh.run(64 if transport == "p2p" else 20)
            npv = 4 * sp if transport == "p2p" else sp
            per_transport[transport] = max_over_ranks(timed(lambda: sh.run(npv))) / npv
            if transport == "p2p":
                r_s = sh.run(0)
                its = [None] * world
                dist.all_gather_object(its, (r_s["iters"], r_s["status"], float(r_s["z"]).hex()))
                ranks_agree = all(v == its[0] for v in its)
            barrier()               # p2p: nobody unmaps while a peer may still store into its region
            sh.close()
        t_s = per_transport["p2p"]
        s_bytes = 16 * (sm_ + 1) * (sn_ + 1)
        extra["column_sharded"] = {"metric": "pivots_per_sec", "value": 1.0 / t_s, "unit": "pivots/s",
                                   "workload": f"one dense tableau {sm_}x{sn_} fp64 column-sharded over {world} GPUs",
                                   "path": "blocked: per pivot one cluster kernel per rank (candidates and the owner's lazily "
                                           "updated entering column exchanged as NVLink peer stores inside it), one local pass "
                                           "over the rank's column block per 32 pivots",
                                   "pivots": 4 * sp, "us_per_pivot": 1e6 * t_s, "iters_after": r_s["iters"], "ranks_agree": ranks_agree,
                                   "rank1_pipeline": {"us_per_pivot_peer_memory": 1e6 * per_transport["p2p_rank1"],
                                                      "us_per_pivot_nccl_allgather": 1e6 * per_transport["nccl"],
                                                      "pivots_per_sec_peer_memory": 1.0 / per_transport["p2p_rank1"],
                                                      "hbm_GBs_algorithmic_total": s_bytes / per_transport["p2p_rank1"] / 1e9,
                                                      "frac_of_measured_peak_per_gpu": s_bytes / per_transport["p2p_rank1"] / 1e9 / world / peak}}

    # ---- the secondary results BASELINE's metric asks for, inside a dict the driver keeps whole (short keys, rounded)
    def r3(v):
        return None if v is None else float(f"{v:.4g}")
    sec = {}
    if rank1 is not None:
        sec["rank1_kernel"] = {"frac": r3(rank1["roofline_frac"]), "ms": r3(rank1["avg_launch_ms"]), "pivots_per_s": r3(rank1["pivots_per_sec"])}
    if inplace is not None:
        sec["blocks_in_place"] = {"pivots_per_s": r3(inplace["pivots_per_sec"]), "pass_ms": r3(inplace["avg_launch_ms"])}
    if "batched" in extra:
        sec["batched_cfg3"] = {"lps_per_s": r3(extra["batched"]["value"]), "frac": r3(extra["batched"]["frac_of_measured_peak_per_gpu"]), "n_lp": world * BN_LP}
    if "batched_m64_n128" in extra:
        b5x = extra["batched_m64_n128"]
        sec["batched_cfg5a"] = {"lps_per_s": r3(b5x["value"]), "n_lp": n5, "frac": r3(b5x["hbm_GBs_algorithmic"] / world / peak)}
    if big_n1 is not None or world > 1:
        bn1 = [big_n1]
        if world > 1:
            dist.broadcast_object_list(bn1, src=0)
        big_n1 = bn1[0]
    if big_n1 is not None:
        sec["cfg5b_one_gpu"] = {"pivots_per_s": r3(big_n1)}
    if "dual_simplex" in extra:
        sec["dual_3000x5000"] = {"pivots_per_s": r3(extra["dual_simplex"]["value"]), "iters": extra["dual_simplex"]["iters"], "status": extra["dual_simplex"]["status"]}
    if "column_sharded" in extra:
        cs = extra["column_sharded"]
        sec["column_sharded"] = {"pivots_per_s": r3(cs["value"]), "us_per_pivot": r3(cs["us_per_pivot"]),
                                 "strong_eff_vs_n1": r3(cs["value"] / (world * big_n1)) if big_n1 else None,
                                 "ranks_agree": cs["ranks_agree"], "nccl_rank1_us": r3(cs["rank1_pipeline"]["us_per_pivot_nccl_allgather"])}
    roofline["secondary"] = sec

    config = {"workload": f"dense tableau {m}x{n} fp64 random feasible LP per GPU (BASELINE config 4)",
              "m": m, "n": n, "tableau_bytes": (m + 1) * (n + 1) * 8, "pivots_per_step": P, "seed": SEED,
              "variant": variant, "path": ((f"blocked, pipelined: {block_k} pivots decided on lazy slices while the pass that applies "
                                            "the previous block runs out of place on a second stream") if pipelined else
                                           f"blocked: {block_k} pivots decided on lazy slices, then one pass over the tableau"
                                           if blocked else "rank-1: one pass over the tableau per pivot"),
              "l2": "inputs larger than L2 (1.07 GB tableau vs 126 MB)",
              "sharding": "independent LP per rank, no collective" if world > 1 else "single GPU",
              "full_solve_parity": full_parity}
    if sharded_parity is not None:
        config["sharded_parity"] = sharded_parity
    line = {
        "metric": "pivots_per_sec", "value": value, "unit": "pivots/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": 1e3 * t_value / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": config,
        "roofline": roofline, "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "extra": extra,
    }
    if rank1 is not None:
        extra["rank1_path"] = rank1
    if inplace is not None:
        extra["blocks_in_place"] = inplace
    if rank == 0 and world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(m, ns)
    if rank == 0:
        emit(line)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    failed = [v for v in (sharded_parity, full_parity) if isinstance(v, str) and v.startswith("FAILED")]
    if failed:
        sys.stderr.write("bench.py: parity check failed: " + " | ".join(failed) + "\n")
        sys.exit(3)


def _reserve_stdout():
    """stdout carries exactly ONE JSON line: everything else that writes to fd 1 (the NCCL version
    banner, library chatter) is sent to stderr; emit() writes to the real stdout."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line):
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()


def main():
    args = parse_args()
    _reserve_stdout()
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()

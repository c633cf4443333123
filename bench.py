#!/usr/bin/env python
"""Benchmark of the bulk phase-transition path: pod events/s classified on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3|c2|c4] [--impl ours|reference]

One step = one full replay (pte_run) of the workload's log: clear the uid dictionary, filter +
intern + pack, LSD radix sort on the interned key, segmented adjacent-diff into the emit bitmap,
bitmap scan + compaction into the payload columns (replay order comes from the bitmap, there is no
second sort).  Default workload: BASELINE.json config 3 (the one the metric is quoted on): 1 B events /
10 M pods, uniform phase churn, staging filters (nothing dropped), per GPU.
For N > 1 (launched under torchrun) every rank replays its own uid-hash shard of an N x larger
log (weak scaling); the only collective is an NCCL all-gather of four counters per rank per step.
--total-events T replays BASELINE config 5 literally (T = 8e9 events over the N GPUs: T / 1e9 uid-hash
shards, each rank replays its share one after the other on one engine; generation between shards is untimed).

Prints ONE JSON line (rank 0).  `value` is device time (CUDA events inside pte_run, max over
ranks); `e2e` is PodWatcher.replay() on a log in pinned host memory: pipelined host->device copies,
replay and the emit bitmap on its way back, wall clock.  After the timed loop every rank compares the
count and the order-sensitive digest of its emitted records with tests/golden/digests.json (written by
the C oracle): `parity`; a mismatch makes the run exit 1.
--impl reference times the reference's own per-event loop on the host cores instead: the unmodified
watcher staged in oracle/_ref when it is there (kind "reference"), else the oracle's restatement ("port").
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED = 0x6B38735F77617463
WORKLOADS = {
    # name: (events per GPU, pods per GPU, dist, p_del, description)
    "c2": (10_000_000, 100_000, 0, 0.0, "BASELINE config 2: synthetic replay 10M events / 100k pods, uniform phase churn"),
    "c3": (1_000_000_000, 10_000_000, 0, 0.0, "BASELINE config 3: synthetic replay 1B events / 10M pods, uniform phase churn"),
    "c4": (1_000_000_000, 10_000_000, 1, 0.05, "BASELINE config 4: skewed replay 1B events, octave-Zipf pod activity + 5% DELETED"),
}
METRIC = "pod events/sec classified"
UNIT = "events/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--events", type=float, default=None, help="override events per GPU (parity/debug only)")
    ap.add_argument("--pods", type=float, default=None)
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-packed-meta", action="store_true",
                    help="the e2e log carries type << 6 | phase as ONE byte per event (ReplayLog.pack_meta, done while the log is built: untimed): 17 instead of 18 B/event over PCIe")
    ap.add_argument("--total-events", type=float, default=None,
                    help="BASELINE config 5 literally: this many events in total over the N GPUs (e.g. 8e9)")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md recipe)
# ----------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------------------------
# CPU arm: the oracle's restatement of the reference's per-event loop, on the host cores
# ----------------------------------------------------------------------------------------------
def _cpu_worker(args):
    """One process = one uid-hash shard of the sample, replayed the way the reference would:
    per event should_process_event -> namespace list -> _extract_pod_data (pod_watcher.py:214-233),
    then the last-known-phase layer.  Logging is off and event_timestamp is a constant."""
    seed, n, pods, dist, p_del, shard, n_shards = args
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import fake_kubernetes as fk
    from oracle import c_oracle, make_ref
    from oracle import podwatch_oracle as po
    cols = c_oracle.generate(seed, n, pods, dist=dist, p_del=p_del, shard=shard, n_shards=n_shards)
    events = [(e["type"], e["object"]) for e in fk.events_from_columns(cols, rich=False)]
    ref = make_ref.load_reference()
    if ref is not None:
        # the unmodified reference: PodWatcher.handle_pod_event per event (pod_watcher.py:214-241), payloads taken at
        # the POST site by wrapping _extract_pod_data (the POST itself is commented out upstream), logging off
        import logging
        os.chdir(make_ref.DST)                     # its config paths are CWD-relative (:22,:25)
        logging.disable(logging.CRITICAL)
        w = ref.PodWatcher(environment="staging")
        payloads, orig = [], w._extract_pod_data

        def rec(pod):
            d = orig(pod)
            payloads.append(d)                     # handle_pod_event adds event_type to this same dict (:233)
            return d
        w._extract_pod_data = rec
        t0 = time.perf_counter()
        for ty, pod in events:
            w.handle_pod_event(ty, pod)
        keep = po.transition_filter(payloads)
        dt = time.perf_counter() - t0
        return n, len(keep), dt, "reference"
    t0 = time.perf_counter()
    payloads = po.reference_trace("staging", {}, events, "1970-01-01T00:00:00")
    keep = po.transition_filter(payloads)
    dt = time.perf_counter() - t0
    return n, len(keep), dt, "port"


def cpu_reference_arm(wl, sample_per_core: int = 600_000):
    """-> dict for `cpu_baseline` (Python port on all cores + C port on one core)."""
    import multiprocessing as mp
    from oracle import c_oracle
    n_ev, n_pods, dist, p_del, _ = wl
    cores = os.cpu_count() or 1
    # same pod density as the workload: events/pods ratio kept, shards by uid hash like the GPU path
    pods_per = max(1, int(sample_per_core * n_pods / n_ev))
    jobs = [(SEED, sample_per_core, pods_per, dist, p_del, g, cores) for g in range(cores)]
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(cores) as pool:
        res = pool.map(_cpu_worker, jobs)
    wall = time.perf_counter() - t0
    busy = max(r[2] for r in res)                    # slowest worker's replay time (set-up excluded)
    total = sum(r[0] for r in res)
    # C port, one thread, larger sample of the very same stream
    cn = min(n_ev, 20_000_000)
    cp = max(1, int(cn * n_pods / n_ev))
    t1 = time.perf_counter()
    c_oracle.generate_replay(SEED, cn, cp, dist=dist, p_del=p_del)
    c_dt = time.perf_counter() - t1
    kind = res[0][3]
    what = ("the unmodified reference PodWatcher.handle_pod_event (oracle/_ref) per event" if kind == "reference"
            else "Python restatement of handle_pod_event")
    return {
        "value": total / busy, "unit": UNIT, "cores": cores, "kind": kind,
        "sample": "%d events (%d per core x %d uid-hash shards, %d pods each), %s + transition layer, "
                  "logging off; pool wall %.1fs" % (total, sample_per_core, cores, pods_per, what, wall),
        "python_1core_events_per_s": res[0][0] / res[0][2],
        "c_port_1thread_events_per_s": cn / c_dt,
        "c_port_sample": "%d events / %d pods incl. generation" % (cn, cp),
    }


def make_config(wl, world, total_events=None):
    """The workload description both arms print (identical keys and values for the same command line)."""
    cfg = {"workload": wl[4], "events_per_gpu": wl[0], "pods_per_gpu": wl[1], "filters": "staging (none)",
           "dedupe": True, "shards": world, "seed": hex(SEED),
           "l2": "inputs (30 GB log, 8 GB record buffers) far larger than the 126 MB L2"}
    if total_events:
        cfg["total_events"] = int(total_events)
        cfg["shards"] = int(total_events) // wl[0]
        cfg["workload"] = ("BASELINE config 5: uid-hash-sharded replay %.0fB events / %.0fM pods across %d B200 (%d shards of %s)"
                           % (total_events / 1e9, total_events / 1e8, world, cfg["shards"], wl[4].split(": ", 1)[1]))
    return cfg


def run_reference(args, wl, rank):
    if rank != 0:
        return
    cfg = make_config(wl, int(os.environ.get("WORLD_SIZE", "1")), args.total_events)
    vals, last = [], None
    for s in range(args.warmup + args.steps):
        last = cpu_reference_arm(wl, sample_per_core=60_000)
        if s >= args.warmup:
            vals.append(last["value"])
    v = sum(vals) / len(vals)
    last["value"] = v
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "python objects", "data": "synthetic", "config": cfg, "cpu_baseline": last,
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference arm = the reference's per-event Python loop on a bounded sample per step, one uid-hash shard per "
                "host core: the unmodified watcher staged in oracle/_ref by oracle/make_ref.py when present (kind "
                "'reference'), else the oracle's restatement (kind 'port')",
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    ev, pods, dist, p_del, desc = WORKLOADS[args.workload]
    if args.events:
        ev = int(args.events)
    if args.pods:
        pods = int(args.pods)
    wl = (ev, pods, dist, p_del, desc)

    if args.impl == "reference":
        run_reference(args, wl, rank)
        return

    import numpy as np
    from k8s_watcher_b200 import PhaseTransitionEngine
    from k8s_watcher_b200.engine import PinnedBuffer

    from k8s_watcher_b200 import sharding
    dist_t = sharding.init_process_group("nccl", local) if world > 1 else None

    def barrier():
        if dist_t is not None:
            dist_t.barrier()

    # shards: one per GPU (weak scaling), or --total-events / events-per-shard of them dealt to the ranks in order
    shards_total = world
    if args.total_events:
        shards_total = int(args.total_events) // ev
        if shards_total * ev != int(args.total_events) or shards_total % world:
            raise SystemExit("--total-events must be a multiple of %d events x %d GPUs" % (ev, world))
    per_rank = shards_total // world
    my_shards = [rank * per_rank + j for j in range(per_rank)]
    multi = per_rank > 1

    eng = PhaseTransitionEngine(ev, pods, device=local)
    if world > 1:
        uid = sharding.broadcast_bytes(dist_t, PhaseTransitionEngine.comm_unique_id() if rank == 0 else None)
        eng.comm_init(uid, rank, world)
    eng.set_filters(None)                                   # staging regime: every event reaches the sort

    def gen(g):
        eng.generate(SEED, ev, pods, dist=dist, p_del=p_del, shard=g, n_shards=shards_total)
    gen(my_shards[0])

    def step():
        """One pass over this rank's shards.  With several shards per rank the next shard's log is generated in place
        between the timed replays (inputs are resident in HBM whenever a timed region starts)."""
        sts, g_ms, counts = [], 0.0, None
        for g in my_shards:
            if multi:
                gen(g)
            sts.append(eng.run(dedupe=True))
            if world > 1:
                t0 = time.perf_counter()
                c = eng.gather_counts()
                g_ms += (time.perf_counter() - t0) * 1e3
                counts = c if counts is None else counts + c
        return sts, g_ms, counts

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    dev_ms = gather_ms = 0.0
    pass_ms = pass_n = 0.0
    stage_sum = {k: 0.0 for k in ("ms_total", "ms_intern", "ms_sort_key", "ms_diff", "ms_assemble")}
    launches = 0
    wall0 = time.perf_counter()
    for _ in range(args.steps):
        sts, g_ms, counts = step()
        gather_ms += g_ms
        for st in sts:
            dev_ms += st["ms_total"]
            pass_ms += st["ms_pass_avg"] * st["n_passes"]; pass_n += st["n_passes"]
            for k in stage_sum:
                stage_sum[k] += st[k]
            launches += st["n_launches"]
    barrier()
    wall_ms = (time.perf_counter() - wall0) * 1e3
    clocks = sampler.stop() if rank == 0 else None

    step_ms = (dev_ms + gather_ms) / args.steps
    if world > 1:
        step_ms, wall_step_ms = sharding.max_over_ranks(dist_t, [step_ms, wall_ms / args.steps])
        total_events = int(counts[:, 0].sum())
        total_emitted = int(counts[:, 2].sum())
    else:
        wall_step_ms = wall_ms / args.steps
        total_events, total_emitted = sum(int(x["n_in"]) for x in sts), sum(int(x["n_emitted"]) for x in sts)

    # ---- parity: every shard this rank replays against the C oracle's table (count + order-sensitive digest)
    parity = _parity(eng, args, wl, my_shards, shards_total, multi, gen, sts, dist_t, world)

    # ---- rooflines (all HBM-bound integer kernels).  Algorithmic bytes per launch are DESIGN.md section 3's
    # per-event figures x the events one launch processes; durations are CUDA-event times inside pte_run.
    peak, peak_src = _peak()
    traffic, traffic_src = _traffic()
    n_surv, n_emit, n_in = float(st["n_survived"]), float(st["n_emitted"]), float(st["n_in"])
    avg = {k: v / (args.steps * per_rank) for k, v in stage_sum.items()}       # per replay of one shard
    kp = int(st["n_passes"])
    rounds = max(1, int(st["n_diff_rounds"]))
    per_kernel = {
        # kernel: (launches per step, algorithmic bytes per launch, ms per launch)
        # the unfiltered kernel reads uid_hi, uid_lo, rv, phase, type (26 B; ns_id is not touched), writes the record,
        # and touches one 16-byte dictionary slot
        "intern_pack_kernel": (1, 26.0 * n_in + 8.0 * n_surv + 16.0 * n_surv, avg["ms_intern"]),
        "radix_pass_kernel": (kp, 16.0 * n_surv, pass_ms / pass_n if pass_n else 0.0),
        "diff_flag_kernel": (rounds, 8.0 * n_surv, avg["ms_diff"] / rounds),
        # bitmap read + event type gathered + arrival index and type written (rv / ns / phase are on-demand columns now)
        "assemble_bitmap_kernel": (1, n_in / 8.0 + 1.0 * n_emit + 5.0 * n_emit, avg["ms_assemble"]),
    }
    by_kernel = {}
    for name, (per_step, nbytes, ms) in per_kernel.items():
        if ms <= 0:
            continue
        ach = nbytes / (ms * 1e-3) / 1e9
        by_kernel[name] = {"launches_per_step": per_step, "avg_launch_ms": ms, "algorithmic_bytes_per_launch": nbytes,
                           "achieved": ach, "frac": ach / peak, "share_of_step": per_step * ms / avg["ms_total"],
                           "traffic": traffic.get(name),
                           # real DRAM bytes (ncu capture of the same workload) over the live launch time
                           "traffic_frac": (traffic[name] / (ms * 1e-3) / 1e9 / peak) if traffic.get(name) else None}
    dom = max(by_kernel, key=lambda k: by_kernel[k]["share_of_step"]) if by_kernel else None
    roof = None
    if dom:
        d = by_kernel[dom]
        roof = {"bound": "hbm", "kernel": dom, "achieved": d["achieved"], "peak": peak, "unit": "GB/s", "frac": d["frac"],
                "peak_source": peak_src, "traffic": d["traffic"], "traffic_frac": d["traffic_frac"],
                "launches_per_step": d["launches_per_step"],
                "avg_launch_ms": d["avg_launch_ms"], "algorithmic_bytes_per_launch": d["algorithmic_bytes_per_launch"],
                "share_of_step": d["share_of_step"], "frac_of_8TBps_nominal": d["achieved"] / 8000.0,
                # the whole step against its algorithmic bytes and against the compulsory I/O (30 B in + 16 B per emitted record)
                "step_algorithmic_bytes": sum(v["launches_per_step"] * v["algorithmic_bytes_per_launch"] for v in by_kernel.values()),
                "step_achieved": sum(v["launches_per_step"] * v["algorithmic_bytes_per_launch"] for v in by_kernel.values()) / (avg["ms_total"] * 1e-3) / 1e9,
                "step_dram_traffic": (sum(v["launches_per_step"] * v["traffic"] for v in by_kernel.values()) if all(v["traffic"] for v in by_kernel.values()) else None),
                "compulsory_io_bytes": 30.0 * n_in + 16.0 * n_emit,
                "compulsory_io_frac": (30.0 * n_in + 16.0 * n_emit) / (avg["ms_total"] * 1e-3) / 1e9 / peak,
                "traffic_source": traffic_src,
                "note": "ms of intern/diff/assemble are stage times (kernel + its memsets)", "by_kernel": by_kernel}

    # ---- end to end through host buffers (rank-local; max over ranks)
    e2e = None
    if not args.no_e2e:
        import psutil
        budget = 0.5 * psutil.virtual_memory().available / world          # all ranks share one host
        m = ev if 27.0 * ev <= budget else max(1 << 20, int(budget / 27.0) // (1 << 20) * (1 << 20))
        if world > 1:                                                      # every rank uses the same size
            m = int(sharding.gather_counts_pg(dist_t, [m, 0, 0, 0])[:, 0].min())
        numa = sharding.bind_to_gpu_numa(local) if world > 1 else None     # host thread + first-touch pinned pages next to the GPU
        e2e = _e2e(eng, args, np, PinnedBuffer, ev, barrier, wl, rank, world, m, my_shards[0], shards_total)
        e2e["numa"] = numa
        if world > 1:
            e2e["ms_per_step"] = sharding.max_over_ranks(dist_t, [e2e["ms_per_step"]])[0]
            b = sharding.gather_counts_pg(dist_t, [e2e["h2d_bytes_per_step"], e2e["d2h_bytes_per_step"], 0, 0])
            e2e["h2d_bytes_per_step"], e2e["d2h_bytes_per_step"] = int(b[:, 0].sum()), int(b[:, 1].sum())
        e2e_events = e2e["events_per_gpu"]
        if world > 1:
            e2e_events = int(sharding.gather_counts_pg(dist_t, [e2e_events, 0, 0, 0])[:, 0].sum())
        e2e["value"] = sharding.throughput(e2e_events, e2e["ms_per_step"])
        if world > 1:
            r = sharding.gather_counts_pg(dist_t, [int(e2e["h2d_only_gbs_this_rank"] * 1000), 0, 0, 0])
            e2e["h2d_only_gbs_all_ranks"] = float(r[:, 0].sum()) / 1000.0      # the host's measured copy ceiling at this N

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_reference_arm(wl)

    if rank == 0:
        line = {
            "metric": METRIC, "value": sharding.throughput(total_events, step_ms), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": make_config(wl, world, args.total_events),
            "run": {"events_total": total_events, "emitted_total": total_emitted, "wall_ms_per_step": wall_step_ms,
                    "shards_per_gpu": per_rank, "path": int(sts[-1]["path"]),
                    "between_shards": "next shard generated in place, untimed" if multi else None},
            "parity": parity,
            "stage_ms": {k: avg[k] for k in ("ms_intern", "ms_sort_key", "ms_diff", "ms_assemble")},
            "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    eng.close()
    if dist_t is not None:
        dist_t.destroy_process_group()
    if parity["checked"] and not parity["match"]:
        sys.exit(1)


def _parity(eng, args, wl, my_shards, shards_total, multi, gen, last_stats, dist_t, world):
    """Count + order-sensitive digest of the emitted records of every shard this rank replays, against
    tests/golden/digests.json (rows written by oracle/oracle.c: orc_generate_replay via tests/golden/make_digests.py).
    Untimed.  -> {"checked", "match", ...}; counts cover all ranks."""
    from k8s_watcher_b200 import sharding
    ev, pods, dist, p_del, _ = wl
    try:
        with open(os.path.join(ROOT, "tests", "golden", "digests.json")) as f:
            table = json.load(f)
    except Exception:
        table = {"workloads": {}, "rows": {}}
    w = table["workloads"].get(args.workload)
    same = bool(w) and (w["events_per_shard"], w["pods_per_shard"], w["dist"], w["p_del"]) == (ev, pods, dist, p_del)
    mine, ok, checked = {}, 0, 0
    for j, g in enumerate(my_shards):
        key = "%s/%d/%d" % (args.workload, shards_total, g)
        row = table["rows"].get(key) if same else None
        if multi:
            gen(g)
            st = eng.run(dedupe=True)
        else:
            st = last_stats[j]
        dig = "%016x" % eng.digest()
        good = None
        if row is not None:
            checked += 1
            good = (row["digest"] == dig and all(int(st[k]) == row[k] for k in ("n_in", "n_survived", "n_emitted", "n_tombstones")))
            ok += bool(good)
        mine[key] = {"digest": dig, "n_emitted": int(st["n_emitted"]), "match": good}
    tot_checked, tot_ok = checked, ok
    if world > 1:
        c = sharding.gather_counts_pg(dist_t, [checked, ok, 0, 0])
        tot_checked, tot_ok = int(c[:, 0].sum()), int(c[:, 1].sum())
    return {"checked": tot_checked > 0, "match": (tot_checked == tot_ok) if tot_checked else None,
            "shards_checked": tot_checked, "shards_matching": tot_ok, "shards_total": shards_total,
            "oracle": "tests/golden/digests.json (oracle/oracle.c orc_generate_replay)", "rank0": mine}


def _key_passes(pods: int) -> int:
    cap = 256
    while cap < int(pods / 0.6) + 1:
        cap <<= 1
    return ((cap - 1).bit_length() + 7) // 8


def _peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def _traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch, per kernel, from the committed ncu --set full capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            doc = json.load(f)
        return doc.get("dram_bytes_per_launch", {}), ("static: one ncu --set full capture of this workload, not measured in this run; "
                                                       + str(doc.get("source", "profiles/traffic.json")))
    except Exception:
        return {}, None


def _e2e(eng, args, np, PinnedBuffer, ev, barrier, wl, rank, world, m, shard=None, n_shards=None):
    """The call a user of the drop-in makes: PodWatcher.replay(log) on a log that sits in pinned host memory
    (ReplayLog.pinned).  Per step, inside the timed region: host -> device copies of the columns the replay reads
    (18 B/event: uid 16, phase 1, type 1 -- the namespace column stays at home while no allow-list is configured, the
    resourceVersion column because the log is declared to be in arrival order, which is all the reference's loop knows),
    the replay, and the emit bitmap back to
    the host (1 bit/event; the host holds the log, so that is the whole result) -- pipelined chunk by chunk
    (pte_replay_pipelined).  Also measured, for context: the plain host -> device copy rate of the same columns.
    All ranks share one host: if the pinned log does not fit in half of the free host memory, every rank runs this
    leg on a shorter log of the same shape (same events-per-pod ratio) and says so."""
    from k8s_watcher_b200 import PodWatcher, ReplayLog, codes
    shard = rank if shard is None else shard
    n_shards = world if n_shards is None else n_shards
    _, pods, dist, p_del, _ = wl
    eng.generate(SEED, m, max(1, int(pods * (m / ev))), dist=dist, p_del=p_del, shard=shard, n_shards=n_shards)
    log = ReplayLog.pinned(m, list(codes.DEFAULT_NAMESPACES))
    eng.read_log(into=log.columns)                                            # set-up, untimed
    log.arrival_ordered = True            # the synthetic log is generated in replay order: the rv column stays on the host
    if args.e2e_packed_meta:
        log.pack_meta()
    w = PodWatcher("staging", engine=eng, config_dir=os.path.join(ROOT, "config"), keep_batches=1)
    w.logger.disabled = True
    batch = w.replay(log)                                                     # warm-up
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        batch = w.replay(log)
    ms = (time.perf_counter() - t0) * 1e3 / args.e2e_steps
    barrier()
    # the host -> device ceiling of this box for these buffers (all ranks copying at once), untimed leg
    t1 = time.perf_counter()
    eng.ingest(log.columns, with_namespaces=False)                            # 26 B/event (this one copies rv too)
    h2d_s = time.perf_counter() - t1
    barrier()
    res = {"value": None, "unit": UNIT, "ms_per_step": ms, "steps": args.e2e_steps, "events_per_gpu": m,
           "h2d_bytes_per_step": (17 if args.e2e_packed_meta else 18) * m, "d2h_bytes_per_step": 4 * ((m + 31) // 32),
           "n_emitted": len(batch), "device_ms_per_step": batch.stats["ms_total"],
           "h2d_only_gbs_this_rank": 26.0 * m / h2d_s / 1e9,
           "api": "PodWatcher.replay(ReplayLog in pinned memory) -> Batch (emit bitmap): pte_replay_pipelined underneath, "
                  "chunks of %d events%s" % (w.pipeline_chunk_events, "; type and phase travel as one packed byte (--e2e-packed-meta)" if args.e2e_packed_meta else "")}
    if m < ev:
        res["reduced"] = "host memory: %d ranks x 26 B x %d events do not fit in half of the free RAM; same shape, %d events per GPU" % (world, ev, m)
    log.free()
    return res


if __name__ == "__main__":
    main()

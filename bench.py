#!/usr/bin/env python
"""bench.py — weighted pair-updates/s of the pairwise hot path on N B200s (one process per GPU).

A "step" = one pass of the hot path over the whole synthetic index with the binned CSR already
resident in HBM: clear the owned triangle block, pair-emission kernels (K1/K2/Kz, K4 beside them),
containment + record compaction (K5), threshold + union-find (K6/K7) and, for N > 1, the label
all-gather + merge.
`value` = U (weighted pair-updates of the index) x steps / time, max over ranks.
`e2e`   = the same through the host-facing call (ksb_pairwise on the pinned host CSR: H2D of the
          blob, kernels, D2H of the records) — wall clock around the call; best of 5 rounds of
          min(steps, 5) calls after 3 warm-up calls, every round listed in `e2e.rounds_ms`.
`--impl reference` times the UNMODIFIED reference (oracle/_ref/pairwise_ref = src/pairwise.cpp,
OpenMP + parallel-hashmap) on the host cores on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "weighted_pair_updates_per_s"
UNIT = "pair-updates/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg2_small", "cfg3", "cfg4", "cfg5"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cutoff", type=float, default=0.3)
    ap.add_argument("--dense-budget-gb", type=float, default=0.0)
    ap.add_argument("--spill", default="auto", choices=["auto", "always", "never"], help="K3 sort-reduce spill path")
    ap.add_argument("--n-giant", type=int, default=16, help="cfg5: number of giant colours")
    ap.add_argument("--giant-size", type=int, default=4000, help="cfg5: members per giant colour")
    return ap.parse_args()


def make_workload(name: str, args=None):
    from kspider_b200 import synth
    if name == "cfg2":
        return synth.zipf_index(1000, 10_000_000, seed=12345, name="cfg2")
    if name == "cfg2_small":
        return synth.zipf_index(1000, 1_000_000, seed=12345, name="cfg2_small")
    if name == "cfg3":
        return synth.cfg3()
    if name == "cfg4":
        return synth.cfg4()
    if name == "cfg5":
        return synth.giant_index(n_giant=args.n_giant if args else 16, giant_size=args.giant_size if args else 4000)
    raise ValueError(name)


WORKLOAD_DESC = {
    "cfg2": "synthetic kProcessor-style colour table: 1k samples, 10M colours, Zipf(1.5) colour sizes 2..64 (BASELINE configs[1])",
    "cfg2_small": "cfg2 at 1/10 scale: 1k samples, 1M colours",
    "cfg3": "synthetic FracMinHash index: 10k genomes (200 clades x 50), sketch 5000, dense u64 triangle (BASELINE configs[2])",
    "cfg4": "synthetic 100k-sample index (2000 clades x 50), row-block sharded triangle (BASELINE configs[3])",
    "cfg5": "cfg2-style background + 16 giant colours of 4000 members, N = 10k (BASELINE configs[4])",
}


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region (B200_PROFILING.md's clocks line).  The
    timed region of this bench is tens of milliseconds, shorter than nvidia-smi's start-up, so the
    sampler reads NVML in-process every ~4 ms (the same counters nvidia-smi prints); nvidia-smi -lms
    is the fallback when pynvml is missing.  mark() brackets the timed region: the reported median is
    over the samples inside it (over the whole run, warm-up included, if none fell inside)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    BITS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
            0x80: "hw_power_brake_slowdown"}

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index
        self.samples, self.marks, self.nvml, self.stop_flag, self.t = [], [], None, False, None
        self.max_mhz = None

    def start(self):
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(self.gpu).uuid)
            self.handle = pynvml.nvmlDeviceGetHandleByUUID(uuid if uuid.startswith("GPU-") else "GPU-" + uuid)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        if shutil.which("nvidia-smi") is None:
            return
        self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                                      "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        self.t = threading.Thread(target=self._read, daemon=True)
        self.t.start()

    def _poll(self):
        nv = self.nvml
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag:
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM))
                self.samples.append((time.perf_counter(), mhz, int(get_reasons(self.handle))))
            except Exception:
                pass
            time.sleep(0.004)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def mark(self):
        self.marks.append(time.perf_counter())

    def stop(self) -> dict:
        if self.nvml is not None:
            self.stop_flag = True
            self.t.join(timeout=2)
            inside = [x for x in self.samples if len(self.marks) >= 2 and self.marks[0] <= x[0] <= self.marks[1]]
            use = inside or self.samples
            reasons = set()
            for _, _, bits in use:
                for b, nme in self.BITS.items():
                    if bits & b:
                        reasons.add(nme)
            return {"sm_mhz": float(np.median([x[1] for x in use])) if use else None, "sm_max_mhz": self.max_mhz,
                    "reasons": sorted(reasons), "samples": len(use), "samples_in_timed_region": len(inside),
                    "source": "NVML in-process, 4 ms period"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except Exception:
                continue
            for k, nme in enumerate(names):
                if len(r) > 5 + k and r[5 + k].lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi -lms 20"}


# ----------------------------------------------------------------------------- reference / CPU arm

def reference_sample(name: str):
    """Bounded sample of the workload for the CPU reference: the same generator and seed at a colour
    count the reference finishes in ~10-30 s on the box's cores."""
    from kspider_b200 import synth
    if name in ("cfg2", "cfg2_small"):
        return synth.zipf_index(1000, 1_000_000, seed=12345, name="cfg2_sample"), \
            "first 1/10 of the workload: same generator/seed at 1M colours (U = 1.21e8 pair-updates)"
    if name == "cfg3":
        return synth.fracminhash_index(n_clades=20, seed=31, name="cfg3_sample"), "20 of the 200 clades (2 of 10 families)"
    if name == "cfg4":
        return synth.fracminhash_index(n_clades=40, clade_size=50, sketch=1000, core=600, family_pool=400, seed=100,
                                       name="cfg4_sample"), "40 of the 2000 clades"
    return synth.giant_index(n_background=200_000, n_giant=4, giant_size=2000, name="cfg5_sample"), \
        "200k background colours + 4 giant colours of 2000"


def run_reference_once(prefix: str, threads: int) -> float:
    from oracle import reftools
    r = reftools.run_pairwise_ref(prefix, threads=threads, timeout=3600)
    if r["construct_secs"] is None:
        raise RuntimeError("reference did not print its 'pairwise hashmap construction' timer")
    return r["construct_secs"]


def reference_arm(args, steps: int, warmup: int) -> dict:
    """Times oracle/_ref/pairwise_ref (unmodified src/pairwise.cpp) with all host threads; the
    reference's own timer around its OpenMP pair loop (src/pairwise.cpp:191-239) is the clock."""
    from kspider_b200 import synth
    from oracle import reftools
    if not reftools.have_ref():
        raise RuntimeError("oracle/_ref/pairwise_ref is missing (built by __graft_entry__.build() where /root/reference exists)")
    sample, desc = reference_sample(args.workload)
    threads = max(1, min(os.cpu_count() or 1, 128))
    td = tempfile.mkdtemp(prefix="ksb_ref_")
    try:
        prefix = os.path.join(td, sample.name)
        synth.write_index_files(sample, prefix, with_names=False)
        secs = []
        for i in range(warmup + steps):
            s = run_reference_once(prefix, threads)
            if i >= warmup:
                secs.append(s)
    finally:
        shutil.rmtree(td, ignore_errors=True)
    t = float(np.mean(secs))
    return {"value": sample.n_updates / t, "unit": UNIT, "cores": threads, "kind": "reference",
            "sample": f"{desc}; {len(secs)} timed run(s) of the reference's own 'pairwise hashmap construction' timer, "
                      f"{t:.2f} s each", "secs": t, "updates": sample.n_updates}


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, 3))      # each step is a ~10-30 s CPU run: keep the arm within minutes
    warmup = 1 if args.warmup > 0 else 0
    cb = reference_arm(args, steps, warmup)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": cb["secs"] * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": WORKLOAD_DESC[args.workload], "name": args.workload, "sample": cb["sample"]},
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------- B200 arm

def main_b200(args):
    import torch
    import torch.distributed as dist
    from kspider_b200 import Index, bench_red
    from kspider_b200 import distributed as kd
    from kspider_b200.capi import check, lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != max(1, args.gpus) and rank == 0:
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE", file=sys.stderr)
    check(lib().ksb_cuda_available())
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    budget = int(args.dense_budget_gb * 1e9)

    # ---- synthetic index: generated and parsed on rank 0 only, blob broadcast once
    idx = None
    shape = np.zeros(13, dtype=np.int64)
    if rank == 0:
        s = make_workload(args.workload, args)
        idx = Index.from_csr(s.offsets, s.members, s.weights, s.total_ids, s.total_vals,
                             host_threads=min(32, os.cpu_count() or 1))
        inf = idx.info
        M_small, U_small, C_small = int(inf.n_small_members), int(inf.n_small_updates), int(inf.n_small_colors)
        shape[:] = [inf.n_samples, inf.n_colors, inf.n_members, inf.n_updates, U_small, M_small, C_small,
                    inf.payload_bytes, inf.n_gram_colors, inf.gram_planes, inf.n_gram_updates, inf.n_gram_members, inf.n_gram_macs]
        del s
    cpus = None
    if world > 1:
        t = torch.from_numpy(shape).to(device)
        dist.broadcast(t, 0)
        shape = t.cpu().numpy()
        # every rank receives the row-block shard of its own rows (cut on the host once, part of loading)
        shards = kd.make_shards(idx, world, host_threads=min(32, os.cpu_count() or 1)) if rank == 0 else None
        cpus = None if os.environ.get("KSPIDER_NO_NUMA_BIND") else kd.bind_to_gpu_numa(local_rank)
        my = kd.host_shard(shards, device, 0)     # one-off: this rank's shard in its own pinned host memory
        del shards
        dev = my.upload(device=local_rank, dense_budget_bytes=budget, spill=args.spill)
        payload_bytes_rank = torch.tensor([int(my.info.payload_bytes)], dtype=torch.int64, device=device)
        dist.all_reduce(payload_bytes_rank)
        shape[7] = int(payload_bytes_rank.cpu())  # bytes moved host -> GPUs per e2e step: all shards, each over its rank's link
    else:
        dev = idx.upload(device=local_rank, dense_budget_bytes=budget, spill=args.spill)
    N, C_, M, U, U_small, M_small, C_small, payload_bytes, C_gram, P_gram, U_gram, M_gram, MAC_gram = (int(v) for v in shape)

    host_ms = {"accumulate+post": 0.0, "cluster+merge": 0.0}

    def step():
        t0 = time.perf_counter()
        res = dev.run_all_passes(want_records=False)
        t1 = time.perf_counter()
        if world > 1:
            labels, ne = kd.cluster_merged(res, N, "max_cont", args.cutoff, device, to_host=True)  # the step's result reaches the host
        else:
            labels, ne = res.cluster(N, "max_cont", args.cutoff)
        t2 = time.perf_counter()
        host_ms["accumulate+post"] += (t1 - t0) * 1e3
        host_ms["cluster+merge"] += (t2 - t1) * 1e3
        launches = res.info.gpu_launches
        rows = res.info.n_rows
        res.close()
        return launches, rows, ne

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()   # sampled from the warm-up on: the timed region alone can be shorter than one sample
    for _ in range(args.warmup):
        step()
    t_before = dev.info.t
    l_before = dev.info.gpu_launches
    before = {k: getattr(t_before, k) for k in ("zero_ms", "small_ms", "mid_ms", "zero_touch_ms", "gram_ms", "spill_emit_ms", "spill_sort_ms", "accumulate_ms", "post_ms")}
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    host_ms.update({k: 0.0 for k in host_ms})
    sampler.mark()
    ev0.record()
    w0 = time.perf_counter()
    launches = 0
    for _ in range(args.steps):
        l, rows, n_edges = step()
        launches = l
    ev1.record()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - w0) * 1e3
    sampler.mark()
    dev_ms = ev0.elapsed_time(ev1)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    t_after = dev.info.t
    phase = {k: (getattr(t_after, k) - before[k]) / args.steps for k in before}
    # launches: the result's counter is cumulative over the device handle (+ the cluster kernels of the last step)
    per_step_launches = (dev.info.gpu_launches - l_before) / args.steps + (launches - dev.info.gpu_launches)
    # roofline leg: the same passes with K4 serialised after K1/K2, so that each kernel's library
    # timer is that kernel running alone (in the timed steps above K4 runs beside K1/K2 on a 2nd stream)
    n_alone = 3
    dev.set_overlap(False)
    dev.run_all_passes(want_records=False).close()
    t0 = dev.info.t
    b0 = {k: getattr(t0, k) for k in before}
    for _ in range(n_alone):
        dev.run_all_passes(want_records=False).close()
    t1 = dev.info.t
    alone = {k: (getattr(t1, k) - b0[k]) / n_alone for k in before}
    dev.set_overlap(True)
    ms = torch.tensor([dev_ms, wall_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms = (float(v) for v in ms.cpu())
    ms_per_step = dev_ms / args.steps
    value = U / (ms_per_step * 1e-3)

    # ---- end to end through the host-facing call (host CSR in pinned memory -> records on the host)
    e2e = None
    if not args.no_e2e:
        n_e2e = max(1, min(args.steps, 5))
        d2h = 0
        e2e_lib = {}

        def e2e_step():
            nonlocal d2h
            src = idx if world == 1 else my   # N>1: every rank uploads its own row-block shard
            res = src.pairwise(device=local_rank, want_records=True, dense_budget_bytes=budget, spill=args.spill)
            if world > 1:
                labels, _ = kd.cluster_merged(res, N, "max_cont", args.cutoff, device, to_host=True)
            else:
                labels, _ = res.cluster(N, "max_cont", args.cutoff)
            d2h = res.info.n_rows * 32 + 4 * N
            tt = res.info.t
            e2e_lib.update({k: round(float(getattr(tt, k)), 3) for k in ("h2d_ms", "accumulate_ms", "post_ms", "d2h_ms", "cluster_ms", "host_setup_ms", "host_accumulate_ms", "host_post_ms", "host_free_ms")})
            res.close()

        for _ in range(3):
            e2e_step()
        # five rounds of n_e2e steps, the best round reported (all five listed in `rounds_ms`): the host-facing
        # call is a few ms, so a single host hiccup (a 14 ms stall inside one cudaMemcpyAsync was seen on a
        # shared box) would otherwise double a 5-step mean
        rounds = []
        for _ in range(5):
            barrier()
            w0 = time.perf_counter()
            for _ in range(n_e2e):
                e2e_step()
            barrier()
            rounds.append((time.perf_counter() - w0) * 1e3 / n_e2e)
        r_ms = torch.tensor(rounds, dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(r_ms, op=dist.ReduceOp.MAX)
        rounds = [float(v) for v in r_ms.cpu()]
        e_ms = torch.tensor([min(rounds)], dtype=torch.float64, device=device)
        d2h_t = torch.tensor([d2h], dtype=torch.int64, device=device)
        if world > 1:
            dist.all_reduce(d2h_t, op=dist.ReduceOp.SUM)
        e2e = {"value": U / (float(e_ms.cpu()) * 1e-3), "unit": UNIT, "h2d_bytes_per_step": payload_bytes,
               "d2h_bytes_per_step": int(d2h_t.cpu()), "ms_per_step": float(e_ms.cpu()), "steps": n_e2e,
               "rounds_ms": [round(v, 4) for v in rounds],
               "timer": "host wall clock around ksb_pairwise + ksb_cluster (synchronous calls); best of 5 rounds of `steps` steps, 3 warm-up steps, max over ranks per round",
               "library_phases_ms_last_step": e2e_lib,
               "host_numa": (f"rank bound to the {len(cpus)} CPUs next to its GPU before its pinned buffers were allocated" if cpus else "unbound")}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (rank 0's launch; algorithmic bytes per SURVEY 8d)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    own_frac = dev.info.n_updates / U if U else 1.0
    peak_tensor = 2.0 * float(peaks.get("bf16_tflops", 2250.0))  # int8 dense = 2 x bf16 on the same tensor pipe
    gram_roof = None
    gram_traffic = None
    try:
        gram_traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload, {}).get("k4_gram")
    except Exception:
        pass
    if C_gram:
        # int8 ops the Gram launches of one step must evaluate: per window, colours x cells of the window's
        # upper triangle x byte planes (the loader's n_gram_macs), x 2, x this rank's share of the rows
        g_ops = 2.0 * MAC_gram * own_frac
        g_ms = alone["gram_ms"]
        gram_roof = {"bound": "tensor", "kernel": "k4_gram (+kb_build, memsets)", "achieved": g_ops / (g_ms * 1e-3) / 1e12,
                     "peak": peak_tensor, "unit": "TFLOP/s", "frac": g_ops / (g_ms * 1e-3) / 1e12 / peak_tensor,
                     "traffic": gram_traffic, "kernel_ms": g_ms, "algorithmic_ops": g_ops, "ops": "u8 x u8 -> s32 multiply-adds x 2",
                     "peak_source": "2 x MEASURED_PEAKS.json bf16_tflops (burst); no int8 figure in that file",
                     "updates_per_s": U_gram * own_frac / (g_ms * 1e-3) if g_ms > 0 else 0.0,
                     "colors": C_gram, "planes": P_gram}
    if alone["small_ms"] >= alone["mid_ms"]:
        kname, k_ms = "k1_small", alone["small_ms"]
        k_bytes = 16 * U_small * own_frac + 4 * M_small + 4 * C_small
        k_upd = U_small * own_frac
    else:
        kname, k_ms = "k2_mid", alone["mid_ms"]
        k_bytes = 16 * (U - U_small - U_gram) * own_frac + 4 * (M - M_small - M_gram) + 12 * (C_ - C_small - C_gram)
        k_upd = (U - U_small - U_gram) * own_frac
    achieved = k_bytes / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload, {}).get(kname)
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "kernel_ms": k_ms, "algorithmic_bytes": k_bytes,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (burst copy)" if peaks else "fallback 6650 GB/s",
                "updates_per_s": k_upd / (k_ms * 1e-3) if k_ms > 0 else 0.0}
    if gram_roof is not None and alone["gram_ms"] > k_ms:
        gram_roof["other"] = roofline       # the atomic kernel's HBM roofline rides along
        roofline = gram_roof
    elif gram_roof is not None:
        roofline["other"] = gram_roof
    # in-run red.global.add.u64 ceiling on a buffer of the owned block's size
    red = None
    try:
        cells = max(dev.info.owned_cells, 1024)
        n_ops = 1 << 28
        red = {}
        for pat, nm in ((0, "random"), (1, "warp_contiguous")):
            red[nm + "_ops_per_s"] = n_ops / (bench_red(cells, n_ops, 64, pat, 3) * 1e-3)
        red["cells"] = cells
        red["kernel_frac_of_warp_contiguous"] = k_upd / (k_ms * 1e-3) / red["warp_contiguous_ops_per_s"] if k_ms > 0 else None
        # the atomic kernel's updates land on member-random columns of a row: its honest ceiling is the RANDOM rate
        red["kernel_frac_of_random"] = k_upd / (k_ms * 1e-3) / red["random_ops_per_s"] if k_ms > 0 else None
        red["kernel"] = kname
    except Exception as e:  # the micro-benchmark is informative only
        red = {"error": str(e)}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            cb = reference_arm(args, 1, 0)
            cpu = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        except Exception as e:
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"unavailable: {e}"}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": WORKLOAD_DESC[args.workload], "name": args.workload, "samples": N, "colors": C_,
                       "members": M, "pair_updates": U, "rows": rows, "edges": int(n_edges), "cutoff": args.cutoff,
                       "parallelism": f"row-blocks x{world}", "passes": dev.n_passes, "spill": bool(dev.is_spill),
                       "l2": f"inputs larger than L2 (CSR blob {payload_bytes / 1e6:.0f} MB streamed every step; the triangle block is cleared every step)"},
            "phases_ms": phase, "kernels_alone_ms": alone, "host_ms_per_step": {k: v / args.steps for k, v in host_ms.items()}, "wall_ms_per_step": wall_ms / args.steps,
            "timer": "torch.cuda.Event pair on the current stream around the K steps (every library call is stream-synchronous); phases: library CUDA events on the launching stream",
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(round(per_step_launches * args.steps)),
            "roofline": roofline, "red_roofline": red, "cpu_baseline": cpu}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        main_reference(a)
    else:
        main_b200(a)

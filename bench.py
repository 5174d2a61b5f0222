#!/usr/bin/env python
"""bench.py -- weighted pair-updates/s of the pairwise hot path on N B200s (one process per GPU).

A "step" = one CLUSTER-ONLY pass of the hot path over the whole synthetic index with the binned CSR already resident
in HBM (north_star 3: only the labels return to the host): clear the owned block of the banded triangle,
pair-emission kernels (K1/K2/Kz, K4 beside them), ONE fused scan (K56: containment + the reference's threshold +
append of the kept pairs only), unions over the kept-edge list, labels to the host; for N > 1 the label all-gather +
merge.  One host synchronisation per step.
`value`  = U (weighted pair-updates of the index) x steps / device time, max over ranks.
`e2e`    = the same through the host-facing call ksb_pairwise_cluster on the pinned HOST index: H2D of the blob
           (pipelined), the same kernels, labels D2H; wall clock around the call; MEDIAN of 5 rounds of min(steps, 5)
           calls after 3 warm-up calls (all rounds listed, min / max beside it).  `e2e_records` = the TSV-producing
           variant (ksb_pairwise with the 32-byte records brought to the host + ksb_cluster on them).
`wall`   = pairwise wall-time (the other half of BASELINE.json's metric): the drop-in binary kspider_b200/bin/pairwise
           (= kSpider::pairwise(prefix, threads): index files -> both TSVs) as a subprocess on the full config's files,
           split into load / GPU / TSV phases; and both binaries (ours, the unmodified reference) on the same files
           of a bounded sample.
`--impl reference` times the UNMODIFIED reference (oracle/_ref/pairwise_ref = src/pairwise.cpp, OpenMP +
parallel-hashmap) on the host cores on the FULL config, once (its own 'pairwise hashmap construction' timer = the
throughput clock, subprocess wall beside it), and at one thread on a stated sample.
"""
from __future__ import annotations

import argparse
import json
import os
import re
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
DEFAULT_WORKLOAD = "cfg4"   # the config north_star's multi-GPU target names; it fits one GPU (banded accumulator: 0.8 GB)
# cut-offs that split the config non-trivially (max containment): cfg2's containments sit at 0.022-0.027,
# the clade configs have within-clade containments ~0.6 and ~0.1 across clades of a family
DEFAULT_CUTOFF = {"cfg2": 0.0258, "cfg2_small": 0.0258, "cfg3": 0.3, "cfg4": 0.3, "cfg5": 0.3}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=["cfg2", "cfg2_small", "cfg3", "cfg4", "cfg5"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-wall", action="store_true")
    ap.add_argument("--cutoff", type=float, default=None, help="clustering cut-off on max containment (default: per workload)")
    ap.add_argument("--dense-budget-gb", type=float, default=0.0)
    ap.add_argument("--spill", default="auto", choices=["auto", "always", "never"], help="K3 sort-reduce spill path")
    ap.add_argument("--n-giant", type=int, default=16, help="cfg5: number of giant colours")
    ap.add_argument("--giant-size", type=int, default=4000, help="cfg5: members per giant colour")
    a = ap.parse_args()
    if a.cutoff is None:
        a.cutoff = DEFAULT_CUTOFF[a.workload]
    return a


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

def host_threads() -> int:
    return max(1, min(os.cpu_count() or 1, 128))


def reference_sample(name: str, one_thread: bool = False):
    """Bounded sample of the workload for the CPU legs of the B200 arm: the same generator and seed at a size the
    reference finishes in ~10-30 s on the box's cores (one_thread: at a size ONE thread finishes in that time)."""
    from kspider_b200 import synth
    if one_thread and name == "cfg4":
        return synth.fracminhash_index(n_clades=40, clade_size=50, sketch=1000, core=600, family_pool=400, seed=100,
                                       name="cfg4_sample_t1"), "40 of the 2000 clades (2 of 100 families)"
    if one_thread and name == "cfg3":
        return synth.fracminhash_index(n_clades=20, seed=31, name="cfg3_sample_t1"), "20 of the 200 clades (1 of 10 families)"
    if name in ("cfg2", "cfg2_small"):
        return synth.zipf_index(1000, 1_000_000, seed=12345, name="cfg2_sample"), \
            "1/10 of the workload: same generator/seed at 1M colours (U = 1.21e8 pair-updates)"
    if name == "cfg3":
        return synth.fracminhash_index(n_clades=40, seed=31, name="cfg3_sample"), "40 of the 200 clades (2 of 10 families)"
    if name == "cfg4":
        return synth.fracminhash_index(n_clades=200, clade_size=50, sketch=1000, core=600, family_pool=400, seed=100,
                                       name="cfg4_sample"), "200 of the 2000 clades (10 of 100 families)"
    return synth.giant_index(n_background=200_000, n_giant=4, giant_size=2000, name="cfg5_sample"), \
        "200k background colours + 4 giant colours of 2000"


def write_files_with_reference_writer(s, prefix: str) -> None:
    """The three index files through oracle/_ref/index_writer_ref (the reference's own vendored phmap does the
    dumping): the reference arm's process tree never maps the product library."""
    from oracle import reftools
    n, c = s.n_samples, s.n_colors
    color_ids = np.concatenate([np.arange(1, n + 1, dtype=np.uint64), np.arange(n + 1, n + 1 + c, dtype=np.uint64)])
    offsets = np.concatenate([np.arange(0, n + 1, dtype=np.uint64), s.offsets[1:] + np.uint64(n)])
    members = np.concatenate([np.arange(1, n + 1, dtype=np.uint32), s.members])
    count_vals = np.concatenate([np.ones(n, dtype=np.uint64), s.weights.astype(np.uint64)])
    raw = prefix + ".raw"
    reftools.write_raw_arrays(raw, s.total_ids, s.total_vals, color_ids, count_vals, color_ids, offsets, members)
    reftools.run_index_writer_ref(raw, prefix)
    os.remove(raw)


_T = {"map": r"mapping colors to groups: ([0-9.eE+-]+) secs", "parse": r"parsing index colors: ([0-9.eE+-]+) secs",
      "count": r"kmer counting: ([0-9.eE+-]+) secs", "construct": r"pairwise hashmap construction: ([0-9.eE+-]+) secs"}


def run_reference_binary(prefix: str, threads: int) -> dict:
    """oracle/_ref/pairwise_ref <prefix> <threads> as a subprocess: wall clock + its own stdout timers
    (src/pairwise.cpp:133,155,181,239); what is left of the wall is the serial TSV writer (:242-275) and process start."""
    from oracle import reftools
    t0 = time.perf_counter()
    p = subprocess.run([reftools.PAIRWISE_REF, prefix, str(int(threads))], check=True, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=7200)
    wall = time.perf_counter() - t0
    t = {}
    for k, pat in _T.items():
        m = re.search(pat, p.stdout)
        t[k] = float(m.group(1)) if m else None
    if t["construct"] is None:
        raise RuntimeError("reference did not print its 'pairwise hashmap construction' timer")
    load = sum(v for k, v in t.items() if k != "construct" and v is not None)
    return {"wall_s": wall, "load_s": load, "construct_s": t["construct"], "write_and_start_s": wall - load - t["construct"],
            "tsv_bytes": os.path.getsize(prefix + "_kSpider_pairwise.tsv"), "threads": threads}


def run_dropin_binary(prefix: str, threads: int, env_extra=None) -> dict:
    """kspider_b200/bin/pairwise <prefix> <threads> (csrc/pairwise_main.cpp -> kSpider::pairwise) as a subprocess:
    wall clock (process start and CUDA context creation included) + its phases (KSPIDER_VERBOSE line on stderr)."""
    exe = os.path.join(ROOT, "kspider_b200", "bin", "pairwise")
    env = dict(os.environ, KSPIDER_VERBOSE="1", **(env_extra or {}))
    t0 = time.perf_counter()
    p = subprocess.run([exe, prefix, str(int(threads))], check=True, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                       text=True, env=env, timeout=3600)
    wall = time.perf_counter() - t0
    ph = {k: float(v) for k, v in re.findall(r"(\w+_s)=([0-9.eE+-]+)", p.stderr)}
    out = {"wall_s": wall, "threads": threads, "tsv_bytes": os.path.getsize(prefix + "_kSpider_pairwise.tsv")}
    out.update(ph)
    out["start_and_exit_s"] = wall - sum(ph.values())
    return out


def index_file_bytes(prefix: str) -> int:
    return sum(os.path.getsize(prefix + suf) for suf in ("_color_to_sources.bin", "_color_count.bin", "_groupID_to_kmerCount.bin"))


def cpu_baseline_on_sample(args, also_dropin: bool) -> dict:
    """The unmodified reference on a bounded sample of the workload (all host threads), and -- same files -- the drop-in."""
    from oracle import reftools
    if not reftools.have_ref():
        raise RuntimeError("oracle/_ref/pairwise_ref is missing (built by __graft_entry__.build() where /root/reference exists)")
    sample, desc = reference_sample(args.workload)
    threads = host_threads()
    td = tempfile.mkdtemp(prefix="ksb_ref_")
    try:
        prefix = os.path.join(td, sample.name)
        write_files_with_reference_writer(sample, prefix)
        r = run_reference_binary(prefix, threads)
        d = None
        if also_dropin:
            for f in ("_kSpider_pairwise.tsv", "_kSpider_seqToKmersNo.tsv"):
                os.replace(prefix + f, prefix + f + ".ref")
            d = run_dropin_binary(prefix, threads)
        fbytes = index_file_bytes(prefix)
    finally:
        shutil.rmtree(td, ignore_errors=True)
    cb = {"value": sample.n_updates / r["construct_s"], "unit": UNIT, "cores": threads, "kind": "reference",
          "sample": f"{desc}; one run of the unmodified reference binary, its own 'pairwise hashmap construction' timer: "
                    f"{r['construct_s']:.2f} s for {sample.n_updates} updates"}
    return {"cpu_baseline": cb, "reference": r, "dropin": d, "index_file_bytes": fbytes, "updates": sample.n_updates,
            "sample": desc}


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import reftools
    if not reftools.have_ref():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/pairwise_ref is not built on this box"}), flush=True)
        return
    threads = host_threads()
    s = make_workload(args.workload, args)
    td = tempfile.mkdtemp(prefix="ksb_ref_")
    try:
        prefix = os.path.join(td, s.name)
        t0 = time.perf_counter()
        write_files_with_reference_writer(s, prefix)
        t_files = time.perf_counter() - t0
        fbytes = index_file_bytes(prefix)
        full = run_reference_binary(prefix, threads)     # ONE run of the full config: tens of seconds to minutes of 16+ cores
        shutil.rmtree(td, ignore_errors=True)
        # one thread, on the bounded sample (a full-config run at T = 1 is several minutes to hours)
        os.makedirs(td, exist_ok=True)
        sample, desc = reference_sample(args.workload, one_thread=True)
        sp = os.path.join(td, sample.name)
        write_files_with_reference_writer(sample, sp)
        t1 = run_reference_binary(sp, 1)
    finally:
        shutil.rmtree(td, ignore_errors=True)
    value = s.n_updates / full["construct_s"]
    cb = {"value": value, "unit": UNIT, "cores": threads, "kind": "reference",
          "sample": f"the FULL config, one run: {s.n_updates} updates in {full['construct_s']:.2f} s of the reference's own "
                    f"'pairwise hashmap construction' timer (src/pairwise.cpp:191-239)"}
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": 1, "warmup": 0, "steps_requested": args.steps, "ms_per_step": full["construct_s"] * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "same_config": True,
            "config": {"workload": WORKLOAD_DESC[args.workload], "name": args.workload, "samples": s.n_samples,
                       "colors": s.n_colors, "pair_updates": s.n_updates,
                       "inputs": f"index files written by oracle/_ref/index_writer_ref ({fbytes} bytes, {t_files:.1f} s, untimed)"},
            "cpu_baseline": cb,
            "wall": {"reference_full": full, "index_file_bytes": fbytes,
                     "one_thread_on_sample": {"sample": desc, "updates": sample.n_updates, **t1,
                                              "value": sample.n_updates / t1["construct_s"]}},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------- B200 arm

PHASES = ("zero_ms", "small_ms", "mid_ms", "zero_touch_ms", "gram_ms", "spill_emit_ms", "spill_sort_ms", "accumulate_ms",
          "post_ms", "cluster_ms")


def traffic_of(workload: str, kernel: str):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture (profiles/traffic.json names the
    commit and the command of every capture); None when there is none for this workload."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        e = t.get(workload, {}).get(kernel)
        if e is None:
            return None, None
        if isinstance(e, dict):
            return e.get("bytes"), e.get("source")
        return e, t.get("_source", {}).get(workload)
    except Exception:
        return None, None


def main_b200(args):
    import torch
    import torch.distributed as dist
    from kspider_b200 import Index, bench_red
    from kspider_b200 import distributed as kd
    from kspider_b200.api import bench_imma
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
    threads = min(32, os.cpu_count() or 1)

    # ---- synthetic index: generated and parsed on rank 0 only
    idx = None
    s = None
    shape = np.zeros(15, dtype=np.int64)
    t_build = 0.0
    if rank == 0:
        s = make_workload(args.workload, args)
        t0 = time.perf_counter()
        idx = Index.from_csr(s.offsets, s.members, s.weights, s.total_ids, s.total_vals, host_threads=threads)
        t_build = time.perf_counter() - t0
        inf = idx.info
        sizes = np.diff(s.offsets).astype(np.int64)
        wsum = int((s.weights.astype(np.uint64) * (sizes * (sizes - 1) // 2).astype(np.uint64)).sum(dtype=np.uint64))  # < 2^63 on every config
        shape[:] = [inf.n_samples, inf.n_colors, inf.n_members, inf.n_updates, inf.n_small_updates, inf.n_small_members,
                    inf.n_small_colors, inf.payload_bytes, inf.n_gram_colors, inf.gram_planes, inf.n_gram_updates,
                    inf.n_gram_members, inf.n_gram_macs, wsum, inf.n_gram_cells]
    cpus = None
    my = None
    if world > 1:
        t = torch.from_numpy(shape).to(device)
        dist.broadcast(t, 0)
        shape = t.cpu().numpy()
        # every rank receives the row-block shard of its own rows (cut on the host once, part of loading)
        shards = kd.make_shards(idx, world, host_threads=threads) if rank == 0 else None
        cpus = None if os.environ.get("KSPIDER_NO_NUMA_BIND") else kd.bind_to_gpu_numa(local_rank)
        my = kd.host_shard(shards, device, 0)     # one-off: this rank's shard in its own pinned host memory
        del shards
        dev = my.upload(device=local_rank, dense_budget_bytes=budget, spill=args.spill)
        payload_bytes_rank = torch.tensor([int(my.info.payload_bytes)], dtype=torch.int64, device=device)
        dist.all_reduce(payload_bytes_rank)
        shape[7] = int(payload_bytes_rank.cpu())  # bytes moved host -> GPUs per e2e step: all shards, each over its rank's link
    else:
        dev = idx.upload(device=local_rank, dense_budget_bytes=budget, spill=args.spill)
    N, C_, M, U, U_small, M_small, C_small, payload_bytes, C_gram, P_gram, U_gram, M_gram, MAC_gram, WSUM, CELLS_gram = (int(v) for v in shape)
    spill = bool(dev.is_spill)
    gram_form = int((my if my is not None else idx).info.gram_form)   # bit 0: bitmaps in the blob; >> 8: bits of a window-relative member

    host_ms = {"step": 0.0}
    last = {}

    def step():
        t0 = time.perf_counter()
        if spill:      # K3 has no fused scan: sort-reduce run, then K6 + K7 over its records
            res = dev.run_all_passes(want_records=False)
            if world > 1:
                labels, ne = kd.cluster_merged(res, N, "max_cont", args.cutoff, device, to_host=True)
            else:
                labels, ne = res.cluster(N, "max_cont", args.cutoff)
            nr = res.info.n_rows
            res.close()
        elif world > 1:
            labels, ne, nr = kd.cluster_step(dev, N, "max_cont", args.cutoff, device)  # the step's result reaches the host
        else:
            labels, ne, nr = dev.cluster(N, "max_cont", args.cutoff)
        host_ms["step"] += (time.perf_counter() - t0) * 1e3
        last.update(labels=labels, edges=ne, rows=nr)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timers():
        t = dev.info.t
        return {k: float(getattr(t, k)) for k in PHASES}

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()   # sampled from the warm-up on: the timed region alone can be shorter than one sample
    for _ in range(max(3, args.warmup)):
        step()
    before = timers()
    l_before = dev.info.gpu_launches
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    host_ms["step"] = 0.0
    sampler.mark()
    ev0.record()
    w0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    ev1.record()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - w0) * 1e3
    host_step_ms = host_ms["step"] / args.steps
    sampler.mark()
    dev_ms = ev0.elapsed_time(ev1)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    # the timed steps replay ONE CUDA graph per step (events inside a graph carry no timestamps), so the per-phase
    # timers come from two extra legs of the same step enqueued kernel by kernel:
    #   phases leg   -- K4 beside K1/K2 on its own stream, as in the timed steps;
    #   roofline leg -- K4 serialised after K1/K2, so that each kernel's timer is that kernel running alone
    per_step_launches = (dev.info.gpu_launches - l_before) / args.steps + (2 if world > 1 else 0)  # + the K7 merge kernels
    labels_step = np.array(last["labels"], copy=True)
    rows_l = torch.tensor([int(last["rows"]), int(last["edges"])], dtype=torch.int64, device=device)
    if world > 1:
        dist.all_reduce(rows_l)
    rows, n_edges = (int(v) for v in rows_l.cpu())
    phase = {k: 0.0 for k in PHASES}
    alone = {k: 0.0 for k in PHASES}
    if spill:
        after = timers()
        phase = {k: (after[k] - before[k]) / args.steps for k in PHASES}
    else:
        n_leg = 3
        dev.set_graph(False)
        step()
        b0 = timers()
        for _ in range(n_leg):
            step()
        b1 = timers()
        phase = {k: (b1[k] - b0[k]) / n_leg for k in PHASES}
        dev.set_overlap(False)
        step()
        b0 = timers()
        for _ in range(n_leg):
            step()
        b1 = timers()
        alone = {k: (b1[k] - b0[k]) / n_leg for k in PHASES}
        dev.set_overlap(True)
        dev.set_graph(True)
    ms = torch.tensor([dev_ms, wall_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms = (float(v) for v in ms.cpu())
    ms_per_step = dev_ms / args.steps
    value = U / (ms_per_step * 1e-3)
    own_frac = dev.info.n_updates / U if U else 1.0
    owned_cells = int(dev.info.owned_cells)

    # ---- end to end through the host-facing call (host index in pinned memory -> labels on the host)
    e2e = None
    e2e_rec = None
    check_blk = None
    if not args.no_e2e:
        n_e2e = max(1, min(args.steps, 5))
        e2e_lib = {}
        e2e_labels = {}

        def e2e_step():
            if world > 1:  # every rank uploads its own row-block shard over its own PCIe link
                labels, ne, nr, info = kd.pairwise_cluster_step(my, N, "max_cont", args.cutoff, device, dense_budget_bytes=budget, spill=args.spill)
            else:
                labels, ne, nr, info = idx.pairwise_cluster("max_cont", args.cutoff, device=local_rank, dense_budget_bytes=budget, spill=args.spill)
            e2e_labels["labels"] = labels
            tt = info.t
            e2e_lib.update({k: round(float(getattr(tt, k)), 3) for k in ("h2d_ms", "accumulate_ms", "post_ms", "cluster_ms", "host_setup_ms", "host_accumulate_ms", "host_free_ms")})

        def timed_rounds(fn, n_rounds=5):
            for _ in range(3):
                fn()
            rounds = []
            for _ in range(n_rounds):
                barrier()
                w0 = time.perf_counter()
                for _ in range(n_e2e):
                    fn()
                barrier()
                rounds.append((time.perf_counter() - w0) * 1e3 / n_e2e)
            r_ms = torch.tensor(rounds, dtype=torch.float64, device=device)
            if world > 1:
                dist.all_reduce(r_ms, op=dist.ReduceOp.MAX)
            return [float(v) for v in r_ms.cpu()]

        rounds = timed_rounds(e2e_step)
        med = float(np.median(rounds))
        e2e = {"value": U / (med * 1e-3), "unit": UNIT, "h2d_bytes_per_step": payload_bytes,
               "d2h_bytes_per_step": (4 * N + 16) * world, "ms_per_step": med, "steps": n_e2e,
               "statistic": "median of the rounds", "min_ms": min(rounds), "max_ms": max(rounds),
               "rounds_ms": [round(v, 4) for v in rounds], "result": "cluster labels (uint32[N]) + kept-edge and row counts",
               "timer": "host wall clock around ksb_pairwise_cluster (a synchronous call: pipelined H2D of the pinned blob, kernels, labels D2H; the call keeps its device handle -- allocations and plans, no data -- between calls on one index, KSPIDER_E2E_CACHE=0 disables); 5 rounds of `steps` calls after 3 warm-up calls, max over ranks per round",
               "library_phases_ms_last_step": dict(e2e_lib),
               "host_numa": (f"rank bound to the {len(cpus)} CPUs next to its GPU before its pinned buffers were allocated" if cpus else "unbound")}
        if world == 1:
            # the TSV-producing variant: 32-byte records of every touched pair brought to the host, clustered from them
            rec = {}

            def rec_step():
                res = idx.pairwise(device=local_rank, want_records=True, dense_budget_bytes=budget, spill=args.spill)
                lab, ne = res.cluster(N, "max_cont", args.cutoff)
                rec.update(labels=lab, n_rows=int(res.info.n_rows), edges=int(ne), res=res)
                old = rec.pop("prev", None)
                if old is not None:
                    old.close()
                rec["prev"] = res

            rr = timed_rounds(rec_step, n_rounds=3)
            rmed = float(np.median(rr))
            e2e_rec = {"value": U / (rmed * 1e-3), "unit": UNIT, "ms_per_step": rmed, "rounds_ms": [round(v, 4) for v in rr],
                       "h2d_bytes_per_step": payload_bytes, "d2h_bytes_per_step": rec["n_rows"] * 32 + 4 * N,
                       "result": "ksb_pairwise with the 32-byte records of every touched pair on the host (what ksb_write_tsv formats) + ksb_cluster"}
            # ---- in-run parity checks (cheap, size-independent): the records' shared counts add up to the closed form
            # sum_c w_c s_c (s_c - 1) / 2 of the input; the fused cluster-only step, the end-to-end call and the record
            # path agree on rows, kept edges and every label
            recs = rec["res"].records()
            got_sum = int(recs["shared_kmers"].sum(dtype=np.uint64))
            check_blk = {"sum_shared": got_sum, "sum_shared_expected": WSUM, "rows_records": rec["n_rows"], "rows_fused": rows,
                         "edges_records": rec["edges"], "edges_fused": n_edges,
                         "labels_equal": bool(np.array_equal(rec["labels"], labels_step) and np.array_equal(rec["labels"], e2e_labels["labels"])),
                         "clusters": int(np.unique(labels_step).size)}
            check_blk["ok"] = bool(got_sum == WSUM and rec["n_rows"] == rows and rec["edges"] == n_edges and check_blk["labels_equal"])
            rec["res"].close()
            del recs

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- rooflines of the step's kernels (rank 0's launches; algorithmic bytes per SURVEY 8d), the slowest one first
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (burst copy)" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    kernels = []
    if alone["small_ms"] > 0:
        kb = 16 * U_small * own_frac + 4 * M_small + 4 * C_small
        tr, trs = traffic_of(args.workload, "k1_small")
        kernels.append({"bound": "hbm", "kernel": "k1_small", "kernel_ms": alone["small_ms"], "achieved": kb / (alone["small_ms"] * 1e-3) / 1e9,
                        "peak": peak, "unit": "GB/s", "algorithmic_bytes": kb, "traffic": tr, "traffic_source": trs, "peak_source": peak_src,
                        "updates_per_s": U_small * own_frac / (alone["small_ms"] * 1e-3)})
    if alone["mid_ms"] > 0:
        um = (U - U_small - U_gram) * own_frac
        kb = 16 * um + 4 * (M - M_small - M_gram) + 12 * (C_ - C_small - C_gram)
        tr, trs = traffic_of(args.workload, "k2_mid")
        kernels.append({"bound": "hbm", "kernel": "k2_mid", "kernel_ms": alone["mid_ms"], "achieved": kb / (alone["mid_ms"] * 1e-3) / 1e9,
                        "peak": peak, "unit": "GB/s", "algorithmic_bytes": kb, "traffic": tr, "traffic_source": trs, "peak_source": peak_src,
                        "updates_per_s": um / (alone["mid_ms"] * 1e-3)})
    if alone["post_ms"] > 0:
        kb = 8 * owned_cells + 8 * int(last["edges"])
        tr, trs = traffic_of(args.workload, "k56")
        kernels.append({"bound": "hbm", "kernel": "k56_rows / k56_scan (cluster-only post-pass)", "kernel_ms": alone["post_ms"], "achieved": kb / (alone["post_ms"] * 1e-3) / 1e9,
                        "peak": peak, "unit": "GB/s", "algorithmic_bytes": kb, "traffic": tr, "traffic_source": trs, "peak_source": peak_src,
                        "bytes": "8 B per cell of the owned block (read once) + 8 B per kept edge (algorithmic: u64 cells; with u32 cells the scan reads half of it)"})
    imma = None
    if C_gram and alone["gram_ms"] > 0:
        try:   # int8 tensor peak measured in this run, in K4's own instruction shape
            pms, pmacs = bench_imma(2048, 5)
            imma = {"tops": 2.0 * pmacs / (pms * 1e-3) / 1e12, "ms": pms, "macs": pmacs,
                    "kernel": "k_imma_peak: tcgen05.mma kind::i8 cta_group::1 M128 N256 K32 on resident tiles, one CTA per SM"}
        except Exception as e:
            imma = {"error": str(e)}
        g_ops = 2.0 * MAC_gram * own_frac
        p_t = imma["tops"] if imma and "tops" in imma else 2.0 * float(peaks.get("bf16_tflops", 2250.0))
        tr, trs = traffic_of(args.workload, "k4_gram")
        g_ms = alone["gram_ms"]
        tensor = {"bound": "tensor", "achieved": g_ops / (g_ms * 1e-3) / 1e12, "peak": p_t, "unit": "TFLOP/s", "algorithmic_ops": g_ops,
                  "ops": "u8 x u8 -> s32 multiply-adds x 2 (loader's n_gram_macs: per window, colours x width(width-1)/2 x byte planes)",
                  "peak_source": "int8 tcgen05 peak measured in this run (ksb_bench_imma)" if imma and "tops" in imma else "2 x MEASURED_PEAKS.json bf16_tflops"}
        tensor["frac"] = tensor["achieved"] / p_t
        # the other bound of the same launches: every cell a window can update takes one red.global.add.u64 (8 B read +
        # 8 B write of the accumulator) and every member is read once by the bitmap build
        g_bytes = 16.0 * CELLS_gram * own_frac + 4.0 * M_gram
        hbm = {"bound": "hbm", "achieved": g_bytes / (g_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "algorithmic_bytes": g_bytes,
               "bytes": "16 B per cell a Gram window can update (n_gram_cells) + 4 B per member of the Gram colours", "peak_source": peak_src,
               "note": "ALGORITHMIC bytes (SURVEY 8d: a u64 cell read and written per update, members as u32).  With u32 cells and the "
                       "bitmap form of the blob the launches move less than half of that (`traffic`), so this fraction can exceed 1."}
        hbm["frac"] = hbm["achieved"] / peak
        binding, other = (tensor, hbm) if tensor["frac"] >= hbm["frac"] else (hbm, tensor)
        entry = {"kernel": "k4_gram (+ kb_build)" if (gram_form & 1) == 0 else "k4_gram (+ kb_weights; the bitmaps are part of the resident blob)", "kernel_ms": g_ms, "traffic": tr, "traffic_source": trs,
                 "updates_per_s": U_gram * own_frac / (g_ms * 1e-3), "colors": C_gram, "planes": P_gram,
                 "binding": "the larger of the two fractions: the bound the launches are closer to", "other_bound": other}
        entry.update(binding)
        kernels.append(entry)
    for k in kernels:
        k["frac"] = k["achieved"] / k["peak"] if k["peak"] else None
    kernels.sort(key=lambda k: -k["kernel_ms"])
    roofline = dict(kernels[0]) if kernels else {"bound": "hbm", "kernel": "k3 spill (cub sort + reduce)", "achieved": 16 * U * own_frac / (max(phase["accumulate_ms"], 1e-9) * 1e-3) / 1e9,
                                                 "peak": peak, "unit": "GB/s", "frac": 16 * U * own_frac / (max(phase["accumulate_ms"], 1e-9) * 1e-3) / 1e9 / peak, "traffic": None,
                                                 "kernel_ms": phase["accumulate_ms"], "peak_source": peak_src}
    if len(kernels) > 1:
        roofline["others"] = kernels[1:]
    # in-run red.global.add.u64 ceiling on a buffer of the owned block's size
    red = None
    try:
        cells = max(owned_cells, 1024)
        n_ops = 1 << 28
        red = {}
        for pat, nm in ((0, "random"), (1, "warp_contiguous")):
            red[nm + "_ops_per_s"] = n_ops / (bench_red(cells, n_ops, 64, pat, 3) * 1e-3)
        red["cells"] = cells
        for k in kernels:
            if k["kernel"] in ("k1_small", "k2_mid"):
                red[k["kernel"] + "_frac_of_random"] = k["updates_per_s"] / red["random_ops_per_s"]
                red[k["kernel"] + "_frac_of_warp_contiguous"] = k["updates_per_s"] / red["warp_contiguous_ops_per_s"]
    except Exception as e:  # the micro-benchmark is informative only
        red = {"error": str(e)}

    cpu = None
    wall = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            cbs = cpu_baseline_on_sample(args, also_dropin=not args.no_wall)
            cpu = cbs["cpu_baseline"]
            wall = {"sample": {"what": cbs["sample"], "updates": cbs["updates"], "index_file_bytes": cbs["index_file_bytes"],
                               "reference": cbs["reference"], "dropin": cbs["dropin"],
                               "wall_ratio": (cbs["reference"]["wall_s"] / cbs["dropin"]["wall_s"]) if cbs["dropin"] else None}}
        except Exception as e:
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"unavailable: {e}"}
    if world == 1 and not args.no_wall:
        # pairwise wall-time of the FULL config through the drop-in binary: index files -> both TSVs (+ side-car)
        td = tempfile.mkdtemp(prefix="ksb_wall_")
        try:
            from kspider_b200 import synth
            prefix = os.path.join(td, s.name)
            synth.write_index_files(s, prefix, with_names=False)
            fbytes = index_file_bytes(prefix)
            d = run_dropin_binary(prefix, host_threads())
            d["index_file_bytes"] = fbytes
            d["loader_GB_per_s"] = fbytes / 1e9 / d["load_s"] if d.get("load_s") else None
            d["tsv_rows_per_s"] = rows / d["tsv_s"] if d.get("tsv_s") else None
            wall = dict(wall or {})
            wall["dropin_full"] = d
        except Exception as e:
            wall = dict(wall or {})
            wall["dropin_full"] = {"error": str(e)}
        finally:
            shutil.rmtree(td, ignore_errors=True)

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": f"u{dev.cell_bits}", "data": "synthetic",
            "config": {"workload": WORKLOAD_DESC[args.workload], "name": args.workload, "samples": N, "colors": C_,
                       "members": M, "pair_updates": U, "rows": rows, "edges": n_edges, "cutoff": args.cutoff,
                       "clusters": int(np.unique(labels_step).size), "step": "cluster-only (ksb_device_cluster), one CUDA graph launch per step" if not spill else "spill: sort-reduce records + K6/K7",
                       "parallelism": f"row-blocks x{world}", "passes": dev.n_passes, "spill": spill,
                       "index_build_s": t_build,
                       "blob": {"bytes": payload_bytes, "gram_section": "bitmaps" if (gram_form & 1) else ("member lists" + (f", {gram_form >> 8}-bit offsets inside the window" if gram_form >> 8 else ""))},
                       "cell_bits": dev.cell_bits,
                       "l2": f"inputs larger than L2 (blob {payload_bytes / 1e6:.0f} MB streamed every step; the owned block of the triangle ({owned_cells * dev.cell_bits / 8 / 1e6:.0f} MB of u{dev.cell_bits} cells) is cleared every step)"},
            "phases_ms": phase, "phases_note": "per-kernel timers of extra legs run kernel by kernel (the timed steps are graph replays)", "kernels_alone_ms": alone, "host_ms_per_step": host_step_ms, "wall_ms_per_step": wall_ms / args.steps,
            "timer": "torch.cuda.Event pair on the current stream around the K steps (every step ends with a stream synchronisation); phases: library CUDA events on the launching stream",
            "clocks": clocks, "e2e": e2e, "e2e_records": e2e_rec, "check": check_blk, "wall": wall,
            "gpu_launches": int(round(per_step_launches * args.steps)),
            "roofline": roofline, "int8_peak": imma, "red_roofline": red, "cpu_baseline": cpu}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        main_reference(a)
    else:
        main_b200(a)

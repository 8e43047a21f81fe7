#!/usr/bin/env python
"""bench.py -- PROCLU (+flank check) throughput of libvsk on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload proclu|flank|sweep]

Default workload `proclu` (BASELINE.json configs[1]): synthetic 1 M read-TR profiles vs 250 K
hg38-shaped reference TRs per GPU.  One step = one pass of the whole per-read pipeline (seeds ->
dedup -> LCS -> profile DP -> flank edit distances -> links) over the batch.  N > 1: the read TRs
are sharded (each rank scans its own reads against the replicated reference index, weak scaling)
and the per-rank link tables are merged with one NCCL all-gather of device buffers, as
join_clusters would merge the per-file .clu outputs.

Prints ONE JSON line (rank 0).  `value` = read-TRs/s with the reads resident in HBM; `e2e` = the
same through vsk_proclu_run with host buffers (H2D of the reads and D2H of the links inside the
timed region).  `roofline` is the binding one (integer ALU/issue) for the dominant kernel
(vsk_dp_dir_kernel, timed by itself); the HBM figures sit beside it as evidence of non-boundness.
`parity_at_scale`: the first reads of the very workload, against all references, through the
unmodified reference psearch.exe and through libvsk -- counters and canonical .clu must agree
(the process exits non-zero if they do not).

`--workload flank` = BASELINE.json configs[4] (flank-alignment-only batch through
vsk_flankalign_run), `--workload sweep` = configs[3] (pattern-size sweep of the profile DP).
`--impl reference` times the unmodified reference binaries (oracle/_ref) on the host cores.
"""
import argparse
import ctypes as C
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

# bytes: dram read+write of one vsk_dp_dir_kernel launch at the default workload (ncu --set full,
# profiles/): refreshed whenever the kernel or the batch size changes
DIR_TRAFFIC = {"bytes": 11.74e9, "source": "profiles/r02_summary.md section F (launch 0: a 12 GiB batch of the first chunk)"}
METRIC = "proclu_read_TRs_mapped_per_sec"
UNIT = "read-TRs/s"
OPS_PER_CELL = 14        # algorithmic integer operations per profile-DP cell
ISSUE_LIMIT_PER_CLK_SM = 128  # thread-instructions per clock per SM (4 schedulers x 32 lanes)


def load_table():
    return np.load(os.path.join(ROOT, "tests", "golden", "dist_table.npz"))["sm"]


def write_dst(path):
    with open(path, "wb") as f:
        f.write(np.array([777, 10, 5], np.int32).tobytes())
        f.write(load_table().tobytes())


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


def host_info():
    model = "unknown"
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    return {"host_cores": os.cpu_count() or 1, "cpu_model": model, "mem_available_gb": round(available_mem_gb(), 1)}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (the profiling recipe's line: ONE process
    that samples every GPU of the box every 200 ms).  It is started before the warm-up steps -- nvidia-smi takes
    a second or two to attach to eight GPUs, and eight of them starting inside a one-second timed region slowed
    the kernels they were meant to observe -- and only the samples stamped inside [mark(), stop()] count.
    Per sample and GPU it asks for two values only (SM clock, the bit mask of active clock-event reasons): on an
    eight-GPU box the recipe's nine fields five times a second cost the ranks 5 % of their step; the constant
    clocks.max.sm is read once before, the power draw once after."""
    Q = "timestamp,index,clocks.sm,clocks_event_reasons.active"
    BITS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
            0x80: "hw_power_brake_slowdown"}

    @staticmethod
    def _once(fields):
        try:
            txt = subprocess.run(["nvidia-smi", "--query-gpu=index," + fields, "--format=csv,noheader,nounits"],
                                 stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=30).stdout
        except (OSError, subprocess.SubprocessError):
            return {}
        out = {}
        for line in txt.splitlines():
            f = [x.strip() for x in line.split(",")]
            try:
                out[int(f[0])] = float(f[1])
            except (ValueError, IndexError):
                pass
        return out

    def __init__(self, period_ms=200, enabled=True):
        self.p = None
        self.t_mark = None
        self.first = ""
        self.max_mhz = {}
        if period_ms <= 0 or not enabled:
            return
        self.max_mhz = self._once("clocks.max.sm")
        try:
            self.p = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-lms", str(int(period_ms))],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            import select
            ready, _, _ = select.select([self.p.stdout], [], [], 60.0)
            if ready:
                self.first = self.p.stdout.readline()  # attached: the first sample is out
            else:  # never produced a line: do without it rather than hang the bench
                self.p.kill()
                self.p = None
        except OSError:
            self.p = None

    def mark(self):
        """start of the timed region"""
        import datetime
        self.t_mark = datetime.datetime.now()

    def stop(self, gpus=None):
        """gpus: nvidia-smi indices to report (None = all).  sm_mhz = the lowest per-GPU median under load."""
        import datetime
        t_end = datetime.datetime.now()
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if not self.p:
            return out
        self.p.terminate()
        try:
            txt = self.first + self.p.communicate(timeout=5)[0]
        except Exception:
            return out
        per, reasons = {}, set()
        for line in txt.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 4:
                continue
            try:
                ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f")
                gi = int(f[1])
                if self.t_mark is not None and not (self.t_mark <= ts <= t_end):
                    continue
                if gpus is not None and gi not in gpus:
                    continue
                per.setdefault(gi, []).append(float(f[2]))
                mask = int(f[3], 16)
            except ValueError:
                continue
            for bit, name in self.BITS.items():
                if mask & bit:
                    reasons.add(name)
        if per:
            med = {g: float(np.median(x)) for g, x in per.items()}
            out["sm_mhz"] = min(med.values())
            mx = [v for g, v in self.max_mhz.items() if g in med]
            out["sm_max_mhz"] = float(max(mx)) if mx else None
            out["samples"] = int(sum(len(x) for x in per.values()))
            if len(med) > 1:
                out["per_gpu_sm_mhz"] = [med[g] for g in sorted(med)]
        out["reasons"] = sorted(reasons)
        return out


def visible_gpus(world):
    """nvidia-smi indices of the GPUs this job runs on"""
    cvd = os.environ.get("CUDA_VISIBLE_DEVICES")
    if cvd:
        try:
            ids = [int(x) for x in cvd.split(",") if x.strip() != ""]
            return set(ids[:world]) if world > 1 else set(ids[:1])
        except ValueError:
            return None
    return set(range(world))


# --------------------------------------------------------------------------- reference arm
def available_mem_gb():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable"):
                return int(line.split()[1]) / 1e6
    except OSError:
        pass
    return 16.0


def run_reference_processes(workdir, dst, nproc, reads_files, n_marks):
    """Runs `nproc` reference psearch.exe processes (one per read file, the way
    run_proclu.pl fans out) and timestamps their stderr progress lines.  Returns per
    process: scan_start (start -> 'Scanning reads...'), the arrival time of every
    progress percent line, and the whole stderr text."""
    exe = os.path.join(ROOT, "oracle", "_ref", "psearch.exe")
    procs, results = [], [None] * nproc

    def watch(i, p, t0):
        marks, scan_t, buf, log = {}, None, "", []
        while True:
            ch = p.stderr.read(1)
            if not ch:
                break
            log.append(ch)
            if ch == "\n":
                now = time.perf_counter() - t0
                m = re.match(r"(\d+)% \(ProfileAlignments", buf)
                if m:
                    marks[int(m.group(1))] = now
                buf = ""
            else:
                buf += ch
                if buf == "Scanning reads...":
                    scan_t = time.perf_counter() - t0
        p.wait()
        results[i] = {"scan_start": scan_t, "marks": marks, "end": time.perf_counter() - t0, "rc": p.returncode,
                      "log": "".join(log)}

    t0 = time.perf_counter()
    threads = []
    for i in range(nproc):
        p = subprocess.Popen([exe, "reference.leb36", os.path.basename(reads_files[i]), dst, "88", "0", "50"],
                             cwd=workdir, stderr=subprocess.PIPE, stdout=subprocess.DEVNULL, text=True, bufsize=0)
        th = threading.Thread(target=watch, args=(i, p, t0))
        th.start()
        threads.append(th)
        procs.append(p)
    for th in threads:
        th.join()
    return results


def reference_procs(args):
    """process count of the reference arm: NPROC = nproc as BASELINE.md section 3 has it, unless the
    box cannot hold that many copies of the reference's index (about 13.5 GB each at 250 K refs)"""
    cores = os.cpu_count() or 1
    per_proc_gb = 13.5 * max(args.nref, 1) / 250000.0 + 0.3
    by_mem = int(available_mem_gb() * 0.85 / per_proc_gb)
    want = cores if args.ref_procs <= 0 else args.ref_procs
    return max(1, min(want, by_mem)), cores, by_mem


def reference_arm(args, rank):
    """The unmodified reference psearch.exe on the host cores, same config and metric.
    Every process rebuilds the reference index (that is the reference's design); the
    scan phase of each is cut into W+K steps at its own stderr progress marks."""
    if rank != 0:
        return
    from vntrseek_b200 import synth
    W, K = args.warmup, args.steps
    nproc, cores, by_mem = reference_procs(args)
    per_step = args.ref_reads_per_step
    nreads = per_step * (W + K)
    work = tempfile.mkdtemp(prefix="vsk_ref_")
    dst = os.path.join(work, "eucledian.dst")
    write_dst(dst)
    refs = synth.fast_refs(args.nref, seed=1)
    refs.write_leb36(os.path.join(work, "reference.leb36"))
    reads = synth.fast_reads(refs, nreads * nproc, seed=2)
    files = []
    for i in range(nproc):
        f = os.path.join(work, "reads%d.leb36" % i)
        reads.slice(i * nreads, (i + 1) * nreads).write_leb36(f)
        files.append(f)
    res = run_reference_processes(work, dst, nproc, files, W + K)
    shutil.rmtree(work, ignore_errors=True)
    if any(r is None or r["rc"] != 0 for r in res):
        print(json.dumps({"impl": "reference", "unavailable": "reference psearch.exe failed"}))
        return
    # step boundaries: percent marks at k*100/(W+K)
    step_ms = []
    for k in range(W, W + K):
        lo = int(round(100.0 * k / (W + K)))
        hi = int(round(100.0 * (k + 1) / (W + K)))
        dt = 0.0
        for r in res:
            a = r["scan_start"] if lo == 0 else r["marks"].get(lo)
            b = r["marks"].get(hi, r["end"])
            dt = max(dt, b - a)
        step_ms.append(dt * 1e3)
    ms = float(np.mean(step_ms))
    value = nproc * per_step / (ms * 1e-3)
    index_s = float(np.mean([r["scan_start"] for r in res]))
    cpu = {"value": value, "unit": UNIT, "cores": nproc, "kind": "reference",
           "sample": "%d x oracle/_ref/psearch.exe (unmodified reference, one process per core as run_proclu.pl "
                     "does; %d cores on the box, memory allows %d processes), %d reads/step each, scan phase timed "
                     "between stderr progress marks" % (nproc, cores, by_mem, per_step)}
    cpu.update(host_info())
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": K, "warmup": W, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": "synthetic %d read-TR profiles per step per process vs %d hg38-shaped reference TRs"
                   % (per_step, args.nref), "refs": args.nref, "reads_per_step": nproc * per_step,
                   "note": "scan phase only; every process first rebuilds the reference index (index_build_s)"},
        "index_build_s": index_s,
        "cpu_baseline": cpu,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# --------------------------------------------------------------------------- our arm
def cpu_baseline_single(args, refs, reads, out):
    """one reference process on the full reference set and a small read sample;
    scan phase only (the 1-core baseline reported beside the GPU number).  Its .clu and
    counters are kept: they are the reference side of `parity_at_scale`."""
    try:
        from vntrseek_b200.clu import canonical_clu, parse_counters
        exe = os.path.join(ROOT, "oracle", "_ref", "psearch.exe")
        if not os.path.exists(exe):
            out["err"] = "oracle/_ref/psearch.exe missing"
            return
        work = tempfile.mkdtemp(prefix="vsk_cpu_")
        dst = os.path.join(work, "eucledian.dst")
        write_dst(dst)
        refs.write_leb36(os.path.join(work, "reference.leb36"))
        n = args.cpu_sample_reads
        f = os.path.join(work, "reads0.leb36")
        reads.slice(0, n).write_leb36(f)
        r = run_reference_processes(work, dst, 1, [f], 1)[0]
        if r["rc"] != 0:
            out["err"] = "psearch.exe rc=%d" % r["rc"]
            return
        scan = r["marks"].get(100, r["end"]) - r["scan_start"]
        out["_ref_counters"] = parse_counters(r["log"])
        out["_ref_clu"] = canonical_clu(f + ".clu")
        shutil.rmtree(work, ignore_errors=True)
        out.update({"value": n / scan, "unit": UNIT, "cores": 1, "kind": "reference",
                    "sample": "oracle/_ref/psearch.exe (unmodified reference, 1 process): first %d reads of the "
                              "workload vs all %d refs; scan phase %.1f s (index build %.1f s not counted)"
                              % (n, args.nref, scan, r["scan_start"])})
        out.update(host_info())
    except Exception as e:  # noqa: BLE001
        out["err"] = repr(e)


def parity_at_scale(args, v, ctx, idx, reads, cpu_out):
    """the same first reads of the workload through libvsk against the same reference index:
    six counters + canonical .clu against what the reference binary produced"""
    from vntrseek_b200.clu import canonical_clu, stats_as_counters
    n = args.cpu_sample_reads
    ts = v.TRSetHandle.from_packed(reads.slice(0, n), is_read=True, maxflank=50)
    links, st = v.proclu(ctx, idx, ts)
    work = tempfile.mkdtemp(prefix="vsk_par_")
    clu = os.path.join(work, "ours.clu")
    trcount = v.write_clusters(links, clu)
    ours_c = stats_as_counters(st, trcount)
    ours_clu = canonical_clu(clu)
    shutil.rmtree(work, ignore_errors=True)
    del links
    ts.close()
    ref_c, ref_clu = cpu_out.pop("_ref_counters"), cpu_out.pop("_ref_clu")
    return {"reads": n, "refs": args.nref, "counters_equal": ours_c == ref_c, "clu_equal": ours_clu == ref_clu,
            "clusters": len(ref_clu), "counters": ours_c,
            "reference_counters": None if ours_c == ref_c else ref_c,
            "checker": "oracle/_ref/psearch.exe (unmodified reference) on the same files"}


def ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    import vntrseek_b200 as v
    from vntrseek_b200 import synth
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    W, K = args.warmup, args.steps
    t = time.perf_counter()
    refs = synth.fast_refs(args.nref, seed=1)
    reads = synth.fast_reads(refs, args.nreads, seed=2 + rank, key0=1 + rank * args.nreads)
    gen_s = time.perf_counter() - t
    cpu_out = {}
    cpu_thread = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_thread = threading.Thread(target=cpu_baseline_single, args=(args, refs, reads, cpu_out))
        cpu_thread.start()
    ctx = v.Context(load_table(), local_rank)
    t = time.perf_counter()
    refs_ts = v.TRSetHandle.from_packed(refs, is_read=False)
    reads_ts = v.TRSetHandle.from_packed(reads, is_read=True, maxflank=50)
    reads_ts.pin()  # e2e leg copies the step's inputs from pinned host memory
    pack_s = time.perf_counter() - t
    t = time.perf_counter()
    idx = v.RefIndex(ctx, refs_ts)
    index_s = time.perf_counter() - t
    resident = v.ResidentReads(ctx, reads_ts)
    parity = None
    if cpu_thread:
        cpu_thread.join()  # keep the host quiet during the timed regions
        if "_ref_counters" in cpu_out:
            parity = parity_at_scale(args, v, ctx, idx, reads, cpu_out)

    def barrier():
        if world > 1:
            if pending[0] is not None:  # the merged table of the last step is on the host before the clock stops
                pending[0].wait()
                pending[0] = None
            dist.barrier()
        torch.cuda.synchronize()
        ctx.sync()

    gatherer = None
    if world > 1:
        from vntrseek_b200.dist import LinkGatherer
        gatherer = LinkGatherer("cuda:%d" % local_rank, root=0)
    ag_ev = [torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)] if world > 1 else None
    ag_ms = [0.0]
    pending = [None]

    def step(reads_arg, timed):
        """one pass; N > 1: the links stay in HBM and go straight into the all-gather, the root
        alone copies the merged table to the host (it would write the cluster file)"""
        if world == 1:
            links, st = v.proclu(ctx, idx, reads_arg)
            return st
        dl, st = v.proclu(ctx, idx, reads_arg, links_on_device=True)
        if timed:
            ag_ev[0].record()
        if pending[0] is not None:
            pending[0].wait()           # the root's copy of the previous step's table (it overlapped this scan)
        pending[0] = gatherer.gather(dl.tensor())
        if timed:
            ag_ev[1].record()
            ag_ev[1].synchronize()
            ag_ms[0] += ag_ev[0].elapsed_time(ag_ev[1])
        return st

    # ---- device-resident leg ------------------------------------------------
    sampler = ClockSampler(args.clock_ms, enabled=(rank == 0))  # attached before the warm-up, see the class
    # untimed pass with exact cell accounting (the same batch is scanned every step)
    _, stats_cells = v.proclu(ctx, idx, resident, count_cells=True)
    stats = None
    for _ in range(W):
        stats = step(resident, False)
    barrier()
    sampler.mark()
    ctx.profile(True)
    l0 = ctx.launches
    ctx.timer_start()
    t0 = time.perf_counter()
    for _ in range(K):
        stats = step(resident, True)
    barrier()
    dev_ms = ctx.timer_stop()
    wall_ms = (time.perf_counter() - t0) * 1e3
    launches = ctx.launches - l0
    dp_ms, dp_launches = ctx.profile_read()
    sweep_ms, sweep_launches, sweep_busy_ms = ctx.profile_read_sweep()
    stages = ctx.profile_stages()
    ctx.profile(False)
    clocks = sampler.stop(visible_gpus(world))
    ag_resident_ms = ag_ms[0] / K
    ag_ms[0] = 0.0
    ag_phases = {k: val / (K + W) for k, val in gatherer.phase_ms.items()} if gatherer else {}
    # ---- end-to-end leg (host buffers in, links out) ----------------------------
    st2 = None
    for _ in range(min(W, 1)):
        st2 = step(reads_ts, False)
    barrier()
    ctx.profile(True)
    ctx.timer_start()
    t0 = time.perf_counter()
    for _ in range(K):
        st2 = step(reads_ts, True)
    barrier()
    e2e_ms = max(ctx.timer_stop(), (time.perf_counter() - t0) * 1e3)
    e2e_stages = ctx.profile_stages()
    ctx.profile(False)
    int_peak = ctx.int_peak_gops()

    def allmax(x):
        if world == 1:
            return x
        tt = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    def allsum(x):
        if world == 1:
            return x
        tt = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.SUM)
        return float(tt.item())

    def allgather_floats(x):
        if world == 1:
            return [x]
        tt = torch.tensor([x], device="cuda", dtype=torch.float64)
        outt = torch.zeros(world, device="cuda", dtype=torch.float64)
        dist.all_gather_into_tensor(outt, tt)
        return [float(v) for v in outt.tolist()]

    rank_step_ms = allgather_floats(max(dev_ms, wall_ms) / K)
    rank_ag_ms = allgather_floats(ag_resident_ms)
    rank_scan_ms = allgather_floats(sum(val for k, val in stages.items() if not k.startswith("host_")) / K)
    rank_cells = allgather_floats(float(stats_cells["dp_cells"]))
    rank_dp_ms = allgather_floats(dp_ms / K)
    step_ms = allmax(max(dev_ms, wall_ms)) / K
    e2e_step_ms = allmax(e2e_ms) / K
    total_reads = allsum(float(len(reads)))
    stats["dp_cells"] = stats_cells["dp_cells"]
    stats["dp_cells_sweep"] = stats_cells["dp_cells_sweep"]
    cells = allsum(float(stats["dp_cells"]))
    ag_resident_ms = allmax(ag_resident_ms)
    d2h_links = st2["d2h_bytes"]
    if world > 1:  # the root's copy of the merged table
        d2h_links = int(28 * allsum(float(st2["confirmed"])))
    ok = True
    if rank == 0:
        peaks, peak_src = measured_peaks()
        # the dominant kernel by itself: vsk_dp_dir_kernel, events around each of its launches
        # (its launches overlap across the two side streams, so the average duration of a launch is its share
        # of the time during which the kernel was running, not its own start-to-end time)
        sw_avg_s = (sweep_busy_ms / max(sweep_launches, 1)) * 1e-3
        sw_per_step = max(sweep_launches / K, 1)
        sw_cells_per_launch = stats["dp_cells_sweep"] / sw_per_step
        sw_bytes_per_launch = stats["dp_alg_bytes_sweep"] / sw_per_step
        sw_gcups = sw_cells_per_launch / sw_avg_s / 1e9 if sw_avg_s > 0 else 0.0
        sw_ops = OPS_PER_CELL * sw_gcups  # Gop/s
        sw_gbs = sw_bytes_per_launch / sw_avg_s / 1e9 if sw_avg_s > 0 else 0.0
        sm_mhz = clocks.get("sm_mhz") or 1965.0
        issue_limit = ISSUE_LIMIT_PER_CLK_SM * 148 * sm_mhz * 1e6 / 1e9  # G thread-instructions/s
        stage_s = dp_ms / K * 1e-3
        stage_gcups = stats["dp_cells"] / stage_s / 1e9 if stage_s > 0 else 0.0
        default_wl = args.nreads == 1000000 and args.nref == 250000
        line = {
            "metric": METRIC, "value": total_reads / (step_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int32", "data": "synthetic",
            "config": {"workload": "synthetic %d read-TR profiles per GPU vs %d hg38-shaped reference TRs "
                                   "(BASELINE.json configs[1])" % (args.nreads, args.nref),
                       "refs": args.nref, "reads_per_gpu": args.nreads, "cutoff": 88, "maxerrors": 0, "maxflank": 50,
                       "l2": "inputs larger than L2 (read set %.0f MB, index resident in HBM)"
                             % ((reads.ids.nbytes + reads.flanks.nbytes) / 1e6),
                       "parallelism": "reads sharded over %d GPU(s), reference index replicated, link tables merged "
                                      "by one all-gather of device buffers" % world},
            "gcups": cells / (step_ms * 1e-3) / 1e9,
            "dp_cells_per_step": cells,
            "counters_rank0": {k: stats[k] for k in ("candidates_considered", "lcs_alignments", "profile_alignments",
                                                      "profile_alignments_success", "confirmed", "chunks")},
            "stage_ms_per_step": {k: val / K for k, val in stages.items() if not k.startswith("host_")},
            "host_ms_per_step": {k[5:]: val / K for k, val in stages.items() if k.startswith("host_")},
            "index_build_s": index_s, "datagen_s": gen_s, "pack_s": pack_s,
            "e2e": {"value": total_reads / (e2e_step_ms * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": st2["h2d_bytes"], "d2h_bytes_per_step": d2h_links,
                    "ms_per_step": e2e_step_ms,
                    "stage_ms_per_step": {k: val / K for k, val in e2e_stages.items() if not k.startswith("host_")}},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {
                "bound": "int_alu", "kernel": "vsk_dp_dir_kernel",
                "achieved": sw_ops, "peak": int_peak, "unit": "Gop/s",
                "frac": sw_ops / int_peak if int_peak else None,
                "ops_per_cell": OPS_PER_CELL,
                "kernel_gcups": sw_gcups,
                "kernel_ms_per_step": sweep_busy_ms / K, "kernel_launches_per_step": sweep_launches / K,
                "kernel_avg_launch_ms": sw_avg_s * 1e3,
                "kernel_sum_of_launch_durations_ms_per_step": sweep_ms / K,
                "kernel_share_of_step": (sweep_busy_ms / K) / step_ms,
                "cells_per_launch": sw_cells_per_launch,
                "timing": "CUDA events around every launch of this kernel on the side stream it runs on; batches "
                          "alternate between two streams and overlap, so kernel_ms_per_step is the time during which "
                          "at least one launch was running (union of the event intervals) and the average launch "
                          "duration is that time / launches",
                "peak_source": "vsk_int_peak microkernel on this GPU: chains of VIADDMNMX (add+min in one DPX "
                               "instruction, ALU pipe) + IMAD (FMA pipe) = 3 integer operations per 2 issued "
                               "instructions (profiles/r02_int_peak_sass.txt)",
                "frac_of_issue_limit": sw_ops / issue_limit,
                "issue_limit_gops": issue_limit,
                "issue_limit_note": "128 thread-instructions per clock per SM x 148 SMs x the SM clock sampled under load",
                "hbm": {"achieved": sw_gbs, "peak": peaks.get("hbm_gbs"), "unit": "GB/s",
                        "frac": sw_gbs / peaks.get("hbm_gbs", 1.0), "alg_bytes_per_launch": sw_bytes_per_launch,
                        "peak_source": peak_src,
                        "note": "2*(L1+L2) id bytes + 16 result bytes per pair (SURVEY 8d): evidence that HBM does not "
                                "bound the kernel"},
                "traffic": DIR_TRAFFIC["bytes"] if default_wl else None,
                "traffic_note": "dram__bytes_read+write of one launch, ncu --set full, " + DIR_TRAFFIC["source"] +
                                "; the direction words (0.25 B per cell) are all of it",
                "dp_stage": {"ms_per_step": dp_ms / K, "gcups": stage_gcups,
                             "achieved": OPS_PER_CELL * stage_gcups,
                             "frac": OPS_PER_CELL * stage_gcups / int_peak if int_peak else None,
                             "kernels": "vsk_dp_dir_kernel + vsk_dp_q8dir_kernel + vsk_dp_tinydir_kernel + "
                                        "vsk_dp_trace_kernel (+ vsk_dp_kernel for handed-back pairs), two side streams"},
            },
        }
        if world > 1:
            line["allgather_ms_per_step"] = ag_resident_ms
            line["allgather_note"] = ("CUDA events around the merge on every rank, max over ranks: it includes the time a "
                                      "rank waits for the slowest rank's scan to finish (per_rank below)")
            line["per_rank"] = {"ms_per_step": rank_step_ms, "scan_ms_per_step": rank_scan_ms,
                                "allgather_ms_per_step": rank_ag_ms, "dp_cells_per_step": rank_cells,
                                "dp_stage_ms_per_step": rank_dp_ms,
                                "rank0_merge_phases_host_ms": ag_phases}
        if parity is not None:
            line["parity_at_scale"] = parity
            ok = parity["counters_equal"] and parity["clu_equal"]
        if cpu_out:
            line["cpu_baseline"] = cpu_out
        elif world == 1:
            line["cpu_baseline"] = {"err": "skipped"}
        print(json.dumps(line))
    resident.close()
    idx.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    if not ok:
        sys.stderr.write("bench.py: parity_at_scale FAILED (libvsk and the reference binary disagree)\n")
        sys.exit(3)


# --------------------------------------------------------------------------- flank workload (configs[4])
def write_flank_workload(path, nblocks, nref, nread, seed=5, first_block=1):
    """`.flanks` blocks of the shape run_flankcomp.pl writes (perl/run_flankcomp.pl:229-268): `nref`
    references and `nread` reads per block, every read inside the +-15 % pattern-size window of every
    reference of its block, read flanks = 10-50 nt of a reference's 60-nt flanks with 3 % substitutions.
    Returns the number of read x reference pairs."""
    rng = np.random.default_rng(seed)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    with open(path, "w") as f:
        for b in range(nblocks):
            L = int(rng.integers(10, 60))
            pat = acgt[rng.integers(0, 4, L)].tobytes().decode()
            fl = acgt[rng.integers(0, 4, (nref, 120))]
            f.write("@(%d_1):\n" % (first_block + b))
            f.write("*" * 70 + "\n")
            for r in range(nref):
                s = fl[r].tobytes().decode()
                f.write("-%d',%s,%s,%s,%s\n" % (1000 * (first_block + b) + r + 1, s[:60], pat * 2, s[60:], pat))
            src = rng.integers(0, nref, nread)
            ll, rl = rng.integers(10, 51, nread), rng.integers(10, 51, nread)
            rd = fl[src].copy()
            m = rng.random(rd.shape) < 0.03
            rd[m] = acgt[rng.integers(0, 4, int(m.sum()))]
            rep = pat * 2
            base = (first_block + b) * 1000000
            for q in range(nread):
                row = rd[q].tobytes().decode()
                left, right = row[60 - ll[q]:60], row[60:60 + rl[q]]
                f.write("%d',%d,%d,%s%s%s,%s\n" % (base + q, ll[q] + 1, ll[q] + len(rep), left, rep, right, pat))
    return nblocks * nref * nread


def flank_reference_time(flanks, nproc):
    """oracle/_ref/flankalign.exe OUT RES FLANKS 0 50 NPROC 15 (vntrseek.pl.in:956-962): wall seconds"""
    exe = os.path.join(ROOT, "oracle", "_ref", "flankalign.exe")
    work = tempfile.mkdtemp(prefix="vsk_fref_")
    os.makedirs(os.path.join(work, "res"))
    t = time.perf_counter()
    p = subprocess.run([exe, os.path.join(work, "out"), os.path.join(work, "res"), flanks, "0", "50", str(nproc), "15"],
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    dt = time.perf_counter() - t
    nmaps = len(os.listdir(os.path.join(work, "out"))) if os.path.isdir(os.path.join(work, "out")) else 0
    shutil.rmtree(work, ignore_errors=True)
    return dt, p.returncode, nmaps


class FlankStats(C.Structure):
    _fields_ = [("blocks", C.c_int64), ("pairs", C.c_int64), ("accepted", C.c_int64), ("h2d_bytes", C.c_int64),
                ("d2h_bytes", C.c_int64), ("kernel_ms", C.c_double)]


def flank_workload(args, rank, world, local_rank):
    """configs[4]: flank-alignment-only batch.  A step = vsk_flankalign_run over the whole `.flanks`
    file (parse, pair enumeration, H2D, 4 edit distances per pair on the GPU, D2H, `.map` files):
    that is the end-to-end number; `value` = the pairs over the kernel time alone (CUDA events)."""
    if rank != 0:
        return
    import vntrseek_b200 as v
    from vntrseek_b200._lib import lib
    W, K = args.warmup, args.steps
    nref, nread = 50, 4000
    nblocks = max(1, args.flank_pairs // (nref * nread))
    work = tempfile.mkdtemp(prefix="vsk_flank_")
    flanks = os.path.join(work, "all.flanks")
    t = time.perf_counter()
    npairs = write_flank_workload(flanks, nblocks, nref, nread)
    gen_s = time.perf_counter() - t
    cores = os.cpu_count() or 1
    sub_blocks = max(1, min(nblocks, args.flank_cpu_pairs // (nref * nread)))
    if args.impl == "reference":
        sub = os.path.join(work, "sub.flanks")
        step_ms = []
        for k in range(W + K):
            write_flank_workload(sub, sub_blocks, nref, nread, seed=100 + k)
            dt, rc, nmaps = flank_reference_time(sub, cores)
            if rc != 0 or nmaps != sub_blocks:
                print(json.dumps({"impl": "reference", "unavailable": "reference flankalign.exe failed (rc %d)" % rc}))
                return
            if k >= W:
                step_ms.append(dt * 1e3)
        ms = float(np.mean(step_ms))
        val = sub_blocks * nref * nread / (ms * 1e-3)
        cpu = {"value": val, "unit": "flank pairs/s", "cores": cores, "kind": "reference",
               "sample": "oracle/_ref/flankalign.exe OUT RES FLANKS 0 50 %d 15 on %d blocks (%d pairs) per step, wall "
                         "time of the whole process (fork per block, %d at a time)" % (cores, sub_blocks,
                                                                                      sub_blocks * nref * nread, cores)}
        cpu.update(host_info())
        print(json.dumps({"impl": "reference", "metric": "flankalign_pairs_per_sec", "value": val,
                          "unit": "flank pairs/s", "n_gpus": args.gpus, "steps": K, "warmup": W, "ms_per_step": ms,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64 bit-vectors",
                          "data": "synthetic",
                          "config": {"workload": "flank-alignment-only batch (BASELINE.json configs[4]), %d pairs per "
                                                 "step" % (sub_blocks * nref * nread)},
                          "cpu_baseline": cpu,
                          "e2e": {"value": val, "unit": "flank pairs/s", "h2d_bytes_per_step": 0,
                                  "d2h_bytes_per_step": 0}, "gpu_launches": 0}))
        shutil.rmtree(work, ignore_errors=True)
        return
    # CPU baseline on a bounded subset, before the GPU legs (host quiet afterwards)
    cpu = {"err": "skipped"}
    if not args.no_cpu_baseline and os.path.exists(os.path.join(ROOT, "oracle", "_ref", "flankalign.exe")):
        sub = os.path.join(work, "sub.flanks")
        write_flank_workload(sub, sub_blocks, nref, nread)
        dt, rc, nmaps = flank_reference_time(sub, cores)
        cpu = {"value": sub_blocks * nref * nread / dt, "unit": "flank pairs/s", "cores": cores, "kind": "reference",
               "sample": "oracle/_ref/flankalign.exe OUT RES FLANKS 0 50 %d 15 on the first %d blocks (%d pairs): %.1f s "
                         "wall, rc %d, %d .map files" % (cores, sub_blocks, sub_blocks * nref * nread, dt, rc, nmaps)}
        cpu.update(host_info())
    ctx = v.Context(load_table(), local_rank)
    L = lib()
    L.vsk_flankalign_run.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int64, C.POINTER(FlankStats)]
    outdir = os.path.join(work, "out")
    os.makedirs(outdir)
    cwd = os.getcwd()
    os.chdir(outdir)
    st = FlankStats()
    sampler = ClockSampler()
    e2e, kern = [], []
    l0 = 0
    try:
        for k in range(W + K):
            if k == W:
                sampler.mark()
                l0 = ctx.launches
            t = time.perf_counter()
            rc = L.vsk_flankalign_run(ctx.h, flanks.encode(), 0, 50, 15, 0, C.byref(st))
            dt = time.perf_counter() - t
            ctx._check(rc)
            if k >= W:
                e2e.append(dt * 1e3)
                kern.append(st.kernel_ms)
    finally:
        os.chdir(cwd)
    clocks = sampler.stop(visible_gpus(1))
    launches = ctx.launches - l0
    int_peak = ctx.int_peak_gops()
    ms_k, ms_e = float(np.mean(kern)), float(np.mean(e2e))
    # bit-vector work: one 64-bit word per row while the reference flank has at most 63 characters
    # (60 here), rows = read flank length, 4 edit distances per pair, 20 operations per word-row
    # (SURVEY 8d); read flanks are 10-50 nt (mean 30)
    word_rows = 4.0 * 30.0 * st.pairs
    ops = 20.0 * word_rows / (ms_k * 1e-3) / 1e9
    peaks, peak_src = measured_peaks()
    line = {"metric": "flankalign_pairs_per_sec", "value": st.pairs / (ms_k * 1e-3), "unit": "flank pairs/s",
            "n_gpus": 1, "steps": K, "warmup": W, "ms_per_step": ms_k, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64 bit-vectors", "data": "synthetic",
            "config": {"workload": "flank-alignment-only batch (BASELINE.json configs[4]): %d read-flank x reference-flank "
                                   "pairs in %d blocks of %d references x %d reads, PATLEN_SIZE_ERR 15, "
                                   "MAX_FLANK_CONSIDERED 50" % (st.pairs, st.blocks, nref, nread),
                       "l2": "inputs larger than L2 (pair lists %.0f MB per step)" % (8.0 * st.pairs / 1e6)},
            "accepted_pairs": int(st.accepted), "datagen_s": gen_s,
            "e2e": {"value": st.pairs / (ms_e * 1e-3), "unit": "flank pairs/s", "h2d_bytes_per_step": int(st.h2d_bytes),
                    "d2h_bytes_per_step": int(st.d2h_bytes), "ms_per_step": ms_e,
                    "note": "vsk_flankalign_run from the .flanks file: parsing, pair enumeration, copies, kernel, .map files"},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "int_alu", "kernel": "vsk_flankalign_kernel", "achieved": ops, "peak": int_peak,
                         "unit": "Gop/s", "frac": ops / int_peak if int_peak else None, "ops_per_word_row": 20,
                         "bit_gcups": 4.0 * 60 * 30 * st.pairs / (ms_k * 1e-3) / 1e9,
                         "peak_source": "vsk_int_peak microkernel (32-bit operations; the kernel's 64-bit word operations "
                                        "count once each, as SURVEY 8d counts them)",
                         "hbm": {"achieved": 12.0 * st.pairs / (ms_k * 1e-3) / 1e9,
                                 "peak": peaks.get("hbm_gbs"), "unit": "GB/s", "peak_source": peak_src,
                                 "note": "12 B per pair (two indices in, one result word out)"},
                         "traffic": None},
            "cpu_baseline": cpu}
    print(json.dumps(line))
    ctx.close()
    shutil.rmtree(work, ignore_errors=True)


# --------------------------------------------------------------------------- pattern-size sweep (configs[3])
SWEEP_SIZES = (7, 10, 17, 25, 30, 39, 59, 89, 134, 201, 303, 474, 715, 1111, 2000)  # psearch.c:286-315


def sweep_workload(args, rank, world, local_rank):
    """configs[3]: pairs of equal-size profiles at the TRANGE boundaries, full matrix and narrowband,
    through vsk_dp_batch (pair list resident in HBM).  A step = the whole sweep."""
    if rank != 0:
        return
    from vntrseek_b200 import synth
    rng = np.random.default_rng(0)
    pure, mix = synth._PURE, synth._MIX[synth._MIX > 0]

    def profiles(lens):
        tot = int(lens.sum())
        ids = pure[rng.integers(0, 4, tot)]
        m = rng.random(tot) < 0.15
        ids[m] = mix[rng.integers(0, len(mix), int(m.sum()))]
        off = np.zeros(len(lens) + 1, np.int64)
        np.cumsum(lens, out=off[1:])
        return off, ids.astype(np.uint16)

    W, K = args.warmup, args.steps
    cases = []
    for L in SWEEP_SIZES:
        n = int(max(2000, min(200000, 2e9 / (L * L))))
        o1, i1 = profiles(np.full(n, L, np.int64))
        i2 = i1.copy()
        mut = rng.random(len(i2)) < 0.03
        i2[mut] = pure[rng.integers(0, 4, int(mut.sum()))]
        cases.append((L, n, o1, i1, i2))
    if args.impl == "reference":
        from oracle.pyoracle import Reference, have_reference
        if not have_reference():
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref_harness.so not built"}))
            return
        ref = Reference()
        per, tot_cells, tot_s = [], 0.0, 0.0
        for L, n, o1, i1, i2 in cases:
            m = int(max(2, min(n, 4e7 / (L * L))))
            ar = np.arange(m, dtype=np.int32)
            t = time.perf_counter()
            ref.dp_batch(o1, i1, o1, i2, ar, ar, np.zeros(m, np.uint8), 88)
            dt = time.perf_counter() - t
            per.append({"pattern": L, "pairs": m, "gcups_full": m * L * L / dt / 1e9})
            tot_cells += m * L * L
            tot_s += dt
        val = tot_cells / tot_s / 1e9
        cpu = {"value": val, "unit": "GCUPS", "cores": 1, "kind": "reference",
               "sample": "GetFixedProfileAP of the unmodified reference (oracle/_ref/libref_harness.so), one thread, "
                         "about 4e7 cells per pattern size"}
        cpu.update(host_info())
        print(json.dumps({"impl": "reference", "metric": "profile_dp_gcups_sweep", "value": val, "unit": "GCUPS",
                          "n_gpus": args.gpus, "steps": 1, "warmup": 0, "ms_per_step": tot_s * 1e3,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
                          "data": "synthetic", "config": {"workload": "pattern-size sweep 7-2000 (BASELINE.json configs[3])"},
                          "sweep": per, "cpu_baseline": cpu,
                          "e2e": {"value": val, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                          "gpu_launches": 0}))
        return
    import vntrseek_b200 as v
    ctx = v.Context(load_table(), local_rank)
    int_peak = ctx.int_peak_gops()
    per = []
    tot_cells = tot_ms = tot_e2e_ms = 0.0
    h2d = d2h = 0
    sampler = ClockSampler()
    sampler.mark()
    l0 = ctx.launches
    for L, n, o1, i1, i2 in cases:
        pa, pb = ctx.profiles(off=o1, ids=i1), ctx.profiles(off=o1, ids=i2)
        idx = np.arange(n, dtype=np.int32)
        row = {"pattern": L, "pairs": n}
        for mode, name in ((0, "full"), (1, "narrowband")):
            mvec = np.full(n, mode, np.uint8)
            b = ctx.dp_batch(pa, pb, idx, idx, mvec, 88)
            for _ in range(W):
                b.run()
            ctx.sync()
            ctx.timer_start()
            for _ in range(K):
                b.run()
            dt = ctx.timer_stop() / K
            row["gcups_" + name] = b.cells / (dt * 1e-3) / 1e9
            row["frac_int_alu_" + name] = OPS_PER_CELL * row["gcups_" + name] / int_peak if int_peak else None
            if mode == 0:
                tot_cells += b.cells
                tot_ms += dt
            b.close()
            if mode == 0:   # end to end: host pair list in, results out (vsk_pair_dp)
                t = time.perf_counter()
                ctx.pair_dp(pa, pb, idx, idx, mvec, 88)
                tot_e2e_ms += (time.perf_counter() - t) * 1e3
                h2d += 9 * n
                d2h += 16 * n
        pa.close()
        pb.close()
        per.append(row)
    launches = ctx.launches - l0
    clocks = sampler.stop(visible_gpus(1))
    val = tot_cells / (tot_ms * 1e-3) / 1e9
    cpu = {"err": "skipped"}
    if not args.no_cpu_baseline:
        try:
            from oracle.pyoracle import Reference, have_reference
            if have_reference():
                ref = Reference()
                c, s = 0.0, 0.0
                for L, n, o1, i1, i2 in cases:
                    m = int(max(2, min(n, 2e7 / (L * L))))
                    ar = np.arange(m, dtype=np.int32)
                    t = time.perf_counter()
                    ref.dp_batch(o1, i1, o1, i2, ar, ar, np.zeros(m, np.uint8), 88)
                    s += time.perf_counter() - t
                    c += m * L * L
                cpu = {"value": c / s / 1e9, "unit": "GCUPS", "cores": 1, "kind": "reference",
                       "sample": "GetFixedProfileAP of the unmodified reference (oracle/_ref/libref_harness.so), one "
                                 "thread, about 2e7 cells per pattern size (%.1f s)" % s}
                cpu.update(host_info())
        except Exception as e:  # noqa: BLE001
            cpu = {"err": repr(e)}
    print(json.dumps({"metric": "profile_dp_gcups_sweep", "value": val, "unit": "GCUPS", "n_gpus": 1, "steps": K,
                      "warmup": W, "ms_per_step": tot_ms, "higher_is_better": True, "scaling": "weak",
                      "vs_baseline": None, "dtype": "int32", "data": "synthetic",
                      "config": {"workload": "pattern-size sweep 7-2000 at the TRANGE boundaries (BASELINE.json configs[3]): "
                                             "pairs of equal-size profiles (3 % substitutions), full matrix and narrowband, "
                                             "pair lists resident in HBM (vsk_dp_batch)",
                                 "l2": "each batch is re-run back to back; the id arrays of the large batches exceed L2 "
                                       "only for the short patterns (the kernels are integer-bound, not HBM-bound)"},
                      "sweep": per,
                      "e2e": {"value": tot_cells / (tot_e2e_ms * 1e-3) / 1e9, "unit": "GCUPS", "h2d_bytes_per_step": h2d,
                              "d2h_bytes_per_step": d2h, "note": "vsk_pair_dp: host pair lists in, results out, full matrix"},
                      "gpu_launches": int(launches), "clocks": clocks,
                      "roofline": {"bound": "int_alu", "kernel": "profile DP kernels by size class", "achieved":
                                   OPS_PER_CELL * val, "peak": int_peak, "unit": "Gop/s",
                                   "frac": OPS_PER_CELL * val / int_peak if int_peak else None, "traffic": None,
                                   "ops_per_cell": OPS_PER_CELL},
                      "cpu_baseline": cpu}))
    ctx.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--workload", default="proclu", choices=["proclu", "flank", "sweep"])
    ap.add_argument("--nref", type=int, default=250000)
    ap.add_argument("--nreads", type=int, default=1000000)
    ap.add_argument("--cpu-sample-reads", type=int, default=3000)
    ap.add_argument("--ref-reads-per-step", type=int, default=1500)
    ap.add_argument("--ref-procs", type=int, default=0, help="reference processes (0 = one per host core, memory permitting)")
    ap.add_argument("--flank-pairs", type=int, default=50000000)
    ap.add_argument("--flank-cpu-pairs", type=int, default=1000000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--clock-ms", type=int, default=200, help="nvidia-smi sampling period during the timed region (0 = off)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.workload == "flank":
        flank_workload(args, rank, world, local_rank)
        return
    if args.workload == "sweep":
        sweep_workload(args, rank, world, local_rank)
        return
    if args.impl == "reference":
        reference_arm(args, rank)
        return
    ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py -- PROCLU (+flank check) throughput of libvsk on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[1]): synthetic 1 M read-TR profiles vs 250 K
hg38-shaped reference TRs per GPU.  One step = one pass of the whole per-read
pipeline (seeds -> dedup -> LCS -> profile DP -> flank edit distances -> links)
over the batch.  N > 1: the read TRs are sharded (each rank scans its own 1 M
reads against the replicated reference index, weak scaling) and the per-rank
link tables are merged with one NCCL all-gather, as join_clusters would merge
the per-file .clu outputs.

Prints ONE JSON line (rank 0).  `value` = read-TRs/s with the reads resident in
HBM; `e2e` = the same through vsk_proclu_run with host buffers (H2D of the reads
and D2H of the links inside the timed region).  `--impl reference` times the
unmodified reference psearch.exe (oracle/_ref) on the host cores instead.
"""
import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

DIR_TRAFFIC = 11.01e9  # bytes: dram read+write of the first vsk_dp_dir_kernel launch, profiles/r01_summary.md (r01g)
METRIC = "proclu_read_TRs_mapped_per_sec"
UNIT = "read-TRs/s"


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


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, dev):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(dev), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "200"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            pass

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if not self.p:
            return out
        self.p.terminate()
        try:
            txt = self.p.communicate(timeout=5)[0]
        except Exception:
            return out
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in txt.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(max(mx))
        out["reasons"] = sorted(reasons)
        out["samples"] = len(sm)
        return out


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
    process: index_s (start -> 'Scanning reads...') and the arrival time of every
    progress percent line."""
    exe = os.path.join(ROOT, "oracle", "_ref", "psearch.exe")
    procs, results = [], [None] * nproc

    def watch(i, p, t0):
        marks, scan_t, buf = {}, None, ""
        while True:
            ch = p.stderr.read(1)
            if not ch:
                break
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
        results[i] = {"scan_start": scan_t, "marks": marks, "end": time.perf_counter() - t0, "rc": p.returncode}

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


def reference_arm(args, rank):
    """The unmodified reference psearch.exe on the host cores, same config and metric.
    Every process rebuilds the reference index (that is the reference's design); the
    scan phase of each is cut into W+K steps at its own stderr progress marks."""
    if rank != 0:
        return
    exe = os.path.join(ROOT, "oracle", "_ref", "psearch.exe")
    from vntrseek_b200 import synth
    W, K = args.warmup, args.steps
    cores = os.cpu_count() or 1
    nproc = max(1, min(cores, int(available_mem_gb() * 0.8 / 13.5), args.ref_procs))
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
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": K, "warmup": W, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": "synthetic %d read-TR profiles per step per process vs %d hg38-shaped reference TRs"
                   % (per_step, args.nref), "refs": args.nref, "reads_per_step": nproc * per_step,
                   "note": "scan phase only; every process first rebuilds the reference index (index_build_s)"},
        "index_build_s": index_s,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": nproc, "kind": "reference",
                         "sample": "%d x oracle/_ref/psearch.exe (unmodified reference), %d reads/step each, "
                                   "scan phase timed between stderr progress marks" % (nproc, per_step)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# --------------------------------------------------------------------------- our arm
def cpu_baseline_single(args, refs, reads, out):
    """one reference process on the full reference set and a small read sample;
    scan phase only (the 1-core baseline reported beside the GPU number)"""
    try:
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
        out.update({"value": n / scan, "unit": UNIT, "cores": 1, "kind": "reference",
                    "sample": "oracle/_ref/psearch.exe (unmodified reference, 1 process): first %d reads of the "
                              "workload vs all %d refs; scan phase %.1f s (index build %.1f s not counted)"
                              % (n, args.nref, scan, r["scan_start"])})
    except Exception as e:  # noqa: BLE001
        out["err"] = repr(e)


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
    if cpu_thread:
        cpu_thread.join()  # keep the host quiet during the timed regions

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.sync()

    from vntrseek_b200.dist import allgather_links

    def merge_links(links):
        """per-rank link tables -> every rank (one NCCL all-gather over NVLink)"""
        return allgather_links(links, device="cuda") if world > 1 else links

    # ---- device-resident leg ------------------------------------------------
    # untimed pass with exact cell accounting (the same batch is scanned every step)
    _, stats_cells = v.proclu(ctx, idx, resident, count_cells=True)
    stats = None
    for _ in range(W):
        links, stats = v.proclu(ctx, idx, resident)
        merge_links(links)
    barrier()
    sampler = ClockSampler(local_rank)
    ctx.profile(True)
    l0 = ctx.launches
    ctx.timer_start()
    t0 = time.perf_counter()
    for _ in range(K):
        links, stats = v.proclu(ctx, idx, resident)
        merge_links(links)
    barrier()
    dev_ms = ctx.timer_stop()
    wall_ms = (time.perf_counter() - t0) * 1e3
    launches = ctx.launches - l0
    dp_ms, dp_launches = ctx.profile_read()
    stages = ctx.profile_stages()
    ctx.profile(False)
    clocks = sampler.stop()
    # ---- end-to-end leg (host buffers in, links out) ----------------------------
    for _ in range(min(W, 1)):
        v.proclu(ctx, idx, reads_ts)
    barrier()
    ctx.timer_start()
    for _ in range(K):
        links, st2 = v.proclu(ctx, idx, reads_ts)
        merge_links(links)
    barrier()
    e2e_ms = ctx.timer_stop()
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

    step_ms = allmax(max(dev_ms, wall_ms)) / K
    e2e_step_ms = allmax(e2e_ms) / K
    total_reads = allsum(float(len(reads)))
    stats["dp_cells"] = stats_cells["dp_cells"]
    cells = allsum(float(stats["dp_cells"]))
    if rank == 0:
        peaks, peak_src = measured_peaks()
        dp_avg_s = (dp_ms / max(dp_launches, 1)) * 1e-3
        cells_per_launch = stats["dp_cells"] / max(dp_launches / K, 1)
        # algorithmic HBM bytes of the DP kernel per launch: 2*(L1+L2) id bytes + 16 result bytes per pair
        # (SURVEY 8d), summed by the pipeline over the job list it built
        alg_bytes_per_launch = stats["dp_alg_bytes"] / max(dp_launches / K, 1)
        achieved_gbs = alg_bytes_per_launch / dp_avg_s / 1e9 if dp_avg_s > 0 else 0.0
        int_ops = 14.0 * cells_per_launch / dp_avg_s / 1e9 if dp_avg_s > 0 else 0.0
        line = {
            "metric": METRIC, "value": total_reads / (step_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int32", "data": "synthetic",
            "config": {"workload": "synthetic %d read-TR profiles per GPU vs %d hg38-shaped reference TRs "
                                   "(BASELINE.json configs[1])" % (args.nreads, args.nref),
                       "refs": args.nref, "reads_per_gpu": args.nreads, "cutoff": 88, "maxerrors": 0, "maxflank": 50,
                       "l2": "inputs larger than L2 (read set %.0f MB, index resident in HBM)"
                             % ((reads.ids.nbytes + reads.flanks.nbytes) / 1e6),
                       "parallelism": "reads sharded over %d GPU(s), reference index replicated, links all-gathered" % world},
            "gcups": cells / (step_ms * 1e-3) / 1e9,
            "dp_cells_per_step": cells,
            "counters_rank0": {k: stats[k] for k in ("candidates_considered", "lcs_alignments", "profile_alignments",
                                                      "profile_alignments_success", "confirmed", "chunks")},
            "stage_ms_per_step": {k: val / K for k, val in stages.items()},
            "index_build_s": index_s, "datagen_s": gen_s, "pack_s": pack_s,
            "e2e": {"value": total_reads / (e2e_step_ms * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": st2["h2d_bytes"], "d2h_bytes_per_step": st2["d2h_bytes"],
                    "ms_per_step": e2e_step_ms},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved_gbs, "peak": peaks.get("hbm_gbs"), "unit": "GB/s",
                         "frac": achieved_gbs / peaks.get("hbm_gbs", 1.0),
                         "traffic": DIR_TRAFFIC if args.nreads == 1000000 and args.nref == 250000 else None,
                         "traffic_note": "dram__bytes_read+write of one vsk_dp_dir_kernel launch (first 12 GiB batch of the "
                                         "first chunk), ncu --set full, profiles/r01_summary.md; the direction words "
                                         "(0.25 B per cell) are all of it; a step has about 9 such launches, while "
                                         "`achieved` is over the whole DP stage of a chunk",
                         "kernel": "vsk_dp_dir_kernel (forward sweep of the profile DP; the DP stage timed here also holds "
                                   "vsk_dp_q8dir_kernel, vsk_dp_tiny_kernel and vsk_dp_trace_kernel)",
                         "kernel_ms_per_step": dp_ms / K,
                         "kernel_share_of_step": (dp_ms / K) / step_ms, "peak_source": peak_src,
                         "note": "integer DP: issue/ALU bound, HBM fraction is evidence of non-boundness; see int_alu"},
            "int_alu": {"achieved": int_ops, "peak": int_peak, "unit": "Gop/s", "frac": int_ops / int_peak if int_peak else None,
                        "ops_per_cell": 14, "kernel_gcups": cells_per_launch / dp_avg_s / 1e9 if dp_avg_s > 0 else 0.0,
                        "peak_source": "vsk_int_peak microkernel (IADD3/VIMNMX/IMAD chains) on this GPU"},
        }
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--nref", type=int, default=250000)
    ap.add_argument("--nreads", type=int, default=1000000)
    ap.add_argument("--cpu-sample-reads", type=int, default=3000)
    ap.add_argument("--ref-reads-per-step", type=int, default=1500)
    ap.add_argument("--ref-procs", type=int, default=16)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        reference_arm(args, rank)
        return
    ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py — k-mers inserted per second by the counter workload (Brisk<uint8_t>, insert-or-increment).

Workload (BASELINE.json configs[2], the synthetic configuration at the metric's k=63 m=21):
100 Mbp random genome, 250 bp error-free reads, k=63 m=21 b=14.  One step = one batch of 2^20 reads
pushed through the hot path (pack -> enumerate superkmers -> insert/count into the bucket store),
followed by the batched query of the previous step's batch (the config's "batched get queries
interleaved with inserts").  Successive steps take successive batches of the read set, so the store
grows as in the real job.  The metric counts the INSERTED k-mer occurrences only.

  python bench.py [--gpus N] [--steps K] [--warmup W]          the CUDA path, through the C ABI
  python bench.py --impl reference ...                         the reference's own CPU code
                                                               (oracle/_ref/ref_dump, all host threads,
                                                               a bounded sample of the same workload)
One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

K, M, B = 63, 21, 14
GENOME, READ_LEN, COVERAGE, SEED = 100_000_000, 250, 30, 3
BATCH_READS = 1 << 20
METRIC = "k-mers inserted/sec (counter, k=63 m=21)"
UNIT = "k-mer occurrences/s"
WORKLOAD = ("cfg3: synthetic 100 Mbp genome, 250 bp reads at 30x, k=63 m=21 b=14, uint8 counts; step = count a batch of "
            "2^20 reads, then batched query of the previous batch")
# algorithmic bytes of the insert/count kernel = 18 B per occurrence (16-B key + DATA read,
# DATA written) + 16 B per distinct k-mer (key claim)
INSERT_B_PER_OCC, INSERT_B_PER_NEW = 18.0, 16.0


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def read_batch(gen, batch_index, n_reads=BATCH_READS):
    from brisk_b200 import synth
    total = synth.n_reads_for(GENOME, READ_LEN, COVERAGE)
    first = (batch_index * n_reads) % max(1, total - n_reads)
    arr = synth.reads(gen, READ_LEN, SEED, first, n_reads)
    return arr.reshape(-1)


class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.path = tempfile.mktemp(suffix=".csv")
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.QUERY,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if not self.proc:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in open(self.path):
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(max(mx))
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def cpu_reference(steps, warmup, sample_reads, threads, gen=None):
    """Times oracle/_ref/ref_dump (the unmodified reference library behind our driver TU) on the host cores:
    step j = count sample chunk j + query chunk j-1, exactly the step of the GPU arm on a bounded sample."""
    from brisk_b200 import synth
    binary = os.path.join(ROOT, "oracle", "_ref", "ref_dump")
    if not os.path.exists(binary):
        return None
    if gen is None:
        gen = synth.genome(GENOME, SEED)
    n_steps = steps + warmup
    with tempfile.TemporaryDirectory() as t:
        path = os.path.join(t, "reads.txt")
        with open(path, "wb") as f:
            for j in range(n_steps):
                arr = synth.reads(gen, READ_LEN, SEED, j * sample_reads, sample_reads)
                lines = np.concatenate([arr, np.full((sample_reads, 1), 10, np.uint8)], axis=1)
                f.write(lines.tobytes())
        out = subprocess.run([binary, "steps", path, str(K), str(M), str(B), str(threads), str(sample_reads), str(n_steps)],
                             check=True, capture_output=True, text=True).stdout
    rows = [ln.split() for ln in out.splitlines() if ln.startswith("step ")]
    rows = rows[warmup:]
    secs = sum(float(r[2]) + float(r[3]) for r in rows)
    occ = sum(int(r[4]) for r in rows)
    return dict(value=occ / secs, seconds=secs, occ=occ, steps=len(rows),
                ms_per_step=1000.0 * secs / max(1, len(rows)))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = args.sample_reads
    res = cpu_reference(args.steps, args.warmup, sample, threads)
    if res is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_dump is not built"}))
        return
    desc = "%d reads (%d k-mer occurrences) per step of the same read set" % (sample, sample * (READ_LEN - K + 1))
    line = {
        "impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "k": K, "m": M, "b": B, "sample": desc},
        "cpu_baseline": {"value": res["value"], "unit": UNIT, "cores": threads, "kind": "reference", "sample": desc},
        "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def run_ours(args):
    import torch
    import torch.distributed as dist
    from brisk_b200 import Brisk, Parameters, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly ONE JSON line: anything a library prints to fd 1 meanwhile (NCCL announces its version
    # there at communicator creation) is sent to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    steps, warmup = args.steps, max(3, args.warmup)
    n_batches = steps + 1
    occ_per_batch = BATCH_READS * (READ_LEN - K + 1)

    gen = synth.genome(GENOME, SEED)
    offs = np.arange(BATCH_READS + 1, dtype=np.uint64) * np.uint64(READ_LEN)
    # pinned host batches (e2e arm) and their device-resident copies (value arm); rank r takes every world-th batch
    host_batches, dev_batches = [], []
    for j in range(n_batches):
        t = torch.empty(BATCH_READS * READ_LEN, dtype=torch.uint8, pin_memory=True)
        t.numpy()[:] = read_batch(gen, j * world + rank)
        host_batches.append(t)
        dev_batches.append(t.cuda())
    offs_host = torch.from_numpy(offs.astype(np.int64)).pin_memory()
    offs_dev = offs_host.cuda()
    stream = torch.cuda.Stream()  # a real (non-null) stream handle the library can be pointed at
    torch.cuda.set_stream(stream)
    hint = 2 * GENOME + (1 << 20)  # k > 32: about 1.93 stored keys per genomic k-mer

    def make_index():
        idx = Brisk(Parameters(K, M, B), capacity_hint=hint if world == 1 else int(hint * 1.1 / world), device=local)
        idx.set_stream(stream.cuda_stream)
        if world > 1:
            idx.comm_init_torch(dist, rank, world)
        return idx

    def one_step(idx, batches, offsets, j, acc, stage_times=True):
        # stage_times: read the library's per-stage CUDA-event times after every call.  Reading them waits for the
        # call's store kernel, which count_reads itself does not (it returns once its input is consumed), so the
        # host-buffer arm leaves them alone: there the next call's copy overlaps the store kernel.
        idx.count_reads(batches[j], offsets)
        if stage_times:
            t = idx.last_timing()
            acc["store_ms"] += t["store_ms"]
            acc["enum_ms"] += t["enumerate_ms"]
            acc["pack_ms"] += t["pack_ms"]
            acc["part_ms"] += t["partition_ms"]
            acc["launches"] += t["launches"]
        if j > 0:
            hits, total = idx.query_reads(batches[j - 1], offsets)
            if stage_times:
                tq = idx.last_timing()
                acc["launches"] += tq["launches"]
                acc["q_enum_ms"] += tq["enumerate_ms"]
                acc["q_store_ms"] += tq["store_ms"]
                acc["part_ms"] += tq["partition_ms"]
            acc["hits"] += hits

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(batches, offsets, stage_times=True):
        idx = make_index()
        acc = dict(store_ms=0.0, enum_ms=0.0, pack_ms=0.0, part_ms=0.0, launches=0, hits=0, q_enum_ms=0.0, q_store_ms=0.0)
        wacc = dict(acc)
        # warm-up on the first batches, then a fresh store for the timed steps
        for j in range(min(warmup, n_batches)):
            one_step(idx, batches, offsets, j, wacc)
        idx.close()
        idx = make_index()
        one_step(idx, batches, offsets, 0, dict(wacc))  # step 0 (no previous batch to query) stays untimed
        barrier()
        sampler = ClockSampler(local) if rank == 0 else None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for j in range(1, steps + 1):
            one_step(idx, batches, offsets, j, acc, stage_times)
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        clocks = sampler.stop() if sampler else None
        st = idx.stats()
        idx.close()
        if world > 1:
            tms = torch.tensor([ms], device="cuda")
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            ms = float(tms.item())
        return ms, acc, clocks, st

    probe = make_index()
    n_recs_batch = probe.count_superkmers(dev_batches[0], offs_dev)  # superkmer records of one batch (roofline bytes)
    probe.close()
    ms_dev, acc_dev, clocks, st_dev = timed(dev_batches, offs_dev)
    ms_e2e, acc_e2e, _, _ = timed(host_batches, offs_host, stage_times=False) if not args.skip_e2e else (float("nan"), None, None, None)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    total_occ = steps * occ_per_batch * world
    value = total_occ / (ms_dev / 1000.0)
    e2e = total_occ / (ms_e2e / 1000.0)
    peak, peak_src = peaks()
    # Kernels of a step: enumerate_kernel twice (count + query), records_kernel<COUNT> once, records_kernel<GET> once.
    # The roofline object describes the one with the largest share of the step; the other goes to "roofline_other".
    # Algorithmic bytes per launch follow SURVEY.md §8(d):
    #   enumerate: sum of read lengths + per superkmer record 8 + ceil((n + k - 1) / 4) bytes
    #   insert/count: 18 B per occurrence + 16 B per newly stored k-mer
    new_keys = st_dev["nb_kmers"]
    ins_bytes = occ_per_batch * INSERT_B_PER_OCC + new_keys * INSERT_B_PER_NEW / (steps + 1)
    ins_ms = acc_dev["store_ms"] / steps
    enum_bytes = BATCH_READS * READ_LEN + n_recs_batch * (8 + 0.375) + (occ_per_batch + n_recs_batch * (K - 1)) / 4.0
    enum_ms = (acc_dev["enum_ms"] + acc_dev["q_enum_ms"]) / (2 * steps)

    def roof(kernel, nbytes, ms, launches_per_step, traffic, note):
        a = nbytes / (ms / 1000.0) / 1e9 if ms > 0 else None
        return {"bound": "hbm", "kernel": kernel, "achieved": a, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                "frac": (a / peak) if a else None, "traffic": traffic, "algorithmic_bytes_per_launch": nbytes,
                "ms_per_launch": ms, "launches_per_step": launches_per_step,
                "share_of_step": launches_per_step * ms / (ms_dev / steps), "note": note}

    r_enum = roof("enumerate_kernel<4,ringless> (pack excluded)", enum_bytes, enum_ms, 2, 0.2942e9,
                  "issue-bound (81 % issue-active, 3.05 G warp instructions per launch in profiles/r1i_summary.md), not "
                  "HBM-bound; traffic from the ncu capture of the same 2^20-read batch shape")
    r_ins = roof("records_kernel<4,2,COUNT> (insert/count)", ins_bytes, ins_ms, 1, 12.33e9,
                 "one 32-byte slot read and written back per occurrence: 11.9-12.8 GB of DRAM traffic per launch at the full "
                 "17 GB table (profiles/r1i_full_table.csv, 2.73-2.86 ms per launch = 4.4 TB/s, 67 % of the measured HBM "
                 "peak); the 18 B per occurrence of SURVEY.md 8(d) assumes byte-granular access, a 32-byte sector is what "
                 "HBM moves")
    dominant, other = (r_enum, r_ins) if r_enum["share_of_step"] >= r_ins["share_of_step"] else (r_ins, r_enum)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": ms_dev / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "k": K, "m": M, "b": B, "reads_per_step_per_gpu": BATCH_READS,
                   "occurrences_per_step": occ_per_batch * world, "l2": "inputs larger than L2 (262 MB of reads per step, "
                   "%.1f GB store)" % (st_dev["memory_usage"] / 1e9), "stored_kmers": st_dev["nb_kmers"]},
        "clocks": clocks,
        "e2e": {"value": e2e, "unit": UNIT, "ms_per_step": ms_e2e / steps,
                "h2d_bytes_per_step": world * 2 * (BATCH_READS * READ_LEN + 8 * (BATCH_READS + 1)),
                "d2h_bytes_per_step": world * (32 + 16 + 8 + 8)},
        "gpu_launches": acc_dev["launches"],
        "stage_ms_per_step": {"pack": acc_dev["pack_ms"] / steps, "enumerate": acc_dev["enum_ms"] / steps,
                              "region_order": acc_dev["part_ms"] / steps,
                              "insert": acc_dev["store_ms"] / steps, "query_enumerate": acc_dev["q_enum_ms"] / steps,
                              "query_get": acc_dev["q_store_ms"] / steps},
        "roofline": dominant, "roofline_other": other,
    }
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        res = cpu_reference(6, 2, args.sample_reads, threads, gen)
        if res:
            line["cpu_baseline"] = {"value": res["value"], "unit": UNIT, "cores": threads, "kind": "reference",
                                    "sample": "oracle/_ref/ref_dump steps: %d reads per step, 6 timed steps after 2 warm-up, "
                                              "same read set" % args.sample_reads}
    os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--sample-reads", type=int, default=40000, help="reads per step of the CPU reference runs")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="profiling runs: skip the CPU reference leg")
    ap.add_argument("--skip-e2e", action="store_true", help="profiling runs: skip the host-buffer arm")
    ap.add_argument("--genome", type=int, default=GENOME, help="profiling runs: smaller genome (the default is the config's)")
    args = ap.parse_args()
    globals()["GENOME"] = args.genome
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py — k-mers inserted per second by the counter workload (Brisk<uint8_t>, insert-or-increment).

Default workload = BASELINE.json configs[2] ("cfg3", the synthetic configuration at the metric's k=63 m=21):
100 Mbp random genome, 250 bp error-free reads, k=63 m=21 b=14.  One step = one batch of 2^20 reads pushed through the
hot path (pack -> enumerate superkmers -> insert/count into the bucket store), followed by the batched query of the
previous step's batch (the config's "batched get queries interleaved with inserts").  Successive steps take successive
batches of the read set, so the store grows as in the real job.  The metric counts the INSERTED k-mer occurrences only.

  python bench.py [--gpus N] [--steps K] [--warmup W]          the CUDA path, through the C ABI
  python bench.py --config cfg2|cfg3|cfg4|cfg5                 the other synthetic configurations of BASELINE.json
  python bench.py --scaling weak|strong|weak-fixed-table       N > 1: what is held fixed as N grows
  python bench.py --impl reference ...                         the reference's own CPU code
                                                               (oracle/_ref/ref_dump, all host threads,
                                                               a bounded sample of the same workload)
One JSON line on stdout (rank 0).  Reads are generated on the device (libbrisk_b200's generator, bit-identical to
brisk_b200/synth.py: tests/test_gpu_mid.py); the host-buffer arm copies them to pinned memory before its timed region.

Every run checks itself (`parity_check` in the line): each query step must find every occurrence it asks for, the sum
of the stored counts must equal the occurrences inserted, and the checksum of the stored set (brisk_checksum, summed
over the ranks) must equal the checksum of an independent index over the same reads — at N = 1 an index enumerated by
the run-time-(k, m) kernel with the LITERAL decycling sums, at N > 1 an unsharded index on rank 0 (so the NCCL exchange
is witnessed at every N the driver runs).  The host-buffer (e2e) arm counts the same batches in the same order: its final
store must have the device arm's checksum, shard by shard (`e2e_store_equals_device_arm`).
"""
import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "k-mers inserted/sec (counter, k=63 m=21)"
UNIT = "k-mer occurrences/s"
BATCH_READS = 1 << 20
# algorithmic bytes of the insert/count kernel = 18 B per occurrence (16-B key + DATA read,
# DATA written) + 16 B per distinct k-mer (key claim)
INSERT_B_PER_OCC, INSERT_B_PER_NEW = 18.0, 16.0

CONFIGS = {
    #        k   m   b   genome         L    cov seed keys/genomic k-mer  description
    "cfg2": (31, 15, 12, 5_000_000, 150, 50, 2, 1.0, "cfg2: synthetic 5 Mbp genome, 150 bp reads at 50x, k=31 m=15 b=12, uint8 counts"),
    "cfg3": (63, 21, 14, 100_000_000, 250, 30, 3, 2.0, "cfg3: synthetic 100 Mbp genome, 250 bp reads at 30x, k=63 m=21 b=14, uint8 counts"),
    "cfg4": (64, 23, 14, 100_000_000, 250, 10, 4, 2.0, "cfg4: synthetic 1 Gbp of reads (100 Mbp genome, 250 bp at 10x), k=64 m=23 b=14, uint8 counts"),
    "cfg5": (31, 19, 16, 3_000_000_000, 150, 10, 5, 1.0, "cfg5: synthetic 3 Gbp genome, 150 bp reads at 10x, k=31 m=19 b=16, uint8 counts"),
}
STEP_TEXT = "; step = count a batch of %d reads%s, then batched query of the previous batch"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class Workload:
    def __init__(self, args, world):
        k, m, b, G, L, cov, seed, keys_per, text = CONFIGS[args.config]
        self.name = args.config
        self.k, self.m, self.b, self.L, self.cov, self.seed = k, m, b, L, cov, seed
        self.G = args.genome or G
        note = ""
        if self.name == "cfg5" and world == 1 and not args.genome:
            # 3e9 keys need a 2^33-slot table (137 GB) plus its 34 GB count array: more than one GPU holds.  One GPU
            # runs the largest power-of-two subset that fits; the full genome runs sharded (N >= 2).
            self.G = 1_500_000_000
            note = " [one GPU: 1.5 Gbp subset of the genome, the full table does not fit 180 GB]"
        if args.genome:
            note = " [genome shortened to %d bp]" % self.G
        self.keys_per = keys_per
        self.scaling = args.scaling
        self.world = world
        self.reads_per_rank = BATCH_READS // world if args.scaling == "strong" else BATCH_READS
        self.total_reads = (self.G * cov) // L
        self.occ_per_read = L - k + 1
        self.text = text + note + STEP_TEXT % (self.reads_per_rank, " per GPU" if world > 1 else "")
        # expected stored keys of the whole job (k > 32: about 2-2.8 per genomic k-mer)
        self.hint_total = int(args.hint_scale * keys_per * self.G) + (1 << 20)

    def hint(self):
        if self.world == 1 or self.scaling == "weak-fixed-table":
            return self.hint_total
        return int(self.hint_total * 1.1 / self.world)

    def first_read(self, batch_index):
        n = self.reads_per_rank
        return (batch_index * n) % max(1, self.total_reads - n)


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


# ------------------------------------------------------------------------------------------------ CPU reference legs
def host_reads(W, first, count):
    """(count, L) uint8 reads of the workload for the CPU legs: generated on the device when there is one (bit-identical
    to the numpy generator, tests/test_gpu_mid.py, and much faster for the 3 Gbp genome), else with numpy"""
    from brisk_b200 import synth
    try:
        import torch
        if torch.cuda.is_available():
            if not hasattr(W, "_dgen"):
                W._dgen = synth.device_genome(W.G, W.seed)
            out = synth.device_reads(W._dgen, W.L, W.seed, first, count)
            torch.cuda.synchronize()
            return out.cpu().numpy().reshape(count, W.L)
    except Exception:
        pass
    if not hasattr(W, "_gen"):
        W._gen = synth.genome(W.G, W.seed)
    return synth.reads(W._gen, W.L, W.seed, first, count)


def cpu_reference(W, steps, warmup, sample_reads, threads):
    """Times oracle/_ref/ref_dump (the unmodified reference library behind our driver TU) on the host cores:
    step j = count sample chunk j + query chunk j-1, exactly the step of the GPU arm on a bounded sample."""
    binary = os.path.join(ROOT, "oracle", "_ref", "ref_dump_k64" if W.k == 64 else "ref_dump")
    if not os.path.exists(binary):
        return None
    n_steps = steps + warmup
    with tempfile.TemporaryDirectory() as t:
        path = os.path.join(t, "reads.txt")
        with open(path, "wb") as f:
            for j in range(n_steps):
                arr = host_reads(W, j * sample_reads, sample_reads)
                f.write(np.concatenate([arr, np.full((sample_reads, 1), 10, np.uint8)], axis=1).tobytes())
        out = subprocess.run([binary, "steps", path, str(W.k), str(W.m), str(W.b), str(threads), str(sample_reads), str(n_steps)],
                             check=True, capture_output=True, text=True).stdout
    rows = [ln.split() for ln in out.splitlines() if ln.startswith("step ")]
    rows = rows[warmup:]
    secs = sum(float(r[2]) + float(r[3]) for r in rows)
    occ = sum(int(r[4]) for r in rows)
    ctor = float(re.search(r"ctor_s ([0-9.]+)", out).group(1))
    return dict(value=occ / secs, seconds=secs, occ=occ, steps=len(rows), ctor_s=ctor,
                ms_per_step=1000.0 * secs / max(1, len(rows)))


def cpu_counter_app(W, n_reads, threads_list):
    """the unmodified apps/counter on a FASTA of the first n_reads reads, `-t 1` and `-t nproc`:
    program-reported elapsed (constructor + file reading + counting, apps/counter.cpp:375-381), its `nb kmers`, and the
    `-t N` result compared with `-t 1`."""
    binary = os.path.join(ROOT, "oracle", "_ref", "counter")
    if not os.path.exists(binary) or W.k > 63:  # the app's CLI stops at k = 63 (apps/counter.cpp:32)
        return None
    res = {}
    with tempfile.TemporaryDirectory() as t:
        path = os.path.join(t, "reads.fa")
        arr = host_reads(W, 0, n_reads)
        with open(path, "wb") as f:
            hdr = np.frombuffer(b">r\n", np.uint8)
            rows = np.concatenate([np.tile(hdr, (n_reads, 1)), arr, np.full((n_reads, 1), 10, np.uint8)], axis=1)
            f.write(rows.tobytes())
        occ = n_reads * W.occ_per_read
        for th in threads_list:
            out = subprocess.run([binary, "-f", path, "-k", str(W.k), "-m", str(W.m), "-b", str(W.b), "--mode", "1", "-t", str(th)],
                                 check=True, capture_output=True, text=True).stdout
            el = float(re.search(r"Kmer counted elapsed time: ([0-9.eE+-]+)s", out).group(1))
            nk = int(re.search(r"nb kmers: ([0-9 ,.]+)", out).group(1).replace(" ", "").replace(",", "").replace(".", ""))
            res["t%d" % th] = {"threads": th, "elapsed_s": el, "occurrences_per_s": occ / el, "distinct_per_s": nk / el, "nb_kmers": nk}
    keys = sorted(res)
    res["reads"] = n_reads
    res["same_nb_kmers_all_thread_counts"] = len({res[k_]["nb_kmers"] for k_ in keys}) == 1
    return res


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    W = Workload(args, 1)
    threads = os.cpu_count() or 1
    sample = args.sample_reads
    res = cpu_reference(W, args.steps, args.warmup, sample, threads)
    if res is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_dump is not built"}))
        return
    desc = ("%d reads (%d k-mer occurrences) per step of the same read set (reads 0 .. of the GPU arm's first batch), index "
            "constructed before the timed steps (%.2f s, excluded), reads pre-loaded in RAM; the GPU arm runs 2^20 reads per "
            "step on an index that already holds its warm-up batches" % (sample, sample * W.occ_per_read, res["ctor_s"]))
    line = {
        "impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": W.text, "k": W.k, "m": W.m, "b": W.b, "sample": desc},
        "cpu_baseline": {"value": res["value"], "unit": UNIT, "cores": threads, "kind": "reference", "sample": desc},
        "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------ the CUDA path
def run_ours(args):
    import torch
    import torch.distributed as dist
    from brisk_b200 import Brisk, Parameters, synth
    from brisk_b200.brisk import F_GENERIC_ENUM, F_LITERAL_DECYCLE, F_SERIAL_EXCHANGE

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
    W = Workload(args, world)
    K, M, B, L = W.k, W.m, W.b, W.L
    n_reads = W.reads_per_rank
    steps, warmup = args.steps, max(3, args.warmup)
    n_batches = steps + 1
    occ_per_batch = n_reads * W.occ_per_read

    stream = torch.cuda.Stream()  # a real (non-null) stream handle the library can be pointed at
    torch.cuda.set_stream(stream)
    gen = synth.device_genome(W.G, W.seed)
    # device-resident batches (value arm) and their pinned host copies (e2e arm); rank r takes every world-th batch
    dev_batches, host_batches = [], []
    for j in range(n_batches):
        d = synth.device_reads(gen, L, W.seed, W.first_read(j * world + rank), n_reads)
        dev_batches.append(d)
        if not args.skip_e2e:
            t = torch.empty(n_reads * L, dtype=torch.uint8, pin_memory=True)
            t.copy_(d)
            host_batches.append(t)
    offs_host = torch.from_numpy(np.arange(n_reads + 1, dtype=np.int64) * L).pin_memory()
    offs_dev = offs_host.cuda()
    torch.cuda.synchronize()

    base_flags = F_SERIAL_EXCHANGE if args.serial_exchange else 0

    def make_index(flags=0, sharded=True):
        flags |= base_flags
        idx = Brisk(Parameters(K, M, B), capacity_hint=W.hint() if sharded else W.hint_total, device=local, flags=flags)
        idx.set_stream(stream.cuda_stream)
        if world > 1 and sharded:
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
                acc["q_pack_ms"] += tq["pack_ms"]
                acc["q_store_ms"] += tq["store_ms"]
                acc["part_ms"] += tq["partition_ms"]
            acc["hits"] += hits
            acc["query_steps"] += 1
            acc["query_ok"] += int(hits == occ_per_batch * world and total >= hits)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def new_acc():
        return dict(store_ms=0.0, enum_ms=0.0, pack_ms=0.0, part_ms=0.0, launches=0, hits=0, q_enum_ms=0.0, q_store_ms=0.0,
                    q_pack_ms=0.0, query_steps=0, query_ok=0)

    def timed(batches, offsets, stage_times=True, keep=False):
        idx = make_index()
        acc, wacc = new_acc(), new_acc()
        # warm-up on the first batches, then a fresh store for the timed steps
        for j in range(min(warmup, n_batches)):
            one_step(idx, batches, offsets, j, wacc)
        idx.close()
        idx = make_index()
        one_step(idx, batches, offsets, 0, new_acc())  # step 0 (no previous batch to query) stays untimed
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
        if world > 1:
            tms = torch.tensor([ms], device="cuda")
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            ms = float(tms.item())
        if not keep:
            idx.close()
            idx = None
        return ms, acc, clocks, st, idx

    probe = make_index(sharded=False)
    n_recs_batch = probe.count_superkmers(dev_batches[0], offs_dev)  # superkmer records of one batch (roofline bytes)
    probe.close()
    ms_dev, acc_dev, clocks, st_dev, idx_dev = timed(dev_batches, offs_dev, keep=True)

    # ---- parity check of the timed run's final store (outside every timed region)
    cs = idx_dev.checksum()
    idx_dev.close()
    sums = torch.tensor([cs[0] & 0xffffffff, cs[0] >> 32, cs[1], cs[2]], dtype=torch.int64, device="cuda")
    if world > 1:
        # 64-bit wrap-around sum of the mixes: the two 32-bit halves are added separately
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    mix_sum = (int(sums[0].item()) + (int(sums[1].item()) << 32)) & ((1 << 64) - 1)
    parity = {"query_steps": acc_dev["query_steps"], "query_steps_all_found": acc_dev["query_ok"],
              "sum_counts": int(sums[2].item()), "occurrences_inserted": (steps + 1) * occ_per_batch * world,
              "nb_kmers": int(sums[3].item()), "nb_kmers_stats": st_dev["nb_kmers"], "checksum": "%016x" % mix_sum}
    fits_one_gpu = not (W.name == "cfg5" and W.G > 1_500_000_000)  # the unsharded table of the full 3 Gbp genome does not
    if rank == 0 and not args.no_parity and not fits_one_gpu:
        parity["reference_index"] = "skipped: an unsharded index over the whole 3 Gbp genome does not fit one GPU"
    if rank == 0 and not args.no_parity and fits_one_gpu:
        # the independent index over the same reads: every rank's batches, in any order (the sums are order-independent)
        flags = (F_GENERIC_ENUM | F_LITERAL_DECYCLE) if world == 1 else 0
        ref = make_index(flags=flags, sharded=False)
        buf = torch.empty(n_reads * L, dtype=torch.uint8, device="cuda")
        for j in range(steps + 1):
            for r in range(world):
                synth.device_reads(gen, L, W.seed, W.first_read(j * world + r), n_reads, out=buf)
                ref.count_reads(buf, offs_dev)
        rcs = ref.checksum()
        ref.close()
        parity["reference_index"] = ("run-time-(k,m) kernel, literal decycling sums, one GPU" if world == 1 else
                                     "unsharded index on rank 0 over all ranks' reads")
        parity["reference_checksum"] = "%016x" % rcs[0]
        parity["reference_nb_kmers"] = rcs[2]
        # while the read set is not walked more than once and the coverage is moderate the counts stay below the uint8
        # wrap and their sum is the number of occurrences inserted (beyond that only the reference index can tell)
        wrap_free = (steps + 1) * world * n_reads <= W.total_reads and W.cov <= 60
        parity["ok"] = bool(rcs[0] == mix_sum and rcs[2] == parity["nb_kmers"] == parity["nb_kmers_stats"] and
                            rcs[1] == parity["sum_counts"] and
                            (not wrap_free or parity["sum_counts"] == parity["occurrences_inserted"]) and
                            parity["query_steps_all_found"] == parity["query_steps"])
    elif rank == 0:
        parity["ok"] = bool(parity["query_steps_all_found"] == parity["query_steps"] and parity["nb_kmers"] == parity["nb_kmers_stats"])

    ms_e2e, acc_e2e = float("nan"), None
    if not args.skip_e2e:
        ms_e2e, acc_e2e, _, _, idx_e2e = timed(host_batches, offs_host, stage_times=False, keep=True)
        # the host-buffer arm counts the same batches in the same order: its store must be the device arm's, shard by shard
        cs_e2e = idx_e2e.checksum()
        idx_e2e.close()
        same = int(cs_e2e == cs and acc_e2e["query_ok"] == acc_e2e["query_steps"])
        if world > 1:
            t_same = torch.tensor([same], device="cuda")
            dist.all_reduce(t_same, op=dist.ReduceOp.MIN)
            same = int(t_same.item())
        parity["e2e_store_equals_device_arm"] = bool(same)
        if "ok" in parity:
            parity["ok"] = bool(parity["ok"] and same)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    total_occ = steps * occ_per_batch * world
    value = total_occ / (ms_dev / 1000.0)
    e2e = total_occ / (ms_e2e / 1000.0) if not args.skip_e2e else None
    peak, peak_src = peaks()
    # Kernels of a step: pack + enumerate twice (count + query), records_kernel<COUNT> once, records_kernel<GET> once.
    # Algorithmic bytes per launch follow SURVEY.md §8(d):
    #   stage A+B (pack + enumerate): sum of read lengths (ASCII in) + per superkmer record 8 + ceil((n + k - 1) / 4) bytes
    #   stage C (partition / region order): 2 x the record bytes
    #   insert/count: 18 B per occurrence + 16 B per newly stored k-mer
    new_keys = st_dev["nb_kmers"]
    ins_bytes = occ_per_batch * INSERT_B_PER_OCC + new_keys / max(1, world) * INSERT_B_PER_NEW / (steps + 1)
    ins_ms = acc_dev["store_ms"] / steps
    rec_bytes = n_recs_batch * (8 + 0.375) + (occ_per_batch + n_recs_batch * (K - 1)) / 4.0
    ab_bytes = n_reads * L + rec_bytes
    enum_ms = (acc_dev["enum_ms"] + acc_dev["q_enum_ms"]) / (2 * steps)
    pack_ms = (acc_dev["pack_ms"] + acc_dev["q_pack_ms"]) / (2 * steps)
    part_ms = acc_dev["part_ms"] / (2 * steps)
    step_ms = ms_dev / steps

    def roof(kernel, nbytes, ms, launches_per_step, note):
        a = nbytes / (ms / 1000.0) / 1e9 if ms > 0 else None
        return {"bound": "hbm", "kernel": kernel, "achieved": a, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                "frac": (a / peak) if a else None, "traffic": None, "algorithmic_bytes_per_launch": nbytes,
                "ms_per_launch": ms, "launches_per_step": launches_per_step,
                "share_of_step": launches_per_step * ms / step_ms, "note": note}

    r_enum = roof("pack_reads_kernel + enumerate_kernel<%d,%d> (stage A+B of SURVEY.md 8(d))" % (K, M), ab_bytes, pack_ms + enum_ms, 2,
                  "bound by the integer ALU pipe (79.5 %% of its peak), not by HBM (ncu: profiles/r2_final_kernels.json); the bytes are the stage's ASCII input "
                  "(read by pack_reads_kernel) plus the packed records written; pack %.3f ms + enumerate %.3f ms per launch; "
                  "traffic: see traffic_source" % (pack_ms, enum_ms))
    # DRAM traffic of the two kernels from the committed ncu --set full captures of this command (cfg 3 at full size):
    # only for the workload they were taken on
    kfile = os.path.join(ROOT, "profiles", "r2_final_kernels.json")
    if W.name == "cfg3" and not args.genome and n_reads == BATCH_READS and os.path.exists(kfile):
        kk = json.load(open(kfile))["kernels"]
        e, pk = kk["enumerate_kernel<63,21,4>"], kk["pack_reads_kernel"]
        r_enum["traffic"] = e["dram_bytes_read"] + e["dram_bytes_write"] + pk["dram_bytes_read"] + pk["dram_bytes_write"]
        r_enum["traffic_source"] = ("ncu --set full captures of `bench.py --steps 2 --warmup 3`, one launch each, "
                                    "profiles/r2_final_kernels.json (+ r2_final_enumerate_full.csv, r2_final_pack_full.csv): "
                                    "enumerate_kernel %.1f MB (tree 6e2412a, kernel unchanged since) + pack_reads_kernel %.1f MB "
                                    "(captured on 6e2412a, before the kernel's last rewrite: a streaming kernel, ASCII in, 2-bit "
                                    "words out); not re-measured by this run" %
                                    ((e["dram_bytes_read"] + e["dram_bytes_write"]) / 1e6, (pk["dram_bytes_read"] + pk["dram_bytes_write"]) / 1e6))
    r_ins = roof("records_kernel<COUNT> (insert/count)", ins_bytes, ins_ms, 1,
                 "one 32-byte slot sector read per occurrence plus the count word of the side array (red.add, one sector per "
                 "superkmer); the 18 B per occurrence of SURVEY.md 8(d) assumes byte-granular access")
    dominant, other = (r_enum, r_ins) if r_enum["share_of_step"] >= r_ins["share_of_step"] else (r_ins, r_enum)
    r_ep = roof("stage A+B+C: pack + enumerate + partition (the north star's enumerate+partition target)", ab_bytes + 2 * rec_bytes,
                pack_ms + enum_ms + part_ms, 2, "partition = region order of the records before the store kernel (and the owner "
                "partition at N > 1)")
    line = {
        "metric": METRIC if W.name == "cfg3" else "k-mers inserted/sec (counter, k=%d m=%d)" % (K, M),
        "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": step_ms, "higher_is_better": True,
        "scaling": "strong" if W.scaling == "strong" else "weak", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": {"workload": W.text, "k": K, "m": M, "b": B, "reads_per_step_per_gpu": n_reads,
                   "occurrences_per_step": occ_per_batch * world, "scaling_mode": W.scaling,
                   "capacity_hint_per_gpu": W.hint(),
                   "l2": "inputs larger than L2 (%d MB of reads per step, %.1f GB store)" % (n_reads * L >> 20, st_dev["memory_usage"] / 1e9),
                   "stored_kmers": st_dev["nb_kmers"], "distinct_per_s": st_dev["nb_kmers"] / ((steps + 1) * step_ms / 1000.0)},
        "clocks": clocks,
        "gpu_launches": acc_dev["launches"],
        "stage_ms_per_step": {"pack": acc_dev["pack_ms"] / steps, "enumerate": acc_dev["enum_ms"] / steps,
                              "region_order": acc_dev["part_ms"] / steps,
                              "insert": acc_dev["store_ms"] / steps, "query_pack": acc_dev["q_pack_ms"] / steps,
                              "query_enumerate": acc_dev["q_enum_ms"] / steps, "query_get": acc_dev["q_store_ms"] / steps},
        "roofline": dominant, "roofline_other": other, "roofline_enumerate_partition": r_ep,
        "parity_check": parity,
    }
    if not args.skip_e2e:
        line["e2e"] = {"value": e2e, "unit": UNIT, "ms_per_step": ms_e2e / steps,
                       "h2d_bytes_per_step": world * 2 * (n_reads * L + 8 * (n_reads + 1)),
                       "d2h_bytes_per_step": world * (32 + 16 + 8 + 8)}
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        res = cpu_reference(W, 6, 2, args.sample_reads, threads)
        if res:
            line["cpu_baseline"] = {"value": res["value"], "unit": UNIT, "cores": threads, "kind": "reference",
                                    "sample": "oracle/_ref/ref_dump steps (the unmodified reference library, insert_superkmer / "
                                              "get_superkmer under OpenMP): %d reads per step, 6 timed steps after 2 warm-up, same "
                                              "read set, constructor (%.2f s) excluded" % (args.sample_reads, res["ctor_s"])}
            app = cpu_counter_app(W, 4 * args.sample_reads, [1, threads])
            if app:
                line["cpu_baseline"]["counter_app"] = app
    os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg3", choices=sorted(CONFIGS))
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong", "weak-fixed-table"],
                    help="N > 1: weak = 2^20 reads per GPU and step, table hint divided by N; strong = 2^20 reads per step in total; "
                         "weak-fixed-table = weak with every rank's table held at the one-GPU size")
    ap.add_argument("--sample-reads", type=int, default=40000, help="reads per step of the CPU reference runs")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="profiling runs: skip the CPU reference leg")
    ap.add_argument("--skip-e2e", action="store_true", help="profiling runs: skip the host-buffer arm")
    ap.add_argument("--no-parity", action="store_true", help="profiling runs: skip the independent index of the parity check")
    ap.add_argument("--serial-exchange", action="store_true", help="experiments: N > 1 without the chunk-pipelined exchange")
    ap.add_argument("--hint-scale", type=float, default=1.0, help="experiments: multiply the capacity hint (table size)")
    ap.add_argument("--genome", type=int, default=0, help="profiling runs: smaller genome (the default is the config's)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

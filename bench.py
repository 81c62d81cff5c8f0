#!/usr/bin/env python
"""bench.py — aligned reads/sec of the BSF read-alignment hot path (BASELINE.json metric) on N B200s of one node.

Workload (BASELINE.json configs[1]): synthetic 100 Mbp single-contig i.i.d. genome (seed 2), 1 000 000 single-end 100 bp reads with
e ~ U{0..3} edits (70 % substitution / 15 % insertion / 15 % deletion), `-k 3`, per GPU. A step = one pass of the whole path
(quick scan -> search/locate/verify rounds -> hit records) over that batch. Reads are sharded over GPUs (each holds a replicated
index; no collective on the data path), so N GPUs process N batches: weak scaling.

  value      device-resident: reads and result buffers already in HBM (gw_align_batch_device)
  e2e        through the reference-facing C-ABI call with pinned HOST buffers (gw_align_batch): H2D of reads, D2H of hit records
  roofline   the dominant FM kernel family: occurrence-table lookups x 64 B / CUDA-event time of that family, against the
             measured HBM copy bandwidth (MEASURED_PEAKS.json)
  cpu_baseline / --impl reference: the C++ oracle (line-by-line restatement of the JVM reference, which cannot run here) on the host cores
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "aligned reads/sec (100bp, k<=3 edits)"
UNIT = "reads/s"


def workload(args):
    return dict(genome_bp=args.genome_bp, reads=args.reads, read_len=args.read_len, k=args.k, genome_seed=2, reads_seed=12)


def cache_dir(args):
    return os.path.join(args.cache, "gw_bench_%d" % args.genome_bp)


def make_genome(args):
    from genome_weaver_align_b200 import synth
    return synth.random_genome(args.genome_bp, 2)


def ensure_index(args, genome, use_gpu, device=0):
    """writes the FASTA and the `genome-weaver bwt` files once per box (setup, untimed)"""
    from genome_weaver_align_b200 import synth
    d = cache_dir(args)
    os.makedirs(d, exist_ok=True)
    fa = os.path.join(d, "genome.fa")
    done = os.path.join(d, "index.done")
    if os.path.exists(done):
        return fa, 0.0, "cached"
    synth.write_fasta(fa, [("chr1", genome)])
    t = time.time()
    how = None
    if use_gpu:
        try:
            import genome_weaver_align_b200 as gw
            gw.bwt(fa, device)
            how = "device builder (gw_index_build)"
        except Exception as e:  # noqa: BLE001
            sys.stderr.write("device index build unavailable (%s); using the oracle's CPU builder\n" % e)
    if how is None:
        import oracle_lib as O
        O.OracleIndex.build_files(fa)
        how = "oracle CPU builder"
    open(done, "w").write(how)
    return fa, time.time() - t, how


def make_reads(args, genome, rank):
    from genome_weaver_align_b200 import synth
    arr = synth.simulate_reads_fast(genome, args.reads, args.read_len, args.k, 12 + rank)
    return np.ascontiguousarray(arr)


class ClockSampler(threading.Thread):
    def __init__(self, device):
        super().__init__(daemon=True)
        self.device, self.stop_flag, self.sm, self.reasons, self.max_mhz = device, False, [], set(), None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.device)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "hw_power_brake": 0x80}
            while not self.stop_flag:
                self.sm.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:  # noqa: BLE001
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for n, bit in names.items():
                    if r & bit:
                        self.reasons.add(n)
                time.sleep(0.05)
        except Exception as e:  # noqa: BLE001
            self.reasons.add("sampler_error:%s" % type(e).__name__)

    def summary(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.sm)}


def oracle_baseline(args, fa, reads_arr, n_sample, threads):
    import oracle_lib as O
    oi = O.OracleIndex.load(fa)
    reads = [r.tobytes().decode() for r in reads_arr[:n_sample]]
    names = ["r%d" % i for i in range(len(reads))]
    res = oi.align(names, reads, None, nthreads=threads, k=args.k)
    oi.close()
    return len(reads) / res["elapsed"], res


def parity_check(lib, L, idx, bsf, reads_arr, h_hits, n_hits, h_cig, oracle_res, n_s):
    """the timed GPU batch against the oracle on the first n_s reads of the same workload: whole SAM lines and the four per-read
    debug counters (reads on which the reference throws: status only). Returns the `parity` object of the JSON line."""
    hits = h_hits.numpy().view(L.HIT_DTYPE)[:n_hits]
    cut = int(np.searchsorted(hits["read_index"], n_s, side="left"))
    sub = np.ascontiguousarray(hits[:cut])
    m = reads_arr.shape[1]
    names = ["r%d" % i for i in range(n_s)]
    nb, noff = L.concat(names)
    bases = np.ascontiguousarray(reads_arr[:n_s]).reshape(-1)
    boff = np.arange(n_s + 1, dtype=np.uint64) * np.uint64(m)
    lines = np.zeros(max(n_s, 1), dtype=np.uint32)
    cap = C.c_uint64(n_s * (2 * m + 200) + 4096)
    buf = C.create_string_buffer(int(cap.value))
    cig = h_cig.numpy().view(np.uint32)
    L.check(lib.gw_sam_format(idx._h, L.ptr(nb), L.ptr(noff), L.ptr(bases), L.ptr(boff), None, None, n_s, L.ptr(sub), len(sub), L.ptr(cig),
                              C.cast(buf, C.c_void_p), C.byref(cap), L.ptr(lines)))
    got_lines = buf.raw[:cap.value].decode().split("\n")[:-1]
    ctr = np.zeros((n_s, 4), dtype=np.uint32)
    L.check(lib.gw_aligner_read_counters(bsf._h, 0, n_s, L.ptr(ctr)))
    status = np.full(n_s, -1, dtype=np.int32)
    heads = sub[sub["is_head"] == 1]
    status[heads["read_index"][::-1]] = heads["status"][::-1]
    mism, cmism, first, p = 0, 0, None, 0
    for i in range(n_s):
        g = got_lines[p:p + int(lines[i])]
        p += int(lines[i])
        w = oracle_res["sam"][i]
        ws = oracle_res["status"][i]
        bad = g != w or int(status[i]) != ws
        badc = ws != 2 and [int(x) for x in ctr[i]] != list(oracle_res["counters"][i])
        if bad:
            mism += 1
        if badc:
            cmism += 1
        if (bad or badc) and first is None:
            first = {"read": i, "got": g, "want": w, "got_counters": [int(x) for x in ctr[i]], "want_counters": list(oracle_res["counters"][i])}
    out = {"checked": n_s, "mismatches": mism, "counter_mismatches": cmism,
           "what": "whole SAM lines + status + numFMIndexSearches/numSW/numCutOff/numFiltered per read, GPU (timed e2e batch) vs C++ oracle"}
    if first:
        out["first_mismatch"] = first
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    genome = make_genome(args)
    fa, _, how = ensure_index(args, genome, use_gpu=True)
    reads = make_reads(args, genome, 0)
    threads = os.cpu_count() or 1
    n_sample = min(args.reads, args.ref_sample)
    import oracle_lib as O
    oi = O.OracleIndex.load(fa)
    sample = [r.tobytes().decode() for r in reads[:n_sample]]
    names = ["r%d" % i for i in range(n_sample)]
    times = []
    for it in range(args.warmup + args.steps):
        res = oi.align(names, sample, None, nthreads=threads, k=args.k)
        if it >= args.warmup:
            times.append(res["elapsed"])
    v = n_sample * len(times) / sum(times)
    sample_desc = "first %d reads of the workload per step, %d threads sharing one index, output discarded" % (n_sample, threads)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1000.0 * sum(times) / len(times), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64",
        "data": "synthetic", "config": dict(workload="configs[1]: 100 Mbp synthetic genome, 100 bp reads, k=3 (CPU oracle; the JVM reference cannot run here)",
                                                      **dict(workload(args), reads_per_step=n_sample, reads_in_workload=args.reads)),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample_desc},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "index_built_by": how}))


def run_ours(args):
    import torch
    import torch.distributed as dist
    import genome_weaver_align_b200 as gw
    from genome_weaver_align_b200 import _lib as L

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    genome = make_genome(args)
    t_build, how = 0.0, "cached"
    if rank == 0:
        fa, t_build, how = ensure_index(args, genome, use_gpu=True, device=local)
    barrier()
    fa = os.path.join(cache_dir(args), "genome.fa")
    idx = gw.FMIndexOnGenome.load(fa, local)
    if args.locate_step is not None:
        idx.locateTable(args.locate_step)
    reads = make_reads(args, genome, rank)
    n, m = reads.shape
    offsets = (np.arange(n + 1, dtype=np.uint64) * np.uint64(m))
    bsf = gw.BidirectionalSuffixFilter(idx, gw.AlignmentConfig(k=args.k))
    for opt in args.opt:
        kname, v = opt.split("=")
        bsf.set_option(kname, int(v))
    lib = L.lib()
    dev = torch.device("cuda", local)

    # ---- device-resident arm
    d_bases = torch.from_numpy(reads.reshape(-1)).to(dev)
    d_off = torch.from_numpy(offsets.view(np.int64)).to(dev)
    hits_cap, cig_cap = 2 * n + 16, 16 * n + 64
    d_hits = torch.empty(hits_cap * 44, dtype=torch.uint8, device=dev)
    d_cig = torch.empty(cig_cap, dtype=torch.int32, device=dev)
    nh, nc = C.c_uint64(0), C.c_uint64(0)
    cnt = L.Counters()
    ms = (C.c_float * 6)()

    def step_device():
        L.check(lib.gw_align_batch_device(bsf._h, d_bases.data_ptr(), d_off.data_ptr(), n, n * m, m, d_hits.data_ptr(), hits_cap, d_cig.data_ptr(),
                                          cig_cap, C.byref(nh), C.byref(nc), C.byref(cnt)))

    for _ in range(args.warmup):
        step_device()
    fam = np.zeros(6)
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_device()
        lib.gw_aligner_last_timing(bsf._h, ms)
        fam += np.array(list(ms))
    barrier()
    t1 = time.perf_counter()
    sampler.stop_flag = True
    sampler.join()
    dt = torch.tensor([t1 - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    dev_time = float(dt.item())
    counters = cnt.as_dict()
    fam_ms = fam / args.steps

    # ---- end-to-end arm: pinned host buffers through gw_align_batch
    h_bases = torch.from_numpy(reads.reshape(-1)).pin_memory()
    h_off = torch.from_numpy(offsets.view(np.int64)).pin_memory()
    h_hits = torch.empty(hits_cap * 44, dtype=torch.uint8).pin_memory()
    h_cig = torch.empty(cig_cap, dtype=torch.int32).pin_memory()

    def step_host():
        a, b = C.c_uint64(hits_cap), C.c_uint64(cig_cap)
        L.check(lib.gw_align_batch(bsf._h, h_bases.data_ptr(), h_off.data_ptr(), n, h_hits.data_ptr(), C.byref(a), h_cig.data_ptr(), C.byref(b), None))
        return a.value, b.value

    for _ in range(min(args.warmup, 1)):
        step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e_nh, e_nc = step_host()
    barrier()
    t1 = time.perf_counter()
    de = torch.tensor([t1 - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(de, op=dist.ReduceOp.MAX)
    e2e_time = float(de.item())

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:  # noqa: BLE001
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        # Kernel families of one step. Search, locate and verification run inside ONE kernel (gw_flow_kernel; further launches of it only
        # for reads that need a still larger work area), so their lookups are charged to the same CUDA-event time: 64 B per occurrence-table lookup
        # (search: 2 per bidirectional step; locate: 1 per LF step) + 4 B per SA sample.
        fams = {
            "scan": (counters["occ_lookups_scan"] * 64.0, fam_ms[0]),
            "search+locate+verify": ((counters["occ_lookups_search"] + counters["occ_lookups_locate"]) * 64.0 + counters["sw_calls"] * 4.0,
                                     fam_ms[1] + fam_ms[2] + fam_ms[3]),
        }
        roof_all = {k: {"bytes": b, "ms": float(t), "gbs": (b / (t * 1e-3) / 1e9 if t > 0 else None)} for k, (b, t) in fams.items()}
        roof_all["search+locate+verify"].update(occ_lookups_search=counters["occ_lookups_search"], occ_lookups_locate=counters["occ_lookups_locate"],
                                                sw_calls=counters["sw_calls"], fm_searches=counters["fm_searches"])
        roof_all["emit"] = {"ms": float(fam_ms[4])}
        dom = max(fams, key=lambda k: fams[k][1])
        ach = roof_all[dom]["gbs"] or 0.0
        # the random-access bound, measured in this job: independent 64 B gathers over the same occurrence table
        gbs, gms = C.c_double(0), C.c_double(0)
        gather = None
        if lib.gw_gather_benchmark(idx._h, 1, 1 << 28, 3, C.byref(gbs), C.byref(gms)) == 0:
            gather = float(gbs.value)
        traffic = None
        try:   # DRAM bytes per step of the dominant kernel family from the committed ncu --set full capture (profiles/)
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            if tj.get("reads") == n and tj.get("genome_bp") == args.genome_bp:
                traffic = tj["dram_bytes_per_step"]
        except Exception:  # noqa: BLE001
            pass
        value = world * n * args.steps / dev_time
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000.0 * dev_time / args.steps, "device_event_ms_per_step": float(fam_ms[5]), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32/u64 integer", "data": "synthetic",
            "config": dict(workload="configs[1]: 100 Mbp synthetic genome, 1M x 100 bp reads per GPU, k=3, best hit", **workload(args),
                           l2="working set (index %.2f GB + per-read work areas of several GB rewritten every step) exceeds the 126 MB L2; no explicit flush" % (
                               (genome.size / 128 * 64 * 2 + genome.size / 8 + genome.size * 3 / 8) / 1e9),
                           parallelism="reads sharded over %d GPU(s), index replicated, no collective" % world, index_built_by=how,
                           index_build_s=round(t_build, 2), locate_table_step=idx.locateStep()),
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic,
                         "peak_source": peak_src, "random_gather_gbs": gather, "frac_of_random_gather": (ach / gather if gather else None),
                         "note": "algorithmic bytes = occurrence-table lookups x 64 B (+4 B per SA sample) over the CUDA-event time of the kernel family; "
                                 "random_gather_gbs = independent 64 B gathers over the same table measured in this job (the bound a dependent FM "
                                 "step can approach); traffic = DRAM bytes of the family per step from the ncu capture in profiles/ (work areas, "
                                 "DP history and thread stacks, not the index, dominate it)",
                         "families": roof_all},
            "e2e": {"value": world * n * args.steps / e2e_time, "unit": UNIT, "h2d_bytes_per_step": int(n * m + (n + 1) * 8),
                    "d2h_bytes_per_step": int(e_nh * 44 + e_nc * 4), "ms_per_step": 1000.0 * e2e_time / args.steps},
            "gpu_launches": int(counters["gpu_launches"]) * args.steps, "counters_per_step": counters,
            "clocks": sampler.summary(),
        }
        parity_bad = False
        if not args.no_cpu:
            threads = os.cpu_count() or 1
            n_s = min(n, args.ref_sample if world == 1 else min(args.ref_sample, 20_000))
            v, res = oracle_baseline(args, fa, reads, n_s, threads)
            if world == 1:
                v1, _ = oracle_baseline(args, fa, reads, max(1, n_s // max(threads, 1)), 1)
                line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                        "sample": "first %d reads of the workload, %d threads sharing one index (C++ oracle; JVM reference unavailable on this box)" % (n_s, threads),
                                        "single_thread_value": v1}
            line["parity"] = parity_check(lib, L, idx, bsf, reads, h_hits, e_nh, h_cig, res, n_s)
            parity_bad = line["parity"]["mismatches"] != 0 or line["parity"]["counter_mismatches"] != 0
        print(json.dumps(line))
        if parity_bad:
            sys.stderr.write("bench.py: the GPU output differs from the oracle on the checked sample\n")
            bsf.close()
            idx.close()
            sys.exit(3)
    bsf.close()
    idx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--genome-bp", type=int, default=100_000_000)
    ap.add_argument("--reads", type=int, default=1_000_000)
    ap.add_argument("--read-len", type=int, default=100)
    ap.add_argument("--k", type=int, default=3)
    ap.add_argument("--ref-sample", type=int, default=200_000)
    ap.add_argument("--cache", default="/tmp")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--locate-step", type=int, default=None, help="density of the locate table (1..32; default: what gw_index_open chooses)")
    ap.add_argument("--opt", action="append", default=[], help="aligner option name=value (e.g. scan_coop=0)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py — aligned reads/sec of the BSF read-alignment hot path (BASELINE.json metric) on N B200s of one node.

Default workload (BASELINE.json configs[1], `--config 2`): synthetic 100 Mbp single-contig i.i.d. genome (seed 2), 1 000 000 single-end
100 bp reads with e ~ U{0..3} edits (70 % substitution / 15 % insertion / 15 % deletion), `-k 3`, per GPU. `--config 1|3|4|5` run the
other configs of BASELINE.json at their stated sizes; they are parity cases and roofline references, not the line the
driver reads. A step = one pass of the whole path (quick scan -> search / locate / verify -> hit records) over one batch. Reads are
sharded over GPUs (each holds a replicated index; no collective on the data path), so N GPUs process N batches: weak scaling.

  value      device-resident: reads and result buffers already in HBM (gw_align_batch_device)
  e2e        through the reference-facing C-ABI call with pinned HOST buffers (gw_align_batch): H2D of reads, D2H of hit records
  roofline   the dominant FM kernel family: occurrence-table lookups x 64 B (+ 4 B per locate) / CUDA-event time of that family,
             against the measured HBM copy bandwidth (MEASURED_PEAKS.json)
  parity     the timed e2e batch against the C++ oracle on a sample of the same workload: whole SAM lines, status, per-read counters
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

# BASELINE.json configs, with the sizes / seeds of SURVEY.md 8(d). reads = per GPU and step; sample = reads checked against the oracle.
CONFIGS = {
    1: dict(name="configs[0]: 4.6 Mbp i.i.d. genome, 10 000 x 100 bp reads, k=2, best hit", genome="random", genome_bp=4_600_000, genome_seed=1,
            reads=10_000, lengths=[100], edits=2, k=2, reads_seed=11, report="besthit", extended=False, sample=10_000),
    2: dict(name="configs[1]: 100 Mbp synthetic genome, 1M x 100 bp reads per GPU, k=3, best hit", genome="random", genome_bp=100_000_000,
            genome_seed=2, reads=1_000_000, lengths=[100], edits=3, k=3, reads_seed=12, report="besthit", extended=False, sample=200_000),
    3: dict(name="configs[2]: 3.1 Gbp genome in 24 contigs with injected repeats and N runs, 150 bp reads (10M over 8 GPUs = 1.25M per GPU), k=3",
            genome="human", genome_bp=3_100_000_000, genome_seed=3, reads=1_250_000, lengths=[150], edits=3, k=3, reads_seed=13, report="besthit",
            extended=False, sample=20_000),
    4: dict(name="configs[3]: 3.1 Gbp genome of config 3, 1M x 250 bp reads with 6 edits (10 % with 0..5), k=6, extended domain "
                 "(the reference throws on part of these reads: SURVEY.md A.6)", genome="human", genome_bp=3_100_000_000, genome_seed=3,
            reads=1_000_000, lengths=[250], edits=6, k=6, reads_seed=14, report="besthit", extended=True, sample=5_000),
    5: dict(name="configs[4]: 50 Mbp low-complexity genome, all best hits, 200 000 reads per length 36..250 bp in one batch, k=3, extended domain "
                 "for the lengths >= 156", genome="lowcomplexity", genome_bp=50_000_000, genome_seed=5, reads=1_600_000,
            lengths=[36, 50, 75, 100, 125, 150, 200, 250], edits=3, k=3, reads_seed=15, report="allhits", extended=True, sample=16_000),
}
REPORT = {"besthit": 0, "allhits": 1, "topl": 2}


_T0 = time.time()


def log(msg):
    """stage log on stderr (the big configs spend minutes in set-up: a run that is cut short must say where it was)"""
    sys.stderr.write("[bench %7.1fs] %s\n" % (time.time() - _T0, msg))
    sys.stderr.flush()


def spec_of(args):
    s = dict(CONFIGS[args.config])
    if args.genome_bp:
        s["genome_bp"] = args.genome_bp
    if args.reads:
        s["reads"] = args.reads
    if args.ref_sample is not None:
        s["sample"] = args.ref_sample
    s["sample"] = min(s["sample"], s["reads"])
    return s


def workload(spec):
    return dict(genome_bp=spec["genome_bp"], reads=spec["reads"], read_len=spec["lengths"][0] if len(spec["lengths"]) == 1 else spec["lengths"],
                k=spec["k"], genome_seed=spec["genome_seed"], reads_seed=spec["reads_seed"], report=spec["report"], extended=spec["extended"])


def cache_dir(args, spec):
    return os.path.join(args.cache, "gw_bench_%s_%d_%d" % (spec["genome"], spec["genome_bp"], spec["genome_seed"]))


def make_genome(args, spec):
    """(contigs, concatenated genome); big genomes are cached as .npy next to the index files (they take a minute to draw)"""
    from genome_weaver_align_b200 import synth
    d = cache_dir(args, spec)
    os.makedirs(d, exist_ok=True)
    npy = os.path.join(d, "genome.npy")
    n, seed = spec["genome_bp"], spec["genome_seed"]
    if os.path.exists(npy + ".done"):
        g = np.load(npy)
    else:
        if spec["genome"] == "random":
            g = synth.random_genome(n, seed)
        elif spec["genome"] == "lowcomplexity":
            g = synth.low_complexity_genome(n, seed)
        else:   # 24 contigs, 20 repeat families of 300 bp..6 kb with 10^2..10^4 copies and 0..10 % divergence, 0.5 % N, last 1 kb unique
            g = synth.random_genome(n, seed, n_frac=0.005)
            g = synth.inject_repeats(g, seed, families=20, copies=(100, max(200, min(10000, n // 300000))))
            g[-1000:] = synth.random_genome(1000, seed + 30)
        if int(os.environ.get("RANK", "0")) == 0 and n >= 500_000_000:
            np.save(npy, g)
            open(npy + ".done", "w").write("ok")
    if spec["genome"] == "human":
        w = np.arange(24, 0, -1, dtype=np.float64)
        cuts = np.concatenate([[0], np.cumsum((w / w.sum() * n).astype(np.int64))])
        cuts[-1] = n
        contigs = [("chr%d" % (i + 1), g[cuts[i]:cuts[i + 1]]) for i in range(24)]
    else:
        contigs = [("chr1", g)]
    return contigs, g


def ensure_index(args, spec, contigs, use_gpu, device=0):
    """writes the FASTA and the `genome-weaver bwt` files once per box (setup, untimed)"""
    from genome_weaver_align_b200 import synth
    d = cache_dir(args, spec)
    os.makedirs(d, exist_ok=True)
    fa = os.path.join(d, "genome.fa")
    done = os.path.join(d, "index.done")
    if os.path.exists(done):
        return fa, 0.0, open(done).read()
    synth.write_fasta(fa, contigs)
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


def make_reads(spec, genome, rank):
    """list of (n, m) uint8 arrays, one per read length (equal shares of spec['reads'])"""
    from genome_weaver_align_b200 import synth
    per = spec["reads"] // len(spec["lengths"])
    out = []
    for j, m in enumerate(spec["lengths"]):
        arr = synth.simulate_reads_fast(genome, per, m, spec["edits"], spec["reads_seed"] + rank + 1000 * j, exact_edits=6 if spec["edits"] == 6 else None)
        if spec["edits"] == 6:   # config 4: e = 6 for 90 % of the reads, e ~ U{0..5} for the rest
            few = synth.simulate_reads_fast(genome, per, m, 5, spec["reads_seed"] + rank + 77)
            pick = np.random.default_rng(spec["reads_seed"] + 5).random(per) < 0.10
            arr[pick] = few[pick]
        out.append(np.ascontiguousarray(arr))
    return out


def flatten(groups):
    """read groups -> (concatenated bases uint8, offsets uint64[n+1], max length)"""
    bases = np.concatenate([g.reshape(-1) for g in groups])
    lens = np.concatenate([np.full(g.shape[0], g.shape[1], dtype=np.uint64) for g in groups])
    off = np.zeros(len(lens) + 1, dtype=np.uint64)
    off[1:] = np.cumsum(lens)
    return np.ascontiguousarray(bases), off, int(max(g.shape[1] for g in groups))


def sample_indices(n, n_s):
    """the reads checked against the oracle: a prefix for a one-length batch would do, but a batch of several read lengths is sampled
    evenly, so every length is covered"""
    return np.unique(np.linspace(0, n - 1, n_s).astype(np.int64)) if n_s < n else np.arange(n, dtype=np.int64)
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


def oracle_cfg(spec):
    return dict(k=spec["k"], report_type=REPORT[spec["report"]], extended=1 if spec["extended"] else 0)


def oracle_run(fa, seqs, threads, spec, oi=None):
    import oracle_lib as O
    own = oi is None
    if own:
        oi = O.OracleIndex.load(fa)
    names = ["r%d" % i for i in range(len(seqs))]
    res = oi.align(names, seqs, None, nthreads=threads, **oracle_cfg(spec))
    if own:
        oi.close()
    return len(seqs) / res["elapsed"], res


def parity_check(lib, L, idx, bsf, bases, off, pick, h_hits, n_hits, h_cig, oracle_res):
    """the timed GPU batch against the oracle on the reads `pick` (ascending indices) of the same workload: whole SAM lines, status and
    the four per-read debug counters (reads on which the reference throws: status only). Returns the `parity` object of the JSON line."""
    n_s = len(pick)
    hits = h_hits.numpy().view(L.HIT_DTYPE)[:n_hits]
    pos = np.searchsorted(pick, hits["read_index"])
    pos[pos >= n_s] = 0
    keep = pick[pos] == hits["read_index"]
    sub = np.ascontiguousarray(hits[keep])
    sub["read_index"] = pos[keep].astype(np.uint32)
    names = ["r%d" % i for i in range(n_s)]
    nb, noff = L.concat(names)
    lens = (off[pick + 1] - off[pick]).astype(np.int64)
    boff = np.zeros(n_s + 1, dtype=np.uint64)
    boff[1:] = np.cumsum(lens)
    sb = np.concatenate([bases[int(off[i]):int(off[i + 1])] for i in pick]) if n_s else np.zeros(0, dtype=np.uint8)
    sb = np.ascontiguousarray(sb)
    lines = np.zeros(max(n_s, 1), dtype=np.uint32)
    cap = C.c_uint64(int(boff[-1]) * 2 * max(1, len(sub) // max(n_s, 1) + 1) + 300 * (len(sub) + n_s) + 4096)
    buf = C.create_string_buffer(int(cap.value))
    cig = h_cig.numpy().view(np.uint32)
    L.check(lib.gw_sam_format(idx._h, L.ptr(nb), L.ptr(noff), L.ptr(sb), L.ptr(boff), None, None, n_s, L.ptr(sub), len(sub), L.ptr(cig),
                              C.cast(buf, C.c_void_p), C.byref(cap), L.ptr(lines)))
    got_lines = buf.raw[:cap.value].decode().split("\n")[:-1]
    ctr_all = np.zeros((int(pick[-1]) + 1 if n_s else 1, 4), dtype=np.uint32)
    L.check(lib.gw_aligner_read_counters(bsf._h, 0, len(ctr_all), L.ptr(ctr_all)))
    ctr = ctr_all[pick]
    status = np.full(n_s, -1, dtype=np.int32)
    heads = sub[sub["is_head"] == 1]
    status[heads["read_index"][::-1]] = heads["status"][::-1]
    mism, cmism, first, p = 0, 0, None, 0
    counts = {}
    for i in range(n_s):
        g = got_lines[p:p + int(lines[i])]
        p += int(lines[i])
        w = oracle_res["sam"][i]
        ws = oracle_res["status"][i]
        counts[ws] = counts.get(ws, 0) + 1
        bad = g != w or int(status[i]) != ws
        badc = ws != 2 and [int(x) for x in ctr[i]] != list(oracle_res["counters"][i])
        mism += bool(bad)
        cmism += bool(badc)
        if (bad or badc) and first is None:
            first = {"read": int(pick[i]), "got": g, "want": w, "got_status": int(status[i]), "want_status": ws,
                     "got_counters": [int(x) for x in ctr[i]], "want_counters": list(oracle_res["counters"][i])}
    out = {"checked": n_s, "mismatches": mism, "counter_mismatches": cmism, "oracle_status_counts": {str(k): v for k, v in sorted(counts.items())},
           "what": "whole SAM lines + status + numFMIndexSearches/numSW/numCutOff/numFiltered per read, GPU (timed e2e batch) vs C++ oracle"}
    if first:
        out["first_mismatch"] = first
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    spec = spec_of(args)
    contigs, genome = make_genome(args, spec)
    fa, _, how = ensure_index(args, spec, contigs, use_gpu=True)
    groups = make_reads(spec, genome, 0)
    bases, off, _ = flatten(groups)
    threads = os.cpu_count() or 1
    pick = sample_indices(spec["reads"], spec["sample"])
    sample = [bases[int(off[i]):int(off[i + 1])].tobytes().decode() for i in pick]
    import oracle_lib as O
    oi = O.OracleIndex.load(fa)
    times = []
    for it in range(args.warmup + args.steps):
        _, res = oracle_run(fa, sample, threads, spec, oi)
        if it >= args.warmup:
            times.append(res["elapsed"])
    n_sample = len(sample)
    v = n_sample * len(times) / sum(times)
    sample_desc = "%d reads of the workload (evenly spaced) per step, %d threads sharing one index, output discarded" % (n_sample, threads)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1000.0 * sum(times) / len(times), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64",
        "data": "synthetic", "config": dict(workload=spec["name"], **dict(workload(spec), reads_per_step=n_sample, reads_in_workload=spec["reads"],
                                                                          note="C++ oracle of the reference path on the host cores, a bounded sample of the "
                                                                               "workload per step (the JVM reference cannot run here)")),
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

    spec = spec_of(args)
    log("config %d: drawing the genome" % args.config)
    contigs, genome = make_genome(args, spec)
    t_build, how = 0.0, "cached"
    if rank == 0:
        log("genome ready; FASTA + index files")
        fa, t_build, how = ensure_index(args, spec, contigs, use_gpu=True, device=local)
        log("index files ready (%s, %.1f s)" % (how, t_build))
    barrier()
    fa = os.path.join(cache_dir(args, spec), "genome.fa")
    t_open = time.time()
    idx = gw.FMIndexOnGenome.load(fa, local)
    t_open = time.time() - t_open
    log("index open on the device in %.1f s (locate table step %d)" % (t_open, idx.locateStep()))
    if args.locate_step is not None:
        idx.locateTable(args.locate_step)
        log("locate table step %d" % idx.locateStep())
    groups = make_reads(spec, genome, rank)
    log("reads drawn")
    index_gb = (genome.size / 128 * 64 * 2 + genome.size / 8 + genome.size * 3 / 8 + (genome.size * 8 / idx.locateStep() if idx.locateStep() < 32 else 0)) / 1e9
    del genome, contigs
    bases, offsets, m_max = flatten(groups)
    n = len(offsets) - 1
    total_bases = int(offsets[-1])
    bsf = gw.BidirectionalSuffixFilter(idx, gw.AlignmentConfig(k=spec["k"], reportType=spec["report"], extended=spec["extended"]))
    for opt in args.opt:
        kname, v = opt.split("=")
        bsf.set_option(kname, int(v))
    lib = L.lib()
    dev = torch.device("cuda", local)

    # ---- device-resident arm
    d_bases = torch.from_numpy(bases).to(dev)
    d_off = torch.from_numpy(offsets.view(np.int64)).to(dev)
    hits_cap, cig_cap = (2 * n + 16, 16 * n + 64) if spec["report"] == "besthit" else (40 * n + 16, 160 * n + 64)
    d_hits = torch.empty(hits_cap * 44, dtype=torch.uint8, device=dev)
    d_cig = torch.empty(cig_cap, dtype=torch.int32, device=dev)
    nh, nc = C.c_uint64(0), C.c_uint64(0)
    cnt = L.Counters()
    ms = (C.c_float * 6)()

    def step_device():
        return lib.gw_align_batch_device(bsf._h, d_bases.data_ptr(), d_off.data_ptr(), n, total_bases, m_max, d_hits.data_ptr(), hits_cap,
                                         d_cig.data_ptr(), cig_cap, C.byref(nh), C.byref(nc), C.byref(cnt))

    rc = step_device()   # sizes the result buffers (all-hits configs report a data-dependent number of records)
    if rc == L.GW_E_CAPACITY:
        hits_cap, cig_cap = int(nh.value) + 16, int(nc.value) + 64
        d_hits = torch.empty(hits_cap * 44, dtype=torch.uint8, device=dev)
        d_cig = torch.empty(cig_cap, dtype=torch.int32, device=dev)
        rc = step_device()
    L.check(rc)
    log("first device-resident batch done")
    for _ in range(max(0, args.warmup - 1)):
        L.check(step_device())
    fam = np.zeros(6)
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        L.check(step_device())
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
    del d_hits, d_cig

    # ---- end-to-end arm: pinned host buffers through gw_align_batch
    h_bases = torch.from_numpy(bases).pin_memory()
    h_off = torch.from_numpy(offsets.view(np.int64)).pin_memory()
    h_hits = torch.empty(hits_cap * 44, dtype=torch.uint8).pin_memory()
    h_cig = torch.empty(cig_cap, dtype=torch.int32).pin_memory()

    def step_host():
        a, b = C.c_uint64(hits_cap), C.c_uint64(cig_cap)
        L.check(lib.gw_align_batch(bsf._h, h_bases.data_ptr(), h_off.data_ptr(), n, h_hits.data_ptr(), C.byref(a), h_cig.data_ptr(), C.byref(b), None))
        return a.value, b.value

    log("device-resident arm timed: %.1f ms per step" % (1000.0 * dev_time / args.steps))
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
        # for reads that need a still larger work area), so their lookups are charged to the same CUDA-event time: 64 B per occurrence-table
        # lookup actually done (search: 2 per bidirectional step; locate: 1 per LF step, none with the dense locate table) + 4 B per locate.
        # `reference_algorithm_bytes` is the same work priced as SURVEY.md 8d prices the REFERENCE's algorithm (a locate = 64 x E[s] + 4 B
        # with E[s] = 15.5 LF steps of the .ssa walk, whatever the device did instead).
        locates = counters["sw_calls"]
        b_search = (counters["occ_lookups_search"] + counters["occ_lookups_locate"]) * 64.0 + locates * 4.0
        b_refalg = counters["occ_lookups_search"] * 64.0 + locates * (64.0 * 15.5 + 4.0)
        fams = {"scan": (counters["occ_lookups_scan"] * 64.0, fam_ms[0]), "search+locate+verify": (b_search, fam_ms[1] + fam_ms[2] + fam_ms[3])}
        roof_all = {k: {"bytes": b, "ms": float(t), "gbs": (b / (t * 1e-3) / 1e9 if t > 0 else None)} for k, (b, t) in fams.items()}
        t_flow = fams["search+locate+verify"][1]
        roof_all["search+locate+verify"].update(occ_lookups_search=counters["occ_lookups_search"], occ_lookups_locate=counters["occ_lookups_locate"],
                                                locates=locates, fm_searches=counters["fm_searches"], reference_algorithm_bytes=b_refalg,
                                                reference_algorithm_gbs=(b_refalg / (t_flow * 1e-3) / 1e9 if t_flow > 0 else None))
        roof_all["emit"] = {"ms": float(fam_ms[4])}
        dom = max(fams, key=lambda k: fams[k][1])
        ach = roof_all[dom]["gbs"] or 0.0
        # the random-access bounds, measured in this job: independent 64 B gathers over the occurrence table of this index (HBM- or
        # L2-resident, whatever the index is)
        gbs, gms = C.c_double(0), C.c_double(0)
        gather = None
        if lib.gw_gather_benchmark(idx._h, 1, 1 << 28, 3, C.byref(gbs), C.byref(gms)) == 0:
            gather = float(gbs.value)
        traffic, traffic_at = None, None
        try:   # DRAM bytes per step of the dominant kernel family from the committed ncu --set full capture (profiles/)
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            if tj.get("reads") == n and tj.get("genome_bp") == spec["genome_bp"]:
                traffic, traffic_at = tj["dram_bytes_per_step"], tj.get("git_head")   # a capture, not measured in this run: it says of which build
        except Exception:  # noqa: BLE001
            pass
        value = world * n * args.steps / dev_time
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1000.0 * dev_time / args.steps, "device_event_ms_per_step": float(fam_ms[5]), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32/u64 integer", "data": "synthetic",
            "config": dict(workload=spec["name"], **workload(spec),
                           l2="working set (index %.2f GB + per-read work areas of several GB rewritten every step) exceeds the 126 MB L2; no explicit flush" % index_gb,
                           parallelism="reads sharded over %d GPU(s), index replicated, no collective" % world, index_built_by=how,
                           index_build_s=round(t_build, 2), index_open_s=round(t_open, 2), locate_table_step=idx.locateStep()),
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "traffic_captured_at_git_head": traffic_at,
                         "peak_source": peak_src, "random_gather_gbs": gather, "frac_of_random_gather": (ach / gather if gather else None),
                         "note": "algorithmic bytes = occurrence-table lookups the kernel did x 64 B (+4 B per locate) over the CUDA-event time of the "
                                 "kernel family; with the dense locate table a locate is one 4-byte gather, so the lookups of the reference's "
                                 ".ssa walks are not done at all (families.*.reference_algorithm_* prices them as SURVEY.md 8d does); "
                                 "random_gather_gbs = independent 64 B gathers over the same table measured in this job; traffic = DRAM bytes "
                                 "of the family per step from the ncu capture in profiles/",
                         "families": roof_all},
            "e2e": {"value": world * n * args.steps / e2e_time, "unit": UNIT, "h2d_bytes_per_step": int(total_bases + (n + 1) * 8),
                    "d2h_bytes_per_step": int(e_nh * 44 + e_nc * 4), "ms_per_step": 1000.0 * e2e_time / args.steps},
            "gpu_launches": int(counters["gpu_launches"]) * args.steps, "counters_per_step": counters,
            "clocks": sampler.summary(),
        }
        parity_bad = False
        if not args.no_cpu:
            threads = os.cpu_count() or 1
            n_s = spec["sample"] if world == 1 else min(spec["sample"], 20_000)
            pick = sample_indices(n, n_s)
            seqs = [bases[int(offsets[i]):int(offsets[i + 1])].tobytes().decode() for i in pick]
            import oracle_lib as O
            log("end-to-end arm timed; opening the oracle's index")
            t_oo = time.time()
            oi = O.OracleIndex.load(fa)
            t_oo = time.time() - t_oo
            log("oracle index open in %.1f s; aligning %d sample reads on %d threads" % (t_oo, len(seqs), threads))
            v, res = oracle_run(fa, seqs, threads, spec, oi)
            log("oracle sample done")
            if world == 1:
                v1, _ = oracle_run(fa, seqs[: max(1, len(seqs) // max(threads, 1))], 1, spec, oi)
                line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                        "sample": "%d reads of the workload (evenly spaced), %d threads sharing one index (C++ oracle; JVM reference "
                                                  "unavailable on this box); oracle index open %.0f s" % (len(seqs), threads, t_oo),
                                        "single_thread_value": v1}
            oi.close()
            line["parity"] = parity_check(lib, L, idx, bsf, bases, offsets, pick, h_hits, e_nh, h_cig, res)
            parity_bad = line["parity"]["mismatches"] != 0 or line["parity"]["counter_mismatches"] != 0
        print(json.dumps(line))
        sys.stdout.flush()
        if parity_bad:
            sys.stderr.write("bench.py: the GPU output differs from the oracle on the checked sample\n")
            bsf.close()
            idx.close()
            os._exit(3)
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
    ap.add_argument("--config", type=int, default=2, choices=sorted(CONFIGS), help="BASELINE.json config (1-based); 2 is the metric's config")
    ap.add_argument("--genome-bp", type=int, default=0, help="override the genome size of the config")
    ap.add_argument("--reads", type=int, default=0, help="override the reads per GPU and step")
    ap.add_argument("--ref-sample", type=int, default=None, help="reads checked against / timed on the oracle (default: per config)")
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

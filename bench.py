#!/usr/bin/env python
"""Benchmark of the approximate-matching hot path (see BASELINE.json).

  python bench.py --gpus N --steps K --warmup W [--impl reference]

Workload (config.workload): BASELINE.json configs[2] -- synthetic 3.1 Gbp
GRCh38-sized reference with injected repeats (24 sequences; SINE-, LINE-like,
1 kbp families, segmental duplications, microsatellites, tandem arrays:
SURVEY.md section 8d), single-end 150 bp reads, edit distance k=4, ALL mode,
default `columba` strategy (dynamic scheme selection), dynamic partitioning.
The reference text and its index (the reference's own file set + the .gidx
sidecar of the device layout) are produced once per box by the standalone
`columba_b200/bin/clb_gpu_build` executable (not the product library; about
15 s for 3.1 Gbp) and cached under $TMPDIR.  One step = one pass of the whole
path over one batch of `--reads` reads; with N GPUs every rank maps its own
batches (weak scaling, index replicated in every GPU's HBM, no data-path
collective; the 17 counters are reduced with one NCCL all-reduce per run).

The ours-arm prints `value` (reads resident in HBM, occurrences left in HBM),
`e2e` (host buffers through clb_map_batch: H2D of the reads and D2H of the
occurrences inside the timed region), a roofline object for the dominant
kernel, the path's own unit (rank lookups/s, 32 B each) against 8 TB/s, the
measured HBM copy peak and a measured random-32-B-gather ceiling, and a
cpu_baseline object (the unmodified reference's multithreaded CPU build on a
bounded sample, rank 0, N=1 only).  `--impl reference` runs only the
reference's CPU build (no product code is imported on that arm).
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
sys.path.insert(0, os.path.join(ROOT, "tools"))

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "columba")
BUILDER = os.path.join(ROOT, "columba_b200", "bin", "clb_gpu_build")
MAPPER = os.path.join(ROOT, "columba_b200", "bin", "clb_map")
METRIC_NAME = "reads/sec (150bp, k=4 edit, ALL) on 3.1 Gbp ref"
METRIC_NAME_C5 = "reads/sec (250bp, k=5 Hamming, ALL, custom pigeon scheme) on 3.1 Gbp ref"


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


# --------------------------------------------------------------------------- data
def data_dir(args):
    d = os.path.join(tempfile.gettempdir(), "clb_bench_%s_%d_%d" % (args.profile, args.ref_bp, args.ref_seed))
    os.makedirs(d, exist_ok=True)
    return d


def make_reference(args, device=0):
    """Synthetic reference + index files (the reference's formats and the .gidx sidecar), written by the
    standalone builder executable.  Returns the index basename."""
    d = data_dir(args)
    base = os.path.join(d, "ref")
    done = base + ".done"
    if not os.path.exists(done):
        if not os.path.exists(BUILDER):
            raise SystemExit("missing %s: run `python -c 'import __graft_entry__ as g; g.build()'`" % BUILDER)
        cmd = [BUILDER, "--synthetic", args.profile, "--bp", str(args.ref_bp), "--seed", str(args.ref_seed),
               "--out", base, "--device", str(device), "--quiet"]
        t0 = time.perf_counter()
        subprocess.check_call(cmd)
        with open(done, "w") as f:
            f.write("%.1f\n" % (time.perf_counter() - t0))
    return base


class RefText:
    """The concatenated reference text (memory-mapped .txt.bin) and the sequence starts (.pos)."""

    def __init__(self, base):
        self.n = int(np.fromfile(base + ".txt.bin", dtype=np.uint32, count=1)[0])
        self.cat = np.memmap(base + ".txt.bin", dtype=np.uint8, mode="r", offset=4, shape=(self.n,))
        self.starts = np.fromfile(base + ".pos", dtype=np.uint32).astype(np.int64)


def make_reads(args, ref, rank, batch):
    import gen_data
    return gen_data.gen_reads_cat(ref.cat, ref.starts, args.reads, args.read_len,
                                  (250 if args.config == 5 else args.read_seed) + 1000 * batch + rank, args.metric,
                                  args.k, over_e=args.k + 2)


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu = gpu
        self.proc = None
        self.path = None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile(prefix="clocks_", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=f,
                                         stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not self.proc:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            a = [x.strip() for x in line.split(",")]
            if len(a) < 9:
                continue
            try:
                sm.append(float(a[1]))
                mx.append(float(a[2]))
            except ValueError:
                continue
            for nm, v in zip(names, a[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.path)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons),
                       samples=len(sm))
        return out


# --------------------------------------------------------------------------- reference arm
def write_fastq_matrix(path, reads, first=0):
    n, L = reads.shape
    qual = b"I" * L
    with open(path, "wb") as f:
        for c0 in range(0, n, 1 << 16):
            blk = reads[c0:c0 + (1 << 16)]
            f.write(b"".join(b"@r%d\n%s\n+\n%s\n" % (first + c0 + i, blk[i].tobytes(), qual)
                             for i in range(len(blk))))


REF_EXTRA = []  # extra command-line arguments of the reference for the selected config


def run_reference_cpu(base, reads, k, threads, out_dir, tag="s"):
    """The reference's own CPU implementation (oracle/_ref/columba, compiled from the unmodified sources) on
    `reads`; returns (reads/s, mapping seconds from its own log, kind)."""
    n = len(reads)
    if os.path.exists(REF_BIN):
        fq = os.path.join(out_dir, "sample_%s_%d_%d.fq" % (tag, n, reads.shape[1]))
        if not os.path.exists(fq):
            write_fastq_matrix(fq, reads)
        shm = "/dev/shm" if os.path.isdir("/dev/shm") else out_dir
        sam = os.path.join(shm, "clb_ref_out_%d.sam" % os.getpid())
        cmd = [REF_BIN, "-r", base, "-f", fq, "-a", "all", "-e", str(k), "-t", str(threads), "-o", sam] + REF_EXTRA
        p = subprocess.run(cmd, capture_output=True, text=True)
        if os.path.exists(sam):
            os.unlink(sam)
        if p.returncode != 0:
            raise RuntimeError("reference failed: " + p.stdout[-500:] + p.stderr[-500:])
        m = re.search(r"(\d+) mapping threads finished mapping in ([0-9.]+)s", p.stdout)
        secs = float(m.group(2))
        return n / secs, secs, "reference"
    # the reference binary did not travel: time the oracle port (1 thread)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_util as ou
    orc = ou.Oracle(base)
    t0 = time.perf_counter()
    for i in range(n):
        orc.map_read(reads[i].tobytes().decode(), k)
    secs = time.perf_counter() - t0
    return n / secs, secs, "port"


def cpu_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def sized_cpu_sample(base, reads, k, threads, out_dir, target_s, max_n):
    """Pick a sample size that gives about target_s seconds of mapping: probe with 5 000 reads first."""
    n0 = min(5000, len(reads))
    rate, _, kind = run_reference_cpu(base, reads[:n0], k, threads, out_dir, "probe")  # also warms the page cache
    n = int(min(max_n, len(reads), max(n0, rate * target_s)))
    return n, kind


# --------------------------------------------------------------------------- main
def progress(msg):
    """Phase marks on stderr (stdout carries the one JSON line)."""
    mem = ""
    try:
        import torch
        if torch.cuda.is_available() and torch.cuda.is_initialized():
            free, total = torch.cuda.mem_get_info()
            mem = " [HBM in use %.1f GB]" % ((total - free) / 1e9)
    except Exception:  # noqa: BLE001
        pass
    sys.stderr.write("[bench %7.1f s] %s%s\n" % (time.perf_counter() - T_START, msg, mem))
    sys.stderr.flush()


T_START = time.perf_counter()


def main():
    if os.environ.get("BENCH_TRACE_AFTER"):  # dump the Python stacks if the run is still going after that many seconds
        import faulthandler
        faulthandler.dump_traceback_later(float(os.environ["BENCH_TRACE_AFTER"]), exit=False)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--reads", type=int, default=1_000_000, help="reads per step per GPU")
    ap.add_argument("--ref-bp", type=int, default=3_100_000_000)
    ap.add_argument("--profile", default="human", choices=["human", "ecoli", "none"])
    ap.add_argument("--ref-seed", type=int, default=3100)
    ap.add_argument("--read-seed", type=int, default=150)
    ap.add_argument("--k", type=int, default=None, help="maximal distance (default: 4, or 5 with --config 5)")
    ap.add_argument("--config", type=int, default=3, choices=[3, 5],
                    help="BASELINE.json configs[] entry (1-based as SURVEY.md counts them): 3 = the metric's workload "
                         "(150 bp, edit distance k=4, ALL); 5 = 250 bp reads, Hamming distance k=5, custom pigeon scheme "
                         "of the reference's search_schemes/ (or --scheme-dir), dynamic partitioning")
    ap.add_argument("--scheme-dir", default=None, help="config 5: search scheme directory (-c DIR of the reference)")
    ap.add_argument("--batches", type=int, default=2, help="distinct read batches the steps rotate through")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="mapping time the cpu_baseline sample is sized for")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e-sam", action="store_true", help="skip the FASTQ-to-SAM leg (batch driver clb_map)")
    ap.add_argument("--no-numa-bind", action="store_true", help="do not bind the process to the GPU's local CPUs")
    ap.add_argument("--host-threads", type=int, default=2, help="host threads (one context each) of the e2e leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    hamming = args.config == 5
    if args.k is None:
        args.k = 5 if hamming else 4
    args.read_len = 250 if hamming else 150
    args.metric = "hamming" if hamming else "edit"
    if hamming and not args.scheme_dir:
        args.scheme_dir = os.path.join(ROOT, "tests", "golden", "schemes", "ref", "pigeon")
    if hamming and not args.scheme_dir.endswith("/"):
        args.scheme_dir += "/"

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if hamming:
        REF_EXTRA.extend(["-m", "hamming", "-c", args.scheme_dir, "-p", "dynamic"])
    config = {"workload": ("configs[4]: synthetic %.2f Gbp GRCh38-sized reference (%s repeat profile, seed %d), %d SE "
                           "250 bp reads per step per GPU, Hamming distance k=%d, ALL mode, custom scheme %s, "
                           "dynamic partitioning" % (args.ref_bp / 1e9, args.profile, args.ref_seed, args.reads, args.k,
                                                     os.path.basename(args.scheme_dir.rstrip("/")))) if hamming else
                          ("configs[2]: synthetic %.2f Gbp GRCh38-sized reference (%s repeat profile, seed %d), %d SE "
                           "150 bp reads per step per GPU, edit distance k=%d, ALL mode, dynamic scheme selection, "
                           "dynamic partitioning" % (args.ref_bp / 1e9, args.profile, args.ref_seed, args.reads, args.k)),
              "ref_bp": args.ref_bp, "repeat_profile": args.profile, "reads_per_step_per_gpu": args.reads,
              "read_len": args.read_len, "k": args.k, "mode": "ALL",
              "scheme": ("custom (dynamic selection): " + args.scheme_dir) if hamming else "columba (dynamic selection)",
              "distance": args.metric,
              "partitioning": "dynamic", "in_text_switch": 4, "sa_sparseness": 4, "kmer": 10,
              "l2": "the index (about 2.5 B per reference base in HBM) exceeds the 126 MB L2 for any ref_bp above "
                    "50 Mbp; inputs per step (150 B per read) exceed it from 840 k reads on",
              "parallelism": "reads sharded over %d GPU(s), index replicated" % args.gpus}

    if args.impl == "reference":
        if rank != 0:
            return
        base = make_reference(args, 0)
        ref = RefText(base)
        threads = cpu_threads()
        reads = make_reads(args, ref, 0, 0)
        # bounded sample: the whole --steps K --warmup W run is sized for about two minutes of mapping
        per_step = max(2.0, min(args.cpu_seconds, 120.0 / max(1, args.steps + min(args.warmup, 1))))
        n, kind = sized_cpu_sample(base, reads, args.k, threads, data_dir(args), per_step, len(reads))
        reads = reads[:n]
        for _ in range(min(args.warmup, 1)):  # the probe above already warmed the page cache
            run_reference_cpu(base, reads, args.k, threads, data_dir(args))
        secs = []
        for _ in range(args.steps):
            _, s, kind = run_reference_cpu(base, reads, args.k, threads, data_dir(args))
            secs.append(s)
        ms = 1e3 * float(np.mean(secs))
        val = n / (ms / 1e3)
        sample = "first %d reads of the first batch per step, all %d host threads (%s), mapping time from the " \
                 "reference's own log line (it includes FASTQ parsing and SAM formatting/writing to /dev/shm, which " \
                 "the GPU arm's e2e does not)" % (n, threads, cpu_model())
        line = {"impl": "reference", "metric": METRIC_NAME_C5 if hamming else METRIC_NAME, "value": val, "unit": "reads/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": val, "unit": "reads/s", "cores": threads if kind == "reference" else 1,
                                 "kind": kind, "sample": sample, "cpu_model": cpu_model()},
                "e2e": {"value": val, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    import threading as _th

    import torch
    import torch.distributed as dist
    import columba_b200 as cb
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the product has no CPU fallback)")
    torch.cuda.set_device(local)
    from columba_b200.parallel import bind_to_gpu_numa_node, reduce_counters
    numa_cpus = None if args.no_numa_bind else bind_to_gpu_numa_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- data: rank 0 writes the index files, everybody loads them
    t_build = time.perf_counter()
    if rank == 0:
        base = make_reference(args, local)
    barrier()
    base = make_reference(args, local)
    t_build = time.perf_counter() - t_build
    progress("reference + index files ready")
    ref = RefText(base)
    t_load = time.perf_counter()
    ix = cb.Index(base, 4, 10, 4, device=local)
    t_load = time.perf_counter() - t_load
    progress("index in HBM")
    nb = max(1, args.batches)
    batches = [make_reads(args, ref, rank, b) for b in range(nb)]
    n, L = batches[0].shape
    progress("read batches generated")
    off = (np.arange(n + 1, dtype=np.int64) * L).astype(np.uint32)

    mapper_args = ((("custom_dynamic", args.scheme_dir.rstrip("/")), "hamming", "dynamic") if hamming else ())
    m = cb.Mapper(ix, *mapper_args)
    stream = torch.cuda.current_stream()
    m.set_stream(stream.cuda_stream)
    h_reads = [torch.from_numpy(b.reshape(-1)).pin_memory() for b in batches]
    h_off = torch.from_numpy(off.view(np.int32)).pin_memory()
    d_reads = [h.cuda(non_blocking=True) for h in h_reads]
    d_off = h_off.cuda(non_blocking=True)
    torch.cuda.synchronize()
    step_no = [0]

    def step_device():
        b = step_no[0] % nb
        step_no[0] += 1
        return m.map_batch_device(d_reads[b].data_ptr(), d_off.data_ptr(), n, L, args.k, 0, fetch=False)

    def timed(fn, steps, warmup):
        for w in range(warmup):
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            progress("warm-up step %d: %.2f s" % (w, time.perf_counter() - t0))
        step_no[0] = 0
        barrier()
        clk = ClockSampler(local)
        clk.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        results = [fn() for _ in range(steps)]
        e1.record(stream)
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        clocks = clk.stop()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), wall, results, clocks

    # ---- value: inputs resident in HBM
    ms, wall, results, clocks = timed(step_device, args.steps, args.warmup)
    ms_per_step = ms / args.steps
    progress("device-resident leg done: %.1f ms per step" % ms_per_step)
    value = world * n * args.steps / (ms / 1e3)
    launches = int(sum(r.gpu_launches for r in results))
    stage = {s: float(np.mean([r.stage_ms[s] for r in results])) for s in cb.STAGE_NAMES}
    stats = {s: float(np.mean([r.stats[s] for r in results])) for s in cb.STAT_NAMES}
    counters = {s: int(sum(r.counters[s] for r in results)) for s in cb.COUNTER_NAMES}
    # the path's one collective: counters of all ranks (NCCL all-reduce over NVLink)
    counters_all = reduce_counters(counters)[0] if world > 1 else dict(counters)

    # ---- e2e: host buffers through the C ABI, copies inside the timed region.  As in the reference, where
    # every worker thread maps its own chunks (src/parallel.cpp:88-115), the host side is `--host-threads`
    # threads (default 2) with one context each, taking the steps' batches in turn: the copies of one
    # context overlap the kernels of the other.  Every step still moves its reads host->device and its
    # occurrences device->host inside the timed region.
    n_thr = max(1, min(args.host_threads, args.steps))
    mappers = [m] + [cb.Mapper(ix, *mapper_args) for _ in range(n_thr - 1)]
    m.set_stream(None)  # every context on its own stream

    def host_call(mp, b):
        r = cb.api.ClbResult()
        rc = mp.lib.clb_map_batch(mp.h, h_reads[b].data_ptr(), h_off.data_ptr(), n, args.k, 0, cb.api.C.byref(r))
        if rc:
            raise cb.ClbError(rc, mp.lib.clb_last_error(mp.h).decode())
        return (r.n_occ, r.cigar_stride, int(r.gpu_launches))

    def timed_host(steps, warmup):
        for mp in mappers:
            for w in range(warmup):
                host_call(mp, w % nb)
        barrier()
        clk = ClockSampler(local)
        clk.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        out = [None] * steps
        errs = []

        def work(tid):
            try:
                for s_ in range(tid, steps, n_thr):
                    out[s_] = host_call(mappers[tid], s_ % nb)
            except Exception as e:  # noqa: BLE001
                errs.append(e)
        e0.record(stream)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ths = [_th.Thread(target=work, args=(i,)) for i in range(n_thr)]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        e1.record(stream)
        torch.cuda.synchronize()
        clocks = clk.stop()
        if errs:
            raise errs[0]
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), wall, out, clocks

    try:
        ms2, wall2, results2, clocks2 = timed_host(args.steps, 2)
    except cb.ClbError as e:
        if "out of memory" not in str(e) or n_thr == 1:
            raise
        # a second context does not fit beside the first (every context owns its lists and scratch): one host thread
        progress("end-to-end leg: second context out of memory, falling back to one host thread")
        for mp in mappers[1:]:
            mp.close()
        mappers[:] = mappers[:1]
        n_thr = 1
        ms2, wall2, results2, clocks2 = timed_host(args.steps, 2)
    progress("end-to-end leg done: %.1f ms per step" % (ms2 / args.steps))
    n_occ_mean = float(np.mean([r[0] for r in results2]))
    cig_stride = results2[-1][1]
    h2d = int(h_reads[0].numel() + h_off.numel() * 4)
    d2h = int(n_occ_mean * (24 + 2 * cig_stride) + 8 * n)
    e2e_val = world * n * args.steps / (ms2 / 1e3)

    # ---- roofline of the dominant kernel (CUDA events around that stage's launches)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    dom = max(stage, key=lambda s: stage[s])
    part_lookups = stats["RANK_LOOKUPS"] - stats["SEARCH_LOOKUPS"]
    # algorithmic bytes per launch: 32 B per rank lookup; a search node = 2 lookups + its
    # 64-byte frontier record written once and read once; locate = 32 B BWT record + 16 B marker record per LF
    # step + 16 B marker + 4 B sample per row; verification reads 2-bit text (1 B per 4 rows) and the read's
    # bit planes (96 B); the filter reads/writes its 24 B records + CIGARs
    alg = {
        "prep": n * 2 * (L + L + 2 * 4 * 4 * 8),
        "partition": part_lookups * 32 + n * 2 * (L + 2 * 5 * 20),
        "search": stats["SEARCH_LOOKUPS"] * 32 + stats["SEARCH_ROWS"] * 128,
        "verify": stats["CANDIDATES"] * (24 + 20 + 96) + stats["LF_STEPS"] * 48 + stats["VERIFY_ROWS"] / 4.0
        + stats["TEXT_OCCS"] * 44,
        "fm_expand": stats["FM_OCCS"] * 24 * 3,
        "text_filter": stats["TEXT_OCCS"] * (24 + 8) * 2 + counters["TOTAL_UNIQUE_MATCHES"] / args.steps * 44,
    }
    achieved = alg[dom] / (stage[dom] / 1e3) / 1e9 if stage[dom] > 0 else 0.0
    # dram__bytes_read.sum + dram__bytes_write.sum per launch of the stage's kernels, from the committed
    # `ncu --set full` capture of this same command (profiles/r2_traffic.json; null when there is none)
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
        if tj.get("reads_per_step") == n and tj.get("ref_bp") == args.ref_bp:
            traffic = tj["stage_dram_bytes"].get(dom)
    except (OSError, ValueError, KeyError):
        pass
    gather_rate, gather_gbs = m.gather_ceiling(1 << 28)
    lookups_per_s = stats["RANK_LOOKUPS"] / (ms_per_step / 1e3)
    roofline = {"kernel": dom, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "kernel_ms": stage[dom], "algorithmic_bytes": alg[dom],
                "rank_lookups_per_s": lookups_per_s,
                "rank_GBps": lookups_per_s * 32 / 1e9,
                "rank_roofline_frac_of_8TBs": lookups_per_s * 32 / 8e12,
                "rank_frac_of_measured_copy_peak": lookups_per_s * 32 / 1e9 / peak,
                "random_32B_gather_ceiling_GBps": gather_gbs,
                "random_32B_gather_ceiling_lookups_per_s": gather_rate,
                "rank_frac_of_gather_ceiling": lookups_per_s / gather_rate if gather_rate else None,
                "search_stage_lookups_per_s": stats["SEARCH_LOOKUPS"] / (stage["search"] / 1e3) if stage["search"] > 0 else None}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    R_total = world * n * args.steps
    line = {"metric": METRIC_NAME_C5 if hamming else METRIC_NAME, "value": value, "unit": "reads/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": config,
            "e2e": {"value": e2e_val, "unit": "reads/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms2 / args.steps, "host_threads": n_thr, "wall_ms_per_step": wall2 / args.steps,
                    "host_cpus_bound_to_gpu_node": len(numa_cpus) if numa_cpus else 0,
                    "ends_at": "binary occurrence arrays (position, strand, distance, CIGAR runs) in host memory; "
                               "no FASTQ parsing or SAM text formatting"},
            "gpu_launches": launches, "clocks": clocks, "clocks_e2e": clocks2, "roofline": roofline,
            "stage_ms": stage, "stats_per_step": stats,
            "counters_all_ranks": counters_all,
            "searches_started_per_read": counters_all["SEARCH_STARTED"] / R_total,
            "nodes_per_read": counters_all["NODE_COUNTER"] / R_total,
            "matches_per_read": counters_all["TOTAL_UNIQUE_MATCHES"] / R_total,
            "wall_ms_per_step": wall / args.steps, "index_device_bytes": ix.device_bytes,
            "index_build_or_cache_s": t_build, "index_load_s": t_load, "distinct_batches": nb}
    # the contexts of this process own tens of GB of lists and scratch: release them before other processes map
    for mp in mappers:
        mp.close()
    ix.close()
    torch.cuda.empty_cache()
    if world == 1 and not args.no_e2e_sam and not hamming and os.path.exists(MAPPER):
        # the same timer as the reference's: FASTQ parsing, mapping, SAM formatting and writing (to /dev/shm), index
        # load excluded, no warm-up (the first batch of every context grows its lists)
        try:
            d = data_dir(args)
            fq = os.path.join(d, "e2e_sam_%d.fq" % n)
            if not os.path.exists(fq):
                write_fastq_matrix(fq, batches[0])
            shm = "/dev/shm" if os.path.isdir("/dev/shm") else d
            sam = os.path.join(shm, "clb_map_out_%d.sam" % os.getpid())
            p = subprocess.run([MAPPER, "-r", base, "-f", fq, "-o", sam, "-e", str(args.k), "-t", str(max(2, n_thr)),
                                "-b", "262144", "-d", str(local), "-D", cb.SCHEMES_DIR], capture_output=True, text=True)
            mt = re.search(r"(\d+) mapping threads finished mapping in ([0-9.]+)s", p.stderr)
            sam_bytes = os.path.getsize(sam) if os.path.exists(sam) else 0
            if os.path.exists(sam):
                os.unlink(sam)
            if p.returncode == 0 and mt:
                mf = re.search(r"formatting threads per mapping thread (\d+)", p.stderr)
                line["e2e_sam"] = {"value": n / float(mt.group(2)), "unit": "reads/s", "seconds": float(mt.group(2)),
                                   "reads": n, "sam_bytes": sam_bytes, "host_threads": int(mt.group(1)),
                                   "format_threads_per_host_thread": int(mf.group(1)) if mf else 1,
                                   "includes": "FASTQ parsing, H2D, kernels, D2H, SAM formatting and writing to /dev/shm by "
                                               "the batch driver columba_b200/bin/clb_map (one context per host thread, "
                                               "batches of 262144 reads, a batch's records formatted and written by "
                                               "several threads, no warm-up); index load excluded, as in the "
                                               "reference's 'mapping threads finished mapping in' timer"}
            else:
                line["e2e_sam"] = {"value": None, "error": (p.stderr or "")[-200:]}
        except Exception as e:  # informative leg
            line["e2e_sam"] = {"value": None, "error": str(e)[:200]}
        progress("FASTQ-to-SAM leg done")
    if world == 1 and not args.no_cpu_baseline:
        threads = cpu_threads()
        try:
            ns, kind = sized_cpu_sample(base, batches[0], args.k, threads, data_dir(args), args.cpu_seconds, n)
            v, secs, kind = run_reference_cpu(base, batches[0][:ns], args.k, threads, data_dir(args))
            cb_line = {"value": v, "unit": "reads/s", "cores": threads if kind == "reference" else 1, "kind": kind,
                       "seconds": secs, "cpu_model": cpu_model(),
                       "sample": "first %d reads of the first batch, %s" % (
                           ns, "unmodified reference binary (oracle/_ref/columba -t %d), mapping time from its own "
                               "log (includes FASTQ parsing and SAM writing)" % threads
                           if kind == "reference" else "oracle port, 1 thread")}
            if kind == "reference" and threads > 1:
                n1 = max(1000, ns // (2 * threads))
                v1, s1, _ = run_reference_cpu(base, batches[0][:n1], args.k, 1, data_dir(args), "one")
                cb_line["one_thread"] = {"value": v1, "seconds": s1, "reads": n1}
            line["cpu_baseline"] = cb_line
        except Exception as e:  # the baseline is informative; never lose the GPU numbers over it
            line["cpu_baseline"] = {"value": None, "unit": "reads/s", "cores": 0, "kind": "reference",
                                    "sample": "failed: %s" % str(e)[:200]}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

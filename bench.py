#!/usr/bin/env python
"""bench.py — headline benchmark of the PEnG-motif hot path on B200.

Metric (BASELINE.json): k-mer count Gbases/s.  A "step" = one pass of BasePattern::count_patterns over one
resident synthetic sequence set: k-mer start marking + exact non-overlap counting into the 4^W table + fold
(+ mirror copy / + NCCL all-reduce of the histogram for N > 1).

  workload at N = 1 : BASELINE configs[2] — 1 M sequences x 200 bp, planted MafK-like motif, W = 10, both strands
  workload at N > 1 : the same shard shape PER GPU (weak scaling): disjoint shards of one global synthetic set,
                      per-GPU histograms + ltot combined with one NCCL all_reduce
  --workload cfg4   : BASELINE configs[3] — 50 M x 500 bp (25 Gbp) split over the N GPUs (strong scaling)

  value     whole-job Gbases/s with the packed set resident in HBM (device-timed, max over ranks)
  e2e       the same metric through the C-ABI calls a reference binding makes, from HOST buffers:
            peng_cuda_ingest_codes (pinned 1 B/base codes -> H2D -> pack) + peng_cuda_count + peng_cuda_get_counts
  roofline  the dominant kernel (k_count_global): algorithmic bytes / its CUDA-event duration vs measured HBM peak
  cpu_baseline  the reference's count_patterns (oracle/_ref, else the oracle port) on the host cores
  extra     EM PWM-iterations/s (17 PWMs, FP64 and float32-faithful), stats-pass time, ingest rate

`--impl reference` times the reference's own CPU implementation on the same config / metric.
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MOTIF = "TGCTGASTCAGCA"  # the reference's own MafK result, S = C|G
ALGO_BYTES_PER_BASE = 0.375  # 2-bit codes + 1-bit k-mer-start mask


def workload_shape(name, n_gpus):
    if name == "cfg3":
        return dict(n_seq_per_gpu=1_000_000, seq_len=200, W=10, strand="BOTH", scaling="weak",
                    label="configs[2]: synthetic 1M x 200 bp, planted MafK-like motif, W=10, both strands (per GPU)")
    if name == "cfg4":
        return dict(n_seq_per_gpu=50_000_000 // n_gpus, seq_len=500, W=10, strand="BOTH", scaling="strong",
                    label="configs[3]: synthetic 50M x 500 bp (25 Gbp) sharded over the GPUs, W=10, both strands")
    if name == "cfg5":
        return dict(n_seq_per_gpu=10_000_000 // n_gpus, seq_len=300, W=14, strand="PLUS", scaling="strong",
                    label="configs[4]: synthetic 10M x 300 bp, --strand PLUS, W=14")
    if name == "small":
        return dict(n_seq_per_gpu=20_000, seq_len=200, W=10, strand="BOTH", scaling="weak", label="smoke-size set")
    raise SystemExit(f"unknown workload {name}")


def bench_config(shape, world):
    """The `config` object of the JSON line — the same keys and values in both arms (`--impl b200` and `--impl reference`)."""
    return {"workload": shape["label"], "W": shape["W"], "strand": shape["strand"], "n_seq_per_gpu": shape["n_seq_per_gpu"],
            "seq_len": shape["seq_len"], "seed": 1234, "motif": MOTIF, "plant_fraction": 0.3, "n_gpus": world}


# ---------------------------------------------------------------------------------------------------------
# clocks sampling (B200_PROFILING.md "clocks DURING the timed region")
# ---------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                          str(self.index), "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                                         text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=3)
            except Exception:
                self.proc.kill()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        reasons = set()
        for r in self.rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------
# CPU arms
# ---------------------------------------------------------------------------------------------------------
def write_fasta_from_codes(codes, offs, path):
    import numpy as np

    lut = np.frombuffer(b"NACGT", np.uint8)
    letters = lut[codes]
    with open(path, "wb") as f:
        for s in range(len(offs) - 1):
            f.write(b">s%d\n" % s)
            f.write(letters[int(offs[s]):int(offs[s + 1])].tobytes())
            f.write(b"\n")


class CpuReference:
    """Times the reference's count_patterns on the host: oracle/_ref/ref_dump (the unmodified reference, timed
    around a second call of its own count_patterns) if it was built, else the oracle port."""

    def __init__(self, codes, offs, W, strand):
        self.codes, self.offs, self.W, self.strand = codes, offs, W, strand
        self.cores = os.cpu_count() or 1
        self.ref_dump = os.path.join(ROOT, "oracle", "_ref", "ref_dump")
        self.dir = None
        if os.path.exists(self.ref_dump):
            self.dir = tempfile.mkdtemp(prefix="peng_ref_")
            write_fasta_from_codes(codes, offs, os.path.join(self.dir, "sample.fa"))

    def run(self):
        """-> (seconds of count_patterns, kind, cores)"""
        if self.dir:
            env = dict(os.environ, OMP_NUM_THREADS=str(self.cores))
            r = subprocess.run([self.ref_dump, os.path.join(self.dir, "sample.fa"), str(self.W), self.strand, self.dir,
                                "--counts-only"], capture_output=True, text=True, env=env)
            if r.returncode == 0:
                meta = dict(l.split() for l in open(os.path.join(self.dir, "meta.txt")))
                return float(meta["t_count_patterns_s"]), "reference", self.cores
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import _oracle as O  # cpu_baseline leg: the one place bench.py executes oracle/

        t = time.perf_counter()
        O.count(self.codes, self.offs, self.W, self.strand == "BOTH")
        return time.perf_counter() - t, "port", 1

    def close(self):
        if self.dir:
            shutil.rmtree(self.dir, ignore_errors=True)


def sample_codes_host(shape, n_sample, seed=1234):
    """The first n_sample sequences of the synthetic workload, generated WITHOUT a GPU for the reference arm
    when none is present (kernel bodies compiled for the host, tests/sim) — else exported from the device."""
    import numpy as np

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _sim as S

    motif = ["ACGTS".index(c) for c in MOTIF]
    st = S.synth(n_sample, shape["seq_len"], seed, 0, motif, 0.3, 0.0)
    offs = (np.arange(n_sample + 1, dtype=np.uint64) * np.uint64(shape["seq_len"]))
    return S.unpack(st, offs), offs


def run_reference_arm(args, shape):
    import numpy as np

    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_sample = min(shape["n_seq_per_gpu"], args.ref_sample)
    # the reference arm never loads libpeng_cuda.so: the sample comes from the host build of the generator (tests/sim,
    # the same counter-based generator the device runs: identical sequences)
    codes, offs = sample_codes_host(shape, n_sample)
    bases = int(offs[-1])
    times, kind, cores = [], None, 1
    ref = CpuReference(codes, offs, shape["W"], shape["strand"])
    for i in range(args.warmup + args.steps):
        t, kind, cores = ref.run()
        if i >= args.warmup:
            times.append(t)
    ref.close()
    tot = sum(times)
    value = bases * len(times) / tot / 1e9
    sample = (f"first {n_sample} sequences x {shape['seq_len']} bp of the workload ({bases / 1e6:.0f} Mbases) per step; "
              f"count_patterns is serial in the reference (src/peng.cpp:60-61, no pragma on src/base_pattern.cpp:331-441)")
    line = {"impl": "reference", "metric": "kmer_count_gbases_per_s", "value": value, "unit": "Gbases/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot / len(times), "higher_is_better": True,
            "scaling": shape["scaling"], "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": bench_config(shape, args.gpus),
            "cpu_baseline": {"value": value, "unit": "Gbases/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "Gbases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------
def end_to_end_cli():
    import tempfile
    fasta = os.path.join(ROOT, "tests", "golden", "data", "MafK.fasta")
    ours = os.path.join(ROOT, "peng-motif_b200", "host", "peng_motif")
    ref = os.path.join(ROOT, "oracle", "_ref", "peng_motif_ref")
    out = {"input": "configs[1]: MafK.fasta (5000 x 205 bp), -w 10, default thresholds + EM + merging"}
    with tempfile.TemporaryDirectory() as d:
        times = []
        for i in range(3):  # the first run pays the page-in of the binary and the CUDA libraries
            t0 = time.perf_counter()
            r = subprocess.run([ours, fasta, "-w", "10", "-o", os.path.join(d, "ours.meme"), "--timings"], capture_output=True, text=True, timeout=600)
            times.append(time.perf_counter() - t0)
            if r.returncode != 0:
                return {"error": r.stderr[-300:]}
        out["b200_seconds"] = min(times)
        out["b200_phases_ms"] = {l.split("]")[1].rsplit(None, 2)[0].strip(): float(l.split()[-2]) for l in r.stderr.splitlines()
                                 if l.startswith("[timing]")}
        if os.path.exists(ref):
            nthreads = os.cpu_count() or 1
            t0 = time.perf_counter()
            rr = subprocess.run([ref, fasta, "-w", "10", "--threads", str(nthreads), "-o", os.path.join(d, "ref.meme")],
                                capture_output=True, text=True)
            out["reference_cpu_seconds"] = time.perf_counter() - t0
            out["reference_threads"] = nthreads
            out["meme_identical"] = open(os.path.join(d, "ours.meme")).read() == open(os.path.join(d, "ref.meme")).read()
            # with --threads > 1 the reference's "Calculating PWMs" loop is an OpenMP loop whose lines interleave, so the console
            # report is compared byte for byte against a second, single-threaded run of the reference
            r1 = subprocess.run([ref, fasta, "-w", "10", "--threads", "1", "-o", os.path.join(d, "ref1.meme")], capture_output=True, text=True)
            out["stdout_identical_to_reference_threads_1"] = r.stdout == r1.stdout
            out["stdout_same_lines"] = sorted(r.stdout.splitlines()) == sorted(r1.stdout.splitlines())
    return out


def fp64_peak_tflops():
    """B200 FP64 vector peak: 148 SMs x 64 FMA lanes x 2 flop x the SM clock of MEASURED_PEAKS.json (no measured FP64 number
    exists in that file, so this is the nominal pipe rate at the measured clock — said so in the line)."""
    mhz = 1965.0
    try:
        mhz = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["sm_max_mhz"])
    except Exception:
        pass
    return 148 * 64 * 2 * mhz * 1e6 / 1e12


def em_roofline(g, pkg, W, pw, hbm_peak_gbs, fp64_peak):
    """One EM iteration of all PWMs = ONE pass over {count, bgprob} (16 B per bin) + ~10.6 FP64 instructions (~17 flop: the
    fused multiply-adds count twice) per (bin, PWM).  Which roof binds depends on the number of PWMs: HBM (or L2, W <= 10)
    below ~4 PWMs, the FP64 pipe above."""
    import torch

    P = len(pw)
    g.em(pw, pkg.F64, 1e4, 0.0, 2)
    torch.cuda.synchronize()
    its_n = 10
    t0 = time.perf_counter()
    _, its = g.em(pw, pkg.F64, 1e4, 0.0, its_n)
    dt = time.perf_counter() - t0
    # the call also pays the one-off pair table (k_em_pairs) and allocation: time a 1-iteration call and take the difference
    t0 = time.perf_counter()
    g.em(pw, pkg.F64, 1e4, 0.0, 1)
    dt1 = time.perf_counter() - t0
    per_it = max(1e-9, (dt - dt1) / (its_n - 1))
    bins = 4 ** W
    flop = 17.0 * bins * P
    return {"ms_per_iteration_all_pwms": 1e3 * per_it, "pwm_iterations_per_s_steady": P / per_it,
            "roofline": {"bound": "fp64" if P >= 4 else "hbm", "table_bytes_per_iteration": 16 * bins,
                         "table_GBps": 16 * bins / per_it / 1e9, "hbm_peak_GBps": hbm_peak_gbs, "hbm_frac": 16 * bins / per_it / 1e9 / hbm_peak_gbs,
                         "fp64_tflops": flop / per_it / 1e12, "fp64_peak_tflops": fp64_peak, "fp64_frac": flop / per_it / 1e12 / fp64_peak,
                         "fp64_peak_source": "nominal: 148 SMs x 64 FMA/clk x SM clock of MEASURED_PEAKS.json",
                         "flop_per_bin_and_pwm": 17}}


def em_cpu_baseline():
    """The reference's own Peng::em_optimize_pwms (src/peng.cpp:48-178), timed inside oracle/_ref/ref_dump on configs[1]
    (MafK.fasta, W = 10, the PWMs its own hill climb produced), EM threshold 0 so that every PWM runs all 10 iterations."""
    ref_dump = os.path.join(ROOT, "oracle", "_ref", "ref_dump")
    if not os.path.exists(ref_dump):
        return {"unavailable": "oracle/_ref/ref_dump not built"}
    d = tempfile.mkdtemp(prefix="peng_em_")
    try:
        cores = os.cpu_count() or 1
        r = subprocess.run([ref_dump, os.path.join(ROOT, "tests", "golden", "data", "MafK.fasta"), "10", "BOTH", d, "--em-threshold", "0"],
                           capture_output=True, text=True, env=dict(os.environ, OMP_NUM_THREADS=str(cores)))
        if r.returncode != 0:
            return {"unavailable": r.stderr[-200:]}
        meta = dict(l.split() for l in open(os.path.join(d, "meta.txt")))
        n, t = int(meta["n_em_pwms"]), float(meta["t_em_s"])
        return {"kind": "reference", "cores": cores, "pwms": n, "iterations": 10, "W": 10, "seconds": t,
                "pwm_iterations_per_s": n * 10 / t, "sample": "configs[1]: MafK.fasta, W = 10, both strands; the loop over PWMs is an OpenMP "
                "loop in the reference (src/peng.cpp:60-61), the table pass inside a PWM is serial"}
    finally:
        shutil.rmtree(d, ignore_errors=True)


def configs3_run(pkg, torch, dist, rank, world, local, steps=3):
    """BASELINE configs[3]: 50 M x 500 bp (25 Gbp) split over the GPUs (strong scaling): > 2^30 positions per GPU (several
    chunks of the partition path), 2.5e10 > 2^32 in total -> 64-bit exchange (ncclAllReduce of the canonical uint64 table
    + ltot, mirror afterwards).  Device-timed, max over ranks; ltot checked in closed form."""
    n_seq, L, W = 50_000_000, 500, 10
    first, n_local = pkg.shard_synthetic(n_seq, rank, world)
    g = pkg.PengCuda(local)
    g.set_stream(torch.cuda.current_stream().cuda_stream)  # main() made a non-default stream current
    g.ingest_synthetic(n_local, L, 1234, first, MOTIF, 0.3, 0.0)

    def step():
        if world == 1:
            return g.count(W, pkg.STRAND_BOTH)
        g.count_local(W, pkg.STRAND_BOTH)
        dist.all_reduce(torch.as_tensor(g.count_buffer(), device="cuda"))
        return g.count_finish()

    lt = step()
    ok = int(lt) == n_seq * (L - W + 1)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        step()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / steps
    g.close()
    return {"workload": "configs[3]: synthetic 50M x 500 bp (25 Gbp) split over the GPUs, W=10, both strands", "n_gpus": world,
            "steps": steps, "ms_per_step": ms, "gbases_per_s": n_seq * L / (ms * 1e-3) / 1e9, "scaling": "strong",
            "exchange": "ncclAllReduce of 4^10 uint64 + ltot (the set has more than 2^32 positions)" if world > 1 else None,
            "ltot_ok": ok}


def run_cfg5_sliced(args, pkg, torch, dist, rank, world, local):
    """BASELINE configs[4] as north_star states it: 10 M x 300 bp, --strand PLUS, the MAXIMUM pattern length W = 16
    (src/peng.cpp:325-330 of the reference) with the long-k-mer tables, on 8 GPUs.  The 4^16 tables do not fit one GPU (six FP64
    tables of 32 GiB), so every rank keeps 1/N of the id range: count_local32 (deep partition path) -> ncclReduceScatter of
    the uint32 table (+ all-reduce of the ltot word) -> count_finish32_sliced -> stats on the slice -> EM over the slice,
    accumulators all-reduced.  Checked: ltot in closed form, and 2048 occupied bins against an independent path (the
    L2-atomic kernel per rank + an all-reduce of the gathered bins)."""
    import numpy as np

    W, L, n_seq = int(os.environ.get("PENG_CFG5_W", "16")), 300, int(os.environ.get("PENG_CFG5_NSEQ", "10000000"))  # env: functional checks of this flow at a small scale
    first, n_local = pkg.shard_synthetic(n_seq, rank, world)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    g = pkg.PengCuda(local)
    g.set_stream(stream.cuda_stream)
    g.set_table_slice(rank, world)
    g.ingest_synthetic(n_local, L, 1234, first, MOTIF, 0.3, 0.0)
    n = 4 ** W

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    # ---- reference bins through another path: first W-mers of 256 sequences per rank, counted by the L2-atomic kernel ----
    codes, offs = None, None
    sample_n = 256
    gs = pkg.PengCuda(local)
    gs.ingest_synthetic(sample_n, L, 1234, first, MOTIF, 0.3, 0.0)
    sc, so = gs.export_codes()
    gs.close()
    ids_local = np.array([sum((int(c) - 1) << (2 * p) for p, c in enumerate(sc[int(o):int(o) + W])) for o in so[:-1]], np.int64)
    ids_t = torch.from_numpy(ids_local).cuda()
    all_ids = [torch.empty_like(ids_t) for _ in range(world)]
    dist.all_gather(all_ids, ids_t)
    ids = torch.cat(all_ids).cpu().numpy().astype(np.uint64)
    g.set_table_slice(0, 1)
    g.set_count_path(1)
    g.count(W, pkg.STRAND_PLUS)
    ref = torch.from_numpy(g.gather_counts(ids).astype(np.int64)).cuda()
    dist.all_reduce(ref)
    ref = ref.cpu().numpy().astype(np.uint64)
    g.set_count_path(0)
    g.set_table_slice(rank, world)

    scratch = None

    def step():
        nonlocal scratch
        g.count_local32(W, pkg.STRAND_PLUS)
        buf = torch.as_tensor(g.count_buffer32(), device="cuda")
        lo, hi = g.table_slice()
        if scratch is None:
            scratch = torch.empty(hi - lo, dtype=torch.int32, device="cuda")
        dist.reduce_scatter_tensor(scratch, buf[:n])
        buf[lo:hi].copy_(scratch)
        dist.all_reduce(buf[n:])
        return g.count_finish32_sliced()

    lt = step()
    lo, hi = g.table_slice()
    mine = (ids >= lo) & (ids < hi)
    ok_bins = bool(np.array_equal(g.gather_counts(ids[mine]), ref[mine])) and bool(ref.min() >= 1)
    ok = torch.tensor([1 if (ok_bins and int(lt) == n_seq * (L - W + 1)) else 0], device="cuda")
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    barrier()
    steps = max(2, min(args.steps, 5))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        step()
    b.record()
    barrier()
    t = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / steps
    # ---- table pass + one EM iteration on the slices ----
    v64 = g.background_model(2, precision=pkg.F64)
    vt = torch.from_numpy(v64).cuda()
    dist.broadcast(vt, 0)  # one model for all ranks (rank 0's shard; the timing does not depend on the values)
    v64 = vt.cpu().numpy()
    barrier()
    t0 = time.perf_counter()
    g.stats(None, v64, K=2, k=2, max_k=2)
    g.synchronize()
    stats_ms = 1e3 * (time.perf_counter() - t0)
    rng = np.random.RandomState(0)
    pw = rng.dirichlet(np.ones(4), (17, W)) * 0.3
    for m in range(17):
        for p in range(W):
            pw[m, p, "ACGT".index("TGCTGACTCAGCA"[(m + p) % 13])] += 0.7
    pw /= pw.sum(2, keepdims=True)
    acc = torch.as_tensor(g.em_begin(pw, 1e4, 0.0, 3, 2, rank, world), device="cuda")
    em_ms = []
    for it in range(3):
        barrier()
        t0 = time.perf_counter()
        g.em_accumulate()
        g.synchronize()
        dist.all_reduce(acc)
        g.em_update(False)
        barrier()
        em_ms.append(1e3 * (time.perf_counter() - t0))
    out, its = g.em_end()
    em_ok = bool(np.isfinite(out).all()) and bool(np.allclose(out.sum(2), 1.0))
    if rank == 0:
        line = {"metric": "kmer_count_gbases_per_s", "value": n_seq * L / (ms * 1e-3) / 1e9, "unit": "Gbases/s", "n_gpus": world,
                "steps": steps, "warmup": 1, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "u32", "data": "synthetic",
                "config": {"workload": f"configs[4]: synthetic {n_seq} x 300 bp, --strand PLUS, " + ("MAXIMUM pattern length W=16" if W == 16 else f"W={W}") + ", id-sliced long-k-mer tables",
                           "W": W, "strand": "PLUS", "n_seq_per_gpu": n_local, "seq_len": L, "seed": 1234, "motif": MOTIF, "plant_fraction": 0.3,
                           "n_gpus": world},
                "timing": {"exchange": f"ncclReduceScatter of 4^{W} uint32 + all-reduce of the ltot word, sliced finish",
                           "result_checks": {"ltot_and_2048_bins_vs_l2_atomic_path_all_ranks": bool(int(ok.item()))}},
                "extra": {"stats_slice_ms_fp64": stats_ms, "em_fp64": {"pwms": 17, "W": W, "ms_per_iteration_all_pwms": min(em_ms),
                                                                       "iterations_timed": em_ms, "pwms_valid": em_ok,
                                                                       "protocol": "em_accumulate on the id slice -> ncclAllReduce of 17 x 16 x 4 doubles -> em_update"},
                          "slice_ids_per_gpu": int(hi - lo)}}
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    g.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg3")
    ap.add_argument("--ref-sample", type=int, default=250_000, help="sequences per CPU-reference step")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N > 1, sets below 2^32 positions: fused sum + fold over NVLink peer memory (k_peer_exchange32) or "
                         "ncclAllReduce of the uint32 table + fold")
    ap.add_argument("--kernels-only", action="store_true",
                    help="short run for ncu: resident steps only (no time-based warm-up, e2e, CPU baseline, extras)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    shape = workload_shape(args.workload, args.gpus)
    if args.impl == "reference":
        return run_reference_arm(args, shape)

    import numpy as np
    import torch

    import _pkg

    pkg = _pkg.load()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the B200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    if args.workload == "cfg5" and (world >= 8 or (world > 1 and "PENG_CFG5_W" in os.environ)):
        return run_cfg5_sliced(args, pkg, torch, dist, rank, world, local)

    W, strand = shape["W"], pkg.STRAND_BOTH if shape["strand"] == "BOTH" else pkg.STRAND_PLUS
    n_local, L = shape["n_seq_per_gpu"], shape["seq_len"]
    first_seq = rank * n_local
    bases_local = n_local * L
    g = pkg.PengCuda(local)
    # One stream for everything that is timed: torch's DEFAULT stream is the NULL stream, which peng_cuda_set_stream takes
    # as "the handle's own stream" — the library's kernels would then run beside the L2 flush and outside the torch events.
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    g.set_stream(stream.cuda_stream)  # the library launches on torch's current stream: torch events see its kernels
    g.ingest_synthetic(n_local, L, 1234, first_seq, MOTIF, 0.3, 0.0)
    g.enable_timing(True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    # the whole set has fewer than 2^32 positions and a rank's share at most 2^30: the ranks exchange the unfolded uint32
    # table (half the bytes of the uint64 one) and fold once, after the sum
    exchange32 = world > 1 and (bases_local + 2 * n_local) * world < 2 ** 32 - 2 ** 20 and bases_local + 2 * n_local < 2 ** 30 - 2 ** 12

    def count_multi_nccl():
        if exchange32:
            g.count_local32(W, strand)
            buf = torch.as_tensor(g.count_buffer32(), device="cuda")
            dist.all_reduce(buf)  # per-GPU uint32 tables + ltot -> global table (NCCL over NVLink)
            return g.count_finish32()
        g.count_local(W, strand)
        buf = torch.as_tensor(g.count_buffer(), device="cuda")
        dist.all_reduce(buf)  # per-GPU histograms + ltot -> global table (NCCL over NVLink)
        return g.count_finish()

    # Peer group: the ranks exchange CUDA IPC handles of their tables once (NCCL is the bootstrap only); every step then
    # runs ONE kernel per rank that sums the uint32 tables over NVLink peer memory, folds, mirrors and publishes the
    # finished table on every rank (k_peer_exchange32) — no library collective in the data path, no host synchronisation.
    peer, peer_note = False, None
    if world > 1 and exchange32 and args.exchange == "peer" and W >= 8:
        try:
            blob = torch.from_numpy(g.peer_export(W)).cuda()
            blobs = [torch.empty_like(blob) for _ in range(world)]
            dist.all_gather(blobs, blob)
            g.peer_attach(rank, world, torch.stack(blobs).cpu().numpy())
            peer = True
        except Exception as e:  # e.g. no IPC between the ranks' containers: the NCCL path still measures the step
            peer_note = f"peer attach failed ({e!r}): NCCL all-reduce path"
        ok = torch.tensor([1 if peer else 0], device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if peer and not int(ok.item()):
            g.peer_detach()
            peer = False

    def count_multi(want_ltot=False):
        if peer:
            g.count_local32(W, strand)
            return g.count_exchange32(want_ltot)
        return count_multi_nccl()

    def step_resident():
        if world > 1:
            return count_multi()
        return g.count(W, strand, want_ltot=False)  # ltot stays on the device: no read-back, no host synchronisation per step

    # ---- the counted result is checked, not only timed: ltot in closed form (no N in the set: every sequence gives
    # L - W + 1 W-mers), and for N > 1 the peer exchange against the NCCL path on 2^16 bins + the sum of all bins ----
    expect_ltot = n_local * world * (L - W + 1)
    checks = {}
    if world > 1:
        ids = np.random.RandomState(5).randint(0, 4 ** W, 1 << 16, dtype=np.int64).astype(np.uint64)
        lt_n = count_multi_nccl()
        ref_bins, ref_total = g.gather_counts(ids), g.counts_total()
        checks["ltot_nccl_path"] = int(lt_n) == expect_ltot
        if peer:
            lt_p = count_multi(True)
            checks["ltot_peer_path"] = int(lt_p) == expect_ltot
            checks["peer_equals_nccl_bins"] = bool(np.array_equal(g.gather_counts(ids), ref_bins)) and g.counts_total() == ref_total
    else:
        checks["ltot"] = int(g.count(W, strand)) == expect_ltot
        checks["bins_sum_le_ltot"] = g.counts_total() <= (2 if strand == pkg.STRAND_BOTH else 1) * expect_ltot
    if not all(checks.values()):
        raise SystemExit(f"bench.py: counted result failed its checks on rank {rank}: {checks}")

    # ---- resident throughput ("value") ----
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()  # nvidia-smi needs a few 100 ms to deliver its first sample: start before the warm-up
    t_load = time.perf_counter()
    for i in range(10 ** 9):
        flush.zero_()
        step_resident()
        if i + 1 >= args.warmup and (time.perf_counter() - t_load > 0.6 or world > 1 or args.kernels_only):
            break  # keep the GPU under load until the clock sampler has data (single GPU only: ranks must agree)
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kern_ms, phase_ms = [], []
    partition = g.last_count_path == 2  # W = 10 on >= 4 M positions: k_count_part / k_count_slow / k_hist_part
    launches0 = g.launches
    t_wall = time.perf_counter()
    for i, (a, b) in enumerate(ev):
        flush.zero_()  # L2 flush between timed iterations (packed set = 75 MB < L2)
        a.record()
        step_resident()
        b.record()
        # the library's own events of this step (kernel durations for the roofline), read on every 8th step only: reading
        # them waits for the step to finish, and a wait per step would leave the GPU idle for a launch latency each time
        if i % 8 == 7 or i == len(ev) - 1:
            kern_ms.append(g.get_timing((6 if W > 10 else 3) if partition else 1))
            phase_ms.append([g.get_timing(p) for p in range(7)] + ([g.get_timing(7)] if peer else [0.0]))
    barrier()
    t_wall = time.perf_counter() - t_wall
    launches = g.launches - launches0
    checks["ltot_after_timed_steps"] = int(g.count_finish()) == expect_ltot
    peer_trace = [round(float(x) / 1e3, 2) for x in g.peer_trace()] if peer else None  # us, last timed step, this rank
    if not checks["ltot_after_timed_steps"]:
        raise SystemExit(f"bench.py: ltot after the timed steps is wrong on rank {rank}")
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    t = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms = float(t.item())
    total_bases = bases_local * world
    value = total_bases * args.steps / (dev_ms * 1e-3) / 1e9
    # per-rank medians of stage P / stage H / exchange kernel: how far apart the ranks run (the exchange waits for the slowest)
    rank_phases = None
    if dist is not None:
        med = np.median(np.array(phase_ms), axis=0)
        mine = torch.tensor([med[3], med[5], med[7], sum(a.elapsed_time(b) for a, b in ev) / len(ev)], dtype=torch.float64, device="cuda")
        allr = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        rank_phases = {"stage_P_ms": [float(x[0]) for x in allr], "stage_H_ms": [float(x[1]) for x in allr],
                       "exchange_kernel_ms": [float(x[2]) for x in allr], "step_ms": [float(x[3]) for x in allr]}

    if args.kernels_only:
        if rank == 0:
            sampler.stop()
            print(json.dumps({"metric": "kmer_count_gbases_per_s", "value": value, "unit": "Gbases/s", "n_gpus": world,
                              "steps": args.steps, "ms_per_step": dev_ms / args.steps, "kernel_ms": sorted(kern_ms)[len(kern_ms) // 2],
                              "gpu_launches": int(launches), "note": "kernels-only short run (profiling)"}), flush=True)
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- end to end from host buffers through the C ABI ("e2e") ----
    codes_np, offs = g.export_codes()
    codes_pin = torch.from_numpy(codes_np).pin_memory()
    # the set has fewer than 2^32 positions: the finished table travels back as uint32 (peng_cuda_get_counts32), half the bytes
    counts_host = torch.empty(4 ** W, dtype=torch.int32).pin_memory()
    d2h = counts_host.numel() * 4
    g.set_count_prefetch(W, strand)  # what a binding knows before it hands the sequences over: stage P runs under the ingest's DMA

    def step_e2e():
        g.ingest_codes(codes_pin.numpy(), offs)  # H2D of 1 B/base codes + device packing
        if world > 1:
            count_multi()
        else:
            g.count(W, strand, want_ltot=False)
        g._ck(g.lib.peng_cuda_get_counts32(g.h, counts_host.data_ptr()))  # D2H of the 4^W table (fetches ltot with it)

    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(2):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    barrier()
    e2e_s = time.perf_counter() - t0
    # bytes that crossed PCIe per step: the sequences the host threads packed travel as 0.375 B/base (2-bit codes + validity),
    # the rest as 1 B/base for the device to pack, plus the record offsets
    packed_bases = int(offs[g.last_ingest_host_packed])
    h2d = g.last_ingest_h2d_bytes  # counted by the library copy by copy (2-bit codes, validity words where they are not implied, raw rest, offsets)
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = total_bases * e2e_steps / float(t.item()) / 1e9
    # ---- the same end-to-end step from what a binding inside the reference really has: one PAGEABLE array per sequence
    # (Sequence::getSequence() / getL()), handed over as a pointer table — peng_cuda_ingest_sequences packs them on the host ----
    seq_ptrs = (np.uint64(codes_np.ctypes.data) + offs[:-1]).astype(np.uint64)
    seq_lens = np.diff(offs).astype(np.uint32)

    def step_e2e_pageable():
        g.ingest_sequences(seq_ptrs, seq_lens)
        if world > 1:
            count_multi()
        else:
            g.count(W, strand, want_ltot=False)
        g._ck(g.lib.peng_cuda_get_counts32(g.h, counts_host.data_ptr()))

    step_e2e_pageable()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e_pageable()
    barrier()
    t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_pageable = {"value": total_bases * e2e_steps / float(t.item()) / 1e9, "unit": "Gbases/s", "steps": e2e_steps,
                    "h2d_bytes_per_step": int(g.last_ingest_h2d_bytes), "d2h_bytes_per_step": int(d2h),
                    "path": "peng_cuda_ingest_sequences (pointer table over pageable per-sequence arrays; every sequence packed by the "
                            "host threads) -> peng_cuda_count -> peng_cuda_get_counts32"}
    clocks = sampler.stop() if rank == 0 else None  # sampled over warm-up + timed region + e2e region
    # where the end-to-end time goes (rank 0, one more step, host clock around each C-ABI call)
    e2e_phases = {}
    t0 = time.perf_counter(); g.ingest_codes(codes_pin.numpy(), offs); g.synchronize()
    e2e_phases["ingest_codes_ms"] = 1e3 * (time.perf_counter() - t0)
    t0 = time.perf_counter(); g.count(W, strand) if world == 1 else g.count_local(W, strand); g.synchronize()
    e2e_phases["count_ms"] = 1e3 * (time.perf_counter() - t0)
    if world == 1:
        t0 = time.perf_counter(); g._ck(g.lib.peng_cuda_get_counts32(g.h, counts_host.data_ptr()))
        e2e_phases["get_counts_ms"] = 1e3 * (time.perf_counter() - t0)
    g.set_count_prefetch(0)
    # the e2e result is checked too: the table that came back is the resident one
    if world == 1:
        g.count(W, strand)
        checks["e2e_table_equals_resident"] = bool(np.array_equal(counts_host.numpy().view(np.uint32), g.get_counts().astype(np.uint32)))

    # ---- the EM protocol on N real GPUs: every rank accumulates its slice of the id range, the P x W x 4 FP64 accumulators are
    # all-reduced over NVLink, every rank updates; asserted <= 1e-9 against the single-GPU peng_cuda_em of the same PWMs ----
    em_sharded = None
    if world > 1 and not args.no_extras:
        try:
            count_multi(True)  # every rank holds the global table
            v64g = g.background_model(2, precision=pkg.F64)
            vt = torch.from_numpy(v64g).cuda()
            dist.broadcast(vt, 0)
            v64g = vt.cpu().numpy()
            g.stats(None, v64g, K=2)
            rng = np.random.RandomState(3)
            pwm = rng.dirichlet(np.ones(4), (17, W)) * 0.3
            for m in range(17):
                for p in range(W):
                    pwm[m, p, "ACGT".index("TGCTGACTCAGCA"[(m + p) % 13])] += 0.7
            pwm /= pwm.sum(2, keepdims=True)
            whole, its_whole = g.em(pwm, pkg.F64, 1e4, 0.0, 5)
            acc = torch.as_tensor(g.em_begin(pwm, 1e4, 0.0, 5, None, rank, world), device="cuda")
            barrier()
            t0 = time.perf_counter()
            for _ in range(5):
                g.em_accumulate()
                g.synchronize()  # the handle's stream and NCCL's are ordered through the host here
                dist.all_reduce(acc)
                torch.cuda.synchronize()
                g.em_update(False)
            barrier()
            dt = time.perf_counter() - t0
            sharded, its = g.em_end()
            rel = float(np.max(np.abs(sharded - whole) / np.maximum(np.abs(whole), 1e-300)))
            em_sharded = {"pwms": 17, "W": W, "iterations": 5, "n_gpus": world, "max_rel_diff_vs_single_gpu": rel, "within_1e-9": rel <= 1e-9,
                          "ms_per_iteration": 1e3 * dt / 5, "exchange": "ncclAllReduce of 17 x W x 4 doubles per iteration"}
        except Exception as e:
            em_sharded = {"error": repr(e)}

    if peer:  # every rank closes its mappings of the peers' tables before anybody frees or re-allocates them
        g.peer_detach()
        barrier()

    # ---- BASELINE configs[3] (the set north_star quotes the scaling target on): every rank takes part ----
    cfg3_extra = None
    if not args.no_extras and args.workload == "cfg3":
        try:
            cfg3_extra = configs3_run(pkg, torch, dist, rank, world, local)
        except Exception as e:
            cfg3_extra = {"error": repr(e)}

    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ----
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json, burst copy)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    k_ms = sorted(kern_ms)[len(kern_ms) // 2]
    ph = np.median(np.array(phase_ms), axis=0)
    # one launch of the dominant kernel covers the whole local set, or its first chunk of 2^30 stream positions when the
    # partition path splits a larger set (peng_cuda.cu: launch_count_partition); the timed launch is that first chunk
    launch_bases = min(bases_local, 2 ** 30) if partition else bases_local
    algo_bytes = ALGO_BYTES_PER_BASE * launch_bases + 4.0 * 4 ** W
    achieved = algo_bytes / (k_ms * 1e-3) / 1e9
    if partition and W > 10:
        # deep partition path: the largest kernel is stage P2, a pure 256-way split of the 32-bit ids of one chunk:
        # algorithmic bytes = 4 B read + 2 B written per id
        algo_bytes = 6.0 * launch_bases
        achieved = algo_bytes / (k_ms * 1e-3) / 1e9
        kname = f"k_part2 (stage P2 of the deep partition path, W = {W})"
        note = ("stage P2 splits the 32-bit ids of a stage-P class 256 ways into 16-bit leaf entries through shared-memory queues; "
                "bound by shared-memory wavefronts (bank conflicts of the slot atomics / scattered stores) and issue, see DESIGN.md")
    elif partition:
        kname = f"k_count_part<{W},{shape['strand']}> (stage P of the partition path)"
        note = ("stage P streams the packed set once (codes + validity), derives k-mer starts and conflict candidates and "
                "partitions 16-bit keys into class buckets; bound by the ALU pipe and issue, not by bytes (counters: the ncu extract "
                "named in roofline.traffic_source), see DESIGN.md")
    else:
        kname = f"k_count_global<{W},{shape['strand']}>" if W >= 10 else "k_count_smem"
        note = "one L2 atomic per counted k-mer: atomic-throughput-bound far below the HBM roof (DESIGN.md)"
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "kernel": kname, "kernel_ms": k_ms, "algorithmic_bytes": algo_bytes, "peak_source": peak_src, "note": note,
                "step_phases_ms": {"mark": float(ph[0]), "count_total": float(ph[1]), "fold": float(ph[2]),
                                   "stage_P": float(ph[3]), "slow_units": float(ph[4]), "stage_H": float(ph[5]),
                                   "stage_P2": float(ph[6]), "peer_exchange_kernel": float(ph[7])},
                "whole_step_frac": (ALGO_BYTES_PER_BASE * bases_local + 4.0 * 4 ** W) / (dev_ms / args.steps * 1e-3) / 1e9 / peak}
    tr = os.path.join(ROOT, "profiles", "count_kernel_traffic.json")
    if os.path.exists(tr):
        try:
            tj = json.load(open(tr))
            roofline["traffic"] = tj.get(args.workload)
            roofline["traffic_source"] = tj.get("source", {}).get(args.workload)
            # distance to the ceiling that actually binds the two-stage design: one shared-memory atomic per counted W-mer in
            # each stage (2.4e12/s chip-wide, profiles/r1_atomic_microbench.txt) -> bases / 2.4e12 s per stage
            if partition and W == 10:
                roofline["smem_atomic_ceiling_frac"] = (bases_local / 2.4e12) / (dev_ms / args.steps * 1e-3)
        except Exception:
            pass

    # ---- CPU baseline on a bounded sample of the same workload ----
    n_sample = min(n_local, args.ref_sample)
    s_codes, s_offs = codes_np[: n_sample * L], offs[: n_sample + 1]
    ref = CpuReference(s_codes, s_offs, W, shape["strand"])
    cpu_s, kind, cores = ref.run()
    ref.close()
    cpu_baseline = {"value": n_sample * L / cpu_s / 1e9, "unit": "Gbases/s", "cores": cores, "kind": kind,
                    "sample": f"first {n_sample} sequences x {L} bp of the workload ({n_sample * L / 1e6:.0f} Mbases), "
                              f"count_patterns only; serial in the reference, so effective parallelism is 1"}

    # ---- extras: the other two BASELINE metrics on the same table ----
    extra = {}
    if cfg3_extra is not None:
        extra["configs3_25gbp"] = cfg3_extra
    if em_sharded is not None:
        extra["em_fp64_sharded"] = em_sharded
    if not args.no_extras:
        try:
            t0 = time.perf_counter()
            v32 = g.background_model(2, precision=pkg.F32REF)
            v64 = g.background_model(2, precision=pkg.F64)
            g.synchronize()
            extra["background_model_ms"] = 1e3 * (time.perf_counter() - t0) / 2
            t0 = time.perf_counter()
            g.stats(v32, v64, K=2)
            extra["stats_pass_ms_both_arms"] = 1e3 * (time.perf_counter() - t0)
            rng = np.random.RandomState(0)
            pw = rng.dirichlet(np.ones(4), (17, W)) * 0.3
            for m in range(17):
                for p in range(W):
                    pw[m, p, "ACGT".index("TGCTGACTCAGCA"[(m + p) % 13])] += 0.7
            pw /= pw.sum(2, keepdims=True)
            for prec, name in ((pkg.F64, "em_fp64"), (pkg.F32REF, "em_f32_reference_faithful")):
                g.em(pw, prec, 1e4, 0.0, 2)  # warm-up
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                _, its = g.em(pw, prec, 1e4, 0.0, 10)  # threshold 0: every PWM runs all 10 iterations
                dt = time.perf_counter() - t0
                extra[name] = {"pwm_iterations_per_s": float(its.sum()) / dt, "pwms": 17, "iterations": 10, "W": W,
                               "seconds": dt, "bytes_per_bin_per_iteration": 16}
            extra["em_fp64"].update(em_roofline(g, pkg, W, pw, peak, fp64_peak_tflops()))
            # the long-k-mer point of the EM metric (BASELINE configs[4]): W = 14, PLUS, the same set
            g.count(14, pkg.STRAND_PLUS)
            g.stats(None, v64, K=2)
            pw14 = rng.dirichlet(np.ones(4), (17, 14)) * 0.3
            for m in range(17):
                for p in range(14):
                    pw14[m, p, "ACGT".index("TGCTGACTCAGCA"[(m + p) % 13])] += 0.7
            pw14 /= pw14.sum(2, keepdims=True)
            extra["em_fp64_W14"] = {"pwms": 17, "W": 14, **em_roofline(g, pkg, 14, pw14, peak, fp64_peak_tflops())}
            g.count(W, strand)  # back to the headline table
            extra["em_cpu_baseline"] = em_cpu_baseline()
            # third BASELINE metric: end-to-end seconds of the whole CLI run on configs[1] (MafK.fasta, W = 10, defaults),
            # the host driver on this GPU vs the reference binary on this box's host cores, MEME files compared
            if rank == 0:
                extra["end_to_end_cli"] = end_to_end_cli()
        except Exception as e:  # extras never invalidate the headline line
            extra["error"] = repr(e)

    line = {"metric": "kmer_count_gbases_per_s", "value": value, "unit": "Gbases/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": shape["scaling"],
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": bench_config(shape, world),
            "timing": {"l2": "256 MiB buffer written between timed iterations (inputs smaller than L2)",
                       "parallelism": f"sequence-sharded x{world}" + (
                           (", per step ONE fused kernel per rank: sum of the uint32 tables over NVLink peer memory + fold + mirror + "
                            "publication on every rank (k_peer_exchange32; NCCL only bootstraps the IPC handles)" if peer else
                            ", 1 NCCL all_reduce of 4^W u32 + ltot per step (32-bit exchange)" if exchange32 else
                            ", 1 NCCL all_reduce of 4^W u64 + ltot per step") if world > 1 else ""),
                       "data_plane": ("peer memory (own kernel)" if peer else "nccl") if world > 1 else None,
                       "note": peer_note, "result_checks": checks, "per_rank": rank_phases,
                       "peer_exchange_trace_us": {"arrived": peer_trace[0], "cta0_units_done": peer_trace[1], "cta0_stores_visible": peer_trace[2],
                                                  "all_ctas_done": peer_trace[3], "departed": peer_trace[4]} if peer_trace else None},
            "e2e": {"value": e2e_value, "unit": "Gbases/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "steps": e2e_steps, "path": "peng_cuda_set_count_prefetch + peng_cuda_ingest_codes (pinned host codes; hybrid: host threads pack part of the set "
                            "into the 2-bit layout, the DMA engine moves the raw rest for the device to pack; stage P of the count runs on "
                            "the unit ranges that have landed) -> peng_cuda_count -> peng_cuda_get_counts32",
                    "host_packed_fraction": packed_bases / max(1, codes_pin.numel()), "host_bytes_per_step": int(codes_pin.numel() + offs.nbytes),
                    "phases_ms": e2e_phases, "pageable_per_sequence": e2e_pageable},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline,
            "wall_s_timed_region": t_wall, "extra": extra}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

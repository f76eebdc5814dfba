#!/usr/bin/env python
"""End-to-end seconds of the whole command line on a large synthetic set: BASELINE.json's third metric on configs[2]
(1 M x 200 bp, planted MafK-like motif, -w 10, defaults) — `peng_motif` of this repo on the GPU against the unmodified
reference binary (oracle/_ref/peng_motif_ref) on the same box's host cores, both from the same FASTA file, with the
MEME files and console reports compared.  This is also the full-size parity check of the whole path against the
reference itself.

    python tools/e2e_cli_large.py [--n-seq 1000000] [--seq-len 200] [--skip-reference] > profiles/...json
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
MOTIF = "TGCTGASTCAGCA"


def write_fasta(codes, n_seq, seq_len, path):
    """one header line + one sequence line per record, written in one vectorised pass"""
    lut = np.frombuffer(b"NACGT", np.uint8)
    letters = lut[codes].reshape(n_seq, seq_len)
    head = np.array([b">s%07d\n" % i for i in range(n_seq)], dtype="S10")  # 10 bytes with the newline
    head = np.frombuffer(head.tobytes(), np.uint8).reshape(n_seq, 10)
    rec = np.empty((n_seq, 10 + seq_len + 1), np.uint8)
    rec[:, :10] = head
    rec[:, 10:10 + seq_len] = letters
    rec[:, -1] = ord("\n")
    rec.tofile(path)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n-seq", type=int, default=1_000_000)
    ap.add_argument("--seq-len", type=int, default=200)
    ap.add_argument("--skip-reference", action="store_true")
    ap.add_argument("--gpus", type=int, nargs="*", default=[], help="also run peng_motif --gpus N for these N and compare with the one-GPU run")
    args = ap.parse_args()
    import _pkg

    pkg = _pkg.load()
    g = pkg.PengCuda(0)
    g.ingest_synthetic(args.n_seq, args.seq_len, 1234, 0, MOTIF, 0.3, 0.0)
    codes, _ = g.export_codes()
    g.close()
    ours = os.path.join(ROOT, "peng-motif_b200", "host", "peng_motif")
    ref = os.path.join(ROOT, "oracle", "_ref", "peng_motif_ref")
    out = {"input": f"synthetic {args.n_seq} x {args.seq_len} bp, seed 1234, motif {MOTIF} in 30 %, FASTA file", "args": "-w 10"}
    with tempfile.TemporaryDirectory() as d:
        fasta = os.path.join(d, "set.fa")
        write_fasta(codes, args.n_seq, args.seq_len, fasta)
        out["fasta_bytes"] = os.path.getsize(fasta)
        assert open(fasta, "rb").read(11)[:10] == b">s0000000\n"
        times = []
        for _ in range(3):
            t0 = time.perf_counter()
            r = subprocess.run([ours, fasta, "-w", "10", "-o", os.path.join(d, "ours.meme"), "--timings"], capture_output=True, text=True)
            times.append(time.perf_counter() - t0)
            assert r.returncode == 0, r.stderr[-500:]
        out["b200_seconds"] = min(times)
        out["b200_seconds_all_runs"] = times
        out["b200_phases_ms"] = {l.split("]")[1].rsplit(None, 2)[0].strip(): float(l.split()[-2]) for l in r.stderr.splitlines()
                                 if l.startswith("[timing]")}
        for n in args.gpus:  # the sharded front (FASTA shards, summed background counts, fused exchange): same bytes out
            tn = []
            for _ in range(3):
                t0 = time.perf_counter()
                rn = subprocess.run([ours, fasta, "-w", "10", "-o", os.path.join(d, f"ours{n}.meme"), "--timings", "--gpus", str(n)],
                                    capture_output=True, text=True)
                tn.append(time.perf_counter() - t0)
                assert rn.returncode == 0, rn.stderr[-500:]
            out[f"gpus_{n}"] = {"seconds": min(tn), "stdout_identical_to_one_gpu": rn.stdout == r.stdout,
                                "meme_identical_to_one_gpu": open(os.path.join(d, f"ours{n}.meme")).read() == open(os.path.join(d, "ours.meme")).read(),
                                "phases_ms": {l.split("]")[1].rsplit(None, 2)[0].strip(): float(l.split()[-2]) for l in rn.stderr.splitlines()
                                              if l.startswith("[timing]")}}
        if not args.skip_reference and os.path.exists(ref):
            n = os.cpu_count() or 1
            t0 = time.perf_counter()
            rr = subprocess.run([ref, fasta, "-w", "10", "--threads", str(n), "-o", os.path.join(d, "ref.meme")], capture_output=True, text=True)
            out["reference_cpu_seconds"] = time.perf_counter() - t0
            out["reference_threads"] = n
            out["reference_returncode"] = rr.returncode
            if rr.returncode == 0:
                out["meme_identical"] = open(os.path.join(d, "ours.meme")).read() == open(os.path.join(d, "ref.meme")).read()
                out["speedup"] = out["reference_cpu_seconds"] / out["b200_seconds"]
                # byte-exact console comparison needs the reference single-threaded: its "Calculating PWMs" loop is an
                # OpenMP loop whose output lines interleave with more threads
                t0 = time.perf_counter()
                r1 = subprocess.run([ref, fasta, "-w", "10", "--threads", "1", "-o", os.path.join(d, "ref1.meme")], capture_output=True, text=True)
                out["reference_cpu_seconds_1_thread"] = time.perf_counter() - t0
                out["stdout_identical_vs_1_thread"] = r.stdout == r1.stdout
                out["meme_identical_vs_1_thread"] = open(os.path.join(d, "ours.meme")).read() == open(os.path.join(d, "ref1.meme")).read()
                out["stdout_lines"] = len(r.stdout.splitlines())
                if not out["stdout_identical_vs_1_thread"]:
                    import difflib
                    out["stdout_diff"] = list(difflib.unified_diff(r1.stdout.splitlines(), r.stdout.splitlines(), "reference", "b200",
                                                                   lineterm="", n=0))[:60]
    print(json.dumps(out))


if __name__ == "__main__":
    main()

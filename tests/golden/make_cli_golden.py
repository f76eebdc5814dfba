#!/usr/bin/env python
"""Console report + MEME + JSON of the UNMODIFIED reference CLI (oracle/_ref/peng_motif_ref, built by `make -C oracle ref`)
for a set of option combinations on the small shipped input.  Run in the build container only; tests read the files
written under tests/golden/cli/ and never the reference tree.

    python tests/golden/make_cli_golden.py
"""
import json
import os
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
BIN = os.path.join(ROOT, "oracle", "_ref", "peng_motif_ref")
DATA = os.path.join(HERE, "data")

# name -> (fasta, extra args); every case also writes -o and -j (MEME first, so the JSON shows the doubled no_zero_pwm)
CASES = {
    "w8_default": ("MafK_100seqs.fasta", ["-w", "8"]),
    "w8_plus": ("MafK_100seqs.fasta", ["-w", "8", "--strand", "PLUS"]),
    "w8_logpval": ("MafK_100seqs.fasta", ["-w", "8", "--optimization_score", "LOGPVAL"]),
    "w8_enrichment": ("MafK_100seqs.fasta", ["-w", "8", "--optimization_score", "ENRICHMENT", "--enrich_pseudocount_factor", "0.1"]),
    "w8_no_em": ("MafK_100seqs.fasta", ["-w", "8", "--no-em"]),
    "w8_no_merging": ("MafK_100seqs.fasta", ["-w", "8", "--no-merging"]),
    "w8_default_pwm": ("MafK_100seqs.fasta", ["-w", "8", "--use-default-pwm", "--no-em"]),
    "w8_thresholds": ("MafK_100seqs.fasta", ["-w", "8", "-t", "5", "--count-threshold", "2", "--no-neighbor-filtering",
                                             "--max-optimized-patterns", "12", "-b", "0.6", "--pseudo-counts", "5"]),
    "w8_em_opts": ("MafK_100seqs.fasta", ["-w", "8", "-a", "100", "--em-threshold", "0.01", "--em-max-iterations", "4",
                                          "--minimum-processed-patterns", "5", "--max_merged_length", "9"]),
    "w6_bg1": ("MafK_100seqs.fasta", ["-w", "6", "--bg-model-order", "1", "-t", "6"]),
    "w4_too_short_to_merge": ("MafK_100seqs.fasta", ["-w", "4", "-t", "3"]),
    "w8_other_background": ("MafK_100seqs.fasta", ["-w", "8", "--background-sequences", os.path.join(DATA, "default_sequence_set.fa")]),
    "w8_nothing_found": ("default_sequence_set.fa", ["-w", "8"]),
    "w10_unknown_option": ("MafK_100seqs.fasta", ["-w", "10", "--frobnicate", "-t", "8"]),
}


def main():
    index = {}
    for name, (fasta, extra) in CASES.items():
        with tempfile.TemporaryDirectory() as d:
            meme, js = os.path.join(d, "o.meme"), os.path.join(d, "o.json")
            args = [BIN, os.path.join(DATA, fasta), *extra, "-o", meme, "-j", js]
            r = subprocess.run(args, capture_output=True, text=True)
            open(os.path.join(HERE, "cli", name + ".stdout"), "w").write(r.stdout)
            for path, suf in ((meme, ".meme"), (js, ".json")):
                if os.path.exists(path):
                    open(os.path.join(HERE, "cli", name + suf), "w").write(open(path).read())
            index[name] = {"fasta": fasta, "args": [a.replace(DATA + os.sep, "@DATA@/") for a in extra], "returncode": r.returncode}
            print(name, r.returncode, len(r.stdout))
    json.dump(index, open(os.path.join(HERE, "cli", "index.json"), "w"), indent=1)


if __name__ == "__main__":
    main()

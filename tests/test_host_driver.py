"""The `peng_motif` host driver (peng-motif_b200/host) against the reference CLI's recorded console report, MEME and
JSON files (tests/golden/cfg*_stdout.txt / *.meme, tests/golden/cli/*: written by running the UNMODIFIED reference,
tests/golden/make_golden.py and make_cli_golden.py).

* not gpu: the driver's HOST logic (hill climb bookkeeping, filters, PWM arithmetic, merging, writers, option
  parser) linked against tests/sim/abi_shim.cpp, which answers the C-ABI calls with the CPU oracle — test
  infrastructure, never shipped;
* gpu: the shipped binary (links libpeng_cuda.so) on the B200, same comparisons byte for byte.
"""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "peng-motif_b200", "host")
GOLDEN = os.path.join(ROOT, "tests", "golden")
DATA = os.path.join(GOLDEN, "data")
SIM_BIN = os.path.join(ROOT, "tests", "sim", "peng_motif_sim")
GPU_BIN = os.path.join(HOST, "peng_motif")
SOURCES = [os.path.join(HOST, f) for f in ("main.cpp", "options.cpp", "motif.cpp", "pipeline.cpp", "writers.cpp")]

CLI_CASES = json.load(open(os.path.join(GOLDEN, "cli", "index.json")))
CONFIG_CASES = {  # BASELINE.json configs[0] (both strands + PLUS) and configs[1]
    "cfg1_W8_BOTH": ("MafK_100seqs.fasta", ["-w", "8", "--strand", "BOTH"]),
    "cfg1_W8_PLUS": ("MafK_100seqs.fasta", ["-w", "8", "--strand", "PLUS"]),
    "cfg2_W10_BOTH": ("MafK.fasta", ["-w", "10", "--strand", "BOTH"]),
}


@pytest.fixture(scope="session")
def sim_binary():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "port"])
    shim = os.path.join(ROOT, "tests", "sim", "abi_shim.cpp")
    deps = SOURCES + [shim, os.path.join(HOST, "peng_host.h"), os.path.join(ROOT, "oracle", "libpeng_oracle.so")]
    if not os.path.exists(SIM_BIN) or os.path.getmtime(SIM_BIN) < max(os.path.getmtime(d) for d in deps):
        subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", *SOURCES, shim, "-o",
                               SIM_BIN, "-L" + os.path.join(ROOT, "oracle"), "-lpeng_oracle",
                               "-Wl,-rpath," + os.path.join(ROOT, "oracle")])
    return SIM_BIN


@pytest.fixture(scope="session")
def gpu_binary():
    if not os.path.exists(GPU_BIN):
        subprocess.check_call(["make", "-s", "-C", HOST])
    return GPU_BIN


def run_cli(binary, fasta, args, tmp_path):
    meme, js = str(tmp_path / "o.meme"), str(tmp_path / "o.json")
    r = subprocess.run([binary, os.path.join(DATA, fasta), *[a.replace("@DATA@", DATA) for a in args], "-o", meme, "-j", js],
                       capture_output=True, text=True, timeout=600)
    read = lambda p: open(p).read() if os.path.exists(p) else None
    return r, read(meme), read(js)


def check_cli_case(binary, name, tmp_path):
    c = CLI_CASES[name]
    r, meme, js = run_cli(binary, c["fasta"], c["args"], tmp_path)
    g = lambda suf: open(os.path.join(GOLDEN, "cli", name + suf)).read()
    assert r.returncode == c["returncode"], r.stderr
    assert r.stdout == g(".stdout")
    assert meme == g(".meme")
    assert js == g(".json")


def check_config_case(binary, name, tmp_path):
    fasta, args = CONFIG_CASES[name]
    r, meme, _ = run_cli(binary, fasta, args, tmp_path)
    assert r.returncode == 0, r.stderr
    assert r.stdout == open(os.path.join(GOLDEN, name + "_stdout.txt")).read()
    assert meme == open(os.path.join(GOLDEN, name + ".meme")).read()


# ---- host logic on the CPU shim --------------------------------------------------------------------------
@pytest.mark.parametrize("name", sorted(CLI_CASES))
def test_host_logic_cli_case(sim_binary, name, tmp_path):
    check_cli_case(sim_binary, name, tmp_path)


@pytest.mark.parametrize("name", sorted(CONFIG_CASES))
def test_host_logic_config(sim_binary, name, tmp_path):
    check_config_case(sim_binary, name, tmp_path)


def test_help_and_version(sim_binary):
    """Help text, version lines and exit codes of Global::readArguments (src/Global.cpp:81-96,296-303)."""
    r = subprocess.run([sim_binary, "-h"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and " Usage: peng_motif SEQFILE [options] " in r.stdout
    for flag in ("-o,", "-j,", "--background-sequences,", "-t,", "-w,", "--bg-model-order,", "--count-threshold,", "--strand,",
                 "--optimization_score,", "--enrich_pseudocount_factor,", "-b,", "--no-em", "-a,", "--em-threshold,",
                 "--em-max-iterations,", "--no-merging", "--max_merged_length", "--use-default-pwm", "--pseudo-counts,",
                 "--threads,", "--no-neighbor-filtering", "--minimum-processed-patterns", "--version", "--max-optimized-patterns"):
        assert "\n      " + flag in r.stdout, flag
    r = subprocess.run([sim_binary, "-version"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and r.stdout == "peng_motif version 1.0.0\n"
    r = subprocess.run([sim_binary, "x.fa", "--version"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and r.stdout == "peng_motif 1.0.0\n"
    r = subprocess.run([sim_binary], capture_output=True, text=True, timeout=600)
    assert r.returncode == 255 and "Error: Arguments are missing!" in r.stderr
    r = subprocess.run([sim_binary, "x.fa", "-w"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 4 and "No expression following -w" in r.stderr
    r = subprocess.run([sim_binary, "x.fa", "-w", "7"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 4 and "multiple of 2" in r.stderr
    r = subprocess.run([sim_binary, "x.fa", "--strand", "MINUS"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 4 and "Unknown expression following --strand" in r.stderr


def test_shipped_binary_has_no_cpu_path(gpu_binary, pkg):
    """The shipped driver links libpeng_cuda.so only (no oracle, no shim) and stops without a CUDA device."""
    out = subprocess.run(["ldd", gpu_binary], capture_output=True, text=True).stdout
    assert "libpeng_cuda.so" in out and "oracle" not in out and "sim" not in out
    if pkg.load_library().peng_cuda_device_count() > 0:
        pytest.skip("a GPU is present")
    r = subprocess.run([gpu_binary, os.path.join(DATA, "MafK_100seqs.fasta"), "-w", "8"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 1 and "no CPU fallback" in r.stderr and "[STATUS]" not in r.stdout


# ---- the shipped binary on the GPU -----------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CLI_CASES))
def test_gpu_cli_case(gpu_binary, name, tmp_path):
    check_cli_case(gpu_binary, name, tmp_path)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CONFIG_CASES))
def test_gpu_config(gpu_binary, name, tmp_path):
    check_config_case(gpu_binary, name, tmp_path)


# ---- --gpus N: sharded ingest + counting, combined by the fused peer exchange (or on the host), same bytes out ----------
def run_gpus(binary, name, n, tmp_path, same_device):
    fasta, args = CONFIG_CASES[name]
    env = dict(os.environ)
    if same_device:
        env["PENG_GPUS_SAME_DEVICE"] = "1"
    meme = str(tmp_path / "o.meme")
    r = subprocess.run([binary, os.path.join(DATA, fasta), *args, "--gpus", str(n), "-o", meme], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr
    assert r.stdout == open(os.path.join(GOLDEN, name + "_stdout.txt")).read()
    assert open(meme).read() == open(os.path.join(GOLDEN, name + ".meme")).read()


@pytest.mark.parametrize("n", [2, 5])
@pytest.mark.parametrize("name", ["cfg1_W8_BOTH", "cfg1_W8_PLUS"])
def test_host_logic_multi_gpu(sim_binary, name, n, tmp_path):
    """The driver's --gpus bookkeeping (one thread per handle, FASTA shards, summed background counts, exchange protocol)
    on the CPU shim: same console report and MEME as the single-handle run and the reference."""
    run_gpus(sim_binary, name, n, tmp_path, False)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [2, 3])
@pytest.mark.parametrize("name", sorted(CONFIG_CASES))
def test_gpu_multi_handle_one_device(gpu_binary, name, n, tmp_path):
    """N handles on ONE device (PENG_GPUS_SAME_DEVICE): FASTA shards, summed background counts, count_local32 +
    peer exchange (W = 8, 10) — console report and MEME byte-identical to the reference's."""
    run_gpus(gpu_binary, name, n, tmp_path, True)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CONFIG_CASES))
def test_gpu_multi_device(gpu_binary, pkg, name, tmp_path):
    n = pkg.load_library().peng_cuda_device_count()
    if n < 2:
        pytest.skip("needs >= 2 CUDA devices (the multi-handle protocol itself runs in test_gpu_multi_handle_one_device)")
    run_gpus(gpu_binary, name, min(n, 8), tmp_path, False)


@pytest.mark.gpu
def test_gpu_multi_handle_small_w_host_sum(gpu_binary, tmp_path):
    """W = 6 is outside the fused exchange: per-shard tables are added on the host; output identical to one GPU."""
    fasta = os.path.join(DATA, "MafK_100seqs.fasta")
    outs = []
    for extra in ([], ["--gpus", "2"]):
        r = subprocess.run([gpu_binary, fasta, "-w", "6", *extra, "-o", str(tmp_path / "x.meme")], capture_output=True, text=True,
                           env=dict(os.environ, PENG_GPUS_SAME_DEVICE="1"), timeout=600)
        assert r.returncode == 0, r.stderr
        outs.append((r.stdout, open(tmp_path / "x.meme").read()))
    assert outs[0] == outs[1]

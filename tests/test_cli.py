"""construct_sa CLI (host/construct_sa.cpp) vs the reference's command-line behaviour
(src/main.cpp:48-246): option parsing on CPU, bit-exact .sa5 files on the GPU."""
import os
import subprocess
import numpy as np
import pytest
from golden_cases import golden_cases, golden_text

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "host", "construct_sa")


@pytest.fixture(scope="module")
def cli():
    if not os.path.exists(BIN):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "host")], stdout=subprocess.DEVNULL)
    return BIN


def run(cli, args, stdin=None, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([cli] + args, input=stdin, capture_output=True, text=True, env=e, timeout=240)


def test_cli_argument_errors(cli, tmp_path):
    f = tmp_path / "t.txt"
    f.write_bytes(b"banana")
    r = run(cli, [])
    assert r.returncode == 1 and "Error: FILE not provided" in r.stderr and "Usage:" in r.stdout
    r = run(cli, ["-h"])
    assert r.returncode == 1 and "Construct the suffix array of text stored in FILE." in r.stdout
    r = run(cli, [str(f), "-m", "3.5gi"])
    assert r.returncode == 1 and "Error: parsing RAM limit (3.5gi) failed" in r.stderr
    r = run(cli, [str(f), "-m", "0"])
    assert r.returncode == 1 and "Error: invalid RAM limit (0)" in r.stderr
    r = run(cli, [str(tmp_path / "missing")])
    assert r.returncode == 1 and "does not exist" in r.stderr
    (tmp_path / "t.txt.sa5").write_bytes(b"x")
    r = run(cli, [str(f)], stdin="n\n")
    assert r.returncode == 1 and "exists. Overwrite? [y/n]: " in r.stdout
    assert (tmp_path / "t.txt.sa5").read_bytes() == b"x"


@pytest.mark.gpu
def test_cli_bit_exact_sa5(cli, tmp_path):
    for name, gen, kw, n, sa5 in golden_cases()[:4]:
        f = tmp_path / (name + ".txt")
        golden_text(gen, kw, n).tofile(str(f))
        r = run(cli, [str(f), "-m", "3584Mi", "-v"])
        assert r.returncode == 0, r.stderr
        assert "Computation finished. Summary:" in r.stderr and "speed:" in r.stderr
        assert np.array_equal(np.fromfile(str(f) + ".sa5", dtype=np.uint8), sa5)
        out = tmp_path / (name + ".out")
        r = run(cli, [str(f), "-o", str(out), "-g", str(tmp_path / "gap")], env={"PSASCAN_MAX_BLOCK": "1999"})
        assert r.returncode == 0, r.stderr
        assert np.array_equal(np.fromfile(str(out), dtype=np.uint8), sa5)
        r = run(cli, [str(f)], stdin="y\n")  # overwrite prompt answered yes
        assert r.returncode == 0 and "Overwrite? [y/n]" in r.stdout


@pytest.mark.gpu
def test_cli_rejects_255(cli, tmp_path):
    f = tmp_path / "bad.txt"
    f.write_bytes(b"abc\xffdef" * 50)
    r = run(cli, [str(f)])
    assert r.returncode == 1 and "byte with value 255 was detected in the input text!" in r.stdout


@pytest.mark.gpu
def test_cli_all_visible_gpus(cli, tmp_path):
    """construct_sa uses every visible GPU for a text of several blocks (one host thread per GPU inside one
    process, NCCL set up by the library); the file must equal the golden .sa5 whatever the GPU count.
    PSASCAN_GPUS=1 pins the single-GPU pipeline on the same input."""
    import torch
    ngpu = torch.cuda.device_count()
    for name, gen, kw, n, sa5 in golden_cases()[-3:]:
        f = tmp_path / (name + ".txt")
        golden_text(gen, kw, n).tofile(str(f))
        for env in ({"PSASCAN_MAX_BLOCK": str(n // 5 + 3)}, {"PSASCAN_MAX_BLOCK": "1000", "PSASCAN_GPUS": "1"},
                    {"PSASCAN_MAX_BLOCK": str(n // 2 + 1), "PSASCAN_GPUS": "2"}):
            out = tmp_path / (name + ".out")
            if out.exists():
                out.unlink()
            r = run(cli, [str(f), "-o", str(out)], env=env)
            assert r.returncode == 0, r.stderr[-2000:]
            assert np.array_equal(np.fromfile(str(out), dtype=np.uint8), sa5), (name, env)
            want = min(ngpu, int(env.get("PSASCAN_GPUS", ngpu)), -(-n // int(env["PSASCAN_MAX_BLOCK"])))
            assert ("GPU = %d x" % want) in r.stderr, r.stderr[:600]


def test_cli_memory_error_text(cli, tmp_path):
    """ram_use below 6 bytes: the reference's message and exit status (psascan.hpp:57-60)"""
    f = tmp_path / "t.txt"
    f.write_bytes(b"banana")
    r = run(cli, [str(f), "-m", "5"])
    assert r.returncode == 1 and "Error: not enough memory to run pSAscan." in r.stderr


@pytest.mark.gpu
def test_cli_external_memory_mode(cli, tmp_path):
    """PSASCAN_EM=1: partial suffix arrays next to OUTFILE, v-byte gap files next to the -g name, streaming
    merge; the result equals the golden file and no temporary file survives."""
    gap_dir = tmp_path / "gapdisk"
    gap_dir.mkdir()
    for name, gen, kw, n, sa5 in golden_cases()[-3:]:
        f = tmp_path / (name + ".txt")
        golden_text(gen, kw, n).tofile(str(f))
        out = tmp_path / (name + ".sa5")
        r = run(cli, [str(f), "-o", str(out), "-g", str(gap_dir / name), "-m", "200k"],
                env={"PSASCAN_EM": "1", "PSASCAN_MAX_BLOCK": str(n // 3 + 1)})
        assert r.returncode == 0, r.stderr[-2000:]
        assert "external-memory mode" in r.stderr
        assert np.array_equal(np.fromfile(str(out), dtype=np.uint8), sa5), name
        assert sorted(os.listdir(str(gap_dir))) == [] and not [x for x in os.listdir(str(tmp_path)) if "distrfile" in x]

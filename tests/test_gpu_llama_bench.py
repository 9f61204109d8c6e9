"""The reference's own llama-bench (tools/llama-bench/llama-bench.cpp, compiled unmodified by `make -C oracle llama` against the
reference libraries of oracle/_ref) loads a synthetic-weight Llama GGUF and runs prompt processing and token generation on
the B200 plugin through GGML_BACKEND_PATH — the drop-in boundary exercised by the consumer it was made for."""
import os
import re
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _exe():
    d = os.path.join(ROOT, "oracle", "_ref")
    with open("/proc/cpuinfo") as f:
        if "avx512_vnni" in f.read() and os.path.exists(os.path.join(d, "avx512", "llama-bench")):
            d = os.path.join(d, "avx512")
    return os.path.join(d, "llama-bench")


@pytest.mark.parametrize("ftype", ["q4_k_m", "q8_0", "gpt-oss"])
def test_llama_bench_runs_on_the_plugin(tmp_path, ftype):
    exe = _exe()
    plugin = os.path.join(ROOT, "llama.cpp-gfx906_b200", "lib", "libggml-b200.so")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/llama-bench not built (make -C oracle llama)")
    gguf = str(tmp_path / "synth.gguf")
    # "gpt-oss": the MoE architecture of BASELINE config 4 (MXFP4 experts, attention sinks, sliding window, biases through ADD_ID)
    kind = ["--arch", "gpt-oss"] if ftype == "gpt-oss" else ["--ftype", ftype]
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_synth_gguf.py"), "--out", gguf, "--layers", "4", "--vocab", "32000"] + kind,
                   check=True, capture_output=True, timeout=300)
    env = dict(os.environ, GGML_BACKEND_PATH=plugin)
    r = subprocess.run([exe, "-m", gguf, "-ngl", "99", "-fa", "0", "-p", "64", "-n", "16", "-r", "2"], env=env, capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-2000:]
    assert "loaded B200 backend" in r.stdout + r.stderr
    rows = re.findall(r"\|\s*(\w+)\s*\|\s*99\s*\|\s*(pp64|tg16)\s*\|\s*([0-9.]+)\s*±", r.stdout)
    got = {t: (be, float(v)) for be, t, v in rows}
    assert set(got) == {"pp64", "tg16"}, r.stdout[-1500:]
    assert got["pp64"][0] == "B200" and got["tg16"][0] == "B200"
    # a 4-layer model: the CPU backend does ~300 / ~60 tok/s on these shapes; far below that means ops bounced off the device
    assert got["pp64"][1] > 2000 and got["tg16"][1] > 500, got
    print(f"\n[llama-bench {ftype}, 4 layers] pp64 {got['pp64'][1]:.0f} t/s, tg16 {got['tg16'][1]:.0f} t/s")
